// The handful of packed 16-bit operations K_fill is made of, with plain C++ stand-ins so that tests/emu can run the
// kernel's lane code on the CPU (32 lane objects stepped in lock step) and compare it with the oracle without a GPU.
// On sm_100a each is one instruction (checked with cuobjdump -sass, profiles/):
//   s16x2 add                 VIADD.16x2            (FMA pipe)
//   s16x2 max / max3, relu    VIMNMX(3).S16x2.RELU  (ALU pipe, DPX)
//   fp16x2 "not equal"        HSET2.BF.NEU          (ALU pipe)      on the 16-bit PATTERNS of the scores: 0..0x7bff are
//   fp16x2 fused multiply-add HFMA2                 (FMA pipe)      finite fp16 values (denormals included, not flushed) that
//                                                                   compare like the integers; negative sums are NaN / -inf
//                                                                   patterns that can only meet a non-negative partner,
//                                                                   where "unordered or not equal" is the right answer.
#pragma once
#include <stdint.h>
#if defined(__CUDACC__)
#include <cuda_fp16.h>
#endif

namespace bt2 {

#define BT2_FL_BASE 0x64006400u                                  // fp16 1024.0 | 1024.0: the pattern of 1024 + x is 0x6400 | x
#define BT2_FL_W(log2bit) ((0x3c00u + 0x400u * (log2bit)) * 0x00010001u)   // 2^log2bit in both halves
// (hi - BASE) * 256 + (lo - BASE) for two flag accumulators: bytes [lo_A, hi_A, lo_B, hi_B] (arithmetic mod 2^32)
#define BT2_FL_PAIRFIX 0x64646400u

#if defined(__CUDA_ARCH__)
__device__ __forceinline__ __half2 as_h2(uint32_t a) { return *reinterpret_cast<const __half2*>(&a); }
__device__ __forceinline__ uint32_t as_u32(__half2 a) { return *reinterpret_cast<const uint32_t*>(&a); }
__device__ __forceinline__ uint32_t v_add2(uint32_t a, uint32_t b) { return __vadd2(a, b); }
__device__ __forceinline__ uint32_t v_max3r(uint32_t a, uint32_t b, uint32_t c) { return __vimax3_s16x2_relu(a, b, c); }
__device__ __forceinline__ uint32_t v_maxr(uint32_t a, uint32_t b) { return __vimax_s16x2_relu(a, b); }
// 0xffff in each half where the 16-bit patterns differ / are equal
__device__ __forceinline__ uint32_t v_ne2(uint32_t a, uint32_t b) { return __hneu2_mask(as_h2(a), as_h2(b)); }
__device__ __forceinline__ uint32_t v_eq2(uint32_t a, uint32_t b) { return __heq2_mask(as_h2(a), as_h2(b)); }
// acc + 2^LOG2BIT in each half where the 16-bit patterns of a and b differ (acc: fp16 pair BASE | flags so far)
template<int LOG2BIT> __device__ __forceinline__ uint32_t v_flag_ne(uint32_t a, uint32_t b, uint32_t acc) {
	return as_u32(__hfma2(__hneu2(as_h2(a), as_h2(b)), as_h2(BT2_FL_W(LOG2BIT)), as_h2(acc)));
}
__device__ __forceinline__ uint32_t v_prmt(uint32_t a, uint32_t b, uint32_t sel) { return __byte_perm(a, b, sel); }
#define BT2_SIMD __device__ __forceinline__
#else
inline int16_t h_lo(uint32_t a) { return (int16_t)(a & 0xffffu); }
inline int16_t h_hi(uint32_t a) { return (int16_t)(a >> 16); }
inline uint32_t h_pack(int lo, int hi) { return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16); }
inline uint32_t v_add2(uint32_t a, uint32_t b) { return h_pack(h_lo(a) + h_lo(b), h_hi(a) + h_hi(b)); }
inline int h_relu(int x) { return x < 0 ? 0 : x; }
inline int h_max(int a, int b) { return a > b ? a : b; }
inline uint32_t v_max3r(uint32_t a, uint32_t b, uint32_t c) {
	return h_pack(h_relu(h_max(h_max(h_lo(a), h_lo(b)), h_lo(c))), h_relu(h_max(h_max(h_hi(a), h_hi(b)), h_hi(c))));
}
inline uint32_t v_maxr(uint32_t a, uint32_t b) { return h_pack(h_relu(h_max(h_lo(a), h_lo(b))), h_relu(h_max(h_hi(a), h_hi(b)))); }
// fp16 semantics of one half: NaN compares unordered, +0 == -0
inline bool h_isnan(uint32_t x) { return (x & 0x7c00u) == 0x7c00u && (x & 0x3ffu) != 0; }
inline bool h_neu(uint32_t x, uint32_t y) {
	x &= 0xffffu; y &= 0xffffu;
	if(h_isnan(x) || h_isnan(y)) return true;
	if(((x | y) & 0x7fffu) == 0) return false;
	return x != y;
}
inline uint32_t v_ne2(uint32_t a, uint32_t b) { return (h_neu(a, b) ? 0xffffu : 0u) | (h_neu(a >> 16, b >> 16) ? 0xffff0000u : 0u); }
inline uint32_t v_eq2(uint32_t a, uint32_t b) { return ~v_ne2(a, b); }   // (no NaN operand where it is used: codes 0..7)
template<int LOG2BIT> inline uint32_t v_flag_ne(uint32_t a, uint32_t b, uint32_t acc) {
	return acc + ((h_neu(a, b) ? (1u << LOG2BIT) : 0u) | (h_neu(a >> 16, b >> 16) ? (0x10000u << LOG2BIT) : 0u));
}
inline uint32_t v_prmt(uint32_t a, uint32_t b, uint32_t sel) {
	const uint64_t w = (uint64_t)a | ((uint64_t)b << 32); uint32_t r = 0;
	for(int k = 0; k < 4; k++) r |= (uint32_t)((w >> (8 * ((sel >> (4 * k)) & 7u))) & 0xffu) << (8 * k);
	return r;
}
#define BT2_SIMD inline
#endif

BT2_SIMD uint32_t v_pack2(int lo, int hi) { return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16); }
BT2_SIMD uint32_t v_flag_pair(uint32_t lo, uint32_t hi) { return hi * 256u + lo - BT2_FL_PAIRFIX; }

} // namespace bt2
