// Measures the issue rate of the packed-integer / half2 instructions the DP fill is made of
// ("report GCUPS against a MEASURED int-pipe peak").  For each instruction: 8
// independent dependency chains per thread, 1024 threads per SM resident, all SMs; result in
// warp-instructions per clock per SM (4 = one per scheduler per clock).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o pipe_bench tools/pipe_bench.cu && ./pipe_bench
#include <cstdio>
#include <cstdint>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define CHAINS 8
#define INNER 32
#define OUTER 256

template<int OP> __device__ __forceinline__ uint32_t op(uint32_t a, uint32_t b, uint32_t c) {
	if(OP == 0) return __vadd2(a, b);                          // VIADD.16x2
	if(OP == 1) return __vmaxs2(a, b);                         // VIMNMX.S16x2
	if(OP == 2) return __vimax3_s16x2_relu(a, b, c);           // VIMNMX3.S16x2.RELU
	if(OP == 3) return __viaddmax_s16x2_relu(a, b, c);         // VIADDMNMX.S16x2.RELU
	if(OP == 4) return (a & b) ^ c;                            // LOP3
	if(OP == 5) return a + b + c;                              // IADD3
	if(OP == 6) return a * b + c;                              // IMAD
	if(OP == 7) { __half2 r = __hneu2(*(__half2*)&a, *(__half2*)&b); return *(uint32_t*)&r; }          // HSET2.BF
	if(OP == 8) { __half2 r = __hfma2(*(__half2*)&a, *(__half2*)&b, *(__half2*)&c); return *(uint32_t*)&r; } // HFMA2
	if(OP == 9) return __byte_perm(a, b, c);                   // PRMT (register selector)
	if(OP == 10) return __hneu2_mask(*(__half2*)&a, *(__half2*)&b);                                     // HSET2 (mask)
	if(OP == 11) return __shfl_sync(0xffffffffu, a, (b + 31) & 31);                                     // SHFL.IDX
	if(OP == 12) return __popc(a) + b;                         // POPC (+ add)
	if(OP == 13) return a + b;                                 // IADD (two-input add; may become IMAD.IADD)
	if(OP == 14) { __half2 r = __hadd2(*(__half2*)&a, *(__half2*)&b); return *(uint32_t*)&r; }          // HADD2
	if(OP == 15) { __half2 r = __hmax2(*(__half2*)&a, *(__half2*)&b); return *(uint32_t*)&r; }          // HMNMX2
	if(OP == 16) return __vimax_s16x2_relu(a, b);              // VIMNMX.S16x2.RELU
	if(OP == 17) { uint32_t r; asm volatile("mad.lo.u32 %0, %1, 1, %2;" : "=r"(r) : "r"(a), "r"(b)); return r; }  // IMAD used as an add
	if(OP == 18) return __float_as_uint(__uint_as_float(a) + __uint_as_float(b));                       // FADD
	if(OP == 19) return __float_as_uint(fmaxf(__uint_as_float(a), __uint_as_float(b)));                 // FMNMX
	return a;
}

template<int OP> __global__ void k(uint32_t* out, uint32_t seed) {
	uint32_t x[CHAINS];
	for(int i = 0; i < CHAINS; i++) x[i] = seed + threadIdx.x * 7 + i;
	
	for(int o = 0; o < OUTER; o++) {
		#pragma unroll
		for(int j = 0; j < INNER; j++) {
			#pragma unroll
			for(int i = 0; i < CHAINS; i++) x[i] = op<OP>(x[i], x[(i + 3) % CHAINS], x[(i + 5) % CHAINS]);
		}
	}
	uint32_t s = 0; for(int i = 0; i < CHAINS; i++) s ^= x[i];
	if(s == 0x12345678u) out[0] = s;
}

// two instructions interleaved (even chains: A, odd chains: B): the combined rate shows whether they share a pipe
template<int A, int B> __global__ void k2(uint32_t* out, uint32_t seed) {
	uint32_t x[CHAINS];
	for(int i = 0; i < CHAINS; i++) x[i] = seed + threadIdx.x * 7 + i;
	for(int o = 0; o < OUTER; o++) {
		#pragma unroll
		for(int j = 0; j < INNER; j++) {
			#pragma unroll
			for(int i = 0; i < CHAINS; i++) x[i] = (i & 1) ? op<B>(x[i], x[(i + 2) % CHAINS], x[(i + 4) % CHAINS]) : op<A>(x[i], x[(i + 2) % CHAINS], x[(i + 4) % CHAINS]);
		}
	}
	uint32_t s = 0; for(int i = 0; i < CHAINS; i++) s ^= x[i];
	if(s == 0x12345678u) out[0] = s;
}

template<int A, int B> void run2(const char* name, int nsm, double mhz) {
	uint32_t* d; cudaMalloc(&d, 4);
	cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
	k2<A, B><<<nsm * 4, 256>>>(d, 1); cudaDeviceSynchronize();
	cudaEventRecord(a);
	k2<A, B><<<nsm * 4, 256>>>(d, 2);
	cudaEventRecord(b); cudaEventSynchronize(b);
	float ms; cudaEventElapsedTime(&ms, a, b);
	const double winst = (double)nsm * 32 /*warps*/ * OUTER * INNER * CHAINS;
	const double clk = ms * 1e-3 * mhz * 1e6;
	printf("%-28s %8.3f ms  %6.2f warp-inst/clk/SM  (both instructions together)\n", name, ms, winst / clk / nsm * 1.0);
	cudaFree(d);
}

template<int OP> void run(const char* name, int nsm, double mhz) {
	uint32_t* d; cudaMalloc(&d, 4);
	cudaEvent_t a, b; cudaEventCreate(&a); cudaEventCreate(&b);
	k<OP><<<nsm * 4, 256>>>(d, 1); cudaDeviceSynchronize();
	cudaEventRecord(a);
	k<OP><<<nsm * 4, 256>>>(d, 2);
	cudaEventRecord(b); cudaEventSynchronize(b);
	float ms; cudaEventElapsedTime(&ms, a, b);
	const double winst = (double)nsm * 32 /*warps*/ * OUTER * INNER * CHAINS;
	const double clk = ms * 1e-3 * mhz * 1e6;
	printf("%-28s %8.3f ms  %6.2f warp-inst/clk/SM  (%5.1f lanes/clk/SM)\n", name, ms, winst / clk / nsm * 1.0, winst * 32 / clk / nsm);
	cudaFree(d);
}

int main() {
	cudaDeviceProp p; cudaGetDeviceProperties(&p, 0);
	int mhz = 0; cudaDeviceGetAttribute(&mhz, cudaDevAttrClockRate, 0);
	const double f = mhz / 1000.0;
	printf("%s, %d SMs, %.0f MHz (max SM clock; rates assume the run holds it)\n", p.name, p.multiProcessorCount, f);
	const int n = p.multiProcessorCount;
	run<0>("VIADD.16x2 (__vadd2)", n, f);
	run<1>("VIMNMX.S16x2 (__vmaxs2)", n, f);
	run<16>("VIMNMX.S16x2.RELU", n, f);
	run<2>("VIMNMX3.S16x2.RELU", n, f);
	run<3>("VIADDMNMX.S16x2.RELU", n, f);
	run<4>("LOP3", n, f);
	run<5>("IADD3", n, f);
	run<13>("IADD 2-input", n, f);
	run<6>("IMAD", n, f);
	run<7>("HSET2.BF.NEU", n, f);
	run<10>("HSET2.NEU (mask)", n, f);
	run<8>("HFMA2", n, f);
	run<14>("HADD2", n, f);
	run<15>("HMNMX2", n, f);
	run<9>("PRMT", n, f);
	run<11>("SHFL.IDX", n, f);
	run<12>("POPC+IADD", n, f);
	run<17>("IMAD x*1+y (mad.lo)", n, f);
	run<18>("FADD", n, f);
	run<19>("FMNMX", n, f);
	printf("-- pairs, half of the chains each\n");
	run2<1, 7>("VIMNMX + HSET2.BF", n, f);
	run2<1, 8>("VIMNMX + HFMA2", n, f);
	run2<7, 8>("HSET2.BF + HFMA2", n, f);
	run2<1, 15>("VIMNMX + HMNMX2", n, f);
	run2<8, 15>("HFMA2 + HMNMX2", n, f);
	run2<1, 14>("VIMNMX + HADD2", n, f);
	run2<8, 14>("HFMA2 + HADD2", n, f);
	run2<0, 1>("VIADD.16x2 + VIMNMX", n, f);
	run2<0, 8>("VIADD.16x2 + HFMA2", n, f);
	run2<1, 17>("VIMNMX + IMAD x*1+y", n, f);
	run2<8, 17>("HFMA2 + IMAD x*1+y", n, f);
	run2<1, 13>("VIMNMX + IADD", n, f);
	run2<8, 13>("HFMA2 + IADD", n, f);
	run2<1, 4>("VIMNMX + LOP3", n, f);
	run2<8, 4>("HFMA2 + LOP3", n, f);
	run2<1, 9>("VIMNMX + PRMT", n, f);
	run2<8, 9>("HFMA2 + PRMT", n, f);
	run2<7, 4>("HSET2.BF + LOP3", n, f);
	run2<8, 18>("HFMA2 + FADD", n, f);
	run2<1, 19>("VIMNMX + FMNMX", n, f);
	return 0;
}
