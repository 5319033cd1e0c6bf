// K_fill: the end-to-end affine-gap DP fill of one read-window problem per warp.
//   SwAligner::alignEnd2EndSseU8 / EEU8_alignNucleotides / EEU8_alignOne / EEU8_lazyF
//       aligner_swsse_ee_u8.cpp:250-671   (recurrence, unsigned saturating 8-bit, start 0xff)
//   SwAligner::buildQueryProfileEnd2EndSseU8   aligner_swsse_ee_u8.cpp:71-134
//   candidate gathering + DpBtCandidate order  aligner_swsse_ee_u8.cpp:600-671, aligner_sw_nuc.h
//
// The reference computes the full rdlen x rflen rectangle in a Farrar-striped SSE2
// layout; results do not depend on the evaluation order, so here the rectangle is
// swept as a wavefront: the rows are cut into 64 strips of R = ceil(rdlen/64) rows,
// strip v works on column t-v at step t.  Two strips share a 32-bit register as two
// signed 16-bit lanes (strip L in the low half, strip 32+L in the high half of lane L),
// because on sm_100a the 16x2 DPX instructions (VIADDMNMX.S16x2, VIMNMX3.S16x2) are
// native while the 8x4 video intrinsics (__vmaxu4/__vsubus4) are emulated with 4-5
// LOP3/PRMT each (SASS evidence).  u8 saturation is exact in 16 bits:
// subs_u8(a,b) == max(a-b,0) for 0<=a,b<=255.
// Strip boundaries travel down with one rotating __shfl_sync per value.
//
// The reference keeps H, E and F of every cell for its backtrace; the backtrace only ever
// asks which of its score equalities hold at a cell, so the fill answers those while the
// operands are in registers and stores ONE byte of flags per cell (dev_common.cuh, CF_*).
// The kernel is bound by instruction issue, with the integer ALU pipe (half rate) next, so
// the flag work is kept off the ALU pipe: each equality is one HSET2.BF.NEU on the two 16-bit
// halves (values 0..255 and the small negative unsaturated sums compared as fp16 bit patterns:
// denormals compare exactly, negative values are NaN/-inf patterns and can only meet a
// non-negative partner, where "unordered or not equal" is the right answer) giving 1.0 / 0.0,
// and one HFMA2 that adds the flag's weight to an accumulator that starts at 1024.0: the fp16
// pattern of 1024+x is 0x6400|x, so the low byte of each half IS the flag byte, partial
// accumulators merge with a plain OR, and no conversion is ever needed.
// The substitution score needs no table either: score = mismatch mask AND the row's penalty
// (a second penalty when the reference character is N; that variant of the loop is only
// run for windows that contain an N).
#pragma once
#include <cuda_fp16.h>
#include "dev_common.cuh"

namespace bt2 {

struct FillArgs {
	Problem*         probs;     // [nslots]
	uint32_t         nprob;
	const ReadState* states;
	const uint8_t*   seq;
	const uint8_t*   qual;
	uint8_t*         cells;     // [nslots * BT2_CELL_BYTES]
	uint32_t*        pcache;    // [nslots * BT2_PC_WORDS] character cache for the backtrace, or null
	uint32_t*        hef_dbg;   // test hook only: [nslots * BT2_HEF_WORDS] H|E<<8|F<<16, or null
	unsigned long long* stats;  // [0] problems [1] cells
};

#define BT2_FILL_WARPS 4
#define BT2_FILL_REFPAD 80             // columns read past the window's end by strips that are done (<= 63 + R)
#define BT2_FILL_REFOFF 32             // columns before the window's start asked for by strips that have not started
struct FillShared {
	uint8_t  refc[BT2GPU_MAX_COLS + 8 + BT2_FILL_REFPAD]; // reference code 0..4 per column, 4 beyond the window
	uint32_t refw[BT2_FILL_REFOFF + BT2GPU_MAX_COLS + 8 + BT2_FILL_REFPAD]; // [REFOFF + c] = code(c) | code(c-32) << 16, 4 outside the window
	uint32_t hl[BT2GPU_MAX_COLS + 64 + 8];  // [step]: H of the strip that owns the last row (both halves of its lane), step = col + strip
	uint8_t  rdc[64 * 3];               // read code per row (0..3, 7 = N or beyond the read)
	uint8_t  pen[64 * 3], penN[64 * 3]; // penalty of the row against a mismatching / an N reference character
	uint16_t ckey[BT2GPU_MAX_COLS + 8];
};

__device__ __forceinline__ uint32_t pack2(int lo, int hi) { return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16); }
__device__ __forceinline__ __half2 as_h2(uint32_t a) { return *reinterpret_cast<const __half2*>(&a); }
__device__ __forceinline__ uint32_t as_u32(__half2 a) { return *reinterpret_cast<const uint32_t*>(&a); }
// 0xffff in each half where the 16-bit patterns differ
__device__ __forceinline__ uint32_t ne2(uint32_t a, uint32_t b) { return __hneu2_mask(as_h2(a), as_h2(b)); }
// acc + w in each half where the 16-bit patterns of a and b differ (acc, w: fp16 pairs)
__device__ __forceinline__ uint32_t flag_ne(uint32_t a, uint32_t b, uint32_t w, uint32_t acc) {
	return as_u32(__hfma2(__hneu2(as_h2(a), as_h2(b)), as_h2(w), as_h2(acc)));
}
#define BT2_FL_BASE 0x64006400u                                  // 1024.0 | 1024.0
#define BT2_FL_W(log2bit) ((0x3c00u + 0x400u * (log2bit)) * 0x00010001u)   // 2^log2bit in both halves

// The steps [t0, t_end) of the sweep, four per iteration (one 16-byte store per lane and half).
// MASK: some strips have not reached column 0 yet (steps < 64): their H/E must stay 0.  REFN: the window contains an N.
template<int R, int KS, bool DBG, bool MASK, bool REFN>
__device__ __forceinline__ void fill_steps(int t0, const int t_end, const int cols, const uint32_t lane, FillShared& sh,
		const bool strip_lo, const bool strip_hi, const bool own_any,
		const uint32_t (&rdc2)[R], const uint32_t (&negpen)[R], const uint32_t (&negpenN)[R],
		const uint32_t (&neg_rdgo)[R], const uint32_t (&neg_rfgo)[R], const uint32_t (&neg_rfge)[R], const uint32_t neg_rdge,
		const uint32_t nx_rfgo, const uint32_t nx_rfge,
		uint32_t (&Hp)[R], uint32_t (&Ep)[R], uint32_t (&Ecar)[R], uint32_t& sendH, uint32_t& sendF, uint32_t& prevRecvH,
		uint32_t& pipe1, uint32_t& pipe2, uint4* __restrict__ out, const int g_lo, const int g_hi, uint32_t* __restrict__ hef) {
	const uint32_t W_D = BT2_FL_W(0), W_HE = BT2_FL_W(1), W_HF = BT2_FL_W(2), W_EO = BT2_FL_W(3), W_EX = BT2_FL_W(4), W_FO = BT2_FL_W(5), W_FX = BT2_FL_W(6);
	const uint32_t rot = (lane + 31u) & 31u;
	for(; t0 < t_end; t0 += 4) {
		uint32_t accl[4], acch[4];
		#pragma unroll
		for(int tj = 0; tj < 4; tj++) {
			const int t = t0 + tj;
			const int c_lo = t - (int)lane, c_hi = c_lo - 32;
			// values produced by the strip above in the previous step (lane 0: the strip above its high half is lane 31's low half)
			uint32_t recvH = __shfl_sync(0xffffffffu, sendH, rot), recvF = __shfl_sync(0xffffffffu, sendF, rot);
			if(lane == 0) { recvH = 0x00ffu | (recvH << 16); recvF = ((uint32_t)(CF_FO | CF_FX) << 8) | (recvF << 16); } // row 0: H(-1,*) = 0xff, F = 0
			uint32_t am = 0xffffffffu;
			bool act_lo = true, act_hi = true;
			if(MASK) {
				act_lo = strip_lo && c_lo >= 0 && c_lo < cols; act_hi = strip_hi && c_hi >= 0 && c_hi < cols;
				am = (act_lo ? 0xffffu : 0u) | (act_hi ? 0xffff0000u : 0u);
			}
			const uint32_t xpk = sh.refw[BT2_FILL_REFOFF + c_lo];  // reference characters of the two halves' columns
			const uint32_t nmask = REFN ? __heq2_mask(as_h2(xpk), as_h2(0x00040004u)) : 0u;
			uint32_t diag = prevRecvH;
			uint32_t f = recvF & 0x00ff00ffu;                              // F(r,c)
			// the cell's flag accumulator: the two comparisons of its E (made one step ago) and of its F (made one row up)
			uint32_t car = __byte_perm(recvF, 0u, 0x4341) | Ecar[0];
			uint32_t flg[R];
			uint32_t wlo[R], whi[R];
			#pragma unroll
			for(int k = 0; k < R; k++) {
				const uint32_t mm = ne2(rdc2[k], xpk);
				const uint32_t pen = REFN ? ((negpen[k] & ~nmask) | (negpenN[k] & nmask)) : negpen[k];
				const uint32_t s = mm & pen;
				const uint32_t hd = diag; diag = Hp[k];
				const uint32_t u = __vadd2(hd, s);            // unsaturated H(r-1,c-1) + sc
				const uint32_t e = Ep[k];
				const uint32_t h = __vimax3_s16x2_relu(u, e, f);
				// the comparisons of the backtrace at this cell
				uint32_t fl = car | (mm & (CF_MM * 0x00010001u));
				fl = flag_ne(h, u, W_D, fl);
				fl = flag_ne(h, e, W_HE, fl);
				fl = flag_ne(h, f, W_HF, fl);
				flg[k] = fl;
				if(DBG) {
					wlo[k] = __byte_perm(__byte_perm(h, e, 0x0040), f, 0x5410);
					whi[k] = __byte_perm(__byte_perm(h, e, 0x0062), f, 0x7610);
				}
				// E of the next column in this row and its two comparisons
				const uint32_t be = __vadd2(h, neg_rdgo[k]), ae = __vadd2(e, neg_rdge);
				const uint32_t en = __vimax_s16x2_relu(ae, be);
				Ecar[k] = flag_ne(en, be, W_EO, flag_ne(en, ae, W_EX, BT2_FL_BASE));
				if(MASK) { Hp[k] = h & am; Ep[k] = en & am; } else { Hp[k] = h; Ep[k] = en; }
				// F of the next row in this column and its two comparisons
				const uint32_t bf = __vadd2(h, k + 1 < R ? neg_rfgo[(k + 1) % R] : nx_rfgo), af = __vadd2(f, k + 1 < R ? neg_rfge[(k + 1) % R] : nx_rfge);
				f = __vimax_s16x2_relu(af, bf);
				car = flag_ne(f, bf, W_FO, flag_ne(f, af, W_FX, k + 1 < R ? Ecar[(k + 1) % R] : BT2_FL_BASE));
			}
			if(own_any) sh.hl[t] = Hp[KS];   // the last row is row KS of one half of one lane (the reader picks the half)
			{ // a word of the cell array = the strip's R cells on one diagonal: byte k comes from the step R-1-k before this one
				uint32_t blo, bhi;
				if(R == 3) {
					blo = __byte_perm(pipe2, flg[2 % R], 0x4410); bhi = __byte_perm(pipe2, flg[2 % R], 0x6632);
					pipe2 = __byte_perm(pipe1, flg[1 % R], 0x6240); pipe1 = flg[0];
				} else if(R == 2) {
					blo = __byte_perm(pipe1, flg[1 % R], 0x4440); bhi = __byte_perm(pipe1, flg[1 % R], 0x6662);
					pipe1 = flg[0];
				} else { blo = flg[0] & 0xffu; bhi = (flg[0] >> 16) & 0xffu; }
				accl[tj] = blo; acch[tj] = bhi;
				if(DBG) {
					uint32_t* dlo = hef + ((size_t)lane * BT2_HEF_T + (uint32_t)t) * 4u;
					uint32_t* dhi = hef + ((size_t)(32 + lane) * BT2_HEF_T + (uint32_t)t) * 4u;
					#pragma unroll
					for(int k = 0; k < R; k++) { if(act_lo) dlo[k] = wlo[k]; if(act_hi) dhi[k] = whi[k]; }
				}
			}
			sendH = Hp[R - 1];
			sendF = __byte_perm(f, car, 0x6240);   // F in the low byte of each half, its two flags in the high byte
			if(MASK) sendF &= am;
			prevRecvH = recvH;
		}
		// strip v is active at steps v .. v+cols-1 and its last words complete R-1 steps later
		if(strip_lo && (!MASK || t0 + 3 >= (int)lane) && t0 - (R - 1) < (int)lane + cols) out[(g_lo + (t0 >> 2)) * 64 + (int)lane] = make_uint4(accl[0], accl[1], accl[2], accl[3]);
		if(strip_hi && (!MASK || t0 + 3 >= 32 + (int)lane) && t0 - (R - 1) < 32 + (int)lane + cols) out[(g_hi + (t0 >> 2)) * 64 + 32 + (int)lane] = make_uint4(acch[0], acch[1], acch[2], acch[3]);
	}
}

template<int R, int KS, bool DBG>
__device__ __forceinline__ void fill_body(const DevParams& P, Problem& pb, FillShared& sh, uint8_t* __restrict__ cells,
                                          uint32_t* __restrict__ hef, uint32_t lane, bool refHasN) {
	const int rows = pb.rows, cols = pb.cols;
	const int nv = (rows + R - 1) / R;                 // active strips
	// per-row gap penalties; inside the gap barrier (aligner_swsse_ee_u8.cpp:101-115: first/last gapbar
	// rows) they are large enough to keep E and F at 0
	const uint32_t neg_rdge = pack2(-P.rdGapExt, -P.rdGapExt);
	uint32_t neg_rdgo[R], neg_rfgo[R], neg_rfge[R];    // [k]: read gaps in row k; ref gaps INTO row k (from the row above)
	uint32_t rdc2[R], negpen[R], negpenN[R];           // read character and (negated) penalties of row k, both halves
	#pragma unroll
	for(int k = 0; k < R; k++) {
		const int rlo = (int)lane * R + k, rhi = (32 + (int)lane) * R + k;
		const bool alo = !(rlo < P.gapbar || rows - 1 - rlo < P.gapbar), ahi = !(rhi < P.gapbar || rows - 1 - rhi < P.gapbar);
		neg_rdgo[k] = pack2(alo ? -P.rdGapOpen : -1024, ahi ? -P.rdGapOpen : -1024);
		neg_rfgo[k] = pack2(alo ? -P.rfGapOpen : -1024, ahi ? -P.rfGapOpen : -1024);
		neg_rfge[k] = pack2(alo ? -P.rfGapExt : -1024, ahi ? -P.rfGapExt : -1024);
		rdc2[k] = (uint32_t)sh.rdc[rlo] | ((uint32_t)sh.rdc[rhi] << 16);
		negpen[k] = pack2(-(int)sh.pen[rlo], -(int)sh.pen[rhi]);
		negpenN[k] = pack2(-(int)sh.penN[rlo], -(int)sh.penN[rhi]);
	}
	uint32_t nx_rfgo, nx_rfge;                         // ref gaps into the first row of the next lane's strips
	{
		const int rlo = ((int)lane + 1) * R, rhi = (33 + (int)lane) * R;
		const bool alo = !(rlo < P.gapbar || rows - 1 - rlo < P.gapbar), ahi = !(rhi < P.gapbar || rows - 1 - rhi < P.gapbar);
		nx_rfgo = pack2(alo ? -P.rfGapOpen : -1024, ahi ? -P.rfGapOpen : -1024);
		nx_rfge = pack2(alo ? -P.rfGapExt : -1024, ahi ? -P.rfGapExt : -1024);
	}
	const bool strip_lo = (int)lane < nv, strip_hi = 32 + (int)lane < nv;
	// which (lane, half, k) owns the last row
	const int vstar = (rows - 1) / R;                  // (rows - 1) % R == KS
	const bool own_any = (vstar & 31) == (int)lane;
	uint32_t Hp[R], Ep[R], Ecar[R];
	#pragma unroll
	for(int k = 0; k < R; k++) { Hp[k] = 0; Ep[k] = 0; Ecar[k] = BT2_FL_BASE | ((CF_EO | CF_EX) * 0x00010001u); }
	// sendF: F for the row below in bits 0-7 of each half, its two comparison flags in bits 8-15
	uint32_t sendH = 0, sendF = 0, prevRecvH = (lane == 0) ? 0x000000ffu : 0u;
	// A word of the cell array holds the R cells of a strip on one diagonal (dev_common.cuh): byte k comes from
	// the step R-1-k before the one that completes it, so the flag bytes pass through a short register pipeline
	// (pipe1: row 0 of the previous step; pipe2: rows 0,1 of the two previous steps) and the sweep runs R-1 steps longer.
	const int nsteps = cols + nv - 1 + (R - 1);
	uint32_t pipe1 = 0, pipe2 = 0;
	uint4* const out = reinterpret_cast<uint4*>(cells);
	const int g_lo = BT2_CELL_DOFF / 4 - (int)((lane * (R + 1) + R - 1) >> 2), g_hi = BT2_CELL_DOFF / 4 - (int)(((32 + lane) * (R + 1) + R - 1) >> 2);
	const int t_mask = DBG ? nsteps : (nsteps < 64 ? nsteps : 64);   // every strip has started after step 63
	#define BT2_FILL_ARGS cols, lane, sh, strip_lo, strip_hi, own_any, rdc2, negpen, negpenN, neg_rdgo, neg_rfgo, neg_rfge, neg_rdge, \
		nx_rfgo, nx_rfge, Hp, Ep, Ecar, sendH, sendF, prevRecvH, pipe1, pipe2, out, g_lo, g_hi, hef
	if(refHasN) {
		fill_steps<R, KS, DBG, true, true>(0, t_mask, BT2_FILL_ARGS);
		if(!DBG) fill_steps<R, KS, false, false, true>((t_mask + 3) & ~3, nsteps, BT2_FILL_ARGS);
	} else {
		fill_steps<R, KS, DBG, true, false>(0, t_mask, BT2_FILL_ARGS);
		if(!DBG) fill_steps<R, KS, false, false, false>((t_mask + 3) & ~3, nsteps, BT2_FILL_ARGS);
	}
	#undef BT2_FILL_ARGS
}

template<bool DBG>
__global__ void __launch_bounds__(BT2_FILL_WARPS * 32) k_fill(const DevParams* __restrict__ Pp, const FillArgs a) {
	__shared__ FillShared shs[BT2_FILL_WARPS];
	const DevParams& P = *Pp;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t slot = blockIdx.x * BT2_FILL_WARPS + warp;
	if(slot >= a.nprob) return;
	Problem& pb = a.probs[slot];
	if(pb.read == BT2_NIL) return;
	FillShared& sh = shs[warp];
	const ReadState& rs = a.states[pb.read];
	const int rows = pb.rows, cols = pb.cols, R = pb.R;
	const bool fw = pb.fw != 0;
	const uint8_t* rseq = a.seq + rs.seq_off; const uint8_t* rqual = a.qual + rs.seq_off;
	// query profile: per row its character and its penalty against a mismatching / an N reference character
	for(int r = (int)lane; r < R * 64; r += 32) {
		int rdc = 4, q = 0;
		if(r < rows) {
			if(fw) { rdc = rseq[r]; q = rqual[r]; }
			else { const int c0 = rseq[rows - 1 - r]; rdc = c0 > 3 ? 4 : 3 - c0; q = rqual[rows - 1 - r]; }
			q -= 33; if(q < 0) q = 0; if(q > 255) q = 255;
		}
		sh.rdc[r] = (uint8_t)(rdc > 3 ? 7 : rdc);
		sh.pen[r] = r < rows ? (rdc > 3 ? P.npens[q] : P.mmpens[q]) : (uint8_t)0;
		sh.penN[r] = r < rows ? P.npens[q] : (uint8_t)0;
	}
	bool hasN = false;
	for(int c = (int)lane; c < cols + BT2_FILL_REFPAD; c += 32) {
		uint32_t code = 4;
		if(c < cols) { const uint32_t m = pb.rf[c]; code = m > 15 ? 4u : (uint32_t)(31 - __clz(m)); hasN = hasN || code == 4u; }
		sh.refc[c] = (uint8_t)code;
	}
	const bool refHasN = __any_sync(0xffffffffu, hasN);
	__syncwarp();
	for(int i = (int)lane; i < BT2_FILL_REFOFF + cols + BT2_FILL_REFPAD; i += 32) {
		const int c = i - BT2_FILL_REFOFF;
		const uint32_t lo = c >= 0 ? (uint32_t)sh.refc[c] : 4u, hi = c >= 32 ? (uint32_t)sh.refc[c - 32] : 4u;
		sh.refw[i] = lo | (hi << 16);
	}
	__syncwarp();
	if(a.pcache) { // the problem's characters as 4-bit codes, for the backtrace's edits (dev_common.cuh)
		uint32_t* pcw = a.pcache + (size_t)slot * BT2_PC_WORDS;
		uint32_t wr = 0, wf = 0;
		#pragma unroll
		for(int i = 0; i < 8; i++) {
			const int r = (int)lane * 8 + i, c = (int)lane * 8 + i;
			const uint32_t rc = r < rows ? (uint32_t)sh.rdc[r] : 0u;
			wr |= (rc > 3u ? 4u : rc) << (4 * i);
			wf |= (c < cols ? (uint32_t)sh.refc[c] : 0u) << (4 * i);
		}
		if(lane < BT2_PC_RD) pcw[lane] = wr;
		if(lane < BT2_PC_RF) pcw[BT2_PC_RD + lane] = wf;
	}
	uint8_t* cells = a.cells + (size_t)slot * BT2_CELL_BYTES;
	uint32_t* hef = DBG ? a.hef_dbg + (size_t)slot * BT2_HEF_WORDS : nullptr;
	const int vstar = (rows - 1) / R, kstar = (rows - 1) - vstar * R;   // the strip and the row within it that hold the last row
	if(R == 1) fill_body<1, 0, DBG>(P, pb, sh, cells, hef, lane, refHasN);
	else if(R == 2) { if(kstar == 0) fill_body<2, 0, DBG>(P, pb, sh, cells, hef, lane, refHasN); else fill_body<2, 1, DBG>(P, pb, sh, cells, hef, lane, refHasN); }
	else if(kstar == 0) fill_body<3, 0, DBG>(P, pb, sh, cells, hef, lane, refHasN);
	else if(kstar == 1) fill_body<3, 1, DBG>(P, pb, sh, cells, hef, lane, refHasN);
	else fill_body<3, 2, DBG>(P, pb, sh, cells, hef, lane, refHasN);
	__syncwarp();
	// ---- last-row candidates: H-0xff >= minsc, sorted by (score desc, col desc)
	const int minsc = pb.minsc;
	int lrmax = 0; uint32_t ncand = 0;
	for(int c0 = 0; c0 < cols; c0 += 32) {
		const int c = c0 + (int)lane;
		const int hv = c < cols ? (int)((sh.hl[c + vstar] >> (vstar & 32 ? 16 : 0)) & 0xffu) : 0;
		lrmax = max(lrmax, hv);
		const bool isc = c < cols && (hv - 255) >= minsc;
		const uint32_t bal = __ballot_sync(0xffffffffu, isc);
		if(isc) sh.ckey[ncand + __popc(bal & ((1u << lane) - 1))] = (uint16_t)((hv << 8) | c);
		ncand += __popc(bal);
	}
	#pragma unroll
	for(int o = 16; o > 0; o >>= 1) lrmax = max(lrmax, __shfl_xor_sync(0xffffffffu, lrmax, o));
	__syncwarp();
	for(uint32_t j = lane; j < ncand; j += 32) {
		const uint32_t key = sh.ckey[j];
		uint32_t rank = 0;
		for(uint32_t i = 0; i < ncand; i++) rank += (sh.ckey[i] > key) ? 1u : 0u;
		pb.cand[rank] = (uint16_t)key;
	}
	if(lane == 0) {
		pb.lrmax = lrmax; pb.ncand = ncand;
		atomicAdd(&a.stats[0], 1ull); atomicAdd(&a.stats[1], (unsigned long long)rows * cols);
	}
}

} // namespace bt2
