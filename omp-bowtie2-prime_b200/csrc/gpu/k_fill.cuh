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
// Strip boundaries travel down with __shfl_up_sync.  The reference keeps H, E and F of
// every cell for its backtrace; the backtrace only ever asks which of its score
// equalities hold at a cell, so the fill answers those while the operands are in
// registers and stores ONE byte of flags per cell (dev_common.cuh, CF_*): 1 B/cell of
// HBM traffic instead of 3-4.  Each equality is one HSET2.NEU on the two 16-bit halves
// (the values 0..255 and the small negative unsaturated sums are compared as fp16 bit
// patterns: denormals compare exactly, negative values are NaN/-inf patterns and can
// only meet a non-negative partner, where "unordered or not equal" is the right answer)
// plus one LOP3 that drops the mask into its bit.
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
struct FillShared {
	uint32_t tab[3 * 5 * 64];          // negated penalties, pre-positioned per half
	uint8_t  refc[BT2GPU_MAX_COLS + 8]; // reference code 0..4 per column
	uint8_t  hlast[BT2GPU_MAX_COLS + 8];
	uint8_t  rdc[64 * 3];               // read code per row (0..3, 7 = N or beyond the read)
	uint16_t ckey[BT2GPU_MAX_COLS + 8];
};

__device__ __forceinline__ uint32_t pack2(int lo, int hi) { return ((uint32_t)lo & 0xffffu) | ((uint32_t)hi << 16); }
// 0xffff in each half where the 16-bit patterns differ
__device__ __forceinline__ uint32_t ne2(uint32_t a, uint32_t b) {
	return __hneu2_mask(*reinterpret_cast<const __half2*>(&a), *reinterpret_cast<const __half2*>(&b));
}

template<int R, bool DBG>
__device__ __forceinline__ void fill_body(const DevParams& P, Problem& pb, FillShared& sh, uint8_t* __restrict__ cells,
                                          uint32_t* __restrict__ hef, uint32_t lane) {
	const int rows = pb.rows, cols = pb.cols;
	const int nv = (rows + R - 1) / R;                 // active strips
	// per-row gap penalties; inside the gap barrier (aligner_swsse_ee_u8.cpp:101-115: first/last gapbar
	// rows) they are large enough to keep E and F at 0
	const uint32_t neg_rdge = pack2(-P.rdGapExt, -P.rdGapExt);
	uint32_t neg_rdgo[R], neg_rfgo[R], neg_rfge[R];    // [k]: read gaps in row k; ref gaps INTO row k (from the row above)
	#pragma unroll
	for(int k = 0; k < R; k++) {
		const int rlo = (int)lane * R + k, rhi = (32 + (int)lane) * R + k;
		const bool alo = !(rlo < P.gapbar || rows - 1 - rlo < P.gapbar), ahi = !(rhi < P.gapbar || rows - 1 - rhi < P.gapbar);
		neg_rdgo[k] = pack2(alo ? -P.rdGapOpen : -1024, ahi ? -P.rdGapOpen : -1024);
		neg_rfgo[k] = pack2(alo ? -P.rfGapOpen : -1024, ahi ? -P.rfGapOpen : -1024);
		neg_rfge[k] = pack2(alo ? -P.rfGapExt : -1024, ahi ? -P.rfGapExt : -1024);
	}
	uint32_t nx_rfgo, nx_rfge;                         // ref gaps into the first row of the next lane's strips
	{
		const int rlo = ((int)lane + 1) * R, rhi = (33 + (int)lane) * R;
		const bool alo = !(rlo < P.gapbar || rows - 1 - rlo < P.gapbar), ahi = !(rhi < P.gapbar || rows - 1 - rhi < P.gapbar);
		nx_rfgo = pack2(alo ? -P.rfGapOpen : -1024, ahi ? -P.rfGapOpen : -1024);
		nx_rfge = pack2(alo ? -P.rfGapExt : -1024, ahi ? -P.rfGapExt : -1024);
	}
	uint32_t rdc2[R];
	#pragma unroll
	for(int k = 0; k < R; k++) rdc2[k] = (uint32_t)sh.rdc[(int)lane * R + k] | ((uint32_t)sh.rdc[(32 + (int)lane) * R + k] << 16);
	const bool strip_lo = (int)lane < nv, strip_hi = 32 + (int)lane < nv;
	// which (lane, half, k) owns the last row
	const int vstar = (rows - 1) / R, kstar = (rows - 1) - vstar * R;
	const bool own_lo = vstar == (int)lane, own_hi = vstar == 32 + (int)lane;
	uint32_t Hp[R], Ep[R], Eflag[R];
	#pragma unroll
	for(int k = 0; k < R; k++) { Hp[k] = 0; Ep[k] = 0; Eflag[k] = (CF_EO | CF_EX) * 0x00010001u; }
	// sendF: F for the row below in bits 0-7 of each half, the stored F of this strip's last row in bits 8-15
	uint32_t sendH = 0, sendF = 0, prevRecvH = (lane == 0) ? 0x000000ffu : 0u;
	// A word of the cell array holds the R cells of a strip on one diagonal (dev_common.cuh): byte k comes from
	// the step R-1-k before the one that completes it, so the flag bytes pass through a short register pipeline
	// (pipe1: row 0 of the previous step; pipe2: rows 0,1 of the two previous steps) and the sweep runs R-1 steps longer.
	const int nsteps = cols + nv - 1 + (R - 1);
	uint32_t pipe1 = 0, pipe2 = 0;
	uint4* const out = reinterpret_cast<uint4*>(cells);
	const int g_lo = BT2_CELL_DOFF / 4 - (int)((lane * (R + 1) + R - 1) >> 2), g_hi = BT2_CELL_DOFF / 4 - (int)(((32 + lane) * (R + 1) + R - 1) >> 2);
	for(int t0 = 0; t0 < nsteps; t0 += 4) {
	uint32_t accl[4], acch[4];
	#pragma unroll
	for(int tj = 0; tj < 4; tj++) {
		const int t = t0 + tj;
		const int c_lo = t - (int)lane, c_hi = c_lo - 32;
		const bool act_lo = strip_lo && c_lo >= 0 && c_lo < cols, act_hi = strip_hi && c_hi >= 0 && c_hi < cols;
		const uint32_t am = (act_lo ? 0xffffu : 0u) | (act_hi ? 0xffff0000u : 0u);
		// values produced by the strip above in the previous step
		uint32_t recvH = __shfl_up_sync(0xffffffffu, sendH, 1), recvF = __shfl_up_sync(0xffffffffu, sendF, 1);
		const uint32_t wrapH = __shfl_sync(0xffffffffu, sendH, 31), wrapF = __shfl_sync(0xffffffffu, sendF, 31);
		if(lane == 0) { recvH = 0x00ffu | (wrapH << 16); recvF = (wrapF << 16); } // row 0: H(-1,*) = 0xff, F = 0
		const uint32_t x_lo = act_lo ? sh.refc[c_lo] : 4u, x_hi = act_hi ? sh.refc[c_hi] : 4u;
		const uint32_t xpk = x_lo | (x_hi << 16);
		uint32_t diag = prevRecvH;
		uint32_t f = recvF & 0x00ff00ffu;                 // F(r,c) before the barrier rule (already 0 there)
		uint32_t hup = recvH;                             // H(r-1,c)
		uint32_t fup = (recvF >> 8) & 0x00ff00ffu;        // F(r-1,c) as stored
		uint32_t flg[R];
		uint32_t wlo[R], whi[R];
		#pragma unroll
		for(int k = 0; k < R; k++) {
			const uint32_t s = sh.tab[(k * 5 + x_lo) * 64 + lane] | sh.tab[(k * 5 + x_hi) * 64 + 32 + lane];
			const uint32_t hd = diag; diag = Hp[k];
			const uint32_t u = __vadd2(hd, s);            // unsaturated H(r-1,c-1) + sc
			const uint32_t e = Ep[k];
			const uint32_t h = __vimax3_s16x2_relu(u, e, f);
			// the comparisons of the backtrace at this cell
			const uint32_t bf = __vadd2(hup, neg_rfgo[k]), af = __vadd2(fup, neg_rfge[k]);
			uint32_t fl = Eflag[k];
			fl |= ne2(h, u) & (CF_D * 0x00010001u);
			fl |= ne2(h, e) & (CF_HE * 0x00010001u);
			fl |= ne2(h, f) & (CF_HF * 0x00010001u);
			fl |= ne2(f, bf) & (CF_FO * 0x00010001u);
			fl |= ne2(f, af) & (CF_FX * 0x00010001u);
			fl |= ne2(rdc2[k], xpk) & (CF_MM * 0x00010001u);
			flg[k] = fl;
			if(DBG) {
				wlo[k] = __byte_perm(__byte_perm(h, e, 0x0040), f, 0x5410);
				whi[k] = __byte_perm(__byte_perm(h, e, 0x0062), f, 0x7610);
			}
			if(k == kstar) {
				if(own_lo && act_lo) sh.hlast[c_lo] = (uint8_t)(h & 0xffu);
				if(own_hi && act_hi) sh.hlast[c_hi] = (uint8_t)((h >> 16) & 0xffu);
			}
			// E of the next column in this row and its two comparisons
			const uint32_t be = __vadd2(h, neg_rdgo[k]), ae = __vadd2(e, neg_rdge);
			const uint32_t en = __vimax3_s16x2_relu(ae, be, 0u);
			Eflag[k] = (ne2(en, be) & (CF_EO * 0x00010001u)) | (ne2(en, ae) & (CF_EX * 0x00010001u));
			Hp[k] = h & am; Ep[k] = en & am;
			// F of the next row in this column
			hup = h; fup = f;
			f = __vimax3_s16x2_relu(__vadd2(f, k + 1 < R ? neg_rfge[(k + 1) % R] : nx_rfge), __vadd2(h, k + 1 < R ? neg_rfgo[(k + 1) % R] : nx_rfgo), 0u);
		}
		{ // the strip's R flag bytes of this column as one word; four steps make one 16-byte store
			uint32_t blo, bhi;
			if(R == 3) {
				blo = __byte_perm(pipe2, flg[2 % R], 0x4410); bhi = __byte_perm(pipe2, flg[2 % R], 0x6632);
				pipe2 = __byte_perm(pipe1, flg[1 % R], 0x6240); pipe1 = flg[0];
			} else if(R == 2) {
				blo = __byte_perm(pipe1, flg[1 % R], 0x4440); bhi = __byte_perm(pipe1, flg[1 % R], 0x6662);
				pipe1 = flg[0];
			} else { blo = flg[0] & 0xffu; bhi = flg[0] >> 16; }
			accl[tj] = blo; acch[tj] = bhi;
			if(DBG) {
				uint32_t* dlo = hef + ((size_t)lane * BT2_HEF_T + (uint32_t)t) * 4u;
				uint32_t* dhi = hef + ((size_t)(32 + lane) * BT2_HEF_T + (uint32_t)t) * 4u;
				#pragma unroll
				for(int k = 0; k < R; k++) { if(act_lo) dlo[k] = wlo[k]; if(act_hi) dhi[k] = whi[k]; }
			}
		}
		sendH = Hp[R - 1]; sendF = (f | (fup << 8)) & am;
		prevRecvH = (lane == 0) ? (0x00ffu | (recvH & 0xffff0000u)) : recvH;
	}
	// strip v is active at steps v .. v+cols-1 and its last words complete R-1 steps later
	if(strip_lo && t0 + 3 >= (int)lane && t0 - (R - 1) < (int)lane + cols) out[(g_lo + (t0 >> 2)) * 64 + (int)lane] = make_uint4(accl[0], accl[1], accl[2], accl[3]);
	if(strip_hi && t0 + 3 >= 32 + (int)lane && t0 - (R - 1) < 32 + (int)lane + cols) out[(g_hi + (t0 >> 2)) * 64 + 32 + (int)lane] = make_uint4(acch[0], acch[1], acch[2], acch[3]);
	}
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
	// query profile: tab[(k*5+x)*64 + v] = -(penalty of row v*R+k against reference code x), positioned in its half
	for(int idx = (int)lane; idx < R * 64; idx += 32) {
		const int k = idx / 64, v = idx - k * 64;
		const int r = v * R + k;
		int rdc = 4, q = 0;
		if(r < rows) {
			if(fw) { rdc = rseq[r]; q = rqual[r]; }
			else { const int c0 = rseq[rows - 1 - r]; rdc = c0 > 3 ? 4 : 3 - c0; q = rqual[rows - 1 - r]; }
			q -= 33; if(q < 0) q = 0; if(q > 255) q = 255;
		}
		sh.rdc[r] = (uint8_t)(rdc > 3 ? 7 : rdc);
		#pragma unroll
		for(int x = 0; x < 5; x++) {
			int pen = 0;
			if(r < rows) pen = (x == 4 || rdc > 3) ? (int)P.npens[q] : (x == rdc ? 0 : (int)P.mmpens[q]);
			const uint32_t neg = (uint32_t)(-pen) & 0xffffu;
			sh.tab[(k * 5 + x) * 64 + v] = v < 32 ? neg : (neg << 16);
		}
	}
	for(int c = (int)lane; c < cols; c += 32) {
		const uint32_t m = pb.rf[c];
		sh.refc[c] = (uint8_t)(m > 15 ? 4 : (31 - __clz(m)));
		sh.hlast[c] = 0;
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
	if(R == 1) fill_body<1, DBG>(P, pb, sh, cells, hef, lane);
	else if(R == 2) fill_body<2, DBG>(P, pb, sh, cells, hef, lane);
	else fill_body<3, DBG>(P, pb, sh, cells, hef, lane);
	__syncwarp();
	// ---- last-row candidates: H-0xff >= minsc, sorted by (score desc, col desc)
	const int minsc = pb.minsc;
	int lrmax = 0; uint32_t ncand = 0;
	for(int c0 = 0; c0 < cols; c0 += 32) {
		const int c = c0 + (int)lane;
		const int hv = c < cols ? (int)sh.hlast[c] : 0;
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
