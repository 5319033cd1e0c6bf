// K_fill: the end-to-end affine-gap DP fill, two read-window problems per warp.
//   SwAligner::alignEnd2EndSseU8 / EEU8_alignNucleotides / EEU8_alignOne / EEU8_lazyF
//       aligner_swsse_ee_u8.cpp:250-671   (recurrence, unsigned saturating 8-bit, start 0xff)
//   SwAligner::buildQueryProfileEnd2EndSseU8   aligner_swsse_ee_u8.cpp:71-134
//   candidate gathering + DpBtCandidate order  aligner_swsse_ee_u8.cpp:600-671, aligner_sw_nuc.h
//
// The reference computes the full rdlen x rflen rectangle in a Farrar-striped SSE2
// layout; results do not depend on the evaluation order, so here the rectangle is
// swept as a wavefront: the rows are cut into 32 strips of R = ceil(rdlen/32) rows (R <= 5),
// strip v (lane v) works on column t-v at step t.  A 32-bit register holds two signed 16-bit
// lanes: the low half belongs to problem A, the high half to problem B of the warp's pair, so
// one instruction stream fills two rectangles (both halves see the same strip index, the same
// column index and the same neighbour lane).  On sm_100a the 16x2 DPX instructions
// (VIADDMNMX.S16x2, VIMNMX3.S16x2) are native while the 8x4 video intrinsics
// (__vmaxu4/__vsubus4) are emulated with 4-5 LOP3/PRMT each (SASS evidence).
// u8 saturation is exact in 16 bits: subs_u8(a,b) == max(a-b,0) for 0<=a,b<=255.
// Strip boundaries travel down with one rotating __shfl_sync per value.
// The rows are aligned to the END of the strips: `shift` = nv*R - rdlen virtual rows precede
// row 0 in strip 0 (they hold H = 0xff, E = F = 0, i.e. they are copies of row -1), so the last
// row is always row R-1 of strip nv-1 and its H is the value that lane hands down anyway.
//
// The reference keeps H, E and F of every cell for its backtrace; the backtrace only ever
// asks which of its score equalities hold at a cell, so the fill answers those while the
// operands are in registers and stores ONE byte of flags per cell (dev_common.cuh, CF_*).
// The kernel is bound by the integer ALU pipe (VIMNMX, HSET2, LOP3, PRMT; measured in
// tools/pipe_bench.cu), so work is kept off it where the FMA pipe can do it: each equality is one
// HSET2.BF.NEU on the two 16-bit halves (values 0..255 and the small negative unsaturated sums
// compared as fp16 bit patterns: denormals compare exactly, negative values are NaN/-inf patterns
// and can only meet a non-negative partner, where "unordered or not equal" is the right answer)
// giving 1.0 / 0.0, and one HFMA2 (FMA pipe) that adds the flag's weight to an accumulator that
// starts at 1024.0: the fp16 pattern of 1024+x is 0x6400|x, so the low byte of each half IS the
// flag byte.  The flag bytes of a strip's R cells on one diagonal are gathered into a word with
// integer multiply-adds (FMA pipe) and two PRMTs per step.
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
#define BT2_FILL_REFPAD 48             // columns read past the window's end by strips that are done (<= 31 + R + 3)
#define BT2_FILL_REFOFF 32             // columns before the window's start asked for by strips that have not started
#define BT2_FILL_ROWS (BT2_STRIPS * BT2_MAX_R)
struct FillShared {
	uint32_t refw[BT2_FILL_REFOFF + BT2GPU_MAX_COLS + 8 + BT2_FILL_REFPAD]; // [REFOFF + c] = code_A(c) | code_B(c) << 16, 4 outside the window
	uint32_t hl[2][BT2GPU_MAX_COLS + BT2_STRIPS + 16];  // [problem][step]: H handed down by the strip that owns the last row, step = col + strip
	uint8_t  rdc[2][BT2_FILL_ROWS];     // read code per (shifted) row (0..3, 7 = N, virtual or beyond the read)
	uint8_t  pen[2][BT2_FILL_ROWS], penN[2][BT2_FILL_ROWS]; // penalty of the row against a mismatching / an N reference character
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
// (hi - BASE) * 256 + (lo - BASE) for two flag accumulators: bytes [lo_A, hi_A, lo_B, hi_B] (arithmetic mod 2^32)
#define BT2_FL_PAIRFIX 0x64646400u
__device__ __forceinline__ uint32_t flag_pair(uint32_t lo, uint32_t hi) { return hi * 256u + lo - BT2_FL_PAIRFIX; }

struct FillHalf { // one problem of the pair
	int rows, cols, nv, shift;
	bool on;
};

// The steps [t0, t_end) of the sweep, four per iteration.
// MASK: some strips have not reached column 0 yet (steps < 32): their H/E must stay 0.  REFN: a window contains an N.
template<int R, bool DBG, bool MASK, bool REFN>
__device__ __forceinline__ void fill_steps(int t0, const int t_end, const FillHalf& A, const FillHalf& B, const uint32_t lane, FillShared& sh,
		const bool strip_a, const bool strip_b, const bool own_a, const bool own_b,
		const uint32_t (&rdc2)[R], const uint32_t (&negpen)[R], const uint32_t (&negpenN)[R],
		const uint32_t (&neg_rdgo)[R], const uint32_t (&neg_rfgo)[R], const uint32_t (&neg_rfge)[R], const uint32_t neg_rdge,
		const uint32_t nx_rfgo, const uint32_t nx_rfge,
		uint32_t (&Hp)[R], uint32_t (&Ep)[R], uint32_t (&Ecar)[R], uint32_t& sendH, uint32_t& sendF, uint32_t& sendC, uint32_t& prevRecvH,
		uint32_t (&dl)[6], uint4* __restrict__ out_a, uint4* __restrict__ out_b, uint32_t* __restrict__ hef_a, uint32_t* __restrict__ hef_b) {
	const uint32_t W_D = BT2_FL_W(0), W_HE = BT2_FL_W(1), W_HF = BT2_FL_W(2), W_EO = BT2_FL_W(3), W_EX = BT2_FL_W(4), W_FO = BT2_FL_W(5), W_FX = BT2_FL_W(6);
	const uint32_t rot = (lane + 31u) & 31u;
	constexpr int LG = R > 4 ? 1 : 2;      // log2(diagonal words per 16-byte chunk)
	for(; t0 < t_end; t0 += 4) {
		uint32_t wa[4], wb[4], xa[4], xb[4];  // words of the four steps (x*: the fifth byte of R = 5)
		#pragma unroll
		for(int tj = 0; tj < 4; tj++) {
			const int t = t0 + tj;
			const int c = t - (int)lane;
			// values produced by the strip above in the previous step; lane 0: row -1 (H = 0xff, F = 0, neither F comparison holds)
			uint32_t recvH = __shfl_sync(0xffffffffu, sendH, rot), recvF = __shfl_sync(0xffffffffu, sendF, rot), recvC = __shfl_sync(0xffffffffu, sendC, rot);
			if(lane == 0) { recvH = 0x00ff00ffu; recvF = 0u; recvC = BT2_FL_BASE | ((uint32_t)(CF_FO | CF_FX) * 0x00010001u); }
			uint32_t am = 0xffffffffu;
			bool act_a = true, act_b = true;
			if(MASK) {
				act_a = strip_a && c >= 0 && c < A.cols; act_b = strip_b && c >= 0 && c < B.cols;
				am = (act_a ? 0xffffu : 0u) | (act_b ? 0xffff0000u : 0u);
			}
			const uint32_t xpk = sh.refw[BT2_FILL_REFOFF + c];  // reference characters of the two problems' column c
			const uint32_t nmask = REFN ? __heq2_mask(as_h2(xpk), as_h2(0x00040004u)) : 0u;
			uint32_t diag = prevRecvH;
			uint32_t f = recvF;                                            // F(r,c)
			// the cell's flag accumulator: the two comparisons of its E (made one step ago) and of its F (made one row up)
			uint32_t car = recvC | Ecar[0];
			uint32_t flg[R];
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
				uint32_t fl = car;                            // (bit 7 stays 0: not visited)
				fl = flag_ne(h, u, W_D, fl);
				fl = flag_ne(h, e, W_HE, fl);
				fl = flag_ne(h, f, W_HF, fl);
				flg[k] = fl;
				if(DBG) { // H|E<<8|F<<16 of the two problems' cells
					const uint32_t he = __byte_perm(h, e, 0x6240);   // H_A E_A H_B E_B
					const uint32_t wA = (he & 0xffffu) | ((f & 0xffu) << 16), wB = (he >> 16) | (((f >> 16) & 0xffu) << 16);
					if(act_a) hef_a[((size_t)lane * BT2_HEF_T + (uint32_t)t) * 8u + k] = wA;
					if(act_b) hef_b[((size_t)lane * BT2_HEF_T + (uint32_t)t) * 8u + k] = wB;
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
			sendH = Hp[R - 1];
			sendF = MASK ? (f & am) : f;
			sendC = car;
			prevRecvH = recvH;
			// the last row is row R-1 of one strip: what that lane hands down is the last row's H
			if(own_a) sh.hl[0][t] = sendH;
			if(own_b) sh.hl[1][t] = sendH;
			// A word of the cell array = the strip's R cells on one diagonal: byte k comes from the step R-1-k before this
			// one.  dl[]: delay lines of the flag accumulators / byte pairs still waiting for their word.
			if(R == 1) {
				const uint32_t m = flg[0] - BT2_FL_BASE;
				wa[tj] = m & 0xffffu; wb[tj] = m >> 16;
			} else if(R == 2) {
				const uint32_t p = flag_pair(dl[0], flg[1 % R]); dl[0] = flg[0];
				wa[tj] = p & 0xffffu; wb[tj] = p >> 16;
			} else if(R == 3) {
				const uint32_t p = flag_pair(dl[0], flg[1 % R]); dl[0] = flg[0];      // rows 0 (previous step) and 1 (now): their word is done next step
				const uint32_t m2 = flg[2 % R] - BT2_FL_BASE;
				wa[tj] = __byte_perm(dl[1], m2, 0x5410); wb[tj] = __byte_perm(dl[1], m2, 0x7632);
				dl[1] = p;
			} else if(R == 4) {
				const uint32_t p01 = flag_pair(dl[0], flg[1 % R]); dl[0] = flg[0];
				const uint32_t p23 = flag_pair(dl[3], flg[3 % R]); dl[3] = flg[2 % R];
				wa[tj] = __byte_perm(dl[2], p23, 0x5410); wb[tj] = __byte_perm(dl[2], p23, 0x7632);
				dl[2] = dl[1]; dl[1] = p01;
			} else {
				const uint32_t p01 = flag_pair(dl[0], flg[1 % R]); dl[0] = flg[0];
				const uint32_t p23 = flag_pair(dl[3], flg[3 % R]); dl[3] = flg[2 % R];
				const uint32_t m4 = flg[4 % R] - BT2_FL_BASE;
				wa[tj] = __byte_perm(dl[4], dl[5], 0x5410); wb[tj] = __byte_perm(dl[4], dl[5], 0x7632);
				xa[tj] = m4 & 0xffffu; xb[tj] = m4 >> 16;
				dl[4] = dl[2]; dl[2] = dl[1]; dl[1] = p01; dl[5] = p23;
			}
			if(R > 4 && (tj & 1)) { // two diagonal words (8 bytes each) make a chunk
				const bool in_a = strip_a && (!MASK || t >= (int)lane) && t - 1 - (R - 1) < (int)lane + A.cols;
				const bool in_b = strip_b && (!MASK || t >= (int)lane) && t - 1 - (R - 1) < (int)lane + B.cols;
				const int g = (t >> LG) * BT2_STRIPS + (int)lane;
				if(in_a) out_a[g] = make_uint4(wa[(tj + 3) & 3], xa[(tj + 3) & 3], wa[tj], xa[tj]);
				if(in_b) out_b[g] = make_uint4(wb[(tj + 3) & 3], xb[(tj + 3) & 3], wb[tj], xb[tj]);
			}
		}
		if(R <= 4) { // strip v is active at steps v .. v+cols-1 and its last words complete R-1 steps later
			const bool in_a = strip_a && (!MASK || t0 + 3 >= (int)lane) && t0 - (R - 1) < (int)lane + A.cols;
			const bool in_b = strip_b && (!MASK || t0 + 3 >= (int)lane) && t0 - (R - 1) < (int)lane + B.cols;
			const int g = (t0 >> LG) * BT2_STRIPS + (int)lane;
			if(in_a) out_a[g] = make_uint4(wa[0], wa[1], wa[2], wa[3]);
			if(in_b) out_b[g] = make_uint4(wb[0], wb[1], wb[2], wb[3]);
		}
	}
}

template<int R, bool DBG>
__device__ __forceinline__ void fill_body(const DevParams& P, const FillHalf& A, const FillHalf& B, FillShared& sh, uint8_t* __restrict__ cells_a, uint8_t* __restrict__ cells_b,
                                          uint32_t* __restrict__ hef_a, uint32_t* __restrict__ hef_b, uint32_t lane, bool refHasN) {
	// per-row gap penalties; inside the gap barrier (aligner_swsse_ee_u8.cpp:101-115: first/last gapbar rows), in the
	// virtual rows and (reference gaps) into row 0 they are large enough to keep E and F at 0
	const uint32_t neg_rdge = pack2(-P.rdGapExt, -P.rdGapExt);
	uint32_t neg_rdgo[R], neg_rfgo[R], neg_rfge[R];    // [k]: read gaps in row k; ref gaps INTO row k (from the row above)
	uint32_t rdc2[R], negpen[R], negpenN[R];           // read character and (negated) penalties of row k, both problems
	auto gaps_ok = [&](const FillHalf& X, int rp) { const int r = rp - X.shift; return r >= 0 && !(r < P.gapbar || X.rows - 1 - r < P.gapbar); };
	auto into_ok = [&](const FillHalf& X, int rp) { const int r = rp - X.shift; return r > 0 && !(r < P.gapbar || X.rows - 1 - r < P.gapbar); };
	#pragma unroll
	for(int k = 0; k < R; k++) {
		const int rp = (int)lane * R + k;
		neg_rdgo[k] = pack2(gaps_ok(A, rp) ? -P.rdGapOpen : -1024, gaps_ok(B, rp) ? -P.rdGapOpen : -1024);
		neg_rfgo[k] = pack2(into_ok(A, rp) ? -P.rfGapOpen : -1024, into_ok(B, rp) ? -P.rfGapOpen : -1024);
		neg_rfge[k] = pack2(into_ok(A, rp) ? -P.rfGapExt : -1024, into_ok(B, rp) ? -P.rfGapExt : -1024);
		rdc2[k] = (uint32_t)sh.rdc[0][rp] | ((uint32_t)sh.rdc[1][rp] << 16);
		negpen[k] = pack2(-(int)sh.pen[0][rp], -(int)sh.pen[1][rp]);
		negpenN[k] = pack2(-(int)sh.penN[0][rp], -(int)sh.penN[1][rp]);
	}
	uint32_t nx_rfgo, nx_rfge;                         // ref gaps into the first row of the next lane's strip
	{
		const int rp = ((int)lane + 1) * R;
		nx_rfgo = pack2(into_ok(A, rp) ? -P.rfGapOpen : -1024, into_ok(B, rp) ? -P.rfGapOpen : -1024);
		nx_rfge = pack2(into_ok(A, rp) ? -P.rfGapExt : -1024, into_ok(B, rp) ? -P.rfGapExt : -1024);
	}
	const bool strip_a = A.on && (int)lane < A.nv, strip_b = B.on && (int)lane < B.nv;
	const bool own_a = A.on && (int)lane == A.nv - 1, own_b = B.on && (int)lane == B.nv - 1;
	uint32_t Hp[R], Ep[R], Ecar[R];
	#pragma unroll
	for(int k = 0; k < R; k++) {
		// virtual rows (strip 0 only) are copies of row -1: H = 0xff in every column, including column -1
		const int rp = (int)lane * R + k;
		Hp[k] = pack2(rp < A.shift ? 0xff : 0, rp < B.shift ? 0xff : 0);
		Ep[k] = 0; Ecar[k] = BT2_FL_BASE | ((CF_EO | CF_EX) * 0x00010001u);
	}
	uint32_t sendH = 0, sendF = 0, sendC = BT2_FL_BASE, prevRecvH = (lane == 0) ? 0x00ff00ffu : 0u;
	// the sweep runs R-1 steps longer than the last column needs: the last diagonal words complete then
	const int nst_a = A.on ? A.cols + A.nv - 1 + (R - 1) : 0, nst_b = B.on ? B.cols + B.nv - 1 + (R - 1) : 0;
	const int nsteps = nst_a > nst_b ? nst_a : nst_b;
	uint32_t dl[6] = { BT2_FL_BASE, 0, 0, BT2_FL_BASE, 0, 0 };
	uint4* const out_a = reinterpret_cast<uint4*>(cells_a); uint4* const out_b = reinterpret_cast<uint4*>(cells_b);
	const int t_mask = DBG ? nsteps : (nsteps < BT2_STRIPS ? nsteps : BT2_STRIPS);   // every strip has started after step 31
	#define BT2_FILL_ARGS A, B, lane, sh, strip_a, strip_b, own_a, own_b, rdc2, negpen, negpenN, neg_rdgo, neg_rfgo, neg_rfge, neg_rdge, \
		nx_rfgo, nx_rfge, Hp, Ep, Ecar, sendH, sendF, sendC, prevRecvH, dl, out_a, out_b, hef_a, hef_b
	if(refHasN) {
		fill_steps<R, DBG, true, true>(0, t_mask, BT2_FILL_ARGS);
		if(!DBG) fill_steps<R, false, false, true>((t_mask + 3) & ~3, nsteps, BT2_FILL_ARGS);
	} else {
		fill_steps<R, DBG, true, false>(0, t_mask, BT2_FILL_ARGS);
		if(!DBG) fill_steps<R, false, false, false>((t_mask + 3) & ~3, nsteps, BT2_FILL_ARGS);
	}
	#undef BT2_FILL_ARGS
}

template<bool DBG>
__global__ void __launch_bounds__(BT2_FILL_WARPS * 32) k_fill(const DevParams* __restrict__ Pp, const FillArgs a) {
	__shared__ FillShared shs[BT2_FILL_WARPS];
	const DevParams& P = *Pp;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t slot0 = (blockIdx.x * BT2_FILL_WARPS + warp) * 2u;
	if(slot0 >= a.nprob) return;
	FillShared& sh = shs[warp];
	FillHalf hv[2];
	int R = 1;
	#pragma unroll
	for(int x = 0; x < 2; x++) {
		const uint32_t slot = slot0 + x;
		hv[x].on = slot < a.nprob && a.probs[slot].read != BT2_NIL;
		hv[x].rows = hv[x].on ? a.probs[slot].rows : 0; hv[x].cols = hv[x].on ? a.probs[slot].cols : 0;
		const int r = (int)fill_R((uint32_t)hv[x].rows);
		if(r > R) R = r;
	}
	if(!hv[0].on && !hv[1].on) return;
	bool hasN = false;
	#pragma unroll
	for(int x = 0; x < 2; x++) {
		FillHalf& X = hv[x];
		X.nv = X.on ? (X.rows + R - 1) / R : 0; X.shift = X.on ? X.nv * R - X.rows : 0;
		if(!X.on) { // an absent problem: rows that never match, no penalties
			for(int r = (int)lane; r < BT2_FILL_ROWS; r += 32) { sh.rdc[x][r] = 7; sh.pen[x][r] = 0; sh.penN[x][r] = 0; }
			continue;
		}
		Problem& pb = a.probs[slot0 + x];
		const ReadState& rs = a.states[pb.read];
		const bool fw = pb.fw != 0;
		const uint8_t* rseq = a.seq + rs.seq_off; const uint8_t* rqual = a.qual + rs.seq_off;
		// query profile: per (shifted) row its character and its penalty against a mismatching / an N reference character
		for(int rp = (int)lane; rp < R * BT2_STRIPS; rp += 32) {
			const int r = rp - X.shift;
			int rdc = 4, q = 0;
			const bool real = r >= 0 && r < X.rows;
			if(real) {
				if(fw) { rdc = rseq[r]; q = rqual[r]; }
				else { const int c0 = rseq[X.rows - 1 - r]; rdc = c0 > 3 ? 4 : 3 - c0; q = rqual[X.rows - 1 - r]; }
				q -= 33; if(q < 0) q = 0; if(q > 255) q = 255;
			}
			sh.rdc[x][rp] = (uint8_t)(rdc > 3 ? 7 : rdc);
			sh.pen[x][rp] = real ? (rdc > 3 ? P.npens[q] : P.mmpens[q]) : (uint8_t)0;
			sh.penN[x][rp] = real ? P.npens[q] : (uint8_t)0;
		}
		if(lane == 0) pb.R = (uint8_t)R;   // the layout of this problem's cells (the pair's R)
	}
	for(int i = (int)lane; i < BT2_FILL_REFOFF + BT2GPU_MAX_COLS + 8 + BT2_FILL_REFPAD; i += 32) {
		const int c = i - BT2_FILL_REFOFF;
		uint32_t w = 0;
		#pragma unroll
		for(int x = 0; x < 2; x++) {
			uint32_t code = 4;
			if(hv[x].on && c >= 0 && c < hv[x].cols) { const uint32_t m = a.probs[slot0 + x].rf[c]; code = m > 15 ? 4u : (uint32_t)(31 - __clz(m)); hasN = hasN || code == 4u; }
			w |= code << (16 * x);
		}
		sh.refw[i] = w;
	}
	const bool refHasN = __any_sync(0xffffffffu, hasN);
	__syncwarp();
	if(a.pcache) { // the problems' characters as 4-bit codes, for the backtrace's edits (dev_common.cuh)
		#pragma unroll
		for(int x = 0; x < 2; x++) {
			if(!hv[x].on) continue;
			uint32_t* pcw = a.pcache + (size_t)(slot0 + x) * BT2_PC_WORDS;
			uint32_t wr = 0, wf = 0;
			#pragma unroll
			for(int i = 0; i < 8; i++) {
				const int r = (int)lane * 8 + i, c = (int)lane * 8 + i;
				const uint32_t rc = r < hv[x].rows ? (uint32_t)sh.rdc[x][r + hv[x].shift] : 0u;
				wr |= (rc > 3u ? 4u : rc) << (4 * i);
				wf |= (c < hv[x].cols ? ((sh.refw[BT2_FILL_REFOFF + c] >> (16 * x)) & 0xffu) : 0u) << (4 * i);
			}
			if(lane < BT2_PC_RD) pcw[lane] = wr;
			if(lane < BT2_PC_RF) pcw[BT2_PC_RD + lane] = wf;
		}
	}
	uint8_t* cells_a = a.cells + (size_t)slot0 * BT2_CELL_BYTES; uint8_t* cells_b = cells_a + BT2_CELL_BYTES;
	uint32_t* hef_a = DBG ? a.hef_dbg + (size_t)slot0 * BT2_HEF_WORDS : nullptr; uint32_t* hef_b = DBG ? hef_a + BT2_HEF_WORDS : nullptr;
	if(R == 1) fill_body<1, DBG>(P, hv[0], hv[1], sh, cells_a, cells_b, hef_a, hef_b, lane, refHasN);
	else if(R == 2) fill_body<2, DBG>(P, hv[0], hv[1], sh, cells_a, cells_b, hef_a, hef_b, lane, refHasN);
	else if(R == 3) fill_body<3, DBG>(P, hv[0], hv[1], sh, cells_a, cells_b, hef_a, hef_b, lane, refHasN);
	else if(R == 4) fill_body<4, DBG>(P, hv[0], hv[1], sh, cells_a, cells_b, hef_a, hef_b, lane, refHasN);
	else fill_body<5, DBG>(P, hv[0], hv[1], sh, cells_a, cells_b, hef_a, hef_b, lane, refHasN);
	__syncwarp();
	// ---- last-row candidates: H-0xff >= minsc, sorted by (score desc, col desc)
	for(int x = 0; x < 2; x++) {
		if(!hv[x].on) continue;
		Problem& pb = a.probs[slot0 + x];
		const int cols = hv[x].cols, vstar = hv[x].nv - 1;
		const int minsc = pb.minsc;
		int lrmax = 0; uint32_t ncand = 0;
		for(int c0 = 0; c0 < cols; c0 += 32) {
			const int c = c0 + (int)lane;
			const int hv8 = c < cols ? (int)((sh.hl[x][c + vstar] >> (16 * x)) & 0xffu) : 0;
			lrmax = max(lrmax, hv8);
			const bool isc = c < cols && (hv8 - 255) >= minsc;
			const uint32_t bal = __ballot_sync(0xffffffffu, isc);
			if(isc) sh.ckey[ncand + __popc(bal & ((1u << lane) - 1))] = (uint16_t)((hv8 << 8) | c);
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
		__syncwarp();
		if(lane == 0) {
			pb.lrmax = lrmax; pb.ncand = ncand;
			atomicAdd(&a.stats[0], 1ull); atomicAdd(&a.stats[1], (unsigned long long)hv[x].rows * cols);
		}
	}
}

} // namespace bt2
