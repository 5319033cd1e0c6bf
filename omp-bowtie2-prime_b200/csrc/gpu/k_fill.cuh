// K_fill: the end-to-end affine-gap DP fill, four read-window problems per warp, banded.
//   SwAligner::alignEnd2EndSseU8 / EEU8_alignNucleotides / EEU8_alignOne / EEU8_lazyF
//       aligner_swsse_ee_u8.cpp:250-671   (recurrence, unsigned saturating 8-bit, start 0xff)
//   SwAligner::buildQueryProfileEnd2EndSseU8   aligner_swsse_ee_u8.cpp:71-134
//   candidate gathering + DpBtCandidate order  aligner_swsse_ee_u8.cpp:600-671, aligner_sw_nuc.h
//
// The reference computes the full rdlen x rflen rectangle in a Farrar-striped SSE2 layout.  Its results do not depend
// on the evaluation order, and nothing it reports depends on cells that no alignment scoring >= minsc can pass through
// (dev_common.cuh, "Which cells are filled"): only the band of diagonals -g <= col - row <= cols - rows + g is computed,
// about half of the rectangle for 150 bp reads at the default thresholds.
//
// Mapping.  Lane v owns K consecutive diagonals of the band and sweeps the ROWS: in one step it computes the K cells of one
// row that lie on its diagonals, left to right.  Of a cell's three neighbours, the diagonal one is the lane's own cell of the
// previous row (a register), the left one (E) is the lane's previous cell of the same row or, for its first cell, the last
// cell lane v-1 computed in that row, and the upper one (F) is the next diagonal's cell of the previous row: the lane's own,
// or for its last cell the first cell lane v+1 computed in the previous row.  Lane v therefore works on row r at step
// 2 r + v: one step after lane v-1 did that row and one step after lane v+1 did the row above.  A lane would idle every other
// step, so the warp interleaves TWO problem sets, the second one step behind the first (even lanes do set X on even steps and
// set Y on odd ones, odd lanes the other way round); a set is a PAIR of problems, one in the low and one in the high 16-bit
// half of every register (s16x2 DPX instructions).  Four problems per warp, every lane busy at every step.
// Two shuffles per step carry E (and the outcome of E's two comparisons) to the right, two carry F to the left.
//
// The reference keeps H, E and F of every cell for its backtrace; the backtrace only ever asks which of its score
// equalities hold at a cell, so the fill answers those while the operands are in registers and stores ONE byte of flags per
// cell (dev_common.cuh, CF_*): each equality is one HSET2.BF.NEU on the two halves and one HFMA2 (FMA pipe) that adds the
// flag's weight to an fp16 accumulator whose low byte is the flag byte (simd16.cuh).  The K flag bytes of a step go to one
// unit, 16 / K steps of a problem to one 16-byte chunk; all lanes store the chunk of the same iterations together, one
// 512-byte row per problem.
// Rows before row 0 ("virtual rows": a lane starts (v + y) / 2 iterations early so that the loop has no per-lane start) are
// copies of row -1: H = 0xff, no penalties, no gaps, so they hand the right boundary values to row 0 by themselves.
// Columns left of the window need nothing: their diagonals start with H = 0 and stay there (see fill_step).
#pragma once
#include "dev_common.cuh"
#include "simd16.cuh"

namespace bt2 {

struct FillArgs {
	Problem*         probs;     // [nslots]
	uint32_t         nprob;
	const ReadState* states;
	const uint8_t*   seq;
	const uint8_t*   qual;
	uint8_t*         cells;     // [nslots * cell_stride]
	size_t           cell_stride;
	uint32_t*        pcache;    // [nslots * BT2_PC_WORDS] character cache for the backtrace, or null
	uint32_t*        hef_dbg;   // test hook only: [nslots * BT2_HEF_WORDS] H|E<<8|F<<16, or null
	unsigned long long* stats;  // [0] problems [1] cells of the rectangles [2] cells computed
	uint32_t*        error;     // |= 64 if a problem's band does not fit the lanes (host-side K too small: internal error)
	int32_t          exact;     // fill the whole rectangle (bt2gpu_params.exactBestLtMinsc)
};

#if BT2GPU_MAX_ROWS > 160
#define BT2_FILL_WARPS 1               // the long-read build: a warp's tables take ~70 KB of (dynamic) shared memory
#define BT2_FILL_DYNSMEM 1
#else
#ifndef BT2_FILL_WARPS
#define BT2_FILL_WARPS 2
#endif
#define BT2_FILL_DYNSMEM 0
#endif
#ifndef BT2_FILL_UNROLL
#define BT2_FILL_UNROLL 4              // iterations of the sweep per loop trip (a divisor of 16 / K; at most the 16 / K of a chunk).  Measured for
                                       // K = 4 (B200, 1 M x 150 bp): 1 -> 15.6 ms, 2 -> 14.35 ms, 4 -> 14.05 ms (the chunk's shift register and the rotating
                                       // state become register renaming); while the body still held the masked variant and the one-by-one prologue,
                                       // 4 iterations (1 430 instructions) did not fit the instruction cache and 1 was the fastest
#endif
#define BT2_FILL_ROWPAD BT2_SKEW                       // virtual rows before row 0 (and rows past the end the last lanes touch)
#define BT2_FILL_NROW (BT2_FILL_ROWPAD + BT2GPU_MAX_ROWS + BT2_SKEW + 4)

// "No gap here": a penalty no score survives.  H + NOGAP is negative for every H in 0 .. BT2_I16_START, and its 16-bit pattern
// (0x8c00 .. 0xfc00) is a negative fp16 number, -inf or NaN, never -0: the comparison with the (non-negative) maximum says "differs".
#define BT2_FILL_NOGAP (-0x7400)
struct FillRow { uint32_t rdc2, negpen, neg_rdgo, neg_rfgo_nx; };   // row r: characters, mismatch penalty, read-gap open; ref-gap open INTO row r+1
struct FillRow2 { uint32_t neg_rfge_nx, negpenN; };                 // ref-gap extension into row r+1; penalty against an N reference character

template<int K> struct FillShared {
	FillRow  row[2][BT2_FILL_NROW];                 // [set][BT2_FILL_ROWPAD + r], both problems of the set packed (A low half, B high half)
	FillRow2 row2[2][BT2_FILL_NROW];
	uint32_t refw[2][BT2_FILL_ROWPAD + BT2GPU_MAX_ROWS + BT2_SKEW + BT2_LANES * K + 24];   // [set][BT2_FILL_ROWPAD + t]: code_A(t + d0_A) | code_B(t + d0_B) << 16
	uint16_t hl[4][BT2GPU_MAX_COLS + 8];            // [problem][col]: H of the last row (0 outside the band)
	// (the candidate keys of the epilogue reuse row2[0]: BT2GPU_MAX_COLS + 8 words, dead after the sweep)
};
static_assert(sizeof(FillRow2) * BT2_FILL_NROW >= sizeof(uint32_t) * (BT2GPU_MAX_COLS + 8), "candidate keys overlay row2[0]");

struct FillProb { int rows, cols, d0, start, minsc; bool on, fw; };   // start: H of row -1 (0xff, or BT2_I16_START on the reference's 16-bit path)

// What one lane holds of one problem set.
template<int K> struct FillSet {
	uint32_t Hd[K];     // H(r-1, c-1) of the cell this lane computes next on diagonal j
	uint32_t Fdn[K];    // F of the cell below the one computed last on diagonal j (consumed by diagonal j-1 in the next row)
	uint32_t Fc[K];     // ... and the flag accumulator holding the outcome of that F's two comparisons
	uint32_t ca[4], cb[4];   // the 16-byte chunk being assembled for problem A / B
	const FillRow* row; const FillRow2* row2; const uint32_t* refw;   // shared-memory rows of the set, biased so that [i] is this lane's iteration i
	int tA, tB;         // column of the lane's first cell at iteration i is i + t
	int delta;          // row at iteration i is i - delta
	int colsA, colsB;
	uint16_t* hlA; uint16_t* hlB;
	uint32_t* hefA; uint32_t* hefB;   // debug
	int rowsA, rowsB;
};

// One step of one lane: the K cells of row i - delta on the lane's diagonals.
// recvE/recvCE: E of the first cell and the accumulator with E's comparisons, from the lane on the left (previous step);
// recvF/recvCF: F of the last cell, from the lane on the right.  Leaves what the neighbours need next in send*.
// (Cells of real rows left of the window -- col < 0, on the diagonals below 0 -- need no special case: their diagonal starts with H = 0
// (fill_set_init), scores are <= 0, E enters lane 0 as 0 and F comes down from cells of the same kind, so they compute to H = E = F = 0,
// which is what the reference's first column sees to its left.)
template<int K, bool REFN, bool DBG>
BT2_SIMD void fill_step(FillSet<K>& S, const int i, const uint32_t lane, const uint32_t neg_rdge,
                        const uint32_t recvE, const uint32_t recvCE, const uint32_t recvF, const uint32_t recvCF,
                        uint32_t& sendE, uint32_t& sendCE, uint32_t& sendF, uint32_t& sendCF) {
	const FillRow rw = S.row[i];
	const FillRow2 rw2 = S.row2[i];
	const uint32_t pen = rw.negpen;
	uint32_t e = recvE, ce = recvCE;
	uint32_t P2[K / 2];   // flag bytes of cells 2m, 2m+1: [A lo, A hi, B lo, B hi]
	uint32_t prevfl = 0;
	#pragma unroll
	for(int j = 0; j < K; j++) {
		const uint32_t xpk = S.refw[i + j];
		const uint32_t mm = v_ne2(rw.rdc2, xpk);
		uint32_t s;
		if(REFN) { const uint32_t nmask = v_eq2(xpk, 0x00040004u); s = mm & ((pen & ~nmask) | (rw2.negpenN & nmask)); }
		else s = mm & pen;
		const uint32_t u = v_add2(S.Hd[j], s);                        // unsaturated H(r-1,c-1) + score
		const uint32_t f = j + 1 < K ? S.Fdn[(j + 1) % K] : recvF;
		const uint32_t cf = j + 1 < K ? S.Fc[(j + 1) % K] : recvCF;
		const uint32_t h = v_max3r(u, e, f);
		// the comparisons of the backtrace at this cell (bit 7 stays 0: not visited)
		uint32_t fl = ce | cf;
		fl = v_flag_ne<0>(h, u, fl);
		fl = v_flag_ne<1>(h, e, fl);
		fl = v_flag_ne<2>(h, f, fl);
		if(DBG) {
			const int r = i - S.delta, cA = i + S.tA + j, cB = i + S.tB + j;
			const uint32_t he = v_prmt(h, e, 0x6240);   // H_A E_A H_B E_B
			if(r >= 0 && r < S.rowsA && cA >= 0 && cA < S.colsA) S.hefA[(size_t)r * (BT2_LANES * K) + lane * K + j] = (he & 0xffffu) | ((f & 0xffu) << 16);
			if(r >= 0 && r < S.rowsB && cB >= 0 && cB < S.colsB) S.hefB[(size_t)r * (BT2_LANES * K) + lane * K + j] = (he >> 16) | (((f >> 16) & 0xffu) << 16);
		}
		// E of the next column in this row and its two comparisons
		const uint32_t be = v_add2(h, rw.neg_rdgo), ae = v_add2(e, neg_rdge);
		const uint32_t en = v_maxr(ae, be);
		ce = v_flag_ne<3>(en, be, v_flag_ne<4>(en, ae, BT2_FL_BASE));
		e = en;
		// F of the next row in this column and its two comparisons
		const uint32_t bf = v_add2(h, rw.neg_rfgo_nx), af = v_add2(f, rw2.neg_rfge_nx);
		const uint32_t fn = v_maxr(af, bf);
		S.Fc[j] = v_flag_ne<5>(fn, bf, v_flag_ne<6>(fn, af, BT2_FL_BASE));
		S.Fdn[j] = fn;
		S.Hd[j] = h;
		if(j & 1) P2[j >> 1] = v_flag_pair(prevfl, fl); else prevfl = fl;
	}
	sendE = e; sendCE = ce; sendF = S.Fdn[0]; sendCF = S.Fc[0];
	// the step's K flag bytes into the chunk: a 16-byte shift register per problem, oldest step in the lowest bytes (after
	// 16 / K steps it holds the chunk; when the loop is unrolled the moves are register renaming)
	if constexpr(K == 2) {
		#pragma unroll
		for(int m = 0; m < 3; m++) { S.ca[m] = v_prmt(S.ca[m], S.ca[m + 1], 0x5432); S.cb[m] = v_prmt(S.cb[m], S.cb[m + 1], 0x5432); }
		S.ca[3] = v_prmt(S.ca[3], P2[0], 0x5432); S.cb[3] = v_prmt(S.cb[3], P2[0], 0x7632);
	} else {
		constexpr int NW = K / 4;   // words per step and problem
		#pragma unroll
		for(int m = 0; m + NW < 4; m++) { S.ca[m] = S.ca[m + NW]; S.cb[m] = S.cb[m + NW]; }
		#pragma unroll
		for(int m = 0; m < NW; m++) {
			S.ca[4 - NW + m] = v_prmt(P2[2 * m], P2[2 * m + 1], 0x5410);
			S.cb[4 - NW + m] = v_prmt(P2[2 * m], P2[2 * m + 1], 0x7632);
		}
	}
}

// After the sweep.  Rows past a problem's end have no penalties and no gaps, like the virtual rows before row 0: E and F stay 0
// and H = max(H(r-1, c-1) + 0, 0, 0), i.e. the last row's H travels down its diagonal unchanged.  When the sweep ends, Hd[j]
// therefore still is H(last row, diagonal j) of either problem: it goes to shared memory for the candidate scan.
template<int K> BT2_SIMD void fill_last_row(const FillSet<K>& S) {
	#pragma unroll
	for(int j = 0; j < K; j++) {
		const int cA = S.rowsA - 1 + S.tA + S.delta + j, cB = S.rowsB - 1 + S.tB + S.delta + j;   // column of diagonal j in the last row
		if(S.rowsA > 0 && (unsigned)cA < (unsigned)S.colsA) S.hlA[cA] = (uint16_t)(S.Hd[j] & 0xffffu);
		if(S.rowsB > 0 && (unsigned)cB < (unsigned)S.colsB) S.hlB[cB] = (uint16_t)(S.Hd[j] >> 16);
	}
}

// per-lane setup of one set: q0 = index of the set's first problem among the warp's four
template<int K> BT2_SIMD void fill_set_init(FillSet<K>& S, FillShared<K>& sh, const FillProb (&pr)[4], const int set, const uint32_t lane, const int delta) {
	const FillProb& A = pr[2 * set]; const FillProb& B = pr[2 * set + 1];
	S.delta = delta;
	S.row = sh.row[set] + (BT2_FILL_ROWPAD - delta);
	S.row2 = sh.row2[set] + (BT2_FILL_ROWPAD - delta);
	S.refw = sh.refw[set] + (BT2_FILL_ROWPAD - delta + (int)lane * K);
	S.tA = A.d0 + (int)lane * K - delta; S.tB = B.d0 + (int)lane * K - delta;
	S.rowsA = A.rows; S.rowsB = B.rows; S.colsA = A.cols; S.colsB = B.cols;
	S.hlA = sh.hl[2 * set]; S.hlB = sh.hl[2 * set + 1];
	S.hefA = nullptr; S.hefB = nullptr;
	#pragma unroll
	for(int j = 0; j < K; j++) {
		// diagonals that start in row 0 take H(-1, c-1) = 0xff; the others start left of the window (H = 0)
		const int dA = A.d0 + (int)lane * K + j, dB = B.d0 + (int)lane * K + j;
		S.Hd[j] = v_pack2(dA >= 0 ? A.start : 0, dB >= 0 ? B.start : 0);
		S.Fdn[j] = 0; S.Fc[j] = BT2_FL_BASE | ((uint32_t)(CF_FO | CF_FX) * 0x00010001u);
	}
	#pragma unroll
	for(int m = 0; m < 4; m++) { S.ca[m] = 0; S.cb[m] = 0; }
}

// query profile of one set in shared memory, entries idx, idx + 32, ... of this lane: per row its characters and its penalty
// against a mismatching / an N reference character, and the (negated) gap penalties; inside the gap barrier
// (aligner_swsse_ee_u8.cpp:101-115: first/last gapbar rows), in the virtual rows and (reference gaps) into row 0 the gap
// penalties are large enough to keep E and F at 0
// pentab: Scoring::mmpens then Scoring::npens, 256 bytes each (DevParams holds them back to back; the kernel reads its shared-memory copy)
template<int K> BT2_SIMD void fill_build_rows(const DevParams& P, const uint8_t* pentab, FillShared<K>& sh, const int set, const FillProb& A, const FillProb& B,
                                              const uint8_t* seqA, const uint8_t* qualA, bool fwA, const uint8_t* seqB, const uint8_t* qualB, bool fwB,
                                              const uint32_t lane, const int nrow) {
	for(int idx = (int)lane; idx < nrow; idx += BT2_LANES) {
		const int r = idx - BT2_FILL_ROWPAD;
		int rdc[2], pen[2], penN[2], rdgo[2], rfgo[2], rfge[2];
		#pragma unroll
		for(int x = 0; x < 2; x++) {
			const FillProb& X = x ? B : A; const uint8_t* rseq = x ? seqB : seqA; const uint8_t* rqual = x ? qualB : qualA; const bool fw = x ? fwB : fwA;
			const bool real = X.on && r >= 0 && r < X.rows;
			int c = 4, q = 0;
			if(real) {
				if(fw) { c = rseq[r]; q = rqual[r]; }
				else { const int c0 = rseq[X.rows - 1 - r]; c = c0 > 3 ? 4 : 3 - c0; q = rqual[X.rows - 1 - r]; }
				q -= 33; if(q < 0) q = 0; if(q > 255) q = 255;
			}
			rdc[x] = c > 3 ? 7 : c;
			pen[x] = real ? (c > 3 ? pentab[256 + q] : pentab[q]) : 0;
			penN[x] = real ? pentab[256 + q] : 0;
			const bool gaps_ok = real && !(r < P.gapbar || X.rows - 1 - r < P.gapbar);
			const int r1 = r + 1;
			const bool into_ok = X.on && r1 > 0 && r1 < X.rows && !(r1 < P.gapbar || X.rows - 1 - r1 < P.gapbar);
			rdgo[x] = gaps_ok ? -P.rdGapOpen : BT2_FILL_NOGAP;
			rfgo[x] = into_ok ? -P.rfGapOpen : BT2_FILL_NOGAP;
			rfge[x] = into_ok ? -P.rfGapExt : BT2_FILL_NOGAP;
		}
		FillRow w; w.rdc2 = v_pack2(rdc[0], rdc[1]); w.negpen = v_pack2(-pen[0], -pen[1]); w.neg_rdgo = v_pack2(rdgo[0], rdgo[1]); w.neg_rfgo_nx = v_pack2(rfgo[0], rfgo[1]);
		FillRow2 w2; w2.neg_rfge_nx = v_pack2(rfge[0], rfge[1]); w2.negpenN = v_pack2(-penN[0], -penN[1]);
		sh.row[set][idx] = w; sh.row2[set][idx] = w2;
	}
}

// reference characters of one set by diagonal-relative column t = col - d0 (code 4 = N or outside the window); returns whether
// a window holds an N
template<int K> BT2_SIMD bool fill_build_ref(FillShared<K>& sh, const int set, const FillProb& A, const FillProb& B, const uint8_t* rfA, const uint8_t* rfB,
                                             const uint32_t lane, const int nref) {
	bool hasN = false;
	for(int idx = (int)lane; idx < nref; idx += BT2_LANES) {
		const int t = idx - BT2_FILL_ROWPAD;
		uint32_t w = 0;
		#pragma unroll
		for(int x = 0; x < 2; x++) {
			const FillProb& X = x ? B : A; const uint8_t* rf = x ? rfB : rfA;
			const int c = t + X.d0;
			uint32_t code = 4;
			if(X.on && c >= 0 && c < X.cols) {
				const uint32_t m = rf[c];
				code = m > 15 ? 4u : (m >= 8 ? 3u : (m >= 4 ? 2u : (m >= 2 ? 1u : 0u)));
				hasN = hasN || code == 4u;
			}
			w |= code << (16 * x);
		}
		sh.refw[set][idx] = w;
	}
	return hasN;
}

#if defined(__CUDACC__)

// the sweep: iterations [i0, i1), 16 / K per chunk
template<int K, bool REFN, bool DBG>
__device__ __forceinline__ void fill_sweep(int i0, const int i1, FillSet<K>& S0, FillSet<K>& S1, const uint32_t lane, const uint32_t neg_rdge,
                                           uint32_t& sendE, uint32_t& sendCE, uint32_t& sendF, uint32_t& sendCF,
                                           uint4* __restrict__ o0a, uint4* __restrict__ o0b, uint4* __restrict__ o1a, uint4* __restrict__ o1b) {
	constexpr int PER = 16 / K;
	constexpr int UNR = BT2_FILL_UNROLL < PER ? BT2_FILL_UNROLL : PER;
	const uint32_t E_EDGE = BT2_FL_BASE | ((uint32_t)(CF_EO | CF_EX) * 0x00010001u), F_EDGE = BT2_FL_BASE | ((uint32_t)(CF_FO | CF_FX) * 0x00010001u);
	for(; i0 < i1; i0 += PER) {
		// (BT2_FILL_UNROLL iterations per trip; see there for what the instruction cache made of the alternatives)
		#pragma unroll UNR
		for(int u = 0; u < PER; u++) {
			const int i = i0 + u;
			#pragma unroll
			for(int slot = 0; slot < 2; slot++) {
				uint32_t rE = __shfl_up_sync(0xffffffffu, sendE, 1), rCE = __shfl_up_sync(0xffffffffu, sendCE, 1);
				uint32_t rF = __shfl_down_sync(0xffffffffu, sendF, 1), rCF = __shfl_down_sync(0xffffffffu, sendCF, 1);
				if(lane == 0) { rE = 0u; rCE = E_EDGE; }                   // nothing left of the band
				if(lane == BT2_LANES - 1) { rF = 0u; rCF = F_EDGE; }       // nothing right of it
				if(slot == 0) fill_step<K, REFN, DBG>(S0, i, lane, neg_rdge, rE, rCE, rF, rCF, sendE, sendCE, sendF, sendCF);
				else fill_step<K, REFN, DBG>(S1, i, lane, neg_rdge, rE, rCE, rF, rCF, sendE, sendCE, sendF, sendCF);
			}
		}
		const int g = (i0 / PER) * BT2_LANES;
		if(o0a) o0a[g] = make_uint4(S0.ca[0], S0.ca[1], S0.ca[2], S0.ca[3]);
		if(o0b) o0b[g] = make_uint4(S0.cb[0], S0.cb[1], S0.cb[2], S0.cb[3]);
		if(o1a) o1a[g] = make_uint4(S1.ca[0], S1.ca[1], S1.ca[2], S1.ca[3]);
		if(o1b) o1b[g] = make_uint4(S1.cb[0], S1.cb[1], S1.cb[2], S1.cb[3]);
	}
}

template<int K, bool DBG>
__global__ void __launch_bounds__(BT2_FILL_WARPS * 32, K <= 4 ? 8 : 4) k_fill(const DevParams* __restrict__ Pp, const FillArgs a) {
#if BT2_FILL_DYNSMEM
	extern __shared__ __align__(16) unsigned char fill_smem[];
	FillShared<K>* const shs = reinterpret_cast<FillShared<K>*>(fill_smem);
#else
	__shared__ FillShared<K> shs[BT2_FILL_WARPS];
#endif
	const DevParams& P = *Pp;
	const uint32_t warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
	const uint32_t slot0 = (blockIdx.x * BT2_FILL_WARPS + warp) * 4u;
	if(slot0 >= a.nprob) return;
	FillShared<K>& sh = shs[warp];
	// ---- the warp's four problems: set 0 = slots 0, 1 (y = 0), set 1 = slots 2, 3 (y = 1).
	// Phase A: everything the warp reads from global memory, issued together: the problems' headers (one 16-byte load each), the window
	// masks (8 bytes per lane) and the reads' bases and qualities (a byte per lane and load), staged as they are in shared memory that
	// the sweep's tables take over later.  (ncu r2: a third of the warps' stall samples sat in this prologue, waiting on dependent loads
	// it had issued one by one.)
	static_assert(offsetof(Problem, minsc) == 12 && offsetof(Problem, rf) % 8 == 0, "header is read as one uint4, the masks as uint2");
	FillProb pr[4];
	uint4 hd[4]; uint32_t so[4];
	#pragma unroll
	for(int q = 0; q < 4; q++) {
		const bool valid = slot0 + q < a.nprob;
		hd[q] = valid ? *reinterpret_cast<const uint4*>(&a.probs[slot0 + q]) : make_uint4(BT2_NIL, 0, 0, 0);
		so[q] = valid ? a.probs[slot0 + q].seq_off : 0u;
	}
	uint8_t* const raw_rf = reinterpret_cast<uint8_t*>(&sh.hl[0][0]);        // [4][BT2GPU_MAX_COLS + 8] masks as the extension step left them
	uint8_t* const raw_sq = reinterpret_cast<uint8_t*>(&sh.refw[0][0]);      // [4][2][BT2GPU_MAX_ROWS] bases, qualities (forward read)
	uint8_t* const pentab = raw_rf + 4 * (BT2GPU_MAX_COLS + 8);             // [512] mismatch and N penalties by quality (a lane copies 16 bytes)
	static_assert(sizeof(sh.hl) >= 4 * (BT2GPU_MAX_COLS + 8) + 512 && sizeof(sh.refw) >= 8 * BT2GPU_MAX_ROWS, "staging fits the arrays it borrows");
	static_assert(offsetof(DevParams, mmpens) % 16 == 0 && offsetof(DevParams, npens) == offsetof(DevParams, mmpens) + 256 && offsetof(FillShared<K>, hl) % 16 == 0
	              && (4 * (BT2GPU_MAX_COLS + 8)) % 16 == 0 && sizeof(FillShared<K>) % 16 == 0, "the penalty tables are copied as one uint4 per lane");
	const uint4 pent = *reinterpret_cast<const uint4*>(P.mmpens + 16u * lane);
	#pragma unroll
	for(int q = 0; q < 4; q++) {
		FillProb& X = pr[q];
		X.on = hd[q].x != BT2_NIL;
		X.rows = X.on ? (int)(hd[q].y & 0xffffu) : 0; X.cols = X.on ? (int)(hd[q].y >> 16) : 0; X.d0 = 0;
		X.fw = (hd[q].z & 0xffu) != 0; X.start = (X.on && (hd[q].z & 0x4000u)) ? BT2_I16_START : 0xff; X.minsc = (int)hd[q].w;
	}
#if BT2GPU_MAX_ROWS <= 160 && BT2GPU_MAX_COLS + 8 <= 256
	{ // loads first, all of them; the stores into the staging area (which wait for their loads) after
		constexpr int NIT = (BT2GPU_MAX_ROWS + BT2_LANES - 1) / BT2_LANES;
		uint2 mk[4]; uint8_t sb[4][NIT], qb[4][NIT];
		#pragma unroll
		for(int q = 0; q < 4; q++) {
			const FillProb& X = pr[q];
			mk[q] = (X.on && lane * 8u < (uint32_t)X.cols + 1u) ? *reinterpret_cast<const uint2*>(a.probs[slot0 + q].rf + lane * 8u) : make_uint2(0, 0);
			#pragma unroll
			for(int it = 0; it < NIT; it++) {
				const int r = (int)lane + it * BT2_LANES;
				const bool in = X.on && r < X.rows;
				sb[q][it] = in ? a.seq[so[q] + r] : (uint8_t)0; qb[q][it] = in ? a.qual[so[q] + r] : (uint8_t)0;
			}
		}
		#pragma unroll
		for(int q = 0; q < 4; q++) {
			const FillProb& X = pr[q];
			if(!X.on) continue;
			if(lane * 8u < (uint32_t)X.cols + 1u) *reinterpret_cast<uint2*>(raw_rf + q * (BT2GPU_MAX_COLS + 8) + lane * 8u) = mk[q];
			#pragma unroll
			for(int it = 0; it < NIT; it++) {
				const int r = (int)lane + it * BT2_LANES;
				if(r < X.rows) { raw_sq[(2 * q) * BT2GPU_MAX_ROWS + r] = sb[q][it]; raw_sq[(2 * q + 1) * BT2GPU_MAX_ROWS + r] = qb[q][it]; }
			}
		}
	}
#else
	#pragma unroll
	for(int q = 0; q < 4; q++) {   // (the long-read build: plain loops, a warp's prologue is a small part of a 1000-row sweep)
		const FillProb& X = pr[q];
		if(!X.on) continue;
		for(uint32_t c = lane * 8u; c < (uint32_t)X.cols + 1u; c += BT2_LANES * 8u)
			*reinterpret_cast<uint2*>(raw_rf + q * (BT2GPU_MAX_COLS + 8) + c) = *reinterpret_cast<const uint2*>(a.probs[slot0 + q].rf + c);
		for(int r = (int)lane; r < X.rows; r += BT2_LANES) { raw_sq[(2 * q) * BT2GPU_MAX_ROWS + r] = a.seq[so[q] + r]; raw_sq[(2 * q + 1) * BT2GPU_MAX_ROWS + r] = a.qual[so[q] + r]; }
	}
#endif
	*reinterpret_cast<uint4*>(pentab + 16u * lane) = pent;
	int rowsW = 0, widthW = 1; bool any = false;
	#pragma unroll
	for(int q = 0; q < 4; q++) {
		FillProb& X = pr[q];
		if(!X.on) continue;
		any = true;
		int dlo, dhi; fill_band(P, X.rows, X.cols, X.minsc, a.exact != 0, dlo, dhi);
		if(dhi - dlo + 1 > BT2_LANES * K) { if(lane == 0) atomicOr(a.error, 64u); dhi = dlo + BT2_LANES * K - 1; }
		X.d0 = dlo;
		if(X.rows > rowsW) rowsW = X.rows;
		if(dhi - dlo + 1 > widthW) widthW = dhi - dlo + 1;
		if(lane == 0) { a.probs[slot0 + q].lay = (uint8_t)(((hd[q].z >> 8) & 0x40u) | K | ((q >> 1) << 7)); a.probs[slot0 + q].d0 = (int16_t)dlo; }
	}
	if(!any) return;
	constexpr int PER = 16 / K;
	// lanes past the widest band hold no cell anyone reads: the sweep ends when the last lane that does has done its last row
	const int vlast = (widthW - 1) / K;
	int iters = ((rowsW + ((vlast + 2) >> 1) + PER - 1) / PER) * PER;
	if(iters > (int)fill_iters((uint32_t)rowsW, K)) iters = (int)fill_iters((uint32_t)rowsW, K);
	__syncwarp();
	// ---- phase B: the sweep's tables out of the staged bytes: row tables first (they read the bases the window table will overwrite)
	#pragma unroll
	for(int set = 0; set < 2; set++)
		fill_build_rows<K>(P, pentab, sh, set, pr[2 * set], pr[2 * set + 1],
		                   raw_sq + (4 * set) * BT2GPU_MAX_ROWS, raw_sq + (4 * set + 1) * BT2GPU_MAX_ROWS, pr[2 * set].fw,
		                   raw_sq + (4 * set + 2) * BT2GPU_MAX_ROWS, raw_sq + (4 * set + 3) * BT2GPU_MAX_ROWS, pr[2 * set + 1].fw, lane, BT2_FILL_ROWPAD + iters + 1);
	__syncwarp();
	bool hasN = false;
	#pragma unroll
	for(int set = 0; set < 2; set++)
		hasN = fill_build_ref<K>(sh, set, pr[2 * set], pr[2 * set + 1], raw_rf + (2 * set) * (BT2GPU_MAX_COLS + 8), raw_rf + (2 * set + 1) * (BT2GPU_MAX_COLS + 8),
		                         lane, BT2_FILL_ROWPAD + iters + BT2_LANES * K + 1) || hasN;
	const bool refHasN = __any_sync(0xffffffffu, hasN);
	if(a.pcache) { // the problems' characters as 4-bit codes, for the backtrace's edits (dev_common.cuh)
		#pragma unroll
		for(int q = 0; q < 4; q++) {
			if(!pr[q].on) continue;
			const int set = q >> 1, x = q & 1;
			uint32_t* pcw = a.pcache + (size_t)(slot0 + q) * BT2_PC_WORDS;
			#pragma unroll
			for(int base = 0; base < (BT2_PC_RF > BT2_PC_RD ? BT2_PC_RF : BT2_PC_RD); base += BT2_LANES) {
				const int w = base + (int)lane;
				uint32_t wr = 0, wf = 0;
				#pragma unroll
				for(int k = 0; k < 8; k++) {
					const int r = w * 8 + k, c = w * 8 + k;
					const uint32_t rc = r < pr[q].rows ? ((sh.row[set][BT2_FILL_ROWPAD + r].rdc2 >> (16 * x)) & 0xffu) : 0u;
					wr |= (rc > 3u ? 4u : rc) << (4 * k);
					uint32_t fc = 0;
					if(c < pr[q].cols) { const uint32_t m = raw_rf[q * (BT2GPU_MAX_COLS + 8) + c]; fc = m > 15 ? 4u : (m >= 8 ? 3u : (m >= 4 ? 2u : (m >= 2 ? 1u : 0u))); }
					wf |= fc << (4 * k);
				}
				if(w < BT2_PC_RD) pcw[w] = wr;
				if(w < BT2_PC_RF) pcw[BT2_PC_RD + w] = wf;
			}
		}
	}
	__syncwarp();
	for(int idx = (int)lane; idx < 4 * (BT2GPU_MAX_COLS + 8); idx += 32) (&sh.hl[0][0])[idx] = 0;   // (the staged masks are done with)
	__syncwarp();
	// ---- per-lane state: slot 0 of every iteration is set (lane & 1), slot 1 the other one
	const int set0 = (int)(lane & 1u), set1 = set0 ^ 1;
	FillSet<K> S0, S1;
	fill_set_init<K>(S0, sh, pr, set0, lane, (int)((lane + 1u) >> 1));
	fill_set_init<K>(S1, sh, pr, set1, lane, (int)(lane >> 1));
	if(DBG) {
		S0.hefA = a.hef_dbg + (size_t)(slot0 + 2 * set0) * BT2_HEF_WORDS; S0.hefB = S0.hefA + BT2_HEF_WORDS;
		S1.hefA = a.hef_dbg + (size_t)(slot0 + 2 * set1) * BT2_HEF_WORDS; S1.hefB = S1.hefA + BT2_HEF_WORDS;
	}
	uint8_t* const cb0 = a.cells + (size_t)slot0 * a.cell_stride;
	auto outp = [&](int q) -> uint4* { return pr[q].on ? reinterpret_cast<uint4*>(cb0 + (size_t)q * a.cell_stride) + lane : nullptr; };
	uint4* const o0a = outp(2 * set0); uint4* const o0b = outp(2 * set0 + 1); uint4* const o1a = outp(2 * set1); uint4* const o1b = outp(2 * set1 + 1);
	const uint32_t neg_rdge = v_pack2(-P.rdGapExt, -P.rdGapExt);
	uint32_t sendE = 0u, sendCE = BT2_FL_BASE | ((uint32_t)(CF_EO | CF_EX) * 0x00010001u), sendF = 0u, sendCF = BT2_FL_BASE | ((uint32_t)(CF_FO | CF_FX) * 0x00010001u);
	#define BT2_SWEEP_ARGS S0, S1, lane, neg_rdge, sendE, sendCE, sendF, sendCF, o0a, o0b, o1a, o1b
	if(refHasN) fill_sweep<K, true, DBG>(0, iters, BT2_SWEEP_ARGS);
	else fill_sweep<K, false, DBG>(0, iters, BT2_SWEEP_ARGS);
	#undef BT2_SWEEP_ARGS
	fill_last_row<K>(S0); fill_last_row<K>(S1);
	__syncwarp();
	// ---- last-row candidates: H-0xff >= minsc, sorted by (score desc, col desc)
	uint32_t* const ckey = reinterpret_cast<uint32_t*>(sh.row2[0]);
	#pragma unroll
	for(int q = 0; q < 4; q++) {
		if(!pr[q].on) continue;
		Problem& pb = a.probs[slot0 + q];
		const int cols = pr[q].cols;
		const int minsc = pr[q].minsc;
		int lrmax = 0; uint32_t ncand = 0;
		for(int c0 = 0; c0 < cols; c0 += 32) {
			const int c = c0 + (int)lane;
			const int hv8 = c < cols ? (int)sh.hl[q][c] : 0;
			lrmax = max(lrmax, hv8);
			const bool isc = c < cols && (hv8 - pr[q].start) >= minsc;
			const uint32_t bal = __ballot_sync(0xffffffffu, isc);
			if(isc) ckey[ncand + __popc(bal & ((1u << lane) - 1))] = ((uint32_t)hv8 << BT2_CAND_SHIFT) | (uint32_t)c;
			ncand += __popc(bal);
		}
		#pragma unroll
		for(int o = 16; o > 0; o >>= 1) lrmax = max(lrmax, __shfl_xor_sync(0xffffffffu, lrmax, o));
		__syncwarp();
		for(uint32_t j = lane; j < ncand; j += 32) {
			const uint32_t key = ckey[j];
			uint32_t rank = 0;
			for(uint32_t k = 0; k < ncand; k++) rank += (ckey[k] > key) ? 1u : 0u;
			pb.cand[rank] = key;
		}
		__syncwarp();
		if(lane == 0) {
			pb.lrmax = lrmax; pb.ncand = ncand;
			atomicAdd(&a.stats[0], 1ull); atomicAdd(&a.stats[1], (unsigned long long)pr[q].rows * cols);
			atomicAdd(&a.stats[2], (unsigned long long)pr[q].rows * (BT2_LANES * K));
		}
	}
}

#endif // __CUDACC__

} // namespace bt2
