// Data model and scalar primitives shared by the sm_100a kernels of libbt2gpu.so.
//
// Everything here is plain integer code marked BT2_HD so that tests/emu can
// compile the per-read state machines with g++ and step them against the
// oracle without a GPU; the product library only ever runs them inside CUDA
// kernels (kernels.cu).
//
// Reference semantics followed (file:line into sfiligoi/omp-bowtie2-prime):
//   RandomSource                      random_source.h:33-150
//   Ebwt::countBt2Side / rowL / mapLF bt2_idx.h:1811-1840, 2306, 2511-2564
//   Ebwt::tryOffset / getOffset       bt2_idx.h:1621-1634, bt2_idx.cpp:147-172
//   Ebwt::joinedToTextOff             bt2_idx.cpp:54-124
//   BitPairReference::getStretch      reference.cpp:330+
//   Scoring::maxReadGaps/maxRefGaps   scoring.cpp:42-98
//   DynProgFramer::frameSeedExtensionRect dp_framer.cpp:81-125
#pragma once
#include <cstddef>
#include <stdint.h>
#include "bt2gpu.h"

#if defined(__CUDACC__)
#define BT2_HD __host__ __device__ __forceinline__
#define BT2_D  __device__ __forceinline__
#else
#define BT2_HD inline
#endif

namespace bt2 {

static const uint64_t OFF_MASK_U64 = ~0ull;
#define BT2_I32_MIN (-2147483647 - 1)

// ---------------------------------------------------------------- parameters
struct DevParams {
	int32_t seedLen, seedMms, nSeedRounds, seedBoostThresh, maxhalf, mxIter, mxUg, mxDp, streak;
	int32_t doExtend, enable8, tighten, khits, mhits, msample, nofw, norc, gapbar;
	int32_t rdGapOpen, rdGapExt, rfGapOpen, rfGapExt, matchBonus, i16;
	uint8_t mmpens[256]; // Scoring::mmpens (<= 255 in u8 mode)
	uint8_t npens[256];
};

// ---------------------------------------------------------------- index in HBM
struct DevIndex {
	const uint8_t*  ebwt;     // sides of 128 B: 96 B BWT (384 bp, 2 bit) + 4 x u64 occ
	const uint64_t* ftab;
	const uint64_t* eftab;
	const uint64_t* offs;
	const uint64_t* rstarts;  // 3 per fragment
	const uint64_t* plen;
	const uint32_t* sa32;     // device-built full suffix array (one entry per BWT row), or null (see k_index.cuh)
	const uint32_t* ktab;     // device-built prefix table: first row whose suffix starts with a ktabChars-mer >= the key, or null
	int32_t  ktabChars;
	uint64_t fchr[5];
	uint64_t len, zOff, nFrag, nPat;
	int32_t  offRate, ftabChars;
};
#define BT2_SIDE_SZ      128u
#define BT2_SIDE_BWT_SZ   96u
#define BT2_SIDE_BWT_LEN 384u

struct DevRef {
	const uint8_t*  buf;        // 2-bit unambiguous bases
	const uint64_t* rec_off;    // [nrecs]
	const uint64_t* rec_len;    // [nrecs]
	const uint64_t* rec_tstart; // [nrecs] text position (within its reference) where the record's N-gap starts
	const uint64_t* rec_cum;    // [nrecs] # unambiguous bases before the record (index into buf)
	const uint64_t* ref_rec0;   // [nPat+1] first record of each reference
	uint64_t nrecs;
};

// ---------------------------------------------------------------- RNG
struct Rng {
	uint32_t last, lastOff;
	BT2_HD void init(uint32_t seed) { last = seed; lastOff = 30; }
	BT2_HD uint32_t nextU32() {
		last = 1664525u * last + 1013904223u;
		uint32_t ret = last >> 16;
		last = 1664525u * last + 1013904223u;
		ret ^= last;
		lastOff = 0;
		return ret;
	}
	BT2_HD bool nextBool() {
		if(lastOff > 31) nextU32();
		uint32_t ret = (last >> lastOff) & 1;
		lastOff++;
		return ret != 0;
	}
	BT2_HD float nextFloat() {
#if defined(__CUDA_ARCH__)
		return __fdiv_rn((float)nextU32(), (float)0xffffffff);
#else
		return (float)nextU32() / (float)0xffffffff;
#endif
	}
};

// ---------------------------------------------------------------- FM primitives (scalar)
BT2_HD uint64_t load_u64(const uint8_t* p) {
	return *reinterpret_cast<const uint64_t*>(p); // sides are 128-byte aligned, occ at +96
}

// # of code c among the first nb (<=32) bases of a 64-bit word of 2-bit codes
BT2_HD uint32_t count_in_word(uint64_t w, int c, uint32_t nb) {
	if(nb == 0) return 0;
	const uint64_t x = w ^ (0x5555555555555555ull * (uint64_t)c); // 00 where the base equals c
	uint64_t m = ~(x | (x >> 1)) & 0x5555555555555555ull;
	if(nb < 32) m &= ((1ull << (2 * nb)) - 1);
#if defined(__CUDA_ARCH__)
	return (uint32_t)__popcll(m);
#else
	return (uint32_t)__builtin_popcountll(m);
#endif
}

BT2_HD int row_L(const DevIndex& ix, uint64_t row) {
	const uint64_t side = row / BT2_SIDE_BWT_LEN;
	const uint32_t co = (uint32_t)(row - side * BT2_SIDE_BWT_LEN);
	const uint8_t b = ix.ebwt[side * BT2_SIDE_SZ + (co >> 2)];
	return (b >> ((co & 3) * 2)) & 3;
}

// rank of c before `row` (not counting '$') + fchr[c]  == Ebwt::countBt2Side
BT2_HD uint64_t rank_c(const DevIndex& ix, int c, uint64_t row) {
	const uint64_t side = row / BT2_SIDE_BWT_LEN;
	const uint32_t co = (uint32_t)(row - side * BT2_SIDE_BWT_LEN);
	const uint8_t* s = ix.ebwt + side * BT2_SIDE_SZ;
	uint32_t cnt = 0;
	const uint32_t nw = (co + 31) >> 5;
	for(uint32_t w = 0; w < nw; w++) {
		const uint32_t rem = co - w * 32;
		cnt += count_in_word(load_u64(s + 8 * w), c, rem < 32 ? rem : 32);
	}
	if(c == 0) {
		const uint64_t sideStart = side * BT2_SIDE_BWT_LEN;
		if(ix.zOff >= sideStart && ix.zOff < row) cnt--;
	}
	return (uint64_t)cnt + load_u64(s + BT2_SIDE_BWT_SZ + 8 * c) + ix.fchr[c];
}

// SA[row] by LF-walking to the next sampled row (== what GroupWalk2S::advanceElement resolves)
BT2_HD uint64_t sa_offset(const DevIndex& ix, uint64_t row, uint32_t* steps) {
	if(ix.sa32) return ix.sa32[row];
	uint64_t jumps = 0;
	const uint64_t offMask = OFF_MASK_U64 << ix.offRate;
	while(true) {
		if(row == ix.zOff) { *steps += (uint32_t)jumps; return jumps; }
		if((row & offMask) == row) { *steps += (uint32_t)jumps; return jumps + ix.offs[row >> ix.offRate]; }
		row = rank_c(ix, row_L(ix, row), row);
		jumps++;
	}
}

BT2_HD void joined_to_text(const DevIndex& ix, uint64_t off, uint64_t& tidx, uint64_t& textoff, uint64_t& tlen) {
	uint64_t top = 0, bot = ix.nFrag;
	while(true) {
		const uint64_t elt = top + ((bot - top) >> 1);
		const uint64_t lower = ix.rstarts[elt * 3];
		const uint64_t upper = (elt == ix.nFrag - 1) ? ix.len : ix.rstarts[(elt + 1) * 3];
		if(lower <= off) {
			if(upper > off) {
				tidx = ix.rstarts[elt * 3 + 1];
				textoff = (off - lower) + ix.rstarts[elt * 3 + 2];
				break;
			}
			top = elt;
		} else bot = elt;
	}
	tlen = ix.plen[tidx];
}

// dst[0..count) <- reference masks (1<<base, N -> 16) of reference tidx from toff
BT2_HD void ref_masks(const DevRef& rf, uint8_t* dst, uint64_t tidx, uint64_t toff, uint32_t count) {
	uint64_t lo = rf.ref_rec0[tidx], hi = rf.ref_rec0[tidx + 1];
	const uint64_t last = hi;
	// first record whose end lies beyond toff
	while(lo < hi) {
		const uint64_t mid = (lo + hi) >> 1;
		const uint64_t recEnd = rf.rec_tstart[mid] + rf.rec_off[mid] + rf.rec_len[mid];
		if(recEnd > toff) hi = mid; else lo = mid + 1;
	}
	uint32_t k = 0;
	const uint64_t end = toff + count;
#if defined(__CUDA_ARCH__)
	// The usual window lies inside one record: 16 bases per step -- two words of the packed text funnel-shifted to the window's phase, four
	// codes -> four mask bytes by one PRMT (byte c of 0x08040201 = 1 << c), one 16-byte store.  (The byte loop below, 200 loads and 200
	// one-byte stores per problem, was most of k_step.)  Writes up to 15 bytes past `count` inside the caller's array: Problem::rf holds
	// a multiple of 16 bytes beyond every window; nobody reads past the window.  upload() pads the packed text by 16 bytes.
	static_assert((BT2GPU_MAX_COLS + 15) / 16 * 16 <= BT2GPU_MAX_COLS + 8, "a window rounded up to 16 masks fits Problem::rf");
	if(lo < last && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0u) {
		const uint64_t gapEnd = rf.rec_tstart[lo] + rf.rec_off[lo], recEnd = gapEnd + rf.rec_len[lo];
		if(toff >= gapEnd && end <= recEnd) {
			const uint32_t* __restrict__ w = reinterpret_cast<const uint32_t*>(rf.buf);
			const uint64_t bi = rf.rec_cum[lo] + (toff - gapEnd);
			const uint32_t sh = (uint32_t)(bi & 15u) * 2u;
			uint64_t wi = bi >> 4;
			uint32_t w0 = w[wi];
			for(; k < count; k += 16, wi++) {
				const uint32_t w1 = sh ? w[wi + 1] : 0u;
				const uint32_t codes = __funnelshift_r(w0, w1, sh);
				uint4 m;
				#define BT2_M4(q) __byte_perm(0x08040201u, 0u, ((codes >> (8 * (q))) & 3u) | (((codes >> (8 * (q) + 2)) & 3u) << 4) | (((codes >> (8 * (q) + 4)) & 3u) << 8) | (((codes >> (8 * (q) + 6)) & 3u) << 12))
				m.x = BT2_M4(0); m.y = BT2_M4(1); m.z = BT2_M4(2); m.w = BT2_M4(3);
				#undef BT2_M4
				*reinterpret_cast<uint4*>(dst + k) = m;
				w0 = sh ? w1 : w[wi + 1];
			}
			return;
		}
	}
#endif
	for(uint64_t r = lo; r < last && k < count; r++) {
		const uint64_t gapEnd = rf.rec_tstart[r] + rf.rec_off[r];
		const uint64_t recEnd = gapEnd + rf.rec_len[r];
		for(uint64_t t = toff + k; t < gapEnd && t < end; t++) dst[k++] = 16;
		for(uint64_t t = toff + k; t < recEnd && t < end; t++) {
			const uint64_t bi = rf.rec_cum[r] + (t - gapEnd);
			dst[k++] = (uint8_t)(1u << ((rf.buf[bi >> 2] >> ((bi & 3) * 2)) & 3));
		}
	}
	while(k < count) dst[k++] = 16;
}

// ---------------------------------------------------------------- scoring scalars
BT2_HD int max_read_gaps(const DevParams& P, int64_t minsc, uint32_t rdlen) {
	int64_t sc = (int64_t)rdlen * P.matchBonus; bool first = true; int num = 0;
	while(sc >= minsc) { if(first) { first = false; sc -= P.rdGapOpen; } else sc -= P.rdGapExt; num++; }
	return num - 1;
}
BT2_HD int max_ref_gaps(const DevParams& P, int64_t minsc, uint32_t rdlen) {
	int64_t sc = (int64_t)rdlen * P.matchBonus; bool first = true; int num = 0;
	while(sc >= minsc) { sc -= P.matchBonus; if(first) { first = false; sc -= P.rfGapOpen; } else sc -= P.rfGapExt; num++; }
	return num - 1;
}

// ---------------------------------------------------------------- per-read records in HBM
// K_seed output, one per (read, strand, offset index)
struct SeedHit {
	uint64_t top;     // BWT range [top, top+len)
	uint64_t len;     // 0 = no hit
	uint64_t key;     // SAKey::seq (2-bit packed seed, first base most significant)
	uint32_t nlex;    // left-extension length (<=255)
	uint32_t valid;   // seed was searched (no N) and has a non-empty range
};

// one entry of SwDriver::satpos_ (+ its Random1toN of at most nsm=5 rows)
struct SatEntry {
	uint64_t topf;
	uint32_t size;    // rows in this entry (<=5, or 1 for a sampled row)
	uint16_t rdoff;   // seed offset in the read (5' end, fw coordinates)
	uint8_t  fw;
	uint8_t  cur;     // Random1toN::cur_
	uint32_t perm;    // Random1toN swap list, 3 bits per slot, valid once cur>0 || inited
	uint32_t pad;
};

struct Interval16 { int64_t off; uint32_t ref; int32_t len_fw; }; // len_fw = len*2 + fw
struct ExRange16  { uint64_t sz; uint32_t off; uint32_t len; };

// generic chunked list in a pool: 7 items of 16 B + link
struct Chunk { uint64_t item[7][2]; uint32_t next; uint32_t n; uint32_t pad[2]; };
#define BT2_CHUNK_ITEMS 7
#define BT2_NIL 0xffffffffu

struct AlnRec {           // one reported alignment (AlnRes fields the sink consumes)
	int64_t  refoff;
	uint32_t tidx;
	int32_t  score, minsc, ns, gaps, edits, basesAligned, refns;
	uint32_t edit_off;    // into the edit pool
	uint16_t n_edits;
	uint8_t  fw, pad;
	uint32_t prev;        // previous alignment of the same read (BT2_NIL = none)
	uint32_t pad2;
};

// packed Edit: pos[0:9] chr[10:17] qchr[18:25] type[26:27]
BT2_HD uint32_t pack_edit(uint32_t pos, uint32_t chr, uint32_t qchr, uint32_t type) {
	return (pos & 1023u) | (chr << 10) | (qchr << 18) | (type << 26);
}
BT2_HD uint32_t edit_pos(uint32_t e)  { return e & 1023u; }
BT2_HD uint32_t edit_chr(uint32_t e)  { return (e >> 10) & 255u; }
BT2_HD uint32_t edit_qchr(uint32_t e) { return (e >> 18) & 255u; }
BT2_HD uint32_t edit_type(uint32_t e) { return (e >> 26) & 3u; }
BT2_HD uint32_t edit_with_pos(uint32_t e, uint32_t pos) { return (e & ~1023u) | (pos & 1023u); }
#define BT2_SCRATCH_EDITS (((BT2GPU_MAX_ROWS + 96 + 63) / 64) * 64)   // >= DList<Edit,ALN_MAX_ROWS> capacity actually reachable (256 for 160 rows)
static_assert(BT2GPU_MAX_ROWS <= 1008 && BT2GPU_MAX_COLS <= 4080, "a packed edit holds 10 bits of read position, a candidate 12 bits of column");

enum { RS_IDLE = 0, RS_EXTEND = 1, RS_ROUND_DONE = 2, RS_FINISHED = 3 };

struct ReadState {
	Rng      rnd;
	int32_t  minsc;          // current (tightened) minimum score
	int32_t  nceil;
	int32_t  interval;
	uint32_t seq_off;        // into the batch's seq/qual arrays
	uint16_t rdlen;
	uint8_t  phase;          // RS_*
	uint8_t  round;          // seed round being run
	uint8_t  status, flags, roundsRun, donePolicy;
	// sink state read back by extendSeeds (ReportingState + best/second best)
	uint32_t nunpair;
	int32_t  best, best2;    // BT2_I32_MIN = invalid
	uint8_t  doneUnpair, pad0[3];
	// PerReadMetrics
	uint32_t nExIters, nExDps, nExDpSuccs, nExDpFails, nDpFail, nDpFailStreak, nDpLastSucc, nRedundants;
	int32_t  bestLtMinsc;
	// seed summary of the last round
	uint32_t nonzTot, numRanges, nSeedsTried;
	uint64_t numElts;
	// extension progress (SwDriver::extendSeeds loop variables)
	uint32_t sat_off, sat_n; // entries in the SatEntry pool
	uint32_t cur_i;          // index into satpos; == sat_n means "start of a pass"
	uint32_t neltLeft;
	uint8_t  first, passStarted, pad1[2];
	// lists
	uint32_t seen_head;      // Chunk list of Interval16 (EIvalMergeListBinned seenDiags1_)
	uint32_t exfw_head, exrc_head; // Chunk lists of ExRange16 (seedExRangeFw_/Rc_)
	uint32_t aln_tail, n_alns;
	// F14 scratch bookkeeping (contents in a separate array)
	uint16_t scr_cur, scr_hw;
	// work counters
	uint32_t lfResolve, saRows, btAttempts, btSteps;
	uint64_t dpCells;
};

// DPRect + SwAligner::initRef outputs for one DP problem, and what the fill leaves for the backtrace
struct alignas(16) Problem {
	uint32_t read;
	uint16_t rows, cols;       // rdlen, rflen (# DP columns)
	uint8_t  fw, lay;          // lay = K | i16 << 6 | y << 7: the fill's layout (below): K diagonals per lane, y = which of the warp's two interleaved
	                           // pairs; i16 (set by the extension step): the reference takes its 16-bit DP path for this problem (prob_i16)
	int16_t  d0;               // leftmost diagonal (col - row) of the band the fill computed
	int32_t  minsc;            // SwAligner::minsc_ (clamped to >= -254 on the 8-bit path)
	int64_t  refl;             // rect.refl
	uint32_t tidx;
	int32_t  triml, corel, corer;
	// outputs of K_fill
	int32_t  lrmax;            // max raw H in the last row
	uint32_t ncand;
	uint32_t seq_off, padq[3]; // the read's offset into the batch's seq / qual arrays (ReadState::seq_off: one load less on the fill's way to them)
	uint32_t cand[BT2GPU_MAX_COLS + 8]; // sorted best-first: (rawH << 8) | col; at a 16-byte boundary (read in groups of 4 by the backtrace)
	uint8_t  rf[BT2GPU_MAX_COLS + 8];   // reference masks for columns 0..cols (cols+1 fetched)
};
BT2_HD uint32_t prob_K(const Problem& p) { return p.lay & 0x3fu; }
// The reference built with -DENABLE_I16 runs a problem through its 16-bit fill when 8-bit SSE is off (--no-sse8) or the minimum score is
// below -254 (aligner_sw.cpp:77-83).  The recurrence is the same with another bias and floor (start 0x7fff, floor -0x8000 instead of 0xff
// and 0); here the scores live in signed 16-bit lanes either way, so the 16-bit path only changes H of row -1 -- BT2_I16_START instead of
// 0xff: the largest bias whose scores still compare exactly as fp16 bit patterns (simd16.cuh), floor at a score of -28 672, far below
// any minimum score a read of <= 160 bp can have -- the candidate threshold, and the protocol around the fill: nextAlignment reseeds the
// generator only after the backtrace (aligner_sw.cpp:336-348) and `best` is the raw 16-bit word minus 0xff (aligner_swsse_ee_i16.cpp:640).
#define BT2_I16_START 0x7000
BT2_HD bool prob_i16(const Problem& p) { return (p.lay & 0x40u) != 0; }
BT2_HD int prob_start(const Problem& p) { return prob_i16(p) ? BT2_I16_START : 0xff; }
BT2_HD uint32_t prob_y(const Problem& p) { return p.lay >> 7; }
static_assert(offsetof(Problem, cand) % 16 == 0 && sizeof(Problem) % 16 == 0, "Problem::cand is copied in 16-byte groups");

// What the fill leaves per DP cell for the backtrace: one byte of "not equal" flags, i.e. the
// outcome of every score comparison backtraceNucleotidesEnd2EndSseU8 makes at that cell
// (aligner_swsse_ee_u8.cpp:820-1100), evaluated while H/E/F and their neighbours are in registers.
// A flag bit is 0 when the equality holds.
//   CF_D   H(r,c) == H(r-1,c-1) + sc(r,c)         CF_HE  H(r,c) == E(r,c)     CF_HF  H(r,c) == F(r,c)
//   CF_EO  E(r,c) == H(r,c-1) - rdGapOpen         CF_EX  E(r,c) == E(r,c-1) - rdGapExtend
//   CF_FO  F(r,c) == H(r-1,c) - rfGapOpen         CF_FX  F(r,c) == F(r-1,c) - rfGapExtend
//   CF_VIS (bit 7) is not a comparison: the fill writes 0, the backtrace sets it when it visits the cell (dev_bt.cuh)
// The H-cell's gap-move comparisons follow from these: in a row where gaps are allowed
// F = max(F_up - ext, H_up - open, 0), so H == H_up - open iff (H == F and F == H_up - open), and
// H != F implies H > both; likewise for E.  (Rows inside the gap barrier never look at them.)
// E/F flags of rows inside the gap barrier are unspecified: no backtrace can be in E or F state
// there (E/F are 0 in those rows and every cell on a backtrace path has a raw score >= 1).
enum { CF_D = 1, CF_HE = 2, CF_HF = 4, CF_EO = 8, CF_EX = 16, CF_FO = 32, CF_FX = 64, CF_VIS = 128 };
// Which cells are filled.  A backtrace only ever stands on cells that lie on a full path (row 0 .. last row inside the window)
// scoring >= minsc: it starts from a last-row cell with H >= minsc and every move follows a score equality, so the value at
// the cell it stands on plus the penalties already walked is that candidate's score.  A path scoring >= minsc holds at most
// g = fill_gap_bound() gap positions, starts in a column >= 0 and ends in a column < cols, hence never leaves the band of
// diagonals  -g <= col - row <= cols - rows + g.  Values outside the band are taken as 0 (the saturation floor): every value inside
// is then <= the full rectangle's and equal to it wherever a full path scoring >= minsc passes, and so is every comparison a
// backtrace reads (an equality that holds in the full rectangle continues such a path; one that fails compares with a value that
// can only have become smaller).  Candidates, alignments, edits, the visited marks and the random draws are those of the full
// rectangle; the one observable difference is the best score of a window WITHOUT any candidate (prm.bestLtMinscMate1), which
// becomes a lower bound -- bt2gpu_params.exactBestLtMinsc fills the whole rectangle instead.
//
// Layout.  The fill's lane v owns the K consecutive diagonals d0 + v K .. d0 + v K + K-1 and sweeps the rows; lane v works on
// row r at iteration i = r + ((v + y) >> 1) (two steps per row and lane: the cell above-right comes from lane v+1, the cell to
// the left from lane v-1; a warp interleaves two pairs of problems, y = 0 / 1, on the alternate steps).  The K cells of a lane
// and iteration form a unit of K bytes, 16 / K consecutive iterations a 16-byte chunk, the chunks of the 32 lanes are adjacent:
//   dd = col - row - d0,  v = dd / K,  j = dd % K,  i = row + ((v + y) >> 1),  per = 16 / K
//   byte index = ((i / per) * 32 + v) * 16 + (i % per) * K + j
// The fill writes whole 512-byte rows (every lane of the warp stores its chunk of the same iterations at the same time); a
// backtrace that follows a diagonal stays with one lane and crosses into a new chunk every 16 / K rows.
#define BT2_PK_WORDS ((BT2GPU_MAX_ROWS + 15) / 16 + 2)   // 2-bit packed read: 16 bases per word (+ words a 32-base window may touch); 12 for 160 rows
#define BT2_NM_WORDS ((BT2GPU_MAX_ROWS + 31) / 32 + 1)   // its N mask: 32 bases per word; 6 for 160 rows
#define BT2_LANES 32
#define BT2_SKEW  (BT2_LANES / 2)           // iterations the last lane lags behind the first
#define BT2_KMAX  16
BT2_HD uint32_t fill_lgK(uint32_t K) { return K == 2 ? 1u : (K == 4 ? 2u : (K == 8 ? 3u : 4u)); }
// iterations of the sweep / bytes of cell flags of one problem slot
BT2_HD uint32_t fill_iters(uint32_t rows, uint32_t K) { const uint32_t per = 16u / K; return ((rows + BT2_SKEW + per - 1u) / per) * per; }
BT2_HD size_t fill_cell_bytes(uint32_t rows, uint32_t K) { return (size_t)(fill_iters(rows, K) / (16u / K)) * BT2_LANES * 16u; }
BT2_HD uint32_t cell_index(uint32_t row, uint32_t col, int d0, uint32_t K, uint32_t y) {
	const uint32_t lg = fill_lgK(K), dd = (uint32_t)((int)col - (int)row - d0), v = dd >> lg, j = dd & (K - 1u);
	const uint32_t i = row + ((v + y) >> 1);
	return ((((i >> (4u - lg)) << 5) + v) << 4) + ((i & ((16u >> lg) - 1u)) << lg) + j;
}
// most gap positions a path scoring >= minsc can hold (perfect score 0: end-to-end, no match bonus)
BT2_HD int fill_gap_bound(const DevParams& P, int minsc) {
	const int omin = P.rdGapOpen < P.rfGapOpen ? P.rdGapOpen : P.rfGapOpen;
	int emin = P.rdGapExt < P.rfGapExt ? P.rdGapExt : P.rfGapExt; if(omin < emin) emin = omin;
	const int budget = -minsc;
	if(budget < omin) return 0;
	if(emin <= 0) return 1 << 20;
	return 1 + (budget - omin) / emin;
}
// band of diagonals [dlo, dhi] of a rows x cols problem
BT2_HD void fill_band(const DevParams& P, int rows, int cols, int minsc, bool exact, int& dlo, int& dhi) {
	dlo = -(rows - 1); dhi = cols - 1;
	if(exact) return;
	const int g = fill_gap_bound(P, minsc);
	if(-g > dlo) dlo = -g;
	if((cols - rows) + g < dhi) dhi = (cols - rows) + g;
	if(dhi < dlo) dhi = dlo;       // no path fits the window: nothing can be a candidate, any band will do
}
BT2_HD uint32_t fill_K_for(int width) { return width <= 64 ? 2u : (width <= 128 ? 4u : (width <= 256 ? 8u : 16u)); }
// debug-only H|E<<8|F<<16 words of the fill's test instantiation: [row][dd], dd = col - row - d0 < 32 K (bt2gpu_test_fill)
#define BT2_HEF_WORDS (BT2GPU_MAX_ROWS * BT2_LANES * BT2_KMAX)

// Per-problem character cache handed from the fill to the backtrace: 4-bit codes, 8 per word.
//   words [0, BT2_PC_RD)            read characters in the orientation of the problem's strand (0..3, 4 = N)
//   words [BT2_PC_RD, BT2_PC_WORDS) reference characters of the window's columns (0..3, 4 = N)
#define BT2_PC_RD    ((BT2GPU_MAX_ROWS + 7) / 8)
#define BT2_PC_RF    ((BT2GPU_MAX_COLS + 8 + 7) / 8)
#define BT2_PC_WORDS (((BT2_PC_RD + BT2_PC_RF + 3) / 4) * 4)   // (copied in 16-byte groups; 48 for 160 x 200)
#define BT2_CAND_SHIFT (BT2GPU_MAX_COLS > 255 ? 12 : 8)       // a candidate key: raw H << BT2_CAND_SHIFT | column
BT2_HD uint32_t pc_read_code(const uint32_t* pc, uint32_t stride, uint32_t row) { return (pc[(row >> 3) * stride] >> ((row & 7u) * 4u)) & 15u; }
BT2_HD uint32_t pc_ref_code(const uint32_t* pc, uint32_t stride, uint32_t col) { return (pc[(BT2_PC_RD + (col >> 3)) * stride] >> ((col & 7u) * 4u)) & 15u; }

// pools shared by the per-read kernels
struct Pools {
	Chunk*    chunks;   uint32_t* chunk_top;  uint32_t chunk_cap;
	SatEntry* sats;     uint32_t* sat_top;    uint32_t sat_cap;
	AlnRec*   alns;     uint32_t* aln_top;    uint32_t aln_cap;
	uint32_t* edits;    uint32_t* edit_top;   uint32_t edit_cap;
	uint32_t* error;    // sticky: 1 chunk pool, 2 sat pool, 4 aln pool, 8 edit pool exhausted
};

BT2_HD uint32_t atomic_bump(uint32_t* ctr, uint32_t n) {
#if defined(__CUDA_ARCH__)
	return atomicAdd(ctr, n);
#else
	uint32_t v = *ctr; *ctr = v + n; return v;
#endif
}
BT2_HD void set_error(const Pools& p, uint32_t bit) {
#if defined(__CUDA_ARCH__)
	atomicOr(p.error, bit);
#else
	*p.error |= bit;
#endif
}

// append a 16-byte item to a chunk list; returns false if the pool is exhausted
BT2_HD bool list_push(const Pools& p, uint32_t& head, uint64_t a, uint64_t b) {
	if(head == BT2_NIL || p.chunks[head].n == BT2_CHUNK_ITEMS) {
		const uint32_t c = atomic_bump(p.chunk_top, 1);
		if(c >= p.chunk_cap) { set_error(p, 1); return false; }
		p.chunks[c].next = head; p.chunks[c].n = 0;
		head = c;
	}
	Chunk& ch = p.chunks[head];
	ch.item[ch.n][0] = a; ch.item[ch.n][1] = b; ch.n++;
	return true;
}

} // namespace bt2
