// What follows a DP fill inside SwDriver::extendSeeds:
//   result bookkeeping                               aligner_sw_driver.cpp:505-531
//   SwAligner::nextAlignment                         aligner_sw.cpp:252-375
//   SwAligner::backtraceNucleotidesEnd2EndSseU8      aligner_swsse_ee_u8.cpp:746-1352
//   RedundantAlns::overlap/add                       aligner_result.cpp:821-925
//   AlnSinkWrap::report + ReportingState::foundUnpaired  aln_sink.cpp:1395-1453, 107-140
//   minsc tightening                                 aligner_sw_driver.cpp:588-618
// The per-read edit scratch reproduces SwDriver::resGap_.alres.ned_.
//
// Written as a flat state machine so that the lanes of a warp, each working on its own
// problem and pulling a new one when done, stay in one loop.  BtMachine::hot() does one
// unit of the frequent work (start a candidate / one backtrace cell) with ONE memory read,
// the 16-byte chunk of the cell array the walk enters (copied to shared memory, the chunk it
// is predicted to enter next copied ahead); everything a cell needs besides that (read and
// reference characters for the edits) comes from a per-lane copy of the problem in shared
// memory.  The paths of the state machine that differ between lanes (which move, whether it
// makes an edit) are computed with selects and share one copy of the code that follows.
//
// What the reference's per-cell backtrace state boils down to: the reference marks a cell
// "reported through" on EVERY visit (aligner_swsse_ee_u8.cpp:1023), before it even knows
// whether the walk can continue, and a marked cell rejects any later visit before its
// H/E/F option masks are looked at (:801-804).  The option masks are therefore only ever
// read on a cell's first visit, where they still hold the fill's comparison outcome, and
// popping a branch frame returns to a cell that is already marked: the whole branch stack
// unwinds one frame per step without trying anything.  So the state is ONE bit per cell
// (visited), and a failed walk costs `number of branch frames` extra steps.  Nothing of a
// popped frame is observable: the edit count after unwinding is discarded by res.reset().
// That bit is bit 7 of the cell's own flag byte (CF_VIS; the fill writes it as 0, so every
// fill of a slot starts with nothing visited): the visit test comes with the flags, and
// marking is one byte store to the chunk's copy and one to the cell array.
#pragma once
#include "dev_common.cuh"

#if defined(__CUDA_ARCH__)
#define BT2_POPC(x) __popc(x)
#define BT2_CLZ(x) __clz(x)
#else
#define BT2_POPC(x) __builtin_popcount(x)
#define BT2_CLZ(x) ((x) ? __builtin_clz(x) : 32)
#endif

// host-only work counters of the emulator build (tests/emu, -DBT2_BT_PROFILE): what the lane iterations of k_bt are spent on
#if defined(BT2_BT_PROFILE) && !defined(__CUDA_ARCH__)
#define BT2_PROF(i, n) (bt2::g_bt_prof[i] += (n))
namespace bt2 { static unsigned long long g_bt_prof[32]; }
#else
#define BT2_PROF(i, n) ((void)0)
#endif

namespace bt2 {

enum { BT_CONTINUE = 0, BT_ROUND_END = 1 };
enum { CT_E = 0, CT_F = 1, CT_H = 2 };
enum { BP_CAND = 0, BP_STEP = 1, BP_POST = 2, BP_DONE = 3 };

struct ReadView {
	const uint8_t* seq; const uint8_t* qual; uint32_t rdlen; bool fw;
	BT2_HD int rd(uint32_t row) const {
		if(fw) return seq[row];
		const uint8_t c = seq[rdlen - 1 - row];
		return c > 3 ? 4 : 3 - c;
	}
	BT2_HD int q(uint32_t row) const { return fw ? qual[row] : qual[rdlen - 1 - row]; }
};

BT2_HD int sc_score(const DevParams& P, int rdc, int refm, int q) {
	if(q < 0) q = 0; if(q > 255) q = 255;
	if(rdc > 3 || refm > 15) return -(int)P.npens[q];
	if((refm & (1 << rdc)) != 0) return P.matchBonus;
	return -(int)P.mmpens[q];
}
// character of a 0..4 code ("ACGTN"), no memory access
BT2_HD uint32_t code_char(uint32_t c) {
#if defined(__CUDA_ARCH__)
	return __byte_perm(0x54474341u, 0x4eu, c > 4u ? 4u : c) & 255u;   // one PRMT
#else
	return (uint32_t)((0x4e54474341ull >> (8u * (c > 4u ? 4u : c))) & 255u);
#endif
}

// left-end-based view of an edit list that may be stored 5'-based for a rc read
struct EditView {
	const uint32_t* e; uint32_t n; bool inv; uint32_t rdlen;
	BT2_HD uint32_t get(uint32_t i) const {
		if(!inv) return e[i];
		const uint32_t x = e[n - 1 - i];
		const uint32_t pos = rdlen - edit_pos(x) - (edit_type(x) == BT2GPU_EDIT_READ_GAP ? 0u : 1u);
		return edit_with_pos(x, pos);
	}
};

// Reference-column interval [left,right) covered by read row i, advancing a cursor
// (RedundantAlns::overlap/add walk, aligner_result.cpp:821-925)
struct CellWalk {
	EditView ev; int64_t left; uint32_t nedidx; uint32_t len;
	BT2_HD void init(const EditView& v, int64_t refoff, uint32_t rdlen) { ev = v; left = refoff; nedidx = 0; len = rdlen; }
	BT2_HD void row(uint32_t i, int64_t& l, int64_t& r) {
		int diff = 1;
		int64_t right = left + 1;
		while(nedidx < ev.n && edit_pos(ev.get(nedidx)) == i) { if(edit_type(ev.get(nedidx)) == BT2GPU_EDIT_REF_GAP) diff = 0; nedidx++; }
		if(i < len - 1) {
			uint32_t nx = nedidx;
			while(nx < ev.n && edit_pos(ev.get(nx)) == i + 1) { if(edit_type(ev.get(nx)) == BT2GPU_EDIT_READ_GAP) right++; nx++; }
		}
		l = left; r = right;
		left = right + diff - 1;
	}
};

BT2_HD bool alns_overlap(const EditView& a, int64_t arefoff, const EditView& b, int64_t brefoff, uint32_t rdlen) {
	CellWalk wa, wb; wa.init(a, arefoff, rdlen); wb.init(b, brefoff, rdlen);
	for(uint32_t i = 0; i < rdlen; i++) {
		int64_t al, ar, bl, br;
		wa.row(i, al, ar); wb.row(i, bl, br);
		if(al < br && bl < ar) return true;
	}
	return false;
}

// launch-wide context of the backtrace stage (device: one kernel parameter, i.e. constant bank)
struct BtCtx {
	DevParams P;
	Pools pools;
	ReadState* states; const Problem* probs;
	uint8_t* cells;          // [nslots * cell_stride] CF_* flag bytes (cell_index); bit 7 = visited, set here
	size_t cell_stride;      // bytes of cell flags per problem slot (fill_cell_bytes of the batch's longest read and its K)
	const uint32_t* pcache;  // [nslots * BT2_PC_WORDS]
	uint32_t* scratch;       // [nreads * BT2_SCRATCH_EDITS] per-read edit scratch
	const uint8_t* seq; const uint8_t* qual; // batch arrays
	int32_t pen[8];          // penalty of a move by its option bit: read-gap extend, open, ref-gap extend, open, diagonal (0)
	void set_pens() { pen[0] = P.rdGapExt; pen[1] = P.rdGapOpen; pen[2] = P.rfGapExt; pen[3] = P.rfGapOpen; pen[4] = pen[5] = pen[6] = pen[7] = 0; }
};

struct BtMachine {
	// ---- the current problem
	uint32_t slot, read;
	const uint32_t* pc; uint32_t pcs; // character cache of the problem: word j at pc[j * pcs]
	// two 16-byte chunks of the cell array: the one the walk is in and the one it is predicted to enter next
	// (asynchronous copy issued when a chunk is entered, so that its latency overlaps the steps in between)
	// (buffers 2 and 3: the group of 8 candidates the loop is in and the next one)
	uint8_t* cbuf; uint32_t cstride;  // buffer b of this lane at cbuf + b * cstride
	uint32_t cbuf_sa;                 // cbuf as a shared-window address (device)
	uint32_t ctag0, ctag1, ccur;
	// ---- cached per problem
	uint8_t* cellp; uint32_t* scr;    // the problem's cells, the read's edit scratch
	int rows, dlo;                    // core diagonals: dlo <= col-row < dlo + dspan
	uint32_t dspan;
	int d0; uint32_t lgK, yy;         // fill layout (dev_common.cuh)
	bool pi16;                        // the problem runs on the reference's 16-bit path (prob_i16)
	uint32_t ncand;
	Rng rnd; uint32_t hw, nsteps;     // nsteps: backtrace steps in the low 22 bits, attempts in the high 10 (one per candidate at most)
	// ---- nextAlignment state
	int phase, result;
	uint32_t cural, reseed;
	// ---- backtrace state (one attempt)
	int row, col, ct, score;
	uint32_t cidx;                    // cell_index(row, col)
	uint32_t ned, nbranch;
	bool coreHit;

	BT2_HD const Problem& pb(const BtCtx& cx) const { return cx.probs[slot]; }
	BT2_HD ReadState& rs(const BtCtx& cx) const { return cx.states[read]; }
	BT2_HD uint32_t* scratch(const BtCtx&) const { return scr; }
	BT2_HD ReadView view(const BtCtx& cx) const {
		const ReadState& r = rs(cx);
		ReadView v; v.seq = cx.seq + r.seq_off; v.qual = cx.qual + r.seq_off; v.rdlen = r.rdlen; v.fw = pb(cx).fw != 0; return v;
	}
	// cell_index() with the problem's layout constants in registers
	BT2_HD uint32_t cell_at(int r, int c) const {
		const uint32_t dd = (uint32_t)(c - r - d0), v = dd >> lgK, j = dd & ((1u << lgK) - 1u);
		const uint32_t i = (uint32_t)r + ((v + yy) >> 1);
		return ((((i >> (4u - lgK)) << 5) + v) << 4) + ((i & ((16u >> lgK) - 1u)) << lgK) + j;
	}
	// the next candidate's start cell (last row), or the end of the candidate loop
	BT2_HD void seek_cand() {
		if(cural >= ncand) { phase = BP_DONE; return; }
		row = rows - 1; col = (int)(cand_at(cural) & ((1u << BT2_CAND_SHIFT) - 1u));
		cidx = cell_at(row, col);
		ct = CT_H;   // (the chunk prediction of the start cell reads it)
		phase = BP_CAND;
	}
	BT2_HD void relocate() { cidx = cell_at(row, col); }
	BT2_HD void push_edit(uint32_t x) { if(ned < BT2_SCRATCH_EDITS) scr[ned] = x; ned++; }

	// candidate i of the problem: group i / 4 sits in buffer 2 + (i / 4) % 2
	BT2_HD uint32_t cand_at(uint32_t i) const {
		return *(const volatile uint32_t*)(cbuf + (2u + ((i >> 2) & 1u)) * cstride + (i & 3u) * 4u);
	}
	BT2_HD void cand_fetch(const BtCtx& cx, uint32_t g) {
		const uint32_t* src = cx.probs[slot].cand + g * 4u;
		const uint32_t b = 2u + (g & 1u);
#if defined(__CUDA_ARCH__)
		const uint32_t sa = cbuf_sa + b * cstride;
		asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(sa), "l"(src) : "memory");
#else
		memcpy(cbuf + b * cstride, src, 16);
#endif
	}
	BT2_HD void chunk_fetch(uint32_t b, uint32_t ch) {
		const uint8_t* src = cellp + ch * 16u;
#if defined(__CUDA_ARCH__)
		const uint32_t sa = cbuf_sa + b * cstride;
		asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" :: "r"(sa), "l"(src) : "memory");   // .cg: marks of earlier walks are read from L2
#else
		for(int i = 0; i < 16; i++) cbuf[b * cstride + i] = src[i];
#endif
	}
	BT2_HD static void chunk_wait() {
#if defined(__CUDA_ARCH__)
		asm volatile("cp.async.wait_all;" ::: "memory");
#endif
	}
	// chunk the walk enters next if it keeps its direction: up the column in a reference-gap run (F), along the row in a
	// read-gap run (E), else up the diagonal.  Inside a chunk (16 / K iterations x K diagonals of one lane) a diagonal move
	// keeps j and goes one iteration back, a column move goes to j + 1 one iteration back, a row move to j - 1.
	BT2_HD uint32_t predict_chunk() const {
		const uint32_t K = 1u << lgK, j = cidx & (K - 1u), iw = (cidx & 15u) >> lgK;   // position inside the chunk
		int r2 = row, c2 = col;
		if(ct == CT_E) c2 -= (int)j + 1;                                             // leaves through the chunk's left edge
		else if(ct == CT_F) { const uint32_t n = iw + 1u < K - j ? iw + 1u : K - j; r2 -= (int)n; }   // through its top or right edge
		else { r2 -= (int)iw + 1; c2 -= (int)iw + 1; }
		if(r2 < 0 || c2 < 0 || (uint32_t)(c2 - r2 - d0) >= (K << 5)) return BT2_NIL;
		return cell_at(r2, c2) >> 4;
	}

	// SwDriver::extendSeeds right after swa.align(): counters, then either the candidate loop or done
	BT2_HD void begin(const BtCtx& cx) {
		const DevParams& Pr = cx.P; const Problem& p = pb(cx); ReadState& r = rs(cx);
		rows = p.rows; ncand = p.ncand;
		d0 = p.d0; lgK = fill_lgK(prob_K(p)); yy = prob_y(p); pi16 = prob_i16(p);
		dlo = (p.corel > 0 ? p.corel : 0) - p.triml; { const int dhi = p.corer - p.triml; dspan = dhi >= dlo ? (uint32_t)(dhi - dlo) + 1u : 0u; }
		cellp = cx.cells + (size_t)slot * cx.cell_stride; scr = cx.scratch + (size_t)read * BT2_SCRATCH_EDITS;
		chunk_wait();   // a chunk copy of the previous problem may still be on its way into this lane's buffers
		nsteps = 0; ctag0 = BT2_NIL; ctag1 = BT2_NIL; ccur = 0;
		rnd = r.rnd; hw = r.scr_hw;
		BT2_PROF(16, 1); BT2_PROF(17, ncand);
		r.dpCells += (uint64_t)p.rows * p.cols;
		// 8-bit: best = lrmax - 0xff, found iff best >= minsc (then the last row holds a candidate).  16-bit: the reference subtracts 0xff from
		// the RAW word (bias 0x7fff) as well (aligner_swsse_ee_i16.cpp:640), so its test never fails and found = some candidate exists
		const int32_t bestCell = pi16 ? (0x7fff + (p.lrmax - BT2_I16_START)) - 255 : p.lrmax - 255;
		const bool found = bestCell >= p.minsc && ncand > 0;
		r.nExDps++;
		result = BT_CONTINUE;
		if(!found) {
			r.nExDpFails++; r.nDpFail++;
			if(r.nDpFail >= (uint32_t)Pr.streak) { r.status = BT2GPU_ST_SOFT_LIMIT; result = BT_ROUND_END; }
			else if(bestCell > r.bestLtMinsc) r.bestLtMinsc = bestCell;
			phase = BP_DONE;
			return;
		}
		r.nExDpSuccs++; r.nDpLastSucc = r.nExDps - 1;
		if(r.nDpFail > r.nDpFailStreak) r.nDpFailStreak = r.nDpFail;
		r.nDpFail = 0;
		cural = 0;
		cand_fetch(cx, 0); if(ncand > 4u) cand_fetch(cx, 1);
		chunk_wait();
		seek_cand();
	}
	BT2_HD void finish(const BtCtx& cx) {
		ReadState& r = rs(cx);
		r.btSteps += nsteps & 0x3fffffu; r.btAttempts += nsteps >> 22;
		r.rnd = rnd; r.scr_hw = (uint16_t)hw;
	}
	// entering group g = cural / 4: group g + 1 is copied into the buffer group g - 1 leaves (waited for by the next cell's chunk_wait)
	BT2_HD void next_cand(const BtCtx& cx) { cural++; if((cural & 3u) == 0u && cural + 4u < ncand) cand_fetch(cx, (cural >> 2) + 1u); }

	BT2_HD void attempt_failed(const BtCtx& cx) { // backtrace returned false: scratch keeps what was written (res.reset() only drops the count)
		if(ned > hw) hw = ned;
		rnd.init(pi16 ? reseed : reseed + 1);
		next_cand(cx);
		seek_cand();
	}

	// One unit of the frequent work: SwAligner::nextAlignment picking the next candidate whose start
	// cell has not been reported through, and one cell of backtraceNucleotidesEnd2EndSseU8.
	// (phase BP_CAND: positioned on a candidate's start cell, attempt not begun; BP_STEP: inside a walk.)
	BT2_HD void hot(const BtCtx& cx) {
		BT2_PROF(0, 1);
		const bool starting = phase == BP_CAND;
		// ---- memory stage: the chunk holding the cell
		const uint32_t ch = cidx >> 4;
		const uint32_t curtag = ccur ? ctag1 : ctag0, othtag = ccur ? ctag0 : ctag1;
		const bool newC = ch != curtag, needC = newC && ch != othtag;
		if(needC) chunk_fetch(ccur, ch);
		chunk_wait();
		if(newC) {
			if(needC) { if(ccur) ctag1 = ch; else ctag0 = ch; } else ccur ^= 1u;
			const uint32_t nx = predict_chunk();
			if(nx != BT2_NIL) chunk_fetch(ccur ^ 1u, nx);
			if(ccur) ctag0 = nx; else ctag1 = nx;
		}
		volatile uint8_t* const cp = cbuf + ccur * cstride + (cidx & 15u);
		const uint32_t raw = *cp;
		const bool was = (raw & CF_VIS) != 0;
		if(!was) { *cp = (uint8_t)(raw | CF_VIS); cellp[cidx] = (uint8_t)(raw | CF_VIS); }
		const uint32_t fl = ~raw;   // bit set = the comparison holds
		// ---- candidate start
		if(starting) {
			if(was) { BT2_PROF(2, 1); next_cand(cx); seek_cand(); return; }
			BT2_PROF(3, 1);
			reseed = rnd.nextU32() + 1;
			if(!pi16) rnd.init(reseed);   // u8: reseed before and reseed+1 after the backtrace; i16: reseed after only (aligner_sw.cpp:324-348)
			ct = CT_H;
			score = 0; nbranch = 0; ned = 0; coreHit = false;
			nsteps += 1u << 22;
			phase = BP_STEP;
		}
		nsteps++;
		if(was) { nsteps += nbranch; BT2_PROF(4, 1); attempt_failed(cx); return; } // every branch frame pops onto a marked cell
		// options of the cell, highest priority first: bit4 diagonal, bit3 ref-gap open, bit2 ref-gap extend,
		// bit1 read-gap open, bit0 read-gap extend (H); E and F cells have only their own open/extend pair
		const bool hasLeft = col > 0;
		const uint32_t eopts = hasLeft ? (((fl >> 2) & 2u) | ((fl >> 4) & 1u)) : 0u;   // EO -> bit1, EX -> bit0
		const uint32_t fopts = ((fl >> 2) & 8u) | ((fl >> 4) & 4u);                      // FO -> bit3, FX -> bit2
		uint32_t hopts = (hasLeft && (fl & CF_D)) ? 16u : 0u;
		if(row >= cx.P.gapbar && row < rows - cx.P.gapbar) {   // gaps allowed in this row
			if(fl & CF_HF) hopts |= fopts;
			if(fl & CF_HE) hopts |= eopts;
		}
		uint32_t opts = ct == CT_E ? eopts : (ct == CT_F ? fopts : hopts);
		if(row == 0) opts = 0;
		coreHit = coreHit || (uint32_t)(col - row - dlo) < dspan; // core-diagonal test (aligner_swsse_ee_u8.cpp:1229-1262) folded in
		if(opts == 0) { phase = BP_POST; BT2_PROF(6, 1); return; } // row 0, or a cell nothing leads into: the walk ends here
		if(opts & (opts - 1u)) nbranch++;
		const uint32_t sel = 31u - (uint32_t)BT2_CLZ(opts);
		BT2_PROF(8 + sel, 1);
		// ---- the move and its edit: diagonal = mismatch edit unless the characters match (its penalty and N count are taken
		// in post()), ref-gap (sel 2,3) and read-gap (sel 0,1) edits always
		const bool isDiag = sel == 4u, isRead = sel < 2u, isRef = !isDiag && !isRead;
		const uint32_t rc = pc_read_code(pc, pcs, (uint32_t)row), fc = pc_ref_code(pc, pcs, (uint32_t)col);
		if(!(isDiag && rc == fc && rc < 4u)) {
			const uint32_t chr = isRef ? (uint32_t)'-' : code_char(fc), qchr = isRead ? (uint32_t)'-' : code_char(rc);
			const uint32_t type = isDiag ? (uint32_t)BT2GPU_EDIT_MM : (isRef ? (uint32_t)BT2GPU_EDIT_REF_GAP : (uint32_t)BT2GPU_EDIT_READ_GAP);
			push_edit(pack_edit((uint32_t)row + (isRead ? 1u : 0u), chr, qchr, type));
		}
		score -= cx.pen[sel];
		ct = sel == 2u ? CT_F : (sel == 0u ? CT_E : CT_H);
		if(sel >= 2u) row--;
		if(!isRef) col--;
		relocate();
	}

	// end of a backtrace that reached row 0 / an empty cell, then the rest of nextAlignment and of
	// the extendSeeds body for one alignment: redundancy, report, policy, tighten
	BT2_HD void post(const BtCtx& cx) {
		const DevParams& Pr = cx.P; const Problem& p = pb(cx); const Pools* pools = &cx.pools;
		uint32_t* const scratch = this->scratch(cx);
		if(!coreHit) { attempt_failed(cx); return; }
		{ // the last cell of the path: mismatch edit unless the characters match
			const uint32_t rc = pc_read_code(pc, pcs, (uint32_t)row), fc = pc_ref_code(pc, pcs, (uint32_t)col);
			if(!(rc == fc && rc < 4u)) push_edit(pack_edit((uint32_t)row, code_char(fc), code_char(rc), BT2GPU_EDIT_MM));
		}
		ReadState& rs = this->rs(cx);
		// mismatch penalties, N count and gap counts of the path, from its edits
		int ns = 0, gaps = 0;
		if(ned <= BT2_SCRATCH_EDITS) {
			const ReadView rv = view(cx);
			for(uint32_t i = 0; i < ned; i++) {
				const uint32_t x = scratch[i];
				if(edit_type(x) != BT2GPU_EDIT_MM) { gaps++; continue; }
				const bool isn = edit_chr(x) == 'N' || edit_qchr(x) == 'N';
				int q = rv.q(edit_pos(x)) - 33; if(q < 0) q = 0; if(q > 255) q = 255;
				score -= isn ? (int)Pr.npens[q] : (int)Pr.mmpens[q];
				if(isn) ns++;
			}
		}
		if(ns > rs.nceil || ned > BT2_SCRATCH_EDITS) { attempt_failed(cx); return; } // (> capacity: beyond the reference's own DList<Edit,160>)
		const uint32_t ne = ned;
		for(uint32_t i = 0; i < ne / 2; i++) { const uint32_t t = scratch[i]; scratch[i] = scratch[ne - 1 - i]; scratch[ne - 1 - i] = t; }
		if(ned > hw) hw = ned;
		rnd.init(pi16 ? reseed : reseed + 1);
		AlnRec a;
		a.fw = p.fw; a.tidx = p.tidx; a.refoff = (int64_t)col + p.refl;
		a.score = score; a.ns = ns; a.gaps = gaps; a.edits = (int32_t)ne;
		a.basesAligned = (int32_t)p.rows - (int32_t)ne;   // trimEnd is 0: candidates come from the last row only
		int refns = 0;
		const int origCol = (int)(cand_at(cural) & ((1u << BT2_CAND_SHIFT) - 1u));   // the walk's start column
		for(int i = col; i <= origCol; i++) if(pc_ref_code(pc, pcs, (uint32_t)i) > 3u) refns++;
		a.refns = refns; a.n_edits = (uint16_t)ne;
		if(!p.fw) { // AlnRes::invertEdits: scratch becomes 5'-based
			for(uint32_t i = 0; i < ne / 2; i++) { const uint32_t t = scratch[i]; scratch[i] = scratch[ne - 1 - i]; scratch[ne - 1 - i] = t; }
			for(uint32_t i = 0; i < ne; i++) {
				const uint32_t x = scratch[i];
				scratch[i] = edit_with_pos(x, (uint32_t)rs.rdlen - edit_pos(x) - (edit_type(x) == BT2GPU_EDIT_READ_GAP ? 0u : 1u));
			}
		}
		// ---- RedundantAlns::overlap against every alignment already reported for this read
		EditView nv; nv.e = scratch; nv.n = ne; nv.inv = !p.fw; nv.rdlen = rs.rdlen;
		bool olap = false;
		for(uint32_t q = rs.aln_tail; q != BT2_NIL && !olap; q = pools->alns[q].prev) {
			const AlnRec& o = pools->alns[q];
			if(o.tidx != a.tidx || o.fw != a.fw) continue;
			EditView ov; ov.e = pools->edits + o.edit_off; ov.n = o.n_edits; ov.inv = !o.fw; ov.rdlen = rs.rdlen;
			olap = alns_overlap(nv, a.refoff, ov, o.refoff, rs.rdlen);
		}
		if(olap) { next_cand(cx); seek_cand(); return; }
		// ---- report
		const uint32_t slot = atomic_bump(pools->aln_top, 1);
		const uint32_t eoff = ne ? atomic_bump(pools->edit_top, ne) : 0;
		if(slot >= pools->aln_cap) { set_error(*pools, 4); rs.status = BT2GPU_ST_HARD_LIMIT; result = BT_ROUND_END; phase = BP_DONE; return; }
		if(ne && eoff + ne > pools->edit_cap) { set_error(*pools, 8); rs.status = BT2GPU_ST_HARD_LIMIT; result = BT_ROUND_END; phase = BP_DONE; return; }
		for(uint32_t i = 0; i < ne; i++) pools->edits[eoff + i] = scratch[i];
		a.edit_off = eoff; a.minsc = rs.minsc; a.prev = rs.aln_tail; a.pad = 0; a.pad2 = 0;
		pools->alns[slot] = a;
		rs.aln_tail = slot; rs.n_alns++;
		// ReportingState::foundUnpaired + best / second best
		rs.nunpair++;
		const bool mhitsSet = Pr.mhits != 0;
		if(!rs.doneUnpair) {
			if(rs.nunpair >= (uint32_t)Pr.khits && !mhitsSet) rs.doneUnpair = 1;
			else if(mhitsSet && rs.nunpair > (uint32_t)Pr.mhits) rs.doneUnpair = 1;
			if(rs.doneUnpair) rs.donePolicy = 1;
		}
		if(a.score > rs.best) { rs.best2 = rs.best; rs.best = a.score; } else if(a.score > rs.best2) rs.best2 = a.score;
		if(rs.donePolicy) { rs.status = BT2GPU_ST_POLICY_FULFILLED; result = BT_ROUND_END; phase = BP_DONE; return; }
		if(Pr.tighten > 0 && mhitsSet && rs.best2 != BT2_I32_MIN) {
			const int32_t perfect = (int32_t)rs.rdlen * Pr.matchBonus;
			const int32_t best = rs.best, sec = rs.best2;
			if(Pr.tighten == 1) {
				if(best >= rs.minsc) { rs.minsc = best; if(rs.minsc < perfect && best == sec) rs.minsc++; }
			} else if(Pr.tighten == 2) {
				if(sec >= rs.minsc) { rs.minsc = sec; if(rs.minsc < perfect) rs.minsc++; }
			} else {
				const int32_t diff = best - sec;
				const int32_t bot = sec + ((diff * 3) / 4);
				if(bot >= rs.minsc) { rs.minsc = bot; if(rs.minsc < perfect) rs.minsc++; }
			}
		}
		next_cand(cx); seek_cand();
	}

	// one unit of work; returns false once the problem is finished (result / the read's state are final)
	BT2_HD bool step(const BtCtx& cx) {
		if(phase == BP_STEP || phase == BP_CAND) hot(cx);
		else if(phase == BP_POST) post(cx);
		return phase != BP_DONE;
	}
};

} // namespace bt2
