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
// unit of the frequent work (start a candidate / one backtrace cell) with exactly two
// independent memory reads (the cell's flag byte and, when the walk leaves it, the next
// word of the visited map); everything a cell needs besides that (read and reference
// characters for the edits) comes from a per-lane copy of the problem in shared memory.
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
#pragma once
#include "dev_common.cuh"

#if defined(__CUDA_ARCH__)
#define BT2_POPC(x) __popc(x)
#define BT2_CLZ(x) __clz(x)
#else
#define BT2_POPC(x) __builtin_popcount(x)
#define BT2_CLZ(x) ((x) ? __builtin_clz(x) : 32)
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
	const uint8_t* cells;    // [nslots * BT2_CELL_BYTES] CF_* flag bytes (cell_index)
	uint32_t* vis;           // [nslots * BT2_VIS_WORDS] visited maps (vis_index)
	const uint32_t* pcache;  // [nslots * BT2_PC_WORDS]
	uint32_t* scratch;       // [nreads * BT2_SCRATCH_EDITS] per-read edit scratch
	const uint8_t* seq; const uint8_t* qual; // batch arrays
	uint32_t gtag;           // generation << 16
};

struct BtMachine {
	// ---- the current problem
	uint32_t slot, read;
	const uint32_t* pc; uint32_t pcs; // character cache of the problem: word j at pc[j * pcs]
	// two 16-byte chunks of the cell array: the one the walk is in and the one it is predicted to enter next
	// (asynchronous copy issued when a chunk is entered, so that its latency overlaps the steps in between)
	uint8_t* cbuf; uint32_t cstride;  // chunk b of this lane at cbuf + b * cstride
	uint32_t cbuf_sa;                 // cbuf as a shared-window address (device)
	uint32_t ctag0, ctag1, ccur;
	// the word of the visited map the walk is in (written back when the walk leaves it)
	uint32_t vwi, vwv;
	// ---- cached per problem
	int rows, R, shift, dlo, dhi;     // fill layout (dev_common.cuh); core diagonals: dlo <= col-row <= dhi
	uint32_t ncand;
	Rng rnd; uint32_t hw, attempts, nsteps;
	// ---- nextAlignment state
	int phase, result;
	uint32_t cural, reseed, candNext, candAhead;
	// ---- backtrace state (one attempt)
	int row, col, origCol, ct, score;
	uint32_t cidx, sv, sk;            // cell_index(row + shift, col) and the strip / row within the strip it lies in
	uint32_t ned, nbranch, lastfl;
	bool coreHit;

	BT2_HD const Problem& pb(const BtCtx& cx) const { return cx.probs[slot]; }
	BT2_HD ReadState& rs(const BtCtx& cx) const { return cx.states[read]; }
	BT2_HD uint32_t* scratch(const BtCtx& cx) const { return cx.scratch + (size_t)read * BT2_SCRATCH_EDITS; }
	BT2_HD ReadView view(const BtCtx& cx) const {
		const ReadState& r = rs(cx);
		ReadView v; v.seq = cx.seq + r.seq_off; v.qual = cx.qual + r.seq_off; v.rdlen = r.rdlen; v.fw = pb(cx).fw != 0; return v;
	}
	BT2_HD void locate() {   // (row, col) set from scratch: a candidate's start cell
		const uint32_t rp = (uint32_t)(row + shift);
		sv = strip_of(rp, (uint32_t)R); sk = rp - sv * (uint32_t)R;
		cidx = cell_index_vk(sv, sk, (uint32_t)col, (uint32_t)R);
	}
	BT2_HD void row_up() {   // row-- : the strip / row-in-strip pair follows without a division
		row--;
		if(sk == 0u) { sk = (uint32_t)R - 1u; sv--; } else sk--;
	}
	BT2_HD void relocate() { cidx = cell_index_vk(sv, sk, (uint32_t)col, (uint32_t)R); }
	BT2_HD void push_edit(const BtCtx& cx, uint32_t x) { if(ned < BT2_SCRATCH_EDITS) scratch(cx)[ned] = x; ned++; }
	BT2_HD uint32_t read_char(int r) const { return code_char(pc_read_code(pc, pcs, (uint32_t)r)); }
	BT2_HD uint32_t ref_char(int c) const { return code_char(pc_ref_code(pc, pcs, (uint32_t)c)); }

	BT2_HD void chunk_fetch(const BtCtx& cx, uint32_t b, uint32_t ch) {
		const uint8_t* src = cx.cells + (size_t)slot * BT2_CELL_BYTES + (size_t)ch * 16u;
#if defined(__CUDA_ARCH__)
		const uint32_t sa = cbuf_sa + b * cstride;
		asm volatile("cp.async.ca.shared.global [%0], [%1], 16;" :: "r"(sa), "l"(src) : "memory");
#else
		for(int i = 0; i < 16; i++) cbuf[b * cstride + i] = src[i];
#endif
	}
	BT2_HD static void chunk_wait() {
#if defined(__CUDA_ARCH__)
		asm volatile("cp.async.wait_all;" ::: "memory");
#endif
	}
	// chunk the walk most likely enters after the one holding (row, col): the next strip up the diagonal
	BT2_HD uint32_t predict_chunk() const {
		const int c2 = col - (int)sk - 1;
		if(sv == 0u || c2 < 0) return BT2_NIL;
		return cell_index_vk(sv - 1u, (uint32_t)R - 1u, (uint32_t)c2, (uint32_t)R) >> 4;   // row R-1 of the strip above, same diagonal
	}

	// SwDriver::extendSeeds right after swa.align(): counters, then either the candidate loop or done
	BT2_HD void begin(const BtCtx& cx) {
		const DevParams& Pr = cx.P; const Problem& p = pb(cx); ReadState& r = rs(cx);
		rows = p.rows; R = p.R; shift = (int)fill_shift((uint32_t)p.rows, (uint32_t)p.R); ncand = p.ncand;
		dlo = (p.corel > 0 ? p.corel : 0) - p.triml; dhi = p.corer - p.triml;
		nsteps = 0; attempts = 0; vwi = BT2_NIL; vwv = 0; ctag0 = BT2_NIL; ctag1 = BT2_NIL; ccur = 0;
		rnd = r.rnd; hw = r.scr_hw;
		r.dpCells += (uint64_t)p.rows * p.cols;
		const int32_t bestCell = p.lrmax - 255;
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
		cural = 0; candNext = p.cand[0]; candAhead = ncand > 1 ? p.cand[1] : 0u;
		phase = BP_CAND;
	}
	BT2_HD void finish(const BtCtx& cx) {
		ReadState& r = rs(cx);
		r.btSteps += nsteps; r.btAttempts += attempts;
		r.rnd = rnd; r.scr_hw = (uint16_t)hw;
	}
	// candNext = cand[cural]; candAhead = cand[cural + 1], loaded one candidate early
	BT2_HD void next_cand(const BtCtx& cx) { cural++; candNext = candAhead; candAhead = cural + 1 < ncand ? pb(cx).cand[cural + 1] : 0u; }

	BT2_HD void attempt_failed(const BtCtx& cx) { // backtrace returned false: scratch keeps what was written (res.reset() only drops the count)
		if(ned > hw) hw = ned;
		rnd.init(cx.P.i16path ? reseed : reseed + 1);
		next_cand(cx);
		phase = BP_CAND;
	}

	// the diagonal move / the last cell of a path: mismatch edit (its penalty and N count are taken in post())
	BT2_HD void take_cell(const BtCtx& cx, uint32_t fl) {
		if(fl & CF_MM) return;    // (fl holds the equalities that are TRUE: bit set = match; matchBonus is 0 end-to-end)
		push_edit(cx, pack_edit((uint32_t)row, ref_char(col), read_char(row), BT2GPU_EDIT_MM));
	}

	// One unit of the frequent work: SwAligner::nextAlignment picking the next candidate whose start
	// cell has not been reported through, and one cell of backtraceNucleotidesEnd2EndSseU8.
	// All memory the cell needs is requested up front and waited for at ONE point, so that the lanes
	// of a warp (each in its own problem) pay one memory latency per call, not one per kind of miss.
	BT2_HD void hot(const BtCtx& cx) {
		bool starting = false;
		if(phase == BP_CAND) {
			if(cural >= ncand) { phase = BP_DONE; return; }
			row = rows - 1; col = (int)(candNext & 255u);
			locate();
			starting = true;
		}
		// ---- memory stage
		uint32_t* const vis = cx.vis + (size_t)slot * BT2_VIS_WORDS;
		const uint32_t wi = vis_index((uint32_t)row, (uint32_t)col);
		const uint32_t ch = cidx >> 4;
		const uint32_t curtag = ccur ? ctag1 : ctag0, othtag = ccur ? ctag0 : ctag1;
		const bool needV = wi != vwi, newC = ch != curtag, needC = newC && ch != othtag;
		uint32_t mv = 0;
		if(needV) { if(vwi != BT2_NIL) vis[vwi] = vwv; mv = vis[wi]; }
		if(needC) chunk_fetch(cx, ccur, ch);
		chunk_wait();
		if(needV) { vwv = ((mv & 0xffff0000u) == cx.gtag) ? mv : cx.gtag; vwi = wi; }
		if(newC) {
			if(needC) { if(ccur) ctag1 = ch; else ctag0 = ch; } else ccur ^= 1u;
			const uint32_t nx = predict_chunk();
			if(nx != BT2_NIL) chunk_fetch(cx, ccur ^ 1u, nx);
			if(ccur) ctag0 = nx; else ctag1 = nx;
		}
		const uint32_t fl = ~(uint32_t)*(volatile const uint8_t*)(cbuf + ccur * cstride + (cidx & 15u));
		const uint32_t bit = vis_bit((uint32_t)row, (uint32_t)col);
		const bool was = (vwv & bit) != 0;
		vwv |= bit;
		// ---- candidate start
		if(starting) {
			if(was) { next_cand(cx); return; }
			reseed = rnd.nextU32() + 1;
			if(!cx.P.i16path) rnd.init(reseed);   // u8: reseed before and reseed+1 after the backtrace; i16: reseed after only (aligner_sw.cpp:324-348)
			origCol = col; ct = CT_H;
			score = 0; nbranch = 0; ned = 0; coreHit = false;
			attempts++;
			phase = BP_STEP;
		}
		nsteps++;
		if(was) { nsteps += nbranch; attempt_failed(cx); return; } // every branch frame pops onto a marked cell
		// options of the cell, highest priority first: bit4 diagonal, bit3 ref-gap open, bit2 ref-gap extend,
		// bit1 read-gap open, bit0 read-gap extend (H); E and F cells have only their own open/extend pair
		uint32_t opts = 0;
		if(row > 0) {
			const bool hasLeft = col > 0;
			const uint32_t eopts = hasLeft ? (((fl >> 2) & 2u) | ((fl >> 4) & 1u)) : 0u;   // EO -> bit1, EX -> bit0
			const uint32_t fopts = ((fl >> 2) & 8u) | ((fl >> 4) & 4u);                      // FO -> bit3, FX -> bit2
			if(ct == CT_E) opts = eopts;
			else if(ct == CT_F) opts = fopts;
			else {
				if(row >= cx.P.gapbar && row < rows - cx.P.gapbar) { // gaps allowed in this row
					if(fl & CF_HF) opts |= fopts;
					if(fl & CF_HE) opts |= eopts;
				}
				if(hasLeft && (fl & CF_D)) opts |= 16u;
			}
		}
		const int dg = col - row;
		coreHit = coreHit || (dg >= dlo && dg <= dhi); // core-diagonal test (aligner_swsse_ee_u8.cpp:1229-1262) folded in
		if(opts == 0) { phase = BP_POST; lastfl = fl; return; } // row 0, or a cell nothing leads into: the walk ends here
		if(opts & (opts - 1u)) nbranch++;
		const uint32_t sel = 31u - (uint32_t)BT2_CLZ(opts);
		if(sel == 4u) {
			ct = CT_H;
			take_cell(cx, fl);
			row_up(); col--;
		} else if(sel >= 2u) {
			push_edit(cx, pack_edit((uint32_t)row, '-', read_char(row), BT2GPU_EDIT_REF_GAP));
			row_up();
			ct = (sel == 3u) ? CT_H : CT_F;
			score -= (sel == 3u) ? cx.P.rfGapOpen : cx.P.rfGapExt;
		} else {
			push_edit(cx, pack_edit((uint32_t)row + 1, ref_char(col), '-', BT2GPU_EDIT_READ_GAP));
			col--;
			ct = (sel == 1u) ? CT_H : CT_E;
			score -= (sel == 1u) ? cx.P.rdGapOpen : cx.P.rdGapExt;
		}
		relocate();
	}

	// end of a backtrace that reached row 0 / an empty cell, then the rest of nextAlignment and of
	// the extendSeeds body for one alignment: redundancy, report, policy, tighten
	BT2_HD void post(const BtCtx& cx) {
		const DevParams& Pr = cx.P; const Problem& p = pb(cx); const Pools* pools = &cx.pools;
		uint32_t* const scratch = this->scratch(cx);
		if(!coreHit) { attempt_failed(cx); return; }
		take_cell(cx, lastfl);
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
		rnd.init(cx.P.i16path ? reseed : reseed + 1);
		next_cand(cx);
		phase = BP_CAND;
		AlnRec a;
		a.fw = p.fw; a.tidx = p.tidx; a.refoff = (int64_t)col + p.refl;
		a.score = score; a.ns = ns; a.gaps = gaps; a.edits = (int32_t)ne;
		a.basesAligned = (int32_t)p.rows - (int32_t)ne;   // trimEnd is 0: candidates come from the last row only
		int refns = 0;
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
		if(olap) return;
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
	}

	// one unit of work; returns false once the problem is finished (result / the read's state are final)
	BT2_HD bool step(const BtCtx& cx) {
		if(phase == BP_STEP || phase == BP_CAND) hot(cx);
		else if(phase == BP_POST) post(cx);
		return phase != BP_DONE;
	}
};

} // namespace bt2
