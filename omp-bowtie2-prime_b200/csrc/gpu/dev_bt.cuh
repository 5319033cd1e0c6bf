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
BT2_HD uint32_t code_char(uint32_t c) { return (0x4e54474341ull >> (8u * (c > 4u ? 4u : c))) & 255u; }

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

struct BtMachine {
	// ---- context of the current problem
	const DevParams* P; const Pools* pools;
	const Problem* pb;
	const uint8_t* cells;  // CF_* flag bytes of this problem (cell_index)
	uint32_t* vis;         // visited map of this problem slot (vis_word)
	uint32_t* scratch;     // per-read edit scratch
	const uint8_t* seqbase; const uint8_t* qualbase; // batch arrays
	ReadState* rs;         // the read's state; rnd / scratch marks are held here while the problem runs
	const uint32_t* pc; uint32_t pcs; // character cache of the problem: word j at pc[j * pcs]
	uint32_t gtag;         // generation << 16
	int rdGapOpen, rdGapExt, rfGapOpen, rfGapExt;
	// ---- cached per problem
	int rows, R, glo, ghi, dlo, dhi; // gaps allowed in rows [glo,ghi); core diagonals: dlo <= col-row <= dhi
	uint32_t ncand;
	Rng rnd; uint32_t hw, attempts, nsteps;
	// ---- nextAlignment state
	int phase, result;
	uint32_t cural, reseed, candNext;
	// ---- backtrace state (one attempt)
	int row, col, origCol, ct, score, readGaps, refGaps;
	uint32_t cidx;         // cell_index(row,col)
	uint32_t ned, nbranch;
	bool coreHit;
	// ---- the word of the visited map the walk is in (written back when the walk leaves it)
	uint32_t vwi, vwv;

	BT2_HD ReadView view() const { ReadView v; v.seq = seqbase + rs->seq_off; v.qual = qualbase + rs->seq_off; v.rdlen = rs->rdlen; v.fw = pb->fw != 0; return v; }
	BT2_HD void locate() { cidx = cell_index((uint32_t)row, (uint32_t)col, (uint32_t)R); }
	BT2_HD void push_edit(uint32_t x) { if(ned < BT2_SCRATCH_EDITS) scratch[ned] = x; ned++; if(ned > hw) hw = ned; }
	BT2_HD uint32_t read_char(int r) const { return code_char(pc_read_code(pc, pcs, (uint32_t)r)); }
	BT2_HD uint32_t ref_char(int c) const { return code_char(pc_ref_code(pc, pcs, (uint32_t)c)); }

	// visited(row,col); marks the cell
	BT2_HD bool visit() {
		const uint32_t wi = vis_word((uint32_t)row, (uint32_t)col);
		if(wi != vwi) {
			if(vwi != BT2_NIL) vis[vwi] = vwv;
			const uint32_t m = vis[wi];
			vwv = ((m & 0xffff0000u) == gtag) ? m : gtag;
			vwi = wi;
		}
		const uint32_t bit = 1u << ((uint32_t)row & 15u);
		const bool was = (vwv & bit) != 0;
		vwv |= bit;
		return was;
	}

	// SwDriver::extendSeeds right after swa.align(): counters, then either the candidate loop or done
	BT2_HD void begin() {
		const DevParams& Pr = *P;
		rows = pb->rows; R = pb->R; ncand = pb->ncand;
		glo = Pr.gapbar; ghi = rows - Pr.gapbar;
		dlo = (pb->corel > 0 ? pb->corel : 0) - pb->triml; dhi = pb->corer - pb->triml;
		rdGapOpen = Pr.rdGapOpen; rdGapExt = Pr.rdGapExt; rfGapOpen = Pr.rfGapOpen; rfGapExt = Pr.rfGapExt;
		nsteps = 0; attempts = 0; vwi = BT2_NIL; vwv = 0;
		rnd = rs->rnd; hw = rs->scr_hw;
		rs->dpCells += (uint64_t)pb->rows * pb->cols;
		const int32_t bestCell = pb->lrmax - 255;
		const bool found = bestCell >= pb->minsc && ncand > 0;
		rs->nExDps++;
		result = BT_CONTINUE;
		if(!found) {
			rs->nExDpFails++; rs->nDpFail++;
			if(rs->nDpFail >= (uint32_t)Pr.streak) { rs->status = BT2GPU_ST_SOFT_LIMIT; result = BT_ROUND_END; }
			else if(bestCell > rs->bestLtMinsc) rs->bestLtMinsc = bestCell;
			phase = BP_DONE;
			return;
		}
		rs->nExDpSuccs++; rs->nDpLastSucc = rs->nExDps - 1;
		if(rs->nDpFail > rs->nDpFailStreak) rs->nDpFailStreak = rs->nDpFail;
		rs->nDpFail = 0;
		cural = 0; candNext = pb->cand[0];
		phase = BP_CAND;
	}
	BT2_HD void finish() {
		rs->btSteps += nsteps; rs->btAttempts += attempts;
		rs->rnd = rnd; rs->scr_hw = (uint16_t)hw;
	}
	BT2_HD void next_cand() { cural++; candNext = cural < ncand ? pb->cand[cural] : 0u; }

	BT2_HD void attempt_failed() { // backtrace returned false: scratch keeps what was written (res.reset() only drops the count)
		rnd.init(reseed + 1);
		next_cand();
		phase = BP_CAND;
	}

	// the diagonal move / the last cell of a path: mismatch edit (its penalty and N count are taken in post())
	BT2_HD void take_cell(uint32_t fl) {
		if(fl & CF_MM) return;    // (fl holds the equalities that are TRUE: bit set = match; matchBonus is 0 end-to-end)
		push_edit(pack_edit((uint32_t)row, ref_char(col), read_char(row), BT2GPU_EDIT_MM));
	}

	// One unit of the frequent work: SwAligner::nextAlignment picking the next candidate whose start
	// cell has not been reported through, and one cell of backtraceNucleotidesEnd2EndSseU8.
	BT2_HD void hot() {
		bool starting = false;
		if(phase == BP_CAND) {
			if(cural >= ncand) { phase = BP_DONE; return; }
			row = rows - 1; col = (int)(candNext & 255u);
			locate();
			starting = true;
		}
		const uint32_t fl = ~(uint32_t)cells[cidx];
		const bool was = visit();
		if(starting) {
			if(was) { next_cand(); return; }
			reseed = rnd.nextU32() + 1;
			rnd.init(reseed);
			origCol = col; ct = CT_H;
			score = 0; readGaps = 0; refGaps = 0; nbranch = 0; ned = 0; coreHit = false;
			attempts++;
			phase = BP_STEP;
		}
		nsteps++;
		if(was) { nsteps += nbranch; attempt_failed(); return; } // every branch frame pops onto a marked cell
		// options of the cell, highest priority first: bit4 diagonal, bit3 ref-gap open, bit2 ref-gap extend,
		// bit1 read-gap open, bit0 read-gap extend (H); E and F cells have only their own open/extend pair
		uint32_t opts = 0;
		if(row > 0) {
			const bool hasLeft = col > 0;
			if(ct == CT_E) {
				if(hasLeft) opts = ((fl >> 2) & 2u) | ((fl >> 4) & 1u);            // EO -> bit1, EX -> bit0
			} else if(ct == CT_F) {
				opts = ((fl >> 2) & 8u) | ((fl >> 4) & 4u);                      // FO -> bit3, FX -> bit2
			} else {
				if(row >= glo && row < ghi) {
					if(fl & CF_HF) opts |= ((fl >> 2) & 8u) | ((fl >> 4) & 4u);
					if(hasLeft && (fl & CF_HE)) opts |= ((fl >> 2) & 2u) | ((fl >> 4) & 1u);
				}
				if(hasLeft && (fl & CF_D)) opts |= 16u;
			}
		}
		const int dg = col - row;
		coreHit = coreHit || (dg >= dlo && dg <= dhi); // core-diagonal test (aligner_swsse_ee_u8.cpp:1229-1262) folded in
		if(opts == 0) { phase = BP_POST; cidx = fl; return; } // row 0, or a cell nothing leads into: the walk ends here
		if(opts & (opts - 1u)) nbranch++;
		const uint32_t sel = 31u - (uint32_t)BT2_CLZ(opts);
		if(sel == 4u) {
			ct = CT_H;
			take_cell(fl);
			row--; col--;
		} else if(sel >= 2u) {
			push_edit(pack_edit((uint32_t)row, '-', read_char(row), BT2GPU_EDIT_REF_GAP));
			row--;
			ct = (sel == 3u) ? CT_H : CT_F;
			score -= (sel == 3u) ? rfGapOpen : rfGapExt;
			refGaps++;
		} else {
			push_edit(pack_edit((uint32_t)row + 1, ref_char(col), '-', BT2GPU_EDIT_READ_GAP));
			col--;
			ct = (sel == 1u) ? CT_H : CT_E;
			score -= (sel == 1u) ? rdGapOpen : rdGapExt;
			readGaps++;
		}
		locate();
	}

	// end of a backtrace that reached row 0 / an empty cell, then the rest of nextAlignment and of
	// the extendSeeds body for one alignment: redundancy, report, policy, tighten
	BT2_HD void post() {
		const DevParams& Pr = *P; const Problem& p = *pb;
		if(!coreHit) { attempt_failed(); return; }
		take_cell(cidx); // (hot() left the last cell's flags in cidx)
		ReadState& rs = *this->rs;
		// mismatch penalties and N count of the path, from its edits
		int ns = 0;
		if(ned <= BT2_SCRATCH_EDITS) {
			const ReadView rv = view();
			for(uint32_t i = 0; i < ned; i++) {
				const uint32_t x = scratch[i];
				if(edit_type(x) != BT2GPU_EDIT_MM) continue;
				const bool isn = edit_chr(x) == 'N' || edit_qchr(x) == 'N';
				int q = rv.q(edit_pos(x)) - 33; if(q < 0) q = 0; if(q > 255) q = 255;
				score -= isn ? (int)Pr.npens[q] : (int)Pr.mmpens[q];
				if(isn) ns++;
			}
		}
		if(ns > rs.nceil || ned > BT2_SCRATCH_EDITS) { attempt_failed(); return; } // (> capacity: beyond the reference's own DList<Edit,160>)
		const uint32_t ne = ned;
		for(uint32_t i = 0; i < ne / 2; i++) { const uint32_t t = scratch[i]; scratch[i] = scratch[ne - 1 - i]; scratch[ne - 1 - i] = t; }
		rnd.init(reseed + 1);
		next_cand();
		phase = BP_CAND;
		AlnRec a;
		a.fw = p.fw; a.tidx = p.tidx; a.refoff = (int64_t)col + p.refl;
		a.score = score; a.ns = ns; a.gaps = readGaps + refGaps; a.edits = (int32_t)ne;
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

	// one unit of work; returns false once the problem is finished (result / rs are final)
	BT2_HD bool step() {
		if(phase == BP_STEP || phase == BP_CAND) hot();
		else if(phase == BP_POST) post();
		return phase != BP_DONE;
	}
};

} // namespace bt2
