// What follows a DP fill inside SwDriver::extendSeeds:
//   result bookkeeping                               aligner_sw_driver.cpp:505-531
//   SwAligner::nextAlignment                         aligner_sw.cpp:252-375
//   SwAligner::backtraceNucleotidesEnd2EndSseU8      aligner_swsse_ee_u8.cpp:746-1352
//   RedundantAlns::overlap/add                       aligner_result.cpp:821-925
//   AlnSinkWrap::report + ReportingState::foundUnpaired  aln_sink.cpp:1395-1453, 107-140
//   minsc tightening                                 aligner_sw_driver.cpp:588-618
// The per-read edit scratch reproduces SwDriver::resGap_.alres.ned_.
//
// Written as a flat state machine (BtMachine::step does one unit of work: take a
// candidate, one backtrace cell, or finish an alignment) so that the lanes of a warp,
// each working on its own problem and pulling a new one when done, stay in one loop.
#pragma once
#include "dev_common.cuh"

#if defined(__CUDA_ARCH__)
#define BT2_POPC(x) __popc(x)
#define BT2_CTZ(x) (__ffs(x) - 1)
#else
#define BT2_POPC(x) __builtin_popcount(x)
#define BT2_CTZ(x) __builtin_ctz(x)
#endif

namespace bt2 {

enum { BT_CONTINUE = 0, BT_ROUND_END = 1 };
enum { CT_E = 0, CT_F = 1, CT_H = 2 };
enum { MV_DIAG, MV_REF_OPEN, MV_READ_OPEN, MV_RFGAP_EXT, MV_RDGAP_EXT };
enum { BP_CAND = 0, BP_STEP = 1, BP_POST = 2, BP_DONE = 3 };

struct BtFrame { uint32_t a, b, c, d; }; // packed branch-stack frame
#define BT2_BT_STACK 400
// Backtrace state per cell (what SSEMatrix::masks_ holds, aligner_swsse.h:140-420), one byte,
// stored like the cell flags: one 32-bit word per (strip, step) = the bytes of its R rows plus
// an 8-bit generation tag in byte 3.  A word whose tag differs from the launch's generation
// reads as "all untouched", which replaces the reference's lazy per-row memset
// (SSEMatrix::resetRow).  The byte:
//   bit 0      reportedThru
//   bits 1-3   H: 0 = mask not set, n+1 = set with n options still to try
//   bits 4-5   E: 0 = not set, 1 = set / nothing left, 2 = set / gap extension left
//   bits 6-7   F: likewise
// The reference stores the remaining options as a bit set; since options are always consumed in
// the same priority order (diagonal, ref open, ref extend, read open, read extend) and the set
// of options at a cell is a function of the fill alone, "how many are left" is the same
// information: the remaining ones are the lowest-priority n of the original set.
#define BT2_MASK_GEN_MAX 255u
#define BT2_MASK_WORDS (BT2_CELL_T * 64)

struct ReadView {
	const uint8_t* seq; const uint8_t* qual; uint32_t rdlen; bool fw;
	BT2_HD int rd(uint32_t row) const {
		if(fw) return seq[row];
		const uint8_t c = seq[rdlen - 1 - row];
		return c > 3 ? 4 : 3 - c;
	}
	BT2_HD int q(uint32_t row) const { return fw ? qual[row] : qual[rdlen - 1 - row]; }
};

BT2_HD int mask_to_char(int m) { return m > 15 ? 'N' : "?ACMGRSVTWYHKDBN"[m]; }
BT2_HD int sc_mm(const DevParams& P, int rdc, int refm, int q) {
	if(q < 0) q = 0; if(q > 255) q = 255;
	return (rdc > 3 || refm > 15) ? P.npens[q] : P.mmpens[q];
}
BT2_HD int sc_score(const DevParams& P, int rdc, int refm, int q) {
	if(q < 0) q = 0; if(q > 255) q = 255;
	if(rdc > 3 || refm > 15) return -(int)P.npens[q];
	if((refm & (1 << rdc)) != 0) return P.matchBonus;
	return -(int)P.mmpens[q];
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

// the n lowest-priority options of an H-cell option set (priority high -> low: 16, 1, 4, 2, 8)
BT2_HD int lowest_options(int orig, int n) {
	int out = 0;
	if(n > 0 && (orig & 8))  { out |= 8;  n--; }
	if(n > 0 && (orig & 2))  { out |= 2;  n--; }
	if(n > 0 && (orig & 4))  { out |= 4;  n--; }
	if(n > 0 && (orig & 1))  { out |= 1;  n--; }
	if(n > 0 && (orig & 16)) { out |= 16; n--; }
	return out;
}

struct BtMachine {
	// ---- context of the current problem
	const DevParams* P; const Pools* pools;
	const Problem* pb;
	const uint8_t* cells;  // CF_* flag bytes of this problem (cell_index)
	uint32_t* masks;       // backtrace state words of this problem slot (same indexing / 4)
	BtFrame* stack;        // [BT2_BT_STACK]
	uint32_t* scratch;     // per-read edit scratch
	const uint8_t* seqbase; const uint8_t* qualbase; // batch arrays
	ReadState* rs;         // the read's state, updated in place
	uint32_t gtag;         // generation << 24
	// ---- cached per problem
	int rows, R, glo, ghi, dlo, dhi; // gaps allowed in rows [glo,ghi); core diagonals: dlo <= col-row <= dhi
	// ---- nextAlignment state
	int phase, result;
	uint32_t cural, reseed;
	// ---- backtrace state (one attempt)
	int row, col, origCol, ct, score, ns, readGaps, refGaps;
	uint32_t cidx;         // cell_index(row,col); cidx & 3 = row % R, cidx >> 2 = state word
	uint32_t nstack, ned, hw, fl, nsteps;
	bool coreHit;

	BT2_HD ReadView view() const { ReadView v; v.seq = seqbase + rs->seq_off; v.qual = qualbase + rs->seq_off; v.rdlen = rs->rdlen; v.fw = pb->fw != 0; return v; }
	BT2_HD uint32_t state_word(uint32_t widx) const { const uint32_t m = masks[widx]; return ((m & 0xff000000u) == gtag) ? m : gtag; }
	BT2_HD void locate() { cidx = cell_index((uint32_t)row, (uint32_t)col, (uint32_t)R); } // (row, col) changed arbitrarily
	BT2_HD void row_dec() { row--; if(cidx & 3u) cidx--; else cidx -= (BT2_CELL_T + 1) * 4u - (uint32_t)(R - 1); }
	BT2_HD void col_dec() { col--; cidx -= 4u; }
	BT2_HD void push_edit(uint32_t x) { if(ned < BT2_SCRATCH_EDITS) scratch[ned] = x; ned++; if(ned > hw) hw = ned; }

	// SwDriver::extendSeeds right after swa.align(): counters, then either the candidate loop or done
	BT2_HD void begin() {
		const DevParams& Pr = *P;
		rows = pb->rows; R = pb->R;
		glo = Pr.gapbar; ghi = rows - Pr.gapbar;
		dlo = (pb->corel > 0 ? pb->corel : 0) - pb->triml; dhi = pb->corer - pb->triml;
		nsteps = 0;
		rs->dpCells += (uint64_t)pb->rows * pb->cols;
		const int32_t bestCell = pb->lrmax - 255;
		const bool found = bestCell >= pb->minsc && pb->ncand > 0;
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
		cural = 0;
		phase = BP_CAND;
	}
	BT2_HD void finish() { rs->btSteps += nsteps; }

	// SwAligner::nextAlignment: next candidate whose start cell has not been reported through
	BT2_HD void cand() {
		if(cural >= pb->ncand) { phase = BP_DONE; return; }
		const int lastRow = rows - 1;
		const int c0 = pb->cand[cural] & 255;
		const uint32_t ci = cell_index((uint32_t)lastRow, (uint32_t)c0, (uint32_t)R);
		if((state_word(ci >> 2) >> ((ci & 3u) * 8u)) & 1u) { cural++; return; }
		Rng rnd = rs->rnd;
		reseed = rnd.nextU32() + 1;
		rnd.init(reseed);
		rs->rnd = rnd;
		row = lastRow; col = c0; origCol = c0; ct = CT_H;
		locate();
		score = 0; ns = 0; readGaps = 0; refGaps = 0; nstack = 0; ned = 0; hw = rs->scr_hw; coreHit = false;
		rs->btAttempts++;
		phase = BP_STEP;
	}

	BT2_HD void attempt_failed() { // backtrace returned false: scratch keeps what was written (res.reset() only drops the count)
		rs->scr_cur = (uint16_t)ned; rs->scr_hw = (uint16_t)hw;
		rs->rnd.init(reseed + 1);
		cural++;
		phase = BP_CAND;
	}

	// the diagonal move / the last cell of a path: mismatch edit, N count, score
	BT2_HD void take_cell() {
		if(fl & CF_MM) return;    // (fl holds the equalities that are TRUE: bit set = match; matchBonus is 0 end-to-end)
		const ReadView rv = view();
		const int readc = rv.rd((uint32_t)row), refm = pb->rf[col];
		push_edit(pack_edit((uint32_t)row, (uint32_t)mask_to_char(refm), (uint32_t)"ACGTN"[readc], BT2GPU_EDIT_MM));
		score -= sc_mm(*P, readc, refm, rv.q((uint32_t)row) - 33);
		if(readc > 3 || refm > 15) ns++;
	}

	// one cell of backtraceNucleotidesEnd2EndSseU8
	BT2_HD void bt_step() {
		nsteps++;
		fl = ~(uint32_t)cells[cidx];
		bool empty = false, canMoveThru = true, branch = false;
		int cur = MV_DIAG;
		const uint32_t widx = cidx >> 2, sh8 = (cidx & 3u) * 8u;
		const uint32_t word = state_word(widx);
		uint32_t st = (word >> sh8) & 0xffu;
		if(st & 1u) {
			canMoveThru = false;
		} else if(row > 0) {
			if(ct == CT_E) {
				const int origMask = col > 0 ? (int)((fl >> 3) & 3u) : 0;
				const uint32_t es = (st >> 4) & 3u;
				const int mask = es ? (es == 2u ? 2 : 0) : origMask;
				if(mask == 3)      { cur = MV_READ_OPEN; st = (st & ~(3u << 4)) | (2u << 4); branch = true; }
				else if(mask == 2) { cur = MV_RDGAP_EXT; st = (st & ~(3u << 4)) | (1u << 4); }
				else if(mask == 1) { cur = MV_READ_OPEN; st = (st & ~(3u << 4)) | (1u << 4); }
				else { empty = true; canMoveThru = (origMask == 0); }
			} else if(ct == CT_F) {
				const int origMask = (int)((fl >> 5) & 3u);
				const uint32_t fs = (st >> 6) & 3u;
				const int mask = fs ? (fs == 2u ? 2 : 0) : origMask;
				if(mask == 3)      { cur = MV_REF_OPEN;  st = (st & ~(3u << 6)) | (2u << 6); branch = true; }
				else if(mask == 2) { cur = MV_RFGAP_EXT; st = (st & ~(3u << 6)) | (1u << 6); }
				else if(mask == 1) { cur = MV_REF_OPEN;  st = (st & ~(3u << 6)) | (1u << 6); }
				else { empty = true; canMoveThru = (origMask == 0); }
			} else {
				const bool hasLeft = col > 0;
				int origMask = 0;
				if(row >= glo && row < ghi) { // gaps allowed in this row
					if(fl & CF_HF) origMask |= (int)(((fl >> 5) & 1u) | (((fl >> 6) & 1u) << 2));             // H_up - rfgo (1), F_up - rfge (4)
					if(hasLeft && (fl & CF_HE)) origMask |= (int)((((fl >> 3) & 1u) << 1) | (((fl >> 4) & 1u) << 3)); // H_left - rdgo (2), E_left - rdge (8)
				}
				if(hasLeft && (fl & CF_D)) origMask |= 16;
				const uint32_t hs = (st >> 1) & 7u;
				int mask = hs ? lowest_options(origMask, (int)hs - 1) : origMask;
				int select = -1;
				const int opts = BT2_POPC((unsigned)mask);
				if(opts == 1) {
					select = BT2_CTZ((unsigned)mask);
					st = (st & ~(7u << 1)) | (1u << 1);
				} else if(opts > 1) {
					if(mask & 16) select = 4; else if(mask & 1) select = 0; else if(mask & 4) select = 2;
					else if(mask & 2) select = 1; else select = 3;
					st = (st & ~(7u << 1)) | ((uint32_t)opts << 1);   // opts - 1 options left
					branch = true;
				}
				if(select != -1) {
					cur = select == 4 ? MV_DIAG : select == 0 ? MV_REF_OPEN : select == 1 ? MV_READ_OPEN : select == 2 ? MV_RFGAP_EXT : MV_RDGAP_EXT;
				} else { empty = true; canMoveThru = (origMask == 0); }
			}
		}
		masks[widx] = (word & ~(0xffu << sh8)) | ((st | 1u) << sh8);
		if(!canMoveThru) {
			if(nstack > 0) {
				const BtFrame f = stack[--nstack];
				ned = f.a & 1023u; row = (int)((f.a >> 10) & 255u); col = (int)((f.a >> 18) & 255u);
				ct = (int)((f.a >> 26) & 3u); coreHit = ((f.a >> 28) & 1u) != 0;
				readGaps = (int)(f.b & 1023u); refGaps = (int)((f.b >> 10) & 1023u); ns = (int)((f.b >> 20) & 1023u);
				score = (int)f.c;
				locate();
				return; // retry the same cell with its remaining options
			}
			attempt_failed();
			return;
		}
		// this cell joins the path: core-diagonal test (aligner_swsse_ee_u8.cpp:1229-1262) folded in
		const int dg = col - row;
		const bool cellCore = dg >= dlo && dg <= dhi;
		if(empty || row == 0) { coreHit = coreHit || cellCore; phase = BP_POST; return; }
		if(branch && nstack < BT2_BT_STACK) {
			BtFrame f;
			f.a = (ned & 1023u) | ((uint32_t)row << 10) | ((uint32_t)col << 18) | ((uint32_t)ct << 26) | ((coreHit ? 1u : 0u) << 28);
			f.b = ((uint32_t)readGaps & 1023u) | (((uint32_t)refGaps & 1023u) << 10) | (((uint32_t)ns & 1023u) << 20);
			f.c = (uint32_t)score; f.d = 0;
			stack[nstack++] = f;
		}
		coreHit = coreHit || cellCore;
		if(cur == MV_DIAG) {
			ct = CT_H;
			take_cell();
			row_dec(); col_dec();
		} else if(cur == MV_REF_OPEN || cur == MV_RFGAP_EXT) {
			const int readc = view().rd((uint32_t)row);
			push_edit(pack_edit((uint32_t)row, '-', (uint32_t)"ACGTN"[readc], BT2GPU_EDIT_REF_GAP));
			row_dec();
			ct = (cur == MV_REF_OPEN) ? CT_H : CT_F;
			score -= (cur == MV_REF_OPEN) ? P->rfGapOpen : P->rfGapExt;
			refGaps++;
		} else {
			const int refm = pb->rf[col];
			push_edit(pack_edit((uint32_t)row + 1, (uint32_t)mask_to_char(refm), '-', BT2GPU_EDIT_READ_GAP));
			col_dec();
			ct = (cur == MV_READ_OPEN) ? CT_H : CT_E;
			score -= (cur == MV_READ_OPEN) ? P->rdGapOpen : P->rdGapExt;
			readGaps++;
		}
	}

	// end of a backtrace that reached row 0 / an empty cell, then the rest of nextAlignment and of
	// the extendSeeds body for one alignment: redundancy, report, policy, tighten
	BT2_HD void post() {
		const DevParams& Pr = *P; const Problem& p = *pb;
		if(!coreHit) { attempt_failed(); return; }
		take_cell();
		ReadState& rs = *this->rs;
		if(ns > rs.nceil || ned > BT2_SCRATCH_EDITS) { attempt_failed(); return; } // (> capacity: beyond the reference's own DList<Edit,160>)
		const uint32_t ne = ned;
		for(uint32_t i = 0; i < ne / 2; i++) { const uint32_t t = scratch[i]; scratch[i] = scratch[ne - 1 - i]; scratch[ne - 1 - i] = t; }
		rs.scr_cur = (uint16_t)ned; rs.scr_hw = (uint16_t)hw;
		rs.rnd.init(reseed + 1);
		cural++;
		phase = BP_CAND;
		AlnRec a;
		a.fw = p.fw; a.tidx = p.tidx; a.refoff = (int64_t)col + p.refl;
		a.score = score; a.ns = ns; a.gaps = readGaps + refGaps; a.edits = (int32_t)ne;
		a.basesAligned = (int32_t)p.rows - (int32_t)ne;   // trimEnd is 0: candidates come from the last row only
		int refns = 0;
		for(int i = col; i <= origCol; i++) if(p.rf[i] > 15) refns++;
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
		if(phase == BP_STEP) bt_step();
		else if(phase == BP_CAND) cand();
		else if(phase == BP_POST) post();
		return phase != BP_DONE;
	}
};

} // namespace bt2
