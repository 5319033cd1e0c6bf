// What follows a DP fill inside SwDriver::extendSeeds, one thread per problem:
//   result bookkeeping                               aligner_sw_driver.cpp:505-531
//   SwAligner::nextAlignment                         aligner_sw.cpp:252-375
//   SwAligner::backtraceNucleotidesEnd2EndSseU8      aligner_swsse_ee_u8.cpp:746-1352
//   RedundantAlns::overlap/add                       aligner_result.cpp:821-925
//   AlnSinkWrap::report + ReportingState::foundUnpaired  aln_sink.cpp:1395-1453, 107-140
//   minsc tightening                                 aligner_sw_driver.cpp:588-618
// The per-read edit scratch reproduces SwDriver::resGap_.alres.ned_.
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

struct BtFrame { uint32_t a, b, c, d; }; // packed branch-stack frame
#define BT2_BT_STACK 400

struct ReadView {
	const uint8_t* seq; const uint8_t* qual; uint32_t rdlen; bool fw;
	BT2_HD int rd(uint32_t row) const {
		if(fw) return seq[row];
		const uint8_t c = seq[rdlen - 1 - row];
		return c > 3 ? 4 : 3 - c;
	}
	BT2_HD int q(uint32_t row) const { return fw ? qual[row] : qual[rdlen - 1 - row]; }
};

struct BtCtx {
	const DevParams* P;
	const Problem* prob;
	const uint32_t* hef;   // skewed H|E<<8|F<<16 words of this problem
	uint16_t* masks;       // [rows][cols] of this problem slot: SSEMatrix::masks_ (low 13 bits) tagged with `gen`
	uint32_t gen;          // generation of this problem (1..7, 3 bits); a cell whose tag differs reads as 0, which
	                       // replaces the reference's lazy per-row memset (resetRow, aligner_swsse.h)
	BtFrame* stack;        // [BT2_BT_STACK] of this problem slot
	uint32_t* scratch;     // per-read edit scratch
	ReadView rv;
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

// One backtrace attempt from (row, col).  Edits are appended to the scratch 3'->5' and
// reversed on success.  Returns true and fills `out` on success.
BT2_HD bool backtrace(BtCtx& cx, ReadState& rs, AlnRec& out, int row, int col) {
	const DevParams& P = *cx.P; const Problem& pb = *cx.prob;
	const int rows = pb.rows, cols = pb.cols; const uint32_t R = pb.R;
	int score = 0, ns = 0, readGaps = 0, refGaps = 0;
	const int origCol = col;
	uint32_t nstack = 0;
	uint32_t ned = 0; // scratch count
	uint32_t hw = rs.scr_hw;
	#define PUSH_EDIT(x) do { if(ned < BT2_SCRATCH_EDITS) { cx.scratch[ned] = (x); } ned++; if(ned > hw) hw = ned; } while(0)
	bool coreHit = false;
	const int trimEnd = rows - row - 1;
	int ct = CT_H;
	rs.btAttempts++;
	const int rdgo = P.rdGapOpen, rdge = P.rdGapExt, rfgo = P.rfGapOpen, rfge = P.rfGapExt;
	#define HEF(r, c) cx.hef[hef_index((uint32_t)(r), (uint32_t)(c), R)]
	#define HV(w) ((int)((w) & 255u) - 255)
	#define EV(w) ((int)(((w) >> 8) & 255u) - 255)
	#define FV(w) ((int)(((w) >> 16) & 255u) - 255)
	#define MK(r, c) cx.masks[(uint32_t)(r) * (uint32_t)cols + (uint32_t)(c)]
	const uint32_t gtag = cx.gen << 13;
	while(row >= 0) {
		rs.btSteps++;
		const int readc = cx.rv.rd(row), refm = pb.rf[col], readq = cx.rv.q(row);
		bool empty = false, canMoveThru = true, branch = false;
		int cur = MV_DIAG;
		uint32_t m16 = MK(row, col);
		m16 = ((m16 & ~0x1fffu) == gtag) ? (m16 & 0x1fffu) : 0u;
		const bool reportedThru = (m16 & 1) != 0;
		if(reportedThru) {
			canMoveThru = false;
		} else if(row > 0) {
			const bool gapsAllowed = !(row < P.gapbar || (rows - row - 1) < P.gapbar);
			if(ct == CT_E) {
				const uint32_t wc = HEF(row, col), wl = col > 0 ? HEF(row, col - 1) : 0u;
				const int sc_cur = EV(wc);
				int mask = 0;
				if(HV(wl) - rdgo == sc_cur) mask |= 1;
				if(EV(wl) - rdge == sc_cur) mask |= 2;
				const int origMask = mask;
				if(m16 & (1u << 7)) mask = (m16 >> 8) & 3;
				if(mask == 3)      { cur = MV_READ_OPEN; m16 = (m16 & ~(7u << 7)) | (1u << 7) | (2u << 8); branch = true; }
				else if(mask == 2) { cur = MV_RDGAP_EXT; m16 = (m16 & ~(7u << 7)) | (1u << 7); }
				else if(mask == 1) { cur = MV_READ_OPEN; m16 = (m16 & ~(7u << 7)) | (1u << 7); }
				else { empty = true; canMoveThru = (origMask == 0); }
			} else if(ct == CT_F) {
				const uint32_t wc = HEF(row, col), wu = HEF(row - 1, col);
				const int sc_cur = FV(wc);
				int mask = 0;
				if(HV(wu) - rfgo == sc_cur) mask |= 1;
				if(FV(wu) - rfge == sc_cur) mask |= 2;
				const int origMask = mask;
				if(m16 & (1u << 10)) mask = (m16 >> 11) & 3;
				if(mask == 3)      { cur = MV_REF_OPEN;  m16 = (m16 & ~(7u << 10)) | (1u << 10) | (2u << 11); branch = true; }
				else if(mask == 2) { cur = MV_RFGAP_EXT; m16 = (m16 & ~(7u << 10)) | (1u << 10); }
				else if(mask == 1) { cur = MV_REF_OPEN;  m16 = (m16 & ~(7u << 10)) | (1u << 10); }
				else { empty = true; canMoveThru = (origMask == 0); }
			} else {
				const uint32_t wc = HEF(row, col), wu = HEF(row - 1, col);
				const bool hasLeft = col > 0;
				const uint32_t wl = hasLeft ? HEF(row, col - 1) : 0xffffffu, wul = hasLeft ? HEF(row - 1, col - 1) : 0xffffffu;
				const int sc_cur = HV(wc);
				const int sc_diag = sc_score(P, readc, refm, readq - 33);
				int mask = 0;
				if(gapsAllowed) {
					if(sc_cur == HV(wu) - rfgo) mask |= 1;
					if(hasLeft && sc_cur == HV(wl) - rdgo) mask |= 2;
					if(sc_cur == FV(wu) - rfge) mask |= 4;
					if(hasLeft && sc_cur == EV(wl) - rdge) mask |= 8;
				}
				if(hasLeft && sc_cur == HV(wul) + sc_diag) mask |= 16;
				const int origMask = mask;
				if(m16 & (1u << 1)) mask = (m16 >> 2) & 31;
				int select = -1;
				const int opts = BT2_POPC((unsigned)mask);
				if(opts == 1) {
					select = BT2_CTZ((unsigned)mask);
					m16 = (m16 & ~(31u << 1)) | (1u << 1);
				} else if(opts > 1) {
					if(mask & 16) select = 4; else if(mask & 1) select = 0; else if(mask & 4) select = 2;
					else if(mask & 2) select = 1; else select = 3;
					mask &= ~(1 << select);
					m16 = (m16 & ~(31u << 1)) | (1u << 1) | ((uint32_t)mask << 2);
					branch = true;
				}
				if(select != -1) {
					cur = select == 4 ? MV_DIAG : select == 0 ? MV_REF_OPEN : select == 1 ? MV_READ_OPEN : select == 2 ? MV_RFGAP_EXT : MV_RDGAP_EXT;
				} else { empty = true; canMoveThru = (origMask == 0); }
			}
		}
		m16 |= 1;
		MK(row, col) = (uint16_t)(gtag | m16);
		if(!canMoveThru) {
			if(nstack > 0) {
				const BtFrame f = cx.stack[--nstack];
				ned = f.a & 1023u; row = (int)((f.a >> 10) & 255u); col = (int)((f.a >> 18) & 255u);
				ct = (int)((f.a >> 26) & 3u); coreHit = ((f.a >> 28) & 1u) != 0;
				readGaps = (int)(f.b & 1023u); refGaps = (int)((f.b >> 10) & 1023u); ns = (int)((f.b >> 20) & 1023u);
				score = (int)f.c;
				continue;
			}
			rs.scr_cur = (uint16_t)ned; rs.scr_hw = (uint16_t)hw;
			return false;
		}
		// this cell joins the path: core-diagonal test (aligner_swsse_ee_u8.cpp:1229-1262) folded in
		const int diag = col - row + pb.triml;
		const bool cellCore = diag >= 0 && diag >= pb.corel && diag <= pb.corer;
		if(empty || row == 0) { coreHit = coreHit || cellCore; break; }
		if(branch) {
			if(nstack < BT2_BT_STACK) {
				BtFrame f;
				f.a = (ned & 1023u) | ((uint32_t)row << 10) | ((uint32_t)col << 18) | ((uint32_t)ct << 26) | ((coreHit ? 1u : 0u) << 28);
				f.b = ((uint32_t)readGaps & 1023u) | (((uint32_t)refGaps & 1023u) << 10) | (((uint32_t)ns & 1023u) << 20);
				f.c = (uint32_t)score; f.d = 0;
				cx.stack[nstack++] = f;
			}
		}
		coreHit = coreHit || cellCore;
		switch(cur) {
		case MV_DIAG: {
			const int m = (readc > 3 || refm > 15) ? -1 : ((refm & (1 << readc)) ? 1 : 0);
			ct = CT_H;
			if(m != 1) {
				PUSH_EDIT(pack_edit((uint32_t)row, (uint32_t)mask_to_char(refm), (uint32_t)"ACGTN"[readc], BT2GPU_EDIT_MM));
				score -= sc_mm(P, readc, refm, readq - 33);
			} else score += P.matchBonus;
			if(m == -1) ns++;
			row--; col--;
			break; }
		case MV_REF_OPEN: case MV_RFGAP_EXT: {
			PUSH_EDIT(pack_edit((uint32_t)row, '-', (uint32_t)"ACGTN"[readc], BT2GPU_EDIT_REF_GAP));
			row--;
			ct = (cur == MV_REF_OPEN) ? CT_H : CT_F;
			score -= (cur == MV_REF_OPEN) ? rfgo : rfge;
			refGaps++;
			break; }
		default: {
			PUSH_EDIT(pack_edit((uint32_t)row + 1, (uint32_t)mask_to_char(refm), '-', BT2GPU_EDIT_READ_GAP));
			col--;
			ct = (cur == MV_READ_OPEN) ? CT_H : CT_E;
			score -= (cur == MV_READ_OPEN) ? rdgo : rdge;
			readGaps++;
			break; }
		}
	}
	#undef HEF
	#undef HV
	#undef EV
	#undef FV
	#undef MK
	rs.scr_hw = (uint16_t)hw;
	rs.scr_cur = (uint16_t)ned;
	if(!coreHit) return false;
	{
		const int readc = cx.rv.rd(row), refm = pb.rf[col];
		const int m = (readc > 3 || refm > 15) ? -1 : ((refm & (1 << readc)) ? 1 : 0);
		if(m != 1) {
			PUSH_EDIT(pack_edit((uint32_t)row, (uint32_t)mask_to_char(refm), (uint32_t)"ACGTN"[readc], BT2GPU_EDIT_MM));
			score -= sc_mm(P, readc, refm, cx.rv.q(row) - 33);
		} else score += P.matchBonus;
		if(m == -1) ns++;
		rs.scr_cur = (uint16_t)ned; rs.scr_hw = (uint16_t)hw;
	}
	#undef PUSH_EDIT
	if(ns > rs.nceil) return false;
	if(ned > BT2_SCRATCH_EDITS) return false; // beyond the reference's own DList<Edit,160> capacity
	for(uint32_t i = 0; i < ned / 2; i++) { const uint32_t t = cx.scratch[i]; cx.scratch[i] = cx.scratch[ned - 1 - i]; cx.scratch[ned - 1 - i] = t; }
	out.fw = pb.fw; out.tidx = pb.tidx; out.refoff = (int64_t)col + pb.refl;
	out.score = score; out.ns = ns; out.gaps = readGaps + refGaps; out.edits = (int32_t)ned;
	out.basesAligned = rows - trimEnd - (int32_t)ned;
	int refns = 0;
	for(int i = col; i <= origCol; i++) if(pb.rf[i] > 15) refns++;
	out.refns = refns;
	out.n_edits = (uint16_t)ned;
	return true;
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

// Everything extendSeeds does with one filled DP problem.  Returns BT_ROUND_END with
// rs.status set when the round is over, BT_CONTINUE when the element loop goes on.
BT2_HD int after_fill(BtCtx& cx, const Pools& pools, ReadState& rs) {
	const DevParams& P = *cx.P; const Problem& pb = *cx.prob;
	const int32_t perfect = (int32_t)rs.rdlen * P.matchBonus;
	rs.dpCells += (uint64_t)pb.rows * pb.cols;
	const int32_t bestCell = pb.lrmax - 255;
	const bool found = bestCell >= pb.minsc && pb.ncand > 0;
	rs.nExDps++;
	if(!found) {
		rs.nExDpFails++; rs.nDpFail++;
		if(rs.nDpFail >= (uint32_t)P.streak) { rs.status = BT2GPU_ST_SOFT_LIMIT; return BT_ROUND_END; }
		if(bestCell > rs.bestLtMinsc) rs.bestLtMinsc = bestCell;
		return BT_CONTINUE;
	}
	rs.nExDpSuccs++; rs.nDpLastSucc = rs.nExDps - 1;
	if(rs.nDpFail > rs.nDpFailStreak) rs.nDpFailStreak = rs.nDpFail;
	rs.nDpFail = 0;
	uint32_t cural = 0;
	const uint32_t ncand = pb.ncand;
	const int lastRow = pb.rows - 1;
	Rng rnd = rs.rnd;
	int result = BT_CONTINUE;
	while(true) {
		if(cural == ncand) break;
		// ---- SwAligner::nextAlignment
		AlnRec a; bool got = false;
		while(cural < ncand) {
			const int col = pb.cand[cural] & 255;
			{ const uint32_t m = cx.masks[(uint32_t)lastRow * pb.cols + col];
			  if((m & ~0x1fffu) == (cx.gen << 13) && (m & 1u)) { cural++; continue; } } // start cell already reportedThrough
			const uint32_t reseed = rnd.nextU32() + 1;
			rnd.init(reseed);
			got = backtrace(cx, rs, a, lastRow, col);
			rnd.init(reseed + 1);
			if(got) break;
			cural++;
		}
		if(!got) break;
		cural++;
		const uint32_t ne = a.n_edits;
		if(!pb.fw) { // AlnRes::invertEdits: scratch becomes 5'-based
			for(uint32_t i = 0; i < ne / 2; i++) { const uint32_t t = cx.scratch[i]; cx.scratch[i] = cx.scratch[ne - 1 - i]; cx.scratch[ne - 1 - i] = t; }
			for(uint32_t i = 0; i < ne; i++) {
				const uint32_t x = cx.scratch[i];
				cx.scratch[i] = edit_with_pos(x, (uint32_t)rs.rdlen - edit_pos(x) - (edit_type(x) == BT2GPU_EDIT_READ_GAP ? 0u : 1u));
			}
		}
		// ---- RedundantAlns::overlap against every alignment already reported for this read
		EditView nv; nv.e = cx.scratch; nv.n = ne; nv.inv = !pb.fw; nv.rdlen = rs.rdlen;
		bool olap = false;
		for(uint32_t p = rs.aln_tail; p != BT2_NIL && !olap; p = pools.alns[p].prev) {
			const AlnRec& o = pools.alns[p];
			if(o.tidx != a.tidx || o.fw != a.fw) continue;
			EditView ov; ov.e = pools.edits + o.edit_off; ov.n = o.n_edits; ov.inv = !o.fw; ov.rdlen = rs.rdlen;
			olap = alns_overlap(nv, a.refoff, ov, o.refoff, rs.rdlen);
		}
		if(olap) continue;
		// ---- report
		const uint32_t slot = atomic_bump(pools.aln_top, 1);
		const uint32_t eoff = ne ? atomic_bump(pools.edit_top, ne) : 0;
		if(slot >= pools.aln_cap) { set_error(pools, 4); rs.status = BT2GPU_ST_HARD_LIMIT; result = BT_ROUND_END; break; }
		if(ne && eoff + ne > pools.edit_cap) { set_error(pools, 8); rs.status = BT2GPU_ST_HARD_LIMIT; result = BT_ROUND_END; break; }
		for(uint32_t i = 0; i < ne; i++) pools.edits[eoff + i] = cx.scratch[i];
		a.edit_off = eoff; a.minsc = rs.minsc; a.prev = rs.aln_tail; a.pad = 0; a.pad2 = 0;
		pools.alns[slot] = a;
		rs.aln_tail = slot; rs.n_alns++;
		// ReportingState::foundUnpaired + best / second best
		rs.nunpair++;
		const bool mhitsSet = P.mhits != 0;
		if(!rs.doneUnpair) {
			if(rs.nunpair >= (uint32_t)P.khits && !mhitsSet) rs.doneUnpair = 1;
			else if(mhitsSet && rs.nunpair > (uint32_t)P.mhits) rs.doneUnpair = 1;
			if(rs.doneUnpair) rs.donePolicy = 1;
		}
		if(a.score > rs.best) { rs.best2 = rs.best; rs.best = a.score; } else if(a.score > rs.best2) rs.best2 = a.score;
		if(rs.donePolicy) { rs.status = BT2GPU_ST_POLICY_FULFILLED; result = BT_ROUND_END; break; }
		if(P.tighten > 0 && mhitsSet && rs.best2 != BT2_I32_MIN) {
			const int32_t best = rs.best, sec = rs.best2;
			if(P.tighten == 1) {
				if(best >= rs.minsc) { rs.minsc = best; if(rs.minsc < perfect && best == sec) rs.minsc++; }
			} else if(P.tighten == 2) {
				if(sec >= rs.minsc) { rs.minsc = sec; if(rs.minsc < perfect) rs.minsc++; }
			} else {
				const int32_t diff = best - sec;
				const int32_t bot = sec + ((diff * 3) / 4);
				if(bot >= rs.minsc) { rs.minsc = bot; if(rs.minsc < perfect) rs.minsc++; }
			}
		}
	}
	rs.rnd = rnd;
	return result;
}

} // namespace bt2
