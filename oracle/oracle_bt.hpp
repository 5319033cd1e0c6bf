// TEST INFRASTRUCTURE ONLY -- restatement of the reference's backtrace:
// SwAligner::nextAlignment (aligner_sw.cpp:252-375) and
// SwAligner::backtraceNucleotidesEnd2EndSseU8 (aligner_swsse_ee_u8.cpp:746-1352;
// the i16 twin is identical up to the score offset).  See oracle_fm.hpp for
// usage rules.
#pragma once
#include "oracle_dp.hpp"

namespace orc {

enum { CT_E = 0, CT_F = 1, CT_H = 2 }; // SSEMatrixConsts (aligner_swsse.h:104-113)
enum { BT_DIAG, BT_REF_OPEN, BT_READ_OPEN, BT_RFGAP_EXT, BT_RDGAP_EXT };

// `ned` is the shared per-read edit scratch (SwDriver::resGap_.alres.ned_); it
// is cleared (count -> 0) on entry and left as it is on failure.
inline bool SwProblem::backtrace(EditScratch& ned, AlnOut& out, int row, int col, WorkCounters& wc) {
	struct Frame { size_t nedsz, celsz; int row, col; size_t gaps, readGaps, refGaps; int64_t score; int ns; int ct; };
	std::vector<Frame> stack;
	std::vector<std::pair<int,int> > btcells;
	int64_t score = 0; int ns = 0;
	const int origCol = col;
	size_t gaps = 0, readGaps = 0, refGaps = 0;
	ned.clear();
	const int rows = rdlen;
	const int trimEnd = rows - row - 1;
	int ct = CT_H;
	wc.btAttempts++;
	auto M = [&](int r, int c) -> uint16_t& { return masks[(size_t)r * rflen + c]; };
	while(row >= 0) {
		wc.btSteps++;
		const int readc = rd[row], refm = rf[col], readq = qu[row];
		bool empty = false, reportedThru, canMoveThru, branch = false;
		int cur = BT_DIAG;
		if(!rowReset[row]) { for(int c = 0; c < rflen; c++) M(row, c) = 0; rowReset[row] = 1; }
		reportedThru = (M(row, col) & 1) != 0;
		canMoveThru = true;
		if(reportedThru) {
			canMoveThru = false;
		} else if(row > 0) {
			const bool gapsAllowed = !(row < P->gapbar || (rows - row - 1) < P->gapbar);
			if(ct == CT_E) {
				const int64_t sc_cur = E[ix(row, col)];
				int origMask = 0, mask = 0;
				const int64_t sc_h_left = H[ix(row, col - 1)];
				if(sc_h_left - sc.rdGapOpen() == sc_cur) mask |= 1;
				const int64_t sc_e_left = E[ix(row, col - 1)];
				if(sc_e_left - sc.rdGapExt() == sc_cur) mask |= 2;
				origMask = mask;
				if(M(row, col) & (1 << 7)) mask = (M(row, col) >> 8) & 3;
				auto eSet = [&](int m) { M(row, col) &= ~(7 << 7); M(row, col) |= (1 << 7 | m << 8); };
				if(mask == 3) { cur = BT_READ_OPEN; eSet(2); branch = true; }
				else if(mask == 2) { cur = BT_RDGAP_EXT; eSet(0); }
				else if(mask == 1) { cur = BT_READ_OPEN; eSet(0); }
				else { empty = true; canMoveThru = (origMask == 0); }
			} else if(ct == CT_F) {
				const int64_t sc_h_up = H[ix(row - 1, col)], sc_f_up = F[ix(row - 1, col)], sc_cur = F[ix(row, col)];
				int origMask = 0, mask = 0;
				if(sc_h_up - sc.rfGapOpen() == sc_cur) mask |= 1;
				if(sc_f_up - sc.rfGapExt() == sc_cur) mask |= 2;
				origMask = mask;
				if(M(row, col) & (1 << 10)) mask = (M(row, col) >> 11) & 3;
				auto fSet = [&](int m) { M(row, col) &= ~(7 << 10); M(row, col) |= (1 << 10 | m << 11); };
				if(mask == 3) { cur = BT_REF_OPEN; fSet(2); branch = true; }
				else if(mask == 2) { cur = BT_RFGAP_EXT; fSet(0); }
				else if(mask == 1) { cur = BT_REF_OPEN; fSet(0); }
				else { empty = true; canMoveThru = (origMask == 0); }
			} else {
				const int64_t sc_cur = H[ix(row, col)];
				const int64_t sc_f_up = F[ix(row - 1, col)], sc_h_up = H[ix(row - 1, col)];
				const bool hasLeft = col > 0;
				const int64_t sc_h_left = hasLeft ? H[ix(row, col - 1)] : 0;
				const int64_t sc_e_left = hasLeft ? E[ix(row, col - 1)] : 0;
				const int64_t sc_h_upleft = hasLeft ? H[ix(row - 1, col - 1)] : 0;
				const int64_t sc_diag = sc.score(readc, refm, readq - 33);
				int origMask = 0, mask = 0;
				if(gapsAllowed) {
					if(sc_cur == sc_h_up - sc.rfGapOpen()) mask |= 1;
					if(hasLeft && sc_cur == sc_h_left - sc.rdGapOpen()) mask |= 2;
					if(sc_cur == sc_f_up - sc.rfGapExt()) mask |= 4;
					if(hasLeft && sc_cur == sc_e_left - sc.rdGapExt()) mask |= 8;
				}
				if(hasLeft && sc_cur == sc_h_upleft + sc_diag) mask |= 16;
				origMask = mask;
				if(M(row, col) & (1 << 1)) mask = (M(row, col) >> 2) & 31;
				auto hSet = [&](int m) { M(row, col) &= ~(31 << 1); M(row, col) |= (1 << 1 | m << 2); };
				const int opts = alts5[mask];
				int select = -1;
				if(opts == 1) { select = firsts5[mask]; hSet(0); }
				else if(opts > 1) {
					if(mask & 16) select = 4; else if(mask & 1) select = 0; else if(mask & 4) select = 2;
					else if(mask & 2) select = 1; else if(mask & 8) select = 3;
					mask &= ~(1 << select);
					hSet(mask);
					branch = true;
				}
				if(select != -1) {
					if(select == 4) cur = BT_DIAG; else if(select == 0) cur = BT_REF_OPEN;
					else if(select == 1) cur = BT_READ_OPEN; else if(select == 2) cur = BT_RFGAP_EXT;
					else cur = BT_RDGAP_EXT;
				} else { empty = true; canMoveThru = (origMask == 0); }
			}
		}
		M(row, col) |= 1; // setReportedThrough
		if(!canMoveThru) {
			if(!stack.empty()) {
				const Frame& f = stack.back();
				btcells.resize(f.celsz); ned.resize(f.nedsz);
				row = f.row; col = f.col; gaps = f.gaps; readGaps = f.readGaps; refGaps = f.refGaps;
				score = f.score; ns = f.ns; ct = f.ct;
				stack.pop_back();
				continue;
			}
			return false; // res.reset(): scratch contents stay
		}
		if(empty || row == 0) { btcells.push_back(std::make_pair(row, col)); break; }
		if(branch) {
			Frame f; f.nedsz = ned.size(); f.celsz = btcells.size(); f.row = row; f.col = col; f.gaps = gaps;
			f.readGaps = readGaps; f.refGaps = refGaps; f.score = score; f.ns = ns; f.ct = ct;
			stack.push_back(f);
		}
		btcells.push_back(std::make_pair(row, col));
		switch(cur) {
		case BT_DIAG: {
			const int m = (readc > 3 || refm > 15) ? -1 : ((refm & (1 << readc)) ? 1 : 0); // matchesEx (mask.h:37)
			ct = CT_H;
			if(m != 1) {
				Edit e; e.pos = (uint32_t)row; e.chr = (uint8_t)refMaskToChar(refm); e.qchr = (uint8_t)"ACGTN"[readc]; e.type = BT2GPU_EDIT_MM;
				ned.push_back(e);
				score -= sc.mm(readc, refm, readq - 33);
			} else score += P->matchBonus;
			if(m == -1) ns++;
			row--; col--;
			break; }
		case BT_REF_OPEN: case BT_RFGAP_EXT: {
			Edit e; e.pos = (uint32_t)row; e.chr = '-'; e.qchr = (uint8_t)"ACGTN"[readc]; e.type = BT2GPU_EDIT_REF_GAP;
			ned.push_back(e);
			row--;
			ct = (cur == BT_REF_OPEN) ? CT_H : CT_F;
			score -= (cur == BT_REF_OPEN) ? sc.rfGapOpen() : sc.rfGapExt();
			gaps++; refGaps++;
			break; }
		case BT_READ_OPEN: case BT_RDGAP_EXT: {
			Edit e; e.pos = (uint32_t)row + 1; e.chr = (uint8_t)refMaskToChar(refm); e.qchr = '-'; e.type = BT2GPU_EDIT_READ_GAP;
			ned.push_back(e);
			col--;
			ct = (cur == BT_READ_OPEN) ? CT_H : CT_E;
			score -= (cur == BT_READ_OPEN) ? sc.rdGapOpen() : sc.rdGapExt();
			gaps++; readGaps++;
			break; }
		}
	}
	// must overlap a core diagonal (aligner_swsse_ee_u8.cpp:1229-1262)
	bool overlappedCoreDiag = false;
	for(size_t i = 0; i < btcells.size(); i++) {
		int64_t diagi = (int64_t)btcells[i].second - btcells[i].first + (int64_t)rect.triml;
		if(diagi >= 0 && (size_t)diagi >= rect.corel && (size_t)diagi <= rect.corer) { overlappedCoreDiag = true; break; }
	}
	if(!overlappedCoreDiag) return false;
	{
		const int readc = rd[row], refm = rf[col];
		const int m = (readc > 3 || refm > 15) ? -1 : ((refm & (1 << readc)) ? 1 : 0);
		if(m != 1) {
			Edit e; e.pos = (uint32_t)row; e.chr = (uint8_t)refMaskToChar(refm); e.qchr = (uint8_t)"ACGTN"[readc]; e.type = BT2GPU_EDIT_MM;
			ned.push_back(e);
			score -= sc.mm(readc, refm, (int)qu[row] - 33);
		} else score += P->matchBonus;
		if(m == -1) ns++;
	}
	if(ns > nceil) return false;
	std::reverse(ned.begin(), ned.end()); // res.reverse()
	out.fw = fw; out.tidx = tidx; out.refoff = (int64_t)col + rect.refl;
	out.score = score; out.ns = ns; out.gaps = (int)gaps; out.edits = (int)ned.size();
	out.basesAligned = rdlen - 0 - trimEnd - out.edits;
	int refns = 0;
	for(int i = col; i <= origCol; i++) if(rf[i] > 15) refns++;
	out.refns = refns;
	return true;
}

// Edit::invertPoss with sort=false (edit.cpp:50-88) as called by
// AlnRes::invertEdits (aligner_result.h:852): reverse the list and flip every
// position to the other end of the read (read gaps sit between bases).
template<typename L> inline void invertEdits(L& ned, size_t sz) {
	std::reverse(ned.begin(), ned.end());
	for(size_t i = 0; i < ned.size(); i++) {
		ned[i].pos = (uint32_t)(sz - ned[i].pos - (ned[i].type == BT2GPU_EDIT_READ_GAP ? 0 : 1));
	}
}

inline bool SwProblem::nextAlignment(EditScratch& scratch, AlnOut& out, int64_t minscArg, Rng& rnd, WorkCounters& wc) {
	if(done()) return false;
	const size_t candsz = cands.size();
	while(cural < candsz) {
		if(cands[cural].score < minscArg) { cural++; continue; }
		const int row = cands[cural].row, col = cands[cural].col;
		if(rowReset[row] && (masks[(size_t)row * rflen + col] & 1)) { cural++; continue; }
		bool ret;
		const uint32_t reseed = rnd.nextU32() + 1;
		if(!i16) rnd.init(reseed);
		ret = backtrace(scratch, out, row, col, wc);
		rnd.init(i16 ? reseed : reseed + 1);
		if(ret) break;
		cural++;
	}
	if(cural == candsz) return false;
	if(!fw) invertEdits(scratch, (size_t)rdlen);
	out.ned.assign(scratch.begin(), scratch.end());
	cural++;
	return true;
}

} // namespace orc
