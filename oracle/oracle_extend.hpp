// TEST INFRASTRUCTURE ONLY -- restatement of SwDriver::extendSeeds
// (aligner_sw_driver.cpp:278-631), the sink state it reads back
// (AlnSinkWrap::report aln_sink.cpp:1395-1453, ReportingState aln_sink.cpp:107-330)
// and the per-read round loop of multiseedSearchWorker (bt2_search.cpp:2414-2862).
// See oracle_fm.hpp for usage rules.
#pragma once
#include "oracle_bt.hpp"

namespace orc {

struct Interval { uint64_t ref; int orient; int64_t off, len; };

// EIvalMergeListBinned (ival_list.h): observable behaviour = membership in the
// union of the added intervals of the same (ref, orientation).
struct SeenDiags {
	std::vector<Interval> iv;
	void reset() { iv.clear(); }
	void add(uint64_t ref, bool fw, int64_t off, int64_t len) { Interval i; i.ref = ref; i.orient = fw ? 1 : 0; i.off = off; i.len = len; iv.push_back(i); }
	bool locusPresent(uint64_t ref, bool fw, int64_t off) const {
		const int o = fw ? 1 : 0;
		for(size_t i = 0; i < iv.size(); i++)
			if(iv[i].ref == ref && iv[i].orient == o && off >= iv[i].off && off < iv[i].off + iv[i].len) return true;
		return false;
	}
};

// RedundantAlns (aligner_result.cpp:821-925): set of (ref, fw, refcol, row) cells
struct RedundantAlns {
	struct Cell { uint64_t rfid; bool fw; int64_t rfoff; size_t rdoff;
		bool operator==(const Cell& c) const { return rfid == c.rfid && fw == c.fw && rfoff == c.rfoff && rdoff == c.rdoff; } };
	std::vector<std::vector<Cell> > cells;
	void init(size_t n) { cells.assign(n, std::vector<Cell>()); }
	// walk the cells of an alignment; edits must be w.r.t. the upstream end, ascending
	template<typename FN> static void walk(const AlnOut& a, const std::vector<Edit>& ned, size_t len, FN fn) {
		int64_t left = a.refoff, right;
		size_t nedidx = 0;
		for(size_t i = 0; i < len; i++) {
			size_t diff = 1;
			right = left + 1;
			while(nedidx < ned.size() && ned[nedidx].pos == i) { if(ned[nedidx].type == BT2GPU_EDIT_REF_GAP) diff = 0; nedidx++; }
			if(i < len - 1) {
				size_t nx = nedidx;
				while(nx < ned.size() && ned[nx].pos == i + 1) { if(ned[nx].type == BT2GPU_EDIT_READ_GAP) right++; nx++; }
			}
			for(int64_t j = left; j < right; j++) if(fn(j, i)) return;
			left = right + (int64_t)diff - 1;
		}
	}
	bool overlap(const AlnOut& a, const std::vector<Edit>& nedUp, size_t len) const {
		bool olap = false;
		walk(a, nedUp, len, [&](int64_t j, size_t i) {
			Cell c; c.rfid = a.tidx; c.fw = a.fw; c.rfoff = j; c.rdoff = i;
			for(size_t k = 0; k < cells[i].size(); k++) if(cells[i][k] == c) { olap = true; return true; }
			return false; });
		return olap;
	}
	void add(const AlnOut& a, const std::vector<Edit>& nedUp, size_t len) {
		walk(a, nedUp, len, [&](int64_t j, size_t i) {
			Cell c; c.rfid = a.tidx; c.fw = a.fw; c.rfoff = j; c.rdoff = i; cells[i].push_back(c); return false; });
	}
};

// The sink state extendSeeds reads back (aln_sink.h:1157-1210, aln_sink.cpp:107-330)
struct SinkState {
	uint64_t nunpair = 0; bool doneUnpair = false, done = false;
	int64_t best = LLONG_MIN, best2 = LLONG_MIN;
	int khits, mhits; bool msample;
	void nextRead(const bt2gpu_params& P) { nunpair = 0; doneUnpair = done = false; best = best2 = LLONG_MIN;
		khits = P.khits; mhits = P.mhits; msample = P.msample != 0; }
	bool mhitsSet() const { return msample; } // ReportingParams::mhitsSet (aln_sink.h)
	bool report(int64_t score) {   // foundUnpaired(true) + best/second-best update
		nunpair++;
		if(!doneUnpair) {
			if(nunpair >= (uint64_t)khits && !mhitsSet()) doneUnpair = true;
			else if(mhitsSet() && nunpair > (uint64_t)mhits) doneUnpair = true;
			if(doneUnpair) done = true; // updateDone(): concordant/discordant/unpair2 are done from the start
		}
		if(score > best) { best2 = best; best = score; } else if(score > best2) best2 = score;
		return done;
	}
	bool Mmode() const { return msample; }
	bool hasSecondBest() const { return best2 != LLONG_MIN; }
	bool doneWithMate() const { return doneUnpair; }
};

struct PerRead {
	uint32_t nExIters = 0, nExDps = 0, nExDpSuccs = 0, nExDpFails = 0, nDpFail = 0, nDpFailStreak = 0, nDpLastSucc = 0, nRedundants = 0;
	int64_t bestLtMinsc = LLONG_MIN;
};

struct ReadOutcome {
	std::vector<AlnOut> alns; std::vector<int64_t> alnMinsc;
	EditScratch scratch;             // final contents of the aliased edit scratch (F14)
	Rng rnd; int64_t minsc = 0; int status = 0; unsigned flags = 0; int rounds = 0; bool donePolicy = false;
	PerRead prm; SeedState lastSeeds; uint32_t seedsTried = 0;
};

// DynProgFramer::frameSeedExtensionRect with trimToRef (dp_framer.cpp:81-125)
inline bool frameRect(int64_t off, size_t rdlen, int64_t reflen, size_t maxrdgap, size_t maxrfgap, size_t maxhalf, DPRect& rect) {
	size_t maxgap = std::max(maxrdgap, maxrfgap); maxgap = std::min(maxgap, maxhalf);
	int64_t refl = off - 2 * (int64_t)maxgap, refr = off + ((int64_t)rdlen - 1) + 2 * (int64_t)maxgap;
	size_t triml = 0, trimr = 0;
	const int64_t maxns = 0;
	if(refr >= reflen + maxns) trimr = (size_t)(refr - (reflen + maxns - 1));
	if(refl < -maxns) triml = (size_t)(-refl) - (size_t)maxns;
	rect.refl_pretrim = refl; rect.refr_pretrim = refr;
	rect.refl = refl + (int64_t)triml; rect.refr = refr - (int64_t)trimr;
	rect.triml = triml; rect.trimr = trimr; rect.maxgap = maxgap;
	rect.corel = maxgap; rect.corer = rect.corel + 2 * maxgap;
	return !(rect.refr < rect.refl);
}

struct ExtendCtx {
	const Fm* fm; const Ref* ref; const bt2gpu_params* P; const ReadView* rd;
	SeenDiags seen; RedundantAlns red; SinkState sink; DriverState ds;
};

// SwDriver::extendSeeds; returns EXTEND_* (== BT2GPU_ST_*)
inline int extendSeeds(ExtendCtx& cx, ReadOutcome& out, int64_t& minsc, int nceilSw, int seedival, WorkCounters& wc) {
	const bt2gpu_params& P = *cx.P; const Fm& fm = *cx.fm; Scor sc; sc.P = &P;
	PerRead& prm = out.prm; Rng& rnd = out.rnd;
	prm.nDpFail = 0;
	const size_t nsm = 5;
	const size_t rdlen = (size_t)cx.rd->len, rows = rdlen;
	const int64_t perfectScore = (int64_t)rdlen * P.matchBonus * (P.matchBonus != 0); // Scoring::perfectScore, monotone => 0
	const size_t gws = cx.ds.satpos.size();
	std::vector<Random1toN> rands(gws);
	for(size_t i = 0; i < gws; i++) rands[i].init((uint16_t)cx.ds.rand_ns[i], false);
	size_t neltLeft = cx.ds.nelt;
	const size_t maxDp = (size_t)P.mxDp, maxIters = (size_t)P.mxIter, maxDpStreak = (size_t)P.streak;
	SwProblem sw; sw.P = &P; sw.sc = sc;
	while(neltLeft > 0) {
		if(minsc == perfectScore) return BT2GPU_ST_PERFECT_SCORE;
		for(size_t i = 0; i < gws; i++) {
			const SatPos& sp = cx.ds.satpos[i];
			const bool is_small = sp.size < nsm;
			const bool fw = sp.fw;
			uint32_t rdoff = sp.rdoff; const uint32_t seedhitlen = sp.seedlen;
			if(!fw) rdoff = (uint32_t)(rdlen - rdoff - seedhitlen);
			bool first = true;
			while(!rands[i].done() && (first || is_small)) {
				if(minsc == perfectScore) return BT2GPU_ST_PERFECT_SCORE;
				if(prm.nExDps >= maxDp) return BT2GPU_ST_HARD_LIMIT;
				if(0 >= (size_t)P.mxUg) return BT2GPU_ST_HARD_LIMIT; // prm.nExUgs is never incremented
				if(prm.nExIters >= maxIters) return BT2GPU_ST_HARD_LIMIT;
				prm.nExIters++;
				first = false;
				const size_t elt = rands[i].next(rnd);
				wc.saRows++;
				const uint64_t joff = fm.saOffset(sp.topf + elt, &wc.lfResolve);
				neltLeft--;
				uint64_t tidx = 0, toff = 0, tlen = 0; bool straddled = false;
				fm.joinedToTextOff(sp.key.len, joff, tidx, toff, tlen, straddled);
				const int64_t refoff = (int64_t)toff - rdoff;
				if(cx.seen.locusPresent(tidx, fw, refoff)) { prm.nRedundants++; continue; }
				const int readGaps = sc.maxReadGaps(minsc, rdlen), refGaps = sc.maxRefGaps(minsc, rdlen);
				DPRect rect;
				bool found = frameRect(refoff, rows, (int64_t)tlen, (size_t)readGaps, (size_t)refGaps, (size_t)P.maxhalf, rect);
				cx.seen.add(tidx, fw, refoff, 1);
				if(!found) continue;
				// SwAligner::initRef (aligner_sw.cpp:117-229)
				{
					const int64_t rfi = rect.refl, rff = rect.refr + 2;
					const size_t rflenFull = (size_t)(rff - rfi);
					if(rflenFull > BT2GPU_MAX_COLS) { prm.nExDpFails++; prm.nDpFail++; return BT2GPU_ST_HARD_LIMIT; }
					size_t leftNs = rfi >= 0 ? 0 : (size_t)(-rfi); leftNs = std::min(leftNs, rflenFull);
					size_t rightNs = rff <= (int64_t)tlen ? 0 : (size_t)(rff - (int64_t)tlen); rightNs = std::min(rightNs, rflenFull);
					const size_t inner = rflenFull - (leftNs + rightNs);
					std::vector<uint8_t> buf(rflenFull, 4);
					if(inner > 0) cx.ref->getStretch(buf.data() + leftNs, tidx, rfi < 0 ? 0 : (uint64_t)rfi, inner);
					sw.rf.resize(rflenFull);
					for(size_t k = 0; k < rflenFull; k++) sw.rf[k] = (uint8_t)(1 << buf[k]);
					sw.rflen = (int)(rflenFull - 1);
					// second initRef overload (aligner_sw.cpp:66-110)
					int64_t ms = minsc;
					sw.i16 = P.i16 ? !(P.enable8 && ms >= -254) : false;
					if(!P.i16) ms = std::max<int64_t>(ms, -254);
					sw.minsc = ms;
					sw.fw = fw; sw.rd = fw ? cx.rd->fw.data() : cx.rd->rc.data(); sw.qu = fw ? cx.rd->qfw.data() : cx.rd->qrc.data();
					sw.rdlen = (int)rdlen; sw.nceil = nceilSw; sw.tidx = tidx; sw.reflen = (int64_t)tlen; sw.rect = rect;
					if((size_t)sw.rflen > BT2GPU_MAX_COLS) { prm.nExDpFails++; prm.nDpFail++; return BT2GPU_ST_HARD_LIMIT; }
				}
				cx.seen.add(tidx, fw, rect.refl_pretrim + (int64_t)rect.corel, (int64_t)(rect.corer - rect.corel + 1));
				int64_t bestCell = LLONG_MIN;
				found = sw.align(bestCell, wc);
				prm.nExDps++;
				if(!found) {
					prm.nExDpFails++; prm.nDpFail++;
					if(prm.nDpFail >= maxDpStreak) return BT2GPU_ST_SOFT_LIMIT;
					if(bestCell > LLONG_MIN && bestCell > prm.bestLtMinsc) prm.bestLtMinsc = bestCell;
					continue;
				}
				prm.nExDpSuccs++; prm.nDpLastSucc = prm.nExDps - 1;
				if(prm.nDpFail > prm.nDpFailStreak) prm.nDpFailStreak = prm.nDpFail;
				prm.nDpFail = 0;
				while(true) {
					// resGap_.reset(): count -> 0, contents kept (F14)
					if(sw.done()) break;
					AlnOut a;
					if(!sw.nextAlignment(out.scratch, a, minsc, rnd, wc)) break;
					// refival / overhang tests: reportOverhangs=false and trimToRef => always inside the reference
					// RedundantAlns::overlap/add want upstream-relative edits; invert-and-restore for rc
					std::vector<Edit> up = a.ned;
					if(!fw) invertEdits(up, rdlen);
					if(cx.red.overlap(a, up, rdlen)) continue;
					cx.red.add(a, up, rdlen);
					out.alns.push_back(a); out.alnMinsc.push_back(minsc);
					wc.alns++;
					if(cx.sink.report(a.score)) return BT2GPU_ST_POLICY_FULFILLED;
					if(P.tighten > 0 && cx.sink.Mmode() && cx.sink.hasSecondBest()) {
						const int64_t best = cx.sink.best, sec = cx.sink.best2;
						if(P.tighten == 1) {
							if(best >= minsc) { minsc = best; if(minsc < perfectScore && best == sec) minsc++; }
						} else if(P.tighten == 2) {
							if(sec >= minsc) { minsc = sec; if(minsc < perfectScore) minsc++; }
						} else {
							const int64_t diff = best - sec;
							const int64_t bot = sec + ((diff * 3) / 4);
							if(bot >= minsc) { minsc = bot; if(minsc < perfectScore) minsc++; }
						}
					}
				}
			}
		}
	}
	return BT2GPU_ST_EXHAUSTED;
}

// One read through all its seed rounds (bt2_search.cpp:2485-2586, 2784-2826)
inline void alignRead(const Fm& fm, const Ref& ref, const bt2gpu_params& P, const uint8_t* seq, const uint8_t* qual, int rdlen,
                      uint32_t seed, int32_t minsc0, int32_t nceilSw, int32_t interval, ReadOutcome& out, WorkCounters& wc) {
	out = ReadOutcome();
	out.rnd.init(seed);
	out.minsc = minsc0;
	if(rdlen == 0) { out.flags |= BT2GPU_RF_EMPTY; return; }
	if(rdlen > BT2GPU_MAX_ROWS) { out.flags |= BT2GPU_RF_TOO_LONG; return; } // SwAligner::initRead fails; read ends unaligned
	ReadView rd; rd.init(seq, qual, rdlen);
	ExtendCtx cx; cx.fm = &fm; cx.ref = &ref; cx.P = &P; cx.rd = &rd;
	cx.red.init((size_t)rdlen); cx.sink.nextRead(P);
	int64_t minsc = minsc0;
	const int nrounds = std::min<int>(P.nSeedRounds, interval);
	for(int roundi = 0; roundi < nrounds; roundi++) {
		const int offset = (interval * roundi) / nrounds;
		if(offset > 0 && P.seedLen + offset > rdlen) break;
		out.rounds++;
		SeedState ss;
		seedRound(fm, P, rd, offset, interval, ss, wc);
		out.lastSeeds = ss; out.seedsTried += (uint32_t)(2 * ss.numOffs);
		if(ss.numRanges == 0 || ss.nonzTot == 0) { out.flags |= BT2GPU_RF_NO_SEED_HITS; break; }
		rankSeedHits(ss, out.rnd);
		out.lastSeeds = ss;
		prioritizeSATups(ss, P, cx.ds, out.rnd);
		const int ret = extendSeeds(cx, out, minsc, nceilSw, interval, wc);
		out.status = ret;
		if(ret == BT2GPU_ST_EXHAUSTED || ret == BT2GPU_ST_SOFT_LIMIT || ret == BT2GPU_ST_POLICY_FULFILLED) {
			if(ss.averageHitsPerSeed() < (uint64_t)P.seedBoostThresh) break;
			if(cx.sink.doneWithMate()) break;
		} else break;
	}
	out.minsc = minsc;
	out.donePolicy = cx.sink.done;
}

} // namespace orc
