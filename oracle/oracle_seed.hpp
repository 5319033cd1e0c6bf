// TEST INFRASTRUCTURE ONLY -- restatement of the reference's seed stage:
// seed layout, exact FM search, left extension, per-read hit bookkeeping,
// ranking and SA-tuple prioritisation.  See oracle_fm.hpp for usage rules.
#pragma once
#include <cstdint>
#include <vector>
#include <algorithm>
#include "oracle_fm.hpp"
#include "oracle_rng.hpp"

namespace orc {

struct WorkCounters {
	uint64_t seeds = 0, lfSearch = 0, lfExtend = 0, sides = 0, saRows = 0, lfResolve = 0,
	         dps = 0, cells = 0, btAttempts = 0, btSteps = 0, alns = 0;
};

// One read with both strands (Read::patFw/patRc/qual/qualRev, read.h:39-270)
struct ReadView {
	int len = 0;
	std::vector<uint8_t> fw, rc, qfw, qrc;
	void init(const uint8_t* s, const uint8_t* q, int n) {
		len = n; fw.assign(s, s + n); qfw.assign(q, q + n);
		rc.resize(n); qrc.resize(n);
		for(int i = 0; i < n; i++) {
			uint8_t c = s[n - 1 - i];
			rc[i] = c > 3 ? 4 : (uint8_t)(3 - c);
			qrc[i] = q[n - 1 - i];
		}
	}
};

// SAKey / QKey::init (aligner_cache.h:123-150)
struct SAKey {
	uint64_t seq = 0; uint32_t len = 0xffffffffu;
	bool operator==(const SAKey& o) const { return seq == o.seq && len == o.len; }
};

// One (strand, offset index) seed hit: QVal + the SAVal its key maps to
struct SeedHit {
	bool valid = false;      // QVal::valid() (rangen_ != 0)
	SAKey key;
	uint64_t elts = 0;       // QVal::eltn_ = botf-topf of THIS seed
	int savIdx = -1;         // index into SeedState::savs (first insert of the key)
};

struct SAVal { SAKey key; uint64_t topf; uint64_t len; uint8_t nlex; };

// SeedResults (aligner_seed.h:560-1230) for one round of one read
struct SeedState {
	int rdlen = 0, off = 0, per = 1, numOffs = 0, seqLen = 0;
	std::vector<SeedHit> hitsFw, hitsRc;
	std::vector<bool> sortedFw, sortedRc;
	std::vector<SAVal> savs;            // AlignmentCache::samap_ in insertion order
	uint64_t nonzTot = 0, numElts = 0, numRanges = 0;
	std::vector<uint32_t> rankOffs; std::vector<bool> rankFws;
	int idx2off(int idx) const { return per * idx + off; }
	uint64_t averageHitsPerSeed() const { return nonzTot == 0 ? ~0ull : numElts / nonzTot; }
};

// SeedAligner::searchSeedBi for one seed (aligner_seed.cpp:757-957).
// Returns true iff the seed must be reported (need_reporting).
inline bool searchSeed(const Fm& fm, const uint8_t* seq, int n_steps, uint64_t& top, uint64_t& bot, WorkCounters& wc) {
	const bt2gpu_index* ix = fm.ix;
	int step = 0;
	top = bot = 0;
	const int ftabLen = ix->ftabChars;
	if(ftabLen > 1 && ftabLen <= n_steps) {
		uint64_t k = 0; // Ebwt::ftabSeqToInt(ftabLen, true, seq, off-ftabLen+1, false), bt2_idx.h:1384-1408
		for(int i = 0; i < ftabLen; i++) k = (k << 2) | seq[n_steps - ftabLen + i];
		top = fm.ftabHi(k); bot = fm.ftabLo(k + 1);
		if(bot - top == 0) return false;
		step += ftabLen;
	} else if(n_steps > 0) {
		int c = seq[n_steps - 1];
		top = ix->fchr[c]; bot = ix->fchr[c + 1];
		if(bot - top == 0) return false;
		step++;
	} else {
		top = 0; bot = ix->fchr[4];
	}
	if(step == n_steps) return true;
	while(step < n_steps) {
		const int c = seq[n_steps - step - 1];
		step++;
		uint64_t t, b;
		if(bot - top != 1) { // bloc valid: mapBiLFEx (bt2_idx.h:2474)
			wc.lfSearch++;
			if(c == 4) return false;
			t = fm.rank(c, top); b = fm.rank(c, bot);
		} else {             // mapLF1(row, l, c) (bt2_idx.h:2511)
			wc.lfSearch++;
			if(c == 4) return false;
			t = fm.mapLF1c(top, c);
			if(t == OFF_MASK) return false;
			b = t + 1;
		}
		if(b == t) return false;
		top = t; bot = b;
		if(step == n_steps) return true;
	}
	return false;
}

// MultiSeedAligner::extend (aligner_seed.cpp:603-672).  buf points at the
// strand buffer, lim = # characters available to the left of the seed, which
// are consumed from buf[lim-1] downwards.
inline uint32_t extendLeft(const Fm& fm, const uint8_t* buf, int lim, uint64_t top, uint64_t bot, WorkCounters& wc) {
	uint32_t nlex = 0;
	for(int ii = 0; ii < lim; ii++) {
		const int rdc = buf[lim - 1 - ii];
		if(bot - top != 1) {
			wc.lfExtend++;
			uint64_t t[4], b[4];
			for(int j = 0; j < 4; j++) { t[j] = fm.rank(j, top); b[j] = fm.rank(j, bot); }
			int nonz = -1; bool abort = false;
			const uint64_t origSz = bot - top;
			for(int j = 0; j < 4; j++) {
				if(b[j] > t[j]) {
					if(nonz >= 0) { abort = true; break; }
					nonz = j; top = t[j]; bot = b[j];
				}
			}
			if(abort || (nonz != rdc && rdc <= 3) || bot - top < origSz) break;
		} else {
			wc.lfExtend++;
			int c = fm.mapLF1(top);
			if(c != rdc && rdc <= 3) break;
			bot = top + 1;
		}
		if(++nlex == 255) break;
	}
	return nlex;
}

// Phases P1-P4 for one read and one seed round: prepareSeed, instantiateSeed,
// searchAllSeedsDoAll, searchAllSeedsFinalize + SeedResults::add.
inline void seedRound(const Fm& fm, const bt2gpu_params& P, const ReadView& rd, int offset, int interval,
                      SeedState& ss, WorkCounters& wc) {
	const int rdlen = rd.len, len = P.seedLen;
	int nseeds = 1; // SeedAligner::prepareSeed (aligner_seed.cpp:321-341)
	if(rdlen - offset > len) nseeds += (rdlen - offset - len) / interval;
	ss = SeedState();
	ss.rdlen = rdlen; ss.off = offset; ss.per = interval; ss.numOffs = nseeds;
	ss.seqLen = std::min(len, rdlen);
	ss.hitsFw.assign(nseeds, SeedHit()); ss.hitsRc.assign(nseeds, SeedHit());
	ss.sortedFw.assign(nseeds, false); ss.sortedRc.assign(nseeds, false);
	const int L = ss.seqLen;
	for(int fwi = 0; fwi < 2; fwi++) {
		const bool fw = (fwi == 0);
		if((fw && P.nofw) || (!fw && P.norc)) continue;
		const std::vector<uint8_t>& pat = fw ? rd.fw : rd.rc;
		for(int i = 0; i < nseeds; i++) {
			const int depth = i * interval + offset;
			const int so = fw ? depth : (rdlen - depth - L); // SeedAligner::instantiateSeq (:300-314)
			if(so < 0 || so + L > rdlen) continue;           // cannot happen given the round logic
			const uint8_t* seq = pat.data() + so;
			bool valid = true;                               // InstantiatedSeed::instantiateExact (:236-263)
			for(int k = 0; k < L; k++) if(seq[k] == 4) { valid = false; break; }
			if(!valid) continue;
			wc.seeds++;
			uint64_t top = 0, bot = 0;
			if(!searchSeed(fm, seq, L, top, bot, wc)) continue;
			uint32_t nlex = 0;
			if(P.doExtend) {
				const int rdoff = ss.idx2off(i);
				const int lim = fw ? rdoff : (rdlen - L - rdoff); // SeedResults::get_rel_offs (aligner_seed.h:1155-1167)
				if(lim > 0) nlex = extendLeft(fm, pat.data(), lim, top, bot, wc);
			}
			// searchAllSeedsFinalize -> AlignmentCache::addOnTheFly -> SeedResults::add
			SAKey key; key.len = (uint32_t)L; key.seq = 0;
			if(L > 32) { key.len = 0xffffffffu; }
			else for(int k = 0; k < L; k++) key.seq = (key.seq << 2) | seq[k];
			SeedHit& h = (fw ? ss.hitsFw : ss.hitsRc)[i];
			h.valid = true; h.key = key; h.elts = bot - top;
			int found = -1;
			for(size_t s = 0; s < ss.savs.size(); s++) if(ss.savs[s].key == key) { found = (int)s; break; }
			if(found < 0) {
				SAVal sv; sv.key = key; sv.topf = top; sv.len = bot - top; sv.nlex = (uint8_t)nlex;
				ss.savs.push_back(sv); found = (int)ss.savs.size() - 1;
			}
			h.savIdx = found;
			ss.numElts += h.elts; ss.numRanges += 1; ss.nonzTot++;
		}
	}
}

// SeedResults::rankSeedHits, non-`all` branch (aligner_seed.h:1000-1061)
inline void rankSeedHits(SeedState& ss, Rng& rnd) {
	ss.rankOffs.clear(); ss.rankFws.clear();
	const uint32_t numOffs = (uint32_t)ss.numOffs;
	while(ss.rankOffs.size() < ss.nonzTot) {
		uint64_t minsz = 0xffffffffull;
		uint32_t minidx = 0; bool minfw = true;
		const bool rb = rnd.nextBool();
		for(int fwi = 0; fwi <= 1; fwi++) {
			const bool fw = (fwi == (rb ? 1 : 0));
			std::vector<SeedHit>& rrs = fw ? ss.hitsFw : ss.hitsRc;
			std::vector<bool>& sorted = fw ? ss.sortedFw : ss.sortedRc;
			uint32_t i = rnd.nextU32() % numOffs;
			for(uint32_t ii = 0; ii < numOffs; ii++) {
				if(rrs[i].valid && rrs[i].elts > 0 && !sorted[i] && rrs[i].elts < minsz) {
					minsz = rrs[i].elts; minidx = i; minfw = fw;
				}
				if(++i == numOffs) i = 0;
			}
		}
		if(minfw) ss.sortedFw[minidx] = true; else ss.sortedRc[minidx] = true;
		ss.rankOffs.push_back(minidx); ss.rankFws.push_back(minfw);
	}
}

// SATupleAndPos (aligner_sw_driver.h:160-176) flattened
struct SatPos {
	SAKey key; uint8_t nlex = 0; uint64_t topf = 0; uint64_t size = 0; int savIdx = -1;
	bool fw = true; uint32_t offidx = 0, rdoff = 0, seedlen = 0; // SeedPos
	uint64_t origSz = 0;
	bool operator<(const SatPos& o) const { // SATuple::operator< then SeedPos::operator<
		if(size < o.size) return true; if(size > o.size) return false;
		if(topf < o.topf) return true; if(topf > o.topf) return false;
		if(offidx < o.offidx) return true; if(offidx > o.offidx) return false;
		if(rdoff < o.rdoff) return true; if(rdoff > o.rdoff) return false;
		if(seedlen < o.seedlen) return true; if(seedlen > o.seedlen) return false;
		if(fw && !o.fw) return true;
		return false;
	}
};

struct ExtendRange { uint64_t off, len, sz; };

// Per-read SwDriver state that outlives a seed round (aligner_sw_driver.h:492-536)
struct DriverState {
	std::vector<ExtendRange> seedExRangeFw, seedExRangeRc;
	std::vector<SatPos> satpos;      // satpos_
	std::vector<uint32_t> rand_ns;   // rand_ns_
	uint64_t nelt = 0;               // nelt_
};

// RowSampler (aligner_sw_driver.h:182-273)
struct RowSampler {
	double mass = 0.0; std::vector<bool> elim; std::vector<double> masses;
	void init(const std::vector<SatPos>& sl, size_t sai, size_t saf) {
		elim.assign(saf - sai, false); masses.assign(saf - sai, 0.0); mass = 0.0f;
		for(size_t i = sai; i < saf; i++) {
			double num = (double)((size_t)sl[i].nlex + 1); num *= num;
			double denom = (double)sl[i].size; denom *= denom;
			masses[i - sai] = num / denom;
			mass += masses[i - sai];
		}
	}
	void finishedRange(size_t i) { elim[i] = true; mass -= masses[i]; }
	size_t next(Rng& rnd) {
		double rd = rnd.nextFloat() * mass;
		double sofar = 0.0f;
		size_t last = ~(size_t)0;
		for(size_t i = 0; i < masses.size(); i++) {
			if(!elim[i]) { last = i; sofar += masses[i]; if(rd < sofar) return i; }
		}
		return last;
	}
};

// SwDriver::prioritizeSATups (aligner_sw_driver.cpp:61-260), all=false
inline void prioritizeSATups(const SeedState& ss, const bt2gpu_params& P, DriverState& ds, Rng& rnd) {
	const size_t nsm = 5, maxelt = (size_t)P.mxIter;
	ds.satpos.clear(); ds.rand_ns.clear();
	std::vector<SatPos> sp2;
	size_t nelt = 0, nsmall = 0, nsmall_elts = 0;
	for(size_t i = 0; i < ss.nonzTot; i++) {
		const bool fw = ss.rankFws[i];
		const uint32_t offidx = ss.rankOffs[i], rdoff = (uint32_t)ss.idx2off(offidx), seedlen = (uint32_t)ss.seqLen;
		const SeedHit& h = (fw ? ss.hitsFw : ss.hitsRc)[offidx];
		const SAVal& sv = ss.savs[h.savIdx];
		if(sv.len == 0) continue;  // satup invalid
		const size_t sz = sv.len;
		bool skip = false;
		if(P.seedMms == 0) {
			const std::vector<ExtendRange>& range = fw ? ds.seedExRangeFw : ds.seedExRangeRc;
			for(size_t k = 0; k < range.size(); k++) {
				if(range[k].off <= rdoff && range[k].off + range[k].len >= (uint64_t)(rdoff + seedlen)) {
					if(sz <= range[k].sz) { skip = true; break; }
				}
			}
		}
		if(!skip) for(size_t k = 0; k < sp2.size(); k++) skip |= (sp2[k].savIdx == h.savIdx);
		if(skip) continue;
		nelt += sz;
		SatPos s; s.key = sv.key; s.nlex = sv.nlex; s.topf = sv.topf; s.size = sz; s.savIdx = h.savIdx;
		s.fw = fw; s.offidx = offidx; s.rdoff = rdoff; s.seedlen = seedlen; s.origSz = sz;
		sp2.push_back(s);
		if(sz <= nsm) { nsmall++; nsmall_elts += sz; }
		if(P.seedMms == 0 && sv.nlex > 0) {
			ExtendRange er; er.off = rdoff - (fw ? sv.nlex : 0); er.len = seedlen + sv.nlex; er.sz = sz;
			(fw ? ds.seedExRangeFw : ds.seedExRangeRc).push_back(er);
		}
	}
	ds.nelt = nelt;
	std::sort(sp2.begin(), sp2.end());
	size_t nelt_added = 0;
	for(size_t j = 0; j < nsmall && nelt_added < maxelt; j++) {
		ds.satpos.push_back(sp2[j]);
		ds.rand_ns.push_back((uint32_t)sp2[j].size);
		nelt_added += sp2[j].size;
	}
	if(nelt_added >= maxelt || nsmall == sp2.size()) { ds.nelt = nelt_added; return; }
	RowSampler rowsamp; rowsamp.init(sp2, nsmall, sp2.size());
	std::vector<Random1toN> rands(sp2.size());
	while(nelt_added < maxelt && nelt_added < nelt) {
		size_t ri = rowsamp.next(rnd) + nsmall;
		if(!rands[ri].inited()) rands[ri].init((uint16_t)sp2[ri].size, false);
		size_t r = rands[ri].next(rnd);
		if(rands[ri].done()) rowsamp.finishedRange(ri - nsmall);
		SatPos s = sp2[ri];
		s.topf = sp2[ri].topf + r; s.nlex = 0; s.size = 1;
		ds.satpos.push_back(s);
		ds.rand_ns.push_back(1);
		nelt_added++;
	}
	ds.nelt = nelt_added;
}

} // namespace orc
