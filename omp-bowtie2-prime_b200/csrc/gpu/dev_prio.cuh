// Per-read glue between the seed kernel and the extension loop, one thread per read:
//   SeedAligner::searchAllSeedsFinalize + SeedResults::add   aligner_seed.cpp:508-540, aligner_seed.h:603-636
//   AlignmentCache::addOnTheFly (first insert of a key wins)  aligner_cache.h
//   SeedResults::rankSeedHits                                 aligner_seed.h:1000-1061
//   SwDriver::prioritizeSATups, RowSampler                    aligner_sw_driver.cpp:61-260, aligner_sw_driver.h:182-273
//   Random1toN                                                random_util.h:31-232
#pragma once
#include "dev_common.cuh"

namespace bt2 {

// per-read scratch of the glue kernels, W = 2*maxOffs entries each (hit index h = strand*maxOffs + offidx)
struct PrioWork {
	uint16_t* sav;     // [W] hit index of the first hit with the same key (AlignmentCache entry)
	uint8_t*  sorted;  // [W]
	uint16_t* rank;    // [W] ranked hit indices
	uint16_t* sp2;     // [W] hit indices that survive the dedup, later sorted
	ExRange16* exr;    // [W] this round's seedExRange additions
	uint8_t*  exrfw;   // [W]
	uint32_t  nsp, nsmall, nexr;
	uint64_t  nelt;    // total rows over sp2
	uint32_t  nelt_added;
};

struct SeedGeom { uint32_t numOffs, off, per, seqLen, maxOffs; };

BT2_HD SeedGeom seed_geom(const DevParams& P, uint32_t rdlen, uint32_t offset, uint32_t interval, uint32_t maxOffs) {
	SeedGeom g; // SeedAligner::prepareSeed (aligner_seed.cpp:321-341)
	g.numOffs = 1;
	if((int)rdlen - (int)offset > P.seedLen) g.numOffs += (rdlen - offset - (uint32_t)P.seedLen) / interval;
	g.off = offset; g.per = interval;
	g.seqLen = (uint32_t)P.seedLen < rdlen ? (uint32_t)P.seedLen : rdlen;
	g.maxOffs = maxOffs;
	return g;
}

// Stage 1: bookkeeping + ranking + dedup + sort + the "small" ranges.  Returns true
// if row sampling (stage 2) is still needed.
BT2_HD bool prio_stage1(const DevParams& P, const Pools& pools, ReadState& rs, const SeedHit* hits,
                        const SeedGeom& g, PrioWork& w) {
	const uint32_t n = g.numOffs, M = g.maxOffs;
	// ---- SeedResults::add in search order: fw seeds by offset, then rc seeds
	uint32_t nonz = 0, nranges = 0; uint64_t nelts = 0;
	for(uint32_t s = 0; s < 2; s++) {
		for(uint32_t i = 0; i < n; i++) {
			const uint32_t h = s * M + i;
			w.sorted[h] = 0;
			if(!hits[h].valid) continue;
			uint32_t found = h;
			for(uint32_t s2 = 0; s2 <= s && found == h; s2++) {
				const uint32_t lim = (s2 == s) ? i : n;
				for(uint32_t j = 0; j < lim; j++) {
					const uint32_t h2 = s2 * M + j;
					if(hits[h2].valid && hits[h2].key == hits[h].key) { found = w.sav[h2]; break; }
				}
			}
			w.sav[h] = (uint16_t)found;
			nelts += hits[h].len; nranges++; nonz++;
		}
	}
	rs.nonzTot = nonz; rs.numRanges = nranges; rs.numElts = nelts;
	rs.nSeedsTried += 2 * n;
	w.nsp = 0; w.nsmall = 0; w.nexr = 0; w.nelt = 0; w.nelt_added = 0;
	if(nranges == 0 || nonz == 0) return false;
	// ---- rankSeedHits (non-`all` branch)
	Rng rnd = rs.rnd;
	uint32_t nrank = 0;
	while(nrank < nonz) {
		uint64_t minsz = 0xffffffffull; uint32_t minidx = 0; bool minfw = true;
		const bool rb = rnd.nextBool();
		for(int fwi = 0; fwi <= 1; fwi++) {
			const bool fw = (fwi == (rb ? 1 : 0));
			const uint32_t base = fw ? 0 : M;
			uint32_t i = rnd.nextU32() % n;
			for(uint32_t ii = 0; ii < n; ii++) {
				const SeedHit& hh = hits[base + i];
				if(hh.valid && hh.len > 0 && !w.sorted[base + i] && hh.len < minsz) { minsz = hh.len; minidx = i; minfw = fw; }
				if(++i == n) i = 0;
			}
		}
		const uint32_t h = (minfw ? 0 : M) + minidx;
		w.sorted[h] = 1;
		w.rank[nrank++] = (uint16_t)h;
	}
	// ---- prioritizeSATups: dedup
	const uint32_t nsm = 5, seedlen = g.seqLen;
	uint64_t nelt = 0; uint32_t nsp = 0, nsmall = 0, nexr = 0;
	for(uint32_t i = 0; i < nonz; i++) {
		const uint32_t h = w.rank[i];
		const bool fw = h < M;
		const uint32_t offidx = fw ? h : h - M;
		const uint32_t rdoff = g.per * offidx + g.off;
		const SeedHit& sv = hits[w.sav[h]];
		const uint64_t sz = sv.len;
		if(sz == 0) continue;
		bool skip = false;
		if(P.seedMms == 0) {
			// ranges recorded by earlier rounds (pool lists) and by this round (w.exr)
			uint32_t c = fw ? rs.exfw_head : rs.exrc_head;
			while(c != BT2_NIL && !skip) {
				const Chunk& ch = pools.chunks[c];
				for(uint32_t k = 0; k < ch.n; k++) {
					const uint64_t esz = ch.item[k][0]; const uint32_t eoff = (uint32_t)ch.item[k][1], elen = (uint32_t)(ch.item[k][1] >> 32);
					if(eoff <= rdoff && eoff + elen >= rdoff + seedlen && sz <= esz) { skip = true; break; }
				}
				c = ch.next;
			}
			for(uint32_t k = 0; k < nexr && !skip; k++) {
				if((w.exrfw[k] != 0) != fw) continue;
				if(w.exr[k].off <= rdoff && w.exr[k].off + w.exr[k].len >= rdoff + seedlen && sz <= w.exr[k].sz) skip = true;
			}
		}
		if(!skip) for(uint32_t k = 0; k < nsp; k++) if(w.sav[w.sp2[k]] == w.sav[h]) { skip = true; break; }
		if(skip) continue;
		nelt += sz;
		w.sp2[nsp++] = (uint16_t)h;
		if(sz <= nsm) nsmall++;
		if(P.seedMms == 0 && sv.nlex > 0) {
			w.exr[nexr].off = rdoff - (fw ? sv.nlex : 0); w.exr[nexr].len = seedlen + sv.nlex; w.exr[nexr].sz = sz;
			w.exrfw[nexr] = fw ? 1 : 0; nexr++;
		}
	}
	rs.rnd = rnd;
	w.nsp = nsp; w.nsmall = nsmall; w.nexr = nexr; w.nelt = nelt;
	// a later seed round can only happen if averageHitsPerSeed() >= seedBoostThresh (bt2_search.cpp:2792);
	// only then must this round's extended ranges outlive the round
	if(nexr > 0 && (nelts / nonz) >= (uint64_t)P.seedBoostThresh && (uint32_t)rs.round + 1 < (uint32_t)P.nSeedRounds) {
		for(uint32_t k = 0; k < nexr; k++) {
			uint32_t& head = w.exrfw[k] ? rs.exfw_head : rs.exrc_head;
			list_push(pools, head, w.exr[k].sz, (uint64_t)w.exr[k].off | ((uint64_t)w.exr[k].len << 32));
		}
	}
	// ---- sort by (size, topf, offidx, rdoff, seedlen, fw first): insertion sort
	for(uint32_t a = 1; a < nsp; a++) {
		const uint16_t h = w.sp2[a];
		const SeedHit& sa = hits[w.sav[h]];
		const bool fwa = h < M; const uint32_t oa = fwa ? h : h - M;
		uint32_t b = a;
		while(b > 0) {
			const uint16_t hb = w.sp2[b - 1];
			const SeedHit& sb = hits[w.sav[hb]];
			const bool fwb = hb < M; const uint32_t ob = fwb ? hb : hb - M;
			bool less; // is a < b ?
			if(sa.len != sb.len) less = sa.len < sb.len;
			else if(sa.top != sb.top) less = sa.top < sb.top;
			else if(oa != ob) less = oa < ob;   // rdoff = per*offidx+off is monotone in offidx; seedlen is equal
			else less = (fwa && !fwb);
			if(!less) break;
			w.sp2[b] = hb; b--;
		}
		w.sp2[b] = h;
	}
	// ---- how many entries will satpos_ hold?
	const uint32_t maxelt = (uint32_t)P.mxIter;
	uint32_t added = 0, nent = 0;
	for(uint32_t j = 0; j < nsmall && added < maxelt; j++) { added += (uint32_t)hits[w.sav[w.sp2[j]]].len; nent++; }
	bool sample = !(added >= maxelt || nsmall == nsp);
	uint32_t nsampled = 0;
	if(sample) {
		const uint64_t lim = nelt < (uint64_t)maxelt ? nelt : (uint64_t)maxelt;
		nsampled = lim > added ? (uint32_t)(lim - added) : 0;
		if(nsampled == 0) sample = false;
	}
	const uint32_t tot = nent + nsampled;
	uint32_t base = 0;
	if(tot > 0) {
		base = atomic_bump(pools.sat_top, tot);
		if(base + tot > pools.sat_cap) { set_error(pools, 2); rs.sat_off = 0; rs.sat_n = 0; rs.neltLeft = 0; return false; }
	}
	rs.sat_off = base; rs.sat_n = nent; // stage 2 appends the sampled rows
	for(uint32_t j = 0; j < nent; j++) {
		const uint16_t h = w.sp2[j];
		const SeedHit& sv = hits[w.sav[h]];
		const bool fw = h < M; const uint32_t offidx = fw ? h : h - M;
		SatEntry& e = pools.sats[base + j];
		e.topf = sv.top; e.size = (uint32_t)sv.len; e.rdoff = (uint16_t)(g.per * offidx + g.off); e.fw = fw ? 1 : 0;
		e.cur = 0; e.perm = 0; e.pad = 0;
	}
	w.nelt_added = added;
	rs.neltLeft = added;
	return sample;
}

// ---- Stage 2: row sampling over the large ranges (reads that fall into repeats).
// Sparse Random1toN: the swap list of a range is the identity (or, once "converted", the sorted complement of the values
// seen so far) plus an overlay of the swapped positions.  A read makes at most BT2GPU_MAX_ITER draws over all its ranges, a
// range may have 65 535 rows: the overlay and the "seen" sets live in ONE open-addressing hash table per read (global memory,
// cleared by the host before the launch), so that a draw costs a handful of probes whatever came before it.
#define BT2_SAMPLE_HASH 4096u             // >= 3 * (BT2GPU_MAX_ITER + 8): two overlay entries and one seen entry per draw at most
#define BT2_SAMPLE_EMPTY 0xffffffffu
struct SampleWork {
	uint32_t* hkey; uint16_t* hval;                                          // [BT2_SAMPLE_HASH]: key = range << 16 | position (overlay), bit 31 | range << 16 | value (seen)
	uint16_t* seen_rng; uint16_t* seen_val; uint32_t nseen;                  // [MAX_ITER] unsorted, per range in draw order (read back when a range converts)
	uint16_t* conv_val; uint32_t nconv;                                      // [MAX_ITER] sorted seen values of converted ranges
	// per large range (index = position in sp2 minus nsmall), [W]
	uint16_t* r_n; uint16_t* r_cur; uint16_t* r_thresh; uint8_t* r_flags;    // flags: 1 inited, 2 swaplist, 4 converted
	uint32_t* r_conv_off; uint16_t* r_conv_n; uint16_t* r_nseen;
	double*   mass; uint8_t* elim;                                           // RowSampler
};

BT2_HD uint32_t sample_slot(uint32_t key) { return (key * 2654435761u) >> 20; }   // 12 bits
// slot holding `key`, or the empty slot where it belongs
BT2_HD uint32_t sample_find(const SampleWork& s, uint32_t key) {
	uint32_t h = sample_slot(key);
	while(s.hkey[h] != BT2_SAMPLE_EMPTY && s.hkey[h] != key) h = (h + 1u) & (BT2_SAMPLE_HASH - 1u);
	return h;
}

BT2_HD uint16_t r1n_base(const SampleWork& s, uint32_t ri, uint16_t j) {
	if(!(s.r_flags[ri] & 4)) return j;
	// j-th smallest value not among the sorted converted-seen values of this range: j + #{k : cv[k] - k <= j}
	const uint16_t* cv = s.conv_val + s.r_conv_off[ri];
	uint32_t lo = 0, hi = s.r_conv_n[ri];
	while(lo < hi) { const uint32_t mid = (lo + hi) >> 1; if((uint32_t)cv[mid] - mid <= (uint32_t)j) lo = mid + 1; else hi = mid; }
	return (uint16_t)(j + lo);
}
BT2_HD uint16_t r1n_get(const SampleWork& s, uint32_t ri, uint16_t pos) {
	const uint32_t key = (ri << 16) | pos, h = sample_find(s, key);
	return s.hkey[h] == key ? s.hval[h] : r1n_base(s, ri, pos);
}
BT2_HD void r1n_set(SampleWork& s, uint32_t ri, uint16_t pos, uint16_t val) {
	const uint32_t key = (ri << 16) | pos, h = sample_find(s, key);
	s.hkey[h] = key; s.hval[h] = val;
}

BT2_HD uint16_t r1n_next(SampleWork& s, uint32_t ri, Rng& rnd) {
	uint16_t& n = s.r_n[ri]; uint16_t& cur = s.r_cur[ri];
	if(cur == 0 && !(s.r_flags[ri] & 4)) {
		if(n == 1) { cur = 1; return 0; }
	}
	if(s.r_flags[ri] & 2) { // swap list
		const uint16_t r = (uint16_t)(cur + (rnd.nextU32() % (uint32_t)(n - cur)));
		if(r != cur) {
			const uint16_t a = r1n_get(s, ri, cur), b = r1n_get(s, ri, r);
			r1n_set(s, ri, cur, b); r1n_set(s, ri, r, a);
		}
		const uint16_t ret = r1n_get(s, ri, cur);
		cur++;
		return ret;
	}
	// "seen" list mode
	uint16_t rn = 0; uint32_t hs = 0;
	while(true) {
		rn = (uint16_t)(rnd.nextU32() % (uint32_t)n);
		const uint32_t key = 0x80000000u | (ri << 16) | rn;
		hs = sample_find(s, key);
		if(s.hkey[hs] != key) { s.hkey[hs] = key; s.hval[hs] = 0; break; }   // not seen before
	}
	s.seen_rng[s.nseen] = (uint16_t)ri; s.seen_val[s.nseen] = rn; s.nseen++;
	s.r_nseen[ri]++;
	cur++;
	if(s.r_nseen[ri] >= s.r_thresh[ri] && cur < n) {
		// convert: sorted seen values become the holes of the new swap list
		const uint32_t off = s.nconv; uint32_t nc = 0;
		for(uint32_t k = 0; k < s.nseen; k++) {
			if(s.seen_rng[k] != ri) continue;
			const uint16_t v = s.seen_val[k];
			uint32_t p = nc;
			while(p > 0 && s.conv_val[off + p - 1] > v) { s.conv_val[off + p] = s.conv_val[off + p - 1]; p--; }
			s.conv_val[off + p] = v; nc++;
			s.seen_rng[k] = 0xffff; // consumed
		}
		s.nconv += nc;
		s.r_conv_off[ri] = off; s.r_conv_n[ri] = (uint16_t)nc;
		s.r_nseen[ri] = 0;
		n = (uint16_t)(n - cur);
		cur = 0;
		s.r_flags[ri] |= (4 | 2);
	}
	return rn;
}

BT2_HD void prio_stage2(const DevParams& P, const Pools& pools, ReadState& rs, const SeedHit* hits,
                        const SeedGeom& g, PrioWork& w, SampleWork& s) {
	const uint32_t M = g.maxOffs, nsmall = w.nsmall, nl = w.nsp - w.nsmall;
	const uint32_t maxelt = (uint32_t)P.mxIter;
	Rng rnd = rs.rnd;
	s.nseen = s.nconv = 0;
	double mass = 0.0f;
	for(uint32_t i = 0; i < nl; i++) {
		const SeedHit& sv = hits[w.sav[w.sp2[nsmall + i]]];
		double num = (double)((uint64_t)sv.nlex + 1); num *= num;
		double denom = (double)sv.len; denom *= denom;
		s.mass[i] = num / denom; s.elim[i] = 0;
		mass += s.mass[i];
		s.r_flags[i] = 0; s.r_n[i] = 0; s.r_cur[i] = 0; s.r_nseen[i] = 0; s.r_conv_n[i] = 0; s.r_conv_off[i] = 0;
	}
	uint32_t added = w.nelt_added, nent = rs.sat_n;
	while(added < maxelt && (uint64_t)added < w.nelt) {
		// RowSampler::next
		const double rd = rnd.nextFloat() * mass;
		double sofar = 0.0f;
		uint32_t pick = 0xffffffffu, lastok = 0xffffffffu;
		for(uint32_t i = 0; i < nl; i++) {
			if(!s.elim[i]) { lastok = i; sofar += s.mass[i]; if(rd < sofar) { pick = i; break; } }
		}
		if(pick == 0xffffffffu) pick = lastok;
		const uint32_t ri = pick;
		const uint16_t h = w.sp2[nsmall + ri];
		const SeedHit& sv = hits[w.sav[h]];
		if(!(s.r_flags[ri] & 1) || s.r_n[ri] == 0) { // Random1toN::init((uint16_t)size, false)
			const uint16_t nn = (uint16_t)sv.len;
			s.r_n[ri] = nn; s.r_cur[ri] = 0; s.r_nseen[ri] = 0;
			s.r_flags[ri] = (uint8_t)(1 | (nn < 128 ? 2 : 0));
			uint16_t th = (uint16_t)(0.10f * nn); if(th < 16) th = 16;
			s.r_thresh[ri] = th;
		}
		const uint16_t r = r1n_next(s, ri, rnd);
		if(s.r_n[ri] > 0 && s.r_cur[ri] >= s.r_n[ri]) { s.elim[ri] = 1; mass -= s.mass[ri]; }
		const bool fw = h < M; const uint32_t offidx = fw ? h : h - M;
		SatEntry& e = pools.sats[rs.sat_off + nent];
		e.topf = sv.top + r; e.size = 1; e.rdoff = (uint16_t)(g.per * offidx + g.off); e.fw = fw ? 1 : 0;
		e.cur = 0; e.perm = 0; e.pad = 0;
		nent++; added++;
	}
	rs.rnd = rnd;
	rs.sat_n = nent;
	rs.neltLeft = added;
}

} // namespace bt2
