// SwDriver::extendSeeds (aligner_sw_driver.cpp:278-631) as a resumable per-read state
// machine: advance() runs the loop from where the read left off up to the point where a
// dynamic-programming problem is needed (the fill runs in its own warp-per-problem kernel),
// or to the end of the seed round.  Also the round logic of multiseedSearchWorker
// (bt2_search.cpp:2571-2586, 2784-2799, 2819-2826).
#pragma once
#include "dev_common.cuh"

namespace bt2 {

enum { ADV_NEED_DP = 0, ADV_ROUND_END = 1 };

// Random1toN for n <= 5 (always a swap list): state lives in the SatEntry
BT2_HD uint32_t sat_next(SatEntry& e, Rng& rnd) {
	const uint32_t n = e.size;
	if(e.cur == 0) {
		if(n == 1) { e.cur = 1; return 0; }
		e.perm = 0u | (1u << 3) | (2u << 6) | (3u << 9) | (4u << 12);
	}
	const uint32_t cur = e.cur;
	const uint32_t r = cur + (rnd.nextU32() % (n - cur));
	uint32_t a = (e.perm >> (3 * cur)) & 7u, b = (e.perm >> (3 * r)) & 7u;
	if(r != cur) {
		e.perm &= ~((7u << (3 * cur)) | (7u << (3 * r)));
		e.perm |= (b << (3 * cur)) | (a << (3 * r));
		a = b;
	}
	e.cur = (uint8_t)(cur + 1);
	return a;
}

BT2_HD bool seen_present(const Pools& p, uint32_t head, uint32_t ref, bool fw, int64_t off) {
	uint32_t c = head;
	while(c != BT2_NIL) {
		const Chunk& ch = p.chunks[c];
		for(uint32_t k = 0; k < ch.n; k++) {
			const int64_t ioff = (int64_t)ch.item[k][0];
			const uint32_t iref = (uint32_t)ch.item[k][1];
			const int32_t lf = (int32_t)(ch.item[k][1] >> 32);
			if(iref == ref && ((lf & 1) != 0) == fw && off >= ioff && off < ioff + (int64_t)(lf >> 1)) return true;
		}
		c = ch.next;
	}
	return false;
}
BT2_HD void seen_add(const Pools& p, uint32_t& head, uint32_t ref, bool fw, int64_t off, int64_t len) {
	const uint32_t lf = (uint32_t)(((int32_t)len << 1) | (fw ? 1 : 0));
	list_push(p, head, (uint64_t)off, (uint64_t)ref | ((uint64_t)lf << 32));
}

// Returns ADV_NEED_DP with `prob` filled in, or ADV_ROUND_END with rs.status = EXTEND_* code.
BT2_HD int advance(const DevParams& P, const DevIndex& ix, const DevRef& rf, const Pools& pools,
                   ReadState& rs, Problem& prob) {
	const uint32_t nsm = 5;
	const uint32_t rdlen = rs.rdlen;
	const int32_t perfect = (int32_t)rdlen * P.matchBonus;
	const uint32_t gws = rs.sat_n;
	const uint32_t seedlen = (uint32_t)P.seedLen < rdlen ? (uint32_t)P.seedLen : rdlen;
	Rng rnd = rs.rnd;
	int ret = ADV_ROUND_END;
	while(true) {
		if(!rs.passStarted) {
			if(rs.neltLeft == 0) { rs.status = BT2GPU_ST_EXHAUSTED; break; }
			if(rs.minsc == perfect) { rs.status = BT2GPU_ST_PERFECT_SCORE; break; }
			rs.cur_i = 0; rs.first = 1; rs.passStarted = 1;
		}
		if(rs.cur_i >= gws) { rs.passStarted = 0; continue; }
		SatEntry& e = pools.sats[rs.sat_off + rs.cur_i];
		const bool is_small = e.size < nsm;
		if(e.cur >= e.size || !(rs.first || is_small)) { rs.cur_i++; rs.first = 1; continue; }
		if(rs.minsc == perfect) { rs.status = BT2GPU_ST_PERFECT_SCORE; break; }
		if(rs.nExDps >= (uint32_t)P.mxDp) { rs.status = BT2GPU_ST_HARD_LIMIT; break; }
		if(0 >= P.mxUg) { rs.status = BT2GPU_ST_HARD_LIMIT; break; }
		if(rs.nExIters >= (uint32_t)P.mxIter) { rs.status = BT2GPU_ST_HARD_LIMIT; break; }
		rs.nExIters++;
		rs.first = 0;
		const uint32_t elt = sat_next(e, rnd);
		const bool fw = e.fw != 0;
		rs.saRows++;
		const uint64_t joff = sa_offset(ix, e.topf + elt, &rs.lfResolve);
		rs.neltLeft--;
		uint64_t tidx = 0, toff = 0, tlen = 0;
		joined_to_text(ix, joff, tidx, toff, tlen);
		const uint32_t rdoff = fw ? e.rdoff : (rdlen - e.rdoff - seedlen);
		const int64_t refoff = (int64_t)toff - (int64_t)rdoff;
		if(seen_present(pools, rs.seen_head, (uint32_t)tidx, fw, refoff)) { rs.nRedundants++; continue; }
		const int readGaps = max_read_gaps(P, rs.minsc, rdlen), refGaps = max_ref_gaps(P, rs.minsc, rdlen);
		// DynProgFramer::frameSeedExtensionRect, trimToRef
		int64_t maxgap = readGaps > refGaps ? readGaps : refGaps;
		if(maxgap > P.maxhalf) maxgap = P.maxhalf;
		const int64_t refl0 = refoff - 2 * maxgap, refr0 = refoff + ((int64_t)rdlen - 1) + 2 * maxgap;
		int64_t triml = 0, trimr = 0;
		if(refr0 >= (int64_t)tlen) trimr = refr0 - ((int64_t)tlen - 1);
		if(refl0 < 0) triml = -refl0;
		const int64_t refl = refl0 + triml, refr = refr0 - trimr;
		const bool found = !(refr < refl);
		seen_add(pools, rs.seen_head, (uint32_t)tidx, fw, refoff, 1);
		if(!found) continue;
		// SwAligner::initRef
		const int64_t rfi = refl, rff = refr + 2;
		const int64_t rflenFull = rff - rfi;
		if(rflenFull > BT2GPU_MAX_COLS) {
			rs.nExDpFails++; rs.nDpFail++;
			rs.status = BT2GPU_ST_HARD_LIMIT; break;
		}
		int64_t leftNs = rfi >= 0 ? 0 : -rfi; if(leftNs > rflenFull) leftNs = rflenFull;
		int64_t rightNs = rff <= (int64_t)tlen ? 0 : rff - (int64_t)tlen; if(rightNs > rflenFull) rightNs = rflenFull;
		const int64_t inner = rflenFull - (leftNs + rightNs);
		for(int64_t k = 0; k < leftNs; k++) prob.rf[k] = 16;
		if(inner > 0) ref_masks(rf, prob.rf + leftNs, tidx, rfi < 0 ? 0 : (uint64_t)rfi, (uint32_t)inner);
		for(int64_t k = leftNs + (inner > 0 ? inner : 0); k < rflenFull; k++) prob.rf[k] = 16;
		prob.rows = (uint16_t)rdlen; prob.cols = (uint16_t)(rflenFull - 1); prob.seq_off = rs.seq_off;
		prob.fw = fw ? 1 : 0;
		prob.d0 = 0;   // set by the fill (k_fill.cuh)
		// default build: 8-bit fill, minsc clamped (bt2_search.cpp:2487-2491, aligner_sw.cpp:84).  -DENABLE_I16 build: --no-sse8 or
		// minsc < -254 selects the 16-bit fill for THIS problem (aligner_sw.cpp:77-83; minsc is tightened as alignments are found, so a
		// read may change paths between two of its problems)
		const bool i16 = P.i16 && (!P.enable8 || rs.minsc < -254);
		if(i16 && rs.minsc < -(BT2_I16_START - 0x100)) set_error(pools, 16);   // beyond the range the fill keeps (never with reads <= 160 bp and sane penalties)
		prob.lay = i16 ? 0x40 : 0;
		prob.minsc = (!i16 && rs.minsc < -254) ? -254 : rs.minsc;
		prob.tidx = (uint32_t)tidx; prob.refl = refl;
		prob.triml = (int32_t)triml; prob.corel = (int32_t)maxgap; prob.corer = (int32_t)(3 * maxgap);
		prob.lrmax = 0; prob.ncand = 0;
		seen_add(pools, rs.seen_head, (uint32_t)tidx, fw, refl0 + maxgap, 2 * maxgap + 1);
		ret = ADV_NEED_DP;
		break;
	}
	rs.rnd = rnd;
	return ret;
}

// Start of a read (bt2_search.cpp:2485-2560): everything the host loop sets when a slot gets a new read.
BT2_HD void read_begin(ReadState& rs, uint32_t seq_off, uint32_t rdlen, uint32_t seed, int32_t minsc, int32_t nceil, int32_t interval) {
	rs.rnd.init(seed);
	rs.minsc = minsc; rs.nceil = nceil; rs.interval = interval; rs.seq_off = seq_off; rs.rdlen = (uint16_t)rdlen;
	rs.phase = RS_IDLE; rs.round = 0; rs.status = 0; rs.flags = 0; rs.roundsRun = 0; rs.donePolicy = 0;
	rs.nunpair = 0; rs.best = BT2_I32_MIN; rs.best2 = BT2_I32_MIN; rs.doneUnpair = 0;
	rs.nExIters = rs.nExDps = rs.nExDpSuccs = rs.nExDpFails = rs.nDpFail = rs.nDpFailStreak = rs.nDpLastSucc = rs.nRedundants = 0;
	rs.bestLtMinsc = BT2_I32_MIN;
	rs.nonzTot = rs.numRanges = rs.nSeedsTried = 0; rs.numElts = 0;
	rs.sat_off = rs.sat_n = 0; rs.cur_i = 0; rs.neltLeft = 0; rs.first = 1; rs.passStarted = 0;
	rs.seen_head = BT2_NIL; rs.exfw_head = BT2_NIL; rs.exrc_head = BT2_NIL; rs.aln_tail = BT2_NIL; rs.n_alns = 0;
	rs.scr_cur = rs.scr_hw = 0;
	rs.lfResolve = rs.saRows = rs.btAttempts = rs.btSteps = 0; rs.dpCells = 0;
	if(rdlen == 0) { rs.flags |= BT2GPU_RF_EMPTY; rs.phase = RS_FINISHED; }
	else if(rdlen > BT2GPU_MAX_ROWS) { rs.flags |= BT2GPU_RF_TOO_LONG; rs.phase = RS_FINISHED; }
}

// Should this read run seed round `round`?  (round loop of oracle alignRead / bt2_search.cpp:2571-2586)
BT2_HD bool round_wanted(const DevParams& P, const ReadState& rs, uint32_t round, uint32_t& offset) {
	const uint32_t interval = (uint32_t)rs.interval;
	const uint32_t nrounds = (uint32_t)P.nSeedRounds < interval ? (uint32_t)P.nSeedRounds : interval;
	if(round >= nrounds) return false;
	offset = (interval * round) / nrounds;
	if(offset > 0 && (uint32_t)P.seedLen + offset > rs.rdlen) return false;
	return true;
}

// After extendSeeds returned rs.status for the round: continue with another seed round?
BT2_HD bool next_round_wanted(const DevParams& P, const ReadState& rs) {
	const int st = rs.status;
	if(!(st == BT2GPU_ST_EXHAUSTED || st == BT2GPU_ST_SOFT_LIMIT || st == BT2GPU_ST_POLICY_FULFILLED)) return false;
	const uint64_t avg = rs.nonzTot == 0 ? ~0ull : rs.numElts / rs.nonzTot;
	if(avg < (uint64_t)P.seedBoostThresh) return false;
	if(rs.doneUnpair) return false;
	uint32_t off;
	return round_wanted(P, rs, (uint32_t)rs.round + 1, off);
}

} // namespace bt2
