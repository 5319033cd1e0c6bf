// (aln_glob, edit_glob): where this batch's arrays start in the caller-visible result arrays (pipelined sub-batches).
// Packs the per-read device state into the C-ABI result arrays (include/bt2gpu.h):
// what AlnSinkWrap::report received, in call order, plus what the host restores
// before AlnSinkWrap::finishRead (RNG state, minsc, PerReadMetrics, F14 scratch).
#pragma once
#include "dev_common.cuh"

namespace bt2 {

BT2_HD void unpack_edit(uint32_t e, bt2gpu_edit& o) {
	o.pos = edit_pos(e); o.chr = (uint8_t)edit_chr(e); o.qchr = (uint8_t)edit_qchr(e); o.type = (uint8_t)edit_type(e); o.pad = 0;
}

// The host only ever reads the scratch through an alignment's own edit count (the AlnRes copies keep
// their count by value and alias the buffer), so the image is cut at the largest n_edits.
BT2_HD uint32_t scratch_count(const ReadState& rs, const Pools& pools) {
	uint32_t m = 0;
	for(uint32_t p = rs.aln_tail; p != BT2_NIL; p = pools.alns[p].prev) if(pools.alns[p].n_edits > m) m = pools.alns[p].n_edits;
	return m < rs.scr_hw ? m : rs.scr_hw;
}
// # of bt2gpu_edit entries this read contributes (its alignments' edits + the scratch image)
BT2_HD uint32_t read_edit_count(const ReadState& rs, const Pools& pools) {
	uint32_t n = scratch_count(rs, pools);
	for(uint32_t p = rs.aln_tail; p != BT2_NIL; p = pools.alns[p].prev) n += pools.alns[p].n_edits;
	return n;
}

BT2_HD void gather_read(const ReadState& rs, const Pools& pools, const uint32_t* scratch,
                        uint32_t aln_base, uint32_t edit_base, uint32_t aln_glob, uint32_t edit_glob,
                        bt2gpu_read_result& rr, bt2gpu_aln* alns, bt2gpu_edit* edits) {
	rr.aln_off = aln_base + aln_glob; rr.n_alns = rs.n_alns;
	rr.rng_last = rs.rnd.last; rr.rng_lastOff = rs.rnd.lastOff;
	rr.minsc = rs.minsc; rr.status = rs.status; rr.flags = rs.flags; rr.rounds = rs.roundsRun; rr.done_policy = rs.donePolicy;
	rr.nExIters = rs.nExIters; rr.nExDps = rs.nExDps; rr.nExDpSuccs = rs.nExDpSuccs; rr.nExDpFails = rs.nExDpFails;
	rr.nDpFail = rs.nDpFail; rr.nDpFailStreak = rs.nDpFailStreak; rr.nDpLastSucc = rs.nDpLastSucc; rr.nRedundants = rs.nRedundants;
	rr.bestLtMinsc = rs.bestLtMinsc;
	rr.nonzTot = rs.nonzTot; rr.numElts = (uint32_t)rs.numElts; rr.numRanges = rs.numRanges; rr.nSeedsTried = rs.nSeedsTried;
	// alignments: the list is linked newest-first; emit in report order
	uint32_t k = rs.n_alns;
	uint32_t eo = edit_base;
	// edits are laid out in report order too: first pass computes each alignment's offset
	uint32_t tot = 0;
	for(uint32_t p = rs.aln_tail; p != BT2_NIL; p = pools.alns[p].prev) tot += pools.alns[p].n_edits;
	uint32_t end = edit_base + tot;
	for(uint32_t p = rs.aln_tail; p != BT2_NIL; p = pools.alns[p].prev) {
		const AlnRec& a = pools.alns[p];
		k--;
		end -= a.n_edits;
		bt2gpu_aln& o = alns[aln_base + k];
		o.refoff = a.refoff; o.tidx = a.tidx; o.score = a.score; o.minsc = a.minsc; o.ns = a.ns; o.gaps = a.gaps;
		o.edits = a.edits; o.basesAligned = a.basesAligned; o.refns = a.refns; o.edit_off = end + edit_glob; o.n_edits = a.n_edits;
		o.fw = a.fw; o.pad = 0;
		for(uint32_t i = 0; i < a.n_edits; i++) unpack_edit(pools.edits[a.edit_off + i], edits[end + i]);
	}
	eo = edit_base + tot;
	const uint32_t ns = scratch_count(rs, pools);
	rr.scratch_off = eo + edit_glob; rr.n_scratch = ns;
	rr.nExUgs = 0; rr.exhaustive = 0;   // extendSeeds runs no ungapped extension and never sets `exhaustive` (aligner_sw_driver.cpp:371, 302)
	for(uint32_t i = 0; i < ns; i++) unpack_edit(scratch[i], edits[eo + i]);
}

} // namespace bt2
