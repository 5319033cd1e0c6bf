// TEST INFRASTRUCTURE ONLY -- C entry points over the oracle restatement, with
// the same result layout as bt2gpu_align_batch() so that tests can compare the
// CUDA path with the oracle struct-for-struct, and bench.py can time it as the
// "port" CPU baseline.  Never linked or loaded by the product library.
#include <cstdio>
#include <cstring>
#include <vector>
#include <omp.h>
#include "oracle_extend.hpp"

namespace {
struct Store {
	std::vector<bt2gpu_read_result> reads;
	std::vector<bt2gpu_aln> alns;
	std::vector<bt2gpu_edit> edits;
	orc::WorkCounters wc;
};
void pack_edit(const orc::Edit& e, bt2gpu_edit& o) { o.pos = e.pos; o.chr = e.chr; o.qchr = e.qchr; o.type = e.type; o.pad = 0; }
}

extern "C" {

struct oracle_ctx {
	const bt2gpu_index* ix; const bt2gpu_ref* rf; bt2gpu_params P;
	orc::Fm* fm; orc::Ref* ref; Store st;
};

oracle_ctx* oracle_create(const bt2gpu_index* ix, const bt2gpu_ref* rf, const bt2gpu_params* p) {
	oracle_ctx* c = new oracle_ctx();
	c->ix = ix; c->rf = rf; c->P = *p;
	c->fm = new orc::Fm(ix); c->ref = new orc::Ref(rf);
	return c;
}
void oracle_destroy(oracle_ctx* c) { if(!c) return; delete c->fm; delete c->ref; delete c; }

// nthreads <= 0: all OpenMP threads. Reads are independent, so this is the same
// data-parallel split the reference makes (bt2_search.cpp:2723).
int oracle_align_batch(oracle_ctx* c, const bt2gpu_reads* in, bt2gpu_results* out, int nthreads) {
	const uint32_t n = in->n;
	std::vector<orc::ReadOutcome> oc(n);
	if(nthreads <= 0) nthreads = omp_get_max_threads();
	std::vector<orc::WorkCounters> wcs(nthreads);
	#pragma omp parallel for schedule(dynamic, 64) num_threads(nthreads)
	for(uint32_t i = 0; i < n; i++) {
		const uint32_t o = in->seq_off[i];
		const int len = (int)(in->seq_off[i + 1] - o);
		orc::alignRead(*c->fm, *c->ref, c->P, in->seq + o, in->qual + o, len, in->rng_seed[i],
		               in->minsc[i], in->nceil[i], in->interval[i], oc[i], wcs[omp_get_thread_num()]);
	}
	orc::WorkCounters& wc = c->st.wc;
	for(int t = 0; t < nthreads; t++) {
		wc.seeds += wcs[t].seeds; wc.lfSearch += wcs[t].lfSearch; wc.lfExtend += wcs[t].lfExtend; wc.saRows += wcs[t].saRows;
		wc.lfResolve += wcs[t].lfResolve; wc.dps += wcs[t].dps; wc.cells += wcs[t].cells; wc.btAttempts += wcs[t].btAttempts;
		wc.btSteps += wcs[t].btSteps; wc.alns += wcs[t].alns;
	}
	Store& st = c->st;
	st.reads.assign(n, bt2gpu_read_result());
	st.alns.clear(); st.edits.clear();
	for(uint32_t i = 0; i < n; i++) {
		const orc::ReadOutcome& r = oc[i];
		bt2gpu_read_result& rr = st.reads[i];
		memset(&rr, 0, sizeof(rr));
		rr.aln_off = (uint32_t)st.alns.size(); rr.n_alns = (uint32_t)r.alns.size();
		rr.rng_last = r.rnd.last; rr.rng_lastOff = r.rnd.lastOff; rr.minsc = (int32_t)r.minsc;
		rr.status = (uint8_t)r.status; rr.flags = (uint8_t)r.flags; rr.rounds = (uint8_t)r.rounds; rr.done_policy = r.donePolicy;
		rr.nExIters = r.prm.nExIters; rr.nExDps = r.prm.nExDps; rr.nExDpSuccs = r.prm.nExDpSuccs; rr.nExDpFails = r.prm.nExDpFails;
		rr.nDpFail = r.prm.nDpFail; rr.nDpFailStreak = r.prm.nDpFailStreak; rr.nDpLastSucc = r.prm.nDpLastSucc; rr.nRedundants = r.prm.nRedundants;
		rr.bestLtMinsc = r.prm.bestLtMinsc == LLONG_MIN ? INT32_MIN : (int32_t)r.prm.bestLtMinsc;
		rr.nonzTot = (uint32_t)r.lastSeeds.nonzTot; rr.numElts = (uint32_t)r.lastSeeds.numElts;
		rr.numRanges = (uint32_t)r.lastSeeds.numRanges; rr.nSeedsTried = r.seedsTried;
		for(size_t k = 0; k < r.alns.size(); k++) {
			const orc::AlnOut& a = r.alns[k];
			bt2gpu_aln o; memset(&o, 0, sizeof(o));
			o.refoff = a.refoff; o.tidx = (uint32_t)a.tidx; o.score = (int32_t)a.score; o.minsc = (int32_t)r.alnMinsc[k];
			o.ns = a.ns; o.gaps = a.gaps; o.edits = a.edits; o.basesAligned = a.basesAligned; o.refns = a.refns;
			o.edit_off = (uint32_t)st.edits.size(); o.n_edits = (uint16_t)a.ned.size(); o.fw = a.fw;
			for(size_t e = 0; e < a.ned.size(); e++) { bt2gpu_edit pe; pack_edit(a.ned[e], pe); st.edits.push_back(pe); }
			st.alns.push_back(o);
		}
		size_t nscr = 0; // the image is cut at the largest edit count any alignment reads it through
		for(size_t k = 0; k < r.alns.size(); k++) nscr = std::max(nscr, r.alns[k].ned.size());
		nscr = std::min(nscr, r.scratch.hw);
		rr.scratch_off = (uint32_t)st.edits.size(); rr.n_scratch = (uint32_t)nscr;
		for(size_t e = 0; e < nscr; e++) { bt2gpu_edit pe; pack_edit(r.scratch.buf[e], pe); st.edits.push_back(pe); }
	}
	out->n = n; out->reads = st.reads.data(); out->alns = st.alns.data(); out->n_alns = st.alns.size();
	out->edits = st.edits.data(); out->n_edits = st.edits.size();
	return 0;
}

// [seeds, lfSearch, lfExtend, saRows, lfResolve, dps, cells, btAttempts, btSteps, alns]
void oracle_counters(oracle_ctx* c, uint64_t out[10]) {
	const orc::WorkCounters& w = c->st.wc;
	out[0] = w.seeds; out[1] = w.lfSearch; out[2] = w.lfExtend; out[3] = w.saRows; out[4] = w.lfResolve;
	out[5] = w.dps; out[6] = w.cells; out[7] = w.btAttempts; out[8] = w.btSteps; out[9] = w.alns;
}

} // extern "C"

// Stage oracle for the DP fill: one SwAligner problem (strand buffer rd/qu already oriented),
// returns the raw u8 H/E/F matrices [rows*cols] and the sorted candidate list (rawH<<8|col).
extern "C" int oracle_fill(const bt2gpu_params* P, const uint8_t* rd, const uint8_t* qu, int rows,
                           const uint8_t* rfmasks, int cols, int minsc, uint8_t* H, uint8_t* E, uint8_t* F,
                           uint16_t* cand, int* ncand, int* lrmax) {
	orc::SwProblem sw; sw.P = P; sw.sc.P = P; sw.i16 = false; sw.fw = true; sw.rd = rd; sw.qu = qu; sw.rdlen = rows; sw.nceil = rows;
	sw.rf.assign(rfmasks, rfmasks + cols + 1); sw.rflen = cols; sw.minsc = minsc;
	int64_t best = 0; orc::WorkCounters wc;
	sw.align(best, wc);
	for(size_t i = 0; i < (size_t)rows * cols; i++) { H[i] = (uint8_t)(sw.H[i] + 255); E[i] = (uint8_t)(sw.E[i] + 255); F[i] = (uint8_t)(sw.F[i] + 255); }
	*ncand = (int)sw.cands.size();
	int mx = 0;
	for(int c = 0; c < cols; c++) mx = std::max(mx, (int)H[(size_t)(rows - 1) * cols + c]);
	*lrmax = mx;
	for(size_t i = 0; i < sw.cands.size(); i++) cand[i] = (uint16_t)((sw.cands[i].score << 8) | sw.cands[i].col);
	return 0;
}
