// TEST INFRASTRUCTURE ONLY -- steps the per-read state machines of the CUDA library
// (dev_prio/dev_step/dev_bt/dev_gather .cuh, compiled here as plain C++) on the CPU so that
// their control flow can be checked against the oracle without a GPU.  The two
// warp-cooperative kernels (seed search, DP fill) are stood in for by the oracle's
// searchSeed/extendLeft and a scalar fill that writes the same skewed layout; those
// kernels themselves are only tested on the GPU (tests/test_gpu_parity.py).
//   emu_pipeline <index_base> <reads.fq> "<options>" <mode> > out.sam
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <algorithm>
#include "bt2host.h"
#include "oracle_extend.hpp"
#include "../../omp-bowtie2-prime_b200/csrc/gpu/dev_prio.cuh"
#include "../../omp-bowtie2-prime_b200/csrc/gpu/dev_step.cuh"
#include "../../omp-bowtie2-prime_b200/csrc/gpu/dev_bt.cuh"
#include "../../omp-bowtie2-prime_b200/csrc/gpu/dev_gather.cuh"

using namespace bt2;

#include "fill_scalar.hpp"

#ifdef BT2_BT_PROFILE
static void print_bt_prof() {
	static const char* nm[32] = {"hot calls", "cand loop ends", "cand start visited", "attempts", "attempts ending on a visited cell", "cells of those (row+col distance)", "walks reaching an end cell", "",
		"read-gap ext", "read-gap open", "ref-gap ext", "ref-gap open", "diagonal", "", "", "", "problems", "candidates"};
	for(int i = 0; i < 18; i++) if(nm[i][0]) fprintf(stderr, "[bt prof] %-36s %llu\n", nm[i], bt2::g_bt_prof[i]);
}
#endif

int main(int argc, char** argv) {
#ifdef BT2_BT_PROFILE
	atexit(print_bt_prof);
#endif
	if(argc < 5) { fprintf(stderr, "usage\n"); return 2; }
	char err[512] = {0};
	bt2host_job* job = bt2host_job_create(argv[1], argv[3], err, sizeof(err));
	if(!job) { fprintf(stderr, "error: %s\n", err); return 1; }
	bt2host_batch* b = bt2host_batch_from_fastq(job, argv[2], 0, 0, err, sizeof(err));
	if(!b) { fprintf(stderr, "error: %s\n", err); return 1; }
	const bt2gpu_index* hix = bt2host_job_index(job); const bt2gpu_ref* hrf = bt2host_job_ref(job);
	const bt2gpu_params* hp = bt2host_job_params(job); const bt2gpu_reads* in = bt2host_batch_reads(b);
	DevParams P; memset(&P, 0, sizeof(P));
	P.seedLen = hp->seedLen; P.seedMms = hp->seedMms; P.nSeedRounds = hp->nSeedRounds; P.seedBoostThresh = hp->seedBoostThresh;
	P.maxhalf = hp->maxhalf; P.mxIter = hp->mxIter; P.mxUg = hp->mxUg; P.mxDp = hp->mxDp; P.streak = hp->streak; P.doExtend = hp->doExtend;
	P.enable8 = hp->enable8; P.tighten = hp->tighten; P.khits = hp->khits; P.mhits = hp->mhits; P.msample = hp->msample; P.nofw = hp->nofw; P.norc = hp->norc;
	P.gapbar = hp->gapbar; P.rdGapOpen = hp->rdGapConst + hp->rdGapLinear; P.rdGapExt = hp->rdGapLinear;
	P.rfGapOpen = hp->rfGapConst + hp->rfGapLinear; P.rfGapExt = hp->rfGapLinear; P.matchBonus = hp->matchBonus; P.i16 = hp->i16;
	for(int i = 0; i < 256; i++) { P.mmpens[i] = (uint8_t)hp->mmpens[i]; P.npens[i] = (uint8_t)hp->npens[i]; }
	DevIndex ix; ix.sa32 = nullptr; ix.ktab = nullptr; ix.ktabChars = 0; ix.ebwt = hix->ebwt; ix.ftab = hix->ftab; ix.eftab = hix->eftab; ix.offs = hix->offs; ix.rstarts = hix->rstarts; ix.plen = hix->plen;
	for(int i = 0; i < 5; i++) ix.fchr[i] = hix->fchr[i];
	ix.len = hix->len; ix.zOff = hix->zOff; ix.nFrag = hix->nFrag; ix.nPat = hix->nPat; ix.offRate = hix->offRate; ix.ftabChars = hix->ftabChars;
	std::vector<uint64_t> tstart(hrf->nrecs), cum(hrf->nrecs), rec0;
	{ uint64_t c = 0, pos = 0;
	  for(uint64_t r = 0; r < hrf->nrecs; r++) { if(hrf->rec_first[r]) { rec0.push_back(r); pos = 0; } tstart[r] = pos; cum[r] = c; pos += hrf->rec_off[r] + hrf->rec_len[r]; c += hrf->rec_len[r]; }
	  rec0.push_back(hrf->nrecs); }
	DevRef rf; rf.buf = hrf->buf; rf.rec_off = hrf->rec_off; rf.rec_len = hrf->rec_len; rf.rec_tstart = tstart.data(); rf.rec_cum = cum.data(); rf.ref_rec0 = rec0.data(); rf.nrecs = hrf->nrecs;
	orc::Fm fm(hix);
	const uint32_t n = in->n;
	// pools
	std::vector<Chunk> chunks(64 * (size_t)n + 1024); std::vector<SatEntry> sats(2048 * (size_t)n + 4096); std::vector<AlnRec> alnp(8 * n + 1024); std::vector<uint32_t> editp(256 * n + 4096);
	uint32_t ctop = 0, stop = 0, atop = 0, etop = 0, perr = 0;
	Pools pools; pools.chunks = chunks.data(); pools.chunk_top = &ctop; pools.chunk_cap = (uint32_t)chunks.size();
	pools.sats = sats.data(); pools.sat_top = &stop; pools.sat_cap = (uint32_t)sats.size();
	pools.alns = alnp.data(); pools.aln_top = &atop; pools.aln_cap = (uint32_t)alnp.size();
	pools.edits = editp.data(); pools.edit_top = &etop; pools.edit_cap = (uint32_t)editp.size(); pools.error = &perr;
	std::vector<ReadState> states(n);
	std::vector<uint32_t> scratch((size_t)n * BT2_SCRATCH_EDITS, 0);
	uint32_t maxOffs = 1;
	for(uint32_t i = 0; i < n; i++) { uint32_t L = in->seq_off[i + 1] - in->seq_off[i]; SeedGeom g = seed_geom(P, L ? L : 1, 0, (uint32_t)in->interval[i], 1); if(g.numOffs > maxOffs) maxOffs = g.numOffs; }
	const uint32_t W = 2 * maxOffs;
	std::vector<SeedHit> hits(W);
	std::vector<uint16_t> wsav(W), wrank(W), wsp2(W); std::vector<uint8_t> wsorted(W), wexrfw(W); std::vector<ExRange16> wexr(W);
	const uint32_t MI = BT2GPU_MAX_ITER + 8;
	std::vector<uint32_t> hkey(BT2_SAMPLE_HASH); std::vector<uint16_t> hval(BT2_SAMPLE_HASH), seen_rng(MI), seen_val(MI), conv_val(MI), r_n(W), r_cur(W), r_thresh(W), r_conv_n(W), r_nseen(W);
	std::vector<uint8_t> r_flags(W), elim(W); std::vector<uint32_t> r_conv_off(W); std::vector<double> mass(W);
	std::vector<uint8_t> cells(fill_cell_bytes(BT2GPU_MAX_ROWS, BT2_KMAX));
	const bool exactFill = getenv("EMU_EXACT_FILL") != nullptr; uint32_t nprob = 0; std::vector<uint32_t> pcache(BT2_PC_WORDS, 0);
	orc::WorkCounters wc;
	for(uint32_t ri = 0; ri < n; ri++) {
		ReadState& rs = states[ri];
		const uint32_t so = in->seq_off[ri], rdlen = in->seq_off[ri + 1] - so;
		read_begin(rs, so, rdlen, in->rng_seed[ri], in->minsc[ri], in->nceil[ri], in->interval[ri]);
		if(rs.phase == RS_FINISHED) continue;
		orc::ReadView orv; orv.init(in->seq + so, in->qual + so, (int)rdlen);
		for(uint32_t round = 0;; round++) {
			uint32_t offset = 0;
			if(!round_wanted(P, rs, round, offset)) break;
			rs.round = (uint8_t)round; rs.roundsRun++;
			const SeedGeom g = seed_geom(P, rdlen, offset, (uint32_t)rs.interval, maxOffs);
			// ---- stand-in for K_seed
			for(uint32_t h = 0; h < W; h++) { hits[h].valid = 0; hits[h].len = 0; }
			for(uint32_t s = 0; s < 2; s++) {
				if((s == 0 && P.nofw) || (s == 1 && P.norc)) continue;
				const std::vector<uint8_t>& pat = s == 0 ? orv.fw : orv.rc;
				for(uint32_t i = 0; i < g.numOffs; i++) {
					const int depth = (int)(i * g.per + g.off), L = (int)g.seqLen;
					const int sofs = s == 0 ? depth : ((int)rdlen - depth - L);
					const uint8_t* seq = pat.data() + sofs;
					bool ok = true; for(int k = 0; k < L; k++) if(seq[k] == 4) ok = false;
					if(!ok) continue;
					uint64_t top = 0, bot = 0;
					if(!orc::searchSeed(fm, seq, L, top, bot, wc)) continue;
					uint32_t nlex = 0;
					if(P.doExtend) { const int lim = s == 0 ? depth : ((int)rdlen - L - depth); if(lim > 0) nlex = orc::extendLeft(fm, pat.data(), lim, top, bot, wc); }
					SeedHit& hh = hits[s * maxOffs + i];
					hh.top = top; hh.len = bot - top; hh.nlex = nlex; hh.valid = 1; hh.key = 0;
					for(int k = 0; k < L; k++) hh.key = (hh.key << 2) | seq[k];
				}
			}
			PrioWork w; w.sav = wsav.data(); w.sorted = wsorted.data(); w.rank = wrank.data(); w.sp2 = wsp2.data(); w.exr = wexr.data(); w.exrfw = wexrfw.data();
			rs.nDpFail = 0; rs.passStarted = 0; rs.cur_i = 0; rs.first = 1; rs.sat_n = 0; rs.neltLeft = 0;
			const bool sample = prio_stage1(P, pools, rs, hits.data(), g, w);
			if(rs.numRanges == 0 || rs.nonzTot == 0) { rs.flags |= BT2GPU_RF_NO_SEED_HITS; break; }
			if(sample) {
				SampleWork s; std::fill(hkey.begin(), hkey.end(), BT2_SAMPLE_EMPTY); s.hkey = hkey.data(); s.hval = hval.data(); s.seen_rng = seen_rng.data(); s.seen_val = seen_val.data();
				s.conv_val = conv_val.data(); s.r_n = r_n.data(); s.r_cur = r_cur.data(); s.r_thresh = r_thresh.data(); s.r_flags = r_flags.data();
				s.r_conv_off = r_conv_off.data(); s.r_conv_n = r_conv_n.data(); s.r_nseen = r_nseen.data(); s.mass = mass.data(); s.elim = elim.data();
				prio_stage2(P, pools, rs, hits.data(), g, w, s);
			}
			// ---- extendSeeds
			while(true) {
				Problem pb; pb.read = ri;
				if(advance(P, ix, rf, pools, rs, pb) == ADV_ROUND_END) break;
				BtCtx cx; cx.P = P; cx.pools = pools; cx.states = states.data(); cx.probs = &pb; cx.cells = cells.data(); cx.cell_stride = 0;
				cx.pcache = pcache.data(); cx.scratch = scratch.data(); cx.seq = in->seq; cx.qual = in->qual; cx.set_pens();
				BtMachine m; m.slot = 0; m.read = ri; m.pc = pcache.data(); m.pcs = 1; uint8_t cbuf[64]; m.cbuf = cbuf; m.cstride = 16;
				fill_emu(P, pb, m.view(cx), cells.data(), pcache.data(), exactFill, (nprob++) & 1u);
				m.begin(cx);
				while(m.phase != BP_DONE && m.step(cx)) { }
				m.finish(cx);
				if(m.result == BT_ROUND_END) break;
			}
			if(!next_round_wanted(P, rs)) break;
		}
	}
	if(perr) { fprintf(stderr, "pool error %u\n", perr); return 1; }
	// gather
	std::vector<bt2gpu_read_result> rres(n); std::vector<bt2gpu_aln> oal(atop + 1); std::vector<bt2gpu_edit> oed;
	{ uint64_t tot = 0; for(uint32_t i = 0; i < n; i++) tot += read_edit_count(states[i], pools); oed.resize(tot + 1); }
	uint32_t ab = 0, eb = 0;
	for(uint32_t i = 0; i < n; i++) {
		memset(&rres[i], 0, sizeof(rres[i]));
		gather_read(states[i], pools, scratch.data() + (size_t)i * BT2_SCRATCH_EDITS, ab, eb, 0, 0, rres[i], oal.data(), oed.data());
		ab += states[i].n_alns; eb += read_edit_count(states[i], pools);
	}
	bt2gpu_results res; res.n = n; res.reads = rres.data(); res.alns = oal.data(); res.n_alns = ab; res.edits = oed.data(); res.n_edits = eb;
	char* h = bt2host_sam_header(job); fputs(h, stdout); bt2host_free(h);
	size_t len = 0; char* s = bt2host_sam_records(job, b, &res, atoi(argv[4]), &len);
	fwrite(s, 1, len, stdout); bt2host_free(s);
	uint64_t dps = 0, bts = 0, bta = 0; for(uint32_t i = 0; i < n; i++) { dps += states[i].nExDps; bts += states[i].btSteps; bta += states[i].btAttempts; }
	fprintf(stderr, "emu: reads %u dps %lu btAttempts %lu btSteps %lu alns %u cells in band %zu of %zu\n", n, dps, bta, bts, atop, g_cells_in_band, g_cells_rect);
	return 0;
}
