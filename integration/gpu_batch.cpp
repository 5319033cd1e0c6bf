// gpu_batch.cpp -- the reference-side binding of libbt2gpu.so (include/bt2gpu.h).
//
// NOT a stand-alone translation unit: bt2_search_gpu.patch makes the reference's bt2_search.cpp #include this file right
// before multiseedSearchWorker() (bt2_search.cpp:2297), inside the reference's unity build, so everything bt2_search.cpp
// sees (its file-static option globals :69-246, msWorkerConsts :2194, msWorkerObjs, AlnSinkWrapOne :1970) is in scope.
// Nothing of the reference is copied here; the patch itself adds three `friend struct Bt2GpuBridge;` lines
// (RandomSource, SwDriver, BitPairReference) and one `if(!gpu) { ... }` around phases P1-P6 of the batch loop.
//
// With BT2GPU=1 in the environment (BT2GPU_DEVICE picks the GPU) the patched bowtie2-align-l runs P0 ("read") and
// P7 ("finishReadOne") unchanged and replaces P1-P6 (bt2_search.cpp:2591-2805: instantiateSeeds ... extendSeeds, all seed rounds of
// a read at once) by ONE bt2gpu_align_batch() per batch iteration, then replays what those phases leave behind in the host
// objects: AlnSinkWrap::report() (aln_sink.cpp:1395) for every alignment in device order through the slot's own
// SwDriver::resGap_ (so that the sink's AlnRes copies alias the same edit buffer as in the reference), the
// per-slot RandomSource state (consumed again by finishRead -> selectByScore, aln_sink.cpp:1477), the tightened minsc and the
// PerReadMetrics counters.
#include "bt2gpu.h"

struct Bt2GpuBridge {
	bt2gpu_ctx* ctx = nullptr;
	bool on = false;
	// SoA staging of one batch (bt2gpu_reads)
	std::vector<uint32_t> seq_off, rng_seed, slot_of;
	std::vector<int32_t> minsc32, nceil32, interval32;
	std::vector<uint8_t> seq, qual;
	double ms_gpu = 0; uint64_t nbatches = 0, nreads = 0;

	static bool wanted() { const char* e = getenv("BT2GPU"); return e && atoi(e) != 0; }

	// once per run: index, reference and every option the path reads, handed over as plain arrays (the library copies them to HBM)
	void init(const msWorkerConsts& k) {
		if(!wanted()) return;
		const Ebwt& e = k.ebwtFw; const BitPairReference& r = k.ref; const Scoring& sc = k.sc;
		bt2gpu_index ix; memset(&ix, 0, sizeof(ix));                          // accessors bt2_idx.h:1252-1271
		ix.len = e.eh().len(); ix.lineRate = e.eh().lineRate(); ix.offRate = e.eh().offRate(); ix.ftabChars = e.eh().ftabChars();
		ix.nPat = e.nPat(); ix.plen = (const uint64_t*)e.plen(); ix.nFrag = e.nFrag(); ix.rstarts = (const uint64_t*)e.rstarts();
		ix.ebwt = e.ebwt(); ix.ebwtTotLen = e.eh().ebwtTotLen(); ix.zOff = e.zOff(); ix.fchr = (const uint64_t*)e.fchr();
		ix.ftab = (const uint64_t*)e.ftab(); ix.ftabLen = e.eh().ftabLen(); ix.eftab = (const uint64_t*)e.eftab(); ix.eftabLen = e.eh().eftabLen();
		ix.offs = (const uint64_t*)e.offs(); ix.offsLen = e.eh().offsLen();
		std::vector<uint64_t> roff(r.recs_.size()), rlen(r.recs_.size()); std::vector<uint8_t> rfirst(r.recs_.size());   // reference.h:106-121
		uint64_t bases = 0;
		for(size_t i = 0; i < r.recs_.size(); i++) { roff[i] = r.recs_[i].off; rlen[i] = r.recs_[i].len; rfirst[i] = r.recs_[i].first ? 1 : 0; bases += r.recs_[i].len; }
		bt2gpu_ref rf; memset(&rf, 0, sizeof(rf));
		rf.nrecs = r.recs_.size(); rf.rec_off = roff.data(); rf.rec_len = rlen.data(); rf.rec_first = rfirst.data(); rf.buf = r.buf_; rf.bufBases = bases;
		bt2gpu_params p; memset(&p, 0, sizeof(p)); p.abi_version = BT2GPU_ABI_VERSION;
		p.seedLen = multiseedLen; p.seedMms = multiseedMms; p.nSeedRounds = (int32_t)nSeedRounds; p.seedBoostThresh = (int32_t)seedBoostThresh;
		p.maxhalf = (int32_t)k.maxHalf; p.mxIter = (int32_t)k.mxIter; p.mxUg = (int32_t)k.mxUg; p.mxDp = (int32_t)k.mxDp; p.streak = (int32_t)k.streak;
		p.doExtend = k.extend ? 1 : 0; p.enable8 = k.doEnable8 ? 1 : 0; p.tighten = k.doTighten;
		p.khits = (int32_t)khits; p.mhits = (int32_t)mhits; p.msample = msample ? 1 : 0; p.nofw = k.nofw ? 1 : 0; p.norc = k.norc ? 1 : 0;
		p.gapbar = sc.gapbar; p.rdGapConst = sc.rdGapConst; p.rdGapLinear = sc.rdGapLinear; p.rfGapConst = sc.rfGapConst; p.rfGapLinear = sc.rfGapLinear;
		p.matchBonus = (int32_t)sc.match(30);
#ifdef ENABLE_I16
		p.i16 = 1;
#endif
		for(int q = 0; q < 256; q++) { p.mmpens[q] = sc.mmpens[q]; p.npens[q] = sc.npens[q]; }
		char err[512] = {0};
		const char* dv = getenv("BT2GPU_DEVICE");
		ctx = bt2gpu_create(&ix, &rf, &p, dv ? atoi(dv) : 0, err, sizeof(err));
		if(!ctx) { cerr << "bt2gpu: " << err << endl; throw 1; }
		on = true;
	}

	// One batch iteration: every slot with mate_idx >= 0 holds a read that P0 has just prepared (bt2_search.cpp:2429-2590).
	// done_code = MATE_DONE of the caller.
	template<typename TSinkWrap, typename TObjs>
	void alignBatch(uint32_t nslots, int32_t* mate_idx, Read** rds, TAlScore* minsc, const int* nceil, const uint16_t* intervals,
	                bool* exhaustive, TObjs* msobjs, TSinkWrap* sinks, const msWorkerConsts& k, int32_t done_code) {
		seq_off.clear(); rng_seed.clear(); slot_of.clear(); minsc32.clear(); nceil32.clear(); interval32.clear(); seq.clear(); qual.clear();
		for(uint32_t m = 0; m < nslots; m++) {
			if(mate_idx[m] < 0) continue;
			const Read& rd = *rds[m];
			const size_t len = rd.length();
			seq_off.push_back((uint32_t)seq.size()); slot_of.push_back(m);
			seq.insert(seq.end(), (const uint8_t*)rd.patFw.buf(), (const uint8_t*)rd.patFw.buf() + len);      // codes 0..4
			qual.insert(qual.end(), (const uint8_t*)rd.qual.buf(), (const uint8_t*)rd.qual.buf() + len);      // Phred+33
			rng_seed.push_back(rd.seed); minsc32.push_back((int32_t)minsc[m]); interval32.push_back(intervals[m]);
			nceil32.push_back(k.sc.nCeil.f<int>((double)len));   // what SwAligner::initRead takes (aligner_sw.cpp:44), not the caller's min(.., rdlen) (:2555)
		}
		const uint32_t n = (uint32_t)slot_of.size();
		if(n == 0) return;
		seq_off.push_back((uint32_t)seq.size());
		bt2gpu_reads in; memset(&in, 0, sizeof(in));
		in.n = n; in.seq_off = seq_off.data(); in.seq = seq.data(); in.qual = qual.data(); in.rng_seed = rng_seed.data();
		in.minsc = minsc32.data(); in.nceil = nceil32.data(); in.interval = interval32.data();
		bt2gpu_results out;
		const auto t0 = std::chrono::steady_clock::now();
		if(bt2gpu_align_batch(ctx, &in, &out) != 0) { cerr << "bt2gpu: " << bt2gpu_last_error(ctx) << endl; throw 1; }
		ms_gpu += std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t0).count(); nbatches++; nreads += n;
		const TIndexOffU* plen = k.ebwtFw.plen();
		// replay, slot by slot (slots are disjoint: same parallelism as the reference's own phases)
#pragma omp parallel for default(shared)
		for(uint32_t j = 0; j < n; j++) {
			const uint32_t m = slot_of[j];
			const bt2gpu_read_result& r = out.reads[j];
			TSinkWrap& sink = sinks[m];
			SwResult& rg = msobjs[m].sd.resGap_;            // aligner_sw_driver.h:531: the buffer every AlnRes copy of this slot aliases
			const size_t rdlen = rds[m]->length();
			for(uint32_t a = 0; a < r.n_alns; a++) {
				const bt2gpu_aln& al = out.alns[r.aln_off + a];
				rg.reset();                                 // as SwDriver::extendSeeds does before every nextAlignment (:533)
				AlnRes& res = rg.alres;
				for(uint32_t e = 0; e < al.n_edits; e++) { const bt2gpu_edit& x = out.edits[al.edit_off + e]; res.ned().push_back(Edit(x.pos, x.chr, x.qchr, x.type)); }
				res.setScore(AlnScore(al.score, al.basesAligned, al.edits, al.ns, al.gaps));          // aligner_swsse_ee_u8.cpp:1297-1320
				res.setShape(al.tidx, al.refoff, plen[al.tidx], al.fw != 0, rdlen, true, 0, 0, true, 0, 0);
				res.setRefNs(al.refns);
				res.setParams(multiseedMms, multiseedLen, intervals[m], al.minsc);                    // aligner_sw_driver.cpp:571-575
				sink.report(0, &res, NULL);                 // aln_sink.cpp:1395 (copies res: count by value, edits through the alias)
			}
			// what the LAST backtrace attempt of the read left in that buffer (the SAM writer reads the copies' edits through it)
			rg.reset();
			for(uint32_t e = 0; e < r.n_scratch; e++) { const bt2gpu_edit& x = out.edits[r.scratch_off + e]; rg.alres.ned().push_back(Edit(x.pos, x.chr, x.qchr, x.type)); }
			rg.reset();
			RandomSource& rnd = msobjs[m].rnd;
			rnd.last = r.rng_last; rnd.lastOff = r.rng_lastOff; rnd.inited_ = true;
			minsc[m] = r.minsc; exhaustive[m] = r.exhaustive != 0;
			PerReadMetrics& prm = sink.prm;
			prm.nExIters = r.nExIters; prm.nExDps = r.nExDps; prm.nExDpSuccs = r.nExDpSuccs; prm.nExDpFails = r.nExDpFails; prm.nExUgs = r.nExUgs;
			prm.nDpFail = r.nDpFail; prm.nDpFailStreak = r.nDpFailStreak; prm.nDpLastSucc = r.nDpLastSucc; prm.nRedundants = r.nRedundants;
			if(r.bestLtMinsc != INT32_MIN) prm.bestLtMinscMate1 = r.bestLtMinsc;
			prm.nSeedRanges = r.numRanges; prm.nSeedElts = r.numElts;
			mate_idx[m] = done_code;                        // every seed round of the read ran on the device
		}
	}

	void finish() {
		if(!ctx) return;
		if(getenv("BT2GPU_TRACE")) cerr << "bt2gpu: " << nbatches << " batches, " << nreads << " reads, " << ms_gpu << " ms in bt2gpu_align_batch" << endl;
		bt2gpu_destroy(ctx); ctx = nullptr; on = false;
	}
};
