// libbt2gpu.so -- C-ABI (include/bt2gpu.h) and the host-side orchestration of the
// sm_100a kernels.  One context = one CUDA device, one stream; index + reference are
// copied to HBM at create time and stay resident.
//
// Per batch the host drives a short loop of phases (the device analogue of the
// reference's "one omp parallel for + barrier per phase", bt2_search.cpp:2591-2860):
//   begin -> [round select -> K_seed -> glue -> { step -> K_fill -> backtrace }* -> round end]* -> gather
// The host only reads back list lengths (a few words) between phases.
#include <cuda_runtime.h>
#include <cub/cub.cuh>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <string>
#include <vector>
#include <future>
#include <thread>
#include <mutex>
#include <algorithm>
#include "bt2gpu.h"
#include "dev_common.cuh"
#include "k_index.cuh"
#include "k_seed.cuh"
#include "k_fill.cuh"
#include "k_scalar.cuh"

using namespace bt2;

#define CK(call) do { cudaError_t e_ = (call); if(e_ != cudaSuccess) { \
	char b_[512]; snprintf(b_, sizeof(b_), "%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
	throw std::string(b_); } } while(0)

enum { PH_SEED = 0, PH_GLUE, PH_RESOLVE, PH_FILL, PH_BT, PH_H2D, PH_D2H, PH_RUN, PH_N };
enum { C_CAND0 = 0, C_CAND1, C_ACTIVE, C_EXT0, C_EXT1, C_ROUNDEND, C_SAMP, C_SEEDTASK, C_BTTASK, C_SEEDNT, C_XTASKN, C_XTASK, C_BTTASK1, C_N = 16 };

struct DBuf { // growable device buffer
	void* p = nullptr; size_t cap = 0;
	void ensure(size_t bytes) {
		if(bytes <= cap) return;
		if(p) CK(cudaFree(p));
		p = nullptr; cap = 0;
		const size_t want = bytes + bytes / 8 + 256;
		CK(cudaMalloc(&p, want)); cap = want;
	}
	void release() { if(p) cudaFree(p); p = nullptr; cap = 0; }
	template<typename T> T* as() const { return reinterpret_cast<T*>(p); }
};
struct HBuf { // growable pinned host buffer
	void* p = nullptr; size_t cap = 0;
	void ensure(size_t bytes) {
		if(bytes <= cap) return;
		if(p) CK(cudaFreeHost(p));
		p = nullptr; cap = 0;
		const size_t want = bytes + bytes / 8 + 256;
		CK(cudaMallocHost(&p, want)); cap = want;
	}
	void release() { if(p) cudaFreeHost(p); p = nullptr; cap = 0; }
	template<typename T> T* as() const { return reinterpret_cast<T*>(p); }
};

struct bt2gpu_ctx {
	int device = 0;
	cudaStream_t stream = nullptr;
	// the second slot set's stream: while one chunk of DP problems is in k_bt the next chunk's k_step + k_fill run beside it
	cudaStream_t stream2 = nullptr; cudaEvent_t evFork = nullptr, evJoin = nullptr, evFill[2] = { nullptr, nullptr };
	bool overlap = false; int btBlocksOverlap = 3;   // off: measured 37.6 vs 38.3 ms per 1M reads (both kernels scale with their resident warps); BT2GPU_OVERLAP=1 turns it on
	std::string err;
	bt2gpu_params hp;
	DevParams P; DevParams* dP = nullptr;
	DevIndex ix; DevRef rf;
	std::vector<void*> resident;      // index / reference allocations
	uint32_t nslots = 0, sampSlots = 0, maxBatch = 0;
	uint32_t fillK = 4; size_t cellStride = 0; uint32_t chunkSlots = 0;   // of the resident batch: the fill's K, bytes of cell flags per problem slot, slots usable
	bool textOk = false;              // the 2-bit reference is the index's joined text (same length)
	// problem-slot arrays
	DBuf probs, cells, pcache;
	DBuf pk, nm;                      // 2-bit packed reads and their N masks (k_pack_reads)
	// per-batch arrays
	// two sets of input arrays and of packed results: while one sub-batch is being aligned the next one is
	// uploaded and the previous one's results are copied back (bt2gpu_align_batch, separate copy streams)
	struct InSet { DBuf seq, qual, seq_off, rng_seed, minsc, nceil, interval; } in[2];
	struct OutSet { DBuf o_rr, o_alns, o_edits; } out[2];
	struct CopySpan { cudaEvent_t a = nullptr, b = nullptr; bool pending = false; } h2dspan[2], d2hspan[2];
	cudaStream_t h2d = nullptr, d2h = nullptr;
	int cur = 0;                      // set the resident batch lives in
	uint32_t seq_base = 0;            // seq_off[0] of the resident batch (offsets are rebased on the device)
	uint32_t pipeChunk = 0;           // reads per sub-batch of the pipelined bt2gpu_align_batch
	DBuf states, scratch, hits, seedtasks, exttasks, lists[7], counters, stats;
	DBuf p_sav, p_sorted, p_rank, p_sp2, p_exr, p_exrfw, p_nsp, p_nsmall, p_nexr, p_nelt, p_nadd;
	DBuf s_ovr, s_ovp, s_ovv, s_sr, s_sv, s_cv, s_rn, s_rcur, s_rth, s_rfl, s_rco, s_rcn, s_rns, s_mass, s_elim;
	DBuf chunks, sats, alnp, editp, pooltops, poolerr;
	DBuf nal, ned, albase, edbase, cubtmp;
	// two sets of pinned result arrays, taken in turn by successive calls: the results of a call stay valid while the next one runs
	struct HostSet { HBuf rr, alns, edits; } hset[2]; int hcur = 0;
	// bt2gpu_submit / bt2gpu_wait: the batch in flight on the context's worker thread
	std::thread worker; bool inflight = false; int flight_rc = 0; bt2gpu_results flight_out;
	volatile HostBox* hbox = nullptr; HostBox* dbox = nullptr;   // mapped pinned mailbox (k_scalar.cuh)
	// state of the current resident batch
	uint32_t n = 0, W = 0, maxOffs = 0; uint64_t nbases = 0;
	uint64_t out_alns = 0, out_edits = 0;
	Pools pools;
	double create_ms[4] = { 0, 0, 0, 0 }; uint64_t create_bytes[4] = { 0, 0, 0, 0 };   // bt2gpu_get_create_info
	uint32_t poolScale[4] = { 1, 1, 1, 1 };   // chunk / sat / aln / edit pool: doubled when a batch overflows it (the batch is then run again)
	uint64_t poolRetries = 0;
	bt2gpu_counters ctr;
	std::vector<cudaEvent_t> evpool; size_t evused = 0;
	struct Span { int ph; cudaEvent_t a, b; }; std::vector<Span> spans;
};

namespace {

// the K (diagonals per lane) K_fill needs for reads of length L with minimum score minsc: the widest band any of their windows
// can have (dev_common.cuh: fill_band; windows are framed as in dev_step.cuh: cols = L + 4 maxgap, at most 199)
uint32_t fill_K_of(const DevParams& P, uint32_t L, int32_t minsc, bool exact) {
	if(L == 0 || L > BT2GPU_MAX_ROWS) return 2;
	const int rg = max_read_gaps(P, minsc, L), fg = max_ref_gaps(P, minsc, L);
	int64_t maxgap = rg > fg ? rg : fg; if(maxgap > P.maxhalf) maxgap = P.maxhalf; if(maxgap < 0) maxgap = 0;
	int64_t cols = (int64_t)L + 4 * maxgap; if(cols > BT2GPU_MAX_COLS - 1) cols = BT2GPU_MAX_COLS - 1;
	const bool i16 = P.i16 && (!P.enable8 || minsc < -254);      // the problem's path (dev_step.cuh)
	int dlo, dhi; fill_band(P, (int)L, (int)cols, (!i16 && minsc < -254) ? -254 : minsc, exact, dlo, dhi);
	return fill_K_for(dhi - dlo + 1);
}

template<typename T> T* upload(bt2gpu_ctx* c, const T* src, size_t n) {
	T* d = nullptr;
	CK(cudaMalloc((void**)&d, std::max<size_t>(n, 1) * sizeof(T) + 16));
	if(n) CK(cudaMemcpy(d, src, n * sizeof(T), cudaMemcpyHostToDevice));
	c->resident.push_back(d); c->create_bytes[0] += n * sizeof(T);
	return d;
}

cudaEvent_t get_event(bt2gpu_ctx* c) {
	if(c->evused == c->evpool.size()) { cudaEvent_t e; CK(cudaEventCreate(&e)); c->evpool.push_back(e); }
	return c->evpool[c->evused++];
}
struct Phase {
	bt2gpu_ctx* c; size_t idx;
	cudaStream_t st;
	Phase(bt2gpu_ctx* c_, int ph, cudaStream_t st_ = nullptr) : c(c_), st(st_ ? st_ : c_->stream) {
		bt2gpu_ctx::Span s; s.ph = ph; s.a = get_event(c); s.b = get_event(c);
		CK(cudaEventRecord(s.a, st)); c->spans.push_back(s); idx = c->spans.size() - 1;
	}
	~Phase() { cudaEventRecord(c->spans[idx].b, st); }
};
void flush_spans(bt2gpu_ctx* c) {
	double* acc[PH_N] = { &c->ctr.ms_seed, &c->ctr.ms_glue, &c->ctr.ms_resolve, &c->ctr.ms_fill, &c->ctr.ms_bt, &c->ctr.ms_h2d, &c->ctr.ms_d2h, &c->ctr.ms_run };
	for(auto& s : c->spans) { float ms = 0; if(cudaEventElapsedTime(&ms, s.a, s.b) == cudaSuccess) *acc[s.ph] += ms; }
	c->spans.clear(); c->evused = 0;
}

inline uint32_t nblk(uint64_t n, uint32_t t) { return (uint32_t)((n + t - 1) / t); }

// copy the device counters to the host and wait
const uint32_t* read_counters(bt2gpu_ctx* c) {
	k_publish_counters<<<1, 32, 0, c->stream>>>(c->dbox, c->counters.as<uint32_t>());
	CK(cudaStreamSynchronize(c->stream));
	return const_cast<const uint32_t*>(c->hbox->counters);
}
void zero_counters(bt2gpu_ctx* c, uint32_t mask, cudaStream_t st = nullptr) { k_zero<<<1, 32, 0, st ? st : c->stream>>>(c->counters.as<uint32_t>(), mask); }
void zero_counter(bt2gpu_ctx* c, int which, cudaStream_t st = nullptr) { zero_counters(c, 1u << which, st); }

void setup_params(bt2gpu_ctx* c, const bt2gpu_params* p) {
	DevParams& P = c->P; memset(&P, 0, sizeof(P));
	P.seedLen = p->seedLen; P.seedMms = p->seedMms; P.nSeedRounds = p->nSeedRounds; P.seedBoostThresh = p->seedBoostThresh;
	P.maxhalf = p->maxhalf; P.mxIter = p->mxIter; P.mxUg = p->mxUg; P.mxDp = p->mxDp; P.streak = p->streak; P.doExtend = p->doExtend;
	P.enable8 = p->enable8; P.tighten = p->tighten; P.khits = p->khits; P.mhits = p->mhits; P.msample = p->msample;
	P.nofw = p->nofw; P.norc = p->norc; P.gapbar = p->gapbar;
	P.rdGapOpen = p->rdGapConst + p->rdGapLinear; P.rdGapExt = p->rdGapLinear;
	P.rfGapOpen = p->rfGapConst + p->rfGapLinear; P.rfGapExt = p->rfGapLinear;
	P.matchBonus = p->matchBonus; P.i16 = p->i16;
	for(int i = 0; i < 256; i++) {
		P.mmpens[i] = (uint8_t)std::min(255, std::max(0, p->mmpens[i]));
		P.npens[i] = (uint8_t)std::min(255, std::max(0, p->npens[i]));
	}
}

// scratch of the row sampler (k_prio2: RowSampler + Random1toN of SwDriver::prioritizeSATups), one slot per read sampled at a time
void ensure_sample(bt2gpu_ctx* c, uint32_t W) {
	const size_t S = c->sampSlots, MI = BT2_SAMPLE_MI;
	c->s_ovr.ensure(S * BT2_SAMPLE_HASH * 4); c->s_ovv.ensure(S * BT2_SAMPLE_HASH * 2);   // hash table of the sparse Random1toN (dev_prio.cuh)
	c->s_sr.ensure(S * MI * 2); c->s_sv.ensure(S * MI * 2); c->s_cv.ensure(S * MI * 2);
	c->s_rn.ensure(S * W * 2); c->s_rcur.ensure(S * W * 2); c->s_rth.ensure(S * W * 2); c->s_rfl.ensure(S * W);
	c->s_rco.ensure(S * W * 4); c->s_rcn.ensure(S * W * 2); c->s_rns.ensure(S * W * 2); c->s_mass.ensure(S * W * 8); c->s_elim.ensure(S * W);
}

void ensure_batch(bt2gpu_ctx* c, uint32_t n, uint64_t nbases, uint32_t maxOffs) {
	const uint32_t W = 2 * maxOffs;
	const size_t nn = (size_t)n + 1;
	c->states.ensure(nn * sizeof(ReadState));
	c->scratch.ensure(nn * BT2_SCRATCH_EDITS * 4);
	c->pk.ensure(nn * BT2_PK_WORDS * 4); c->nm.ensure(nn * BT2_NM_WORDS * 4);
	c->hits.ensure(nn * W * sizeof(SeedHit)); c->seedtasks.ensure(nn * W * sizeof(SeedTask)); c->exttasks.ensure(nn * W * sizeof(SeedTask));
	for(int i = 0; i < 7; i++) c->lists[i].ensure(nn * 4);
	c->p_sav.ensure(nn * W * 2); c->p_sorted.ensure(nn * W); c->p_rank.ensure(nn * W * 2); c->p_sp2.ensure(nn * W * 2);
	c->p_exr.ensure(nn * W * sizeof(ExRange16)); c->p_exrfw.ensure(nn * W);
	c->p_nsp.ensure(nn * 4); c->p_nsmall.ensure(nn * 4); c->p_nexr.ensure(nn * 4); c->p_nelt.ensure(nn * 8); c->p_nadd.ensure(nn * 4);
	ensure_sample(c, W);
	// pools
	Pools& p = c->pools;
	static const size_t pdiv = getenv("BT2GPU_POOL_DIV") ? (size_t)std::max(1, atoi(getenv("BT2GPU_POOL_DIV"))) : 1;   // test knob: start with pools this much too small
	p.chunk_cap = (uint32_t)std::min<size_t>((nn * 6 + 4096) / pdiv * c->poolScale[0] + 8, 0x7fffffff);
	p.sat_cap   = (uint32_t)std::min<size_t>((nn * 48 + 65536) / pdiv * c->poolScale[1] + 8, 0x7fffffff);
	p.aln_cap   = (uint32_t)std::min<size_t>((nn * 6 + 4096) / pdiv * c->poolScale[2] + 8, 0x7fffffff);
	p.edit_cap  = (uint32_t)std::min<size_t>((nn * 96 + 65536) / pdiv * c->poolScale[3] + 8, 0x7fffffff);
	c->chunks.ensure((size_t)p.chunk_cap * sizeof(Chunk)); c->sats.ensure((size_t)p.sat_cap * sizeof(SatEntry));
	c->alnp.ensure((size_t)p.aln_cap * sizeof(AlnRec)); c->editp.ensure((size_t)p.edit_cap * 4);
	p.chunks = c->chunks.as<Chunk>(); p.sats = c->sats.as<SatEntry>(); p.alns = c->alnp.as<AlnRec>(); p.edits = c->editp.as<uint32_t>();
	uint32_t* tops = c->pooltops.as<uint32_t>();
	p.chunk_top = tops + 0; p.sat_top = tops + 1; p.aln_top = tops + 2; p.edit_top = tops + 3;
	p.error = c->poolerr.as<uint32_t>();
	c->nal.ensure(nn * 4); c->ned.ensure(nn * 4); c->albase.ensure(nn * 4); c->edbase.ensure(nn * 4);
	size_t tmp = 0; cub::DeviceScan::ExclusiveSum(nullptr, tmp, (uint32_t*)nullptr, (uint32_t*)nullptr, (int)nn);
	c->cubtmp.ensure(tmp + 256);
}

// K_fill for the batch's K (template parameter: the cells of a step live in registers); dbg = the test instantiation that also
// dumps H, E and F
void launch_fill(uint32_t K, bool dbg, const DevParams* dP, const FillArgs& fa, cudaStream_t st) {
	const dim3 grid(nblk(fa.nprob, BT2_FILL_WARPS * 4)), block(BT2_FILL_WARPS * 32);
#if BT2_FILL_DYNSMEM
	// (the long-read build: the tables of a 1000-row problem do not fit the 48 KB a kernel gets without asking)
	#define BT2_LF1(KK, DD) { static bool once[64] = { false }; const size_t smem = BT2_FILL_WARPS * sizeof(FillShared<KK>); \
		int dev_ = 0; CK(cudaGetDevice(&dev_)); dev_ &= 63;   /* (the attribute is per device) */ \
		if(!once[dev_]) { CK(cudaFuncSetAttribute(k_fill<KK, DD>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)); once[dev_] = true; } \
		k_fill<KK, DD><<<grid, block, smem, st>>>(dP, fa); }
	#define BT2_LF(KK) { if(dbg) BT2_LF1(KK, true) else BT2_LF1(KK, false) }
#else
	#define BT2_LF(KK) { if(dbg) k_fill<KK, true><<<grid, block, 0, st>>>(dP, fa); else k_fill<KK, false><<<grid, block, 0, st>>>(dP, fa); }
#endif
	switch(K) { case 2: BT2_LF(2) break; case 4: BT2_LF(4) break; case 8: BT2_LF(8) break; default: BT2_LF(16) break; }
	#undef BT2_LF
	#undef BT2_LF1
}

struct Staged { uint32_t n = 0, maxOffs = 1, base = 0; uint64_t nb = 0; int set = 0; uint32_t fillK = 2, maxRows = 1; };

void settle(bt2gpu_ctx::CopySpan& sp, double& acc) { // the copy a span brackets is over (or is waited for here): add its time
	if(!sp.pending) return;
	CK(cudaEventSynchronize(sp.b));
	float ms = 0; if(cudaEventElapsedTime(&ms, sp.a, sp.b) == cudaSuccess) acc += ms;
	sp.pending = false;
}

// validates reads [lo, hi) and starts their copy into input set `set` on stream `st` (asynchronous when the
// caller's buffers are pinned); the batch becomes the resident one with bind_staged()
Staged stage_upload(bt2gpu_ctx* c, const bt2gpu_reads* in, uint32_t lo, uint32_t hi, int set, cudaStream_t st) {
	Staged g; g.n = hi - lo; g.set = set; g.base = in->seq_off[lo]; g.nb = (uint64_t)in->seq_off[hi] - g.base;
	const uint32_t seedLen = (uint32_t)c->P.seedLen;
	uint32_t lastL = 0; int32_t lastIv = 1, lastMin = 1;   // most batches have one read length: divide only when (L, interval, minsc) changes
	const bool exact = c->hp.exactBestLtMinsc != 0;
	for(uint32_t i = lo; i < hi; i++) {
		const int32_t iv = in->interval[i], ms = in->minsc[i];
		const uint32_t o0 = in->seq_off[i], o1 = in->seq_off[i + 1];
		if(iv < 1) throw std::string("bt2gpu: interval must be >= 1");
		if(o1 < o0) throw std::string("bt2gpu: seq_off must be non-decreasing");
		const uint32_t L = o1 - o0;
		if(L == lastL && iv == lastIv && ms == lastMin) continue;
		lastL = L; lastIv = iv; lastMin = ms;
		if(L > seedLen) { const uint32_t k = 1 + (L - seedLen) / (uint32_t)iv; if(k > g.maxOffs) g.maxOffs = k; }
		if(L <= BT2GPU_MAX_ROWS) { const uint32_t K = fill_K_of(c->P, L, ms, exact); if(K > g.fillK) g.fillK = K; if(L > g.maxRows) g.maxRows = L; }
	}
	if(2 * g.maxOffs > 65535u) throw std::string("bt2gpu: too many seeds per read");
	bt2gpu_ctx::InSet& I = c->in[set];
	const size_t nn = (size_t)g.n + 1;
	I.seq.ensure(g.nb + 16); I.qual.ensure(g.nb + 16);
	I.seq_off.ensure(nn * 4); I.rng_seed.ensure(nn * 4); I.minsc.ensure(nn * 4); I.nceil.ensure(nn * 4); I.interval.ensure(nn * 4);
	bt2gpu_ctx::CopySpan& sp = c->h2dspan[set];
	settle(sp, c->ctr.ms_h2d);
	CK(cudaEventRecord(sp.a, st));
	CK(cudaMemcpyAsync(I.seq_off.p, in->seq_off + lo, nn * 4, cudaMemcpyHostToDevice, st)); // rebased by k_begin
	if(g.nb) {
		CK(cudaMemcpyAsync(I.seq.p, in->seq + g.base, g.nb, cudaMemcpyHostToDevice, st));
		CK(cudaMemcpyAsync(I.qual.p, in->qual + g.base, g.nb, cudaMemcpyHostToDevice, st));
	}
	if(g.n) {
		CK(cudaMemcpyAsync(I.rng_seed.p, in->rng_seed + lo, (size_t)g.n * 4, cudaMemcpyHostToDevice, st));
		CK(cudaMemcpyAsync(I.minsc.p, in->minsc + lo, (size_t)g.n * 4, cudaMemcpyHostToDevice, st));
		CK(cudaMemcpyAsync(I.nceil.p, in->nceil + lo, (size_t)g.n * 4, cudaMemcpyHostToDevice, st));
		CK(cudaMemcpyAsync(I.interval.p, in->interval + lo, (size_t)g.n * 4, cudaMemcpyHostToDevice, st));
	}
	CK(cudaEventRecord(sp.b, st)); sp.pending = true;
	return g;
}

// makes a staged batch the resident one (no kernel of the context may be running) and orders the compute
// stream after its copy
void bind_staged(bt2gpu_ctx* c, const Staged& g) {
	c->cur = g.set; c->n = g.n; c->nbases = g.nb; c->maxOffs = g.maxOffs; c->W = 2 * g.maxOffs; c->seq_base = g.base;
	{ // the layout of the cell flags follows the batch: K from its lowest minimum score, rows from its longest read
	  static const char* ek = getenv("BT2GPU_FILL_K");
	  uint32_t K = g.fillK; if(ek && atoi(ek) > (int)K) K = (uint32_t)atoi(ek);
	  if(K != 2 && K != 4 && K != 8 && K != 16) K = 16;
	  c->fillK = K; c->cellStride = fill_cell_bytes(g.maxRows, K);
	  c->chunkSlots = (uint32_t)std::min<size_t>(c->nslots, c->cells.cap / c->cellStride) & ~3u;
	  if(c->chunkSlots < 4) throw std::string("bt2gpu: not enough device memory for the DP cell flags"); }
	ensure_batch(c, g.n, g.nb, g.maxOffs);
	CK(cudaStreamWaitEvent(c->stream, c->h2dspan[g.set].b, 0));
}

#define LAUNCH(c) (c)->ctr.launches++

struct PoolOverflow { uint32_t mask; };

// aligns the resident batch; the packed results stay in out[cur] with offsets shifted by (aln_glob, edit_glob)
void do_run(bt2gpu_ctx* c, uint32_t aln_glob = 0, uint32_t edit_glob = 0) {
	const uint32_t n = c->n, W = c->W, maxOffs = c->maxOffs;
	bt2gpu_ctx::InSet& I = c->in[c->cur]; bt2gpu_ctx::OutSet& O = c->out[c->cur];
	cudaStream_t st = c->stream;
	c->out_alns = c->out_edits = 0;
	if(n == 0) return;
	bt2gpu_ctx::Span runspan; runspan.ph = PH_RUN; runspan.a = get_event(c); runspan.b = get_event(c);
	CK(cudaEventRecord(runspan.a, st));
	k_reset<<<1, 32, 0, st>>>(c->counters.as<uint32_t>(), c->pooltops.as<uint32_t>(), c->poolerr.as<uint32_t>(), c->stats.as<unsigned long long>());
	k_clear16<<<148 * 8, 256, 0, st>>>(c->scratch.as<uint4>(), (size_t)n * BT2_SCRATCH_EDITS / 4);
	uint32_t* cnt = c->counters.as<uint32_t>();
	unsigned long long* stats = c->stats.as<unsigned long long>();
	ReadState* states = c->states.as<ReadState>();
	uint32_t* L[7]; for(int i = 0; i < 7; i++) L[i] = c->lists[i].as<uint32_t>();
	// L0/L1 round candidates, L2 active, L3/L4 extension ping-pong, L5 round end, L6 sampling
	BatchArrays ba; ba.n = n; ba.seq_base = c->seq_base; ba.seq_off = I.seq_off.as<uint32_t>(); ba.seq = I.seq.as<uint8_t>(); ba.qual = I.qual.as<uint8_t>();
	ba.rng_seed = I.rng_seed.as<uint32_t>(); ba.minsc = I.minsc.as<int32_t>(); ba.nceil = I.nceil.as<int32_t>(); ba.interval = I.interval.as<int32_t>();
	PrioArrays pa; pa.sav = c->p_sav.as<uint16_t>(); pa.sorted = c->p_sorted.as<uint8_t>(); pa.rank = c->p_rank.as<uint16_t>(); pa.sp2 = c->p_sp2.as<uint16_t>();
	pa.exr = c->p_exr.as<ExRange16>(); pa.exrfw = c->p_exrfw.as<uint8_t>(); pa.nsp = c->p_nsp.as<uint32_t>(); pa.nsmall = c->p_nsmall.as<uint32_t>();
	pa.nexr = c->p_nexr.as<uint32_t>(); pa.nelt = c->p_nelt.as<uint64_t>(); pa.nelt_added = c->p_nadd.as<uint32_t>();
	BtCtx bcx; bcx.P = c->P; bcx.pools = c->pools; bcx.states = states; bcx.probs = c->probs.as<Problem>();
	bcx.cells = c->cells.as<uint8_t>(); bcx.cell_stride = c->cellStride; bcx.pcache = c->pcache.as<uint32_t>();
	bcx.scratch = c->scratch.as<uint32_t>(); bcx.seq = ba.seq; bcx.qual = ba.qual; bcx.set_pens();
	Problem* probs = c->probs.as<Problem>();
	const Pools pools = c->pools;
	int seedBlocks = 0, btMaxBlocks = 0, btOvBlocks = 0; { int nsm = 148; cudaDeviceGetAttribute(&nsm, cudaDevAttrMultiProcessorCount, c->device); seedBlocks = nsm * 8; btMaxBlocks = nsm * BT2_BT_BLOCKS_PER_SM; btOvBlocks = nsm * c->btBlocksOverlap; }

	k_begin<<<nblk(n, 256), 256, 0, st>>>(ba, states, L[0], cnt + C_CAND0); LAUNCH(c);
	k_pack_reads<<<nblk(n, 128), 128, 0, st>>>(ba, states, c->pk.as<uint32_t>(), c->nm.as<uint32_t>()); LAUNCH(c);
	uint32_t ncand = read_counters(c)[C_CAND0];
	int cand = 0;
	for(uint32_t round = 0; ncand > 0; round++) {
		zero_counters(c, (1u << C_ACTIVE) | (1u << C_EXT0) | (1u << C_EXT1) | (1u << C_ROUNDEND) | (1u << C_SAMP) | (1u << C_SEEDTASK) | (1u << C_SEEDNT)
		               | (1u << C_XTASKN) | (1u << C_XTASK) | (1u << (cand ? C_CAND0 : C_CAND1)));
		k_round_select<<<nblk(ncand, 256), 256, 0, st>>>(c->dP, states, L[cand], ncand, round, L[2], cnt + C_ACTIVE); LAUNCH(c);
		const uint32_t nactive = read_counters(c)[C_ACTIVE];
		if(nactive == 0) break;
		{
			Phase ph(c, PH_SEED);
			SeedKernelArgs sk; sk.active = L[2]; sk.nactive = nactive; sk.W = W; sk.maxOffs = maxOffs; sk.states = states; sk.seq = ba.seq;
			sk.pk = c->pk.as<uint32_t>(); sk.nm = c->nm.as<uint32_t>(); sk.hits = c->hits.as<SeedHit>(); sk.tasks = c->seedtasks.as<SeedTask>(); sk.ntasks = cnt + C_SEEDNT; sk.counter = cnt + C_SEEDTASK; sk.stats = stats;
			sk.xtasks = c->exttasks.as<SeedTask>(); sk.nxtasks = cnt + C_XTASKN; sk.xcounter = cnt + C_XTASK; sk.text = (c->ix.sa32 && c->textOk) ? c->rf.buf : nullptr;
			const uint64_t nslots = (uint64_t)nactive * W;
			k_seed_init<<<(uint32_t)((nslots + 255) / 256), 256, 0, st>>>(c->ix, c->dP, sk); LAUNCH(c);
			const uint32_t blocks = (uint32_t)std::min<uint64_t>((uint64_t)seedBlocks, (nslots + 31) / 32);
			if(c->ix.len < 0xfffffff0ull) k_seed_walk<uint32_t><<<blocks, 256, 0, st>>>(c->ix, c->dP, sk);
			else k_seed_walk<uint64_t><<<blocks, 256, 0, st>>>(c->ix, c->dP, sk);
			LAUNCH(c); c->ctr.n_seed++;
			if(sk.text) { k_seed_ext1<<<blocks, 256, 0, st>>>(c->ix, sk); LAUNCH(c); }
		}
		uint32_t ncur, nsamp;
		{
			Phase ph(c, PH_GLUE);
			k_prio1<<<nblk(nactive, 128), 128, 0, st>>>(c->dP, pools, states, L[2], nactive, c->hits.as<SeedHit>(), W, maxOffs, pa,
			                                            L[3], cnt + C_EXT0, L[6], cnt + C_SAMP); LAUNCH(c);
			const uint32_t* hc = read_counters(c);
			ncur = hc[C_EXT0]; nsamp = hc[C_SAMP];
			// reads with large seed ranges (repeats) are sampled row by row, a thread per read with ~16 KB of scratch each: the
			// slots grow with the demand (a repeat-rich batch sends a tenth of its reads here), up to 128 Ki (2 GB)
			if(nsamp > c->sampSlots && c->sampSlots < (1u << 17)) {
				c->sampSlots = std::min<uint32_t>((nsamp + 4095u) & ~4095u, 1u << 17);
				ensure_sample(c, W);
			}
			SampleArrays sa; sa.hkey = c->s_ovr.as<uint32_t>(); sa.hval = c->s_ovv.as<uint16_t>();
			sa.seen_rng = c->s_sr.as<uint16_t>(); sa.seen_val = c->s_sv.as<uint16_t>(); sa.conv_val = c->s_cv.as<uint16_t>();
			sa.r_n = c->s_rn.as<uint16_t>(); sa.r_cur = c->s_rcur.as<uint16_t>(); sa.r_thresh = c->s_rth.as<uint16_t>(); sa.r_flags = c->s_rfl.as<uint8_t>();
			sa.r_conv_off = c->s_rco.as<uint32_t>(); sa.r_conv_n = c->s_rcn.as<uint16_t>(); sa.r_nseen = c->s_rns.as<uint16_t>();
			sa.mass = c->s_mass.as<double>(); sa.elim = c->s_elim.as<uint8_t>();
			for(uint32_t b = 0; b < nsamp; b += c->sampSlots) {
				const uint32_t k = std::min(c->sampSlots, nsamp - b);
				CK(cudaMemsetAsync(sa.hkey, 0xff, (size_t)k * BT2_SAMPLE_HASH * 4, st));
				k_prio2<<<nblk(k, 64), 64, 0, st>>>(c->dP, pools, states, L[2], c->hits.as<SeedHit>(), W, maxOffs, pa, sa, L[6], b, k); LAUNCH(c);
			}
		}
		int cur = 0; // L[3+cur] is the current extension list
		while(ncur > 0) {
			const int nxt = cur ^ 1;
			zero_counter(c, nxt ? C_EXT1 : C_EXT0);
			// A list longer than half the problem slots is cut into chunks that alternate between two slot sets on two streams,
			// staggered: chunk i+1's fill starts when chunk i's fill is over, so the ALU-bound fill shares the SMs with the
			// latency-bound backtrace of the chunk before (k_bt then takes fewer blocks per SM, leaving registers for the fill).
			const uint32_t half = (c->chunkSlots / 2) & ~3u;
			const bool ov = c->overlap && half >= 4 && ncur > half;
			const uint32_t per = ov ? half : c->chunkSlots;
			if(ov) { CK(cudaEventRecord(c->evFork, st)); CK(cudaStreamWaitEvent(c->stream2, c->evFork, 0)); }
			uint32_t ci = 0;
			for(uint32_t b = 0; b < ncur; b += per, ci++) {
				const uint32_t k = std::min(per, ncur - b);
				const uint32_t set = ov ? (ci & 1u) : 0u;
				cudaStream_t cs = set ? c->stream2 : st;
				Problem* sprobs = probs + (size_t)set * half;
				uint8_t* scells = c->cells.as<uint8_t>() + (size_t)set * half * c->cellStride;
				uint32_t* spc = c->pcache.as<uint32_t>() + (size_t)set * half * BT2_PC_WORDS;
				{ Phase ph(c, PH_RESOLVE, cs);
				  k_step<<<nblk(k, 128), 128, 0, cs>>>(c->dP, c->ix, c->rf, pools, states, L[3 + cur], b, k, sprobs, L[5], cnt + C_ROUNDEND); LAUNCH(c); c->ctr.n_step++; }
				if(ov && ci > 0) CK(cudaStreamWaitEvent(cs, c->evFill[set ^ 1u], 0));
				{ Phase ph(c, PH_FILL, cs);
				  FillArgs fa; fa.probs = sprobs; fa.nprob = k; fa.states = states; fa.seq = ba.seq; fa.qual = ba.qual; fa.cells = scells; fa.cell_stride = c->cellStride;
				  fa.pcache = spc; fa.hef_dbg = nullptr; fa.stats = stats + 8; fa.error = c->poolerr.as<uint32_t>(); fa.exact = c->hp.exactBestLtMinsc;
				  // (one launch over 1 Mi problems measured 3 % slower than four over 256 Ki -- 14.5 vs 14.0 ms: BT2GPU_FILL_SPLIT problems per launch, 0 = one launch)
				  static const uint32_t fsplit = getenv("BT2GPU_FILL_SPLIT") ? (atoi(getenv("BT2GPU_FILL_SPLIT")) > 0 ? (uint32_t)std::max(4, atoi(getenv("BT2GPU_FILL_SPLIT"))) & ~3u : 0u) : 262144u;
				  if(fsplit && k > fsplit) {
					for(uint32_t p0 = 0; p0 < k; p0 += fsplit) {
						FillArgs fb = fa; fb.probs = sprobs + p0; fb.nprob = std::min(fsplit, k - p0);
						fb.cells = scells + (size_t)p0 * c->cellStride; fb.pcache = spc + (size_t)p0 * BT2_PC_WORDS;
						launch_fill(c->fillK, false, c->dP, fb, cs); LAUNCH(c); c->ctr.n_fill++;
					}
				  } else { launch_fill(c->fillK, false, c->dP, fa, cs); LAUNCH(c); c->ctr.n_fill++; } }
				if(ov) CK(cudaEventRecord(c->evFill[set], cs));
				{ Phase ph(c, PH_BT, cs);
				  const int tc = set ? C_BTTASK1 : C_BTTASK;
				  zero_counter(c, tc, cs);
				  BtCtx sb = bcx; sb.probs = sprobs; sb.cells = scells; sb.pcache = spc;
				  const uint32_t btBlocks = std::min<uint32_t>(nblk(k, BT2_BT_THREADS), (uint32_t)(ov ? btOvBlocks : btMaxBlocks));
				  k_bt<<<btBlocks, BT2_BT_THREADS, 0, cs>>>(sb, k, cnt + tc, L[3 + nxt], cnt + (nxt ? C_EXT1 : C_EXT0), L[5], cnt + C_ROUNDEND); LAUNCH(c); c->ctr.n_bt++; }
			}
			if(ov) { CK(cudaEventRecord(c->evJoin, c->stream2)); CK(cudaStreamWaitEvent(st, c->evJoin, 0)); }
			ncur = read_counters(c)[nxt ? C_EXT1 : C_EXT0];
			cur = nxt;
		}
		const uint32_t nend = read_counters(c)[C_ROUNDEND];
		const int ncandIdx = cand ^ 1;
		if(nend) { k_round_end<<<nblk(nend, 256), 256, 0, st>>>(c->dP, states, L[5], nend, L[ncandIdx], cnt + (ncandIdx ? C_CAND1 : C_CAND0)); LAUNCH(c); }
		ncand = read_counters(c)[ncandIdx ? C_CAND1 : C_CAND0];
		cand = ncandIdx;
	}
	// ---- pack results
	{
		Phase ph(c, PH_GLUE);
		k_count<<<nblk(n, 256), 256, 0, st>>>(states, pools, n, c->nal.as<uint32_t>(), c->ned.as<uint32_t>(), stats + 16); LAUNCH(c);
		size_t tmp = c->cubtmp.cap;
		CK(cub::DeviceScan::ExclusiveSum(c->cubtmp.p, tmp, c->nal.as<uint32_t>(), c->albase.as<uint32_t>(), (int)n, st)); LAUNCH(c);
		tmp = c->cubtmp.cap;
		CK(cub::DeviceScan::ExclusiveSum(c->cubtmp.p, tmp, c->ned.as<uint32_t>(), c->edbase.as<uint32_t>(), (int)n, st)); LAUNCH(c);
		k_publish_tail<<<1, 32, 0, st>>>(c->dbox, c->albase.as<uint32_t>(), c->nal.as<uint32_t>(), c->edbase.as<uint32_t>(), c->ned.as<uint32_t>(), n, c->poolerr.as<uint32_t>());
		CK(cudaStreamSynchronize(st));
		const volatile uint32_t* last = c->hbox->last; const uint32_t perr = c->hbox->perr;
		if(perr & 32u) throw std::string("bt2gpu: internal error: a backtrace lane did not finish (k_bt watchdog)");
		if(perr & 64u) throw std::string("bt2gpu: internal error: a DP problem's band is wider than the fill's lanes hold");
		if(perr & 16u) throw std::string("bt2gpu: a read's minimum score is below -28416, beyond the score range of the 16-bit DP path");
		if(perr & 15u) throw PoolOverflow{perr & 15u};   // a bump-allocated pool ran out: do_run_grow() enlarges it and runs the batch again
		if(perr) { char b[128]; snprintf(b, sizeof(b), "bt2gpu: internal error (device error mask %u)", perr); throw std::string(b); }
		c->out_alns = (uint64_t)last[0] + last[1]; c->out_edits = (uint64_t)last[2] + last[3];
		settle(c->d2hspan[c->cur], c->ctr.ms_d2h);   // the previous results in this set have been copied out
		O.o_rr.ensure(((size_t)n + 1) * sizeof(bt2gpu_read_result));
		O.o_alns.ensure((c->out_alns + 1) * sizeof(bt2gpu_aln)); O.o_edits.ensure((c->out_edits + 1) * sizeof(bt2gpu_edit));
		k_gather<<<nblk(n, 128), 128, 0, st>>>(states, pools, c->scratch.as<uint32_t>(), n, c->albase.as<uint32_t>(), c->edbase.as<uint32_t>(), aln_glob, edit_glob,
		                                       O.o_rr.as<bt2gpu_read_result>(), O.o_alns.as<bt2gpu_aln>(), O.o_edits.as<bt2gpu_edit>()); LAUNCH(c);
	}
	CK(cudaEventRecord(runspan.b, st)); c->spans.push_back(runspan);
	// ---- counters
	k_publish_stats<<<1, 32, 0, st>>>(c->dbox, c->stats.as<unsigned long long>());
	CK(cudaStreamSynchronize(st));
	const volatile unsigned long long* hs = c->hbox->stats;
	CK(cudaGetLastError());
	c->ctr.reads += n; c->ctr.seeds_searched += hs[0]; c->ctr.lf_steps_search += hs[1]; c->ctr.lf_steps_extend += hs[2]; c->ctr.lf_sides_touched += hs[3]; c->ctr.ext_text_steps += hs[4]; c->ctr.sa_probes += hs[5];
	c->ctr.dp_problems += hs[8]; c->ctr.dp_cells += hs[9]; c->ctr.dp_cells_computed += hs[10];
	c->ctr.sa_rows_resolved += hs[16]; c->ctr.lf_steps_resolve += hs[17]; c->ctr.bt_attempts += hs[18]; c->ctr.bt_steps += hs[19]; c->ctr.alns_reported += hs[20];
	flush_spans(c);
}

// do_run; a batch that overflows one of the bump-allocated pools (seen-diagonal / seed-range chunks, prioritised SA rows, alignments,
// edits: sized for typical reads, -k 50 on a repeat-rich reference needs more) is run again from its inputs with that pool doubled --
// a run starts from the uploaded reads alone (k_begin), so nothing of the aborted attempt survives.
void do_run_grow(bt2gpu_ctx* c, uint32_t aln_glob = 0, uint32_t edit_glob = 0) {
	for(int attempt = 0;; attempt++) {
		try { do_run(c, aln_glob, edit_glob); return; }
		catch(const PoolOverflow& po) {
			c->spans.clear(); c->evused = 0;
			if(attempt >= 8) throw std::string("bt2gpu: device pools still exhausted after growing them 8 times; use smaller batches");
			for(int k = 0; k < 4; k++) if(po.mask & (1u << k)) c->poolScale[k] *= 2;
			c->poolRetries++;
			if(getenv("BT2GPU_TRACE")) fprintf(stderr, "[bt2gpu] pool overflow (mask %u): scales now %u %u %u %u, running the batch again\n", po.mask, c->poolScale[0], c->poolScale[1], c->poolScale[2], c->poolScale[3]);
			ensure_batch(c, c->n, c->nbases, c->maxOffs);
		}
	}
}

// grows a pinned result buffer, keeping what is already in it
void grow_keep(HBuf& h, size_t keep, size_t want) {
	if(want <= h.cap) return;
	HBuf nb; nb.ensure(std::max(want, h.cap * 2));
	if(keep) memcpy(nb.p, h.p, keep);
	h.release(); h = nb;
}

// starts the copy of the resident batch's packed results to the pinned host arrays at (read, alignment, edit)
// offsets (rb, ab, eb) on stream `st`; the compute stream has been synchronised by do_run
void start_fetch(bt2gpu_ctx* c, size_t rb, size_t ab, size_t eb, cudaStream_t st) {
	const uint32_t n = c->n;
	bt2gpu_ctx::OutSet& O = c->out[c->cur];
	bt2gpu_ctx::CopySpan& sp = c->d2hspan[c->cur];
	settle(sp, c->ctr.ms_d2h);
	CK(cudaEventRecord(sp.a, st));
	if(n) CK(cudaMemcpyAsync(c->hset[c->hcur].rr.as<bt2gpu_read_result>() + rb, O.o_rr.p, (size_t)n * sizeof(bt2gpu_read_result), cudaMemcpyDeviceToHost, st));
	if(c->out_alns) CK(cudaMemcpyAsync(c->hset[c->hcur].alns.as<bt2gpu_aln>() + ab, O.o_alns.p, c->out_alns * sizeof(bt2gpu_aln), cudaMemcpyDeviceToHost, st));
	if(c->out_edits) CK(cudaMemcpyAsync(c->hset[c->hcur].edits.as<bt2gpu_edit>() + eb, O.o_edits.p, c->out_edits * sizeof(bt2gpu_edit), cudaMemcpyDeviceToHost, st));
	CK(cudaEventRecord(sp.b, st)); sp.pending = true;
}

void do_fetch(bt2gpu_ctx* c, bt2gpu_results* out) {
	const uint32_t n = c->n;
	c->hset[c->hcur].rr.ensure(((size_t)n + 1) * sizeof(bt2gpu_read_result));
	c->hset[c->hcur].alns.ensure((c->out_alns + 1) * sizeof(bt2gpu_aln)); c->hset[c->hcur].edits.ensure((c->out_edits + 1) * sizeof(bt2gpu_edit));
	start_fetch(c, 0, 0, 0, c->stream);
	settle(c->d2hspan[c->cur], c->ctr.ms_d2h);
	out->n = n; out->reads = c->hset[c->hcur].rr.as<bt2gpu_read_result>(); out->alns = c->hset[c->hcur].alns.as<bt2gpu_aln>(); out->n_alns = c->out_alns;
	out->edits = c->hset[c->hcur].edits.as<bt2gpu_edit>(); out->n_edits = c->out_edits;
}

// bt2gpu_align_batch for more reads than one sub-batch: sub-batch s+1 is uploaded and sub-batch s-1's results are
// copied back while sub-batch s is aligned (input / result sets alternate; copies run on their own streams)
void do_align_pipelined(bt2gpu_ctx* c, const bt2gpu_reads* in, bt2gpu_results* out, uint32_t chunk) {
	const uint32_t n = in->n;
	c->hset[c->hcur].rr.ensure(((size_t)n + 1) * sizeof(bt2gpu_read_result));
	grow_keep(c->hset[c->hcur].alns, 0, ((size_t)n + (n >> 2) + 1024) * sizeof(bt2gpu_aln));
	grow_keep(c->hset[c->hcur].edits, 0, ((size_t)n * 4 + 4096) * sizeof(bt2gpu_edit));
	size_t rb = 0, ab = 0, eb = 0;
	static const bool trace = getenv("BT2GPU_TRACE") != nullptr;
	auto now = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; };
	const double t00 = now();
	// cut points: a short first sub-batch (its upload is the only one nothing overlaps); a short last one (for its download,
	// BT2GPU_PIPE_TAIL=1) measured slower: a sub-batch costs ~1.4 ms of fixed time, the last download less
	std::vector<uint32_t> cut(1, 0u);
	{
		static const char* er = getenv("BT2GPU_PIPE_RAMP");
		const uint32_t small = std::max<uint32_t>(chunk / (er ? (uint32_t)std::max(1, atoi(er)) : 8u), 1u);
		uint32_t pos = std::min(n, small);
		cut.push_back(pos);
		const uint32_t lastMax = std::max(chunk, std::min(chunk + small, c->maxBatch));   // no sub-batch beyond BT2GPU_BATCH
		while(n - pos > lastMax) { pos += chunk; cut.push_back(pos); }
		static const char* et = getenv("BT2GPU_PIPE_TAIL");
		if((et ? atoi(et) != 0 : false) && n - pos > small && small < chunk) { pos = n - small; cut.push_back(pos); }
		if(pos < n) cut.push_back(n);
	}
	Staged g = stage_upload(c, in, cut[0], cut[1], 0, c->h2d);
	for(size_t ci = 0; ci + 1 < cut.size(); ci++) {
		const uint32_t lo = cut[ci], hi = cut[ci + 1];
		Staged nx;
		const double t0 = now();
		// the next sub-batch is staged by a helper thread: a copy from pageable memory blocks its caller, and this thread has
		// kernels to launch (the other set's batch finished before this point)
		std::future<Staged> fut;
		if(hi < n) {
			const uint32_t nlo = hi, nhi = cut[ci + 2]; const int nset = g.set ^ 1; const int dev = c->device;
			fut = std::async(std::launch::async, [c, in, nlo, nhi, nset, dev]() { CK(cudaSetDevice(dev)); return stage_upload(c, in, nlo, nhi, nset, c->h2d); });
		}
		struct Join { std::future<Staged>& f; ~Join() { if(f.valid()) { try { f.get(); } catch(...) { } } } } join{fut};   // never leave the helper running
		const double t1 = now();
		bind_staged(c, g);
		if(ab > 0xf0000000ull || eb > 0xf0000000ull) throw std::string("bt2gpu_align_batch: more than 2^32 alignments or edits in one call; use smaller batches");
		do_run_grow(c, (uint32_t)ab, (uint32_t)eb);
		const double t2 = now();
		if(trace) fprintf(stderr, "[bt2gpu] sub-batch at %u: +%.2f ms  stage next %.2f ms  run %.2f ms\n", lo, t0 - t00, t1 - t0, t2 - t1);
		if((ab + c->out_alns + 1) * sizeof(bt2gpu_aln) > c->hset[c->hcur].alns.cap || (eb + c->out_edits + 1) * sizeof(bt2gpu_edit) > c->hset[c->hcur].edits.cap) {
			CK(cudaStreamSynchronize(c->d2h)); // copies into the old arrays are over before they move
			grow_keep(c->hset[c->hcur].alns, ab * sizeof(bt2gpu_aln), (ab + c->out_alns + 1) * sizeof(bt2gpu_aln));
			grow_keep(c->hset[c->hcur].edits, eb * sizeof(bt2gpu_edit), (eb + c->out_edits + 1) * sizeof(bt2gpu_edit));
		}
		start_fetch(c, rb, ab, eb, c->d2h);
		rb += c->n; ab += c->out_alns; eb += c->out_edits;
		if(fut.valid()) nx = fut.get();   // (rethrows what the helper threw)
		g = nx;
	}
	const double t3 = now();
	for(int k = 0; k < 2; k++) { settle(c->d2hspan[k], c->ctr.ms_d2h); settle(c->h2dspan[k], c->ctr.ms_h2d); }
	if(trace) fprintf(stderr, "[bt2gpu] drain %.2f ms, total %.2f ms\n", now() - t3, now() - t00);
	out->n = n; out->reads = c->hset[c->hcur].rr.as<bt2gpu_read_result>(); out->alns = c->hset[c->hcur].alns.as<bt2gpu_aln>(); out->n_alns = ab;
	out->edits = c->hset[c->hcur].edits.as<bt2gpu_edit>(); out->n_edits = eb;
}

// ---------------------------------------------------------------- DP fill without any allocation (bt2gpu_fill_noalloc)
// No counterpart in the reference; added for callers that cannot call cudaMalloc
// (code under CUDA graph capture, tight serving loops).  The scratch k_fill needs -- the Problem / ReadState records it reads
// and the cell flags it writes -- lives in a file-scope device array that every GPU loading this library holds until exit.
#define BT2_NA_K 16                                                     // every band fits (32 x 16 diagonals >= 160 + 199)
#define BT2_NA_STRIDE ((size_t)((BT2GPU_MAX_ROWS + BT2_SKEW) * BT2_LANES * BT2_NA_K))   // fill_cell_bytes(160, 16)
struct NoAllocScratch {
	unsigned long long stats[32];
	uint32_t  error[4];
	Problem   probs[BT2GPU_NOALLOC_MAX_PROBLEMS];
	ReadState states[BT2GPU_NOALLOC_MAX_PROBLEMS];
	alignas(256) uint8_t cells[(size_t)BT2GPU_NOALLOC_MAX_PROBLEMS * BT2_NA_STRIDE];
};
// (a byte array: no dynamic initialisation of device objects; ~11.6 MB)
static __device__ __align__(256) unsigned char g_noalloc_scratch[sizeof(NoAllocScratch)];

// builds the records k_fill reads from the caller's arrays; a problem with a size out of range is switched off (read = NIL)
__global__ void k_na_setup(NoAllocScratch* sc, uint32_t nprob, const uint32_t* seq_off, const uint16_t* rows, const uint16_t* cols,
                           const uint32_t* rf_off, const uint8_t* rf, const int32_t* minsc) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i < 32) sc->stats[i] = 0;   // the array keeps the previous call's contents
	if(i < 4) sc->error[i] = 0;
	if(i >= nprob) return;
	Problem& pb = sc->probs[i];
	const uint32_t r = rows[i], c = cols[i];
	const bool ok = r >= 1 && r <= BT2GPU_MAX_ROWS && c >= 1 && c < BT2GPU_MAX_COLS;
	pb.read = ok ? i : BT2_NIL; pb.rows = (uint16_t)r; pb.cols = (uint16_t)c; pb.fw = 1; pb.lay = 0; pb.d0 = 0;
	pb.minsc = minsc[i]; pb.refl = 0; pb.tidx = 0; pb.triml = pb.corel = pb.corer = 0; pb.lrmax = -1; pb.ncand = 0;
	if(ok) for(uint32_t k = 0; k <= c; k++) pb.rf[k] = rf[rf_off[i] + k];
	ReadState& rs = sc->states[i];
	rs.seq_off = seq_off[i]; rs.rdlen = (uint16_t)r; pb.seq_off = seq_off[i];
}
// one block per problem: dense flags (row-major rows x cols; 0x80 outside the band) out of the cell layout, candidates, counts
__global__ void k_na_export(const DevParams* __restrict__ Pp, uint32_t nprob, const NoAllocScratch* sc, const uint64_t* flags_off, uint8_t* flags, uint16_t* cand, uint32_t* ncand, int32_t* lrmax) {
	const uint32_t i = blockIdx.x;
	const Problem& pb = sc->probs[i];
	const bool ok = pb.read != BT2_NIL;
	if(threadIdx.x == 0) { ncand[i] = ok ? pb.ncand : 0u; lrmax[i] = ok ? pb.lrmax : -1; }
	if(!ok) return;
	for(uint32_t k = threadIdx.x; k < pb.ncand; k += blockDim.x) cand[256u * i + k] = pb.cand[k];
	const uint32_t rows = pb.rows, cols = pb.cols, K = prob_K(pb), y = prob_y(pb); const int d0 = pb.d0;
	// diagonals the kernel completes for this problem: the lanes up to the widest band of its warp (problems 4w .. 4w+3)
	int widthW = 1;
	for(uint32_t q = i & ~3u; q < (i & ~3u) + 4u && q < nprob; q++) {
		if(sc->probs[q].read == BT2_NIL) continue;
		int lo, hi; fill_band(*Pp, sc->probs[q].rows, sc->probs[q].cols, sc->probs[q].minsc, false, lo, hi);
		if(hi - lo + 1 > widthW) widthW = hi - lo + 1;
	}
	const int ndiag = ((widthW - 1) / (int)K + 1) * (int)K;
	const uint8_t* cells = sc->cells + (size_t)i * BT2_NA_STRIDE;
	uint8_t* out = flags + flags_off[i];
	for(uint32_t k = threadIdx.x; k < rows * cols; k += blockDim.x) {
		const uint32_t r = k / cols, c = k % cols; const int dd = (int)c - (int)r - d0;
		out[k] = (dd >= 0 && dd < ndiag) ? cells[cell_index(r, c, d0, K, y)] : (uint8_t)0x80;
	}
}

// device address of the scratch array, resolved once per device with that device current: in host code the array's name is a
// host-side stand-in, not a device address
struct NoAllocDevice { void* ptr = nullptr; cudaEvent_t done = nullptr; bool used = false; };
static std::mutex g_na_mu;
static NoAllocDevice g_na_dev[64];
// Set-up, once per device, with that device current (called by bt2gpu_create, so that the first bt2gpu_fill_noalloc of a device does
// nothing but enqueue): the address of the scratch array, the modules of the three kernels (lazy loading), and the event that orders
// consecutive calls on different streams.
NoAllocDevice& noalloc_device(int device) {
	if(device < 0 || device >= 64) throw std::string("bt2gpu_fill_noalloc: device ordinal out of range");
	std::lock_guard<std::mutex> lk(g_na_mu);
	NoAllocDevice& d = g_na_dev[device];
	if(!d.ptr) {
		void* p = nullptr;
		CK(cudaGetSymbolAddress(&p, g_noalloc_scratch));
		if(!p) throw std::string("bt2gpu_fill_noalloc: cudaGetSymbolAddress returned no address");
		cudaFuncAttributes fa;
		CK(cudaFuncGetAttributes(&fa, k_na_setup)); CK(cudaFuncGetAttributes(&fa, k_na_export)); CK(cudaFuncGetAttributes(&fa, k_fill<BT2_NA_K, false>));
		CK(cudaEventCreateWithFlags(&d.done, cudaEventDisableTiming));
		d.ptr = p;
	}
	return d;
}

} // namespace

extern "C" {

int bt2gpu_abi_version(void) { return BT2GPU_ABI_VERSION; }

bt2gpu_ctx* bt2gpu_create(const bt2gpu_index* hix, const bt2gpu_ref* hrf, const bt2gpu_params* p, int device, char* err, size_t errlen) {
	bt2gpu_ctx* c = nullptr;
	try {
		if(!hix || !hrf || !p) throw std::string("bt2gpu_create: null argument");
		if(p->abi_version != BT2GPU_ABI_VERSION) throw std::string("bt2gpu_create: ABI version mismatch");
		if(hix->lineRate != 7) throw std::string("bt2gpu_create: only 128-byte sides (.bt2l, lineRate 7) are supported");
		if(p->matchBonus != 0) throw std::string("bt2gpu_create: end-to-end mode only (matchBonus must be 0)");
		if(p->seedLen < 1 || p->seedLen > 32) throw std::string("bt2gpu_create: seedLen must be within [1,32]");
		if(p->mxIter > BT2GPU_MAX_ITER) throw std::string("bt2gpu_create: mxIter too big");
		int ndev = 0;
		if(cudaGetDeviceCount(&ndev) != cudaSuccess || ndev == 0) throw std::string("bt2gpu_create: no CUDA device (this library has no CPU path)");
		if(device < 0 || device >= ndev) throw std::string("bt2gpu_create: bad device ordinal");
		CK(cudaSetDevice(device));
		c = new bt2gpu_ctx();
		c->device = device; c->hp = *p;
		memset(&c->ctr, 0, sizeof(c->ctr));
		CK(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&c->stream2, cudaStreamNonBlocking));
		for(cudaEvent_t* e : { &c->evFork, &c->evJoin, &c->evFill[0], &c->evFill[1] }) CK(cudaEventCreateWithFlags(e, cudaEventDisableTiming));
		if(const char* e = getenv("BT2GPU_OVERLAP")) c->overlap = atoi(e) != 0;                 // 0: one stream, one slot set (what the per-kernel times are read from)
		if(const char* e = getenv("BT2GPU_BT_BLOCKS")) c->btBlocksOverlap = std::max(1, std::min(BT2_BT_BLOCKS_PER_SM, atoi(e)));
		CK(cudaStreamCreateWithFlags(&c->h2d, cudaStreamNonBlocking)); CK(cudaStreamCreateWithFlags(&c->d2h, cudaStreamNonBlocking));
		for(int k = 0; k < 2; k++) {
			CK(cudaEventCreate(&c->h2dspan[k].a)); CK(cudaEventCreate(&c->h2dspan[k].b)); CK(cudaEventCreate(&c->d2hspan[k].a)); CK(cudaEventCreate(&c->d2hspan[k].b));
		}
		setup_params(c, p);
		auto now_ms = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; };
		const double t_start = now_ms(); double t_sa = 0, t_kt = 0;
		CK(cudaMalloc((void**)&c->dP, sizeof(DevParams)));
		CK(cudaMemcpy(c->dP, &c->P, sizeof(DevParams), cudaMemcpyHostToDevice));
		// ---- index
		DevIndex& ix = c->ix;
		ix.ebwt = upload(c, hix->ebwt, hix->ebwtTotLen);
		ix.ftab = upload(c, hix->ftab, hix->ftabLen); ix.eftab = upload(c, hix->eftab, hix->eftabLen);
		ix.offs = upload(c, hix->offs, hix->offsLen); ix.rstarts = upload(c, hix->rstarts, hix->nFrag * 3); ix.plen = upload(c, hix->plen, hix->nPat);
		for(int i = 0; i < 5; i++) ix.fchr[i] = hix->fchr[i];
		ix.len = hix->len; ix.zOff = hix->zOff; ix.nFrag = hix->nFrag; ix.nPat = hix->nPat; ix.offRate = hix->offRate; ix.ftabChars = hix->ftabChars;
		ix.sa32 = nullptr; ix.ktab = nullptr; ix.ktabChars = 0;
		// ---- full suffix array (k_index.cuh); BT2GPU_FULL_SA=0 keeps the sampled one (LF walks), e.g. to save 4 B/row
		{ const char* efs = getenv("BT2GPU_FULL_SA");
		  if(!(efs && atoi(efs) == 0) && hix->len < 0xfffffff0ull && hix->offRate >= 0 && hix->offRate < 31) {
			const uint64_t nrows = hix->len + 1;
			uint32_t* sa = nullptr;
			const double t0 = now_ms();
			CK(cudaMalloc((void**)&sa, nrows * 4 + 16)); c->resident.push_back(sa); c->create_bytes[1] = nrows * 4;
			CK(cudaMemset(sa, 0xff, nrows * 4));
			const uint64_t nsample = (nrows + ((1ull << hix->offRate) - 1)) >> hix->offRate;
			if(nsample > hix->offsLen) throw std::string("bt2gpu_create: SA sample shorter than the index needs");
			k_build_sa<<<(uint32_t)((nsample + 127) / 128), 128>>>(ix, sa, nsample);
			unsigned long long* dbad = nullptr; unsigned long long hbad = 0;
			CK(cudaMalloc((void**)&dbad, 8)); CK(cudaMemset(dbad, 0, 8));
			k_fix_sa<<<(uint32_t)((nrows + 255) / 256), 256>>>(ix, sa, nrows, 1, dbad);
			k_fix_sa<<<(uint32_t)((nrows + 255) / 256), 256>>>(ix, sa, nrows, 0, dbad);
			CK(cudaMemcpy(&hbad, dbad, 8, cudaMemcpyDeviceToHost)); CK(cudaFree(dbad));
			CK(cudaGetLastError());
			if(hbad) throw std::string("bt2gpu_create: inconsistent index (suffix-array expansion left rows unresolved)");
			ix.sa32 = sa; t_sa = now_ms() - t0;
		  } }
		// ---- reference
		std::vector<uint64_t> tstart(hrf->nrecs), cum(hrf->nrecs), rec0;
		{ uint64_t cu = 0, pos = 0;
		  for(uint64_t r = 0; r < hrf->nrecs; r++) {
			if(hrf->rec_first[r]) { rec0.push_back(r); pos = 0; }
			tstart[r] = pos; cum[r] = cu; pos += hrf->rec_off[r] + hrf->rec_len[r]; cu += hrf->rec_len[r];
		  }
		  rec0.push_back(hrf->nrecs);
		  if(rec0.size() != hix->nPat + 1) throw std::string("bt2gpu_create: reference records do not match the index (nPat)");
		}
		DevRef& rf = c->rf;
		rf.buf = upload(c, hrf->buf, (hrf->bufBases + 3) / 4);
		c->textOk = hrf->bufBases == hix->len;
		// ---- prefix table over the suffix array (k_index.cuh); BT2GPU_KTAB=0 keeps the index's own 10-mer ftab for the seed search
		{ const char* ek = getenv("BT2GPU_KTAB");
		  if(ix.sa32 && c->textOk && !(ek && atoi(ek) == 0) && hix->len > 0) {
			int K = 1; while(K < 16 && (1ull << (2 * K)) <= hix->len) K++;          // 4^K > len: about one row per key or fewer
			K = std::max(K, (int)hix->ftabChars); K = std::min(K, hix->len > (1ull << 28) ? 14 : 13);   // 268 MB / 1 GB; measured: below ~3 probes per seed a bigger table gains nothing
			if(ek && atoi(ek) > 0) K = std::min(std::max(atoi(ek), 2), 15);
			if(K > hix->ftabChars || (ek && atoi(ek) > 0)) {
				const uint64_t nk = 1ull << (2 * K), nrows = hix->len + 1;
				uint32_t* kt = nullptr;
				const double t0 = now_ms();
				CK(cudaMalloc((void**)&kt, (nk + 1) * 4 + 16)); c->resident.push_back(kt); c->create_bytes[2] = (nk + 1) * 4;
				const uint32_t* tw = reinterpret_cast<const uint32_t*>(rf.buf);
				k_build_ktab<<<(uint32_t)((nrows + 255) / 256), 256>>>(ix, tw, kt, nrows, (uint32_t)K);
				k_ktab_tail<<<(uint32_t)((nk + 255) / 256), 256>>>(ix, tw, kt, nrows, (uint32_t)K);
				CK(cudaGetLastError()); CK(cudaDeviceSynchronize());
				if(getenv("BT2GPU_CHECK_INDEX")) {   // self-test: every row lies in the interval of its own K-mer, the table is monotone
					std::vector<uint32_t> hk(nk + 1), hs(nrows);
					CK(cudaMemcpy(hk.data(), kt, (nk + 1) * 4, cudaMemcpyDeviceToHost)); CK(cudaMemcpy(hs.data(), ix.sa32, nrows * 4, cudaMemcpyDeviceToHost));
					const uint8_t* tb = hrf->buf;
					auto base = [&](uint64_t j) { return (uint64_t)((tb[j >> 2] >> ((j & 3) * 2)) & 3); };
					if(hk[0] != 0 || hk[nk] != nrows) { char b[160]; snprintf(b, sizeof(b), "bt2gpu_create: prefix table ends wrong (%u, %u of %llu rows)", hk[0], hk[nk], (unsigned long long)nrows); throw std::string(b); }
					for(uint64_t k = 0; k < nk; k++) if(hk[k] > hk[k + 1]) throw std::string("bt2gpu_create: prefix table not monotone");
					for(uint64_t i = 0; i < nrows; i++) {
						uint64_t key = 0; const uint64_t p = hs[i];
						for(int q = 0; q < K; q++) key = (key << 2) | (p + q < hix->len ? base(p + q) : 3);   // the end of the text sorts after T
						if(!(hk[key] <= i && i < hk[key + 1])) {
							char b[200]; snprintf(b, sizeof(b), "bt2gpu_create: prefix table wrong at row %llu (sa %llu key %llu: interval [%u, %u))", (unsigned long long)i, (unsigned long long)p, (unsigned long long)key, hk[key], hk[key + 1]);
							throw std::string(b);
						}
					}
				}
				ix.ktab = kt; ix.ktabChars = K; t_kt = now_ms() - t0;
			}
		  } }
		rf.rec_off = upload(c, hrf->rec_off, hrf->nrecs); rf.rec_len = upload(c, hrf->rec_len, hrf->nrecs);
		rf.rec_tstart = upload(c, tstart.data(), tstart.size()); rf.rec_cum = upload(c, cum.data(), cum.size());
		rf.ref_rec0 = upload(c, rec0.data(), rec0.size()); rf.nrecs = hrf->nrecs;
		// ---- problem slots
		CK(cudaDeviceSynchronize());
		const double t_slots = now_ms();
		const char* es = getenv("BT2GPU_SLOTS"); const char* eb = getenv("BT2GPU_BATCH");
		size_t freeB = 0, totB = 0; CK(cudaMemGetInfo(&freeB, &totB));
		const size_t slotCells = fill_cell_bytes(BT2GPU_MAX_ROWS, 4);   // K = 4 holds every band of the default scoring scheme; a batch that needs more gets fewer slots
		const size_t perSlot = sizeof(Problem) + slotCells + (size_t)BT2_PC_WORDS * 4;
		// 1 Mi slots (24 GB of cell flags at 150 bp): an extension pass over a 1 M-read batch is then ONE launch each of k_step / K_fill / k_bt,
		// and k_bt's persistent lanes take 9 problems each instead of 2.3 -- its tail, when most lanes have run out of problems, cost 2.4 ms of
		// 14.1 per 1 M reads with 262 144 slots (measured: 35.8 / 34.3 / 33.4 ms per step with 256 Ki / 512 Ki / 1 Mi slots)
		uint32_t slots = es ? (uint32_t)atoi(es) : 1048576u;
		const size_t budget = std::min<size_t>((freeB / 20) * 9, (size_t)32 << 30);   // (the long-read build's slots are 137 KB each: 32 GB of them is 245 000 problems in flight)
		if((size_t)slots * perSlot > budget) slots = (uint32_t)(budget / perSlot);
		if(slots < 64) throw std::string("bt2gpu_create: not enough device memory for the DP problem slots");
		c->nslots = slots; c->sampSlots = 2048;
		c->maxBatch = eb ? (uint32_t)atoi(eb) : (1u << 20);
		if(c->maxBatch < 1) c->maxBatch = 1;
		{ const char* ep = getenv("BT2GPU_PIPE"); c->pipeChunk = ep ? (uint32_t)std::max(1, atoi(ep)) : 524288u; }   // (measured through the C-ABI at 1 M reads: 24.6 M reads/s with 262 144, 25.8 M with 524 288)
		c->probs.ensure((size_t)slots * sizeof(Problem)); c->cells.ensure((size_t)slots * slotCells);
		c->pcache.ensure((size_t)slots * BT2_PC_WORDS * 4);
		c->counters.ensure(C_N * 4); c->stats.ensure(32 * 8); c->pooltops.ensure(16); c->poolerr.ensure(4);
		{ void* hb = nullptr; CK(cudaHostAlloc(&hb, sizeof(HostBox), cudaHostAllocMapped)); memset(hb, 0, sizeof(HostBox));
		  c->hbox = reinterpret_cast<volatile HostBox*>(hb); void* db = nullptr; CK(cudaHostGetDevicePointer(&db, hb, 0)); c->dbox = reinterpret_cast<HostBox*>(db); }
		noalloc_device(device);   // bt2gpu_fill_noalloc's per-device set-up happens here, not inside its first call
		CK(cudaDeviceSynchronize());
		c->create_bytes[3] = c->probs.cap + c->cells.cap + c->pcache.cap;
		{ const double t_end = now_ms(); c->create_ms[1] = t_sa; c->create_ms[2] = t_kt; c->create_ms[3] = t_end - t_slots;
		  c->create_ms[0] = (t_slots - t_start) - t_sa - t_kt; }
	} catch(const std::string& e) {
		if(err && errlen) snprintf(err, errlen, "%s", e.c_str());
		if(c) bt2gpu_destroy(c);
		return nullptr;
	}
	return c;
}

// every entry point runs with the context's device current and puts the caller's device back before it returns
struct DeviceScope { int prev = -1; explicit DeviceScope(int dev) { cudaGetDevice(&prev); if(prev != dev) cudaSetDevice(dev); else prev = -1; } ~DeviceScope() { if(prev >= 0) cudaSetDevice(prev); } };
#define API_GUARD(...) \
	if(!ctx) return -1; \
	try { DeviceScope ds_(ctx->device); __VA_ARGS__; } catch(const std::string& e) { ctx->err = e; ctx->spans.clear(); ctx->evused = 0; return -2; } \
	return 0;

#define NO_FLIGHT(what) if(ctx->inflight) throw std::string(what ": a batch submitted with bt2gpu_submit is still in flight; call bt2gpu_wait first")

int bt2gpu_upload(bt2gpu_ctx* ctx, const bt2gpu_reads* in) {
	API_GUARD({
		if(!in) throw std::string("bt2gpu_upload: null batch");
		NO_FLIGHT("bt2gpu_upload");
		if(in->n > ctx->maxBatch) throw std::string("bt2gpu_upload: batch larger than BT2GPU_BATCH; use bt2gpu_align_batch");
		bind_staged(ctx, stage_upload(ctx, in, 0, in->n, 0, ctx->stream));
		settle(ctx->h2dspan[0], ctx->ctr.ms_h2d);
	})
}
int bt2gpu_run_resident(bt2gpu_ctx* ctx) { API_GUARD({ NO_FLIGHT("bt2gpu_run_resident"); do_run_grow(ctx); }) }
int bt2gpu_fetch(bt2gpu_ctx* ctx, bt2gpu_results* out) { API_GUARD({ if(!out) throw std::string("bt2gpu_fetch: null out"); NO_FLIGHT("bt2gpu_fetch"); do_fetch(ctx, out); }) }

static void align_into_next_set(bt2gpu_ctx* ctx, const bt2gpu_reads* in, bt2gpu_results* out) {
	ctx->hcur ^= 1;   // the other set of pinned result arrays: what the previous call returned stays as it is
	const uint32_t chunk = std::min(ctx->pipeChunk, ctx->maxBatch);
	if(in->n <= chunk) { bind_staged(ctx, stage_upload(ctx, in, 0, in->n, 0, ctx->stream)); do_run_grow(ctx); do_fetch(ctx, out); }
	else do_align_pipelined(ctx, in, out, chunk);
}
int bt2gpu_align_batch(bt2gpu_ctx* ctx, const bt2gpu_reads* in, bt2gpu_results* out) {
	API_GUARD({
		if(!in || !out) throw std::string("bt2gpu_align_batch: null argument");
		NO_FLIGHT("bt2gpu_align_batch");
		align_into_next_set(ctx, in, out);
	})
}

// bt2gpu_align_batch split in two for a caller with one thread (the reference's batch loop): submit returns at once, the batch is
// aligned by a worker thread of the context while the caller finishes the reads of the batch before (P7, the SAM sink)
int bt2gpu_submit(bt2gpu_ctx* ctx, const bt2gpu_reads* in) {
	API_GUARD({
		if(!in) throw std::string("bt2gpu_submit: null batch");
		NO_FLIGHT("bt2gpu_submit");
		if(ctx->worker.joinable()) ctx->worker.join();
		ctx->inflight = true; ctx->flight_rc = 0; memset(&ctx->flight_out, 0, sizeof(ctx->flight_out));
		ctx->worker = std::thread([ctx, in]() {
			try { DeviceScope ds(ctx->device); align_into_next_set(ctx, in, &ctx->flight_out); }
			catch(const std::string& e) { ctx->err = e; ctx->spans.clear(); ctx->evused = 0; ctx->flight_rc = -2; }
			catch(...) { ctx->err = "bt2gpu_submit: unexpected exception on the worker thread"; ctx->flight_rc = -2; }
		});
	})
}
int bt2gpu_wait(bt2gpu_ctx* ctx, bt2gpu_results* out) {
	if(!ctx) return -1;
	if(!ctx->inflight) { ctx->err = "bt2gpu_wait: nothing was submitted"; return -2; }
	if(ctx->worker.joinable()) ctx->worker.join();
	ctx->inflight = false;
	if(ctx->flight_rc == 0 && out) *out = ctx->flight_out;
	return ctx->flight_rc;
}

int bt2gpu_test_fill(bt2gpu_ctx* ctx, uint32_t nprob, const uint32_t* seq_off, const uint8_t* seq, const uint8_t* qual,
                     const uint16_t* rows, const uint16_t* cols, const uint32_t* rf_off, const uint8_t* rf, const int32_t* minsc,
                     const uint64_t* hef_off, uint8_t* H, uint8_t* E, uint8_t* F, uint8_t* flags, uint16_t* cand, uint32_t* ncand, int32_t* lrmax,
                     int32_t mode, int32_t* band) {
	API_GUARD({
		if(nprob > 4096) throw std::string("bt2gpu_test_fill: more than 4096 problems");
		const bool exact = mode == 0;
		std::vector<ReadState> hs(nprob); std::vector<Problem> hp(nprob);
		uint64_t nb = 0; uint32_t K = 2, maxRows = 1;
		for(uint32_t i = 0; i < nprob; i++) {
			if(rows[i] < 1 || rows[i] > BT2GPU_MAX_ROWS || cols[i] < 1 || cols[i] >= BT2GPU_MAX_COLS) throw std::string("bt2gpu_test_fill: bad problem size");
			memset(&hs[i], 0, sizeof(ReadState)); hs[i].seq_off = seq_off[i]; hs[i].rdlen = rows[i];
			memset(&hp[i], 0, sizeof(Problem));
			hp[i].read = i; hp[i].rows = rows[i]; hp[i].cols = cols[i]; hp[i].fw = 1; hp[i].minsc = minsc[i]; hp[i].seq_off = seq_off[i];
			memcpy(hp[i].rf, rf + rf_off[i], (size_t)cols[i] + 1);
			nb = std::max<uint64_t>(nb, (uint64_t)seq_off[i] + rows[i]);
			int dlo, dhi; fill_band(ctx->P, rows[i], cols[i], minsc[i], exact, dlo, dhi);
			K = std::max(K, fill_K_for(dhi - dlo + 1)); maxRows = std::max<uint32_t>(maxRows, rows[i]);
		}
		if(mode == 2) K = BT2_NA_K;   // the layout bt2gpu_fill_noalloc uses
		const size_t stride = fill_cell_bytes(maxRows, K);
		DBuf cells, probs, states, dbg;
		struct Rel { DBuf* b[4]; ~Rel() { for(DBuf* x : b) x->release(); } } rel{{&cells, &probs, &states, &dbg}};
		const uint32_t np4 = (nprob + 3u) & ~3u;
		cells.ensure((size_t)np4 * stride); probs.ensure((size_t)np4 * sizeof(Problem)); states.ensure((size_t)np4 * sizeof(ReadState));
		if(exact) dbg.ensure((size_t)np4 * BT2_HEF_WORDS * 4);
		ctx->in[0].seq.ensure(nb + 16); ctx->in[0].qual.ensure(nb + 16);
		cudaStream_t st = ctx->stream;
		CK(cudaMemcpyAsync(ctx->in[0].seq.p, seq, nb, cudaMemcpyHostToDevice, st)); CK(cudaMemcpyAsync(ctx->in[0].qual.p, qual, nb, cudaMemcpyHostToDevice, st));
		CK(cudaMemcpyAsync(states.p, hs.data(), (size_t)nprob * sizeof(ReadState), cudaMemcpyHostToDevice, st));
		CK(cudaMemcpyAsync(probs.p, hp.data(), (size_t)nprob * sizeof(Problem), cudaMemcpyHostToDevice, st));
		CK(cudaMemsetAsync(ctx->stats.p, 0, 32 * 8, st)); CK(cudaMemsetAsync(ctx->poolerr.p, 0, 4, st));
		CK(cudaMemsetAsync(cells.p, 0x80, (size_t)np4 * stride, st));
		FillArgs fa; fa.probs = probs.as<Problem>(); fa.nprob = nprob; fa.states = states.as<ReadState>(); fa.seq = ctx->in[0].seq.as<uint8_t>();
		fa.qual = ctx->in[0].qual.as<uint8_t>(); fa.cells = cells.as<uint8_t>(); fa.cell_stride = stride; fa.pcache = nullptr; fa.hef_dbg = exact ? dbg.as<uint32_t>() : nullptr;
		fa.stats = ctx->stats.as<unsigned long long>() + 8; fa.error = ctx->poolerr.as<uint32_t>(); fa.exact = exact ? 1 : 0;
		launch_fill(K, exact, ctx->dP, fa, st);
		CK(cudaMemcpyAsync(hp.data(), probs.p, (size_t)nprob * sizeof(Problem), cudaMemcpyDeviceToHost, st));
		uint32_t ferr = 0; CK(cudaMemcpyAsync(&ferr, ctx->poolerr.p, 4, cudaMemcpyDeviceToHost, st));
		std::vector<uint32_t> w(exact ? BT2_HEF_WORDS : 1); std::vector<uint8_t> wc(stride);
		for(uint32_t i = 0; i < nprob; i++) {
			if(exact) CK(cudaMemcpyAsync(w.data(), dbg.as<uint32_t>() + (size_t)i * BT2_HEF_WORDS, (size_t)BT2_HEF_WORDS * 4, cudaMemcpyDeviceToHost, st));
			CK(cudaMemcpyAsync(wc.data(), cells.as<uint8_t>() + (size_t)i * stride, stride, cudaMemcpyDeviceToHost, st));
			CK(cudaStreamSynchronize(st));
			if(ferr) throw std::string("bt2gpu_test_fill: a band did not fit the fill's lanes");
			const uint32_t pk = prob_K(hp[i]), py = prob_y(hp[i]); const int d0 = hp[i].d0;
			if(pk != K) throw std::string("bt2gpu_test_fill: the kernel did not record its layout");
			// diagonals the kernel completes for this problem: the lanes up to the widest band of its warp (problems 4w .. 4w+3)
			int widthW = 1;
			for(uint32_t q = i & ~3u; q < std::min(nprob, (i & ~3u) + 4u); q++) { int lo, hi; fill_band(ctx->P, rows[q], cols[q], minsc[q], exact, lo, hi); widthW = std::max(widthW, hi - lo + 1); }
			const int ndiag = ((widthW - 1) / (int)pk + 1) * (int)pk;
			if(band) { band[2 * i] = d0; band[2 * i + 1] = ndiag; }
			for(uint32_t r = 0; r < rows[i]; r++) for(uint32_t cc = 0; cc < cols[i]; cc++) {
				const uint64_t o = hef_off[i] + (uint64_t)r * cols[i] + cc;
				const int dd = (int)cc - (int)r - d0;
				const bool in = dd >= 0 && dd < ndiag;
				if(exact) {
					const uint32_t x = in ? w[(size_t)r * (BT2_LANES * pk) + dd] : 0u;
					H[o] = (uint8_t)x; E[o] = (uint8_t)(x >> 8); F[o] = (uint8_t)(x >> 16);
				}
				flags[o] = in ? wc[cell_index(r, cc, d0, pk, py)] : (uint8_t)0x80;
			}
			ncand[i] = hp[i].ncand; lrmax[i] = hp[i].lrmax;
			for(uint32_t k = 0; k < hp[i].ncand && k < 256; k++) cand[256 * (size_t)i + k] = hp[i].cand[k];
		}
		CK(cudaGetLastError());
	})
}

int bt2gpu_get_create_info(const bt2gpu_ctx* ctx, double ms[4], uint64_t bytes[4]) {
	if(!ctx || !ms || !bytes) return -1;
	for(int i = 0; i < 4; i++) { ms[i] = ctx->create_ms[i]; bytes[i] = ctx->create_bytes[i]; }
	return 0;
}
int bt2gpu_get_counters(bt2gpu_ctx* ctx, bt2gpu_counters* c) { if(!ctx || !c) return -1; *c = ctx->ctr; return 0; }
int bt2gpu_reset_counters(bt2gpu_ctx* ctx) { if(!ctx) return -1; memset(&ctx->ctr, 0, sizeof(ctx->ctr)); return 0; }
void* bt2gpu_host_alloc(size_t bytes) { void* p = nullptr; if(cudaMallocHost(&p, bytes ? bytes : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; } return p; }
void bt2gpu_host_free(void* p) { if(p && cudaFreeHost(p) != cudaSuccess) cudaGetLastError(); }
int bt2gpu_max_rows(void) { return BT2GPU_MAX_ROWS; }
int bt2gpu_max_cols(void) { return BT2GPU_MAX_COLS; }
const char* bt2gpu_last_error(const bt2gpu_ctx* ctx) { return ctx ? ctx->err.c_str() : "null context"; }

int bt2gpu_fill_noalloc(bt2gpu_ctx* ctx, void* stream, uint32_t nprob, const uint32_t* d_seq_off, const uint8_t* d_seq, const uint8_t* d_qual,
                        const uint16_t* d_rows, const uint16_t* d_cols, const uint32_t* d_rf_off, const uint8_t* d_rf, const int32_t* d_minsc,
                        const uint64_t* d_flags_off, uint8_t* d_flags, uint16_t* d_cand, uint32_t* d_ncand, int32_t* d_lrmax) {
	API_GUARD({
		if(nprob > BT2GPU_NOALLOC_MAX_PROBLEMS) {
			char b[160]; snprintf(b, sizeof(b), "bt2gpu_fill_noalloc: %u problems, the static scratch holds at most %d", nprob, BT2GPU_NOALLOC_MAX_PROBLEMS);
			throw std::string(b);
		}
		if(!d_seq_off || !d_seq || !d_qual || !d_rows || !d_cols || !d_rf_off || !d_rf || !d_minsc || !d_flags_off || !d_flags || !d_cand || !d_ncand || !d_lrmax)
			throw std::string("bt2gpu_fill_noalloc: null argument");
		if(nprob == 0) return 0;
		NoAllocDevice& nd = noalloc_device(ctx->device);   // (resolved by bt2gpu_create already)
		NoAllocScratch* sc = reinterpret_cast<NoAllocScratch*>(nd.ptr);
		cudaStream_t st = reinterpret_cast<cudaStream_t>(stream);
		// all calls on a device share the scratch: a call waits (on the device, stream-ordered) for the previous call's kernels, whatever
		// stream they ran on.  Not under stream capture, where an event recorded outside the capture cannot be waited for.
		cudaStreamCaptureStatus cap = cudaStreamCaptureStatusNone;
		if(cudaStreamIsCapturing(st, &cap) != cudaSuccess) { cudaGetLastError(); cap = cudaStreamCaptureStatusNone; }
		std::lock_guard<std::mutex> lk(g_na_mu);
		if(cap == cudaStreamCaptureStatusNone && nd.used) CK(cudaStreamWaitEvent(st, nd.done, 0));
		k_na_setup<<<nblk(std::max(nprob, 32u), 128), 128, 0, st>>>(sc, nprob, d_seq_off, d_rows, d_cols, d_rf_off, d_rf, d_minsc);
		FillArgs fa; fa.probs = sc->probs; fa.nprob = nprob; fa.states = sc->states; fa.seq = d_seq; fa.qual = d_qual; fa.cells = sc->cells; fa.cell_stride = BT2_NA_STRIDE;
		fa.pcache = nullptr; fa.hef_dbg = nullptr; fa.stats = sc->stats; fa.error = sc->error; fa.exact = 0;
		launch_fill(BT2_NA_K, false, ctx->dP, fa, st);   // the production kernel, banded per problem; K = 16 holds any band
		k_na_export<<<nprob, 256, 0, st>>>(ctx->dP, nprob, sc, d_flags_off, d_flags, d_cand, d_ncand, d_lrmax);
		CK(cudaPeekAtLastError());
		if(cap == cudaStreamCaptureStatusNone) { CK(cudaEventRecord(nd.done, st)); nd.used = true; }
	})
}

void bt2gpu_destroy(bt2gpu_ctx* c) {
	if(!c) return;
	if(c->worker.joinable()) c->worker.join();   // (a batch still in flight is finished first)
	cudaSetDevice(c->device);
	cudaDeviceSynchronize();
	for(void* p : c->resident) cudaFree(p);
	if(c->dP) cudaFree(c->dP);
	DBuf* all[] = { &c->probs, &c->cells, &c->pcache, &c->pk, &c->nm,
		&c->states, &c->scratch, &c->hits, &c->seedtasks, &c->exttasks, &c->lists[0], &c->lists[1], &c->lists[2], &c->lists[3], &c->lists[4], &c->lists[5], &c->lists[6], &c->counters, &c->stats,
		&c->p_sav, &c->p_sorted, &c->p_rank, &c->p_sp2, &c->p_exr, &c->p_exrfw, &c->p_nsp, &c->p_nsmall, &c->p_nexr, &c->p_nelt, &c->p_nadd,
		&c->s_ovr, &c->s_ovp, &c->s_ovv, &c->s_sr, &c->s_sv, &c->s_cv, &c->s_rn, &c->s_rcur, &c->s_rth, &c->s_rfl, &c->s_rco, &c->s_rcn, &c->s_rns, &c->s_mass, &c->s_elim,
		&c->chunks, &c->sats, &c->alnp, &c->editp, &c->pooltops, &c->poolerr, &c->nal, &c->ned, &c->albase, &c->edbase, &c->cubtmp };
	for(DBuf* b : all) b->release();
	for(int k = 0; k < 2; k++) {
		DBuf* io[] = { &c->in[k].seq, &c->in[k].qual, &c->in[k].seq_off, &c->in[k].rng_seed, &c->in[k].minsc, &c->in[k].nceil, &c->in[k].interval,
			&c->out[k].o_rr, &c->out[k].o_alns, &c->out[k].o_edits };
		for(DBuf* b : io) b->release();
		cudaEvent_t ev[] = { c->h2dspan[k].a, c->h2dspan[k].b, c->d2hspan[k].a, c->d2hspan[k].b };
		for(cudaEvent_t e : ev) if(e) cudaEventDestroy(e);
	}
	if(c->h2d) cudaStreamDestroy(c->h2d);
	if(c->d2h) cudaStreamDestroy(c->d2h);
	if(c->hbox) cudaFreeHost(const_cast<HostBox*>(c->hbox));
	for(int k = 0; k < 2; k++) { c->hset[k].rr.release(); c->hset[k].alns.release(); c->hset[k].edits.release(); }
	for(cudaEvent_t e : c->evpool) cudaEventDestroy(e);
	if(c->stream) cudaStreamDestroy(c->stream);
	if(c->stream2) cudaStreamDestroy(c->stream2);
	for(cudaEvent_t e : { c->evFork, c->evJoin, c->evFill[0], c->evFill[1] }) if(e) cudaEventDestroy(e);
	delete c;
}

} // extern "C"
