// Thread-per-read / thread-per-problem kernels around the two warp-cooperative ones:
// batch setup, round selection, the glue (dev_prio.cuh), the extension step
// (dev_step.cuh), the post-fill stage (dev_bt.cuh) and result packing (dev_gather.cuh).
// Work lists are compact arrays of read ids appended to with one atomic per entry.
#pragma once
#include "dev_common.cuh"
#include "dev_prio.cuh"
#include "dev_step.cuh"
#include "dev_bt.cuh"
#include "dev_gather.cuh"

namespace bt2 {

struct BatchArrays {
	uint32_t n, seq_base;      // seq_off[i] - seq_base indexes seq/qual
	const uint32_t* seq_off; const uint8_t* seq; const uint8_t* qual;
	const uint32_t* rng_seed; const int32_t* minsc; const int32_t* nceil; const int32_t* interval;
};

// What the host reads between phases: written by tiny kernels straight into mapped pinned host memory, so the
// compute stream never queues a transfer behind the bulk copies of the neighbouring sub-batches (copy engines).
struct HostBox { uint32_t counters[16]; uint32_t last[4]; uint32_t perr; uint32_t pad[3]; unsigned long long stats[32]; };

__global__ void k_zero(uint32_t* cnt, uint32_t mask) { if(threadIdx.x < 16 && ((mask >> threadIdx.x) & 1u)) cnt[threadIdx.x] = 0; }
__global__ void k_reset(uint32_t* cnt, uint32_t* pooltops, uint32_t* poolerr, unsigned long long* stats) {
	const uint32_t t = threadIdx.x;
	if(t < 16) cnt[t] = 0;
	if(t < 4) pooltops[t] = 0;
	if(t == 0) *poolerr = 0;
	if(t < 32) stats[t] = 0;
}
__global__ void k_clear16(uint4* p, size_t n16) { // zero n16 16-byte words
	for(size_t i = (size_t)blockIdx.x * blockDim.x + threadIdx.x; i < n16; i += (size_t)gridDim.x * blockDim.x) p[i] = make_uint4(0, 0, 0, 0);
}
__global__ void k_publish_counters(HostBox* box, const uint32_t* cnt) { if(threadIdx.x < 16) box->counters[threadIdx.x] = cnt[threadIdx.x]; }
__global__ void k_publish_tail(HostBox* box, const uint32_t* albase, const uint32_t* nal, const uint32_t* edbase, const uint32_t* ned, uint32_t n, const uint32_t* poolerr) {
	if(threadIdx.x == 0) { box->last[0] = albase[n - 1]; box->last[1] = nal[n - 1]; box->last[2] = edbase[n - 1]; box->last[3] = ned[n - 1]; box->perr = *poolerr; }
}
__global__ void k_publish_stats(HostBox* box, const unsigned long long* stats) { if(threadIdx.x < 32) box->stats[threadIdx.x] = stats[threadIdx.x]; }

__device__ __forceinline__ void list_append(uint32_t* list, uint32_t* cnt, uint32_t v) { list[atomicAdd(cnt, 1u)] = v; }

__global__ void k_begin(const BatchArrays b, ReadState* states, uint32_t* list, uint32_t* cnt) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= b.n) return;
	ReadState rs;
	const uint32_t so = b.seq_off[i];
	read_begin(rs, so - b.seq_base, b.seq_off[i + 1] - so, b.rng_seed[i], b.minsc[i], b.nceil[i], b.interval[i]);
	states[i] = rs;
	if(rs.phase != RS_FINISHED) list_append(list, cnt, i);
}

// 2-bit copy of every read (forward strand; N -> 0 plus a bit in the N mask) for the seed kernels: a seed's key is then a
// funnel shift instead of a loop of byte loads.  Base j: bits 2 (j % 16) of word j / 16; N mask bit j % 32 of word j / 32.
__global__ void k_pack_reads(const BatchArrays b, const ReadState* __restrict__ states, uint32_t* __restrict__ pk, uint32_t* __restrict__ nm) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= b.n) return;
	uint32_t* po = pk + (size_t)i * BT2_PK_WORDS; uint32_t* no = nm + (size_t)i * BT2_NM_WORDS;
	const uint32_t len = states[i].rdlen <= BT2GPU_MAX_ROWS ? states[i].rdlen : 0u;   // (longer reads never reach the seed search)
	const uint8_t* s = b.seq + states[i].seq_off;
	uint32_t nw = 0;
	for(uint32_t w = 0; w < BT2_PK_WORDS; w++) {
		uint32_t v = 0;
		for(uint32_t j = w * 16u; j < w * 16u + 16u && j < len; j++) {
			const uint32_t c = s[j];
			if(c > 3u) nw |= 1u << (j & 31u); else v |= c << ((j & 15u) * 2u);
		}
		po[w] = v;
		if(w & 1u) { no[w >> 1] = nw; nw = 0; }
	}
}

// reads that want seed round `round` go to `out`; the others are finished
__global__ void k_round_select(const DevParams* __restrict__ Pp, ReadState* states, const uint32_t* in, uint32_t nin,
                               uint32_t round, uint32_t* out, uint32_t* cnt) {
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if(j >= nin) return;
	ReadState& rs = states[in[j]];
	uint32_t off;
	if(!round_wanted(*Pp, rs, round, off)) { rs.phase = RS_FINISHED; return; }
	rs.round = (uint8_t)round; rs.roundsRun++;
	rs.phase = RS_EXTEND;
	rs.nDpFail = 0; rs.passStarted = 0; rs.cur_i = 0; rs.first = 1; rs.sat_n = 0; rs.neltLeft = 0; // extendSeeds entry (:310-312)
	list_append(out, cnt, in[j]);
}

struct PrioArrays { // per active read, W entries each
	uint16_t* sav; uint8_t* sorted; uint16_t* rank; uint16_t* sp2; ExRange16* exr; uint8_t* exrfw;
	uint32_t* nsp; uint32_t* nsmall; uint32_t* nexr; uint64_t* nelt; uint32_t* nelt_added; // [nactive]
};

__device__ __forceinline__ void prio_work(const PrioArrays& pa, uint32_t ai, uint32_t W, PrioWork& w) {
	const size_t o = (size_t)ai * W;
	w.sav = pa.sav + o; w.sorted = pa.sorted + o; w.rank = pa.rank + o; w.sp2 = pa.sp2 + o; w.exr = pa.exr + o; w.exrfw = pa.exrfw + o;
}

#ifndef BT2_STEP_BLOCKS_PER_SM
#define BT2_STEP_BLOCKS_PER_SM 5
#endif
#ifndef BT2_PRIO_BLOCKS_PER_SM
#define BT2_PRIO_BLOCKS_PER_SM 6
#endif
#ifndef BT2_PRIO_CACHE
#define BT2_PRIO_CACHE 48   // seed hits (both strands) a thread copies: 2 x 9 for 150 bp --sensitive, 2 x 19 for --very-sensitive
#endif
static_assert(sizeof(SeedHit) == 32, "SeedHit is copied as two uint4");
__global__ void __launch_bounds__(128, BT2_PRIO_BLOCKS_PER_SM) k_prio1(const DevParams* __restrict__ Pp, const Pools pools, ReadState* states, const uint32_t* active, uint32_t nactive,
                        const SeedHit* hits, uint32_t W, uint32_t maxOffs, const PrioArrays pa,
                        uint32_t* ext_list, uint32_t* ext_cnt, uint32_t* samp_list, uint32_t* samp_cnt) {
	const uint32_t ai = blockIdx.x * blockDim.x + threadIdx.x;
	if(ai >= nactive) return;
	const DevParams& P = *Pp;
	const uint32_t read = active[ai];
	ReadState rs = states[read];
	uint32_t offset = 0; round_wanted(P, rs, rs.round, offset);
	const SeedGeom g = seed_geom(P, rs.rdlen, offset, (uint32_t)rs.interval, maxOffs);
	PrioWork w; prio_work(pa, ai, W, w);
	// The read's seed hits are scanned O(n^2) times (duplicate keys, rankSeedHits, the sort): a thread-local copy turns those scans from
	// 32-byte records 32 W bytes apart across the warp into local-memory accesses, which the hardware interleaves across the lanes.
	const SeedHit* hp = hits + (size_t)ai * W;
	SeedHit lh[BT2_PRIO_CACHE];
	if(W <= BT2_PRIO_CACHE) {
		const uint4* src = reinterpret_cast<const uint4*>(hp); uint4* dst = reinterpret_cast<uint4*>(lh);
		for(uint32_t k = 0; k < 2u * W; k++) dst[k] = src[k];
		hp = lh;
	}
	// (the same for the stage's index arrays measured no further gain: they stay in global memory, where the row sampler reads them)
	const bool sample = prio_stage1(P, pools, rs, hp, g, w);
	if(rs.numRanges == 0 || rs.nonzTot == 0) {
		rs.flags |= BT2GPU_RF_NO_SEED_HITS; rs.phase = RS_FINISHED;
	} else {
		pa.nsp[ai] = w.nsp; pa.nsmall[ai] = w.nsmall; pa.nexr[ai] = w.nexr; pa.nelt[ai] = w.nelt; pa.nelt_added[ai] = w.nelt_added;
		if(sample) list_append(samp_list, samp_cnt, ai);
		list_append(ext_list, ext_cnt, read);
	}
	states[read] = rs;
}

struct SampleArrays { // per sampling slot
	uint32_t* hkey; uint16_t* hval;                                  // [slots * BT2_SAMPLE_HASH], keys cleared to BT2_SAMPLE_EMPTY before the launch
	uint16_t* seen_rng; uint16_t* seen_val; uint16_t* conv_val;      // [slots * MI]
	uint16_t* r_n; uint16_t* r_cur; uint16_t* r_thresh; uint8_t* r_flags; uint32_t* r_conv_off; uint16_t* r_conv_n; uint16_t* r_nseen;
	double* mass; uint8_t* elim; // [slots * W]
};
#define BT2_SAMPLE_MI (BT2GPU_MAX_ITER + 8)

__global__ void k_prio2(const DevParams* __restrict__ Pp, const Pools pools, ReadState* states, const uint32_t* active,
                        const SeedHit* hits, uint32_t W, uint32_t maxOffs, const PrioArrays pa, const SampleArrays sa,
                        const uint32_t* samp_list, uint32_t base, uint32_t count) {
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if(j >= count) return;
	const DevParams& P = *Pp;
	const uint32_t ai = samp_list[base + j];
	const uint32_t read = active[ai];
	ReadState rs = states[read];
	uint32_t offset = 0; round_wanted(P, rs, rs.round, offset);
	const SeedGeom g = seed_geom(P, rs.rdlen, offset, (uint32_t)rs.interval, maxOffs);
	PrioWork w; prio_work(pa, ai, W, w);
	w.nsp = pa.nsp[ai]; w.nsmall = pa.nsmall[ai]; w.nexr = pa.nexr[ai]; w.nelt = pa.nelt[ai]; w.nelt_added = pa.nelt_added[ai];
	SampleWork s;
	const size_t oh = (size_t)j * BT2_SAMPLE_HASH, o1 = (size_t)j * BT2_SAMPLE_MI, ow = (size_t)j * W;
	s.hkey = sa.hkey + oh; s.hval = sa.hval + oh;
	s.seen_rng = sa.seen_rng + o1; s.seen_val = sa.seen_val + o1; s.conv_val = sa.conv_val + o1;
	s.r_n = sa.r_n + ow; s.r_cur = sa.r_cur + ow; s.r_thresh = sa.r_thresh + ow; s.r_flags = sa.r_flags + ow;
	s.r_conv_off = sa.r_conv_off + ow; s.r_conv_n = sa.r_conv_n + ow; s.r_nseen = sa.r_nseen + ow; s.mass = sa.mass + ow; s.elim = sa.elim + ow;
	prio_stage2(P, pools, rs, hits + (size_t)ai * W, g, w, s);
	states[read] = rs;
}

// one extension step per read of the chunk; problem slot = position in the chunk
__global__ void __launch_bounds__(128, BT2_STEP_BLOCKS_PER_SM) k_step(const DevParams* __restrict__ Pp, const DevIndex ix, const DevRef rf, const Pools pools, ReadState* states,
                       const uint32_t* list, uint32_t base, uint32_t count, Problem* probs,
                       uint32_t* roundend_list, uint32_t* roundend_cnt) {
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if(j >= count) return;
	const uint32_t read = list[base + j];
	ReadState rs = states[read];
	Problem& pb = probs[j];
	pb.read = read;
	if(advance(*Pp, ix, rf, pools, rs, pb) == ADV_ROUND_END) {
		pb.read = BT2_NIL;
		rs.phase = RS_ROUND_DONE;
		list_append(roundend_list, roundend_cnt, read);
	}
	states[read] = rs;
}

#define BT2_BT_SPIN_LIMIT (1u << 22)
#ifndef BT2_BT_POST_MIN
#define BT2_BT_POST_MIN 16  // lanes that must want the rare paths before the warp takes them while others still walk
#endif                      // (k_bt per 1 M reads with POST_MIN / DONE_MIN = 4 / 4: 13.4 ms, 8 / 8: 11.7, 16 / 16: 11.2, 24 / 24: 11.2, 16 / 4: 13.9 --
#ifndef BT2_BT_DONE_MIN     //  the fetch of a new problem is a chain of dependent global loads, best paid by many lanes at once)
#define BT2_BT_DONE_MIN 16
#endif
#if BT2GPU_MAX_ROWS > 160      // the long-read build: a lane's copy of its problem's characters would be ~1 KB of shared memory;
#define BT2_BT_PC_GLOBAL 1     // it reads them where K_fill left them instead (two L1-cached loads per backtrace cell)
#else
#define BT2_BT_PC_GLOBAL 0
#endif
#ifndef BT2_BT_THREADS
#define BT2_BT_THREADS 128
#endif
#ifndef BT2_BT_BLOCKS_PER_SM
#define BT2_BT_BLOCKS_PER_SM 6
#endif

// One lane per problem, persistent: a lane that finishes its problem takes the next one.
// The lanes of a warp are in different phases (candidate pick, backtrace cell, alignment
// post-processing, fetch); the cheap and frequent ones (candidate, cell: BtMachine::hot) run
// every iteration, the long and rare ones (post-processing, fetch + begin) are held back until
// a quarter of the warp wants them or nothing else is left, so that they are executed
// for several lanes at once instead of stalling 31 lanes for one.
// Shared memory holds each lane's copy of its problem's characters (word j of lane t at
// [j * BT2_BT_THREADS + t]: conflict-free whatever j the lanes ask for) and its two cell chunks.
__global__ void __launch_bounds__(BT2_BT_THREADS, BT2_BT_BLOCKS_PER_SM) k_bt(const __grid_constant__ BtCtx cx, uint32_t count,
                     uint32_t* task, uint32_t* next_list, uint32_t* next_cnt, uint32_t* roundend_list, uint32_t* roundend_cnt) {
#if !BT2_BT_PC_GLOBAL
	__shared__ uint32_t pcs[BT2_PC_WORDS * BT2_BT_THREADS];
#endif
	__shared__ __align__(16) uint8_t chunks[4 * 16 * BT2_BT_THREADS];   // two cell chunks and two candidate groups per lane
	BtMachine m;
#if BT2_BT_PC_GLOBAL
	m.pc = cx.pcache; m.pcs = 1;
#else
	m.pc = pcs + threadIdx.x; m.pcs = BT2_BT_THREADS;
#endif
	m.cbuf = chunks + 16 * threadIdx.x; m.cstride = 16 * BT2_BT_THREADS; m.cbuf_sa = (uint32_t)__cvta_generic_to_shared(m.cbuf);
	m.phase = BP_DONE; m.result = BT_CONTINUE; m.ct = CT_H;
	bool have = false, drained = false;
	uint32_t spins = 0;   // iterations since this lane last took a problem: a problem needs far fewer (cells visited + candidates)
	while(true) {
		if(__any_sync(0xffffffffu, ++spins > BT2_BT_SPIN_LIMIT)) {   // never expected: fail loudly (error bit 32) instead of hanging the stream
#ifdef BT2_BT_DEBUG
			if(have && spins > BT2_BT_SPIN_LIMIT)
				printf("bt2gpu k_bt: lane stuck: slot %u read %u phase %d row %d col %d cural %u ncand %u ct %d cidx %u\n", m.slot, m.read, m.phase, m.row, m.col, m.cural, m.ncand, m.ct, m.cidx);
#endif
			set_error(cx.pools, 32);
			break;
		}
		if(m.phase == BP_CAND || m.phase == BP_STEP) m.hot(cx);
		const uint32_t hot = __ballot_sync(0xffffffffu, m.phase == BP_CAND || m.phase == BP_STEP);
		const uint32_t wpost = __ballot_sync(0xffffffffu, m.phase == BP_POST);
		if(wpost && (hot == 0 || __popc(wpost) >= BT2_BT_POST_MIN)) { if(m.phase == BP_POST) m.post(cx); }
		const uint32_t wdone = __ballot_sync(0xffffffffu, m.phase == BP_DONE && !drained);
		if(wdone && (hot == 0 || __popc(wdone) >= BT2_BT_DONE_MIN)) {
			if(m.phase == BP_DONE && !drained) {
				if(have) {
					m.finish(cx);
					if(m.result == BT_ROUND_END) { cx.states[m.read].phase = RS_ROUND_DONE; list_append(roundend_list, roundend_cnt, m.read); }
					else list_append(next_list, next_cnt, m.read);
					have = false;
				}
				while(!have) {
					const uint32_t j = atomicAdd(task, 1u);
					if(j >= count) { drained = true; break; }
					const uint32_t rd = cx.probs[j].read;
					if(rd == BT2_NIL) continue;
					m.slot = j; m.read = rd;
#if BT2_BT_PC_GLOBAL
					m.pc = cx.pcache + (size_t)j * BT2_PC_WORDS;
#else
					const uint4* src = reinterpret_cast<const uint4*>(cx.pcache + (size_t)j * BT2_PC_WORDS);
					#pragma unroll
					for(int q = 0; q < BT2_PC_WORDS / 4; q++) {
						const uint4 w = src[q];
						pcs[(4 * q + 0) * BT2_BT_THREADS + threadIdx.x] = w.x; pcs[(4 * q + 1) * BT2_BT_THREADS + threadIdx.x] = w.y;
						pcs[(4 * q + 2) * BT2_BT_THREADS + threadIdx.x] = w.z; pcs[(4 * q + 3) * BT2_BT_THREADS + threadIdx.x] = w.w;
					}
#endif
					m.begin(cx);
					have = true; spins = 0;
				}
			}
		}
		if(__ballot_sync(0xffffffffu, !drained || have) == 0) break;
	}
}

// reads whose round is over: another seed round, or finished
__global__ void k_round_end(const DevParams* __restrict__ Pp, ReadState* states, const uint32_t* in, uint32_t nin,
                            uint32_t* out, uint32_t* cnt) {
	const uint32_t j = blockIdx.x * blockDim.x + threadIdx.x;
	if(j >= nin) return;
	ReadState& rs = states[in[j]];
	if(next_round_wanted(*Pp, rs)) list_append(out, cnt, in[j]);
	else rs.phase = RS_FINISHED;
}

__global__ void k_count(const ReadState* states, const Pools pools, uint32_t n, uint32_t* nal, uint32_t* ned, unsigned long long* wstats) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	unsigned long long v[6] = { 0, 0, 0, 0, 0, 0 };
	if(i < n) {
		const ReadState& rs = states[i];
		nal[i] = rs.n_alns; ned[i] = read_edit_count(rs, pools);
		v[0] = rs.saRows; v[1] = rs.lfResolve; v[2] = rs.btAttempts; v[3] = rs.btSteps; v[4] = rs.n_alns; v[5] = rs.nExDps;
	}
	// work counters for the roofline arithmetic (warp-aggregated)
	#pragma unroll
	for(int k = 0; k < 6; k++) {
		unsigned long long x = v[k];
		for(int o = 16; o > 0; o >>= 1) x += __shfl_xor_sync(0xffffffffu, x, o);
		if((threadIdx.x & 31) == 0 && x) atomicAdd(&wstats[k], x);
	}
}

__global__ void k_gather(const ReadState* states, const Pools pools, const uint32_t* scratch, uint32_t n,
                         const uint32_t* aln_base, const uint32_t* edit_base, uint32_t aln_glob, uint32_t edit_glob,
                         bt2gpu_read_result* rr, bt2gpu_aln* alns, bt2gpu_edit* edits) {
	const uint32_t i = blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= n) return;
	bt2gpu_read_result r;
	gather_read(states[i], pools, scratch + (size_t)i * BT2_SCRATCH_EDITS, aln_base[i], edit_base[i], aln_glob, edit_glob, r, alns, edits);
	rr[i] = r;
}

} // namespace bt2
