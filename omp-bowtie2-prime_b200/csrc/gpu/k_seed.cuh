// K_seed: exact FM backward search of every seed of every active read, followed by
// the left extension of the hits (nlex).
//   SeedAligner::prepareSeed / instantiateSeed / instantiateExact aligner_seed.cpp:236-341
//   SeedAligner::searchSeedBi / startSearchSeedBi / nextLocsBi   aligner_seed.cpp:732-957
//   MultiSeedAligner::extend                                     aligner_seed.cpp:603-672
//   Ebwt::countBt2SideEx / mapBiLFEx / mapLF1 / ftabHi / ftabLo  bt2_idx.h:1940-1972, 2474-2564, 1430-1607
//
// Left extension of a hit whose range is a single row is not walked through the FM index: with the
// device's full suffix array (k_index.cuh) the row's text offset is one gather, and "the BWT
// character of the row equals the read character, step to LF(row)" is the same statement as "the
// text character before the offset equals the read character, step the offset down" (the row of
// offset 0 is the '$' row).  k_seed_ext1 does that comparison against the 2-bit reference, one
// thread per hit: no dependent random reads.  Ranges of several rows keep the LF walk.
//
// With the full suffix array and the joined text resident (k_index.cuh), the exact search itself needs no LF walk
// either: the rows whose suffix starts with the seed are one interval of the suffix array, the ftab gives the interval
// of the seed's FIRST ftabChars bases (the table is symmetric: "k-mer -> rows whose suffix starts with it"), and two
// binary searches over that interval compare the rest of the seed with the text behind each row's offset (one 4-byte
// gather from the suffix array + 12 bytes of text per probe instead of a dependent 128-byte side per base).  The
// interval is the same (top, bot) the reference's backward search ends with.  Seeds shorter than the ftab, indexes
// without the device suffix array (>= 2^32 rows) and left extension of multi-row ranges keep the LF walk below.
//
// Three kernels.  k_seed_init (one thread per seed): validity (no N), the 2-bit key,
// the suffix-array search or (fallback) the ftab jump over the last ftabChars bases; seeds that still need LF steps are
// appended to a task list.  k_seed_walk: a group of 8 lanes owns one task.  One LF
// step = the group reads the 128-byte side(s) of top and bot with one 16-byte load per
// lane (a full line per group, four lines per warp instruction), each of the six BWT
// lanes popcounts its 64 bases, counts are summed with three xor-shuffles, and the
// occ[] words held by lanes 6-7 are broadcast.  Groups are persistent: when a seed is
// finished the group takes the next task from a global counter, so the four groups of
// a warp stay in the same "one LF step per iteration" loop although their seeds are in
// different phases (search vs. extend).  Row arithmetic is 32-bit when the index is
// shorter than 2^32 (template TRow), which removes the 64-bit divisions by 384.
#pragma once
#include "dev_common.cuh"
#include "dev_prio.cuh"
#include "dev_step.cuh"

namespace bt2 {

struct SeedTask {          // a seed that survived the ftab jump
	uint32_t hit;          // index into hits[]
	uint32_t seq_off;      // read's bases in the batch
	uint16_t rdlen;
	int16_t  pos;          // next strand-buffer position to consume (descending)
	int16_t  seed_lo;      // strand-buffer offset of the seed's first base
	uint16_t rc;
};

struct SeedKernelArgs {
	const uint32_t*  active;    // [nactive] read ids
	uint32_t         nactive;
	uint32_t         W, maxOffs; // W = 2*maxOffs hit slots per active read
	const ReadState* states;
	const uint8_t*   seq;       // batch sequences, codes 0..4
	SeedHit*         hits;      // [nactive*W]
	SeedTask*        tasks;     // [nactive*W]
	uint32_t*        ntasks;    // tasks appended by k_seed_init
	SeedTask*        xtasks;    // [nactive*W] single-row hits waiting for their left extension (k_seed_ext1)
	uint32_t*        nxtasks;
	uint32_t*        xcounter;  // task counter of k_seed_ext1
	const uint32_t*  pk;        // [n * BT2_PK_WORDS] 2-bit packed reads, by read id (k_pack_reads)
	const uint32_t*  nm;        // [n * BT2_NM_WORDS] their N masks
	const uint8_t*   text;      // 2-bit joined text (DevRef::buf), or null: no text-compare extension
	uint32_t*        counter;   // task counter of k_seed_walk
	unsigned long long* stats;  // [0] seeds searched [1] LF steps search [2] LF steps extend [3] sides loaded [4] text-compare extension steps [5] suffix-array probes
};

__device__ __forceinline__ uint4 ldg_u4(const uint8_t* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

__device__ __forceinline__ uint64_t ftab_hi(const DevIndex& ix, uint64_t i) {
	const uint64_t v = __ldg(ix.ftab + i);
	if(v <= ix.len) return v;
	return __ldg(ix.eftab + (v ^ OFF_MASK_U64) * 2 + 1);
}
__device__ __forceinline__ uint64_t ftab_lo(const DevIndex& ix, uint64_t i) {
	const uint64_t v = __ldg(ix.ftab + i);
	if(v <= ix.len) return v;
	return __ldg(ix.eftab + (v ^ OFF_MASK_U64) * 2);
}
__device__ __forceinline__ uint64_t fchr_at(const DevIndex& ix, int c) { // no dynamic indexing of the kernel parameter
	return c == 0 ? ix.fchr[0] : c == 1 ? ix.fchr[1] : c == 2 ? ix.fchr[2] : c == 3 ? ix.fchr[3] : ix.fchr[4];
}

// up to 32 bases of the 2-bit text starting at base x, first base in the lowest bits (16 bases per 32-bit word)
__device__ __forceinline__ uint64_t text_bases32(const uint32_t* __restrict__ tw, uint64_t x) {
	const uint64_t w = x >> 4; const uint32_t sh = (uint32_t)(x & 15u) * 2u;
	const uint32_t w0 = __ldg(tw + w), w1 = __ldg(tw + w + 1), w2 = __ldg(tw + w + 2);
	return (uint64_t)__funnelshift_r(w0, w1, sh) | ((uint64_t)__funnelshift_r(w1, w2, sh) << 32);
}
// reverses the order of the 32 base pairs of v: the first base becomes the most significant one
__device__ __forceinline__ uint64_t rev_pairs64(uint64_t v) {
	v = __brevll(v);
	return ((v & 0x5555555555555555ull) << 1) | ((v >> 1) & 0x5555555555555555ull);
}
// Compares the T bases that follow the first `fl` bases of row r's suffix with the pattern tail `pv` (T bases, first base
// most significant): -1 suffix < pattern, 0 suffix starts with the pattern, +1 suffix > pattern.  A suffix that ends
// inside the pattern is GREATER: the index sorts the end of the text after every base ("ACG$" follows "ACGT...").
__device__ __forceinline__ int sa_cmp(const DevIndex& ix, const uint32_t* __restrict__ tw, uint32_t r, uint32_t fl, uint32_t T, uint64_t pv) {
	const uint64_t x = (uint64_t)__ldg(ix.sa32 + r) + fl;
	if(x >= ix.len) return 1;                          // a suffix of at most fl bases
	const uint64_t avail = ix.len - x;
	const uint32_t m = avail < (uint64_t)T ? (uint32_t)avail : T;
	const uint64_t sv = rev_pairs64(text_bases32(tw, x)) >> (64u - 2u * m);
	const uint64_t pm = pv >> (2u * (T - m));
	if(sv < pm) return -1;
	if(sv > pm) return 1;
	return m < T ? 1 : 0;
}

__global__ void __launch_bounds__(256) k_seed_init(const DevIndex ix, const DevParams* __restrict__ Pp, const SeedKernelArgs a) {
	const DevParams& P = *Pp;
	const uint64_t tk = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	const bool in_range = tk < (uint64_t)a.nactive * a.W;
	bool searched = false, live = false, livex = false;
	uint32_t probes = 0;
	SeedTask t; t.hit = 0; t.seq_off = 0; t.rdlen = 0; t.pos = 0; t.seed_lo = 0; t.rc = 0;
	if(in_range) {
		const uint32_t ai = (uint32_t)(tk / a.W), h = (uint32_t)(tk - (uint64_t)ai * a.W);
		const uint32_t s = h / a.maxOffs, i = h - s * a.maxOffs;
		const ReadState& rs = a.states[a.active[ai]];
		const uint32_t rdlen = rs.rdlen; const bool rc = (s == 1);
		SeedHit* out = a.hits + tk;
		uint32_t offset = 0; round_wanted(P, rs, rs.round, offset);
		const SeedGeom g = seed_geom(P, rdlen, offset, (uint32_t)rs.interval, a.maxOffs);
		bool ok = i < g.numOffs && !((s == 0 && P.nofw) || (s == 1 && P.norc));
		const int L = (int)g.seqLen;
		const int depth = (int)(i * g.per + g.off);
		const int seed_lo = rc ? ((int)rdlen - depth - L) : depth;
		if(ok && (seed_lo < 0 || seed_lo + L > (int)rdlen)) ok = false;
		uint64_t key = 0, top = 0, bot = 0;
		if(ok) { // InstantiatedSeed::instantiateExact: no N inside the seed; the key from the packed read
			const uint32_t read = a.active[ai];
			const uint32_t* pk = a.pk + (size_t)read * BT2_PK_WORDS; const uint32_t* nm = a.nm + (size_t)read * BT2_NM_WORDS;
			const uint32_t f0 = rc ? rdlen - (uint32_t)seed_lo - (uint32_t)L : (uint32_t)seed_lo;   // the seed's first base on the forward strand
			const uint64_t maskL = L >= 32 ? ~0ull : ((1ull << (2 * L)) - 1ull);
			const uint32_t w = f0 >> 4, sh = (f0 & 15u) * 2u;
			const uint32_t w0 = pk[w], w1 = pk[w + 1], w2 = pk[w + 2];
			const uint64_t v = ((uint64_t)__funnelshift_r(w0, w1, sh) | ((uint64_t)__funnelshift_r(w1, w2, sh) << 32)) & maskL;   // base f0 + j at bits 2j
			const uint32_t nb = __funnelshift_r(nm[f0 >> 5], nm[(f0 >> 5) + 1], f0 & 31u) & (L >= 32 ? ~0u : ((1u << L) - 1u));
			if(nb) ok = false;
			// first strand base most significant; the reverse-complement strand reads the same window backwards, complemented
			key = rc ? (~v & maskL) : (rev_pairs64(v) >> (64 - 2 * L));
		}
		bool hit = false; int pos = 0;
		if(ok) {
			searched = true;
			const bool kt = a.text && ix.ktab && ix.ktabChars < L;        // the device's prefix table (k_index.cuh), else the index's ftab
			const int fl = kt ? ix.ktabChars : ix.ftabChars;
			int step = 0;
			if(a.text && fl > 1 && fl <= L && L - fl <= 32) {
				// suffix-array search: interval of the first fl bases from the table, the other L - fl bases by binary search
				const uint64_t k = key >> (2 * (L - fl));   // the first fl bases
				uint32_t lo = kt ? __ldg(ix.ktab + k) : (uint32_t)ftab_hi(ix, k), hi = kt ? __ldg(ix.ktab + k + 1) : (uint32_t)ftab_lo(ix, k + 1);
				const uint32_t T = (uint32_t)(L - fl);
				if(T > 0 && hi > lo) {
					const uint32_t* tw = reinterpret_cast<const uint32_t*>(a.text);
					const uint64_t pv = key & (T >= 32u ? ~0ull : ((1ull << (2u * T)) - 1ull));   // the last T bases of the seed
					uint32_t l0 = lo, h0 = hi;
					while(l0 < h0) { const uint32_t mid = l0 + ((h0 - l0) >> 1); probes++; if(sa_cmp(ix, tw, mid, (uint32_t)fl, T, pv) < 0) l0 = mid + 1; else h0 = mid; }
					uint32_t l1 = l0, h1 = hi;
					while(l1 < h1) { const uint32_t mid = l1 + ((h1 - l1) >> 1); probes++; if(sa_cmp(ix, tw, mid, (uint32_t)fl, T, pv) <= 0) l1 = mid + 1; else h1 = mid; }
					lo = l0; hi = l1;
				}
				top = lo; bot = hi;
				step = L;
			} else if(fl > 1 && fl <= L) { // startSearchSeedBi: ftab jump over the last ftabChars bases
				const uint64_t k = key & ((1ull << (2 * fl)) - 1ull);   // the last fl bases
				top = ftab_hi(ix, k); bot = ftab_lo(ix, k + 1);
				step = fl;
			} else if(L > 0) {
				const int ch = (int)(key & 3ull);
				top = fchr_at(ix, ch); bot = fchr_at(ix, ch + 1);
				step = 1;
			} else { top = 0; bot = ix.fchr[4]; }
			hit = bot != top;
			pos = seed_lo + L - 1 - step;
		}
		// (top, len) is the range reported when the whole seed was consumed here; k_seed_walk
		// overwrites it with the range at the end of the search and fills nlex / valid
		const bool whole = hit && pos < seed_lo;
		const bool done = !hit || (whole && (!P.doExtend || pos < 0));
		out->top = hit ? top : 0; out->len = hit ? bot - top : 0; out->key = key; out->nlex = 0;
		out->valid = (hit && done) ? 1u : 0u;
		if(hit && !done) {
			// a single-row range that is already complete goes straight to the text-compare extension
			if(whole && bot - top == 1 && a.text) livex = true; else live = true;
			t.hit = (uint32_t)tk; t.seq_off = rs.seq_off; t.rdlen = (uint16_t)rdlen; t.pos = (int16_t)pos; t.seed_lo = (int16_t)seed_lo; t.rc = rc ? 1 : 0;
		}
	}
	// warp-aggregated append + statistics
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t bal = __ballot_sync(0xffffffffu, live);
	uint32_t base = 0;
	if(lane == 0 && bal) base = atomicAdd(a.ntasks, (uint32_t)__popc(bal));
	base = __shfl_sync(0xffffffffu, base, 0);
	if(live) a.tasks[base + __popc(bal & ((1u << lane) - 1))] = t;
	const uint32_t balx = __ballot_sync(0xffffffffu, livex);
	uint32_t basex = 0;
	if(lane == 0 && balx) basex = atomicAdd(a.nxtasks, (uint32_t)__popc(balx));
	basex = __shfl_sync(0xffffffffu, basex, 0);
	if(livex) a.xtasks[basex + __popc(balx & ((1u << lane) - 1))] = t;
	const uint32_t sb = __ballot_sync(0xffffffffu, searched);
	if(lane == 0 && sb) atomicAdd(&a.stats[0], (unsigned long long)__popc(sb));
	#pragma unroll
	for(int o = 16; o > 0; o >>= 1) probes += __shfl_xor_sync(0xffffffffu, probes, o);
	if(lane == 0 && probes) atomicAdd(&a.stats[5], (unsigned long long)probes);
}

// # of code c among the first n (<= 0 .. >= 64) bases held in the four 32-bit words of w (16 bases each)
__device__ __forceinline__ uint32_t count_u4(const uint4& w, uint32_t cmask, int n) {
	uint32_t cnt = 0;
	const uint32_t ws[4] = { w.x, w.y, w.z, w.w };
	#pragma unroll
	for(int j = 0; j < 4; j++) {
		const int nj = n - 16 * j;
		const uint32_t x = ws[j] ^ cmask;
		uint32_t m = ~(x | (x >> 1)) & 0x55555555u;
		if(nj < 16) m = nj > 0 ? (m << (32 - 2 * nj)) : 0u;
		cnt += __popc(m);
	}
	return cnt;
}

// ranks of code c before rows `top` and `bot` (both + occ + fchr), computed by the 8-lane group.
// If c < 0 the character is taken from the BWT at row `top` (Ebwt::rowL) and returned in c.
template<typename TRow>
__device__ __forceinline__ void group_rank2(const DevIndex& ix, uint32_t gmask, uint32_t gbase, uint32_t l8, int& c,
                                            TRow top, TRow bot, TRow& rt, TRow& rb, uint32_t& sides) {
	const TRow side_t = top / (TRow)BT2_SIDE_BWT_LEN, side_b = bot / (TRow)BT2_SIDE_BWT_LEN;
	const uint32_t co_t = (uint32_t)(top - side_t * (TRow)BT2_SIDE_BWT_LEN), co_b = (uint32_t)(bot - side_b * (TRow)BT2_SIDE_BWT_LEN);
	const uint4 wt = ldg_u4(ix.ebwt + (uint64_t)side_t * BT2_SIDE_SZ + 16u * l8);
	uint4 wb = wt;
	if(side_b != side_t) { wb = ldg_u4(ix.ebwt + (uint64_t)side_b * BT2_SIDE_SZ + 16u * l8); sides += 2; } else sides += 1;
	if(c < 0) { // rowL(top): lane co_t/64 holds it, 16 bases per 32-bit word
		const uint32_t wi = (co_t >> 4) & 3u;
		const uint32_t wsel = wi == 0 ? wt.x : wi == 1 ? wt.y : wi == 2 ? wt.z : wt.w;
		c = (int)__shfl_sync(gmask, (wsel >> ((co_t & 15u) * 2)) & 3u, gbase + (co_t >> 6));
	}
	uint32_t ct = 0, cb = 0;
	if(l8 < 6) {
		const uint32_t cmask = 0x55555555u * (uint32_t)c;
		ct = count_u4(wt, cmask, (int)co_t - 64 * (int)l8);
		cb = count_u4(wb, cmask, (int)co_b - 64 * (int)l8);
	}
	uint32_t packed = ct | (cb << 16);
	packed += __shfl_xor_sync(gmask, packed, 1);
	packed += __shfl_xor_sync(gmask, packed, 2);
	packed += __shfl_xor_sync(gmask, packed, 4);
	// occ[c]: lane 6 holds occ[0],occ[1]; lane 7 holds occ[2],occ[3] (u64 each, low word first)
	const uint32_t src = gbase + 6 + ((uint32_t)c >> 1);
	const uint32_t tlo = (c & 1) ? wt.z : wt.x, blo = (c & 1) ? wb.z : wb.x;
	TRow occ_t = (TRow)__shfl_sync(gmask, tlo, src), occ_b = (TRow)__shfl_sync(gmask, blo, src);
	if(sizeof(TRow) == 8) {
		const uint32_t thi = (c & 1) ? wt.w : wt.y, bhi = (c & 1) ? wb.w : wb.y;
		occ_t |= (TRow)((uint64_t)__shfl_sync(gmask, thi, src) << 32);
		occ_b |= (TRow)((uint64_t)__shfl_sync(gmask, bhi, src) << 32);
	}
	TRow cnt_t = (TRow)(packed & 0xffffu), cnt_b = (TRow)(packed >> 16);
	if(c == 0) { // '$' is stored as an A
		const TRow z = (TRow)ix.zOff;
		if(z >= side_t * (TRow)BT2_SIDE_BWT_LEN && z < top) cnt_t--;
		if(z >= side_b * (TRow)BT2_SIDE_BWT_LEN && z < bot) cnt_b--;
	}
	const TRow fc = (TRow)fchr_at(ix, c);
	rt = cnt_t + occ_t + fc;
	rb = cnt_b + occ_b + fc;
}

#define BT2_SEED_GRAB 4

template<typename TRow>
__global__ void __launch_bounds__(256) k_seed_walk(const DevIndex ix, const DevParams* __restrict__ Pp, const SeedKernelArgs a) {
	const bool doExtend = Pp->doExtend != 0;
	const uint32_t lane = threadIdx.x & 31, l8 = lane & 7, gbase = lane & ~7u;
	const uint32_t gmask = 0xffu << gbase;
	const uint32_t ntasks = *a.ntasks;
	const TRow zOff = (TRow)ix.zOff;
	uint32_t task = 0, task_end = 0;
	// per-seed state (uniform within the group)
	bool busy = false, extending = false;
	TRow top = 0, bot = 0;
	int pos = 0, seed_lo = 0;
	uint32_t nlex = 0;
	bool rc = false;
	const uint8_t* rseq = nullptr; uint32_t rdlen = 0;
	SeedHit* out = nullptr;
	uint32_t st_search = 0, st_extend = 0, st_sides = 0;
	while(true) {
		if(!busy) {
			if(task == task_end) {
				uint32_t t0 = 0;
				if(l8 == 0) t0 = atomicAdd(a.counter, (uint32_t)BT2_SEED_GRAB);
				t0 = __shfl_sync(gmask, t0, gbase);
				task = t0; task_end = t0 + BT2_SEED_GRAB;
				if(task_end > ntasks) task_end = ntasks;
				if(task >= ntasks) break;
			}
			const SeedTask t = a.tasks[task++];
			out = a.hits + t.hit;
			rseq = a.seq + t.seq_off; rdlen = t.rdlen; rc = t.rc != 0; pos = t.pos; seed_lo = t.seed_lo;
			top = (TRow)out->top; bot = top + (TRow)out->len;
			busy = true; nlex = 0;
			extending = pos < seed_lo;   // whole seed consumed by the ftab jump: (top,len) stays as k_seed_init wrote it
			continue;
		}
		// ---- one LF step at strand position `pos`: a single rank query per iteration
		int ch;
		{ const uint32_t p = rc ? (rdlen - 1 - (uint32_t)pos) : (uint32_t)pos; const int b = rseq[p]; ch = rc ? (b > 3 ? 4 : 3 - b) : b; }
		bool finished = false, failed = false;
		const bool range1 = (bot - top == 1);
		if(extending && !range1 && ch > 3) {
			// rare: read N while extending a range: any single child of the same size (aligner_seed.cpp:639-655)
			st_extend++;
			const TRow origSz = bot - top;
			int nonz = -1; bool abort_ = false; TRow nt = top, nb = bot;
			for(int j = 0; j < 4; j++) {
				int cj = j; TRow t, b; group_rank2<TRow>(ix, gmask, gbase, l8, cj, top, bot, t, b, st_sides);
				if(b > t) { if(nonz >= 0) { abort_ = true; break; } nonz = j; nt = t; nb = b; }
			}
			if(abort_ || nonz < 0 || nb - nt < origSz) finished = true;
			else { top = nt; bot = nb; nlex++; pos--; if(nlex == 255 || pos < 0) finished = true; }
		} else {
			int qc = range1 ? -1 : ch;          // range of one row: follow the BWT char (mapLF1)
			const bool atZ = range1 && top == zOff;
			TRow t = 0, b = 0;
			if(!atZ) group_rank2<TRow>(ix, gmask, gbase, l8, qc, top, range1 ? top : bot, t, b, st_sides);
			if(!extending) {
				st_search++;
				if(range1) { if(atZ || qc != ch) failed = true; b = t + 1; }
				if(!failed && b == t) failed = true;
				if(!failed) {
					top = t; bot = b; pos--;
					if(pos < seed_lo) { // seed fully matched: this is the range that is reported
						extending = true;
						if(l8 == 0) { out->top = (uint64_t)top; out->len = (uint64_t)(bot - top); }
						if(!doExtend || pos < 0) finished = true;
						else if(bot - top == 1 && a.text) { // single row: hand over to k_seed_ext1
							if(l8 == 0) {
								SeedTask x; x.hit = (uint32_t)(out - a.hits); x.seq_off = (uint32_t)(rseq - a.seq); x.rdlen = (uint16_t)rdlen;
								x.pos = (int16_t)pos; x.seed_lo = (int16_t)seed_lo; x.rc = rc ? 1 : 0;
								a.xtasks[atomicAdd(a.nxtasks, 1u)] = x;
							}
							finished = true;
						}
					}
				}
			} else {
				st_extend++;
				bool stop = false;
				if(range1) {
					const int c = atZ ? -1 : qc;
					if(!atZ) top = t;
					if(c != ch && ch <= 3) stop = true;
					bot = top + 1;
				} else {
					if(b - t == bot - top) { top = t; bot = b; } else stop = true;
				}
				if(stop) finished = true;
				else { nlex++; pos--; if(nlex == 255 || pos < 0) finished = true; }
			}
		}
		if(failed) {
			if(l8 == 0) { out->top = 0; out->len = 0; out->nlex = 0; out->valid = 0; }
			busy = false;
		} else if(finished) {
			if(l8 == 0) { out->nlex = nlex; out->valid = 1; }
			busy = false;
		}
	}
	if(l8 == 0) {
		atomicAdd(&a.stats[1], (unsigned long long)st_search);
		atomicAdd(&a.stats[2], (unsigned long long)st_extend); atomicAdd(&a.stats[3], (unsigned long long)st_sides);
	}
}

// Left extension of single-row hits by text comparison (MultiSeedAligner::extend, aligner_seed.cpp:603-672,
// range == 1 branch: mapLF1 on the row, stop at the first read character that is not the BWT character;
// an N in the read never stops it; at the '$' row only Ns keep counting).  One thread per hit, a warp
// takes 32 tasks at a time.
__global__ void __launch_bounds__(256) k_seed_ext1(const DevIndex ix, const SeedKernelArgs a) {
	const uint32_t lane = threadIdx.x & 31;
	const uint32_t ntasks = *a.nxtasks;
	unsigned long long steps = 0;
	while(true) {
		uint32_t t0 = 0;
		if(lane == 0) t0 = atomicAdd(a.xcounter, 32u);
		t0 = __shfl_sync(0xffffffffu, t0, 0);
		if(t0 >= ntasks) break;
		const uint32_t ti = t0 + lane;
		if(ti < ntasks) {
			const SeedTask t = a.xtasks[ti];
			SeedHit* out = a.hits + t.hit;
			const uint32_t read = a.active[t.hit / a.W];
			const uint32_t* pk = a.pk + (size_t)read * BT2_PK_WORDS; const uint32_t* nm = a.nm + (size_t)read * BT2_NM_WORDS;
			const uint32_t* tw = reinterpret_cast<const uint32_t*>(a.text);
			const uint32_t rdlen = t.rdlen; const bool rc = t.rc != 0;
			uint64_t off = ix.sa32[out->top];
			int pos = t.pos; uint32_t nlex = 0;
			// The reference compares one base per LF step: strand base `pos` with the text base before `off`, both stepping down;
			// an N in the read matches anything, and at text offset 0 only Ns go on (without moving).  Here up to 16 steps at a
			// time: the packed read window against the packed text window, the first mismatch in step order ends the run.
			while(true) {
				if(off == 0) {   // the '$' row: nothing precedes it
					steps++;
					const uint32_t q = rc ? (rdlen - 1u - (uint32_t)pos) : (uint32_t)pos;
					if(!((nm[q >> 5] >> (q & 31u)) & 1u)) break;
					nlex++; pos--;
					if(nlex == 255 || pos < 0) break;
					continue;
				}
				uint32_t cnt = 16u;
				if(off < cnt) cnt = (uint32_t)off;
				if((uint32_t)pos + 1u < cnt) cnt = (uint32_t)pos + 1u;
				if(255u - nlex < cnt) cnt = 255u - nlex;
				const uint32_t cmask = cnt >= 16u ? ~0u : ((1u << (2u * cnt)) - 1u);
				const uint64_t x = off - cnt;                                        // text bases [x, off), base x + j at bits 2j
				const uint32_t tv = __funnelshift_r(__ldg(tw + (x >> 4)), __ldg(tw + (x >> 4) + 1), (uint32_t)(x & 15u) * 2u);
				const uint32_t q0 = rc ? (rdlen - 1u - (uint32_t)pos) : ((uint32_t)pos + 1u - cnt);   // forward-strand bases [q0, q0 + cnt)
				const uint32_t rv = __funnelshift_r(pk[q0 >> 4], pk[(q0 >> 4) + 1], (q0 & 15u) * 2u);
				uint32_t nb = __funnelshift_r(nm[q0 >> 5], nm[(q0 >> 5) + 1], q0 & 31u) & 0xffffu;
				nb = (nb | (nb << 8)) & 0x00ff00ffu; nb = (nb | (nb << 4)) & 0x0f0f0f0fu; nb = (nb | (nb << 2)) & 0x33333333u; nb = (nb | (nb << 1)) & 0x55555555u;
				uint32_t matched;
				if(!rc) {   // step i compares forward base pos - i with text base off - 1 - i: both windows ascending, steps run from the top
					uint32_t d = (tv ^ rv) & cmask; d = (d | (d >> 1)) & 0x55555555u & ~nb;
					matched = d ? cnt - 1u - ((31u - (uint32_t)__clz(d)) >> 1) : cnt;
				} else {    // step i compares the complement of forward base q0 + i with text base off - 1 - i: the text window reversed
					uint32_t tr = __brev(tv & cmask); tr = ((tr & 0x55555555u) << 1) | ((tr >> 1) & 0x55555555u); tr >>= (32u - 2u * cnt);
					uint32_t d = (tr ^ ~rv) & cmask; d = (d | (d >> 1)) & 0x55555555u & ~nb;
					matched = d ? ((uint32_t)__ffs(d) - 1u) >> 1 : cnt;
				}
				nlex += matched; pos -= (int)matched; off -= matched; steps += matched;
				if(matched < cnt) { steps++; break; }                                 // the comparison that failed was a step too
				if(nlex == 255 || pos < 0) break;
			}
			out->nlex = nlex; out->valid = 1;
		}
	}
	for(int o = 16; o > 0; o >>= 1) steps += __shfl_xor_sync(0xffffffffu, steps, o);
	if(lane == 0 && steps) atomicAdd(&a.stats[4], steps);
}

} // namespace bt2
