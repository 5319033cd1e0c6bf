// K_seed: exact FM backward search of every seed of every active read, followed by
// the left extension of the hits (nlex).
//   SeedAligner::searchSeedBi / startSearchSeedBi / nextLocsBi   aligner_seed.cpp:732-957
//   MultiSeedAligner::extend                                     aligner_seed.cpp:603-672
//   Ebwt::countBt2SideEx / mapBiLFEx / mapLF1 / ftabHi / ftabLo  bt2_idx.h:1940-1972, 2474-2564, 1430-1607
//
// Mapping onto the GPU: a group of 8 lanes owns one seed.  One LF step = the
// group reads the 128-byte side(s) of top and bot with one 16-byte load per lane
// (a full line per group, four lines per warp instruction), each of the six BWT
// lanes popcounts its 64 bases, counts are summed with three xor-shuffles, and
// the occ[] words held by lanes 6-7 are broadcast.  Groups are persistent: when a
// seed is finished the group takes the next task from a global counter, so the
// four groups of a warp stay in the same "one LF step per iteration" loop even
// though their seeds are in different phases (search vs. extend).
#pragma once
#include "dev_common.cuh"
#include "dev_prio.cuh"
#include "dev_step.cuh"

namespace bt2 {

struct SeedKernelArgs {
	const uint32_t*  active;    // [nactive] read ids
	uint32_t         nactive;
	uint32_t         W, maxOffs; // W = 2*maxOffs hit slots per active read
	const ReadState* states;
	const uint8_t*   seq;       // batch sequences, codes 0..4
	SeedHit*         hits;      // [nactive*W]
	uint32_t*        counter;   // task counter
	unsigned long long* stats;  // [0] seeds searched [1] LF steps search [2] LF steps extend [3] sides loaded
};

__device__ __forceinline__ uint4 ldg_u4(const uint8_t* p) { return __ldg(reinterpret_cast<const uint4*>(p)); }

// ranks of code c before rows `top` and `bot` (both + occ + fchr), computed by the 8-lane group
// If c < 0 the character is taken from the BWT at row `top` (Ebwt::rowL) and returned in c.
__device__ __forceinline__ void group_rank2(const DevIndex& ix, uint32_t gmask, uint32_t gbase, uint32_t l8, int& c,
                                            uint64_t top, uint64_t bot, uint64_t& rt, uint64_t& rb, uint32_t& sides) {
	const uint64_t side_t = top / BT2_SIDE_BWT_LEN, side_b = bot / BT2_SIDE_BWT_LEN;
	const uint32_t co_t = (uint32_t)(top - side_t * BT2_SIDE_BWT_LEN), co_b = (uint32_t)(bot - side_b * BT2_SIDE_BWT_LEN);
	const uint4 wt = ldg_u4(ix.ebwt + side_t * BT2_SIDE_SZ + 16u * l8);
	uint4 wb = wt;
	if(side_b != side_t) { wb = ldg_u4(ix.ebwt + side_b * BT2_SIDE_SZ + 16u * l8); sides += 2; } else sides += 1;
	if(c < 0) { // rowL(top): lane co_t/64 holds it, 16 bases per 32-bit word
		const uint32_t wi = (co_t >> 4) & 3u;
		const uint32_t wsel = wi == 0 ? wt.x : wi == 1 ? wt.y : wi == 2 ? wt.z : wt.w;
		c = (int)__shfl_sync(gmask, (wsel >> ((co_t & 15u) * 2)) & 3u, gbase + (co_t >> 6));
	}
	uint32_t ct = 0, cb = 0;
	if(l8 < 6) {
		const int nt = (int)co_t - 64 * (int)l8, nb = (int)co_b - 64 * (int)l8;
		const uint64_t t0 = (uint64_t)wt.x | ((uint64_t)wt.y << 32), t1 = (uint64_t)wt.z | ((uint64_t)wt.w << 32);
		const uint64_t b0 = (uint64_t)wb.x | ((uint64_t)wb.y << 32), b1 = (uint64_t)wb.z | ((uint64_t)wb.w << 32);
		if(nt > 0) ct = count_in_word(t0, c, nt < 32 ? nt : 32) + (nt > 32 ? count_in_word(t1, c, nt - 32 < 32 ? nt - 32 : 32) : 0);
		if(nb > 0) cb = count_in_word(b0, c, nb < 32 ? nb : 32) + (nb > 32 ? count_in_word(b1, c, nb - 32 < 32 ? nb - 32 : 32) : 0);
	}
	uint32_t packed = ct | (cb << 16);
	packed += __shfl_xor_sync(gmask, packed, 1);
	packed += __shfl_xor_sync(gmask, packed, 2);
	packed += __shfl_xor_sync(gmask, packed, 4);
	// occ[c]: lane 6 holds occ[0],occ[1]; lane 7 holds occ[2],occ[3]
	const uint32_t src = gbase + 6 + ((uint32_t)c >> 1);
	const uint32_t tlo = (c & 1) ? wt.z : wt.x, thi = (c & 1) ? wt.w : wt.y;
	const uint32_t blo = (c & 1) ? wb.z : wb.x, bhi = (c & 1) ? wb.w : wb.y;
	const uint64_t occ_t = (uint64_t)__shfl_sync(gmask, tlo, src) | ((uint64_t)__shfl_sync(gmask, thi, src) << 32);
	const uint64_t occ_b = (uint64_t)__shfl_sync(gmask, blo, src) | ((uint64_t)__shfl_sync(gmask, bhi, src) << 32);
	uint64_t cnt_t = packed & 0xffffu, cnt_b = packed >> 16;
	if(c == 0) { // '$' is stored as an A
		if(ix.zOff >= side_t * BT2_SIDE_BWT_LEN && ix.zOff < top) cnt_t--;
		if(ix.zOff >= side_b * BT2_SIDE_BWT_LEN && ix.zOff < bot) cnt_b--;
	}
	rt = cnt_t + occ_t + ix.fchr[c];
	rb = cnt_b + occ_b + ix.fchr[c];
}

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

#define BT2_SEED_GRAB 4

template<int DUMMY>
__global__ void __launch_bounds__(256) k_seed(const DevIndex ix, const DevParams* __restrict__ Pp, const SeedKernelArgs a) {
	const DevParams& P = *Pp;
	const uint32_t lane = threadIdx.x & 31, l8 = lane & 7, gbase = lane & ~7u;
	const uint32_t gmask = 0xffu << gbase;
	const uint64_t ntasks = (uint64_t)a.nactive * a.W;
	uint64_t task = 0, task_end = 0;
	// per-seed state (uniform within the group)
	bool busy = false, extending = false;
	uint64_t top = 0, bot = 0, seed_top = 0, seed_len = 0; // (seed_top, seed_len): the range reported, i.e. before the left extension
	int pos = 0;            // next strand-buffer position to consume (descending)
	int seed_lo = 0;        // strand-buffer offset of the seed's first base
	uint32_t nlex = 0;
	bool rc = false;
	const uint8_t* rseq = nullptr; uint32_t rdlen = 0;
	SeedHit* out = nullptr; uint64_t key = 0;
	unsigned long long st_seeds = 0, st_search = 0, st_extend = 0; uint32_t st_sides = 0;
	#define STRAND_CH(p) (rc ? (rseq[rdlen - 1 - (p)] > 3 ? 4 : 3 - rseq[rdlen - 1 - (p)]) : (int)rseq[(p)])
	while(true) {
		if(!busy) {
			if(task == task_end) {
				uint64_t t0 = 0;
				if(l8 == 0) t0 = atomicAdd(a.counter, (uint32_t)BT2_SEED_GRAB);
				t0 = __shfl_sync(gmask, (uint32_t)t0, gbase);
				task = t0; task_end = t0 + BT2_SEED_GRAB;
				if(task_end > ntasks) task_end = ntasks;
				if(task >= ntasks) break;
			}
			const uint64_t tk = task++;
			const uint32_t ai = (uint32_t)(tk / a.W), h = (uint32_t)(tk - (uint64_t)ai * a.W);
			const uint32_t s = h / a.maxOffs, i = h - s * a.maxOffs;
			const uint32_t read = a.active[ai];
			const ReadState& rs = a.states[read];
			rdlen = rs.rdlen; rseq = a.seq + rs.seq_off; rc = (s == 1);
			out = a.hits + (uint64_t)ai * a.W + h;
			const uint32_t interval = (uint32_t)rs.interval;
			uint32_t offset = 0; round_wanted(P, rs, rs.round, offset);
			const SeedGeom g = seed_geom(P, rdlen, offset, interval, a.maxOffs);
			bool ok = i < g.numOffs && !((s == 0 && P.nofw) || (s == 1 && P.norc));
			const int L = (int)g.seqLen;
			const int depth = (int)(i * g.per + g.off);
			seed_lo = rc ? ((int)rdlen - depth - L) : depth;
			if(ok && (seed_lo < 0 || seed_lo + L > (int)rdlen)) ok = false;
			key = 0;
			if(ok) { // InstantiatedSeed::instantiateExact: no N inside the seed
				for(int k = 0; k < L; k++) { const int ch = STRAND_CH(seed_lo + k); if(ch > 3) ok = false; key = (key << 2) | (uint64_t)(ch & 3); }
			}
			bool hit = false;
			if(ok) {
				st_seeds++;
				// startSearchSeedBi: ftab jump over the last ftabChars bases
				const int fl = ix.ftabChars;
				int step = 0;
				if(fl > 1 && fl <= L) {
					uint64_t k = 0;
					for(int q = 0; q < fl; q++) k = (k << 2) | (uint64_t)STRAND_CH(seed_lo + L - fl + q);
					top = ftab_hi(ix, k); bot = ftab_lo(ix, k + 1);
					step = fl;
				} else if(L > 0) {
					const int ch = STRAND_CH(seed_lo + L - 1);
					top = ix.fchr[ch]; bot = ix.fchr[ch + 1];
					step = 1;
				} else { top = 0; bot = ix.fchr[4]; }
				hit = bot != top;
				pos = seed_lo + L - 1 - step;
			}
			if(!hit) {
				if(l8 == 0) { out->top = 0; out->len = 0; out->key = key; out->nlex = 0; out->valid = 0; }
				continue;
			}
			busy = true; extending = false; nlex = 0;
			if(pos < seed_lo) { // whole seed consumed by the ftab jump
				extending = true; seed_top = top; seed_len = bot - top;
				if(!P.doExtend || pos < 0) { if(l8 == 0) { out->top = top; out->len = bot - top; out->key = key; out->nlex = 0; out->valid = 1; } busy = false; }
			}
			continue;
		}
		// ---- one LF step at strand position `pos`: a single rank query per iteration
		const int ch = STRAND_CH(pos);
		bool finished = false, failed = false;
		const bool range1 = (bot - top == 1);
		if(extending && !range1 && ch > 3) {
			// rare: read N while extending a range: any single child of the same size (aligner_seed.cpp:639-655)
			st_extend++;
			const uint64_t origSz = bot - top;
			int nonz = -1; bool abort_ = false; uint64_t nt = top, nb = bot;
			for(int j = 0; j < 4; j++) {
				int cj = j; uint64_t t, b; group_rank2(ix, gmask, gbase, l8, cj, top, bot, t, b, st_sides);
				if(b > t) { if(nonz >= 0) { abort_ = true; break; } nonz = j; nt = t; nb = b; }
			}
			if(abort_ || nonz < 0 || nb - nt < origSz) finished = true;
			else { top = nt; bot = nb; nlex++; pos--; if(nlex == 255 || pos < 0) finished = true; }
		} else {
			int qc = range1 ? -1 : ch;          // range of one row: follow the BWT char (mapLF1)
			const bool atZ = range1 && top == ix.zOff;
			uint64_t t = 0, b = 0;
			if(!atZ) group_rank2(ix, gmask, gbase, l8, qc, top, range1 ? top : bot, t, b, st_sides);
			if(!extending) {
				st_search++;
				if(range1) { if(atZ || qc != ch) failed = true; b = t + 1; }
				if(!failed && b == t) failed = true;
				if(!failed) {
					top = t; bot = b; pos--;
					if(pos < seed_lo) { extending = true; seed_top = top; seed_len = bot - top; if(!P.doExtend || pos < 0) finished = true; }
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
			if(l8 == 0) { out->top = 0; out->len = 0; out->key = key; out->nlex = 0; out->valid = 0; }
			busy = false;
		} else if(finished) {
			if(l8 == 0) { out->top = seed_top; out->len = seed_len; out->key = key; out->nlex = nlex; out->valid = 1; }
			busy = false;
		}
	}
	#undef STRAND_CH
	if(l8 == 0) {
		atomicAdd(&a.stats[0], st_seeds); atomicAdd(&a.stats[1], st_search);
		atomicAdd(&a.stats[2], st_extend); atomicAdd(&a.stats[3], (unsigned long long)st_sides);
	}
}

} // namespace bt2
