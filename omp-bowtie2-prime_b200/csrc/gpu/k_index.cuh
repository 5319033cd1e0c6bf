// Device-side re-layout of the index, done once at bt2gpu_create.
//
// The .bt2l index samples the suffix array every 2^offRate rows (Ebwt::offs, bt2_idx.h:1252;
// offRate = 4 by default) and the reference resolves any other row by LF-walking to the next
// sampled one (GroupWalk2S::advanceElement group_walk.h:969, Ebwt::tryOffset/getOffset
// bt2_idx.h:1621, bt2_idx.cpp:147): ~2^offRate dependent random 128-byte reads per row.
// A B200 has room for the whole array (4 bytes per row: 12.4 GB for a 3.1 Gbp genome), so the
// library expands the sample into the full suffix array here and every later resolution is one
// 4-byte gather.  Each sampled row r0 owns the rows LF(r0), LF(LF(r0)), ... up to the next
// sampled row (their text offsets are SA[r0]-1, SA[r0]-2, ...), so the expansion costs one LF
// step per row of the index in total.  Only built when the text is shorter than 2^32.
#pragma once
#include "dev_common.cuh"

namespace bt2 {

__global__ void k_build_sa(const DevIndex ix, uint32_t* __restrict__ sa, uint64_t nsample) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= nsample) return;
	uint64_t row = i << ix.offRate;
	if(row > ix.len) return;
	uint64_t off = ix.offs[i];
	const uint64_t low = (1ull << ix.offRate) - 1;
	while(true) {
		sa[row] = (uint32_t)off;
		if(row == ix.zOff) break;             // text offset 0: nothing precedes it
		row = rank_c(ix, row_L(ix, row), row); // LF
		off--;
		if((row & low) == 0) break;           // the next sampled row has its own thread
	}
}

// Rows no chain reached: the suffixes at the very end of the text (the chain that starts at the
// empty suffix's row, which is not a sampled row in general).  They are resolved the reference's
// way, by walking to a sampled row.  `fix` = resolve, else just count the unresolved ones.
__global__ void k_fix_sa(const DevIndex ix, uint32_t* __restrict__ sa, uint64_t nrows, int fix, unsigned long long* bad) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= nrows || sa[i] != 0xffffffffu) return;
	if(!fix) { atomicAdd(bad, 1ull); return; }
	uint64_t row = i, jumps = 0;
	const uint64_t low = (1ull << ix.offRate) - 1;
	while(true) {
		if(row == ix.zOff) { sa[i] = (uint32_t)jumps; return; }
		if((row & low) == 0) { sa[i] = (uint32_t)(jumps + ix.offs[row >> ix.offRate]); return; }
		row = rank_c(ix, row_L(ix, row), row);
		jumps++;
	}
}

// Prefix table over the full suffix array: ktab[key] = first row whose suffix, cut to K bases, is >= key (key = K bases, first
// base most significant); ktab[4^K] = number of rows.  The rows of [ktab[key], ktab[key + 1]) are those whose suffix starts
// with the K-mer -- what Ebwt::ftabHi/ftabLo give for ftabChars = 10 (bt2_idx.h:1384-1607), for a K chosen from the text
// length so that the interval holds about one row and the seed search needs one or two comparisons instead of 2 log2(n / 4^10).
// The index sorts the end of the text AFTER every base ("ACG$" follows "ACGT..."; the empty suffix is the last row), so a suffix
// shorter than K counts as padded with T: it sits at the end of the interval of its padded key and compares greater than
// any pattern there (sa_cmp).
__device__ __forceinline__ uint64_t ktab_key(const uint32_t* __restrict__ tw, uint64_t len, uint32_t p, uint32_t K) {
	const uint64_t avail = len - p;                     // p <= len
	if(avail == 0) return (1ull << (2u * K)) - 1ull;
	const uint64_t w = p >> 4; const uint32_t sh = (p & 15u) * 2u;
	const uint32_t w0 = __ldg(tw + w), w1 = __ldg(tw + w + 1);
	uint32_t v = __funnelshift_r(w0, w1, sh);           // 16 bases, first base in the lowest bits
	v = __brev(v); v = ((v & 0x55555555u) << 1) | ((v >> 1) & 0x55555555u);   // first base most significant
	const uint32_t m = avail < (uint64_t)K ? (uint32_t)avail : K;
	return ((uint64_t)(v >> (32u - 2u * m)) << (2u * (K - m))) | ((1ull << (2u * (K - m))) - 1ull);
}
__global__ void k_build_ktab(const DevIndex ix, const uint32_t* __restrict__ tw, uint32_t* __restrict__ ktab, uint64_t nrows, uint32_t K) {
	const uint64_t i = (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if(i >= nrows) return;
	const uint64_t ki = ktab_key(tw, ix.len, ix.sa32[i], K);
	uint64_t k = i ? ktab_key(tw, ix.len, ix.sa32[i - 1], K) + 1 : 0;
	for(; k <= ki; k++) ktab[k] = (uint32_t)i;          // keys no earlier row reaches
	if(i == nrows - 1) ktab[(1ull << (2u * K))] = (uint32_t)nrows;
}
// keys above the last row's: empty intervals at the end
__global__ void k_ktab_tail(const DevIndex ix, const uint32_t* __restrict__ tw, uint32_t* __restrict__ ktab, uint64_t nrows, uint32_t K) {
	const uint64_t last = ktab_key(tw, ix.len, ix.sa32[nrows - 1], K);
	const uint64_t k = last + 1 + (uint64_t)blockIdx.x * blockDim.x + threadIdx.x;
	if(k < (1ull << (2u * K))) ktab[k] = (uint32_t)nrows;
}

} // namespace bt2
