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

} // namespace bt2
