// TEST INFRASTRUCTURE ONLY -- scalar restatement of what K_fill computes (band, cell layout, candidates), shared by the
// emulated pipeline (emu_pipeline.cpp) and the lane-level emulation of the kernel (emu_fill.cpp).
#pragma once
#include <cstdio>
#include <cstdlib>
#include <vector>
#include <algorithm>
#include "../../omp-bowtie2-prime_b200/csrc/gpu/dev_common.cuh"
#include "../../omp-bowtie2-prime_b200/csrc/gpu/dev_bt.cuh"

using namespace bt2;

// Scalar stand-in for K_fill: the same band of diagonals (values outside it are 0), the same cell layout.
static size_t g_cells_in_band = 0, g_cells_rect = 0;
// Kforce: 0 = the smallest K that holds the band; else the K to lay the cells out with (the band then is the 32 K diagonals
// from d0, as in the kernel).  HEF (optional): H | E << 8 | F << 16 per cell of the rectangle, row-major.
static void fill_emu(const DevParams& P, Problem& pb, const ReadView& rv, uint8_t* cells, uint32_t* pcache, bool exact, uint32_t y, uint32_t Kforce = 0, uint32_t* HEF = nullptr) {
	const int rows = pb.rows, cols = pb.cols;
	int dlo, dhi; fill_band(P, rows, cols, pb.minsc, exact, dlo, dhi);
	const uint32_t K = Kforce ? Kforce : fill_K_for(dhi - dlo + 1);
	dhi = dlo + (int)(BT2_LANES * K) - 1;
	pb.lay = (uint8_t)((pb.lay & 0x40u) | K | (y << 7)); pb.d0 = (int16_t)dlo;
	auto inband = [&](int r, int c) { return c - r >= dlo && c - r <= dhi; };
	auto subs = [](int a, int b) { int v = a - b; return v < 0 ? 0 : v; };
	std::vector<int> Hprev(rows, 0), Eprev(rows, 0), Fprev(rows, 0), Hcur(rows), Ecur(rows), Fcur(rows), Eleft(rows, 0), Ehere(rows, 0);
	int lrmax = 0; std::vector<uint32_t> cand; const int S0 = prob_start(pb);
	g_cells_rect += (size_t)rows * cols;
	for(int j = 0; j < cols; j++) {
		const int refm = pb.rf[j];
		int Fpre = 0;
		for(int r = 0; r < rows; r++) {
			if(!inband(r, j)) { Hcur[r] = 0; Ecur[r] = 0; Fcur[r] = 0; Fpre = 0; Ehere[r] = 0; continue; }   // E of (r, j+1) and F of (r+1, j) see 0
			g_cells_in_band++;
			const bool gb = (r < P.gapbar) || (rows - 1 - r < P.gapbar);
			const int s0 = -sc_score(P, rv.rd(r), refm, rv.q(r) - 33);
			int f = (r == 0) ? 0 : Fpre; if(gb) f = 0;
			const int hd = (r == 0) ? S0 : Hprev[r - 1];
			const int e = Eprev[r];
			int h = subs(hd, s0); if(e > h) h = e; if(f > h) h = f;
			Hcur[r] = h;
			const int e2 = subs(e, P.rdGapExt), ho = gb ? 0 : subs(h, P.rdGapOpen);
			Ecur[r] = e2 > ho ? e2 : ho;
			const int f2 = subs(f, P.rfGapExt), fo = subs(h, P.rfGapOpen);
			Fpre = f2 > fo ? f2 : fo;
			Fcur[r] = f;
			if(HEF) HEF[(size_t)r * cols + j] = (uint32_t)h | ((uint32_t)e << 8) | ((uint32_t)f << 16);
			// the backtrace's comparisons at this cell, literally (raw values: the 0xff bias cancels)
			uint32_t ne = 0;
			const int hup = (r == 0) ? S0 : Hcur[r - 1], fup = (r == 0) ? 0 : Fcur[r - 1];
			if(h != hd - s0) ne |= CF_D;                 // s0 = penalty >= 0 (column 0 compares with H(r-1,-1) = 0; never read)
			if(h != e) ne |= CF_HE;
			if(h != f) ne |= CF_HF;
			if(!(j > 0 && e == Hprev[r] - P.rdGapOpen)) ne |= CF_EO;
			if(!(j > 0 && e == Eleft[r] - P.rdGapExt)) ne |= CF_EX;
			Ehere[r] = e;
			if(!(f == hup - P.rfGapOpen)) ne |= CF_FO;
			if(!(f == fup - P.rfGapExt)) ne |= CF_FX;
			const uint32_t ci = cell_index((uint32_t)r, (uint32_t)j, dlo, K, y);
			if(ci >= fill_cell_bytes((uint32_t)rows, K)) { fprintf(stderr, "emu: cell index out of range\n"); exit(1); }
			cells[ci] = (uint8_t)ne;
		}
		const int lr = Hcur[rows - 1];
		if(lr > lrmax) lrmax = lr;
		if(lr - S0 >= pb.minsc) cand.push_back(((uint32_t)lr << BT2_CAND_SHIFT) | (uint32_t)j);
		Hprev.swap(Hcur); Eprev.swap(Ecur); Fprev.swap(Fcur); Eleft = Ehere;
	}
	// character cache (what K_fill leaves for the backtrace's edits)
	for(int j = 0; j < BT2_PC_WORDS; j++) pcache[j] = 0;
	for(int r = 0; r < rows; r++) { const uint32_t c = (uint32_t)rv.rd(r); pcache[r >> 3] |= (c > 3 ? 4u : c) << ((r & 7) * 4); }
	for(int j = 0; j < cols; j++) { const uint32_t m = pb.rf[j]; const uint32_t c = m > 15 ? 4u : (uint32_t)__builtin_ctz(m); pcache[BT2_PC_RD + (j >> 3)] |= c << ((j & 7) * 4); }
	std::sort(cand.begin(), cand.end(), [](uint32_t a, uint32_t b) { return a > b; });
	pb.lrmax = lrmax; pb.ncand = (uint32_t)cand.size();
	for(size_t i = 0; i < cand.size(); i++) pb.cand[i] = cand[i];
}

