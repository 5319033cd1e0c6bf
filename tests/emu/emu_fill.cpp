// TEST INFRASTRUCTURE ONLY -- runs the LANE CODE of K_fill (k_fill.cuh: fill_step, fill_set_init, fill_build_rows,
// fill_build_ref, compiled here as plain C++ through the stand-ins of simd16.cuh) for 32 lane objects in lock step, with the
// kernel's shuffles done by hand, and compares what the warp would store with the scalar restatement (fill_scalar.hpp):
// every flag byte of the band (specified bits), the last row, the candidates, and in whole-rectangle mode H, E and F too.
// The kernel's own prologue / epilogue (a few lines around these calls) is only exercised on the GPU (tests/test_gpu_fill.py).
//   emu_fill [nproblems] [seed]
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <random>
#include "fill_scalar.hpp"
#include "../../omp-bowtie2-prime_b200/csrc/gpu/k_fill.cuh"

struct TProb { std::vector<uint8_t> rd, qual, rf; int rows, cols, minsc; bool i16; };

static TProb make_problem(std::mt19937& g, int idx) {
	static const int fr[9] = {1, 2, 63, 64, 65, 128, 129, 150, 160}, fc[9] = {1, 5, 199, 100, 70, 199, 150, 199, 199};
	auto rnd = [&](int lo, int hi) { return lo + (int)(g() % (uint32_t)(hi - lo + 1)); };
	TProb p;
	if(idx < 9) { p.rows = fr[idx]; p.cols = fc[idx]; }
	else if(idx % 3 == 0) { p.rows = 150; p.cols = 199; }                      // the benchmark's shape
	else { p.rows = rnd(1, 160); p.cols = rnd(1, 199); }
	std::vector<uint8_t> ref(p.cols + 1); for(auto& c : ref) c = (uint8_t)(g() & 3);
	const int st = p.cols - p.rows + 1 > 1 ? rnd(0, p.cols - p.rows) : 0;
	p.rd.resize(p.rows);
	for(int r = 0; r < p.rows; r++) { p.rd[r] = ref[(st + r) % (p.cols + 1)]; if(g() % 100 < 5) p.rd[r] = (uint8_t)(g() & 3); }
	if(p.rows > 30 && (g() & 1)) { const int q = rnd(5, p.rows - 6); p.rd.erase(p.rd.begin() + q); p.rd.push_back((uint8_t)(g() & 3)); }
	if(p.rows > 30 && (g() % 4 == 0)) { const int q = rnd(5, p.rows - 6); p.rd.insert(p.rd.begin() + q, (uint8_t)(g() & 3)); p.rd.pop_back(); }
	if(g() % 10 < 3) p.rd[rnd(0, p.rows - 1)] = 4;
	p.rf.resize(p.cols + 1); for(int c = 0; c <= p.cols; c++) p.rf[c] = (uint8_t)(1u << ref[c]);
	if(g() % 10 < 3) p.rf[rnd(0, p.cols - 1)] = 16;
	p.qual.resize(p.rows); for(auto& q : p.qual) q = (uint8_t)rnd(35, 73);
	p.minsc = -(int)(g() % 120);
	p.i16 = (g() % 4) == 0;                                   // the reference's 16-bit path: another bias, minimum scores below -254 allowed
	if(p.i16 && (g() & 1)) p.minsc = -(int)(g() % 400);
	return p;
}

static int g_fail = 0; static long g_warps = 0, g_cells = 0, g_skipped = 0;
#define CHECK(cond, ...) do { if(!(cond)) { if(g_fail < 20) { fprintf(stderr, __VA_ARGS__); fprintf(stderr, "\n"); } g_fail++; } } while(0)

// one warp = four problems (absent ones: nullptr)
template<int K, bool DBG> static void run_warp(const DevParams& P, const TProb* tp[4], bool exact) {
	constexpr int PER = 16 / K;
	static FillShared<K> sh; memset(&sh, 0xcd, sizeof(sh));
	FillProb pr[4]; int rowsW = 0, widthW = 1;
	for(int q = 0; q < 4; q++) {
		FillProb& X = pr[q]; X.on = tp[q] != nullptr; X.rows = X.on ? tp[q]->rows : 0; X.cols = X.on ? tp[q]->cols : 0; X.d0 = 0; X.start = (X.on && tp[q]->i16) ? BT2_I16_START : 0xff;
		if(!X.on) continue;
		int dlo, dhi; fill_band(P, X.rows, X.cols, tp[q]->minsc, exact, dlo, dhi);
		if(dhi - dlo + 1 > BT2_LANES * K) { g_skipped++; return; }   // this K cannot hold the problem (the host picks K so that it can)
		X.d0 = dlo; if(X.rows > rowsW) rowsW = X.rows; if(dhi - dlo + 1 > widthW) widthW = dhi - dlo + 1;
	}
	for(int q = 0; q < 4; q++) if(tp[q]) { int dlo, dhi; fill_band(P, tp[q]->rows, tp[q]->cols, tp[q]->minsc, exact, dlo, dhi); if(dhi - dlo + 1 > BT2_LANES * K) return; }
	// as the kernel: the sweep ends when the last lane that holds band cells has done its last row
	const int vlast = (widthW - 1) / K;
	int iters = ((rowsW + ((vlast + 2) >> 1) + PER - 1) / PER) * PER;
	if(iters > (int)fill_iters((uint32_t)rowsW, K)) iters = (int)fill_iters((uint32_t)rowsW, K);
	bool hasN = false; g_warps++;
	static const uint8_t dummy[256] = {0};
	for(int set = 0; set < 2; set++) {
		const TProb* A = tp[2 * set]; const TProb* B = tp[2 * set + 1];
		for(uint32_t lane = 0; lane < 32; lane++) {
			fill_build_rows<K>(P, P.mmpens, sh, set, pr[2 * set], pr[2 * set + 1], A ? A->rd.data() : dummy, A ? A->qual.data() : dummy, true,
			                   B ? B->rd.data() : dummy, B ? B->qual.data() : dummy, true, lane, BT2_FILL_ROWPAD + iters + 1);
			hasN = fill_build_ref<K>(sh, set, pr[2 * set], pr[2 * set + 1], A ? A->rf.data() : nullptr, B ? B->rf.data() : nullptr, lane, BT2_FILL_ROWPAD + iters + BT2_LANES * K + 1) || hasN;
		}
	}
	memset(sh.hl, 0, sizeof(sh.hl));
	std::vector<FillSet<K>> S0(32), S1(32);
	std::vector<std::vector<uint32_t>> hef(4);
	for(int q = 0; q < 4; q++) hef[q].assign(DBG ? BT2_HEF_WORDS : 1, 0xffffffffu);
	for(uint32_t lane = 0; lane < 32; lane++) {
		const int set0 = (int)(lane & 1u), set1 = set0 ^ 1;
		fill_set_init<K>(S0[lane], sh, pr, set0, lane, (int)((lane + 1u) >> 1));
		fill_set_init<K>(S1[lane], sh, pr, set1, lane, (int)(lane >> 1));
		if(DBG) { S0[lane].hefA = hef[2 * set0].data(); S0[lane].hefB = hef[2 * set0 + 1].data(); S1[lane].hefA = hef[2 * set1].data(); S1[lane].hefB = hef[2 * set1 + 1].data(); }
	}
	const uint32_t neg_rdge = v_pack2(-P.rdGapExt, -P.rdGapExt);
	const uint32_t E_EDGE = BT2_FL_BASE | ((uint32_t)(CF_EO | CF_EX) * 0x00010001u), F_EDGE = BT2_FL_BASE | ((uint32_t)(CF_FO | CF_FX) * 0x00010001u);
	uint32_t sE[32], sCE[32], sF[32], sCF[32];
	for(int l = 0; l < 32; l++) { sE[l] = 0; sCE[l] = E_EDGE; sF[l] = 0; sCF[l] = F_EDGE; }
	std::vector<std::vector<uint8_t>> cells(4, std::vector<uint8_t>(fill_cell_bytes(BT2GPU_MAX_ROWS, K), 0xee));
	for(int i0 = 0; i0 < iters; i0 += PER) {
		for(int u = 0; u < PER; u++) {
			const int i = i0 + u;
			for(int slot = 0; slot < 2; slot++) {
				uint32_t rE[32], rCE[32], rF[32], rCF[32];
				for(int l = 0; l < 32; l++) {
					rE[l] = l ? sE[l - 1] : 0u; rCE[l] = l ? sCE[l - 1] : E_EDGE;
					rF[l] = l < 31 ? sF[l + 1] : 0u; rCF[l] = l < 31 ? sCF[l + 1] : F_EDGE;
				}
				for(uint32_t l = 0; l < 32; l++) {
					FillSet<K>& S = slot ? S1[l] : S0[l];
					#define CALL(N) fill_step<K, N, DBG>(S, i, l, neg_rdge, rE[l], rCE[l], rF[l], rCF[l], sE[l], sCE[l], sF[l], sCF[l])
					if(hasN) CALL(true); else CALL(false);
					#undef CALL
				}
			}
		}
		for(uint32_t l = 0; l < 32; l++) {
			const int set0 = (int)(l & 1u), set1 = set0 ^ 1;
			const size_t g = ((size_t)(i0 / PER) * BT2_LANES + l) * 16;
			memcpy(&cells[2 * set0][g], S0[l].ca, 16); memcpy(&cells[2 * set0 + 1][g], S0[l].cb, 16);
			memcpy(&cells[2 * set1][g], S1[l].ca, 16); memcpy(&cells[2 * set1 + 1][g], S1[l].cb, 16);
		}
	}
	for(uint32_t l = 0; l < 32; l++) { fill_last_row<K>(S0[l]); fill_last_row<K>(S1[l]); }
	// ---- compare with the scalar restatement
	for(int q = 0; q < 4; q++) {
		if(!tp[q]) continue;
		const TProb& t = *tp[q];
		Problem pb; memset(&pb, 0, sizeof(pb)); pb.rows = (uint16_t)t.rows; pb.cols = (uint16_t)t.cols; pb.fw = 1; pb.minsc = t.minsc; pb.lay = t.i16 ? 0x40 : 0;
		memcpy(pb.rf, t.rf.data(), t.cols + 1);
		ReadView rv; rv.seq = t.rd.data(); rv.qual = t.qual.data(); rv.rdlen = (uint32_t)t.rows; rv.fw = true;
		std::vector<uint8_t> want(fill_cell_bytes(BT2GPU_MAX_ROWS, K), 0); std::vector<uint32_t> pc(BT2_PC_WORDS), HEF((size_t)t.rows * t.cols, 0);
		const uint32_t y = (uint32_t)(q >> 1);
		fill_emu(P, pb, rv, want.data(), pc.data(), exact, y, K, HEF.data());
		CHECK(pb.d0 == pr[q].d0, "d0 differs");
		const int d0 = pr[q].d0;
		int nbad = 0;
		for(int r = 0; r < t.rows; r++) for(int c = 0; c < t.cols; c++) {
			const int dd = c - r - d0;
			if(dd < 0 || dd >= (vlast + 1) * K) continue;   // (lanes past the warp's widest band do not finish their rows)
			uint8_t care = 0x7f;
			if(c == 0) care &= ~(CF_EO | CF_EX);
			if(r < P.gapbar || t.rows - 1 - r < P.gapbar) care &= ~(CF_EO | CF_EX | CF_FO | CF_FX | CF_HE | CF_HF);
			if(r == 0) care = CF_D;
			const uint32_t ci = cell_index((uint32_t)r, (uint32_t)c, d0, K, y);
			const uint8_t got = cells[q][ci], exp = want[ci]; g_cells++;
			if(((got ^ exp) & care) != 0 || (got & CF_VIS)) { if(nbad++ < 3) CHECK(false, "K=%d exact=%d problem %dx%d minsc %d d0 %d set %d: flags differ at row %d col %d: lanes %02x scalar %02x (care %02x)", K, exact, t.rows, t.cols, t.minsc, d0, (int)y, r, c, got, exp, care); }
			if(DBG && !t.i16) {   // (the debug words hold 8-bit values)
				const uint32_t hg = hef[q][(size_t)r * (BT2_LANES * K) + dd], hw = HEF[(size_t)r * t.cols + c];
				// E at column 0 and F in row 0 are boundary values no comparison reads; inside the gap barrier the scalar code forces F to 0 as the lanes do
				if(hg != hw) { if(nbad++ < 3) CHECK(false, "K=%d problem %dx%d: H/E/F differ at row %d col %d: lanes %06x scalar %06x", K, t.rows, t.cols, r, c, hg, hw); }
			}
		}
		// last row -> candidates
		int lrmax = 0; std::vector<uint32_t> cand;
		for(int c = 0; c < t.cols; c++) { const int h = sh.hl[q][c]; if(h > lrmax) lrmax = h; if(h - pr[q].start >= t.minsc) cand.push_back(((uint32_t)h << BT2_CAND_SHIFT) | (uint32_t)c); }
		std::sort(cand.begin(), cand.end(), [](uint32_t a, uint32_t b) { return a > b; });
		CHECK(lrmax == pb.lrmax, "K=%d exact=%d problem %dx%d minsc %d: lrmax %d vs %d", K, exact, t.rows, t.cols, t.minsc, lrmax, pb.lrmax);
		CHECK(cand.size() == pb.ncand, "K=%d problem %dx%d: ncand %zu vs %u", K, t.rows, t.cols, cand.size(), pb.ncand);
		for(size_t k = 0; k < cand.size() && k < pb.ncand; k++) CHECK(cand[k] == pb.cand[k], "cand %zu differs", k);
	}
}

template<int K> static void run_all(const DevParams& P, const std::vector<TProb>& probs, std::mt19937& g) {
	for(size_t b = 0; b + 4 <= probs.size(); b += 4) {
		const TProb* tp[4] = { &probs[b], &probs[b + 1], &probs[b + 2], &probs[b + 3] };
		if(b % 16 == 4) tp[g() % 4] = nullptr;          // an absent problem in the warp
		run_warp<K, false>(P, tp, false);
		if(b % 8 == 0) run_warp<K, true>(P, tp, false);
		if(K == 16) { run_warp<K, false>(P, tp, true); if(b % 8 == 0) run_warp<K, true>(P, tp, true); }
	}
}

// The reference's own known-answer values for Scoring::maxReadGaps / maxRefGaps (scoring.cpp:194-280, its SCORING_MAIN self-test):
// case 1 = Scoring::base1() (match bonus 1, gaps 11 + 4 per position), case 2 = match bonus 4, gaps 25 + 10; minimum score 0.
static void scoring_kats() {
	DevParams P; memset(&P, 0, sizeof(P));
	P.matchBonus = 1; P.rdGapOpen = P.rfGapOpen = 15; P.rdGapExt = P.rfGapExt = 4;
	static const int rf1[12] = {0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 1, 2}, rd1[12] = {0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2};
	for(int i = 0; i < 12; i++) { CHECK(max_ref_gaps(P, 0, 10 + i) == rf1[i], "maxRefGaps(0, %d)", 10 + i); CHECK(max_read_gaps(P, 0, 10 + i) == rd1[i], "maxReadGaps(0, %d)", 10 + i); }
	P.matchBonus = 4; P.rdGapOpen = P.rfGapOpen = 35; P.rdGapExt = P.rfGapExt = 10;
	static const int rf2[7] = {0, 0, 1, 1, 1, 1, 2}, rd2[7] = {0, 1, 1, 1, 2, 2, 3};
	for(int i = 0; i < 7; i++) { CHECK(max_ref_gaps(P, 0, 8 + i) == rf2[i], "case 2 maxRefGaps(0, %d)", 8 + i); CHECK(max_read_gaps(P, 0, 8 + i) == rd2[i], "case 2 maxReadGaps(0, %d)", 8 + i); }
}

int main(int argc, char** argv) {
	const int n = argc > 1 ? atoi(argv[1]) : 200; const uint32_t seed = argc > 2 ? (uint32_t)atoi(argv[2]) : 99u;
	scoring_kats();
	DevParams P; memset(&P, 0, sizeof(P));
	P.gapbar = 4; P.rdGapOpen = 8; P.rdGapExt = 3; P.rfGapOpen = 8; P.rfGapExt = 3; P.matchBonus = 0;
	for(int q = 0; q < 256; q++) { const int qq = q < 40 ? q : 40; P.mmpens[q] = (uint8_t)(2 + (int)((qq / 40.0f) * 4)); P.npens[q] = 1; }
	std::mt19937 g(seed);
	std::vector<TProb> probs; for(int i = 0; i < n; i++) probs.push_back(make_problem(g, i));
	run_all<2>(P, probs, g); run_all<4>(P, probs, g); run_all<8>(P, probs, g); run_all<16>(P, probs, g);
	// other gap penalties (asymmetric)
	P.rdGapOpen = 6; P.rdGapExt = 2; P.rfGapOpen = 9; P.rfGapExt = 4; P.gapbar = 1;
	run_all<8>(P, probs, g); run_all<16>(P, probs, g);
	if(g_fail) { fprintf(stderr, "emu_fill: %d mismatches\n", g_fail); return 1; }
	printf("emu_fill: %d problems, %ld warp runs (%ld skipped: band wider than 32 K), %ld cells compared: lanes == scalar restatement for K = 2, 4, 8, 16\n", n, g_warps, g_skipped, g_cells);
	return 0;
}
