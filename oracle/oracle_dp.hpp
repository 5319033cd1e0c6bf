// TEST INFRASTRUCTURE ONLY -- restatement of the reference's end-to-end
// dynamic-programming aligner: SwAligner::initRef/align/nextAlignment and the
// u8 / i16 striped fills + backtraces, written as the plain cell recurrence
// (the reference's own -DSSE_SCALAR build computes exactly this).
// See oracle_fm.hpp for usage rules.
#pragma once
#include <cstdint>
#include <vector>
#include <algorithm>
#include <climits>
#include "oracle_seed.hpp"

namespace orc {

struct Edit { uint32_t pos; uint8_t chr, qchr, type; };

// The per-slot edit list SwDriver::resGap_.alres.ned_ (DList<Edit,160>, ds.h:980-1001):
// fixed storage whose contents survive reset()/clear() (only the count drops), which is
// what the reference's SAM writer later reads through aliased copies.
struct EditScratch {
	std::vector<Edit> buf; size_t cur = 0, hw = 0; // hw = high-water mark of entries ever written for this read
	EditScratch() : buf(1024, Edit{0, 0, 0, 0}) {}
	void clear() { cur = 0; }
	size_t size() const { return cur; }
	void resize(size_t n) { cur = n; if(n > hw) hw = n; }
	void push_back(const Edit& e) { buf[cur++] = e; if(cur > hw) hw = cur; }
	Edit* begin() { return buf.data(); }
	Edit* end() { return buf.data() + cur; }
	Edit& operator[](size_t i) { return buf[i]; }
};

// DPRect (dp_framer.h:40-122)
struct DPRect { int64_t refl, refr, refl_pretrim, refr_pretrim; size_t triml, trimr, corel, corer, maxgap; };

struct AlnOut {
	bool fw; uint64_t tidx; int64_t refoff; int64_t score; int ns, gaps, edits, basesAligned, refns;
	std::vector<Edit> ned; // after reverse()/invertEdits()
};

// Scoring::score / mm (scoring.h:150-185)
struct Scor {
	const bt2gpu_params* P;
	inline int score(int rdc, int refm, int q) const {
		if(q < 0) q = 0; if(q > 255) q = 255;
		if(rdc > 3 || refm > 15) return -P->npens[q];
		if((refm & (1 << rdc)) != 0) return P->matchBonus;
		return -P->mmpens[q];
	}
	inline int mm(int rdc, int refm, int q) const {
		if(q < 0) q = 0; if(q > 255) q = 255;
		return (rdc > 3 || refm > 15) ? P->npens[q] : P->mmpens[q];
	}
	int rdGapOpen() const { return P->rdGapConst + P->rdGapLinear; }
	int rfGapOpen() const { return P->rfGapConst + P->rfGapLinear; }
	int rdGapExt() const { return P->rdGapLinear; }
	int rfGapExt() const { return P->rfGapLinear; }
	// Scoring::maxReadGaps / maxRefGaps (scoring.cpp:42-98)
	int maxReadGaps(int64_t minsc, size_t rdlen) const {
		int64_t sc = (int64_t)rdlen * P->matchBonus; bool first = true; int num = 0;
		while(sc >= minsc) { if(first) { first = false; sc -= rdGapOpen(); } else sc -= rdGapExt(); num++; }
		return num - 1;
	}
	int maxRefGaps(int64_t minsc, size_t rdlen) const {
		int64_t sc = (int64_t)rdlen * P->matchBonus; bool first = true; int num = 0;
		while(sc >= minsc) { sc -= P->matchBonus; if(first) { first = false; sc -= rfGapOpen(); } else sc -= rfGapExt(); num++; }
		return num - 1;
	}
};

static const int firsts5[32] = { -1,0,1,0,2,0,1,0,3,0,1,0,2,0,1,0,4,0,1,0,2,0,1,0,3,0,1,0,2,0,1,0 }; // mask.cpp
static const int alts5[32]   = { 0,1,1,2,1,2,2,3,1,2,2,3,2,3,3,4,1,2,2,3,2,3,3,4,2,3,3,4,3,4,4,5 };
static const char mask2dna[] = "?ACMGRSVTWYHKDBN"; // alphabet.cpp; index 16 (N) handled below
inline char refMaskToChar(int m) { return m > 15 ? 'N' : mask2dna[m]; }

// One SwAligner problem (aligner_sw.h) for a fixed (read strand, reference window)
struct SwProblem {
	const bt2gpu_params* P; Scor sc;
	bool i16 = false;          // which fill produced the matrices (aligner_sw.cpp:77-83)
	bool fw = true; const uint8_t* rd = nullptr; const uint8_t* qu = nullptr;
	int rdlen = 0; int nceil = 0;
	uint64_t tidx = 0; int64_t reflen = 0; DPRect rect;
	std::vector<uint8_t> rf;   // reference masks (1<<c, N=16), rflen_+1 entries fetched
	int rflen = 0;             // # DP columns
	int64_t minsc = 0;
	// matrices, scores already offset so that "perfect" = 0 (value - 0xff / - 0x7fff)
	std::vector<int32_t> H, E, F;
	std::vector<uint16_t> masks; std::vector<uint8_t> rowReset;
	struct Cand { int row, col; int64_t score; };
	std::vector<Cand> cands; size_t cural = 0;
	inline size_t ix(int r, int c) const { return (size_t)r * rflen + c; }

	// EEU8_alignNucleotides / EEI16 equivalents as the plain recurrence.
	// u8 : unsigned saturating, start 0xff, offset -0xff (aligner_swsse_ee_u8.cpp:250-540)
	// i16: signed saturating, floor -0x8000, start 0x7fff, offset -0x7fff (aligner_swsse_ee_i16.cpp)
	bool align(int64_t& best, WorkCounters& wc) {
		const int rows = rdlen, cols = rflen;
		wc.dps++; wc.cells += (uint64_t)rows * cols;
		H.assign((size_t)rows * cols, 0); E.assign((size_t)rows * cols, 0); F.assign((size_t)rows * cols, 0);
		cands.clear(); cural = 0;
		const int lo = i16 ? -0x8000 : 0, hi = i16 ? 0x7fff : 0xff;
		const int rfgapo = sc.rfGapOpen(), rfgape = sc.rfGapExt(), rdgapo = sc.rdGapOpen(), rdgape = sc.rdGapExt();
		auto subs = [lo](int a, int b) { int v = a - b; return v < lo ? lo : v; };
		std::vector<int> Hprev(rows, lo), Eprev(rows, lo), Hcur(rows), Ecur(rows), Fcur(rows);
		// column -1: H = lo ("vlo"), E = lo  (u8: 0; i16: 0x8000)
		int lrmax = lo;
		for(int j = 0; j < cols; j++) {
			const int refc = firsts5[rf[j]];
			int Fpre = lo;
			for(int r = 0; r < rows; r++) {
				const bool gb = (r < P->gapbar) || (rows - 1 - r < P->gapbar);
				const int s0 = -sc.score(rd[r], 1 << refc, (int)qu[r] - 33);
				int f;
				if(r == 0) f = lo; // vf starts at "low"; then subs(vf, vs1)
				else f = Fpre;
				if(gb) f = lo;     // subtracting the 0xff (u8) / 0x8000 twice (i16) barrier saturates to the floor
				const int hd = (r == 0) ? hi : Hprev[r - 1];
				int e = Eprev[r]; // E for this column was computed from the previous column
				int h = subs(hd, s0);
				if(e > h) h = e;
				if(f > h) h = f;
				Hcur[r] = h; Fcur[r] = f;
				// E for the next column
				int e2 = subs(e, rdgape);
				int hopen = gb ? lo : subs(h, rdgapo);
				Ecur[r] = e2 > hopen ? e2 : hopen;
				// F for the next row
				int f2 = subs(f, rfgape);
				int fopen = subs(h, rfgapo);
				Fpre = f2 > fopen ? f2 : fopen;
				H[ix(r, j)] = h - hi; E[ix(r, j)] = e - hi; F[ix(r, j)] = f - hi;
			}
			const int lr = Hcur[rows - 1];
			if(lr > lrmax) lrmax = lr;
			if((int64_t)(lr - hi) >= minsc) {
				// both fills store the RAW last-row word as the candidate's score (aligner_swsse_ee_u8.cpp:~520,
				// aligner_swsse_ee_i16.cpp:528 `init(nrow-1, i, lr)`), so nextAlignment's `score < minsc` filter never fires
				Cand c; c.row = rows - 1; c.col = j; c.score = (int64_t)lr;
				cands.push_back(c);
			}
			Hprev.swap(Hcur); Eprev.swap(Ecur);
		}
		// the 16-bit fill also subtracts 0xff here, not 0x7fff (aligner_swsse_ee_i16.cpp:640): with it the test below never
		// fires and `best` is the raw word minus 255; reproduced as is
		const int64_t score = (int64_t)lrmax - 0xff;
		if(score < minsc) { best = score; return false; }
		if(lrmax == lo) { best = LLONG_MIN; return false; }
		if(!cands.empty()) {
			masks.assign((size_t)rows * cols, 0); rowReset.assign(rows, 0);
			std::sort(cands.begin(), cands.end(), [](const Cand& a, const Cand& b) {
				if(a.score != b.score) return a.score > b.score;
				if(a.row != b.row) return a.row > b.row;
				return a.col > b.col;
			});
		}
		best = score;
		return !cands.empty();
	}
	bool done() const { return cural == cands.size(); }
	bool backtrace(EditScratch& ned, AlnOut& out, int row, int col, WorkCounters& wc);
	bool nextAlignment(EditScratch& scratch, AlnOut& out, int64_t minscArg, Rng& rnd, WorkCounters& wc);
};

} // namespace orc
