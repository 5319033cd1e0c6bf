// TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference's FM-index and
// reference-fetch primitives (sfiligoi/omp-bowtie2-prime).  Not shipped, never
// linked by the product; only tests/, __graft_entry__.smoke() and bench.py's
// cpu_baseline leg may use anything under oracle/.
//
// Parity status: PINNED -- the whole oracle pipeline is checked against SAM
// produced by the unmodified reference binary built into oracle/_ref (see
// tests/test_oracle_vs_reference.py and tests/golden/).
//
// Written from the semantics of the reference, not its code shape: ranks are
// computed with a plain per-character loop instead of the popcount/LUT tricks.
#pragma once
#include <cstdint>
#include <cstddef>
#include <vector>
#include "bt2gpu.h"

namespace orc {

static const uint64_t OFF_MASK = ~0ull; // btypes.h (64-bit index)

struct Fm {
	const bt2gpu_index* ix;
	uint64_t sideSz, sideBwtSz, sideBwtLen; // EbwtParams::init, bt2_idx.h:136-169
	explicit Fm(const bt2gpu_index* i) : ix(i) {
		sideSz = 1ull << ix->lineRate;
		sideBwtSz = sideSz - 8 * 4;
		sideBwtLen = sideBwtSz * 4;
	}
	// BWT character of a row: Ebwt::rowL (bt2_idx.h:2306); '$' is stored as A.
	inline int rowL(uint64_t row) const {
		const uint64_t side = row / sideBwtLen, co = row % sideBwtLen;
		const uint8_t b = ix->ebwt[side * sideSz + (co >> 2)];
		return (b >> ((co & 3) * 2)) & 3;
	}
	// Cumulative occ[] stored after the BWT bytes of a side (bt2_idx.h:1957-1964)
	inline uint64_t sideOcc(uint64_t side, int c) const {
		uint64_t v; const uint8_t* p = ix->ebwt + side * sideSz + sideBwtSz + 8 * c;
		__builtin_memcpy(&v, p, 8); return v;
	}
	// # of character c in BWT[0,row) not counting '$', plus fchr[c]:
	// Ebwt::countBt2Side (bt2_idx.h:1811-1840) == one lane of countBt2SideEx (:1940-1972).
	inline uint64_t rank(int c, uint64_t row) const {
		const uint64_t side = row / sideBwtLen, co = row % sideBwtLen;
		const uint8_t* s = ix->ebwt + side * sideSz;
		uint64_t cnt = 0;
		for(uint64_t k = 0; k < co; k++) {
			if(((s[k >> 2] >> ((k & 3) * 2)) & 3) == c) cnt++;
		}
		// '$' looks like an A: subtract it if it lies in this side before the locus (:1819-1825)
		if(c == 0) {
			const uint64_t sideStart = side * sideBwtLen;
			if(ix->zOff >= sideStart && ix->zOff < row) cnt--;
		}
		return cnt + sideOcc(side, c) + ix->fchr[c];
	}
	// Ebwt::mapLF1(row, l, c) (bt2_idx.h:2511-2535)
	inline uint64_t mapLF1c(uint64_t row, int c) const {
		if(rowL(row) != c || row == ix->zOff) return OFF_MASK;
		return rank(c, row);
	}
	// Ebwt::mapLF1(row&, l) (bt2_idx.h:2542-2564): returns char or -1 at '$'
	inline int mapLF1(uint64_t& row) const {
		if(row == ix->zOff) return -1;
		int c = rowL(row);
		row = rank(c, row);
		return c;
	}
	// Ebwt::ftabHi / ftabLo (bt2_idx.h:1430-1463, 1591-1607)
	inline uint64_t ftabHi(uint64_t i) const {
		if(ix->ftab[i] <= ix->len) return ix->ftab[i];
		return ix->eftab[(ix->ftab[i] ^ OFF_MASK) * 2 + 1];
	}
	inline uint64_t ftabLo(uint64_t i) const {
		if(ix->ftab[i] <= ix->len) return ix->ftab[i];
		return ix->eftab[(ix->ftab[i] ^ OFF_MASK) * 2];
	}
	// SA[row]: Ebwt::getOffset (bt2_idx.cpp:147-172); this is what
	// GroupWalk2S::advanceElement resolves (group_walk.h:430-432, 969-1015).
	inline uint64_t saOffset(uint64_t row, uint64_t* steps = nullptr) const {
		uint64_t jumps = 0;
		const uint64_t offMask = OFF_MASK << ix->offRate;
		while(true) {
			if(row == ix->zOff) { if(steps) *steps += jumps; return jumps; }
			if((row & offMask) == row) { if(steps) *steps += jumps; return jumps + ix->offs[row >> ix->offRate]; }
			row = rank(rowL(row), row);
			jumps++;
		}
	}
	// Ebwt::joinedToTextOff with rejectStraddle=false (bt2_idx.cpp:54-124)
	inline void joinedToTextOff(uint64_t qlen, uint64_t off, uint64_t& tidx, uint64_t& textoff,
	                            uint64_t& tlen, bool& straddled) const {
		uint64_t top = 0, bot = ix->nFrag;
		while(true) {
			uint64_t elt = top + ((bot - top) >> 1);
			uint64_t lower = ix->rstarts[elt * 3];
			uint64_t upper = (elt == ix->nFrag - 1) ? ix->len : ix->rstarts[(elt + 1) * 3];
			if(lower <= off) {
				if(upper > off) {
					if(off + qlen > upper) straddled = true;
					tidx = ix->rstarts[elt * 3 + 1];
					textoff = (off - lower) + ix->rstarts[elt * 3 + 2];
					break;
				}
				top = elt;
			} else {
				bot = elt;
			}
		}
		tlen = ix->plen[tidx];
	}
};

// BitPairReference (reference.cpp): per-reference record ranges, and the base at
// a text position (4 = N inside a gap): getStretch semantics (reference.cpp:330+).
struct Ref {
	const bt2gpu_ref* rf;
	std::vector<uint64_t> refRecOffs; // first record of each reference (+ sentinel)
	std::vector<uint64_t> cumUnambig; // # unambiguous bases before each record
	explicit Ref(const bt2gpu_ref* r) : rf(r) {
		uint64_t cum = 0;
		for(uint64_t i = 0; i < rf->nrecs; i++) {
			if(rf->rec_first[i]) refRecOffs.push_back(i);
			cumUnambig.push_back(cum);
			cum += rf->rec_len[i];
		}
		refRecOffs.push_back(rf->nrecs);
		cumUnambig.push_back(cum);
	}
	// fill dst[0..count) with bases of reference tidx starting at toff (must lie inside the reference)
	void getStretch(uint8_t* dst, uint64_t tidx, uint64_t toff, uint64_t count) const {
		uint64_t pos = 0; // text position at the start of the current record's gap
		uint64_t k = 0;
		const uint64_t end = toff + count;
		for(uint64_t r = refRecOffs[tidx]; r < refRecOffs[tidx + 1] && k < count; r++) {
			const uint64_t gapEnd = pos + rf->rec_off[r];
			const uint64_t recEnd = gapEnd + rf->rec_len[r];
			// Ns of the gap
			for(uint64_t t = (toff + k > pos ? toff + k : pos); t < gapEnd && t < end; t++) dst[k++] = 4;
			for(uint64_t t = (toff + k > gapEnd ? toff + k : gapEnd); t < recEnd && t < end; t++) {
				const uint64_t bi = cumUnambig[r] + (t - gapEnd);
				dst[k++] = (rf->buf[bi >> 2] >> ((bi & 3) * 2)) & 3;
			}
			pos = recEnd;
		}
		while(k < count) dst[k++] = 4; // trailing Ns are not recorded; defensive
	}
};

} // namespace orc
