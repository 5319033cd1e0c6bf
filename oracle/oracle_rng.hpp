// TEST INFRASTRUCTURE ONLY -- restatement of the reference's pseudo-random
// sources (random_source.h:28-175, random_util.h:31-232).  See oracle_fm.hpp
// for the rules on who may use oracle/.
#pragma once
#include <cstdint>
#include <vector>
#include <algorithm>

namespace orc {

// RandomSource (random_source.h:33-150): 32-bit LCG a=1664525 c=1013904223.
struct Rng {
	uint32_t last = 0, lastOff = 0;
	void init(uint32_t seed) { last = seed; lastOff = 30; }
	uint32_t nextU32() {
		last = 1664525u * last + 1013904223u;
		uint32_t ret = last >> 16;
		last = 1664525u * last + 1013904223u;
		ret ^= last;
		lastOff = 0;
		return ret;
	}
	uint64_t nextU64() { uint64_t a = nextU32(); a <<= 32; return a | nextU32(); }
	bool nextBool() {
		if(lastOff > 31) nextU32();
		uint32_t ret = (last >> lastOff) & 1;
		lastOff++;
		return ret;
	}
	float nextFloat() { return (float)nextU32() / (float)0xffffffff; }
};

// Random1toN (random_util.h:31-232): values 0..n-1 without replacement.
struct Random1toN {
	uint16_t n = 0, cur = 0, thresh = 0;
	bool swaplist = false, converted = false;
	std::vector<uint16_t> list, seen;
	void reset() { n = cur = 0; swaplist = converted = false; list.clear(); seen.clear(); thresh = 0; }
	void init(uint16_t n_, bool withoutReplacement) {
		n = n_; converted = false;
		swaplist = n_ < 128 || withoutReplacement;
		cur = 0; list.clear(); seen.clear();
		thresh = std::max<uint16_t>(16, (uint16_t)(0.10f * n_));
	}
	bool inited() const { return n > 0; }
	bool done() const { return inited() && cur >= n; }
	uint16_t next(Rng& rnd) {
		if(cur == 0 && !converted) {
			if(n == 1) { cur = 1; return 0; }
			if(swaplist) { list.resize(n); for(uint16_t i = 0; i < n; i++) list[i] = i; }
		}
		if(swaplist) {
			uint16_t r = cur + (rnd.nextU32() % (n - cur));
			if(r != cur) std::swap(list[cur], list[r]);
			return list[cur++];
		}
		bool again = true;
		uint16_t rn = 0;
		const uint16_t seenSz = (uint16_t)seen.size();
		while(again) {
			rn = rnd.nextU32() % (uint16_t)n;
			again = false;
			for(uint16_t i = 0; i < seenSz; i++) if(seen[i] == rn) { again = true; break; }
		}
		seen.push_back(rn);
		cur++;
		if(seen.size() >= thresh && cur < n) {
			std::sort(seen.begin(), seen.end());
			list.resize(n - cur);
			uint16_t prev = 0, c = 0;
			for(uint16_t i = 0; i <= seenSz; i++) {
				for(uint16_t j = prev; j < seen[i]; j++) list[c++] = j;
				prev = seen[i] + 1;
			}
			for(uint16_t j = prev; j < n; j++) list[c++] = j;
			seen.clear();
			cur = 0;
			n = (uint16_t)list.size();
			converted = true;
			swaplist = true;
		}
		return rn;
	}
};

} // namespace orc
