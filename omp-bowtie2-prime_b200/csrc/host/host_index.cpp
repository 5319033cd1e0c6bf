// .bt2l index + 2-bit reference loader.
// Mirrors Ebwt::readIntoMemory (bt2_io.cpp:134-600) and the BitPairReference
// constructor (reference.cpp:104-150) for the 64-bit-offset ("large") format,
// the only one the reference builds.  Little-endian hosts only.
#include "host_internal.hpp"
#include <cstdio>
#include <memory>

namespace bt2host {

namespace {
struct File {
	FILE* f = nullptr;
	explicit File(const std::string& p) { f = fopen(p.c_str(), "rb"); }
	~File() { if(f) fclose(f); }
	bool rd(void* dst, size_t n) { return n == 0 || fread(dst, 1, n, f) == n; }
	template<typename T> bool get(T& v) { return rd(&v, sizeof(T)); }
};
}

bool load_index(Job& job, const std::string& base, std::string& err) {
	// ---- <base>.1.bt2l : header, plen, rstarts, ebwt, zOff, fchr, ftab, eftab, names
	{
		File in(base + ".1.bt2l");
		if(!in.f) { err = "cannot open " + base + ".1.bt2l"; return false; }
		uint32_t one = 0;
		uint64_t len = 0;
		int32_t lineRate = 0, linesPerSide = 0, offRate = 0, ftabChars = 0, flags = 0;
		if(!in.get(one) || one != 1) { err = "bad endian sentinel in .1.bt2l"; return false; }
		if(!in.get(len) || !in.get(lineRate) || !in.get(linesPerSide) || !in.get(offRate) ||
		   !in.get(ftabChars) || !in.get(flags)) { err = "truncated header"; return false; }
		bt2gpu_index& ix = job.ix;
		memset(&ix, 0, sizeof(ix));
		ix.len = len; ix.lineRate = lineRate; ix.offRate = offRate; ix.ftabChars = ftabChars;
		// EbwtParams::init (bt2_idx.h:136-169)
		const uint64_t bwtSz = len / 4 + 1;
		const uint64_t sideSz = 1ull << lineRate;
		const uint64_t sideBwtSz = sideSz - 8 * 4;
		const uint64_t numSides = (bwtSz + sideBwtSz - 1) / sideBwtSz;
		ix.ebwtTotLen = numSides * sideSz;
		ix.ftabLen = (1ull << (ftabChars * 2)) + 1;
		ix.eftabLen = (uint64_t)ftabChars * 2;
		ix.offsLen = ((len + 1) + (1ull << offRate) - 1) >> offRate;
		if(!in.get(ix.nPat)) { err = "truncated nPat"; return false; }
		job.plen.resize(ix.nPat);
		if(!in.rd(job.plen.data(), ix.nPat * 8)) { err = "truncated plen"; return false; }
		if(!in.get(ix.nFrag)) { err = "truncated nFrag"; return false; }
		job.rstarts.resize(ix.nFrag * 3);
		if(!in.rd(job.rstarts.data(), ix.nFrag * 24)) { err = "truncated rstarts"; return false; }
		job.ebwt.resize(ix.ebwtTotLen);
		if(!in.rd(job.ebwt.data(), ix.ebwtTotLen)) { err = "truncated ebwt"; return false; }
		if(!in.get(ix.zOff)) { err = "truncated zOff"; return false; }
		job.fchr.resize(5);
		if(!in.rd(job.fchr.data(), 40)) { err = "truncated fchr"; return false; }
		job.ftab.resize(ix.ftabLen);
		if(!in.rd(job.ftab.data(), ix.ftabLen * 8)) { err = "truncated ftab"; return false; }
		job.eftab.resize(ix.eftabLen);
		if(!in.rd(job.eftab.data(), ix.eftabLen * 8)) { err = "truncated eftab"; return false; }
		// reference names: '\n'-separated, '\0'-terminated (bt2_io.cpp:496-512)
		job.refnames.clear();
		while(true) {
			int c = fgetc(in.f);
			if(c == EOF || c == '\0') break;
			if(c == '\n') { job.refnames.push_back(""); }
			else {
				if(job.refnames.empty()) job.refnames.push_back("");
				job.refnames.back().push_back((char)c);
			}
		}
		// a trailing '\n' leaves one empty name behind
		while(job.refnames.size() > ix.nPat && job.refnames.back().empty()) job.refnames.pop_back();
		while(job.refnames.size() < ix.nPat) job.refnames.push_back(std::to_string(job.refnames.size()));
	}
	// ---- <base>.2.bt2l : sampled suffix array
	{
		File in(base + ".2.bt2l");
		if(!in.f) { err = "cannot open " + base + ".2.bt2l"; return false; }
		uint32_t one = 0;
		if(!in.get(one) || one != 1) { err = "bad endian sentinel in .2.bt2l"; return false; }
		job.offs.resize(job.ix.offsLen);
		if(!in.rd(job.offs.data(), job.ix.offsLen * 8)) { err = "truncated offs"; return false; }
	}
	// ---- <base>.3.bt2l : RefRecords ; <base>.4.bt2l : 2-bit bases
	{
		File in(base + ".3.bt2l");
		if(!in.f) { err = "cannot open " + base + ".3.bt2l"; return false; }
		uint32_t one = 0; uint64_t nrecs = 0;
		if(!in.get(one) || one != 1) { err = "bad endian sentinel in .3.bt2l"; return false; }
		if(!in.get(nrecs) || nrecs == 0) { err = "no reference records"; return false; }
		job.rec_off.resize(nrecs); job.rec_len.resize(nrecs); job.rec_first.resize(nrecs);
		uint64_t cum = 0;
		job.reflens.clear();
		for(uint64_t i = 0; i < nrecs; i++) {
			uint8_t first = 0;
			if(!in.get(job.rec_off[i]) || !in.get(job.rec_len[i]) || !in.get(first)) {
				err = "truncated RefRecord"; return false;
			}
			job.rec_first[i] = first ? 1 : 0;
			if(first) job.reflens.push_back(0);
			if(job.reflens.empty()) { err = "first RefRecord not marked first"; return false; }
			job.reflens.back() += job.rec_off[i] + job.rec_len[i];
			cum += job.rec_len[i];
		}
		File in4(base + ".4.bt2l");
		if(!in4.f) { err = "cannot open " + base + ".4.bt2l"; return false; }
		const uint64_t nbytes = (cum + 3) / 4;
		job.refbuf.assign(nbytes + 16, 0);
		if(!in4.rd(job.refbuf.data(), nbytes)) { err = "truncated .4.bt2l"; return false; }
		job.rf.nrecs = nrecs;
		job.rf.bufBases = cum;
	}
	bt2gpu_index& ix = job.ix;
	ix.plen = job.plen.data(); ix.rstarts = job.rstarts.data(); ix.ebwt = job.ebwt.data();
	ix.fchr = job.fchr.data(); ix.ftab = job.ftab.data(); ix.eftab = job.eftab.data();
	ix.offs = job.offs.data();
	job.rf.rec_off = job.rec_off.data(); job.rf.rec_len = job.rec_len.data();
	job.rf.rec_first = job.rec_first.data(); job.rf.buf = job.refbuf.data();
	return true;
}

} // namespace bt2host
