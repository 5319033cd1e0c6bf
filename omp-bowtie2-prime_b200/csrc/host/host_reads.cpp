// Option parsing, FASTQ input and the per-read scalars of phase P0.
#include "host_internal.hpp"
#include <algorithm>
#include <cctype>
#ifdef _OPENMP
#include <omp.h>
#endif
#include <cstdio>
#include <unistd.h>
#include <ctime>
#include <cstdlib>
#include <sstream>
#include <algorithm>

namespace bt2host {

// ---------------------------------------------------------------- options
static int func_type(const std::string& s) { // SimpleFunc::parseType
	if(s == "C") return 1; if(s == "L") return 2; if(s == "S") return 3; if(s == "G") return 4;
	return 0;
}

// PARSE_FUNC (aligner_seed_policy.cpp:47-75): type, const, coeff, min, (5th token also sets min)
static bool parse_func(SimpleFunc& f, const std::string& arg) {
	std::vector<std::string> t; std::string cur; std::istringstream ss(arg);
	while(std::getline(ss, cur, ',')) t.push_back(cur);
	if(t.empty() || t.size() > 5) return false;
	int ty = func_type(t[0]); if(!ty) return false;
	f.type = ty;
	if(t.size() >= 2) f.C = atof(t[1].c_str());
	if(t.size() >= 3) f.L = atof(t[2].c_str());
	if(t.size() >= 4) f.I = atof(t[3].c_str());
	if(t.size() >= 5) f.I = atof(t[4].c_str());
	return true;
}

// Scoring::initPens (scoring.h:43-66)
static void init_pens(int32_t* pens, int type, int consMin, int consMax) {
	for(int i = 0; i < 256; i++) {
		if(type == 2) { // COST_MODEL_QUAL
			int ii = std::min(i, 40);
			float frac = (float)ii / 40.0f;
			pens[i] = consMin + (int)(frac * (consMax - consMin));
		} else { // COST_MODEL_CONSTANT
			pens[i] = consMax;
		}
	}
}

bool parse_options(Options& o, const std::string& args, std::string& err) {
	const double DMAX = std::numeric_limits<double>::max();
	bt2gpu_params& p = o.p;
	memset(&p, 0, sizeof(p));
	p.abi_version = BT2GPU_ABI_VERSION;
	// resetOptions (bt2_search.cpp:265-444) + preset "sensitive" (presets.cpp:47-53)
	p.seedLen = 22; p.seedMms = 0; p.nSeedRounds = 2; p.seedBoostThresh = 300;
	p.maxhalf = 15; p.doExtend = 1; p.enable8 = 1; p.tighten = 3;
	p.khits = 1; p.mhits = 50; p.msample = 1; p.nofw = 0; p.norc = 0; p.gapbar = 4;
	p.rdGapConst = 5; p.rdGapLinear = 3; p.rfGapConst = 5; p.rfGapLinear = 3; p.matchBonus = 0;
	int maxIters = 400, maxUg = 300, maxDp = 300, maxDpStreak = 15;
	const int maxItersIncr = 20, maxStreakIncr = 10;
	int mmType = 2, mmMax = 6, mmMin = 2, nType = 3, penN = 1;
	bool ignoreQuals = false, sawK = false;
	o.scoreMin.init(2, -0.6f, -0.6f);
	o.nCeil.init(2, 0.0, DMAX, 0.0f, 0.15f);
	o.msIval.init(3, 1.0, DMAX, 1.0f, 1.15f);
	o.cmdline = args;
	std::vector<std::string> tok; { std::istringstream ss(args); std::string t; while(ss >> t) tok.push_back(t); }
	auto need = [&](size_t& i) -> const std::string* {
		if(i + 1 >= tok.size()) { err = "option " + tok[i] + " needs an argument"; return nullptr; }
		return &tok[++i];
	};
	for(size_t i = 0; i < tok.size(); i++) {
		const std::string& t = tok[i];
		const std::string* a;
		if(t == "--very-fast")           { maxDpStreak = 5;  p.nSeedRounds = 1; p.seedLen = 22; o.msIval.init(3, 1.0, DMAX, 0.0, 2.50); }
		else if(t == "--fast")           { maxDpStreak = 10; p.nSeedRounds = 2; p.seedLen = 22; o.msIval.init(3, 1.0, DMAX, 0.0, 2.50); }
		else if(t == "--sensitive")      { maxDpStreak = 15; p.nSeedRounds = 2; p.seedLen = 22; o.msIval.init(3, 1.0, DMAX, 1.0, 1.15); }
		else if(t == "--very-sensitive") { maxDpStreak = 20; p.nSeedRounds = 3; p.seedLen = 20; o.msIval.init(3, 1.0, DMAX, 1.0, 0.50); }
		else if(t == "--end-to-end") { }
		else if(t == "--exact-xs") { p.exactBestLtMinsc = 1; }   // (not a reference option) bt2gpu_params.exactBestLtMinsc: fill whole DP rectangles
		else if(t == "-D") { if(!(a = need(i))) return false; maxDpStreak = atoi(a->c_str()); }
		else if(t == "-R") { if(!(a = need(i))) return false; p.nSeedRounds = atoi(a->c_str());
			if(p.nSeedRounds < 1) { err = "-R must be at least 1 (the reference divides by it: bowtie2-align dies of SIGFPE with -R 0)"; return false; } }
		else if(t == "-L") { if(!(a = need(i))) return false; p.seedLen = atoi(a->c_str());
			if(p.seedLen < 1 || p.seedLen > 32) { err = "-L must be within [1,32]"; return false; } }
		else if(t == "-N") { if(!(a = need(i))) return false; p.seedMms = atoi(a->c_str()); }
		else if(t == "-i") { if(!(a = need(i))) return false; if(!parse_func(o.msIval, *a)) { err = "bad -i"; return false; } }
		else if(t == "--score-min" || t == "--min-score") { if(!(a = need(i))) return false; if(!parse_func(o.scoreMin, *a)) { err = "bad --score-min"; return false; } }
		else if(t == "--n-ceil") { if(!(a = need(i))) return false; if(!parse_func(o.nCeil, *a)) { err = "bad --n-ceil"; return false; } }
		else if(t == "--dpad") { if(!(a = need(i))) return false; p.maxhalf = atoi(a->c_str()); }
		else if(t == "--gbar") { if(!(a = need(i))) return false; p.gapbar = std::max(1, atoi(a->c_str())); }
		else if(t == "-M") { if(!(a = need(i))) return false; p.mhits = (int32_t)std::max(0, atoi(a->c_str())); p.msample = 1; p.khits = 1; sawK = false; }   // bt2_search.cpp:1177-1190
		else if(t == "-k") { if(!(a = need(i))) return false; p.khits = std::max(1, atoi(a->c_str())); sawK = true; }
		else if(t == "--mp") { if(!(a = need(i))) return false; // "MX,MN" (bt2_search.cpp ARG "mp")
			int mx = 6, mn = 2; if(sscanf(a->c_str(), "%d,%d", &mx, &mn) < 2) mn = std::min(mn, mx); mmMax = mx; mmMin = std::min(mn, mx); }
		else if(t == "--np") { if(!(a = need(i))) return false; penN = atoi(a->c_str()); }
		else if(t == "--rdg") { if(!(a = need(i))) return false; sscanf(a->c_str(), "%d,%d", &p.rdGapConst, &p.rdGapLinear); }
		else if(t == "--rfg") { if(!(a = need(i))) return false; sscanf(a->c_str(), "%d,%d", &p.rfGapConst, &p.rfGapLinear); }
		else if(t == "--ignore-quals") ignoreQuals = true;
		else if(t == "--nofw") p.nofw = 1;
		else if(t == "--norc") p.norc = 1;
		else if(t == "--seed") { if(!(a = need(i))) return false; o.seed = (uint32_t)strtoul(a->c_str(), nullptr, 10); }
		else if(t == "--trim5" || t == "-5") { if(!(a = need(i))) return false; o.trim5 = atoi(a->c_str()); }
		else if(t == "--trim3" || t == "-3") { if(!(a = need(i))) return false; o.trim3 = atoi(a->c_str()); }
		else if(t == "--phred64") o.phred64 = true;
		else if(t == "--phred33" || t == "-q" || t == "--reorder" || t == "-t") { }
		else if(t == "--no-unal") o.noUnal = true;
		else if(t == "--no-yf-quirk") o.yfQuirk = false;
		else if(t == "--i16") p.i16 = 1;            // emulate a reference built with -DENABLE_I16
		else if(t == "--no-sse8") p.enable8 = 0;    // bt2_search.cpp:1353: only that build honours it (the default one warns and goes on)
		else if(t == "-p" || t == "--threads") { if(!(a = need(i))) return false; } // batch multiplier in the reference (F8); no meaning here
		else { err = "unsupported option: " + t; return false; }
	}
	if(ignoreQuals) mmType = 3;
	init_pens(p.mmpens, mmType, mmMin, mmMax);
	init_pens(p.npens, nType, penN, penN);
	if(sawK) { p.msample = 0; p.mhits = 0; } // bt2_search.cpp:1681-1687
	const int mul = (p.khits > 1) ? (p.khits - 1) : 0; // msWorkerConsts, bt2_search.cpp:2207-2211
	p.streak = maxDpStreak + mul * maxStreakIncr;
	p.mxDp = maxDp + mul * maxItersIncr;
	p.mxUg = maxUg + mul * maxItersIncr;
	p.mxIter = maxIters + mul * maxItersIncr;
	if(p.mxIter > BT2GPU_MAX_ITER) { err = "FATAL: mxIter too big"; return false; } // bt2_search.cpp:2213
	return true;
}

// ---------------------------------------------------------------- reads
// genRandSeed (pat.cpp:45-83)
uint32_t gen_rand_seed(const uint8_t* seq, const uint8_t* qual, size_t len,
                       const char* name, size_t name_len, uint32_t seed) {
	uint32_t rseed = (seed + 101) * 59 * 61 * 67 * 71 * 73 * 79 * 83;
	// base i lands at bit 2*(i mod 16), quality i at bit 8*(i mod 4): whole groups are folded as words
	size_t i = 0;
	for(; i + 16 <= len; i += 16) {
		uint32_t w = 0;
		for(int k = 0; k < 16; k++) w ^= (uint32_t)seq[i + k] << (2 * k);
		rseed ^= w;
	}
	for(; i < len; i++) rseed ^= (uint32_t)seq[i] << ((i & 15) << 1);
	for(i = 0; i + 4 <= len; i += 4) { uint32_t w; memcpy(&w, qual + i, 4); rseed ^= w; }   // little-endian: byte k at bit 8k
	for(; i < len; i++) rseed ^= (uint32_t)qual[i] << ((i & 3) << 3);
	for(i = 0; i < name_len; i++) {
		const int p = (int)name[i];
		if(p == '/') break;
		rseed ^= (uint32_t)(p << ((i & 3) << 3));
	}
	return rseed;
}

static inline uint8_t asc2dna(int c) { // alphabet.cpp:142
	switch(c) { case 'A': case 'a': return 0; case 'C': case 'c': return 1;
	            case 'G': case 'g': return 2; case 'T': case 't': return 3; default: return 4; }
}

char* Job::file_buffer(size_t n) {
	if(n > filecap) { free(filemem); filemem = (char*)malloc(n + (n >> 3)); filecap = filemem ? n + (n >> 3) : 0; }
	return filemem;
}
Job::~Job() { free(filemem); }

void Batch::finalize() {
	view.n = (uint32_t)(off.size() - 1);
	view.seq_off = off.data(); view.seq = seq.data(); view.qual = qual.data();
	view.rng_seed = seed.data(); view.minsc = minsc.data(); view.nceil = nceil.data();
	view.interval = interval.data();
}

// bt2_search.cpp:2484-2493, 2551-2555; aligner_sw.cpp:44
void fill_p0(const Job& job, Batch& b) {
	const size_t n = b.off.size() - 1;
	b.minsc.resize(n); b.nceil.resize(n); b.interval.resize(n);
	#pragma omp parallel for schedule(static) if(n > 65536)
	for(size_t i = 0; i < n; i++) {
		const size_t rdlen = b.off[i + 1] - b.off[i];
		int64_t ms = job.opt.scoreMin.f<int64_t>((double)rdlen);
		if(ms > 0) ms = 0;
		else if(!job.opt.p.i16 && ms < -254) ms = -254;
		b.minsc[i] = (int32_t)ms;
		b.nceil[i] = job.opt.nCeil.f<int>((double)rdlen);
		b.interval[i] = std::max((int)job.opt.msIval.f<int>((double)rdlen), 1);
	}
}

static void add_read(const Job& job, Batch& b, std::string name, const std::string& s, const std::string& q) {
	// trimming as FastqPatternSource::parse (pat.cpp:1148-1213)
	size_t t5 = std::min<size_t>((size_t)job.opt.trim5, s.size());
	size_t rem = s.size() - t5;
	size_t t3 = std::min<size_t>((size_t)job.opt.trim3, rem);
	size_t len = rem - t3;
	if(name.empty()) name = std::to_string(b.first_id + b.nnames());   // the reference names it by its global read id
	const size_t base = b.seq.size();
	for(size_t i = 0; i < len; i++) {
		int c = s[t5 + i]; if(c == '.') c = 'N';
		b.seq.push_back(asc2dna(c));
		int qc = (t5 + i < q.size()) ? (unsigned char)q[t5 + i] : 'I';
		if(job.opt.phred64) qc -= 31;
		b.qual.push_back((uint8_t)qc);
	}
	b.off.push_back((uint32_t)b.seq.size());
	b.seed.push_back(gen_rand_seed(b.seq.data() + base, b.qual.data() + base, len, name.data(), name.size(), job.opt.seed));
	b.add_name(name.data(), name.size());
}

static bool getline_trim(FILE* f, std::string& line) {
	line.clear();
	int c;
	bool any = false;
	while((c = fgetc(f)) != EOF) {
		any = true;
		if(c == '\n') break;
		if(c != '\r') line.push_back((char)c);
	}
	return any;
}

// per-character tables of the fast reader: kept[c] = c is '.' or alphabetic (pat.cpp:1151-1163), code[c] = asc2dna with '.' as N
struct FqTables {
	uint8_t kept[256], code[256];
	FqTables() { for(int c = 0; c < 256; c++) { kept[c] = (c == '.' || isalpha(c)) ? 1 : 0; code[c] = asc2dna(c == '.' ? 'N' : c); } }
};
static const FqTables kFq;

// the serial parser: handles every layout the fast path below declines (blank lines, CR line ends)
static bool batch_from_fastq_serial(const Job& job, Batch& b, const std::string& path, uint64_t skip, uint64_t maxn, std::string& err) {
	FILE* f = fopen(path.c_str(), "rb");
	if(!f) { err = "cannot open " + path; return false; }
	b.off.assign(1, 0);
	std::string h, s, plus, q;
	uint64_t rec = 0;
	while(true) {
		if(!getline_trim(f, h)) break;
		if(h.empty()) continue;
		if(h[0] != '@') { err = "FASTQ record does not start with @ in " + path; fclose(f); return false; }
		if(!getline_trim(f, s) || !getline_trim(f, plus) || !getline_trim(f, q)) break;
		// sequence keeps alphabetic characters only (pat.cpp:1151-1163)
		std::string s2; for(char c : s) { if(c == '.') c = 'N'; if(isalpha((unsigned char)c)) s2.push_back(c); }
		if(rec >= skip) add_read(job, b, h.substr(1), s2, q);
		rec++;
		if(maxn && rec >= skip + maxn) break;
	}
	fclose(f);
	fill_p0(job, b);
	b.finalize();
	return true;
}

// FastqPatternSource light + heavy parse (pat.cpp:1027-1213) for a whole file at once: the file is read into memory, line
// starts are found by OpenMP threads over equal slices, and records (four lines each) are converted straight into the batch's
// SoA arrays in two passes (lengths -> offsets -> bases/qualities/seed/name).  Same results as the serial parser; files
// with CR line ends, blank lines or a truncated last record take the serial parser.
// The fast reader on the bytes [foff, foff + max_bytes) of the file (max_bytes 0: to the end).  A range that stops before the end of
// the file is cut back to its last complete record; *next_off receives where the next range starts.  Returns 1, 0 when the layout
// needs the serial parser (nothing read into `b`), -1 on error.
static int fastq_fast(const Job& job, Batch& b, const std::string& path, uint64_t foff, uint64_t max_bytes, uint64_t skip, uint64_t maxn,
                      uint64_t* next_off, std::string& err) {
	FILE* f = fopen(path.c_str(), "rb");
	if(!f) { err = "cannot open " + path; return -1; }
	struct View { char* p; size_t n; char* data() const { return p; } size_t size() const { return n; } char& operator[](size_t i) const { return p[i]; } } buf{nullptr, 0};
	bool to_eof = true;
	{
		fseek(f, 0, SEEK_END); const long fsz = ftell(f); fseek(f, 0, SEEK_SET);
		if(fsz < 0) { fclose(f); return 0; }
		if(foff > (uint64_t)fsz) foff = (uint64_t)fsz;
		uint64_t sz = (uint64_t)fsz - foff;
		if(max_bytes && max_bytes < sz) { sz = max_bytes; to_eof = false; }
		buf.p = const_cast<Job&>(job).file_buffer((size_t)sz + 1); buf.n = (size_t)sz;   // the job's buffer keeps its pages between files
		if(!buf.p) { fclose(f); err = "out of memory reading " + path; return -1; }
		// slices of the range are read by several threads (pread: no shared file position)
		const int fd = fileno(f);
		int nr = 1;
#ifdef _OPENMP
		nr = std::max(1, std::min(job.reader_threads > 0 ? job.reader_threads : omp_get_max_threads(), (int)((size_t)sz >> 22) + 1));
#endif
		bool ok = true;
		#pragma omp parallel for num_threads(nr) schedule(static, 1) reduction(&&: ok)
		for(int t = 0; t < nr; t++) {
			size_t lo = (size_t)sz * (size_t)t / (size_t)nr; const size_t hi = (size_t)sz * (size_t)(t + 1) / (size_t)nr;
			while(lo < hi) { const ssize_t g = pread(fd, buf.data() + lo, hi - lo, (off_t)(foff + lo)); if(g <= 0) { ok = false; break; } lo += (size_t)g; }
		}
		fclose(f);
		if(!ok) { err = "cannot read " + path; return -1; }
	}
	static const bool trace = getenv("BT2HOST_TRACE") != nullptr;
	auto now = [] { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; };
	const double tA = now();
	size_t N = buf.size();
	if(next_off) *next_off = foff + N;
	if(N == 0) { if(foff == 0) return 0; b.off.assign(1, 0); b.name_off.assign(1, 0); fill_p0(job, b); b.finalize(); return 1; }   // past the last record: an empty batch
	// ---- line starts: every thread lists the '\n' of its slice (and looks for a CR, which sends the file to the serial parser)
	int nt = 1;
#ifdef _OPENMP
	nt = std::max(1, std::min(job.reader_threads > 0 ? job.reader_threads : omp_get_max_threads(), (int)(N >> 20) + 1));
#endif
	std::vector<std::vector<size_t>>& nl = job.scan_nl;   // kept by the job: no allocation from the second range on
	if(nl.size() < (size_t)nt) nl.resize((size_t)nt);
	bool has_cr = false;
	#pragma omp parallel for num_threads(nt) schedule(static, 1) reduction(||: has_cr)
	for(int t = 0; t < nt; t++) {
		const size_t lo = N * (size_t)t / (size_t)nt, hi = N * (size_t)(t + 1) / (size_t)nt;
		std::vector<size_t> v = std::move(nl[(size_t)t]);   // (on this thread's stack while it grows: neighbouring vector headers share cache lines)
		v.clear();
		const char* p = buf.data() + lo; const char* e = buf.data() + hi;
		if(memchr(p, '\r', (size_t)(e - p))) has_cr = true;
		else while(p < e) { const char* q = (const char*)memchr(p, '\n', (size_t)(e - p)); if(!q) break; v.push_back((size_t)(q - buf.data()) + 1); p = q + 1; }
		nl[(size_t)t] = std::move(v);
	}
	if(has_cr) return 0;
	std::vector<size_t>& ls = job.scan_ls;               // ls[i] = start of line i; ls[nlines] = one past the last line's '\n'
	{
		std::vector<size_t> at((size_t)nt + 1, 1);
		for(int t = 0; t < nt; t++) at[(size_t)t + 1] = at[(size_t)t] + nl[(size_t)t].size();
		ls.resize(at[(size_t)nt] + 1);
		ls[0] = 0;
		#pragma omp parallel for num_threads(nt) schedule(static, 1)
		for(int t = 0; t < nt; t++) if(!nl[(size_t)t].empty()) memcpy(&ls[at[(size_t)t]], nl[(size_t)t].data(), nl[(size_t)t].size() * sizeof(size_t));
		ls.resize(at[(size_t)nt]);
		if(!to_eof) {                                    // cut the range back to its last complete record
			const size_t full = ((ls.size() - 1) / 4) * 4;
			if(full == 0) { err = "FASTQ record longer than the read range in " + path; return -1; }
			ls.resize(full + 1); N = ls[full];
			if(next_off) *next_off = foff + N;
		} else if(ls.back() != N) ls.push_back(N + 1);   // last line without '\n': pretend one follows
	}
	const double tB = now();
	const size_t nlines = ls.size() - 1;
	if(nlines % 4 != 0) return 0;
	const size_t nrec = nlines / 4;
	const size_t r0 = std::min<size_t>((size_t)skip, nrec), r1 = maxn ? std::min<size_t>(nrec, r0 + (size_t)maxn) : nrec;
	const size_t n = r1 - r0;
	auto line_len = [&](size_t i) { return ls[i + 1] - 1 - ls[i]; };
	// ---- pass 1: every record starts with '@' on a non-empty line; trimmed read lengths; name lengths
	std::vector<uint32_t>& len = job.scan_len; len.resize(2 * n + 2);
	uint32_t* nmlen = len.data() + n;                    // bytes of name i incl. its NUL
	bool regular = true;
	const size_t t5o = (size_t)job.opt.trim5, t3o = (size_t)job.opt.trim3;
	#pragma omp parallel for num_threads(nt) schedule(static) reduction(&&: regular)
	for(size_t r = 0; r < nrec; r++) {
		const size_t h = 4 * r;
		if(line_len(h) == 0 || buf[ls[h]] != '@') regular = false;
		if(r < r0 || r >= r1) continue;
		const unsigned char* sq = (const unsigned char*)buf.data() + ls[h + 1]; const size_t sl = line_len(h + 1);
		uint32_t alpha = 0;                              // characters kept by the reference: letters and '.' (pat.cpp:1151-1163)
		for(size_t i = 0; i < sl; i++) alpha += (uint32_t)(((uint8_t)((sq[i] | 0x20) - 'a') < 26) | (sq[i] == '.'));
		const size_t t5 = std::min<size_t>(t5o, alpha), rem = alpha - t5, t3 = std::min<size_t>(t3o, rem);
		len[r - r0] = (uint32_t)(rem - t3);
		size_t nmb = line_len(h) - 1;
		if(nmb == 0) { char t[24]; nmb = (size_t)snprintf(t, sizeof(t), "%llu", (unsigned long long)(b.first_id + (r - r0))); }
		nmlen[r - r0] = ((uint32_t)nmb + 1) | (alpha == sl ? 0x80000000u : 0u);   // top bit: nothing on the sequence line is dropped
	}
	const double tC = now();
	if(!regular) return 0;
	b.off.resize(n + 1); b.name_off.resize(n + 1);
	b.off[0] = 0; b.name_off[0] = 0;
	{ uint64_t so = 0, no = 0;
	  for(size_t i = 0; i < n; i++) { so += len[i]; no += nmlen[i] & 0x7fffffffu; b.off[i + 1] = (uint32_t)so; b.name_off[i + 1] = (uint32_t)no; }
	  if(so > 0xffffffffull || no > 0xffffffffull) { err = "more than 4 G bases or name bytes in one batch: read " + path + " in ranges"; return -1; } }
	b.seq.resize(b.off[n]); b.qual.resize(b.off[n]); b.seed.resize(n); b.name_blob.resize(b.name_off[n]);
	// ---- pass 2: bases -> codes, qualities, name, seed
	const bool p64 = job.opt.phred64;
	#pragma omp parallel for num_threads(nt) schedule(static)
	for(size_t i = 0; i < n; i++) {
		const size_t h = 4 * (r0 + i);
		const unsigned char* sq = (const unsigned char*)buf.data() + ls[h + 1]; const size_t sl = line_len(h + 1);
		const unsigned char* ql = (const unsigned char*)buf.data() + ls[h + 3]; const size_t qn = line_len(h + 3);
		uint8_t* oseq = b.seq.data() + b.off[i]; uint8_t* oq = b.qual.data() + b.off[i];
		const size_t L = len[i];
		const size_t t5 = std::min<size_t>(t5o, sl);
		if((nmlen[i] & 0x80000000u) && qn >= t5 + L) {
			// the usual record: nothing but letters / '.' on the sequence line and a quality for every base
			const unsigned char* s = sq + t5; const unsigned char* q = ql + t5;
			for(size_t k = 0; k < L; k++) {
				const uint8_t u = (uint8_t)(s[k] & 0xDF);    // upper case; '.' and every other letter fall through to N
				oseq[k] = (uint8_t)(4 - 4 * (u == 'A') - 3 * (u == 'C') - 2 * (u == 'G') - (u == 'T'));
			}
			if(p64) { for(size_t k = 0; k < L; k++) oq[k] = (uint8_t)(q[k] - 31); }
			else memcpy(oq, q, L);
		} else {
			size_t a = 0, w = 0;                          // a: index among the kept characters of the line
			for(size_t k = 0; k < sl && w < L; k++) {
				const unsigned char c = sq[k];
				if(!kFq.kept[c]) continue;
				if(a >= t5o) {
					oseq[w] = kFq.code[c];
					int qc = (a < qn) ? ql[a] : 'I';
					if(p64) qc -= 31;
					oq[w] = (uint8_t)qc;
					w++;
				}
				a++;
			}
		}
		char* nm = (char*)b.name_blob.data() + b.name_off[i]; const size_t nmb = (nmlen[i] & 0x7fffffffu) - 1;
		if(line_len(h) > 1) memcpy(nm, buf.data() + ls[h] + 1, nmb);
		else snprintf(nm, nmb + 1, "%llu", (unsigned long long)(b.first_id + i));   // the reference names it by its global read id
		nm[nmb] = 0;
		b.seed[i] = gen_rand_seed(oseq, oq, L, nm, nmb, job.opt.seed);
	}
	const double tD = now();
	fill_p0(job, b);
	b.finalize();
	if(trace) fprintf(stderr, "[bt2host] fastq: lines %.1f ms, lengths %.1f ms, convert %.1f ms, P0 %.1f ms (%d threads)\n", tB - tA, tC - tB, tD - tC, now() - tD, nt);
	return 1;
}

bool batch_from_fastq(const Job& job, Batch& b, const std::string& path, uint64_t skip, uint64_t maxn, std::string& err) {
	b.first_id = skip;
	const int r = fastq_fast(job, b, path, 0, 0, skip, maxn, nullptr, err);
	if(r == 0) return batch_from_fastq_serial(job, b, path, skip, maxn, err);
	return r > 0;
}

// One range of a FASTQ file for a streaming caller: the records that lie completely inside [offset, offset + max_bytes).
bool batch_from_fastq_range(const Job& job, Batch& b, const std::string& path, uint64_t offset, uint64_t max_bytes, uint64_t* next_off, std::string& err) {
	if(offset == 0) job.range_next_id = 0;         // ranges are taken in file order (stream.py); unnamed reads get their global index
	b.first_id = job.range_next_id;
	const int r = fastq_fast(job, b, path, offset, max_bytes, 0, 0, next_off, err);
	if(r > 0) job.range_next_id += b.nnames();
	if(r == 0) err = "FASTQ layout needs the serial parser (CR line ends, blank lines or a truncated record): read the whole file with bt2host_batch_from_fastq";
	return r > 0;
}

bool batch_from_arrays(const Job& job, Batch& b, uint32_t n, const uint32_t* off, const char* seq,
                       const char* qual, const char* const* names) {
	b.off.assign(1, 0);
	for(uint32_t i = 0; i < n; i++) {
		std::string s(seq + off[i], seq + off[i + 1]);
		std::string q(qual + off[i], qual + off[i + 1]);
		add_read(job, b, names ? std::string(names[i]) : std::string(), s, q);
	}
	fill_p0(job, b);
	b.finalize();
	return true;
}

} // namespace bt2host
