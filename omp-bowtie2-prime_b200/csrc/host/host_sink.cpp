// Unpaired SAM sink: turns the per-read results of bt2gpu_align_batch() into
// the SAM records the reference's host sink prints for them.
//
// Mirrors (unpaired, end-to-end, default SAM flags):
//   ReportingState::finish/getReport     aln_sink.cpp:149-306
//   AlnSinkWrap::finishRead              aln_sink.cpp:643-1384 (unpaired branch :1057-1330)
//   AlnSinkWrap::selectByScore           aln_sink.cpp:1477-1627 (+ EList::shufflePortion ds.h:876)
//   AlnSink::reportHits                  aln_sink.h:667-745
//   BowtieMapq2::mapq                    unique.h:170-330
//   AlnSinkSam::appendMate               aln_sink.cpp:1889-2124
//   SamConfig::printAlignedOptFlags      sam.cpp:121-340
//   AlnRes::initStacked, StackedAln::*   aligner_result.h:1526, aligner_result.cpp:448-765
//   Edit::invertPoss                     edit.cpp:50-88
#include "host_internal.hpp"
#include <algorithm>
#include <cstdlib>

namespace bt2host {

namespace {

struct Rng { // RandomSource (random_source.h:33-150)
	uint32_t last, lastOff;
	uint32_t nextU32() {
		last = 1664525u * last + 1013904223u;
		uint32_t ret = last >> 16;
		last = 1664525u * last + 1013904223u;
		ret ^= last;
		lastOff = 0;
		return ret;
	}
	uint64_t nextU64() { uint64_t a = nextU32(); a <<= 32; return a | nextU32(); }
	size_t nextSizeT() { return (size_t)nextU64(); } // sizeof(size_t)==8 branch, random_source.h
};

// decimal text of an integer appended without a temporary string
static inline void put_int(std::string& o, long long v) {
	char t[24]; int n = 0;
	unsigned long long u = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
	do { t[n++] = (char)('0' + u % 10); u /= 10; } while(u);
	if(v < 0) t[n++] = '-';
	const size_t at = o.size(); o.resize(at + (size_t)n);
	for(int i = 0; i < n; i++) o[at + (size_t)i] = t[n - 1 - i];
}

struct Ed { uint32_t pos; uint8_t chr, qchr, type; };

// Edit::invertPoss(edits, sz, 0, n, sort=false) on the first n entries
void invert_poss(Ed* e, size_t n, size_t sz) {
	for(size_t i = 0; i < n / 2; i++) std::swap(e[i], e[n - i - 1]);
	for(size_t i = 0; i < n; i++)
		e[i].pos = (uint32_t)(sz - e[i].pos - (e[i].type == BT2GPU_EDIT_READ_GAP ? 0 : 1));
}

struct Stacked { // StackedAln
	std::string ref, rel, read;
	void init(const uint8_t* s, size_t slen, const Ed* ed, size_t n) {
		ref.clear(); rel.clear(); read.clear();
		size_t rdoff = 0;
		auto rdc = [&](size_t i) -> char { return (i < slen && s[i] <= 4) ? "ACGTN"[s[i]] : 'A'; }; // past-the-end reads are UB in the reference; its buffers are zero-filled in practice
		auto matched = [&](size_t upto) { // '=' columns for the read offsets [rdoff, upto)
			if(upto <= rdoff) return;
			const size_t cnt = upto - rdoff, at = ref.size();
			ref.resize(at + cnt); read.resize(at + cnt); rel.append(cnt, '=');
			for(size_t k = 0; k < cnt; k++) { const char c = rdc(rdoff + k); ref[at + k] = c; read[at + k] = c; }
			rdoff = upto;
		};
		for(size_t i = 0; i < n; i++) {
			const size_t pos = ed[i].pos;
			if(rdoff < pos) matched(std::min(pos, slen + 1024)); // (guard only bites on the reference's aliased garbage)
			if(ed[i].type == BT2GPU_EDIT_MM) {
				char c = rdc(rdoff++);
				ref.push_back((char)ed[i].chr); rel.push_back('X'); read.push_back(c);
			} else if(ed[i].type == BT2GPU_EDIT_REF_GAP) {
				char c = rdc(rdoff++);
				ref.push_back('-'); rel.push_back('I'); read.push_back(c);
			} else if(ed[i].type == BT2GPU_EDIT_READ_GAP) {
				ref.push_back((char)ed[i].chr); rel.push_back('D'); read.push_back('-');
			}
		}
		matched(slen);
	}
	void leftAlign() { // pastMms=false
		const size_t ln = ref.size();
		for(size_t i = 0; i < ln; i++) {
			const char r = rel[i];
			if(r != '=' && r != 'X') {
				size_t glen = 1;
				for(size_t j = i + 1; j < ln; j++) { if(r != rel[j]) break; glen++; }
				size_t l = i - 1; // wraps when i==0 exactly like the reference; then l>0 but ngp[l] is out of range
				size_t rr = l + glen;
				std::string& gp = (r == 'I') ? ref : read;
				const std::string& ngp = (r == 'I') ? read : ref;
				while(l > 0 && l < ln && rr < ln && ngp[l] == ngp[rr]) {
					if(rel[l] == 'X') break;
					std::swap(gp[l], gp[rr]);
					std::swap(rel[l], rel[rr]);
					l--; rr--;
				}
				i += (glen - 1);
			}
		}
	}
	void cigar(std::string& o) const {
		const size_t ln = rel.size();
		for(size_t i = 0; i < ln; i++) {
			char op = rel[i]; if(op == 'X' || op == '=') op = 'M';
			size_t run = 1;
			for(; i + run < ln; run++) {
				char op2 = rel[i + run]; if(op2 == 'X' || op2 == '=') op2 = 'M';
				if(op2 != op) break;
			}
			i += (run - 1);
			put_int(o, (long long)run); o.push_back(op);
		}
	}
	void mdz(std::string& o) const {
		bool mm_last = false, rdgap_last = false, first_print = true;
		const size_t ln = rel.size();
		for(size_t i = 0; i < ln; i++) {
			const char op = rel[i];
			if(op == '=') {
				size_t run = 1, nins = 0;
				for(; i + run < ln; run++) {
					if(rel[i + run] == '=') { } else if(rel[i + run] == 'I') nins++; else break;
				}
				i += (run - 1);
				if(run - nins > 0) {
					put_int(o, (long long)(run - nins));
					first_print = false; mm_last = false; rdgap_last = false;
				}
			} else if(op == 'X') {
				if(rdgap_last || mm_last || first_print) o.push_back('0');
				o.push_back(ref[i]);
				first_print = false; mm_last = true; rdgap_last = false;
			} else if(op == 'D') {
				if(mm_last || first_print) o.push_back('0');
				if(!rdgap_last) o.push_back('^');
				o.push_back(ref[i]);
				first_print = false; mm_last = false; rdgap_last = true;
			}
		}
		if(mm_last || rdgap_last) o.push_back('0');
	}
};

// BowtieMapq2::mapq, end-to-end (monotone) branch
int mapq_v2(bool hasSecbest, int64_t best, int64_t secbestIn, int64_t scMin, int64_t scPer,
            bool primary, bool canMax, bool exhausted) {
	if(!primary || (!canMax && !exhausted && !hasSecbest)) return 255;
	const int64_t diff = std::max<int64_t>(1, scPer - scMin);
	const int64_t bestOver = best - scMin;
	int ret = 0;
	if(!hasSecbest) {
		if     (bestOver >= diff * (double)0.8f) ret = 42;
		else if(bestOver >= diff * (double)0.7f) ret = 40;
		else if(bestOver >= diff * (double)0.6f) ret = 24;
		else if(bestOver >= diff * (double)0.5f) ret = 23;
		else if(bestOver >= diff * (double)0.4f) ret = 8;
		else if(bestOver >= diff * (double)0.3f) ret = 3;
		else ret = 0;
	} else {
		const int64_t secbest = secbestIn;
		const int64_t bestdiff = std::labs(std::labs((long)best) - std::labs((long)secbest));
		if(bestdiff >= diff * (double)0.9f)      ret = (bestOver == diff) ? 39 : 33;
		else if(bestdiff >= diff * (double)0.8f) ret = (bestOver == diff) ? 38 : 27;
		else if(bestdiff >= diff * (double)0.7f) ret = (bestOver == diff) ? 37 : 26;
		else if(bestdiff >= diff * (double)0.6f) ret = (bestOver == diff) ? 36 : 22;
		else if(bestdiff >= diff * (double)0.5f) {
			if(bestOver == diff) ret = 35; else if(bestOver >= diff * (double)0.84f) ret = 25;
			else if(bestOver >= diff * (double)0.68f) ret = 16; else ret = 5;
		} else if(bestdiff >= diff * (double)0.4f) {
			if(bestOver == diff) ret = 34; else if(bestOver >= diff * (double)0.84f) ret = 21;
			else if(bestOver >= diff * (double)0.68f) ret = 14; else ret = 4;
		} else if(bestdiff >= diff * (double)0.3f) {
			if(bestOver == diff) ret = 32; else if(bestOver >= diff * (double)0.88f) ret = 18;
			else if(bestOver >= diff * (double)0.67f) ret = 15; else ret = 3;
		} else if(bestdiff >= diff * (double)0.2f) {
			if(bestOver == diff) ret = 31; else if(bestOver >= diff * (double)0.88f) ret = 17;
			else if(bestOver >= diff * (double)0.67f) ret = 11; else ret = 0;
		} else if(bestdiff >= diff * (double)0.1f) {
			if(bestOver == diff) ret = 30; else if(bestOver >= diff * (double)0.88f) ret = 12;
			else if(bestOver >= diff * (double)0.67f) ret = 7; else ret = 0;
		} else if(bestdiff > 0) {
			ret = (bestOver >= diff * (double)0.67f) ? 6 : 2;
		} else {
			ret = (bestOver >= diff * (double)0.67f) ? 1 : 0;
		}
	}
	return ret;
}

} // namespace

// the sink's per-thread scratch (no allocation per read; the caller keeps one per thread)
struct SinkScratch::Impl {
	std::vector<std::pair<int64_t, size_t> > buf;
	std::vector<Ed> ed; Stacked st; std::vector<uint8_t> rcbuf; std::string md, cig;
};
SinkScratch::SinkScratch() : impl(new Impl) {}
SinkScratch::~SinkScratch() { delete impl; }

namespace {
struct Wr { // writes through a pointer into room the caller reserved
	char* p;
	void c(char x) { *p++ = x; }
	void s(const char* x, size_t n) { memcpy(p, x, n); p += n; }
	template<size_t N> void lit(const char (&x)[N]) { memcpy(p, x, N - 1); p += N - 1; }
	void i(long long v) {
		char t[24]; int n = 0;
		unsigned long long u = v < 0 ? 0ull - (unsigned long long)v : (unsigned long long)v;
		do { t[n++] = (char)('0' + u % 10); u /= 10; } while(u);
		if(v < 0) *p++ = '-';
		while(n) *p++ = t[--n];
	}
};
}

// One read -> SAM text appended to `o`.
void sam_for_read(const Job& job, uint64_t met[4], const Batch& b, const bt2gpu_results& res, uint32_t ri, int mode, std::string& o, SinkScratch& scratch) {
	SinkScratch::Impl& sc = *scratch.impl;
	const Options& opt = job.opt;
	const bt2gpu_read_result& rr = res.reads[ri];
	const uint8_t* seq = b.seq.data() + b.off[ri];
	const uint8_t* qual = b.qual.data() + b.off[ri];
	const size_t rdlen = b.off[ri + 1] - b.off[ri];
	const char* name = b.name_ptr(ri); const size_t name_n = b.name_len(ri);
	const uint64_t n = rr.n_alns;
	const bool mhitsSet = opt.p.mhits != 0;
	// ReportingState::getReport
	uint64_t nrep = 0; bool maxed = false;
	if(n > 0) {
		if(!mhitsSet && n >= (uint64_t)opt.p.khits) nrep = (uint64_t)opt.p.khits;
		else if(mhitsSet && n > (uint64_t)opt.p.mhits) { maxed = true; nrep = 1; }
		else nrep = std::min<uint64_t>(n, (uint64_t)opt.p.khits);
	}
	met[0]++;
	if(nrep == 0) met[1]++; else if(maxed || n > 1) met[3]++; else met[2]++;
	auto qname = [&]() {
		for(size_t i = 0; i < name_n && i < 255; i++) { if(isspace((unsigned char)name[i])) break; o.push_back(name[i]); }
	};
	auto seqqual = [&](bool fw) {
		if(rdlen == 0) { o += "*\t*"; return; }
		const size_t at = o.size(); o.resize(at + 2 * rdlen + 1);
		char* w = &o[at];
		if(fw) { for(size_t i = 0; i < rdlen; i++) w[i] = "ACGTN"[seq[i]]; }
		else   { for(size_t i = 0; i < rdlen; i++) w[i] = "TGCAN"[seq[rdlen - 1 - i]]; }
		w[rdlen] = '\t'; w += rdlen + 1;
		if(fw) memcpy(w, qual, rdlen);
		else   { for(size_t i = 0; i < rdlen; i++) w[i] = (char)qual[rdlen - 1 - i]; }
	};
	if(nrep == 0) {
		if(opt.noUnal) return;
		qname(); o += "\t4\t*\t0\t0\t*\t*\t0\t0\t";
		seqqual(true);
		o += "\tYT:Z:UU";
		if(opt.yfQuirk) o += "\tYF:Z:LN";
		o.push_back('\n');
		return;
	}
	// selectByScore
	const bt2gpu_aln* al = res.alns + rr.aln_off;
	std::vector<std::pair<int64_t, size_t> >& buf = sc.buf;
	buf.resize(n);
	for(size_t i = 0; i < n; i++) { buf[i].first = al[i].score; buf[i].second = i; }
	if(n > 1) { std::sort(buf.begin(), buf.end()); std::reverse(buf.begin(), buf.end()); }
	Rng rnd; rnd.last = rr.rng_last; rnd.lastOff = rr.rng_lastOff;
	auto shuffle = [&](size_t begin, size_t num) {
		if(num < 2) return;
		size_t left = num;
		for(size_t i = begin; i < begin + num - 1; i++) {
			size_t rndi = rnd.nextSizeT() % left;
			if(rndi > 0) std::swap(buf[i], buf[i + rndi]);
			left--;
		}
	};
	size_t streak = 0;
	for(size_t i = 1; i < buf.size(); i++) {
		if(buf[i].first == buf[i - 1].first) { if(streak == 0) streak = 1; streak++; }
		else { if(streak > 1) shuffle(i - streak, streak); streak = 0; }
	}
	if(streak > 1) shuffle(buf.size() - streak, streak);
	const int64_t best = al[buf[0].second].score;
	const bool hasSec = n > 1;
	const int64_t secbest = hasSec ? al[buf[1].second].score : 0;
	const int64_t scMin = opt.scoreMin.f<int64_t>((double)(float)rdlen);
	const int64_t scPer = (int64_t)rdlen * opt.p.matchBonus;
	// the aliased edit scratch
	std::vector<Ed>& ed = sc.ed; Stacked& st = sc.st; std::vector<uint8_t>& rcbuf = sc.rcbuf; std::string& md = sc.md; std::string& cig = sc.cig;
	bool have_rc = false;
	// the reference sequence's name up to the first whitespace (SamConfig::printRefNameFromIndex)
	for(uint64_t k = 0; k < nrep; k++) {
		const bt2gpu_aln& a = al[buf[k].second];
		const bool primary = (k == 0);
		const size_t ne = a.n_edits;
		ed.resize(ne);
		for(size_t i = 0; i < ne; i++) {
			const bt2gpu_edit* src;
			static const bt2gpu_edit zero = {0, 0, 0, 0, 0};
			if(mode == 1) src = (i < rr.n_scratch) ? &res.edits[rr.scratch_off + i] : &zero;
			else src = &res.edits[a.edit_off + i];
			ed[i].pos = src->pos; ed[i].chr = src->chr; ed[i].qchr = src->qchr; ed[i].type = src->type;
		}
		// AlnRes::initStacked: positions w.r.t. the left end of the aligned strand
		if(!a.fw) invert_poss(ed.data(), ne, rdlen);
		// No gap among the edits (nine alignments in ten; an entry of the aliased buffer that is no edit at all -- type 0 -- only moves
		// the read offset): the stacked alignment is a row of '=' and 'X' columns, left-alignment has nothing to move, the CIGAR is one
		// M run, and MD:Z follows from the edits alone.  StackedAln::init and buildMdz walked over the edit list without the strings:
		bool plain = true;
		for(size_t i = 0; i < ne && plain; i++) plain = ed[i].type != BT2GPU_EDIT_READ_GAP && ed[i].type != BT2GPU_EDIT_REF_GAP;
		md.clear(); cig.clear();
		if(plain) {
			size_t rdoff = 0, run = 0, cols = 0; bool mm_last = false, first_print = true;
			for(size_t i = 0; i < ne; i++) {
				const size_t pos = ed[i].pos;
				if(rdoff < pos) { const size_t upto = std::min(pos, rdlen + 1024); run += upto - rdoff; rdoff = upto; }
				if(ed[i].type == BT2GPU_EDIT_MM) {
					if(run > 0) { put_int(md, (long long)run); cols += run; run = 0; first_print = false; mm_last = false; }
					if(mm_last || first_print) md.push_back('0');
					md.push_back((char)ed[i].chr);
					first_print = false; mm_last = true; rdoff++; cols++;
				}
			}
			if(rdoff < rdlen) run += rdlen - rdoff;
			if(run > 0) { put_int(md, (long long)run); cols += run; mm_last = false; }
			if(mm_last) md.push_back('0');
			if(cols > 0) { put_int(cig, (long long)cols); cig.push_back('M'); }
		} else {
			if(!a.fw && !have_rc) {
				rcbuf.resize(rdlen);
				for(size_t i = 0; i < rdlen; i++) { uint8_t c = seq[rdlen - 1 - i]; rcbuf[i] = c > 3 ? 4 : (uint8_t)(3 - c); }
				have_rc = true;
			}
			st.init(a.fw ? seq : rcbuf.data(), rdlen, ed.data(), ne);
			st.leftAlign();
			st.cigar(cig); st.mdz(md);
		}
		if(!a.fw) invert_poss(ed.data(), ne, rdlen);
		size_t num_mm = 0, num_go = 0, num_gx = 0;
		for(size_t i = 0; i < ne; i++) {
			if(ed[i].type == BT2GPU_EDIT_MM) num_mm++;
			else if(ed[i].type == BT2GPU_EDIT_READ_GAP) {
				num_go++; num_gx++;
				while(i < ne - 1 && ed[i + 1].pos == ed[i].pos && ed[i + 1].type == BT2GPU_EDIT_READ_GAP) { i++; num_gx++; }
			} else if(ed[i].type == BT2GPU_EDIT_REF_GAP) {
				num_go++; num_gx++;
				while(i < ne - 1 && ed[i + 1].pos == ed[i].pos + 1 && ed[i + 1].type == BT2GPU_EDIT_REF_GAP) { i++; num_gx++; }
			}
		}
		// the record, written through a pointer into room reserved for its longest possible form
		const std::string* rn = a.tidx < job.refnames.size() ? &job.refnames[a.tidx] : nullptr;
		const size_t at = o.size();
		o.resize(at + 512 + std::min<size_t>(name_n, 255) + (rn ? rn->size() : 1) + 2 * rdlen + cig.size() + md.size());
		Wr w{&o[at]};
		for(size_t i = 0; i < name_n && i < 255; i++) { if(isspace((unsigned char)name[i])) break; w.c(name[i]); }
		w.c('\t'); w.i((primary ? 0 : 256) | (a.fw ? 0 : 16)); w.c('\t');
		if(rn) { for(size_t i = 0; i < rn->size(); i++) { if(isspace((unsigned char)(*rn)[i])) break; w.c((*rn)[i]); } } else w.c('*');
		w.c('\t'); w.i((long long)a.refoff + 1); w.c('\t');
		w.i(mapq_v2(hasSec, best, secbest, scMin, scPer, primary, mhitsSet, false));
		w.c('\t'); w.s(cig.data(), cig.size()); w.lit("\t*\t0\t0\t");
		if(rdlen == 0) w.lit("*\t*");
		else {
			if(a.fw) { for(size_t i = 0; i < rdlen; i++) w.p[i] = "ACGTN"[seq[i]]; }
			else     { for(size_t i = 0; i < rdlen; i++) w.p[i] = "TGCAN"[seq[rdlen - 1 - i]]; }
			w.p[rdlen] = '\t'; w.p += rdlen + 1;
			if(a.fw) memcpy(w.p, qual, rdlen);
			else     { for(size_t i = 0; i < rdlen; i++) w.p[i] = (char)qual[rdlen - 1 - i]; }
			w.p += rdlen;
		}
		w.lit("\tAS:i:"); w.i(a.score);
		if(hasSec) { w.lit("\tXS:i:"); w.i((long long)secbest); }
		w.lit("\tXN:i:"); w.i(a.refns);
		w.lit("\tXM:i:"); w.i((long long)num_mm);
		w.lit("\tXO:i:"); w.i((long long)num_go);
		w.lit("\tXG:i:"); w.i((long long)num_gx);
		w.lit("\tNM:i:"); w.i((long long)ne);
		w.lit("\tMD:Z:"); w.s(md.data(), md.size());
		w.lit("\tYT:Z:UU");
		if(opt.yfQuirk) w.lit("\tYF:Z:LN");
		w.c('\n');
		o.resize((size_t)(w.p - o.data()));
	}
}

std::string sam_header(const Job& job) {
	std::string h = "@HD\tVN:1.5\tSO:unsorted\tGO:query\n";
	for(size_t i = 0; i < job.refnames.size(); i++) {
		h += "@SQ\tSN:";
		const std::string& nm = job.refnames[i];
		for(size_t k = 0; k < nm.size(); k++) { if(isspace((unsigned char)nm[k])) break; h.push_back(nm[k]); }
		h += "\tLN:" + std::to_string(i < job.reflens.size() ? job.reflens[i] : 0) + "\n";
	}
	h += "@PG\tID:bowtie2\tPN:bowtie2\tVN:2.5.4\tCL:\"bt2gpu " + job.opt.cmdline + "\"\n";
	return h;
}

} // namespace bt2host
