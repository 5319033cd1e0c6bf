// C-ABI of libbt2host.so (include/bt2host.h).
#include <unistd.h>
#include <fcntl.h>
#include <sys/mman.h>
#include <sys/stat.h>
#include <cerrno>
#include "host_internal.hpp"
#include <cstdio>
#include <ctime>
#include <cstdlib>
#include <algorithm>
#include <vector>
#ifdef _OPENMP
#include <omp.h>
#endif

namespace bt2host {
void sam_for_read(const Job& job, uint64_t met[4], const Batch& b, const bt2gpu_results& res, uint32_t ri, int mode, std::string& o, SinkScratch& scratch);
std::string sam_header(const Job& job);
}

namespace bt2host { ArrayAllocator g_array_allocator = { nullptr, nullptr }; }

static void set_err(char* err, size_t errlen, const std::string& msg) {
	if(err && errlen) { snprintf(err, errlen, "%s", msg.c_str()); }
}

extern "C" {

bt2host_job* bt2host_job_create(const char* index_base, const char* options, char* err, size_t errlen) {
	bt2host_job* job = new bt2host_job();
	job->j.pool = std::make_shared<bt2host::BatchPool>();
	std::string e;
	if(!bt2host::parse_options(job->j.opt, options ? options : "", e) ||
	   !bt2host::load_index(job->j, index_base ? index_base : "", e)) {
		set_err(err, errlen, e);
		delete job;
		return nullptr;
	}
	return job;
}
void bt2host_job_destroy(bt2host_job* job) { delete job; }
const bt2gpu_index*  bt2host_job_index(const bt2host_job* job)  { return &job->j.ix; }
const bt2gpu_ref*    bt2host_job_ref(const bt2host_job* job)    { return &job->j.rf; }
const bt2gpu_params* bt2host_job_params(const bt2host_job* job) { return &job->j.opt.p; }
uint64_t bt2host_job_nrefs(const bt2host_job* job) { return job->j.refnames.size(); }
const char* bt2host_job_refname(const bt2host_job* job, uint64_t i) {
	return i < job->j.refnames.size() ? job->j.refnames[i].c_str() : nullptr;
}

bt2host_batch* bt2host_batch_from_fastq(const bt2host_job* job, const char* path, uint64_t skip, uint64_t max_reads,
                                        char* err, size_t errlen) {
	bt2host_batch* b = new bt2host_batch();
	std::string e;
	if(!bt2host::batch_from_fastq(job->j, b->b, path, skip, max_reads, e)) { set_err(err, errlen, e); delete b; return nullptr; }
	return b;
}
bt2host_batch* bt2host_batch_from_fastq_range(const bt2host_job* job, const char* path, uint64_t offset, uint64_t max_bytes,
                                              uint64_t* next_offset, char* err, size_t errlen) {
	bt2host_batch* b = nullptr;
	{ bt2host::BatchPool& pl = *job->j.pool; std::lock_guard<std::mutex> g(pl.mu); if(!pl.idle.empty()) { b = pl.idle.back(); pl.idle.pop_back(); } }
	if(!b) b = new bt2host_batch();
	b->b.pool = job->j.pool;
	std::string e;
	if(!bt2host::batch_from_fastq_range(job->j, b->b, path, offset, max_bytes, next_offset, e)) { set_err(err, errlen, e); delete b; return nullptr; }
	return b;
}
bt2host_batch* bt2host_batch_from_arrays(const bt2host_job* job, uint32_t n, const uint32_t* off, const char* seq,
                                         const char* qual, const char* const* names, char* err, size_t errlen) {
	bt2host_batch* b = new bt2host_batch();
	if(!bt2host::batch_from_arrays(job->j, b->b, n, off, seq, qual, names)) { set_err(err, errlen, "bad arrays"); delete b; return nullptr; }
	return b;
}
void bt2host_batch_destroy(bt2host_batch* b) {
	if(!b) return;
	if(std::shared_ptr<bt2host::BatchPool> pl = std::move(b->b.pool)) {   // a range reader's batch: keep up to eight sets of arrays for the next ranges (a pipeline has six batches in flight)
		std::lock_guard<std::mutex> g(pl->mu);
		if(pl->idle.size() < 8) { pl->idle.push_back(b); return; }
	}
	delete b;
}
const bt2gpu_reads* bt2host_batch_reads(const bt2host_batch* b) { return &b->b.view; }
const char* bt2host_batch_name(const bt2host_batch* b, uint32_t i) {
	return i < b->b.nnames() ? b->b.name_ptr(i) : nullptr;
}

static char* dup_out(const std::string& s, size_t* len) {
	char* p = (char*)malloc(s.size() + 1);
	if(!p) return nullptr;
	memcpy(p, s.data(), s.size()); p[s.size()] = 0;
	if(len) *len = s.size();
	return p;
}
char* bt2host_sam_header(const bt2host_job* job) { return dup_out(bt2host::sam_header(job->j), nullptr); }
// AlnSinkWrap::finishRead -> AlnSinkSam::append per read (aln_sink.cpp:643,1477): reads are independent (their RNG state
// comes back with the results), so contiguous blocks of reads are formatted by OpenMP threads; the blocks are in read order
static std::vector<std::string>& sam_blocks(const bt2host_job* job, const bt2host_batch* b, const bt2gpu_results* res, int mode, int set = 0) {
	const uint32_t n = res->n;
	bt2host::Job& j = const_cast<bt2host_job*>(job)->j;
	int nt = 1;
#ifdef _OPENMP
	nt = std::max(1, std::min(j.sink_threads > 0 ? j.sink_threads : omp_get_max_threads(), (int)(n / 2048) + 1));
#endif
	std::vector<std::string>& parts = j.sinkparts[set & 1];   // the blocks' strings keep their capacity (and pages) from call to call
	if(parts.size() < (size_t)nt) parts.resize((size_t)nt);
	for(std::string& o : parts) o.clear();
	std::vector<uint64_t> mets((size_t)nt * 4, 0);
	#pragma omp parallel for num_threads(nt) schedule(static, 1)
	for(int t = 0; t < nt; t++) {
		const uint32_t lo = (uint32_t)((uint64_t)n * t / nt), hi = (uint32_t)((uint64_t)n * (t + 1) / nt);
		// the block's string and counters live on this thread's stack while it appends: the std::string headers (and the counters) of
		// neighbouring blocks share cache lines, and every append writes the length -- measured on 16 cores: formatting took as long
		// as on one thread
		std::string o = std::move(parts[(size_t)t]);
		uint64_t met[4] = { 0, 0, 0, 0 };
		o.clear();
		// (only ever grown: before C++20 std::string::reserve(n) with n below the capacity reallocates to n -- 6 MB of fresh pages per
		// block and call, which was most of the sink's time when the ranges' read counts differed by a few reads)
		if(o.capacity() < (size_t)(hi - lo) * 480) o.reserve((size_t)(hi - lo) * 512);
		bt2host::SinkScratch scratch;
		for(uint32_t i = lo; i < hi; i++) bt2host::sam_for_read(j, met, b->b, *res, i, mode, o, scratch);
		parts[(size_t)t] = std::move(o);
		for(int k = 0; k < 4; k++) mets[(size_t)t * 4 + k] = met[k];
	}
	for(int t = 0; t < nt; t++) for(int k = 0; k < 4; k++) j.met[k] += mets[(size_t)t * 4 + k];
	return parts;
}
char* bt2host_sam_records(const bt2host_job* job, const bt2host_batch* b, const bt2gpu_results* res, int mode, size_t* len) {
	std::vector<std::string>& parts = sam_blocks(job, b, res, mode);
	const int nt = (int)parts.size();
	size_t tot = 0; for(const std::string& o : parts) tot += o.size();
	char* out = (char*)malloc(tot + 1);
	if(!out) return nullptr;
	std::vector<size_t> offs((size_t)nt + 1, 0);
	for(int t = 0; t < nt; t++) offs[(size_t)t + 1] = offs[(size_t)t] + parts[(size_t)t].size();
	#pragma omp parallel for num_threads(nt) schedule(static, 1)
	for(int t = 0; t < nt; t++) memcpy(out + offs[(size_t)t], parts[(size_t)t].data(), parts[(size_t)t].size());
	out[tot] = 0;
	if(len) *len = tot;
	return out;
}
static double now_ms() { timespec ts; clock_gettime(CLOCK_MONOTONIC, &ts); return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6; }
// The sink in two halves, for a caller that formats batch i+1 while batch i is being written: bt2host_sam_format fills block set `set`
// (0 or 1) and needs the batch and its results only until it returns; bt2host_sam_flush writes that set to the descriptor.
int bt2host_sam_format(const bt2host_job* job, const bt2host_batch* b, const bt2gpu_results* res, int mode, int set, size_t* len) {
	static const bool trace = getenv("BT2HOST_TRACE") != nullptr;
	const double t0 = now_ms();
	std::vector<std::string>& parts = sam_blocks(job, b, res, mode, set);
	size_t tot = 0; for(const std::string& o : parts) tot += o.size();
	if(len) *len = tot;
	if(trace) fprintf(stderr, "[bt2host] sam: %u reads formatted in %.1f ms (%zu blocks, set %d)\n", res->n, now_ms() - t0, parts.size(), set & 1);
	return 0;
}
int bt2host_sam_write(const bt2host_job* job, const bt2host_batch* b, const bt2gpu_results* res, int mode, int fd, size_t* len) {
	if(bt2host_sam_format(job, b, res, mode, 0, nullptr) != 0) return -1;
	return bt2host_sam_flush(job, 0, fd, len);
}
int bt2host_sam_flush(const bt2host_job* job, int set, int fd, size_t* len) {
	static const bool trace = getenv("BT2HOST_TRACE") != nullptr;
	const double t1 = now_ms();
	bt2host::Job& j = const_cast<bt2host_job*>(job)->j;
	std::vector<std::string>& parts = j.sinkparts[set & 1];
	struct Tr { bool on; double t1; size_t np; ~Tr() { if(on) fprintf(stderr, "[bt2host] sam: written in %.1f ms (%zu blocks)\n", now_ms() - t1, np); } } tr{trace, t1, parts.size()};
	int ncopy = 1;   // threads that copy / write (the blocks may be more: they are what the formatting call's threads made)
#ifdef _OPENMP
	ncopy = std::max(1, j.sink_threads > 0 ? j.sink_threads : omp_get_max_threads());
#endif
	const int nt = (int)parts.size();
	std::vector<size_t> offs((size_t)nt + 1, 0);
	for(int t = 0; t < nt; t++) offs[(size_t)t + 1] = offs[(size_t)t] + parts[(size_t)t].size();
	const size_t tot = offs[(size_t)nt];
	auto put = [](int fd, const char* p, size_t n, off_t at) -> bool { // at < 0: write at the descriptor's position
		size_t done = 0;
		while(done < n) {
			const ssize_t w = at < 0 ? write(fd, p + done, n - done) : pwrite(fd, p + done, n - done, at + (off_t)done);
			if(w < 0) { if(errno == EINTR) continue; return false; }
			done += (size_t)w;
		}
		return true;
	};
	const off_t base = lseek(fd, 0, SEEK_CUR);
	const int fl = fcntl(fd, F_GETFL);
	const bool append = fl >= 0 && (fl & O_APPEND);   // pwrite ignores its offset on an O_APPEND descriptor (Linux): blocks would land in completion order
	bool ok = true;
	// A regular file opened for reading and writing takes the blocks through a shared mapping: the threads copy into the file's
	// pages side by side.  (write()/pwrite() to one file are serialised by the inode lock; on tmpfs that is one core's page
	// allocation + copy rate, ~1.5 GB/s, however many threads call it.)  BT2HOST_SAM_WRITE=pwrite keeps the system-call path.
	static const bool no_map = getenv("BT2HOST_SAM_WRITE") && !strcmp(getenv("BT2HOST_SAM_WRITE"), "pwrite");
	struct stat sb;
	if(base >= 0 && !append && !no_map && tot >= (1u << 16) && fl >= 0 && (fl & O_ACCMODE) == O_RDWR && fstat(fd, &sb) == 0 && S_ISREG(sb.st_mode)) {
		const off_t pgsz = (off_t)sysconf(_SC_PAGESIZE), from = base & ~(pgsz - 1);
		const size_t span = (size_t)(base + (off_t)tot - from);
		if((sb.st_size >= base + (off_t)tot || ftruncate(fd, base + (off_t)tot) == 0)) {
			void* m = mmap(nullptr, span, PROT_READ | PROT_WRITE, MAP_SHARED, fd, from);
			if(m != MAP_FAILED) {
				char* dst = (char*)m + (base - from);
				const size_t piece = 2u << 20;                      // pieces of 2 MB: blocks differ in size
				static const bool populate = getenv("BT2HOST_SAM_POPULATE") && atoi(getenv("BT2HOST_SAM_POPULATE")) != 0;   // (off: measured no faster than faulting the pages in from the copy)
				std::vector<std::pair<int, size_t>> work;
				for(int t = 0; t < nt; t++) for(size_t o = 0; o < parts[(size_t)t].size(); o += piece) work.push_back(std::make_pair(t, o));
				#pragma omp parallel for num_threads(ncopy) schedule(dynamic, 1)
				for(size_t w = 0; w < work.size(); w++) {
					const std::string& src = parts[(size_t)work[w].first]; const size_t o = work[w].second;
					char* to = dst + offs[(size_t)work[w].first] + o; const size_t nb = std::min(piece, src.size() - o);
#ifdef MADV_POPULATE_WRITE
					// the piece's pages in one call instead of a fault per page (Linux 5.14+; an older kernel says EINVAL and the copy faults them in)
					{ char* p0 = (char*)((uintptr_t)to & ~(uintptr_t)(pgsz - 1)); char* p1 = (char*)(((uintptr_t)(to + nb) + (uintptr_t)pgsz - 1) & ~(uintptr_t)(pgsz - 1));
					  if(p1 > (char*)m + span) p1 = (char*)m + span;
					  if(populate) (void)madvise(p0, (size_t)(p1 - p0), MADV_POPULATE_WRITE); }
#endif
					memcpy(to, src.data() + o, nb);
				}
				munmap(m, span);
				if(lseek(fd, base + (off_t)tot, SEEK_SET) < 0) ok = false;
				if(len) *len = ok ? tot : 0;
				return ok ? 0 : -1;
			}
		}
	}
	if(base >= 0 && nt > 1 && !append) {   // a seekable file: the blocks are written side by side at their final offsets
		#pragma omp parallel for num_threads(ncopy) schedule(static, 1) reduction(&&: ok)
		for(int t = 0; t < nt; t++) ok = ok && put(fd, parts[(size_t)t].data(), parts[(size_t)t].size(), base + (off_t)offs[(size_t)t]);
		if(ok && lseek(fd, base + (off_t)tot, SEEK_SET) < 0) ok = false;
	} else {
		for(int t = 0; t < nt && ok; t++) ok = put(fd, parts[(size_t)t].data(), parts[(size_t)t].size(), -1);
	}
	if(len) *len = ok ? tot : 0;
	return ok ? 0 : -1;
}
void bt2host_free(void* p) { free(p); }
int bt2host_set_array_allocator(void* (*alloc)(size_t), void (*release)(void*)) {
	static bool fixed = false;               // once: arrays remember nothing about who allocated them
	if(fixed || !alloc || !release) return -1;
	bt2host::g_array_allocator.release = release; bt2host::g_array_allocator.alloc = alloc; fixed = true;
	return 0;
}
void bt2host_job_set_threads(bt2host_job* job, int reader_threads, int sink_threads) {
	job->j.reader_threads = reader_threads > 0 ? reader_threads : 0; job->j.sink_threads = sink_threads > 0 ? sink_threads : 0;
}
void bt2host_job_metrics(const bt2host_job* job, uint64_t out[4]) { for(int i = 0; i < 4; i++) out[i] = job->j.met[i]; }

} // extern "C"
