// Internal structures of libbt2host.so (see include/bt2host.h for what each
// piece mirrors in the reference).
#pragma once
#include <cstdint>
#include <cstring>
#include <string>
#include <vector>
#include <memory>
#include <mutex>
#include <utility>
#include <new>
#include <cstdlib>
#include <cmath>
#include <limits>
#include "bt2host.h"

namespace bt2host {

// SimpleFunc (simple_func.h:60-110): f(x) = max(I, min(X, C + L*g(x))), cast by truncation.
struct SimpleFunc {
	int type = 0; // 1 const, 2 linear, 3 sqrt, 4 log  (SIMPLE_FUNC_*)
	double I = -std::numeric_limits<double>::max();
	double X = std::numeric_limits<double>::max();
	double C = 0.0, L = 0.0;
	void init(int t, double c, double l) { type = t; C = c; L = l;
		I = -std::numeric_limits<double>::max(); X = std::numeric_limits<double>::max(); }
	void init(int t, double i, double x, double c, double l) { type = t; I = i; X = x; C = c; L = l; }
	template<typename T> T f(double x) const {
		double g;
		if(type == 1) g = 0.0; else if(type == 2) g = x;
		else if(type == 3) g = std::sqrt(x); else g = std::log(x);
		double ret = std::max(I, std::min(X, C + L * g));
		if(ret == std::numeric_limits<double>::max()) return std::numeric_limits<T>::max();
		if(ret == std::numeric_limits<double>::min()) return std::numeric_limits<T>::min();
		return (T)ret;
	}
	// "S,1,1.15" style (bt2_search.cpp:1436-1474, simple_func.cpp parse)
	bool parse(const std::string& s, double defC, double defL, double defI, double defX);
};

struct Options {
	bt2gpu_params p;
	SimpleFunc scoreMin, nCeil, msIval;
	uint32_t seed = 0;       // --seed
	int trim5 = 0, trim3 = 0;
	bool phred64 = false;
	int mapqv = 2;
	bool yfQuirk = true;     // print YF:Z:LN on every record like the reference does in practice
	bool noUnal = false;
	std::string cmdline;
};

struct BatchPool;
struct Job {
	// index arrays (owned)
	std::vector<uint64_t> plen, rstarts, fchr, ftab, eftab, offs;
	std::vector<uint8_t> ebwt;
	std::vector<uint64_t> rec_off, rec_len;
	std::vector<uint8_t> rec_first, refbuf;
	std::vector<std::string> refnames;
	std::vector<uint64_t> reflens; // full reference lengths incl. Ns (for @SQ LN)
	bt2gpu_index ix;
	bt2gpu_ref rf;
	Options opt;
	uint64_t met[4] = {0, 0, 0, 0};
	// buffers that keep their pages between calls (a job is used by one caller thread at a time)
	std::vector<std::string> sinkparts[2];   // SAM text of the sink's blocks, two sets: one can be written while the other is formatted
	char* file_buffer(size_t n);          // >= n bytes, not initialised, owned by the job
	char* filemem = nullptr; size_t filecap = 0;
	mutable std::vector<std::vector<size_t>> scan_nl;   // the reader's per-thread newline positions
	mutable std::vector<size_t> scan_ls; mutable std::vector<uint32_t> scan_len;
	std::shared_ptr<BatchPool> pool;      // set by bt2host_job_create
	int reader_threads = 0, sink_threads = 0;   // bt2host_job_set_threads: 0 = all the OpenMP threads there are
	mutable uint64_t range_next_id = 0;   // reads handed out by batch_from_fastq_range so far (ranges are taken in file order, from offset 0)
	~Job();
};

// vector storage that is not zero-filled by resize(): the readers overwrite every element (and do so on several threads)
// Where the large per-batch arrays (bases, qualities) come from: malloc, or a pair of functions the caller registered once with
// bt2host_set_array_allocator -- libbt2gpu's bt2gpu_host_alloc / bt2gpu_host_free hand out page-locked memory, from which
// bt2gpu_align_batch copies to the device without a staging pass.  (Process-wide, set before the first batch; this library itself
// stays free of CUDA.)
struct ArrayAllocator { void* (*alloc)(size_t); void (*release)(void*); };
extern ArrayAllocator g_array_allocator;
template<typename T> struct NoInitAlloc {
	typedef T value_type;
	NoInitAlloc() = default;
	template<typename U> NoInitAlloc(const NoInitAlloc<U>&) {}
	template<typename U> struct rebind { typedef NoInitAlloc<U> other; };
	// (a 64-byte header in front of every array remembers who handed it out: arrays made before the hook was set go back to free())
	T* allocate(size_t n) {
		const bool hook = g_array_allocator.alloc != nullptr;
		char* p = (char*)(hook ? g_array_allocator.alloc(n * sizeof(T) + 64) : malloc(n * sizeof(T) + 64));
		if(!p) throw std::bad_alloc();
		*(uint64_t*)p = hook ? 1u : 0u;
		return (T*)(p + 64);
	}
	void deallocate(T* q, size_t) { char* p = (char*)q - 64; if(*(uint64_t*)p) g_array_allocator.release(p); else free(p); }
	template<typename U> void construct(U* p) noexcept { ::new((void*)p) U; }
	template<typename U, typename... A> void construct(U* p, A&&... a) { ::new((void*)p) U(std::forward<A>(a)...); }
	template<typename U> bool operator==(const NoInitAlloc<U>&) const { return true; }
	template<typename U> bool operator!=(const NoInitAlloc<U>&) const { return false; }
};
typedef std::vector<uint8_t, NoInitAlloc<uint8_t>> ByteVec;

struct BatchPool;
struct Batch {
	std::shared_ptr<BatchPool> pool;   // where the batch's arrays go when it is destroyed (a streaming reader's batches: no fresh pages per range)
	std::vector<uint32_t> off;
	ByteVec seq, qual;
	std::vector<uint32_t> seed;
	std::vector<int32_t> minsc, nceil, interval;
	// read names, NUL-terminated, back to back: name i is name_blob[name_off[i] .. name_off[i+1] - 1)
	ByteVec name_blob; std::vector<uint32_t> name_off;
	size_t nnames() const { return name_off.empty() ? 0 : name_off.size() - 1; }
	const char* name_ptr(size_t i) const { return (const char*)name_blob.data() + name_off[i]; }
	size_t name_len(size_t i) const { return name_off[i + 1] - name_off[i] - 1; }
	void add_name(const char* p, size_t n) {
		if(name_off.empty()) name_off.push_back(0);
		name_blob.insert(name_blob.end(), (const uint8_t*)p, (const uint8_t*)p + n); name_blob.push_back(0);
		name_off.push_back((uint32_t)name_blob.size());
	}
	uint64_t first_id = 0;   // global index (Read::rdid) of the batch's first read: a read without a name is named by its rdid
	bt2gpu_reads view;
	void finalize();
};

} // namespace bt2host
struct bt2host_batch { bt2host::Batch b; };
namespace bt2host {
// batches given back by bt2host_batch_destroy, for the next bt2host_batch_from_fastq_range of the same job (their vectors keep
// their capacity, so a steady stream of ranges allocates nothing); shared by the job and its batches, whichever goes last
struct BatchPool { std::mutex mu; std::vector<bt2host_batch*> idle; ~BatchPool() { for(bt2host_batch* b : idle) delete b; } };

struct SinkScratch { struct Impl; Impl* impl; SinkScratch(); ~SinkScratch(); SinkScratch(const SinkScratch&) = delete; SinkScratch& operator=(const SinkScratch&) = delete; };
bool load_index(Job& job, const std::string& base, std::string& err);
bool parse_options(Options& o, const std::string& args, std::string& err);
uint32_t gen_rand_seed(const uint8_t* seq, const uint8_t* qual, size_t len,
                       const char* name, size_t name_len, uint32_t seed);
void fill_p0(const Job& job, Batch& b); // minsc / nceil / interval per read
bool batch_from_fastq(const Job& job, Batch& b, const std::string& path, uint64_t skip, uint64_t maxn, std::string& err);
bool batch_from_fastq_range(const Job& job, Batch& b, const std::string& path, uint64_t offset, uint64_t max_bytes, uint64_t* next_off, std::string& err);
bool batch_from_arrays(const Job& job, Batch& b, uint32_t n, const uint32_t* off, const char* seq,
                       const char* qual, const char* const* names);

} // namespace bt2host

struct bt2host_job   { bt2host::Job j; };
