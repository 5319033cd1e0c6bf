/*
 * bt2host.h -- C-ABI of libbt2host.so: the host-side pieces that sit either
 * side of the GPU hot path when it is used WITHOUT the reference's own host
 * code (bench.py, the tests, the bt2gpu-align CLI).  They mirror, for the
 * unpaired end-to-end mode the reference supports:
 *
 *   index files    Ebwt::readIntoMemory            bt2_io.cpp:39-633
 *                  BitPairReference ctor           reference.cpp:31-181
 *   options        parseOption / presets           bt2_search.cpp:1137-1578, presets.cpp:28-66
 *                  SeedAlignmentPolicy::parseString aligner_seed_policy.cpp
 *                  SimpleFunc::f                   simple_func.h:89-110
 *   read input     FastqPatternSource::parse       pat.cpp:1124-1235, genRandSeed pat.cpp:45-83
 *   per-read P0    minsc / nceil / interval        bt2_search.cpp:2484-2558
 *   sink           AlnSinkWrap::finishRead         aln_sink.cpp:643-1384
 *                  AlnSinkWrap::selectByScore      aln_sink.cpp:1477-1627
 *                  BowtieMapq2::mapq               unique.h:170-330
 *                  AlnSinkSam::appendMate          aln_sink.cpp:1889-2124
 *                  SamConfig::printAlignedOptFlags sam.cpp:121-470
 *                  StackedAln (CIGAR / MD:Z)       aligner_result.cpp
 *
 * In a real drop-in these stay the reference's own C++; this
 * library exists so that the GPU path can be driven and checked end to end
 * (FASTQ in, SAM out) on a box that has no reference sources.
 */
#ifndef BT2HOST_H_
#define BT2HOST_H_

#include "bt2gpu.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct bt2host_job   bt2host_job;
typedef struct bt2host_batch bt2host_batch;

/* Load <base>.{1,2,3,4}.bt2l and parse a bowtie2-style option string
 * (e.g. "--sensitive --dpad 12 -k 3"). Returns NULL and fills err on failure. */
bt2host_job *bt2host_job_create(const char *index_base, const char *options,
                                char *err, size_t errlen);
void bt2host_job_destroy(bt2host_job *job);

const bt2gpu_index  *bt2host_job_index(const bt2host_job *job);
const bt2gpu_ref    *bt2host_job_ref(const bt2host_job *job);
const bt2gpu_params *bt2host_job_params(const bt2host_job *job);
/* Number of reference sequences and their names (as printed in @SQ / RNAME). */
uint64_t    bt2host_job_nrefs(const bt2host_job *job);
const char *bt2host_job_refname(const bt2host_job *job, uint64_t i);

/* Parse up to max_reads FASTQ records starting at record `skip` (0-based) of
 * `path`; fills the per-read scalars of phase P0.  max_reads==0 means all. */
bt2host_batch *bt2host_batch_from_fastq(const bt2host_job *job, const char *path,
                                        uint64_t skip, uint64_t max_reads,
                                        char *err, size_t errlen);
/* Streaming variant: the records that lie completely inside the byte range
 * [offset, offset + max_bytes) of the file (offset must be a record start; 0 to begin).
 * *next_offset receives the offset of the first record not read; at the end of the
 * file the batch is empty (n = 0).  Files the block reader declines (CR line ends, blank
 * lines, a truncated record) return NULL with a message: read them whole instead. */
bt2host_batch *bt2host_batch_from_fastq_range(const bt2host_job *job, const char *path,
                                              uint64_t offset, uint64_t max_bytes,
                                              uint64_t *next_offset, char *err, size_t errlen);
/* Build a batch from in-memory reads (names may be NULL -> "0","1",...). seq is
 * ASCII (ACGTN), qual Phred+33 ASCII. */
bt2host_batch *bt2host_batch_from_arrays(const bt2host_job *job, uint32_t n,
                                         const uint32_t *off, const char *seq,
                                         const char *qual, const char *const *names,
                                         char *err, size_t errlen);
void bt2host_batch_destroy(bt2host_batch *b);
const bt2gpu_reads *bt2host_batch_reads(const bt2host_batch *b);
const char *bt2host_batch_name(const bt2host_batch *b, uint32_t i);

/* SAM header (@HD, @SQ..., @PG) into a malloc'd string the caller frees with
 * bt2host_free. */
char *bt2host_sam_header(const bt2host_job *job);
/* Run the sink over a batch's results: returns a malloc'd buffer holding one
 * SAM record per read, in read order.  *len receives its length.
 * mode: 0 = well-formed CIGAR/MD from each alignment's own edits (what the
 *           reference prints once its DList shallow-copy bug is fixed, SURVEY F14, oracle E)
 *       1 = byte parity with the unmodified reference: edits are read through
 *           the per-read aliased scratch (oracle A). */
char *bt2host_sam_records(const bt2host_job *job, const bt2host_batch *b,
                          const bt2gpu_results *res, int mode, size_t *len);
/* The same records written to an open file descriptor (AlnSinkSam -> OutputQueue::flush,
 * aln_sink.cpp:1477, outq.cpp): blocks of reads are formatted in parallel and written in
 * read order (side by side with pwrite when the descriptor is seekable), without building one
 * contiguous buffer.  Returns 0, or -1 when write(2) fails
 * (errno is write's); *len receives the number of bytes written. */
int bt2host_sam_write(const bt2host_job *job, const bt2host_batch *b,
                      const bt2gpu_results *res, int mode, int fd, size_t *len);
/* bt2host_sam_write in two halves, for a sink that formats batch i+1 while batch i is being written (the reference's AlnSinkSam::append
 * and OutputQueue::flush are two stages as well): bt2host_sam_format fills the job's block set `set` (0 or 1) and needs the batch and
 * its results only until it returns (*len: bytes formatted); bt2host_sam_flush writes that set to the descriptor like
 * bt2host_sam_write does.  One thread at a time per set. */
int bt2host_sam_format(const bt2host_job *job, const bt2host_batch *b, const bt2gpu_results *res, int mode, int set, size_t *len);
int bt2host_sam_flush(const bt2host_job *job, int set, int fd, size_t *len);
void bt2host_free(void *p);

/* Summary counters of the last bt2host_sam_records call sequence on this job
 * (ReportingMetrics, aln_sink.h:44-233): [0]=reads [1]=unaligned [2]=aligned once [3]=aligned >1 */
void bt2host_job_metrics(const bt2host_job *job, uint64_t out[4]);
/* OpenMP threads the FASTQ reader and the SAM sink of this job take (0: all there are, the default).  A pipeline that runs the
 * two at the same time on neighbouring batches (stream.py) splits the cores between them instead of letting both ask for all. */
void bt2host_job_set_threads(bt2host_job *job, int reader_threads, int sink_threads);
/* Where the batches' large arrays (bases, qualities) are allocated from now on: e.g. libbt2gpu's bt2gpu_host_alloc / bt2gpu_host_free
 * (page-locked memory: bt2gpu_align_batch then copies to the device without a staging pass).  Process-wide; to be called once, before
 * the first batch is made (a second call, or a call after batches exist, is refused with -1 / undefined respectively). */
int bt2host_set_array_allocator(void *(*alloc)(size_t), void (*release)(void *));

#ifdef __cplusplus
}
#endif
#endif /* BT2HOST_H_ */
