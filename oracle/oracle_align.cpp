// TEST INFRASTRUCTURE ONLY -- command-line driver: .bt2l index + FASTQ -> SAM
// through the oracle restatement and the bt2host sink.  Used to pin the oracle
// against the unmodified reference binary (oracle/_ref/bowtie2-align-l).
//   oracle_align <index_base> <reads.fq> "<options>" <mode 0|1> [nthreads] > out.sam
#include <cstdio>
#include <cstdlib>
#include <string>
#include "bt2host.h"

extern "C" {
struct oracle_ctx;
oracle_ctx* oracle_create(const bt2gpu_index*, const bt2gpu_ref*, const bt2gpu_params*);
int oracle_align_batch(oracle_ctx*, const bt2gpu_reads*, bt2gpu_results*, int);
void oracle_destroy(oracle_ctx*);
void oracle_counters(oracle_ctx*, uint64_t out[10]);
}

int main(int argc, char** argv) {
	if(argc < 5) { fprintf(stderr, "usage: oracle_align <index_base> <reads.fq> \"<options>\" <mode> [nthreads]\n"); return 2; }
	char err[512] = {0};
	bt2host_job* job = bt2host_job_create(argv[1], argv[3], err, sizeof(err));
	if(!job) { fprintf(stderr, "error: %s\n", err); return 1; }
	bt2host_batch* b = bt2host_batch_from_fastq(job, argv[2], 0, 0, err, sizeof(err));
	if(!b) { fprintf(stderr, "error: %s\n", err); return 1; }
	oracle_ctx* c = oracle_create(bt2host_job_index(job), bt2host_job_ref(job), bt2host_job_params(job));
	bt2gpu_results res;
	oracle_align_batch(c, bt2host_batch_reads(b), &res, argc > 5 ? atoi(argv[5]) : 0);
	char* h = bt2host_sam_header(job); fputs(h, stdout); bt2host_free(h);
	size_t len = 0;
	char* s = bt2host_sam_records(job, b, &res, atoi(argv[4]), &len);
	fwrite(s, 1, len, stdout); bt2host_free(s);
	uint64_t m[4]; bt2host_job_metrics(job, m);
	uint64_t w[10]; oracle_counters(c, w);
	fprintf(stderr, "%lu reads; unaligned %lu; aligned once %lu; aligned >1 %lu\n", m[0], m[1], m[2], m[3]);
	fprintf(stderr, "seeds %lu lfSearch %lu lfExtend %lu saRows %lu lfResolve %lu dps %lu cells %lu btAttempts %lu btSteps %lu alns %lu\n",
	        w[0], w[1], w[2], w[3], w[4], w[5], w[6], w[7], w[8], w[9]);
	oracle_destroy(c); bt2host_batch_destroy(b); bt2host_job_destroy(job);
	return 0;
}
