/*
 * bt2gpu.h -- C-ABI of libbt2gpu.so: the B200 (sm_100a) implementation of the
 * seed-and-extend hot path of sfiligoi/omp-bowtie2-prime.
 *
 * The reference has no plugin/FFI boundary for this path; the stages are plain
 * C++ member calls made from the batch loop of multiseedSearchWorker()
 * (bt2_search.cpp:2414-2862).  This header introduces the boundary exactly
 * around the phases that loop calls between "reads are in their slots" (P0) and
 * "AlnRes objects are in the sink" (P7):
 *
 *   bt2gpu_align_batch() replaces, for a whole batch of reads and for ALL seed
 *   rounds of each read:
 *     MultiSeedResults::prepareOneSeed / instantiateSeeds   bt2_search.cpp:2584,2596   aligner_seed.cpp:321,411
 *     MultiSeedAligner::reserveBuffers / prepareSearchAllSeedsOne          :2607,2624   aligner_seed.cpp:561,478
 *     MultiSeedAligner::searchAllSeedsDoAll (searchSeedBi<8> + extend)      :2638        aligner_seed.cpp:674,820,603
 *     MultiSeedAligner::searchAllSeedsOneFinalize, SeedResults::rankSeedHits:2658,2668   aligner_seed.cpp:508 aligner_seed.h:1000
 *     AlignmentMultiCache::finalize                                         :2674        aligner_cache.h
 *     SwDriver::prioritizeSATups                                            :2701        aligner_sw_driver.cpp:61
 *     SwAligner::initRead, SwDriver::extendSeeds                            :2741,2758   aligner_sw.cpp:35 aligner_sw_driver.cpp:278
 *     the re-seed decision                                                  :2784-2799, 2819-2826
 *   and returns what those calls leave behind in the host objects: the
 *   alignments handed to AlnSinkWrap::report() (aln_sink.cpp:1395) in call
 *   order, the per-read RandomSource state (consumed again by
 *   AlnSinkWrap::finishRead -> selectByScore, aln_sink.cpp:1477), the final
 *   minsc, and the PerReadMetrics counters that steer control flow.
 *
 * Plain pointers and sizes only; no C++ or torch types cross this boundary.
 * All functions return 0 on success and a negative code on error; the message
 * is available through bt2gpu_last_error().  Nothing throws across the ABI.
 */
#ifndef BT2GPU_H_
#define BT2GPU_H_

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define BT2GPU_ABI_VERSION 4

/* Limits inherited from the reference's fixed DP buffers
 * (aligner_result.h:42-43 ALN_MAX_ROWS/ALN_MAX_COLS, aligner_sw_driver.h:95 ALN_MAX_ITER). */
#ifndef BT2GPU_MAX_ROWS      /* a build-time constant, as in the reference: libbt2gpu_long.so is the same source compiled with   */
#define BT2GPU_MAX_ROWS 160  /* -DBT2GPU_MAX_ROWS=1000 -DBT2GPU_MAX_COLS=1072 (reads up to 1 kbp; SURVEY 8f rank 3, beyond the     */
#endif                       /* reference's own limits).  bt2gpu_max_rows() / bt2gpu_max_cols() tell which build was loaded.        */
#ifndef BT2GPU_MAX_COLS
#define BT2GPU_MAX_COLS 200
#endif
#define BT2GPU_MAX_ITER 800

/* Return codes of SwDriver::extendSeeds (aligner_sw_driver.h:274-291) plus the
 * ways a read can leave the batch loop without ever reaching it. */
enum {
	BT2GPU_ST_NONE               = 0, /* never reached extendSeeds (see flags)          */
	BT2GPU_ST_EXHAUSTED          = 1, /* EXTEND_EXHAUSTED_CANDIDATES                    */
	BT2GPU_ST_POLICY_FULFILLED   = 2, /* EXTEND_POLICY_FULFILLED                        */
	BT2GPU_ST_PERFECT_SCORE      = 3, /* EXTEND_PERFECT_SCORE                           */
	BT2GPU_ST_SOFT_LIMIT         = 4, /* EXTEND_EXCEEDED_SOFT_LIMIT                     */
	BT2GPU_ST_HARD_LIMIT         = 5  /* EXTEND_EXCEEDED_HARD_LIMIT                     */
};

/* bt2gpu_read_result.flags */
#define BT2GPU_RF_NO_SEED_HITS  1u /* searchAllSeedsOneFinalize returned false / nonz==0 (bt2_search.cpp:2658-2665) */
#define BT2GPU_RF_TOO_LONG      2u /* SwAligner::initRead failed: rdlen > ALN_MAX_ROWS (aligner_sw.cpp:59)          */
#define BT2GPU_RF_EMPTY         4u /* zero-length read                                                          */

/* Edit types, same numeric values as edit.h (EDIT_TYPE_*). */
enum {
	BT2GPU_EDIT_READ_GAP = 1,
	BT2GPU_EDIT_REF_GAP  = 2,
	BT2GPU_EDIT_MM       = 3
};

/*
 * The forward FM index exactly as Ebwt::loadIntoMemory leaves it in host
 * memory (accessors bt2_idx.h:1252-1271; file layout bt2_io.cpp:134-600).
 * Offsets are 64-bit (the reference only builds the -l binary, btypes.h:24-31).
 * The library copies (and may re-lay out) everything into HBM at create time;
 * the caller keeps ownership and may free the arrays after bt2gpu_create().
 */
typedef struct bt2gpu_index {
	uint64_t        len;        /* EbwtParams::_len (joined text length, no '$')   */
	int32_t         lineRate;   /* side = 1<<lineRate bytes                        */
	int32_t         offRate;    /* SA sample every 1<<offRate rows                 */
	int32_t         ftabChars;
	int32_t         reserved0;
	uint64_t        nPat;       /* # reference sequences                           */
	const uint64_t *plen;       /* [nPat]     lengths of the references            */
	uint64_t        nFrag;      /* # N-free fragments                              */
	const uint64_t *rstarts;    /* [3*nFrag]  (joined off, text id, text off)      */
	const uint8_t  *ebwt;       /* [ebwtTotLen] sides: BWT bytes + 4 x u64 occ     */
	uint64_t        ebwtTotLen;
	uint64_t        zOff;       /* BWT row of '$'                                  */
	const uint64_t *fchr;       /* [5]                                             */
	const uint64_t *ftab;       /* [ftabLen = 4^ftabChars + 1]                     */
	uint64_t        ftabLen;
	const uint64_t *eftab;      /* [eftabLen = 2*ftabChars]                        */
	uint64_t        eftabLen;
	const uint64_t *offs;       /* [offsLen]  sampled suffix array                 */
	uint64_t        offsLen;
} bt2gpu_index;

/*
 * The 2-bit reference exactly as BitPairReference holds it (reference.h:169-178;
 * file layout reference.cpp:104-150 for .3.bt2l, .4.bt2l is buf).
 */
typedef struct bt2gpu_ref {
	uint64_t        nrecs;      /* # RefRecords                                    */
	const uint64_t *rec_off;    /* [nrecs] # Ns preceding the stretch              */
	const uint64_t *rec_len;    /* [nrecs] # unambiguous bases in the stretch      */
	const uint8_t  *rec_first;  /* [nrecs] 1 iff first record of a reference       */
	const uint8_t  *buf;        /* 2-bit packed unambiguous bases, 4 per byte      */
	uint64_t        bufBases;   /* # bases in buf (sum of rec_len)                 */
} bt2gpu_ref;

/*
 * Everything the path reads from the reference's file-static option globals
 * (captured in msWorkerConsts, bt2_search.cpp:2194-2237) and from Scoring
 * (scoring.h).  Immutable for the lifetime of a context.
 */
typedef struct bt2gpu_params {
	int32_t abi_version;     /* must be BT2GPU_ABI_VERSION                              */
	int32_t seedLen;         /* multiseedLen   (-L)                                     */
	int32_t seedMms;         /* multiseedMms   (-N); searched exactly either way (F3)   */
	int32_t nSeedRounds;     /* 1 + -R                                                  */
	int32_t seedBoostThresh; /* 300 (bt2_search.cpp:325)                                */
	int32_t maxhalf;         /* --dpad                                                  */
	int32_t mxIter;          /* maxIters + (k-1)*maxItersIncr, <= BT2GPU_MAX_ITER       */
	int32_t mxUg;
	int32_t mxDp;
	int32_t streak;          /* maxDpStreak (-D) + (k-1)*maxStreakIncr                  */
	int32_t doExtend;        /* seed left-extension (nlex) on                           */
	int32_t enable8;         /* u8 DP allowed; --no-sse8 clears it, which matters only with i16 = 1 (the default build has no other DP) */
	int32_t tighten;         /* --tighten, default 3                                    */
	int32_t khits;           /* -k                                                      */
	int32_t mhits;           /* -M (0 = unset)                                          */
	int32_t msample;
	int32_t nofw;
	int32_t norc;
	int32_t gapbar;
	int32_t rdGapConst, rdGapLinear, rfGapConst, rfGapLinear;
	int32_t matchBonus;      /* 0 in end-to-end mode (the only mode, F5)                */
	int32_t i16;             /* 1: behave as the reference built with -DENABLE_I16: with enable8 == 0 its 16-bit
	                          * DP path (aligner_sw.cpp:239-241, 336-348) is taken; reads whose minsc < -254 make
	                          * bt2gpu_align_batch fail (16-bit score range not implemented)              */
	int32_t exactBestLtMinsc;/* 0 (default): K_fill computes only the band of DP diagonals an alignment scoring >= minsc can lie in
	                          * (every alignment, edit, counter and random draw is the full rectangle's); the best score of a
	                          * window WITHOUT any valid alignment, prm.bestLtMinscMate1 (printed only as Xs:i / ZT:Z with
	                          * --sam-opt-config, sam.cpp:287-296,544), is then a lower bound.  1: fill the whole rectangle
	                          * like EEU8_alignNucleotides (aligner_swsse_ee_u8.cpp:398) so that it is exact as well.    */
	int32_t reserved[6];
	int32_t mmpens[256];     /* Scoring::mmpens                                         */
	int32_t npens[256];      /* Scoring::npens                                          */
} bt2gpu_params;

/*
 * One batch of reads, SoA, host-owned, valid for the duration of the call.
 * What Read{patFw,qual,seed} (read.h:39-270) plus the per-slot scalars of
 * bt2_search.cpp:2485-2558 hold when phase P0 ends.
 */
typedef struct bt2gpu_reads {
	uint32_t        n;
	const uint32_t *seq_off;   /* [n+1] offsets into seq/qual                          */
	const uint8_t  *seq;       /* base codes 0..4 (A,C,G,T,N), forward strand          */
	const uint8_t  *qual;      /* Phred+33 ASCII                                       */
	const uint32_t *rng_seed;  /* [n] Read::seed (pat.cpp:45 genRandSeed)              */
	const int32_t  *minsc;     /* [n] scoreMin.f(rdlen), clamped as bt2_search.cpp:2484-2493 */
	const int32_t  *nceil;     /* [n] nCeil.f<int>(rdlen) (aligner_sw.cpp:44)          */
	const int32_t  *interval;  /* [n] max(1, msIval.f<int>(rdlen)) (:2551)             */
} bt2gpu_reads;

/* One Edit (edit.h): pos is the 0-based offset from the read's 5' end, chr the
 * reference character ('-' for a reference gap), qchr the read character
 * ('-' for a read gap). */
typedef struct bt2gpu_edit {
	uint32_t pos;
	uint8_t  chr;
	uint8_t  qchr;
	uint8_t  type;
	uint8_t  pad;
} bt2gpu_edit;

/* One AlnRes as handed to AlnSinkWrap::report (fields consumed by
 * AlnRes::setScore/setShape/setRefNs/setParams, aligner_swsse_ee_u8.cpp:1292-1320,
 * aligner_sw_driver.cpp:571-575). */
typedef struct bt2gpu_aln {
	int64_t  refoff;       /* 0-based offset of leftmost ref base in reference tidx */
	uint32_t tidx;
	int32_t  score;
	int32_t  minsc;        /* minsc in force when reported (AlnRes::setParams)      */
	int32_t  ns;
	int32_t  gaps;
	int32_t  edits;
	int32_t  basesAligned;
	int32_t  refns;
	uint32_t edit_off;     /* index of first edit in bt2gpu_results.edits           */
	uint16_t n_edits;
	uint8_t  fw;
	uint8_t  pad;
} bt2gpu_aln;

/* Per-read outcome: what the host must restore into its per-slot objects
 * before AlnSinkWrap::finishRead. */
typedef struct bt2gpu_read_result {
	uint32_t aln_off;      /* index of first alignment in bt2gpu_results.alns       */
	uint32_t n_alns;       /* alignments reported, in report() order                */
	uint32_t rng_last;     /* RandomSource::last                                    */
	uint32_t rng_lastOff;  /* RandomSource::lastOff                                 */
	int32_t  minsc;        /* final (tightened) minsc                               */
	uint8_t  status;       /* BT2GPU_ST_* of the last extendSeeds call              */
	uint8_t  flags;        /* BT2GPU_RF_*                                           */
	uint8_t  rounds;       /* seed rounds actually run                              */
	uint8_t  done_policy;  /* ReportingState::done() after the last report          */
	/* PerReadMetrics (read.h:339+) counters that the path maintains */
	uint32_t nExIters, nExDps, nExDpSuccs, nExDpFails;
	uint32_t nDpFail, nDpFailStreak, nDpLastSucc, nRedundants;
	int32_t  bestLtMinsc;  /* prm.bestLtMinscMate1, INT32_MIN if unset              */
	/* SeedResults summary of the last round (aligner_seed.h:1200-1213) */
	uint32_t nonzTot, numElts, numRanges, nSeedsTried;
	/* F14: the reference's per-slot edit scratch (SwDriver::resGap_.alres.ned_)
	 * as the last backtrace attempt of the read left it; the reference's SAM
	 * writer reads alignment edits through it. */
	uint32_t scratch_off;  /* index of first scratch edit in bt2gpu_results.edits; the image is
	                        * cut at the largest n_edits of the read's alignments     */
	uint32_t n_scratch;
	/* the rest of what bt2_search.cpp:2846-2854 hands to finishReadOne / what sam.cpp may print */
	uint32_t nExUgs;       /* prm.nExUgs: ungapped extensions run; always 0 (extendSeeds never runs one, aligner_sw_driver.cpp:371) */
	uint32_t exhaustive;   /* exhaustive[mate] of bt2_search.cpp:2399,2527: extendSeeds never sets it, always 0 */
} bt2gpu_read_result;

/* Library-owned (pinned) output, valid until the next call on the same ctx. */
typedef struct bt2gpu_results {
	uint32_t                  n;
	const bt2gpu_read_result *reads;   /* [n]       */
	const bt2gpu_aln         *alns;    /* [n_alns]  */
	uint64_t                  n_alns;
	const bt2gpu_edit        *edits;   /* [n_edits] */
	uint64_t                  n_edits;
} bt2gpu_results;

/* Work counters since create / last reset; feed the roofline arithmetic. */
typedef struct bt2gpu_counters {
	uint64_t reads;
	uint64_t seeds_searched;
	uint64_t lf_steps_search;   /* mapBiLFEx / mapLF1 steps in searchSeedBi        */
	uint64_t lf_steps_extend;   /* ... in MultiSeedAligner::extend                 */
	uint64_t lf_sides_touched;  /* distinct 128-byte sides read by those steps     */
	uint64_t sa_rows_resolved;
	uint64_t lf_steps_resolve;
	uint64_t dp_problems;
	uint64_t dp_cells;          /* sum rdlen*rflen of the DPs the control flow ran (what the reference computes) */
	uint64_t bt_attempts;
	uint64_t bt_steps;
	uint64_t alns_reported;
	/* CUDA-event totals (ms) on the library's stream, per phase: K_seed, glue kernels,
	 * extension-step kernel (SA resolve + framing + window fetch), K_fill, backtrace kernel,
	 * H2D, D2H, and whole bt2gpu_run_resident calls */
	double   ms_seed, ms_glue, ms_resolve, ms_fill, ms_bt, ms_h2d, ms_d2h, ms_run;
	uint64_t launches;          /* kernels launched by the library                 */
	uint64_t n_seed, n_step, n_fill, n_bt; /* launches of the four main kernels    */
	uint64_t ext_text_steps;    /* MultiSeedAligner::extend steps of single-row ranges, done as
	                             * direct text comparisons through the device's full suffix array
	                             * (each replaces one mapLF1 step of the reference)      */
	uint64_t dp_cells_computed; /* cells K_fill actually computed for them: rdlen x (32 lanes x K diagonals of the band)   */
	uint64_t sa_probes;         /* searchSeedBi done as binary searches over the device's full suffix array: probes
	                             * (one 4-byte suffix-array gather + 12 bytes of text each); lf_steps_search then only
	                             * counts the seeds that took the LF walk                                   */
} bt2gpu_counters;

typedef struct bt2gpu_ctx bt2gpu_ctx;

/* Create a context bound to CUDA device `device`; uploads index and reference. */
bt2gpu_ctx *bt2gpu_create(const bt2gpu_index *ix, const bt2gpu_ref *rf,
                          const bt2gpu_params *p, int device,
                          char *err, size_t errlen);

/* Align one batch (host buffers in, pinned host results out; H2D and D2H inside).
 * A context holds two sets of result arrays and successive calls take them in turn: what a call returned stays valid until the
 * next-but-one call (bt2gpu_align_batch or bt2gpu_submit) on the context starts -- a sink can still be reading batch i while
 * batch i+1 is aligned. */
int bt2gpu_align_batch(bt2gpu_ctx *ctx, const bt2gpu_reads *in, bt2gpu_results *out);

/* The same call in two halves, for a caller whose batch loop has one thread (the reference's: bt2_search.cpp:2414-2862, where
 * P7 finishReadOne of batch i would otherwise wait for the device): bt2gpu_submit returns at once and the batch is aligned on a
 * worker thread of the context; bt2gpu_wait blocks until it is done and hands out the results (same lifetime as above).  `in`
 * and its arrays must stay as they are until bt2gpu_wait returns.  One batch in flight per context; every other call on the
 * context except bt2gpu_last_error fails with -2 until bt2gpu_wait has been called. */
int bt2gpu_submit(bt2gpu_ctx *ctx, const bt2gpu_reads *in);
int bt2gpu_wait(bt2gpu_ctx *ctx, bt2gpu_results *out);

/* Split form used by bench.py to time the kernels with inputs already resident:
 * upload() copies a batch to HBM, run_resident() runs all stages on it (no
 * copies), fetch() brings the results back. align_batch == the three in order. */
int bt2gpu_upload(bt2gpu_ctx *ctx, const bt2gpu_reads *in);
int bt2gpu_run_resident(bt2gpu_ctx *ctx);
int bt2gpu_fetch(bt2gpu_ctx *ctx, bt2gpu_results *out);

/* What bt2gpu_create spent and holds (the index load path; the host side of it is Ebwt::readIntoMemory,
 * bt2_io.cpp:39-633, and BitPairReference's constructor, reference.cpp:50-181):
 *   ms[0] copies of the caller's index and reference arrays to HBM, ms[1] expansion of the sampled suffix array to a full one on the
 *   device, ms[2] K-mer prefix table over it, ms[3] the rest (problem slots, streams);  bytes[0] index + reference as uploaded,
 *   bytes[1] full suffix array, bytes[2] prefix table, bytes[3] per-batch working memory allocated at create. */
int bt2gpu_get_create_info(const bt2gpu_ctx *ctx, double ms[4], uint64_t bytes[4]);

int bt2gpu_get_counters(bt2gpu_ctx *ctx, bt2gpu_counters *c);
int bt2gpu_reset_counters(bt2gpu_ctx *ctx);
const char *bt2gpu_last_error(const bt2gpu_ctx *ctx);
void bt2gpu_destroy(bt2gpu_ctx *ctx);
/* Page-locked host memory (cudaMallocHost / cudaFreeHost) for a caller's input arrays: bt2gpu_align_batch copies from it without a
 * staging pass.  NULL when it cannot be had.  (bt2host_set_array_allocator takes this pair.) */
void *bt2gpu_host_alloc(size_t bytes);
void bt2gpu_host_free(void *p);
int bt2gpu_abi_version(void);
int bt2gpu_max_rows(void);   /* BT2GPU_MAX_ROWS / BT2GPU_MAX_COLS this library was built with */
int bt2gpu_max_cols(void);

/* Stage-level test hook (not part of the reference boundary): runs the DP-fill kernel
 * on nprob independent problems.  Problem i uses rows[i] read bases at seq/qual +
 * seq_off[i] (already oriented: the strand buffer the fill sees), cols[i] reference
 * columns whose masks (1<<base, N=16; cols[i]+1 entries) start at rf + rf_off[i], and
 * minsc[i].  mode 0: the test instantiation of the kernel over the WHOLE rectangle; it also
 * dumps dense raw-u8 H/E/F matrices (row-major rows x cols) at hef_off[i] in H/E/F.
 * mode 1: the production instantiation over the band of diagonals bt2gpu_align_batch would
 * compute (H/E/F untouched); mode 2: the same with the widest lane layout (K = 16, the one
 * bt2gpu_fill_noalloc uses).  Outputs, per problem: the byte of backtrace comparison flags
 * stored per cell in `flags` (same indexing as H; 0x80 for a cell outside the computed band; bit
 * meanings: DESIGN.md "cell flags"), the sorted candidate list ((rawH<<8)|col) at
 * cand + 256*i, ncand[i], lrmax[i], and band[2*i], band[2*i+1] = first diagonal (col - row)
 * and number of diagonals computed.  nprob <= 4096. */
int bt2gpu_test_fill(bt2gpu_ctx *ctx, uint32_t nprob, const uint32_t *seq_off, const uint8_t *seq,
                     const uint8_t *qual, const uint16_t *rows, const uint16_t *cols,
                     const uint32_t *rf_off, const uint8_t *rf, const int32_t *minsc,
                     const uint64_t *hef_off, uint8_t *H, uint8_t *E, uint8_t *F, uint8_t *flags,
                     uint16_t *cand, uint32_t *ncand, int32_t *lrmax, int32_t mode, int32_t *band);

/* DP fill without any device allocation.  NO COUNTERPART IN THE REFERENCE: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The same production fill kernel as bt2gpu_align_batch
 * runs, on the same per-problem inputs and with the same per-problem outputs as bt2gpu_test_fill
 * (minus the debug H/E/F matrices), but
 *   - every pointer is a DEVICE pointer supplied by the caller (inputs and outputs alike), and the
 *     work is only ENQUEUED on `stream` (a cudaStream_t; NULL = the legacy default stream): the
 *     outputs are valid once the caller has synchronised that stream;
 *   - the kernel's scratch (problem records, cell flags) is a file-scope device array of fixed
 *     size, ~11.6 MB, held by every GPU that loads the library until the process exits: at most
 *     BT2GPU_NOALLOC_MAX_PROBLEMS problems per call, more is rejected with an error;
 *   - the call itself performs no cudaMalloc / cudaFree and calls no allocating library routine
 *     (the context it is given was allocated by bt2gpu_create beforehand).
 * All calls on a device share that one array: two calls whose kernels overlap on the GPU (different
 * streams, different contexts or processes' threads) race; issue them on one stream or order them
 * with events.  What has to start at zero is cleared by the call.  A problem whose size is out of
 * range (rows 1..160, cols 1..199) gets ncand = 0, lrmax = -1 and no flags.
 * flags: rows[i] x cols[i] bytes, row-major, at flags + flags_off[i] (0x80 for a cell outside the band of
 * diagonals the kernel computes for the problem's minsc); cand: 256 entries per problem. */
#if BT2GPU_MAX_ROWS > 160
#define BT2GPU_NOALLOC_MAX_PROBLEMS 16    /* the long-read build: a problem's cell flags take 32 times the room */
#else
#define BT2GPU_NOALLOC_MAX_PROBLEMS 128
#endif
int bt2gpu_fill_noalloc(bt2gpu_ctx *ctx, void *stream, uint32_t nprob,
                        const uint32_t *d_seq_off, const uint8_t *d_seq, const uint8_t *d_qual,
                        const uint16_t *d_rows, const uint16_t *d_cols,
                        const uint32_t *d_rf_off, const uint8_t *d_rf, const int32_t *d_minsc,
                        const uint64_t *d_flags_off, uint8_t *d_flags,
                        uint16_t *d_cand, uint32_t *d_ncand, int32_t *d_lrmax);

#ifdef __cplusplus
}
#endif
#endif /* BT2GPU_H_ */
