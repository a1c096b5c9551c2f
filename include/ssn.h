/*
 * ssn.h — C ABI of the B200 tumor/normal pileup-column scorer (libssn.so).
 *
 * The reference (genome/somatic-sniper) has no plugin/FFI layer: its hot path sits behind the
 * in-process per-column callback
 *
 *     typedef int (*bam_sspileup_f)(uint32_t tid, uint32_t pos, int n1, int n2,
 *                                   const bam_pileup1_t *pl1, const bam_pileup1_t *pl2,
 *                                   void *data, FILE *snp_fh);          -- src/lib/sniper/somatic_sniper.h:42
 *
 * installed at src/exe/bam-somaticsniper/main.c:146 (glf_somatic) and invoked once per column at
 * src/lib/sniper/sniper_pileup.c:258.  This library replaces that callback at BATCH granularity:
 * the host packer writes both samples' columns as structure-of-arrays batches into pinned memory,
 * submits them, and drains `ssn_site` records (one per site that glf_somatic would have handed to
 * output_formatter_write, somatic_sniper.c:264) in column order.
 *
 * Plain C types only; no CUDA or torch types appear in any signature.  All functions return
 * SSN_OK (0) or a negative ssn_status; ssn_last_error() gives the message of the calling thread's
 * last failure.  The library never exits the process and has NO CPU fallback: without a usable
 * CUDA device ssn_create() fails with SSN_ERR_CUDA.
 *
 * Threading: one ssn_ctx per GPU; one submitting thread per ctx (the reference's callback is
 * single-threaded and non-reentrant, somatic_sniper.c:13-16).
 */
#ifndef SSN_H
#define SSN_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define SSN_ABI_VERSION 2

typedef enum {
    SSN_OK = 0,
    SSN_ERR_ARG = -1,      /* bad argument / batch contents */
    SSN_ERR_CUDA = -2,     /* CUDA runtime failure (fatal for the ctx) */
    SSN_ERR_NOMEM = -3,
    SSN_ERR_STATE = -4,    /* call out of order (e.g. results_next with nothing submitted) */
    SSN_ERR_OVERFLOW = -5  /* more passing sites than the batch's site ring holds */
} ssn_status;

/* Every knob of pu_data2_t (somatic_sniper.h:21-40) and sniper_maqcns_t (sniper_maqcns.h:13-21)
 * that affects scoring.  Defaults: main.c:69-78 and sniper_maqcns_init (sniper_maqcns.c:102-113). */
typedef struct {
    float theta;                  /* -T, 0.85  (float field in the reference) */
    float het_rate;               /* -r, 0.001 (float field) */
    float eta;                    /* 0.03, not settable from the reference CLI */
    int32_t n_hap;                /* -N, 2 */
    int32_t cap_mapq;             /* 60 */
    int32_t min_somatic_qual;     /* -Q, 15 */
    int32_t include_loh;          /* 1; -L clears */
    int32_t include_gor;          /* 1; -G clears */
    int32_t use_priors;           /* 1; -p clears (prior[][] stays all zero, main.c:122-125) */
    int32_t use_joint_priors;     /* 0; -J / -s set */
    double somatic_mutation_rate; /* -s, 0.01 */
} ssn_params;

/* Host views of the model tables the ctx built (for parity checks against the reference's
 * sniper_cal_coef / sniper_cal_het / makeSoloPrior / make_joint_prior / qAddTableInit). */
typedef struct {
    const double *coef;        /* [q<<16 | n<<8 | k], 64*256*256   sniper_maqcns.c:59-100 */
    const double *lhet;        /* [n1<<8 | n2], 256*256             sniper_maqcns.c:27-53  */
    const double *fk;          /* 256                               sniper_maqcns.c:70-74  */
    float q_r;                 /*                                   sniper_maqcns.c:54-55  */
    const int32_t *prior;      /* [16][10]                          somatic_sniper.c:29-45 */
    const int32_t *jointprior; /* [16][10][10]                      somatic_sniper.c:47-77 */
    const int32_t *qadd;       /* [1024]                            somatic_sniper.c:101-107 */
} ssn_tables;

/* Per-sample part of a site: sample_data_t + dqstats_t (output_format.h:13-23, dqstats.h:14-27).
 * rms_mapq is the RMS mapping quality packed in the consensus word (sniper_maqcns.c:176,268),
 * computed by the reference but never printed. */
typedef struct {
    int32_t genotype;                 /* independent consensus genotype, 4-bit allele mask */
    int32_t joint_genotype;           /* 0 unless -J */
    int32_t joint_consensus_quality;
    int32_t consensus_quality;
    int32_t variant_allele_quality;
    int32_t somatic_score;            /* -1 for the normal sample */
    int32_t variant_status;           /* variant_status_t, allele_util.h:16-22 */
    uint32_t mean_baseQ[4];
    uint32_t mean_mapQ[4];
    uint32_t base_occ[4];
    uint32_t dp4[4];
    uint32_t total_depth;
    uint32_t total_mean_mapQ;
    int32_t rms_mapq;
} ssn_sample;

/* One passing site = one sniper_output_t (output_format.h:25-35) minus the strings. */
typedef struct {
    uint32_t column;     /* column index inside the batch */
    int32_t ref_base4;   /* bam_nt16_table[ref char] */
    ssn_sample tumor;
    ssn_sample normal;
} ssn_site;

/* Optional per-column dump (SSN_SUBMIT_COLUMN_DUMP): both samples' quantised GLFs and the
 * value glf_somatic returns (somatic_sniper.c:268-272). 30 uint32 per column:
 *   [0..9] tumor lk, [10] min_lk, [11] rms mapQ, [12] depth, [13] consensus word,
 *   [14..27] the same for the normal sample, [28] (int32) return value,
 *   [29] 1 if the hom-reference pre-filter would have dropped the column before the full model (audit; the
 *        dump itself always comes from the full model). */
#define SSN_COLUMN_DUMP_WORDS 30

/* Pinned host arrays of one batch, filled by the packer.  Column c owns entries
 * off[2c]..off[2c+1] (tumor) and off[2c+1]..off[2c+2] (normal) of bs/bq/mq.  Deleted bases and
 * reads rejected by the flag/mapQ filter are not entries (sniper_maqcns.c:147, sniper_pileup.c:208). */
typedef struct {
    uint32_t *off;       /* 2*max_columns+1 */
    uint8_t *bs;         /* nt16 base code | strand<<4 */
    uint8_t *bq;         /* base quality */
    uint8_t *mq;         /* mapping quality */
    uint8_t *ref_nt16;   /* bam_nt16_table[reference char] per column; 15 => no call (somatic_sniper.c:127,156) */
    uint32_t max_columns;
    uint64_t max_entries;
} ssn_batch_view;

typedef struct ssn_ctx ssn_ctx;
typedef struct ssn_batch ssn_batch;

#define SSN_SUBMIT_COLUMN_DUMP 1u  /* also produce the per-column dump (parity tests) */
#define SSN_SUBMIT_NO_PREFILTER 2u /* run every column through the full model (measurement; results are identical) */

int ssn_abi_version(void);
const char *ssn_last_error(void);
ssn_status ssn_params_default(ssn_params *p);

/* Builds the model tables on the host (x87 long double where the reference uses it), uploads them,
 * pins coef+lhet in L2 with a persisting access-policy window, creates `n_slots` batch slots of the
 * given capacity (pinned host + device buffers, one stream each). */
ssn_status ssn_create(const ssn_params *p, int device, uint32_t n_slots, uint32_t max_columns,
                      uint64_t max_entries, ssn_ctx **out);
void ssn_destroy(ssn_ctx *ctx);
ssn_status ssn_get_tables(const ssn_ctx *ctx, ssn_tables *out);
/* The same tables without a device (host-only; used by parity tests and by tools that only need the
 * model): *handle owns the memory behind *view until ssn_host_tables_destroy. */
ssn_status ssn_host_tables_create(const ssn_params *p, ssn_tables *view, void **handle);
void ssn_host_tables_destroy(void *handle);
/* bam_nt16_table[ch] (samtools-0.1.6 bam_import.c:23-40): reference character -> ref_nt16 */
int ssn_nt16_of_char(int ch);

/* Slot ring: acquire a free batch (blocks on the oldest in-flight one only if none is free and
 * its results were already drained), fill the view, set its size, submit. */
ssn_status ssn_batch_acquire(ssn_ctx *ctx, ssn_batch **out);
ssn_status ssn_batch_view_get(ssn_batch *b, ssn_batch_view *out);
ssn_status ssn_batch_set_size(ssn_batch *b, uint32_t n_columns, uint64_t n_entries);
/* Asynchronous: H2D copy of the batch, the scoring kernel, D2H copy of the site ring. */
ssn_status ssn_batch_submit(ssn_ctx *ctx, ssn_batch *b, uint32_t flags);
/* Blocks until the OLDEST submitted batch is done; returns its sites sorted by column.  The
 * pointers stay valid until that batch is acquired again.  *batch (optional) receives the batch. */
ssn_status ssn_results_next(ssn_ctx *ctx, const ssn_site **sites, uint32_t *n_sites, ssn_batch **batch);
/* Per-column dump of a finished batch that was submitted with SSN_SUBMIT_COLUMN_DUMP. */
ssn_status ssn_batch_column_dump(ssn_batch *b, const uint32_t **words);
ssn_status ssn_batch_release(ssn_ctx *ctx, ssn_batch *b);

/* Device-resident path (kernel-only measurement): upload once, score many times. */
ssn_status ssn_batch_upload(ssn_ctx *ctx, ssn_batch *b);                 /* H2D + sync */
ssn_status ssn_batch_score_resident(ssn_ctx *ctx, ssn_batch *b, uint32_t flags); /* kernel only, async */
ssn_status ssn_batch_fetch(ssn_ctx *ctx, ssn_batch *b, const ssn_site **sites, uint32_t *n_sites); /* D2H + sync */

/* Device-side timing across all of the ctx's streams (CUDA events), and launch accounting. */
ssn_status ssn_timer_start(ssn_ctx *ctx);
ssn_status ssn_timer_stop(ssn_ctx *ctx, float *elapsed_ms);
ssn_status ssn_sync(ssn_ctx *ctx);
uint64_t ssn_kernel_launches(const ssn_ctx *ctx);   /* kernels launched by this ctx so far */
/* Sum of the scoring kernel's own durations (events on its launch stream) since the last reset. */
ssn_status ssn_kernel_time(ssn_ctx *ctx, float *total_ms, uint64_t *n_launches, int reset);
/* The same split into the pre-filter kernel and the scoring kernel; n_batches = batches timed. */
ssn_status ssn_kernel_time_split(ssn_ctx *ctx, float *prefilter_ms, float *score_ms, uint64_t *n_batches, int reset);
/* 1 if this ctx runs the exact hom-reference pre-filter in front of the full model (default; the parameter set must
 * satisfy the proof's preconditions, and SSN_NO_PREFILTER must be unset in the environment). */
int ssn_prefilter_enabled(const ssn_ctx *ctx);
/* Columns of a finished batch that the pre-filter handed to the full model (0 when it did not run). */
ssn_status ssn_batch_survivors(const ssn_batch *b, uint32_t *n_columns);
/* Host view of the pre-filter's bound tables (parity tests re-evaluate the filter on the CPU against the oracle):
 * ab[(n*N + m)*4 + {0,1,2}] = {cmin_all, cmax_run, g2_run}, h[a*N + b], F[0..N], vt; see ssn_prefilter_tables.c. */
typedef struct {
    const float *ab;
    const float *h;
    const float *F;
    const float *vt;  /* [(n*8 + m-1)*256 + s], m in 1..8, s in 0..255 */
    int32_t n;        /* N = 256 */
    int32_t enabled;
} ssn_prefilter_view;
ssn_status ssn_host_prefilter_create(const ssn_params *p, ssn_prefilter_view *view, void **handle);
void ssn_host_prefilter_destroy(void *handle);
/* ---- device-side pileup: reads in, sites out ----------------------------------------------------------------
 * Replaces, on the GPU, the column construction of get_next_pos / resolve_cigar
 * (src/lib/sniper/sniper_pileup.c:57-104,186-205): instead of packed columns (3 bytes per read-base) the host ships
 * the filtered reads themselves (≈1.6 bytes per read-base), filed under windows of SSN_READ_WINDOW positions, and the
 * device builds the very same column batch (ssn_batch_view layout, same entry order) before scoring it.  Which reads
 * are kept (flag/mapQ filter, keep rule, sortedness abort: sniper_pileup.c:207-222) stays with the caller.
 *
 * A record = ssn_read_rec header, then cigar[n_cigar] (uint32, BAM encoding), seq[(l_qseq+1)/2] (4-bit codes, BAM
 * order), qual[l_qseq]; `bytes` = record size rounded up to a multiple of 4.  Within one window and sample the
 * records must be in non-decreasing `beg` order (a coordinate-sorted BAM's order); a read appears in every window it
 * overlaps.  CIGAR semantics: M adds a read-base per position, D/N skip positions, I/S skip query bases, everything
 * else (H, P, =, X) advances nothing (samtools-0.1.6, QUIRK B6 in SURVEY.md). */
#define SSN_READ_WINDOW_SHIFT 13
#define SSN_READ_WINDOW (1u << SSN_READ_WINDOW_SHIFT)
typedef struct {
    int32_t beg, end;          /* 0-based start; bam_calend() end (informational: the device walks the CIGAR) */
    uint32_t l_qseq;
    uint16_t n_cigar;
    uint8_t mapq, strand;      /* strand 1 = reverse */
    uint32_t bytes;
} ssn_read_rec;
typedef struct {
    uint32_t rec_first[2];     /* index into rec_off of the window's first read: tumor, normal */
    uint32_t n_reads[2];
    uint32_t lo;               /* first position of the window (a multiple of SSN_READ_WINDOW) */
    int32_t tid;
} ssn_read_window;
/* Pinned host arrays of one read batch (owned by the slot; fill them in place). */
typedef struct {
    uint8_t *recs;             /* records back to back */
    uint32_t *rec_off;         /* byte offset / 4 of every record, window after window, tumor reads before normal */
    ssn_read_window *win;
    uint8_t *ref_nt16;         /* [max_windows][SSN_READ_WINDOW]: bam_nt16_table[reference char]; 15 => no column */
    uint64_t max_rec_bytes;
    uint32_t max_reads, max_windows;
} ssn_reads_view;
/* Capacity of the read buffers every slot gets on its first ssn_batch_reads_view (call before that). */
ssn_status ssn_reads_reserve(ssn_ctx *ctx, uint64_t max_rec_bytes, uint32_t max_reads, uint32_t max_windows);
ssn_status ssn_batch_reads_view(ssn_ctx *ctx, ssn_batch *b, ssn_reads_view *out);
/* Asynchronous: H2D of the reads, the pileup kernels (count, column scan, fill).  The scoring kernels are queued when
 * the batch reaches the head of ssn_results_next (their launch geometry needs the column count the device found).
 * The columns must fit the slot (max_columns / max_entries of ssn_create), else results_next fails with
 * SSN_ERR_OVERFLOW. */
ssn_status ssn_batch_submit_reads(ssn_ctx *ctx, ssn_batch *b, uint32_t n_windows, uint32_t n_reads, uint64_t rec_bytes,
                                  uint32_t flags);
/* After ssn_results_next: genome position of every column the device built, pos[c] = window index << 13 | offset. */
ssn_status ssn_batch_column_positions(ssn_batch *b, const uint32_t **pos, uint32_t *n_columns, uint64_t *n_entries);
/* Copies the device-built column batch into the slot's pinned ssn_batch_view arrays (parity tests). */
ssn_status ssn_batch_download_columns(ssn_ctx *ctx, ssn_batch *b);
/* Sum of the pileup kernels' durations since the last reset. */
ssn_status ssn_pileup_time(ssn_ctx *ctx, float *total_ms, uint64_t *n_batches, int reset);

/* ---- BGZF blocks inflated on the device ------------------------------------------------------------------------
 * Replaces, for the BAM readers, the per-block inflateInit2 / inflate / inflateEnd of samtools-0.1.6 bgzf.c:330-353
 * (under bgzf_read :489 and bam_read1, bam.c:181-212): one warp per block (csrc/ssn_inflate_dev.cuh).  A reader thread
 * acquires a LANE (pinned input and output buffers), lays the deflate payloads of up to max_blocks blocks into the input
 * buffer back to back, describes them, and calls ssn_inflater_run: synchronous for the caller, no CPU while it waits;
 * thread-safe, one lane per calling thread at a time.  Behind the call the requests of all threads waiting at that moment
 * go to the device together (one launch over all their blocks, csrc/ssn_inflate_api.cu).  refused[i] = 1 marks a block
 * the device decoder gave up on (nothing usable at its out_off): the caller inflates it on the host, so the service
 * never changes what is read. */
typedef struct { uint32_t in_off, in_len, out_off, out_len; } ssn_inflate_block;
typedef struct ssn_inflater ssn_inflater;
ssn_status ssn_inflater_create(int device, uint32_t n_lanes, uint64_t in_cap, uint64_t out_cap, uint32_t max_blocks, ssn_inflater **out);
void ssn_inflater_destroy(ssn_inflater *f);
ssn_status ssn_inflater_limits(const ssn_inflater *f, uint64_t *in_cap, uint64_t *out_cap, uint32_t *max_blocks);
/* blocks until a lane is free; *in / *out are its pinned buffers (in_cap / out_cap bytes, valid until the release) */
ssn_status ssn_inflater_lane_acquire(ssn_inflater *f, int32_t *lane, uint8_t **in, uint8_t **out);
void ssn_inflater_lane_release(ssn_inflater *f, int32_t lane);
ssn_status ssn_inflater_run(ssn_inflater *f, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes,
                            uint64_t out_bytes, uint8_t *refused);
/* The same, the inflated bytes delivered to h_dst instead of the lane's output buffer (NULL: the lane's).  h_dst must be
 * page-locked memory the device can write: ssn_inflater_host_alloc hands such buffers out (usable with the lanes of any
 * inflater of the process; given back with ssn_inflater_host_free, they are kept for the next caller and released with
 * the inflater).  A reader that owns its output buffer this way saves one copy of everything it reads. */
ssn_status ssn_inflater_run_to(ssn_inflater *f, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes,
                               uint64_t out_bytes, uint8_t *refused, uint8_t *h_dst);
ssn_status ssn_inflater_host_alloc(ssn_inflater *f, uint64_t bytes, void **out);
void ssn_inflater_host_free(ssn_inflater *f, void *p);
ssn_status ssn_inflater_stats(ssn_inflater *f, uint64_t *blocks, uint64_t *refused, uint64_t *runs, uint64_t *in_bytes, uint64_t *out_bytes);

/* The same decoder for callers that cannot allocate (no counterpart in the reference; added:
 * code under CUDA graph capture, tight serving loops, harnesses that forbid cudaMalloc).  The call allocates
 * NOTHING, on the host or on the device: the block table and the warps' work counter go into a file-scope __device__
 * array of the library (64.25 KB per GPU that loads it, held until exit), which limits one call to
 * SSN_INFLATE_STATIC_MAX_BLOCKS blocks — more is refused with SSN_ERR_ARG and a message naming the limit.
 * d_in (in_bytes of payloads, readable 64 bytes further), d_out (out_bytes) and d_status (n words; 0 = block inflated,
 * else refused: nothing usable at its out_off) are DEVICE memory of `device` owned by the caller; blk[0, n) is host memory and
 * may be reused when the call returns.  The work is queued on `stream` (a cudaStream_t, NULL = the default stream) and the
 * call does not wait for it.  Every call shares the one array: two calls whose work may overlap on the GPU (different
 * streams, different threads) race — order them on one stream or with events. */
#define SSN_INFLATE_STATIC_MAX_BLOCKS 4096u
ssn_status ssn_inflate_static(int device, const uint8_t *d_in, uint64_t in_bytes, const ssn_inflate_block *blk, uint32_t n,
                              uint8_t *d_out, uint64_t out_bytes, uint32_t *d_status, void *stream);

/* Writes a buffer larger than L2 on the ctx's device (cold-cache timing). */
ssn_status ssn_flush_l2(ssn_ctx *ctx);

#ifdef __cplusplus
}
#endif
#endif /* SSN_H */
