/*
 * ssn_host.h — C ABI of the host packer (libssn_host.so; no CUDA dependency).
 *
 * Replaces the reference's dual-BAM pileup walk
 *     bam_sspileup_file(fp1, fp2, mask, mapQ, glf_somatic, d, snp_fh)      src/lib/sniper/sniper_pileup.c:226-266
 *     get_next_pos / resolve_cigar                                          src/lib/sniper/sniper_pileup.c:57-104,172-224
 * and the per-column reference-base lookup of glf_somatic (src/lib/sniper/somatic_sniper.c:112-117):
 * instead of one callback per column it emits structure-of-arrays batches in the layout of
 * ssn_batch_view (include/ssn.h), straight into the pinned arrays of a scorer batch.
 *
 * A column is emitted when both samples hold at least one non-deleted read base there and the
 * reference base is callable (bam_nt16_table[ref] != 15): exactly the columns for which glf_somatic can
 * produce output (somatic_sniper.c:127,156).  Read filter, sortedness check and the keep rule that drops
 * the first read of every later target follow the reference (see csrc/host/ssn_bamread.h).
 */
#ifndef SSN_HOST_H
#define SSN_HOST_H

#include "ssn.h"

#ifdef __cplusplus
extern "C" {
#endif

typedef struct ssn_packer ssn_packer;

/* column -> genome map of one batch; arrays of view.max_columns elements owned by the caller */
typedef struct {
    int32_t *tid;        /* target index (tumor header numbering, as the reference compares tids numerically) */
    uint32_t *pos;       /* 0-based position */
    uint8_t *ref_char;   /* raw FASTA character (case preserved), what the writers print */
} ssn_colmap;

typedef struct {
    uint64_t records[2];      /* BAM records framed so far: tumor, normal */
    uint64_t columns, entries, batches;
    double t_wait, t_count, t_fill;   /* seconds inside ssn_packer_next: waiting for readers, counting, filling */
} ssn_packer_stats;

/* tumor may be "-" (stdin), like the reference (main.c:128).  fasta may be NULL (every base 'N': nothing is
 * emitted, main.c requires -f).  min_mapq = -q.  n_threads = worker threads shared by inflate and packing
 * (0: everything on the calling thread plus the two reader threads). */
ssn_status ssn_packer_open(const char *tumor_bam, const char *normal_bam, const char *fasta, int min_mapq,
                           int n_threads, ssn_packer **out);
void ssn_packer_close(ssn_packer *p);
/* header of the tumor (which = 0) or normal (1) BAM */
int32_t ssn_packer_n_targets(const ssn_packer *p, int which);
const char *ssn_packer_target_name(const ssn_packer *p, int which, int32_t tid);
/* Fills the next batch (at most view->max_columns columns / view->max_entries read bases).  *n_columns = 0
 * at the end of the walk. */
ssn_status ssn_packer_next(ssn_packer *p, const ssn_batch_view *view, const ssn_colmap *map,
                           uint32_t *n_columns, uint64_t *n_entries);
/* Device pileup (ssn_batch_submit_reads, include/ssn.h): moves the next joined windows, whole, into the read arrays of a
 * scorer slot — the records as the BAM reader filed them, their index, the reference slice of every window.  Takes
 * windows while the read arrays and the slot (max_columns / max_entries, against the windows' upper bounds) can hold
 * them.  ref_chars (optional, [max_windows][SSN_READ_WINDOW]) receives the raw FASTA characters the writers print.
 * *n_windows = 0 at the end of the walk.  Do not mix with ssn_packer_next on one packer. */
ssn_status ssn_packer_next_reads(ssn_packer *p, const ssn_reads_view *rv, uint32_t max_columns, uint64_t max_entries,
                                 uint8_t *ref_chars, uint32_t *n_windows, uint32_t *n_reads, uint64_t *rec_bytes);
/* Packs whole read windows — the input of the device pileup, ssn_batch_submit_reads (include/ssn.h) — with the packer's
 * own count/fill code, single-threaded: the CPU side of the device pileup's parity tests.  The records of one window
 * and sample must lie back to back in `recs`.  pos[c] (optional) = window index << 13 | offset inside the window. */
ssn_status ssn_pack_read_windows(const ssn_reads_view *in, uint32_t n_windows, const ssn_batch_view *view,
                                 uint32_t *pos, uint32_t *n_columns, uint64_t *n_entries);
void ssn_packer_get_stats(const ssn_packer *p, ssn_packer_stats *out);
const char *ssn_host_last_error(void);
/* Device-pileup fall-back.  ssn_packer_next_reads returns SSN_ERR_OVERFLOW (with *n_windows = 0) when the next window
 * alone is too deep for a read batch; nothing is consumed.  The caller then packs that window into columns:
 * ssn_packer_one_window(p, 1), ssn_packer_next until ssn_packer_partial(p) is 0, ssn_packer_one_window(p, 0), and goes
 * back to ssn_packer_next_reads. */
void ssn_packer_one_window(ssn_packer *p, int on);
int ssn_packer_partial(const ssn_packer *p);

/* ---- inflate service -----------------------------------------------------------------------------------------------
 * libssn_host.so has no CUDA dependency; a caller that owns GPUs hands the readers an inflate service (the CLI wraps
 * ssn_inflater_* of include/ssn.h).  Per batch of BGZF blocks a reader takes a lane (lane_acquire may block until one is
 * free: the number of lanes is the number of batches in flight), lays the deflate payloads of up to max_blocks blocks /
 * in_cap bytes into the lane's input buffer, calls run (blocking; refused[i] != 0: inflate block i on the host), copies
 * the inflated bytes out of the lane's output buffer and gives the lane back — a reader held up by its queue holds no
 * lane.  All lanes of a service have the same capacities.  Without a service the readers inflate on the host. */
typedef struct {
    void *ctx;
    int32_t (*lane_acquire)(void *ctx, uint8_t **in, uint64_t *in_cap, uint8_t **out, uint64_t *out_cap, uint32_t *max_blocks);  /* < 0: none */
    void (*lane_release)(void *ctx, int32_t lane);
    int (*run)(void *ctx, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes, uint64_t out_bytes, uint8_t *refused);  /* 0 ok */
    /* optional, all three or none (NULL): memory the service can deliver into directly.  A reader then owns its output
     * buffer (host_alloc, for its lifetime) and run_to writes the inflated bytes there, which saves it the copy out of
     * the lane's buffer. */
    void *(*host_alloc)(void *ctx, uint64_t bytes);
    void (*host_free)(void *ctx, void *p);
    int (*run_to)(void *ctx, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes, uint64_t out_bytes, uint8_t *refused, uint8_t *dst);
} ssn_inflate_service;

/* ---- region shards: the genome cut into contiguous intervals, one packer each ---------------------
 * Replaces the single front-to-back walk of bam_sspileup_file (src/lib/sniper/sniper_pileup.c:226-266) by many walks
 * over disjoint intervals whose union yields exactly the columns of the one walk: interval edges are multiples of the
 * packer's window size, every interval's readers seek to their first record by bisection (no index file), start a
 * look-back early to catch the reads reaching in, and take over the state of the keep rule (:216, QUIRK B1) from the last
 * passing read in front of them.  Several intervals are read ahead concurrently; callers (one per GPU) take them in genome
 * order with ssn_shardset_next, drive the returned packer with ssn_packer_next / ssn_packer_next_reads until it is
 * exhausted, and hand it back with ssn_shardset_done.  Thread-safe.  Inputs that cannot be cut (stdin, a one-interval
 * genome) come back as a single shard over the front-to-back reader. */
typedef struct ssn_shardset ssn_shardset;
typedef struct {
    int n_threads;              /* host threads to plan for (0: online cores - 2) */
    uint32_t stride_windows;    /* interval length in windows of 8192 positions (0: from the genome length; SSN_SHARD_WINDOWS) */
    uint32_t lookback_windows;  /* how far in front of its interval a reader starts (0: 4; SSN_SHARD_LOOKBACK) */
    int max_reading;            /* intervals whose readers run at once (0: from n_threads; SSN_SHARD_READERS) */
    int max_open;               /* intervals opened but not yet handed back (0: 2 * max_reading + 2) */
    int serial;                 /* 1: one interval, front to back */
} ssn_shard_opts;
ssn_status ssn_shardset_open(const char *tumor_bam, const char *normal_bam, const char *fasta, int min_mapq,
                             const ssn_shard_opts *opts, ssn_shardset **out);
void ssn_shardset_close(ssn_shardset *s);
int32_t ssn_shardset_n_shards(const ssn_shardset *s);
int32_t ssn_shardset_n_targets(const ssn_shardset *s, int which);
const char *ssn_shardset_target_name(const ssn_shardset *s, int which, int32_t tid);
/* interval [ (lo_tid, lo_pos), (hi_tid, hi_pos) ) of a shard; hi_tid = INT32_MAX: to the end */
ssn_status ssn_shardset_range(const ssn_shardset *s, int32_t index, int32_t *lo_tid, uint32_t *lo_pos, int32_t *hi_tid, uint32_t *hi_pos);
/* next shard in genome order; *index = -1 after the last one */
ssn_status ssn_shardset_next(ssn_shardset *s, int32_t *index, ssn_packer **packer);
ssn_status ssn_shardset_done(ssn_shardset *s, int32_t index);
/* 1: a read longer than the look-back crossed an interval edge, so a later interval missed it: the results are
 * incomplete and the caller must run the pair again with opts.serial = 1 (checked after the last ssn_shardset_done) */
int ssn_shardset_long_span(const ssn_shardset *s);
/* readers of intervals opened from now on inflate through the service (NULL: on the host); *svc must outlive the set */
void ssn_shardset_set_inflate_service(ssn_shardset *s, const ssn_inflate_service *svc);
void ssn_shardset_get_stats(const ssn_shardset *s, ssn_packer_stats *out);

/* ---- writers: one passing site -> one line, byte-identical to the reference ---------------------------
 * output_classic.c:9-55, output_vcf.c:135-192, output_bed.c:5-19; format names as in output_format.c:9-17 */
typedef struct ssn_writer ssn_writer;
/* fh is a FILE*; returns NULL for an unknown format name (caller prints the reference's message) */
ssn_writer *ssn_writer_create(const char *format, void *fh);
void ssn_writer_header(ssn_writer *w, const char *refseq, const char *normal_id, const char *tumor_id);
void ssn_writer_site(ssn_writer *w, const char *seq_name, uint32_t pos, int ref_char, const ssn_site *site);
void ssn_writer_destroy(ssn_writer *w);
int ssn_writer_n_formats(void);
const char *ssn_writer_format_name(int i);

#ifdef __cplusplus
}
#endif
#endif /* SSN_HOST_H */
