/*
 * bam-somaticsniper (B200 build) — same command line, messages, exit codes and output bytes as the reference
 * driver (src/exe/bam-somaticsniper/main.c:64-162); the per-column work runs on the GPU(s) through libssn.so.
 *
 *   reference                                           here
 *   getopt "n:t:vf:T:N:r:I:q:Q:pLGJs:F:" (:80-99)        identical option handling
 *   fai_load / table + prior construction (:106-127)    ssn_packer_open (FASTA index) / ssn_create (tables, upload)
 *   bam_sspileup_file(..., glf_somatic, ...) (:146)     ssn_packer_next -> ssn_batch_submit -> ssn_results_next
 *   output_formatter (:141-145, somatic_sniper.c:264)   ssn_writer_*
 *
 *   one front-to-back walk on one thread                region shards (ssn_shardset_*): the genome is cut into intervals,
 *                                                       several are read ahead at once, one feeder thread per GPU takes them
 *                                                       in genome order, the main thread writes their output in that order
 *
 * Not in the reference (environment only, the command line is unchanged):
 *   SSN_GPUS     number of GPUs (default 1); every GPU gets a feeder thread and takes the next region shard when it is free
 *   SSN_DEVICES  comma-separated device ordinals, one scoring context each (overrides SSN_GPUS; an ordinal may repeat)
 *   SSN_THREADS  host threads to plan for: readers, inflate, packing (default: online cores - 2, at least 1)
 *   SSN_SLOTS    batches in flight per GPU (default 3)
 *   SSN_DEVICE_PILEUP  1 (default): ship the filtered reads, build the columns on the GPU; 0: pack columns on the host
 *   SSN_GPU_INFLATE  1 (default below 24 planned threads): the readers hand BGZF blocks to the GPU(s) to inflate (ssn_inflater_*, one warp per
 *                block) and only frame the records; 0: inflate on the host cores.  SSN_INFLATE_LANES batches in flight
 *                per GPU (24), SSN_INFLATE_BLOCKS blocks per batch (96)
 *   SSN_SERIAL   1: one interval, front to back (what a pair that cannot be cut gets anyway)
 *   SSN_SHARD_WINDOWS / SSN_SHARD_LOOKBACK / SSN_SHARD_READERS  interval length and look-back in windows of 8192
 *                positions, intervals read at once (ssn_host.h)
 *   SSN_STATS    file that receives one JSON line of run statistics
 * There is no CPU scoring path: without a usable sm_100 device the program fails.
 */
#include <ctype.h>
#include <execinfo.h>
#include <malloc.h>
#include <signal.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include "ssn.h"
#include "ssn_host.h"

#define SSN_CLI_VERSION "b200-r2"
#define MAX_GPUS 16
#define MAX_SLOTS 8
#define BATCH_COLUMNS (16u << 13)            /* 16 packer windows */
#define BATCH_ENTRIES ((uint64_t)32 << 20)      /* one window of 8192 positions up to ~4000x (tumor + normal) */

static const char *default_normal_id = "NORMAL";
static const char *default_tumor_id = "TUMOR";
static const char *default_format = "classic";

static void version_info(void)
{
    printf("Somatic Sniper version (%s) (commit %s)", SSN_CLI_VERSION, "nogit");
    printf(" (%s)", "release");
    printf("\n");
}

static void usage(const char *progname, const ssn_params *p, int mapq)
{
    const char *pn = strrchr(progname, '/');
    int i;
    pn = pn ? pn + 1 : progname;
    fprintf(stderr, "\n");
    fprintf(stderr, "\n");
    fprintf(stderr, "%s [options] -f <ref.fasta> <tumor.bam> <normal.bam> <snp_output_file>\n\n", pn);
    fprintf(stderr, "Required Option: \n");
    fprintf(stderr, "        -f FILE   REQUIRED reference sequence in the FASTA format\n\n");
    fprintf(stderr, "Options: \n");
    fprintf(stderr, "        -v        Display version information\n\n");
    fprintf(stderr, "        -q INT    filtering reads with mapping quality less than INT [%d]\n", mapq);
    fprintf(stderr, "        -Q INT    filtering somatic snv output with somatic quality less than  INT [%d]\n", p->min_somatic_qual);
    fprintf(stderr, "        -L FLAG   do not report LOH variants as determined by genotypes\n");
    fprintf(stderr, "        -G FLAG   do not report Gain of Reference variants as determined by genotypes\n");
    fprintf(stderr, "        -p FLAG   disable priors in the somatic calculation. Increases sensitivity for solid tumors\n");
    fprintf(stderr, "        -J FLAG   Use prior probabilities accounting for the somatic mutation rate\n");
    fprintf(stderr, "        -s FLOAT  prior probability of a somatic mutation (implies -J) [%f]\n", p->somatic_mutation_rate);
    fprintf(stderr, "        -T FLOAT  theta in maq consensus calling model (for -c/-g) [%f]\n", p->theta);
    fprintf(stderr, "        -N INT    number of haplotypes in the sample (for -c/-g) [%d]\n", p->n_hap);
    fprintf(stderr, "        -r FLOAT  prior of a difference between two haplotypes (for -c/-g) [%f]\n", p->het_rate);
    fprintf(stderr, "        -n STRING normal sample id (for VCF header) [%s]\n", default_normal_id);
    fprintf(stderr, "        -t STRING tumor sample id (for VCF header) [%s]\n", default_tumor_id);
    fprintf(stderr, "        -F STRING select output format [%s]\n", default_format);
    fprintf(stderr, "           Available formats:\n");
    for (i = 0; i < ssn_writer_n_formats(); ++i) fprintf(stderr, "             %s\n", ssn_writer_format_name(i));
    fprintf(stderr, "\n");
}

typedef struct {
    ssn_batch *batch;
    ssn_colmap map;
    /* device pileup: the slot's read arrays and the raw reference characters of its windows */
    ssn_reads_view rv;
    uint8_t *ref_chars;
    int from_reads;
    int32_t shard;
} slot_meta;

/* output of one region shard: lines are formatted into memory by the feeder that scored it and written by main in
 * shard order */
typedef struct {
    char *buf; size_t len;
    FILE *fp;
    ssn_writer *wr;
    int in_flight, exhausted, complete;
} shard_out;

typedef struct run run_t;

typedef struct {
    run_t *run;
    int index, device;
    ssn_ctx *ctx;
    slot_meta meta[MAX_SLOTS];
    int n_meta, in_flight;
    slot_meta *fifo[MAX_SLOTS];
    int f_head, f_count;
    pthread_t th;
    int th_started, setup_rc;
    /* statistics */
    uint64_t n_sites, n_batches, survivors, dev_columns, dev_entries, col_batches;
    double t_gpu_wait, t_setup;
} gpu_t;

struct run {
    ssn_params prm;
    int n_gpus, n_slots, n_threads, device_pileup;
    const char *format;
    gpu_t gpus[MAX_GPUS];
    ssn_shardset *set;
    shard_out *outs;
    int32_t n_shards;
    pthread_mutex_t mu;
    pthread_cond_t cv;
    int error, gpus_ready, go;
};

/* a crash leaves a trace: the faulting address, the call stack (module + offset: addr2line -e <module> <offset>) */
static void crash_handler(int sig, siginfo_t *si, void *uc)
{
    void *frames[48];
    int n;
    (void)uc;
    fprintf(stderr, "[bam-somaticsniper] fatal signal %d at address %p; call stack:\n", sig, si ? si->si_addr : 0);
    n = backtrace(frames, 48);
    backtrace_symbols_fd(frames, n, 2);
    signal(sig, SIG_DFL);
    raise(sig);
}

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static int env_int(const char *name, int dflt, int lo, int hi)
{
    const char *s = getenv(name);
    int v = s && *s ? atoi(s) : dflt;
    if (v < lo) v = lo;
    if (v > hi) v = hi;
    return v;
}

static void run_fail(run_t *R)
{
    pthread_mutex_lock(&R->mu);
    R->error = 1;
    pthread_cond_broadcast(&R->cv);
    pthread_mutex_unlock(&R->mu);
}

static slot_meta *meta_of(gpu_t *g, ssn_batch *b)
{
    int i;
    for (i = 0; i < g->n_meta; ++i)
        if (g->meta[i].batch == b) return &g->meta[i];
    if (g->n_meta == MAX_SLOTS) return 0;
    memset(&g->meta[g->n_meta], 0, sizeof(slot_meta));
    g->meta[g->n_meta].batch = b;
    g->meta[g->n_meta].ref_chars = (uint8_t*)malloc(BATCH_COLUMNS);
    return &g->meta[g->n_meta++];
}

static int meta_colmap(slot_meta *m)
{
    if (m->map.tid) return 0;
    m->map.tid = (int32_t*)malloc(sizeof(int32_t) * BATCH_COLUMNS);
    m->map.pos = (uint32_t*)malloc(sizeof(uint32_t) * BATCH_COLUMNS);
    m->map.ref_char = (uint8_t*)malloc(BATCH_COLUMNS);
    return m->map.tid && m->map.pos && m->map.ref_char ? 0 : -1;
}

static void shard_complete(run_t *R, int32_t k)
{
    shard_out *o = &R->outs[k];
    ssn_writer_destroy(o->wr); o->wr = 0;
    fclose(o->fp); o->fp = 0;               /* finalises buf / len */
    pthread_mutex_lock(&R->mu);
    o->complete = 1;
    pthread_cond_broadcast(&R->cv);
    pthread_mutex_unlock(&R->mu);
}

/* waits for the GPU's oldest batch in flight and formats its sites into its shard's buffer */
static int drain_one(gpu_t *g)
{
    run_t *R = g->run;
    const ssn_site *sites;
    uint32_t n, i;
    ssn_batch *b;
    slot_meta *m = g->fifo[g->f_head];
    shard_out *o = &R->outs[m->shard];
    const double t0 = now_s();
    if (ssn_results_next(g->ctx, &sites, &n, &b) != SSN_OK) {
        fprintf(stderr, "[bam-somaticsniper] GPU scoring failed: %s\n", ssn_last_error());
        return -1;
    }
    g->t_gpu_wait += now_s() - t0;
    if (b != m->batch) { fprintf(stderr, "[bam-somaticsniper] batches came back out of order\n"); return -1; }
    if (m->from_reads) {
        const uint32_t *cpos;
        uint32_t ncol;
        uint64_t nent;
        if (ssn_batch_column_positions(b, &cpos, &ncol, &nent) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error());
            return -1;
        }
        g->dev_columns += ncol; g->dev_entries += nent;
        for (i = 0; i < n; ++i) {
            const uint32_t pw = cpos[sites[i].column], w = pw >> SSN_READ_WINDOW_SHIFT, off = pw & (SSN_READ_WINDOW - 1u);
            const char *name = ssn_shardset_target_name(R->set, 0, m->rv.win[w].tid);   /* names from the tumor header (QUIRK B9) */
            ssn_writer_site(o->wr, name ? name : "", m->rv.win[w].lo + off, m->ref_chars[pw], &sites[i]);
        }
    } else
    for (i = 0; i < n; ++i) {
        const uint32_t c = sites[i].column;
        const char *name = ssn_shardset_target_name(R->set, 0, m->map.tid[c]);
        ssn_writer_site(o->wr, name ? name : "", m->map.pos[c], m->map.ref_char[c], &sites[i]);
    }
    g->n_sites += n;
    {
        uint32_t surv = 0;
        if (ssn_batch_survivors(b, &surv) == SSN_OK) g->survivors += surv;
    }
    g->in_flight -= 1;
    g->f_head = (g->f_head + 1) % MAX_SLOTS;
    g->f_count -= 1;
    if (--o->in_flight == 0 && o->exhausted) shard_complete(R, m->shard);
    return 0;
}

static void push_flight(gpu_t *g, slot_meta *m, int32_t shard, int from_reads)
{
    m->shard = shard; m->from_reads = from_reads;
    g->fifo[(g->f_head + g->f_count) % MAX_SLOTS] = m;
    g->f_count += 1; g->in_flight += 1; g->n_batches += 1;
    g->run->outs[shard].in_flight += 1;
}

/* one region shard through this GPU: batches are submitted while earlier ones (of this or the previous shard) score */
static int feed_shard(gpu_t *g, int32_t k, ssn_packer *pk)
{
    run_t *R = g->run;
    int columns_mode = !R->device_pileup;       /* or: inside a window too deep for a read batch */
    for (;;) {
        ssn_batch *b;
        slot_meta *m;
        while (g->in_flight == R->n_slots)
            if (drain_one(g) != 0) return -1;
        if (ssn_batch_acquire(g->ctx, &b) != SSN_OK) { fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error()); return -1; }
        m = meta_of(g, b);
        if (!m) { fprintf(stderr, "[bam-somaticsniper] slot table overflow\n"); return -1; }
        if (!columns_mode) {
            uint32_t nw = 0, nr = 0;
            uint64_t nb = 0;
            ssn_status st;
            if (ssn_batch_reads_view(g->ctx, b, &m->rv) != SSN_OK) { fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error()); return -1; }
            st = ssn_packer_next_reads(pk, &m->rv, BATCH_COLUMNS, BATCH_ENTRIES, m->ref_chars, &nw, &nr, &nb);
            if (st == SSN_ERR_OVERFLOW) {
                /* this window alone is too deep for a read batch: the host packs it into columns (any depth), then the
                 * reads route resumes */
                columns_mode = 1;
                ssn_packer_one_window(pk, 1);
            } else if (st != SSN_OK) {
                fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_host_last_error());
                return -1;
            } else {
                if (nw == 0) { ssn_batch_release(g->ctx, b); return 0; }
                if (ssn_batch_submit_reads(g->ctx, b, nw, nr, nb, 0) != SSN_OK) {
                    fprintf(stderr, "[bam-somaticsniper] GPU submit failed: %s\n", ssn_last_error());
                    return -1;
                }
                push_flight(g, m, k, 1);
                continue;
            }
        }
        {
            ssn_batch_view view;
            uint32_t ncol = 0;
            uint64_t nent = 0;
            if (ssn_batch_view_get(b, &view) != SSN_OK) { fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error()); return -1; }
            if (meta_colmap(m) != 0) { fprintf(stderr, "[bam-somaticsniper] out of memory\n"); return -1; }
            if (ssn_packer_next(pk, &view, &m->map, &ncol, &nent) != SSN_OK) {
                fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_host_last_error());
                return -1;
            }
            if (ncol == 0) { ssn_batch_release(g->ctx, b); return 0; }
            if (ssn_batch_set_size(b, ncol, nent) != SSN_OK || ssn_batch_submit(g->ctx, b, 0) != SSN_OK) {
                fprintf(stderr, "[bam-somaticsniper] GPU submit failed: %s\n", ssn_last_error());
                return -1;
            }
            push_flight(g, m, k, 0);
            g->col_batches += 1;
            if (R->device_pileup && !ssn_packer_partial(pk)) { columns_mode = 0; ssn_packer_one_window(pk, 0); }
        }
    }
}

static int gpu_setup(gpu_t *g)
{
    run_t *R = g->run;
    const double t0 = now_s();
    if (ssn_create(&R->prm, g->device, (uint32_t)R->n_slots, BATCH_COLUMNS, BATCH_ENTRIES, &g->ctx) != SSN_OK) {
        fprintf(stderr, "[bam-somaticsniper] cannot set up GPU %d: %s\n", g->device, ssn_last_error());
        fprintf(stderr, "[bam-somaticsniper] this build scores on sm_100 GPUs only; there is no CPU path\n");
        return -1;
    }
    if (R->device_pileup) {
        ssn_batch *held[MAX_SLOTS];
        ssn_reads_view rv;
        int k, ok = 1;
        if (ssn_reads_reserve(g->ctx, 2 * BATCH_ENTRIES, (uint32_t)(BATCH_ENTRIES / 24), BATCH_COLUMNS >> SSN_READ_WINDOW_SHIFT) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error());
            return -1;
        }
        /* the slots' read arrays are allocated on first use: do it here, not inside the stream */
        for (k = 0; k < R->n_slots; ++k) ok = ok && ssn_batch_acquire(g->ctx, &held[k]) == SSN_OK && ssn_batch_reads_view(g->ctx, held[k], &rv) == SSN_OK;
        if (!ok) { fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error()); return -1; }
        for (k = 0; k < R->n_slots; ++k) ssn_batch_release(g->ctx, held[k]);
    }
    g->t_setup = now_s() - t0;
    return 0;
}

/* feeder thread of one GPU: sets the context up, then takes region shards in genome order until none is left */
static void *gpu_main(void *arg)
{
    gpu_t *g = (gpu_t*)arg;
    run_t *R = g->run;
    if (!g->ctx) g->setup_rc = gpu_setup(g);
    pthread_mutex_lock(&R->mu);
    R->gpus_ready += 1;
    if (g->setup_rc != 0) R->error = 1;
    pthread_cond_broadcast(&R->cv);
    while (!R->go && !R->error) pthread_cond_wait(&R->cv, &R->mu);
    pthread_mutex_unlock(&R->mu);
    if (g->setup_rc != 0) return 0;
    for (;;) {
        int32_t k = -1;
        ssn_packer *pk = 0;
        shard_out *o;
        int rc;
        if (R->error) break;
        if (ssn_shardset_next(R->set, &k, &pk) != SSN_OK) { fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_host_last_error()); run_fail(R); break; }
        if (k < 0) break;
        o = &R->outs[k];
        o->fp = open_memstream(&o->buf, &o->len);
        o->wr = o->fp ? ssn_writer_create(R->format, o->fp) : 0;
        if (!o->wr) { fprintf(stderr, "[bam-somaticsniper] out of memory\n"); ssn_shardset_done(R->set, k); run_fail(R); break; }
        rc = feed_shard(g, k, pk);
        if (ssn_shardset_done(R->set, k) != SSN_OK && rc == 0) {
            /* the seek could not place this interval: main runs the pair again as one interval */
            pthread_mutex_lock(&R->mu); R->error |= 2; pthread_cond_broadcast(&R->cv); pthread_mutex_unlock(&R->mu);
            break;
        }
        if (rc != 0) { run_fail(R); break; }
        o->exhausted = 1;
        if (o->in_flight == 0) shard_complete(R, k);
    }
    while (g->f_count > 0 && !(R->error & 1))
        if (drain_one(g) != 0) { run_fail(R); break; }
    return 0;
}

/* ---- the readers' inflate service: lanes of one ssn_inflater per GPU, handed out round robin ---------------------- */
#define INFL_BLOCKS 96u
typedef struct {
    ssn_inflater *inf[MAX_GPUS];
    int n;
    unsigned next;
} infl_ctx;

static int32_t svc_acquire(void *ctx, uint8_t **in, uint64_t *in_cap, uint8_t **out, uint64_t *out_cap, uint32_t *max_blocks)
{
    infl_ctx *c = (infl_ctx*)ctx;
    const int g = (int)(__atomic_fetch_add(&c->next, 1u, __ATOMIC_RELAXED) % (unsigned)c->n);
    int32_t lane = -1;
    if (ssn_inflater_lane_acquire(c->inf[g], &lane, in, out) != SSN_OK) return -1;
    ssn_inflater_limits(c->inf[g], in_cap, out_cap, max_blocks);
    return g * 4096 + lane;
}
static void svc_release(void *ctx, int32_t lane)
{
    infl_ctx *c = (infl_ctx*)ctx;
    ssn_inflater_lane_release(c->inf[lane / 4096], lane % 4096);
}
static int svc_run(void *ctx, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes, uint64_t out_bytes, uint8_t *refused)
{
    infl_ctx *c = (infl_ctx*)ctx;
    return ssn_inflater_run(c->inf[lane / 4096], lane % 4096, blk, n, in_bytes, out_bytes, refused) == SSN_OK ? 0 : -1;
}

static int svc_run_to(void *ctx, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes, uint64_t out_bytes, uint8_t *refused, uint8_t *dst)
{
    infl_ctx *c = (infl_ctx*)ctx;
    return ssn_inflater_run_to(c->inf[lane / 4096], lane % 4096, blk, n, in_bytes, out_bytes, refused, dst) == SSN_OK ? 0 : -1;
}
/* the readers' own output buffers: page-locked, from the first inflater's pool (portable: any GPU's lane writes into them) */
static void *svc_host_alloc(void *ctx, uint64_t bytes)
{
    void *p = 0;
    return ssn_inflater_host_alloc(((infl_ctx*)ctx)->inf[0], bytes, &p) == SSN_OK ? p : 0;
}
static void svc_host_free(void *ctx, void *p) { ssn_inflater_host_free(((infl_ctx*)ctx)->inf[0], p); }

static int start_feeders(run_t *R)
{
    int i;
    R->error = 0; R->gpus_ready = 0; R->go = 0;
    for (i = 0; i < R->n_gpus; ++i) {
        gpu_t *g = &R->gpus[i];
        g->f_head = g->f_count = g->in_flight = 0; g->th_started = 0;
        if (pthread_create(&g->th, 0, gpu_main, g) != 0) {
            fprintf(stderr, "[bam-somaticsniper] cannot start a feeder thread\n");
            pthread_mutex_lock(&R->mu); R->error = 1; R->gpus_ready += 1; pthread_cond_broadcast(&R->cv); pthread_mutex_unlock(&R->mu);
            continue;
        }
        g->th_started = 1;
    }
    /* contexts are set up in parallel (model tables, device and pinned memory) */
    pthread_mutex_lock(&R->mu);
    while (R->gpus_ready < R->n_gpus) pthread_cond_wait(&R->cv, &R->mu);
    i = R->error;
    pthread_mutex_unlock(&R->mu);
    return i ? -1 : 0;
}

static void join_feeders(run_t *R)
{
    int i;
    pthread_mutex_lock(&R->mu);
    R->go = 1;
    pthread_cond_broadcast(&R->cv);
    pthread_mutex_unlock(&R->mu);
    for (i = 0; i < R->n_gpus; ++i)
        if (R->gpus[i].th_started) { pthread_join(R->gpus[i].th, 0); R->gpus[i].th_started = 0; }
}

/* Releases the feeders and writes the shards' output in genome order as they complete.
 * 0: done; 1: failed; 2: the pair has to be walked as one interval (a read longer than the look-back crossed an interval
 * edge, or the seek could not place an interval): the caller runs it again with serial = 1. */
static int stream_pass(run_t *R, FILE *out)
{
    int32_t k;
    int err = 0;
    pthread_mutex_lock(&R->mu);
    R->go = 1;
    pthread_cond_broadcast(&R->cv);
    for (k = 0; k < R->n_shards; ++k) {
        shard_out *o = &R->outs[k];
        while (!o->complete && !R->error) pthread_cond_wait(&R->cv, &R->mu);
        if (R->error) break;
        pthread_mutex_unlock(&R->mu);
        if (o->len) fwrite(o->buf, 1, o->len, out);
        free(o->buf); o->buf = 0; o->len = 0;
        pthread_mutex_lock(&R->mu);
    }
    err = R->error;
    pthread_mutex_unlock(&R->mu);
    join_feeders(R);
    err |= R->error;
    if (err & 1) return 1;
    if ((err & 2) || ssn_shardset_long_span(R->set)) return 2;
    return 0;
}

static void free_outs(run_t *R)
{
    int32_t k;
    for (k = 0; k < R->n_shards && R->outs; ++k) {
        shard_out *o = &R->outs[k];
        if (o->wr) ssn_writer_destroy(o->wr);
        if (o->fp) fclose(o->fp);
        free(o->buf);
    }
    free(R->outs); R->outs = 0;
}

static int open_set(run_t *R, const char *tumor, const char *normal, const char *fasta, int mapq, int serial)
{
    ssn_shard_opts so;
    memset(&so, 0, sizeof(so));
    so.n_threads = R->n_threads; so.serial = serial;
    if (ssn_shardset_open(tumor, normal, fasta, mapq, &so, &R->set) != SSN_OK) {
        fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_host_last_error());
        return -1;
    }
    R->n_shards = ssn_shardset_n_shards(R->set);
    R->outs = (shard_out*)calloc((size_t)R->n_shards, sizeof(shard_out));
    return R->outs ? 0 : -1;
}

int main(int argc, char *argv[])
{
    int c, use_priors = 1, mapq = 0, i, rc = 0, passes = 1;
    const char *normal_id = default_normal_id, *tumor_id = default_tumor_id, *fn_fa = 0, *format = default_format;
    static run_t R;
    static infl_ctx infl;
    static ssn_inflate_service svc;
    int gpu_inflate = 0, gpu_inflate_wanted = 0;
    uint64_t infl_blocks = 0, infl_refused = 0;
    ssn_writer *wr;
    FILE *out;
    ssn_packer_stats st;
    int32_t n_shards_first;
    double t_start = now_s(), t_setup;

    /* window buffers of a megabyte or so are allocated and freed by the dozen per second on many threads: keep them on the
     * malloc arenas instead of one mmap/munmap system call each (those serialise on the process's address-space lock) */
    {
        struct sigaction sa;
        memset(&sa, 0, sizeof(sa));
        sa.sa_sigaction = crash_handler; sa.sa_flags = SA_SIGINFO;
        sigaction(SIGSEGV, &sa, 0); sigaction(SIGBUS, &sa, 0); sigaction(SIGABRT, &sa, 0);
    }
    mallopt(M_MMAP_THRESHOLD, 32 << 20);
    mallopt(M_TRIM_THRESHOLD, 1 << 30);
    memset(&R, 0, sizeof(R));
    memset(&st, 0, sizeof(st));
    ssn_params_default(&R.prm);
    while ((c = getopt(argc, argv, "n:t:vf:T:N:r:I:q:Q:pLGJs:F:")) >= 0) {
        switch (c) {
        case 'f': fn_fa = optarg; break;
        case 'T': R.prm.theta = (float)atof(optarg); break;
        case 'N': R.prm.n_hap = atoi(optarg); break;
        case 'r': R.prm.het_rate = (float)atof(optarg); break;
        case 'q': mapq = atoi(optarg); break;
        case 'Q': R.prm.min_somatic_qual = atoi(optarg); break;
        case 'F': format = optarg; break;
        case 'p': use_priors = 0; break;
        case 'J': R.prm.use_joint_priors = 1; break;
        case 's': R.prm.somatic_mutation_rate = atof(optarg); R.prm.use_joint_priors = 1; break;
        case 'v': version_info(); exit(0); break;
        case 't': tumor_id = optarg; break;
        case 'n': normal_id = optarg; break;
        case 'L': R.prm.include_loh = 0; break;
        case 'G': R.prm.include_gor = 0; break;
        default: fprintf(stderr, "Unrecognizd option '-%c'.\n", c); return 1;
        }
    }
    R.prm.use_priors = use_priors;
    if (optind == argc) { usage(argv[0], &R.prm, mapq); return 1; }
    if (!fn_fa) {
        fprintf(stderr, "You MUST specify a reference sequence. It isn't optional.\n");
        exit(1);
    }
    if (argc - optind < 3) {
        /* the reference reads argv past its end here; ask for the missing arguments instead */
        usage(argv[0], &R.prm, mapq);
        return 1;
    }
    R.n_gpus = env_int("SSN_GPUS", 1, 1, MAX_GPUS);
    for (i = 0; i < MAX_GPUS; ++i) R.gpus[i].device = i;
    {
        const char *dv = getenv("SSN_DEVICES");
        if (dv && *dv) {
            R.n_gpus = 0;
            while (*dv && R.n_gpus < MAX_GPUS) {
                R.gpus[R.n_gpus++].device = atoi(dv);
                while (*dv && *dv != ',') ++dv;
                if (*dv == ',') ++dv;
            }
            if (R.n_gpus == 0) R.n_gpus = 1;
        }
    }
    if (!getenv("CUDA_VISIBLE_DEVICES")) {
        /* the driver initialises every visible GPU at the first CUDA call (about a second each): show it the ones this run
         * uses and renumber them */
        char list[16 * MAX_GPUS] = "";
        int seen[MAX_GPUS], n_seen = 0, k;
        for (i = 0; i < R.n_gpus; ++i) {
            for (k = 0; k < n_seen; ++k) if (seen[k] == R.gpus[i].device) break;
            if (k == n_seen) {
                seen[n_seen++] = R.gpus[i].device;
                snprintf(list + strlen(list), sizeof(list) - strlen(list), "%s%d", n_seen > 1 ? "," : "", R.gpus[i].device);
            }
            R.gpus[i].device = k;
        }
        setenv("CUDA_VISIBLE_DEVICES", list, 1);
    }
    for (i = 0; i < R.n_gpus; ++i) { R.gpus[i].run = &R; R.gpus[i].index = i; }
    R.n_slots = env_int("SSN_SLOTS", 3, 1, MAX_SLOTS);
    {
        long cores = sysconf(_SC_NPROCESSORS_ONLN);
        R.n_threads = env_int("SSN_THREADS", cores > 3 ? (int)cores - 2 : 1, 0, 256);
    }
    /* Inflating on the GPU pays where host cores are scarce: measured on B200, 14 planned threads stream 4.5e7 columns/s
     * with it against 3.5e7 without; 30 threads stream 6.5e7 either way (framing and copying bound them then), and the
     * extra hardware queues cost set-up time on every GPU.  So: on below 24 threads, unless SSN_GPU_INFLATE says otherwise. */
    gpu_inflate_wanted = env_int("SSN_GPU_INFLATE", R.n_threads < 24, 0, 1);
    /* 8 inflate streams + the scoring streams: one hardware queue each (see ssn_inflate_api.cu) */
    if (gpu_inflate_wanted) setenv("CUDA_DEVICE_MAX_CONNECTIONS", "16", 0);
    R.device_pileup = env_int("SSN_DEVICE_PILEUP", 1, 0, 1);
    R.format = format;
    pthread_mutex_init(&R.mu, 0);
    pthread_cond_init(&R.cv, 0);

    /* The readers of the first region shards start inside open_set and read ahead while the GPU contexts are built.
     * Same order of stderr lines as main.c:115-131. */
    if (open_set(&R, argv[optind], argv[optind + 1], fn_fa, mapq, env_int("SSN_SERIAL", 0, 0, 1)) != 0) return 1;
    n_shards_first = R.n_shards;
    if (R.prm.use_joint_priors)
        fprintf(stderr, "Using priors accounting for somatic mutation rate. Prior probability of a somatic mutation is %f\n", R.prm.somatic_mutation_rate);
    if (start_feeders(&R) != 0) {
        join_feeders(&R);
        free_outs(&R);
        ssn_shardset_close(R.set);
        return 1;
    }
    fprintf(stderr, "Preparing to snipe some somatics\n");
    if (use_priors) fprintf(stderr, "Using prior probabilities\n");
    fprintf(stderr, "Normal bam is %s\n", argv[optind + 1]);
    fprintf(stderr, "Tumor bam is %s\n", argv[optind]);
    out = fopen(argv[optind + 2], "w");
    wr = ssn_writer_create(format, out);
    if (!wr) {
        fprintf(stderr, "unknown output format: '%s'. Abort!\n", format);     /* output_format.c:31-32 */
        exit(1);
    }
    if (!out) {
        fprintf(stderr, "Unable to open snp file!!!!!!!!!\n");
        exit(1);
    }
    ssn_writer_header(wr, fn_fa, normal_id, tumor_id);
    if (gpu_inflate_wanted && ssn_shardset_n_shards(R.set) > 1) {
        /* lanes = inflate batches in flight per GPU; a lane's buffers are allocated when it is first taken */
        const uint32_t lanes = (uint32_t)env_int("SSN_INFLATE_LANES", 24, 1, 256);
        const uint32_t nblk = (uint32_t)env_int("SSN_INFLATE_BLOCKS", (int)INFL_BLOCKS, 4, 512);
        int ok = 1;
        for (i = 0; i < R.n_gpus && ok; ++i)
            ok = ssn_inflater_create(R.gpus[i].device, lanes, (uint64_t)nblk * 32768u + 65600u, (uint64_t)nblk * 65536u, nblk, &infl.inf[i]) == SSN_OK;
        if (ok) {
            infl.n = R.n_gpus;
            svc.ctx = &infl; svc.lane_acquire = svc_acquire; svc.lane_release = svc_release; svc.run = svc_run;
            /* measured on B200 + 16 cores: no gain (5.1 against 4.9e7 columns/s), so off unless asked for */
            if (env_int("SSN_INFLATE_DIRECT", 0, 0, 1)) { svc.host_alloc = svc_host_alloc; svc.host_free = svc_host_free; svc.run_to = svc_run_to; }
            ssn_shardset_set_inflate_service(R.set, &svc);
            gpu_inflate = 1;
        } else fprintf(stderr, "[bam-somaticsniper] no GPU inflate (%s); the readers inflate on the host\n", ssn_last_error());
    }
    t_setup = now_s() - t_start;

    rc = stream_pass(&R, out);
    ssn_shardset_get_stats(R.set, &st);
    if (rc == 2) {
        /* second pass, one interval front to back: exact for any input the reference accepts */
        fprintf(stderr, "[bam-somaticsniper] a read longer than the shards' look-back crosses an interval edge; walking the pair as one interval\n");
        free_outs(&R);
        ssn_shardset_close(R.set); R.set = 0;
        fflush(out);
        if (ftruncate(fileno(out), 0) != 0 || fseek(out, 0, SEEK_SET) != 0) { fprintf(stderr, "[bam-somaticsniper] cannot rewind the output file\n"); return 1; }
        ssn_writer_header(wr, fn_fa, normal_id, tumor_id);
        for (i = 0; i < R.n_gpus; ++i) { gpu_t *g = &R.gpus[i]; g->n_sites = g->n_batches = g->survivors = g->dev_columns = g->dev_entries = g->col_batches = 0; }
        passes = 2;
        if (open_set(&R, argv[optind], argv[optind + 1], fn_fa, mapq, 1) != 0) return 1;
        if (start_feeders(&R) != 0) { join_feeders(&R); return 1; }
        rc = stream_pass(&R, out);
        ssn_shardset_get_stats(R.set, &st);
        if (rc == 2) rc = 1;
    }
    fflush(out);
    for (i = 0; i < infl.n; ++i) {
        uint64_t b = 0, r = 0;
        if (ssn_inflater_stats(infl.inf[i], &b, &r, 0, 0, 0) == SSN_OK) { infl_blocks += b; infl_refused += r; }
    }
    {
        const char *sp = getenv("SSN_STATS");
        if (sp && *sp) {
            FILE *sf = fopen(sp, "w");
            if (sf) {
                uint64_t launches = 0, n_sites = 0, n_batches = 0, survivors = 0, dcol = 0, dent = 0, colb = 0;
                double t_gpu_wait = 0, t_ctx = 0;
                const double wall = now_s() - t_start;
                for (i = 0; i < R.n_gpus; ++i) {
                    const gpu_t *g = &R.gpus[i];
                    if (g->ctx) launches += ssn_kernel_launches(g->ctx);
                    n_sites += g->n_sites; n_batches += g->n_batches; survivors += g->survivors; dcol += g->dev_columns; dent += g->dev_entries;
                    colb += g->col_batches; t_gpu_wait += g->t_gpu_wait;
                    if (g->t_setup > t_ctx) t_ctx = g->t_setup;
                }
                /* columns / read bases: the device counts what it built, the host packer what it packed */
                st.columns += dcol; st.entries += dent;
                fprintf(sf, "{\"columns\": %llu, \"read_bases\": %llu, \"batches\": %llu, \"sites\": %llu, \"records_tumor\": %llu, "
                        "\"records_normal\": %llu, \"wall_s\": %.6f, \"setup_s\": %.6f, \"gpu_setup_s\": %.6f, \"pack_wait_s\": %.6f, \"pack_count_s\": %.6f, "
                        "\"pack_fill_s\": %.6f, \"gpu_wait_s\": %.6f, \"gpus\": %d, \"threads\": %d, \"slots\": %d, \"kernel_launches\": %llu, "
                        "\"prefilter\": %d, \"prefilter_survivors\": %llu, \"device_pileup\": %d, \"column_route_batches\": %llu, \"shards\": %d, "
                        "\"passes\": %d, \"gpu_inflate\": %d, \"gpu_inflate_blocks\": %llu, \"gpu_inflate_refused\": %llu, \"columns_per_s\": %.1f}\n",
                        (unsigned long long)st.columns, (unsigned long long)st.entries, (unsigned long long)n_batches,
                        (unsigned long long)n_sites, (unsigned long long)st.records[0], (unsigned long long)st.records[1], wall, t_setup, t_ctx,
                        st.t_wait, st.t_count, st.t_fill, t_gpu_wait, R.n_gpus, R.n_threads, R.n_slots, (unsigned long long)launches,
                        R.gpus[0].ctx ? ssn_prefilter_enabled(R.gpus[0].ctx) : 0, (unsigned long long)survivors, R.device_pileup,
                        (unsigned long long)colb, (int)n_shards_first, passes, gpu_inflate, (unsigned long long)infl_blocks, (unsigned long long)infl_refused, wall > t_setup ? (double)st.columns / (wall - t_setup) : 0.0);
                fclose(sf);
            }
        }
    }
    free_outs(&R);
    ssn_shardset_close(R.set);                      /* joins the readers: their lanes are back */
    for (i = 0; i < infl.n; ++i) ssn_inflater_destroy(infl.inf[i]);
    for (i = 0; i < R.n_gpus; ++i) {
        int k;
        gpu_t *g = &R.gpus[i];
        if (g->ctx) ssn_destroy(g->ctx);
        for (k = 0; k < g->n_meta; ++k) { free(g->meta[k].map.tid); free(g->meta[k].map.pos); free(g->meta[k].map.ref_char); free(g->meta[k].ref_chars); }
    }
    ssn_writer_destroy(wr);
    fclose(out);
    return rc;
}
