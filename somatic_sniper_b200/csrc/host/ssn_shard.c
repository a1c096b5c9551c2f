/*
 * ssn_shard.c — region shards of one tumor/normal run (include/ssn_host.h).
 *
 * The reference walks both BAMs front to back on one thread (bam_sspileup_file, src/lib/sniper/sniper_pileup.c:226-266).
 * Columns are independent, so the genome is cut into contiguous intervals at multiples of the packer's window size; every
 * interval gets its own pair of readers (seek by bisection, no index file: ssn_bamseek.c) and its own packer, several
 * intervals are read ahead at once, and the callers (one feeder per GPU) take them in genome order.  What couples
 * neighbours in the reference's walk is reproduced exactly:
 *   - reads that begin in front of an interval and reach into it: every reader starts `lookback` positions early; a read
 *     that begins even earlier and still reaches in raises the long-span flag (the caller then runs the pair again as one
 *     interval) — ssn_bamread.c;
 *   - the keep rule of sniper_pileup.c:216 (QUIRK B1): a reader takes over the walk's cursor from the last passing read
 *     in front of its first record.
 * Inputs that cannot be cut (tumor on stdin, a genome of one interval, a BAM whose records the seek cannot recognise)
 * run as one interval with the front-to-back reader.
 */
#include <limits.h>
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

#include "ssn_pack_internal.h"

enum { SH_NEW = 0, SH_OPENING, SH_OPEN, SH_CLAIMED, SH_DONE };

typedef struct {
    ssn_bam_range rg;
    ssn_packer *pk;
    int state, readers_left;
    struct ssn_shardset *set;
} shard;

struct ssn_shardset {
    ssn_bamfile *bf[2];
    ssn_refcache *rc;
    ssn_pool *pool;
    ssn_packer *serial;            /* one interval, front-to-back readers (owns its pool, files and sequences) */
    int min_mapq;
    shard *sh;
    int32_t n, next_open, next_claim, n_done;
    int reading, max_reading, max_open, max_reading_user;
    const ssn_inflate_service *svc;   /* readers of intervals opened from now on inflate through it */
    int started;                   /* a caller has asked for a shard: until then only two intervals are read ahead, so that the
                                    * readers do not crowd out the caller's own set-up (GPU contexts) */
    int long_span, failed;
    pthread_mutex_t mu;
    pthread_cond_t cv;
    ssn_packer_stats st;
    char err[256];
};

static void reader_finished(void *arg)
{
    shard *s = (shard*)arg;
    ssn_shardset *set = s->set;
    pthread_mutex_lock(&set->mu);
    if (--s->readers_left == 0) { set->reading -= 1; pthread_cond_broadcast(&set->cv); }
    pthread_mutex_unlock(&set->mu);
}

/* mu held on entry and exit; opens shards while the read-ahead limits allow */
static void open_more(ssn_shardset *set)
{
    while (set->next_open < set->n && set->reading < (set->started ? set->max_reading : (set->max_reading < 2 ? set->max_reading : 2)) &&
           set->next_open - set->n_done < set->max_open) {
        shard *s = &set->sh[set->next_open++];
        ssn_packer *pk;
        char err[256] = "";
        s->state = SH_OPENING; s->readers_left = 2; set->reading += 1;
        s->rg.inflate_service = set->svc;
        pthread_mutex_unlock(&set->mu);
        pk = ssn_packer_create_shard(set->pool, 0, set->rc, set->bf[0], set->bf[1], &s->rg, set->min_mapq, reader_finished, s, err, sizeof(err));
        pthread_mutex_lock(&set->mu);
        if (!pk) {
            set->failed = 1; snprintf(set->err, sizeof(set->err), "%s", err);
            /* the readers that did start have been joined by the failed constructor */
            if (s->readers_left) { s->readers_left = 0; set->reading -= 1; }
        }
        s->pk = pk; s->state = SH_OPEN;
        pthread_cond_broadcast(&set->cv);
    }
}

static uint32_t env_u32(const char *name, uint32_t dflt)
{
    const char *e = getenv(name);
    return e && *e ? (uint32_t)strtoul(e, 0, 10) : dflt;
}

ssn_status ssn_shardset_open(const char *tumor_bam, const char *normal_bam, const char *fasta, int min_mapq,
                             const ssn_shard_opts *opts, ssn_shardset **out)
{
    ssn_shardset *set;
    ssn_shard_opts o;
    char err[256];
    int cuttable;
    if (!tumor_bam || !normal_bam || !out) { ssn_host_set_error("ssn_shardset_open: null argument"); return SSN_ERR_ARG; }
    memset(&o, 0, sizeof(o));
    if (opts) o = *opts;
    if (o.n_threads <= 0) { long c = sysconf(_SC_NPROCESSORS_ONLN); o.n_threads = c > 3 ? (int)c - 2 : 1; }
    if (!o.stride_windows) o.stride_windows = env_u32("SSN_SHARD_WINDOWS", 0);
    if (!o.lookback_windows) o.lookback_windows = env_u32("SSN_SHARD_LOOKBACK", 4);
    if (!o.max_reading) o.max_reading = (int)env_u32("SSN_SHARD_READERS", 0);
    set = (ssn_shardset*)calloc(1, sizeof(*set));
    if (!set) { ssn_host_set_error("out of memory"); return SSN_ERR_NOMEM; }
    pthread_mutex_init(&set->mu, 0);
    pthread_cond_init(&set->cv, 0);
    set->min_mapq = min_mapq;
    cuttable = strcmp(tumor_bam, "-") != 0 && strcmp(normal_bam, "-") != 0 && !o.serial;
    if (cuttable) {
        set->bf[0] = ssn_bamfile_open(tumor_bam, err, sizeof(err));
        if (set->bf[0]) set->bf[1] = ssn_bamfile_open(normal_bam, err, sizeof(err));
        if (!set->bf[0] || !set->bf[1]) cuttable = 0;            /* the front-to-back reader reports what is wrong */
    }
    if (cuttable) {
        /* cut points: every `stride` positions of the concatenated targets (tumor header; the reference compares target
         * ids numerically, QUIRK B9), at window multiples inside a target */
        const ssn_bam_header *h = ssn_bamfile_header(set->bf[0]);
        uint64_t genome = 0, stride, next_cut, g = 0;
        int32_t t, cap = 16, n = 0;
        ssn_bam_range *cuts;
        set->max_reading_user = o.max_reading;
        if (!o.max_reading) o.max_reading = o.n_threads >= 3 ? (2 * o.n_threads + 2) / 3 : 1;
        if (o.max_reading > 48) o.max_reading = 48;              /* two threads and up to ~75 MB of finished windows each */
        if (!o.max_open) o.max_open = o.max_reading + (o.max_reading < 8 ? o.max_reading + 2 : 8);
        for (t = 0; t < h->n_targets; ++t) genome += h->target_len[t];
        if (o.stride_windows) stride = (uint64_t)o.stride_windows << SSN_WINDOW_SHIFT;
        else {
            /* An interval should (a) leave every reader pair several intervals to take and (b) hold about 48 MB of
             * records in its larger file, so that its windows fit the reader's queue and read-ahead never stalls on
             * back-pressure (records inflate to roughly 3.5 x their share of the file). */
            const uint64_t big = ssn_bamfile_size(set->bf[0]) > ssn_bamfile_size(set->bf[1]) ? ssn_bamfile_size(set->bf[0]) : ssn_bamfile_size(set->bf[1]);
            const double per_pos = genome ? 3.5 * (double)big / (double)genome : 0.0;
            uint64_t by_bytes = per_pos > 0.0 ? (uint64_t)(((double)((uint64_t)48 << 20)) / per_pos) : genome;
            stride = genome / (4u * (uint64_t)o.max_reading + 1u);
            if (stride > by_bytes) stride = by_bytes;
            if (stride < ((uint64_t)8 << SSN_WINDOW_SHIFT)) stride = (uint64_t)8 << SSN_WINDOW_SHIFT;
            if (stride > ((uint64_t)1024 << SSN_WINDOW_SHIFT)) stride = (uint64_t)1024 << SSN_WINDOW_SHIFT;
            stride &= ~(uint64_t)(SSN_WINDOW - 1u);
        }
        cuts = (ssn_bam_range*)calloc((size_t)cap, sizeof(*cuts));
        cuts[0].lo_tid = 0; cuts[0].lo_pos = 0; n = 1;
        next_cut = stride;
        for (t = 0; t < h->n_targets; ++t) {
            const uint64_t len = h->target_len[t];
            while (next_cut < g + len) {
                const uint32_t pos = (uint32_t)(next_cut - g) & ~(SSN_WINDOW - 1u);
                if (!(cuts[n - 1].lo_tid == t && cuts[n - 1].lo_pos == pos) && !(t == 0 && pos == 0)) {
                    if (n == cap) { cap *= 2; cuts = (ssn_bam_range*)realloc(cuts, sizeof(*cuts) * (size_t)cap); }
                    cuts[n].lo_tid = t; cuts[n].lo_pos = pos; ++n;
                }
                next_cut += stride;
            }
            g += len;
        }
        if (n > 1) {
            int32_t i;
            set->n = n;
            set->sh = (shard*)calloc((size_t)n, sizeof(shard));
            for (i = 0; i < n; ++i) {
                shard *s = &set->sh[i];
                s->set = set;
                s->rg.lo_tid = cuts[i].lo_tid; s->rg.lo_pos = cuts[i].lo_pos;
                s->rg.hi_tid = i + 1 < n ? cuts[i + 1].lo_tid : INT32_MAX;
                s->rg.hi_pos = i + 1 < n ? cuts[i + 1].lo_pos : 0;
                s->rg.lookback = o.lookback_windows << SSN_WINDOW_SHIFT;
                s->rg.queue_max_bytes = (size_t)192 << 20;
            }
            set->max_reading = o.max_reading; set->max_open = o.max_open;
            /* the readers inflate by themselves (one core each); the pool serves the packers' count / fill / copy tasks */
            {
                const char *pe = getenv("SSN_POOL_THREADS");                  /* copy workers of the packers (measurement knob) */
                int np = o.n_threads >= 4 ? o.n_threads / 2 : (o.n_threads >= 2 ? 2 : 0);
                if (pe && *pe && atoi(pe) >= 0 && atoi(pe) <= 256) np = atoi(pe);
                set->pool = ssn_pool_create(np);
            }
            set->rc = ssn_refcache_create(fasta, h);
        }
        free(cuts);
    }
    if (set->n == 0) {
        ssn_status st;
        if (set->bf[0]) { ssn_bamfile_close(set->bf[0]); set->bf[0] = 0; }
        if (set->bf[1]) { ssn_bamfile_close(set->bf[1]); set->bf[1] = 0; }
        st = ssn_packer_open(tumor_bam, normal_bam, fasta, min_mapq, o.n_threads, &set->serial);
        if (st != SSN_OK) { ssn_shardset_close(set); return st; }
        set->n = 1;
        set->sh = (shard*)calloc(1, sizeof(shard));
        set->sh[0].set = set; set->sh[0].rg.hi_tid = INT32_MAX;
    } else {
        pthread_mutex_lock(&set->mu);
        open_more(set);
        pthread_mutex_unlock(&set->mu);
    }
    *out = set;
    return SSN_OK;
}

void ssn_shardset_close(ssn_shardset *set)
{
    int32_t i;
    if (!set) return;
    for (i = 0; i < set->n && set->sh; ++i)
        if (set->sh[i].pk && set->sh[i].pk != set->serial) ssn_packer_close(set->sh[i].pk);
    if (set->serial) ssn_packer_close(set->serial);
    free(set->sh);
    ssn_refcache_destroy(set->rc);
    ssn_pool_destroy(set->pool);
    ssn_bamfile_close(set->bf[0]); ssn_bamfile_close(set->bf[1]);
    pthread_mutex_destroy(&set->mu);
    pthread_cond_destroy(&set->cv);
    free(set);
}

int32_t ssn_shardset_n_shards(const ssn_shardset *set) { return set->n; }

int32_t ssn_shardset_n_targets(const ssn_shardset *set, int which)
{
    if (set->serial) return ssn_packer_n_targets(set->serial, which);
    return ssn_bamfile_header(set->bf[which ? 1 : 0])->n_targets;
}

const char *ssn_shardset_target_name(const ssn_shardset *set, int which, int32_t tid)
{
    const ssn_bam_header *h;
    if (set->serial) return ssn_packer_target_name(set->serial, which, tid);
    h = ssn_bamfile_header(set->bf[which ? 1 : 0]);
    return (tid >= 0 && tid < h->n_targets) ? h->target_name[tid] : 0;
}

ssn_status ssn_shardset_range(const ssn_shardset *set, int32_t index, int32_t *lo_tid, uint32_t *lo_pos, int32_t *hi_tid, uint32_t *hi_pos)
{
    if (!set || index < 0 || index >= set->n) { ssn_host_set_error("ssn_shardset_range: no such shard"); return SSN_ERR_ARG; }
    if (lo_tid) *lo_tid = set->sh[index].rg.lo_tid;
    if (lo_pos) *lo_pos = set->sh[index].rg.lo_pos;
    if (hi_tid) *hi_tid = set->sh[index].rg.hi_tid;
    if (hi_pos) *hi_pos = set->sh[index].rg.hi_pos;
    return SSN_OK;
}

ssn_status ssn_shardset_next(ssn_shardset *set, int32_t *index, ssn_packer **packer)
{
    shard *s;
    if (!set || !index || !packer) { ssn_host_set_error("ssn_shardset_next: null argument"); return SSN_ERR_ARG; }
    *index = -1; *packer = 0;
    pthread_mutex_lock(&set->mu);
    if (set->next_claim >= set->n) { pthread_mutex_unlock(&set->mu); return SSN_OK; }
    set->started = 1;
    /* the index is taken before anything can wait: a second caller arriving while this one waits for its shard's readers
     * must get the NEXT shard, never the same one */
    const int32_t k = set->next_claim++;
    s = &set->sh[k];
    if (set->serial) { s->pk = set->serial; s->state = SH_OPEN; }
    while (s->state < SH_OPEN && !set->failed) {
        open_more(set);
        if (s->state < SH_OPEN && !set->failed) pthread_cond_wait(&set->cv, &set->mu);
    }
    if (set->failed || !s->pk) {
        ssn_host_set_error("region shard %d: %s", (int)k, set->err[0] ? set->err : "cannot start its readers");
        pthread_mutex_unlock(&set->mu);
        return SSN_ERR_STATE;
    }
    s->state = SH_CLAIMED;
    *index = k; *packer = s->pk;
    pthread_mutex_unlock(&set->mu);
    return SSN_OK;
}

ssn_status ssn_shardset_done(ssn_shardset *set, int32_t index)
{
    shard *s;
    ssn_packer *pk;
    ssn_packer_stats st;
    int ls, fl;
    if (!set || index < 0 || index >= set->n || set->sh[index].state != SH_CLAIMED) { ssn_host_set_error("ssn_shardset_done: shard was not claimed"); return SSN_ERR_ARG; }
    s = &set->sh[index];
    pk = s->pk;
    ssn_packer_get_stats(pk, &st);
    if (pk != set->serial) {
        ls = ssn_packer_long_span(pk); fl = ssn_packer_failed(pk);
        ssn_packer_close(pk);          /* stops and joins the readers */
    } else { ls = 0; fl = 0; }
    pthread_mutex_lock(&set->mu);
    s->pk = 0; s->state = SH_DONE; set->n_done += 1;
    set->long_span |= ls; set->failed |= fl;
    set->st.records[0] += st.records[0]; set->st.records[1] += st.records[1];
    set->st.columns += st.columns; set->st.entries += st.entries; set->st.batches += st.batches;
    set->st.t_wait += st.t_wait; set->st.t_count += st.t_count; set->st.t_fill += st.t_fill;
    if (!set->serial) open_more(set);
    pthread_cond_broadcast(&set->cv);
    pthread_mutex_unlock(&set->mu);
    if (fl) { ssn_host_set_error("region shard %d: the seek did not find its records (or memory ran out)", (int)index); return SSN_ERR_STATE; }
    return SSN_OK;
}

int ssn_shardset_long_span(const ssn_shardset *set) { return set ? set->long_span : 0; }

void ssn_shardset_set_inflate_service(ssn_shardset *set, const ssn_inflate_service *svc)
{
    if (!set) return;
    pthread_mutex_lock(&set->mu);
    set->svc = svc;
    /* with the inflating off the cores a reader mostly waits: more of them may run at once */
    if (svc && !set->serial && set->max_reading_user == 0) {
        set->max_reading = set->max_reading * 2 > 48 ? 48 : set->max_reading * 2;
        if (set->max_open < set->max_reading + 8) set->max_open = set->max_reading + 8;
    }
    pthread_mutex_unlock(&set->mu);
}

void ssn_shardset_get_stats(const ssn_shardset *set, ssn_packer_stats *out)
{
    ssn_shardset *s = (ssn_shardset*)set;
    pthread_mutex_lock(&s->mu);
    *out = s->st;
    pthread_mutex_unlock(&s->mu);
}
