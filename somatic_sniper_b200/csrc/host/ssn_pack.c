/*
 * ssn_pack.c — joins the tumor and normal window streams on (tid, window) and packs the pileup columns of
 * both samples into SoA batches (include/ssn_host.h).  Two parallel phases over the windows of a batch:
 *   count  per-position read-base counts of each sample (difference array over the CIGAR match runs)
 *   fill   per-column offsets, then every read base is written at its column's cursor
 * The pileup contents follow get_next_pos/resolve_cigar (src/lib/sniper/sniper_pileup.c:57-104,186-205): a
 * kept read contributes a base at every position covered by a CIGAR M run; positions inside D yield the
 * is_del entries every consumer skips (sniper_maqcns.c:147, dqstats.c:20) and positions inside N yield
 * nothing (:98); I and S advance the query only; H, P and the operators samtools-0.1.6 does not know
 * (=, X) advance nothing (QUIRK B6).
 */
#include <pthread.h>
#include <stdarg.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "ssn_bamread.h"
#include "ssn_bamseek.h"
#include "ssn_fasta.h"
#include "ssn_host.h"
#include "ssn_nt16.h"
#include "ssn_pack_internal.h"
#include "ssn_pool.h"

#define GROUP_MAX 16           /* windows pulled per batch at most */
#define SEG_MAX 1024           /* fill tasks per batch at most */

static __thread char h_err[512] = "";
const char *ssn_host_last_error(void) { return h_err; }
void ssn_host_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(h_err, sizeof(h_err), fmt, ap);
    va_end(ap);
}
#define set_err ssn_host_set_error

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

/* ---- target sequences: one cache for all packers of a run ----------------------------------------------------
 * A sequence lives as long as a joined window points at it (reference count) and, idle, until a few newer ones have
 * pushed it out; packers of different region shards share the cache (one fetch per target, not one per shard). */
typedef struct ssn_refseq {
    int32_t tid;
    char *seq; int len;
    int refs, loading;
    struct ssn_refseq *next;
} ssn_refseq;

struct ssn_refcache {
    pthread_mutex_t mu;
    pthread_cond_t cv;
    ssn_fasta *fa;                 /* owned; NULL: every base reads as 'N' */
    const ssn_bam_header *names;   /* target names come from the tumor header (QUIRK B9) */
    ssn_refseq *head;
};

#define REF_IDLE_MAX 2

ssn_refcache *ssn_refcache_create(const char *fasta, const ssn_bam_header *names)
{
    ssn_refcache *rc = (ssn_refcache*)calloc(1, sizeof(*rc));
    if (!rc) return 0;
    pthread_mutex_init(&rc->mu, 0);
    pthread_cond_init(&rc->cv, 0);
    /* like the reference (main.c:106-108, somatic_sniper.c:112-117) a FASTA that cannot be loaded is not fatal:
     * every reference base then reads as 'N' and no column is ever emitted */
    if (fasta) rc->fa = ssn_fasta_load(fasta);
    rc->names = names;
    return rc;
}

void ssn_refcache_destroy(ssn_refcache *rc)
{
    if (!rc) return;
    while (rc->head) { ssn_refseq *e = rc->head; rc->head = e->next; free(e->seq); free(e); }
    ssn_fasta_destroy(rc->fa);
    pthread_mutex_destroy(&rc->mu);
    pthread_cond_destroy(&rc->cv);
    free(rc);
}

/* mu held: drops idle sequences beyond REF_IDLE_MAX, lowest target first (targets arrive in ascending order) */
static void refcache_trim(ssn_refcache *rc)
{
    for (;;) {
        ssn_refseq **pp, **victim = 0;
        int idle = 0;
        for (pp = &rc->head; *pp; pp = &(*pp)->next)
            if ((*pp)->refs == 0 && !(*pp)->loading) { ++idle; if (!victim || (*pp)->tid < (*victim)->tid) victim = pp; }
        if (idle <= REF_IDLE_MAX) return;
        { ssn_refseq *e = *victim; *victim = e->next; free(e->seq); free(e); }
    }
}

static ssn_refseq *refcache_get(ssn_refcache *rc, int32_t tid)
{
    ssn_refseq *e;
    pthread_mutex_lock(&rc->mu);
    for (e = rc->head; e; e = e->next) if (e->tid == tid) break;
    if (e) {
        e->refs += 1;
        while (e->loading) pthread_cond_wait(&rc->cv, &rc->mu);
        pthread_mutex_unlock(&rc->mu);
        return e;
    }
    e = (ssn_refseq*)calloc(1, sizeof(*e));
    e->tid = tid; e->refs = 1; e->loading = 1;
    e->next = rc->head; rc->head = e;
    pthread_mutex_unlock(&rc->mu);
    if (rc->fa && tid >= 0 && tid < rc->names->n_targets) e->seq = ssn_fasta_fetch(rc->fa, rc->names->target_name[tid], &e->len);
    if (!e->seq) { e->seq = (char*)calloc(1, 1); e->len = 0; }
    pthread_mutex_lock(&rc->mu);
    e->loading = 0;
    pthread_cond_broadcast(&rc->cv);
    refcache_trim(rc);
    pthread_mutex_unlock(&rc->mu);
    return e;
}

static void refcache_put(ssn_refcache *rc, ssn_refseq *e)
{
    if (!rc || !e) return;
    pthread_mutex_lock(&rc->mu);
    e->refs -= 1;
    refcache_trim(rc);
    pthread_mutex_unlock(&rc->mu);
}

/* a joined window: both samples' reads plus the per-position counts */
typedef struct pair {
    ssn_window *w[2];             /* 0 tumor, 1 normal */
    uint32_t *cnt[2];             /* [SSN_WINDOW] read bases per position */
    uint32_t next_i;              /* first position index not yet emitted */
    const char *ref; int ref_len; /* target sequence */
    ssn_refcache *rc; ssn_refseq *rs;   /* where it came from: released with the window */
} pair;

typedef struct {                  /* one fill task */
    pair *p;
    uint32_t i0, i1;              /* position indices inside the window */
    uint32_t col_base;
    uint64_t ent_base;
} segment;

struct ssn_packer {
    ssn_pool *pool;
    int own_pool;
    ssn_bamreader *rd[2];
    ssn_window *head[2];          /* next window of each stream (look-ahead of the join) */
    int ended;
    ssn_refcache *rc;
    int own_rc;
    int one_window;               /* ssn_packer_next takes one new window per call (device-pileup fall-back) */
    pair *pending[GROUP_MAX + 1];
    int n_pending;
    ssn_packer_stats st;
    /* the batch being filled */
    const ssn_batch_view *view;
    const ssn_colmap *map;
    segment seg[SEG_MAX];
};

/* ---- CIGAR walk -------------------------------------------------------------------------------------------- */

static inline const uint32_t *read_cigar(const ssn_read *h) { return (const uint32_t*)((const uint8_t*)h + sizeof(*h)); }

/* adds +1/-1 marks for every M run of the read clipped to [lo, hi) */
static void count_read(const ssn_read *h, uint32_t lo, uint32_t hi, int32_t *diff)
{
    const uint32_t *cg = read_cigar(h);
    uint32_t x = (uint32_t)h->beg, k;
    for (k = 0; k < h->n_cigar && x < hi; ++k) {
        const uint32_t op = cg[k] & 0xf, l = cg[k] >> 4;
        if (op == 0) {
            const uint32_t a = x > lo ? x : lo, b = x + l < hi ? x + l : hi;
            if (a < b) { diff[a - lo] += 1; diff[b - lo] -= 1; }
            x += l;
        } else if (op == 2 || op == 3) x += l;
    }
}

static void count_task(void *arg, int idx)
{
    pair *p = ((pair**)arg)[idx];
    int s;
    for (s = 0; s < 2; ++s) {
        const ssn_window *w = p->w[s];
        const uint32_t lo = w->widx << SSN_WINDOW_SHIFT, hi = lo + SSN_WINDOW;
        int32_t *diff = (int32_t*)calloc(SSN_WINDOW + 1, sizeof(int32_t));
        uint32_t *cnt = (uint32_t*)malloc(sizeof(uint32_t) * SSN_WINDOW);
        size_t o = 0;
        uint32_t i;
        int32_t run = 0;
        while (o < w->len) {
            const ssn_read *h = (const ssn_read*)(w->buf + o);
            count_read(h, lo, hi, diff);
            o += h->bytes;
        }
        for (i = 0; i < SSN_WINDOW; ++i) { run += diff[i]; cnt[i] = (uint32_t)run; }
        free(diff);
        p->cnt[s] = cnt;
    }
    /* a column exists where both samples have a read base and the reference base is callable */
    {
        const uint32_t lo = p->w[0]->widx << SSN_WINDOW_SHIFT;
        uint32_t i;
        for (i = 0; i < SSN_WINDOW; ++i) {
            const uint32_t pos = lo + i;
            const int rb = (p->ref && (int)pos < p->ref_len) ? (unsigned char)p->ref[pos] : 'N';
            if (p->cnt[0][i] == 0 || p->cnt[1][i] == 0 || ssn_nt16_table(rb) == 15) { p->cnt[0][i] = 0; p->cnt[1][i] = 0; }
        }
    }
}

/* writes the read's bases that fall on live columns of [lo + i0, lo + i1) at their cursors */
static void fill_read(const ssn_read *h, uint32_t lo, uint32_t a_lo, uint32_t a_hi, uint32_t *cur,
                      uint8_t *bs, uint8_t *bq, uint8_t *mq)
{
    const uint32_t *cg = read_cigar(h);
    const uint8_t *seq = (const uint8_t*)(cg + h->n_cigar), *qual = seq + ((h->l_qseq + 1) >> 1);
    const uint8_t strand = (uint8_t)(h->strand << 4), mapq = h->mapq;
    uint32_t x = (uint32_t)h->beg, y = 0, k;
    for (k = 0; k < h->n_cigar && x < a_hi; ++k) {
        const uint32_t op = cg[k] & 0xf, l = cg[k] >> 4;
        if (op == 0) {
            const uint32_t a = x > a_lo ? x : a_lo, b = x + l < a_hi ? x + l : a_hi;
            uint32_t pos;
            for (pos = a; pos < b; ++pos) {
                const uint32_t c = cur[pos - lo];
                if (c != 0xffffffffu) {
                    const uint32_t q = y + (pos - x);
                    if (q < h->l_qseq) {
                        bs[c] = (uint8_t)(((seq[q >> 1] >> ((~q & 1) << 2)) & 0xf) | strand);   /* bam1_seqi */
                        bq[c] = qual[q];
                        mq[c] = mapq;
                    } else { bs[c] = (uint8_t)(15 | strand); bq[c] = 0; mq[c] = mapq; }      /* CIGAR longer than the read */
                    cur[pos - lo] = c + 1;
                }
            }
            x += l; y += l;
        } else if (op == 2 || op == 3) x += l;
        else if (op == 1 || op == 4) y += l;
    }
}

static void fill_task(void *arg, int idx)
{
    ssn_packer *pk = (ssn_packer*)arg;
    const segment *sg = &pk->seg[idx];
    pair *p = sg->p;
    const ssn_batch_view *v = pk->view;
    const ssn_colmap *m = pk->map;
    const uint32_t lo = p->w[0]->widx << SSN_WINDOW_SHIFT;
    uint32_t col = sg->col_base, i;
    uint64_t e = sg->ent_base;
    int s;
    /* offsets and cursors; the counts of [i0, i1) are consumed */
    for (i = sg->i0; i < sg->i1; ++i) {
        const uint32_t ct = p->cnt[0][i], cn = p->cnt[1][i];
        if (ct == 0) { p->cnt[0][i] = 0xffffffffu; p->cnt[1][i] = 0xffffffffu; continue; }
        {
            const uint32_t pos = lo + i;
            const int rb = (unsigned char)p->ref[pos];
            v->off[2 * (size_t)col] = (uint32_t)e;
            v->off[2 * (size_t)col + 1] = (uint32_t)(e + ct);
            v->ref_nt16[col] = (uint8_t)ssn_nt16_table(rb);
            m->tid[col] = p->w[0]->tid; m->pos[col] = pos; m->ref_char[col] = (uint8_t)rb;
            p->cnt[0][i] = (uint32_t)e; p->cnt[1][i] = (uint32_t)(e + ct);
            e += (uint64_t)ct + cn;
            ++col;
        }
    }
    for (s = 0; s < 2; ++s) {
        const ssn_window *w = p->w[s];
        size_t o = 0;
        while (o < w->len) {
            const ssn_read *h = (const ssn_read*)(w->buf + o);
            if ((uint32_t)h->end > lo + sg->i0 && (uint32_t)h->beg < lo + sg->i1)
                fill_read(h, lo, lo + sg->i0, lo + sg->i1, p->cnt[s], v->bs, v->bq, v->mq);
            o += h->bytes;
        }
    }
}

/* ---- join ---------------------------------------------------------------------------------------------------- */

static void pair_free(pair *p)
{
    if (!p) return;
    refcache_put(p->rc, p->rs);
    ssn_window_free(p->w[0]); ssn_window_free(p->w[1]);
    free(p->cnt[0]); free(p->cnt[1]);
    free(p);
}

static int win_cmp(const ssn_window *a, const ssn_window *b)
{
    if (a->tid != b->tid) return a->tid < b->tid ? -1 : 1;
    if (a->widx != b->widx) return a->widx < b->widx ? -1 : 1;
    return 0;
}

/* next window present in both streams, NULL when either stream is exhausted (sniper_pileup.c:245) */
static pair *next_pair(ssn_packer *pk)
{
    pair *p;
    if (pk->ended) return 0;
    for (;;) {
        int s, c;
        for (s = 0; s < 2; ++s)
            if (!pk->head[s]) {
                pk->head[s] = ssn_bamreader_pop(pk->rd[s]);
                if (!pk->head[s]) { pk->ended = 1; return 0; }
            }
        c = win_cmp(pk->head[0], pk->head[1]);
        if (c == 0) break;
        s = c < 0 ? 0 : 1;                       /* the stream that is behind has no partner for this window */
        ssn_window_free(pk->head[s]);
        pk->head[s] = 0;
    }
    p = (pair*)calloc(1, sizeof(*p));
    p->w[0] = pk->head[0]; p->w[1] = pk->head[1];
    pk->head[0] = pk->head[1] = 0;
    return p;
}

/* target sequence of the window, shared and reference-counted (released in pair_free) */
static void attach_ref(ssn_packer *pk, pair *p)
{
    p->rc = pk->rc;
    p->rs = refcache_get(pk->rc, p->w[0]->tid);
    p->ref = p->rs->seq; p->ref_len = p->rs->len;
}

/* ---- API -------------------------------------------------------------------------------------------------------- */

ssn_status ssn_packer_open(const char *tumor_bam, const char *normal_bam, const char *fasta, int min_mapq,
                           int n_threads, ssn_packer **out)
{
    ssn_packer *pk;
    char err[256];
    if (!tumor_bam || !normal_bam || !out) { set_err("ssn_packer_open: null argument"); return SSN_ERR_ARG; }
    pk = (ssn_packer*)calloc(1, sizeof(*pk));
    if (!pk) { set_err("out of memory"); return SSN_ERR_NOMEM; }
    pk->pool = ssn_pool_create(n_threads); pk->own_pool = 1;
    pk->rd[0] = ssn_bamreader_open(tumor_bam, -1, min_mapq, pk->pool, err, sizeof(err));
    if (!pk->rd[0]) { set_err("tumor BAM: %s", err); ssn_packer_close(pk); return SSN_ERR_ARG; }
    pk->rd[1] = ssn_bamreader_open(normal_bam, -1, min_mapq, pk->pool, err, sizeof(err));
    if (!pk->rd[1]) { set_err("normal BAM: %s", err); ssn_packer_close(pk); return SSN_ERR_ARG; }
    pk->rc = ssn_refcache_create(fasta, ssn_bamreader_header(pk->rd[0])); pk->own_rc = 1;
    *out = pk;
    return SSN_OK;
}

/* one region shard of a run (ssn_shard.c): pool, target sequences and the two open files belong to the shard set */
ssn_packer *ssn_packer_create_shard(ssn_pool *pool, ssn_pool *reader_pool, ssn_refcache *rc, const ssn_bamfile *tumor,
                                    const ssn_bamfile *normal, const ssn_bam_range *range, int min_mapq,
                                    void (*on_finish)(void*), void *on_finish_arg, char *err, size_t errlen)
{
    ssn_packer *pk = (ssn_packer*)calloc(1, sizeof(*pk));
    if (!pk) { snprintf(err, errlen, "out of memory"); return 0; }
    pk->pool = pool; pk->rc = rc;
    pk->rd[0] = ssn_bamreader_open_range(tumor, range, -1, min_mapq, reader_pool, on_finish, on_finish_arg, err, errlen);
    if (!pk->rd[0]) { ssn_packer_close(pk); return 0; }
    pk->rd[1] = ssn_bamreader_open_range(normal, range, -1, min_mapq, reader_pool, on_finish, on_finish_arg, err, errlen);
    if (!pk->rd[1]) { ssn_packer_close(pk); return 0; }
    return pk;
}

/* Device-pileup fall-back: a window too deep for a read batch is packed into columns by ssn_packer_next.  With the limit
 * set, ssn_packer_next works on one window at a time, so that ssn_packer_partial() turns 0 as soon as that window is used
 * up and the caller can go back to ssn_packer_next_reads. */
void ssn_packer_one_window(ssn_packer *pk, int on) { pk->one_window = on; }
int ssn_packer_partial(const ssn_packer *pk) { return pk->n_pending > 0 && pk->pending[0]->next_i > 0; }

int ssn_packer_long_span(const ssn_packer *pk) { return ssn_bamreader_long_span(pk->rd[0]) || ssn_bamreader_long_span(pk->rd[1]); }
int ssn_packer_failed(const ssn_packer *pk) { return ssn_bamreader_failed(pk->rd[0]) || ssn_bamreader_failed(pk->rd[1]); }

void ssn_packer_close(ssn_packer *pk)
{
    int i;
    if (!pk) return;
    if (pk->rd[0]) ssn_bamreader_close(pk->rd[0]);
    if (pk->rd[1]) ssn_bamreader_close(pk->rd[1]);
    for (i = 0; i < pk->n_pending; ++i) pair_free(pk->pending[i]);
    ssn_window_free(pk->head[0]); ssn_window_free(pk->head[1]);
    if (pk->own_rc) ssn_refcache_destroy(pk->rc);
    if (pk->own_pool) ssn_pool_destroy(pk->pool);
    free(pk);
}

int32_t ssn_packer_n_targets(const ssn_packer *pk, int which) { return ssn_bamreader_header(pk->rd[which ? 1 : 0])->n_targets; }

const char *ssn_packer_target_name(const ssn_packer *pk, int which, int32_t tid)
{
    const ssn_bam_header *h = ssn_bamreader_header(pk->rd[which ? 1 : 0]);
    return (tid >= 0 && tid < h->n_targets) ? h->target_name[tid] : 0;
}

void ssn_packer_get_stats(const ssn_packer *pk, ssn_packer_stats *out)
{
    *out = pk->st;
    out->records[0] = ssn_bamreader_records(pk->rd[0]);
    out->records[1] = ssn_bamreader_records(pk->rd[1]);
}

ssn_status ssn_packer_next(ssn_packer *pk, const ssn_batch_view *view, const ssn_colmap *map,
                           uint32_t *n_columns, uint64_t *n_entries)
{
    uint32_t ncol = 0;
    uint64_t nent = 0;
    int nseg = 0, i, fresh0, full = 0;
    double t0, t1;
    if (!pk || !view || !map || !n_columns || !n_entries) { set_err("ssn_packer_next: null argument"); return SSN_ERR_ARG; }
    if (view->max_columns == 0 || view->max_entries == 0 || view->max_entries >= 0xffffffffull) { set_err("ssn_packer_next: bad batch capacity"); return SSN_ERR_ARG; }
    *n_columns = 0; *n_entries = 0;
    pk->view = view; pk->map = map;

    /* 1. pull joined windows while their upper bounds suggest they still fit */
    t0 = now_s();
    fresh0 = pk->n_pending;
    {
        uint64_t bound_ent = 0, bound_col = 0;
        for (i = 0; i < pk->n_pending; ++i) {
            bound_ent += pk->pending[i]->w[0]->n_bases + pk->pending[i]->w[1]->n_bases;
            bound_col += SSN_WINDOW - pk->pending[i]->next_i;
        }
        while (pk->n_pending < (pk->one_window ? 1 : GROUP_MAX) && bound_ent < view->max_entries && bound_col < view->max_columns) {
            pair *p = next_pair(pk);
            if (!p) break;
            attach_ref(pk, p);
            pk->pending[pk->n_pending++] = p;
            bound_ent += p->w[0]->n_bases + p->w[1]->n_bases;
            bound_col += SSN_WINDOW;
        }
    }
    t1 = now_s(); pk->st.t_wait += t1 - t0;
    if (pk->n_pending == 0) return SSN_OK;

    /* 2. count the new windows (a window ssn_packer_next_reads looked at and left behind was never counted) */
    while (fresh0 > 0 && !pk->pending[fresh0 - 1]->cnt[0]) --fresh0;
    if (pk->n_pending > fresh0) ssn_pool_for(pk->pool, pk->n_pending - fresh0, count_task, pk->pending + fresh0);
    t0 = now_s(); pk->st.t_count += t0 - t1;

    /* 3. cut the batch: whole windows while they fit, then the longest prefix of the next one.  Windows are split
     * into position sub-ranges of about `span` so that a batch of few (deep) windows still fills in parallel. */
    {
        const int workers = ssn_pool_threads(pk->pool) + 1;
        uint32_t span = (uint32_t)((uint64_t)SSN_WINDOW * (uint32_t)pk->n_pending / (uint32_t)(2 * workers));
        if (span < 128u) span = 128u;
        if (span > SSN_WINDOW) span = SSN_WINDOW;
        for (i = 0; i < pk->n_pending && !full; ++i) {
            pair *p = pk->pending[i];
            uint32_t j, c = 0, seg_i0 = p->next_i, seg_c = 0;
            uint64_t e = 0, seg_e = 0;
            for (j = p->next_i; j < SSN_WINDOW; ++j) {
                const uint32_t ct = p->cnt[0][j];
                if (j - seg_i0 >= span && nseg < SEG_MAX - 2) {           /* close a sub-range */
                    segment *sg = &pk->seg[nseg++];
                    sg->p = p; sg->i0 = seg_i0; sg->i1 = j; sg->col_base = ncol + seg_c; sg->ent_base = nent + seg_e;
                    seg_i0 = j; seg_c = c; seg_e = e;
                }
                if (ct == 0) continue;
                {
                    const uint64_t add = (uint64_t)ct + p->cnt[1][j];
                    if (ncol + c + 1 > view->max_columns || nent + e + add > view->max_entries) { full = 1; break; }
                    c += 1; e += add;
                }
            }
            if (full && c == 0 && ncol == 0) {
                set_err("column %d:%u holds %llu read bases; the batch capacity is %llu", p->w[0]->tid,
                        (p->w[0]->widx << SSN_WINDOW_SHIFT) + j, (unsigned long long)((uint64_t)p->cnt[0][j] + p->cnt[1][j]),
                        (unsigned long long)view->max_entries);
                return SSN_ERR_OVERFLOW;
            }
            if (j > seg_i0) {
                segment *sg = &pk->seg[nseg++];
                sg->p = p; sg->i0 = seg_i0; sg->i1 = j; sg->col_base = ncol + seg_c; sg->ent_base = nent + seg_e;
            }
            ncol += c; nent += e;
            p->next_i = j;
        }
    }

    /* 4. fill */
    ssn_pool_for(pk->pool, nseg, fill_task, pk);
    view->off[2 * (size_t)ncol] = (uint32_t)nent;
    t1 = now_s(); pk->st.t_fill += t1 - t0;

    /* 5. retire the windows that are used up */
    {
        int k = 0;
        for (i = 0; i < pk->n_pending; ++i) {
            if (pk->pending[i]->next_i >= SSN_WINDOW) pair_free(pk->pending[i]);
            else pk->pending[k++] = pk->pending[i];
        }
        pk->n_pending = k;
    }
    pk->st.columns += ncol; pk->st.entries += nent; pk->st.batches += 1;
    *n_columns = ncol; *n_entries = nent;
    if (ncol == 0) return ssn_packer_next(pk, view, map, n_columns, n_entries);   /* windows without a common column */
    return SSN_OK;
}

/* ---- read batches for the device pileup: whole joined windows, records copied as they are ----------------------- */

typedef struct {
    pair *p;
    int s;                     /* sample; 2 = the window's reference slice */
    uint8_t *dst; uint32_t *off; uint32_t off_base;   /* records -> dst, word offsets (off_base + ...) -> off */
    uint8_t *nt16, *chars;
} reads_task;

/* Copy into a slot's pinned read array: the destination is read next by the GPU's copy engine, never by this host, so the
 * lines are written with streaming stores (no read-for-ownership, nothing of it kept in the caches the readers work in). */
#if defined(__x86_64__)
#include <emmintrin.h>
static int g_stream_copy = -1;
static void copy_streaming(uint8_t *dst, const uint8_t *src, size_t n)
{
    size_t head;
    int on = __atomic_load_n(&g_stream_copy, __ATOMIC_RELAXED);           /* several pool workers get here at once */
    if (on < 0) {
        const char *e = getenv("SSN_STREAM_COPY");                       /* measurement knob */
        on = e && *e ? atoi(e) != 0 : 1;
        __atomic_store_n(&g_stream_copy, on, __ATOMIC_RELAXED);
    }
    if (!on || n < 4096) { memcpy(dst, src, n); return; }
    head = (size_t)(-(uintptr_t)dst & 15u);
    memcpy(dst, src, head); dst += head; src += head; n -= head;
    for (; n >= 64; n -= 64, dst += 64, src += 64) {
        const __m128i a = _mm_loadu_si128((const __m128i*)src), b = _mm_loadu_si128((const __m128i*)(src + 16));
        const __m128i c = _mm_loadu_si128((const __m128i*)(src + 32)), d = _mm_loadu_si128((const __m128i*)(src + 48));
        _mm_stream_si128((__m128i*)dst, a); _mm_stream_si128((__m128i*)(dst + 16), b);
        _mm_stream_si128((__m128i*)(dst + 32), c); _mm_stream_si128((__m128i*)(dst + 48), d);
    }
    memcpy(dst, src, n);
    _mm_sfence();
}
#else
static void copy_streaming(uint8_t *dst, const uint8_t *src, size_t n) { memcpy(dst, src, n); }
#endif

static void reads_copy_task(void *arg, int idx)
{
    const reads_task *t = &((const reads_task*)arg)[idx];
    if (t->s == 2) {
        const uint32_t lo = t->p->w[0]->widx << SSN_WINDOW_SHIFT;
        uint32_t i;
        for (i = 0; i < SSN_WINDOW; ++i) {
            const uint32_t pos = lo + i;
            const int rb = (t->p->ref && (int)pos < t->p->ref_len) ? (unsigned char)t->p->ref[pos] : 'N';
            t->nt16[i] = (uint8_t)ssn_nt16_table(rb);
            if (t->chars) t->chars[i] = (uint8_t)rb;
        }
    } else {
        const ssn_window *w = t->p->w[t->s];
        uint32_t k;
        copy_streaming(t->dst, w->buf, w->len);
        for (k = 0; k < w->n_reads; ++k) t->off[k] = t->off_base + w->rec_word[k];
    }
}

ssn_status ssn_packer_next_reads(ssn_packer *pk, const ssn_reads_view *rv, uint32_t max_columns, uint64_t max_entries,
                                 uint8_t *ref_chars, uint32_t *n_windows, uint32_t *n_reads, uint64_t *rec_bytes)
{
    reads_task tasks[3 * GROUP_MAX];
    pair *took[GROUP_MAX];
    uint32_t nw = 0, nr = 0;
    uint64_t nb = 0, ne = 0;
    int nt = 0, i;
    double t0 = now_s(), t1;
    if (!pk || !rv || !n_windows || !n_reads || !rec_bytes) { set_err("ssn_packer_next_reads: null argument"); return SSN_ERR_ARG; }
    *n_windows = 0; *n_reads = 0; *rec_bytes = 0;
    for (;;) {
        pair *p;
        uint64_t b, r, e;
        if (pk->n_pending == 0) {
            p = next_pair(pk);
            if (!p) break;
            attach_ref(pk, p);
            pk->pending[pk->n_pending++] = p;
        }
        p = pk->pending[0];
        b = (uint64_t)p->w[0]->len + p->w[1]->len;
        r = (uint64_t)p->w[0]->n_reads + p->w[1]->n_reads;
        e = p->w[0]->n_bases + p->w[1]->n_bases;
        if (nw == GROUP_MAX || nw == rv->max_windows || nb + b > rv->max_rec_bytes || nr + r > rv->max_reads ||
            ne + e > max_entries || (uint64_t)(nw + 1) * SSN_WINDOW > max_columns) {
            if (nw == 0) {
                set_err("window %d:%u holds %llu reads / %llu record bytes / %llu read bases; the read batch capacity is %u / %llu / %llu",
                        p->w[0]->tid, p->w[0]->widx << SSN_WINDOW_SHIFT, (unsigned long long)r, (unsigned long long)b, (unsigned long long)e,
                        rv->max_reads, (unsigned long long)rv->max_rec_bytes, (unsigned long long)max_entries);
                return SSN_ERR_OVERFLOW;
            }
            break;
        }
        {
            int s;
            ssn_read_window *W = &rv->win[nw];
            W->lo = p->w[0]->widx << SSN_WINDOW_SHIFT; W->tid = p->w[0]->tid;
            for (s = 0; s < 2; ++s) {
                reads_task *t = &tasks[nt++];
                W->rec_first[s] = nr; W->n_reads[s] = p->w[s]->n_reads;
                t->p = p; t->s = s; t->dst = rv->recs + nb; t->off = rv->rec_off + nr; t->off_base = (uint32_t)(nb >> 2);
                nr += p->w[s]->n_reads; nb += p->w[s]->len;
            }
            tasks[nt].p = p; tasks[nt].s = 2; tasks[nt].nt16 = rv->ref_nt16 + (size_t)nw * SSN_WINDOW;
            tasks[nt].chars = ref_chars ? ref_chars + (size_t)nw * SSN_WINDOW : 0;
            ++nt;
        }
        took[nw++] = p; ne += e;
        pk->n_pending = 0;
    }
    t1 = now_s(); pk->st.t_wait += t1 - t0;
    if (nw == 0) return SSN_OK;
    ssn_pool_for(pk->pool, nt, reads_copy_task, tasks);
    for (i = 0; i < (int)nw; ++i) pair_free(took[i]);
    pk->st.t_fill += now_s() - t1;
    pk->st.batches += 1;
    *n_windows = nw; *n_reads = nr; *rec_bytes = nb;
    return SSN_OK;
}

/* ---- whole read windows (the input of the device pileup) through the very same count/fill code ---------------- */

ssn_status ssn_pack_read_windows(const ssn_reads_view *in, uint32_t n_windows, const ssn_batch_view *view,
                                 uint32_t *pos, uint32_t *n_columns, uint64_t *n_entries)
{
    static const char nt16_char[] = "=ACMGRSVTWYHKDBN";
    ssn_packer *pk;
    ssn_colmap map;
    char *ref;
    uint32_t ncol = 0, w;
    uint64_t nent = 0;
    ssn_status rc = SSN_OK;
    if (!in || !view || !n_columns || !n_entries) { set_err("ssn_pack_read_windows: null argument"); return SSN_ERR_ARG; }
    pk = (ssn_packer*)calloc(1, sizeof(*pk));
    ref = (char*)malloc(SSN_WINDOW);
    map.tid = (int32_t*)malloc(sizeof(int32_t) * view->max_columns);
    map.pos = (uint32_t*)malloc(sizeof(uint32_t) * view->max_columns);
    map.ref_char = (uint8_t*)malloc(view->max_columns);
    pk->view = view; pk->map = &map;
    for (w = 0; w < n_windows && rc == SSN_OK; ++w) {
        const ssn_read_window *W = &in->win[w];
        ssn_window win[2];
        pair p;
        uint32_t i, c = 0;
        uint64_t e = 0;
        int s;
        memset(&p, 0, sizeof(p));
        for (s = 0; s < 2; ++s) {
            uint32_t k, words = 0;
            memset(&win[s], 0, sizeof(win[s]));
            win[s].tid = W->tid; win[s].widx = W->lo >> SSN_WINDOW_SHIFT; win[s].n_reads = W->n_reads[s];
            win[s].buf = in->recs + 4 * (size_t)(W->n_reads[s] ? in->rec_off[W->rec_first[s]] : 0);
            for (k = 0; k < W->n_reads[s]; ++k) {
                const ssn_read *h = (const ssn_read*)(win[s].buf + 4 * (size_t)words);
                if (in->rec_off[W->rec_first[s] + k] != in->rec_off[W->rec_first[s]] + words) { set_err("ssn_pack_read_windows: records of window %u are not contiguous", w); rc = SSN_ERR_ARG; break; }
                words += h->bytes >> 2;
            }
            win[s].len = 4 * (size_t)words;
            p.w[s] = &win[s];
        }
        if (rc != SSN_OK) break;
        for (i = 0; i < SSN_WINDOW; ++i) ref[i] = nt16_char[in->ref_nt16[(size_t)w * SSN_WINDOW + i] & 15];
        p.ref = ref - W->lo; p.ref_len = (int)(W->lo + SSN_WINDOW);
        {
            pair *pp = &p;
            count_task(&pp, 0);
        }
        for (i = 0; i < SSN_WINDOW; ++i) if (p.cnt[0][i]) { c += 1; e += (uint64_t)p.cnt[0][i] + p.cnt[1][i]; }
        if (ncol + c > view->max_columns || nent + e > view->max_entries) { set_err("ssn_pack_read_windows: batch capacity exceeded"); rc = SSN_ERR_OVERFLOW; }
        else {
            segment *sg = &pk->seg[0];
            sg->p = &p; sg->i0 = 0; sg->i1 = SSN_WINDOW; sg->col_base = ncol; sg->ent_base = nent;
            fill_task(pk, 0);
            if (pos) for (i = 0; i < c; ++i) pos[ncol + i] = (w << SSN_WINDOW_SHIFT) | (map.pos[ncol + i] - W->lo);
            ncol += c; nent += e;
        }
        free(p.cnt[0]); free(p.cnt[1]);
    }
    if (rc == SSN_OK) { view->off[2 * (size_t)ncol] = (uint32_t)nent; *n_columns = ncol; *n_entries = nent; }
    free(map.tid); free(map.pos); free(map.ref_char); free(ref); free(pk);
    return rc;
}
