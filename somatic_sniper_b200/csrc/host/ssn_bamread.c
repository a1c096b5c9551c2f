#include "ssn_bamread.h"

#include <pthread.h>
#include <stdio.h>
#include <sys/resource.h>
#include <sys/syscall.h>
#include <unistd.h>
#include <stdlib.h>
#include <string.h>

#include "ssn_bamseek.h"
#include "ssn_bgzf.h"

#define QUEUE_MAX_BYTES ((size_t)256 << 20)   /* finished windows waiting for the packer (back-pressure) */

struct ssn_bamreader {
    ssn_bgzf *z;
    ssn_pool *pool;
    ssn_bam_header hdr;
    int flag_mask, mapq_thresh;
    /* byte stream over inflated chunks */
    const uint8_t *chunk;
    size_t chunk_len, chunk_pos;
    uint8_t *carry;
    size_t carry_cap;
    int stream_end;
    /* reader thread + queue of finished windows */
    pthread_t th;
    int th_started;
    pthread_mutex_t mu;
    pthread_cond_t not_empty, not_full;
    ssn_window *q_head, *q_tail;
    size_t q_bytes;
    int finished, stop;
    uint64_t n_records;
    /* open windows of the current target, ascending widx */
    ssn_window **open;
    int n_open, cap_open;
    size_t hint_bytes; uint32_t hint_reads;      /* capacity of the last finished window (+12 %) */
    /* region shard (ssn_bamreader_open_range): the reader owns the windows in [lo, hi) */
    const ssn_bamfile *bf;          /* not owned; NULL = the whole file from its first record */
    int hdr_borrowed;
    ssn_bam_range rg;
    uint64_t lo_key, hi_key;        /* (tid << 32 | first position) of the first owned / first foreign window */
    int long_span, failed;
    size_t queue_max;
    void (*on_finish)(void *); void *on_finish_arg;
};

/* ---- byte stream -------------------------------------------------------------------------------- */

static int stream_fill(ssn_bamreader *r)
{
    if (r->stream_end) return 0;
    if (!ssn_bgzf_next(r->z, &r->chunk, &r->chunk_len)) { r->stream_end = 1; r->chunk_len = r->chunk_pos = 0; return 0; }
    r->chunk_pos = 0;
    return 1;
}

/* Returns a pointer to the next n bytes (inside the chunk, or assembled in the carry buffer), NULL at end of stream. */
static const uint8_t *stream_take(ssn_bamreader *r, size_t n)
{
    size_t have;
    if (r->chunk_len - r->chunk_pos >= n) {
        const uint8_t *p = r->chunk + r->chunk_pos;
        r->chunk_pos += n;
        return p;
    }
    if (r->carry_cap < n) {
        r->carry_cap = n + (n >> 1) + 64;
        r->carry = (uint8_t*)realloc(r->carry, r->carry_cap);
        if (!r->carry) return 0;
    }
    have = 0;
    while (have < n) {
        size_t avail = r->chunk_len - r->chunk_pos, take;
        if (avail == 0) {
            if (!stream_fill(r)) return 0;          /* truncated: the reference treats any short read as end-of-file */
            continue;
        }
        take = avail < n - have ? avail : n - have;
        memcpy(r->carry + have, r->chunk + r->chunk_pos, take);
        r->chunk_pos += take; have += take;
    }
    return r->carry;
}

static uint32_t le32(const uint8_t *p) { return p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; }

static int read_header(ssn_bamreader *r, char *err, size_t errlen)
{
    const uint8_t *p = stream_take(r, 4);
    int32_t i;
    if (!p || memcmp(p, "BAM\001", 4) != 0) { snprintf(err, errlen, "wrong BAM header"); return -1; }
    if (!(p = stream_take(r, 4))) { snprintf(err, errlen, "truncated BAM header"); return -1; }
    r->hdr.l_text = (int32_t)le32(p);
    if (r->hdr.l_text < 0) { snprintf(err, errlen, "bad BAM header"); return -1; }
    r->hdr.text = (char*)calloc((size_t)r->hdr.l_text + 1, 1);
    if (r->hdr.l_text) {
        if (!(p = stream_take(r, (size_t)r->hdr.l_text))) { snprintf(err, errlen, "truncated BAM header"); return -1; }
        memcpy(r->hdr.text, p, (size_t)r->hdr.l_text);
    }
    if (!(p = stream_take(r, 4))) { snprintf(err, errlen, "truncated BAM header"); return -1; }
    r->hdr.n_targets = (int32_t)le32(p);
    if (r->hdr.n_targets < 0) { snprintf(err, errlen, "bad BAM header"); return -1; }
    r->hdr.target_name = (char**)calloc((size_t)r->hdr.n_targets + 1, sizeof(char*));
    r->hdr.target_len = (uint32_t*)calloc((size_t)r->hdr.n_targets + 1, 4);
    for (i = 0; i < r->hdr.n_targets; ++i) {
        int32_t nl;
        if (!(p = stream_take(r, 4))) { snprintf(err, errlen, "truncated BAM header"); return -1; }
        nl = (int32_t)le32(p);
        if (nl < 0) { snprintf(err, errlen, "bad BAM header"); return -1; }
        r->hdr.target_name[i] = (char*)calloc((size_t)nl + 1, 1);
        if (nl) {
            if (!(p = stream_take(r, (size_t)nl))) { snprintf(err, errlen, "truncated BAM header"); return -1; }
            memcpy(r->hdr.target_name[i], p, (size_t)nl);
        }
        if (!(p = stream_take(r, 4))) { snprintf(err, errlen, "truncated BAM header"); return -1; }
        r->hdr.target_len[i] = le32(p);
    }
    return 0;
}

/* ---- windows --------------------------------------------------------------------------------------- */

void ssn_window_free(ssn_window *w)
{
    if (!w) return;
    free(w->buf);
    free(w->rec_word);
    free(w);
}

static void queue_push(ssn_bamreader *r, ssn_window *w)
{
    pthread_mutex_lock(&r->mu);
    while (!r->stop && r->q_bytes > r->queue_max && r->q_head) pthread_cond_wait(&r->not_full, &r->mu);
    w->next = 0;
    r->hint_bytes = w->len + (w->len >> 3) + 4096; r->hint_reads = w->n_reads + (w->n_reads >> 3) + 64;
    if (r->q_tail) r->q_tail->next = w; else r->q_head = w;
    r->q_tail = w;
    r->q_bytes += w->len;
    pthread_cond_signal(&r->not_empty);
    pthread_mutex_unlock(&r->mu);
}

/* hands over every open window that ends at or before position `upto` of target `tid` (all of them when tid differs) */
static void flush_open(ssn_bamreader *r, int32_t tid, uint32_t first_live_widx, int all)
{
    int i, k = 0;
    for (i = 0; i < r->n_open; ++i) {
        ssn_window *w = r->open[i];
        if (all || w->tid != tid || w->widx < first_live_widx) queue_push(r, w);
        else r->open[k++] = w;
    }
    r->n_open = k;
}

static ssn_window *open_window(ssn_bamreader *r, int32_t tid, uint32_t widx)
{
    int i;
    ssn_window *w;
    for (i = 0; i < r->n_open; ++i)
        if (r->open[i]->widx == widx) return r->open[i];
    if (r->n_open == r->cap_open) {
        const int cap = r->cap_open ? 2 * r->cap_open : 64;
        ssn_window **o = (ssn_window**)realloc(r->open, sizeof(*o) * (size_t)cap);
        if (!o) return 0;
        r->open = o; r->cap_open = cap;
    }
    w = (ssn_window*)calloc(1, sizeof(*w));
    if (!w) return 0;
    w->tid = tid; w->widx = widx;
    if (r->hint_bytes) {
        /* neighbouring windows are about as deep: one allocation instead of a doubling chain (every large realloc is a
         * system call, and with many readers those contend) */
        w->buf = (uint8_t*)malloc(r->hint_bytes);
        w->rec_word = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)r->hint_reads);
        if (w->buf && w->rec_word) { w->cap = r->hint_bytes; w->cap_reads = r->hint_reads; }
        else { free(w->buf); free(w->rec_word); w->buf = 0; w->rec_word = 0; }
    }
    /* keep ascending widx order (reads arrive sorted by start, so new windows normally go last) */
    i = r->n_open++;
    while (i > 0 && r->open[i - 1]->widx > widx) { r->open[i] = r->open[i - 1]; --i; }
    r->open[i] = w;
    return w;
}

static int window_append(ssn_window *w, const ssn_read *h, const uint8_t *cigar, const uint8_t *seq, const uint8_t *qual)
{
    const size_t nc = (size_t)h->n_cigar * 4, ns = ((size_t)h->l_qseq + 1) >> 1, nq = h->l_qseq;
    uint8_t *d;
    if (w->n_reads == w->cap_reads) {
        const uint32_t cap = w->cap_reads ? 2 * w->cap_reads : 1024;
        uint32_t *no = (uint32_t*)realloc(w->rec_word, sizeof(uint32_t) * (size_t)cap);
        if (!no) return -1;
        w->rec_word = no; w->cap_reads = cap;
    }
    w->rec_word[w->n_reads] = (uint32_t)(w->len >> 2);
    if (w->len + h->bytes > w->cap) {
        size_t cap = w->cap ? w->cap * 2 : ((size_t)1 << 16);
        uint8_t *nb;
        while (cap < w->len + h->bytes) cap *= 2;
        nb = (uint8_t*)realloc(w->buf, cap);
        if (!nb) return -1;
        w->buf = nb;
        w->cap = cap;
    }
    d = w->buf + w->len;
    memcpy(d, h, sizeof(*h)); d += sizeof(*h);
    memcpy(d, cigar, nc); d += nc;
    memcpy(d, seq, ns); d += ns;
    memcpy(d, qual, nq);
    w->len += h->bytes;
    w->n_reads += 1;
    {
        /* upper bound of the read-bases this record can add to the window: its M runs, which a malformed record may
         * stretch past the end of the read (such positions become N bases of quality 0) */
        uint64_t m = 0;
        uint32_t k, c;
        for (k = 0; k < h->n_cigar; ++k) { memcpy(&c, cigar + 4 * (size_t)k, 4); if ((c & 0xf) == 0) m += c >> 4; }
        w->n_bases += m > h->l_qseq ? m : h->l_qseq;
    }
    return 0;
}

/* ---- reader thread ----------------------------------------------------------------------------------- */

static void reader_finish(ssn_bamreader *r)
{
    pthread_mutex_lock(&r->mu);
    r->finished = 1;
    pthread_cond_broadcast(&r->not_empty);
    pthread_mutex_unlock(&r->mu);
    if (r->on_finish) r->on_finish(r->on_finish_arg);
}

/* Region shard: where to start reading and the keep rule's state there.  The reader starts `lookback` positions in
 * front of its first window (the reads that begin there may reach into it) at the first record found by bisection, and
 * takes over the cursor of the reference's walk - the start of the last passing read in front of that record
 * (sniper_pileup.c:186,215-216) - from a short backward scan. */
static int range_start(ssn_bamreader *r, int32_t *max_tid, uint32_t *prev_beg)
{
    const ssn_bam_range *g = &r->rg;
    ssn_bam_point pt;
    int rc;
    if (g->lo_tid == 0 && g->lo_pos == 0) {                  /* the first shard reads from the top */
        pt.voff = ssn_bamfile_first_record(r->bf);
    } else {
        const uint32_t from = g->lo_pos > g->lookback ? g->lo_pos - g->lookback : 0u;
        rc = ssn_bamfile_lower_bound(r->bf, g->lo_tid, (int32_t)from, &pt);
        if (rc < 0) return -1;
        if (rc == 0) return 0;                                /* nothing at or behind the shard's start */
        if (ssn_bam_key(pt.tid, pt.pos) >= r->hi_key) return 0;   /* the first such record already belongs to a later shard */
        {
            int32_t pt_tid = -1, pt_pos = 0;
            rc = ssn_bamfile_prev_passing(r->bf, pt.voff, r->flag_mask, r->mapq_thresh, &pt_tid, &pt_pos);
            if (rc < 0) return -1;
            if (rc == 1) { *max_tid = pt_tid; *prev_beg = (uint32_t)pt_pos; }
        }
    }
    r->z = ssn_bgzf_open_at(ssn_bamfile_path(r->bf), r->pool, pt.voff >> 16);
    if (!r->z) return -1;
    if (g->inflate_service) ssn_bgzf_use_service(r->z, g->inflate_service);   /* no lane to be had: stays on the host */
    if ((pt.voff & 0xffff) && !stream_take(r, (size_t)(pt.voff & 0xffff))) return 0;
    return 1;
}

static void *reader_main(void *v)
{
    ssn_bamreader *r = (ssn_bamreader*)v;
    int32_t max_tid = -1;
    uint32_t prev_beg = 0;          /* cursor position at the next fetch (0 before the first passing read) */
    const int ranged = r->bf != 0;
    if (ranged) {
        /* Region shards run more reader threads than there are cores.  They are the bulk of the work but not the critical
         * path: the one thread per GPU that moves finished windows to the device (and its copy workers) must never wait
         * for a time slice behind them, so the readers give way (per-thread nice value, Linux). */
#ifdef SYS_gettid
        setpriority(PRIO_PROCESS, (id_t)syscall(SYS_gettid), 10);
#endif
        const int st = range_start(r, &max_tid, &prev_beg);
        if (st < 0) r->failed = 1;
        if (st <= 0) { reader_finish(r); return 0; }
    }
    for (;;) {
        const uint8_t *p;
        int32_t block_len, tid, pos;
        uint32_t x2, x3, l_qseq, mapq, l_qname, flag, n_cigar, k, end;
        const uint8_t *cigar, *seq, *qual;
        ssn_read h;
        uint32_t w0, w1, wi;
        if (r->stop) break;
        if (!(p = stream_take(r, 4))) break;
        block_len = (int32_t)le32(p);
        if (block_len < 32) break;
        if (!(p = stream_take(r, (size_t)block_len))) break;
        tid = (int32_t)le32(p); pos = (int32_t)le32(p + 4);
        if (ranged && ssn_bam_key(tid, pos) >= r->hi_key) break;      /* sorted input: everything from here on is a later shard's */
        __atomic_add_fetch(&r->n_records, 1, __ATOMIC_RELAXED);
        x2 = le32(p + 8); x3 = le32(p + 12); l_qseq = le32(p + 16);
        mapq = (x2 >> 8) & 0xff; l_qname = x2 & 0xff; flag = x3 >> 16; n_cigar = x3 & 0xffff;
        if ((flag & (uint32_t)r->flag_mask) || (int)mapq < r->mapq_thresh) continue;       /* sniper_pileup.c:208 */
        if ((size_t)32 + l_qname + (size_t)n_cigar * 4 + ((size_t)l_qseq + 1) / 2 + l_qseq > (size_t)block_len) break;
        cigar = p + 32 + l_qname;
        seq = cigar + (size_t)n_cigar * 4;
        qual = seq + ((size_t)l_qseq + 1) / 2;
        end = (uint32_t)pos;                                                                 /* bam_calend, bam.c:37-47 */
        for (k = 0; k < n_cigar; ++k) {
            const uint32_t c = le32(cigar + 4 * k), op = c & 0xf;
            if (op == 0 || op == 2 || op == 3) end += c >> 4;
        }
        if (tid < max_tid) {                                                                 /* sniper_pileup.c:211-214 */
            fprintf(stderr, "[bam_pileup_core] the input is not sorted. Abort!\n");
            abort();
        }
        if (tid == max_tid && (uint32_t)pos < prev_beg) {
            /* positions running backwards inside a target: the reference does not notice (its test at :211 is
             * shadowed by tid >= max_tid) and piles such reads from the cursor onwards only; this reader needs
             * sorted starts, so it stops here like on any other unsorted input (known deviations) */
            fprintf(stderr, "[bam_pileup_core] the input is not sorted. Abort!\n");
            abort();
        }
        {
            const int keep = end > prev_beg;                                                 /* sniper_pileup.c:216 */
            max_tid = tid; prev_beg = (uint32_t)pos;
            if (!keep || tid < 0 || pos < 0 || end <= (uint32_t)pos) continue;
        }
        /* a read that begins in front of the next shard's look-back and reaches into that shard is not seen there */
        if (ranged && tid == r->rg.hi_tid && end > r->rg.hi_pos && (uint64_t)pos + r->rg.lookback < r->rg.hi_pos) r->long_span = 1;
        h.beg = pos; h.end = (int32_t)end; h.l_qseq = l_qseq; h.n_cigar = (uint16_t)n_cigar;
        h.mapq = (uint8_t)mapq; h.strand = (flag & 16) ? 1 : 0;
        h.bytes = (uint32_t)((sizeof(h) + (size_t)n_cigar * 4 + ((size_t)l_qseq + 1) / 2 + l_qseq + 3) & ~(size_t)3);
        w0 = (uint32_t)pos >> SSN_WINDOW_SHIFT; w1 = (end - 1) >> SSN_WINDOW_SHIFT;
        flush_open(r, tid, w0, 0);
        for (wi = w0; wi <= w1; ++wi) {
            ssn_window *w;
            if (ranged) {
                const uint64_t wk = (uint64_t)(uint32_t)tid << 32 | (uint64_t)wi << SSN_WINDOW_SHIFT;
                if (wk < r->lo_key) continue;
                if (wk >= r->hi_key) break;
            }
            w = open_window(r, tid, wi);
            if (!w || window_append(w, &h, cigar, seq, qual) != 0) {
                fprintf(stderr, "[ssn_bamread] out of memory\n");
                r->failed = 1; r->stop = 1;
                break;
            }
        }
    }
    flush_open(r, 0, 0, 1);
    reader_finish(r);
    return 0;
}

/* ---- API ------------------------------------------------------------------------------------------------ */

static ssn_bamreader *reader_new(int flag_mask, int mapq_thresh, ssn_pool *pool, char *err, size_t errlen)
{
    ssn_bamreader *r = (ssn_bamreader*)calloc(1, sizeof(*r));
    if (!r) { snprintf(err, errlen, "out of memory"); return 0; }
    r->pool = pool;
    r->flag_mask = flag_mask < 0 ? SSN_BAM_DEF_MASK : (4 | flag_mask);       /* bam_plbuf_set_mask, sniper_pileup.c:146-150 */
    r->mapq_thresh = mapq_thresh < 0 ? 0 : mapq_thresh;                      /* bam_plbuf_set_mapq_thresh, :139-143 */
    r->queue_max = QUEUE_MAX_BYTES;
    pthread_mutex_init(&r->mu, 0);
    pthread_cond_init(&r->not_empty, 0);
    pthread_cond_init(&r->not_full, 0);
    return r;
}

ssn_bamreader *ssn_bamreader_open(const char *path, int flag_mask, int mapq_thresh, ssn_pool *pool, char *err, size_t errlen)
{
    ssn_bamreader *r = reader_new(flag_mask, mapq_thresh, pool, err, errlen);
    if (!r) return 0;
    r->z = ssn_bgzf_open(path, pool);
    if (!r->z) { snprintf(err, errlen, "cannot open %s", path); ssn_bamreader_close(r); return 0; }
    if (read_header(r, err, errlen) != 0) { ssn_bamreader_close(r); return 0; }
    if (pthread_create(&r->th, 0, reader_main, r) != 0) { snprintf(err, errlen, "cannot start the reader thread"); ssn_bamreader_close(r); return 0; }
    r->th_started = 1;
    return r;
}

ssn_bamreader *ssn_bamreader_open_range(const ssn_bamfile *bf, const ssn_bam_range *range, int flag_mask, int mapq_thresh,
                                        ssn_pool *pool, void (*on_finish)(void*), void *on_finish_arg, char *err, size_t errlen)
{
    ssn_bamreader *r = reader_new(flag_mask, mapq_thresh, pool, err, errlen);
    if (!r) return 0;
    r->bf = bf; r->rg = *range;
    r->hdr = *ssn_bamfile_header(bf); r->hdr_borrowed = 1;
    r->lo_key = (uint64_t)(uint32_t)range->lo_tid << 32 | range->lo_pos;
    r->hi_key = range->hi_tid == INT32_MAX ? UINT64_MAX : ((uint64_t)(uint32_t)range->hi_tid << 32 | range->hi_pos);
    if (range->queue_max_bytes) r->queue_max = range->queue_max_bytes;
    r->on_finish = on_finish; r->on_finish_arg = on_finish_arg;
    if (pthread_create(&r->th, 0, reader_main, r) != 0) { snprintf(err, errlen, "cannot start the reader thread"); r->on_finish = 0; ssn_bamreader_close(r); return 0; }
    r->th_started = 1;
    return r;
}

const ssn_bam_header *ssn_bamreader_header(const ssn_bamreader *r) { return &r->hdr; }
uint64_t ssn_bamreader_records(const ssn_bamreader *r) { return __atomic_load_n(&r->n_records, __ATOMIC_RELAXED); }
int ssn_bamreader_long_span(const ssn_bamreader *r) { return r->long_span; }
int ssn_bamreader_failed(const ssn_bamreader *r) { return r->failed; }

ssn_window *ssn_bamreader_pop(ssn_bamreader *r)
{
    ssn_window *w;
    pthread_mutex_lock(&r->mu);
    while (!r->q_head && !r->finished) pthread_cond_wait(&r->not_empty, &r->mu);
    w = r->q_head;
    if (w) {
        r->q_head = w->next;
        if (!r->q_head) r->q_tail = 0;
        r->q_bytes -= w->len;
        w->next = 0;
        pthread_cond_signal(&r->not_full);
    }
    pthread_mutex_unlock(&r->mu);
    return w;
}

void ssn_bamreader_close(ssn_bamreader *r)
{
    int i;
    if (!r) return;
    if (r->th_started) {
        pthread_mutex_lock(&r->mu);
        r->stop = 1;
        pthread_cond_broadcast(&r->not_full);
        pthread_mutex_unlock(&r->mu);
        pthread_join(r->th, 0);
    }
    while (r->q_head) { ssn_window *w = r->q_head; r->q_head = w->next; ssn_window_free(w); }
    for (i = 0; i < r->n_open; ++i) ssn_window_free(r->open[i]);
    free(r->open);
    if (!r->hdr_borrowed) {
        if (r->hdr.target_name) for (i = 0; i < r->hdr.n_targets; ++i) free(r->hdr.target_name[i]);
        free(r->hdr.target_name); free(r->hdr.target_len); free(r->hdr.text);
    }
    free(r->carry);
    ssn_bgzf_close(r->z);
    pthread_mutex_destroy(&r->mu);
    pthread_cond_destroy(&r->not_empty);
    pthread_cond_destroy(&r->not_full);
    free(r);
}
