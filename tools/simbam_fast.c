/*
 * simbam_fast.c — multi-threaded generator of synthetic tumor/normal BAM pairs for the BAM-level benchmarks and the
 * larger parity tests (test/benchmark infrastructure, not part of the product).  Same ingredients as
 * somatic_sniper_b200/simbam.py, written in C because a 3-Mb-and-up pair is out of reach for a Python loop:
 *   reference  i.i.d. ACGT per contig (seeded), FASTA with 60 columns
 *   variants   germline het every ~1 kb, hom every ~10 kb, somatic every ~5 kb at tumor VAF {0.1, 0.25, 0.4},
 *              LOH (normal het, tumor hom) and gain-of-reference (normal hom alt, tumor het) every ~50 kb
 *   reads      150 bp, coordinate sorted; 0.25 % carry a 1-5 bp insertion, 0.25 % a deletion, 0.2 % a 300 bp N skip,
 *              2 % a soft clip, 0.3 % hard clips; 1 % duplicates (0x400), 0.25 % QC-fail, 0.25 % secondary;
 *              mapQ {60: 85 %, 40-59: 8 %, 1-39: 5 %, 0: 2 %}; base quality {2, 12, 23, 37} at {2, 5, 13, 80} %;
 *              substitution errors at 10^(-Q/10); N calls (Q2) at 0.05 %
 * Only the CIGAR operators samtools-0.1.6 knows (M I D N S H).  Every (contig, 64 kb chunk, sample) draws from its own
 * seeded generator, so the output does not depend on the thread count.
 *
 *   ssn-simbam <outdir> <seed> <depth_tumor> <depth_normal> <level> <threads> name:len [name:len ...]
 * writes <outdir>/ref.fa, tumor.bam, normal.bam and prints one JSON line (reads per file).  With SSN_SIMBAM_SPLIT=1 in the
 * environment it also writes tumor.<i>.bam / normal.<i>.bam holding contig i's reads alone under the full header: the
 * inputs of the region-split CPU baseline (the reference has no region option).
 */
#include <math.h>
#include <pthread.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <zlib.h>

#define READ_LEN 150
#define CHUNK 65536

typedef struct { uint64_t s; } rng_t;
static inline uint64_t rnd(rng_t *r)
{
    uint64_t z = (r->s += 0x9e3779b97f4a7c15ull);
    z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ull;
    z = (z ^ (z >> 27)) * 0x94d049bb133111ebull;
    return z ^ (z >> 31);
}
static inline double rndu(rng_t *r) { return (double)(rnd(r) >> 11) * (1.0 / 9007199254740992.0); }
static inline uint32_t rndn(rng_t *r, uint32_t n) { return (uint32_t)(((rnd(r) >> 32) * (uint64_t)n) >> 32); }

typedef struct { uint32_t pos; uint8_t alt; float vaf[2]; } variant;
typedef struct {
    char name[64];
    uint32_t len;
    uint8_t *seq;                 /* ASCII */
    variant *var; uint32_t nvar;
} contig;

static int n_contigs;
static contig *contigs;
static uint64_t g_seed;
static int g_level;

static const uint8_t nt16_of[256] = { ['A'] = 1, ['C'] = 2, ['G'] = 4, ['T'] = 8, ['N'] = 15 };

/* ---- growable byte buffer -------------------------------------------------------------------------------------- */
typedef struct { uint8_t *p; size_t n, cap; } buf_t;
static void buf_put(buf_t *b, const void *src, size_t n)
{
    if (b->n + n > b->cap) {
        b->cap = b->cap ? b->cap * 2 : (1 << 20);
        while (b->cap < b->n + n) b->cap *= 2;
        b->p = (uint8_t*)realloc(b->p, b->cap);
    }
    memcpy(b->p + b->n, src, n);
    b->n += n;
}
static void put32(uint8_t *p, uint32_t v) { p[0] = (uint8_t)v; p[1] = (uint8_t)(v >> 8); p[2] = (uint8_t)(v >> 16); p[3] = (uint8_t)(v >> 24); }

/* BGZF: raw bytes -> blocks of at most 0xff00 input bytes */
static void bgzf_blocks(const uint8_t *src, size_t n, buf_t *out, int level)
{
    static const uint8_t head[16] = {31, 139, 8, 4, 0, 0, 0, 0, 0, 255, 6, 0, 'B', 'C', 2, 0};
    size_t at = 0;
    uint8_t *tmp = (uint8_t*)malloc(0x10000 + 1024);
    while (at < n || (n == 0 && at == 0)) {
        const size_t take = n - at < 0xff00 ? n - at : 0xff00;
        z_stream zs;
        uint8_t hdr[18], tail[8];
        uint32_t clen;
        memset(&zs, 0, sizeof(zs));
        deflateInit2(&zs, level, Z_DEFLATED, -15, 8, Z_DEFAULT_STRATEGY);
        zs.next_in = (Bytef*)(src + at); zs.avail_in = (uInt)take;
        zs.next_out = tmp; zs.avail_out = 0x10000 + 1024;
        deflate(&zs, Z_FINISH);
        clen = (uint32_t)zs.total_out;
        deflateEnd(&zs);
        memcpy(hdr, head, 16);
        hdr[16] = (uint8_t)((clen + 25) & 0xff); hdr[17] = (uint8_t)((clen + 25) >> 8);
        put32(tail, (uint32_t)crc32(crc32(0, 0, 0), src + at, (uInt)take)); put32(tail + 4, (uint32_t)take);
        buf_put(out, hdr, 18); buf_put(out, tmp, clen); buf_put(out, tail, 8);
        at += take;
        if (n == 0) break;
    }
    free(tmp);
}

/* ---- one chunk of one sample -------------------------------------------------------------------------------------- */

typedef struct {
    int tid, sample;              /* sample 0 tumor, 1 normal */
    uint32_t lo, hi;
    double depth;
    buf_t out;                    /* BGZF blocks */
    uint64_t n_reads;
    int done;
} task;

static int reg2bin(uint32_t beg, uint32_t end)
{
    --end;
    if (beg >> 14 == end >> 14) return 4681 + (int)(beg >> 14);
    if (beg >> 17 == end >> 17) return 585 + (int)(beg >> 17);
    if (beg >> 20 == end >> 20) return 73 + (int)(beg >> 20);
    if (beg >> 23 == end >> 23) return 9 + (int)(beg >> 23);
    if (beg >> 26 == end >> 26) return 1 + (int)(beg >> 26);
    return 0;
}

static int cmp_u32(const void *a, const void *b) { const uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b; return x < y ? -1 : x > y; }

static void make_chunk(task *t)
{
    const contig *c = &contigs[t->tid];
    rng_t r;
    buf_t raw = {0, 0, 0};
    uint32_t n, i, *starts;
    static const uint8_t qbin[4] = {2, 12, 23, 37};
    static const double perr[4] = {0.63095734448, 0.0630957344, 0.00501187234, 0.000199526231};
    r.s = g_seed * 0x2545f4914f6cdd1dull + ((uint64_t)t->tid << 40) + ((uint64_t)t->sample << 36) + t->lo;
    rnd(&r);
    {
        const double mean = t->depth * (double)(t->hi - t->lo) / READ_LEN;
        n = (uint32_t)mean;
        if (rndu(&r) < mean - n) ++n;
    }
    starts = (uint32_t*)malloc(sizeof(uint32_t) * (n + 1));
    for (i = 0; i < n; ++i) starts[i] = t->lo + rndn(&r, t->hi - t->lo);
    qsort(starts, n, sizeof(uint32_t), cmp_u32);
    for (i = 0; i < n; ++i) {
        const uint32_t p = starts[i];
        uint8_t bases[READ_LEN + 8], qual[READ_LEN + 8], rec[36 + 2 + 16 + (READ_LEN + 8) / 2 + 1 + READ_LEN + 8];
        uint32_t cigar[4], ncig = 1, ref_span = READ_LEN, k, flag, mq, lq = READ_LEN;
        double u;
        int kind = 0, il = 0;
        if (p + READ_LEN + 320 >= c->len) { if (p + READ_LEN > c->len) continue; }
        else {
            u = rndu(&r);
            if (u < 0.0025) { kind = 1; il = 1 + (int)rndn(&r, 5); }
            else if (u < 0.005) { kind = 2; il = 1 + (int)rndn(&r, 5); }
            else if (u < 0.007) kind = 3;
            else if (u < 0.017) kind = 4;
            else if (u < 0.027) kind = 5;
            else if (u < 0.03) kind = 6;
        }
        /* reference-consistent bases along the alignment, then variants, then errors */
        {
            uint32_t x = p, y = 0;
            if (kind == 1) { cigar[0] = 70u << 4; cigar[1] = ((uint32_t)il << 4) | 1; cigar[2] = (uint32_t)(READ_LEN - 70 - il) << 4; ncig = 3; ref_span = READ_LEN - (uint32_t)il; }
            else if (kind == 2) { cigar[0] = 80u << 4; cigar[1] = ((uint32_t)il << 4) | 2; cigar[2] = (uint32_t)(READ_LEN - 80) << 4; ncig = 3; ref_span = READ_LEN + (uint32_t)il; }
            else if (kind == 3) { cigar[0] = 60u << 4; cigar[1] = (300u << 4) | 3; cigar[2] = (uint32_t)(READ_LEN - 60) << 4; ncig = 3; ref_span = READ_LEN + 300; }
            else if (kind == 4) { cigar[0] = (10u << 4) | 4; cigar[1] = (uint32_t)(READ_LEN - 10) << 4; ncig = 2; ref_span = READ_LEN - 10; }
            else if (kind == 5) { cigar[0] = (uint32_t)(READ_LEN - 12) << 4; cigar[1] = (12u << 4) | 4; ncig = 2; ref_span = READ_LEN - 12; }
            else if (kind == 6) { cigar[0] = (5u << 4) | 5; cigar[1] = (uint32_t)READ_LEN << 4; cigar[2] = (3u << 4) | 5; ncig = 3; }
            else cigar[0] = (uint32_t)READ_LEN << 4;
            for (k = 0; k < ncig; ++k) {
                const uint32_t op = cigar[k] & 0xf, l = cigar[k] >> 4;
                uint32_t j;
                if (op == 0) { for (j = 0; j < l; ++j) bases[y + j] = c->seq[x + j]; x += l; y += l; }
                else if (op == 1 || op == 4) { for (j = 0; j < l; ++j) bases[y + j] = "ACGT"[rndn(&r, 4)]; y += l; }
                else if (op == 2 || op == 3) x += l;
            }
            lq = y;
        }
        /* variants by allele fraction (per read and site, like simbam.py) */
        if (c->nvar) {
            uint32_t lo = 0, hi = c->nvar;
            while (lo < hi) { const uint32_t m = (lo + hi) / 2; if (c->var[m].pos < p) lo = m + 1; else hi = m; }
            for (; lo < c->nvar && c->var[lo].pos < p + ref_span; ++lo) {
                /* map the reference position to the query through the CIGAR */
                uint32_t x = p, y = 0, q = 0xffffffffu;
                for (k = 0; k < ncig; ++k) {
                    const uint32_t op = cigar[k] & 0xf, l = cigar[k] >> 4;
                    if (op == 0) { if (c->var[lo].pos >= x && c->var[lo].pos < x + l) { q = y + (c->var[lo].pos - x); break; } x += l; y += l; }
                    else if (op == 1 || op == 4) y += l;
                    else if (op == 2 || op == 3) x += l;
                }
                if (q != 0xffffffffu && rndu(&r) < c->var[lo].vaf[t->sample]) bases[q] = c->var[lo].alt;
            }
        }
        for (k = 0; k < lq; ++k) {
            const double uq = rndu(&r);
            const int qi = uq < 0.02 ? 0 : uq < 0.07 ? 1 : uq < 0.20 ? 2 : 3;
            qual[k] = qbin[qi];
            if (rndu(&r) < perr[qi]) bases[k] = "ACGT"[rndn(&r, 4)];
            if (rndu(&r) < 5e-4) { bases[k] = 'N'; qual[k] = 2; }
        }
        u = rndu(&r);
        mq = u < 0.85 ? 60 : u < 0.93 ? 40 + rndn(&r, 20) : u < 0.98 ? 1 + rndn(&r, 39) : 0;
        flag = (rnd(&r) & 1) ? 16u : 0u;
        flag |= 1u | 2u | ((rnd(&r) & 1) ? 64u : 128u);
        u = rndu(&r);
        if (u < 0.01) flag |= 0x400; else if (u < 0.0125) flag |= 0x200; else if (u < 0.015) flag |= 0x100;
        {
            const uint32_t l_name = 2, body = 32 + l_name + 4 * ncig + (lq + 1) / 2 + lq;
            uint8_t *w = rec;
            put32(w, body); w += 4;
            put32(w, (uint32_t)t->tid); put32(w + 4, p);
            put32(w + 8, (uint32_t)reg2bin(p, p + (ref_span ? ref_span : 1)) << 16 | mq << 8 | l_name);
            put32(w + 12, flag << 16 | ncig); put32(w + 16, lq);
            put32(w + 20, (uint32_t)t->tid); put32(w + 24, p); put32(w + 28, 0);
            w += 32;
            *w++ = 'r'; *w++ = 0;
            for (k = 0; k < ncig; ++k) { put32(w, cigar[k]); w += 4; }
            for (k = 0; k < lq; k += 2) *w++ = (uint8_t)(nt16_of[bases[k]] << 4 | (k + 1 < lq ? nt16_of[bases[k + 1]] : 0));
            memcpy(w, qual, lq); w += lq;
            buf_put(&raw, rec, (size_t)(w - rec));
        }
        t->n_reads += 1;
    }
    if (raw.n) bgzf_blocks(raw.p, raw.n, &t->out, g_level);
    free(raw.p); free(starts);
}

/* ---- work queue with ordered output ----------------------------------------------------------------------------- */

static task *tasks;
static long n_tasks, next_task;
static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;
static pthread_cond_t cv = PTHREAD_COND_INITIALIZER;
static long written;              /* tasks written so far (back-pressure: workers stay at most 256 tasks ahead) */

static void *worker(void *arg)
{
    (void)arg;
    for (;;) {
        long i;
        pthread_mutex_lock(&mu);
        while (next_task < n_tasks && next_task - written > 256) pthread_cond_wait(&cv, &mu);
        i = next_task < n_tasks ? next_task++ : -1;
        pthread_mutex_unlock(&mu);
        if (i < 0) break;
        make_chunk(&tasks[i]);
        pthread_mutex_lock(&mu);
        tasks[i].done = 1;
        pthread_cond_broadcast(&cv);
        pthread_mutex_unlock(&mu);
    }
    return 0;
}

static void make_variants(contig *c, int tid)
{
    rng_t r;
    const uint32_t every[5] = {1000, 10000, 5000, 50000, 50000};
    uint32_t cap = c->len / 700 + 16, n = 0, k, i;
    variant *v = (variant*)malloc(sizeof(variant) * cap);
    r.s = g_seed * 31 + 977 * (uint64_t)(tid + 1);
    for (k = 0; k < 5; ++k) {
        const uint32_t cnt = c->len >= every[k] / 4 ? (c->len / every[k] ? c->len / every[k] : 1) : 0;
        for (i = 0; i < cnt && n < cap; ++i) {
            static const float som[3] = {0.1f, 0.25f, 0.4f};
            v[n].pos = rndn(&r, c->len);
            v[n].alt = 0;
            switch (k) {
            case 0: v[n].vaf[0] = 0.5f; v[n].vaf[1] = 0.5f; break;
            case 1: v[n].vaf[0] = 1.0f; v[n].vaf[1] = 1.0f; break;
            case 2: v[n].vaf[0] = som[rndn(&r, 3)]; v[n].vaf[1] = 0.0f; break;
            case 3: v[n].vaf[0] = 1.0f; v[n].vaf[1] = 0.5f; break;
            default: v[n].vaf[0] = 0.5f; v[n].vaf[1] = 1.0f; break;
            }
            ++n;
        }
    }
    /* sort by position, drop duplicates, pick the alternative allele */
    for (i = 1; i < n; ++i) { variant x = v[i]; uint32_t j = i; while (j > 0 && v[j - 1].pos > x.pos) { v[j] = v[j - 1]; --j; } v[j] = x; }
    for (i = 0, k = 0; i < n; ++i) if (k == 0 || v[k - 1].pos != v[i].pos) v[k++] = v[i];
    n = k;
    for (i = 0; i < n; ++i) {
        const char *acgt = "ACGT";
        const char *q = strchr(acgt, c->seq[v[i].pos]);
        v[i].alt = (uint8_t)acgt[((q ? (int)(q - acgt) : 0) + 1 + (int)rndn(&r, 3)) & 3];
    }
    c->var = v; c->nvar = n;
}

int main(int argc, char **argv)
{
    char path[4096];
    double depth[2];
    int threads, s, i;
    FILE *fa;
    uint64_t reads[2] = {0, 0};
    if (argc < 8) { fprintf(stderr, "usage: ssn-simbam <outdir> <seed> <depth_tumor> <depth_normal> <level> <threads> name:len ...\n"); return 1; }
    g_seed = strtoull(argv[2], 0, 10); depth[0] = atof(argv[3]); depth[1] = atof(argv[4]);
    g_level = atoi(argv[5]); threads = atoi(argv[6]);
    if (threads < 1) threads = 1;
    n_contigs = argc - 7;
    contigs = (contig*)calloc((size_t)n_contigs, sizeof(contig));
    snprintf(path, sizeof(path), "%s/ref.fa", argv[1]);
    fa = fopen(path, "wb");
    if (!fa) { fprintf(stderr, "cannot write %s\n", path); return 1; }
    for (i = 0; i < n_contigs; ++i) {
        contig *c = &contigs[i];
        char *colon = strrchr(argv[7 + i], ':');
        rng_t r;
        uint32_t k;
        if (!colon) { fprintf(stderr, "bad contig '%s'\n", argv[7 + i]); return 1; }
        snprintf(c->name, sizeof(c->name), "%.*s", (int)(colon - argv[7 + i]), argv[7 + i]);
        c->len = (uint32_t)strtoul(colon + 1, 0, 10);
        c->seq = (uint8_t*)malloc((size_t)c->len + 1);
        r.s = g_seed + 1000003ull * (uint64_t)(i + 1);
        for (k = 0; k < c->len; ++k) c->seq[k] = (uint8_t)"ACGT"[rnd(&r) >> 62];
        fprintf(fa, ">%s\n", c->name);
        for (k = 0; k < c->len; k += 60) { fwrite(c->seq + k, 1, c->len - k < 60 ? c->len - k : 60, fa); fputc('\n', fa); }
        make_variants(c, i);
    }
    fclose(fa);
    snprintf(path, sizeof(path), "%s/ref.fa.fai", argv[1]);
    remove(path);
    for (s = 0; s < 2; ++s) {
        FILE *out;
        buf_t hdr = {0, 0, 0}, blk = {0, 0, 0};
        pthread_t *th = (pthread_t*)calloc((size_t)threads, sizeof(pthread_t));
        long k, nt = 0;
        uint8_t w[8];
        const int split = getenv("SSN_SIMBAM_SPLIT") != 0;
        int split_tid = -1;
        FILE *part = 0;
        buf_t hdr_blk = {0, 0, 0}, eof_blk = {0, 0, 0};
        bgzf_blocks((const uint8_t*)"", 0, &eof_blk, g_level);
        snprintf(path, sizeof(path), "%s/%s.bam", argv[1], s == 0 ? "tumor" : "normal");
        out = fopen(path, "wb");
        if (!out) { fprintf(stderr, "cannot write %s\n", path); return 1; }
        {   /* header */
            buf_t text = {0, 0, 0};
            char line[256];
            buf_put(&text, "@HD\tVN:1.0\tSO:coordinate\n", 25);
            for (i = 0; i < n_contigs; ++i) { const int l = snprintf(line, sizeof(line), "@SQ\tSN:%s\tLN:%u\n", contigs[i].name, contigs[i].len); buf_put(&text, line, (size_t)l); }
            buf_put(&hdr, "BAM\001", 4);
            put32(w, (uint32_t)text.n); buf_put(&hdr, w, 4); buf_put(&hdr, text.p, text.n);
            put32(w, (uint32_t)n_contigs); buf_put(&hdr, w, 4);
            for (i = 0; i < n_contigs; ++i) {
                const uint32_t l = (uint32_t)strlen(contigs[i].name) + 1;
                put32(w, l); buf_put(&hdr, w, 4); buf_put(&hdr, contigs[i].name, l);
                put32(w, contigs[i].len); buf_put(&hdr, w, 4);
            }
            bgzf_blocks(hdr.p, hdr.n, &blk, g_level);
            fwrite(blk.p, 1, blk.n, out);
            buf_put(&hdr_blk, blk.p, blk.n);
            free(text.p); free(hdr.p); free(blk.p);
        }
        for (i = 0; i < n_contigs; ++i) nt += (contigs[i].len + CHUNK - 1) / CHUNK;
        tasks = (task*)calloc((size_t)nt, sizeof(task));
        n_tasks = 0;
        for (i = 0; i < n_contigs; ++i) {
            uint32_t lo;
            if (contigs[i].len < 2 * READ_LEN + 200) continue;
            for (lo = 0; lo < contigs[i].len; lo += CHUNK) {
                task *t = &tasks[n_tasks++];
                t->tid = i; t->sample = s; t->lo = lo; t->hi = lo + CHUNK < contigs[i].len ? lo + CHUNK : contigs[i].len;
                t->depth = depth[s];
            }
        }
        next_task = 0; written = 0;
        for (i = 0; i < threads; ++i) pthread_create(&th[i], 0, worker, 0);
        for (k = 0; k < n_tasks; ++k) {
            pthread_mutex_lock(&mu);
            while (!tasks[k].done) pthread_cond_wait(&cv, &mu);
            pthread_mutex_unlock(&mu);
            if (tasks[k].out.n) fwrite(tasks[k].out.p, 1, tasks[k].out.n, out);
            if (split) {
                if (tasks[k].tid != split_tid) {
                    if (part) { fwrite(eof_blk.p, 1, eof_blk.n, part); fclose(part); }
                    split_tid = tasks[k].tid;
                    snprintf(path, sizeof(path), "%s/%s.%d.bam", argv[1], s == 0 ? "tumor" : "normal", split_tid);
                    part = fopen(path, "wb");
                    if (part) fwrite(hdr_blk.p, 1, hdr_blk.n, part);
                }
                if (part && tasks[k].out.n) fwrite(tasks[k].out.p, 1, tasks[k].out.n, part);
            }
            free(tasks[k].out.p); tasks[k].out.p = 0;
            reads[s] += tasks[k].n_reads;
            pthread_mutex_lock(&mu);
            written = k + 1;
            pthread_cond_broadcast(&cv);
            pthread_mutex_unlock(&mu);
        }
        for (i = 0; i < threads; ++i) pthread_join(th[i], 0);
        fwrite(eof_blk.p, 1, eof_blk.n, out);            /* end-of-file marker: an empty block */
        if (part) { fwrite(eof_blk.p, 1, eof_blk.n, part); fclose(part); }
        free(hdr_blk.p); free(eof_blk.p);
        fclose(out);
        free(tasks); free(th);
    }
    printf("{\"reads_tumor\": %llu, \"reads_normal\": %llu}\n", (unsigned long long)reads[0], (unsigned long long)reads[1]);
    return 0;
}
