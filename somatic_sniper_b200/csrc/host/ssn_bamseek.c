#include "ssn_bamseek.h"
#include "ssn_inflate.h"

#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/stat.h>
#include <unistd.h>
#include <zlib.h>

#define LOOK_BLOCKS 4                 /* inflated blocks a probe validates its guess against */
#define CHAIN_MAX 64                  /* records of the chain a probe checks at most */

struct ssn_bamfile {
    int fd;
    char *path;
    uint64_t size;
    ssn_bam_header hdr;
    uint64_t first_voff;
};

static uint32_t le32(const uint8_t *p) { return p[0] | (uint32_t)p[1] << 8 | (uint32_t)p[2] << 16 | (uint32_t)p[3] << 24; }

static ssize_t read_at(int fd, uint64_t off, uint8_t *buf, size_t n)
{
    size_t got = 0;
    while (got < n) {
        ssize_t r = pread(fd, buf + got, n - got, (off_t)(off + got));
        if (r < 0) { if (errno == EINTR) continue; return -1; }
        if (r == 0) break;
        got += (size_t)r;
    }
    return (ssize_t)got;
}

/* gzip member with the BGZF extra field at p: total length of the block, 0 when the bytes are not one.
 * *data_off = offset of the deflate stream, avail may be as short as 18 bytes. */
static uint32_t block_total(const uint8_t *p, size_t avail, uint32_t *data_off)
{
    uint32_t xlen, x, total = 0;
    if (avail < 18 || p[0] != 31 || p[1] != 139 || p[2] != 8 || !(p[3] & 4)) return 0;
    xlen = p[10] | (uint32_t)p[11] << 8;
    if (xlen < 6 || avail < 12 + xlen) return 0;
    for (x = 0; x + 4 <= xlen;) {
        const uint8_t *s = p + 12 + x;
        const uint32_t slen = s[2] | (uint32_t)s[3] << 8;
        if (s[0] == 'B' && s[1] == 'C' && slen == 2 && x + 6 <= xlen) total = (s[4] | (uint32_t)s[5] << 8) + 1;
        x += 4 + slen;
    }
    if (total < 12 + xlen + 8) return 0;
    if (data_off) *data_off = 12 + xlen;
    return total;
}

static __thread z_stream t_zs;
static __thread int t_zs_ready = 0;

static int inflate_raw(const uint8_t *src, uint32_t clen, uint8_t *dst, uint32_t ulen)
{
    if (ssn_inflate_raw(src, clen, dst, ulen) == 0) return 0;      /* every caller's buffer extends past the block's footer */
    if (!t_zs_ready) {
        memset(&t_zs, 0, sizeof(t_zs));
        if (inflateInit2(&t_zs, -15) != Z_OK) return -1;
        t_zs_ready = 1;
    } else inflateReset(&t_zs);
    t_zs.next_in = (Bytef*)src; t_zs.avail_in = clen;
    t_zs.next_out = dst; t_zs.avail_out = ulen;
    if (inflate(&t_zs, Z_FINISH) != Z_STREAM_END || t_zs.avail_out != 0) return -1;
    return 0;
}

/* ---- sequential cursor over the inflated stream, block by block --------------------------------------------- */

typedef struct {
    const ssn_bamfile *f;
    uint64_t coff, next_coff;      /* current block and the one behind it */
    uint32_t ulen, upos;
    int end;                       /* no more data (end of file, empty block, damaged block) */
    uint8_t raw[65536 + 64];
    uint8_t u[65536];
} cursor;

static int cursor_load(cursor *c, uint64_t coff)
{
    uint8_t head[64];
    uint32_t doff = 0, total, ulen;
    ssize_t r;
    c->coff = coff; c->ulen = c->upos = 0;
    if (coff >= c->f->size) { c->end = 1; return 0; }
    r = read_at(c->f->fd, coff, head, sizeof(head));
    if (r < 18 || !(total = block_total(head, (size_t)r, &doff)) || total > sizeof(c->raw)) { c->end = 1; return 0; }
    if (read_at(c->f->fd, coff, c->raw, total) != (ssize_t)total) { c->end = 1; return 0; }
    ulen = le32(c->raw + total - 4);
    if (ulen == 0 || ulen > 65536) { c->end = 1; return 0; }       /* an empty block ends the stream (bgzf_read) */
    if (inflate_raw(c->raw + doff, total - doff - 8, c->u, ulen) != 0) { c->end = 1; return 0; }
    c->ulen = ulen; c->next_coff = coff + total;
    return 1;
}

static void cursor_seek(cursor *c, const ssn_bamfile *f, uint64_t voff)
{
    c->f = f; c->end = 0;
    if (cursor_load(c, voff >> 16)) {
        c->upos = (uint32_t)(voff & 0xffff);
        if (c->upos > c->ulen) c->end = 1;
    }
}

/* position of the next byte; a position at the very end of a block is reported as the start of the next one */
static uint64_t cursor_tell(cursor *c)
{
    while (!c->end && c->upos == c->ulen) cursor_load(c, c->next_coff);
    return c->coff << 16 | c->upos;
}

/* copies (dst != NULL) or skips n bytes; 0 when the stream ends first */
static int cursor_take(cursor *c, uint8_t *dst, size_t n)
{
    while (n) {
        size_t k;
        if (c->end) return 0;
        if (c->upos == c->ulen) { if (!cursor_load(c, c->next_coff)) return 0; continue; }
        k = c->ulen - c->upos;
        if (k > n) k = n;
        if (dst) { memcpy(dst, c->u + c->upos, k); dst += k; }
        c->upos += (uint32_t)k; n -= k;
    }
    return 1;
}

typedef struct { uint64_t voff; int32_t tid, pos; uint32_t flag, mapq; } rec_info;

/* 1: a record (the cursor then stands behind it); 0: end of data */
static int cursor_record(cursor *c, rec_info *ri)
{
    uint8_t b[36];
    int32_t bs;
    ri->voff = cursor_tell(c);
    if (c->end) return 0;
    if (!cursor_take(c, b, 36)) return 0;
    bs = (int32_t)le32(b);
    if (bs < 32) return 0;
    ri->tid = (int32_t)le32(b + 4); ri->pos = (int32_t)le32(b + 8);
    ri->mapq = (le32(b + 12) >> 8) & 0xff; ri->flag = le32(b + 16) >> 16;
    if (!cursor_take(c, 0, (size_t)bs - 32)) return 0;          /* a truncated last record does not count (bam_read1) */
    return 1;
}

/* ---- open: the header ----------------------------------------------------------------------------------------- */

static void header_free(ssn_bam_header *h)
{
    int i;
    if (h->target_name) for (i = 0; i < h->n_targets; ++i) free(h->target_name[i]);
    free(h->target_name); free(h->target_len); free(h->text);
    memset(h, 0, sizeof(*h));
}

ssn_bamfile *ssn_bamfile_open(const char *path, char *err, size_t errlen)
{
    ssn_bamfile *f = (ssn_bamfile*)calloc(1, sizeof(*f));
    struct stat st;
    cursor *c = 0;
    uint8_t b[8];
    int32_t i;
    if (!f) { snprintf(err, errlen, "out of memory"); return 0; }
    f->fd = open(path, O_RDONLY);
    if (f->fd < 0 || fstat(f->fd, &st) != 0 || !S_ISREG(st.st_mode)) {
        snprintf(err, errlen, "cannot open %s for random access", path);
        if (f->fd >= 0) close(f->fd);
        free(f); return 0;
    }
    f->size = (uint64_t)st.st_size;
    f->path = strdup(path);
    c = (cursor*)malloc(sizeof(*c));
    cursor_seek(c, f, 0);
    if (!cursor_take(c, b, 8) || memcmp(b, "BAM\001", 4) != 0) { snprintf(err, errlen, "wrong BAM header"); goto fail; }
    f->hdr.l_text = (int32_t)le32(b + 4);
    if (f->hdr.l_text < 0) { snprintf(err, errlen, "bad BAM header"); goto fail; }
    f->hdr.text = (char*)calloc((size_t)f->hdr.l_text + 1, 1);
    if (f->hdr.l_text && !cursor_take(c, (uint8_t*)f->hdr.text, (size_t)f->hdr.l_text)) { snprintf(err, errlen, "truncated BAM header"); goto fail; }
    if (!cursor_take(c, b, 4)) { snprintf(err, errlen, "truncated BAM header"); goto fail; }
    f->hdr.n_targets = (int32_t)le32(b);
    if (f->hdr.n_targets < 0) { snprintf(err, errlen, "bad BAM header"); goto fail; }
    f->hdr.target_name = (char**)calloc((size_t)f->hdr.n_targets + 1, sizeof(char*));
    f->hdr.target_len = (uint32_t*)calloc((size_t)f->hdr.n_targets + 1, 4);
    for (i = 0; i < f->hdr.n_targets; ++i) {
        int32_t nl;
        if (!cursor_take(c, b, 4)) { snprintf(err, errlen, "truncated BAM header"); goto fail; }
        nl = (int32_t)le32(b);
        if (nl < 0) { snprintf(err, errlen, "bad BAM header"); goto fail; }
        f->hdr.target_name[i] = (char*)calloc((size_t)nl + 1, 1);
        if (nl && !cursor_take(c, (uint8_t*)f->hdr.target_name[i], (size_t)nl)) { snprintf(err, errlen, "truncated BAM header"); goto fail; }
        if (!cursor_take(c, b, 4)) { snprintf(err, errlen, "truncated BAM header"); goto fail; }
        f->hdr.target_len[i] = le32(b);
    }
    f->first_voff = cursor_tell(c);
    free(c);
    return f;
fail:
    free(c);
    ssn_bamfile_close(f);
    return 0;
}

void ssn_bamfile_close(ssn_bamfile *f)
{
    if (!f) return;
    if (f->fd >= 0) close(f->fd);
    header_free(&f->hdr);
    free(f->path);
    free(f);
}

const ssn_bam_header *ssn_bamfile_header(const ssn_bamfile *f) { return &f->hdr; }
const char *ssn_bamfile_path(const ssn_bamfile *f) { return f->path; }
uint64_t ssn_bamfile_first_record(const ssn_bamfile *f) { return f->first_voff; }
uint64_t ssn_bamfile_size(const ssn_bamfile *f) { return f->size; }

/* ---- probe ------------------------------------------------------------------------------------------------------ */

/* One plausible record at offset u of the inflated look-ahead U[0, L)?  Returns the offset behind it, 0 when the bytes
 * cannot be a record, 1 when the record's fixed part runs past L (cannot be judged). */
static size_t plausible(const ssn_bam_header *h, const uint8_t *U, size_t L, size_t u, uint64_t *key)
{
    int32_t bs, tid, pos, lseq, mtid, mpos;
    uint32_t x2, x3, lq, nc, k;
    uint64_t need;
    if (u + 36 > L) return 1;
    bs = (int32_t)le32(U + u);
    if (bs < 34 || bs > (1 << 27)) return 0;
    tid = (int32_t)le32(U + u + 4); pos = (int32_t)le32(U + u + 8);
    if (tid < -1 || tid >= h->n_targets || pos < -1) return 0;
    if (tid >= 0 && h->target_len[tid] && (uint32_t)pos > h->target_len[tid]) return 0;
    x2 = le32(U + u + 12); x3 = le32(U + u + 16);
    lq = x2 & 0xff; nc = x3 & 0xffff;
    lseq = (int32_t)le32(U + u + 20); mtid = (int32_t)le32(U + u + 24); mpos = (int32_t)le32(U + u + 28);
    if (lq < 1 || lseq < 0 || lseq > (1 << 27) || mtid < -1 || mtid >= h->n_targets || mpos < -1) return 0;
    need = 32ull + lq + 4ull * nc + ((uint64_t)lseq + 1) / 2 + (uint64_t)lseq;
    if (need > (uint64_t)bs) return 0;
    {   /* the name: printable, NUL-terminated at its stated length */
        const size_t n0 = u + 36, nz = n0 + lq - 1;
        for (k = 0; k + 1 < lq && k < 24 && n0 + k < L; ++k)
            if (U[n0 + k] < 33 || U[n0 + k] > 126) return 0;
        if (nz < L && U[nz] != 0) return 0;
        for (k = 0; k < nc && k < 8 && nz + 1 + 4 * (size_t)k + 4 <= L; ++k)
            if ((U[nz + 1 + 4 * (size_t)k] & 0xf) > 8) return 0;
    }
    *key = ssn_bam_key(tid, pos);
    return u + 4 + (size_t)bs;
}

static int chain_ok(const ssn_bam_header *h, const uint8_t *U, size_t L, size_t u, int at_eof)
{
    uint64_t prev = 0, key = 0;
    int n = 0;
    while (n < CHAIN_MAX) {
        size_t nx;
        if (u == L) return n >= 1 && (at_eof || n >= 3);
        nx = plausible(h, U, L, u, &key);
        if (nx == 0) return 0;
        if (nx == 1) return at_eof ? 0 : n >= 3;                 /* a cut-off record at the end of the file does not exist */
        if (key < prev) return 0;
        prev = key; ++n;
        if (nx > L) return at_eof ? 0 : n >= 3;
        u = nx;
    }
    return 1;
}

int ssn_bamfile_probe(const ssn_bamfile *f, uint64_t coff, ssn_bam_point *out)
{
    const size_t RAWCAP = (size_t)(LOOK_BLOCKS + 3) * 65536 + 4096;
    uint8_t *raw = (uint8_t*)calloc(1, RAWCAP + SSN_INFLATE_SLACK), *U = (uint8_t*)malloc((size_t)LOOK_BLOCKS * 65536);
    int rc = 0;
    if (!raw || !U) { free(raw); free(U); return -1; }
    while (coff < f->size) {
        ssize_t got = read_at(f->fd, coff, raw, RAWCAP);
        size_t i, b0 = (size_t)-1;
        if (got < 18) break;
        /* a block start: a BGZF header whose successor is one too (or the end of the file) */
        for (i = 0; i + 18 <= (size_t)got && i <= 65536 + 18; ++i) {
            uint32_t t;
            if (raw[i] != 31 || raw[i + 1] != 139) continue;
            t = block_total(raw + i, (size_t)got - i, 0);
            if (!t) continue;
            if (coff + i + t == f->size || (i + t + 18 <= (size_t)got && block_total(raw + i + t, (size_t)got - i - t, 0))) { b0 = i; break; }
        }
        if (b0 == (size_t)-1) break;                              /* no block within 64 KiB: end of data */
        {
            /* inflate the look-ahead */
            size_t at = b0, L = 0, first_ulen = 0;
            uint64_t next_coff = 0;
            int nb = 0, at_eof = 0, empty_first = 0;
            while (nb < LOOK_BLOCKS) {
                uint32_t doff = 0, t, ulen;
                if (at + 18 > (size_t)got) { at_eof = coff + at >= f->size; break; }
                t = block_total(raw + at, (size_t)got - at, &doff);
                if (!t || at + t > (size_t)got) { at_eof = 1; break; }
                ulen = le32(raw + at + t - 4);
                if (ulen == 0 || ulen > 65536) { at_eof = 1; if (nb == 0) empty_first = 1; break; }     /* empty block: the reference stops here */
                if (inflate_raw(raw + at + doff, t - doff - 8, U + L, ulen) != 0) { at_eof = 1; break; }
                if (nb == 0) { first_ulen = ulen; next_coff = coff + at + t; }
                L += ulen; at += t; ++nb;
            }
            if (nb == 0) { (void)empty_first; break; }
            for (i = 0; i < first_ulen; ++i) {
                if (chain_ok(&f->hdr, U, L, i, at_eof)) {
                    out->voff = (coff + b0) << 16 | (uint64_t)i;
                    out->tid = (int32_t)le32(U + i + 4); out->pos = (int32_t)le32(U + i + 8);
                    rc = 1; goto done;
                }
            }
            coff = next_coff;                                     /* no record starts in this block */
        }
    }
done:
    free(raw); free(U);
    return rc;
}

/* ---- lower bound -------------------------------------------------------------------------------------------------- */

int ssn_bamfile_lower_bound(const ssn_bamfile *f, int32_t tid, int32_t pos, ssn_bam_point *out)
{
    const uint64_t target = ssn_bam_key(tid, pos);
    cursor *c = (cursor*)malloc(sizeof(*c));
    rec_info ri;
    ssn_bam_point lo_pt, p;
    uint64_t lo, hi;
    int rc = 0;
    if (!c) return -1;
    cursor_seek(c, f, f->first_voff);
    if (!cursor_record(c, &ri)) { free(c); return 0; }
    lo_pt.voff = ri.voff; lo_pt.tid = ri.tid; lo_pt.pos = ri.pos;
    if (ssn_bam_key(ri.tid, ri.pos) >= target) { *out = lo_pt; free(c); return 1; }
    lo = lo_pt.voff >> 16; hi = f->size;
    while (hi - lo > 2 * 65536) {
        const uint64_t mid = lo + (hi - lo) / 2;
        const int r = ssn_bamfile_probe(f, mid, &p);
        if (r < 0) { free(c); return -1; }
        if (r == 0 || ssn_bam_key(p.tid, p.pos) >= target) hi = mid;
        else {
            if ((p.voff >> 16) >= hi) break;                      /* unsorted input; the walk below still terminates */
            lo = p.voff >> 16; lo_pt = p;
        }
    }
    cursor_seek(c, f, lo_pt.voff);
    while (cursor_record(c, &ri)) {
        if (ssn_bam_key(ri.tid, ri.pos) >= target) { out->voff = ri.voff; out->tid = ri.tid; out->pos = ri.pos; rc = 1; break; }
    }
    free(c);
    return rc;
}

/* ---- the last passing record before a position ------------------------------------------------------------------------ */

int ssn_bamfile_prev_passing(const ssn_bamfile *f, uint64_t voff, int flag_mask, int mapq_thresh, int32_t *tid, int32_t *pos)
{
    cursor *c;
    uint64_t back = (uint64_t)1 << 17;
    int rc = -1;
    if (voff <= f->first_voff) return 0;
    c = (cursor*)malloc(sizeof(*c));
    if (!c) return -1;
    for (;;) {
        const uint64_t tc = voff >> 16, fc = f->first_voff >> 16;
        uint64_t start;
        rec_info ri;
        int at_start = 0, found = 0, landed = 0;
        if (tc < fc + back || tc - back <= fc) { start = f->first_voff; at_start = 1; }
        else {
            ssn_bam_point p;
            const int r = ssn_bamfile_probe(f, tc - back, &p);
            if (r < 0) break;
            if (r == 0 || p.voff >= voff) { back <<= 1; continue; }
            start = p.voff;
        }
        cursor_seek(c, f, start);
        for (;;) {
            if (!cursor_record(c, &ri)) { landed = ri.voff == voff; break; }       /* voff may be the end of the data */
            if (ri.voff >= voff) { landed = ri.voff == voff; break; }
            if (!(ri.flag & (uint32_t)flag_mask) && (int)ri.mapq >= mapq_thresh) { found = 1; *tid = ri.tid; *pos = ri.pos; }
        }
        if (!landed) {
            /* the guessed start was not a record boundary (or voff is not one): try from further back, finally from the top */
            if (at_start) break;
            back <<= 2;
            continue;
        }
        if (found) { rc = 1; break; }
        if (at_start) { rc = 0; break; }
        back <<= 2;
    }
    free(c);
    return rc;
}
