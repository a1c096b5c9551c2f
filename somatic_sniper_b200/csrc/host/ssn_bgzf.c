#include "ssn_bgzf.h"
#include "ssn_host.h"
#include "ssn_inflate.h"

#include <errno.h>
#include <fcntl.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>
#include <zlib.h>

#define RAW_CAP (8u << 20)      /* compressed bytes held at once */
#define MAX_BLOCKS 512          /* blocks inflated per call */

typedef struct {
    const uint8_t *cdata;
    uint32_t clen, ulen, total;
    size_t uoff;
    int bad;
} blk;

struct ssn_bgzf {
    int fd, own_fd, eof_file, done;
    ssn_pool *pool;
    uint8_t *raw;
    size_t raw_len, raw_pos;
    uint8_t *out;
    size_t out_cap;
    blk blocks[MAX_BLOCKS];
    uint64_t consumed;
    /* per call: up to 512 blocks with a worker pool behind it (parallel inflate); 16 when the caller inflates by itself
     * (region shards: many readers, one core each — the inflated bytes are framed while they are still in the cache) */
    int max_blocks;
    size_t raw_cap;
    /* inflate service (GPU): a lane (pinned in / out) is taken per batch; the inflated bytes are copied into `out` */
    const ssn_inflate_service *svc;
    uint64_t lane_in_cap;
    int out_direct;             /* `out` came from the service (host_alloc): run_to delivers into it */
    ssn_inflate_block *lane_blk;
    uint8_t *lane_refused;
    int use_zlib;               /* SSN_ZLIB_INFLATE: every block through zlib (measurement / fall-back knob) */
};

ssn_bgzf *ssn_bgzf_open(const char *path, ssn_pool *pool) { return ssn_bgzf_open_at(path, pool, 0); }

ssn_bgzf *ssn_bgzf_open_at(const char *path, ssn_pool *pool, uint64_t coff)
{
    ssn_bgzf *z = (ssn_bgzf*)calloc(1, sizeof(*z));
    if (!z) return 0;
    if (strcmp(path, "-") == 0) {
        if (coff) { free(z); return 0; }
        z->fd = 0; z->own_fd = 0;
    } else {
        z->fd = open(path, O_RDONLY);
        z->own_fd = 1;
        if (z->fd < 0) { free(z); return 0; }
        if (coff && lseek(z->fd, (off_t)coff, SEEK_SET) < 0) { close(z->fd); free(z); return 0; }
#ifdef POSIX_FADV_SEQUENTIAL
        posix_fadvise(z->fd, 0, 0, POSIX_FADV_SEQUENTIAL);
#endif
    }
    z->pool = pool;
    z->use_zlib = getenv("SSN_ZLIB_INFLATE") != 0;
    z->max_blocks = (pool && ssn_pool_threads(pool) > 0) ? MAX_BLOCKS : 16;
    z->raw_cap = (pool && ssn_pool_threads(pool) > 0) ? RAW_CAP : ((size_t)2 << 20);
    z->raw = (uint8_t*)calloc(1, z->raw_cap + SSN_INFLATE_SLACK);
    z->out_cap = (size_t)z->max_blocks * 65536;
    z->out = (uint8_t*)malloc(z->out_cap);
    if (!z->raw || !z->out) { ssn_bgzf_close(z); return 0; }
    return z;
}

int ssn_bgzf_use_service(ssn_bgzf *z, const void *svc_)
{
    const ssn_inflate_service *svc = (const ssn_inflate_service*)svc_;
    uint8_t *raw, *out, *li, *lo;
    uint64_t in_cap = 0, out_cap = 0;
    uint32_t mb = 0;
    int32_t lane;
    int direct = 0;
    if (!z || !svc || z->svc) return 0;
    /* a lane is held only while a batch is on its way (ssn_bgzf_next); all lanes have the same size: ask once */
    lane = svc->lane_acquire(svc->ctx, &li, &in_cap, &lo, &out_cap, &mb);
    if (lane < 0) return 0;
    svc->lane_release(svc->ctx, lane);
    if (mb > MAX_BLOCKS) mb = MAX_BLOCKS;
    if (mb > out_cap / 65536u) mb = (uint32_t)(out_cap / 65536u);                  /* a block inflates to at most 64 KiB */
    if (in_cap > RAW_CAP) in_cap = RAW_CAP;
    if (mb == 0 || in_cap < 65600u) return 0;
    z->lane_blk = (ssn_inflate_block*)malloc(sizeof(ssn_inflate_block) * mb);
    z->lane_refused = (uint8_t*)malloc(mb);
    /* the compressed side of a whole batch has to be resident: grow the read buffer (what is buffered so far moves along) */
    raw = (uint8_t*)calloc(1, RAW_CAP + SSN_INFLATE_SLACK);
    if (svc->host_alloc && svc->host_free && svc->run_to) { out = (uint8_t*)svc->host_alloc(svc->ctx, (uint64_t)mb * 65536); direct = out != 0; }
    if (!direct) out = (uint8_t*)malloc((size_t)mb * 65536);
    if (!z->lane_blk || !z->lane_refused || !raw || !out) {
        free(z->lane_blk); free(z->lane_refused); free(raw); z->lane_blk = 0; z->lane_refused = 0;
        if (direct) svc->host_free(svc->ctx, out); else free(out);
        return 0;
    }
    memcpy(raw, z->raw + z->raw_pos, z->raw_len - z->raw_pos);
    z->raw_len -= z->raw_pos; z->raw_pos = 0;
    free(z->raw); free(z->out);
    z->raw = raw; z->raw_cap = RAW_CAP;
    z->out = out; z->out_cap = (size_t)mb * 65536;
    z->max_blocks = (int)mb;
    z->lane_in_cap = in_cap;
    z->out_direct = direct;
    z->svc = svc;
    return 1;
}

void ssn_bgzf_close(ssn_bgzf *z)
{
    if (!z) return;
    free(z->lane_blk); free(z->lane_refused);
    if (z->own_fd && z->fd >= 0) close(z->fd);
    if (z->out_direct) z->svc->host_free(z->svc->ctx, z->out); else free(z->out);
    free(z->raw); free(z);
}

uint64_t ssn_bgzf_compressed_bytes(const ssn_bgzf *z) { return z->consumed; }

static void refill(ssn_bgzf *z)
{
    if (z->raw_pos > 0) {
        memmove(z->raw, z->raw + z->raw_pos, z->raw_len - z->raw_pos);
        z->raw_len -= z->raw_pos;
        z->raw_pos = 0;
    }
    while (!z->eof_file && z->raw_len < z->raw_cap) {
        ssize_t r = read(z->fd, z->raw + z->raw_len, z->raw_cap - z->raw_len);
        if (r < 0) { if (errno == EINTR) continue; z->eof_file = 1; break; }
        if (r == 0) { z->eof_file = 1; break; }
        z->raw_len += (size_t)r;
    }
}

/* parses one block header at p (avail bytes); 0 = need more bytes, <0 = invalid, else total block length */
static int parse_block(const uint8_t *p, size_t avail, blk *b)
{
    uint32_t xlen, total = 0, x;
    if (avail < 18) return 0;
    if (p[0] != 31 || p[1] != 139 || p[2] != 8 || !(p[3] & 4)) return -1;
    xlen = p[10] | (uint32_t)p[11] << 8;
    if (avail < 12 + xlen) return 0;
    for (x = 0; x + 4 <= xlen;) {
        const uint8_t *s = p + 12 + x;
        const uint32_t slen = s[2] | (uint32_t)s[3] << 8;
        if (s[0] == 'B' && s[1] == 'C' && slen == 2 && x + 6 <= xlen) total = (s[4] | (uint32_t)s[5] << 8) + 1;
        x += 4 + slen;
    }
    if (total == 0 || total < 12 + xlen + 8) return -1;
    if (avail < total) return 0;
    b->cdata = p + 12 + xlen;
    b->clen = total - 12 - xlen - 8;
    b->ulen = p[total - 4] | (uint32_t)p[total - 3] << 8 | (uint32_t)p[total - 2] << 16 | (uint32_t)p[total - 1] << 24;
    b->bad = 0;
    b->total = total;
    return (int)total;
}

static __thread z_stream t_zs;
static __thread int t_zs_ready = 0;

typedef struct { ssn_bgzf *z; } inflate_arg;

static void inflate_one(void *arg, int i)
{
    ssn_bgzf *z = ((inflate_arg*)arg)->z;
    blk *b = &z->blocks[i];
    int st;
    if (!z->use_zlib && ssn_inflate_raw(b->cdata, b->clen, z->out + b->uoff, b->ulen) == 0) return;
    /* zlib reads every stream the reference reads; the fast decoder gave up on this one (or is switched off) */
    if (!t_zs_ready) {
        memset(&t_zs, 0, sizeof(t_zs));
        if (inflateInit2(&t_zs, -15) != Z_OK) { b->bad = 1; return; }
        t_zs_ready = 1;
    } else inflateReset(&t_zs);
    t_zs.next_in = (Bytef*)b->cdata; t_zs.avail_in = b->clen;
    t_zs.next_out = z->out + b->uoff; t_zs.avail_out = b->ulen;
    st = inflate(&t_zs, Z_FINISH);
    if (st != Z_STREAM_END || t_zs.avail_out != 0) b->bad = 1;
}

int ssn_bgzf_next(ssn_bgzf *z, const uint8_t **data, size_t *len)
{
    int n = 0, i, stop = 0;
    size_t uoff = 0;
    uint64_t csum = 0;
    inflate_arg a;
    *data = 0; *len = 0;
    if (z->done) return 0;
    if (z->raw_len - z->raw_pos < (z->raw_cap >> 1)) refill(z);
    while (n < z->max_blocks) {
        blk *b = &z->blocks[n];
        int r = parse_block(z->raw + z->raw_pos, z->raw_len - z->raw_pos, b);
        if (r == 0) {
            if (n > 0) break;
            if (z->eof_file) {
                if (z->raw_len != z->raw_pos) fprintf(stderr, "[ssn_bgzf] truncated block at the end of the file\n");
                stop = 1; break;
            }
            refill(z);
            if (z->eof_file && z->raw_len - z->raw_pos < 18) { stop = 1; if (z->raw_len != z->raw_pos) fprintf(stderr, "[ssn_bgzf] truncated block at the end of the file\n"); break; }
            continue;
        }
        if (r < 0) { fprintf(stderr, "[ssn_bgzf] invalid block header\n"); stop = 1; break; }
        if (z->svc && n > 0 && csum + b->clen > z->lane_in_cap) break;        /* the lane's input buffer is full: next call */
        csum += b->clen;
        z->raw_pos += (size_t)r;
        z->consumed += (uint64_t)r;
        if (b->ulen == 0) { stop = 1; break; }          /* empty block: the reference's reader stops here */
        if (b->ulen > 65536) { fprintf(stderr, "[ssn_bgzf] block larger than 64 KiB\n"); stop = 1; break; }
        b->uoff = uoff; uoff += b->ulen;
        ++n;
    }
    if (stop) z->done = 1;
    if (n == 0) return 0;
    a.z = z;
    if (z->svc) {
        /* A lane for this batch only (so a reader held up by its queue keeps none): payloads back to back into the lane's
         * pinned input, one run, the inflated bytes into this reader's own buffer, lane back.  What the service refused
         * is inflated here, the way the reference reads it. */
        const ssn_inflate_service *svc = z->svc;
        uint8_t *li = 0, *lo = 0;
        uint64_t ic = 0, oc = 0, in_bytes = 0;
        uint32_t mb = 0;
        int k;
        const int32_t lane = svc->lane_acquire(svc->ctx, &li, &ic, &lo, &oc, &mb);
        memset(z->lane_refused, 1, (size_t)(unsigned)n);
        if (lane >= 0) {
            if (csum <= ic && uoff <= oc && (uint32_t)n <= mb) {
                for (k = 0; k < n; ++k) {
                    memcpy(li + in_bytes, z->blocks[k].cdata, z->blocks[k].clen);
                    z->lane_blk[k].in_off = (uint32_t)in_bytes; z->lane_blk[k].in_len = z->blocks[k].clen;
                    z->lane_blk[k].out_off = (uint32_t)z->blocks[k].uoff; z->lane_blk[k].out_len = z->blocks[k].ulen;
                    in_bytes += z->blocks[k].clen;
                }
                int ok;
                if (z->out_direct) ok = svc->run_to(svc->ctx, lane, z->lane_blk, (uint32_t)n, in_bytes, uoff, z->lane_refused, z->out) == 0;
                else if ((ok = svc->run(svc->ctx, lane, z->lane_blk, (uint32_t)n, in_bytes, uoff, z->lane_refused) == 0)) memcpy(z->out, lo, uoff);
                if (!ok) memset(z->lane_refused, 1, (size_t)(unsigned)n);
            }
            svc->lane_release(svc->ctx, lane);
        }
        for (k = 0; k < n; ++k) if (z->lane_refused[k]) inflate_one(&a, k);
    } else
    ssn_pool_for(z->pool, n, inflate_one, &a);
    for (i = 0; i < n; ++i) {
        if (z->blocks[i].bad) {
            fprintf(stderr, "[ssn_bgzf] inflate failed\n");
            z->done = 1;
            uoff = z->blocks[i].uoff;                    /* keep what came before the bad block */
            break;
        }
    }
    if (uoff == 0) return 0;
    *data = z->out; *len = uoff;
    return 1;
}
