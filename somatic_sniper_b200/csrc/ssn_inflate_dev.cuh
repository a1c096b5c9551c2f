/*
 * ssn_inflate_dev.cuh — raw DEFLATE (RFC 1951) decoder for whole BGZF blocks, one WARP per block.
 *
 * Replaces, on the device, what the reference does per block on the host: inflateInit2 / inflate / inflateEnd
 * (samtools-0.1.6 bgzf.c:330-353, called from bgzf_read :489 under bam_read1, bam.c:181-212).  After the region shards made
 * reading parallel, inflating is four fifths of the host's work; the compressed blocks are also four
 * times smaller than the records they hold, so they are what should cross PCIe.
 *
 * Shape.  Huffman decoding is serial inside a block, so the parallelism is across blocks: a persistent grid, every warp
 * pulls block after block.  Inside a warp all 32 lanes run the SAME decode (bit buffer, table look-ups: shared-memory
 * broadcasts, no divergence), which makes every lane know every symbol without shuffles; the lanes differ only where
 * bytes move: literals collect in a 32-byte register window (lane p & 31 keeps byte p) that is stored with one coalesced
 * instruction, matches are copied 32 bytes per step, code tables are filled 32 entries per step.
 *
 * The same source compiles for the host (SSN_INFL_HOST: the lane loops become real loops), which is how the CPU tests
 * check the decoder against zlib without a GPU (tests/test_inflate.py through ssn_inflate_dev_host, ssn_inflate_emul.cpp,
 * in libssn.so; no compute path of the product calls it).
 *
 * Contract: in[] readable up to in_len + 16 bytes; nothing is written outside out[0, out_len).  Returns 0 when the stream
 * decodes to exactly out_len bytes, non-zero otherwise — the caller then inflates that block on the host (zlib reads
 * everything the reference reads).
 */
#pragma once
#include <stdint.h>

#ifdef SSN_INFL_HOST
#define INFL_FN static inline
#define INFL_LANES(l) for (int l = 0; l < 32; ++l)
#define INFL_SYNC() do { } while (0)
#else
#define INFL_FN __device__ __forceinline__
#define INFL_LANES(l) const int l = (int)(threadIdx.x & 31u);
#define INFL_SYNC() __syncwarp()
#endif

namespace infl {

constexpr int kLLBits = 10, kDBits = 8;
constexpr int kLLSub = 512, kDSub = 512;               /* second-level entries a block may need at most (else: refused) */
constexpr int kLLSize = (1 << kLLBits) + kLLSub, kDSize = (1 << kDBits) + kDSub;

/* entry: bits 0-3 code bits to consume, 4-7 extra bits (pointer: width of its second-level index), 8-10 kind,
 * 12-13 literals held, 15 literal flag, 16-31 value (literal(s) / length or distance base / second-level start) */
constexpr uint32_t K_LIT = 0, K_LEN = 1, K_EOB = 2, K_SUB = 3, K_BAD = 4, K_LIT2 = 5;
INFL_FN uint32_t entry(uint32_t val, uint32_t kind, uint32_t extra, uint32_t len)
{
    return val << 16 | kind << 8 | extra << 4 | len | (kind == K_LIT ? 0x9000u : kind == K_LIT2 ? 0xa000u : 0u);
}
#define E_LEN(e) ((e) & 15u)
#define E_EXTRA(e) (((e) >> 4) & 15u)
#define E_KIND(e) (((e) >> 8) & 7u)
#define E_VAL(e) ((e) >> 16)
#define E_ISLIT(e) ((e) & 0x8000u)
#define E_NLIT(e) (((e) >> 12) & 3u)

/* per-warp working memory (shared memory on the device) */
struct Work {
    uint32_t ll[kLLSize];
    uint32_t d[kDSize];
    uint8_t lens[320];
};                                          /* 9536 B: 8 warps per CTA, three CTAs per SM */

INFL_FN uint32_t len_base(uint32_t s)      /* s = symbol - 257 */
{
    const uint32_t b[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
    return b[s];
}
INFL_FN uint32_t len_extra(uint32_t s) { return (s < 8u || s == 28u) ? 0u : (s - 4u) >> 2; }
INFL_FN uint32_t dist_base(uint32_t s)
{
    const uint32_t e = s < 2u ? 0u : (s - 2u) >> 1;
    return s < 2u ? s + 1u : (((2u + (s & 1u)) << e) + 1u);
}
INFL_FN uint32_t dist_extra(uint32_t s) { return s < 2u ? 0u : (s - 2u) >> 1; }

INFL_FN uint32_t symbol_entry(bool is_dist, uint32_t sym, uint32_t len)
{
    if (is_dist) return sym < 30u ? entry(dist_base(sym), K_LEN, dist_extra(sym), len) : entry(0, K_BAD, 0, len);
    if (sym < 256u) return entry(sym, K_LIT, 0, len);
    if (sym == 256u) return entry(0, K_EOB, 0, len);
    if (sym < 286u) return entry(len_base(sym - 257u), K_LEN, len_extra(sym - 257u), len);
    return entry(0, K_BAD, 0, len);
}

INFL_FN uint32_t bitrev(uint32_t code, uint32_t len)
{
#ifdef SSN_INFL_HOST
    uint32_t r = 0;
    for (uint32_t i = 0; i < len; ++i) { r = r << 1 | (code & 1u); code >>= 1; }
    return r;
#else
    return __brev(code) >> (32u - len);
#endif
}

/* Canonical code -> two-level table.  The counting and the code assignment are the same on every lane; the table
 * entries are written 32 at a time.  Returns non-zero for an over-subscribed code or one whose second-level tables do not
 * fit.  Incomplete codes are accepted (unused entries are K_BAD and fail only if they are hit). */
/* sub_bits: 1 << tb bytes of scratch (widths of the second-level tables) that are free while this table is built */
INFL_FN int build(uint8_t *sub_bits, uint32_t *tab, uint32_t sub_cap, const uint8_t *lens, uint32_t nsym, uint32_t tb, bool is_dist)
{
    const uint32_t main_size = 1u << tb;
    uint32_t cnt[16];
    for (int l = 0; l < 16; ++l) cnt[l] = 0;
    for (uint32_t i = 0; i < nsym; ++i) cnt[lens[i]]++;
    cnt[0] = 0;
    {
        int left = 1;
        for (int l = 1; l < 16; ++l) { left = (left << 1) - (int)cnt[l]; if (left < 0) return 1; }
    }
    uint32_t code = 0, nx[16];
    bool long_codes = false;
    nx[0] = 0;
    for (int l = 1; l < 16; ++l) { code = (code + cnt[l - 1]) << 1; nx[l] = code; if ((uint32_t)l > tb && cnt[l]) long_codes = true; }
    INFL_SYNC();
    { INFL_LANES(lane) { for (uint32_t i = (uint32_t)lane; i < main_size; i += 32u) { tab[i] = entry(0, K_BAD, 0, 1); sub_bits[i] = 0; } } }
    INFL_SYNC();
    uint32_t nfree = main_size;
    if (long_codes) {
        /* widths of the second-level tables (uniform, serial: long codes are rare), then their slots */
        uint32_t n2[16];
        for (int l = 0; l < 16; ++l) n2[l] = nx[l];
        for (uint32_t i = 0; i < nsym; ++i) {
            const uint32_t l = lens[i];
            if (l > tb) {
                const uint32_t prefix = bitrev(n2[l]++, l) & (main_size - 1u);
                if (sub_bits[prefix] < l - tb) {
                    INFL_SYNC();
                    { INFL_LANES(lane) { if (lane == 0) sub_bits[prefix] = (uint8_t)(l - tb); } }
                    INFL_SYNC();
                }
            } else if (l) n2[l]++;
        }
        for (uint32_t i = 0; i < main_size; ++i) {
            const uint32_t sb = sub_bits[i];
            if (sb) {
                if (nfree - main_size + (1u << sb) > sub_cap) return 2;
                INFL_SYNC();
                { INFL_LANES(lane) {
                    if (lane == 0) tab[i] = entry(nfree, K_SUB, sb, tb);
                    for (uint32_t k = (uint32_t)lane; k < (1u << sb); k += 32u) tab[nfree + k] = entry(0, K_BAD, 0, 1);
                } }
                INFL_SYNC();
                nfree += 1u << sb;
            }
        }
    }
    for (uint32_t i = 0; i < nsym; ++i) {
        const uint32_t l = lens[i];
        if (!l) continue;
        const uint32_t rev = bitrev(nx[l]++, l);
        if (l <= tb) {
            const uint32_t e = symbol_entry(is_dist, i, l);
            { INFL_LANES(lane) { for (uint32_t k = rev + ((uint32_t)lane << l); k < main_size; k += 32u << l) tab[k] = e; } }
        } else {
            const uint32_t p = tab[rev & (main_size - 1u)], e = symbol_entry(is_dist, i, l - tb);
            const uint32_t start = E_VAL(p), sb = E_EXTRA(p);
            { INFL_LANES(lane) { for (uint32_t k = (rev >> tb) + ((uint32_t)lane << (l - tb)); k < (1u << sb); k += 32u << (l - tb)) tab[start + k] = e; } }
        }
    }
    INFL_SYNC();
    return 0;
}

/* two literals from one look-up where both codes fit the first-level index (ssn_inflate.c has the argument) */
INFL_FN void pair_literals(uint32_t *tab, uint32_t tb)
{
    const uint32_t n = 1u << tb;
    /* the pass reads single entries and writes pairs: read everything a lane needs before anything is written */
    INFL_SYNC();
#ifdef SSN_INFL_HOST
    uint32_t copy[1 << kLLBits];
    for (uint32_t i = 0; i < n; ++i) copy[i] = tab[i];
    for (uint32_t i = 0; i < n; ++i) {
        const uint32_t e1 = copy[i];
        if (E_KIND(e1) == K_LIT && E_LEN(e1) < tb) {
            const uint32_t l1 = E_LEN(e1), e2 = copy[i >> l1];
            if (E_KIND(e2) == K_LIT && E_LEN(e2) <= tb - l1) tab[i] = entry(E_VAL(e1) | E_VAL(e2) << 8, K_LIT2, 0, l1 + E_LEN(e2));
        }
    }
#else
    const int lane = (int)(threadIdx.x & 31u);
    uint32_t mine[(1 << kLLBits) / 32];
#pragma unroll
    for (uint32_t j = 0; j < n / 32u; ++j) {
        const uint32_t i = j * 32u + (uint32_t)lane, e1 = tab[i];
        uint32_t out = e1;
        if (E_KIND(e1) == K_LIT && E_LEN(e1) < tb) {
            const uint32_t l1 = E_LEN(e1), e2 = tab[i >> l1];
            if (E_KIND(e2) == K_LIT && E_LEN(e2) <= tb - l1) out = entry(E_VAL(e1) | E_VAL(e2) << 8, K_LIT2, 0, l1 + E_LEN(e2));
        }
        mine[j] = out;
    }
    __syncwarp();
#pragma unroll
    for (uint32_t j = 0; j < n / 32u; ++j) tab[j * 32u + (uint32_t)lane] = mine[j];
#endif
    INFL_SYNC();
}

/* ---- the bit reader: 32-bit aligned words, the same on every lane ------------------------------------------------ */
/* The input is staged through registers: lane l keeps word (chunk * 32 + l) of the current 32-word chunk and of the
 * next one, both fetched with one coalesced 128-byte load.  A refill is then a shuffle (the word's owner broadcasts it),
 * and the global load of the chunk after next is in flight while two chunks' worth of symbols are decoded: the input
 * latency, a dependent L2/DRAM round trip every 32 bits before, leaves the decode loop. */

struct Bits {
    const uint32_t *w;         /* aligned word pointer at or below the first input byte */
    uint32_t widx;             /* next word to take */
    uint32_t lead;             /* bytes of w[0] that lie in front of the input */
    uint32_t limit;            /* words of w[] that may be read (the caller's buffer is readable 16 bytes past the input) */
    uint64_t buf;
    uint32_t cnt;
#ifdef SSN_INFL_HOST
    uint32_t cur[32], nxt[32];
#else
    uint32_t cur, nxt;
#endif
};

INFL_FN uint32_t load_word(const uint32_t *p)
{
#ifdef SSN_INFL_HOST
    return *p;                 /* (the host test passes aligned, padded buffers) */
#else
    return __ldg(p);
#endif
}

/* lane l <- word first + l (0 behind the readable end) */
#ifdef SSN_INFL_HOST
INFL_FN void bits_fetch(const Bits &b, uint32_t *dst, uint32_t first)
{
    for (uint32_t l = 0; l < 32u; ++l) dst[l] = first + l < b.limit ? load_word(b.w + first + l) : 0u;
}
INFL_FN uint32_t bits_word(const Bits &b, uint32_t idx) { return b.cur[idx & 31u]; }
#else
INFL_FN void bits_fetch(const Bits &b, uint32_t *dst, uint32_t first)
{
    const uint32_t i = first + (threadIdx.x & 31u);
    *dst = i < b.limit ? load_word(b.w + i) : 0u;
}
INFL_FN uint32_t bits_word(const Bits &b, uint32_t idx) { return __shfl_sync(0xffffffffu, b.cur, (int)(idx & 31u)); }
#endif

/* positions the reader on input byte byte_pos; `remaining` = input bytes from there to the end */
INFL_FN void bits_seek(Bits &b, const uint8_t *in, uint32_t byte_pos, uint32_t remaining)
{
    const uintptr_t a = (uintptr_t)(in + byte_pos);
    b.w = (const uint32_t*)(a & ~(uintptr_t)3);
    b.lead = (uint32_t)(a & 3u);
    b.limit = (b.lead + remaining + 16u) >> 2;
#ifdef SSN_INFL_HOST
    bits_fetch(b, b.cur, 0u);
    bits_fetch(b, b.nxt, 32u);
#else
    bits_fetch(b, &b.cur, 0u);
    bits_fetch(b, &b.nxt, 32u);
#endif
    b.widx = 1;
    b.buf = (uint64_t)(bits_word(b, 0u) >> (8u * b.lead));
    b.cnt = 32u - 8u * b.lead;
}
INFL_FN void bits_refill(Bits &b)          /* afterwards at least 33 bits are buffered; the same on every lane */
{
    if (b.cnt <= 32u) {
        b.buf |= (uint64_t)bits_word(b, b.widx) << b.cnt;
        b.widx += 1; b.cnt += 32u;
        if ((b.widx & 31u) == 0u) {
#ifdef SSN_INFL_HOST
            for (int l = 0; l < 32; ++l) b.cur[l] = b.nxt[l];
            bits_fetch(b, b.nxt, b.widx + 32u);
#else
            b.cur = b.nxt;
            bits_fetch(b, &b.nxt, b.widx + 32u);
#endif
        }
    }
}
/* bytes of the input consumed so far, counted from the last seek position (rounded up to whole bytes) */
INFL_FN uint32_t bits_consumed_bytes(const Bits &b) { return (b.widx * 32u - 8u * b.lead - b.cnt + 7u) >> 3; }
#define BITS_PEEK(b, n) ((uint32_t)((b).buf) & ((1u << (n)) - 1u))
#define BITS_TAKE(b, n) ((b).buf >>= (n), (b).cnt -= (n))

/* ---- output: a 32-byte window in registers ------------------------------------------------------------------------- */

struct Out {
    uint8_t *p;                /* start of the block's output */
    uint32_t pos, len;
    uint32_t win_lo;           /* lanes [win_lo, pos & 31) of the current window hold bytes not yet stored */
#ifdef SSN_INFL_HOST
    uint8_t mine[32];
#else
    uint32_t mine;
#endif
};

INFL_FN void out_flush(Out &o, uint32_t hi)          /* stores the window's lanes [win_lo, hi) */
{
    const uint32_t base = (o.pos - 1u) & ~31u;       /* window of the last byte put (hi == 32) or of pos (hi == pos & 31) */
    { INFL_LANES(lane) {
        if ((uint32_t)lane >= o.win_lo && (uint32_t)lane < hi) {
#ifdef SSN_INFL_HOST
            o.p[base + (uint32_t)lane] = o.mine[lane];
#else
            o.p[base + (uint32_t)lane] = (uint8_t)o.mine;
#endif
        }
    } }
}

INFL_FN void out_literal(Out &o, uint32_t v)        /* caller checked pos < len */
{
    { INFL_LANES(lane) {
        if ((uint32_t)lane == (o.pos & 31u)) {
#ifdef SSN_INFL_HOST
            o.mine[lane] = (uint8_t)v;
#else
            o.mine = v;
#endif
        }
    } }
    o.pos += 1u;
    if ((o.pos & 31u) == 0u) { out_flush(o, 32u); o.win_lo = 0u; }
}

/* pending literals to memory, so that a copy may read them and write behind them */
INFL_FN void out_settle(Out &o)
{
    const uint32_t hi = o.pos & 31u;
    if (hi > o.win_lo) {
        const uint32_t base = o.pos & ~31u;
        { INFL_LANES(lane) {
            if ((uint32_t)lane >= o.win_lo && (uint32_t)lane < hi) {
#ifdef SSN_INFL_HOST
                o.p[base + (uint32_t)lane] = o.mine[lane];
#else
                o.p[base + (uint32_t)lane] = (uint8_t)o.mine;
#endif
            }
        } }
    }
    INFL_SYNC();
}

INFL_FN uint32_t load_out(const uint8_t *p)
{
#ifdef SSN_INFL_HOST
    return *p;
#else
    /* An ordinary load: the bytes were stored by lanes of this warp before the last __syncwarp, which orders them for
     * every lane taking part, and one SM has one L1 - so the copy may hit L1 instead of paying an L2 round trip per
     * match (a quarter of all stall samples with ld.cg, profiles/r02/inflate_ncu.md). */
    return *p;
#endif
}

/* some r with (i * r) >> 20 == i / dist for i, dist <= 258: any r in (2^20 / dist, 2^20 / dist + 2] does, since
 * i * dist * 2 < 2^20 */
INFL_FN uint32_t recip20(uint32_t dist)
{
#ifdef SSN_INFL_HOST
    return ((1u << 20) + dist - 1u) / dist;
#else
    return (uint32_t)(__frcp_ru((float)dist) * 1048576.0f) + 1u;       /* rounded up, times 2^20 exactly, truncated, + 1 */
#endif
}

INFL_FN void out_match(Out &o, uint32_t length, uint32_t dist)    /* caller checked dist <= pos, pos + length <= len */
{
    out_settle(o);
    uint8_t *dst = o.p + o.pos;
    if (length <= 32u) {
        /* Nearly every match of a level-1 BAM (19 turns of the symbol loop in 20 are matches, 8 bytes each on average):
         * one predicated load and store, no loop. */
        if (dist >= length) {
            INFL_LANES(lane) { if ((uint32_t)lane < length) dst[lane] = (uint8_t)load_out(dst + lane - dist); }
        } else if (dist == 1u) {                                      /* a run of one byte (qualities) */
            INFL_LANES(lane) { if ((uint32_t)lane < length) dst[lane] = (uint8_t)load_out(dst - 1); }
        } else {
            const uint32_t recip = recip20(dist);
            INFL_LANES(lane) {
                if ((uint32_t)lane < length) { const uint32_t q = ((uint32_t)lane * recip) >> 20; dst[lane] = (uint8_t)load_out(dst - dist + ((uint32_t)lane - q * dist)); }
            }
        }
    } else if (dist >= length) {
#ifndef SSN_INFL_HOST
#pragma unroll 1
#endif
        for (uint32_t base = 0; base < length; base += 32u) {
            INFL_LANES(lane) { const uint32_t i = base + (uint32_t)lane; if (i < length) dst[i] = (uint8_t)load_out(dst + i - dist); }
        }
    } else {
        /* the copy overlaps itself: byte i repeats byte i mod dist of the `dist` bytes in front of the match, all of
         * which are in memory already */
        const uint32_t recip = recip20(dist);
        const uint8_t *src = dst - dist;
#ifndef SSN_INFL_HOST
#pragma unroll 1
#endif
        for (uint32_t base = 0; base < length; base += 32u) {
            INFL_LANES(lane) {
                const uint32_t i = base + (uint32_t)lane;
                if (i < length) { const uint32_t q = (i * recip) >> 20; dst[i] = (uint8_t)load_out(src + (i - q * dist)); }
            }
        }
    }
    INFL_SYNC();
    o.pos += length;
    o.win_lo = o.pos & 31u;
}

/* ---- one block ------------------------------------------------------------------------------------------------------- */

INFL_FN int inflate_block(Work &W, const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t out_len)
{
    Bits b;
    Out o;
    o.p = out; o.pos = 0; o.len = out_len; o.win_lo = 0;
#ifndef SSN_INFL_HOST
    o.mine = 0;
#endif
    bits_seek(b, in, 0, in_len);
    uint32_t seek_base = 0;                 /* input byte the bit reader was last positioned on */
    for (uint32_t guard = 0; guard < 65536u; ++guard) {
        bits_refill(b);
        if (seek_base + bits_consumed_bytes(b) > in_len) return 10;
        const uint32_t last = BITS_PEEK(b, 1); BITS_TAKE(b, 1);
        const uint32_t type = BITS_PEEK(b, 2); BITS_TAKE(b, 2);
        const uint32_t *ll = W.ll, *dd = W.d;
        if (type == 0u) {
            /* stored: to the byte boundary, LEN, NLEN, bytes */
            BITS_TAKE(b, b.cnt & 7u);
            bits_refill(b);
            const uint32_t len = BITS_PEEK(b, 16); BITS_TAKE(b, 16);
            const uint32_t nlen = BITS_PEEK(b, 16); BITS_TAKE(b, 16);
            const uint32_t at = seek_base + (b.widx * 32u - 8u * b.lead - b.cnt) / 8u;      /* first data byte (the reader is byte aligned) */
            if ((len ^ nlen) != 0xffffu || at + len > in_len || o.pos + len > o.len) return 11;
            out_settle(o);
            for (uint32_t base = 0; base < len; base += 32u) {
                INFL_LANES(lane) { const uint32_t i = base + (uint32_t)lane; if (i < len) o.p[o.pos + i] = in[at + i]; }
            }
            INFL_SYNC();
            o.pos += len; o.win_lo = o.pos & 31u;
            seek_base = at + len;
            bits_seek(b, in, seek_base, in_len - seek_base);
            if (last) break;
            continue;
        }
        if (type == 3u) return 12;
        if (type == 1u) {
            INFL_SYNC();
            { INFL_LANES(lane) { for (uint32_t i = (uint32_t)lane; i < 320u; i += 32u) W.lens[i] = i < 144u ? 8 : i < 256u ? 9 : i < 280u ? 7 : i < 288u ? 8 : 5; } }
            INFL_SYNC();
            if (build((uint8_t*)W.d, W.ll, kLLSub, W.lens, 288, kLLBits, false) || build(W.lens, W.d, kDSub, W.lens + 288, 32, kDBits, true)) return 13;
            pair_literals(W.ll, kLLBits);
        } else {
            const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            const uint32_t hlit = BITS_PEEK(b, 5) + 257u; BITS_TAKE(b, 5);
            const uint32_t hdist = BITS_PEEK(b, 5) + 1u; BITS_TAKE(b, 5);
            const uint32_t hclen = BITS_PEEK(b, 4) + 4u; BITS_TAKE(b, 4);
            if (hlit > 286u || hdist > 30u) return 14;
            INFL_SYNC();
            { INFL_LANES(lane) { for (uint32_t i = (uint32_t)lane; i < 320u; i += 32u) W.lens[i] = 0; } }
            INFL_SYNC();
            /* the code-length code's lengths sit in lens[288..307) while its table (in W.d's space) is built */
            for (uint32_t i = 0; i < hclen; ++i) {
                bits_refill(b);
                const uint32_t v = BITS_PEEK(b, 3); BITS_TAKE(b, 3);
                INFL_SYNC();
                { INFL_LANES(lane) { if (lane == 0) W.lens[288 + order[i]] = (uint8_t)v; } }
            }
            INFL_SYNC();
            if (build(W.lens, W.d, 0, W.lens + 288, 19, 7, false)) return 15;
            { INFL_LANES(lane) { if (lane < 19) W.lens[288 + lane] = 0; } }
            INFL_SYNC();
            const uint32_t n = hlit + hdist;
            /* lengths land in lens[0, hlit) and, shifted, in lens[288, 288 + hdist) */
            for (uint32_t i = 0; i < n;) {
                bits_refill(b);
                if (seek_base + bits_consumed_bytes(b) > in_len) return 16;
                const uint32_t e = W.d[BITS_PEEK(b, 7)];
                if (E_KIND(e) != K_LIT) return 17;
                BITS_TAKE(b, E_LEN(e));
                const uint32_t sym = E_VAL(e);
                uint32_t rep, val;
                if (sym < 16u) { rep = 1; val = sym; }
                else if (sym == 16u) {
                    if (i == 0u) return 18;
                    const uint32_t j = i - 1u;
                    val = W.lens[j < hlit ? j : 288u + (j - hlit)]; rep = 3u + BITS_PEEK(b, 2); BITS_TAKE(b, 2);
                }
                else if (sym == 17u) { val = 0; rep = 3u + BITS_PEEK(b, 3); BITS_TAKE(b, 3); }
                else if (sym == 18u) { val = 0; rep = 11u + BITS_PEEK(b, 7); BITS_TAKE(b, 7); }
                else return 19;
                if (i + rep > n) return 20;
                INFL_SYNC();
                { INFL_LANES(lane) {
                    if ((uint32_t)lane < rep || rep > 32u) {
                        for (uint32_t k = (uint32_t)lane; k < rep; k += 32u) { const uint32_t j = i + k; W.lens[j < hlit ? j : 288u + (j - hlit)] = (uint8_t)val; }
                    }
                } }
                INFL_SYNC();
                i += rep;
            }
            if (W.lens[256] == 0) return 21;
            if (build((uint8_t*)W.d, W.ll, kLLSub, W.lens, 288, kLLBits, false) || build(W.lens, W.d, kDSub, W.lens + 288, 32, kDBits, true)) return 22;
            pair_literals(W.ll, kLLBits);
        }
        /* the symbols of the block */
        for (;;) {
            /* no input bound check per symbol: behind the input the reader yields zero words (bits_fetch), every turn of
             * this loop either ends it or produces output, and the output is bounded - a stream that runs off its end is
             * caught by the output checks or by the consumed-bytes check behind the loop */
            bits_refill(b);
            uint32_t e = ll[BITS_PEEK(b, kLLBits)];
            if (E_KIND(e) == K_SUB) { BITS_TAKE(b, kLLBits); e = ll[E_VAL(e) + BITS_PEEK(b, E_EXTRA(e))]; }
            if (E_ISLIT(e)) {
                /* up to two look-ups (four bytes) per refill: 2 x 15 bits <= 33 */
                const uint32_t n1 = E_NLIT(e);
                if (o.pos + n1 > o.len) return 31;
                out_literal(o, E_VAL(e) & 0xffu);
                if (n1 == 2u) out_literal(o, (E_VAL(e) >> 8) & 0xffu);
                BITS_TAKE(b, E_LEN(e));
                e = ll[BITS_PEEK(b, kLLBits)];
                if (E_KIND(e) == K_SUB) { BITS_TAKE(b, kLLBits); e = ll[E_VAL(e) + BITS_PEEK(b, E_EXTRA(e))]; }
                if (E_ISLIT(e)) {
                    const uint32_t n2 = E_NLIT(e);
                    if (o.pos + n2 > o.len) return 31;
                    out_literal(o, E_VAL(e) & 0xffu);
                    if (n2 == 2u) out_literal(o, (E_VAL(e) >> 8) & 0xffu);
                    BITS_TAKE(b, E_LEN(e));
                    continue;
                }
                bits_refill(b);
            }
            BITS_TAKE(b, E_LEN(e));
            if (E_KIND(e) == K_EOB) break;
            if (E_KIND(e) != K_LEN) return 32;
            const uint32_t length = E_VAL(e) + BITS_PEEK(b, E_EXTRA(e)); BITS_TAKE(b, E_EXTRA(e));
            bits_refill(b);
            e = dd[BITS_PEEK(b, kDBits)];
            if (E_KIND(e) == K_SUB) { BITS_TAKE(b, kDBits); e = dd[E_VAL(e) + BITS_PEEK(b, E_EXTRA(e))]; }
            if (E_KIND(e) != K_LEN) return 33;
            BITS_TAKE(b, E_LEN(e));
            const uint32_t dist = E_VAL(e) + BITS_PEEK(b, E_EXTRA(e)); BITS_TAKE(b, E_EXTRA(e));
            if (dist > o.pos || o.pos + length > o.len) return 34;
            out_match(o, length, dist);
        }
        if (seek_base + bits_consumed_bytes(b) > in_len) return 35;
        if (last) break;
    }
    out_settle(o);
    return o.pos == o.len ? 0 : 40;
}

}  // namespace infl
