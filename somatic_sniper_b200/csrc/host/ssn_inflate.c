/*
 * ssn_inflate.c — raw DEFLATE (RFC 1951) decoder for BGZF blocks, written for throughput.
 *
 * Replaces zlib's inflate() in the BAM readers (the reference inflates every block with inflateInit2/inflate/inflateEnd,
 * samtools-0.1.6 bgzf.c:330-353).  After the region shards made reading parallel, inflate is what the host cores spend
 * their time on, so this decoder does what zlib's portable inner loop does not: a 64-bit bit buffer refilled with one
 * unaligned load, an 11-bit first-level table for literals/lengths and an 8-bit one for distances (second-level tables
 * only for longer codes), up to three literals per refill, and match copies in 8-byte words.  One whole block per call,
 * output size known in advance (BGZF's ISIZE).
 *
 * Contract: `in` must be readable for 16 bytes past in_len (BGZF blocks end in an 8-byte footer and the readers pad their
 * buffers); nothing is written outside out[0, out_len).  Returns 0 when the stream decodes to exactly out_len bytes,
 * -1 otherwise (the callers then fall back to zlib for that block, so a stream this decoder rejects is still read the way
 * the reference reads it).
 */
#include "ssn_inflate.h"

#include <string.h>

#define LL_BITS 11
#define D_BITS 8
#define LL_SIZE ((1 << LL_BITS) + 288 * 16)
#define D_SIZE ((1 << D_BITS) + 32 * 128)

/* table entry: bits 0-3 code bits to consume, 4-7 extra bits (sub-table pointer: its index width), 8-10 kind,
 * 12-13 literals held (1 or 2), 15 set for literal entries, 16-31 value */
#define K_LIT 0u
#define K_LEN 1u
#define K_EOB 2u
#define K_SUB 3u
#define K_BAD 4u
#define K_LIT2 5u                       /* two literals whose codes fit the first-level index together: value = first | second << 8 */
#define ENTRY(val, kind, extra, len) ((uint32_t)(val) << 16 | (uint32_t)(kind) << 8 | (uint32_t)(extra) << 4 | (uint32_t)(len) | \
                                      ((kind) == K_LIT ? 0x9000u : (kind) == K_LIT2 ? 0xa000u : 0u))
#define E_ISLIT(e) ((e) & 0x8000u)
#define E_NLIT(e) (((e) >> 12) & 3u)
#define E_LEN(e) ((e) & 15u)
#define E_EXTRA(e) (((e) >> 4) & 15u)
#define E_KIND(e) (((e) >> 8) & 7u)
#define E_VAL(e) ((e) >> 16)

typedef struct {
    uint32_t ll[LL_SIZE];
    uint32_t d[D_SIZE];
    uint8_t lens[288 + 32];
    uint8_t sub_bits[1 << LL_BITS];
    uint32_t single[1 << LL_BITS];
} tables;

static const uint16_t len_base[29] = {3, 4, 5, 6, 7, 8, 9, 10, 11, 13, 15, 17, 19, 23, 27, 31, 35, 43, 51, 59, 67, 83, 99, 115, 131, 163, 195, 227, 258};
static const uint8_t len_extra[29] = {0, 0, 0, 0, 0, 0, 0, 0, 1, 1, 1, 1, 2, 2, 2, 2, 3, 3, 3, 3, 4, 4, 4, 4, 5, 5, 5, 5, 0};
static const uint16_t dist_base[30] = {1, 2, 3, 4, 5, 7, 9, 13, 17, 25, 33, 49, 65, 97, 129, 193, 257, 385, 513, 769, 1025, 1537, 2049, 3073, 4097, 6145, 8193, 12289, 16385, 24577};
static const uint8_t dist_extra[30] = {0, 0, 0, 0, 1, 1, 2, 2, 3, 3, 4, 4, 5, 5, 6, 6, 7, 7, 8, 8, 9, 9, 10, 10, 11, 11, 12, 12, 13, 13};

static inline uint32_t symbol_entry(int is_dist, unsigned sym, unsigned len)
{
    if (is_dist) return sym < 30 ? ENTRY(dist_base[sym], K_LEN, dist_extra[sym], len) : ENTRY(0, K_BAD, 0, len);
    if (sym < 256) return ENTRY(sym, K_LIT, 0, len);
    if (sym == 256) return ENTRY(0, K_EOB, 0, len);
    if (sym < 286) return ENTRY(len_base[sym - 257], K_LEN, len_extra[sym - 257], len);
    return ENTRY(0, K_BAD, 0, len);
}

static inline unsigned bitrev(unsigned code, unsigned len)
{
    unsigned r = 0, i;
    for (i = 0; i < len; ++i) { r = r << 1 | (code & 1); code >>= 1; }
    return r;
}

/* canonical Huffman code -> two-level decode table; -1 for an over-subscribed code.  Incomplete codes are accepted (their
 * unused entries are K_BAD): a distance code with a single symbol is legal, and anything else fails only if it is used. */
static int build(uint32_t *tab, uint8_t *sub_bits, const uint8_t *lens, unsigned nsym, unsigned tb, int is_dist)
{
    unsigned count[16], next[16], i, l, code, nfree;
    const unsigned main_size = 1u << tb;
    int long_codes = 0;
    for (i = 0; i < 16; ++i) count[i] = 0;
    for (i = 0; i < nsym; ++i) count[lens[i]]++;
    count[0] = 0;
    code = 0;
    {
        int left = 1;
        for (l = 1; l < 16; ++l) { left = (left << 1) - (int)count[l]; if (left < 0) return -1; }
    }
    for (l = 1; l < 16; ++l) { code = (code + count[l - 1]) << 1; next[l] = code; if (l > tb && count[l]) long_codes = 1; }
    for (i = 0; i < main_size; ++i) tab[i] = ENTRY(0, K_BAD, 0, 1);
    if (long_codes) {
        unsigned nx[16];
        memset(sub_bits, 0, main_size);
        for (l = 1; l < 16; ++l) nx[l] = next[l];
        for (i = 0; i < nsym; ++i) {
            l = lens[i];
            if (l > tb) {
                const unsigned rev = bitrev(nx[l]++, l), prefix = rev & (main_size - 1);
                if (sub_bits[prefix] < l - tb) sub_bits[prefix] = (uint8_t)(l - tb);
            } else if (l) nx[l]++;
        }
        nfree = main_size;
        for (i = 0; i < main_size; ++i) {
            if (sub_bits[i]) {
                unsigned k;
                tab[i] = ENTRY(nfree, K_SUB, sub_bits[i], tb);
                for (k = 0; k < (1u << sub_bits[i]); ++k) tab[nfree + k] = ENTRY(0, K_BAD, 0, 1);
                nfree += 1u << sub_bits[i];
            }
        }
    }
    for (i = 0; i < nsym; ++i) {
        l = lens[i];
        if (!l) continue;
        {
            const unsigned rev = bitrev(next[l]++, l);
            unsigned k;
            if (l <= tb) {
                const uint32_t e = symbol_entry(is_dist, i, l);
                for (k = rev; k < main_size; k += 1u << l) tab[k] = e;
            } else {
                const uint32_t p = tab[rev & (main_size - 1)], e = symbol_entry(is_dist, i, l - tb);
                const unsigned start = E_VAL(p), sb = E_EXTRA(p);
                for (k = rev >> tb; k < (1u << sb); k += 1u << (l - tb)) tab[start + k] = e;
            }
        }
    }
    return 0;
}

/* Literal-heavy streams (base qualities, packed bases) spend one dependent table look-up per byte.  Where the codes of two
 * consecutive literals fit the first-level index together, one entry yields both: entries replicate with period 2^len, so
 * the entry at (i >> len1) with the upper bits clear is the second symbol's whenever its code is short enough. */
static void pair_literals(uint32_t *tab, uint32_t *single, unsigned tb)
{
    const unsigned n = 1u << tb;
    unsigned i;
    memcpy(single, tab, n * sizeof(uint32_t));
    for (i = 0; i < n; ++i) {
        const uint32_t e1 = single[i];
        if (E_KIND(e1) == K_LIT && E_LEN(e1) < tb) {
            const unsigned l1 = E_LEN(e1);
            const uint32_t e2 = single[i >> l1];
            if (E_KIND(e2) == K_LIT && E_LEN(e2) <= tb - l1) tab[i] = ENTRY(E_VAL(e1) | E_VAL(e2) << 8, K_LIT2, 0, l1 + E_LEN(e2));
        }
    }
}

static inline uint64_t load64(const uint8_t *p) { uint64_t v; memcpy(&v, p, 8); return v; }

int ssn_inflate_raw(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_len)
{
    static __thread tables T;
    static __thread int fixed_ready;
    static __thread uint32_t fixed_ll[1 << LL_BITS], fixed_d[1 << D_BITS];
    const uint8_t *in_next = in, *const in_end = in + in_len;
    uint8_t *out_next = out, *const out_end = out + out_len;
    uint64_t bitbuf = 0;
    unsigned bitcnt = 0;
    int last;

#define REFILL() do { bitbuf |= load64(in_next) << bitcnt; in_next += (63 - bitcnt) >> 3; bitcnt |= 56; } while (0)
#define OVERRUN() ((size_t)(in_next - in) > in_len + (bitcnt >> 3))          /* consumed bits beyond the end of the input */
#define TAKE(n) (bitbuf >>= (n), bitcnt -= (n))
#define BITS(n) ((unsigned)(bitbuf & (((uint64_t)1 << (n)) - 1)))

    do {
        unsigned type;
        const uint32_t *ll, *dd;
        REFILL();
        if (OVERRUN()) return -1;
        last = (int)BITS(1); TAKE(1);
        type = BITS(2); TAKE(2);
        if (type == 0) {
            /* stored: skip to the byte boundary, LEN, NLEN, bytes */
            unsigned len, nlen;
            TAKE(bitcnt & 7);
            /* give the whole bytes still in the bit buffer back to the input */
            in_next -= bitcnt >> 3; bitbuf = 0; bitcnt = 0;
            if (in_next + 4 > in_end) return -1;
            len = in_next[0] | (unsigned)in_next[1] << 8; nlen = in_next[2] | (unsigned)in_next[3] << 8;
            in_next += 4;
            if ((len ^ nlen) != 0xffffu || (size_t)(in_end - in_next) < len || (size_t)(out_end - out_next) < len) return -1;
            memcpy(out_next, in_next, len);
            in_next += len; out_next += len;
            continue;
        }
        if (type == 1) {
            if (!fixed_ready) {
                unsigned i;
                for (i = 0; i < 144; ++i) T.lens[i] = 8;
                for (; i < 256; ++i) T.lens[i] = 9;
                for (; i < 280; ++i) T.lens[i] = 7;
                for (; i < 288; ++i) T.lens[i] = 8;
                for (i = 0; i < 32; ++i) T.lens[288 + i] = 5;
                if (build(T.ll, T.sub_bits, T.lens, 288, LL_BITS, 0) != 0 || build(T.d, T.sub_bits, T.lens + 288, 32, D_BITS, 1) != 0) return -1;
                pair_literals(T.ll, T.single, LL_BITS);
                memcpy(fixed_ll, T.ll, sizeof(fixed_ll)); memcpy(fixed_d, T.d, sizeof(fixed_d));
                fixed_ready = 1;
            }
            ll = fixed_ll; dd = fixed_d;
        } else if (type == 2) {
            static const uint8_t order[19] = {16, 17, 18, 0, 8, 7, 9, 6, 10, 5, 11, 4, 12, 3, 13, 2, 14, 1, 15};
            unsigned hlit, hdist, hclen, i, n;
            uint8_t cl[19];
            uint32_t *pre = T.d;                               /* the code-length code's table lives in the distance table's space */
            hlit = BITS(5) + 257; TAKE(5);
            hdist = BITS(5) + 1; TAKE(5);
            hclen = BITS(4) + 4; TAKE(4);
            if (hlit > 286 || hdist > 30) return -1;
            memset(cl, 0, sizeof(cl));
            REFILL();
            for (i = 0; i < hclen; ++i) {
                if (bitcnt < 3) REFILL();
                cl[order[i]] = (uint8_t)BITS(3); TAKE(3);
            }
            if (OVERRUN()) return -1;
            if (build(pre, T.sub_bits, cl, 19, 7, 0) != 0) return -1;      /* 7-bit codes: one level; values 0..18 come out as literals */
            n = hlit + hdist;
            for (i = 0; i < n;) {
                uint32_t e;
                unsigned sym, rep, val;
                REFILL();
                if (OVERRUN()) return -1;
                e = pre[BITS(7)];
                if (E_KIND(e) != K_LIT) return -1;
                TAKE(E_LEN(e));
                sym = E_VAL(e);
                if (sym < 16) { T.lens[i++] = (uint8_t)sym; continue; }
                if (sym == 16) { if (i == 0) return -1; val = T.lens[i - 1]; rep = 3 + BITS(2); TAKE(2); }
                else if (sym == 17) { val = 0; rep = 3 + BITS(3); TAKE(3); }
                else if (sym == 18) { val = 0; rep = 11 + BITS(7); TAKE(7); }
                else return -1;
                if (i + rep > n) return -1;
                while (rep--) T.lens[i++] = (uint8_t)val;
            }
            if (T.lens[256] == 0) return -1;
            {
                uint8_t dl[32];
                memset(dl, 0, sizeof(dl));
                memcpy(dl, T.lens + hlit, hdist);
                memset(T.lens + hlit, 0, 288 - hlit);
                if (build(T.ll, T.sub_bits, T.lens, 288, LL_BITS, 0) != 0 || build(T.d, T.sub_bits, dl, 32, D_BITS, 1) != 0) return -1;
                pair_literals(T.ll, T.single, LL_BITS);
            }
            ll = T.ll; dd = T.d;
        } else return -1;

        /* the symbols of the block */
        for (;;) {
            uint32_t e;
            unsigned length, dist;
            REFILL();
            if (OVERRUN()) return -1;
            e = ll[BITS(LL_BITS)];
            if (E_KIND(e) == K_SUB) { TAKE(LL_BITS); e = ll[E_VAL(e) + BITS(E_EXTRA(e))]; }
            /* literals: up to three look-ups (six bytes) on one refill: 3 x 15 bits <= 56.  Away from the end of the
             * output both bytes of an entry are stored unconditionally and the pointer moves by the entry's count: no
             * branch on "one literal or two". */
#define LOOKUP() do { e = ll[BITS(LL_BITS)]; if (E_KIND(e) == K_SUB) { TAKE(LL_BITS); e = ll[E_VAL(e) + BITS(E_EXTRA(e))]; } } while (0)
#define PUT_LITERALS() do { const uint16_t v_ = (uint16_t)E_VAL(e); memcpy(out_next, &v_, 2); out_next += E_NLIT(e); TAKE(E_LEN(e)); } while (0)
            if (E_ISLIT(e)) {
                if ((size_t)(out_end - out_next) >= 8) {
                    PUT_LITERALS();
                    LOOKUP();
                    if (E_ISLIT(e)) {
                        PUT_LITERALS();
                        LOOKUP();
                        if (E_ISLIT(e)) { PUT_LITERALS(); continue; }
                    }
                    if (bitcnt < 32) REFILL();                   /* a length code with its extra bits: <= 20 bits */
                } else {
                    const unsigned n = E_NLIT(e);
                    if ((size_t)(out_end - out_next) < n) return -1;
                    out_next[0] = (uint8_t)E_VAL(e);
                    if (n == 2) out_next[1] = (uint8_t)(E_VAL(e) >> 8);
                    out_next += n;
                    TAKE(E_LEN(e));
                    continue;
                }
            }
            TAKE(E_LEN(e));
            if (E_KIND(e) == K_EOB) break;
            if (E_KIND(e) != K_LEN) return -1;
            length = E_VAL(e) + BITS(E_EXTRA(e)); TAKE(E_EXTRA(e));
            if (bitcnt < 32) REFILL();                           /* a distance code with its extra bits: <= 28 bits */
            e = dd[BITS(D_BITS)];
            if (E_KIND(e) == K_SUB) { TAKE(D_BITS); e = dd[E_VAL(e) + BITS(E_EXTRA(e))]; }
            if (E_KIND(e) != K_LEN) return -1;
            TAKE(E_LEN(e));
            dist = E_VAL(e) + BITS(E_EXTRA(e)); TAKE(E_EXTRA(e));
            if (dist > (size_t)(out_next - out) || length > (size_t)(out_end - out_next)) return -1;
            {
                const uint8_t *src = out_next - dist;
                uint8_t *dst = out_next;
                out_next += length;
                if (dist >= 8 && (size_t)(out_end - dst) >= length + 8) {
                    /* whole words; may write up to 7 bytes past the match, still inside the block's output */
                    uint8_t *const stop = dst + length;
                    do { uint64_t v = load64(src); memcpy(dst, &v, 8); src += 8; dst += 8; } while (dst < stop);
                } else if (dist == 1 && (size_t)(out_end - dst) >= length + 8) {
                    const uint64_t v = 0x0101010101010101ull * src[0];
                    uint8_t *const stop = dst + length;
                    do { memcpy(dst, &v, 8); dst += 8; } while (dst < stop);
                } else {
                    while (length--) *dst++ = *src++;
                }
            }
        }
        if (OVERRUN()) return -1;
    } while (!last);
    if (OVERRUN()) return -1;
    return out_next == out_end ? 0 : -1;
#undef REFILL
#undef LOOKUP
#undef PUT_LITERALS
#undef OVERRUN
#undef TAKE
#undef BITS
}
