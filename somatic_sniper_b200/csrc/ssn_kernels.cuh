/*
 * ssn_kernels.cuh — the fused per-column scoring kernel (sm_100a).
 *
 * One launch scores a whole SoA batch.  A persistent grid of warps pulls TILES of 32 consecutive
 * columns from an atomic counter.  Each tile runs in two phases inside one warp:
 *
 *  FRONT (warp-per-column, 32 lanes = 32 read-bases at a time; looped over the tile's columns)
 *    load bs/bq/mq coalesced -> 16-bit sort key per read-base -> keys of the tile to shared memory
 *  BACK (lane-per-column, 32 columns in flight per warp)
 *    sort the column's keys, walk them: the strictly sequential float<-double error-dependency chains
 *    (sniper_maqcns.c:160-176), then the 10 genotype likelihoods, the p[k,k] correction, the
 *    quantisation, the consensus call (both samples), then the rare somatic-score branch
 *    (somatic_sniper.c:156-227) and, for sites that pass the gate, warp-cooperative site statistics
 *    (dqstats.c:6-53) and a compacted ssn_site record.  GLFs never leave registers.
 *
 * The 16-bit key.  The reference sorts x = q<<24 | valid<<21 | strand<<18 | b4<<16 | baseQ<<8 | mapQ
 * ascending and consumes it from the top (sniper_maqcns.c:144-162).  All state touched by an entry
 * (esum/fsum/c of its base, w of its strand/base class) belongs to its base b4, so the four bases'
 * chains are independent and may be walked one after another; inside one base the reference order is
 * q desc, valid desc, strand desc, baseQ desc.  mapQ never changes a contribution, and baseQ only
 * matters for q in {1,2,3}, where entries with baseQ in {64,128,192} keep q while their neighbours
 * are promoted to 4 (:165).  Hence
 *     key = sample<<15 | b4<<13 | q<<5 | valid<<4 | strand<<3 | sub
 * with sub = 1 + rank of baseQ's group in {[0,63], 64, [65,127], 128, [129,191], 192, [193,255]} (1..7, so
 * a contributing key is never 0; even <=> baseQ in {64,128,192}, the entries that keep q < 4).  For q >= 4 or q == 0 sub only reorders
 * entries whose contributions are identical.  Entries that contribute nothing (q_eff == 0) get key 0
 * and sink to the bottom together with the padding.  Descending key order reproduces the reference's
 * chain for every base exactly.  The class bits come from a shared-memory LUT indexed by
 * (sample, reference nt16, bs), sub from a 256-entry LUT: a handful of integer instructions per read-base.
 *
 * FAST TILES (every column of the tile holds <= ~128 read-bases, every staged base quality is in [1,63]; all of
 * WGS 30x/60x): everything is lane-per-column.  Each lane reads its own column's staged bytes four read-bases at a
 * time (aligned 32-bit shared loads + funnel shift), builds the keys two at a time with packed 16-bit arithmetic
 * (PRMT / VIMNMX.U16x2 / one shared 16x16 LUT lookup per read-base for the class bits) straight into the NW
 * registers the sorting network works on (padding behind the column's last read-base), sorts them with a generated odd-even merge-sort network
 * (ssn_sortnet.cuh: VIMNMX.U16x2 only, no shuffles), stores them over the dead staging area in walk order and
 * runs the chain.  With base qualities in [1,63] every read-base contributes, sub is 1 and the :165 promotion is
 * q_eff = max(q, 4), which keeps the walk branch-free apart from the group change.
 * OTHER TILES (panel depths, base quality 0 or >= 64 somewhere in the tile) take the generic path: warp-cooperative
 * bitonic sort of 1024-key runs in registers (SHFL + VIMNMX) and a per-lane k-way merge during the walk.
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include "ssn.h"
#include "ssn_sortnet.cuh"

#ifndef SSN_ROUND_MAGIC
#define SSN_ROUND_MAGIC 1 /* 1: (s+M)-M rounding in the FP64 pipe; 0: F2F.F32.F64 + F2F.F64.F32 */
#endif

#ifndef SSN_PHASE_LOCK
#define SSN_PHASE_LOCK 0   /* 1: the warps of a CTA take 16 consecutive tiles and step through the phases together (i-cache locality) */
#endif
#if SSN_PHASE_LOCK
#define SSN_PL_BAR() asm volatile("bar.sync 0;" ::: "memory")
#else
#define SSN_PL_BAR() do { } while (0)
#endif

namespace ssn {

#ifndef SSN_WARPS_PER_CTA
#define SSN_WARPS_PER_CTA 16
#endif
constexpr int kWarpsPerCta = SSN_WARPS_PER_CTA;   /* one CTA per SM: 16 x (keys 8 KB + slots 3 KB; the raw staging aliases both) + tables */
constexpr int kRunLen = 1024;          /* keys per sorted run of the generic path (R = 16 packed registers) */
constexpr int kMaxRuns = 64;           /* => at most 65536 read-bases per column (tumor + normal) */
constexpr int kSlotWords = 24 * 32;    /* per warp: 8 (sample,base) slots x {esum,fsum,count} x 32 lanes */
constexpr int kFastWords = 64;         /* largest NW of the fast path: 128 key slots per column */
constexpr int kRawBytes = 3200;        /* per array: staged read-bases of one fast tile (16-byte aligned window) */
constexpr int kAuxWords = 3 * kRawBytes / 4;   /* raw staging of a fast tile: aliases the key store and the head of the slots */
static_assert(kAuxWords <= kFastWords * 32 + kSlotWords, "the raw staging must fit in key store + slots");
constexpr int kListSlot = 160;         /* list mode: staged bytes of one gathered column (<= 128 read-bases + 2 x 15 alignment bytes) */
constexpr int kListWidth = 16;         /* list mode: at most this many columns per tile (kListWidth * kListSlot <= kRawBytes) */
static_assert(kListWidth * kListSlot <= kRawBytes, "gathered tiles must fit the staging area");
constexpr uint32_t kPadPair = 0xffff0000u;     /* an empty register of the sorting network: key 0 | ~key 0 << 16 */
constexpr unsigned kFull = 0xffffffffu;

/* static shared memory of a CTA: model tables + key LUTs */
struct __align__(1024) CtaTables {
    uint32_t lut3[16 * 16];      /* [ref nt16][read nt16] -> b4<<13 | valid<<4 | sub 1; 1 KB aligned: address = row | nt*4 */
    double2 fk2[256];            /* {fk[w], fk[w]/32}: the walk multiplies by q*32 taken straight from the key */
    unsigned long long mbar[kWarpsPerCta];   /* one mbarrier per warp for its bulk copies */
    int32_t qadd[1024 + 64];
    int32_t prior[160];
    uint16_t cls[2 * 16 * 32];   /* [sample][ref nt16][bs & 31] -> sample<<15 | b4<<13 | valid<<4 | strand<<3 */
    uint8_t sub[256];            /* [baseQ] -> sub */
    uint2 tail[5];               /* [live read-bases in a 4-chunk, clamped to 0..4] -> keep-masks of its two registers */
    uint32_t tile_base;
};

struct DevModel {
    const double *coef, *lhet, *fk;
    const int32_t *prior, *jointprior, *qadd;
    int32_t q_r_int, cap_mapq, min_somatic_qual, include_loh, include_gor, use_joint;
};

struct DevBatch {
    const uint32_t *off;
    const uint8_t *bs, *bq, *mq, *ref_nt16;
    uint32_t ncol;
    ssn_site *sites;
    uint32_t *site_count;   /* [0] = sites wanted, [1] = tile counter */
    uint32_t site_cap;
    uint32_t *coldump;      /* ncol * SSN_COLUMN_DUMP_WORDS or null */
    uint32_t *scratch;      /* per-warp global key store for tiles that exceed shared memory */
    uint32_t scratch_words_per_warp;
    uint32_t smem_words_per_warp;
    /* list mode (behind the pre-filter, ssn_prefilter.cuh): score columns collist[0 .. *list_count) instead of 0 .. ncol */
    const uint32_t *collist;
    const uint32_t *list_count;
    uint32_t list_width;    /* 0: pick the tile width from the list length; else 4, 8 or 16 */
    /* batches built on the device (ssn_pileup.cuh) are scored before the host knows their shape: ncol is then an upper
     * bound and the true count is read here; the kernels do nothing when the pileup's summary reports an error or a
     * column deeper than the launch assumed (the host then launches again with the real shape) */
    const uint32_t *dev_ncol;
    const uint32_t *guard;          /* pileup summary: [2] error bits, [4] key-store words of the deepest column */
    uint32_t guard_col_words;
    uint32_t list_generic;          /* 1: gathered tiles take the warp-cooperative sorting path (experiment: short code, one column per warp) */
    uint32_t no_hist;               /* 1: deep tiles take the sorting path (measurement / cross-check knob SSN_NO_HIST) */
};
constexpr uint32_t kGuardStop = 2u | 4u | 16u;   /* pileup::kErrStop */

/* -------------------------------------------------------------------------------------------- */
/* key construction (sniper_maqcns.c:144-156 + the :165 promotion rule folded in)              */

__device__ __forceinline__ void build_key_luts(CtaTables &T, int tid, int nthreads)
{
    for (int i = tid; i < 2 * 16 * 32; i += nthreads) {
        const uint32_t sample = i >> 9, ref16 = (i >> 5) & 15u, strand = (i >> 4) & 1u;
        uint32_t nt = i & 15u;
        nt = nt ? nt : ref16;                                               /* '=' -> reference base */
        const uint32_t valid = (nt != 0u && (nt & (nt - 1u)) == 0u) ? 1u : 0u;  /* bam_nt16_nt4_table < 4 */
        const uint32_t b4 = valid ? ((nt >> 1) - (nt >> 3)) : 0u;               /* 1,2,4,8 -> 0,1,2,3 */
        T.cls[i] = (uint16_t)(sample << 15 | b4 << 13 | valid << 4 | strand << 3);
    }
    for (int i = tid; i < 256; i += nthreads)
        T.sub[i] = (uint8_t)(1 + (i < 64 ? 0 : 2 * (i >> 6) - ((i & 63) == 0 ? 1 : 0)));
    for (int i = tid; i < 256; i += nthreads) {
        const uint32_t ref16 = (uint32_t)i >> 4;
        uint32_t nt = (uint32_t)i & 15u;
        nt = nt ? nt : ref16;
        const uint32_t valid = (nt != 0u && (nt & (nt - 1u)) == 0u) ? 1u : 0u;
        const uint32_t b4 = valid ? ((nt >> 1) - (nt >> 3)) : 0u;
        T.lut3[i] = b4 << 13 | valid << 4 | 1u;
    }
    if (tid < 5) {
        const uint32_t k[5][2] = {{0u, 0u}, {0xffffu, 0u}, {0xffffffffu, 0u}, {0xffffffffu, 0xffffu}, {0xffffffffu, 0xffffffffu}};
        T.tail[tid] = make_uint2(k[tid][0], k[tid][1]);
    }
}

/* cls_col = T.cls + (sample*16 + ref16)*32 */
__device__ __forceinline__ uint32_t make_key(const uint16_t *cls_col, const uint8_t *sub_lut, uint32_t bs, uint32_t bq, uint32_t mq)
{
    const uint32_t q = min(bq, mq);
    const uint32_t key = (uint32_t)cls_col[bs & 31u] + q * 32u + (uint32_t)sub_lut[bq];
    return (q | (bq & 63u)) ? key : 0u;        /* q == 0 and baseQ & 0x3f == 0: contributes nothing (:165,:167) */
}

/* -------------------------------------------------------------------------------------------- */
/* generic path: packed bitonic sort over the warp, R registers per lane, 2 keys per register.  */
/* half h, register r, lane l hold element e = r + R*l of half h; low halves sort ascending,    */
/* high halves hold complemented keys (so they sort descending in true order), then one merge.  */

template <int R, int K>   /* one bitonic merge level of block size K over the 32R elements of each half */
__device__ __forceinline__ void merge_level(uint32_t (&v)[R], int lane)
{
#pragma unroll
    for (int j = K >> 1; j > 0; j >>= 1) {
        if (j >= R) {
            const int jl = j / R, kl = K / R;
            const bool keep_min = ((lane & jl) == 0) == ((lane & kl) == 0);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                uint32_t o = __shfl_xor_sync(kFull, v[r], jl);
                v[r] = keep_min ? __vminu2(v[r], o) : __vmaxu2(v[r], o);
            }
        } else {
            const bool lane_asc = (K >= R) ? ((lane & (K / R)) == 0) : true;
#pragma unroll
            for (int r = 0; r < R; ++r) {
                if ((r & j) == 0) {
                    uint32_t a = v[r], b = v[r | j];
                    uint32_t lo = __vminu2(a, b), hi = __vmaxu2(a, b);
                    bool asc = (K >= R) ? lane_asc : ((r & K) == 0);
                    v[r] = asc ? lo : hi;
                    v[r | j] = asc ? hi : lo;
                }
            }
        }
    }
}

template <int R, int K>
struct SortLevels {
    __device__ static __forceinline__ void run(uint32_t (&v)[R], int lane)
    {
        SortLevels<R, K / 2>::run(v, lane);
        merge_level<R, K>(v, lane);
    }
};
template <int R>
struct SortLevels<R, 1> {
    __device__ static __forceinline__ void run(uint32_t (&)[R], int) {}
};

template <int R>
__device__ __forceinline__ void sort_packed(uint32_t (&v)[R], int lane)
{
    SortLevels<R, 32 * R>::run(v, lane);           /* both halves, lockstep */
#pragma unroll
    for (int r = 0; r < R; ++r) {                  /* un-complement, then the half<->half stage */
        uint32_t x = v[r] ^ 0xffff0000u;
        uint32_t sw = __byte_perm(x, x, 0x1032);
        uint32_t mn = __vminu2(x, sw), mx = __vmaxu2(x, sw);
        v[r] = __byte_perm(mn, mx, 0x7610);
    }
    merge_level<R, 32 * R>(v, lane);               /* finish the 64R-element merge inside each half */
}

/* Load one run of `len` (<= 64R) read-bases starting at entry `e0`, build keys, sort, and store the top
 * ceil(len/2) words (ascending) at store32[0..).  tumor_len = how many of the run's entries are tumor. */
template <int R>
__device__ __forceinline__ void sort_run(const DevBatch &B, const CtaTables &T, uint32_t e0, int len, int tumor_len, uint32_t ref16,
                                         uint32_t *store32, int lane, uint32_t &rms_t, uint32_t &rms_n, int cap_mapq)
{
    uint32_t v[R];
    const uint16_t *cls_t = T.cls + (16 + ref16) * 32, *cls_n = T.cls + ref16 * 32;
#pragma unroll
    for (int row = 0; row < 2 * R; ++row) {
        const int i = row * 32 + lane;
        uint32_t key = 0;
        if (i < len) {
            const uint32_t bs = B.bs[e0 + i], bq = B.bq[e0 + i], mq = B.mq[e0 + i];
            const bool is_t = i < tumor_len;
            key = make_key(is_t ? cls_t : cls_n, T.sub, bs, bq, mq);
            const uint32_t m = min(mq & 0x7fu, (uint32_t)cap_mapq);      /* sniper_maqcns.c:173-174 */
            if (is_t) rms_t += m * m; else rms_n += m * m;
        }
        const int r = row % R;
        if (row < R) v[r] = key;
        else v[r] |= (~key) << 16;
    }
    sort_packed<R>(v, lane);
    const int wfirst = (64 * R - len) >> 1;
    if (R == 1) {
        uint16_t *s16 = reinterpret_cast<uint16_t*>(store32);
        const int p0 = lane, p1 = 32 + lane;
        if (p0 >= 2 * wfirst) s16[p0 - 2 * wfirst] = (uint16_t)(v[0] & 0xffffu);
        if (p1 >= 2 * wfirst) s16[p1 - 2 * wfirst] = (uint16_t)(v[0] >> 16);
    } else {
#pragma unroll
        for (int r = 0; r < R; r += 2) {
            const int wi_lo = (lane * R + r) >> 1, wi_hi = 16 * R + wi_lo;
            const uint32_t lo = __byte_perm(v[r], v[(r + 1) % R], 0x5410);
            const uint32_t hi = __byte_perm(v[r], v[(r + 1) % R], 0x7632);
            if (wi_lo >= wfirst) store32[wi_lo - wfirst] = lo;
            if (wi_hi >= wfirst) store32[wi_hi - wfirst] = hi;
        }
    }
}

/* -------------------------------------------------------------------------------------------- */
/* the sequential chain (lane-per-column)                                                       */

/* round a positive double to float precision, staying in double: (s + 2^(e+29)) - 2^(e+29) rounds to
 * nearest-even at the float ulp of s's binade, i.e. equals (double)(float)s for normal floats. */
__device__ __forceinline__ double round_to_f32(double s)
{
#if SSN_ROUND_MAGIC != 0
    const int hi = __double2hiint(s);
    const double m = __hiloint2double((hi & 0x7ff00000) + 0x01d00000, 0);
    return __dsub_rn(__dadd_rn(s, m), m);
#else
    return (double)__double2float_rn(s);
#endif
}

/* the fsum chain: same rounding; SSN_ROUND_MAGIC == 2 sends it through the conversion pipe instead */
__device__ __forceinline__ double round_to_f32f(double s)
{
#if SSN_ROUND_MAGIC == 2
    return (double)__double2float_rn(s);
#else
    return round_to_f32(s);
#endif
}

struct Chain {
    double e, f;        /* float-valued accumulators held in double */
    uint32_t c, w0, w1; /* count; per-strand dependency ranks */
    int sb;             /* current (sample<<2 | base) slot, -1 = none */
};

__device__ __forceinline__ void chain_init(Chain &ch) { ch.e = 0.0; ch.f = 0.0; ch.c = 0; ch.w0 = 0; ch.w1 = 0; ch.sb = -1; }

__device__ __forceinline__ void chain_flush(const Chain &ch, uint32_t *slots, int lane)
{
    if (ch.sb >= 0) {
        uint32_t *s = slots + ch.sb * 96 + lane;
        s[0] = __float_as_uint(__double2float_rn(ch.e));
        s[32] = __float_as_uint(__double2float_rn(ch.f));
        s[64] = ch.c;
    }
}

__device__ __forceinline__ void chain_step(Chain &ch, uint32_t key, const double2 *fk_s, uint32_t *slots, int lane)
{
    const int sb = (int)(key >> 13);
    if (sb != ch.sb) {
        chain_flush(ch, slots, lane);
        ch.sb = sb; ch.e = 0.0; ch.f = 0.0; ch.c = 0; ch.w0 = 0; ch.w1 = 0;
    }
    uint32_t q = (key >> 5) & 0xffu;
    if (q < 4u) q = (q != 0u && !(key & 1u)) ? q : 4u;   /* sniper_maqcns.c:165; sub even <=> baseQ in {64,128,192} */
    const bool rev = (key >> 3) & 1u;
    const uint32_t w = rev ? ch.w1 : ch.w0;
    const double fkw = fk_s[w].x;
    const double qd = __hiloint2double(0x43300000, (int)q) - 4503599627370496.0;   /* exact int -> double */
    ch.e = round_to_f32(__dadd_rn(ch.e, __dmul_rn(fkw, qd)));      /* esum += fk[w]*q  (:168) */
    ch.f = round_to_f32(__dadd_rn(ch.f, fkw));                     /* fsum += fk[w]    (:169) */
    const uint32_t wn = min(w + 1u, 255u);                         /* :170 */
    if (rev) ch.w1 = wn; else ch.w0 = wn;
    ++ch.c;
}

/* -------------------------------------------------------------------------------------------- */
/* fast path: bulk-copied raw tile -> transposed key store -> per-lane sorting network -> walk   */

__device__ __forceinline__ uint32_t smem_u32(const void *p) { return (uint32_t)__cvta_generic_to_shared(p); }

/* One lane stages [a0, a0+bytes) of bs/bq/mq into the warp's raw area with three 1-D bulk copies (TMA unit)
 * that complete on the warp's mbarrier.  a0 and bytes are multiples of 16. */
__device__ __forceinline__ void stage_issue(const DevBatch &B, uint32_t raw_s, uint32_t mbar_s, uint32_t a0, uint32_t bytes)
{
    /* earlier generic-proxy accesses to the area (previous tile's slots) are ordered before the async-proxy writes */
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mbar_s), "r"(3u * bytes) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(raw_s), "l"(B.bs + a0), "r"(bytes), "r"(mbar_s) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(raw_s + (uint32_t)kRawBytes), "l"(B.bq + a0), "r"(bytes), "r"(mbar_s) : "memory");
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                 :: "r"(raw_s + 2u * (uint32_t)kRawBytes), "l"(B.mq + a0), "r"(bytes), "r"(mbar_s) : "memory");
}

__device__ __forceinline__ void stage_wait(uint32_t mbar_s, uint32_t parity)
{
    asm volatile("{\n\t.reg .pred p;\n\tWAIT_%=:\n\t"
                 "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
                 "@p bra DONE_%=;\n\tbra WAIT_%=;\n\tDONE_%=:\n\t}" :: "r"(mbar_s), "r"(parity) : "memory");
}

__device__ __forceinline__ uint32_t lds_u32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
__device__ __forceinline__ uint2 lds_u32x2(uint32_t a) { uint2 v; asm volatile("ld.shared.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "r"(a)); return v; }
__device__ __forceinline__ uint4 lds_u32x4(uint32_t a) { uint4 v; asm volatile("ld.shared.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "r"(a)); return v; }
__device__ __forceinline__ double2 lds_f64x2(uint32_t a) { double2 v; asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(v.x), "=d"(v.y) : "r"(a)); return v; }

/* Is every staged base quality of the tile in [1,63]?  (Then every read-base contributes, sub == 1 and the :165
 * promotion is max(q,4).)  Warp-cooperative scan of the staged bq bytes; slack bytes of neighbouring tiles only make
 * the answer conservative. */
__device__ __forceinline__ bool tile_is_clean(uint32_t raw_s, uint32_t bytes, int lane)
{
    uint32_t bad = 0;
    for (uint32_t i = (uint32_t)lane * 16u; i < bytes; i += 512u) {
        const uint4 w = lds_u32x4(raw_s + (uint32_t)kRawBytes + i);
        bad |= (w.x & 0xc0c0c0c0u) | ((w.x - 0x01010101u) & ~w.x & 0x80808080u);
        bad |= (w.y & 0xc0c0c0c0u) | ((w.y - 0x01010101u) & ~w.y & 0x80808080u);
        bad |= (w.z & 0xc0c0c0c0u) | ((w.z - 0x01010101u) & ~w.z & 0x80808080u);
        bad |= (w.w & 0xc0c0c0c0u) | ((w.w - 0x01010101u) & ~w.w & 0x80808080u);
    }
    return !__any_sync(kFull, bad != 0u);
}

/* The same for the gathered tiles of list mode: every lane checks the n staged base qualities of its own column
 * (bytes [r0, r0+n) of the staged bq array; W = tile-uniform word count). */
__device__ __forceinline__ bool column_is_clean(uint32_t bq_s, uint32_t r0, uint32_t n, int W)
{
    uint32_t bad = 0;
    const uint32_t a = bq_s + (r0 & ~3u);
    int lo = (int)(r0 & 3u) * 8, hi = lo + (int)n * 8;                        /* live bit range relative to the current word */
    for (int i = 0; i < W; ++i, lo -= 32, hi -= 32) {
        const uint32_t keep = __funnelshift_lc(0xffffffffu, 0u, (uint32_t)max(hi, 0)) & ~__funnelshift_lc(0xffffffffu, 0u, (uint32_t)max(lo, 0));
        const uint32_t w = (lds_u32(a + 4u * (uint32_t)i) & keep) | (0x01010101u & ~keep);
        bad |= (w & 0xc0c0c0c0u) | ((w - 0x01010101u) & ~w & 0x80808080u);
    }
    return !__any_sync(kFull, bad != 0u);
}

/* Keys of two consecutive read-bases (bytes SEL of the chunk words) -> one register of the sorting network:
 * lo = key of the first, hi = complemented key of the second; registers past the column's end become padding.
 * keep / tum = 0xffff per half that is a live read-base / a tumor read-base. */
template <int HALF>
__device__ __forceinline__ uint32_t t3_pair(uint32_t BS, uint32_t ST, uint32_t BQ, uint32_t MQ, uint32_t keep, uint32_t tum, uint32_t lutK,
                                            uint32_t cap2, uint32_t &rms_all, uint32_t &rms_t)
{
    constexpr uint32_t sel = HALF ? 0x4342u : 0x4140u;
    const uint32_t bq = __byte_perm(BQ, 0u, sel), mq = __byte_perm(MQ, 0u, sel);
    const uint32_t q = __vminu2(bq, mq);                                                   /* sniper_maqcns.c:151 */
    const uint32_t x0 = HALF ? (BS >> 14) : (BS << 2), x1 = HALF ? (BS >> 22) : (BS >> 6);
    const uint32_t l0 = lds_u32((x0 & 0x3cu) | lutK), l1 = lds_u32((x1 & 0x3cu) | lutK);
    const uint32_t t = __byte_perm(l0, l1, 0x5410u) + (q << 5) + __byte_perm(ST, 0u, sel) + (tum & 0x80008000u);
    const uint32_t mm = __vminu2(mq & 0x007f007fu & keep, cap2);                           /* :173-174 */
    const uint32_t mb = __byte_perm(mm, 0u, 0x4420u);
    rms_all = __dp2a_lo(mm, mb, rms_all);
    rms_t = __dp2a_lo(mm & tum, mb, rms_t);
    return ((t ^ kPadPair) & keep) | (kPadPair & ~keep);
}

/* FRONT of a fast tile, lane-per-column: the column's read-bases (tumor first, normal right behind them in the
 * staged bytes) become the keys of registers [0, ceil(n/2)), the rest is padding.  C = tile-uniform chunk count
 * (4 read-bases = 2 registers per chunk).  raw_s = shared address of the staged tile. */
template <int NW>
__device__ __forceinline__ void t3_build(uint32_t (&v)[NW], uint32_t raw_s, uint32_t r0, uint32_t nT, uint32_t n, int C,
                                         uint32_t lutK, uint32_t tail_s, int cap_mapq, uint32_t &rms_t, uint32_t &rms_n)
{
    const uint32_t cap2 = (uint32_t)cap_mapq * 0x00010001u;
    const uint32_t a = raw_s + (r0 & ~3u), sh = (r0 & 3u) * 8u;
    uint32_t ws = lds_u32(a), wq = lds_u32(a + kRawBytes), wm = lds_u32(a + 2 * kRawBytes);
    uint32_t rall = 0, rt = 0;
#pragma unroll
    for (int c = 0; c < NW / 2; ++c) {
        if (c < C) {
            const uint32_t ns = lds_u32(a + 4 * (c + 1)), nq = lds_u32(a + kRawBytes + 4 * (c + 1)), nm = lds_u32(a + 2 * kRawBytes + 4 * (c + 1));
            const uint32_t BS = __funnelshift_r(ws, ns, sh), BQ = __funnelshift_r(wq, nq, sh), MQ = __funnelshift_r(wm, nm, sh);
            const uint2 keep = lds_u32x2(tail_s + 8u * (uint32_t)min(max((int)n - 4 * c, 0), 4));
            const uint2 tum = lds_u32x2(tail_s + 8u * (uint32_t)min(max((int)nT - 4 * c, 0), 4));
            const uint32_t ST = (BS & 0x10101010u) >> 1;                                  /* strand -> key bit 3 */
            v[2 * c] = t3_pair<0>(BS, ST, BQ, MQ, keep.x, tum.x, lutK, cap2, rall, rt);
            v[2 * c + 1] = t3_pair<1>(BS, ST, BQ, MQ, keep.y, tum.y, lutK, cap2, rall, rt);
            ws = ns; wq = nq; wm = nm;
        } else {
            v[2 * c] = kPadPair; v[2 * c + 1] = kPadPair;
        }
    }
    rms_t = rt; rms_n = rall - rt;
}

/* sort this lane's column and lay the keys out for the walk (over the dead staging area) */
template <int NW>
__device__ __forceinline__ void t3_sort_store(uint32_t (&v)[NW], uint32_t *keys, int lane)
{
    SortNet<NW>::lists(v);
    SortNet<NW>::merge(v);
    __syncwarp();                                   /* every lane is done reading the staged bytes */
    /* ascending sequence lo[0..NW) ++ hi[0..NW), two keys per word, word k of lane l at keys[k*32 + l] */
#pragma unroll
    for (int k = 0; k < NW / 2; ++k) {
        keys[k * 32 + lane] = __byte_perm(v[2 * k], v[2 * k + 1], 0x5410);
        keys[(NW / 2 + k) * 32 + lane] = __byte_perm(v[2 * k], v[2 * k + 1], 0x7632);
    }
}

/* The walk of a fast tile (every key contributes, sub == 1).  Group changes (and the end: key 0) are detected by
 * one compare against the current group's floor; q*32 is taken from the key unshifted and multiplied by fk/32 (exact
 * scaling); the two strands' dependency ranks share one register (rank < 256: a fast column has <= 128 keys). */
struct Walk {
    double e, f;
    uint32_t ww, floor;
    int sb, start;
};

__device__ __forceinline__ void walk_flush(const Walk &ch, int pos, uint32_t *slots, int lane)
{
    if (ch.sb >= 0) {
        uint32_t *s = slots + ch.sb * 96 + lane;
        s[0] = __float_as_uint(__double2float_rn(ch.e));
        s[32] = __float_as_uint(__double2float_rn(ch.f));
        s[64] = (uint32_t)(ch.start - pos);
    }
}

/* pos = position of the key in the descending walk (counts down); returns false when the column's keys are exhausted */
__device__ __forceinline__ bool walk_step(Walk &ch, uint32_t key, int pos, uint32_t fk2_s, uint32_t *slots, int lane)
{
    if (key < ch.floor) {
        walk_flush(ch, pos, slots, lane);
        ch.sb = -1;
        if (key == 0u) return false;
        ch.sb = (int)(key >> 13); ch.floor = max(key & 0xe000u, 1u);   /* group 0's real keys are >= 1 */
        ch.e = 0.0; ch.f = 0.0; ch.ww = 0; ch.start = pos;
    }
    const uint32_t q32 = max(key & 0x1fe0u, 128u);                     /* sniper_maqcns.c:165 with baseQ in [1,63]: q_eff = max(q,4) */
    const uint32_t sh = key & 8u;                                      /* strand bit doubles as the rank's shift */
    const uint32_t w = (ch.ww >> sh) & 0xffu;
    ch.ww += 1u << sh;                                                 /* :170 (no saturation needed below 256) */
    const double2 fk = lds_f64x2(fk2_s + w * 16u);
    const double qd = __hiloint2double(0x43300000, (int)q32) - 4503599627370496.0;   /* exact int -> double */
    ch.e = round_to_f32(__dadd_rn(ch.e, __dmul_rn(fk.y, qd)));        /* esum += fk[w]*q  (:168) */
    ch.f = round_to_f32f(__dadd_rn(ch.f, fk.x));                      /* fsum += fk[w]    (:169) */
    return true;
}

template <int NW>
__device__ __forceinline__ void fast_walk(const uint32_t *keys, uint32_t fk2_s, uint32_t *slots, int lane)
{
    Walk ch;
    ch.e = 0.0; ch.f = 0.0; ch.ww = 0; ch.floor = 0x10000u; ch.sb = -1; ch.start = 0;
    int pos = 2 * NW - 1;
#pragma unroll 1
    for (int k = NW - 1; k >= 0; --k, pos -= 2) {
        const uint32_t wd = keys[k * 32 + lane];
        if (!walk_step(ch, wd >> 16, pos, fk2_s, slots, lane)) break;
        if (!walk_step(ch, wd & 0xffffu, pos - 1, fk2_s, slots, lane)) break;
    }
    walk_flush(ch, -1, slots, lane);
}

/* -------------------------------------------------------------------------------------------- */
/* likelihoods, quantisation, consensus for one sample (lane-local)                             */

struct Glf {
    uint32_t lk[3];   /* 10 bytes packed: lk[i>>2] >> 8*(i&3) */
    uint32_t min_lk, rms, cns;
};

__device__ __forceinline__ uint32_t glf_lk(const Glf &g, int i) { return (g.lk[i >> 2] >> (8 * (i & 3))) & 0xffu; }

__device__ __forceinline__ int bar_error(float t1, float t3)
{
    int bar = (int)((double)__fdiv_rn(t1, t3) + 0.5);    /* sniper_maqcns.c:191-193 */
    return min(max(bar, 4), 63);
}

#ifndef SSN_GLF_NOINLINE
#define SSN_GLF_NOINLINE 0
#endif
#if SSN_GLF_NOINLINE
__device__ __noinline__
#else
__device__
#endif
void sample_glf(const DevModel &M, const uint32_t *slots, int lane, int sample, uint32_t n,
                           uint32_t rms_sum, Glf &g)
{
    float es[4], fs[4];
    int c[4];
#pragma unroll
    for (int b = 0; b < 4; ++b) {
        const uint32_t *s = slots + (sample * 4 + b) * 96 + lane;
        es[b] = __uint_as_float(s[0]); fs[b] = __uint_as_float(s[32]); c[b] = (int)s[64];
    }
    g.rms = (uint32_t)(uint8_t)(int)(sqrt((double)rms_sum / (double)(int)n) + .499);   /* :176 */
    int tot = c[0] + c[1] + c[2] + c[3];
    if (tot > 255) {                                                                  /* :179-182 */
#pragma unroll
        for (int b = 0; b < 4; ++b) c[b] = (int)(254.0 * (double)c[b] / (double)tot + 0.5);
        tot = c[0] + c[1] + c[2] + c[3];
    }
    /* gather the 10 terms' table indices first so the 16 loads are in flight together */
    float t1v[10];
    int cidx[10], hidx[10];
    {
        int g_i = 0;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
#pragma unroll
            for (int k = j; k < 4; ++k) {
                float t1 = 0.0f, t3 = 0.0f;
                int t2 = 0;
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    if (i == j || i == k) continue;
                    t1 = __fadd_rn(t1, es[i]); t2 += c[i]; t3 = __fadd_rn(t3, fs[i]);
                }
                t1v[g_i] = t1;
                cidx[g_i] = t2 ? (bar_error(t1, t3) << 16 | tot << 8 | t2) : -1;
                hidx[g_i] = (j == k) ? -1 : (c[j] << 8 | c[k]);
                ++g_i;
            }
        }
    }
    double cf[10], lh[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        cf[i] = cidx[i] >= 0 ? __ldg(M.coef + cidx[i]) : 0.0;
        lh[i] = hidx[i] >= 0 ? __ldg(M.lhet + hidx[i]) : 0.0;
    }
    float p[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        float v;
        if (hidx[i] < 0) {                                            /* homozygous :186-197 */
            v = cidx[i] >= 0 ? __double2float_rn(__dadd_rn((double)t1v[i], cf[i])) : 0.0f;
        } else {                                                      /* heterozygous :199-210 */
            const double base = __dmul_rn(-4.343, lh[i]);
            v = cidx[i] >= 0 ? __double2float_rn(__dadd_rn(__dadd_rn(base, (double)t1v[i]), cf[i]))
                             : __double2float_rn(base);
        }
        p[i] = v < 0.0f ? 0.0f : v;                                   /* :212-213 */
    }
    /* the p[k,k] correction, :216-233. diagonal positions in GLF order: 0,4,7,9 */
    {
        float max1 = -1.0f, max2 = -1.0f, min1 = 1e30f, min2 = 1e30f;
        int max_k = -1, min_k = -1;
        const float d[4] = {p[0], p[4], p[7], p[9]};
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (es[k] > max1) { max2 = max1; max1 = es[k]; max_k = k; }
            else if (es[k] > max2) max2 = es[k];
        }
#pragma unroll
        for (int k = 0; k < 4; ++k) {
            if (d[k] < min1) { min2 = min1; min1 = d[k]; min_k = k; }
            else if (d[k] < min2) min2 = d[k];
        }
        if (max1 > max2 && (min_k != max_k || (double)min1 + 1.0 > (double)min2)) {
            const float nv = min1 > 1.0f ? __double2float_rn((double)min1 - 1.0) : 0.0f;
            if (max_k == 0) p[0] = nv;
            else if (max_k == 1) p[4] = nv;
            else if (max_k == 2) p[7] = nv;
            else p[9] = nv;
        }
    }
    /* quantisation :235-244 */
    float min_p = 1e30f;
#pragma unroll
    for (int i = 0; i < 10; ++i) min_p = p[i] < min_p ? p[i] : min_p;
    g.min_lk = min_p > 255.0f ? 255u : (uint32_t)(int)((double)min_p + 0.5);
    g.lk[0] = g.lk[1] = g.lk[2] = 0;
    int lkv[10];
#pragma unroll
    for (int i = 0; i < 10; ++i) {
        const float dlt = __fsub_rn(p[i], min_p);
        lkv[i] = dlt > 255.0f ? 255 : (int)((double)dlt + 0.5);
        g.lk[i >> 2] |= (uint32_t)lkv[i] << (8 * (i & 3));
    }
    /* consensus, sniper_glf2cns :250-273: scan order i<<2|j over i<=j == GLF order */
    {
        int mn = 10000, mn2 = 10000, mn3 = 10000, g1 = -1, g2 = -1;
        int gi = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
#pragma unroll
            for (int j = i; j < 4; ++j) {
                const int t = lkv[gi] + (i == j ? 0 : M.q_r_int);
                const int code = (1 << i) | (1 << j);
                if (t < mn) { mn3 = mn2; mn2 = mn; mn = t; g2 = g1; g1 = code; }
                else if (t < mn2) { mn3 = mn2; mn2 = t; g2 = code; }
                else if (t < mn3) mn3 = t;
                ++gi;
            }
        }
        uint32_t x = (uint32_t)(g1 >= 0 ? g1 : 0xf) << 28;
        x |= (uint32_t)(g2 >= 0 ? g2 : 0xf) << 24;
        x |= g.rms << 16;
        x |= (uint32_t)(mn2 < 10000 ? min(mn2 - mn, 255) : 0xff) << 8;
        x |= (uint32_t)((mn2 < 10000 && mn3 < 10000) ? min(mn3 - mn2, 255) : 0xff);
        g.cns = x;
    }
}

/* -------------------------------------------------------------------------------------------- */
/* somatic score (lane-local, rare branch)                                                      */

__device__ __forceinline__ int qadd(const int32_t *qt, int x, int y) { return x + qt[512 + y - x]; }

__device__ __forceinline__ int genotype_mask_of(int i)
{
    /* glfBase, somatic_sniper.c:26 */
    return (int)((0x8C4A629531ull >> (4 * (unsigned)i)) & 0xf);
}

struct Call {
    int score, t_vaq, n_vaq, jT, jN, jgq, tg, ng;
    bool candidate, pass;
};

__device__ void score_column(const DevModel &M, const int32_t *qadd_s, const int32_t *prior_s, int rb4,
                             const Glf &gt, const Glf &gn, Call &out)
{
    const int tb1 = gt.cns >> 28, tb2 = (gt.cns >> 24) & 0xf, ts1 = (gt.cns >> 8) & 0xff, ts2 = gt.cns & 0xff;
    const int nb1 = gn.cns >> 28, nb2 = (gn.cns >> 24) & 0xf, ns1 = (gn.cns >> 8) & 0xff, ns2 = gn.cns & 0xff;
    out.score = 255; out.t_vaq = 0; out.n_vaq = 0; out.jT = 0; out.jN = 0; out.jgq = 255;
    out.tg = tb1; out.ng = nb1; out.pass = false;
    out.candidate = rb4 != 15 && tb1 != 15 && nb1 != 15 && tb1 != nb1;             /* somatic_sniper.c:156 */
    if (!out.candidate) return;
    out.t_vaq = min(tb2 == rb4 ? ts1 : ts1 + ts2, 255);                             /* :157-164 */
    if (nb1 != rb4) out.n_vaq = min(nb2 == rb4 ? ns1 : ns1 + ns2, 255);
    int score = 255;
    if (M.use_joint) {                                                              /* :170-208 */
        int marg = 255, best = 1000, bi = -1, bj = -1;
        const int32_t *jp = M.jointprior + rb4 * 100;
        for (int i = 0; i < 10; ++i) {
            const int ln = (int)glf_lk(gn, i);
            for (int j = 0; j < 10; ++j) {
                int v = min(ln + (int)glf_lk(gt, j) + __ldg(jp + i * 10 + j), 255);
                if (v < best) { best = v; bi = i; bj = j; }
                marg = qadd(qadd_s, marg, v);
            }
        }
        int jgq = 255;
        for (int j = 0; j < 10; ++j) {
            const int v = min((int)glf_lk(gn, j) + (int)glf_lk(gt, j) + __ldg(jp + j * 11), 255);
            const int l = v - marg;
            score = qadd(qadd_s, score, l);
            if (j != bj) jgq = qadd(qadd_s, jgq, l);     /* the reference's other test is always true (:196) */
        }
        out.jgq = min(jgq, 255);
        out.jN = genotype_mask_of(bi); out.jT = genotype_mask_of(bj);
    } else {                                                                        /* :79-99, :209-214 */
        int post_t[10], post_n[10], qs_t = 255, qs_n = 255;
        const int32_t *pr = prior_s + rb4 * 10;
#pragma unroll
        for (int j = 0; j < 10; ++j) {
            post_t[j] = (int)glf_lk(gt, j) + pr[j]; qs_t = qadd(qadd_s, post_t[j], qs_t);
            post_n[j] = (int)glf_lk(gn, j) + pr[j]; qs_n = qadd(qadd_s, post_n[j], qs_n);
        }
#pragma unroll
        for (int j = 0; j < 10; ++j) {
            const int y = min(post_t[j] - qs_t, 255) + min(post_n[j] - qs_n, 255);
            score = qadd(qadd_s, score, y);
        }
    }
    out.score = score;
    const int tg = out.jT ? out.jT : tb1, ng = out.jN ? out.jN : nb1;               /* :216-223 */
    out.tg = tg; out.ng = ng;
    const bool loh = ng != tg && (tg & ng) == tg;                                   /* allele_util.c:19-22 */
    const bool gor = !(ng != rb4 && (rb4 & ng) == rb4) && (tg & ~ng) == rb4;        /* allele_util.c:25-28 */
    out.pass = M.min_somatic_qual <= score && (M.include_loh || !loh) && (M.include_gor || !gor);
}

/* -------------------------------------------------------------------------------------------- */
/* warp-cooperative site statistics for one sample of one column (dqstats.c:6-53)               */

__device__ void site_stats(const DevBatch &B, uint32_t e0, int n, int rb4, uint32_t wanted, int lane, uint32_t *st /*18, smem*/)
{
    uint32_t sum_mq = 0, dp4[4] = {0, 0, 0, 0}, occ[4] = {0, 0, 0, 0}, sbq[4] = {0, 0, 0, 0}, smq[4] = {0, 0, 0, 0};
    for (int i = lane; i < n; i += 32) {
        const uint32_t bs = B.bs[e0 + i], bq = B.bq[e0 + i], mq = B.mq[e0 + i];
        const uint32_t base = bs & 15u, strand = (bs >> 4) & 1u;
        sum_mq += mq;
        const uint32_t d = ((int)base == rb4 ? 0u : 2u) + strand;
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            dp4[j] += d == (uint32_t)j;
            if ((base & (1u << j)) == base) {
                ++occ[j];
                if (wanted & (1u << j)) { sbq[j] += bq; smq[j] += mq; }
            }
        }
    }
    sum_mq = __reduce_add_sync(kFull, sum_mq);
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        dp4[j] = __reduce_add_sync(kFull, dp4[j]); occ[j] = __reduce_add_sync(kFull, occ[j]);
        sbq[j] = __reduce_add_sync(kFull, sbq[j]); smq[j] = __reduce_add_sync(kFull, smq[j]);
    }
    if (lane == 0) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            st[j] = occ[j] ? (uint32_t)((double)sbq[j] / (double)occ[j] + .499) : sbq[j];
            st[4 + j] = occ[j] ? (uint32_t)((double)smq[j] / (double)occ[j] + .499) : smq[j];
            st[8 + j] = occ[j]; st[12 + j] = dp4[j];
        }
        st[16] = (uint32_t)n;
        st[17] = n > 0 ? (uint32_t)((double)sum_mq / (double)n + .499) : 0u;
    }
}

/* -------------------------------------------------------------------------------------------- */

__device__ __forceinline__ int run_R(int len) { return len <= 64 ? 1 : len <= 128 ? 2 : len <= 256 ? 4 : len <= 512 ? 8 : 16; }

/* generic FRONT + walk: any depth up to kRunLen*kMaxRuns per column */
__device__ __noinline__ void generic_tile(const DevModel &M, const DevBatch &B, const CtaTables &T, uint32_t *slots, uint32_t *keys_s,
                                          uint32_t *keys_g, int lane, uint32_t o0, uint32_t nT, uint32_t n, uint32_t ref16,
                                          uint32_t &my_rms_t, uint32_t &my_rms_n)
{
    /* words of key store this column needs: full runs take kRunLen/2, the tail ceil(len/2) */
    const uint32_t nruns = (n + kRunLen - 1) / kRunLen;
    const uint32_t words = (n >> 10) * (kRunLen / 2) + (((n & (kRunLen - 1)) + 1) >> 1);
    uint32_t incl = words;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        uint32_t t = __shfl_up_sync(kFull, incl, d);
        if (lane >= d) incl += t;
    }
    const uint32_t wbase = incl - words;
    const uint32_t tile_words = __shfl_sync(kFull, incl, 31);
    uint32_t *store = tile_words <= B.smem_words_per_warp ? keys_s : keys_g;

    for (int j = 0; j < 32; ++j) {
        const uint32_t nj = __shfl_sync(kFull, n, j);
        if (nj == 0) continue;
        const uint32_t e0 = __shfl_sync(kFull, o0, j);
        const int nTj = (int)__shfl_sync(kFull, nT, j);
        const uint32_t refj = __shfl_sync(kFull, ref16, j);
        uint32_t *dst = store + __shfl_sync(kFull, wbase, j);
        uint32_t rt = 0, rn = 0;
        for (uint32_t done = 0; done < nj; done += kRunLen) {
            const int len = (int)min(nj - done, (uint32_t)kRunLen);
            const int tl = nTj - (int)done;
            switch (run_R(len)) {
            case 1: sort_run<1>(B, T, e0 + done, len, tl, refj, dst, lane, rt, rn, M.cap_mapq); break;
            case 2: sort_run<2>(B, T, e0 + done, len, tl, refj, dst, lane, rt, rn, M.cap_mapq); break;
            case 4: sort_run<4>(B, T, e0 + done, len, tl, refj, dst, lane, rt, rn, M.cap_mapq); break;
            case 8: sort_run<8>(B, T, e0 + done, len, tl, refj, dst, lane, rt, rn, M.cap_mapq); break;
            default: sort_run<16>(B, T, e0 + done, len, tl, refj, dst, lane, rt, rn, M.cap_mapq); break;
            }
            dst += (len + 1) >> 1;
        }
        rt = __reduce_add_sync(kFull, rt);
        rn = __reduce_add_sync(kFull, rn);
        if (lane == j) { my_rms_t = rt; my_rms_n = rn; }
    }
    __syncwarp();

    Chain ch;
    chain_init(ch);
    const uint16_t *k16 = reinterpret_cast<const uint16_t*>(store + wbase);
    if (nruns <= 1) {
        for (int idx = 2 * (int)words - 1; idx >= 0; --idx) {
            const uint32_t key = k16[idx];
            if (key == 0) break;
            chain_step(ch, key, T.fk2, slots, lane);
        }
    } else {
        /* k-way merge of the sorted runs (deep columns): heads live in local memory */
        int hp[kMaxRuns];
        uint16_t hk[kMaxRuns];
        for (uint32_t r = 0; r < nruns; ++r) {
            const uint32_t len = min(n - r * kRunLen, (uint32_t)kRunLen);
            hp[r] = (int)(r * kRunLen + 2 * ((len + 1) >> 1)) - 1;
            hk[r] = k16[hp[r]];
        }
        for (;;) {
            uint32_t best = 0, br = 0;
            for (uint32_t r = 0; r < nruns; ++r)
                if (hk[r] > best) { best = hk[r]; br = r; }
            if (best == 0) break;
            chain_step(ch, best, T.fk2, slots, lane);
            const int np = hp[br] - 1;
            hp[br] = np;
            hk[br] = np >= (int)(br * kRunLen) ? k16[np] : (uint16_t)0;
        }
    }
    chain_flush(ch, slots, lane);
}

/* -------------------------------------------------------------------------------------------- */
/* deep columns (panel depths): a histogram instead of a sort              */
/*
 * What the sorted order is used for: inside one (sample, base) chain the reference touches the read-bases by falling
 * q, then valid, then strand (sniper_maqcns.c:157-176), and two read-bases that agree in (q, valid, strand) contribute
 * the same amount.  With every base quality in [1,63] (checked here; otherwise the tile takes the sorting path) q =
 * min(baseQ, mapQ) <= 63, every read-base contributes and sub == 1, so a column's keys fall into
 *     bin = sample<<10 | b4<<8 | q<<2 | valid<<1 | strand            (2048 bins)
 * and walking the non-empty bins from the top down, each as many times as it holds read-bases, IS the reference's chain.
 * FRONT, warp per column: the bins are counted in shared memory (over the key store; lanes holding the same bin are
 * found with MATCH.ANY and their leader adds the group's size, so a bin hit by the whole warp costs one atomic), then
 * the non-empty bins are compacted, top down, into a run list (bin<<17 | count) in the warp's global scratch.
 * BACK, lane per column: one flat loop of max(n) steps; a lane fetches its next run when the current one is used up
 * and otherwise does exactly chain_step's arithmetic.  No key is loaded per step and no sorted run is merged.
 * Returns false (nothing written that matters) when the tile has to take the sorting path instead.
 */
constexpr int kHistBins = 2048;

__device__ __noinline__ bool hist_tile(const DevModel &M, const DevBatch &B, const CtaTables &T, uint32_t *slots, uint32_t *hist,
                                       uint32_t *runs, uint32_t runs_cap, int lane, uint32_t o0, uint32_t nT, uint32_t n, uint32_t ref16,
                                       uint32_t &my_rms_t, uint32_t &my_rms_n)
{
    uint32_t my_start = 0, my_count = 0, total = 0;
    const uint32_t lt = (1u << lane) - 1u;
    /* the batch descriptor sits in local memory behind the reference: take what the loops need into registers */
    const uint8_t *__restrict__ g_bs = B.bs, *__restrict__ g_bq = B.bq, *__restrict__ g_mq = B.mq;
    const uint32_t cap = (uint32_t)M.cap_mapq;
    for (int j = 0; j < 32; ++j) {
        const uint32_t nj = __shfl_sync(kFull, n, j);
        if (nj == 0) continue;
        const uint32_t e0 = __shfl_sync(kFull, o0, j), nTj = __shfl_sync(kFull, nT, j), refj = __shfl_sync(kFull, ref16, j);
        const uint16_t *cls_t = T.cls + (16 + refj) * 32, *cls_n = T.cls + refj * 32;
        for (int i = lane * 4; i < kHistBins; i += 128) *reinterpret_cast<uint4*>(hist + i) = make_uint4(0u, 0u, 0u, 0u);
        __syncwarp();
        uint32_t rt = 0, rn = 0, bad = 0, g_lo = 0, g_hi = 0;      /* g_*: groups of 32 bins this lane's atomics touched */
        /* One aligned 32-bit word of each array per lane and iteration (four read-bases; 128 per warp), the next
         * iteration's words already in flight while this one's are counted.  The column starts `lead` bytes into its
         * first word; bytes in front of it and behind its end are masked (the arrays carry 64 spare bytes behind the
         * batch).  The body is branch-free, so the four MATCH.ANY run converged, back to back. */
        const uint32_t a0 = e0 & ~3u, lead = e0 - a0, span = lead + nj;
        uint32_t ws = 0u, wq = 0u, wm = 0u;
        {
            const uint32_t wb = 4u * (uint32_t)lane;
            if (wb < span) {
                ws = __ldg(reinterpret_cast<const uint32_t*>(g_bs + a0 + wb));
                wq = __ldg(reinterpret_cast<const uint32_t*>(g_bq + a0 + wb));
                wm = __ldg(reinterpret_cast<const uint32_t*>(g_mq + a0 + wb));
            }
        }
        for (uint32_t w0 = 0; w0 < span; w0 += 128u) {
            const uint32_t wb = w0 + 4u * (uint32_t)lane, nb = wb + 128u;
            const uint32_t cs = ws, cq = wq, cm = wm;
            if (nb < span) {
                ws = __ldg(reinterpret_cast<const uint32_t*>(g_bs + a0 + nb));
                wq = __ldg(reinterpret_cast<const uint32_t*>(g_bq + a0 + nb));
                wm = __ldg(reinterpret_cast<const uint32_t*>(g_mq + a0 + nb));
            }
            uint32_t bin[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const uint32_t idx = wb + (uint32_t)k, i = idx - lead;
                const uint32_t bs = (cs >> (8 * k)) & 0xffu, bq = (cq >> (8 * k)) & 0xffu, mq = (cm >> (8 * k)) & 0xffu;
                const bool valid = idx >= lead && idx < span, is_t = i < nTj;
                const uint32_t cls = (is_t ? cls_t : cls_n)[bs & 31u];                  /* sample<<15 | b4<<13 | valid<<4 | strand<<3 */
                bad |= (valid && (bq - 1u) > 62u) ? 1u : 0u;
                const uint32_t v = (cls >> 15) << 10 | ((cls >> 13) & 3u) << 8 | (min(bq, mq) & 63u) << 2 | ((cls >> 3) & 3u);
                bin[k] = valid ? v : 0xffffffffu;
                const uint32_t m = valid ? min(mq & 0x7fu, cap) : 0u;                   /* sniper_maqcns.c:173-174 */
                rt += is_t ? m * m : 0u;
                rn += is_t ? 0u : m * m;
            }
            uint32_t peers[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) peers[k] = __match_any_sync(kFull, bin[k]);
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                if (bin[k] != 0xffffffffu && (peers[k] & lt) == 0u) {
                    atomicAdd(hist + bin[k], (uint32_t)__popc(peers[k]));
                    const uint32_t gbit = 1u << ((bin[k] >> 5) & 31u);
                    if (bin[k] & 1024u) g_hi |= gbit; else g_lo |= gbit;
                }
            }
        }
        if (__any_sync(kFull, bad != 0u)) return false;
        rt = __reduce_add_sync(kFull, rt);
        rn = __reduce_add_sync(kFull, rn);
        g_lo = __reduce_or_sync(kFull, g_lo);
        g_hi = __reduce_or_sync(kFull, g_hi);
        __syncwarp();
        /* compaction, top down, over the groups of 32 bins that hold anything */
        uint32_t cnt = 0;
        for (int half = 1; half >= 0; --half) {
            uint32_t groups = half ? g_hi : g_lo;
            while (groups) {
                const int gidx = 31 - __clz((int)groups);
                groups &= ~(1u << gidx);
                const uint32_t bin = (uint32_t)(half * 1024 + gidx * 32 + (31 - lane));
                const uint32_t c = hist[bin];
                const uint32_t mask = __ballot_sync(kFull, c != 0u);
                if (total + cnt + (uint32_t)__popc(mask) > runs_cap) return false;  /* more runs than the scratch holds */
                if (c) runs[total + cnt + (uint32_t)__popc(mask & lt)] = bin << 17 | c;  /* c <= 65536 */
                cnt += (uint32_t)__popc(mask);
            }
        }
        if (lane == j) { my_start = total; my_count = cnt; my_rms_t = rt; my_rms_n = rn; }
        total += cnt;
        __syncwarp();
    }
    __syncwarp();
    /* BACK: lane-per-column walk over the runs */
    Chain ch;
    chain_init(ch);
    const uint32_t *mine = runs + my_start;
    const uint32_t steps = __reduce_max_sync(kFull, n);
    uint32_t left = 0, k = 1, rev = 0;
    uint32_t nxt = my_count ? mine[0] : 0u;                 /* the next run is fetched a whole run ahead of its use */
    double qd = 0.0;
    for (uint32_t it = 0; it < steps; ++it) {
        if (it < n) {
            if (left == 0u) {
                const uint32_t r = nxt;
                nxt = k < my_count ? mine[k] : 0u;
                ++k;
                const uint32_t bin = r >> 17;
                left = r & 0x1ffffu;
                const int sb = (int)(bin >> 8);
                if (sb != ch.sb) {
                    chain_flush(ch, slots, lane);
                    ch.sb = sb; ch.e = 0.0; ch.f = 0.0; ch.c = 0; ch.w0 = 0; ch.w1 = 0;
                }
                const uint32_t q = max((bin >> 2) & 63u, 4u);                       /* :165 with baseQ in [1,63] */
                qd = __hiloint2double(0x43300000, (int)q) - 4503599627370496.0;     /* exact int -> double */
                rev = bin & 1u;
            }
            const uint32_t w = rev ? ch.w1 : ch.w0;
            const double fkw = T.fk2[w].x;
            ch.e = round_to_f32(__dadd_rn(ch.e, __dmul_rn(fkw, qd)));               /* esum += fk[w]*q  (:168) */
            ch.f = round_to_f32(__dadd_rn(ch.f, fkw));                              /* fsum += fk[w]    (:169) */
            const uint32_t wn = min(w + 1u, 255u);                                  /* :170 */
            if (rev) ch.w1 = wn; else ch.w0 = wn;
            ++ch.c;
            --left;
        }
    }
    chain_flush(ch, slots, lane);
    return true;
}

template <int NW>
__device__ __noinline__ void fast_tile(uint32_t raw_s, uint32_t r0, uint32_t nT, uint32_t n, int C, uint32_t lutK, uint32_t tail_s,
                                       int cap_mapq, uint32_t *keys_s, uint32_t *slots, uint32_t fk2_s, int lane,
                                       uint32_t &my_rms_t, uint32_t &my_rms_n)
{
    uint32_t v[NW];
    t3_build<NW>(v, raw_s, r0, nT, n, C, lutK, tail_s, cap_mapq, my_rms_t, my_rms_n);
    SSN_PL_BAR();
    t3_sort_store<NW>(v, keys_s, lane);
    for (int i = 0; i < 24; ++i) slots[i * 32 + lane] = 0;      /* the staging area is dead: its tail becomes the slots */
    __syncwarp();
    SSN_PL_BAR();
    fast_walk<NW>(keys_s, fk2_s, slots, lane);
}

template <bool DUMP>
__global__ void __launch_bounds__(kWarpsPerCta * 32, 1)
ssn_score_kernel(const DevModel M, const DevBatch B)
{
    __shared__ CtaTables T;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    for (int i = threadIdx.x; i < 256; i += blockDim.x) {
        const double f = M.fk[i];
        T.fk2[i] = make_double2(f, f * 0.03125);          /* exact power-of-two scaling */
    }
    for (int i = threadIdx.x; i < 1024 + 64; i += blockDim.x) T.qadd[i] = M.qadd[i];
    for (int i = threadIdx.x; i < 160; i += blockDim.x) T.prior[i] = M.prior[i];
    build_key_luts(T, threadIdx.x, blockDim.x);
    const uint32_t mbar_s = smem_u32(&T.mbar[warp]);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mbar_s) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    /* per warp: [key store: smem_words_per_warp words][chain slots: kSlotWords]; the raw staging of a fast tile
     * (kAuxWords) aliases the region from its start and is dead before keys and slots are written */
    uint32_t *keys_s = reinterpret_cast<uint32_t*>(smem_raw) + (size_t)warp * (kSlotWords + B.smem_words_per_warp);
    uint32_t *slots = keys_s + B.smem_words_per_warp;
    const uint32_t raw_s = smem_u32(keys_s);
    const uint32_t fk2_s = smem_u32(T.fk2), lut_s = smem_u32(T.lut3), tail_s = smem_u32(T.tail);
    const uint32_t gwarp = blockIdx.x * kWarpsPerCta + warp;
    uint32_t *keys_g = B.scratch ? B.scratch + (size_t)gwarp * B.scratch_words_per_warp : nullptr;
    /* list mode: few columns, gathered from anywhere in the batch.  Narrow tiles (W columns per warp) spread them over
     * every resident warp; the lanes past W idle.  Gathered columns are not contiguous, so they take the generic path. */
    const bool listed = B.collist != nullptr;
    if (B.guard && ((B.guard[2] & kGuardStop) || B.guard[4] > B.guard_col_words)) return;
    const uint32_t batch_ncol = B.dev_ncol ? min(*B.dev_ncol, B.ncol) : B.ncol;
    uint32_t ncols = batch_ncol, W = 32u;
    if (listed) {
        ncols = min(*B.list_count, batch_ncol);
        const uint32_t warps = gridDim.x * (uint32_t)kWarpsPerCta, per = (ncols + warps - 1u) / warps;
        W = 4u;
        while (W < per && W < (uint32_t)kListWidth) W <<= 1;
        if (B.list_width) W = B.list_width;
    }
    const uint32_t ntiles = (ncols + W - 1u) / W;
    uint32_t phase = 0;

    for (;;) {
        uint32_t tile = 0;
#if SSN_PHASE_LOCK
        __syncthreads();
        if (threadIdx.x == 0) T.tile_base = atomicAdd(B.site_count + 1, (uint32_t)kWarpsPerCta);
        __syncthreads();
        if (T.tile_base >= ntiles) break;
        tile = T.tile_base + (uint32_t)warp;
#else
        if (lane == 0) tile = atomicAdd(B.site_count + 1, 1u);
        tile = __shfl_sync(kFull, tile, 0);
        if (tile >= ntiles) break;
#endif
        const uint32_t idx = tile * W + lane;
        const bool live = (uint32_t)lane < W && idx < ncols;
        const uint32_t col = listed ? (live ? B.collist[idx] : 0u) : idx;
        uint32_t o0 = 0, o1 = 0, o2 = 0, ref16 = 15;
        if (live) {
            o0 = B.off[2 * col]; o1 = B.off[2 * col + 1]; o2 = B.off[2 * col + 2];
            ref16 = B.ref_nt16[col] & 15u;
        }
        const uint32_t nT = o1 - o0, nN = o2 - o1, n = nT + nN;
        const uint32_t need = __reduce_max_sync(kFull, n);
        /* the tile's read-bases are contiguous: [off of its first column, end of its last live column) */
        const uint32_t a0 = __shfl_sync(kFull, o0, 0) & ~15u;
        const uint32_t bytes = (__reduce_max_sync(kFull, o2) - a0 + 15u) & ~15u;
        const bool fast = need <= 2u * kFastWords && (listed || bytes <= (uint32_t)kRawBytes) && B.smem_words_per_warp >= kFastWords * 32u &&
                          !(listed && B.list_generic);

        uint32_t my_rms_t = 0, my_rms_n = 0;
        bool done = false;
        __syncwarp();                                   /* the previous tile is done with slots / keys */
        if (need != 0u && fast) {
            const int C = (int)((need + 3u) >> 2);       /* tile-uniform chunk count (4 read-bases = 2 registers per chunk) */
            uint32_t r0 = 0;
            bool clean;
            if (!listed) {
                if (lane == 0) stage_issue(B, raw_s, mbar_s, a0, bytes);
                stage_wait(mbar_s, phase);
                phase ^= 1u;
                clean = tile_is_clean(raw_s, bytes, lane);
                r0 = live ? o0 - a0 : 0u;
            } else {
                /* gathered tile: every lane bulk-copies its own column's 16-byte aligned window into its slot */
                const uint32_t s0 = o0 & ~15u, sz = n ? ((o2 - s0 + 15u) & ~15u) : 0u;
                const uint32_t total = __reduce_add_sync(kFull, sz);
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                if (lane == 0) asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(mbar_s), "r"(3u * total) : "memory");
                __syncwarp();
                r0 = sz ? (uint32_t)lane * (uint32_t)kListSlot + (o0 & 15u) : 0u;   /* idle lanes read (and mask) slot 0 */
                if (sz) {
                    const uint32_t dst = raw_s + (uint32_t)lane * (uint32_t)kListSlot;
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(dst), "l"(B.bs + s0), "r"(sz), "r"(mbar_s) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(dst + (uint32_t)kRawBytes), "l"(B.bq + s0), "r"(sz), "r"(mbar_s) : "memory");
                    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
                                 :: "r"(dst + 2u * (uint32_t)kRawBytes), "l"(B.mq + s0), "r"(sz), "r"(mbar_s) : "memory");
                }
                stage_wait(mbar_s, phase);
                phase ^= 1u;
                clean = column_is_clean(raw_s + (uint32_t)kRawBytes, r0, n, (int)((need + 6u) >> 2));
            }
            if (clean) {
                const uint32_t lutK = lut_s + ref16 * 64u;
                if (C <= 8) fast_tile<16>(raw_s, r0, nT, n, C, lutK, tail_s, M.cap_mapq, keys_s, slots, fk2_s, lane, my_rms_t, my_rms_n);
                else if (C <= 16) fast_tile<32>(raw_s, r0, nT, n, C, lutK, tail_s, M.cap_mapq, keys_s, slots, fk2_s, lane, my_rms_t, my_rms_n);
                else fast_tile<64>(raw_s, r0, nT, n, C, lutK, tail_s, M.cap_mapq, keys_s, slots, fk2_s, lane, my_rms_t, my_rms_n);
                done = true;
            }
        }
        if (!done) {
            __syncwarp();
            for (int i = 0; i < 24; ++i) slots[i * 32 + lane] = 0;
            __syncwarp();
            SSN_PL_BAR();
            if (need != 0u) {
                /* deep tiles count bins instead of sorting, when the launch has a global scratch for the run lists and
                 * every base quality is in [1,63]; anything else sorts */
                bool counted = false;
                if (keys_g && need > 2u * kFastWords && B.smem_words_per_warp >= (uint32_t)kHistBins && !B.no_hist)
                    counted = hist_tile(M, B, T, slots, keys_s, keys_g, B.scratch_words_per_warp, lane, o0, nT, n, ref16, my_rms_t, my_rms_n);
                if (!counted) {
                    if (keys_g && need > 2u * kFastWords) {         /* the attempt may have left partial chains behind */
                        __syncwarp();
                        for (int i = 0; i < 24; ++i) slots[i * 32 + lane] = 0;
                        my_rms_t = 0; my_rms_n = 0;
                        __syncwarp();
                    }
                    generic_tile(M, B, T, slots, keys_s, keys_g, lane, o0, nT, n, ref16, my_rms_t, my_rms_n);
                }
            }
            SSN_PL_BAR();
        }
        SSN_PL_BAR();
        __syncwarp();

        Glf gt, gn;
        gt.lk[0] = gt.lk[1] = gt.lk[2] = 0; gt.min_lk = 0; gt.rms = 0; gt.cns = 0xff000000u;
        gn = gt;
        if (nT) sample_glf(M, slots, lane, 1, nT, my_rms_t, gt);
        if (nN) sample_glf(M, slots, lane, 0, nN, my_rms_n, gn);
        SSN_PL_BAR();
        Call call;
        call.score = -1; call.pass = false; call.candidate = false;
        if (nT && nN && ref16 != 15u) score_column(M, T.qadd, T.prior, (int)ref16, gt, gn, call);

        if (DUMP && live) {
            uint32_t *o = B.coldump + (size_t)col * SSN_COLUMN_DUMP_WORDS;
#pragma unroll
            for (int i = 0; i < 10; ++i) { o[i] = glf_lk(gt, i); o[14 + i] = glf_lk(gn, i); }
            o[10] = gt.min_lk; o[11] = gt.rms; o[12] = min(nT, 16777215u); o[13] = gt.cns;
            o[24] = gn.min_lk; o[25] = gn.rms; o[26] = min(nN, 16777215u); o[27] = gn.cns;
            o[28] = (uint32_t)call.score;
        }

        /* ---------------- EMIT: compact passing sites (warp ballot), stats warp-cooperatively ---------------- */
        uint32_t pass_mask = __ballot_sync(kFull, call.pass);
        if (pass_mask) {
            uint32_t base_slot = 0;
            if (lane == 0) base_slot = atomicAdd(B.site_count, (uint32_t)__popc(pass_mask));
            base_slot = __shfl_sync(kFull, base_slot, 0);
            uint32_t rank = 0;
            while (pass_mask) {
                const int j = __ffs(pass_mask) - 1;
                pass_mask &= pass_mask - 1;
                uint32_t *rec = slots;   /* 54-word staging area, chains are done with it */
                __syncwarp();
                const uint32_t e0 = __shfl_sync(kFull, o0, j), e1 = __shfl_sync(kFull, o1, j);
                const int nTj = (int)__shfl_sync(kFull, nT, j), nNj = (int)__shfl_sync(kFull, nN, j);
                const int rb4 = (int)__shfl_sync(kFull, ref16, j);
                const uint32_t wanted = (uint32_t)(rb4 | __shfl_sync(kFull, call.tg, j) | __shfl_sync(kFull, call.ng, j));
                site_stats(B, e0, nTj, rb4, wanted, lane, rec + 9);
                site_stats(B, e1, nNj, rb4, wanted, lane, rec + 9 + 26);
                if (lane == j) {
                    const int tb1 = gt.cns >> 28, nb1 = gn.cns >> 28;
                    rec[0] = col; rec[1] = (uint32_t)rb4;
                    uint32_t *t = rec + 2, *nn = rec + 28;
                    t[0] = tb1; t[1] = call.jT; t[2] = call.jgq; t[3] = (gt.cns >> 8) & 0xff; t[4] = call.t_vaq;
                    t[5] = (uint32_t)call.score;
                    const bool loh = call.ng != call.tg && (call.tg & call.ng) == call.tg;
                    t[6] = call.tg == call.ng ? 1u : loh ? 3u : call.score > 0 ? 2u : 4u;   /* somatic_sniper.c:241-248 */
                    t[25] = gt.rms;
                    nn[0] = nb1; nn[1] = call.jN; nn[2] = call.jgq; nn[3] = (gn.cns >> 8) & 0xff; nn[4] = call.n_vaq;
                    nn[5] = (uint32_t)-1;
                    nn[6] = nb1 == rb4 ? 0u : 1u;                                            /* :258-261 */
                    nn[25] = gn.rms;
                }
                __syncwarp();
                const uint32_t slot = base_slot + rank++;
                if (slot < B.site_cap) {
                    uint32_t *dst = reinterpret_cast<uint32_t*>(B.sites + slot);
                    for (int i = lane; i < (int)(sizeof(ssn_site) / 4); i += 32) dst[i] = rec[i];
                }
            }
            __syncwarp();
        }
    }
}

}  // namespace ssn
