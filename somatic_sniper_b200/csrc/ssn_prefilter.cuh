/*
 * ssn_prefilter.cuh — the exact hom-reference pre-filter kernel (sm_100a).  DESIGN.md §4b has the derivation.
 *
 * glf_somatic emits nothing for a column whose two consensus genotypes agree (somatic_sniper.c:156).  This kernel
 * streams over every column of a batch once (TMA bulk-staged tiles of 32 columns, lane-per-column, packed byte
 * arithmetic: ~9 instructions per read-base instead of the ~95 of the full model) and PROVES, for most columns, that
 * sniper_glf2cns (sniper_maqcns.c:250-273) calls the homozygous reference genotype in both samples.  Those columns are
 * finished.  Every other column - real variants, noisy columns, anything the bounds cannot decide, columns with more than
 * 255 read-bases in a sample, every column of a staging pass that holds a base quality >= 64 - is appended to the
 * survivor list and goes through the full model (ssn_score_kernel in list mode).  Tiles deeper than the staging area are
 * staged in passes of consecutive columns; read-bases of quality 0 (invisible to the model) are masked out.  Results are identical to scoring
 * every column; SSN_SUBMIT_NO_PREFILTER switches the filter off, SSN_SUBMIT_COLUMN_DUMP records its decision next to the
 * full model's output so tests can audit it.
 *
 * Per sample the filter needs (all from one pass over bs/bq/mq):
 *   n      read-bases with a non-zero quality m   of those, the ones whose nt16 code differs from the reference's
 *   S      sum over those m of base quality    n1  reference read-bases on the reverse strand
 *   Q1,Q0  sum of min(baseQ, mapQ & 0x7f) over the reference read-bases of the reverse / forward strand
 * and evaluates (tables: ssn_prefilter_tables.c)
 *   t1 = F[n1] * Q1 / (n1 + m) + F[n0] * Q0 / (n0 + m)            lower bound of the reference base's esum chain
 *   SU = (S + 3 m) (1 + 1e-4)                                      upper bound of the other bases' esum chains
 *   U  = max(min(SU + cmax_run[n][m], vt[n][m][ceil SU]), 0) + 1.1 upper bound of p[RR], plus the margin RR must win by
 *   keep scoring unless  t1 > SU  and  t1 + cmin_all[n][m] >= U  and  h[n-m][m] + g2_run[n][m] >= U   (both samples)
 */
#pragma once
#include "ssn_kernels.cuh"

namespace ssn {

#ifndef SSN_PF_WARPS
#define SSN_PF_WARPS 23
#endif
constexpr int kPfWarps = SSN_PF_WARPS;                 /* 23 x 9.6 KB of staging per SM */
constexpr int kPfStageBytes = 3 * kRawBytes;

struct DevPrefilter {
    const float4 *ab;     /* [n*256 + m] = { cmin_all, cmax_run, g2_run, 0 } */
    const float *h;       /* [(n-m)*256 + m] */
    const float *F;       /* 257 */
    const float *vt;      /* [(n*8 + m-1)*256 + ceil(SU)]: tighter upper bound of p[RR] for m <= 8, SU <= 255 */
};

struct PfShared {
    float F[260];
    unsigned long long mbar[kPfWarps];
};

/* 0xff in every byte whose bit 7 is set, else 0x00: PRMT with the sign-replication bit of each selector nibble */
__device__ __forceinline__ uint32_t sign_bytes(uint32_t x)
{
    uint32_t r;
    asm("prmt.b32 %0, %1, %1, 0xba98;" : "=r"(r) : "r"(x));
    return r;
}

struct PfStats { uint32_t m, S, n1, qref, q1, z; };

/* What the staged base qualities of a pass look like: 0 = all in [1,63]; 1 = some are 0 (those read-bases contribute
 * nothing to the model, sniper_maqcns.c:165-167: the filter ignores them and lowers n); 2 = some are >= 64 (baseQ 64/128/
 * 192 keep q < 4 unpromoted, :165): not filtered.  Slack bytes of the neighbouring tiles only make the answer conservative. */
__device__ __forceinline__ int pass_quality_class(uint32_t raw_s, uint32_t bytes, int lane)
{
    uint32_t hi = 0, zero = 0;
    for (uint32_t i = (uint32_t)lane * 16u; i < bytes; i += 512u) {
        const uint4 w = lds_u32x4(raw_s + (uint32_t)kRawBytes + i);
        hi |= (w.x | w.y | w.z | w.w) & 0xc0c0c0c0u;
        zero |= ((w.x - 0x01010101u) & ~w.x) | ((w.y - 0x01010101u) & ~w.y) | ((w.z - 0x01010101u) & ~w.z) | ((w.w - 0x01010101u) & ~w.w);
    }
    if (__any_sync(kFull, hi != 0u)) return 2;
    return __any_sync(kFull, (zero & 0x80808080u) != 0u) ? 1 : 0;
}

struct PfAcc { uint32_t m, n1, S, qref, q1, z; };

/* four read-bases of one sample: BSr/BQr/MQr = one word of each array, keep = live-byte mask (MASKED) */
template <bool ZEROS, bool MASKED>
__device__ __forceinline__ void pf_chunk(uint32_t BSr, uint32_t BQr, uint32_t MQr, uint32_t keep, uint32_t R4, PfAcc &A)
{
    const uint32_t ones = 0x01010101u;
    if (ZEROS) {                                                                /* quality-0 read-bases: dead bytes, counted */
        const uint32_t nzq = sign_bytes(BQr + 0x7f7f7f7fu);                     /* every quality is < 64 here */
        A.z = __dp4a(keep & ~nzq, ones, A.z);
        keep &= nzq;
    }
    /* dead bytes become "reference base, forward strand, quality 0": they add nothing to any sum */
    const uint32_t BS = MASKED ? ((BSr & keep) | (R4 & ~keep)) : BSr;
    const uint32_t BQ = MASKED ? (BQr & keep) : BQr;
    const uint32_t M7 = MQr & 0x7f7f7f7fu;
    /* byte masks by sign replication (PRMT): NM = 0xff where the nt16 code differs from the reference's,
     * RV = 0xff where a reference read-base sits on the reverse strand */
    const uint32_t NM = sign_bytes(((BS ^ R4) & 0x0f0f0f0fu) + 0x7f7f7f7fu);
    const uint32_t RV = sign_bytes(BS << 3) & ~NM;
    A.m = __dp4a(NM, ones, A.m);                                                /* 255 per non-reference read-base */
    A.n1 = __dp4a(RV, ones, A.n1);
    A.S = __dp4a(BQ & NM, ones, A.S);
    /* q = min(baseQ, mapQ & 0x7f) per byte: baseQ <= 63, so (baseQ + 0x80 - m7) keeps bit 7 iff baseQ >= m7 */
    const uint32_t GM = sign_bytes(BQ + 0x80808080u - M7);
    const uint32_t q = ((M7 & GM) | (BQ & ~GM)) & ~NM;                         /* reference read-bases only */
    A.qref = __dp4a(q, ones, A.qref);
    A.q1 = __dp4a(q & RV, ones, A.q1);
}

/* One sample of this lane's column: len staged read-bases starting at byte r of each staged array (len = 0 in lanes
 * outside the pass: mine = false).  The sums do not care where in a word a read-base sits, and the three arrays share
 * their offsets, so the words are taken as they lie — no re-alignment: the sample's first word and its last ones are
 * masked, the words that are completely live in every column of the pass are not.  R4 = the reference's nt16 code in
 * every byte. */
template <bool ZEROS>
__device__ __forceinline__ void pf_sample(uint32_t raw_s, uint32_t r, uint32_t len, bool mine, uint32_t R4, PfStats &st)
{
    const uint32_t a = raw_s + (r & ~3u), h = r & 3u, total = h + len;
    const int C = (int)__reduce_max_sync(kFull, mine ? (total + 3u) >> 2 : 0u);
    const int Cfull = ZEROS ? 0 : (int)__reduce_min_sync(kFull, mine ? total >> 2 : 0xffffu);
    PfAcc A = {0u, 0u, 0u, 0u, 0u, 0u};
    {
        const uint32_t keep = __funnelshift_lc(0xffffffffu, 0u, min(total, 4u) * 8u) & ~__funnelshift_lc(0xffffffffu, 0u, h * 8u);
        pf_chunk<ZEROS, true>(lds_u32(a), lds_u32(a + kRawBytes), lds_u32(a + 2 * kRawBytes), keep, R4, A);
    }
    int c = 1;
    if (!ZEROS) {
#pragma unroll 2
        for (; c < Cfull; ++c)
            pf_chunk<false, false>(lds_u32(a + 4 * c), lds_u32(a + kRawBytes + 4 * c), lds_u32(a + 2 * kRawBytes + 4 * c), 0xffffffffu, R4, A);
    }
    uint32_t bits = total * 8u;                    /* live bits from word c on */
    bits = max(bits, 32u * (uint32_t)c) - 32u * (uint32_t)c;
#pragma unroll 2
    for (; c < C; ++c) {
        const uint32_t keep = __funnelshift_lc(0xffffffffu, 0u, bits);          /* low min(bits,32) bits */
        bits = max(bits, 32u) - 32u;
        pf_chunk<ZEROS, true>(lds_u32(a + 4 * c), lds_u32(a + kRawBytes + 4 * c), lds_u32(a + 2 * kRawBytes + 4 * c), keep, R4, A);
    }
    st.m = A.m / 255u; st.n1 = A.n1 / 255u; st.z = A.z / 255u;
    st.S = A.S; st.qref = A.qref; st.q1 = A.q1;
}

/* true: genotype RR provably wins sniper_glf2cns for this sample */
__device__ __forceinline__ bool pf_eval(const DevPrefilter &P, const float *F, uint32_t n, const PfStats &st)
{
    const uint32_t m = st.m, n1 = st.n1, n0 = n - m - n1;
    const float4 ab = __ldg(P.ab + n * 256u + m);
    const float hs = __ldg(P.h + (n - m) * 256u + m);
    const float t1 = __fdividef(F[n1] * (float)st.q1, (float)max(n1 + m, 1u)) + __fdividef(F[n0] * (float)(st.qref - st.q1), (float)max(n0 + m, 1u));
    const float SU = (float)(st.S + 3u * m) * 1.0001f;
    float U = fmaxf(SU + ab.y, 0.0f) + 1.1f;
    if (m - 1u < 8u && SU <= 255.0f)                                       /* few other read-bases: the tabulated bound is tighter */
        U = fminf(U, __ldg(P.vt + ((n * 8u + m - 1u) << 8) + (uint32_t)__float2int_ru(SU)) + 1.1f);
    return t1 > SU && t1 + ab.x >= U && hs + ab.z >= U;
}

/* counters: [2] = survivors appended, [3] = tile counter.  decision (optional, stride words per column): 1 = dropped. */
__global__ void __launch_bounds__(kPfWarps * 32, 1)
ssn_prefilter_kernel(const DevPrefilter P, const DevBatch B, uint32_t *survivors, uint32_t *counters, uint32_t *decision, uint32_t decision_stride)
{
    __shared__ PfShared T;
    extern __shared__ __align__(128) unsigned char smem_raw[];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    if (B.guard && (B.guard[2] & kGuardStop)) return;                           /* the device pileup refused the batch */
    for (int i = threadIdx.x; i < 257; i += blockDim.x) T.F[i] = P.F[i];
    const uint32_t mbar_s = smem_u32(&T.mbar[warp]);
    if (lane == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(mbar_s) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();
    const uint32_t raw_s = smem_u32(smem_raw + (size_t)warp * kPfStageBytes);
    const uint32_t ncol = B.dev_ncol ? min(*B.dev_ncol, B.ncol) : B.ncol;      /* device-built batch: the count lives on the device */
    const uint32_t ntiles = (ncol + 31u) >> 5;
    uint32_t phase = 0;

    /* software pipeline over tiles: the tile index for the tile after next is being claimed (atomic in flight) and the
     * next tile's offsets are being loaded while the current tile is staged and evaluated */
    struct Hdr { uint32_t o0, o1, o2, ref16; };
    auto load_hdr = [&](uint32_t t, Hdr &h) {
        const uint32_t c = t * 32u + lane;
        h.o0 = h.o1 = h.o2 = 0u; h.ref16 = 15u;
        if (t < ntiles && c < ncol) {
            h.o0 = B.off[2 * c]; h.o1 = B.off[2 * c + 1]; h.o2 = B.off[2 * c + 2];
            h.ref16 = B.ref_nt16[c] & 15u;
        }
    };
    uint32_t tile = 0, claim = 0;
    if (lane == 0) tile = atomicAdd(counters + 3, 1u);
    tile = __shfl_sync(kFull, tile, 0);
    Hdr cur;
    load_hdr(tile, cur);
    if (lane == 0) claim = atomicAdd(counters + 3, 1u);

    while (tile < ntiles) {
        const uint32_t col = tile * 32u + lane;
        const bool live = col < ncol;
        const uint32_t o0 = cur.o0, o1 = cur.o1, o2 = cur.o2, ref16 = cur.ref16;
        const uint32_t nT = o1 - o0, nN = o2 - o1;
        /* a column can only reach the output with read-bases in both samples and a called reference (somatic_sniper.c:127,156) */
        const bool wanted = live && nT != 0u && nN != 0u && ref16 != 15u;
        const bool cand = wanted && nT <= 255u && nN <= 255u && __popc(ref16) == 1;   /* the bound tables cover n <= 255 */
        bool drop = false;
        __syncwarp();
        const uint32_t next_tile = __shfl_sync(kFull, claim, 0);      /* claimed one iteration ago */
        Hdr nxt;
        load_hdr(next_tile, nxt);
        if (lane == 0) claim = atomicAdd(counters + 3, 1u);
        /* The tile is staged in passes of consecutive columns whose read-bases fit the staging area: one pass at WGS
         * 30x/60x, two or three at 100x; the lanes outside the current pass idle. */
        uint32_t first = 0;
        while (__any_sync(kFull, cand && (uint32_t)lane >= first)) {
            const uint32_t a0 = __shfl_sync(kFull, o0, (int)first) & ~15u;
            const bool fits = (uint32_t)lane >= first && live && o2 - a0 <= (uint32_t)kRawBytes - 16u;
            const uint32_t run = ~(__ballot_sync(kFull, fits) >> first);            /* offsets rise with the lane: a prefix fits */
            const uint32_t cnt = run ? (uint32_t)__ffs((int)run) - 1u : 32u - first;
            if (cnt == 0u) { ++first; continue; }                                   /* one column deeper than the staging area */
            const uint32_t last = first + cnt;
            const bool mine = cand && (uint32_t)lane >= first && (uint32_t)lane < last;
            if (__any_sync(kFull, mine)) {
                const uint32_t bytes = (__shfl_sync(kFull, o2, (int)last - 1) - a0 + 15u) & ~15u;
                if (lane == 0) stage_issue(B, raw_s, mbar_s, a0, bytes);
                const uint32_t lT = mine ? nT : 0u, lN = mine ? nN : 0u;
                stage_wait(mbar_s, phase);
                phase ^= 1u;
                const int qc = pass_quality_class(raw_s, bytes, lane);
                if (qc < 2) {
                    const uint32_t r0 = mine ? o0 - a0 : 0u;
                    const uint32_t R4 = ref16 * 0x01010101u;
                    PfStats sT, sN;
                    /* words that are fully live in every column of the pass need no masks; the lanes outside the pass
                     * run them on whatever is staged and their sums are thrown away below */
                    if (qc == 0) {
                        pf_sample<false>(raw_s, r0, lT, mine, R4, sT);
                        pf_sample<false>(raw_s, r0 + lT, lN, mine, R4, sN);
                    } else {
                        pf_sample<true>(raw_s, r0, lT, mine, R4, sT);
                        pf_sample<true>(raw_s, r0 + lT, lN, mine, R4, sN);
                    }
                    if (!mine) { sT.m = sT.S = sT.n1 = sT.qref = sT.q1 = sT.z = 0u; sN = sT; }
                    /* both samples' table lookups are issued together (no short-circuit): lT/lN are 0 outside the pass */
                    const bool eT = pf_eval(P, T.F, lT - sT.z, sT), eN = pf_eval(P, T.F, lN - sN.z, sN);
                    if (mine) drop = eT && eN;
                }
                __syncwarp();                       /* every lane is done with the staged bytes before the next pass overwrites them */
            }
            first = last;
        }
        const bool survive = wanted && !drop;
        const uint32_t mask = __ballot_sync(kFull, survive);
        if (mask) {
            uint32_t base = 0;
            if (lane == 0) base = atomicAdd(counters + 2, (uint32_t)__popc(mask));
            base = __shfl_sync(kFull, base, 0);
            if (survive) survivors[base + (uint32_t)__popc(mask & ((1u << lane) - 1u))] = col;
        }
        if (decision && live) decision[(size_t)col * decision_stride] = drop ? 1u : 0u;
        tile = next_tile; cur = nxt;
    }
}

}  // namespace ssn
