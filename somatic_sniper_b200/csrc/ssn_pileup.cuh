/*
 * ssn_pileup.cuh — reads -> pileup columns on the device (sm_100a).
 *
 * Builds, from the filtered reads of both samples filed under position windows, the same column batch the host
 * packer writes (csrc/host/ssn_pack.c; layout of ssn_batch_view in include/ssn.h), entry for entry: a read adds a
 * read-base at every position of a CIGAR M run, D and N runs skip positions, I and S skip query bases, anything else
 * advances nothing (get_next_pos / resolve_cigar, src/lib/sniper/sniper_pileup.c:57-104,186-205; QUIRK B6).  A
 * column exists where both samples hold a read-base and the reference base is callable (somatic_sniper.c:127,156).
 * Within a column the read-bases keep the order of the reads in the window, tumor before normal.
 *
 *   count    one CTA per (window, sample): +1/-1 marks of the M runs in shared memory, block scan -> depth per position
 *   columns  one CTA per window: live columns, exclusive scans of columns and entries inside the window
 *   bases    one warp: exclusive scan over the windows, totals, capacity check
 *   offsets  one thread per position: off[], ref_nt16[], column -> position map, absolute cursors
 *   fill     one warp per 32 positions of a window, a lane per position: walks the reads that can cover them in file
 *            order (binary search on the sorted starts, bounded below by the longest reference span of the window),
 *            appends each lane's read-bases at its column's cursor in a shared-memory image of the 32 columns, and
 *            writes that out in 16-byte chunks
 *   shape    one thread per 32-column tile: the key-store footprint the scoring kernel's launch needs
 * No atomics decide an order anywhere, so the batch is bit-identical run to run and to the host packer's.
 */
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "ssn.h"

namespace pileup {

constexpr uint32_t kWin = SSN_READ_WINDOW;
constexpr uint32_t kDead = 0xffffffffu;
constexpr int kThreads = 256;
constexpr int kScanThreads = 1024;           /* count / columns: one CTA per window (and sample), so make it a big one */
constexpr int kRecWords = 5;                 /* sizeof(ssn_read_rec) / 4 */
constexpr int kSummaryWords = 8;             /* [0] columns [1] entries [2] error bits [3] max tile words [4] max column words */
constexpr uint32_t kErrUnsorted = 1u, kErrColumns = 2u, kErrEntries = 4u, kErrDepth = 8u, kErrRecord = 16u;
constexpr uint32_t kErrStop = kErrColumns | kErrEntries | kErrRecord;    /* nothing after the count kernel may run */

struct DevReads {
    const uint32_t *recs;
    const uint32_t *rec_off;
    const ssn_read_window *win;
    const uint8_t *wref;
    uint32_t n_win, n_reads, rec_words;
    /* work arrays of the slot */
    uint32_t *A;          /* [n_win][2][kWin]: depth -> entry cursor of the sample at that position (kDead: no column) */
    uint32_t *P;          /* [n_win][kWin]: column index inside the window */
    int32_t *rbeg;        /* [n_reads] start of every read (binary searches of the fill) */
    uint4 *rinfo;         /* [n_reads][2] {record word offset, start, l_qseq, n_cigar|mapq<<16|strand<<24}, {cigar[0], end from the CIGAR, -, -} */
    uint32_t *wspan;      /* [n_win][2] longest reference span */
    uint2 *wtot;          /* [n_win] {columns, entries} of the window, then {first column, first entry} */
    uint32_t *summary;
    /* the column batch being built */
    uint32_t *off; uint8_t *bs, *bq, *mq, *ref_nt16;
    uint32_t *colpos;
    uint32_t max_columns, max_entries;
    uint32_t run_len, max_runs;   /* scoring kernel's sort-run geometry (shape kernel) */
};

__device__ __forceinline__ uint32_t warp_incl_scan(uint32_t v, int lane)
{
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
        const uint32_t t = __shfl_up_sync(0xffffffffu, v, d);
        if (lane >= d) v += t;
    }
    return v;
}

/* Exclusive scan of kWin values held in shared memory (in place) by a CTA of kScanThreads threads; returns the total.
 * Each warp owns a contiguous chunk and walks it 32 values at a time (conflict-free). */
__device__ __forceinline__ uint32_t block_excl_scan(uint32_t *v, uint32_t *wsum)
{
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    constexpr int kWarps = kScanThreads / 32;
    constexpr uint32_t chunk = kWin / kWarps;
    uint32_t sum = 0;
    for (uint32_t j = 0; j < chunk; j += 32) sum += v[warp * chunk + j + lane];
#pragma unroll
    for (int d = 16; d > 0; d >>= 1) sum += __shfl_xor_sync(0xffffffffu, sum, d);
    if (lane == 0) wsum[warp] = sum;
    __syncthreads();
    uint32_t carry = 0, total = 0;
    for (int k = 0; k < kWarps; ++k) { if (k < warp) carry += wsum[k]; total += wsum[k]; }
    for (uint32_t j = 0; j < chunk; j += 32) {
        const uint32_t x = v[warp * chunk + j + lane];
        const uint32_t inc = warp_incl_scan(x, lane);
        v[warp * chunk + j + lane] = carry + inc - x;
        carry += __shfl_sync(0xffffffffu, inc, 31);
    }
    __syncthreads();
    return total;
}

__global__ void __launch_bounds__(kScanThreads) count_kernel(DevReads R)
{
    __shared__ uint32_t diff[kWin];
    __shared__ uint32_t wsum[kScanThreads / 32];
    __shared__ uint32_t span_s, err_s;
    const uint32_t w = blockIdx.x >> 1, s = blockIdx.x & 1;
    const ssn_read_window W = R.win[w];
    const uint32_t first = W.rec_first[s], n = W.n_reads[s], lo = W.lo, hi = lo + kWin;
    for (uint32_t i = threadIdx.x; i < kWin; i += kScanThreads) diff[i] = 0;
    if (threadIdx.x == 0) { span_s = 0; err_s = 0; }
    __syncthreads();
    for (uint32_t r = threadIdx.x; r < n; r += kScanThreads) {
        /* the index and the records come from the caller: a record must lie inside the buffer before anything reads it */
        const uint32_t ro = R.rec_off[first + r];
        if (ro > R.rec_words || R.rec_words - ro < (uint32_t)kRecWords) { atomicOr(&err_s, 2u); continue; }
        const uint32_t *rec = R.recs + ro;
        const int32_t beg = (int32_t)rec[0];
        const uint32_t n_cigar = rec[3] & 0xffffu;
        {
            const unsigned long long need = (unsigned long long)kRecWords + n_cigar + ((((unsigned long long)rec[2] + 1u) >> 1) + rec[2] + 3u) / 4u;
            if (need > R.rec_words - ro) { atomicOr(&err_s, 2u); continue; }
        }
        uint32_t x = (uint32_t)beg;
        for (uint32_t k = 0; k < n_cigar; ++k) {
            const uint32_t c = rec[kRecWords + k], op = c & 0xfu, l = c >> 4;
            if (op == 0u) {
                const uint32_t a = max(x, lo), b = min(x + l, hi);
                if (a < b) {                                   /* same marks as count_read (ssn_pack.c) */
                    atomicAdd(&diff[a - lo], 1u);
                    if (b < hi) atomicAdd(&diff[b - lo], 0xffffffffu);
                }
                x += l;
            } else if (op == 2u || op == 3u) x += l;
        }
        R.rbeg[first + r] = beg;
        R.rinfo[2 * (size_t)(first + r)] = make_uint4(R.rec_off[first + r], (uint32_t)beg, rec[2], rec[3]);
        R.rinfo[2 * (size_t)(first + r) + 1] = make_uint4(n_cigar ? rec[kRecWords] : 0u, x, 0u, 0u);
        atomicMax(&span_s, x - (uint32_t)beg);
        if (r > 0) { const uint32_t po = R.rec_off[first + r - 1]; if (po < R.rec_words && (int32_t)R.recs[po] > beg) atomicOr(&err_s, 1u); }
    }
    __syncthreads();
    /* depth = inclusive scan of the marks = exclusive scan + own mark */
    uint32_t *A = R.A + ((size_t)w * 2 + s) * kWin;
    uint32_t own[kWin / kScanThreads];
    constexpr uint32_t chunk = kWin / (kScanThreads / 32);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
#pragma unroll
    for (uint32_t j = 0; j < chunk / 32; ++j) own[j] = diff[warp * chunk + j * 32 + lane];
    __syncthreads();
    block_excl_scan(diff, wsum);
#pragma unroll
    for (uint32_t j = 0; j < chunk / 32; ++j) A[warp * chunk + j * 32 + lane] = diff[warp * chunk + j * 32 + lane] + own[j];
    if (threadIdx.x == 0) {
        R.wspan[w * 2 + s] = span_s;
        if (err_s & 1u) atomicOr(R.summary + 2, kErrUnsorted);
        if (err_s & 2u) atomicOr(R.summary + 2, kErrRecord);
    }
}

__global__ void __launch_bounds__(kScanThreads) columns_kernel(DevReads R)
{
    __shared__ uint32_t v[kWin];
    __shared__ uint32_t wsum[kScanThreads / 32];
    const uint32_t w = blockIdx.x;
    uint32_t *A0 = R.A + (size_t)w * 2 * kWin, *A1 = A0 + kWin;
    const uint8_t *ref = R.wref + (size_t)w * kWin;
    for (uint32_t i = threadIdx.x; i < kWin; i += kScanThreads)
        v[i] = (A0[i] != 0u && A1[i] != 0u && ref[i] != 15) ? 1u : 0u;
    __syncthreads();
    const uint32_t ncol = block_excl_scan(v, wsum);
    for (uint32_t i = threadIdx.x; i < kWin; i += kScanThreads) {
        const uint32_t ct = A0[i], cn = A1[i];
        R.P[(size_t)w * kWin + i] = v[i];
        v[i] = (ct != 0u && cn != 0u && ref[i] != 15) ? ct + cn : 0u;
    }
    __syncthreads();
    const uint32_t nent = block_excl_scan(v, wsum);
    for (uint32_t i = threadIdx.x; i < kWin; i += kScanThreads) {
        const uint32_t ct = A0[i], cn = A1[i];
        const bool live = ct != 0u && cn != 0u && ref[i] != 15;
        A0[i] = live ? v[i] : kDead;
        A1[i] = live ? v[i] + ct : kDead;
    }
    if (threadIdx.x == 0) R.wtot[w] = make_uint2(ncol, nent);
}

__global__ void __launch_bounds__(32) bases_kernel(DevReads R)
{
    const int lane = threadIdx.x;
    uint32_t col = 0;
    unsigned long long ent = 0;
    for (uint32_t w0 = 0; w0 < R.n_win; w0 += 32) {
        const uint32_t w = w0 + lane;
        const uint2 t = w < R.n_win ? R.wtot[w] : make_uint2(0u, 0u);
        const uint32_t ci = warp_incl_scan(t.x, lane);
        unsigned long long ei = t.y;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1) {
            const unsigned long long u = __shfl_up_sync(0xffffffffu, ei, d);
            if (lane >= d) ei += u;
        }
        if (w < R.n_win) R.wtot[w] = make_uint2(col + ci - t.x, (uint32_t)(ent + ei - t.y));
        col += __shfl_sync(0xffffffffu, ci, 31);
        ent += __shfl_sync(0xffffffffu, ei, 31);
    }
    if (lane == 0) {
        uint32_t err = 0;
        if (col > R.max_columns) err |= kErrColumns;
        if (ent > (unsigned long long)R.max_entries) err |= kErrEntries;
        R.summary[0] = col;
        R.summary[1] = (uint32_t)ent;
        if (err) atomicOr(R.summary + 2, err);
        else R.off[2 * (size_t)col] = (uint32_t)ent;
    }
}

__global__ void __launch_bounds__(kThreads) offsets_kernel(DevReads R)
{
    if (R.summary[2] & kErrStop) return;
    const uint32_t g = blockIdx.x * kThreads + threadIdx.x;        /* w * kWin + i */
    const uint32_t w = g >> SSN_READ_WINDOW_SHIFT, i = g & (kWin - 1u);
    if (g < 16u) R.bq[(size_t)R.summary[1] + g] = 1;                /* the clean-tile check reads up to 16 bytes past the end */
    if (w >= R.n_win) return;
    uint32_t *A0 = R.A + (size_t)w * 2 * kWin, *A1 = A0 + kWin;
    const uint32_t a = A0[i];
    if (a == kDead) return;
    const uint2 base = R.wtot[w];
    const uint32_t col = base.x + R.P[g], eT = base.y + a, eN = base.y + A1[i];
    R.off[2 * (size_t)col] = eT;
    R.off[2 * (size_t)col + 1] = eN;
    R.ref_nt16[col] = R.wref[g];
    R.colpos[col] = g;
    A0[i] = eT; A1[i] = eN;
}

/* One warp per 32 consecutive positions of a window, one lane per position, tumor then normal.  The warp walks the
 * reads that can cover any of its positions in file order, all lanes on the same read: header and CIGAR loads are
 * uniform, the base and quality loads of neighbouring positions coalesce, and every lane appends at its own column's
 * cursor.  The 32 columns are one contiguous stretch of bs/bq/mq: it is assembled in shared memory (scattered byte
 * stores to global memory run at the L2's partial-sector rate) and written out as whole 16-byte chunks.  A stretch
 * longer than kStage bytes (about 128 read-bases per column) is written directly. */
constexpr uint32_t kStage = 4096;
constexpr int kGroup = 4;
constexpr size_t kFillSmem = (size_t)(kThreads / 32) * 3 * kStage;

template <bool STAGED>
__device__ __forceinline__ void fill_sample(const DevReads &R, const ssn_read_window &W, uint32_t ws, uint32_t s, uint32_t up,
                                            uint32_t lane, bool live, uint32_t cur, uint8_t *sm)
{
    const uint32_t first = W.rec_first[s], n = W.n_reads[s], up0 = up - lane;
    const long long p0 = up0, p31 = (long long)up0 + 31, span = R.wspan[ws];
    const int32_t *rb = R.rbeg + first;
    /* reads [r0, r1): beg in (p0 - span, p31] — every read that can cover one of the 32 positions, in file order */
    uint32_t lo = 0, hi = n;
    while (lo < hi) { const uint32_t m = (lo + hi) >> 1; if ((long long)rb[m] <= p31) lo = m + 1; else hi = m; }
    const uint32_t r1 = lo;
    lo = 0; hi = r1;
    while (lo < hi) { const uint32_t m = (lo + hi) >> 1; if ((long long)rb[m] + span <= p0) lo = m + 1; else hi = m; }
    /* Reads go in groups of kGroup: all their facts (compact rinfo array) are loaded together, then their CIGARs are
     * walked, then all base and quality loads are issued, then the lanes append in file order — the loop is bound by
     * load latency, so the loads of a group overlap instead of queueing behind one another. */
    const uint4 *info = R.rinfo + 2 * (size_t)first;
    for (uint32_t r = lo; r < r1; r += kGroup) {
        uint4 i0[kGroup], i1[kGroup];
#pragma unroll
        for (int j = 0; j < kGroup; ++j) {
            const uint32_t rr = min(r + j, r1 - 1u);
            i0[j] = info[2 * (size_t)rr]; i1[j] = info[2 * (size_t)rr + 1];
        }
        uint32_t y[kGroup];
        bool hit[kGroup];
#pragma unroll
        for (int j = 0; j < kGroup; ++j) {
            const uint32_t n_cigar = i0[j].w & 0xffffu;
            uint32_t x = i0[j].y, c = i1[j].x;
            y[j] = 0; hit[j] = false;
            /* up - x < l  <=>  x <= up < x + l (a lane whose position lies before the read never satisfies it) */
            if (n_cigar == 1u) {
                if ((c & 0xfu) == 0u && up - x < (c >> 4)) { y[j] = up - x; hit[j] = true; }
            } else {
                const uint32_t *rec = R.recs + i0[j].x;
                bool done = false;
                for (uint32_t k = 0; k < n_cigar; ++k) {
                    if (k) c = rec[kRecWords + k];
                    const uint32_t op = c & 0xfu, l = c >> 4;
                    if (done) continue;
                    if (op == 0u) {
                        if (up - x < l) { y[j] += up - x; hit[j] = true; done = true; }
                        else { x += l; y[j] += l; }
                    } else if (op == 2u || op == 3u) {
                        if (up - x < l) done = true;
                        else x += l;
                    } else if (op == 1u || op == 4u) y[j] += l;
                }
            }
            hit[j] = hit[j] && live && r + j < r1;
        }
        uint8_t sb[kGroup], qq[kGroup];
#pragma unroll
        for (int j = 0; j < kGroup; ++j) {
            const uint32_t n_cigar = i0[j].w & 0xffffu, l_qseq = i0[j].z;
            const uint8_t *seq = (const uint8_t*)(R.recs + i0[j].x + kRecWords + n_cigar), *qual = seq + ((l_qseq + 1u) >> 1);
            sb[j] = 0xff; qq[j] = 0;                                      /* CIGAR longer than the read: N, quality 0 */
            if (hit[j] && y[j] < l_qseq) { sb[j] = seq[y[j] >> 1]; qq[j] = qual[y[j]]; }
        }
#pragma unroll
        for (int j = 0; j < kGroup; ++j) {
            if (!hit[j]) continue;
            const uint32_t meta = i0[j].w;
            const uint8_t strand = (uint8_t)(((meta >> 24) & 0xffu) << 4), mapq = (uint8_t)((meta >> 16) & 0xffu);
            const uint8_t b = (uint8_t)(((sb[j] >> ((~y[j] & 1u) << 2)) & 0xfu) | strand);
            if (STAGED) { sm[cur] = b; sm[kStage + cur] = qq[j]; sm[2 * kStage + cur] = mapq; }
            else { R.bs[cur] = b; R.bq[cur] = qq[j]; R.mq[cur] = mapq; }
            ++cur;
        }
    }
}

/* bytes [base, end) of one array from the warp's stage (stage offset = address - (base & ~15)) */
__device__ __forceinline__ void flush_stage(uint8_t *g, const uint8_t *sm, uint32_t base, uint32_t end, uint32_t lane)
{
    const uint32_t al = base & ~15u;
    const uint32_t a1 = min(end, (base + 15u) & ~15u), m1 = max(a1, end & ~15u);
    for (uint32_t j = base + lane; j < a1; j += 32u) g[j] = sm[j - al];
    for (uint32_t j = a1 + 16u * lane; j < m1; j += 512u) *reinterpret_cast<uint4*>(g + j) = *reinterpret_cast<const uint4*>(sm + (j - al));
    for (uint32_t j = m1 + lane; j < end; j += 32u) g[j] = sm[j - al];
}

__global__ void __launch_bounds__(kThreads) fill_kernel(DevReads R)
{
    extern __shared__ __align__(16) uint8_t fill_stage[];
    if (R.summary[2] & kErrStop) return;
    const uint32_t g = blockIdx.x * kThreads + threadIdx.x;        /* w * kWin + i */
    const uint32_t w = g >> SSN_READ_WINDOW_SHIFT, i = g & (kWin - 1u), lane = threadIdx.x & 31u;
    if (w >= R.n_win) return;
    const uint32_t eT = R.A[(size_t)w * 2 * kWin + i], eN = R.A[(size_t)w * 2 * kWin + kWin + i];
    const bool live = eT != kDead;
    const uint32_t lives = __ballot_sync(0xffffffffu, live);
    if (!lives) return;
    const ssn_read_window W = R.win[w];
    /* the warp's stretch of the entry arrays: first live column's start .. last live column's end */
    const uint32_t base = __shfl_sync(0xffffffffu, eT, __ffs(lives) - 1);
    uint32_t end = live ? R.off[2 * (size_t)(R.wtot[w].x + R.P[g]) + 2] : 0u;
    end = __shfl_sync(0xffffffffu, end, 31 - __clz(lives));
    const uint32_t al = base & ~15u, up = W.lo + i;
    uint8_t *sm = fill_stage + (size_t)(threadIdx.x >> 5) * 3 * kStage;
    if (end - al <= kStage) {
        fill_sample<true>(R, W, w * 2, 0, up, lane, live, eT - al, sm);
        fill_sample<true>(R, W, w * 2 + 1, 1, up, lane, live, eN - al, sm);
        __syncwarp();
        flush_stage(R.bs, sm, base, end, lane);
        flush_stage(R.bq, sm + kStage, base, end, lane);
        flush_stage(R.mq, sm + 2 * kStage, base, end, lane);
    } else {
        fill_sample<false>(R, W, w * 2, 0, up, lane, live, eT, nullptr);
        fill_sample<false>(R, W, w * 2 + 1, 1, up, lane, live, eN, nullptr);
    }
}

/* what scan_offsets (ssn_api.cu) computes on the host for packed batches */
__global__ void __launch_bounds__(kThreads) shape_kernel(DevReads R)
{
    if (R.summary[2] & kErrStop) return;
    const uint32_t ncol = R.summary[0];
    const uint32_t t = (blockIdx.x * kThreads + threadIdx.x) * 32u;
    if (t >= ncol) return;
    const uint32_t e = min(ncol, t + 32u);
    uint32_t words = 0, maxc = 0, err = 0;
    for (uint32_t c = t; c < e; ++c) {
        const uint32_t n = R.off[2 * (size_t)c + 2] - R.off[2 * (size_t)c];
        if (n > R.run_len * R.max_runs) err = kErrDepth;
        const uint32_t cw = (n / R.run_len) * (R.run_len / 2) + (((n & (R.run_len - 1u)) + 1u) >> 1);
        words += cw;
        maxc = max(maxc, cw);
    }
    atomicMax(R.summary + 3, words);
    atomicMax(R.summary + 4, maxc);
    if (err) atomicOr(R.summary + 2, err);
}

}  // namespace pileup
