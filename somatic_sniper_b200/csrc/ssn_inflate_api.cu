/*
 * ssn_inflate_api.cu — BGZF blocks inflated on the GPU: the kernel around ssn_inflate_dev.cuh (one warp per block) and
 * the thread-safe "lane" service the BAM readers call (include/ssn.h: ssn_inflater_*).
 *
 * A lane is what one reader thread owns while it reads: a pinned input buffer it fills with the compressed payloads of
 * a few dozen blocks, a pinned output buffer it frames records from, their device twins and a stream.  ssn_inflater_run
 * is synchronous for the caller (copy in, kernel, copy out, wait) and costs it no CPU while it waits; dozens of readers
 * do that at once, so the device sees dozens of small launches on independent streams that together fill its SMs.
 * A block the device decoder refuses comes back flagged and the reader inflates it on the host (zlib reads everything
 * the reference reads), so the service can never change what is read.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <mutex>
#include <condition_variable>
#include <vector>

#include "ssn.h"
#include "ssn_inflate_dev.cuh"

extern "C" void ssn_set_error(const char *fmt, ...);

#define CUI(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            ssn_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return SSN_ERR_CUDA;                                                                   \
        }                                                                                          \
    } while (0)

constexpr int kInflWarps = 8;              /* per CTA: 8 x 9.3 KB of tables; three CTAs per SM */
constexpr uint32_t kInflCtasPerSm = 3;

__global__ void __launch_bounds__(kInflWarps * 32)
ssn_inflate_kernel(const uint8_t *__restrict__ in, const ssn_inflate_block *__restrict__ blk, uint32_t n, uint8_t *out,
                   uint32_t *status, uint32_t *counter)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    infl::Work &W = reinterpret_cast<infl::Work*>(smem_raw)[threadIdx.x >> 5];
    const int lane = threadIdx.x & 31;
    for (;;) {
        uint32_t i = 0;
        if (lane == 0) i = atomicAdd(counter, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= n) break;
        const ssn_inflate_block b = blk[i];
        const int rc = infl::inflate_block(W, in + b.in_off, b.in_len, out + b.out_off, b.out_len);
        if (lane == 0) status[i] = (uint32_t)rc;
        __syncwarp();
    }
}

struct Lane {
    bool ready = false, busy = false;
    uint8_t *h_slab = nullptr, *d_slab = nullptr;                 /* one pinned and one device allocation per lane */
    uint8_t *h_in = nullptr, *h_out = nullptr, *d_in = nullptr, *d_out = nullptr;
    ssn_inflate_block *h_blk = nullptr, *d_blk = nullptr;
    uint32_t *h_status = nullptr, *d_status = nullptr, *d_counter = nullptr;
    cudaStream_t stream = nullptr;
    cudaEvent_t done = nullptr;          /* blocking-sync: a waiting reader sleeps, its core goes to another reader */
};

struct ssn_inflater {
    int device, num_sms;
    uint64_t in_cap, out_cap;
    uint32_t max_blocks;
    std::vector<Lane> lanes;
    std::mutex mu;
    std::condition_variable cv;
    uint64_t blocks = 0, refused = 0, runs = 0, in_bytes = 0, out_bytes = 0;
    /* page-locked buffers handed to callers (ssn_inflater_host_alloc); given back ones are kept for the next caller and
     * freed with the inflater: cudaFreeHost waits for the whole device, which a running stream cannot afford */
    struct HostBuf { void *p; uint64_t bytes; bool busy; };
    std::vector<HostBuf> host_bufs;
};

extern "C" ssn_status ssn_inflater_create(int device, uint32_t n_lanes, uint64_t in_cap, uint64_t out_cap, uint32_t max_blocks, ssn_inflater **out)
{
    if (!out || !n_lanes || !in_cap || !out_cap || !max_blocks) { ssn_set_error("ssn_inflater_create: bad argument"); return SSN_ERR_ARG; }
    CUI(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUI(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) { ssn_set_error("ssn_inflater_create: device %d is not sm_100", device); return SSN_ERR_CUDA; }
    CUI(cudaFuncSetAttribute(ssn_inflate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kInflWarps * sizeof(infl::Work))));
    ssn_inflater *f = new ssn_inflater();
    f->device = device; f->num_sms = prop.multiProcessorCount;
    f->in_cap = (in_cap + 63) & ~63ull; f->out_cap = (out_cap + 63) & ~63ull; f->max_blocks = max_blocks;
    f->lanes.resize(n_lanes);
    *out = f;
    return SSN_OK;
}

static void lane_free(Lane &L)
{
    cudaFreeHost(L.h_slab); cudaFree(L.d_slab);
    if (L.done) cudaEventDestroy(L.done);
    if (L.stream) cudaStreamDestroy(L.stream);
    L = Lane();
}

extern "C" void ssn_inflater_destroy(ssn_inflater *f)
{
    if (!f) return;
    cudaSetDevice(f->device);
    for (Lane &L : f->lanes) if (L.ready) { cudaStreamSynchronize(L.stream); lane_free(L); }
    for (auto &h : f->host_bufs) cudaFreeHost(h.p);
    delete f;
}

/* buffers of a lane on first use (pinned allocations are slow: lanes nobody needs cost nothing) */
static ssn_status lane_prepare(ssn_inflater *f, Lane &L)
{
    if (L.ready) return SSN_OK;
    CUI(cudaSetDevice(f->device));
    const size_t n_in = f->in_cap + 64, n_out = f->out_cap + 64;
    const size_t n_blk = (sizeof(ssn_inflate_block) * f->max_blocks + 63) & ~(size_t)63, n_st = (4 * (size_t)f->max_blocks + 63) & ~(size_t)63;
    CUI(cudaMallocHost(&L.h_slab, n_in + n_out + n_blk + n_st));
    CUI(cudaMalloc(&L.d_slab, n_in + n_out + n_blk + n_st + 64));
    L.h_in = L.h_slab; L.h_out = L.h_in + n_in; L.h_blk = (ssn_inflate_block*)(L.h_out + n_out); L.h_status = (uint32_t*)(L.h_out + n_out + n_blk);
    L.d_in = L.d_slab; L.d_out = L.d_in + n_in; L.d_blk = (ssn_inflate_block*)(L.d_out + n_out); L.d_status = (uint32_t*)(L.d_out + n_out + n_blk);
    L.d_counter = (uint32_t*)(L.d_out + n_out + n_blk + n_st);
    /* numerically lowest = highest priority is for the scoring streams; these take what is left */
    int lo = 0, hi = 0;
    CUI(cudaDeviceGetStreamPriorityRange(&lo, &hi));
    CUI(cudaStreamCreateWithPriority(&L.stream, cudaStreamNonBlocking, lo));
    CUI(cudaEventCreateWithFlags(&L.done, cudaEventBlockingSync | cudaEventDisableTiming));
    memset(L.h_in, 0, n_in);
    L.ready = true;
    return SSN_OK;
}

extern "C" ssn_status ssn_inflater_lane_acquire(ssn_inflater *f, int32_t *lane, uint8_t **in, uint8_t **out)
{
    if (!f || !lane || !in || !out) { ssn_set_error("ssn_inflater_lane_acquire: null"); return SSN_ERR_ARG; }
    int pick = -1;
    {
        std::unique_lock<std::mutex> lk(f->mu);
        for (;;) {
            for (size_t i = 0; i < f->lanes.size(); ++i) if (!f->lanes[i].busy && f->lanes[i].ready) { pick = (int)i; break; }   /* a warm one first */
            if (pick < 0) for (size_t i = 0; i < f->lanes.size(); ++i) if (!f->lanes[i].busy) { pick = (int)i; break; }
            if (pick >= 0) break;
            f->cv.wait(lk);
        }
        f->lanes[pick].busy = true;
    }
    ssn_status st = lane_prepare(f, f->lanes[pick]);
    if (st != SSN_OK) {
        std::lock_guard<std::mutex> lk(f->mu);
        f->lanes[pick].busy = false;
        f->cv.notify_one();
        return st;
    }
    *lane = pick; *in = f->lanes[pick].h_in; *out = f->lanes[pick].h_out;
    return SSN_OK;
}

extern "C" void ssn_inflater_lane_release(ssn_inflater *f, int32_t lane)
{
    if (!f || lane < 0 || (size_t)lane >= f->lanes.size()) return;
    std::lock_guard<std::mutex> lk(f->mu);
    f->lanes[lane].busy = false;
    f->cv.notify_one();
}

extern "C" ssn_status ssn_inflater_limits(const ssn_inflater *f, uint64_t *in_cap, uint64_t *out_cap, uint32_t *max_blocks)
{
    if (!f) { ssn_set_error("ssn_inflater_limits: null"); return SSN_ERR_ARG; }
    if (in_cap) *in_cap = f->in_cap;
    if (out_cap) *out_cap = f->out_cap;
    if (max_blocks) *max_blocks = f->max_blocks;
    return SSN_OK;
}

extern "C" ssn_status ssn_inflater_host_alloc(ssn_inflater *f, uint64_t bytes, void **out)
{
    if (!f || !out || !bytes) { ssn_set_error("ssn_inflater_host_alloc: bad argument"); return SSN_ERR_ARG; }
    {
        std::lock_guard<std::mutex> lk(f->mu);
        for (auto &h : f->host_bufs) if (!h.busy && h.bytes >= bytes && h.bytes <= 2 * bytes) { h.busy = true; *out = h.p; return SSN_OK; }
    }
    void *p = nullptr;
    CUI(cudaSetDevice(f->device));
    CUI(cudaHostAlloc(&p, bytes, cudaHostAllocPortable));           /* any GPU's lane may write into it */
    std::lock_guard<std::mutex> lk(f->mu);
    f->host_bufs.push_back({p, bytes, true});
    *out = p;
    return SSN_OK;
}

extern "C" void ssn_inflater_host_free(ssn_inflater *f, void *p)
{
    if (!f || !p) return;
    std::lock_guard<std::mutex> lk(f->mu);
    for (auto &h : f->host_bufs) if (h.p == p) { h.busy = false; return; }
}

extern "C" ssn_status ssn_inflater_run(ssn_inflater *f, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes,
                                       uint64_t out_bytes, uint8_t *refused)
{
    return ssn_inflater_run_to(f, lane, blk, n, in_bytes, out_bytes, refused, nullptr);
}

extern "C" ssn_status ssn_inflater_run_to(ssn_inflater *f, int32_t lane, const ssn_inflate_block *blk, uint32_t n, uint64_t in_bytes,
                                          uint64_t out_bytes, uint8_t *refused, uint8_t *h_dst)
{
    if (!f || lane < 0 || (size_t)lane >= f->lanes.size() || !blk || !refused) { ssn_set_error("ssn_inflater_run: bad argument"); return SSN_ERR_ARG; }
    Lane &L = f->lanes[lane];
    if (!L.busy || !L.ready) { ssn_set_error("ssn_inflater_run: lane was not acquired"); return SSN_ERR_STATE; }
    if (n > f->max_blocks || in_bytes > f->in_cap || out_bytes > f->out_cap) { ssn_set_error("ssn_inflater_run: batch exceeds the lane"); return SSN_ERR_ARG; }
    if (n == 0) return SSN_OK;
    for (uint32_t i = 0; i < n; ++i) {
        if ((uint64_t)blk[i].in_off + blk[i].in_len > in_bytes || (uint64_t)blk[i].out_off + blk[i].out_len > out_bytes) { ssn_set_error("ssn_inflater_run: block %u lies outside the batch", i); return SSN_ERR_ARG; }
        L.h_blk[i] = blk[i];
    }
    CUI(cudaSetDevice(f->device));
    CUI(cudaMemsetAsync(L.d_counter, 0, 4, L.stream));
    CUI(cudaMemcpyAsync(L.d_in, L.h_in, ((size_t)in_bytes + 63) & ~(size_t)63, cudaMemcpyHostToDevice, L.stream));
    CUI(cudaMemcpyAsync(L.d_blk, L.h_blk, sizeof(ssn_inflate_block) * n, cudaMemcpyHostToDevice, L.stream));
    const uint32_t grid = std::min<uint32_t>((n + kInflWarps - 1) / kInflWarps, (uint32_t)f->num_sms * kInflCtasPerSm);
    ssn_inflate_kernel<<<grid, kInflWarps * 32, kInflWarps * sizeof(infl::Work), L.stream>>>(L.d_in, L.d_blk, n, L.d_out, L.d_status, L.d_counter);
    CUI(cudaGetLastError());
    CUI(cudaMemcpyAsync(h_dst ? h_dst : L.h_out, L.d_out, (size_t)out_bytes, cudaMemcpyDeviceToHost, L.stream));
    CUI(cudaMemcpyAsync(L.h_status, L.d_status, 4 * (size_t)n, cudaMemcpyDeviceToHost, L.stream));
    CUI(cudaEventRecord(L.done, L.stream));
    CUI(cudaEventSynchronize(L.done));
    uint32_t bad = 0;
    for (uint32_t i = 0; i < n; ++i) { refused[i] = L.h_status[i] != 0; bad += refused[i]; }
    {
        std::lock_guard<std::mutex> lk(f->mu);
        f->blocks += n; f->refused += bad; f->runs += 1; f->in_bytes += in_bytes; f->out_bytes += out_bytes;
    }
    return SSN_OK;
}

extern "C" ssn_status ssn_inflater_stats(ssn_inflater *f, uint64_t *blocks, uint64_t *refused, uint64_t *runs, uint64_t *in_bytes, uint64_t *out_bytes)
{
    if (!f) { ssn_set_error("ssn_inflater_stats: null"); return SSN_ERR_ARG; }
    std::lock_guard<std::mutex> lk(f->mu);
    if (blocks) *blocks = f->blocks;
    if (refused) *refused = f->refused;
    if (runs) *runs = f->runs;
    if (in_bytes) *in_bytes = f->in_bytes;
    if (out_bytes) *out_bytes = f->out_bytes;
    return SSN_OK;
}

/* ---- ssn_inflate_static: the same decoder with no allocation anywhere ----------------------------------------------------
 * No counterpart in the reference; added for callers that cannot call cudaMalloc
 * (CUDA graph capture, tight serving loops, harnesses that forbid it).  The caller owns the input, output and status
 * buffers on the device; what ssn_inflater_run keeps per lane in cudaMalloc'ed memory — the block table and the work
 * counter the warps pull from — lives in a file-scope __device__ array of fixed size instead (every GPU that loads the
 * library holds its 64.25 KB until exit).  All calls share that array: calls that overlap on the GPU race. */
constexpr uint32_t kStaticHeadWords = 64;                              /* word 0: work counter; the table starts 256 B in */
static __device__ __align__(256) uint32_t g_infl_static[kStaticHeadWords + 4u * SSN_INFLATE_STATIC_MAX_BLOCKS];

__global__ void __launch_bounds__(kInflWarps * 32)
ssn_inflate_static_kernel(const uint8_t *__restrict__ in, uint32_t n, uint8_t *out, uint32_t *status, uint32_t *scratch)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    infl::Work &W = reinterpret_cast<infl::Work*>(smem_raw)[threadIdx.x >> 5];
    const ssn_inflate_block *blk = reinterpret_cast<const ssn_inflate_block*>(scratch + kStaticHeadWords);
    const int lane = threadIdx.x & 31;
    for (;;) {
        uint32_t i = 0;
        if (lane == 0) i = atomicAdd(scratch, 1u);
        i = __shfl_sync(0xffffffffu, i, 0);
        if (i >= n) break;
        const ssn_inflate_block b = blk[i];
        const int rc = infl::inflate_block(W, in + b.in_off, b.in_len, out + b.out_off, b.out_len);
        if (lane == 0) status[i] = (uint32_t)rc;
        __syncwarp();
    }
}

/* per device, set up once with that device current: the array's DEVICE address (its name is only a host-side stand-in
 * here) and the kernel's shared-memory opt-in */
constexpr int kStaticMaxDevices = 64;
static std::mutex g_static_mu;
static uint32_t *g_static_ptr[kStaticMaxDevices];
static int g_static_sms[kStaticMaxDevices];

extern "C" ssn_status ssn_inflate_static(int device, const uint8_t *d_in, uint64_t in_bytes, const ssn_inflate_block *blk, uint32_t n,
                                         uint8_t *d_out, uint64_t out_bytes, uint32_t *d_status, void *stream)
{
    if (device < 0 || device >= kStaticMaxDevices) { ssn_set_error("ssn_inflate_static: device %d out of range", device); return SSN_ERR_ARG; }
    if (n > SSN_INFLATE_STATIC_MAX_BLOCKS) {
        ssn_set_error("ssn_inflate_static: %u blocks in one call, the static scratch holds at most %u (SSN_INFLATE_STATIC_MAX_BLOCKS); split the batch",
                      n, (unsigned)SSN_INFLATE_STATIC_MAX_BLOCKS);
        return SSN_ERR_ARG;
    }
    if (n == 0) return SSN_OK;
    if (!d_in || !blk || !d_out || !d_status) { ssn_set_error("ssn_inflate_static: null argument"); return SSN_ERR_ARG; }
    if (in_bytes > 0xffffffffull || out_bytes > 0xffffffffull) { ssn_set_error("ssn_inflate_static: a batch is addressed with 32-bit offsets"); return SSN_ERR_ARG; }
    for (uint32_t i = 0; i < n; ++i)
        if ((uint64_t)blk[i].in_off + blk[i].in_len > in_bytes || (uint64_t)blk[i].out_off + blk[i].out_len > out_bytes) {
            ssn_set_error("ssn_inflate_static: block %u lies outside the batch", i);
            return SSN_ERR_ARG;
        }
    CUI(cudaSetDevice(device));
    uint32_t *scratch;
    int sms;
    {
        std::lock_guard<std::mutex> lk(g_static_mu);
        if (!g_static_ptr[device]) {
            cudaDeviceProp prop;
            uint32_t *p = nullptr;
            CUI(cudaGetDeviceProperties(&prop, device));
            if (prop.major < 10) { ssn_set_error("ssn_inflate_static: device %d is not sm_100", device); return SSN_ERR_CUDA; }
            CUI(cudaGetSymbolAddress((void**)&p, g_infl_static));
            CUI(cudaFuncSetAttribute(ssn_inflate_static_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kInflWarps * sizeof(infl::Work))));
            g_static_sms[device] = prop.multiProcessorCount;
            g_static_ptr[device] = p;
        }
        scratch = g_static_ptr[device];
        sms = g_static_sms[device];
    }
    cudaStream_t st = (cudaStream_t)stream;
    /* the array keeps the last call's contents: the counter starts at zero, the table is overwritten for [0, n) */
    CUI(cudaMemsetAsync(scratch, 0, kStaticHeadWords * 4, st));
    CUI(cudaMemcpyAsync(scratch + kStaticHeadWords, blk, sizeof(ssn_inflate_block) * (size_t)n, cudaMemcpyHostToDevice, st));
    const uint32_t grid = std::min<uint32_t>((n + kInflWarps - 1) / kInflWarps, (uint32_t)sms * kInflCtasPerSm);
    ssn_inflate_static_kernel<<<grid, kInflWarps * 32, kInflWarps * sizeof(infl::Work), st>>>(d_in, n, d_out, d_status, scratch);
    CUI(cudaGetLastError());
    return SSN_OK;
}
