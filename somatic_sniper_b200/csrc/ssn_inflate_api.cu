/*
 * ssn_inflate_api.cu — BGZF blocks inflated on the GPU: the kernel around ssn_inflate_dev.cuh (one warp per block) and
 * the thread-safe "lane" service the BAM readers call (include/ssn.h: ssn_inflater_*).
 *
 * A lane is what a reader thread holds for one batch: a pinned input buffer it fills with the compressed payloads of
 * a few dozen blocks and a pinned output buffer it takes the inflated bytes from.  ssn_inflater_run is synchronous for the
 * caller and costs it no CPU while it waits.  Behind it one worker thread per inflater gathers the requests of all the
 * readers waiting at that moment into a WAVE — one launch over all their blocks, their copies in front and behind it — on
 * one of three streams; the stream's completion callback wakes the callers.
 * A block the device decoder refuses comes back flagged and the reader inflates it on the host (zlib reads everything
 * the reference reads), so the service can never change what is read.
 */
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <algorithm>
#include <mutex>
#include <thread>
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

/* A lane is a reader's view: a pinned input buffer to fill, a pinned output buffer to read, and one request at a time.
 * Its device twin is a slice of the inflater's arenas, so that one kernel launch can serve the requests of many lanes:
 * block offsets are relative to the arena. */
struct Lane {
    bool ready = false, busy = false;
    uint8_t *h_slab = nullptr, *h_in = nullptr, *h_out = nullptr;
    /* the request in flight (owner: the calling thread until it is queued, the worker / completion until done) */
    std::vector<ssn_inflate_block> blk;
    uint32_t n = 0;
    uint64_t in_bytes = 0, out_bytes = 0;
    uint8_t *refused = nullptr, *h_dst = nullptr;
    bool done = false;
    int error = 0;
};

struct ssn_inflater;

/* A wave is one launch: the requests that were waiting when the worker came round, up to wave_blocks blocks (three
 * requests' worth).  Eight waves on eight streams keep copies in, kernels and copies out overlapped.  Why not a stream per
 * lane: with more streams than the driver has hardware queues (8 by default, CUDA_DEVICE_MAX_CONNECTIONS), scoring
 * launches share a queue with inflate kernels they do not depend on and wait behind them — measured on B200, the CLI's
 * stream fell from 3.5 to 2.6e7 columns/s that way.  Raising the queue count cures it but costs set-up time per queue and
 * GPU, so the number of streams is kept small and fixed (the CLI asks for 16 queues: 8 here + the scoring streams).
 * Measured (64 Mb pair, 1 GPU, 16 cores, columns/s streamed): host inflate 3.5e7; 3 large waves, 8 queues 4.0e7;
 * 8 waves, 16 queues 4.4-4.7e7; a stream per lane (16), 32 queues 4.9-5.1e7. */
struct Wave {
    ssn_inflater *f = nullptr;
    bool busy = false;
    cudaStream_t stream = nullptr;
    ssn_inflate_block *h_blk = nullptr, *d_blk = nullptr;
    uint32_t *h_status = nullptr, *d_status = nullptr, *d_counter = nullptr;
    std::vector<int> lanes;                 /* lanes served, in table order */
    uint32_t n = 0;
};

constexpr int kMaxWaves = 32;

struct ssn_inflater {
    int device, num_sms;
    uint64_t in_cap, out_cap, in_stride, out_stride;
    uint32_t max_blocks, wave_blocks;
    std::vector<Lane> lanes;
    uint8_t *d_in = nullptr, *d_out = nullptr;           /* arenas: lane i owns [i * stride, (i + 1) * stride) */
    Wave waves[kMaxWaves];
    int n_waves = 8;
    std::vector<int> pending;               /* lanes with a request, oldest first */
    std::thread worker;
    bool stop = false, worker_failed = false;
    std::mutex mu;
    std::condition_variable cv_lane, cv_work, cv_done;
    uint64_t blocks = 0, refused = 0, runs = 0, in_bytes = 0, out_bytes = 0, launches = 0;
    /* page-locked buffers handed to callers (ssn_inflater_host_alloc); given back ones are kept for the next caller and
     * freed with the inflater: cudaFreeHost waits for the whole device, which a running stream cannot afford */
    struct HostBuf { void *p; uint64_t bytes; bool busy; };
    std::vector<HostBuf> host_bufs;
};

static void inflater_worker(ssn_inflater *f);

static void inflater_release(ssn_inflater *f)
{
    for (Lane &L : f->lanes) if (L.h_slab) cudaFreeHost(L.h_slab);
    for (Wave &w : f->waves) {
        if (w.stream) cudaStreamDestroy(w.stream);
        cudaFreeHost(w.h_blk); cudaFreeHost(w.h_status);
        cudaFree(w.d_blk); cudaFree(w.d_status); cudaFree(w.d_counter);
    }
    cudaFree(f->d_in); cudaFree(f->d_out);
    for (auto &h : f->host_bufs) cudaFreeHost(h.p);
    delete f;
}

extern "C" ssn_status ssn_inflater_create(int device, uint32_t n_lanes, uint64_t in_cap, uint64_t out_cap, uint32_t max_blocks, ssn_inflater **out)
{
    if (!out || !n_lanes || !in_cap || !out_cap || !max_blocks) { ssn_set_error("ssn_inflater_create: bad argument"); return SSN_ERR_ARG; }
    const uint64_t in_stride = ((in_cap + 63) & ~63ull) + 64, out_stride = ((out_cap + 63) & ~63ull) + 64;
    if (in_stride * n_lanes > 0xffffffffull || out_stride * n_lanes > 0xffffffffull) {
        ssn_set_error("ssn_inflater_create: %u lanes of %llu + %llu bytes exceed the 4 GiB a launch addresses", n_lanes,
                      (unsigned long long)in_cap, (unsigned long long)out_cap);
        return SSN_ERR_ARG;
    }
    CUI(cudaSetDevice(device));
    cudaDeviceProp prop;
    CUI(cudaGetDeviceProperties(&prop, device));
    if (prop.major < 10) { ssn_set_error("ssn_inflater_create: device %d is not sm_100", device); return SSN_ERR_CUDA; }
    CUI(cudaFuncSetAttribute(ssn_inflate_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)(kInflWarps * sizeof(infl::Work))));
    ssn_inflater *f = new ssn_inflater();
    f->device = device; f->num_sms = prop.multiProcessorCount;
    f->in_cap = (in_cap + 63) & ~63ull; f->out_cap = (out_cap + 63) & ~63ull; f->max_blocks = max_blocks;
    f->in_stride = in_stride; f->out_stride = out_stride;
    f->wave_blocks = 3u * max_blocks;
    {
        /* measurement knobs: streams (waves in flight) and the most blocks one launch takes */
        const char *ew = getenv("SSN_INFLATE_WAVES"), *eb = getenv("SSN_INFLATE_WAVE_BLOCKS");
        if (ew && atoi(ew) >= 1 && atoi(ew) <= kMaxWaves) f->n_waves = atoi(ew);
        if (eb && atoi(eb) >= 1) f->wave_blocks = std::max<uint32_t>(max_blocks, (uint32_t)atoi(eb));
    }
    f->lanes.resize(n_lanes);
#define CUF(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) { ssn_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); inflater_release(f); return SSN_ERR_CUDA; } } while (0)
    CUF(cudaMalloc(&f->d_in, in_stride * n_lanes));
    CUF(cudaMalloc(&f->d_out, out_stride * n_lanes));
    int lo = 0, hi = 0;
    CUF(cudaDeviceGetStreamPriorityRange(&lo, &hi));    /* numerically lowest = highest priority is for the scoring streams */
    for (int wi = 0; wi < f->n_waves; ++wi) {
        Wave &w = f->waves[wi];
        w.f = f;
        CUF(cudaStreamCreateWithPriority(&w.stream, cudaStreamNonBlocking, lo));
        CUF(cudaMallocHost(&w.h_blk, sizeof(ssn_inflate_block) * f->wave_blocks));
        CUF(cudaMallocHost(&w.h_status, 4 * (size_t)f->wave_blocks));
        CUF(cudaMalloc(&w.d_blk, sizeof(ssn_inflate_block) * f->wave_blocks));
        CUF(cudaMalloc(&w.d_status, 4 * (size_t)f->wave_blocks));
        CUF(cudaMalloc(&w.d_counter, 256));
    }
#undef CUF
    f->worker = std::thread(inflater_worker, f);
    *out = f;
    return SSN_OK;
}

extern "C" void ssn_inflater_destroy(ssn_inflater *f)
{
    if (!f) return;
    {
        std::unique_lock<std::mutex> lk(f->mu);
        f->stop = true;
        f->cv_work.notify_all();
    }
    if (f->worker.joinable()) f->worker.join();
    cudaSetDevice(f->device);
    for (Wave &w : f->waves) if (w.stream) cudaStreamSynchronize(w.stream);
    inflater_release(f);
}

/* pinned buffers of a lane on first use (pinned allocations are slow: lanes nobody needs cost nothing) */
static ssn_status lane_prepare(ssn_inflater *f, Lane &L)
{
    if (L.ready) return SSN_OK;
    CUI(cudaSetDevice(f->device));
    CUI(cudaMallocHost(&L.h_slab, f->in_stride + f->out_stride));
    L.h_in = L.h_slab; L.h_out = L.h_slab + f->in_stride;
    memset(L.h_in, 0, f->in_stride);
    L.blk.resize(f->max_blocks);
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
            f->cv_lane.wait(lk);
        }
        f->lanes[pick].busy = true;
    }
    ssn_status st = lane_prepare(f, f->lanes[pick]);
    if (st != SSN_OK) {
        std::lock_guard<std::mutex> lk(f->mu);
        f->lanes[pick].busy = false;
        f->cv_lane.notify_one();
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
    f->cv_lane.notify_one();
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

/* stream callback (a driver thread; no CUDA calls here): the wave's copies back have landed */
static void CUDART_CB wave_done(void *arg)
{
    Wave *w = (Wave*)arg;
    ssn_inflater *f = w->f;
    std::lock_guard<std::mutex> lk(f->mu);
    uint32_t at = 0;
    for (int li : w->lanes) {
        Lane &L = f->lanes[li];
        uint32_t bad = 0;
        for (uint32_t i = 0; i < L.n; ++i) { L.refused[i] = w->h_status[at + i] != 0; bad += L.refused[i]; }
        at += L.n;
        f->blocks += L.n; f->refused += bad; f->runs += 1; f->in_bytes += L.in_bytes; f->out_bytes += L.out_bytes;
        L.done = true;
    }
    f->launches += 1;
    w->lanes.clear();
    w->busy = false;
    f->cv_done.notify_all();
    f->cv_work.notify_all();
}

static cudaError_t wave_issue(ssn_inflater *f, Wave *w)
{
#define CUW(call) do { cudaError_t e_ = (call); if (e_ != cudaSuccess) return e_; } while (0)
    cudaStream_t st = w->stream;
    uint32_t at = 0;
    for (int li : w->lanes) {
        Lane &L = f->lanes[li];
        const uint32_t base_in = (uint32_t)(f->in_stride * (uint64_t)li), base_out = (uint32_t)(f->out_stride * (uint64_t)li);
        for (uint32_t i = 0; i < L.n; ++i) {
            const ssn_inflate_block b = L.blk[i];
            w->h_blk[at + i] = ssn_inflate_block{b.in_off + base_in, b.in_len, b.out_off + base_out, b.out_len};
        }
        CUW(cudaMemcpyAsync(f->d_in + base_in, L.h_in, ((size_t)L.in_bytes + 63) & ~(size_t)63, cudaMemcpyHostToDevice, st));
        at += L.n;
    }
    CUW(cudaMemsetAsync(w->d_counter, 0, 4, st));
    CUW(cudaMemcpyAsync(w->d_blk, w->h_blk, sizeof(ssn_inflate_block) * at, cudaMemcpyHostToDevice, st));
    const uint32_t grid = std::min<uint32_t>((at + kInflWarps - 1) / kInflWarps, (uint32_t)f->num_sms * kInflCtasPerSm);
    ssn_inflate_kernel<<<grid, kInflWarps * 32, kInflWarps * sizeof(infl::Work), st>>>(f->d_in, w->d_blk, at, f->d_out, w->d_status, w->d_counter);
    CUW(cudaGetLastError());
    for (int li : w->lanes) {
        Lane &L = f->lanes[li];
        CUW(cudaMemcpyAsync(L.h_dst ? L.h_dst : L.h_out, f->d_out + f->out_stride * (uint64_t)li, (size_t)L.out_bytes, cudaMemcpyDeviceToHost, st));
    }
    CUW(cudaMemcpyAsync(w->h_status, w->d_status, 4 * (size_t)at, cudaMemcpyDeviceToHost, st));
    CUW(cudaLaunchHostFunc(st, wave_done, w));
#undef CUW
    return cudaSuccess;
}

/* one thread per inflater: gathers the waiting requests into a wave and queues it; completion is the stream's callback */
static void inflater_worker(ssn_inflater *f)
{
    cudaSetDevice(f->device);
    int turn = 0;
    for (;;) {
        Wave *w = nullptr;
        {
            std::unique_lock<std::mutex> lk(f->mu);
            for (;;) {
                if (f->stop) return;
                if (!f->pending.empty()) {
                    for (int k = 0; k < f->n_waves && !w; ++k) { Wave &c = f->waves[(turn + k) % f->n_waves]; if (!c.busy) { w = &c; turn = (turn + k + 1) % f->n_waves; } }
                    if (w) break;
                }
                f->cv_work.wait(lk);
            }
            uint32_t nb = 0;
            size_t k = 0;
            while (k < f->pending.size() && nb + f->lanes[f->pending[k]].n <= f->wave_blocks) { nb += f->lanes[f->pending[k]].n; ++k; }
            w->lanes.assign(f->pending.begin(), f->pending.begin() + (long)k);
            f->pending.erase(f->pending.begin(), f->pending.begin() + (long)k);
            w->n = nb;
            w->busy = true;
        }
        const cudaError_t e = wave_issue(f, w);
        if (e != cudaSuccess) {
            /* what was queued on the stream before the failure is waited for, then the wave's callers are told */
            cudaStreamSynchronize(w->stream);
            std::lock_guard<std::mutex> lk(f->mu);
            if (w->busy) {                                   /* the callback did not run */
                for (int li : w->lanes) { f->lanes[li].error = (int)e; f->lanes[li].done = true; }
                w->lanes.clear();
                w->busy = false;
                f->cv_done.notify_all();
            }
        }
    }
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
        L.blk[i] = blk[i];
    }
    L.n = n; L.in_bytes = in_bytes; L.out_bytes = out_bytes; L.refused = refused; L.h_dst = h_dst; L.error = 0;
    {
        /* the caller sleeps until its wave has landed: its core goes to another reader */
        std::unique_lock<std::mutex> lk(f->mu);
        L.done = false;
        f->pending.push_back(lane);
        f->cv_work.notify_all();
        f->cv_done.wait(lk, [&] { return L.done; });
    }
    if (L.error) { ssn_set_error("ssn_inflater_run: %s", cudaGetErrorString((cudaError_t)L.error)); return SSN_ERR_CUDA; }
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
