/*
 * ssn_api.cu — the C ABI of include/ssn.h: context, model upload, pinned SoA batch slots,
 * stream/event pipeline (H2D -> fused scoring kernel -> D2H of the compacted site ring).
 * There is no CPU path in this file or anywhere else in libssn.so.
 */
#include <algorithm>
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <ctime>
#include <thread>
#include <vector>

#include "ssn_internal.h"
#include "ssn_kernels.cuh"
#include "ssn_prefilter.cuh"
#include "ssn_pileup.cuh"

using namespace ssn;

/* ------------------------------------------------------------------------------------------------ */

static thread_local char g_err[512] = "";

extern "C" void ssn_set_error(const char *fmt, ...)
{
    va_list ap;
    va_start(ap, fmt);
    vsnprintf(g_err, sizeof(g_err), fmt, ap);
    va_end(ap);
}
extern "C" const char *ssn_last_error(void) { return g_err; }
extern "C" int ssn_abi_version(void) { return SSN_ABI_VERSION; }

#define CU(call)                                                                                   \
    do {                                                                                           \
        cudaError_t e_ = (call);                                                                   \
        if (e_ != cudaSuccess) {                                                                   \
            ssn_set_error("%s failed: %s (%s:%d)", #call, cudaGetErrorString(e_), __FILE__, __LINE__); \
            return SSN_ERR_CUDA;                                                                   \
        }                                                                                          \
    } while (0)

enum SlotState { SLOT_FREE = 0, SLOT_FILLING, SLOT_SUBMITTED, SLOT_DONE };

struct ssn_batch {
    ssn_ctx *ctx;
    int index;
    SlotState state;
    uint64_t seq;
    uint32_t max_columns;
    uint64_t max_entries;
    uint32_t ncol;
    uint64_t nent;
    uint32_t flags;
    bool resident;            /* device copy matches the host arrays */
    /* pinned host */
    uint32_t *h_off; uint8_t *h_bs, *h_bq, *h_mq, *h_ref;
    uint32_t *h_count;        /* kCountWords: [0] sites, [1] score tiles, [2] pre-filter survivors, [3] pre-filter tiles */
    ssn_site *h_sites; uint32_t h_sites_cap;
    uint32_t *h_dump; size_t h_dump_words;
    /* device */
    uint32_t *d_off; uint8_t *d_bs, *d_bq, *d_mq, *d_ref;
    uint32_t *d_count;
    ssn_site *d_sites; uint32_t site_cap;
    uint32_t *d_dump; size_t d_dump_words;
    uint32_t max_tile_words;
    uint32_t max_col_words;   /* key-store words of the deepest column (list-mode tiles gather arbitrary columns) */
    bool listed;              /* the last launch ran the pre-filter and scored its survivor list */
    uint32_t n_deep;          /* columns with more than 255 read-bases in a sample: the pre-filter cannot settle those */
    uint32_t *d_survivors;    /* column indices the pre-filter hands to the full model */
    uint32_t *d_scratch; size_t scratch_words;   /* global key store of this slot's launches (tiles deeper than shared
                                                   * memory); per slot, because slots score concurrently on their own streams */
    cudaStream_t stream;
    cudaEvent_t done;
    uint32_t n_sites;
    /* read batch: the device builds the columns (ssn_pileup.cuh) */
    bool reads_alloc, reads_pending, from_reads;
    bool spec;                /* scored speculatively from the device-side column count (submit_reads) */
    uint32_t spec_col_words;
    uint32_t n_win;
    uint8_t *h_recs, *h_index; uint32_t *h_rec_off; ssn_read_window *h_win; uint8_t *h_wref; uint32_t *h_summary, *h_colpos;   /* h_index owns win/wref/rec_off */
    size_t index_off_bytes;
    uint8_t *d_recs, *d_index; uint32_t *d_rec_off; ssn_read_window *d_win; uint8_t *d_wref;
    uint32_t *d_A, *d_P, *d_wspan, *d_summary, *d_colpos; int32_t *d_rbeg; uint4 *d_rinfo; uint2 *d_wtot;
    cudaEvent_t piled, pa, pb;
};

struct KernelTimer {
    cudaEvent_t a, m, b;      /* a -> m: pre-filter kernel (when it ran), m -> b: scoring kernel */
    bool pending, filtered;
};

constexpr int kCountWords = 8;

struct ssn_ctx {
    int device;
    ssn_params params;
    ssn_host_tables tab;
    /* device tables: coef and lhet share one allocation so one access-policy window covers both */
    double *d_coef_lhet; double *d_fk;
    int32_t *d_prior, *d_jointprior, *d_qadd;
    DevModel model;
    std::vector<ssn_batch*> slots;
    uint64_t next_seq;
    cudaStream_t main_stream;
    cudaEvent_t t0, t1;
    uint64_t launches;
    int num_sms, ctas_per_sm_cap;
    size_t smem_optin, smem_static;
    uint32_t smem_words_default;
    void *d_flush; size_t flush_bytes;
    std::vector<KernelTimer> timers;
    size_t timer_next;
    double kernel_ms; uint64_t kernel_n;
    double prefilter_ms, score_ms;
    size_t l2_window_bytes;
    /* exact hom-reference pre-filter (ssn_prefilter.cuh) */
    ssn_prefilter_tables pf;
    float *d_pf_ab, *d_pf_h, *d_pf_F, *d_pf_vt;
    DevPrefilter dpf;
    bool pf_on;
    double survivor_frac;     /* survivors / columns of the last collected batch: sizes the list-mode grid */
    int last_slot;            /* slot of the previous launch: another slot's survivors' kernel may still be running */
    uint64_t reads_cap_bytes; uint32_t reads_cap_reads, reads_cap_windows;   /* ssn_reads_reserve */
    double pileup_ms; uint64_t pileup_n;
};

extern "C" ssn_status ssn_params_default(ssn_params *p)
{
    if (!p) { ssn_set_error("null params"); return SSN_ERR_ARG; }
    memset(p, 0, sizeof(*p));
    p->theta = 0.85f; p->het_rate = 0.001f; p->eta = 0.03f; p->n_hap = 2; p->cap_mapq = 60;
    p->min_somatic_qual = 15; p->include_loh = 1; p->include_gor = 1; p->use_priors = 1;
    p->use_joint_priors = 0; p->somatic_mutation_rate = 0.01;
    return SSN_OK;
}

extern "C" int ssn_nt16_of_char(int ch)
{
    static const char rev[] = "=ACMGRSVTWYHKDBN";
    ch &= 0xff;
    if (ch >= '0' && ch <= '3') return 1 << (ch - '0');
    for (int i = 0; i < 16; ++i) {
        if (ch == rev[i]) return i;
        if (rev[i] >= 'A' && rev[i] <= 'Z' && ch == rev[i] + ('a' - 'A')) return i;
    }
    return 15;
}

static size_t pf_smem_bytes() { return (size_t)kPfWarps * kPfStageBytes + 256; }   /* + slack for the words read past a tile's end */

static size_t smem_bytes_for(uint32_t words_per_warp)
{
    /* dynamic part only; the model tables and key LUTs are static shared memory (CtaTables) */
    return (size_t)kWarpsPerCta * (kSlotWords + words_per_warp) * 4 + 64;   /* + slack for the staged words read past a tile's end */
}

static ssn_status slot_alloc(ssn_ctx *ctx, ssn_batch *b, uint32_t max_columns, uint64_t max_entries)
{
    b->ctx = ctx; b->state = SLOT_FREE; b->max_columns = max_columns; b->max_entries = max_entries;
    const size_t ent = (size_t)max_entries + 64;
    /* the pinned column arrays are allocated when a caller first asks for them (host_columns): a run that ships reads
     * (ssn_batch_submit_reads) never does */
    CU(cudaMallocHost(&b->h_count, kCountWords * sizeof(uint32_t)));
    CU(cudaMalloc(&b->d_off, sizeof(uint32_t) * (2 * (size_t)max_columns + 1)));
    CU(cudaMalloc(&b->d_bs, ent));
    CU(cudaMalloc(&b->d_bq, ent));
    CU(cudaMalloc(&b->d_mq, ent));
    CU(cudaMalloc(&b->d_ref, (size_t)max_columns + 64));
    CU(cudaMalloc(&b->d_count, kCountWords * sizeof(uint32_t)));
    CU(cudaMalloc(&b->d_survivors, sizeof(uint32_t) * ((size_t)max_columns + 32)));
    b->site_cap = std::max<uint32_t>(4096u, max_columns / 16u);
    CU(cudaMalloc(&b->d_sites, sizeof(ssn_site) * (size_t)b->site_cap));
    b->h_sites_cap = 4096;
    CU(cudaMallocHost(&b->h_sites, sizeof(ssn_site) * (size_t)b->h_sites_cap));
    {
        /* highest priority: scoring launches go in front of whatever else shares the device (the readers' inflate lanes) */
        int lo = 0, hi = 0;
        static const int flat = getenv("SSN_FLAT_PRIORITY") ? atoi(getenv("SSN_FLAT_PRIORITY")) : 0;      /* measurement knob */
        CU(cudaDeviceGetStreamPriorityRange(&lo, &hi));
        CU(cudaStreamCreateWithPriority(&b->stream, cudaStreamNonBlocking, flat ? lo : hi));
    }
    CU(cudaEventCreateWithFlags(&b->done, cudaEventDisableTiming));
    return SSN_OK;
}

/* pinned host arrays of the column batch (ssn_batch_view), on first use */
static ssn_status host_columns(ssn_batch *b)
{
    if (b->h_off) return SSN_OK;
    const size_t ent = (size_t)b->max_entries + 64;
    CU(cudaSetDevice(b->ctx->device));
    CU(cudaMallocHost(&b->h_off, sizeof(uint32_t) * (2 * (size_t)b->max_columns + 1)));
    CU(cudaMallocHost(&b->h_bs, ent));
    CU(cudaMallocHost(&b->h_bq, ent));
    CU(cudaMallocHost(&b->h_mq, ent));
    CU(cudaMallocHost(&b->h_ref, (size_t)b->max_columns + 64));
    return SSN_OK;
}

static void slot_free(ssn_batch *b)
{
    if (!b) return;
    cudaFreeHost(b->h_off); cudaFreeHost(b->h_bs); cudaFreeHost(b->h_bq); cudaFreeHost(b->h_mq);
    cudaFreeHost(b->h_ref); cudaFreeHost(b->h_count); cudaFreeHost(b->h_sites); cudaFreeHost(b->h_dump);
    cudaFree(b->d_off); cudaFree(b->d_bs); cudaFree(b->d_bq); cudaFree(b->d_mq); cudaFree(b->d_ref);
    cudaFree(b->d_count); cudaFree(b->d_sites); cudaFree(b->d_dump); cudaFree(b->d_survivors); cudaFree(b->d_scratch);
    cudaFreeHost(b->h_recs); cudaFreeHost(b->h_index);
    cudaFreeHost(b->h_summary); cudaFreeHost(b->h_colpos);
    cudaFree(b->d_recs); cudaFree(b->d_index); cudaFree(b->d_A); cudaFree(b->d_P);
    cudaFree(b->d_wspan); cudaFree(b->d_summary); cudaFree(b->d_colpos); cudaFree(b->d_rbeg); cudaFree(b->d_rinfo); cudaFree(b->d_wtot);
    if (b->piled) cudaEventDestroy(b->piled);
    if (b->pa) cudaEventDestroy(b->pa);
    if (b->pb) cudaEventDestroy(b->pb);
    if (b->stream) cudaStreamDestroy(b->stream);
    if (b->done) cudaEventDestroy(b->done);
    delete b;
}

static double wall_now()
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

extern "C" ssn_status ssn_create(const ssn_params *p, int device, uint32_t n_slots, uint32_t max_columns,
                                 uint64_t max_entries, ssn_ctx **out)
{
    const bool trace = getenv("SSN_TRACE_SETUP") != nullptr;
    double tt = wall_now();
#define SSN_TRACE(what) do { if (trace) { double n_ = wall_now(); fprintf(stderr, "[ssn_create] %-28s %.3f s\n", what, n_ - tt); tt = n_; } } while (0)
    if (!p || !out || n_slots == 0 || max_columns == 0) { ssn_set_error("ssn_create: bad argument"); return SSN_ERR_ARG; }
    if (max_entries >= (1ull << 32) - 64) { ssn_set_error("ssn_create: max_entries must stay below 2^32"); return SSN_ERR_ARG; }
    ssn_ctx *ctx = new ssn_ctx();
    ctx->device = device; ctx->params = *p; ctx->next_seq = 1; ctx->launches = 0;
    ctx->d_flush = nullptr; ctx->flush_bytes = 0;
    ctx->timer_next = 0; ctx->kernel_ms = 0; ctx->kernel_n = 0; ctx->l2_window_bytes = 0;
    ctx->prefilter_ms = 0; ctx->score_ms = 0; ctx->d_pf_ab = ctx->d_pf_h = ctx->d_pf_F = ctx->d_pf_vt = nullptr; ctx->pf_on = false;
    memset(&ctx->pf, 0, sizeof(ctx->pf));
    memset(&ctx->tab, 0, sizeof(ctx->tab));
    ctx->survivor_frac = 1.0; ctx->last_slot = -1;
    /* the model tables (x87 long double, ~0.5 s of host arithmetic) are built on a thread of their own while this one
     * creates the CUDA context, the streams and the batch slots */
    int tables_rc = 0, pf_rc = 0;
    std::thread table_thread([&]() {
        tables_rc = ssn_tables_build(p, &ctx->tab);
        if (tables_rc == 0) pf_rc = ssn_prefilter_build(&ctx->tab, (int)(ctx->tab.q_r + 0.5), &ctx->pf);
    });
    struct Joiner { std::thread &t; ~Joiner() { if (t.joinable()) t.join(); } } table_joiner{table_thread};
    auto fail_early = [&](ssn_status st) {       /* before any device object exists */
        table_thread.join();
        ssn_prefilter_free(&ctx->pf);
        if (tables_rc == 0) ssn_tables_free(&ctx->tab);
        delete ctx;
        return st;
    };
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess) { ssn_set_error("ssn_create: no usable CUDA driver/device (%s)", cudaGetErrorString(cudaGetLastError())); return fail_early(SSN_ERR_CUDA); }
    if (device < 0 || device >= ndev) { ssn_set_error("ssn_create: no CUDA device %d (found %d)", device, ndev); return fail_early(SSN_ERR_CUDA); }
    if (cudaSetDevice(device) != cudaSuccess) { ssn_set_error("ssn_create: cudaSetDevice(%d) failed", device); return fail_early(SSN_ERR_CUDA); }
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) { ssn_set_error("ssn_create: cudaGetDeviceProperties failed"); return fail_early(SSN_ERR_CUDA); }
    SSN_TRACE("device query");
    if (prop.major < 10) { ssn_set_error("ssn_create: device %d is sm_%d%d; this library is built for sm_100a only", device, prop.major, prop.minor); return fail_early(SSN_ERR_CUDA); }
    ctx->num_sms = prop.multiProcessorCount; ctx->smem_optin = prop.sharedMemPerBlockOptin;
    const char *env = getenv("SSN_SMEM_WORDS");
    ctx->smem_words_default = env ? (uint32_t)atoi(env) : (uint32_t)kFastWords * 32u;
    ctx->smem_words_default = (ctx->smem_words_default + 31u) / 32u * 32u;

    CU(cudaFree(0));
    SSN_TRACE("CUDA context");
    CU(cudaStreamCreateWithFlags(&ctx->main_stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&ctx->t0));
    CU(cudaEventCreate(&ctx->t1));
    cudaFuncAttributes fa;
    CU(cudaFuncGetAttributes(&fa, ssn_score_kernel<true>));
    ctx->smem_static = fa.sharedSizeBytes;
    const int max_dyn = (int)(ctx->smem_optin - ctx->smem_static);
    CU(cudaFuncSetAttribute(ssn_score_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn));
    CU(cudaFuncSetAttribute(ssn_score_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, max_dyn));
    CU(cudaFuncSetAttribute(ssn_prefilter_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pf_smem_bytes()));
    const size_t coef_n = (size_t)64 * 65536, lhet_n = 65536;
    CU(cudaMalloc(&ctx->d_coef_lhet, sizeof(double) * (coef_n + lhet_n)));

    /* keep coef+lhet resident in L2: persisting carve-out + per-stream access-policy window */
    size_t want = sizeof(double) * (coef_n + lhet_n);
    size_t persist = std::min<size_t>(want, (size_t)prop.persistingL2CacheMaxSize);
    if (getenv("SSN_NO_L2_WINDOW")) persist = 0;               /* measurement knob: the tables compete for L2 like any other line */
    if (persist > 0 && prop.accessPolicyMaxWindowSize > 0) {
        CU(cudaDeviceSetLimit(cudaLimitPersistingL2CacheSize, persist));
        ctx->l2_window_bytes = std::min<size_t>(want, (size_t)prop.accessPolicyMaxWindowSize);
    }
    SSN_TRACE("kernel attributes, L2 limit");
    for (uint32_t i = 0; i < n_slots; ++i) {
        ssn_batch *b = new ssn_batch();
        memset((void*)b, 0, sizeof(*b));
        b->index = (int)i;
        ctx->slots.push_back(b);
        ssn_status st = slot_alloc(ctx, b, max_columns, max_entries);
        if (st != SSN_OK) { table_thread.join(); ssn_destroy(ctx); return st; }
        if (ctx->l2_window_bytes) {
            cudaStreamAttrValue attr;
            memset(&attr, 0, sizeof(attr));
            attr.accessPolicyWindow.base_ptr = ctx->d_coef_lhet;
            attr.accessPolicyWindow.num_bytes = ctx->l2_window_bytes;
            attr.accessPolicyWindow.hitRatio = (float)std::min(1.0, (double)persist / (double)ctx->l2_window_bytes);
            attr.accessPolicyWindow.hitProp = cudaAccessPropertyPersisting;
            attr.accessPolicyWindow.missProp = cudaAccessPropertyStreaming;
            CU(cudaStreamSetAttribute(b->stream, cudaStreamAttributeAccessPolicyWindow, &attr));
        }
    }
    SSN_TRACE("batch slots");
    ctx->timers.resize(64);
    for (auto &t : ctx->timers) {
        CU(cudaEventCreate(&t.a));
        CU(cudaEventCreate(&t.m));
        CU(cudaEventCreate(&t.b));
        t.pending = false; t.filtered = false;
    }

    table_thread.join();
    SSN_TRACE("wait for the host tables");
    if (tables_rc != 0) { ssn_set_error("ssn_create: table construction failed (n_hap=%d)", p->n_hap); ssn_destroy(ctx); return SSN_ERR_ARG; }
    if (pf_rc != 0) { ssn_set_error("ssn_create: out of memory (pre-filter tables)"); ssn_destroy(ctx); return SSN_ERR_NOMEM; }
    CU(cudaMemcpy(ctx->d_coef_lhet, ctx->tab.coef, sizeof(double) * coef_n, cudaMemcpyHostToDevice));
    CU(cudaMemcpy(ctx->d_coef_lhet + coef_n, ctx->tab.lhet, sizeof(double) * lhet_n, cudaMemcpyHostToDevice));
    CU(cudaMalloc(&ctx->d_fk, sizeof(double) * 256));
    CU(cudaMemcpy(ctx->d_fk, ctx->tab.fk, sizeof(double) * 256, cudaMemcpyHostToDevice));
    CU(cudaMalloc(&ctx->d_prior, sizeof(ctx->tab.prior)));
    CU(cudaMemcpy(ctx->d_prior, ctx->tab.prior, sizeof(ctx->tab.prior), cudaMemcpyHostToDevice));
    CU(cudaMalloc(&ctx->d_jointprior, sizeof(ctx->tab.jointprior)));
    CU(cudaMemcpy(ctx->d_jointprior, ctx->tab.jointprior, sizeof(ctx->tab.jointprior), cudaMemcpyHostToDevice));
    CU(cudaMalloc(&ctx->d_qadd, sizeof(ctx->tab.qadd)));
    CU(cudaMemcpy(ctx->d_qadd, ctx->tab.qadd, sizeof(ctx->tab.qadd), cudaMemcpyHostToDevice));
    DevModel &M = ctx->model;
    M.coef = ctx->d_coef_lhet; M.lhet = ctx->d_coef_lhet + coef_n; M.fk = ctx->d_fk;
    M.prior = ctx->d_prior; M.jointprior = ctx->d_jointprior; M.qadd = ctx->d_qadd;
    M.q_r_int = (int)(ctx->tab.q_r + 0.5);                 /* sniper_maqcns_call, sniper_maqcns.c:279 */
    M.cap_mapq = p->cap_mapq; M.min_somatic_qual = p->min_somatic_qual;
    M.include_loh = p->include_loh; M.include_gor = p->include_gor; M.use_joint = p->use_joint_priors;

    ctx->pf_on = ctx->pf.enabled && getenv("SSN_NO_PREFILTER") == nullptr;
    CU(cudaMalloc(&ctx->d_pf_ab, sizeof(float) * 4 * SSN_PF_N * SSN_PF_N));
    CU(cudaMemcpy(ctx->d_pf_ab, ctx->pf.ab, sizeof(float) * 4 * SSN_PF_N * SSN_PF_N, cudaMemcpyHostToDevice));
    CU(cudaMalloc(&ctx->d_pf_h, sizeof(float) * SSN_PF_N * SSN_PF_N));
    CU(cudaMemcpy(ctx->d_pf_h, ctx->pf.h, sizeof(float) * SSN_PF_N * SSN_PF_N, cudaMemcpyHostToDevice));
    CU(cudaMalloc(&ctx->d_pf_F, sizeof(float) * (SSN_PF_N + 1)));
    CU(cudaMemcpy(ctx->d_pf_F, ctx->pf.F, sizeof(float) * (SSN_PF_N + 1), cudaMemcpyHostToDevice));
    CU(cudaMalloc(&ctx->d_pf_vt, sizeof(float) * SSN_PF_N * SSN_PF_VK * SSN_PF_VS));
    CU(cudaMemcpy(ctx->d_pf_vt, ctx->pf.vt, sizeof(float) * SSN_PF_N * SSN_PF_VK * SSN_PF_VS, cudaMemcpyHostToDevice));
    ctx->dpf.ab = reinterpret_cast<const float4*>(ctx->d_pf_ab); ctx->dpf.h = ctx->d_pf_h; ctx->dpf.F = ctx->d_pf_F; ctx->dpf.vt = ctx->d_pf_vt;
    SSN_TRACE("table upload");
    *out = ctx;
    return SSN_OK;
}

extern "C" void ssn_destroy(ssn_ctx *ctx)
{
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaDeviceSynchronize();
    for (ssn_batch *b : ctx->slots) slot_free(b);
    for (auto &t : ctx->timers) { if (t.a) cudaEventDestroy(t.a); if (t.m) cudaEventDestroy(t.m); if (t.b) cudaEventDestroy(t.b); }
    cudaFree(ctx->d_pf_ab); cudaFree(ctx->d_pf_h); cudaFree(ctx->d_pf_F); cudaFree(ctx->d_pf_vt);
    ssn_prefilter_free(&ctx->pf);
    cudaFree(ctx->d_coef_lhet); cudaFree(ctx->d_fk); cudaFree(ctx->d_prior); cudaFree(ctx->d_jointprior);
    cudaFree(ctx->d_qadd); cudaFree(ctx->d_flush);
    if (ctx->main_stream) cudaStreamDestroy(ctx->main_stream);
    if (ctx->t0) cudaEventDestroy(ctx->t0);
    if (ctx->t1) cudaEventDestroy(ctx->t1);
    ssn_tables_free(&ctx->tab);
    delete ctx;
}

extern "C" ssn_status ssn_get_tables(const ssn_ctx *ctx, ssn_tables *out)
{
    if (!ctx || !out) { ssn_set_error("ssn_get_tables: null"); return SSN_ERR_ARG; }
    out->coef = ctx->tab.coef; out->lhet = ctx->tab.lhet; out->fk = ctx->tab.fk; out->q_r = ctx->tab.q_r;
    out->prior = ctx->tab.prior; out->jointprior = ctx->tab.jointprior; out->qadd = ctx->tab.qadd;
    return SSN_OK;
}

extern "C" ssn_status ssn_host_tables_create(const ssn_params *p, ssn_tables *view, void **handle)
{
    if (!p || !view || !handle) { ssn_set_error("ssn_host_tables_create: null"); return SSN_ERR_ARG; }
    ssn_host_tables *t = new ssn_host_tables();
    if (ssn_tables_build(p, t) != 0) { delete t; ssn_set_error("table construction failed (n_hap=%d)", p->n_hap); return SSN_ERR_ARG; }
    view->coef = t->coef; view->lhet = t->lhet; view->fk = t->fk; view->q_r = t->q_r;
    view->prior = t->prior; view->jointprior = t->jointprior; view->qadd = t->qadd;
    *handle = t;
    return SSN_OK;
}

extern "C" void ssn_host_tables_destroy(void *handle)
{
    ssn_host_tables *t = (ssn_host_tables*)handle;
    if (!t) return;
    ssn_tables_free(t);
    delete t;
}

/* ------------------------------------------------------------------------------------------------ */

extern "C" ssn_status ssn_batch_acquire(ssn_ctx *ctx, ssn_batch **out)
{
    if (!ctx || !out) { ssn_set_error("ssn_batch_acquire: null"); return SSN_ERR_ARG; }
    ssn_batch *pick = nullptr;
    for (ssn_batch *b : ctx->slots) if (b->state == SLOT_FREE) { pick = b; break; }
    if (!pick) {
        for (ssn_batch *b : ctx->slots)
            if (b->state == SLOT_DONE && (!pick || b->seq < pick->seq)) pick = b;
    }
    if (!pick) { ssn_set_error("ssn_batch_acquire: all %zu slots are in flight; drain ssn_results_next first", ctx->slots.size()); return SSN_ERR_STATE; }
    pick->state = SLOT_FILLING; pick->ncol = 0; pick->nent = 0; pick->resident = false; pick->n_sites = 0;
    pick->reads_pending = false; pick->from_reads = false; pick->spec = false;
    *out = pick;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_release(ssn_ctx *ctx, ssn_batch *b)
{
    if (!ctx || !b || b->ctx != ctx) { ssn_set_error("ssn_batch_release: bad batch"); return SSN_ERR_ARG; }
    if (b->state == SLOT_SUBMITTED) { ssn_set_error("ssn_batch_release: batch still in flight"); return SSN_ERR_STATE; }
    b->state = SLOT_FREE;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_view_get(ssn_batch *b, ssn_batch_view *out)
{
    if (!b || !out) { ssn_set_error("ssn_batch_view_get: null"); return SSN_ERR_ARG; }
    ssn_status hs = host_columns(b);
    if (hs != SSN_OK) return hs;
    out->off = b->h_off; out->bs = b->h_bs; out->bq = b->h_bq; out->mq = b->h_mq; out->ref_nt16 = b->h_ref;
    out->max_columns = b->max_columns; out->max_entries = b->max_entries;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_set_size(ssn_batch *b, uint32_t n_columns, uint64_t n_entries)
{
    if (!b) { ssn_set_error("ssn_batch_set_size: null"); return SSN_ERR_ARG; }
    if (n_columns > b->max_columns || n_entries > b->max_entries) {
        ssn_set_error("ssn_batch_set_size: %u columns / %llu entries exceed the slot capacity %u / %llu", n_columns,
                      (unsigned long long)n_entries, b->max_columns, (unsigned long long)b->max_entries);
        return SSN_ERR_ARG;
    }
    b->ncol = n_columns; b->nent = n_entries; b->resident = false;
    return SSN_OK;
}

/* Validates the offsets and finds the largest key-store footprint of any 32-column tile. */
static ssn_status scan_offsets(ssn_batch *b)
{
    const uint32_t *off = b->h_off;
    const uint32_t ncol = b->ncol;
    b->max_col_words = 0; b->n_deep = 0;
    if (ncol == 0) { b->max_tile_words = 0; return SSN_OK; }
    if (off[2 * (size_t)ncol] != b->nent) { ssn_set_error("batch: off[2*ncol]=%u but n_entries=%llu", off[2 * (size_t)ncol], (unsigned long long)b->nent); return SSN_ERR_ARG; }
    uint32_t max_tile = 0;
    for (uint32_t t = 0; t < ncol; t += 32) {
        const uint32_t e = std::min(ncol, t + 32);
        uint32_t words = 0;
        for (uint32_t c = t; c < e; ++c) {
            const uint32_t a = off[2 * (size_t)c], m = off[2 * (size_t)c + 1], z = off[2 * (size_t)c + 2];
            if (m < a || z < m) { ssn_set_error("batch: offsets of column %u decrease", c); return SSN_ERR_ARG; }
            const uint32_t n = z - a;
            if (m - a > 255u || z - m > 255u) b->n_deep += 1;
            if (n > (uint32_t)kRunLen * kMaxRuns) { ssn_set_error("batch: column %u holds %u read-bases; this build supports %d", c, n, kRunLen * kMaxRuns); return SSN_ERR_ARG; }
            const uint32_t cw = (n >> 10) * (kRunLen / 2) + (((n & (kRunLen - 1)) + 1) >> 1);
            words += cw;
            b->max_col_words = std::max(b->max_col_words, cw);
        }
        max_tile = std::max(max_tile, words);
    }
    b->max_tile_words = max_tile;
    return SSN_OK;
}

static ssn_status copy_in(ssn_ctx *ctx, ssn_batch *b)
{
    (void)ctx;
    if (b->ncol == 0) return SSN_OK;
    CU(cudaMemcpyAsync(b->d_off, b->h_off, sizeof(uint32_t) * (2 * (size_t)b->ncol + 1), cudaMemcpyHostToDevice, b->stream));
    CU(cudaMemcpyAsync(b->d_bs, b->h_bs, b->nent, cudaMemcpyHostToDevice, b->stream));
    /* the kernels' clean-tile check scans the last tile's base qualities up to the next 16-byte boundary: pad with a
     * legal quality so stale bytes behind the batch cannot send that tile down the slow path (the slot has 64 spare bytes) */
    memset(b->h_bq + b->nent, 1, 16);
    CU(cudaMemcpyAsync(b->d_bq, b->h_bq, b->nent + 16, cudaMemcpyHostToDevice, b->stream));
    CU(cudaMemcpyAsync(b->d_mq, b->h_mq, b->nent, cudaMemcpyHostToDevice, b->stream));
    CU(cudaMemcpyAsync(b->d_ref, b->h_ref, b->ncol, cudaMemcpyHostToDevice, b->stream));
    return SSN_OK;
}

static ssn_status timer_collect(ssn_ctx *ctx, KernelTimer &kt)
{
    if (!kt.pending) return SSN_OK;
    float ms = 0, fm = 0;
    CU(cudaEventSynchronize(kt.b));
    CU(cudaEventElapsedTime(&ms, kt.a, kt.b));
    if (kt.filtered) CU(cudaEventElapsedTime(&fm, kt.a, kt.m));
    ctx->kernel_ms += ms; ctx->prefilter_ms += fm; ctx->score_ms += ms - fm; ctx->kernel_n += 1;
    kt.pending = false;
    return SSN_OK;
}

static ssn_status launch(ssn_ctx *ctx, ssn_batch *b, uint32_t flags)
{
    const bool dump = (flags & SSN_SUBMIT_COLUMN_DUMP) != 0;
    /* the pre-filter runs unless switched off; with a column dump it only records its decisions (audit) */
    /* a batch of panel depth: the filter would hand (nearly) every column on, and contiguous tiles score faster than
     * gathered ones */
    const bool mostly_deep = !b->from_reads && b->ncol && 2ull * b->n_deep > b->ncol;
    const bool filter = ctx->pf_on && !(flags & SSN_SUBMIT_NO_PREFILTER) && (!mostly_deep || (flags & SSN_SUBMIT_COLUMN_DUMP));
    const bool listed = filter && !dump;
    b->flags = flags; b->listed = listed;
    CU(cudaMemsetAsync(b->d_count, 0, kCountWords * sizeof(uint32_t), b->stream));
    if (b->ncol == 0) return SSN_OK;
    if (dump) {
        const size_t need = (size_t)b->ncol * SSN_COLUMN_DUMP_WORDS;
        if (b->d_dump_words < need) {
            CU(cudaStreamSynchronize(b->stream));
            cudaFree(b->d_dump); cudaFreeHost(b->h_dump);
            b->d_dump = nullptr; b->h_dump = nullptr; b->d_dump_words = b->h_dump_words = 0;
            CU(cudaMalloc(&b->d_dump, need * 4));
            CU(cudaMallocHost(&b->h_dump, need * 4));
            b->d_dump_words = b->h_dump_words = need;
        }
        if (!filter) CU(cudaMemsetAsync(b->d_dump, 0, need * 4, b->stream));   /* word 29 = "not dropped" */
    }
    /* shared-memory key store per warp: kFastWords*32 words serve every fast tile (<= 128 key slots per
     * column); deeper tiles use it too when they fit and the global scratch otherwise */
    const uint32_t words = ctx->smem_words_default;
    size_t smem = smem_bytes_for(words);
    if (smem + ctx->smem_static > ctx->smem_optin) { ssn_set_error("shared memory request %zu exceeds %zu", smem, ctx->smem_optin); return SSN_ERR_ARG; }
    int per_sm = 0;
    if (dump) CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ssn_score_kernel<true>, kWarpsPerCta * 32, smem));
    else CU(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, ssn_score_kernel<false>, kWarpsPerCta * 32, smem));
    if (per_sm < 1) { ssn_set_error("kernel does not fit on an SM with %zu bytes of shared memory", smem); return SSN_ERR_CUDA; }
    const uint32_t ntiles = (b->ncol + 31u) / 32u;
    uint32_t grid = (uint32_t)(ctx->num_sms * per_sm);
    const uint32_t need_ctas = (ntiles + kWarpsPerCta - 1) / kWarpsPerCta;
    if (grid > need_ctas && !listed) grid = need_ctas;
    uint32_t list_width = 0, list_generic_auto = 0;
    if (listed) {
        /* The survivor count is only known on the device.  Size the grid from the last collected batch (x1.5, 16-column
         * tiles, one tile per warp) so that the SMs it does not need stay free for the next batch's pre-filter kernel on
         * another stream; the persistent tile loop keeps any estimate correct. */
        static const int tiles_per_warp = getenv("SSN_LIST_TPW") ? std::max(1, atoi(getenv("SSN_LIST_TPW"))) : 1;
        const double est = std::min(1.0, ctx->survivor_frac * 1.25) * (double)b->ncol;
        /* batches alternating between slots: 16-column tiles on few SMs (throughput; the rest filter the next batch).
         * One slot used back to back: nothing overlaps, so spread 4-column tiles over more SMs (latency). */
        const bool alternating = ctx->last_slot >= 0 && ctx->last_slot != b->index;
        list_width = alternating ? 0u : 4u;
        uint32_t cols_per_warp = alternating ? (uint32_t)kListWidth : 4u;
        /* One batch in flight and no more survivors than resident warps: one column per warp through the warp-cooperative
         * sorting path.  Its code is short (no instruction-cache misses on the unrolled networks), which is what a
         * latency-bound launch of a single tile per warp needs: 46 us instead of 72 us on the bench workload. */
        if (!alternating && est <= (double)grid * kWarpsPerCta) { list_width = 1u; cols_per_warp = 1u; list_generic_auto = 1u; }
        { const char *e = getenv("SSN_LIST_W"); const int w = e ? atoi(e) : 0; if (w == 1 || w == 2 || w == 4 || w == 8 || w == 16) cols_per_warp = (uint32_t)w; }
        const uint32_t want = (uint32_t)(est / (cols_per_warp * kWarpsPerCta * tiles_per_warp)) + 1u;
        grid = std::min(grid, std::max(want, 8u));
    }
    /* a list-mode tile gathers up to 32 arbitrary columns: size the spill store for 32 of the deepest one */
    const uint32_t tile_words = listed ? 32u * b->max_col_words : b->max_tile_words;
    if (tile_words > words) {
        const size_t per_warp = ((size_t)tile_words + 63) / 64 * 64;
        const size_t need = per_warp * grid * kWarpsPerCta;
        if (b->scratch_words < need) {
            CU(cudaStreamSynchronize(b->stream));           /* only this slot's launches ever touch its scratch */
            cudaFree(b->d_scratch); b->d_scratch = nullptr; b->scratch_words = 0;
            CU(cudaMalloc(&b->d_scratch, need * 4));
            b->scratch_words = need;
        }
    }
    DevBatch B;
    B.off = b->d_off; B.bs = b->d_bs; B.bq = b->d_bq; B.mq = b->d_mq; B.ref_nt16 = b->d_ref; B.ncol = b->ncol;
    B.sites = b->d_sites; B.site_count = b->d_count; B.site_cap = b->site_cap;
    B.coldump = dump ? b->d_dump : nullptr;
    B.scratch = tile_words > words ? b->d_scratch : nullptr;
    B.scratch_words_per_warp = (uint32_t)(((size_t)tile_words + 63) / 64 * 64);
    B.smem_words_per_warp = words;
    B.dev_ncol = b->spec ? b->d_summary : nullptr; B.guard = b->spec ? b->d_summary : nullptr; B.guard_col_words = b->spec_col_words;
    { static const int nh = getenv("SSN_NO_HIST") != nullptr; B.no_hist = nh ? 1u : 0u; }
    B.collist = listed ? b->d_survivors : nullptr;
    B.list_count = b->d_count + 2;
    { const char *e = getenv("SSN_LIST_W"); const int w = e ? atoi(e) : 0; B.list_width = (w == 1 || w == 2 || w == 4 || w == 8 || w == 16) ? (uint32_t)w : list_width; }
    { static const int lg = getenv("SSN_LIST_GENERIC") ? atoi(getenv("SSN_LIST_GENERIC")) : 0; B.list_generic = lg ? (uint32_t)lg : list_generic_auto; }
    KernelTimer &kt = ctx->timers[ctx->timer_next];
    ctx->timer_next = (ctx->timer_next + 1) % ctx->timers.size();
    ssn_status st = timer_collect(ctx, kt);
    if (st != SSN_OK) return st;
    CU(cudaEventRecord(kt.a, b->stream));
    kt.filtered = filter;
    if (filter) {
        const uint32_t pf_need = (ntiles + kPfWarps - 1) / kPfWarps;
        /* leave the SMs the survivors' kernel of the previous batch (another slot stream) needs, so the two overlap */
        uint32_t pf_sms = (uint32_t)ctx->num_sms;
        if (listed && grid <= pf_sms / 4u && ctx->last_slot >= 0 && ctx->last_slot != b->index) pf_sms -= grid;
        ctx->last_slot = b->index;
        const uint32_t pf_grid = std::min<uint32_t>(pf_sms, pf_need);
        ssn_prefilter_kernel<<<pf_grid, kPfWarps * 32, pf_smem_bytes(), b->stream>>>(
            ctx->dpf, B, b->d_survivors, b->d_count, dump ? b->d_dump + (SSN_COLUMN_DUMP_WORDS - 1) : nullptr, SSN_COLUMN_DUMP_WORDS);
        CU(cudaGetLastError());
        CU(cudaEventRecord(kt.m, b->stream));
        ctx->launches += 1;
    }
    if (dump) ssn_score_kernel<true><<<grid, kWarpsPerCta * 32, smem, b->stream>>>(ctx->model, B);
    else ssn_score_kernel<false><<<grid, kWarpsPerCta * 32, smem, b->stream>>>(ctx->model, B);
    CU(cudaGetLastError());
    CU(cudaEventRecord(kt.b, b->stream));
    kt.pending = true;
    ctx->launches += 1;
    return SSN_OK;
}

static ssn_status copy_out_async(ssn_batch *b)
{
    CU(cudaMemcpyAsync(b->h_count, b->d_count, kCountWords * sizeof(uint32_t), cudaMemcpyDeviceToHost, b->stream));
    if ((b->flags & SSN_SUBMIT_COLUMN_DUMP) && b->ncol)
        CU(cudaMemcpyAsync(b->h_dump, b->d_dump, (size_t)b->ncol * SSN_COLUMN_DUMP_WORDS * 4, cudaMemcpyDeviceToHost, b->stream));
    CU(cudaEventRecord(b->done, b->stream));
    return SSN_OK;
}

/* after the batch's stream work is done: handle ring overflow, copy the sites, sort by column */
static ssn_status collect(ssn_ctx *ctx, ssn_batch *b)
{
    CU(cudaEventSynchronize(b->done));
    uint32_t n = b->h_count[0];
    if (n > b->site_cap) {
        /* ring too small: grow it and score the (still resident) batch again */
        CU(cudaStreamSynchronize(b->stream));
        cudaFree(b->d_sites); b->d_sites = nullptr;
        b->site_cap = n + n / 8 + 64;
        CU(cudaMalloc(&b->d_sites, sizeof(ssn_site) * (size_t)b->site_cap));
        ssn_status st = launch(ctx, b, b->flags);
        if (st != SSN_OK) return st;
        st = copy_out_async(b);
        if (st != SSN_OK) return st;
        CU(cudaEventSynchronize(b->done));
        n = b->h_count[0];
        if (n > b->site_cap) { ssn_set_error("site ring overflow persists (%u > %u)", n, b->site_cap); return SSN_ERR_OVERFLOW; }
    }
    if (n > b->h_sites_cap) {
        cudaFreeHost(b->h_sites); b->h_sites = nullptr;
        b->h_sites_cap = n + n / 8 + 64;
        CU(cudaMallocHost(&b->h_sites, sizeof(ssn_site) * (size_t)b->h_sites_cap));
    }
    if (n) {
        CU(cudaMemcpyAsync(b->h_sites, b->d_sites, sizeof(ssn_site) * (size_t)n, cudaMemcpyDeviceToHost, b->stream));
        CU(cudaStreamSynchronize(b->stream));
        std::sort(b->h_sites, b->h_sites + n, [](const ssn_site &x, const ssn_site &y) { return x.column < y.column; });
    }
    b->n_sites = n;
    if (b->listed && b->ncol) ctx->survivor_frac = (double)b->h_count[2] / (double)b->ncol;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_submit(ssn_ctx *ctx, ssn_batch *b, uint32_t flags)
{
    if (!ctx || !b || b->ctx != ctx) { ssn_set_error("ssn_batch_submit: bad batch"); return SSN_ERR_ARG; }
    if (b->state != SLOT_FILLING) { ssn_set_error("ssn_batch_submit: batch was not acquired"); return SSN_ERR_STATE; }
    if (!b->h_off) { ssn_set_error("ssn_batch_submit: the batch was never filled (ssn_batch_view_get)"); return SSN_ERR_STATE; }
    CU(cudaSetDevice(ctx->device));
    ssn_status st = scan_offsets(b);
    if (st != SSN_OK) return st;
    if ((st = copy_in(ctx, b)) != SSN_OK) return st;
    b->resident = true;
    if ((st = launch(ctx, b, flags)) != SSN_OK) return st;
    if ((st = copy_out_async(b)) != SSN_OK) return st;
    b->state = SLOT_SUBMITTED; b->seq = ctx->next_seq++;
    return SSN_OK;
}

/* ---- read batches: pileup on the device --------------------------------------------------------------- */

extern "C" ssn_status ssn_reads_reserve(ssn_ctx *ctx, uint64_t max_rec_bytes, uint32_t max_reads, uint32_t max_windows)
{
    if (!ctx || !max_rec_bytes || !max_reads || !max_windows) { ssn_set_error("ssn_reads_reserve: bad argument"); return SSN_ERR_ARG; }
    if (max_rec_bytes >= (1ull << 34) || (uint64_t)max_windows * SSN_READ_WINDOW >= (1ull << 32)) { ssn_set_error("ssn_reads_reserve: capacity too large"); return SSN_ERR_ARG; }
    for (ssn_batch *b : ctx->slots) if (b->reads_alloc) { ssn_set_error("ssn_reads_reserve: read buffers already exist"); return SSN_ERR_STATE; }
    ctx->reads_cap_bytes = (max_rec_bytes + 3) & ~3ull; ctx->reads_cap_reads = max_reads; ctx->reads_cap_windows = max_windows;
    return SSN_OK;
}

static ssn_status reads_alloc(ssn_ctx *ctx, ssn_batch *b)
{
    if (b->reads_alloc) return SSN_OK;
    if (!ctx->reads_cap_bytes) { ssn_set_error("call ssn_reads_reserve first"); return SSN_ERR_STATE; }
    const size_t nb = ctx->reads_cap_bytes, nr = ctx->reads_cap_reads, nw = ctx->reads_cap_windows;
    CU(cudaSetDevice(ctx->device));
    CU(cudaMallocHost(&b->h_recs, nb));
    /* window table, reference slices and record index share one block (one copy per batch): [win][wref][rec_off] */
    const size_t o_wref = (nw * sizeof(ssn_read_window) + 255) & ~(size_t)255, o_off = o_wref + nw * SSN_READ_WINDOW;
    b->index_off_bytes = o_off;
    CU(cudaMallocHost(&b->h_index, o_off + nr * 4));
    b->h_win = (ssn_read_window*)b->h_index; b->h_wref = b->h_index + o_wref; b->h_rec_off = (uint32_t*)(b->h_index + o_off);
    CU(cudaMallocHost(&b->h_summary, pileup::kSummaryWords * 4));
    CU(cudaMallocHost(&b->h_colpos, (size_t)b->max_columns * 4));
    CU(cudaMalloc(&b->d_recs, nb));
    CU(cudaMalloc(&b->d_index, o_off + nr * 4));
    b->d_win = (ssn_read_window*)b->d_index; b->d_wref = b->d_index + o_wref; b->d_rec_off = (uint32_t*)(b->d_index + o_off);
    CU(cudaMalloc(&b->d_A, nw * 2 * SSN_READ_WINDOW * 4));
    CU(cudaMalloc(&b->d_P, nw * SSN_READ_WINDOW * 4));
    CU(cudaMalloc(&b->d_wspan, nw * 2 * 4));
    CU(cudaMalloc(&b->d_wtot, nw * sizeof(uint2)));
    CU(cudaMalloc(&b->d_rbeg, nr * 4));
    CU(cudaMalloc(&b->d_rinfo, nr * 2 * sizeof(uint4)));
    CU(cudaMalloc(&b->d_summary, pileup::kSummaryWords * 4));
    CU(cudaMalloc(&b->d_colpos, (size_t)b->max_columns * 4));
    CU(cudaEventCreateWithFlags(&b->piled, cudaEventDisableTiming));
    CU(cudaEventCreate(&b->pa));
    CU(cudaEventCreate(&b->pb));
    CU(cudaFuncSetAttribute(pileup::fill_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)pileup::kFillSmem));
    b->reads_alloc = true;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_reads_view(ssn_ctx *ctx, ssn_batch *b, ssn_reads_view *out)
{
    if (!ctx || !b || b->ctx != ctx || !out) { ssn_set_error("ssn_batch_reads_view: bad argument"); return SSN_ERR_ARG; }
    ssn_status st = reads_alloc(ctx, b);
    if (st != SSN_OK) return st;
    out->recs = b->h_recs; out->rec_off = b->h_rec_off; out->win = b->h_win; out->ref_nt16 = b->h_wref;
    out->max_rec_bytes = ctx->reads_cap_bytes; out->max_reads = ctx->reads_cap_reads; out->max_windows = ctx->reads_cap_windows;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_submit_reads(ssn_ctx *ctx, ssn_batch *b, uint32_t n_windows, uint32_t n_reads, uint64_t rec_bytes,
                                             uint32_t flags)
{
    if (!ctx || !b || b->ctx != ctx) { ssn_set_error("ssn_batch_submit_reads: bad batch"); return SSN_ERR_ARG; }
    if (b->state != SLOT_FILLING) { ssn_set_error("ssn_batch_submit_reads: batch was not acquired"); return SSN_ERR_STATE; }
    if (!b->reads_alloc) { ssn_set_error("ssn_batch_submit_reads: no read buffers (ssn_batch_reads_view)"); return SSN_ERR_STATE; }
    if (n_windows > ctx->reads_cap_windows || n_reads > ctx->reads_cap_reads || rec_bytes > ctx->reads_cap_bytes || (rec_bytes & 3)) {
        ssn_set_error("ssn_batch_submit_reads: %u windows / %u reads / %llu bytes exceed the reserved %u / %u / %llu", n_windows, n_reads,
                      (unsigned long long)rec_bytes, ctx->reads_cap_windows, ctx->reads_cap_reads, (unsigned long long)ctx->reads_cap_bytes);
        return SSN_ERR_ARG;
    }
    /* the device trusts the index: check what is cheap to check */
    {
        uint64_t reads = 0;
        for (uint32_t w = 0; w < n_windows; ++w) {
            const ssn_read_window &W = b->h_win[w];
            if ((W.lo & (SSN_READ_WINDOW - 1u)) || (uint64_t)W.rec_first[0] + W.n_reads[0] > n_reads || (uint64_t)W.rec_first[1] + W.n_reads[1] > n_reads) {
                ssn_set_error("ssn_batch_submit_reads: window %u is malformed", w);
                return SSN_ERR_ARG;
            }
            reads += (uint64_t)W.n_reads[0] + W.n_reads[1];
        }
        if (reads > n_reads) { ssn_set_error("ssn_batch_submit_reads: windows hold %llu reads, the index %u", (unsigned long long)reads, n_reads); return SSN_ERR_ARG; }
    }
    CU(cudaSetDevice(ctx->device));
    b->flags = flags; b->n_win = n_windows; b->from_reads = true; b->ncol = 0; b->nent = 0; b->resident = false;
    cudaStream_t st = b->stream;
    CU(cudaMemsetAsync(b->d_summary, 0, pileup::kSummaryWords * 4, st));
    if (n_windows) {
        CU(cudaMemcpyAsync(b->d_recs, b->h_recs, rec_bytes, cudaMemcpyHostToDevice, st));
        CU(cudaMemcpyAsync(b->d_index, b->h_index, b->index_off_bytes + (size_t)n_reads * 4, cudaMemcpyHostToDevice, st));
        pileup::DevReads R;
        R.recs = (const uint32_t*)b->d_recs; R.rec_off = b->d_rec_off; R.win = b->d_win; R.wref = b->d_wref; R.n_win = n_windows; R.n_reads = n_reads; R.rec_words = (uint32_t)(rec_bytes >> 2);
        R.A = b->d_A; R.P = b->d_P; R.rbeg = b->d_rbeg; R.rinfo = b->d_rinfo; R.wspan = b->d_wspan; R.wtot = b->d_wtot; R.summary = b->d_summary;
        R.off = b->d_off; R.bs = b->d_bs; R.bq = b->d_bq; R.mq = b->d_mq; R.ref_nt16 = b->d_ref; R.colpos = b->d_colpos;
        R.max_columns = b->max_columns; R.max_entries = (uint32_t)std::min<uint64_t>(b->max_entries, 0xfffffff0ull);
        R.run_len = (uint32_t)kRunLen; R.max_runs = (uint32_t)kMaxRuns;
        const uint32_t npos = n_windows * SSN_READ_WINDOW;
        CU(cudaEventRecord(b->pa, st));
        pileup::count_kernel<<<n_windows * 2, pileup::kScanThreads, 0, st>>>(R);
        pileup::columns_kernel<<<n_windows, pileup::kScanThreads, 0, st>>>(R);
        pileup::bases_kernel<<<1, 32, 0, st>>>(R);
        pileup::offsets_kernel<<<npos / pileup::kThreads, pileup::kThreads, 0, st>>>(R);
        pileup::fill_kernel<<<npos / pileup::kThreads, pileup::kThreads, pileup::kFillSmem, st>>>(R);
        pileup::shape_kernel<<<(npos / 32 + pileup::kThreads - 1) / pileup::kThreads, pileup::kThreads, 0, st>>>(R);
        CU(cudaGetLastError());
        CU(cudaEventRecord(b->pb, st));
        ctx->launches += 6;
    }
    CU(cudaMemcpyAsync(b->h_summary, b->d_summary, pileup::kSummaryWords * 4, cudaMemcpyDeviceToHost, st));
    CU(cudaEventRecord(b->piled, st));
    b->reads_pending = true;
    b->state = SLOT_SUBMITTED; b->seq = ctx->next_seq++;
    /* Product path (pre-filter + survivors, no dump): queue the scoring now, from the column count the device holds.
     * The launch assumes no column outgrows the shared-memory key store; the kernels check that (and the pileup's
     * error bits) against the summary and do nothing otherwise — finish_reads then launches with the real shape. */
    /* Measured on B200: with few slots (the CLI's 3) the early launch hides the scoring behind the next batch's copy
     * (streaming +10 %); with 8 slots saturating PCIe it costs 6 % of the copy rate, so scoring then waits for
     * ssn_results_next.  SSN_READS_LAZY_SCORE=0|1 forces either (measurement knob). */
    static const int lazy_env = getenv("SSN_READS_LAZY_SCORE") ? atoi(getenv("SSN_READS_LAZY_SCORE")) : -1;
    const bool lazy = lazy_env >= 0 ? lazy_env != 0 : ctx->slots.size() > 4;
    if (n_windows && !lazy && ctx->pf_on && !(flags & (SSN_SUBMIT_NO_PREFILTER | SSN_SUBMIT_COLUMN_DUMP))) {
        b->spec = true; b->spec_col_words = ctx->smem_words_default / 32u;
        b->ncol = (uint32_t)std::min<uint64_t>((uint64_t)n_windows * SSN_READ_WINDOW, b->max_columns);
        b->max_col_words = b->spec_col_words; b->max_tile_words = 0;
        ssn_status rc = launch(ctx, b, flags);
        if (rc != SSN_OK) return rc;
        CU(cudaMemcpyAsync(b->h_colpos, b->d_colpos, (size_t)b->ncol * 4, cudaMemcpyDeviceToHost, st));
        if ((rc = copy_out_async(b)) != SSN_OK) return rc;
    }
    return SSN_OK;
}

/* the device knows the batch's shape now: queue the scoring kernels and the copies back */
static ssn_status finish_reads(ssn_ctx *ctx, ssn_batch *b)
{
    CU(cudaEventSynchronize(b->piled));
    b->reads_pending = false;
    const uint32_t *S = b->h_summary;
    if (S[2]) {
        if (S[2] & pileup::kErrRecord) { ssn_set_error("read batch: a record or its index entry points outside the record buffer"); return SSN_ERR_ARG; }
        if (S[2] & pileup::kErrUnsorted) { ssn_set_error("read batch: the reads of a window are not sorted by start"); return SSN_ERR_ARG; }
        if (S[2] & (pileup::kErrColumns | pileup::kErrEntries)) {
            ssn_set_error("read batch: %u columns / %u read-bases exceed the slot capacity %u / %llu", S[0], S[1], b->max_columns, (unsigned long long)b->max_entries);
            return SSN_ERR_OVERFLOW;
        }
        ssn_set_error("read batch: a column holds more than %d read-bases", kRunLen * kMaxRuns);
        return SSN_ERR_ARG;
    }
    if (b->n_win) {
        float ms = 0;
        CU(cudaEventElapsedTime(&ms, b->pa, b->pb));
        ctx->pileup_ms += ms; ctx->pileup_n += 1;
    }
    b->ncol = S[0]; b->nent = S[1]; b->max_tile_words = S[3]; b->max_col_words = S[4];
    b->resident = true;
    if (b->spec && S[4] <= b->spec_col_words) return SSN_OK;      /* already scored; the copies back are queued */
    b->spec = false;
    ssn_status st = launch(ctx, b, b->flags);
    if (st != SSN_OK) return st;
    if (b->ncol) CU(cudaMemcpyAsync(b->h_colpos, b->d_colpos, (size_t)b->ncol * 4, cudaMemcpyDeviceToHost, b->stream));
    return copy_out_async(b);
}

extern "C" ssn_status ssn_batch_column_positions(ssn_batch *b, const uint32_t **pos, uint32_t *n_columns, uint64_t *n_entries)
{
    if (!b || !pos || !n_columns) { ssn_set_error("ssn_batch_column_positions: null"); return SSN_ERR_ARG; }
    if (!b->from_reads || b->reads_pending || b->state == SLOT_SUBMITTED) { ssn_set_error("ssn_batch_column_positions: not a finished read batch"); return SSN_ERR_STATE; }
    *pos = b->h_colpos; *n_columns = b->ncol;
    if (n_entries) *n_entries = b->nent;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_download_columns(ssn_ctx *ctx, ssn_batch *b)
{
    if (!ctx || !b || b->ctx != ctx) { ssn_set_error("ssn_batch_download_columns: bad batch"); return SSN_ERR_ARG; }
    if (!b->from_reads || b->reads_pending || !b->resident) { ssn_set_error("ssn_batch_download_columns: not a finished read batch"); return SSN_ERR_STATE; }
    CU(cudaSetDevice(ctx->device));
    ssn_status hs = host_columns(b);
    if (hs != SSN_OK) return hs;
    CU(cudaMemcpyAsync(b->h_off, b->d_off, sizeof(uint32_t) * (2 * (size_t)b->ncol + 1), cudaMemcpyDeviceToHost, b->stream));
    if (b->nent) {
        CU(cudaMemcpyAsync(b->h_bs, b->d_bs, b->nent, cudaMemcpyDeviceToHost, b->stream));
        CU(cudaMemcpyAsync(b->h_bq, b->d_bq, b->nent, cudaMemcpyDeviceToHost, b->stream));
        CU(cudaMemcpyAsync(b->h_mq, b->d_mq, b->nent, cudaMemcpyDeviceToHost, b->stream));
    }
    if (b->ncol) CU(cudaMemcpyAsync(b->h_ref, b->d_ref, b->ncol, cudaMemcpyDeviceToHost, b->stream));
    CU(cudaStreamSynchronize(b->stream));
    return SSN_OK;
}

extern "C" ssn_status ssn_pileup_time(ssn_ctx *ctx, float *total_ms, uint64_t *n_batches, int reset)
{
    if (!ctx) { ssn_set_error("ssn_pileup_time: null"); return SSN_ERR_ARG; }
    if (total_ms) *total_ms = (float)ctx->pileup_ms;
    if (n_batches) *n_batches = ctx->pileup_n;
    if (reset) { ctx->pileup_ms = 0; ctx->pileup_n = 0; }
    return SSN_OK;
}

extern "C" ssn_status ssn_results_next(ssn_ctx *ctx, const ssn_site **sites, uint32_t *n_sites, ssn_batch **batch)
{
    if (!ctx || !sites || !n_sites) { ssn_set_error("ssn_results_next: null"); return SSN_ERR_ARG; }
    ssn_batch *pick = nullptr;
    for (ssn_batch *b : ctx->slots)
        if (b->state == SLOT_SUBMITTED && (!pick || b->seq < pick->seq)) pick = b;
    if (!pick) { ssn_set_error("ssn_results_next: nothing submitted"); return SSN_ERR_STATE; }
    CU(cudaSetDevice(ctx->device));
    ssn_status st = SSN_OK;
    if (pick->reads_pending && (st = finish_reads(ctx, pick)) != SSN_OK) {
        pick->state = SLOT_DONE; pick->n_sites = 0;        /* the batch is over: the caller may release it and go on */
        if (batch) *batch = pick;
        return st;
    }
    if ((st = collect(ctx, pick)) != SSN_OK) return st;
    pick->state = SLOT_DONE;
    *sites = pick->h_sites; *n_sites = pick->n_sites;
    if (batch) *batch = pick;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_column_dump(ssn_batch *b, const uint32_t **words)
{
    if (!b || !words) { ssn_set_error("ssn_batch_column_dump: null"); return SSN_ERR_ARG; }
    if (!(b->flags & SSN_SUBMIT_COLUMN_DUMP) || !b->h_dump) { ssn_set_error("batch was not scored with SSN_SUBMIT_COLUMN_DUMP"); return SSN_ERR_STATE; }
    *words = b->h_dump;
    return SSN_OK;
}

/* ---- device-resident path ------------------------------------------------------------------------- */

extern "C" ssn_status ssn_batch_upload(ssn_ctx *ctx, ssn_batch *b)
{
    if (!ctx || !b || b->ctx != ctx) { ssn_set_error("ssn_batch_upload: bad batch"); return SSN_ERR_ARG; }
    if (b->state != SLOT_FILLING) { ssn_set_error("ssn_batch_upload: batch was not acquired"); return SSN_ERR_STATE; }
    if (!b->h_off) { ssn_set_error("ssn_batch_upload: the batch was never filled (ssn_batch_view_get)"); return SSN_ERR_STATE; }
    CU(cudaSetDevice(ctx->device));
    ssn_status st = scan_offsets(b);
    if (st != SSN_OK) return st;
    if ((st = copy_in(ctx, b)) != SSN_OK) return st;
    CU(cudaStreamSynchronize(b->stream));
    b->resident = true;
    return SSN_OK;
}

extern "C" ssn_status ssn_batch_score_resident(ssn_ctx *ctx, ssn_batch *b, uint32_t flags)
{
    if (!ctx || !b || b->ctx != ctx) { ssn_set_error("ssn_batch_score_resident: bad batch"); return SSN_ERR_ARG; }
    if (!b->resident) { ssn_set_error("ssn_batch_score_resident: call ssn_batch_upload first"); return SSN_ERR_STATE; }
    CU(cudaSetDevice(ctx->device));
    return launch(ctx, b, flags);
}

extern "C" ssn_status ssn_batch_fetch(ssn_ctx *ctx, ssn_batch *b, const ssn_site **sites, uint32_t *n_sites)
{
    if (!ctx || !b || b->ctx != ctx || !sites || !n_sites) { ssn_set_error("ssn_batch_fetch: bad argument"); return SSN_ERR_ARG; }
    if (!b->resident) { ssn_set_error("ssn_batch_fetch: batch is not resident"); return SSN_ERR_STATE; }
    CU(cudaSetDevice(ctx->device));
    ssn_status st = copy_out_async(b);
    if (st != SSN_OK) return st;
    if ((st = collect(ctx, b)) != SSN_OK) return st;
    *sites = b->h_sites; *n_sites = b->n_sites;
    return SSN_OK;
}

/* ---- timing / accounting ----------------------------------------------------------------------------- */

static ssn_status join_streams(ssn_ctx *ctx, cudaEvent_t ev)
{
    /* main_stream waits for everything queued so far on every slot stream, then records ev */
    for (ssn_batch *b : ctx->slots) {
        cudaEvent_t tmp;
        CU(cudaEventCreateWithFlags(&tmp, cudaEventDisableTiming));
        CU(cudaEventRecord(tmp, b->stream));
        CU(cudaStreamWaitEvent(ctx->main_stream, tmp, 0));
        CU(cudaEventDestroy(tmp));
    }
    CU(cudaEventRecord(ev, ctx->main_stream));
    return SSN_OK;
}

extern "C" ssn_status ssn_timer_start(ssn_ctx *ctx)
{
    if (!ctx) { ssn_set_error("null ctx"); return SSN_ERR_ARG; }
    CU(cudaSetDevice(ctx->device));
    ssn_status st = join_streams(ctx, ctx->t0);
    if (st != SSN_OK) return st;
    /* every slot stream starts after t0 */
    for (ssn_batch *b : ctx->slots) CU(cudaStreamWaitEvent(b->stream, ctx->t0, 0));
    return SSN_OK;
}

extern "C" ssn_status ssn_timer_stop(ssn_ctx *ctx, float *elapsed_ms)
{
    if (!ctx || !elapsed_ms) { ssn_set_error("null argument"); return SSN_ERR_ARG; }
    CU(cudaSetDevice(ctx->device));
    ssn_status st = join_streams(ctx, ctx->t1);
    if (st != SSN_OK) return st;
    CU(cudaEventSynchronize(ctx->t1));
    CU(cudaEventElapsedTime(elapsed_ms, ctx->t0, ctx->t1));
    return SSN_OK;
}

extern "C" ssn_status ssn_sync(ssn_ctx *ctx)
{
    if (!ctx) { ssn_set_error("null ctx"); return SSN_ERR_ARG; }
    CU(cudaSetDevice(ctx->device));
    CU(cudaDeviceSynchronize());
    return SSN_OK;
}

extern "C" uint64_t ssn_kernel_launches(const ssn_ctx *ctx) { return ctx ? ctx->launches : 0; }

extern "C" ssn_status ssn_kernel_time(ssn_ctx *ctx, float *total_ms, uint64_t *n_launches, int reset)
{
    if (!ctx) { ssn_set_error("null ctx"); return SSN_ERR_ARG; }
    CU(cudaSetDevice(ctx->device));
    for (auto &kt : ctx->timers) {
        ssn_status st = timer_collect(ctx, kt);
        if (st != SSN_OK) return st;
    }
    if (total_ms) *total_ms = (float)ctx->kernel_ms;
    if (n_launches) *n_launches = ctx->kernel_n;
    if (reset) { ctx->kernel_ms = 0; ctx->kernel_n = 0; ctx->prefilter_ms = 0; ctx->score_ms = 0; }
    return SSN_OK;
}

extern "C" ssn_status ssn_kernel_time_split(ssn_ctx *ctx, float *prefilter_ms, float *score_ms, uint64_t *n_batches, int reset)
{
    if (!ctx) { ssn_set_error("null ctx"); return SSN_ERR_ARG; }
    CU(cudaSetDevice(ctx->device));
    for (auto &kt : ctx->timers) {
        ssn_status st = timer_collect(ctx, kt);
        if (st != SSN_OK) return st;
    }
    if (prefilter_ms) *prefilter_ms = (float)ctx->prefilter_ms;
    if (score_ms) *score_ms = (float)ctx->score_ms;
    if (n_batches) *n_batches = ctx->kernel_n;
    if (reset) { ctx->kernel_ms = 0; ctx->kernel_n = 0; ctx->prefilter_ms = 0; ctx->score_ms = 0; }
    return SSN_OK;
}

extern "C" int ssn_prefilter_enabled(const ssn_ctx *ctx) { return ctx && ctx->pf_on ? 1 : 0; }

extern "C" ssn_status ssn_batch_survivors(const ssn_batch *b, uint32_t *n_columns)
{
    if (!b || !n_columns) { ssn_set_error("ssn_batch_survivors: null"); return SSN_ERR_ARG; }
    *n_columns = b->h_count[2];
    return SSN_OK;
}

extern "C" ssn_status ssn_host_prefilter_create(const ssn_params *p, ssn_prefilter_view *view, void **handle)
{
    if (!p || !view || !handle) { ssn_set_error("ssn_host_prefilter_create: null"); return SSN_ERR_ARG; }
    ssn_host_tables m;
    if (ssn_tables_build(p, &m) != 0) { ssn_set_error("table construction failed (n_hap=%d)", p->n_hap); return SSN_ERR_ARG; }
    ssn_prefilter_tables *t = new ssn_prefilter_tables();
    const int rc = ssn_prefilter_build(&m, (int)(m.q_r + 0.5), t);
    ssn_tables_free(&m);
    if (rc != 0) { delete t; ssn_set_error("out of memory"); return SSN_ERR_NOMEM; }
    view->ab = t->ab; view->h = t->h; view->F = t->F; view->vt = t->vt; view->n = SSN_PF_N; view->enabled = t->enabled;
    *handle = t;
    return SSN_OK;
}

extern "C" void ssn_host_prefilter_destroy(void *handle)
{
    ssn_prefilter_tables *t = (ssn_prefilter_tables*)handle;
    if (!t) return;
    ssn_prefilter_free(t);
    delete t;
}

__global__ void ssn_flush_kernel(uint4 *p, size_t n)
{
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n; i += (size_t)gridDim.x * blockDim.x)
        p[i] = make_uint4((unsigned)i, 1u, 2u, 3u);
}

extern "C" ssn_status ssn_flush_l2(ssn_ctx *ctx)
{
    if (!ctx) { ssn_set_error("null ctx"); return SSN_ERR_ARG; }
    CU(cudaSetDevice(ctx->device));
    if (!ctx->d_flush) {
        ctx->flush_bytes = (size_t)256 << 20;
        CU(cudaMalloc(&ctx->d_flush, ctx->flush_bytes));
    }
    ssn_flush_kernel<<<ctx->num_sms * 4, 256, 0, ctx->main_stream>>>((uint4*)ctx->d_flush, ctx->flush_bytes / 16);
    CU(cudaGetLastError());
    CU(cudaStreamSynchronize(ctx->main_stream));
    return SSN_OK;
}
