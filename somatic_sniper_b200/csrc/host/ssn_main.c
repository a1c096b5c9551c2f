/*
 * bam-somaticsniper (B200 build) — same command line, messages, exit codes and output bytes as the reference
 * driver (src/exe/bam-somaticsniper/main.c:64-162); the per-column work runs on the GPU(s) through libssn.so.
 *
 *   reference                                           here
 *   getopt "n:t:vf:T:N:r:I:q:Q:pLGJs:F:" (:80-99)        identical option handling
 *   fai_load / table + prior construction (:106-127)    ssn_packer_open (FASTA index) / ssn_create (tables, upload)
 *   bam_sspileup_file(..., glf_somatic, ...) (:146)     ssn_packer_next -> ssn_batch_submit -> ssn_results_next
 *   output_formatter (:141-145, somatic_sniper.c:264)   ssn_writer_*
 *
 * Not in the reference (environment only, the command line is unchanged):
 *   SSN_GPUS     number of GPUs to spread batches over (default 1; batch k goes to GPU k mod N)
 *   SSN_DEVICES  comma-separated device ordinals, one scoring context each (overrides SSN_GPUS; an ordinal may repeat)
 *   SSN_THREADS  worker threads for inflate + packing (default: online cores - 2, at least 1)
 *   SSN_SLOTS    batches in flight per GPU (default 3)
 *   SSN_STATS    file that receives one JSON line of run statistics
 * There is no CPU scoring path: without a usable sm_100 device the program fails.
 */
#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <unistd.h>

#include "ssn.h"
#include "ssn_host.h"

#define SSN_CLI_VERSION "b200-r1"
#define MAX_GPUS 16
#define MAX_SLOTS 8
#define BATCH_COLUMNS (16u << 13)            /* 16 packer windows */
#define BATCH_ENTRIES ((uint64_t)24 << 20)

static const char *default_normal_id = "NORMAL";
static const char *default_tumor_id = "TUMOR";
static const char *default_format = "classic";

static void version_info(void)
{
    printf("Somatic Sniper version (%s) (commit %s)", SSN_CLI_VERSION, "nogit");
    printf(" (%s)", "release");
    printf("\n");
}

static void usage(const char *progname, const ssn_params *p, int mapq)
{
    const char *pn = strrchr(progname, '/');
    int i;
    pn = pn ? pn + 1 : progname;
    fprintf(stderr, "\n");
    fprintf(stderr, "\n");
    fprintf(stderr, "%s [options] -f <ref.fasta> <tumor.bam> <normal.bam> <snp_output_file>\n\n", pn);
    fprintf(stderr, "Required Option: \n");
    fprintf(stderr, "        -f FILE   REQUIRED reference sequence in the FASTA format\n\n");
    fprintf(stderr, "Options: \n");
    fprintf(stderr, "        -v        Display version information\n\n");
    fprintf(stderr, "        -q INT    filtering reads with mapping quality less than INT [%d]\n", mapq);
    fprintf(stderr, "        -Q INT    filtering somatic snv output with somatic quality less than  INT [%d]\n", p->min_somatic_qual);
    fprintf(stderr, "        -L FLAG   do not report LOH variants as determined by genotypes\n");
    fprintf(stderr, "        -G FLAG   do not report Gain of Reference variants as determined by genotypes\n");
    fprintf(stderr, "        -p FLAG   disable priors in the somatic calculation. Increases sensitivity for solid tumors\n");
    fprintf(stderr, "        -J FLAG   Use prior probabilities accounting for the somatic mutation rate\n");
    fprintf(stderr, "        -s FLOAT  prior probability of a somatic mutation (implies -J) [%f]\n", p->somatic_mutation_rate);
    fprintf(stderr, "        -T FLOAT  theta in maq consensus calling model (for -c/-g) [%f]\n", p->theta);
    fprintf(stderr, "        -N INT    number of haplotypes in the sample (for -c/-g) [%d]\n", p->n_hap);
    fprintf(stderr, "        -r FLOAT  prior of a difference between two haplotypes (for -c/-g) [%f]\n", p->het_rate);
    fprintf(stderr, "        -n STRING normal sample id (for VCF header) [%s]\n", default_normal_id);
    fprintf(stderr, "        -t STRING tumor sample id (for VCF header) [%s]\n", default_tumor_id);
    fprintf(stderr, "        -F STRING select output format [%s]\n", default_format);
    fprintf(stderr, "           Available formats:\n");
    for (i = 0; i < ssn_writer_n_formats(); ++i) fprintf(stderr, "             %s\n", ssn_writer_format_name(i));
    fprintf(stderr, "\n");
}

typedef struct {
    ssn_batch *batch;
    ssn_colmap map;
    /* SSN_DEVICE_PILEUP=1: the slot's read arrays and the raw reference characters of its windows */
    ssn_reads_view rv;
    uint8_t *ref_chars;
} slot_meta;

typedef struct {
    ssn_ctx *ctx;
    slot_meta meta[MAX_SLOTS];
    int n_meta, in_flight;
} gpu_t;

typedef struct { int gpu; } flight;

static double now_s(void)
{
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double)ts.tv_sec + 1e-9 * (double)ts.tv_nsec;
}

static int env_int(const char *name, int dflt, int lo, int hi)
{
    const char *s = getenv(name);
    int v = s && *s ? atoi(s) : dflt;
    if (v < lo) v = lo;
    if (v > hi) v = hi;
    return v;
}

static slot_meta *meta_of(gpu_t *g, ssn_batch *b)
{
    int i;
    for (i = 0; i < g->n_meta; ++i)
        if (g->meta[i].batch == b) return &g->meta[i];
    if (g->n_meta == MAX_SLOTS) return 0;
    g->meta[g->n_meta].batch = b;
    g->meta[g->n_meta].map.tid = (int32_t*)malloc(sizeof(int32_t) * BATCH_COLUMNS);
    g->meta[g->n_meta].map.pos = (uint32_t*)malloc(sizeof(uint32_t) * BATCH_COLUMNS);
    g->meta[g->n_meta].map.ref_char = (uint8_t*)malloc(BATCH_COLUMNS);
    g->meta[g->n_meta].ref_chars = (uint8_t*)malloc(BATCH_COLUMNS);
    return &g->meta[g->n_meta++];
}

static uint64_t g_survivors;   /* columns the pre-filter handed to the full model (statistics only) */
static int g_device_pileup;    /* SSN_DEVICE_PILEUP=1: ship the reads, build the columns on the GPU (ssn_batch_submit_reads) */
static uint64_t g_dev_columns, g_dev_entries;

/* waits for the oldest batch in flight and writes its sites */
static int drain_one(gpu_t *gpus, flight *fifo, int *head, int *count, int cap, ssn_packer *pk, ssn_writer *wr, uint64_t *n_sites)
{
    gpu_t *g = &gpus[fifo[*head].gpu];
    const ssn_site *sites;
    uint32_t n, i;
    ssn_batch *b;
    slot_meta *m;
    if (ssn_results_next(g->ctx, &sites, &n, &b) != SSN_OK) {
        fprintf(stderr, "[bam-somaticsniper] GPU scoring failed: %s\n", ssn_last_error());
        return -1;
    }
    m = meta_of(g, b);
    if (g_device_pileup) {
        const uint32_t *cpos;
        uint32_t ncol;
        uint64_t nent;
        if (ssn_batch_column_positions(b, &cpos, &ncol, &nent) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error());
            return -1;
        }
        g_dev_columns += ncol; g_dev_entries += nent;
        for (i = 0; i < n; ++i) {
            const uint32_t pw = cpos[sites[i].column], w = pw >> SSN_READ_WINDOW_SHIFT, o = pw & (SSN_READ_WINDOW - 1u);
            const char *name = ssn_packer_target_name(pk, 0, m->rv.win[w].tid);
            ssn_writer_site(wr, name ? name : "", m->rv.win[w].lo + o, m->ref_chars[pw], &sites[i]);
        }
    } else
    for (i = 0; i < n; ++i) {
        const uint32_t c = sites[i].column;
        const char *name = ssn_packer_target_name(pk, 0, m->map.tid[c]);        /* names from the tumor header (QUIRK B9) */
        ssn_writer_site(wr, name ? name : "", m->map.pos[c], m->map.ref_char[c], &sites[i]);
    }
    *n_sites += n;
    {
        uint32_t surv = 0;
        if (ssn_batch_survivors(b, &surv) == SSN_OK) g_survivors += surv;
    }
    g->in_flight -= 1;
    *head = (*head + 1) % cap;
    *count -= 1;
    return 0;
}

int main(int argc, char *argv[])
{
    int c, use_priors = 1, mapq = 0, n_gpus, n_slots, n_threads, i, rc = 0, devices[MAX_GPUS];
    const char *normal_id = default_normal_id, *tumor_id = default_tumor_id, *fn_fa = 0, *format = default_format;
    ssn_params prm;
    ssn_packer *pk = 0;
    ssn_writer *wr;
    FILE *out;
    gpu_t gpus[MAX_GPUS];
    flight fifo[MAX_GPUS * MAX_SLOTS];
    int f_head = 0, f_count = 0, next_gpu = 0;
    uint64_t n_sites = 0, n_batches = 0;
    double t_start = now_s(), t_setup, t_gpu_wait = 0;

    ssn_params_default(&prm);
    while ((c = getopt(argc, argv, "n:t:vf:T:N:r:I:q:Q:pLGJs:F:")) >= 0) {
        switch (c) {
        case 'f': fn_fa = optarg; break;
        case 'T': prm.theta = (float)atof(optarg); break;
        case 'N': prm.n_hap = atoi(optarg); break;
        case 'r': prm.het_rate = (float)atof(optarg); break;
        case 'q': mapq = atoi(optarg); break;
        case 'Q': prm.min_somatic_qual = atoi(optarg); break;
        case 'F': format = optarg; break;
        case 'p': use_priors = 0; break;
        case 'J': prm.use_joint_priors = 1; break;
        case 's': prm.somatic_mutation_rate = atof(optarg); prm.use_joint_priors = 1; break;
        case 'v': version_info(); exit(0); break;
        case 't': tumor_id = optarg; break;
        case 'n': normal_id = optarg; break;
        case 'L': prm.include_loh = 0; break;
        case 'G': prm.include_gor = 0; break;
        default: fprintf(stderr, "Unrecognizd option '-%c'.\n", c); return 1;
        }
    }
    prm.use_priors = use_priors;
    if (optind == argc) { usage(argv[0], &prm, mapq); return 1; }
    if (!fn_fa) {
        fprintf(stderr, "You MUST specify a reference sequence. It isn't optional.\n");
        exit(1);
    }
    if (argc - optind < 3) {
        /* the reference reads argv past its end here; ask for the missing arguments instead */
        usage(argv[0], &prm, mapq);
        return 1;
    }
    n_gpus = env_int("SSN_GPUS", 1, 1, MAX_GPUS);
    for (i = 0; i < MAX_GPUS; ++i) devices[i] = i;
    {
        const char *dv = getenv("SSN_DEVICES");
        if (dv && *dv) {
            n_gpus = 0;
            while (*dv && n_gpus < MAX_GPUS) {
                devices[n_gpus++] = atoi(dv);
                while (*dv && *dv != ',') ++dv;
                if (*dv == ',') ++dv;
            }
            if (n_gpus == 0) n_gpus = 1;
        }
    }
    n_slots = env_int("SSN_SLOTS", 3, 1, MAX_SLOTS);
    {
        long cores = sysconf(_SC_NPROCESSORS_ONLN);
        n_threads = env_int("SSN_THREADS", cores > 3 ? (int)cores - 2 : 1, 0, 256);
    }
    memset(gpus, 0, sizeof(gpus));
    g_device_pileup = env_int("SSN_DEVICE_PILEUP", 0, 0, 1);

    /* same order of stderr lines as main.c:115-131 */
    if (ssn_packer_open(argv[optind], argv[optind + 1], fn_fa, mapq, n_threads, &pk) != SSN_OK) {
        fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_host_last_error());
        return 1;
    }
    if (prm.use_joint_priors)
        fprintf(stderr, "Using priors accounting for somatic mutation rate. Prior probability of a somatic mutation is %f\n", prm.somatic_mutation_rate);
    for (i = 0; i < n_gpus; ++i) {
        if (ssn_create(&prm, devices[i], (uint32_t)n_slots, BATCH_COLUMNS, BATCH_ENTRIES, &gpus[i].ctx) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] cannot set up GPU %d: %s\n", devices[i], ssn_last_error());
            fprintf(stderr, "[bam-somaticsniper] this build scores on sm_100 GPUs only; there is no CPU path\n");
            ssn_packer_close(pk);
            return 1;
        }
        if (g_device_pileup && ssn_reads_reserve(gpus[i].ctx, 2 * BATCH_ENTRIES, (uint32_t)(BATCH_ENTRIES / 24), BATCH_COLUMNS >> SSN_READ_WINDOW_SHIFT) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error());
            ssn_packer_close(pk);
            return 1;
        }
        if (g_device_pileup) {                       /* the slots' read arrays are allocated on first use: do it here */
            ssn_batch *held[MAX_SLOTS];
            ssn_reads_view rv;
            int k, ok = 1;
            for (k = 0; k < n_slots; ++k) ok = ok && ssn_batch_acquire(gpus[i].ctx, &held[k]) == SSN_OK && ssn_batch_reads_view(gpus[i].ctx, held[k], &rv) == SSN_OK;
            if (!ok) { fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error()); ssn_packer_close(pk); return 1; }
            for (k = 0; k < n_slots; ++k) ssn_batch_release(gpus[i].ctx, held[k]);
        }
    }
    fprintf(stderr, "Preparing to snipe some somatics\n");
    if (use_priors) fprintf(stderr, "Using prior probabilities\n");
    fprintf(stderr, "Normal bam is %s\n", argv[optind + 1]);
    fprintf(stderr, "Tumor bam is %s\n", argv[optind]);
    out = fopen(argv[optind + 2], "w");
    wr = ssn_writer_create(format, out);
    if (!wr) {
        fprintf(stderr, "unknown output format: '%s'. Abort!\n", format);     /* output_format.c:31-32 */
        exit(1);
    }
    if (!out) {
        fprintf(stderr, "Unable to open snp file!!!!!!!!!\n");
        exit(1);
    }
    ssn_writer_header(wr, fn_fa, normal_id, tumor_id);
    t_setup = now_s() - t_start;

    for (;;) {
        gpu_t *g = &gpus[next_gpu];
        ssn_batch *b;
        ssn_batch_view view;
        slot_meta *m;
        uint32_t ncol = 0;
        uint64_t nent = 0;
        double t0;
        t0 = now_s();
        while (g->in_flight == n_slots)
            if (drain_one(gpus, fifo, &f_head, &f_count, MAX_GPUS * MAX_SLOTS, pk, wr, &n_sites) != 0) { rc = 1; goto done; }
        t_gpu_wait += now_s() - t0;
        if (ssn_batch_acquire(g->ctx, &b) != SSN_OK || ssn_batch_view_get(b, &view) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error());
            rc = 1; goto done;
        }
        m = meta_of(g, b);
        if (!m) { fprintf(stderr, "[bam-somaticsniper] slot table overflow\n"); rc = 1; goto done; }
        if (g_device_pileup) {
            uint32_t nw = 0, nr = 0;
            uint64_t nb = 0;
            if (ssn_batch_reads_view(g->ctx, b, &m->rv) != SSN_OK) { fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_last_error()); rc = 1; goto done; }
            if (ssn_packer_next_reads(pk, &m->rv, BATCH_COLUMNS, BATCH_ENTRIES, m->ref_chars, &nw, &nr, &nb) != SSN_OK) {
                fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_host_last_error());
                rc = 1; goto done;
            }
            if (nw == 0) { ssn_batch_release(g->ctx, b); break; }
            if (ssn_batch_submit_reads(g->ctx, b, nw, nr, nb, 0) != SSN_OK) {
                fprintf(stderr, "[bam-somaticsniper] GPU submit failed: %s\n", ssn_last_error());
                rc = 1; goto done;
            }
            fifo[(f_head + f_count) % (MAX_GPUS * MAX_SLOTS)].gpu = next_gpu;
            f_count += 1; g->in_flight += 1; n_batches += 1;
            next_gpu = (next_gpu + 1) % n_gpus;
            continue;
        }
        if (ssn_packer_next(pk, &view, &m->map, &ncol, &nent) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] %s\n", ssn_host_last_error());
            rc = 1; goto done;
        }
        if (ncol == 0) { ssn_batch_release(g->ctx, b); break; }
        if (ssn_batch_set_size(b, ncol, nent) != SSN_OK || ssn_batch_submit(g->ctx, b, 0) != SSN_OK) {
            fprintf(stderr, "[bam-somaticsniper] GPU submit failed: %s\n", ssn_last_error());
            rc = 1; goto done;
        }
        fifo[(f_head + f_count) % (MAX_GPUS * MAX_SLOTS)].gpu = next_gpu;
        f_count += 1; g->in_flight += 1; n_batches += 1;
        next_gpu = (next_gpu + 1) % n_gpus;
    }
    {
        double t0 = now_s();
        while (f_count > 0)
            if (drain_one(gpus, fifo, &f_head, &f_count, MAX_GPUS * MAX_SLOTS, pk, wr, &n_sites) != 0) { rc = 1; goto done; }
        t_gpu_wait += now_s() - t0;
    }
done:
    fflush(out);
    {
        const char *sp = getenv("SSN_STATS");
        if (sp && *sp) {
            FILE *sf = fopen(sp, "w");
            if (sf) {
                ssn_packer_stats st;
                uint64_t launches = 0;
                const double wall = now_s() - t_start;
                ssn_packer_get_stats(pk, &st);
                if (g_device_pileup) { st.columns = g_dev_columns; st.entries = g_dev_entries; }
                for (i = 0; i < n_gpus; ++i) launches += ssn_kernel_launches(gpus[i].ctx);
                fprintf(sf, "{\"columns\": %llu, \"read_bases\": %llu, \"batches\": %llu, \"sites\": %llu, \"records_tumor\": %llu, "
                        "\"records_normal\": %llu, \"wall_s\": %.6f, \"setup_s\": %.6f, \"pack_wait_s\": %.6f, \"pack_count_s\": %.6f, "
                        "\"pack_fill_s\": %.6f, \"gpu_wait_s\": %.6f, \"gpus\": %d, \"threads\": %d, \"slots\": %d, \"kernel_launches\": %llu, "
                        "\"prefilter\": %d, \"prefilter_survivors\": %llu, \"device_pileup\": %d, \"columns_per_s\": %.1f}\n",
                        (unsigned long long)st.columns, (unsigned long long)st.entries, (unsigned long long)n_batches,
                        (unsigned long long)n_sites, (unsigned long long)st.records[0], (unsigned long long)st.records[1], wall, t_setup,
                        st.t_wait, st.t_count, st.t_fill, t_gpu_wait, n_gpus, n_threads, n_slots, (unsigned long long)launches,
                        ssn_prefilter_enabled(gpus[0].ctx), (unsigned long long)g_survivors, g_device_pileup,
                        wall > t_setup ? (double)st.columns / (wall - t_setup) : 0.0);
                fclose(sf);
            }
        }
    }
    ssn_packer_close(pk);
    for (i = 0; i < n_gpus; ++i) {
        int k;
        ssn_destroy(gpus[i].ctx);
        for (k = 0; k < gpus[i].n_meta; ++k) { free(gpus[i].meta[k].map.tid); free(gpus[i].meta[k].map.pos); free(gpus[i].meta[k].map.ref_char); free(gpus[i].meta[k].ref_chars); }
    }
    ssn_writer_destroy(wr);
    fclose(out);
    return rc;
}
