/* Host-side stress of the CLI's feed loop without a GPU: region shards -> ssn_packer_next_reads into plain buffers, one or
 * several consumer threads, repeated.  Built with -fsanitize=address / thread by tools/sanitize_host.sh.
 *   shard_stress <tumor.bam> <normal.bam> <ref.fa> <consumers> <threads> <repeats> */
#include <pthread.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include "ssn_host.h"

#define BATCH_COLUMNS (16u << 13)
#define BATCH_ENTRIES ((uint64_t)32 << 20)

static ssn_shardset *set;
static unsigned long long g_windows, g_reads;
static pthread_mutex_t mu = PTHREAD_MUTEX_INITIALIZER;

static double now_s(void) { struct timespec t; clock_gettime(CLOCK_MONOTONIC, &t); return (double)t.tv_sec + 1e-9 * (double)t.tv_nsec; }

static void *consumer(void *arg)
{
    ssn_reads_view rv;
    uint8_t *ref_chars = (uint8_t*)malloc(BATCH_COLUMNS);
    unsigned long long nw_tot = 0, nr_tot = 0;
    (void)arg;
    rv.max_rec_bytes = 2 * BATCH_ENTRIES; rv.max_reads = (uint32_t)(BATCH_ENTRIES / 24); rv.max_windows = 16;
    rv.recs = (uint8_t*)malloc(rv.max_rec_bytes);
    rv.rec_off = (uint32_t*)malloc(4ull * rv.max_reads);
    rv.win = (ssn_read_window*)malloc(sizeof(ssn_read_window) * rv.max_windows);
    rv.ref_nt16 = (uint8_t*)malloc((size_t)rv.max_windows * SSN_READ_WINDOW);
    for (;;) {
        int32_t k;
        ssn_packer *pk;
        if (ssn_shardset_next(set, &k, &pk) != SSN_OK) { fprintf(stderr, "next: %s\n", ssn_host_last_error()); exit(2); }
        if (k < 0) break;
        for (;;) {
            uint32_t nw = 0, nr = 0;
            uint64_t nb = 0;
            ssn_status st = ssn_packer_next_reads(pk, &rv, BATCH_COLUMNS, BATCH_ENTRIES, ref_chars, &nw, &nr, &nb);
            if (st != SSN_OK) { fprintf(stderr, "next_reads: %s\n", ssn_host_last_error()); exit(2); }
            if (nw == 0) break;
            nw_tot += nw; nr_tot += nr;
        }
        if (ssn_shardset_done(set, k) != SSN_OK) { fprintf(stderr, "done: %s\n", ssn_host_last_error()); exit(2); }
    }
    pthread_mutex_lock(&mu); g_windows += nw_tot; g_reads += nr_tot; pthread_mutex_unlock(&mu);
    free(rv.recs); free(rv.rec_off); free(rv.win); free(rv.ref_nt16); free(ref_chars);
    return 0;
}

int main(int argc, char **argv)
{
    int consumers, threads, repeats, r, i;
    unsigned long long first_w = 0, first_r = 0;
    if (argc < 7) return 1;
    consumers = atoi(argv[4]); threads = atoi(argv[5]); repeats = atoi(argv[6]);
    for (r = 0; r < repeats; ++r) {
        pthread_t th[16];
        ssn_shard_opts o;
        memset(&o, 0, sizeof(o));
        o.n_threads = threads;
        if (argc > 7) o.stride_windows = (uint32_t)atoi(argv[7]);
        if (ssn_shardset_open(argv[1], argv[2], argv[3], 0, &o, &set) != SSN_OK) { fprintf(stderr, "open: %s\n", ssn_host_last_error()); return 2; }
        g_windows = g_reads = 0;
        const double t0 = now_s();
        for (i = 0; i < consumers; ++i) pthread_create(&th[i], 0, consumer, 0);
        for (i = 0; i < consumers; ++i) pthread_join(th[i], 0);
        {
            const double dt = now_s() - t0;
            ssn_packer_stats ps;
            ssn_shardset_get_stats(set, &ps);
            printf("repeat %d: shards %d windows %llu reads %llu long_span %d  %.3f s, %.3g positions/s (wait %.2f s, fill %.2f s)\n", r,
                   (int)ssn_shardset_n_shards(set), g_windows, g_reads, ssn_shardset_long_span(set), dt, (double)g_windows * SSN_READ_WINDOW / dt, ps.t_wait, ps.t_fill);
        }
        if (r == 0) { first_w = g_windows; first_r = g_reads; }
        else if (g_windows != first_w || g_reads != first_r) { fprintf(stderr, "run %d differs from run 0\n", r); return 3; }
        ssn_shardset_close(set);
    }
    return 0;
}
