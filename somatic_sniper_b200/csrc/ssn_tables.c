/*
 * ssn_tables.c — host-side construction of the MAQ consensus model tables and the Phred-space
 * prior / log-add tables.  Compiled by gcc (never nvcc) with -ffp-contract=off: the numbers must
 * carry the reference's exact rounding, which mixes IEEE double with x87 80-bit long double and
 * glibc's expl/logl/lgamma/pow.  Built once per ssn_ctx; the coef rows for
 * different base-error levels q are independent and are spread over a few pthreads.
 *
 * Reference behaviour followed (paths relative to /root/reference/src/lib/sniper):
 *   coef, fk      sniper_maqcns.c:59-100  (sniper_cal_coef)
 *   lhet, q_r     sniper_maqcns.c:27-56   (sniper_cal_het)
 *   prior         somatic_sniper.c:29-45  (makeSoloPrior; skipped under -p, main.c:122-125)
 *   jointprior    somatic_sniper.c:47-77  (make_joint_prior; only under -J/-s, main.c:115-118)
 *   qadd          somatic_sniper.c:101-107 (qAddTableInit)
 */
#include "ssn_internal.h"

#include <math.h>
#include <pthread.h>
#include <stdlib.h>
#include <string.h>

#define PHRED_K 4.343

/* 10 genotypes in GLF order AA,AC,AG,AT,CC,CG,CT,GG,GT,TT as 4-bit allele masks */
static const int genotype_mask[10] = {1, 3, 5, 9, 2, 6, 10, 4, 12, 8};

static int popcount4(int m) { return (m & 1) + (m >> 1 & 1) + (m >> 2 & 1) + (m >> 3 & 1); }

/* (int) of +-0.5 - 4.343 ln x with C truncation: the reference's logPhred macro */
static int phred_of(double x)
{
    double v = x < 1 ? 0.5 - PHRED_K * log(x) : -0.5 - PHRED_K * log(x);
    return (int)v;
}

/* ---- coef ------------------------------------------------------------------------------------ */

typedef struct {
    double *coef;
    const double *fk;
    const double *log_binom; /* [n<<8|k] */
    int q_begin, q_end;
} coef_job;

static void *coef_worker(void *arg)
{
    coef_job *job = (coef_job*)arg;
    long double tail[257], ratio[256], cum[256], half_fk[256];
    int q, n, k;
    half_fk[0] = 1.0;
    for (n = 1; n < 256; ++n) half_fk[n] = job->fk[n >> 1];
    for (q = job->q_begin; q < job->q_end; ++q) {
        const double err = pow(10.0, -q / 10.0);
        const double ln_err = log(err), ln_ok = log(1.0 - err);
        for (n = 1; n < 256; ++n) {
            double *dst = job->coef + ((size_t)q << 16 | (size_t)n << 8);
            const double *lb = job->log_binom + (n << 8);
            tail[n + 1] = 0.0;
            for (k = n; k >= 0; --k) {
                /* argument is formed in double; expl and the running tail are 80-bit */
                tail[k] = tail[k + 1] + expl(lb[k] + k * ln_err + (n - k) * ln_ok);
                ratio[k] = tail[k + 1] / tail[k];
                if (ratio[k] > 0.99) ratio[k] = 0.99;
            }
            for (k = 0; k < n; ++k) cum[k] = -4.343 * half_fk[k] * logl(ratio[k] / err);
            for (k = 1; k < n; ++k) cum[k] += cum[k - 1];
            for (k = 0; k <= n; ++k) {
                long double last = -4.343 * logl(1.0 - expl(half_fk[k] * logl(ratio[k])));
                dst[k] = (k ? cum[k - 1] : 0) + last;
            }
        }
    }
    return 0;
}

static int build_coef(const ssn_params *p, ssn_host_tables *t)
{
    enum { NTHREAD = 8 };
    pthread_t th[NTHREAD];
    coef_job job[NTHREAD];
    double *log_binom;
    int n, k, i, started = 0;
    t->coef = (double*)calloc((size_t)64 * 65536, sizeof(double));
    log_binom = (double*)calloc(65536, sizeof(double));
    if (!t->coef || !log_binom) { free(log_binom); return -1; }
    t->fk[0] = 1.0;
    for (n = 1; n < 256; ++n) t->fk[n] = pow(p->theta, n) * (1.0 - p->eta) + p->eta;
    for (n = 1; n < 256; ++n)
        for (k = 1; k <= n; ++k)
            log_binom[n << 8 | k] = lgamma(n + 1) - lgamma(k + 1) - lgamma(n - k + 1);
    for (i = 0; i < NTHREAD; ++i) {
        job[i].coef = t->coef; job[i].fk = t->fk; job[i].log_binom = log_binom;
        job[i].q_begin = 1 + (63 * i) / NTHREAD;
        job[i].q_end = 1 + (63 * (i + 1)) / NTHREAD;
    }
    for (i = 1; i < NTHREAD; ++i) {
        if (pthread_create(&th[i], 0, coef_worker, &job[i]) != 0) break;
        ++started;
    }
    coef_worker(&job[0]);
    for (i = 1; i <= started; ++i) pthread_join(th[i], 0);
    for (i = started + 1; i < NTHREAD; ++i) coef_worker(&job[i]); /* thread creation failed: do it here */
    free(log_binom);
    return 0;
}

/* ---- lhet, q_r ---------------------------------------------------------------------------------- */

static int build_lhet(const ssn_params *p, ssn_host_tables *t)
{
    const int N = p->n_hap;
    double harmonic = 0.0, poly;
    int k, a, b;
    t->lhet = (double*)calloc(65536, sizeof(double));
    if (!t->lhet) return -1;
    for (k = 1; k <= N - 1; ++k) harmonic += 1.0 / k;
    for (a = 0; a < 256; ++a)
        for (b = 0; b < 256; ++b) {
            long double mix = 0.0;
            double lc = lgamma(a + b + 1) - lgamma(a + 1) - lgamma(b + 1);
            for (k = 1; k <= N - 1; ++k) {
                double pk = 1.0 / k / harmonic;
                double lx = log((double)k / N);
                double ly = log(1.0 - (double)k / N);
                mix += pk * 0.5 * (expl(lx * b) * expl(ly * a) + expl(lx * a) * expl(ly * b));
            }
            t->lhet[a << 8 | b] = lc + logl(mix);
        }
    poly = p->het_rate * harmonic;
    t->q_r = -4.343 * log(2.0 * poly / (1.0 - poly));
    return 0;
}

/* ---- priors ---------------------------------------------------------------------------------------- */

static int germline_phred(int ref_mask, int gt_mask)
{
    const double pop_theta = 0.001; /* THETA, somatic_sniper.c:14 — unrelated to -T */
    if ((gt_mask & ~ref_mask) == 0) return 0;
    if (gt_mask & ref_mask) return phred_of(pop_theta);
    if (popcount4(gt_mask) == 1) return phred_of(0.5 * pop_theta);
    return phred_of(pop_theta * pop_theta);
}

static void build_priors(const ssn_params *p, ssn_host_tables *t)
{
    int ref, i, j;
    memset(t->prior, 0, sizeof(t->prior));
    memset(t->jointprior, 0, sizeof(t->jointprior));
    if (p->use_priors)
        for (ref = 0; ref < 16; ++ref)
            for (i = 0; i < 10; ++i) t->prior[ref * 10 + i] = germline_phred(ref, genotype_mask[i]);
    if (p->use_joint_priors) {
        const int one_step = phred_of(p->somatic_mutation_rate);
        const int two_step = phred_of(p->somatic_mutation_rate * p->somatic_mutation_rate);
        for (ref = 0; ref < 16; ++ref)
            for (i = 0; i < 10; ++i) {
                const int normal = genotype_mask[i], g = germline_phred(ref, normal);
                for (j = 0; j < 10; ++j) {
                    const int tumor = genotype_mask[j];
                    /* The reference looks its isHet/isHom tables up with the genotype INDEX j, not the
                     * mask (somatic_sniper.c:66-72).  As masks, indices 1,2,4,8 are homozygous and
                     * 3,5,6,9 heterozygous; indices 0 and 7 are neither, so a tumor AA or GG that shares
                     * an allele with a different normal genotype is charged the two-step penalty. */
                    const int index_is_genotype = popcount4(j) == 1 || popcount4(j) == 2;
                    int v = g;
                    if (normal != tumor) v += ((normal & tumor) && index_is_genotype) ? one_step : two_step;
                    t->jointprior[(ref * 10 + i) * 10 + j] = v;
                }
            }
    }
}

static void build_qadd(ssn_host_tables *t)
{
    int i;
    memset(t->qadd, 0, sizeof(t->qadd));
    /* entries 1000..1023 are never written by the reference and stay 0; SSN_QADD_PAD more zeros
     * keep an index of 1024+ (possible in independent mode when the running sum dips below 0) defined */
    for (i = 0; i < 1000; ++i)
        t->qadd[i] = phred_of(1 + exp((double)(-(i - 512)) / PHRED_K));
}

int ssn_tables_build(const ssn_params *p, ssn_host_tables *t)
{
    memset(t, 0, sizeof(*t));
    if (p->n_hap < 2 || p->n_hap > 1 << 20) return -1;
    if (build_coef(p, t) != 0 || build_lhet(p, t) != 0) { ssn_tables_free(t); return -1; }
    build_priors(p, t);
    build_qadd(t);
    return 0;
}

void ssn_tables_free(ssn_host_tables *t)
{
    free(t->coef); free(t->lhet);
    t->coef = t->lhet = 0;
}
