/*
 * ssn_prefilter_tables.c — bound tables of the exact hom-reference pre-filter.
 *
 * glf_somatic only looks at a column whose two consensus genotypes differ (somatic_sniper.c:156).  For the
 * overwhelming majority of WGS columns both samples are plainly homozygous for the same base R, and that can be
 * PROVEN from a handful of per-sample sums without running the error-dependency chain of sniper_maqcns_glfgen
 * (sniper_maqcns.c:127-248): the tables below turn those sums into rigorous lower/upper bounds of the ten
 * genotype terms p[] (sniper_maqcns.c:186-233).  A column is dropped only when the bounds show that genotype RR is the
 * strict winner of sniper_glf2cns (sniper_maqcns.c:250-273) in BOTH samples; every other column goes through the full
 * model.  The derivation, with every inequality, is in DESIGN.md §4b; tests/test_prefilter.py checks it against the
 * oracle on the CPU and tests/test_gpu_parity.py audits the kernel's decisions against the full model on the GPU.
 *
 * Notation: n = read-bases of the sample (all contribute: base qualities in [1,63]), m = read-bases NOT equal to the
 * reference base (a superset of the read-bases outside R's chain), coef/lhet/fk = the model tables.
 *   ab[n][m] = { cmin_all, cmax_run, g2_run, 0 }
 *     cmin_all = min(0, min_{k in [n-m, n], k >= 1} min_{e in [4,63]} coef[e][n][k])   (p[XX], p[XY] >= chain(R) + cmin_all)
 *     cmax_run = max(0, max_{1 <= k <= m}          max_{e}           coef[e][n][k])   (p[RR] <= sum(others) + cmax_run)
 *     g2_run   = min(0, min_{1 <= k <= m} min_e ((e - 0.5) * F[k] + coef[e][n][k]))    (p[RX] >= h + g2_run)
 *   h[a][b]  = min_{a' >= a, b' <= b, a'+b' <= 255} -4.343 * max(lhet[a'][b'], lhet[b'][a'])
 *   F[k]     = (1 - 1e-4) * sum_{w < k} fk[w]
 *   vt[n][m][s] = max_{1 <= k <= m} max_e ( min(s, (e + 0.5) k (1 + 1e-4)) + coef[e][n][k] )   for m <= 8, s < 256:
 *              a tighter upper bound of p[RR] than SU + cmax_run when the m other read-bases' qualities sum to SU <= s
 *              (esum <= SU, and esum <= (bar_e + 0.5) * fsum <= (bar_e + 0.5) * k because bar_e = round(esum / fsum))
 * Lower bounds are rounded down to float, upper bounds up.
 */
#include "ssn_internal.h"

#include <float.h>
#include <math.h>
#include <stdlib.h>
#include <string.h>

static float f_down(double x) { float f = (float)x; return (double)f > x ? nextafterf(f, -INFINITY) : f; }
static float f_up(double x) { float f = (float)x; return (double)f < x ? nextafterf(f, INFINITY) : f; }

void ssn_prefilter_free(ssn_prefilter_tables *t)
{
    free(t->ab); free(t->h); free(t->vt);
    t->ab = t->h = t->vt = 0; t->enabled = 0;
}

int ssn_prefilter_build(const ssn_host_tables *m, int q_r_int, ssn_prefilter_tables *t)
{
    const int N = SSN_PF_N;
    int n, k, e, a, b;
    double Fd[SSN_PF_N + 1];
    double *cmin, *cmax, *g2, *hh;
    memset(t, 0, sizeof(*t));
    t->ab = (float*)calloc((size_t)N * N * 4, sizeof(float));
    t->h = (float*)calloc((size_t)N * N, sizeof(float));
    t->vt = (float*)calloc((size_t)N * SSN_PF_VK * SSN_PF_VS, sizeof(float));
    cmin = (double*)malloc(sizeof(double) * N * N);
    cmax = (double*)malloc(sizeof(double) * N * N);
    g2 = (double*)malloc(sizeof(double) * N * N);
    hh = (double*)malloc(sizeof(double) * N * N);
    if (!t->ab || !t->h || !t->vt || !cmin || !cmax || !g2 || !hh) { free(cmin); free(cmax); free(g2); free(hh); ssn_prefilter_free(t); return -1; }

    /* preconditions of the proof: dependency weights in (0,1], non-increasing, fk[0] = 1; lhet <= 0; a positive
     * heterozygote penalty.  Parameter sets that break one of them simply run without the pre-filter. */
    t->enabled = q_r_int >= 1 && m->fk[0] == 1.0;
    for (k = 1; k < 256; ++k)
        if (!(m->fk[k] > 0.0) || m->fk[k] > m->fk[k - 1]) t->enabled = 0;
    for (k = 0; k < 65536; ++k)
        if (!(m->lhet[k] <= 1e-9)) t->enabled = 0;

    Fd[0] = 0.0;
    for (k = 1; k <= N; ++k) Fd[k] = Fd[k - 1] + m->fk[k - 1];
    for (k = 0; k <= N; ++k) t->F[k] = f_down(Fd[k] * (1.0 - 1e-4));

    for (n = 0; n < N; ++n)
        for (k = 0; k < N; ++k) {
            double lo = 0.0, hi = 0.0, g = 0.0;
            if (n >= 1 && k >= 1 && k <= n) {
                lo = DBL_MAX; hi = -DBL_MAX; g = DBL_MAX;
                for (e = 4; e < 64; ++e) {
                    const double c = m->coef[(size_t)e << 16 | (size_t)n << 8 | (size_t)k];
                    const double v = ((double)e - 0.5) * (double)t->F[k] + c;
                    if (c < lo) lo = c;
                    if (c > hi) hi = c;
                    if (v < g) g = v;
                }
            }
            cmin[n * N + k] = lo; cmax[n * N + k] = hi; g2[n * N + k] = g;
        }
    for (n = 0; n < N; ++n) {
        double run_max = 0.0, run_g2 = 0.0;
        for (b = 0; b < N; ++b) {          /* b = m, the number of non-reference read-bases */
            float *dst = t->ab + ((size_t)n * N + b) * 4;
            double all = 0.0;
            if (b >= 1 && b <= n) {
                if (cmax[n * N + b] > run_max) run_max = cmax[n * N + b];
                if (g2[n * N + b] < run_g2) run_g2 = g2[n * N + b];
            }
            for (k = n - b < 1 ? 1 : n - b; k <= n; ++k)
                if (cmin[n * N + k] < all) all = cmin[n * N + k];
            dst[0] = f_down(all); dst[1] = f_up(run_max); dst[2] = f_down(run_g2); dst[3] = 0.0f;
        }
    }
    for (n = 1; n < N; ++n) {
        double run[SSN_PF_VS];
        for (k = 0; k < SSN_PF_VS; ++k) run[k] = -DBL_MAX;
        for (k = 1; k <= SSN_PF_VK; ++k) {
            float *dst = t->vt + ((size_t)n * SSN_PF_VK + (k - 1)) * SSN_PF_VS;
            if (k <= n) {
                for (e = 4; e < 64; ++e) {
                    const double c = m->coef[(size_t)e << 16 | (size_t)n << 8 | (size_t)k];
                    const double cap = ((double)e + 0.5) * (double)k * (1.0 + 1e-4);
                    for (b = 0; b < SSN_PF_VS; ++b) {
                        const double v = ((double)b < cap ? (double)b : cap) + c;
                        if (v > run[b]) run[b] = v;
                    }
                }
            }
            for (b = 0; b < SSN_PF_VS; ++b) dst[b] = f_up(run[b] > 0.0 ? run[b] : 0.0);
        }
    }
    /* h: closed under a' >= a, b' <= b inside a+b <= 255 */
    for (a = N - 1; a >= 0; --a)
        for (b = 0; b < N; ++b) {
            double v;
            if (a + b > 255) { hh[a * N + b] = DBL_MAX; continue; }
            {
                const double l1 = m->lhet[a << 8 | b], l2 = m->lhet[b << 8 | a];
                v = -4.343 * (l1 > l2 ? l1 : l2);
            }
            if (a + 1 < N && a + 1 + b <= 255 && hh[(a + 1) * N + b] < v) v = hh[(a + 1) * N + b];
            if (b >= 1 && hh[a * N + b - 1] < v) v = hh[a * N + b - 1];
            hh[a * N + b] = v;
        }
    for (a = 0; a < N; ++a)
        for (b = 0; b < N; ++b)
            t->h[a * N + b] = a + b > 255 ? -FLT_MAX : f_down(hh[a * N + b]);
    free(cmin); free(cmax); free(g2); free(hh);
    return 0;
}
