/* ssn_internal.h — declarations shared by the host C and the CUDA translation units of libssn.so. */
#ifndef SSN_INTERNAL_H
#define SSN_INTERNAL_H

#include "ssn.h"

#ifdef __cplusplus
extern "C" {
#endif

#define SSN_QADD_PAD 64

typedef struct {
    double *coef;               /* 64*256*256 */
    double *lhet;               /* 256*256 */
    double fk[256];
    float q_r;
    int32_t prior[16 * 10];
    int32_t jointprior[16 * 10 * 10];
    int32_t qadd[1024 + SSN_QADD_PAD];
} ssn_host_tables;

int ssn_tables_build(const ssn_params *p, ssn_host_tables *t);

/* bound tables of the exact hom-reference pre-filter (ssn_prefilter_tables.c) */
#define SSN_PF_N 256
#define SSN_PF_VK 8
#define SSN_PF_VS 256
typedef struct {
    float *ab;                  /* [n][m][4] = { cmin_all, cmax_run, g2_run, 0 } */
    float *h;                   /* [a][b] */
    float *vt;                  /* [n][k-1][s], k in 1..SSN_PF_VK, s in 0..SSN_PF_VS-1: upper bound of p[RR] given SU <= s */
    float F[SSN_PF_N + 1];      /* F[k] = (1 - 1e-4) * sum_{w<k} fk[w], rounded down */
    int enabled;                /* 0: the parameter set breaks a precondition of the proof */
} ssn_prefilter_tables;
int ssn_prefilter_build(const ssn_host_tables *m, int q_r_int, ssn_prefilter_tables *t);
void ssn_prefilter_free(ssn_prefilter_tables *t);
void ssn_tables_free(ssn_host_tables *t);

/* bam_nt16_table (samtools-0.1.6 bam_import.c:23-40): reference character -> 4-bit code */
int ssn_nt16_of_char(int ch);

void ssn_set_error(const char *fmt, ...);

#ifdef __cplusplus
}
#endif
#endif
