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
void ssn_tables_free(ssn_host_tables *t);

/* bam_nt16_table (samtools-0.1.6 bam_import.c:23-40): reference character -> 4-bit code */
int ssn_nt16_of_char(int ch);

void ssn_set_error(const char *fmt, ...);

#ifdef __cplusplus
}
#endif
#endif
