/* ssn_pool.h — small pthread pool with a blocking parallel-for; several client threads may call
 * ssn_pool_for() at the same time (BAM readers inflating blocks, the packer counting/filling windows). */
#ifndef SSN_POOL_H
#define SSN_POOL_H
#ifdef __cplusplus
extern "C" {
#endif

typedef struct ssn_pool ssn_pool;
typedef void (*ssn_pool_fn)(void *arg, int index);

ssn_pool *ssn_pool_create(int n_threads);   /* n_threads <= 0: run everything inline in the caller */
void ssn_pool_destroy(ssn_pool *p);
int ssn_pool_threads(const ssn_pool *p);
/* runs fn(arg, i) for i in [0, n); returns when all are done; the caller helps */
void ssn_pool_for(ssn_pool *p, int n, ssn_pool_fn fn, void *arg);

#ifdef __cplusplus
}
#endif
#endif
