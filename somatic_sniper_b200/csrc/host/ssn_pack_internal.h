/* ssn_pack_internal.h — what the region-shard scheduler (ssn_shard.c) needs from the packer (ssn_pack.c). */
#ifndef SSN_PACK_INTERNAL_H
#define SSN_PACK_INTERNAL_H
#include "ssn_bamread.h"
#include "ssn_bamseek.h"
#include "ssn_host.h"
#include "ssn_pool.h"

typedef struct ssn_refcache ssn_refcache;

/* target sequences shared by every packer of a run; `names` (the tumor header, QUIRK B9) must outlive the cache */
ssn_refcache *ssn_refcache_create(const char *fasta, const ssn_bam_header *names);
void ssn_refcache_destroy(ssn_refcache *rc);

/* a packer over one region shard: `pool` serves its count/fill/copy tasks, `reader_pool` its readers' inflate (NULL:
 * each reader inflates by itself); nothing passed in is owned by the packer */
ssn_packer *ssn_packer_create_shard(ssn_pool *pool, ssn_pool *reader_pool, ssn_refcache *rc, const ssn_bamfile *tumor,
                                    const ssn_bamfile *normal, const ssn_bam_range *range, int min_mapq,
                                    void (*on_finish)(void*), void *on_finish_arg, char *err, size_t errlen);
int ssn_packer_long_span(const ssn_packer *pk);
int ssn_packer_failed(const ssn_packer *pk);
void ssn_host_set_error(const char *fmt, ...);

#endif
