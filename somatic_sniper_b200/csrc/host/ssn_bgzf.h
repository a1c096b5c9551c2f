/* ssn_bgzf.h — sequential BGZF reader with block-parallel inflate.
 * Replaces what the reference gets from samtools-0.1.6 bgzf.c (bgzf_read :489, read_block :434,
 * inflate_block :330) for the two input BAMs.  Same stopping rules: the stream ends at end of file,
 * at the first empty block (ISIZE 0 — bgzf_read breaks out there, so bam_read1 sees end-of-file), or
 * at a block that cannot be read/inflated. */
#ifndef SSN_BGZF_H
#define SSN_BGZF_H
#include <stddef.h>
#include <stdint.h>
#include "ssn_pool.h"

typedef struct ssn_bgzf ssn_bgzf;

ssn_bgzf *ssn_bgzf_open(const char *path, ssn_pool *pool);    /* "-" reads stdin; NULL on failure */
/* the same, starting at the block that begins at file offset coff (region shards; not for stdin) */
ssn_bgzf *ssn_bgzf_open_at(const char *path, ssn_pool *pool, uint64_t coff);
/* blocks go through an inflate service (include/ssn_host.h) from now on; a lane is taken here and given back at close.
 * Returns 0 when no lane could be had (the reader then keeps inflating on the host). */
struct ssn_inflate_service_tag;
int ssn_bgzf_use_service(ssn_bgzf *z, const void *svc);
/* Next run of decompressed bytes (a few MB); valid until the next call.  Returns 1 with data, 0 at the end. */
int ssn_bgzf_next(ssn_bgzf *z, const uint8_t **data, size_t *len);
void ssn_bgzf_close(ssn_bgzf *z);
uint64_t ssn_bgzf_compressed_bytes(const ssn_bgzf *z);

#endif
