/* ssn_inflate.h — raw DEFLATE decoder for whole BGZF blocks (ssn_inflate.c). */
#ifndef SSN_INFLATE_H
#define SSN_INFLATE_H
#include <stddef.h>
#include <stdint.h>

#define SSN_INFLATE_SLACK 32   /* bytes that must be readable behind in[in_len) */

/* 0: the stream decoded to exactly out_len bytes; -1: anything else (callers fall back to zlib for that block) */
int ssn_inflate_raw(const uint8_t *in, size_t in_len, uint8_t *out, size_t out_len);

#endif
