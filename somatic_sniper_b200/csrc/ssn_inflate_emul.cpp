/* ssn_inflate_emul.cpp — the device DEFLATE decoder (ssn_inflate_dev.cuh) compiled for the host, lane loops as real
 * loops: lets the CPU test suite check the very source the GPU runs against zlib (tests/test_inflate.py).  Test
 * infrastructure; nothing in the product calls it. */
#define SSN_INFL_HOST 1
#include "ssn_inflate_dev.cuh"

#include <cstdlib>
#include <cstring>

extern "C" int ssn_inflate_dev_host(const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t out_len)
{
    /* the bit reader loads aligned 32-bit words: give it an aligned, padded copy */
    uint8_t *buf = (uint8_t*)calloc(1, (size_t)in_len + 64);
    infl::Work *W = (infl::Work*)malloc(sizeof(infl::Work));
    int rc;
    if (!buf || !W) { free(buf); free(W); return -1; }
    memcpy(buf + 4, in, in_len);          /* deliberately misaligned by 4 + (caller's choice below) */
    rc = infl::inflate_block(*W, buf + 4, in_len, out, out_len);
    free(buf); free(W);
    return rc;
}

/* the same with the input starting `shift` bytes (0..3) past a word boundary */
extern "C" int ssn_inflate_dev_host_shift(const uint8_t *in, uint32_t in_len, uint8_t *out, uint32_t out_len, uint32_t shift)
{
    uint8_t *buf = (uint8_t*)calloc(1, (size_t)in_len + 64);
    infl::Work *W = (infl::Work*)malloc(sizeof(infl::Work));
    int rc;
    if (!buf || !W) { free(buf); free(W); return -1; }
    memcpy(buf + 8 + (shift & 3u), in, in_len);
    rc = infl::inflate_block(*W, buf + 8 + (shift & 3u), in_len, out, out_len);
    free(buf); free(W);
    return rc;
}
