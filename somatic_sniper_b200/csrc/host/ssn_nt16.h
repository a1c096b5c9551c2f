/* ssn_nt16.h — the two character tables of samtools-0.1.6 the path depends on:
 * bam_nt16_table (bam_import.c:23-40: reference/read character -> 4-bit code, 15 for anything unknown,
 * '0'..'3' -> 1,2,4,8) and bam_nt16_rev_table (bam_import.c:61: "=ACMGRSVTWYHKDBN"). */
#ifndef SSN_NT16_H
#define SSN_NT16_H

static const char ssn_nt16_rev[] = "=ACMGRSVTWYHKDBN";

static inline int ssn_nt16_slow(int ch)
{
    int i;
    ch &= 0xff;
    if (ch >= '0' && ch <= '3') return 1 << (ch - '0');
    for (i = 0; i < 16; ++i) {
        const int r = ssn_nt16_rev[i];
        if (ch == r || (r >= 'A' && r <= 'Z' && ch == r + ('a' - 'A'))) return i;
    }
    return 15;
}

/* the same through a 256-entry table (filled on first use; the packer asks once per reference position) */
static inline int ssn_nt16_table(int ch)
{
    static unsigned char lut[256];
    static volatile int ready = 0;
    if (!ready) {
        int i;
        for (i = 0; i < 256; ++i) lut[i] = (unsigned char)ssn_nt16_slow(i);   /* idempotent: a race writes the same bytes */
        __sync_synchronize();
        ready = 1;
    }
    return lut[ch & 0xff];
}

#endif
