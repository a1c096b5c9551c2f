/* ssn_bamseek.h — random access into a coordinate-sorted BAM without an index file.
 *
 * The reference walks each BAM front to back with bam_read1 (src/lib/sniper/sniper_pileup.c:207-222, samtools-0.1.6
 * bam.c:181-212, bgzf.c:434-520) and has no region option.  Region shards need to start reading in
 * the middle of the file, so this module finds record boundaries there:
 *   probe        the first record that starts in the first BGZF block at or after an arbitrary file offset.  A block
 *                start is recognised by its gzip/BGZF header plus a valid successor block; a record start by a chain of
 *                plausible, coordinate-sorted records up to the end of the inflated look-ahead.  Every use below walks
 *                on from such a guess record by record, and the walk that precedes a shard (prev_passing) has to land
 *                on the shard's first record exactly, which proves the guess.
 *   lower_bound  bisection over file offsets for the first record with (tid, pos) >= a target (unplaced reads, tid -1,
 *                sort last), then a short walk.
 *   prev_passing the last record before a virtual offset that passes the reference's read filter
 *                ((flag & mask) == 0 && mapQ >= threshold, sniper_pileup.c:208): the state the keep rule of :216
 *                (QUIRK B1) needs when a reader starts there.
 * All functions are thread-safe (pread, no shared mutable state). */
#ifndef SSN_BAMSEEK_H
#define SSN_BAMSEEK_H
#include <stddef.h>
#include <stdint.h>
#include "ssn_bamread.h"

typedef struct ssn_bamfile ssn_bamfile;

typedef struct {
    uint64_t voff;             /* compressed offset of the block << 16 | offset inside the inflated block */
    int32_t tid, pos;
} ssn_bam_point;

ssn_bamfile *ssn_bamfile_open(const char *path, char *err, size_t errlen);
void ssn_bamfile_close(ssn_bamfile *f);
const ssn_bam_header *ssn_bamfile_header(const ssn_bamfile *f);
const char *ssn_bamfile_path(const ssn_bamfile *f);
uint64_t ssn_bamfile_first_record(const ssn_bamfile *f);       /* virtual offset of the first record */
uint64_t ssn_bamfile_size(const ssn_bamfile *f);

/* 1 + *out: first record starting in a block at or after compressed offset coff; 0: none (end of data); -1: error */
int ssn_bamfile_probe(const ssn_bamfile *f, uint64_t coff, ssn_bam_point *out);
/* 1 + *out: first record with (tid, pos) >= the target; 0: every record sorts before it; -1: error */
int ssn_bamfile_lower_bound(const ssn_bamfile *f, int32_t tid, int32_t pos, ssn_bam_point *out);
/* 1 + (*tid, *pos): last passing record before voff; 0: none; -1: error (the walk did not land on voff) */
int ssn_bamfile_prev_passing(const ssn_bamfile *f, uint64_t voff, int flag_mask, int mapq_thresh, int32_t *tid, int32_t *pos);

/* sort key of a record: unplaced reads (tid < 0) after every target */
static inline uint64_t ssn_bam_key(int32_t tid, int32_t pos)
{
    return (uint64_t)(tid < 0 ? 0x7fffffffu : (uint32_t)tid) << 32 | (uint32_t)(pos < 0 ? 0 : pos);
}

#endif
