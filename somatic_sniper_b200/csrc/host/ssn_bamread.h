/* ssn_bamread.h — one BAM -> stream of position windows holding the reads the reference's pileup buffer
 * would keep.  Replaces, for the packer, samtools-0.1.6 bam_header_read (bam.c:90) / bam_read1 (bam.c:181)
 * and the read-ahead half of get_next_pos (src/lib/sniper/sniper_pileup.c:207-222):
 *   - read filter  (flag & mask) == 0 && mapQ >= threshold                       sniper_pileup.c:208
 *   - abort when the target id decreases among passing reads                     sniper_pileup.c:211-214
 *   - keep rule    end > cursor position at fetch time                           sniper_pileup.c:216
 *     The cursor stands on the start of the previously fetched passing read whenever a read is fetched
 *     (it is fetched exactly when the read-ahead test of :185 fails), so a passing read is kept iff
 *     bam_calend(read) > start of the previous passing read (0 for the first read of the file) — on
 *     sorted input this only ever drops the first read of a later target (SURVEY.md QUIRK B1).
 * A reader thread frames records out of the inflated stream and files every kept read under each
 * window of SSN_WINDOW positions it overlaps; finished windows are handed over in genomic order. */
#ifndef SSN_BAMREAD_H
#define SSN_BAMREAD_H
#include <stddef.h>
#include <stdint.h>
#include "ssn_pool.h"

#define SSN_WINDOW_SHIFT 13
#define SSN_WINDOW (1u << SSN_WINDOW_SHIFT)
#define SSN_BAM_DEF_MASK 0x704     /* BAM_FUNMAP | BAM_FSECONDARY | BAM_FQCFAIL | BAM_FDUP, samtools bam.h:121 */

/* read as stored in a window buffer (4-byte aligned); followed by cigar[n_cigar] (uint32), seq[(l_qseq+1)/2], qual[l_qseq] */
typedef struct {
    int32_t beg, end;          /* 0-based start, bam_calend() end */
    uint32_t l_qseq;
    uint16_t n_cigar;
    uint8_t mapq, strand;
    uint32_t bytes;            /* size of this record including the header, multiple of 4 */
} ssn_read;

typedef struct ssn_window {
    int32_t tid;
    uint32_t widx;             /* positions [widx << SSN_WINDOW_SHIFT, (widx+1) << SSN_WINDOW_SHIFT) */
    uint32_t n_reads;
    uint64_t n_bases;          /* sum of max(l_qseq, M-run length): upper bound of the entries the window can produce */
    uint8_t *buf;
    size_t len, cap;
    uint32_t *rec_word;        /* [n_reads] offset / 4 of every record inside buf (the read batch's index, ready to be rebased) */
    uint32_t cap_reads;
    struct ssn_window *next;
} ssn_window;

typedef struct {
    int32_t n_targets;
    char **target_name;
    uint32_t *target_len;
    char *text;
    int32_t l_text;
} ssn_bam_header;

typedef struct ssn_bamreader ssn_bamreader;

/* Opens the file, parses the header (synchronously) and starts the reader thread. */
ssn_bamreader *ssn_bamreader_open(const char *path, int flag_mask, int mapq_thresh, ssn_pool *pool, char *err, size_t errlen);

/* Region shard: the reader files the windows of [ (lo_tid, lo_pos), (hi_tid, hi_pos) ) only - both multiples of
 * SSN_WINDOW, hi_tid = INT32_MAX for "to the end" - and stops at the first record at or behind hi.  It starts `lookback`
 * positions in front of lo (reads beginning there may reach into the shard) with the keep rule's state taken over
 * from the last passing read before that point, so every keep decision equals the front-to-back walk's
 * (sniper_pileup.c:216, QUIRK B1 across shard edges).  A kept read that begins more than `lookback` in front of hi and
 * reaches behind it is invisible to the next shard: ssn_bamreader_long_span() then reports 1 and the caller must fall
 * back to one shard.  The thread starts at once; the seek happens inside it.  on_finish (optional) is called from the
 * reader thread when it is done. */
struct ssn_bamfile;
typedef struct {
    int32_t lo_tid; uint32_t lo_pos;
    int32_t hi_tid; uint32_t hi_pos;
    uint32_t lookback;
    size_t queue_max_bytes;        /* 0: default back-pressure limit */
    const void *inflate_service;   /* ssn_inflate_service* (include/ssn_host.h) or NULL: inflate on the host */
} ssn_bam_range;
ssn_bamreader *ssn_bamreader_open_range(const struct ssn_bamfile *bf, const ssn_bam_range *range, int flag_mask, int mapq_thresh,
                                        ssn_pool *pool, void (*on_finish)(void*), void *on_finish_arg, char *err, size_t errlen);
int ssn_bamreader_long_span(const ssn_bamreader *r);
int ssn_bamreader_failed(const ssn_bamreader *r);
const ssn_bam_header *ssn_bamreader_header(const ssn_bamreader *r);
/* Next finished window in (tid, widx) order, NULL after the last one.  The caller owns it (ssn_window_free). */
ssn_window *ssn_bamreader_pop(ssn_bamreader *r);
void ssn_window_free(ssn_window *w);
/* Stops the thread (the rest of the file is not read) and frees everything. */
void ssn_bamreader_close(ssn_bamreader *r);
uint64_t ssn_bamreader_records(const ssn_bamreader *r);

#endif
