/*
 * ssn_shim.c — the binding INTEGRATION.md describes, written out and compiled against the UNMODIFIED reference.
 *
 * The reference installs its per-column scorer as a callback (glf_somatic, type bam_sspileup_f,
 * src/lib/sniper/somatic_sniper.h:42) at src/exe/bam-somaticsniper/main.c:146; its walker bam_sspileup_file calls it once
 * per column (src/lib/sniper/sniper_pileup.c:258).  This file supplies a callback of the same type that batches the
 * columns into libssn.so's pinned SoA slots (include/ssn.h), scores them on the GPU and feeds every passing site to the
 * reference's own output formatter (output_formatter_write, somatic_sniper.c:264).  Everything else — option parsing,
 * BAM/FASTA reading, the pileup walk, the writers — stays the reference's code.
 *
 * oracle/build_ref.sh compiles the reference's main.c with
 *     -Dglf_somatic=ssn_shim_column            the callback installed at main.c:146
 *     -DmakeSoloPrior=ssn_shim_solo_prior      main.c:122-125: only called without -p, which pu_data2_t does not record
 *     -Dbam_header_destroy=ssn_shim_finish     main.c:151: the first clean-up call behind the walk that can be redirected -
 *                                              the batches still in flight are drained while the tumor header (target
 *                                              names) and the output file are still there
 * i.e. the three one-line edits a maintainer would make by hand, and links this file, the other reference sources and
 * -lssn into oracle/_ref/bam-somaticsniper-shim.  Test infrastructure (tests/test_cli_gpu.py::test_reference_walker_
 * through_the_shim); the product's own CLI replaces the walker as well.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "ssn.h"

/* the reference's headers, with the names the build line redirects restored for this file */
#undef glf_somatic
#undef makeSoloPrior
#undef bam_header_destroy
#include "sniper/somatic_sniper.h"

#define SHIM_COLUMNS (1u << 16)
#define SHIM_ENTRIES ((uint64_t)12 << 20)
#define SHIM_SLOTS 3

typedef struct { ssn_batch *b; uint32_t *tid, *pos; uint8_t *rb; } shim_slot;

static struct {
    ssn_ctx *ctx;
    pu_data2_t *d;
    int use_priors;
    shim_slot slot[SHIM_SLOTS];
    shim_slot *cur, *fifo[SHIM_SLOTS];
    int f_head, f_count;
    ssn_batch_view v;
    uint32_t ncol;
    uint64_t nent;
} S;

static void die(const char *what)
{
    fprintf(stderr, "[ssn_shim] %s: %s\n", what, ssn_last_error());
    exit(1);                                    /* the reference's error style; there is no CPU path to fall back to */
}

void ssn_shim_solo_prior(void)
{
    S.use_priors = 1;
    makeSoloPrior();
}

static shim_slot *slot_of(ssn_batch *b)
{
    int i;
    for (i = 0; i < SHIM_SLOTS; ++i) if (S.slot[i].b == b) return &S.slot[i];
    for (i = 0; i < SHIM_SLOTS; ++i)
        if (!S.slot[i].b) {
            S.slot[i].b = b;
            S.slot[i].tid = (uint32_t*)malloc(sizeof(uint32_t) * SHIM_COLUMNS);
            S.slot[i].pos = (uint32_t*)malloc(sizeof(uint32_t) * SHIM_COLUMNS);
            S.slot[i].rb = (uint8_t*)malloc(SHIM_COLUMNS);
            return &S.slot[i];
        }
    return 0;
}

/* ssn_site -> sniper_output_t (output_format.h:13-35), one to one, then the reference's formatter */
static void drain_one(void)
{
    const ssn_site *sites;
    uint32_t n, i;
    ssn_batch *b;
    shim_slot *sl = S.fifo[S.f_head];
    if (ssn_results_next(S.ctx, &sites, &n, &b) != SSN_OK) die("ssn_results_next");
    for (i = 0; i < n; ++i) {
        const ssn_site *s = &sites[i];
        sniper_output_t out;
        const ssn_sample *src[2] = {&s->tumor, &s->normal};
        sample_data_t *dst[2] = {&out.tumor, &out.normal};
        int k, j;
        memset(&out, 0, sizeof(out));
        out.seq_name = S.d->h1->target_name[sl->tid[s->column]];
        out.pos = sl->pos[s->column];
        out.ref_base = sl->rb[s->column];
        out.ref_base4 = s->ref_base4;
        for (k = 0; k < 2; ++k) {
            dst[k]->genotype = src[k]->genotype;
            dst[k]->joint_genotype = src[k]->joint_genotype;
            dst[k]->joint_consensus_quality = src[k]->joint_consensus_quality;
            dst[k]->consensus_quality = src[k]->consensus_quality;
            dst[k]->variant_allele_quality = src[k]->variant_allele_quality;
            dst[k]->somatic_score = src[k]->somatic_score;
            dst[k]->variant_status = (variant_status_t)src[k]->variant_status;
            for (j = 0; j < 4; ++j) {
                dst[k]->dqstats.mean_baseQ[j] = src[k]->mean_baseQ[j];
                dst[k]->dqstats.mean_mapQ[j] = src[k]->mean_mapQ[j];
                dst[k]->dqstats.base_occ[j] = src[k]->base_occ[j];
                dst[k]->dqstats.dp4[j] = src[k]->dp4[j];
            }
            dst[k]->dqstats.total_depth = src[k]->total_depth;
            dst[k]->dqstats.total_mean_mapQ = src[k]->total_mean_mapQ;
        }
        output_formatter_write(S.d->output_formatter, &out);
    }
    (void)b;
    S.f_head = (S.f_head + 1) % SHIM_SLOTS;
    S.f_count -= 1;
}

static void next_batch(void)
{
    ssn_batch *b;
    while (S.f_count == SHIM_SLOTS) drain_one();
    if (ssn_batch_acquire(S.ctx, &b) != SSN_OK) die("ssn_batch_acquire");
    S.cur = slot_of(b);
    if (!S.cur || ssn_batch_view_get(b, &S.v) != SSN_OK) die("ssn_batch_view_get");
    S.ncol = 0; S.nent = 0;
}

static void submit_current(void)
{
    if (!S.cur) return;
    if (S.ncol == 0) { ssn_batch_release(S.ctx, S.cur->b); S.cur = 0; return; }
    S.v.off[2 * (size_t)S.ncol] = (uint32_t)S.nent;
    if (ssn_batch_set_size(S.cur->b, S.ncol, S.nent) != SSN_OK || ssn_batch_submit(S.ctx, S.cur->b, 0) != SSN_OK) die("ssn_batch_submit");
    S.fifo[(S.f_head + S.f_count) % SHIM_SLOTS] = S.cur;
    S.f_count += 1;
    S.cur = 0;
}

/* same signature as glf_somatic (bam_sspileup_f, somatic_sniper.h:42) */
int ssn_shim_column(uint32_t tid, uint32_t pos, int n1, int n2, const bam_pileup1_t *pl1, const bam_pileup1_t *pl2, void *data, FILE *snp_fh)
{
    pu_data2_t *d = (pu_data2_t*)data;
    int rb, pass;
    (void)snp_fh;
    if (!S.ctx) {
        ssn_params p;
        ssn_params_default(&p);
        p.theta = d->c->theta; p.het_rate = d->c->het_rate; p.eta = d->c->eta; p.n_hap = d->c->n_hap; p.cap_mapq = d->c->cap_mapQ;
        p.min_somatic_qual = d->min_somatic_qual; p.include_loh = d->include_loh; p.include_gor = d->include_gor;
        p.use_priors = S.use_priors; p.use_joint_priors = d->use_joint_priors; p.somatic_mutation_rate = d->somatic_mutation_rate;
        if (ssn_create(&p, 0, SHIM_SLOTS, SHIM_COLUMNS, SHIM_ENTRIES, &S.ctx) != SSN_OK) die("ssn_create");
        S.d = d;
    }
    if (!S.cur) next_batch();
    if (S.ncol == S.v.max_columns || S.nent + (uint64_t)n1 + (uint64_t)n2 > S.v.max_entries) {
        if (S.ncol == 0) { fprintf(stderr, "[ssn_shim] a column of %d read-bases exceeds the batch\n", n1 + n2); exit(1); }
        submit_current();
        next_batch();
    }
    /* the reference base exactly as glf_somatic finds it (somatic_sniper.c:112-117) */
    if (d->fai && (int)tid != d->tid) {
        free(d->ref);
        d->ref = fai_fetch(d->fai, d->h1->target_name[tid], &d->len);
        d->tid = (int)tid;
    }
    rb = (d->ref && (int)pos < d->len) ? d->ref[pos] : 'N';
    S.v.ref_nt16[S.ncol] = (uint8_t)ssn_nt16_of_char(rb);
    S.v.off[2 * (size_t)S.ncol] = (uint32_t)S.nent;
    for (pass = 0; pass < 2; ++pass) {
        const bam_pileup1_t *pl = pass ? pl2 : pl1;
        const int n = pass ? n2 : n1;
        int i;
        for (i = 0; i < n; ++i) {
            const bam_pileup1_t *p = pl + i;
            if (p->is_del || (p->b->core.flag & BAM_FUNMAP)) continue;              /* sniper_maqcns.c:147, dqstats.c:20 */
            S.v.bs[S.nent] = (uint8_t)(bam1_seqi(bam1_seq(p->b), p->qpos) | (bam1_strand(p->b) ? 16 : 0));
            S.v.bq[S.nent] = bam1_qual(p->b)[p->qpos];
            S.v.mq[S.nent] = (uint8_t)p->b->core.qual;
            ++S.nent;
        }
        S.v.off[2 * (size_t)S.ncol + 1 + pass] = (uint32_t)S.nent;
    }
    S.cur->tid[S.ncol] = tid; S.cur->pos[S.ncol] = pos; S.cur->rb[S.ncol] = (uint8_t)rb;
    ++S.ncol;
    return 0;
}

/* behind the walk, before main() frees the headers and closes the output file (main.c:149-159) */
void ssn_shim_finish(bam_header_t *h)
{
    if (S.ctx) {
        int i;
        submit_current();
        while (S.f_count > 0) drain_one();
        fflush(S.d->output_formatter->fh);
        ssn_destroy(S.ctx);
        S.ctx = 0;
        for (i = 0; i < SHIM_SLOTS; ++i) { free(S.slot[i].tid); free(S.slot[i].pos); free(S.slot[i].rb); }
    }
    bam_header_destroy(h);
}
