/*
 * ref_harness.c — TEST INFRASTRUCTURE ONLY (never linked into the product).
 *
 * Function-level driver around the UNMODIFIED reference sources, compiled where
 * they lie under /root/reference by oracle/build_ref.sh into
 * oracle/_ref/libref_harness.so.  Nothing from the reference is copied into this
 * repository: the two translation units that own file-static state are pulled in
 * with #include so the harness can read `prior`, `jointprior`, `qAddTable`
 * (somatic_sniper.c:13-16) after the reference's own initialisers ran.
 *
 * A pileup column is fabricated exactly as SURVEY.md §8(c) describes: every
 * entry becomes a one-base bam1_t (l_qseq=1, data={nt16<<4, baseQ},
 * core.qual=mapQ, core.flag=strand?16:0) wrapped in a bam_pileup1_t with qpos=0.
 *
 * Entry encoding shared with the product and the oracle port:
 *     bs[i] = nt16 | strand<<4     bq[i] = base quality     mq[i] = mapping quality
 */
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "sniper/somatic_sniper.c"   /* reference TU: glf_somatic + static tables */

/* ---- model state ------------------------------------------------------- */
static pu_data2_t *g_d = 0;
static bam_header_t g_hdr;
static char *g_tname[1];
static char g_name0[] = "chrH";
static FILE *g_null = 0;
static sniper_output_t g_last;
static int g_emitted = 0;
static output_formatter_t g_fmt;

static void capture_fn(FILE *fh, const sniper_output_t *p) { (void)fh; g_last = *p; g_emitted = 1; }
static void capture_hdr(FILE *fh, const header_data_t *h) { (void)fh; (void)h; }

/* Mirrors main.c:64-127 (defaults, option effects, init order). */
int ref_init(float theta, float het_rate, int n_hap, int use_priors, int use_joint,
             double somatic_rate, int min_somatic_qual, int include_loh, int include_gor)
{
    if (g_d) {
        sniper_maqcns_destroy(g_d->c); free(g_d->ref); free(g_d); g_d = 0;
    }
    memset(prior, 0, sizeof(prior));
    memset(jointprior, 0, sizeof(jointprior));
    memset(qAddTable, 0, sizeof(qAddTable));
    g_d = (pu_data2_t*)calloc(1, sizeof(pu_data2_t));
    g_d->min_somatic_qual = min_somatic_qual;
    g_d->tid = 0; g_d->mask = BAM_DEF_MASK; g_d->mapQ = 0;
    g_d->c = sniper_maqcns_init();
    g_d->c->theta = theta; g_d->c->het_rate = het_rate; g_d->c->n_hap = n_hap;
    g_d->include_loh = include_loh; g_d->include_gor = include_gor;
    g_d->use_joint_priors = use_joint; g_d->somatic_mutation_rate = somatic_rate;
    if (use_joint) make_joint_prior(somatic_rate);
    sniper_maqcns_prepare(g_d->c);
    if (use_priors) makeSoloPrior();
    qAddTableInit();
    g_tname[0] = g_name0;
    memset(&g_hdr, 0, sizeof(g_hdr));
    g_hdr.n_targets = 1; g_hdr.target_name = g_tname;
    g_d->h1 = &g_hdr; g_d->h2 = &g_hdr; g_d->fai = 0;
    g_d->ref = (char*)malloc(2); g_d->ref[0] = 'N'; g_d->ref[1] = 0; g_d->len = 1;
    if (!g_null) g_null = fopen("/dev/null", "w");
    g_fmt.fh = g_null; g_fmt.header_fn = capture_hdr; g_fmt.output_fn = capture_fn; g_fmt.data = 0;
    g_d->output_formatter = &g_fmt;
    return 0;
}

/* ---- table dumps --------------------------------------------------------- */
const double *ref_coef(void) { return g_d->c->coef; }   /* 64*256*256 */
const double *ref_lhet(void) { return g_d->c->lhet; }   /* 256*256    */
const double *ref_fk(void)   { return g_d->c->fk; }     /* 256        */
float ref_q_r(void)          { return g_d->c->q_r; }
const int *ref_prior(void)      { return &prior[0][0]; }         /* 16*10    */
const int *ref_jointprior(void) { return &jointprior[0][0][0]; } /* 16*10*10 */
const int *ref_qadd(void)       { return qAddTable; }            /* 1024     */

/* ---- fabricated pileups ---------------------------------------------------- */
typedef struct { bam1_t *b; uint8_t *data; bam_pileup1_t *pl; int cap; } fake_pu_t;
static fake_pu_t g_pu[2];

static const bam_pileup1_t *fabricate(fake_pu_t *f, int n, const uint8_t *bs, const uint8_t *bq, const uint8_t *mq)
{
    int i;
    if (f->cap < n) {
        f->cap = n < 64 ? 64 : n;
        f->b = (bam1_t*)realloc(f->b, sizeof(bam1_t) * f->cap);
        f->data = (uint8_t*)realloc(f->data, 2 * (size_t)f->cap);
        f->pl = (bam_pileup1_t*)realloc(f->pl, sizeof(bam_pileup1_t) * f->cap);
    }
    for (i = 0; i < n; ++i) {
        bam1_t *b = f->b + i;
        memset(b, 0, sizeof(*b));
        b->core.l_qseq = 1; b->core.qual = mq[i];
        b->core.flag = (bs[i] & 0x10) ? BAM_FREVERSE : 0;
        b->data = f->data + 2 * (size_t)i;
        b->data[0] = (uint8_t)((bs[i] & 0xf) << 4); b->data[1] = bq[i];
        b->data_len = b->m_data = 2;
        memset(&f->pl[i], 0, sizeof(bam_pileup1_t));
        f->pl[i].b = b; f->pl[i].qpos = 0;
    }
    return f->pl;
}

/* out[0..9]=lk, [10]=min_lk, [11]=rms mapQ (glf max_mapQ), [12]=depth, [13]=cns word */
void ref_glfgen(int n, const uint8_t *bs, const uint8_t *bq, const uint8_t *mq, int ref_nt16, uint32_t *out)
{
    const bam_pileup1_t *pl = fabricate(&g_pu[0], n, bs, bq, mq);
    glf1_t *g = sniper_maqcns_glfgen(n, pl, (uint8_t)ref_nt16, g_d->c);
    int i;
    for (i = 0; i < 10; ++i) out[i] = g->lk[i];
    out[10] = g->min_lk; out[11] = g->max_mapQ; out[12] = g->depth;
    out[13] = sniper_maqcns_call(n, g, g_d->c);
    free(g);
}

/* Flattened sniper_output_t (output_format.h:13-35): 2 + 2*26 ints.
 * rec[0]=ref_base char, rec[1]=ref_base4, then per sample (tumor, normal):
 * genotype, joint_genotype, joint_consensus_quality, consensus_quality, variant_allele_quality,
 * somatic_score, variant_status, mean_baseQ[4], mean_mapQ[4], base_occ[4], dp4[4], total_depth, total_mean_mapQ */
#define REC_INTS 54
static void flatten_sample(const sample_data_t *s, int32_t *o)
{
    int i;
    o[0] = s->genotype; o[1] = s->joint_genotype; o[2] = s->joint_consensus_quality;
    o[3] = s->consensus_quality; o[4] = s->variant_allele_quality; o[5] = s->somatic_score;
    o[6] = (int)s->variant_status;
    for (i = 0; i < 4; ++i) {
        o[7 + i] = s->dqstats.mean_baseQ[i]; o[11 + i] = s->dqstats.mean_mapQ[i];
        o[15 + i] = s->dqstats.base_occ[i]; o[19 + i] = s->dqstats.dp4[i];
    }
    o[23] = s->dqstats.total_depth; o[24] = s->dqstats.total_mean_mapQ;
    o[25] = 0;
}

/* Runs glf_somatic (somatic_sniper.c:109-273) on one fabricated column.
 * Returns glf_somatic's return value; *emitted says whether a record was written. */
int ref_score_column(int ref_char, int nT, const uint8_t *bsT, const uint8_t *bqT, const uint8_t *mqT,
                     int nN, const uint8_t *bsN, const uint8_t *bqN, const uint8_t *mqN,
                     int *emitted, int32_t *rec)
{
    const bam_pileup1_t *pl1 = fabricate(&g_pu[0], nT, bsT, bqT, mqT);
    const bam_pileup1_t *pl2 = fabricate(&g_pu[1], nN, bsN, bqN, mqN);
    int ret;
    g_d->ref[0] = (char)ref_char; g_d->len = 1; g_d->tid = 0;
    g_emitted = 0;
    ret = glf_somatic(0, 0, nT, nN, pl1, pl2, g_d, g_null);
    *emitted = g_emitted;
    if (g_emitted && rec) {
        rec[0] = g_last.ref_base; rec[1] = g_last.ref_base4;
        flatten_sample(&g_last.tumor, rec + 2);
        flatten_sample(&g_last.normal, rec + 28);
    }
    return ret;
}

/* Batch form used for timing the reference's own per-column path on host cores
 * (bench.py --impl reference) and for bulk parity: columns in the product's SoA layout
 * (off[2c]..off[2c+1] tumor, off[2c+1]..off[2c+2] normal). Returns #emitted; fills
 * score[c] with glf_somatic's return value and, for emitted columns, appends records. */
int ref_score_batch(int ncol, const uint32_t *off, const uint8_t *bs, const uint8_t *bq, const uint8_t *mq,
                    const uint8_t *ref_char, int32_t *score, int32_t *emit_col, int32_t *recs, int max_recs)
{
    int c, n_emit = 0;
    for (c = 0; c < ncol; ++c) {
        uint32_t t0 = off[2 * c], n0 = off[2 * c + 1], e = off[2 * c + 2];
        int emitted = 0;
        int32_t rec[REC_INTS];
        int r = ref_score_column(ref_char[c], (int)(n0 - t0), bs + t0, bq + t0, mq + t0,
                                 (int)(e - n0), bs + n0, bq + n0, mq + n0, &emitted, rec);
        if (score) score[c] = r;
        if (emitted) {
            if (n_emit < max_recs) {
                if (emit_col) emit_col[n_emit] = c;
                if (recs) memcpy(recs + (size_t)n_emit * REC_INTS, rec, sizeof(rec));
            }
            ++n_emit;
        }
    }
    return n_emit;
}

/* dqstats.c:6-53 on a fabricated column: out = mean_baseQ[4], mean_mapQ[4], base_occ[4], dp4[4], depth, mean mapQ */
void ref_dqstats(int n, const uint8_t *bs, const uint8_t *bq, const uint8_t *mq, int ref_base4, uint32_t wanted, uint32_t *out)
{
    const bam_pileup1_t *pl = fabricate(&g_pu[0], n, bs, bq, mq);
    dqstats_t s;
    int i;
    get_dqstats(pl, n, ref_base4, wanted, &s);
    for (i = 0; i < 4; ++i) { out[i] = s.mean_baseQ[i]; out[4 + i] = s.mean_mapQ[i]; out[8 + i] = s.base_occ[i]; out[12 + i] = s.dp4[i]; }
    out[16] = s.total_depth; out[17] = s.total_mean_mapQ;
}

int ref_should_filter_as_loh(int r, int t, int n) { return should_filter_as_loh(r, t, n); }
int ref_should_filter_as_gor(int r, int t, int n) { return should_filter_as_gor(r, t, n); }
int ref_count_alleles(int a) { return count_alleles(a); }
int ref_nt16_of_char(int ch) { return bam_nt16_table[ch & 0xff]; }

/* ---- the reference's own dual-BAM walk with a capturing callback ------------------------------------------
 * Runs bam_sspileup_file (sniper_pileup.c:226-266) over two BAM files and records every column the walker
 * hands to its callback: (tid, pos, n1, n2) plus the non-deleted entries of both samples in the product's
 * entry encoding.  Used to pin the host packer (tests/test_packer.py). */
typedef struct {
    int32_t *cols; size_t n_cols, m_cols;            /* 6 ints per column: tid, pos, n1, n2, entries tumor, entries normal */
    uint8_t *ent; size_t n_ent, m_ent;               /* 3 bytes per entry: bs, bq, mq (tumor entries first) */
} pileup_dump_t;
static pileup_dump_t g_dump;

static void dump_entries(const bam_pileup1_t *pl, int n, int32_t *count)
{
    int i;
    *count = 0;
    for (i = 0; i < n; ++i) {
        const bam_pileup1_t *p = pl + i;
        if (p->is_del || (p->b->core.flag & BAM_FUNMAP)) continue;
        if (g_dump.n_ent + 3 > g_dump.m_ent) {
            g_dump.m_ent = g_dump.m_ent ? g_dump.m_ent * 2 : (1u << 20);
            g_dump.ent = (uint8_t*)realloc(g_dump.ent, g_dump.m_ent);
        }
        g_dump.ent[g_dump.n_ent++] = (uint8_t)(bam1_seqi(bam1_seq(p->b), p->qpos) | (bam1_strand(p->b) ? 16 : 0));
        g_dump.ent[g_dump.n_ent++] = bam1_qual(p->b)[p->qpos];
        g_dump.ent[g_dump.n_ent++] = (uint8_t)p->b->core.qual;
        ++*count;
    }
}

static int dump_cb(uint32_t tid, uint32_t pos, int n1, int n2, const bam_pileup1_t *pl1, const bam_pileup1_t *pl2, void *data, FILE *fh)
{
    int32_t *c;
    (void)data; (void)fh;
    if (g_dump.n_cols + 6 > g_dump.m_cols) {
        g_dump.m_cols = g_dump.m_cols ? g_dump.m_cols * 2 : (1u << 16);
        g_dump.cols = (int32_t*)realloc(g_dump.cols, sizeof(int32_t) * g_dump.m_cols);
    }
    c = g_dump.cols + g_dump.n_cols;
    g_dump.n_cols += 6;
    c[0] = (int32_t)tid; c[1] = (int32_t)pos; c[2] = n1; c[3] = n2;
    dump_entries(pl1, n1, &c[4]);
    c = g_dump.cols + g_dump.n_cols - 6;              /* realloc inside dump_entries does not move cols, but stay safe */
    dump_entries(pl2, n2, &c[5]);
    return 0;
}

/* returns the number of columns; pointers stay valid until the next call */
long ref_pileup_dump(const char *tumor_bam, const char *normal_bam, int mapq, const int32_t **cols, const uint8_t **ent, long *n_ent)
{
    bamFile fp1 = bam_open(tumor_bam, "r"), fp2 = bam_open(normal_bam, "r");
    bam_header_t *h1, *h2;
    if (!fp1 || !fp2) return -1;
    h1 = bam_header_read(fp1); h2 = bam_header_read(fp2);
    g_dump.n_cols = 0; g_dump.n_ent = 0;
    bam_sspileup_file(fp1, fp2, BAM_DEF_MASK, mapq, dump_cb, 0, 0);
    bam_close(fp1); bam_close(fp2);
    bam_header_destroy(h1); bam_header_destroy(h2);
    *cols = g_dump.cols; *ent = g_dump.ent; *n_ent = (long)(g_dump.n_ent / 3);
    return (long)(g_dump.n_cols / 6);
}
