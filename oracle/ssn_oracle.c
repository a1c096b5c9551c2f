/*
 * ssn_oracle.c — CPU ORACLE. TEST INFRASTRUCTURE ONLY.
 *
 * A plain-C restatement of the reference's per-column scoring path, used solely as the
 * checker in tests/, __graft_entry__.smoke() and bench.py's cpu_baseline leg.  The product
 * (somatic_sniper_b200/csrc) never includes, links or calls anything in this directory.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function below against
 *   (a) the reference's own expected.vcf sites (integration-test/data/expected.vcf:19-21),
 *   (b) the known-answer columns KAT0..KAT3 of SURVEY.md App. C.3, and
 *   (c) the unmodified reference compiled from /root/reference (oracle/_ref/libref_harness.so)
 *       on seeded random columns and on the bit patterns of every model table,
 *   with golden vectors committed under tests/golden/ for the GPU box (no /root/reference there).
 *
 * Entry encoding (shared with the product's SoA batches):
 *     bs[i] = nt16 | strand<<4,   bq[i] = base quality,   mq[i] = mapping quality.
 * Deleted bases and unmapped reads never reach this layer (the packer drops them, as
 * sniper_maqcns.c:147 and dqstats.c:20 do).
 *
 * Each function names the reference lines it restates (paths relative to /root/reference).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#define ORC_PHRED 4.343 /* PHRED_CONST, src/lib/sniper/somatic_sniper.h:12 */

typedef struct {
    /* parameters: sniper_maqcns_t (sniper_maqcns.h:13-21), pu_data2_t (somatic_sniper.h:21-40) */
    float theta, het_rate, eta;
    int n_hap, cap_mapq;
    int use_priors, use_joint, min_somatic_qual, include_loh, include_gor;
    double somatic_rate;
    /* tables */
    double *coef;   /* [q<<16 | n<<8 | k], 64*256*256 */
    double *lhet;   /* [n1<<8 | n2], 256*256 */
    double fk[256];
    float q_r;
    int prior[16][10];
    int jointprior[16][10][10];
    int qadd[1024 + 64]; /* 1024 as in the reference + zero slack so a 1023+ index stays defined */
} orc_model;

typedef struct {
    uint8_t lk[10];
    uint8_t min_lk, rms_mapq;
    uint32_t depth;
    uint32_t cns;
    float esum[4], fsum[4];
    uint32_t c[4];
    float p[16];
} orc_glf;

/* flattened sniper_output_t, same layout as oracle/ref_harness.c REC_INTS */
#define ORC_REC_INTS 54

static const int orc_glf_base[10] = {1, 3, 5, 9, 2, 6, 10, 4, 12, 8};          /* somatic_sniper.c:26 */
static const int orc_is_hom[16] = {0,1,1,0,1,0,0,0,1,0,0,0,0,0,0,0};           /* somatic_sniper.c:24 */
static const int orc_is_het[16] = {0,0,0,1,0,1,1,0,0,1,1,0,1,0,0,0};           /* somatic_sniper.c:25 */
static const int orc_nt16_nt4[16] = {4,0,1,4,2,4,4,4,3,4,4,4,4,4,4,4};         /* sniper_maqcns.c:19 */

/* logPhred / expPhred, somatic_sniper.h:13-14 (C truncation toward zero) */
static int orc_log_phred(double x)
{
    return (int)(x < 1 ? (0.5 - ORC_PHRED * log(x)) : (-0.5 - ORC_PHRED * log(x)));
}
static double orc_exp_phred(int x) { return exp((double)(-x) / ORC_PHRED); }

/* ---- tables ------------------------------------------------------------------------------ */

/* sniper_cal_coef, sniper_maqcns.c:59-100. x87 long double where the reference uses it. */
static void orc_build_coef(orc_model *m)
{
    long double suffix[257], beta[256], qc[256], fk2[256];
    double *lchoose = (double*)calloc(256 * 256, sizeof(double));
    int n, k, q;
    m->coef = (double*)calloc((size_t)64 * 256 * 256, sizeof(double));
    m->fk[0] = 1.0; fk2[0] = 1.0;
    for (n = 1; n < 256; ++n) {
        m->fk[n] = pow(m->theta, n) * (1.0 - m->eta) + m->eta;  /* float fields promote to double */
        fk2[n] = m->fk[n >> 1];
    }
    for (n = 1; n < 256; ++n)
        for (k = 1; k <= n; ++k)
            lchoose[n << 8 | k] = lgamma(n + 1) - lgamma(k + 1) - lgamma(n - k + 1);
    for (q = 1; q < 64; ++q) {
        double e = pow(10.0, -q / 10.0), le = log(e), le1 = log(1.0 - e);
        for (n = 1; n < 256; ++n) {
            double *row = m->coef + (q << 16 | n << 8);
            suffix[n + 1] = 0.0;
            for (k = n; k >= 0; --k) {
                suffix[k] = suffix[k + 1] + expl(lchoose[n << 8 | k] + k * le + (n - k) * le1);
                beta[k] = suffix[k + 1] / suffix[k];
                if (beta[k] > 0.99) beta[k] = 0.99;
            }
            for (k = 0; k < n; ++k) qc[k] = -4.343 * fk2[k] * logl(beta[k] / e);
            for (k = 1; k < n; ++k) qc[k] += qc[k - 1];
            for (k = 0; k <= n; ++k) {
                long double t = -4.343 * logl(1.0 - expl(fk2[k] * logl(beta[k])));
                row[k] = (k ? qc[k - 1] : 0) + t;
            }
        }
    }
    free(lchoose);
}

/* sniper_cal_het, sniper_maqcns.c:27-56 */
static void orc_build_het(orc_model *m)
{
    int k, n1, n2;
    double harm = 0.0, poly;
    m->lhet = (double*)calloc(256 * 256, sizeof(double));
    for (k = 1; k <= m->n_hap - 1; ++k) harm += 1.0 / k;
    for (n1 = 0; n1 < 256; ++n1)
        for (n2 = 0; n2 < 256; ++n2) {
            long double s = 0.0;
            double lc = lgamma(n1 + n2 + 1) - lgamma(n1 + 1) - lgamma(n2 + 1);
            for (k = 1; k <= m->n_hap - 1; ++k) {
                double pk = 1.0 / k / harm;
                double l1 = log((double)k / m->n_hap);
                double l2 = log(1.0 - (double)k / m->n_hap);
                s += pk * 0.5 * (expl(l1 * n2) * expl(l2 * n1) + expl(l1 * n1) * expl(l2 * n2));
            }
            m->lhet[n1 << 8 | n2] = lc + logl(s);
        }
    poly = m->het_rate * harm;
    m->q_r = -4.343 * log(2.0 * poly / (1.0 - poly));
}

/* makeSoloPrior / make_joint_prior, somatic_sniper.c:29-77 (THETA = 0.001, :14) */
static int orc_germline_prior(int ref, int b)
{
    const double T = 0.001;
    if (!(b & ~ref)) return 0;
    if (b & ref) return orc_log_phred(T);
    if (orc_is_hom[b]) return orc_log_phred(0.5 * T);
    return orc_log_phred(T * T);
}
static void orc_build_priors(orc_model *m)
{
    int ref, i, j;
    memset(m->prior, 0, sizeof(m->prior));
    memset(m->jointprior, 0, sizeof(m->jointprior));
    if (m->use_priors)
        for (ref = 0; ref < 16; ++ref)
            for (i = 0; i < 10; ++i) m->prior[ref][i] = orc_germline_prior(ref, orc_glf_base[i]);
    if (m->use_joint)
        for (ref = 0; ref < 16; ++ref)
            for (i = 0; i < 10; ++i) {
                int b = orc_glf_base[i], g = orc_germline_prior(ref, b);
                for (j = 0; j < 10; ++j) {
                    int c = orc_glf_base[j];
                    /* the reference indexes isHet/isHom with the genotype INDEX j (:66-72) */
                    if (b == c) m->jointprior[ref][i][j] = g;
                    else if ((b & c) && (orc_is_het[j] || orc_is_hom[j]))
                        m->jointprior[ref][i][j] = g + orc_log_phred(m->somatic_rate);
                    else
                        m->jointprior[ref][i][j] = g + orc_log_phred(m->somatic_rate * m->somatic_rate);
                }
            }
}

/* qAddTableInit, somatic_sniper.c:101-107: entries 1000.. stay zero */
static void orc_build_qadd(orc_model *m)
{
    int i;
    memset(m->qadd, 0, sizeof(m->qadd));
    for (i = 0; i < 1000; ++i) m->qadd[i] = orc_log_phred(1 + orc_exp_phred(i - 512));
}

orc_model *orc_model_new(float theta, float het_rate, int n_hap, int use_priors, int use_joint,
                         double somatic_rate, int min_somatic_qual, int include_loh, int include_gor)
{
    orc_model *m = (orc_model*)calloc(1, sizeof(orc_model));
    m->theta = theta; m->het_rate = het_rate; m->n_hap = n_hap;
    m->eta = 0.03; m->cap_mapq = 60;              /* sniper_maqcns_init, sniper_maqcns.c:110-111 */
    m->use_priors = use_priors; m->use_joint = use_joint; m->somatic_rate = somatic_rate;
    m->min_somatic_qual = min_somatic_qual; m->include_loh = include_loh; m->include_gor = include_gor;
    orc_build_coef(m); orc_build_het(m); orc_build_priors(m); orc_build_qadd(m);
    return m;
}
void orc_model_free(orc_model *m)
{
    if (!m) return;
    free(m->coef); free(m->lhet); free(m);
}
const double *orc_coef(const orc_model *m) { return m->coef; }
const double *orc_lhet(const orc_model *m) { return m->lhet; }
const double *orc_fk(const orc_model *m) { return m->fk; }
float orc_q_r(const orc_model *m) { return m->q_r; }
const int *orc_prior(const orc_model *m) { return &m->prior[0][0]; }
const int *orc_jointprior(const orc_model *m) { return &m->jointprior[0][0][0]; }
const int *orc_qadd(const orc_model *m) { return m->qadd; }

/* ---- per-sample genotype likelihoods ------------------------------------------------------------ */

static int orc_cmp_u32(const void *a, const void *b)
{
    uint32_t x = *(const uint32_t*)a, y = *(const uint32_t*)b;
    return x < y ? -1 : x > y;
}

/* sniper_maqcns_glfgen, sniper_maqcns.c:127-248; the reference's introsort (:157) is any
 * ascending sort of the 32-bit words: equal words are indistinguishable downstream. */
void orc_glfgen(const orc_model *m, int n, const uint8_t *bs, const uint8_t *bq, const uint8_t *mq,
                int ref_nt16, orc_glf *g)
{
    uint32_t stackbuf[512], *key = stackbuf;
    int w[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    float *p = g->p, min_p = 1e30f;
    uint64_t rms = 0;
    int i, j, k, tot;
    memset(g, 0, sizeof(*g));
    if (n == 0) { g->cns = 0xfU << 28 | 0xfU << 24; return; }   /* sniper_maqcns_call :281 */
    if (n > 512) key = (uint32_t*)malloc(sizeof(uint32_t) * (size_t)n);
    /* :144-156 the sort word */
    for (i = 0; i < n; ++i) {
        uint32_t q = bq[i], x = (uint32_t)((bs[i] >> 4) & 1) << 18 | q << 8 | mq[i];
        int nt = bs[i] & 0xf, b4;
        if (mq[i] < q) q = mq[i];
        x |= q << 24;
        b4 = orc_nt16_nt4[nt ? nt : ref_nt16];
        if (b4 < 4) x |= 1u << 21 | (uint32_t)b4 << 16;
        key[i] = x;
    }
    qsort(key, (size_t)n, sizeof(uint32_t), orc_cmp_u32);
    /* :160-176 error-dependency accumulation, highest word first */
    for (j = n - 1; j >= 0; --j) {
        uint32_t x = key[j];
        int cap;
        if (x >> 24 < 4 && (x >> 8 & 0x3f) != 0) x = 4u << 24 | (x & 0xffffff);
        k = x >> 16 & 7;
        if (x >> 24 > 0) {
            g->esum[k & 3] += m->fk[w[k]] * (x >> 24);   /* float += double*uint */
            g->fsum[k & 3] += m->fk[w[k]];
            if (w[k] < 0xff) ++w[k];
            ++g->c[k & 3];
        }
        cap = (int)(x & 0x7f) < m->cap_mapq ? (int)(x & 0x7f) : m->cap_mapq;
        rms += (uint64_t)(cap * cap);
    }
    g->rms_mapq = (uint8_t)(sqrt((double)rms / n) + .499);
    /* :178-182 count rescale */
    for (j = tot = 0; j < 4; ++j) tot += g->c[j];
    if (tot > 255) {
        for (j = 0; j < 4; ++j) g->c[j] = (int)(254.0 * g->c[j] / tot + 0.5);
        for (j = tot = 0; j < 4; ++j) tot += g->c[j];
    }
    /* :184-214 the 4 homozygous + 6 heterozygous terms */
    for (j = 0; j < 4; ++j) {
        float t1, t3; int t2, bar;
        for (k = 0, t1 = t3 = 0.0f, t2 = 0; k < 4; ++k) {
            if (k == j) continue;
            t1 += g->esum[k]; t2 += g->c[k]; t3 += g->fsum[k];
        }
        if (t2) {
            bar = (int)(t1 / t3 + 0.5);
            if (bar < 4) bar = 4;
            if (bar > 63) bar = 63;
            p[j << 2 | j] = t1 + m->coef[bar << 16 | tot << 8 | t2];
        } else p[j << 2 | j] = 0.0f;
        for (k = j + 1; k < 4; ++k) {
            for (i = 0, t2 = 0, t1 = t3 = 0.0f; i < 4; ++i) {
                if (i == j || i == k) continue;
                t1 += g->esum[i]; t2 += g->c[i]; t3 += g->fsum[i];
            }
            if (t2) {
                bar = (int)(t1 / t3 + 0.5);
                if (bar < 4) bar = 4;
                if (bar > 63) bar = 63;
                p[j << 2 | k] = p[k << 2 | j] =
                    -4.343 * m->lhet[g->c[j] << 8 | g->c[k]] + t1 + m->coef[bar << 16 | tot << 8 | t2];
            } else p[j << 2 | k] = p[k << 2 | j] = -4.343 * m->lhet[g->c[j] << 8 | g->c[k]];
        }
        for (k = 0; k < 4; ++k)
            if (p[j << 2 | k] < 0.0) p[j << 2 | k] = 0.0;
    }
    /* :216-233 the p[k,k] correction */
    {
        float max1 = -1.0f, max2 = -1.0f, min1 = 1e30f, min2 = 1e30f;
        int max_k = -1, min_k = -1;
        for (k = 0; k < 4; ++k) {
            if (g->esum[k] > max1) { max2 = max1; max1 = g->esum[k]; max_k = k; }
            else if (g->esum[k] > max2) max2 = g->esum[k];
        }
        for (k = 0; k < 4; ++k) {
            if (p[k << 2 | k] < min1) { min2 = min1; min1 = p[k << 2 | k]; min_k = k; }
            else if (p[k << 2 | k] < min2) min2 = p[k << 2 | k];
        }
        if (max1 > max2 && (min_k != max_k || min1 + 1.0 > min2))
            p[max_k << 2 | max_k] = min1 > 1.0 ? min1 - 1.0 : 0.0;
    }
    /* :235-244 quantisation */
    g->depth = n > 16777215 ? 16777215 : (uint32_t)n;
    for (j = 0; j < 4; ++j)
        for (k = j; k < 4; ++k)
            if (p[j << 2 | k] < min_p) min_p = p[j << 2 | k];
    g->min_lk = min_p > 255.0 ? 255 : (int)(min_p + 0.5);
    for (j = i = 0; j < 4; ++j)
        for (k = j; k < 4; ++k)
            g->lk[i++] = p[j << 2 | k] - min_p > 255.0 ? 255 : (int)(p[j << 2 | k] - min_p + 0.5);
    /* sniper_glf2cns, sniper_maqcns.c:250-273, with q_r rounded as in sniper_maqcns_call :279 */
    {
        int t[16], qr = (int)(m->q_r + 0.5), mn = 10000, mn2 = 10000, mn3 = 10000, g1 = -1, g2 = -1;
        uint32_t x;
        for (i = k = 0; i < 4; ++i)
            for (j = i; j < 4; ++j) { t[j << 2 | i] = -1; t[i << 2 | j] = g->lk[k++] + (i == j ? 0 : qr); }
        for (i = 0; i < 16; ++i) {
            if (t[i] < 0) continue;
            if (t[i] < mn) { mn3 = mn2; mn2 = mn; mn = t[i]; g2 = g1; g1 = i; }
            else if (t[i] < mn2) { mn3 = mn2; mn2 = t[i]; g2 = i; }
            else if (t[i] < mn3) mn3 = t[i];
        }
        x = g1 >= 0 ? (1U << (g1 >> 2 & 3) | 1U << (g1 & 3)) << 28 : 0xfU << 28;
        x |= g2 >= 0 ? (1U << (g2 >> 2 & 3) | 1U << (g2 & 3)) << 24 : 0xfU << 24;
        x |= (uint32_t)g->rms_mapq << 16;
        x |= mn2 < 10000 ? (uint32_t)(mn2 - mn < 256 ? mn2 - mn : 255) << 8 : 0xffU << 8;
        x |= mn2 < 10000 && mn3 < 10000 ? (uint32_t)(mn3 - mn2 < 256 ? mn3 - mn2 : 255) : 0xffU;
        g->cns = x;
    }
    if (key != stackbuf) free(key);
}

/* ---- site statistics ------------------------------------------------------------------------------ */

/* get_dqstats, dqstats.c:6-53. out: mean_baseQ[4], mean_mapQ[4], base_occ[4], dp4[4], depth, mean mapQ */
void orc_dqstats(int n, const uint8_t *bs, const uint8_t *bq, const uint8_t *mq, int ref_base4,
                 uint32_t wanted, uint32_t *out)
{
    int i, j;
    memset(out, 0, 18 * sizeof(uint32_t));
    for (i = 0; i < n; ++i) {
        int base = bs[i] & 0xf, strand = (bs[i] >> 4) & 1;
        ++out[16];
        out[17] += mq[i];
        if (base == ref_base4) ++out[12 + strand]; else ++out[14 + strand];
        for (j = 0; j < 4; ++j) {
            int v = 1 << j;
            if ((base & v) == base) {
                ++out[8 + j];
                if (v & wanted) { out[j] += bq[i]; out[4 + j] += mq[i]; }
            }
        }
    }
    for (j = 0; j < 4; ++j)
        if (out[8 + j] > 0) {
            out[j] = out[j] / (double)out[8 + j] + .499;
            out[4 + j] = out[4 + j] / (double)out[8 + j] + .499;
        }
    if (out[16] > 0) out[17] = out[17] / (double)out[16] + .499;
}

/* ---- the tumor/normal column scorer ------------------------------------------------------------ */

#define ORC_QADD(m, x, y) ((x) + (m)->qadd[512 + (y) - (x)]) /* qAdd, somatic_sniper.c:18 */

/* calculatePosteriors, somatic_sniper.c:79-99 */
static void orc_posteriors(const orc_model *m, const orc_glf *g, int ref_nt16, int *post)
{
    int qsum = 255, j;
    for (j = 0; j < 10; ++j) {
        int x = g->lk[j] + m->prior[ref_nt16][j];
        qsum = ORC_QADD(m, x, qsum);
        post[j] = x;
    }
    for (j = 0; j < 10; ++j) {
        post[j] -= qsum;
        if (post[j] > 255) post[j] = 255;
    }
}

/* glf_somatic, somatic_sniper.c:109-273. ref_nt16 = bam_nt16_table[ref char]; a reference 'N'
 * (nt16 15) yields no call either way (:127 and :156 agree). Returns the reference's return
 * value; *emitted / rec as in oracle/ref_harness.c. gT/gN (optional) receive the two GLFs. */
int orc_score_column(const orc_model *m, int ref_char, int ref_nt16,
                     int nT, const uint8_t *bsT, const uint8_t *bqT, const uint8_t *mqT,
                     int nN, const uint8_t *bsN, const uint8_t *bqN, const uint8_t *mqN,
                     int *emitted, int32_t *rec, orc_glf *gT_out, orc_glf *gN_out)
{
    orc_glf gT, gN;
    int score = 255, postT[10], postN[10];
    int rb4 = ref_nt16, tb1, tb2, ts1, ts2, nb1, nb2, ns1, ns2;
    int t_vaq = 0, n_vaq = 0, jT = 0, jN = 0, jgq = 255, tg, ng, i, j;
    *emitted = 0;
    orc_glfgen(m, nT, bsT, bqT, mqT, ref_nt16, &gT);
    orc_glfgen(m, nN, bsN, bqN, mqN, ref_nt16, &gN);
    if (gT_out) *gT_out = gT;
    if (gN_out) *gN_out = gN;
    if (ref_char == 'N' || gT.depth == 0 || gN.depth == 0) return -1;       /* :127, :272 */
    tb1 = gT.cns >> 28; tb2 = gT.cns >> 24 & 0xf; ts1 = gT.cns >> 8 & 0xff; ts2 = gT.cns & 0xff;
    nb1 = gN.cns >> 28; nb2 = gN.cns >> 24 & 0xf; ns1 = gN.cns >> 8 & 0xff; ns2 = gN.cns & 0xff;
    if (!(rb4 != 15 && tb1 != 15 && nb1 != 15 && tb1 != nb1)) return score;   /* :156 */
    t_vaq = tb2 == rb4 ? ts1 : ts1 + ts2;                                     /* :157-164 */
    if (t_vaq > 255) t_vaq = 255;
    if (nb1 != rb4) { n_vaq = nb2 == rb4 ? ns1 : ns1 + ns2; if (n_vaq > 255) n_vaq = 255; }
    orc_posteriors(m, &gT, ref_nt16, postT);
    orc_posteriors(m, &gN, ref_nt16, postN);
    if (m->use_joint) {                                                        /* :170-208 */
        int J[10][10], marg = 255, best = 1000, bi = -1, bj = -1;
        for (i = 0; i < 10; ++i)
            for (j = 0; j < 10; ++j) {
                int v = (int)gN.lk[i] + (int)gT.lk[j] + m->jointprior[rb4][i][j];
                if (v > 255) v = 255;
                J[i][j] = v;
                if (v < best) { best = v; bi = i; bj = j; }
                marg = ORC_QADD(m, marg, v);
            }
        for (j = 0; j < 10; ++j) {
            int l = J[j][j] - marg;
            score = ORC_QADD(m, score, l);
            /* the reference tests its outer loop index, which is 10 here: always true (:196) */
            if (j != bj) jgq = ORC_QADD(m, jgq, l);
        }
        if (jgq > 255) jgq = 255;
        jN = orc_glf_base[bi]; jT = orc_glf_base[bj];
    } else {
        for (j = 0; j < 10; ++j) { int y = postT[j] + postN[j]; score = ORC_QADD(m, score, y); }
    }
    tg = jT ? jT : tb1; ng = jN ? jN : nb1;                                     /* :216-223 */
    if (m->min_somatic_qual <= score
        && (m->include_loh || !(ng != tg && (tg & ng) == tg))                  /* allele_util.c:19-22 */
        && (m->include_gor || !(!(ng != rb4 && (rb4 & ng) == rb4) && (tg & ~ng) == rb4))) { /* :25-28 */
        int32_t *t = rec + 2, *nn = rec + 28;
        uint32_t st[18];
        *emitted = 1;
        memset(rec, 0, ORC_REC_INTS * sizeof(int32_t));
        rec[0] = ref_char; rec[1] = rb4;
        t[0] = tb1; t[1] = jT; t[2] = jgq; t[3] = ts1; t[4] = t_vaq; t[5] = score;
        if (tg == ng) t[6] = 1;
        else if (ng != tg && (tg & ng) == tg) t[6] = 3;
        else if (score > 0) t[6] = 2;
        else t[6] = 4;
        orc_dqstats(nT, bsT, bqT, mqT, rb4, (uint32_t)(rb4 | tg | ng), st);
        for (i = 0; i < 18; ++i) t[7 + i] = (int32_t)st[i];
        nn[0] = nb1; nn[1] = jN; nn[2] = jgq; nn[3] = ns1; nn[4] = n_vaq; nn[5] = -1;
        nn[6] = nb1 == rb4 ? 0 : 1;
        orc_dqstats(nN, bsN, bqN, mqN, rb4, (uint32_t)(rb4 | ng | tg), st);
        for (i = 0; i < 18; ++i) nn[7 + i] = (int32_t)st[i];
    }
    return score;
}

/* Batch form over the product's SoA layout: off[2c]..off[2c+1] tumor, off[2c+1]..off[2c+2] normal.
 * glf_out (optional): per column 2 x 14 uint32 = lk[10], min_lk, rms, depth, cns (tumor then normal). */
int orc_score_batch(const orc_model *m, int ncol, const uint32_t *off, const uint8_t *bs, const uint8_t *bq,
                    const uint8_t *mq, const uint8_t *ref_char, const uint8_t *ref_nt16, int32_t *score,
                    int32_t *emit_col, int32_t *recs, int max_recs, uint32_t *glf_out)
{
    int c, n_emit = 0, i;
    for (c = 0; c < ncol; ++c) {
        uint32_t t0 = off[2 * c], n0 = off[2 * c + 1], e = off[2 * c + 2];
        int emitted = 0;
        int32_t rec[ORC_REC_INTS];
        orc_glf gT, gN;
        int r = orc_score_column(m, ref_char[c], ref_nt16[c], (int)(n0 - t0), bs + t0, bq + t0, mq + t0,
                                 (int)(e - n0), bs + n0, bq + n0, mq + n0, &emitted, rec, &gT, &gN);
        if (score) score[c] = r;
        if (glf_out) {
            uint32_t *o = glf_out + (size_t)c * 28;
            for (i = 0; i < 10; ++i) { o[i] = gT.lk[i]; o[14 + i] = gN.lk[i]; }
            o[10] = gT.min_lk; o[11] = gT.rms_mapq; o[12] = gT.depth; o[13] = gT.cns;
            o[24] = gN.min_lk; o[25] = gN.rms_mapq; o[26] = gN.depth; o[27] = gN.cns;
        }
        if (emitted) {
            if (n_emit < max_recs) {
                if (emit_col) emit_col[n_emit] = c;
                if (recs) memcpy(recs + (size_t)n_emit * ORC_REC_INTS, rec, sizeof(rec));
            }
            ++n_emit;
        }
    }
    return n_emit;
}

/* single-sample entry for tests: out = lk[10], min_lk, rms, depth, cns ; fl (optional) = esum[4], fsum[4], p[16] */
void orc_glfgen_flat(const orc_model *m, int n, const uint8_t *bs, const uint8_t *bq, const uint8_t *mq,
                     int ref_nt16, uint32_t *out, float *fl)
{
    orc_glf g;
    int i;
    orc_glfgen(m, n, bs, bq, mq, ref_nt16, &g);
    for (i = 0; i < 10; ++i) out[i] = g.lk[i];
    out[10] = g.min_lk; out[11] = g.rms_mapq; out[12] = g.depth; out[13] = g.cns;
    if (fl) { memcpy(fl, g.esum, 16); memcpy(fl + 4, g.fsum, 16); memcpy(fl + 8, g.p, 64); }
}

/* bam_nt16_table, samtools-0.1.6 bam_import.c:23-40, rebuilt from its generating rule:
 * the 16 IUPAC letters of "=ACMGRSVTWYHKDBN" in either case, the digits 0-3 as A,C,G,T, all else 15. */
int orc_nt16_of_char(int ch)
{
    static const char rev[] = "=ACMGRSVTWYHKDBN";
    int i;
    ch &= 0xff;
    if (ch >= '0' && ch <= '3') return 1 << (ch - '0');
    for (i = 0; i < 16; ++i)
        if (ch == rev[i] || (rev[i] >= 'A' && rev[i] <= 'Z' && ch == rev[i] + 32)) return i;
    return 15;
}
