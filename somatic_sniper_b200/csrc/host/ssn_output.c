/*
 * ssn_output.c — classic / vcf / bed line writers over ssn_site records.  The output grammar is the
 * reference's (byte-identical lines): src/lib/sniper/output_classic.c:9-55, output_vcf.c:33-192,
 * output_bed.c:5-19, the shared print helpers of dqstats.c:55-83 and the format registry of
 * output_format.c:9-17.  Lines are buffered by stdio; the reference's per-site fflush is not needed for
 * identical bytes.
 */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include "ssn_host.h"
#include "ssn_nt16.h"

typedef void (*site_fn)(FILE *fh, const char *name, uint32_t pos, int ref_char, const ssn_site *s);
typedef void (*header_fn)(FILE *fh, const char *refseq, const char *normal_id, const char *tumor_id);

struct ssn_writer {
    FILE *fh;
    site_fn site;
    header_fn header;
};

/* values of the bases selected by the 4-bit mask, comma separated; "0" when none (dqstats.c:55-83) */
static void put_masked_or_zero(FILE *fh, int bases, const uint32_t v[4])
{
    int i, need_comma = 0;
    for (i = 0; i < 4; ++i) {
        if (bases & (1 << i)) {
            if (need_comma) fputc(',', fh);
            fprintf(fh, "%d", (int)v[i]);
            need_comma = 1;
        }
    }
    if (!need_comma) fputc('0', fh);
}

/* ---- classic: 26 tab-separated columns ----------------------------------------------------------------------- */

static void classic_header(FILE *fh, const char *refseq, const char *normal_id, const char *tumor_id)
{
    (void)fh; (void)refseq; (void)normal_id; (void)tumor_id;
}

static void classic_sample_tail(FILE *fh, int ref4, const ssn_sample *s, int last)
{
    const int alt = ~ref4 & s->genotype;
    put_masked_or_zero(fh, ref4, s->mean_baseQ); fputc('\t', fh);
    put_masked_or_zero(fh, ref4, s->mean_mapQ); fputc('\t', fh);
    put_masked_or_zero(fh, ref4, s->base_occ); fputc('\t', fh);
    put_masked_or_zero(fh, alt, s->mean_baseQ); fputc('\t', fh);
    put_masked_or_zero(fh, alt, s->mean_mapQ); fputc('\t', fh);
    put_masked_or_zero(fh, alt, s->base_occ); fputc(last ? '\n' : '\t', fh);
}

static void classic_site(FILE *fh, const char *name, uint32_t pos, int ref_char, const ssn_site *p)
{
    fprintf(fh, "%s\t%d\t%c\t%c\t%c\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t%d\t", name, (int)(pos + 1), ref_char,
            ssn_nt16_rev[p->tumor.genotype & 15], ssn_nt16_rev[p->normal.genotype & 15],
            p->tumor.somatic_score, p->tumor.consensus_quality, p->tumor.variant_allele_quality,
            (int)p->tumor.total_mean_mapQ, p->normal.consensus_quality, p->normal.variant_allele_quality,
            (int)p->normal.total_mean_mapQ, (int)p->tumor.total_depth, (int)p->normal.total_depth);
    classic_sample_tail(fh, p->ref_base4, &p->tumor, 0);
    classic_sample_tail(fh, p->ref_base4, &p->normal, 1);
}

/* ---- bed ---------------------------------------------------------------------------------------------------------- */

static void bed_header(FILE *fh, const char *refseq, const char *normal_id, const char *tumor_id)
{
    (void)refseq; (void)normal_id; (void)tumor_id;
    fprintf(fh, "#CHROM\tSTART\tSTOP\tREF/ALT\tSOMATIC_SCORE\tTUMOR_DEPTH\n");
}

static void bed_site(FILE *fh, const char *name, uint32_t pos, int ref_char, const ssn_site *p)
{
    fprintf(fh, "%s\t%d\t%d\t%c/%c\t%d\t%d\n", name, (int)pos, (int)(pos + 1), ref_char,
            ssn_nt16_rev[p->tumor.genotype & 15], p->tumor.somatic_score, (int)p->tumor.total_depth);
}

/* ---- vcf ---------------------------------------------------------------------------------------------------------- */

static const struct { const char *id, *number, *type, *description; } vcf_fields[] = {
    { "GT", "1", "String", "Genotype" },
    { "IGT", "1", "String", "Genotype when called independently (only filled if called in joint prior mode)" },
    { "DP", "1", "Integer", "Total read depth" },
    { "DP4", "4", "Integer", "# high-quality ref-forward bases, ref-reverse, alt-forward and alt-reverse bases" },
    { "BCOUNT", "4", "Integer", "Occurrence count for each base at this site (A,C,G,T)" },
    { "GQ", "1", "Integer", "Genotype quality" },
    { "JGQ", "1", "Integer", "Joint genotype quality (only filled if called in join prior mode)" },
    { "VAQ", "1", "Integer", "Variant allele quality" },
    { "BQ", ".", "Integer", "Average base quality" },
    { "MQ", "1", "Integer", "Average mapping quality across all reads" },
    { "AMQ", ".", "Integer", "Average mapping quality for each allele present in the genotype" },
    { "SS", "1", "Integer", "Variant status relative to non-adjacent Normal, 0=wildtype,1=germline,2=somatic,3=LOH,4=unknown" },
    { "SSC", "1", "Integer", "Somatic Score" },
};

static void vcf_header(FILE *fh, const char *refseq, const char *normal_id, const char *tumor_id)
{
    char filedate[255];
    time_t now = time(NULL);
    size_t i;
    strftime(filedate, sizeof(filedate), "%Y%m%d", localtime(&now));
    fprintf(fh, "##fileformat=VCFv4.1\n");
    fprintf(fh, "##fileDate=%s\n", filedate);
    fprintf(fh, "##phasing=none\n");
    fprintf(fh, "##reference=file://%s\n", refseq);
    for (i = 0; i < sizeof(vcf_fields) / sizeof(vcf_fields[0]); ++i)
        fprintf(fh, "##FORMAT=<ID=%s,Number=%s,Type=%s,Description=\"%s\">\n", vcf_fields[i].id, vcf_fields[i].number,
                vcf_fields[i].type, vcf_fields[i].description);
    fprintf(fh, "#CHROM\tPOS\tID\tREF\tALT\tQUAL\tFILTER\tINFO\tFORMAT\t%s\t%s\n", normal_id, tumor_id);
}

/* genotype as allele indices: 0 = reference, alternates numbered in A,C,G,T order over `alts` (output_vcf.c:46-79) */
static void vcf_gt(FILE *fh, uint32_t ref4, uint32_t alts, uint32_t gt)
{
    const int n_alleles = __builtin_popcount(gt & 15u);       /* count_alleles, allele_util.c:6-8 */
    uint32_t idx = 0, printed = 0, i;
    if (gt & ref4) {
        if (n_alleles == 1) { fputs("0/0", fh); return; }
        fputc('0', fh);
        ++printed;
    }
    gt &= ~ref4;
    for (i = 0; i < 4; ++i) {
        const uint32_t bit = 1u << i;
        if (alts & bit) ++idx;
        if (gt & bit) {
            if (n_alleles == 1) { fprintf(fh, "%d/%d", (int)idx, (int)idx); return; }
            if (printed > 0) fputc('/', fh);
            fprintf(fh, "%d", (int)idx);
            ++printed;
        }
    }
}

/* values of the bases in the mask, comma separated; nothing when the mask is empty (output_vcf.c:33-44) */
static void put_masked(FILE *fh, const uint32_t v[4], uint32_t mask)
{
    uint32_t i, c = 0;
    for (i = 0; i < 4; ++i)
        if (mask & (1u << i)) {
            if (++c > 1) fputc(',', fh);
            fprintf(fh, "%d", (int)v[i]);
        }
}

static void vcf_sample(FILE *fh, int ref4, uint32_t alts, const ssn_sample *s)
{
    vcf_gt(fh, (uint32_t)ref4, alts, (uint32_t)(s->joint_genotype ? s->joint_genotype : s->genotype));
    fputc(':', fh);
    vcf_gt(fh, (uint32_t)ref4, alts, (uint32_t)s->genotype);
    fprintf(fh, ":%d:%d,%d,%d,%d:%d,%d,%d,%d:%d:", (int)s->total_depth, (int)s->dp4[0], (int)s->dp4[1], (int)s->dp4[2], (int)s->dp4[3],
            (int)s->base_occ[0], (int)s->base_occ[1], (int)s->base_occ[2], (int)s->base_occ[3], s->consensus_quality);
    if (s->joint_genotype) fprintf(fh, "%d:", s->joint_consensus_quality);
    else fputs(".:", fh);
    fprintf(fh, "%d:", s->variant_allele_quality);
    put_masked(fh, s->mean_baseQ, (uint32_t)s->genotype);         /* masked by the independent genotype even with -J (QUIRK B8) */
    fputc(':', fh);
    fprintf(fh, "%d:", (int)s->total_mean_mapQ);
    put_masked(fh, s->mean_mapQ, (uint32_t)s->genotype);
    fputc(':', fh);
    fprintf(fh, "%d:", s->variant_status);
    if (s->somatic_score >= 0) fprintf(fh, "%d", s->somatic_score);
    else fputc('.', fh);
}

static void vcf_site(FILE *fh, const char *name, uint32_t pos, int ref_char, const ssn_site *p)
{
    const uint32_t alts = (uint32_t)(p->tumor.genotype | p->normal.genotype) & ~(uint32_t)p->ref_base4;
    uint32_t n_alt = 0, i;
    fprintf(fh, "%s\t%d\t.\t%c\t", name, (int)(pos + 1), ref_char);
    for (i = 0; i < 4; ++i)
        if (alts & (1u << i)) {
            if (++n_alt > 1) fputc(',', fh);
            fputc(ssn_nt16_rev[1u << i], fh);
        }
    if (n_alt == 0) fputc('.', fh);
    fputs("\t.\t.\t.\tGT:IGT:DP:DP4:BCOUNT:GQ:JGQ:VAQ:BQ:MQ:AMQ:SS:SSC\t", fh);
    vcf_sample(fh, p->ref_base4, alts, &p->normal);
    fputc('\t', fh);
    vcf_sample(fh, p->ref_base4, alts, &p->tumor);
    fputc('\n', fh);
}

/* ---- registry ------------------------------------------------------------------------------------------------------ */

static const struct { const char *name; header_fn header; site_fn site; } formats[] = {
    { "classic", classic_header, classic_site },
    { "vcf", vcf_header, vcf_site },
    { "bed", bed_header, bed_site },
};

int ssn_writer_n_formats(void) { return (int)(sizeof(formats) / sizeof(formats[0])); }
const char *ssn_writer_format_name(int i) { return (i >= 0 && i < ssn_writer_n_formats()) ? formats[i].name : 0; }

ssn_writer *ssn_writer_create(const char *format, void *fh)
{
    int i;
    for (i = 0; i < ssn_writer_n_formats(); ++i)
        if (strcmp(format, formats[i].name) == 0) {
            ssn_writer *w = (ssn_writer*)calloc(1, sizeof(*w));
            w->fh = (FILE*)fh; w->header = formats[i].header; w->site = formats[i].site;
            return w;
        }
    return 0;
}

void ssn_writer_header(ssn_writer *w, const char *refseq, const char *normal_id, const char *tumor_id) { w->header(w->fh, refseq, normal_id, tumor_id); }
void ssn_writer_site(ssn_writer *w, const char *seq_name, uint32_t pos, int ref_char, const ssn_site *site) { w->site(w->fh, seq_name, pos, ref_char, site); }
void ssn_writer_destroy(ssn_writer *w) { free(w); }
