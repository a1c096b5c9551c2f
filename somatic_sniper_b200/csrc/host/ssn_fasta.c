#include "ssn_fasta.h"

#include <ctype.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    char *name;
    int len, line_len, line_blen;
    long long offset;
} fa_entry;

struct ssn_fasta {
    char *fn;
    fa_entry *e;
    int n, m;
};

static void fa_add(ssn_fasta *fa, const char *name, int len, int line_len, int line_blen, long long offset)
{
    int i;
    /* same key twice: the reference's hash keeps one slot per name and the later values win */
    for (i = 0; i < fa->n; ++i)
        if (strcmp(fa->e[i].name, name) == 0) break;
    if (i == fa->n) {
        if (fa->n == fa->m) {
            fa->m = fa->m ? fa->m << 1 : 16;
            fa->e = (fa_entry*)realloc(fa->e, sizeof(fa_entry) * (size_t)fa->m);
        }
        fa->e[fa->n].name = strdup(name);
        ++fa->n;
    }
    fa->e[i].len = len; fa->e[i].line_len = line_len; fa->e[i].line_blen = line_blen; fa->e[i].offset = offset;
}

/* Writes <fn>.fai with the layout rules of fai_build_core: offset = first byte after the header line,
 * line_blen = isgraph() bytes per line, line_len = bytes per line including the newline. */
static int fa_build(const char *fn, const char *fai_fn)
{
    FILE *in = fopen(fn, "rb"), *out;
    char *name = 0;
    size_t l_name = 0, m_name = 0;
    int c, len = -1, line_len = -1, line_blen = -1, state = 0, l1, l2;
    long long offset = 0, at = 0;
    ssn_fasta tmp;
    int i;
    memset(&tmp, 0, sizeof(tmp));
    if (!in) { fprintf(stderr, "[fai_build] fail to open the FASTA file.\n"); return -1; }
    while ((c = getc(in)) != EOF) {
        ++at;
        if (c == '\n') {
            if (state == 1) { offset = at; continue; }
            else if ((state == 0 && len < 0) || state == 2) continue;
        }
        if (c == '>') {
            if (len >= 0) fa_add(&tmp, name, len, line_len, line_blen, offset);
            l_name = 0;
            while ((c = getc(in)) != EOF && (++at, !isspace(c))) {
                if (m_name < l_name + 2) { m_name = (l_name + 2) * 2; name = (char*)realloc(name, m_name); }
                name[l_name++] = (char)c;
            }
            if (!name) { name = (char*)malloc(2); m_name = 2; }
            name[l_name] = 0;
            if (c == EOF) {
                fprintf(stderr, "[fai_build_core] the last entry has no sequence\n");
                fclose(in); free(name); return -1;
            }
            if (c != '\n') while ((c = getc(in)) != EOF && (++at, c != '\n')) {}
            state = 1; len = 0; offset = at;
        } else {
            if (state == 3) {
                fprintf(stderr, "[fai_build_core] inlined empty line is not allowed in sequence '%s'.\n", name ? name : "");
                fclose(in); free(name); return -1;
            }
            if (state == 2) state = 3;
            l1 = l2 = 0;
            do {
                ++l1;
                if (isgraph(c)) ++l2;
            } while ((c = getc(in)) != EOF && (++at, c != '\n'));
            if (state == 3 && l2) {
                fprintf(stderr, "[fai_build_core] different line length in sequence '%s'.\n", name ? name : "");
                fclose(in); free(name); return -1;
            }
            ++l1; len += l2;
            if (l2 >= 0x10000) {
                fprintf(stderr, "[fai_build_core] line length exceeds 65535 in sequence '%s'.\n", name ? name : "");
                fclose(in); free(name); return -1;
            }
            if (state == 1) { line_len = l1; line_blen = l2; state = 0; }
            else if (state == 0) { if (l1 != line_len || l2 != line_blen) state = 2; }
        }
    }
    if (len >= 0) fa_add(&tmp, name, len, line_len, line_blen, offset);
    free(name);
    fclose(in);
    out = fopen(fai_fn, "wb");
    if (!out) {
        fprintf(stderr, "[fai_build] fail to write FASTA index.\n");
        for (i = 0; i < tmp.n; ++i) free(tmp.e[i].name);
        free(tmp.e);
        return -1;
    }
    for (i = 0; i < tmp.n; ++i) {
        fprintf(out, "%s\t%d\t%lld\t%d\t%d\n", tmp.e[i].name, tmp.e[i].len, tmp.e[i].offset, tmp.e[i].line_blen, tmp.e[i].line_len);
        free(tmp.e[i].name);
    }
    free(tmp.e);
    fclose(out);
    return 0;
}

ssn_fasta *ssn_fasta_load(const char *fn)
{
    ssn_fasta *fa;
    char *fai_fn = (char*)calloc(strlen(fn) + 5, 1), *buf;
    FILE *fp;
    sprintf(fai_fn, "%s.fai", fn);
    fp = fopen(fai_fn, "rb");
    if (!fp) {
        fprintf(stderr, "[fai_load] build FASTA index.\n");
        fa_build(fn, fai_fn);
        fp = fopen(fai_fn, "r");
        if (!fp) {
            fprintf(stderr, "[fai_load] fail to open FASTA index.\n");
            free(fai_fn);
            return 0;
        }
    }
    fa = (ssn_fasta*)calloc(1, sizeof(*fa));
    buf = (char*)calloc(0x10000, 1);
    while (!feof(fp) && fgets(buf, 0x10000, fp)) {
        char *p;
        int len = 0, line_len = 0, line_blen = 0;
        long long offset = 0;
        for (p = buf; *p && isgraph((unsigned char)*p); ++p) {}
        *p = 0; ++p;
        sscanf(p, "%d%lld%d%d", &len, &offset, &line_blen, &line_len);
        fa_add(fa, buf, len, line_len, line_blen, offset);
    }
    free(buf);
    fclose(fp);
    free(fai_fn);
    fp = fopen(fn, "rb");
    if (!fp) {
        fprintf(stderr, "[fai_load] fail to open FASTA file.\n");
        ssn_fasta_destroy(fa);
        return 0;
    }
    fclose(fp);
    fa->fn = strdup(fn);
    return fa;
}

char *ssn_fasta_fetch(const ssn_fasta *fa, const char *str, int *len)
{
    const size_t l = strlen(str);
    char *s = (char*)malloc(l + 1), *seq;
    size_t i, k = 0;
    const fa_entry *e = 0;
    FILE *fp;
    int n = 0, j, c;
    *len = 0;
    for (i = 0; i < l; ++i)
        if (str[i] != ',' && !isspace((unsigned char)str[i])) s[k++] = str[i];
    s[k] = 0;
    for (i = 0; i < k; ++i) if (s[i] == ':') break;
    if (i < k) {
        /* "name:beg-end": never a whole target; the reference would fetch that sub-range.  Target names with a
         * ':' therefore do not resolve to their own sequence (QUIRK B15); handled as "unknown" apart from the
         * degenerate case where the prefix is itself a target, which no real header produces. */
        s[i] = 0;
    }
    for (j = 0; j < fa->n; ++j)
        if (strcmp(fa->e[j].name, s) == 0) { e = &fa->e[j]; break; }
    if (!e) { free(s); return 0; }
    {
        int beg = 0, end = e->len;
        if (i < k) {                                    /* region arithmetic of fai_fetch :268-279 */
            char *p = s + i + 1;
            size_t t = i + 1;
            for (; t < k; ++t) if (s[t] == '-') break;
            beg = atoi(p);
            end = t < k ? atoi(s + t + 1) : e->len;
            if (beg > 0) --beg;
            if (beg >= e->len) beg = e->len;
            if (end >= e->len) end = e->len;
            if (beg > end) beg = end;
        }
        free(s);
        seq = (char*)malloc((size_t)(end - beg) + 2);
        fp = fopen(fa->fn, "rb");
        if (!fp) { seq[0] = 0; return seq; }
        if (e->line_blen > 0)
            fseeko(fp, (off_t)(e->offset + (long long)(beg / e->line_blen) * e->line_len + beg % e->line_blen), SEEK_SET);
        else fseeko(fp, (off_t)e->offset, SEEK_SET);
        while (n < end - beg && (c = getc_unlocked(fp)) != EOF)
            if (isgraph(c)) seq[n++] = (char)c;
        seq[n] = 0;
        fclose(fp);
    }
    *len = n;
    return seq;
}

void ssn_fasta_destroy(ssn_fasta *fa)
{
    int i;
    if (!fa) return;
    for (i = 0; i < fa->n; ++i) free(fa->e[i].name);
    free(fa->e); free(fa->fn); free(fa);
}
