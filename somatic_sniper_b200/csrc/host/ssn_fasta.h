/* ssn_fasta.h — reference sequence access with the semantics of samtools-0.1.6 faidx.c as the reference
 * uses it (main.c:106-108 fai_load, somatic_sniper.c:112-117 fai_fetch of a whole target):
 *   - <fasta>.fai is read if present, else built beside the FASTA (fai_load :216-240, fai_build_core :58-130);
 *   - the name handed to fetch is treated as a region string: ',' and white space are squeezed out and the
 *     name ends at the first ':' (fai_fetch :256-279, SURVEY.md QUIRK B15);
 *   - a whole-target fetch seeks to the indexed offset and keeps every isgraph() byte until `len` bases were
 *     collected or the file ends (:287-292); an unknown name yields no sequence (every base reads as 'N'). */
#ifndef SSN_FASTA_H
#define SSN_FASTA_H
#include <stdint.h>

typedef struct ssn_fasta ssn_fasta;

ssn_fasta *ssn_fasta_load(const char *fn);                 /* prints the reference's messages; NULL on failure */
/* malloc'd sequence of the target (NUL terminated), *len = bases read; NULL (*len = 0) when the name is unknown */
char *ssn_fasta_fetch(const ssn_fasta *fa, const char *name, int *len);
void ssn_fasta_destroy(ssn_fasta *fa);

#endif
