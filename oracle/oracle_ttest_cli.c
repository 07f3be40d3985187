/* CPU ORACLE — test infrastructure only (see dekupl_oracle.h). Used by bench.py's CPU baseline / --impl reference.
 *
 * ttest_text: the row pass of TtestFilter (share/TtestFilter/TtestFilter.c:373-440) as a PROCESS over the text
 * joinCounts printed, so that the CPU baseline pays what the reference pays: reading the table as text (the reference
 * tokenises with kseq/atoi, :373-383), normalising, the Welch test per row. TtestFilter itself cannot be built here
 * (Boost.Math absent), so this is the port (oracle_ttest.c) behind a main(); BH and the final filter (:456-497) are
 * done by the caller on the concatenated p-values, single-threaded like the reference.
 *
 * usage: ttest_text <matrix.tsv> <S> <nA> <out.p> <out.fc> nf_0 .. nf_{S-1}
 *        columns 0..nA-1 are condition A, nA..S-1 condition B; a first line that does not start with ACGT is a header.
 *        out.p / out.fc receive one native double per row (raw p-value, log2FC). Prints the row count. */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include "dekupl_oracle.h"

#define BATCH 65536

int main(int argc, char **argv)
{
    if (argc < 6) { fprintf(stderr, "usage: ttest_text matrix.tsv S nA out.p out.fc nf...\n"); return 1; }
    const int S = atoi(argv[2]), nA = atoi(argv[3]);
    if (S < 2 || nA < 1 || nA >= S || argc != 6 + S) { fprintf(stderr, "ttest_text: bad arguments\n"); return 1; }
    const int nB = S - nA;
    double *nf = (double *)malloc(sizeof(double) * S);
    int *idx = (int *)malloc(sizeof(int) * S);
    for (int s = 0; s < S; s++) { nf[s] = atof(argv[6 + s]); idx[s] = s; }      /* atof: TtestFilter.c:261 */
    FILE *f = fopen(argv[1], "r"), *fp = fopen(argv[4], "wb"), *ff = fopen(argv[5], "wb");
    if (!f || !fp || !ff) { fprintf(stderr, "ttest_text: cannot open files\n"); return 1; }
    int32_t *c = (int32_t *)malloc(sizeof(int32_t) * (size_t)BATCH * S);
    double *p = (double *)malloc(sizeof(double) * BATCH), *mA = (double *)malloc(sizeof(double) * BATCH),
           *mB = (double *)malloc(sizeof(double) * BATCH), *fc = (double *)malloc(sizeof(double) * BATCH);
    size_t bufn = 1 << 20, nb = 0, R = 0;
    char *line = (char *)malloc(bufn);
    int first = 1;
    for (;;) {
        char *got = fgets(line, (int)bufn, f);
        if (got) {
            if (first) { first = 0; if (!(line[0] == 'A' || line[0] == 'C' || line[0] == 'G' || line[0] == 'T')) continue; }
            char *q = line;
            while (*q && *q != '\t' && *q != ' ' && *q != '\n') q++;              /* the k-mer token */
            int s = 0;
            while (s < S && *q && *q != '\n') {
                while (*q == '\t' || *q == ' ') q++;
                c[nb * S + s++] = atoi(q);                                        /* :380 */
                while (*q && *q != '\t' && *q != ' ' && *q != '\n') q++;
            }
            for (; s < S; s++) c[nb * S + s] = 0;
            nb++;
        }
        if (nb == BATCH || (!got && nb)) {
            orc_ttest_rows(c, nb, S, nf, idx, nA, idx + nA, nB, p, mA, mB, fc);
            fwrite(p, sizeof(double), nb, fp);
            fwrite(fc, sizeof(double), nb, ff);
            R += nb; nb = 0;
        }
        if (!got) break;
    }
    fclose(f); fclose(fp); fclose(ff);
    printf("%zu\n", R);
    return 0;
}
