/* tests/tools/seqprov_main.c -- TEST INFRASTRUCTURE.  A caller of the reference's SeqProvider
 * (ecoz2/src/include/vq.h:150-186, vq/vq.c:162-334) the way hmm_classify.c:396-440 uses it with
 * codebooks + predictors: for every predictor file, the symbol sequence and the distortion of
 * every codebook.  Built twice by oracle/Makefile: with the reference's own quantize.c, and with
 * quantize() coming from libecoz2_vq_b200.so (vq.c, prd.c ... still the reference's, unmodified).
 * usage: seqprov <num_codebooks> <cb files...> <prd files...>
 * output: one line per (file, model): file index, model index, T, distortion (%.12e), FNV-1a of the
 * symbols, first 8 symbols. */
#include <stdio.h>
#include <stdlib.h>

#include "vq.h"

int main(int argc, char **argv) {
    if (argc < 4) return 2;
    const int R = atoi(argv[1]);
    if (R < 1 || argc < 2 + R + 1) return 2;
    char **cbs = argv + 2;
    char **prds = argv + 2 + R;
    const int F = argc - 2 - R;
    SeqProvider *sp = seq_provider_create(R, 0, 0, 0, cbs, R, prds, F);
    if (!sp) return 3;
    int file = 0;
    while (seq_provider_has_next(sp)) {
        NextSeq *ns = seq_provider_get_next(sp);
        if (!ns) return 4;
        for (int r = 0; r < R; r++) {
            unsigned long long h = 1469598103934665603ull;
            for (int t = 0; t < ns->T; t++) {
                h ^= (unsigned long long) ns->sequences[r][t];
                h *= 1099511628211ull;
            }
            printf("%d %d %d %.12e %016llx", file, r, ns->T, (double) ns->dists[r], h);
            for (int t = 0; t < ns->T && t < 8; t++) printf(" %d", (int) ns->sequences[r][t]);
            printf("\n");
        }
        file++;
    }
    seq_provider_destroy(sp);
    return 0;
}
