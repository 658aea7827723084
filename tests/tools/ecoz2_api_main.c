/* tests/tools/ecoz2_api_main.c -- a caller of the reference's PUBLIC C API (ecoz2/src/include/ecoz2.h),
 * i.e. the three calls the Rust crate makes (src/ecoz2_lib/mod.rs:96-122, :252-342).  Linked by
 * oracle/Makefile (target `linkcheck`) against the reference's static `ecoz2_lib` built from the file
 * list of the PATCHED build.rs (integration/build.rs.patch: no vq_learn.c / vq_quantize.c /
 * vq_classify.c / report.c / sigma.c / inertia.c) plus libecoz2_vq_b200.so: the reference's unmodified
 * ecoz2.c forwards ecoz2_vq_learn -> vq_learn, and vq_learn is the CUDA library's.
 *
 *   linked.ecoz2main learn <P> <eps> <class> <prd>...
 *   linked.ecoz2main quantize <cbook> <prd>...                                                  */
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include "ecoz2.h"

static void on_stage(void *target, int M, double avg, double sigma, double inertia) {
    int *n = (int *) target;
    (*n)++;
    printf("[callback %d] M=%d avgDistortion=%.9g sigma=%.9g inertia=%.9g\n", *n, M, avg, sigma, inertia);
}

int main(int argc, char **argv) {
    if (argc >= 6 && strcmp(argv[1], "learn") == 0) {
        int calls = 0;
        const int rc = ecoz2_vq_learn(atoi(argv[2]), atof(argv[3]), argv[4], (const char **) (argv + 5),
                                      argc - 5, &calls, on_stage);
        printf("ecoz2_vq_learn -> %d after %d callbacks (%s)\n", rc, calls, ecoz2_version());
        return rc;
    }
    if (argc >= 4 && strcmp(argv[1], "quantize") == 0)
        return ecoz2_vq_quantize(argv[2], (const char **) (argv + 3), argc - 3, 0);
    fprintf(stderr, "usage: %s learn <P> <eps> <class> <prd>... | quantize <cbook> <prd>...\n", argv[0]);
    return 64;
}
