/* oracle/ref_vq_learn_wrap.c -- TEST INFRASTRUCTURE ONLY.
 *
 * Compiles the REAL reference LBG driver (ecoz2/src/vq/vq_learn.c, which
 * textually includes vq_learn.i, vq_learn_par.c and vq_learn_ser.c) with the
 * stop size MAX_CODEBOOK_SIZE (ecoz2/src/include/vq.h:11) overridden to
 * REF_MAX_M.  The reference has no run-time "target size"; its loop ends only
 * at that constant (ecoz2/src/vq/vq_learn_par.c:147).  No reference source is
 * copied: the include below resolves into /root/reference via -I.
 *
 * It additionally exports thin accessors to the file-static codebook arrays so
 * the tests can read the reference's in-memory state after a run.
 */
#include "vq.h"

#ifndef REF_MAX_M
#error "REF_MAX_M must be defined (see oracle/Makefile)"
#endif
#undef MAX_CODEBOOK_SIZE
#define MAX_CODEBOOK_SIZE REF_MAX_M

#include "vq/vq_learn.c"

int ref_max_codebook_size(void) { return MAX_CODEBOOK_SIZE; }

/* Current raas / reflections of the reference's static state (row stride 1+P). */
const sample_t *ref_codebook_ptr(void) { return codebook; }
const sample_t *ref_reflections_ptr(void) { return reflections; }

/* One teacher-forced centroid step through the reference's own static
 * calculate_reflections (ecoz2/src/vq/vq_learn.i:125-155): the caller provides
 * the previous reflections/raas (they stay untouched for empty cells), the cell
 * sums and cardinalities; the updated arrays are copied back out. */
void ref_calculate_reflections(int P, int num_raas,
                               sample_t *refl_io, sample_t *raas_io,
                               sample_t *cells_io, int *cardd) {
    const long n = (long) num_raas * (1 + P);
    memcpy(reflections, refl_io, n * sizeof(sample_t));
    memcpy(codebook, raas_io, n * sizeof(sample_t));
    calculate_reflections(P, cells_io, cardd, num_raas);
    memcpy(refl_io, reflections, n * sizeof(sample_t));
    memcpy(raas_io, codebook, n * sizeof(sample_t));
}

/* Teacher-forced split through the reference's own static grow_codebook
 * (ecoz2/src/vq/vq_learn.i:47-72). refl_io/raas_out must hold 2*num_raas rows. */
int ref_grow_codebook(int P, int num_raas, sample_t *refl_io, sample_t *raas_out) {
    const long n = (long) num_raas * (1 + P);
    memcpy(reflections, refl_io, n * sizeof(sample_t));
    int m = grow_codebook(num_raas, P);
    memcpy(refl_io, reflections, 2 * n * sizeof(sample_t));
    memcpy(raas_out, codebook, 2 * n * sizeof(sample_t));
    return m;
}

/* The reference's 2-codeword start (ecoz2/src/vq/vq_learn.i:31-44). */
int ref_initial_codebook(int P, sample_t *refl_out, sample_t *raas_out) {
    init_max_codebook_and_reflections(P);
    int m = initial_codebook(P);
    const long n = (long) m * (1 + P);
    memcpy(refl_out, reflections, n * sizeof(sample_t));
    memcpy(raas_out, codebook, n * sizeof(sample_t));
    return m;
}
