/* oracle/ref_stubs.c -- TEST INFRASTRUCTURE ONLY.
 *
 * The reference's C mains (ecoz2/src/x/vq.learn.c:105) call
 * ecoz2_set_random_seed(), which lives in ecoz2/src/ecoz2/ecoz2.c together
 * with the HMM/LPC/signal entry points that are out of scope here.  The VQ
 * path never draws a random number, so a trivial stand-in keeps the link
 * closed without dragging the rest of the system in.
 */
#include <stdlib.h>

unsigned long ecoz2_set_random_seed(long seed) {
    unsigned u = seed < 0 ? 1u : (unsigned) seed;
    srand(u);
    return u;
}
