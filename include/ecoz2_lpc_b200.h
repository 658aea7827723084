/* ecoz2_lpc_b200.h -- C ABI of the LPC front end of the VQ path on B200.
 *
 * The step before `vq learn` / `vq quantize`: a signal is cut into overlapping frames, each
 * frame is mean-removed, pre-emphasised (0.95), Hamming-windowed, its autocorrelation
 * r[0..P] is taken, Levinson-Durbin gives the prediction error, and the vector r / errPred
 * is what a `.prd` file stores and what the VQ kernels consume.  Everything below computes
 * on the GPU in FP64 (sample_t is double in the reference, sgn.h:11); there is no CPU path.
 *
 * Exported by the same library as the VQ entry points (libecoz2_vq_b200.so).
 */
#ifndef ECOZ2_LPC_B200_H
#define ECOZ2_LPC_B200_H

#ifdef __cplusplus
extern "C" {
#endif

/* ---- drop-in entry points (same names and signatures as the reference) ---------------- */

/* Replaces ecoz2_lpc_signals, reference ecoz2/src/include/ecoz2.h:23-33
 * (ecoz2/src/ecoz2/ecoz2.c:56-80 -> ecoz2/src/lpc/lpc_signals.c:38-153): groups the wav files
 * by class (parent directory), shuffles each class with rand() exactly like list_shuffle
 * (utl/list.c:84-92, so ecoz2_set_random_seed gives the same train/test split), analyses
 * every signal and writes data/predictors[/TRAIN|/TEST]/<class>/<name>.prd (prd.c:76-115).
 * Mono RIFF/WAVE, PCM 8/16/24/32 bit or IEEE float, converted to float32 then double like
 * the reference's dr_wav path (sgn/sgn.c:12-52).  Always returns 0, like the reference. */
int ecoz2_lpc_signals(int P, int windowLengthMs, int offsetLengthMs, int minpc, float split,
                      char *sgn_filenames[], int num_signals, float mintrpt, int verbose);

/* Replaces ecoz2_lpca, reference ecoz2/src/include/ecoz2.h:39-47 (ecoz2/src/lpc/lpca.c:9-59):
 * autocorrelation r[0..P] of x[0..N), reflection coefficients rc[1..P], predictor a[0..P],
 * prediction error *pe.  Returns lpca's status (0; 1: r[0] == 0; 2: error <= 0), or -1 when
 * no GPU is present / arguments are out of range (P in 1..255). */
int ecoz2_lpca(double *x, int N, int P, double *r, double *rc, double *a, double *pe);

/* ---- array-level equivalent of lpa_on_signal ------------------------------------------ */

/* Frames that lpa_on_signal (reference ecoz2/src/lpc/lpa_on_signal.c:84-101) makes of a signal:
 * winSize = windowLengthMs * (long) sampleRate / 1000, offset likewise, incomplete tail
 * dropped.  -1 when the signal is shorter than one window ("signal too short"). */
long ecoz2b200_lpa_num_frames(long num_samples, double sample_rate, int window_ms, int offset_ms);

/* Replaces lpa_on_signal(P, windowLengthMs, offsetLengthMs, sgn, verbose), reference
 * ecoz2/src/include/lpc.h:72-82 / ecoz2/src/lpc/lpa_on_signal.c:81-157, with the Sgn and the
 * Predictor flattened to arrays: samples [num_samples] -> vectors [T][P+1] (host memory).
 * Returns T, or -1 (see ecoz2b200_lpc_last_error).  *num_errors (optional) = frames on which
 * lpca reported an error (the reference prints one line per such frame and carries on). */
long ecoz2b200_lpa_on_signal(const double *samples, long num_samples, double sample_rate, int P,
                             int window_ms, int offset_ms, double *vectors, long capacity_frames,
                             int *num_errors);

/* Same with the signal and the vectors in device memory (the vectors can be handed to
 * ecoz2b200 VQ engines without leaving HBM); `stream` is a cudaStream_t or NULL.
 * Synchronises the stream before returning. */
long ecoz2b200_lpa_on_signal_device(const double *d_samples, long num_samples, double sample_rate,
                                    int P, int window_ms, int offset_ms, double *d_vectors,
                                    long capacity_frames, int *num_errors, void *stream);

const char *ecoz2b200_lpc_last_error(void);

/* Measured FP64 FMA throughput of `device` in TFLOP/s (a pure DFMA-chain kernel filling every
 * SM for ~`millis` ms): the denominator of the autocorrelation kernel's roofline. */
double ecoz2b200_dfma_peak_tflops(int device, double millis);

#ifdef __cplusplus
}
#endif
#endif
