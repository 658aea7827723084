// lpc_front.h -- LPC front end (signal -> gain-normalised autocorrelation vectors) on sm_100a.
// Reference: ecoz2/src/lpc/lpa_on_signal.c:81-157, ecoz2/src/lpc/lpca.c:9-59.
#pragma once
#include <cuda_runtime.h>

namespace lpcb200 {

constexpr int kMaxOrder = 255;

const char *last_error();

// frame geometry of lpa_on_signal.c:84-101: window / offset in samples and the number of
// complete frames T; -1 (and last_error) when the signal is shorter than one window
long num_frames(long num_samples, double sample_rate, int window_ms, int offset_ms, int *win_out,
                int *off_out);

// d_signal [num_samples] -> d_vectors [T][P+1] (both device memory), asynchronous on `stream`.
// d_hamming: device scratch of `win` doubles.  d_err_count (optional, device int, zeroed by
// the caller) counts the frames on which lpca reports an error.  Returns false on launch errors.
bool lpa_device(const double *d_signal, long T, int win, int off, int P, double *d_vectors,
                double *d_hamming, int *d_err_count, cudaStream_t stream);

// plain lpca (lpca.c:9-59) of T frames of N samples, frame t at d_x + t * stride (device memory);
// outputs [T][P+1] (rc index 0 untouched), pe [T], status [T]; any output may be null except r
bool lpca_device(const double *d_x, long T, int N, long stride, int P, double *d_r, double *d_rc,
                 double *d_a, double *d_pe, int *d_status, cudaStream_t stream);

// measured FP64 FMA throughput of `device` in TFLOP/s (pure DFMA chains on every SM for ~millis ms)
double dfma_peak_tflops(int device, double millis);

}  // namespace lpcb200
