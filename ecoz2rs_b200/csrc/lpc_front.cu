// lpc_front.cu -- the LPC front end of the VQ path on sm_100a:
// signal -> frames -> mean removal, pre-emphasis, Hamming window -> autocorrelation r[0..P]
// -> Levinson-Durbin prediction error -> gain-normalised vectors r / errPred, i.e. the
// content of a `.prd` file, produced in device memory.
//
// Reference: ecoz2/src/lpc/lpa_on_signal.c:81-157 (framing :96-101, window pipeline :24-80,
// normalisation :143-150) and ecoz2/src/lpc/lpca.c:9-59 (autocorrelation :16-21, recursion
// :27-47).  All arithmetic is FP64 like the reference (sample_t = double, sgn.h:11).
//
// Two kernels:
//   lpc_autocorr<LAGS>   FP64-FMA bound.  A CTA stages the contiguous stretch of signal that
//       its G overlapping frames cover in shared memory once (a 45 ms window every 15 ms
//       re-uses each sample three times) together with the Hamming table.  Eight lanes share
//       a frame: each takes a contiguous eighth of the window and keeps LAGS running sums plus
//       a sliding register window of the windowed samples, so one new sample costs a few
//       shared-memory loads against LAGS FMAs with register operands.  Windowed samples are
//       recomputed on the fly from the raw ones -- no per-frame copy exists anywhere.
//   lpc_levinson<PMAX>   thread = frame, the recursion fully unrolled in registers, same
//       operation order as lpca.c; divides the vector by the final prediction error in place.
#include <cuda_runtime.h>
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#include <cstdlib>
#include <string>
#include <vector>

#include "lpc_front.h"

namespace lpcb200 {

namespace {

constexpr int kLanesPerFrame = 8;
constexpr int kStep = 8;   // window positions per unrolled block

struct AcParams {
    const double *signal;   // [num_samples]
    const double *hamming;  // [win] (windowed mode)
    double *r;              // [T][ld]
    long T;
    int win, off, P, ld;
    int frames_per_cta;     // G
    int staged;             // signal stretch + Hamming table in shared memory
    int windowed;           // 0: plain autocorrelation of each frame (ecoz2_lpca)
    int lane_stride;        // fast kernel: samples between the pieces of consecutive lanes (odd)
    int bulk;               // fast kernel: signal and Hamming table 16-byte aligned -> cp.async.bulk staging
};

// windowed sample n of a frame (0 outside [0, N)): lpa_on_signal.c:42-80
__device__ __forceinline__ double win_sample(const double *x, const double *h, int n, int N, double mean,
                                             bool windowed) {
    if (n >= N) return 0.0;
    const double d0 = x[n];
    if (!windowed) return d0;
    const double c0 = d0 - mean;
    const double c1 = n > 0 ? x[n - 1] - mean : 0.0;
    return (c0 - .95 * c1) * h[n];
}

template <int LAGS>
__global__ void __launch_bounds__(256, 1) lpc_autocorr(const AcParams p) {
    extern __shared__ __align__(16) unsigned char lpc_smem[];
    const int G = p.frames_per_cta;
    const long t0 = (long) blockIdx.x * G;
    const int g = threadIdx.x / kLanesPerFrame, j = threadIdx.x % kLanesPerFrame;
    const int N = p.win;
    const double *xs;   // samples of frame t0 (shared or global)
    const double *hs;
    if (p.staged) {
        double *s_sig = reinterpret_cast<double *>(lpc_smem);
        const long n_frames = p.T - t0 < G ? p.T - t0 : G;
        const int span = (int) (n_frames - 1) * p.off + N;
        double *s_ham = s_sig + (size_t) (G - 1) * p.off + N;
        const double *src = p.signal + t0 * p.off;
        for (int i = threadIdx.x; i < span; i += blockDim.x) s_sig[i] = src[i];
        if (p.windowed)
            for (int i = threadIdx.x; i < N; i += blockDim.x) s_ham[i] = p.hamming[i];
        __syncthreads();
        xs = s_sig;
        hs = s_ham;
    } else {
        xs = p.signal + t0 * p.off;
        hs = p.hamming;
    }
    const long t = t0 + g;
    if (g >= G || t >= p.T) return;   // whole 8-lane groups leave together
    const double *x = xs + (size_t) g * p.off;
    const unsigned gmask = 0xffu << ((threadIdx.x & 31) & ~7);
    const bool windowed = p.windowed != 0;

    // this lane's eighth of the window
    const int L = (N + kLanesPerFrame - 1) / kLanesPerFrame;
    const int c0 = j * L, c1 = c0 + L < N ? c0 + L : N;

    double mean = 0.0;
    if (windowed) {
        double s = 0.0;
        for (int n = c0; n < c1; n++) s += x[n];
        s += __shfl_xor_sync(gmask, s, 4);
        s += __shfl_xor_sync(gmask, s, 2);
        s += __shfl_xor_sync(gmask, s, 1);
        mean = s / N;
    }

    for (int l0 = 0; l0 <= p.P; l0 += LAGS) {
        double acc[LAGS];
#pragma unroll
        for (int i = 0; i < LAGS; i++) acc[i] = 0.0;
        // wb[q] = y(k + l0 + q), q < LAGS + kStep - 1
        double wb[LAGS + kStep - 1];
#pragma unroll
        for (int q = 0; q < LAGS - 1; q++) wb[q] = win_sample(x, hs, c0 + l0 + q, N, mean, windowed);
        for (int k = c0; k < c1; k += kStep) {
            double a[kStep];
#pragma unroll
            for (int u = 0; u < kStep; u++) {
                a[u] = k + u < c1 ? win_sample(x, hs, k + u, N, mean, windowed) : 0.0;
                wb[LAGS - 1 + u] = win_sample(x, hs, k + l0 + LAGS - 1 + u, N, mean, windowed);
            }
#pragma unroll
            for (int u = 0; u < kStep; u++)
#pragma unroll
                for (int i = 0; i < LAGS; i++) acc[i] = fma(a[u], wb[u + i], acc[i]);
#pragma unroll
            for (int q = 0; q < LAGS - 1; q++) wb[q] = wb[q + kStep];
        }
        double *out = p.r + t * p.ld + l0;
#pragma unroll
        for (int i = 0; i < LAGS; i++) {
            double s = acc[i];
            s += __shfl_xor_sync(gmask, s, 4);
            s += __shfl_xor_sync(gmask, s, 2);
            s += __shfl_xor_sync(gmask, s, 1);
            if ((i % kLanesPerFrame) == j && l0 + i <= p.P) out[i] = s;
        }
    }
}

// ---------------------------------------------------------------------------------------
// lpc_autocorr_fast<LAGS>: the windowed case with P < LAGS and a stretch of signal that
// fits shared memory twice per SM -- what lpa_on_signal runs at the usual 45/15 ms framing.
//
// Same decomposition as lpc_autocorr (8 lanes share a frame, each owns a contiguous piece
// of the window and all LAGS running sums) with the inner loop made branch-free:
//   * the staged signal and the Hamming table are zero-extended by kFastPad entries, so the
//     sliding window may run past the end of the frame without bounds checks on the loads
//     (the value is forced to zero by a select, never by arithmetic on what was loaded);
//   * each windowed sample is produced exactly once per lane (3 FP64 ops: centre, pre-
//     emphasise with the previous centred sample kept in a register, Hamming) and enters
//     the LAGS x 8 block of DFMAs both as the "row" and the "column" operand;
//   * raw samples of the next block are loaded before the 8 x LAGS DFMAs of this one;
//   * lane pieces start every Ls samples with Ls odd, so the 8 lanes of a frame hit
//     distinct 8-byte banks; the piece is processed in whole blocks of 8, the rows past
//     the end of the piece (they belong to the next lane) are masked to zero;
//   * 16 frames (128 threads) per CTA, two CTAs per SM: one stages its signal (one
//     cp.async.bulk issued by thread 0 when the pointers are 16-byte aligned, 8-byte cp.async
//     per thread otherwise) while the other keeps the FP64 pipe busy;
//   * all 32 lanes of a warp stay to the end (groups past the last frame recompute frame t0
//     and store nothing): full-mask shuffles, no reconvergence code around them.
// ---------------------------------------------------------------------------------------
constexpr int kFastPad = 112;   // >= (odd-Ls slack 15) + (block round-up 7) + LAGS - 1 + 2 * kStep

template <int LAGS>
__global__ void __launch_bounds__(128, 2) lpc_autocorr_fast(const AcParams p) {
    extern __shared__ __align__(16) unsigned char lpc_smem[];
    const int G = p.frames_per_cta;
    const long t0 = (long) blockIdx.x * G;
    const int N = p.win;
    double *s_sig = reinterpret_cast<double *>(lpc_smem);
    const int sig_cap = ((G - 1) * p.off + N + kFastPad + 1) & ~1;   // even: s_ham stays 16-byte aligned
    double *s_ham = s_sig + sig_cap;
    __shared__ uint64_t s_bar;
    {
        const long n_frames = p.T - t0 < G ? p.T - t0 : G;
        const int span = (int) (n_frames - 1) * p.off + N;
        const double *src = p.signal + t0 * p.off;
        if (p.bulk) {
            // one elected thread hands the two copies to the TMA engine (16-byte granules; an odd
            // last element is copied by hand)
            const uint32_t bar = (uint32_t) __cvta_generic_to_shared(&s_bar);
            if (threadIdx.x == 0) {
                asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
                asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
            }
            __syncthreads();
            if (threadIdx.x == 0) {
                const uint32_t b_sig = (uint32_t) (span & ~1) * 8u, b_ham = (uint32_t) (N & ~1) * 8u;
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(b_sig + b_ham)
                             : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                                 "r"((uint32_t) __cvta_generic_to_shared(s_sig)), "l"(src), "r"(b_sig), "r"(bar)
                             : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                                 "r"((uint32_t) __cvta_generic_to_shared(s_ham)), "l"(p.hamming), "r"(b_ham), "r"(bar)
                             : "memory");
            }
            if (threadIdx.x == 32 && (span & 1)) s_sig[span - 1] = src[span - 1];
            if (threadIdx.x == 64 && (N & 1)) s_ham[N - 1] = p.hamming[N - 1];
        } else {
            for (int i = threadIdx.x; i < span; i += blockDim.x) {
                const unsigned dst = (unsigned) __cvta_generic_to_shared(s_sig + i);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(src + i) : "memory");
            }
            for (int i = threadIdx.x; i < N; i += blockDim.x) {
                const unsigned dst = (unsigned) __cvta_generic_to_shared(s_ham + i);
                asm volatile("cp.async.ca.shared.global [%0], [%1], 8;" ::"r"(dst), "l"(p.hamming + i) : "memory");
            }
            asm volatile("cp.async.commit_group;" ::: "memory");
        }
        for (int i = span + threadIdx.x; i < sig_cap; i += blockDim.x) s_sig[i] = 0.0;
        for (int i = N + threadIdx.x; i < N + kFastPad; i += blockDim.x) s_ham[i] = 0.0;
        if (p.bulk) {
            const uint32_t bar = (uint32_t) __cvta_generic_to_shared(&s_bar);
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n.reg .pred q;\nmbarrier.try_wait.parity.shared::cta.b64 q, [%1], 0;\n"
                             "selp.u32 %0, 1, 0, q;\n}\n"
                             : "=r"(done)
                             : "r"(bar)
                             : "memory");
        } else {
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
        __syncthreads();
    }
    // a warp holds 4 frames; warps past the last frame leave, groups past it compute frame t0 again
    // and store nothing (all 32 lanes stay for the full-mask shuffles)
    if (t0 + (threadIdx.x >> 5) * (32 / kLanesPerFrame) >= p.T) return;
    const int j = threadIdx.x % kLanesPerFrame;
    const long t = t0 + threadIdx.x / kLanesPerFrame;
    const bool valid = t < p.T;
    const int g = valid ? threadIdx.x / kLanesPerFrame : 0;
    constexpr unsigned gmask = 0xffffffffu;
    const int Ls = p.lane_stride;
    const int c0 = j * Ls;
    const int c1 = c0 + Ls < N ? c0 + Ls : N;   // may be <= c0 for tiny windows: nothing to add
    const double *x = s_sig + (size_t) g * p.off;

    double mean;
    {
        double s0 = 0.0, s1 = 0.0, s2 = 0.0, s3 = 0.0;   // four chains: the adds are latency bound
        int n = c0;
        for (; n + 4 <= c1; n += 4) {
            s0 += x[n];
            s1 += x[n + 1];
            s2 += x[n + 2];
            s3 += x[n + 3];
        }
        for (; n < c1; n++) s0 += x[n];
        double s = (s0 + s1) + (s2 + s3);
        s += __shfl_xor_sync(gmask, s, 4);
        s += __shfl_xor_sync(gmask, s, 2);
        s += __shfl_xor_sync(gmask, s, 1);
        mean = s / N;
    }

    const double *xp = x + c0, *hp = s_ham + c0;
    // centred sample before the piece (lpa_on_signal.c:52-60: the first sample has no predecessor)
    double cprev = c0 > 0 ? xp[-1] - mean : 0.0;
    int lim = N - c0;   // windowed sample m of this piece is zero for m >= lim
    // y(m) = ((x[m] - mean) - .95 (x[m-1] - mean)) * h[m]
#define LPC_GEN(dst, xv, hv, m)                                \
    {                                                          \
        const double c_ = (xv) - mean;                         \
        const double v_ = (c_ - .95 * cprev) * (hv);           \
        cprev = c_;                                            \
        (dst) = (m) < lim ? v_ : 0.0;                          \
    }
    double acc[LAGS];
#pragma unroll
    for (int i = 0; i < LAGS; i++) acc[i] = 0.0;
    double wb[LAGS - 1 + kStep];   // wb[q] = y(k + q) relative to the piece
#pragma unroll
    for (int q = 0; q < LAGS - 1; q++) LPC_GEN(wb[q], xp[q], hp[q], q)
    double xr[kStep], hr[kStep];
#pragma unroll
    for (int u = 0; u < kStep; u++) {
        xr[u] = xp[LAGS - 1 + u];
        hr[u] = hp[LAGS - 1 + u];
    }
    const int rows = c1 - c0;   // rows of this piece that count
    for (int k = 0; k < Ls; k += kStep) {
#pragma unroll
        for (int u = 0; u < kStep; u++) LPC_GEN(wb[LAGS - 1 + u], xr[u], hr[u], k + LAGS - 1 + u)
#pragma unroll
        for (int u = 0; u < kStep; u++) {   // next block's raw values, in flight during the DFMAs
            xr[u] = xp[k + kStep + LAGS - 1 + u];
            hr[u] = hp[k + kStep + LAGS - 1 + u];
        }
#pragma unroll
        for (int u = 0; u < kStep; u++) {
            const double a = k + u < rows ? wb[u] : 0.0;
#pragma unroll
            for (int i = 0; i < LAGS; i++) acc[i] = fma(a, wb[u + i], acc[i]);
        }
#pragma unroll
        for (int q = 0; q < LAGS - 1; q++) wb[q] = wb[q + kStep];
    }
#undef LPC_GEN
    double *out = p.r + t * p.ld;
#pragma unroll
    for (int i = 0; i < LAGS; i++) {
        double s = acc[i];
        s += __shfl_xor_sync(gmask, s, 4);
        s += __shfl_xor_sync(gmask, s, 2);
        s += __shfl_xor_sync(gmask, s, 1);
        if ((i % kLanesPerFrame) == j && i <= p.P && valid) out[i] = s;
    }
}

struct LvParams {
    double *r;        // [T][ld], divided by the prediction error in place when normalise != 0
    long T;
    int P, ld, normalise;
    double *rc;       // optional [T][ld]: reflection coefficients (index 1..P)
    double *a;        // optional [T][ld]: predictor coefficients
    double *pe;       // optional [T]
    int *status;      // optional [T]: lpca's return value (0, 1: r0 == 0, 2: error <= 0)
    int *err_count;   // frames with status != 0
};

// lpca.c:23-47, the recursion proper (arrays indexed 0..P)
#define LPC_LEVINSON_BODY(UNROLL_PRAGMA)                                                     \
    int status = 0;                                                                          \
    double pe = rr[0];                                                                       \
    if (rr[0] == 0.0) {                                                                      \
        status = 1;                                                                          \
    } else {                                                                                 \
        a[0] = 1.0;                                                                          \
        bool live = true;                                                                    \
        UNROLL_PRAGMA                                                                        \
        for (int k = 1; k <= KMAX; k++) {                                                    \
            if (k <= P && live) {                                                            \
                double sum = 0.0;                                                            \
                UNROLL_PRAGMA                                                                \
                for (int i = 1; i <= k; i++) sum -= a[k - i] * rr[i];                        \
                const double akk = sum / pe;                                                 \
                if (p.rc) p.rc[t * p.ld + k] = akk;                                          \
                a[k] = akk;                                                                  \
                UNROLL_PRAGMA                                                                \
                for (int i = 1; i <= k >> 1; i++) {                                          \
                    const double ai = a[i], aj = a[k - i];                                   \
                    a[i] = ai + akk * aj;                                                    \
                    a[k - i] = aj + akk * ai;                                                \
                }                                                                            \
                pe *= 1.0 - akk * akk;                                                       \
                if (pe <= 0.0) {                                                             \
                    status = 2;                                                              \
                    live = false;                                                            \
                }                                                                            \
            }                                                                                \
        }                                                                                    \
    }

// thread = frame; everything in registers (loops fully unrolled to PMAX)
template <int PMAX>
__global__ void __launch_bounds__(128) lpc_levinson(const LvParams p) {
    const long t = (long) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.T) return;
    const int P = p.P;
    constexpr int KMAX = PMAX;
    double rr[PMAX + 1], a[PMAX + 1];
    double *row = p.r + t * p.ld;
#pragma unroll
    for (int i = 0; i <= PMAX; i++) {
        rr[i] = i <= P ? row[i] : 0.0;
        a[i] = 0.0;
    }
    LPC_LEVINSON_BODY(_Pragma("unroll"))
    if (p.status) p.status[t] = status;
    if (status != 0 && p.err_count) atomicAdd(p.err_count, 1);
    if (p.pe) p.pe[t] = pe;
#pragma unroll
    for (int i = 0; i <= PMAX; i++) {
        if (i <= P) {
            if (p.a) p.a[t * p.ld + i] = a[i];
            // lpa_on_signal.c:143-150 (an all-zero frame, status 1, has no defined error: left as is)
            if (p.normalise && status != 1 && pe != 0.0) row[i] = rr[i] / pe;
        }
    }
}

// any order up to 255: the same recursion with the arrays in local memory
__global__ void __launch_bounds__(128) lpc_levinson_any(const LvParams p) {
    const long t = (long) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= p.T) return;
    const int P = p.P;
    const int KMAX = P;
    double rr[256], a[256];
    double *row = p.r + t * p.ld;
    for (int i = 0; i <= P; i++) {
        rr[i] = row[i];
        a[i] = 0.0;
    }
    LPC_LEVINSON_BODY(_Pragma("unroll 1"))
    if (p.status) p.status[t] = status;
    if (status != 0 && p.err_count) atomicAdd(p.err_count, 1);
    if (p.pe) p.pe[t] = pe;
    for (int i = 0; i <= P; i++) {
        if (p.a) p.a[t * p.ld + i] = a[i];
        if (p.normalise && status != 1 && pe != 0.0) row[i] = rr[i] / pe;
    }
}
#undef LPC_LEVINSON_BODY

__global__ void lpc_hamming(double *h, int win) {
    const int n = blockIdx.x * blockDim.x + threadIdx.x;
    if (n < win) h[n] = .54 - .46 * cos((n * 2 * M_PI) / (win - 1));   // lpa_on_signal.c:24-31
}

// FP64 DFMA peak probe: 8 independent FMA chains per thread, nothing else -- the measured
// denominator of lpc_autocorr's roofline on the box the numbers are taken on.
__global__ void __launch_bounds__(256) lpc_dfma_probe(double *out, int iters, double a, double b) {
    double c0 = threadIdx.x, c1 = c0 + 1., c2 = c0 + 2., c3 = c0 + 3.;
    double c4 = c0 + 4., c5 = c0 + 5., c6 = c0 + 6., c7 = c0 + 7.;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            c0 = fma(c0, a, b); c1 = fma(c1, a, b); c2 = fma(c2, a, b); c3 = fma(c3, a, b);
            c4 = fma(c4, a, b); c5 = fma(c5, a, b); c6 = fma(c6, a, b); c7 = fma(c7, a, b);
        }
    }
    const double r = ((c0 + c1) + (c2 + c3)) + ((c4 + c5) + (c6 + c7));
    if (r == 123.456) out[0] = r;  // keep the chains alive
}

thread_local std::string g_err;

bool cuda_ok(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return true;
    g_err = std::string(what) + ": " + cudaGetErrorString(e);
    return false;
}

// fast path when the order fits one pass of LAGS and two CTAs of 16 (8, 4) frames fit an SM
template <int LAGS>
bool launch_autocorr_fast(AcParams p, cudaStream_t s, bool *taken) {
    *taken = false;
    if (!p.windowed || p.P >= LAGS || std::getenv("ECOZ2_LPC_GENERIC")) return true;
    int Ls = (p.win + kLanesPerFrame - 1) / kLanesPerFrame;
    if ((Ls & 1) == 0) Ls++;
    int G = 16;
    size_t smem = 0;
    for (; G >= 4; G >>= 1) {
        smem = ((size_t) (G - 1) * p.off + 2 * (size_t) p.win + 2 * kFastPad + 2) * sizeof(double);
        if (smem <= 110 * 1024) break;
    }
    if (G < 4) return true;
    p.frames_per_cta = G;
    p.lane_stride = Ls;
    p.staged = 1;
    // frame t0 = blockIdx * G starts at an even sample (G is even), so alignment of the bases decides
    p.bulk = ((reinterpret_cast<uintptr_t>(p.signal) | reinterpret_cast<uintptr_t>(p.hamming)) & 15) == 0 &&
             !std::getenv("ECOZ2_LPC_NO_BULK");
    // per device (function attributes belong to the context): set on every launch, ~1 us
    if (!cuda_ok(cudaFuncSetAttribute(lpc_autocorr_fast<LAGS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      110 * 1024), "cudaFuncSetAttribute(lpc_autocorr_fast)"))
        return false;
    const long grid = (p.T + G - 1) / G;
    lpc_autocorr_fast<LAGS><<<(unsigned) grid, G * kLanesPerFrame, smem, s>>>(p);
    *taken = true;
    return cuda_ok(cudaGetLastError(), "lpc_autocorr_fast launch");
}

template <int LAGS>
bool launch_autocorr(AcParams p, cudaStream_t s) {
    bool taken = false;
    if (!launch_autocorr_fast<LAGS>(p, s, &taken)) return false;
    if (taken) return true;
    // frames per CTA: as many (32, 16, 8, 4) as the staged stretch of signal allows
    int G = 32;
    size_t smem = 0;
    p.staged = 0;
    for (; G >= 4; G >>= 1) {
        smem = ((size_t) (G - 1) * p.off + 2 * (size_t) p.win) * sizeof(double);
        if (smem <= 200 * 1024) {
            p.staged = 1;
            break;
        }
    }
    if (!p.staged) {
        G = 32;
        smem = 0;
    }
    p.frames_per_cta = G;
    // per device (function attributes belong to the context): set on every launch, ~1 us
    if (!cuda_ok(cudaFuncSetAttribute(lpc_autocorr<LAGS>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                      200 * 1024), "cudaFuncSetAttribute(lpc_autocorr)"))
        return false;
    const long grid = (p.T + G - 1) / G;
    lpc_autocorr<LAGS><<<(unsigned) grid, G * kLanesPerFrame, smem, s>>>(p);
    return cuda_ok(cudaGetLastError(), "lpc_autocorr launch");
}

bool launch_levinson(const LvParams &p, cudaStream_t s) {
    const unsigned grid = (unsigned) ((p.T + 127) / 128);
    if (p.P <= 16) lpc_levinson<16><<<grid, 128, 0, s>>>(p);
    else if (p.P <= 40) lpc_levinson<40><<<grid, 128, 0, s>>>(p);
    else lpc_levinson_any<<<grid, 128, 0, s>>>(p);
    return cuda_ok(cudaGetLastError(), "lpc_levinson launch");
}

}  // namespace

const char *last_error() { return g_err.c_str(); }

double dfma_peak_tflops(int device, double millis) {
    cudaDeviceProp prop;
    if (cudaSetDevice(device) != cudaSuccess || cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 0.0;
    double *out = nullptr;
    if (cudaMalloc(&out, 16) != cudaSuccess) return 0.0;
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    const int blocks = prop.multiProcessorCount * 8;
    int iters = 500;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(e0, 0);
        lpc_dfma_probe<<<blocks, 256>>>(out, iters, 0.999, 0.001);
        cudaEventRecord(e1, 0);
        if (cudaEventSynchronize(e1) != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double tf = 2.0 * 128.0 * iters * 256.0 * blocks / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
        if (ms < millis && iters < (1 << 24)) iters = (int) (iters * (millis / (ms > 0.01 ? ms : 0.01)) * 1.05) + 1;
    }
    cudaEventDestroy(e0);
    cudaEventDestroy(e1);
    cudaFree(out);
    return best;
}

long num_frames(long num_samples, double sample_rate, int window_ms, int offset_ms, int *win_out,
                int *off_out) {
    // lpa_on_signal.c:84-101 (the sample rate is truncated to an integer there)
    const long sr = (long) sample_rate;
    const int win = (int) ((window_ms * sr) / 1000);
    const int off = (int) ((offset_ms * sr) / 1000);
    if (win_out) *win_out = win;
    if (off_out) *off_out = off;
    if (win <= 0 || off <= 0) {
        g_err = "lpa_on_signal: window or offset of zero samples";
        return -1;
    }
    if (win > num_samples) {
        g_err = "lpa_on_signal: signal too short";
        return -1;
    }
    long T = (num_samples - (win - off)) / off;
    if ((T - 1) * off + win > num_samples) T--;
    return T;
}

namespace {
bool run_autocorr(AcParams ap, cudaStream_t s) {
    // exact widths for the orders ECOZ2 is run with (P = 12, 20, 36): no padded lags
    if (ap.P == 12) return launch_autocorr<13>(ap, s);
    if (ap.P <= 15) return launch_autocorr<16>(ap, s);
    if (ap.P == 20) return launch_autocorr<21>(ap, s);
    if (ap.P <= 23) return launch_autocorr<24>(ap, s);
    if (ap.P == 36) return launch_autocorr<37>(ap, s);
    return launch_autocorr<40>(ap, s);   // longer orders: several passes of 40 lags
}
}  // namespace

bool lpa_device(const double *d_signal, long T, int win, int off, int P, double *d_vectors,
                double *d_hamming, int *d_err_count, cudaStream_t stream) {
    if (P < 1 || P > kMaxOrder) {
        g_err = "lpa_on_signal: prediction order out of range (1..255)";
        return false;
    }
    if (T <= 0) return true;
    if (win < 2) {
        g_err = "lpa_on_signal: window shorter than two samples";
        return false;
    }
    lpc_hamming<<<(win + 255) / 256, 256, 0, stream>>>(d_hamming, win);
    AcParams ap{};
    ap.signal = d_signal; ap.hamming = d_hamming; ap.r = d_vectors;
    ap.T = T; ap.win = win; ap.off = off; ap.P = P; ap.ld = P + 1; ap.windowed = 1;
    if (!run_autocorr(ap, stream)) return false;
    LvParams lp{};
    lp.r = d_vectors; lp.T = T; lp.P = P; lp.ld = P + 1; lp.normalise = 1; lp.err_count = d_err_count;
    return launch_levinson(lp, stream);
}

bool lpca_device(const double *d_x, long T, int N, long stride, int P, double *d_r, double *d_rc,
                 double *d_a, double *d_pe, int *d_status, cudaStream_t stream) {
    if (P < 1 || P > kMaxOrder || N < 1 || stride < 1 || stride > 0x7fffffffL) {
        g_err = "lpca: arguments out of range";
        return false;
    }
    if (T <= 0) return true;
    AcParams ap{};
    ap.signal = d_x; ap.hamming = nullptr; ap.r = d_r;
    ap.T = T; ap.win = N; ap.off = (int) stride; ap.P = P; ap.ld = P + 1; ap.windowed = 0;
    if (!run_autocorr(ap, stream)) return false;
    LvParams lp{};
    lp.r = d_r; lp.T = T; lp.P = P; lp.ld = P + 1; lp.normalise = 0;
    lp.rc = d_rc; lp.a = d_a; lp.pe = d_pe; lp.status = d_status;
    return launch_levinson(lp, stream);
}

}  // namespace lpcb200
