// lpc_capi.cpp -- C ABI of the LPC front end (include/ecoz2_lpc_b200.h): the array-level
// equivalent of lpa_on_signal, the drop-in ecoz2_lpca, and the file-level ecoz2_lpc_signals
// (wav -> .prd).  The arithmetic runs in lpc_front.cu; nothing here computes on the CPU.
#include <cuda_runtime.h>
#include <errno.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <algorithm>
#include <string>
#include <vector>

#include "../../include/ecoz2_lpc_b200.h"
#include "lpc_front.h"
#include "vq_engine.h"
#include "vq_io.h"

namespace {

thread_local std::string g_lpc_err;

void set_err(const std::string &m) { g_lpc_err = m; }

bool ck(cudaError_t e, const char *what) {
    if (e == cudaSuccess) return true;
    set_err(std::string(what) + ": " + cudaGetErrorString(e));
    return false;
}

struct DevBuf {
    void *p = nullptr;
    ~DevBuf() { if (p) cudaFree(p); }
    bool alloc(size_t bytes) { return ck(cudaMalloc(&p, bytes ? bytes : 8), "cudaMalloc"); }
    template <typename T> T *as() { return static_cast<T *>(p); }
};

bool have_device() {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess || n <= 0) {
        set_err("no CUDA device: the LPC front end has no CPU path");
        return false;
    }
    return true;
}

}  // namespace

extern "C" {

const char *ecoz2b200_lpc_last_error(void) { return g_lpc_err.c_str(); }

double ecoz2b200_dfma_peak_tflops(int device, double millis) {
    if (!have_device()) return 0.0;
    vqb200::DeviceGuard device_guard_;
    return lpcb200::dfma_peak_tflops(device, millis);
}

long ecoz2b200_lpa_num_frames(long num_samples, double sample_rate, int window_ms, int offset_ms) {
    const long T = lpcb200::num_frames(num_samples, sample_rate, window_ms, offset_ms, nullptr, nullptr);
    if (T < 0) set_err(lpcb200::last_error());
    return T;
}

long ecoz2b200_lpa_on_signal_device(const double *d_samples, long num_samples, double sample_rate, int P,
                                    int window_ms, int offset_ms, double *d_vectors, long capacity_frames,
                                    int *num_errors, void *stream) {
    int win = 0, off = 0;
    const long T = lpcb200::num_frames(num_samples, sample_rate, window_ms, offset_ms, &win, &off);
    if (T < 0) { set_err(lpcb200::last_error()); return -1; }
    if (T > capacity_frames) { set_err("lpa_on_signal: output buffer too small"); return -1; }
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    // scratch (Hamming table + error counter) from the stream-ordered pool: no device-wide
    // synchronisation per call, the block is recycled by the next call on the stream
    static thread_local int pool_ready_for = -1;
    int dev = 0;
    if (cudaGetDevice(&dev) == cudaSuccess && pool_ready_for != dev) {
        // keep freed blocks in the pool across synchronisations (the default hands them back)
        cudaMemPool_t pool;
        if (cudaDeviceGetDefaultMemPool(&pool, dev) == cudaSuccess) {
            unsigned long long keep = 64ull << 20;
            cudaMemPoolSetAttribute(pool, cudaMemPoolAttrReleaseThreshold, &keep);
        }
        pool_ready_for = dev;
    }
    void *scratch = nullptr;
    const size_t ham_bytes = (sizeof(double) * (size_t) win + 15) & ~(size_t) 15;
    if (!ck(cudaMallocAsync(&scratch, ham_bytes + 16, s), "cudaMallocAsync")) return -1;
    double *d_ham = static_cast<double *>(scratch);
    int *d_err = reinterpret_cast<int *>(static_cast<char *>(scratch) + ham_bytes);
    int ne = 0;
    bool ok = ck(cudaMemsetAsync(d_err, 0, sizeof(int), s), "cudaMemsetAsync");
    if (ok && !lpcb200::lpa_device(d_samples, T, win, off, P, d_vectors, d_ham, d_err, s)) {
        set_err(lpcb200::last_error());
        ok = false;
    }
    ok = ok && ck(cudaMemcpyAsync(&ne, d_err, sizeof(int), cudaMemcpyDeviceToHost, s), "cudaMemcpyAsync");
    cudaFreeAsync(scratch, s);
    if (!ck(cudaStreamSynchronize(s), "lpa_on_signal") || !ok) return -1;
    if (num_errors) *num_errors = ne;
    return T;
}

long ecoz2b200_lpa_on_signal(const double *samples, long num_samples, double sample_rate, int P,
                             int window_ms, int offset_ms, double *vectors, long capacity_frames,
                             int *num_errors) {
    if (!have_device()) return -1;
    int win = 0, off = 0;
    const long T = lpcb200::num_frames(num_samples, sample_rate, window_ms, offset_ms, &win, &off);
    if (T < 0) { set_err(lpcb200::last_error()); return -1; }
    if (T > capacity_frames) { set_err("lpa_on_signal: output buffer too small"); return -1; }
    if (T == 0) { if (num_errors) *num_errors = 0; return 0; }
    const long used = (T - 1) * off + win;   // samples the frames actually cover
    DevBuf sig, vec;
    if (!sig.alloc(sizeof(double) * (size_t) used) || !vec.alloc(sizeof(double) * (size_t) T * (P + 1))) return -1;
    if (!ck(cudaMemcpy(sig.p, samples, sizeof(double) * (size_t) used, cudaMemcpyHostToDevice), "H2D")) return -1;
    const long r = ecoz2b200_lpa_on_signal_device(sig.as<double>(), used, sample_rate, P, window_ms, offset_ms,
                                                  vec.as<double>(), T, num_errors, nullptr);
    if (r < 0) return -1;
    if (r != T) { set_err("lpa_on_signal: internal frame count mismatch"); return -1; }
    if (!ck(cudaMemcpy(vectors, vec.p, sizeof(double) * (size_t) T * (P + 1), cudaMemcpyDeviceToHost), "D2H"))
        return -1;
    return T;
}

// drop-in for ecoz2_lpca (reference ecoz2/src/include/ecoz2.h:39-47, ecoz2/src/ecoz2/ecoz2.c)
int ecoz2_lpca(double *x, int N, int P, double *r, double *rc, double *a, double *pe) {
    if (!have_device()) { fprintf(stderr, "ecoz2_lpca: %s\n", g_lpc_err.c_str()); return -1; }
    if (N < 1 || P < 1 || P > lpcb200::kMaxOrder) {
        fprintf(stderr, "ecoz2_lpca: arguments out of range\n");
        return -1;
    }
    const size_t row = sizeof(double) * (size_t) (P + 1);
    DevBuf dx, dr, drc, da, dpe, dst;
    if (!dx.alloc(sizeof(double) * (size_t) N) || !dr.alloc(row) || !drc.alloc(row) || !da.alloc(row) ||
        !dpe.alloc(sizeof(double)) || !dst.alloc(sizeof(int)))
        return -1;
    int status = -1;
    double h_pe = 0;
    std::vector<double> h_rc(P + 1);
    const bool ok =
        ck(cudaMemcpy(dx.p, x, sizeof(double) * (size_t) N, cudaMemcpyHostToDevice), "H2D") &&
        ck(cudaMemset(drc.p, 0, row), "memset") &&
        lpcb200::lpca_device(dx.as<double>(), 1, N, N, P, dr.as<double>(), drc.as<double>(), da.as<double>(),
                             dpe.as<double>(), dst.as<int>(), nullptr) &&
        ck(cudaMemcpy(&status, dst.p, sizeof(int), cudaMemcpyDeviceToHost), "D2H") &&
        ck(cudaMemcpy(r, dr.p, row, cudaMemcpyDeviceToHost), "D2H") &&
        ck(cudaMemcpy(h_rc.data(), drc.p, row, cudaMemcpyDeviceToHost), "D2H") &&
        ck(cudaMemcpy(&h_pe, dpe.p, sizeof(double), cudaMemcpyDeviceToHost), "D2H");
    if (!ok) {
        if (g_lpc_err.empty()) set_err(lpcb200::last_error());
        fprintf(stderr, "ecoz2_lpca: %s\n", g_lpc_err.c_str());
        return -1;
    }
    if (status == 1) return 1;   // lpca.c:23-24 returns before touching rc, a, pe
    // rc[0] is never written by lpca (lpca.c:33); a and pe hold the state at the point of return
    for (int k = 1; k <= P; k++) rc[k] = h_rc[k];
    if (!ck(cudaMemcpy(a, da.p, row, cudaMemcpyDeviceToHost), "D2H")) return -1;
    *pe = h_pe;
    return status;
}

}  // extern "C"
