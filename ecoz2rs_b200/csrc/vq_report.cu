// vq_report.cu -- the per-frame stage report on the device.
//
// reference: report_cbook, ecoz2/src/vq/report.c:109-123 writes "<cbook>.min_dists.csv": one text
// line "%d,%f\n" (codeword, minDistortion) per training vector and stage -- 10M lines per stage at
// configs[1], more host time than the whole LBG run once a pass takes milliseconds.  Here the lines
// are formatted where the values are: one thread per frame computes the exact "%f" digits with
// integer arithmetic (x = m 2^e, round-half-even of m 10^6 2^e in 128 bits: byte for byte what
// glibc prints), a scan turns the line lengths into offsets, a second sweep writes the text; the
// host only copies the finished bytes into the file.
//
//   vq_fmt_lengths   line length per frame (u8) + per-block sums
//   vq_fmt_scan      exclusive scan of the block sums (one CTA), total size, "cannot format" flag
//   vq_fmt_write     the text
#include "vq_kernels.cuh"
#include "vq_engine.h"      // set_last_error, the error codes of the public header

#include <cstdio>
#include <mutex>

namespace vqb200 {

namespace {

constexpr int kFmtBlock = 256;          // frames (threads) per block
constexpr int kFmtMaxLine = 40;         // "-32768," + "-" + 19 digits + "." + 6 digits + "\n" fits

struct U128 {
    unsigned long long hi, lo;
};

// round(|x| * 10^6) as a 128-bit integer for finite |x| < 2^63; false otherwise (inf, nan, huge)
__device__ __forceinline__ bool scaled_micro(double x, U128 &v) {
    const unsigned long long bits = (unsigned long long) __double_as_longlong(x);
    const int ef = (int) ((bits >> 52) & 0x7ff);
    unsigned long long m = bits & ((1ull << 52) - 1);
    if (ef >= 1023 + 62) return false;
    int e;
    if (ef == 0) e = -1074;
    else { m |= 1ull << 52; e = ef - 1075; }
    // m * 10^6 < 2^73
    unsigned long long lo = m * 1000000ull, hi = __umul64hi(m, 1000000ull);
    if (e >= 0) {                         // e <= 9 here: shift left
        hi = (hi << e) | (e ? (lo >> (64 - e)) : 0ull);
        lo <<= e;
    } else if (-e >= 127) {
        hi = lo = 0;
    } else {
        const int sh = -e;                // 1 .. 126: q = v >> sh, round half to even on the exact remainder
        unsigned long long qhi, qlo, rhi, rlo, hhi, hlo;   // quotient, remainder, half
        if (sh >= 64) {
            qhi = 0;
            qlo = sh == 64 ? hi : (hi >> (sh - 64));
            rhi = sh == 64 ? 0ull : (hi & ((1ull << (sh - 64)) - 1));
            rlo = lo;
            hhi = sh == 64 ? 0ull : (1ull << (sh - 65));
            hlo = sh == 64 ? (1ull << 63) : 0ull;
            if (sh == 65) { hhi = 1ull << 0; hlo = 0; }
        } else {
            qhi = hi >> sh;
            qlo = (lo >> sh) | (hi << (64 - sh));
            rhi = 0;
            rlo = lo & ((1ull << sh) - 1);
            hhi = 0;
            hlo = 1ull << (sh - 1);
        }
        const bool gt = rhi > hhi || (rhi == hhi && rlo > hlo);
        const bool eq = rhi == hhi && rlo == hlo;
        if (gt || (eq && (qlo & 1ull))) {
            qlo += 1ull;
            if (qlo == 0ull) qhi += 1ull;
        }
        hi = qhi;
        lo = qlo;
    }
    v.hi = hi;
    v.lo = lo;
    return true;
}

// v / 10^6 (fits 64 bits for |x| < 2^63) and v % 10^6
__device__ __forceinline__ void divmod_micro(const U128 &v, unsigned long long &ip, unsigned &fp) {
    const unsigned long long r2 = v.hi % 1000000ull;            // v.hi < 2^19: quotient part is 0 for |x| < 2^63
    const unsigned long long t1 = (r2 << 32) | (v.lo >> 32);
    const unsigned long long q1 = t1 / 1000000ull, r1 = t1 % 1000000ull;
    const unsigned long long t0 = (r1 << 32) | (v.lo & 0xffffffffull);
    const unsigned long long q0 = t0 / 1000000ull;
    fp = (unsigned) (t0 % 1000000ull);
    ip = (q1 << 32) + q0;
}

__device__ __forceinline__ int dec_digits(unsigned long long v) {
    int n = 1;
    while (v >= 10ull) { v /= 10ull; n++; }
    return n;
}

// length of "%d,%f\n"; 0 = cannot be formatted here (the caller falls back to the host formatter)
__device__ __forceinline__ int line_length(int cw, double md, unsigned long long &ip, unsigned &fp) {
    U128 v;
    if (!scaled_micro(md, v)) return 0;
    divmod_micro(v, ip, fp);
    const unsigned long long acw = cw < 0 ? (unsigned long long) (-(long long) cw) : (unsigned long long) cw;
    return (cw < 0) + dec_digits(acw) + 1 + (__double_as_longlong(md) < 0 ? 1 : 0) + dec_digits(ip) + 1 + 6 + 1;
}

__device__ __forceinline__ char *put_dec(char *p, unsigned long long v, int digits) {
    for (int k = digits - 1; k >= 0; k--) { p[k] = (char) ('0' + (int) (v % 10ull)); v /= 10ull; }
    return p + digits;
}

__global__ void __launch_bounds__(kFmtBlock)
vq_fmt_lengths(const int32_t *__restrict__ cw, const double *__restrict__ md, long n,
               unsigned long long *__restrict__ block_sum, int *__restrict__ bad) {
    const long v = (long) blockIdx.x * kFmtBlock + threadIdx.x;
    int len = 0;
    if (v < n) {
        unsigned long long ip;
        unsigned fp;
        len = line_length(cw[v], md[v], ip, fp);
        if (len == 0) atomicExch(bad, 1);
    }
    // block sum
    __shared__ int s_part[kFmtBlock / 32];
    int s = len;
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) s += __shfl_xor_sync(0xffffffffu, s, off);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = s;
    __syncthreads();
    if (threadIdx.x == 0) {
        int t = 0;
        for (int i = 0; i < kFmtBlock / 32; i++) t += s_part[i];
        block_sum[blockIdx.x] = (unsigned long long) t;
    }
}

// exclusive scan of block_sum[0 .. nb) in place; total -> block_sum[nb]
__global__ void __launch_bounds__(1024) vq_fmt_scan(unsigned long long *__restrict__ block_sum, long nb) {
    __shared__ unsigned long long s_tot[1024];
    const long per = (nb + 1023) / 1024;
    const long lo = threadIdx.x * per;
    const long hi = lo + per < nb ? lo + per : nb;
    unsigned long long t = 0;
    for (long i = lo; i < hi; i++) t += block_sum[i];
    s_tot[threadIdx.x] = t;
    __syncthreads();
    for (int off = 1; off < 1024; off <<= 1) {
        unsigned long long a = 0;
        if ((int) threadIdx.x >= off) a = s_tot[threadIdx.x - off];
        __syncthreads();
        s_tot[threadIdx.x] += a;
        __syncthreads();
    }
    unsigned long long run = s_tot[threadIdx.x] - t;
    for (long i = lo; i < hi; i++) {
        const unsigned long long b = block_sum[i];
        block_sum[i] = run;
        run += b;
    }
    if (threadIdx.x == 1023) block_sum[nb] = s_tot[1023];
}

__global__ void __launch_bounds__(kFmtBlock)
vq_fmt_write(const int32_t *__restrict__ cw, const double *__restrict__ md, long n,
             const unsigned long long *__restrict__ block_off, char *__restrict__ text) {
    const long v = (long) blockIdx.x * kFmtBlock + threadIdx.x;
    unsigned long long ip = 0;
    unsigned fp = 0;
    int c = 0;
    double x = 0.0;
    int len = 0;
    if (v < n) {
        c = cw[v];
        x = md[v];
        len = line_length(c, x, ip, fp);
    }
    // offset inside the block: exclusive scan of the lengths
    __shared__ int s_warp[kFmtBlock / 32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    int incl = len;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int t = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += t;
    }
    if (lane == 31) s_warp[w] = incl;
    __syncthreads();
    int before = 0;
    for (int i = 0; i < w; i++) before += s_warp[i];
    if (len == 0) return;
    char *p = text + block_off[blockIdx.x] + (unsigned long long) (before + incl - len);
    const unsigned long long acw = c < 0 ? (unsigned long long) (-(long long) c) : (unsigned long long) c;
    if (c < 0) *p++ = '-';
    p = put_dec(p, acw, dec_digits(acw));
    *p++ = ',';
    if (__double_as_longlong(x) < 0) *p++ = '-';
    p = put_dec(p, ip, dec_digits(ip));
    *p++ = '.';
    p = put_dec(p, (unsigned long long) fp, 6);
    *p = '\n';
}

}  // namespace

size_t format_scratch_bytes(long n) {
    const long nb = (n + kFmtBlock - 1) / kFmtBlock;
    return sizeof(unsigned long long) * (size_t) (nb + 1) + 16;
}
size_t format_text_capacity(long n) { return (size_t) n * kFmtMaxLine; }

void launch_format_lengths(const int32_t *cw, const double *md, long n, void *scratch, cudaStream_t s) {
    if (n <= 0) return;
    const long nb = (n + kFmtBlock - 1) / kFmtBlock;
    unsigned long long *block_sum = static_cast<unsigned long long *>(scratch);
    int *bad = reinterpret_cast<int *>(block_sum + nb + 1);
    cudaMemsetAsync(bad, 0, sizeof(int), s);
    vq_fmt_lengths<<<(unsigned) nb, kFmtBlock, 0, s>>>(cw, md, n, block_sum, bad);
    vq_fmt_scan<<<1, 1024, 0, s>>>(block_sum, nb);
}
void format_result_ptrs(void *scratch, long n, const unsigned long long **total, const int **bad) {
    const long nb = (n + kFmtBlock - 1) / kFmtBlock;
    const unsigned long long *block_sum = static_cast<const unsigned long long *>(scratch);
    *total = block_sum + nb;
    *bad = reinterpret_cast<const int *>(block_sum + nb + 1);
}
void launch_format_write(const int32_t *cw, const double *md, long n, const void *scratch, char *text,
                         cudaStream_t s) {
    if (n <= 0) return;
    const long nb = (n + kFmtBlock - 1) / kFmtBlock;
    vq_fmt_write<<<(unsigned) nb, kFmtBlock, 0, s>>>(cw, md, n, static_cast<const unsigned long long *>(scratch),
                                                     text);
}

}  // namespace vqb200

// ---------------------------------------------------------------------------------------------
// ecoz2b200_format_lines_noalloc -- the same formatter for callers that must not allocate (code
// under CUDA graph capture, serving loops, harnesses that forbid cudaMalloc).  No counterpart in
// the reference.
//
// The scratch of the three kernels (per-block sums, the total, the "cannot format" flag) lives in
// a file-scope __device__ array: every GPU that loads the library holds these 128.3 KB until the
// process exits.  In HOST code the array's name is only a stand-in for the symbol -- its device
// address comes from cudaGetSymbolAddress, looked up once per device with that device current.
// All calls on one device share the array: calls that overlap on the GPU (different streams, no
// ordering between them) race on it.  Everything a call reads from the array it has written
// itself first (block sums, total) or cleared (the flag), so the previous call's contents do not
// matter.
namespace {
constexpr long kNoAllocMaxLines = ECOZ2B200_NOALLOC_MAX_LINES;
constexpr long kNoAllocMaxBlocks = (kNoAllocMaxLines + vqb200::kFmtBlock - 1) / vqb200::kFmtBlock;
// (blocks + 1) sums + 16 bytes for the flag, rounded up to the array's alignment
constexpr size_t kNoAllocScratchBytes = ((size_t) (kNoAllocMaxBlocks + 1) * 8 + 16 + 255) / 256 * 256;
constexpr int kNoAllocMaxDevices = 64;
}  // namespace

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocScratchBytes];

namespace {

std::mutex g_noalloc_mutex;
void *g_noalloc_ptr[kNoAllocMaxDevices] = {};      // device address of g_noalloc_scratch per device

// *bytes = size of the text, or ~0 when some value could not be formatted
__global__ void vq_fmt_result(const unsigned long long *__restrict__ total, const int *__restrict__ bad,
                              unsigned long long *__restrict__ bytes) {
    *bytes = *bad ? ~0ull : *total;
}

int noalloc_fail(int code, const char *msg) {
    vqb200::set_last_error(msg);
    return code;
}

// device address of the scratch array on the CURRENT device (looked up once per device)
int noalloc_scratch(void **ptr) {
    int dev = -1;
    if (cudaGetDevice(&dev) != cudaSuccess || dev < 0 || dev >= kNoAllocMaxDevices)
        return noalloc_fail(ECOZ2B200_ERR_CUDA, "format_lines_noalloc: no current CUDA device");
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[dev]) {
        void *p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) {
            char msg[160];
            snprintf(msg, sizeof msg, "format_lines_noalloc: cudaGetSymbolAddress failed on device %d: %s", dev,
                     cudaGetErrorString(e));
            return noalloc_fail(ECOZ2B200_ERR_CUDA, msg);
        }
        g_noalloc_ptr[dev] = p;
    }
    *ptr = g_noalloc_ptr[dev];
    return ECOZ2B200_OK;
}

// a pointer the kernels can dereference on device `dev`?  (plain host memory cannot be)
bool noalloc_device_pointer(const void *p, int dev) {
    cudaPointerAttributes a;
    if (cudaPointerGetAttributes(&a, p) != cudaSuccess) { cudaGetLastError(); return false; }
    if (a.type == cudaMemoryTypeDevice) return a.device == dev;
    return a.type == cudaMemoryTypeManaged || a.type == cudaMemoryTypeHost;
}

}  // namespace

extern "C" int ecoz2b200_format_lines_noalloc(const int32_t *d_codeword, const double *d_min_dist, long n,
                                              char *d_text, size_t text_capacity,
                                              unsigned long long *d_bytes, void *stream) {
    char msg[200];
    if (n < 0 || !d_bytes || (n > 0 && (!d_codeword || !d_min_dist || !d_text)))
        return noalloc_fail(ECOZ2B200_ERR_ARG, "format_lines_noalloc: null pointer or negative line count");
    if (n > kNoAllocMaxLines) {
        snprintf(msg, sizeof msg, "format_lines_noalloc: %ld lines exceed the maximum of %ld per call "
                 "(ECOZ2B200_NOALLOC_MAX_LINES; the scratch is a fixed device array)", n, kNoAllocMaxLines);
        return noalloc_fail(ECOZ2B200_ERR_ARG, msg);
    }
    if (text_capacity < vqb200::format_text_capacity(n)) {
        snprintf(msg, sizeof msg, "format_lines_noalloc: text_capacity %zu is below %zu (%d bytes per line)",
                 text_capacity, vqb200::format_text_capacity(n), vqb200::kFmtMaxLine);
        return noalloc_fail(ECOZ2B200_ERR_ARG, msg);
    }
    void *scratch = nullptr;
    const int rc = noalloc_scratch(&scratch);
    if (rc != ECOZ2B200_OK) return rc;
    int dev = 0;
    cudaGetDevice(&dev);
    if (!noalloc_device_pointer(d_bytes, dev) ||
        (n > 0 && (!noalloc_device_pointer(d_codeword, dev) || !noalloc_device_pointer(d_min_dist, dev) ||
                   !noalloc_device_pointer(d_text, dev))))
        return noalloc_fail(ECOZ2B200_ERR_ARG,
                            "format_lines_noalloc: every buffer must be device-accessible memory of the current device");
    cudaStream_t s = static_cast<cudaStream_t>(stream);
    if (n == 0) {
        if (cudaMemsetAsync(d_bytes, 0, sizeof(unsigned long long), s) != cudaSuccess)
            return noalloc_fail(ECOZ2B200_ERR_CUDA, "format_lines_noalloc: cudaMemsetAsync failed");
        return ECOZ2B200_OK;
    }
    const unsigned long long *d_total;
    const int *d_bad;
    vqb200::format_result_ptrs(scratch, n, &d_total, &d_bad);
    vqb200::launch_format_lengths(d_codeword, d_min_dist, n, scratch, s);   // clears the flag first
    vq_fmt_result<<<1, 1, 0, s>>>(d_total, d_bad, d_bytes);
    vqb200::launch_format_write(d_codeword, d_min_dist, n, scratch, d_text, s);
    const cudaError_t e = cudaGetLastError();
    if (e != cudaSuccess) {
        snprintf(msg, sizeof msg, "format_lines_noalloc: launch failed: %s", cudaGetErrorString(e));
        return noalloc_fail(ECOZ2B200_ERR_CUDA, msg);
    }
    return ECOZ2B200_OK;
}
