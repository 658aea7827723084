// vq_models.cu -- quantization of the resident frames against SEVERAL codebooks at once.
//
// reference: cbook_quantize, ecoz2/src/vq/vq.c:154-157 -> quantize, vq/quantize.c:8-34, called once per
// (model, predictor file) by vq_classify.c:265-280 and by the SeqProvider of `hmm classify`
// (vq/vq.c:285-315): R x F serial sweeps.  Here the search of all R codebooks is ONE sweep of the
// tensor-core kernel over the concatenated codebooks (vq_assign_tc2<..., SEG>, vq_tc.cu); this file
// holds what follows it:
//
//   vq_models_recheck      the (model, frame) pairs whose best-two gap is inside the error bound get
//                          the reference's own scan in FP64 (strict <, first index wins)
//   vq_models_distortion   FP64 distortion - 1 of every frame's winner in every model, one read of
//                          the frame for all R models, summed per chunk of frames in a fixed order
//   vq_models_segments     chunk sums -> per (model, file) sums, in chunk order (deterministic)
#include "vq_kernels.cuh"

#include <cstdlib>

namespace vqb200 {

namespace {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// One warp per key.  The frame sits in shared memory (warp-private, factor 2 of distortion.c:9-12
// folded in); lane l evaluates codewords l, l + 32, ... completely (its own row of raas, read
// sequentially: the 32 rows of a step are 9.5 KB that stay in L1), in the reference's order of k;
// first index among equal minima wins, as the strict < of quantize.c:22.
__global__ void __launch_bounds__(256)
vq_models_recheck(const double *__restrict__ x64, int D, const double *__restrict__ raas_all, int K,
                  const unsigned long long *__restrict__ keys, const int32_t *__restrict__ count, int cap,
                  int32_t *__restrict__ idx_out, long idx_stride) {
    __shared__ double s_x[8][kMaxD];
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    int n = *count;
    if (n > cap) n = cap;
    for (long i = warp; i < n; i += n_warps) {
        const unsigned long long key = keys[i];
        const int model = (int) (key >> 40);
        const long row = (long) (key & ((1ull << 40) - 1));
        __syncwarp();
        for (int k = lane; k < D; k += 32) s_x[wib][k] = x64[row * D + k] * (k > 0 ? 2.0 : 1.0);
        __syncwarp();
        const double *w = raas_all + (size_t) model * K * D;
        double best = INFINITY;
        int bidx = 0x7fffffff;
        for (int c0 = 0; c0 < K; c0 += 32) {
            const int c = c0 + lane;
            if (c < K) {
                const double *wc = w + (size_t) c * D;
                double p = 0.0;
                for (int k = 0; k < D; k++) p = fma(s_x[wib][k], wc[k], p);
                if (p < best) { best = p; bidx = c; }       // c increases: first index kept on ties
            }
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, off);
            const int oi = __shfl_xor_sync(0xffffffffu, bidx, off);
            if (ob < best || (ob == best && oi < bidx)) { best = ob; bidx = oi; }
        }
        if (lane == 0) idx_out[(long) model * idx_stride + row] = bidx < K ? bidx : 0;   // all NaN: codeword 0
    }
}

// One warp per chunk of frames.  Lane l holds components l, l + 32, ... of a frame and, for every
// model of a group of 32, the running sum over the chunk's frames of ITS components' products with
// the frame's winner in that model: no cross-lane traffic per frame, 2 R independent loads in
// flight.  One transpose-reduce (31 shuffles) per chunk then leaves lane m with the chunk's sum for
// model m.  The summation order is fixed, so the result is deterministic.
template <int NX>   // components per lane: ceil(D / 32)
__global__ void __launch_bounds__(128)
vq_models_distortion(const double *__restrict__ x64, int D, const double *__restrict__ raas_all, int K, int R,
                     const int32_t *__restrict__ idx_all, long idx_stride, const long *__restrict__ chunks,
                     int num_chunks, double *__restrict__ partial) {
    const int lane = threadIdx.x & 31;
    const int wib = threadIdx.x >> 5;
    __shared__ int32_t s_idx[4][32][33];   // [warp][model of the group][frame of the batch]
    const long warp = (long) blockIdx.x * 4 + wib;
    if (warp >= num_chunks) return;
    const long first = chunks[2 * warp], count = chunks[2 * warp + 1];
    for (int g0 = 0; g0 < R; g0 += 32) {            // groups of 32 models
        const int gm = R - g0 < 32 ? R - g0 : 32;
        double p[32];
#pragma unroll
        for (int m = 0; m < 32; m++) p[m] = 0.0;
        for (long b0 = 0; b0 < count; b0 += 32) {    // batches of 32 frames
            const int bf = count - b0 < 32 ? (int) (count - b0) : 32;
            __syncwarp();
            for (int m = 0; m < gm; m++)             // coalesced: 32 consecutive frames of one model
                s_idx[wib][m][lane] = lane < bf ? idx_all[(long) (g0 + m) * idx_stride + first + b0 + lane] : 0;
            __syncwarp();
            for (int f = 0; f < bf; f++) {
                const long row = first + b0 + f;
                double x[NX];
#pragma unroll
                for (int j = 0; j < NX; j++) {
                    const int k = lane + 32 * j;
                    x[j] = k < D ? x64[row * D + k] * (k > 0 ? 2.0 : 1.0) : 0.0;
                }
#pragma unroll
                for (int m = 0; m < 32; m++) {
                    if (m < gm) {
                        const double *w = raas_all + ((size_t) (g0 + m) * K + s_idx[wib][m][f]) * D;
#pragma unroll
                        for (int j = 0; j < NX; j++) {
                            const int k = lane + 32 * j;
                            if (k < D) p[m] = fma(x[j], w[k], p[m]);
                        }
                    }
                }
            }
        }
        // transpose-reduce: after the step with offset `off`, p[i] (i < off) of a lane holds the sum
        // over the lane pair of the model whose bit `off` equals the lane's bit
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const bool up = (lane & off) != 0;
#pragma unroll
            for (int i = 0; i < off; i++) {
                const double send = up ? p[i] : p[i + off];
                const double keep = up ? p[i + off] : p[i];
                p[i] = keep + __shfl_xor_sync(0xffffffffu, send, off);
            }
        }
        // lane m: model g0 + m; the "- 1" of every frame (quantize.c:27) at once
        if (lane < gm) partial[(size_t) warp * R + g0 + lane] = p[0] - (double) count;
    }
}

// The same sums with the codebooks in SHARED memory: a persistent block of 32 warps per SM takes the
// models in groups of G <= 8 (G K D doubles <= ~200 KB), copies the group's codebooks in once, and
// sweeps its share of the chunks; every codeword operand then comes from shared memory, the frame
// rows straight from global memory (one coalesced row per warp, two frames in flight).  The frames
// are re-read once per group (R / G sweeps of HBM) instead of 296 bytes of L2 traffic per
// (frame, model).  (First version: 8 warps per SM with staged frames -- ncu: 2 warps per scheduler,
// issue slots 23 % busy, 7.3 ms per 10M-frame sweep.)
constexpr int kMdFrames = 16;    // frames whose winners a warp stages at a time
constexpr int kMdMaxG = 8;       // models per group
constexpr int kMdWarps = 32;
template <int NX>
__global__ void __launch_bounds__(kMdWarps * 32, 1)
vq_models_distortion_smem(const double *__restrict__ x64, int D, const double *__restrict__ raas_all, int K, int R,
                          int G, const int32_t *__restrict__ idx_all, long idx_stride,
                          const long *__restrict__ chunks, int num_chunks, double *__restrict__ partial) {
    extern __shared__ __align__(16) unsigned char md_smem[];
    double *s_w = reinterpret_cast<double *>(md_smem);                          // [G K][D]
    int32_t *s_idx = reinterpret_cast<int32_t *>(s_w + (size_t) G * K * D);     // [warps][kMdMaxG][16]
    const int lane = threadIdx.x & 31, wib = threadIdx.x >> 5;
    int32_t *my_idx = s_idx + wib * kMdMaxG * kMdFrames;
    for (int g0 = 0; g0 < R; g0 += G) {
        const int gm = R - g0 < G ? R - g0 : G;
        __syncthreads();
        const size_t w_vals = (size_t) gm * K * D;
        const double *src = raas_all + (size_t) g0 * K * D;
        for (size_t i = threadIdx.x; i < w_vals; i += blockDim.x) s_w[i] = src[i];
        __syncthreads();
        for (long c = (long) blockIdx.x * kMdWarps + wib; c < num_chunks; c += (long) gridDim.x * kMdWarps) {
            const long first = chunks[2 * c], count = chunks[2 * c + 1];
            double p[kMdMaxG];
#pragma unroll
            for (int m = 0; m < kMdMaxG; m++) p[m] = 0.0;
            for (long b0 = 0; b0 < count; b0 += kMdFrames) {
                const int bf = count - b0 < kMdFrames ? (int) (count - b0) : kMdFrames;
                __syncwarp();
                {   // winners of the batch: kMdMaxG x 16 ints, loads before stores
                    int32_t t[(kMdMaxG * kMdFrames) / 32];
#pragma unroll
                    for (int u = 0; u < (kMdMaxG * kMdFrames) / 32; u++) {
                        const int i = 32 * u + lane;
                        const int m = i / kMdFrames, f = i % kMdFrames;
                        t[u] = (m < gm && f < bf) ? idx_all[(long) (g0 + m) * idx_stride + first + b0 + f] : 0;
                    }
#pragma unroll
                    for (int u = 0; u < (kMdMaxG * kMdFrames) / 32; u++) my_idx[32 * u + lane] = t[u];
                }
                __syncwarp();
                const double *xs = x64 + (first + b0) * D;
                for (int f = 0; f < bf; f += 2) {
                    double xa[NX], xb[NX];
                    const bool two = f + 1 < bf;
#pragma unroll
                    for (int j = 0; j < NX; j++) {
                        const int k = lane + 32 * j;
                        xa[j] = k < D ? xs[(size_t) f * D + k] : 0.0;
                        xb[j] = (two && k < D) ? xs[(size_t) (f + 1) * D + k] : 0.0;
                    }
#pragma unroll
                    for (int j = 0; j < NX; j++) {
                        if (lane + 32 * j > 0) { xa[j] *= 2.0; xb[j] *= 2.0; }
                    }
#pragma unroll
                    for (int m = 0; m < kMdMaxG; m++) {
                        if (m < gm) {
                            const double *wa = s_w + ((size_t) m * K + my_idx[m * kMdFrames + f]) * D;
                            const double *wb = s_w + ((size_t) m * K + my_idx[m * kMdFrames + (two ? f + 1 : f)]) * D;
#pragma unroll
                            for (int j = 0; j < NX; j++) {
                                const int k = lane + 32 * j;
                                if (k < D) {
                                    p[m] = fma(xa[j], wa[k], p[m]);
                                    p[m] = fma(xb[j], wb[k], p[m]);
                                }
                            }
                        }
                    }
                }
            }
#pragma unroll
            for (int m = 0; m < kMdMaxG; m++) {
                if (m < gm) {
                    const double t = warp_sum(p[m]);
                    if (lane == 0) partial[(size_t) c * R + g0 + m] = t - (double) count;
                }
            }
        }
    }
}

// sums[seg][m] = sum of partial[c][m] over the chunks of segment seg, in chunk order
__global__ void vq_models_segments(const double *__restrict__ partial, const int *__restrict__ seg_first_chunk,
                                   int num_segments, int R, double *__restrict__ sums) {
    const long t = (long) blockIdx.x * blockDim.x + threadIdx.x;
    if (t >= (long) num_segments * R) return;
    const int seg = (int) (t / R), m = (int) (t % R);
    double s = 0.0;
    for (int c = seg_first_chunk[seg]; c < seg_first_chunk[seg + 1]; c++) s += partial[(size_t) c * R + m];
    sums[(size_t) seg * R + m] = s;
}

// partial[c][m] = sum of md over chunk c (one warp per chunk, fixed order): the serial route, where
// vq_accumulate has already produced the per-frame values of model m
__global__ void __launch_bounds__(256)
vq_chunk_sums(const double *__restrict__ md, const long *__restrict__ chunks, int num_chunks, int R, int m,
              double *__restrict__ partial) {
    const int lane = threadIdx.x & 31;
    const long c = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    if (c >= num_chunks) return;
    const long first = chunks[2 * c], count = chunks[2 * c + 1];
    double t = 0.0;
    for (long i = lane; i < count; i += 32) t += md[first + i];
    t = warp_sum(t);
    if (lane == 0) partial[(size_t) c * R + m] = t;
}

__global__ void vq_idx_to_u16(const int32_t *__restrict__ idx, uint16_t *__restrict__ seq, size_t n) {
    for (size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x; i < n; i += (size_t) gridDim.x * blockDim.x)
        seq[i] = (uint16_t) idx[i];
}

}  // namespace

void launch_idx_to_u16(const int32_t *idx, uint16_t *seq, size_t n, cudaStream_t s) {
    if (n == 0) return;
    size_t blocks = (n + 255) / 256;
    if (blocks > 148 * 16) blocks = 148 * 16;
    vq_idx_to_u16<<<(unsigned) blocks, 256, 0, s>>>(idx, seq, n);
}

void launch_chunk_sums(const double *md, const long *chunks, int num_chunks, int R, int m, double *partial,
                       cudaStream_t s) {
    if (num_chunks <= 0) return;
    vq_chunk_sums<<<(unsigned) ((num_chunks + 7) / 8), 256, 0, s>>>(md, chunks, num_chunks, R, m, partial);
}

void launch_models_recheck(const double *x64, int D, const double *raas_all, int K, const unsigned long long *keys,
                           const int32_t *count, int cap, int32_t *idx_out, long idx_stride, int num_sms,
                           cudaStream_t s) {
    if (cap <= 0) return;
    vq_models_recheck<<<num_sms * 8, 256, 0, s>>>(x64, D, raas_all, K, keys, count, cap, idx_out, idx_stride);
}

template <int NX>
static void launch_md_smem(const double *x64, int D, const double *raas_all, int K, int R, int G, const int32_t *idx_all,
                           long idx_stride, const long *chunks, int num_chunks, double *partial, int num_sms,
                           size_t smem, cudaStream_t s) {
    cudaFuncSetAttribute(vq_models_distortion_smem<NX>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    int blocks = (num_chunks + kMdWarps - 1) / kMdWarps;
    if (blocks > num_sms) blocks = num_sms;
    vq_models_distortion_smem<NX><<<blocks, kMdWarps * 32, smem, s>>>(x64, D, raas_all, K, R, G, idx_all, idx_stride, chunks,
                                                           num_chunks, partial);
}

void launch_models_distortion(const double *x64, int D, const double *raas_all, int K, int R, const int32_t *idx_all,
                              long idx_stride, const long *chunks, int num_chunks, double *partial, int num_sms,
                              cudaStream_t s) {
    if (num_chunks <= 0 || R <= 0) return;
    // codebooks in shared memory when at least one fits next to the staging buffers
    {
        const size_t stage = (size_t) kMdWarps * kMdMaxG * kMdFrames * sizeof(int32_t);
        const size_t budget = 227 * 1024 - stage - 1024;
        const size_t per_model = (size_t) K * D * sizeof(double);
        int G = (int) (budget / per_model);
        if (G > kMdMaxG) G = kMdMaxG;
        if (G > R) G = R;
        static int use = -1;
        if (use < 0) { const char *e = getenv("ECOZ2_VQ_MODELS_L2"); use = (e && e[0] == '1') ? 0 : 1; }
        if (use && G >= 1 && D <= 64) {
            const size_t smem = (size_t) G * per_model + stage;
            if (D <= 32) launch_md_smem<1>(x64, D, raas_all, K, R, G, idx_all, idx_stride, chunks, num_chunks, partial, num_sms, smem, s);
            else launch_md_smem<2>(x64, D, raas_all, K, R, G, idx_all, idx_stride, chunks, num_chunks, partial, num_sms, smem, s);
            return;
        }
    }
    const unsigned blocks = (unsigned) ((num_chunks + 3) / 4);
    const int nx = (D + 31) / 32;
    switch (nx) {
        case 1: vq_models_distortion<1><<<blocks, 128, 0, s>>>(x64, D, raas_all, K, R, idx_all, idx_stride, chunks, num_chunks, partial); break;
        case 2: vq_models_distortion<2><<<blocks, 128, 0, s>>>(x64, D, raas_all, K, R, idx_all, idx_stride, chunks, num_chunks, partial); break;
        case 3: vq_models_distortion<3><<<blocks, 128, 0, s>>>(x64, D, raas_all, K, R, idx_all, idx_stride, chunks, num_chunks, partial); break;
        case 4: vq_models_distortion<4><<<blocks, 128, 0, s>>>(x64, D, raas_all, K, R, idx_all, idx_stride, chunks, num_chunks, partial); break;
        default: vq_models_distortion<kMaxD / 32 + 1><<<blocks, 128, 0, s>>>(x64, D, raas_all, K, R, idx_all, idx_stride, chunks, num_chunks, partial); break;
    }
}

void launch_models_segments(const double *partial, const int *seg_first_chunk, int num_segments, int R, double *sums,
                            cudaStream_t s) {
    const long items = (long) num_segments * R;
    if (items <= 0) return;
    vq_models_segments<<<(unsigned) ((items + 255) / 256), 256, 0, s>>>(partial, seg_first_chunk, num_segments, R, sums);
}

}  // namespace vqb200
