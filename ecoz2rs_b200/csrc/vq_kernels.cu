// vq_kernels.cu -- hand-written sm_100a kernels of the ECOZ2 VQ hot path.
//
// Not a port of the reference's C loops: the reference (ecoz2/src/vq) walks
// frame by frame and codeword by codeword on the CPU in FP64.  Here the search
// is a skinny GEMM with a fused row-argmin epilogue (the N x K distortion
// matrix never exists), the cell sums are run-length aggregated FP64 reductions
// and the centroid / split step is one warp per codeword.
//
// Kernels:
//   vq_assign_tiled<T,...>   (a) fused distortion + argmin (FP32 main pass with
//                                ambiguity flagging, FP64 exact pass / re-evaluation)
//   vq_accumulate<...>       (b) cell sums, counts, distortion, symbols
//   vq_centroids, vq_grow,   (c) Levinson-Durbin centroid, split, step-up
//   vq_refl_to_raas
//   vq_sigma_sum, vq_prepare_codebook, vq_convert_frames
#include "vq_kernels.cuh"

#include <cfloat>
#include <cmath>
#include <type_traits>

namespace vqb200 {

// --------------------------------------------------------------------------------
// small PTX helpers: mbarrier + 1-D bulk async copy (TMA engine, SASS UBLKCP)
// --------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void *p) {
    return (uint32_t) __cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ bool mbar_try_wait(uint64_t *bar, uint32_t parity) {
    uint32_t ok;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(ok)
        : "r"(smem_u32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// global -> shared bulk copy, completion signalled on an mbarrier (bytes % 16 == 0)
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                         uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(smem_u32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

template <typename T>
__device__ __forceinline__ T fma_t(T a, T b, T c);
template <>
__device__ __forceinline__ float fma_t<float>(float a, float b, float c) {
    return fmaf(a, b, c);
}
template <>
__device__ __forceinline__ double fma_t<double>(double a, double b, double c) {
    return fma(a, b, c);
}

__device__ __forceinline__ void ld4(const float *p, float (&v)[4]) {
    const float4 t = *reinterpret_cast<const float4 *>(p);
    v[0] = t.x; v[1] = t.y; v[2] = t.z; v[3] = t.w;
}
__device__ __forceinline__ void ld4(const double *p, double (&v)[4]) {
    const double2 a = *reinterpret_cast<const double2 *>(p);
    const double2 b = *reinterpret_cast<const double2 *>(p + 2);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y;
}

template <typename T> struct Lim;
template <> struct Lim<float> { static __device__ __forceinline__ float inf() { return INFINITY; } };
template <> struct Lim<double> { static __device__ __forceinline__ double inf() { return (double) INFINITY; } };

// --------------------------------------------------------------------------------
// (a) fused distortion + argmin
//
// CTA = 256 threads as a 16 x 16 grid; CTA tile = BM frames x BN codewords with
// BM = 16*4*RM, BN = 16*4*RN; every thread owns a (4*RM) x (4*RN) register tile.
// The frame tile is transposed into shared memory once (Xs[n][frame]) and stays
// there while ALL codeword tiles stream past it: codeword tiles are
// pre-transposed in global memory ([tile][n][BN], built by vq_prepare_codebook)
// so one cp.async.bulk brings a whole tile, double buffered on two mbarriers.
// Inner step n: 2*RM + 2*RN 128-bit shared loads feed (4RM)(4RN) FMAs.
// Epilogue per tile: running (best, second, argbest) per frame in registers;
// after the last tile the 16 threads that share a frame merge with shuffles.
//
// Reference semantics restated: strict '<' scan in index order
// (vq_learn_par.c:66-75, quantize.c:19-26) == lexicographic min of (dd, index).
// --------------------------------------------------------------------------------
enum { kModeMain = 0, kModeExact = 1, kModeFilter = 2 };

template <typename T>
struct AssignParams {
    const T *x;                 // frames [n][D] in the search precision
    long n_frames;
    int D_rt;
    const T *wt;                // codeword tiles [tile][D][BN]
    int K;
    const int32_t *list;        // optional gather list (+ device-side count)
    const int32_t *list_count;
    const float *w1max;         // max_i sum_n |w_i[n]|
    float err_scale;            // kappa * 2^-24
    int32_t *idx_out;
    int32_t *flag_list;         // kModeMain: frames to re-evaluate ...
    int32_t *flag_count;
    float *flag_ub;             // ... with an upper bound of their true minimum,
    unsigned long long *min_key;  //   and their slot of the FP64 minimum (set to "max")
    const float *w1;            // sum_n |w_i[n]| per codeword
    // kModeFilter outputs: (list position, codeword) candidates, or -- when the
    // candidate buffer is full -- the frames of the tile for the all-FP64 search
    int2 *cand;
    int32_t *cand_count;
    int cand_cap;
    int32_t *exact_list;
    int32_t *exact_count;
};

constexpr int kCandBuf = 4096;  // candidates staged per CTA frame tile (shared memory)

// MODE kModeMain  : T = float. argmin + runner-up; frames whose best-two gap is inside
//                   the FP32 error bound go to flag_list.
// MODE kModeExact : plain (FP64) argmin, optional gather list.
// MODE kModeFilter: T = float over the gather list: the FP32 values only SELECT the
//                   candidates (d32_j - c W1_j <= ub, c = err_scale |rx0|, ub = upper bound
//                   of the true minimum); vq_candidates_eval / _select then evaluate every
//                   candidate in FP64 and let the FP64 values decide.
template <typename T, int RM, int RN, int DT, int MODE>
__global__ void __launch_bounds__(256, 2)
vq_assign_tiled(const AssignParams<T> p) {
    constexpr int TM = 4 * RM, TN = 4 * RN;
    constexpr int BM = 16 * TM, BN = 16 * TN;
    constexpr int XS = BM + 4;  // padded row of the transposed frame tile
    constexpr bool TRACK2 = MODE == kModeMain;
    constexpr bool FILTER = MODE == kModeFilter;
    using R = T;
    const int D = DT > 0 ? DT : p.D_rt;
    const T *__restrict__ x = p.x;
    const T *__restrict__ wt = p.wt;
    const int32_t *__restrict__ list = p.list;
    const int K = p.K;

    extern __shared__ __align__(128) unsigned char smem_raw[];
    // layout: [2 mbarriers | pad to 128][Ws0: D*BN][Ws1: D*BN][Xs: D*XS]
    uint64_t *full = reinterpret_cast<uint64_t *>(smem_raw);
    T *Ws0 = reinterpret_cast<T *>(smem_raw + 128);
    T *Ws1 = Ws0 + (size_t) D * BN;
    T *Xs = Ws1 + (size_t) D * BN;
    // row D of Xs: the frame's norm max_k |x[k]| the error bound scales with (for an autocorrelation
    // that is |rx[0]|, but frames_append accepts arbitrary rows)
    // kModeFilter only: [Xs][int2 cbuf[kCandBuf]][int s_count, s_base]
    int2 *cbuf = reinterpret_cast<int2 *>(Xs + (size_t) (D + 1) * XS);
    int *s_ctl = reinterpret_cast<int *>(cbuf + kCandBuf);

    const int tid = threadIdx.x;
    const int tx = tid & 15, ty = tid >> 4;
    const uint32_t tile_bytes = (uint32_t) ((size_t) D * BN * sizeof(T));
    const int n_wtiles = (K + BN - 1) / BN;

    if (tid == 0) {
        mbar_init(&full[0], 1);
        mbar_init(&full[1], 1);
        fence_barrier_init();
        if (FILTER) s_ctl[0] = 0;
    }
    __syncthreads();
    uint32_t phase0 = 0, phase1 = 0;

    const long total = list ? (long) *p.list_count : p.n_frames;
    const long n_ftiles = (total + BM - 1) / BM;
    float w1max = 0.f;
    if (TRACK2 || FILTER) w1max = *p.w1max;

    for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x) {
        const long base = ft * BM;
        const int nvalid = (int) (total - base < BM ? total - base : BM);
        __syncthreads();  // previous tile's epilogue is done with Xs

        // -- stage + transpose the frame tile: Xs[n][f]
        if (list == nullptr) {
            const T *src = x + base * D;
            const int count = nvalid * D;
            for (int e = tid; e < BM * D; e += 256) {
                const int f = e / D, n = e - f * D;
                Xs[n * XS + f] = e < count ? src[e] : (T) 0;
            }
        } else {
            for (int e = tid; e < BM * D; e += 256) {
                const int f = e / D, n = e - f * D;
                T v = (T) 0;
                if (f < nvalid) v = x[(long) list[base + f] * D + n];
                Xs[n * XS + f] = v;
            }
        }
        if (tid == 0) {
            mbar_expect_tx(&full[0], tile_bytes);
            bulk_g2s(Ws0, wt, tile_bytes, &full[0]);
            if (n_wtiles > 1) {
                mbar_expect_tx(&full[1], tile_bytes);
                bulk_g2s(Ws1, wt + (size_t) D * BN, tile_bytes, &full[1]);
            }
        }
        __syncthreads();
        if (TRACK2 || FILTER) {
            for (int f = tid; f < BM; f += 256) {
                T m = (T) 0;
                for (int n = 0; n < D; n++) m = max(m, (T) fabs((double) Xs[n * XS + f]));
                Xs[D * XS + f] = m;
            }
            __syncthreads();
        }

        R best[TM];
        T second[TRACK2 ? TM : 1];
        int bidx[TM];
        float cand_c[FILTER ? TM : 1], cand_b[FILTER ? TM : 1];  // c_i and best32_i + c_i W1max
#pragma unroll
        for (int i = 0; i < TM; i++) {
            best[i] = Lim<R>::inf();
            if (TRACK2) second[i] = Lim<T>::inf();
            bidx[i] = 0;
            if (FILTER) {
                const int f = (i >> 2) * (BM / RM) + ty * 4 + (i & 3);
                const float c = p.err_scale * (float) Xs[D * XS + f];
                cand_c[i] = c;
                // frames beyond the list never produce a candidate
                cand_b[i] = f < nvalid ? p.flag_ub[base + f] : -INFINITY;
            }
        }

        for (int t = 0; t < n_wtiles; t++) {
            const T *Ws = (t & 1) ? Ws1 : Ws0;
            if (t & 1) { mbar_wait(&full[1], phase1); phase1 ^= 1; }
            else       { mbar_wait(&full[0], phase0); phase0 ^= 1; }

            T acc[TM][TN];
#pragma unroll
            for (int i = 0; i < TM; i++)
#pragma unroll
                for (int j = 0; j < TN; j++) acc[i][j] = (T) 0;

#pragma unroll 4
            for (int n = 0; n < D; n++) {
                T xa[TM], wb[TN];
#pragma unroll
                for (int r = 0; r < RM; r++) {
                    T v[4];
                    ld4(Xs + n * XS + r * (BM / RM) + ty * 4, v);
#pragma unroll
                    for (int q = 0; q < 4; q++) xa[r * 4 + q] = v[q];
                }
#pragma unroll
                for (int r = 0; r < RN; r++) {
                    T v[4];
                    ld4(Ws + n * BN + r * (BN / RN) + tx * 4, v);
#pragma unroll
                    for (int q = 0; q < 4; q++) wb[r * 4 + q] = v[q];
                }
#pragma unroll
                for (int i = 0; i < TM; i++)
#pragma unroll
                    for (int j = 0; j < TN; j++) acc[i][j] = fma_t<T>(xa[i], wb[j], acc[i][j]);
            }

            // -- fused argmin epilogue for this codeword tile (index order ascending in j)
            const int c0 = t * BN + tx * 4;
            const bool partial = (t + 1) * BN > K;
#pragma unroll
            for (int j = 0; j < TN; j++) {
                const int c = c0 + (j >> 2) * (BN / RN) + (j & 3);
                if (partial && c >= K) continue;
                if (FILTER) {
                    const float w1j = p.w1[c];
#pragma unroll
                    for (int i = 0; i < TM; i++) {
                        if ((float) acc[i][j] <= fmaf(cand_c[i], w1j, cand_b[i])) {
                            const int f = (i >> 2) * (BM / RM) + ty * 4 + (i & 3);
                            const int slot = atomicAdd(&s_ctl[0], 1);
                            if (slot < kCandBuf) cbuf[slot] = make_int2((int) (base + f), c);
                        }
                    }
                } else {
#pragma unroll
                    for (int i = 0; i < TM; i++) {
                        const T d = acc[i][j];
                        if (TRACK2) second[i] = min(second[i], max(d, (T) best[i]));
                        bidx[i] = d < (T) best[i] ? c : bidx[i];
                        best[i] = min((T) best[i], d);
                    }
                }
            }

            __syncthreads();  // everyone is done reading this buffer
            if (tid == 0 && t + 2 < n_wtiles) {
                uint64_t *bar = (t & 1) ? &full[1] : &full[0];
                mbar_expect_tx(bar, tile_bytes);
                bulk_g2s((t & 1) ? Ws1 : Ws0, wt + (size_t) (t + 2) * D * BN, tile_bytes, bar);
            }
        }

        if (FILTER) {
            // publish this tile's candidates (all or nothing); a full buffer sends the
            // tile's frames to the all-FP64 search instead
            __syncthreads();
            const int m = s_ctl[0];
            if (tid == 0) {
                int gbase = -1;
                if (m <= kCandBuf) {
                    const int g = atomicAdd(p.cand_count, m);
                    if (g + m <= p.cand_cap) gbase = g;
                }
                if (gbase < 0) gbase = -1 - atomicAdd(p.exact_count, nvalid);
                s_ctl[1] = gbase;
            }
            __syncthreads();
            const int gbase = s_ctl[1];
            if (gbase >= 0) {
                for (int e = tid; e < m; e += 256) p.cand[gbase + e] = cbuf[e];
            } else {
                const int ebase = -1 - gbase;
                for (int e = tid; e < nvalid; e += 256) p.exact_list[ebase + e] = list[base + e];
            }
            __syncthreads();
            if (tid == 0) s_ctl[0] = 0;
            continue;
        }

        // -- merge the 16 threads (tx = low 4 lane bits) that share each frame
#pragma unroll
        for (int i = 0; i < TM; i++) {
#pragma unroll
            for (int off = 8; off >= 1; off >>= 1) {
                const R ob = __shfl_xor_sync(0xffffffffu, best[i], off);
                const int oi = __shfl_xor_sync(0xffffffffu, bidx[i], off);
                if (TRACK2) {
                    const T os = __shfl_xor_sync(0xffffffffu, second[i], off);
                    second[i] = min(min(second[i], os), max((T) best[i], (T) ob));
                }
                if (ob < best[i] || (ob == best[i] && oi < bidx[i])) {
                    best[i] = ob;
                    bidx[i] = oi;
                }
            }
        }
        if (tx == 0) {
            const float bound = 2.f * p.err_scale * w1max;
#pragma unroll
            for (int i = 0; i < TM; i++) {
                const int f = (i >> 2) * (BM / RM) + ty * 4 + (i & 3);
                if (f < nvalid) {
                    const long row = list ? (long) list[base + f] : base + f;
                    p.idx_out[row] = bidx[i];
                    if (TRACK2) {
                        const float x0 = (float) Xs[D * XS + f];
                        // NaN-safe: "not clearly separated" => re-evaluate
                        if (!((float) (second[i] - (T) best[i]) >= bound * x0)) {
                            const int pos = atomicAdd(p.flag_count, 1);
                            p.flag_list[pos] = (int32_t) row;
                            // true minimum <= d64(best) <= best32 + (2 err_scale) |rx0| W1_best
                            p.flag_ub[pos] = fmaf(2.f * p.err_scale * x0, p.w1[bidx[i]], (float) best[i]);
                            p.min_key[pos] = ~0ull;
                        }
                    }
                }
            }
        }
    }
}

template <typename T, int RM, int RN>
static size_t assign_smem_bytes(int D, bool filter) {
    constexpr int BM = 64 * RM, BN = 64 * RN;
    return 128 + sizeof(T) * ((size_t) 2 * D * BN + (size_t) (D + 1) * (BM + 4)) +
           (filter ? sizeof(int2) * kCandBuf + 16 : 0);
}

template <typename T, int RM, int RN, int DT, int MODE>
static void launch_assign_inst(const AssignParams<T> &p, int num_sms, cudaStream_t s) {
    auto kern = vq_assign_tiled<T, RM, RN, DT, MODE>;
    const size_t smem = assign_smem_bytes<T, RM, RN>(p.D_rt, MODE == kModeFilter);
    cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
    constexpr int BM = 64 * RM;
    long tiles = (p.n_frames + BM - 1) / BM;
    if (tiles < 1) tiles = 1;
    long grid = tiles;
    if (p.list != nullptr) {
        // device-side count: persistent grid, CTAs loop over however many tiles there are
        const long cap = (long) num_sms * 4;
        if (grid > cap) grid = cap;
    }
    kern<<<(unsigned) grid, 256, smem, s>>>(p);
}

template <typename T, int RM, int RN, int MODE>
static void launch_assign_d(const AssignParams<T> &p, int num_sms, cudaStream_t s) {
    if (p.D_rt == 37) launch_assign_inst<T, RM, RN, 37, MODE>(p, num_sms, s);
    else if (p.D_rt == 13) launch_assign_inst<T, RM, RN, 13, MODE>(p, num_sms, s);
    else launch_assign_inst<T, RM, RN, 0, MODE>(p, num_sms, s);
}

static size_t align256(size_t v) { return (v + 255) & ~(size_t) 255; }
size_t scratch_bytes(long cap) {
    const long cand_cap = cap < 4096 ? 4096 : cap;
    return align256(sizeof(int32_t) * cap) * 2 + align256(sizeof(float) * cap) +
           align256(sizeof(unsigned long long) * cap) + align256(sizeof(int2) * cand_cap) +
           align256(sizeof(double) * cand_cap) + 256;
}
void scratch_carve(void *base, long cap, SearchScratch *out) {
    const long cand_cap = cap < 4096 ? 4096 : cap;
    char *p = static_cast<char *>(base);
    out->counters = reinterpret_cast<int32_t *>(p); p += 256;
    out->min_key = reinterpret_cast<unsigned long long *>(p); p += align256(sizeof(unsigned long long) * cap);
    out->cand_d = reinterpret_cast<double *>(p); p += align256(sizeof(double) * cand_cap);
    out->cand = reinterpret_cast<int2 *>(p); p += align256(sizeof(int2) * cand_cap);
    out->flag_list = reinterpret_cast<int32_t *>(p); p += align256(sizeof(int32_t) * cap);
    out->exact_list = reinterpret_cast<int32_t *>(p); p += align256(sizeof(int32_t) * cap);
    out->flag_ub = reinterpret_cast<float *>(p);
    out->cap = cap;
    out->cand_cap = cand_cap;
}

void launch_assign_f32(const float *x32, long n, int D, const float *w32, int K,
                       const SearchScratch &sc, const float *w1, const float *w1max,
                       float err_scale, int32_t *idx, int num_sms, cudaStream_t s) {
    if (n <= 0) return;
    AssignParams<float> p{};
    p.x = x32; p.n_frames = n; p.D_rt = D; p.wt = w32; p.K = K;
    p.w1 = w1; p.w1max = w1max; p.err_scale = err_scale; p.idx_out = idx;
    p.flag_list = sc.flag_list; p.flag_count = sc.counters + 0; p.flag_ub = sc.flag_ub;
    p.min_key = sc.min_key;
    launch_assign_d<float, 2, 2, kModeMain>(p, num_sms, s);
}

void launch_assign_f64(const double *x64, long n, int D, const double *w64, int K,
                       const int32_t *list, const int32_t *list_count, int32_t *idx, int num_sms,
                       cudaStream_t s) {
    if (n <= 0) return;
    AssignParams<double> p{};
    p.x = x64; p.n_frames = n; p.D_rt = D; p.wt = w64; p.K = K;
    p.list = list; p.list_count = list_count; p.idx_out = idx;
    launch_assign_d<double, 1, 1, kModeExact>(p, num_sms, s);
}

// --------------------------------------------------------------------------------
// FP64 evaluation of the candidates: one warp per (frame, codeword) pair, lanes over
// the D components (coalesced 8 D-byte rows), fixed shuffle tree => bit-identical
// values for identical codewords, so exact ties resolve to the lower index.
// --------------------------------------------------------------------------------
__device__ __forceinline__ unsigned long long order_key(double d) {
    const unsigned long long b = (unsigned long long) __double_as_longlong(d);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);   // monotone in d
}

__global__ void __launch_bounds__(256)
vq_candidates_eval(const double *__restrict__ x64, const double *__restrict__ raas, int D,
                   const int32_t *__restrict__ list, const int2 *__restrict__ cand,
                   const int32_t *__restrict__ cand_count, int cand_cap,
                   double *__restrict__ cand_d, unsigned long long *__restrict__ min_key,
                   int32_t *__restrict__ idx_out) {
    const int lane = threadIdx.x & 31;
    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    long total = *cand_count;
    if (total > cand_cap) total = cand_cap;
    for (long k = warp; k < total; k += n_warps) {
        const int2 cd = cand[k];
        const long row = list[cd.x];
        const double *xr = x64 + row * D, *rr = raas + (long) cd.y * D;
        double part = 0.0;
        for (int m = lane; m < D; m += 32) {
            const double w = m > 0 ? 2.0 * rr[m] : rr[m];
            part = fma(xr[m], w, part);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
        if (lane == 0) {
            cand_d[k] = part;
            atomicMin(min_key + cd.x, order_key(part));
            idx_out[row] = 0x7fffffff;   // every candidate of the row writes the same value
        }
    }
}

__global__ void __launch_bounds__(256)
vq_candidates_select(const int32_t *__restrict__ list, const int2 *__restrict__ cand,
                     const int32_t *__restrict__ cand_count, int cand_cap,
                     const double *__restrict__ cand_d,
                     const unsigned long long *__restrict__ min_key, int32_t *__restrict__ idx_out) {
    long total = *cand_count;
    if (total > cand_cap) total = cand_cap;
    const long stride = (long) gridDim.x * blockDim.x;
    for (long k = (long) blockIdx.x * blockDim.x + threadIdx.x; k < total; k += stride) {
        const int2 cd = cand[k];
        if (order_key(cand_d[k]) == min_key[cd.x]) atomicMin(idx_out + list[cd.x], cd.y);
    }
}

void launch_recheck(const float *x32, const double *x64, long n, int D, const float *w32,
                    const double *w64, const float *w1, const float *w1max, const double *raas,
                    int K, float err_scale, const SearchScratch &sc, int32_t *idx, int num_sms,
                    cudaStream_t s, const float *wtc, float err_tc) {
    if (n <= 0) return;
    if (wtc != nullptr && filter_tc_supported()) {
        // candidates selected by the tensor-core kernel over the flagged frames
        launch_filter_tc(x32, n, D, wtc, K, sc, err_tc, num_sms, s);
        const int blocks = num_sms * 8;
        vq_candidates_eval<<<blocks, 256, 0, s>>>(x64, raas, D, sc.flag_list, sc.cand, sc.counters + 1,
                                                  (int) sc.cand_cap, sc.cand_d, sc.min_key, idx);
        vq_candidates_select<<<blocks, 256, 0, s>>>(sc.flag_list, sc.cand, sc.counters + 1,
                                                    (int) sc.cand_cap, sc.cand_d, sc.min_key, idx);
        launch_assign_f64(x64, n, D, w64, K, sc.exact_list, sc.counters + 2, idx, num_sms, s);
        return;
    }
    AssignParams<float> p{};
    p.x = x32; p.n_frames = n; p.D_rt = D; p.wt = w32; p.K = K;
    p.list = sc.flag_list; p.list_count = sc.counters + 0; p.w1max = w1max; p.err_scale = err_scale;
    p.idx_out = idx; p.flag_ub = sc.flag_ub; p.w1 = w1;
    p.cand = sc.cand; p.cand_count = sc.counters + 1; p.cand_cap = (int) sc.cand_cap;
    p.exact_list = sc.exact_list; p.exact_count = sc.counters + 2;
    launch_assign_d<float, 2, 2, kModeFilter>(p, num_sms, s);
    const int blocks = num_sms * 8;
    vq_candidates_eval<<<blocks, 256, 0, s>>>(x64, raas, D, sc.flag_list, sc.cand, sc.counters + 1,
                                              (int) sc.cand_cap, sc.cand_d, sc.min_key, idx);
    vq_candidates_select<<<blocks, 256, 0, s>>>(sc.flag_list, sc.cand, sc.counters + 1,
                                                (int) sc.cand_cap, sc.cand_d, sc.min_key, idx);
    // tiles whose candidates did not fit: the all-FP64 search (normally an empty list)
    launch_assign_f64(x64, n, D, w64, K, sc.exact_list, sc.counters + 2, idx, num_sms, s);
}

// --------------------------------------------------------------------------------
// data staging
// --------------------------------------------------------------------------------
__global__ void vq_convert_frames(const double *__restrict__ x64, float *__restrict__ x32,
                                  long total) {
    const long stride = (long) gridDim.x * blockDim.x;
    for (long e = (long) blockIdx.x * blockDim.x + threadIdx.x; e < total; e += stride)
        x32[e] = (float) x64[e];
}
void launch_convert_frames(const double *x64, float *x32, long n, int D, cudaStream_t s) {
    const long total = n * D;
    if (total <= 0) return;
    long blocks = (total + 255) / 256;
    if (blocks > 148L * 32) blocks = 148L * 32;
    vq_convert_frames<<<(unsigned) blocks, 256, 0, s>>>(x64, x32, total);
}

// one thread per (codeword slot, n) of the padded tile layouts
__global__ void vq_prepare_codebook(const double *__restrict__ raas, int K, int D,
                                    float *__restrict__ w32, double *__restrict__ w64,
                                    float *__restrict__ w1, float *__restrict__ w1max) {
    const int K32 = (K + kTileF32 - 1) / kTileF32 * kTileF32;
    const long total = (long) K32 * D;
    const long e = (long) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= total) return;
    const int i = (int) (e / D), n = (int) (e - (long) i * D);
    double w = 0.0;
    if (i < K) {
        w = raas[(long) i * D + n];
        if (n > 0) w *= 2.0;
    }
    w32[((long) (i / kTileF32) * D + n) * kTileF32 + (i % kTileF32)] = (float) w;
    const int K64 = (K + kTileF64 - 1) / kTileF64 * kTileF64;
    if (i < K64) w64[((long) (i / kTileF64) * D + n) * kTileF64 + (i % kTileF64)] = w;
    if (n == 0 && i < K) {
        float s1 = 0.f;
        for (int m = 0; m < D; m++) {
            double v = raas[(long) i * D + m];
            if (m > 0) v *= 2.0;
            s1 += fabsf((float) v);
        }
        s1 *= 1.0001f;  // cover the rounding of this sum itself
        w1[i] = s1;
        atomicMax(reinterpret_cast<int *>(w1max), __float_as_int(s1));  // s1 >= 0
    }
}
void launch_prepare_codebook(const double *raas, int K, int D, float *w32, double *w64,
                             float *w1, float *w1max, cudaStream_t s) {
    cudaMemsetAsync(w1max, 0, sizeof(float), s);
    const int K32 = (K + kTileF32 - 1) / kTileF32 * kTileF32;
    const long total = (long) K32 * D;
    vq_prepare_codebook<<<(unsigned) ((total + 255) / 256), 256, 0, s>>>(raas, K, D, w32, w64, w1, w1max);
}

// --------------------------------------------------------------------------------
// (b) cell accumulation: one warp walks a contiguous run of frames, lane l owns
// components l, l+32, ... of the cell sum.  Consecutive frames that fall in the
// same cell are summed in registers and flushed with ONE FP64 atomic per
// component (run-length aggregation); the winner's distortion is recomputed in
// FP64 so DD / discel / minDist do not depend on the FP32 search.
// reference: add_to_cell, vq_learn.i:99-106; DD, vq_learn_par.c:77-79
// --------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

template <int NR, bool CELLS>
__global__ void __launch_bounds__(256)
vq_accumulate(const double *__restrict__ x64, const int32_t *__restrict__ idx, long n, int D,
              const double *__restrict__ raas, PackedLayout L, double *__restrict__ packed,
              double *__restrict__ mindist, uint16_t *__restrict__ seq, int32_t *__restrict__ neg_list,
              int32_t *__restrict__ neg_count) {
    const int lane = threadIdx.x & 31;
    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    const long chunk = (n + n_warps - 1) / n_warps;
    const long lo = warp * chunk;
    const long hi = lo + chunk < n ? lo + chunk : n;

    double run[NR];
#pragma unroll
    for (int r = 0; r < NR; r++) run[r] = 0.0;
    int run_idx = -1;
    int run_cnt = 0;
    double run_dis = 0.0, dd_sum = 0.0, in_sum = 0.0;

    auto flush = [&]() {
        if (run_idx >= 0) {
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const int m = lane + 32 * r;
                if (m < D) atomicAdd(packed + L.cells() + (long) run_idx * D + m, run[r]);
            }
            if (lane == 0) {
                atomicAdd(packed + L.counts() + run_idx, (double) run_cnt);
                atomicAdd(packed + L.discel() + run_idx, run_dis);
            }
        }
    };

    constexpr int U = 4;
    for (long v0 = lo; v0 < hi; v0 += U) {
        int ci[U];
        double xv[U][NR], wv[U][NR];
#pragma unroll
        for (int u = 0; u < U; u++) ci[u] = v0 + u < hi ? idx[v0 + u] : 0;
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const int m = lane + 32 * r;
                xv[u][r] = (v0 + u < hi && m < D) ? x64[(v0 + u) * D + m] : 0.0;
            }
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const int m = lane + 32 * r;
                double w = (v0 + u < hi && m < D) ? raas[(long) ci[u] * D + m] : 0.0;
                wv[u][r] = m > 0 ? 2.0 * w : w;
            }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (v0 + u >= hi) break;
            double part = 0.0;
#pragma unroll
            for (int r = 0; r < NR; r++) part = fma(xv[u][r], wv[u][r], part);
            const double dd = warp_sum(part);
            const double md = dd - 1.0;
            dd_sum += md;
            // inertia.c:21-25 takes min_i (dd_i^2): equal to (min_i dd_i)^2 only when the minimum is
            // >= 0 (every gain-normalised frame).  A frame whose minimum is negative is listed for
            // the exact scan (vq_inertia_neg) instead.
            if (neg_list != nullptr && dd < 0.0) {
                if (lane == 0) neg_list[atomicAdd(neg_count, 1)] = (int32_t) (v0 + u);
            } else {
                in_sum += dd * dd;
            }
            if (lane == 0) {
                if (mindist) mindist[v0 + u] = md;
                if (seq) seq[v0 + u] = (uint16_t) ci[u];
            }
            if (CELLS) {
                if (ci[u] != run_idx) {
                    flush();
                    run_idx = ci[u];
                    run_cnt = 0;
                    run_dis = 0.0;
#pragma unroll
                    for (int r = 0; r < NR; r++) run[r] = 0.0;
                }
#pragma unroll
                for (int r = 0; r < NR; r++) run[r] += xv[u][r];
                run_cnt++;
                run_dis += md;
            }
        }
    }
    if (CELLS) flush();

    __shared__ double s_dd[8], s_in[8];
    const int w = threadIdx.x >> 5;
    if (lane == 0) { s_dd[w] = dd_sum; s_in[w] = in_sum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < (int) (blockDim.x >> 5); i++) { a += s_dd[i]; b += s_in[i]; }
        atomicAdd(packed + L.dd(), a);
        atomicAdd(packed + L.inertia(), b);
    }
}

template <int NR>
static void launch_accumulate_nr(const double *x64, const int32_t *idx, long n, int D,
                                 const double *raas, int K, double *packed, double *mindist,
                                 uint16_t *seq, bool with_cells, int num_sms, cudaStream_t s,
                                 int32_t *neg_list, int32_t *neg_count) {
    PackedLayout L{K, D};
    long warps_wanted = (n + 31) / 32;  // >= 32 frames per warp
    long blocks = (warps_wanted + 7) / 8;
    const long cap = (long) num_sms * 8;
    if (blocks > cap) blocks = cap;
    if (blocks < 1) blocks = 1;
    if (with_cells)
        vq_accumulate<NR, true><<<(unsigned) blocks, 256, 0, s>>>(x64, idx, n, D, raas, L, packed, mindist, seq,
                                                                  neg_list, neg_count);
    else
        vq_accumulate<NR, false><<<(unsigned) blocks, 256, 0, s>>>(x64, idx, n, D, raas, L, packed, mindist, seq,
                                                                   neg_list, neg_count);
}

// inertia.c:15-28 for the frames whose minimum distortion is negative (never for gain-normalised
// input): min_i (dd_i^2) by an exact scan of all K codewords; one warp per listed frame, lanes over
// codewords.
__global__ void __launch_bounds__(256)
vq_inertia_neg(const double *__restrict__ x64, int D, const double *__restrict__ raas, PackedLayout L,
               const int32_t *__restrict__ neg_list, const int32_t *__restrict__ neg_count,
               double *__restrict__ packed) {
    const int lane = threadIdx.x & 31;
    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    const long total = *neg_count;
    double acc = 0.0;
    for (long k = warp; k < total; k += n_warps) {
        const double *xr = x64 + (long) neg_list[k] * D;
        double best = (double) INFINITY;
        for (int c = lane; c < L.K; c += 32) {
            const double *rr = raas + (long) c * D;
            double dd = xr[0] * rr[0];
            for (int m = 1; m < D; m++) dd = fma(xr[m], 2.0 * rr[m], dd);
            best = fmin(best, dd * dd);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) best = fmin(best, __shfl_xor_sync(0xffffffffu, best, off));
        acc += best;
    }
    if (lane == 0 && acc != 0.0) atomicAdd(packed + L.inertia(), acc);
}
void launch_inertia_neg(const double *x64, int D, const double *raas, int K, const int32_t *neg_list,
                        const int32_t *neg_count, double *packed, int num_sms, cudaStream_t s) {
    PackedLayout L{K, D};
    vq_inertia_neg<<<num_sms * 2, 256, 0, s>>>(x64, D, raas, L, neg_list, neg_count, packed);
}

// ECOZ2_VQ_AUDIT=n: every n-th frame is searched again over ALL codewords in FP64 (one warp per
// audited frame, lanes over codewords, strict '<' in index order like quantize.c:19-26) and compared
// with the index the search + re-evaluation produced.
//   out[0] += frames audited
//   out[1] += frames whose index differs from the FP64 scan's
//   out[2] += of those, frames whose chosen codeword is worse than the FP64 minimum by more than
//             1e-6 relative (north_star's exemption does not cover them: a real mis-assignment)
__global__ void __launch_bounds__(256)
vq_audit(const double *__restrict__ x64, long n, int D, const double *__restrict__ raas, int K,
         const int32_t *__restrict__ idx, int every, unsigned long long *__restrict__ out) {
    const int lane = threadIdx.x & 31;
    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    const long n_audit = (n + every - 1) / every;
    unsigned long long audited = 0, differ = 0, wrong = 0;
    for (long a = warp; a < n_audit; a += n_warps) {
        const long v = a * every;
        const double *xr = x64 + v * D;
        double best = (double) INFINITY;
        int bi = 0x7fffffff;
        const int chosen = idx[v];
        double d_chosen = 0.0;
        for (int c = lane; c < K; c += 32) {
            const double *rr = raas + (long) c * D;
            double dd = xr[0] * rr[0];
            for (int m = 1; m < D; m++) dd = fma(xr[m], 2.0 * rr[m], dd);
            if (dd < best) { best = dd; bi = c; }
        }
        {   // the chosen codeword by the same summation order (all lanes: no divergence in the shuffles)
            const double *rr = raas + (long) (chosen >= 0 && chosen < K ? chosen : 0) * D;
            d_chosen = xr[0] * rr[0];
            for (int m = 1; m < D; m++) d_chosen = fma(xr[m], 2.0 * rr[m], d_chosen);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) {
            const double ob = __shfl_xor_sync(0xffffffffu, best, off);
            const int oi = __shfl_xor_sync(0xffffffffu, bi, off);
            if (ob < best || (ob == best && oi < bi)) { best = ob; bi = oi; }
        }
        audited++;
        if (bi != chosen && bi != 0x7fffffff) {     // (a NaN frame keeps index 0 everywhere)
            differ++;
            if (!(d_chosen - best <= 1e-6 * fabs(best))) wrong++;
        }
    }
    if (lane == 0 && audited) {
        atomicAdd(out + 0, audited);
        if (differ) atomicAdd(out + 1, differ);
        if (wrong) atomicAdd(out + 2, wrong);
    }
}
void launch_audit(const double *x64, long n, int D, const double *raas, int K, const int32_t *idx,
                  int every, unsigned long long *out, int num_sms, cudaStream_t s) {
    if (n <= 0 || every < 1) return;
    vq_audit<<<num_sms * 4, 256, 0, s>>>(x64, n, D, raas, K, idx, every, out);
}

// --------------------------------------------------------------------------------
// (b') the same without cell sums (vq quantize, inertia, min_dists): FP64 distortion of every
// frame against its chosen codeword, symbols, DD and inertia.  vq_accumulate spends ~80 warp
// instructions per frame (a 5-level 64-bit shuffle tree per frame) and is issue-bound at
// 2.7 TB/s; here a warp takes 8 consecutive frames (16 for rows of D <= 20) = 8 D consecutive
// doubles with fully coalesced loads (lane = element, not component), multiplies each element
// with its codeword weight, parks the products in shared memory and lets 4 (2) lanes per frame
// add them up.  Sum order per frame: strided partial sums (components q, q + 4, ... without
// component 0), then (q0 + q1) + (q2 + q3), times 2, plus component 0.
// reference: quantize.c:14-30 (distortion of the winner), inertia.c:15-28
// --------------------------------------------------------------------------------
constexpr int kWdMaxD = 40;
// FB frames per warp step: 8 (4 lanes add up a frame), or 16 (2 lanes) for the short rows D <= 20
template <int NL, int FB>
__global__ void __launch_bounds__(256, 4)
vq_winner_distortion(const double *__restrict__ x64, const int32_t *__restrict__ idx, long n, int D,
                     const double *__restrict__ raas, PackedLayout L, double *__restrict__ packed,
                     double *__restrict__ mindist, uint16_t *__restrict__ seq,
                     int32_t *__restrict__ neg_list, int32_t *__restrict__ neg_count) {
    constexpr int LPF = 32 / FB;       // lanes per frame in the sums
    constexpr int RS = LPF * NL + 1;   // row stride of the product rows (odd: spreads the banks)
    __shared__ double s_p[8][FB * RS];
    __shared__ int s_ci[8][FB];
    // element e of a block -> (frame in the block) << 8 | component (a table: registers are short)
    __shared__ int s_efm[32 * NL];
    const int lane = threadIdx.x & 31;
    const int w = threadIdx.x >> 5;
    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    double *sp = s_p[w];
    int *sci = s_ci[w];
    // columns D .. LPF NL - 1 of every row are never written: zero once, the sums below run over LPF NL
    for (int k = lane; k < FB * RS; k += 32) sp[k] = 0.0;
    for (int e = threadIdx.x; e < 32 * NL; e += blockDim.x) s_efm[e] = ((e / D) << 8) | (e % D);
    __syncthreads();
    const int *efm = s_efm + lane;
    const bool last_ok = lane + 32 * (NL - 1) < FB * D;
    const int rf = lane / LPF, rq = lane % LPF;
    const double *srow = sp + rf * RS + rq;
    double dd_sum = 0.0, in_sum = 0.0;

    // sums of the FB product rows and the outputs of nf <= FB frames.  The factor 2 of the components
    // m >= 1 (distortion.c:6-13) is applied to their SUM: scaling by 2 is exact, so this equals the
    // sum of the doubled terms bit for bit.
    auto finish = [&](long v0, int nf) {
        __syncwarp();
        double part = 0.0;
#pragma unroll
        for (int j = 1; j < NL; j++) part += srow[LPF * j];
        const double first = srow[0];          // component rq: the undoubled one in lane rq == 0
        if (rq != 0) part += first;
        part += __shfl_xor_sync(0xffffffffu, part, 1);
        if (LPF == 4) part += __shfl_xor_sync(0xffffffffu, part, 2);
        __syncwarp();
        if (rq == 0 && rf < nf) {
            const double dd = first + 2.0 * part;
            const double md = dd - 1.0;
            dd_sum += md;
            // (inertia.c:21-25: min_i (dd_i^2); a negative minimum goes to the exact scan, see vq_accumulate)
            if (neg_list != nullptr && dd < 0.0) neg_list[atomicAdd(neg_count, 1)] = (int32_t) (v0 + rf);
            else in_sum += dd * dd;
            if (mindist) mindist[v0 + rf] = md;
            if (seq) seq[v0 + rf] = (uint16_t) sci[rf];
        }
    };

    const long n_full = n / FB;
    long b = warp;
    int ci_next = (b < n_full && lane < FB) ? idx[b * FB + lane] : 0;
    for (; b < n_full; b += n_warps) {
        const long v0 = b * FB;
        const double *xb = x64 + v0 * D;
        double xv[NL];
#pragma unroll
        for (int r = 0; r < NL; r++) xv[r] = (r < NL - 1 || last_ok) ? xb[lane + 32 * r] : 0.0;
        if (lane < FB) sci[lane] = ci_next;
        __syncwarp();
        if (b + n_warps < n_full && lane < FB) ci_next = idx[(b + n_warps) * FB + lane];
#pragma unroll
        for (int r = 0; r < NL; r++) {
            if (r < NL - 1 || last_ok) {
                const int fm = efm[32 * r];
                const int f = fm >> 8, m = fm & 255;
                const double wv = raas[sci[f] * D + m];
                sp[f * RS + m] = xv[r] * wv;
            }
        }
        finish(v0, FB);
    }
    if (n_full * FB < n && warp == n_full % n_warps) {   // the partial block at the end
        const long v0 = n_full * FB;
        const int nf = (int) (n - v0);
        if (lane < nf) sci[lane] = idx[v0 + lane];
        __syncwarp();
        for (int e = lane; e < FB * D; e += 32) {
            const int f = e / D, m = e - f * D;
            sp[f * RS + m] = f < nf ? x64[v0 * D + e] * raas[sci[f] * D + m] : 0.0;
        }
        finish(v0, nf);
    }
    dd_sum = warp_sum(dd_sum);
    in_sum = warp_sum(in_sum);
    __shared__ double s_dd[8], s_in[8];
    if (lane == 0) { s_dd[w] = dd_sum; s_in[w] = in_sum; }
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0, b2 = 0.0;
        for (int i = 0; i < (int) (blockDim.x >> 5); i++) { a += s_dd[i]; b2 += s_in[i]; }
        atomicAdd(packed + L.dd(), a);
        atomicAdd(packed + L.inertia(), b2);
    }
}

template <int NL, int FB>
static void launch_winner_distortion_nl(const double *x64, const int32_t *idx, long n, int D, const double *raas,
                                        int K, double *packed, double *mindist, uint16_t *seq, int num_sms,
                                        cudaStream_t s, int32_t *neg_list, int32_t *neg_count) {
    PackedLayout L{K, D};
    const long n_blocks = (n + FB - 1) / FB;
    long blocks = (n_blocks + 7) / 8;
    const long cap = (long) num_sms * 8;
    if (blocks > cap) blocks = cap;
    vq_winner_distortion<NL, FB><<<(unsigned) blocks, 256, 0, s>>>(x64, idx, n, D, raas, L, packed, mindist, seq,
                                                               neg_list, neg_count);
}

static bool winner_distortion_enabled() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("ECOZ2_VQ_WINNER_KERNEL"); on = (e && e[0] == '0') ? 0 : 1; }
    return on == 1;
}

void launch_accumulate(const double *x64, const int32_t *idx, long n, int D, const double *raas,
                       int K, double *packed, double *mindist, uint16_t *seq, bool with_cells,
                       int num_sms, cudaStream_t s, int32_t *neg_list, int32_t *neg_count) {
    if (n <= 0) return;
    if (!with_cells && D <= kWdMaxD && winner_distortion_enabled()) {
#define WD_CASE(NLV) case NLV: if (D <= 20) launch_winner_distortion_nl<NLV, 16>(x64, idx, n, D, raas, K, packed, mindist, seq, num_sms, s, neg_list, neg_count); \
                               else launch_winner_distortion_nl<NLV, 8>(x64, idx, n, D, raas, K, packed, mindist, seq, num_sms, s, neg_list, neg_count); return;
        switch (((D <= 20 ? 16 : 8) * D + 31) / 32) {
            WD_CASE(1) WD_CASE(2) WD_CASE(3) WD_CASE(4) WD_CASE(5) WD_CASE(6) WD_CASE(7) WD_CASE(8) WD_CASE(9) WD_CASE(10)
            default: break;
        }
#undef WD_CASE
    }
    const int nr = (D + 31) / 32;
    switch (nr) {
        case 1: launch_accumulate_nr<1>(x64, idx, n, D, raas, K, packed, mindist, seq, with_cells, num_sms, s, neg_list, neg_count); break;
        case 2: launch_accumulate_nr<2>(x64, idx, n, D, raas, K, packed, mindist, seq, with_cells, num_sms, s, neg_list, neg_count); break;
        case 3: launch_accumulate_nr<3>(x64, idx, n, D, raas, K, packed, mindist, seq, with_cells, num_sms, s, neg_list, neg_count); break;
        case 4: launch_accumulate_nr<4>(x64, idx, n, D, raas, K, packed, mindist, seq, with_cells, num_sms, s, neg_list, neg_count); break;
        default: launch_accumulate_nr<7>(x64, idx, n, D, raas, K, packed, mindist, seq, with_cells, num_sms, s, neg_list, neg_count); break;
    }
}

// --------------------------------------------------------------------------------
// (a+b) fused small-codebook pass, K <= 32: search, cell sums, counts and distortions
// in ONE sweep over the FP64 frames -- all FP64, so there is nothing to re-evaluate.
//   * lane j owns codeword j: its D weights live in registers;
//   * the warp loads frames coalesced (lane l holds components l, l+32), parks them in a
//     small per-warp shared buffer and every lane reads them back as broadcasts;
//   * min over lanes by xor-shuffles, argmin = lowest lane holding the minimum
//     (== strict '<' scan in index order, vq_learn_par.c:66-75);
//   * cell sums go to per-WARP private shared-memory cells (no atomics); counts and
//     cell distortions stay in the owning lane's registers; one flush per CTA at the end.
// Small codebooks are HBM-bound (intensity K/2 flop/byte): this kernel reads 8 D bytes
// per frame once instead of 4 D (search) + 8 D (accumulate) + contended atomics.
// --------------------------------------------------------------------------------
template <int DT, bool CELLS>
__global__ void __launch_bounds__(128)
vq_pass_small(const double *__restrict__ x64, long n, const double *__restrict__ raas, PackedLayout L,
              double *__restrict__ packed, int32_t *__restrict__ idx_out,
              double *__restrict__ mindist, uint16_t *__restrict__ seq) {
    constexpr int D = DT;
    constexpr int NR = (D + 31) / 32;
    constexpr int U = 4;              // frames in flight per warp
    constexpr int XROW = (D + 1) & ~1;  // even: 16-byte aligned rows for the double2 reads
    const int K = L.K;
    extern __shared__ __align__(16) unsigned char smem_small[];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int n_w = blockDim.x >> 5;
    double *xs = reinterpret_cast<double *>(smem_small) + (size_t) w * U * XROW;  // [U][XROW]
    double *cells = reinterpret_cast<double *>(smem_small) + (size_t) n_w * U * XROW +
                    (size_t) w * K * D;                                         // [K][D]
    if (CELLS)
        for (int e = lane; e < K * D; e += 32) cells[e] = 0.0;

    // codeword of this lane (w = (raa0, 2 raa1, ..)); lanes >= K never win
    double wv[D];
    const bool owner = lane < K;
#pragma unroll
    for (int m = 0; m < D; m++) {
        double v = owner ? raas[(long) lane * D + m] : 0.0;
        wv[m] = m > 0 ? 2.0 * v : v;
    }
    __syncwarp();

    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    const long chunk = ((n + n_warps - 1) / n_warps + U - 1) / U * U;
    const long lo = warp * chunk;
    const long hi = lo + chunk < n ? lo + chunk : n;

    int my_cnt = 0;
    double my_dis = 0.0, dd_sum = 0.0, in_sum = 0.0;

    double xr[U][NR];
    auto load = [&](long v0) {
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const int m = lane + 32 * r;
                xr[u][r] = (v0 + u < hi && m < D) ? x64[(v0 + u) * D + m] : 0.0;
            }
    };
    if (lo < hi) load(lo);
    for (long v0 = lo; v0 < hi; v0 += U) {
        // park the U frames in shared memory, then prefetch the next group
#pragma unroll
        for (int u = 0; u < U; u++)
#pragma unroll
            for (int r = 0; r < NR; r++) {
                const int m = lane + 32 * r;
                if (m < XROW) xs[u * XROW + m] = xr[u][r];
            }
        __syncwarp();
        if (v0 + U < hi) load(v0 + U);

        // broadcast reads, software pipelined one step ahead; volatile asm keeps the
        // loads in program order so the unrolled loop does not hoist all of them
        double acc[U];
        double2 cur[U], nxt[U];
        const uint32_t xs_addr = smem_u32(xs);
#pragma unroll
        for (int u = 0; u < U; u++) {
            acc[u] = 0.0;
            asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(cur[u].x), "=d"(cur[u].y)
                         : "r"(xs_addr + (uint32_t) (u * XROW) * 8u));
        }
#pragma unroll
        for (int m = 0; m < D; m += 2) {
            if (m + 2 < D) {
#pragma unroll
                for (int u = 0; u < U; u++)
                    asm volatile("ld.shared.v2.f64 {%0, %1}, [%2];" : "=d"(nxt[u].x), "=d"(nxt[u].y)
                                 : "r"(xs_addr + (uint32_t) (u * XROW + m + 2) * 8u));
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                acc[u] = fma(cur[u].x, wv[m], acc[u]);
                if (m + 1 < D) acc[u] = fma(cur[u].y, wv[m + 1], acc[u]);
            }
#pragma unroll
            for (int u = 0; u < U; u++) cur[u] = nxt[u];
        }
#pragma unroll
        for (int u = 0; u < U; u++) {
            if (v0 + u >= hi) break;
            const double mine = owner ? acc[u] : (double) INFINITY;
            double mn = mine;
#pragma unroll
            for (int off = 16; off >= 1; off >>= 1) mn = fmin(mn, __shfl_xor_sync(0xffffffffu, mn, off));
            const unsigned hit = __ballot_sync(0xffffffffu, mine == mn);
            const int ci = hit ? __ffs(hit) - 1 : 0;    // NaN frame: reference quantize keeps 0
            const double md = mn - 1.0;
            dd_sum += md;
            double m2 = mn * mn;
            if (mn < 0.0) {   // inertia.c:21-25 as written: min_i (dd_i^2), not (min_i dd_i)^2
                m2 = owner ? mine * mine : (double) INFINITY;
#pragma unroll
                for (int off = 16; off >= 1; off >>= 1) m2 = fmin(m2, __shfl_xor_sync(0xffffffffu, m2, off));
            }
            in_sum += m2;
            if (lane == 0) {
                idx_out[v0 + u] = ci;
                if (mindist) mindist[v0 + u] = md;
                if (seq) seq[v0 + u] = (uint16_t) ci;
            }
            if (CELLS) {
                if (lane == ci) { my_cnt++; my_dis += md; }
#pragma unroll
                for (int r = 0; r < NR; r++) {
                    const int m = lane + 32 * r;
                    if (m < D) cells[ci * D + m] += xs[u * XROW + m];
                }
            }
        }
        __syncwarp();
    }

    // ---- flush: per-warp partials -> global packed buffer
    __shared__ double s_dd[8], s_in[8];
    if (lane == 0) { s_dd[w] = dd_sum; s_in[w] = in_sum; }
    if (CELLS && owner && my_cnt > 0) {
        atomicAdd(packed + L.counts() + lane, (double) my_cnt);
        atomicAdd(packed + L.discel() + lane, my_dis);
    }
    __syncthreads();
    if (CELLS) {
        const double *all = reinterpret_cast<double *>(smem_small) + (size_t) n_w * U * XROW;
        for (int e = threadIdx.x; e < K * D; e += blockDim.x) {
            double t = 0.0;
            for (int ww = 0; ww < n_w; ww++) t += all[(size_t) ww * K * D + e];
            if (t != 0.0) atomicAdd(packed + L.cells() + e, t);
        }
    }
    if (threadIdx.x == 0) {
        double a = 0.0, b = 0.0;
        for (int i = 0; i < n_w; i++) { a += s_dd[i]; b += s_in[i]; }
        atomicAdd(packed + L.dd(), a);
        atomicAdd(packed + L.inertia(), b);
    }
}

// --------------------------------------------------------------------------------
// (a+b) fused small-codebook pass, second version: thread = frame.
// The first version above gives a lane to every codeword, so a frame costs D warp-wide
// DFMAs whatever K is (2.9 ms per 10M-frame pass, K-independent).  Here
//   * a CTA of 64 threads takes tiles of 128 consecutive frames: ONE cp.async.bulk of the
//     contiguous 128 x D doubles (37.9 KB at D = 37) into shared memory; up to 5 CTAs per SM
//     keep enough bytes in flight without double buffering;
//   * a thread takes two rows (row stride D is odd: conflict-free 8-byte accesses) and keeps all
//     2 x KT running distortions in registers; the codebook sits in shared memory as [m][KT]
//     (factor 2 folded in), read as 16-byte broadcasts: one load feeds four DFMAs.  (With the
//     codebook in __constant__ memory the kernel ran at a third of the DFMA rate: 5-10 KB cycled
//     once per frame do not stay in the constant cache);
//   * argmin with strict '<' in index order (vq_learn_par.c:66-75, quantize.c:14-30);
//   * cell sums live in REGISTERS for the whole kernel: lane m of a warp owns components m and
//     m + 32 of all KT cells; the warp walks its 64 rows (cell index by shuffle, 8 rows per step)
//     and every cell's accumulator takes the row under a predicate.  KT predicated adds per row
//     cost issue slots but there is no dependent load-add-store chain (per-warp shared-memory
//     tables, tried first, spent 70 cycles per row in that chain and limited the CTAs per SM);
//     counts through match_any + one shared int atomic per distinct cell and warp;
//   * per-cell distortion and DD come from the sums (vq_small_finish: cellsum . w - count),
//     like the sorted path of vq_cells.cu; the inertia pass (CELLS = false) adds DD and
//     sum ddmin^2 directly.
// FP64 throughout: K x D DFMAs per frame under the HBM time of the 8 D bytes per frame.
// --------------------------------------------------------------------------------
constexpr int kSmallMaxK = 16, kSmallTile = 128, kSmallThreads = 64;
constexpr int kSmallRows = kSmallTile / kSmallThreads;   // frames per thread

// sum[cu] += x with the accumulators in registers: a warp-uniform switch whose cases add into
// statically named registers.  The adds are asm volatile on purpose: written as plain C++
// (`if (cu == c) sum[c] += x`) the compiler turns the chain into a dynamically indexed local-memory
// array, and explicitly predicated adds become add + select (K adds per row).
template <int KT, bool HI, int NLO, int NHI>
__device__ __forceinline__ void small_add_to_cell(double (&lo)[NLO], double (&hi)[NHI], int cu, double xa, double xb) {
#define ECOZ2_SMALL_CASE(c)                                                              \
    case c:                                                                              \
        if constexpr (c < KT && c < NLO) {                                               \
            asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(lo[c]) : "d"(xa));              \
            if constexpr (HI && c < NHI) asm volatile("add.rn.f64 %0, %0, %1;" : "+d"(hi[c]) : "d"(xb)); \
        }                                                                                \
        break;
    switch (cu) {
        ECOZ2_SMALL_CASE(0) ECOZ2_SMALL_CASE(1) ECOZ2_SMALL_CASE(2) ECOZ2_SMALL_CASE(3)
        ECOZ2_SMALL_CASE(4) ECOZ2_SMALL_CASE(5) ECOZ2_SMALL_CASE(6) ECOZ2_SMALL_CASE(7)
        ECOZ2_SMALL_CASE(8) ECOZ2_SMALL_CASE(9) ECOZ2_SMALL_CASE(10) ECOZ2_SMALL_CASE(11)
        ECOZ2_SMALL_CASE(12) ECOZ2_SMALL_CASE(13) ECOZ2_SMALL_CASE(14) ECOZ2_SMALL_CASE(15)
        default: break;
    }
#undef ECOZ2_SMALL_CASE
}

template <int DT, int KT, bool CELLS>
__global__ void __launch_bounds__(kSmallThreads, 5)
vq_pass_small2(const double *__restrict__ x64, long n, const double *__restrict__ raas, int K, PackedLayout L,
               double *__restrict__ packed, int32_t *__restrict__ idx_out, double *__restrict__ mindist,
               uint16_t *__restrict__ seq, int bulk_ok) {
    constexpr int D = DT, R = kSmallRows, NW = kSmallThreads / 32;
    constexpr bool HI = D > 32;   // components 32.. exist
    extern __shared__ __align__(16) unsigned char smem_small2[];
    double *tile = reinterpret_cast<double *>(smem_small2);                       // [128][D]
    double *wsm = tile + kSmallTile * D;                                          // [D][KT] weights
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    __shared__ uint64_t s_bar;
    __shared__ int s_cnt[kSmallMaxK];
    __shared__ double s_dd[NW], s_in[NW];
    const uint32_t bar = smem_u32(&s_bar);
    for (int e = threadIdx.x; e < D * KT; e += kSmallThreads) {
        const int m = e / KT, c = e % KT;
        const double v = c < K ? raas[(long) c * D + m] : 0.0;
        wsm[e] = m > 0 ? 2.0 * v : v;   // exact doubling: distortion = r0 a0 + 2 sum r_m a_m (distortion.c:6-13)
    }
    if (CELLS && threadIdx.x < kSmallMaxK) s_cnt[threadIdx.x] = 0;
    if (threadIdx.x == 0) {
        asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" ::"r"(bar));
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    double sum_lo[CELLS ? KT : 1], sum_hi[CELLS && HI ? KT : 1];   // cell sums, components lane / lane + 32
#pragma unroll
    for (int c = 0; c < (CELLS ? KT : 1); c++) sum_lo[c] = 0.0;
#pragma unroll
    for (int c = 0; c < (CELLS && HI ? KT : 1); c++) sum_hi[c] = 0.0;
    double dd_sum = 0.0, in_sum = 0.0;
    uint32_t phase = 0;
    const long n_tiles = (n + kSmallTile - 1) / kSmallTile;
    for (long t = blockIdx.x; t < n_tiles; t += gridDim.x) {
        const long v0 = t * kSmallTile;
        const int rows = (int) (n - v0 < kSmallTile ? n - v0 : kSmallTile);
        if (bulk_ok && rows == kSmallTile) {
            if (threadIdx.x == 0) {
                constexpr uint32_t bytes = kSmallTile * D * 8;
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");   // the tile's readers are done
                asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(bar), "r"(bytes) : "memory");
                asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
                                 "r"(smem_u32(tile)), "l"(x64 + v0 * D), "r"(bytes), "r"(bar)
                             : "memory");
            }
            uint32_t done = 0;
            while (!done)
                asm volatile("{\n.reg .pred q;\nmbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n"
                             "selp.u32 %0, 1, 0, q;\n}\n"
                             : "=r"(done)
                             : "r"(bar), "r"(phase)
                             : "memory");
            phase ^= 1;
        } else {
            for (int e = threadIdx.x; e < rows * D; e += kSmallThreads) tile[e] = x64[v0 * D + e];
            __syncthreads();
        }
        // rows threadIdx.x and threadIdx.x + 64 (rows past the end of a partial tile read stale
        // shared memory and are discarded)
        int best[R];
        {
            const double *xr = tile + threadIdx.x * D;
            double acc[R][KT];
#pragma unroll
            for (int m = 0; m < D; m++) {
                double xv[R];
#pragma unroll
                for (int q = 0; q < R; q++) xv[q] = xr[q * kSmallThreads * D + m];
#pragma unroll
                for (int c = 0; c < KT; c += 2) {
                    const double2 wv = *reinterpret_cast<const double2 *>(wsm + m * KT + c);
#pragma unroll
                    for (int q = 0; q < R; q++) {
                        acc[q][c] = m == 0 ? xv[q] * wv.x : fma(xv[q], wv.x, acc[q][c]);
                        acc[q][c + 1] = m == 0 ? xv[q] * wv.y : fma(xv[q], wv.y, acc[q][c + 1]);
                    }
                }
            }
#pragma unroll
            for (int q = 0; q < R; q++) {
                double bv = acc[q][0];
                best[q] = 0;
#pragma unroll
                for (int c = 1; c < KT; c++)
                    if (c < K && acc[q][c] < bv) {
                        bv = acc[q][c];
                        best[q] = c;
                    }
                const int r = threadIdx.x + q * kSmallThreads;
                if (r < rows) {
                    const double md = bv - 1.0;
                    dd_sum += md;
                    double m2 = bv * bv;
                    if (bv < 0.0) {   // inertia.c:21-25 as written: min_i (dd_i^2), not (min_i dd_i)^2
#pragma unroll
                        for (int c = 0; c < KT; c++)
                            if (c < K) m2 = fmin(m2, acc[q][c] * acc[q][c]);
                    }
                    in_sum += m2;
                    idx_out[v0 + r] = best[q];
                    if (mindist) mindist[v0 + r] = md;
                    if (seq) seq[v0 + r] = (uint16_t) best[q];
                } else {
                    best[q] = -1;   // matches no cell below
                }
            }
        }
        if (CELLS) {
#pragma unroll
            for (int q = 0; q < R; q++) {
                const int first = q * kSmallThreads + w * 32;   // first row of this warp's group
                if (first < rows) {
                    const unsigned live = __ballot_sync(0xffffffffu, best[q] >= 0);
                    if (best[q] >= 0) {
                        const unsigned same = __match_any_sync(live, best[q]);
                        if (lane == __ffs(same) - 1) atomicAdd(&s_cnt[best[q]], __popc(same));
                    }
                    const double *wt = tile + (size_t) first * D;
#pragma unroll 1
                    for (int u = 0; u < 32; u += 8) {
                        int cu[8];
                        double xa[8], xb[8];
#pragma unroll
                        for (int k = 0; k < 8; k++) {
                            cu[k] = __shfl_sync(0xffffffffu, best[q], u + k);
                            xa[k] = lane < D ? wt[(u + k) * D + lane] : 0.0;
                            xb[k] = HI && lane + 32 < D ? wt[(u + k) * D + lane + 32] : 0.0;
                        }
#pragma unroll
                        for (int k = 0; k < 8; k++)
                            if constexpr (HI) {
                                small_add_to_cell<KT, HI>(sum_lo, sum_hi, cu[k], xa[k], xb[k]);
                            } else {
                                // short rows (D <= 32): K add + select pairs are cheaper than the jump
#pragma unroll
                                for (int c = 0; c < KT; c++)
                                    asm("{\n.reg .pred p;\nsetp.eq.s32 p, %2, %3;\n@p add.rn.f64 %0, %0, %1;\n}"
                                        : "+d"(sum_lo[c])
                                        : "d"(xa[k]), "r"(cu[k]), "r"(c));
                            }
                    }
                }
            }
        }
        __syncthreads();   // the tile is free for the next copy
    }

    // ---- flush: register cell sums and partial sums -> global packed buffer
#pragma unroll
    for (int off = 16; off >= 1; off >>= 1) {
        dd_sum += __shfl_xor_sync(0xffffffffu, dd_sum, off);
        in_sum += __shfl_xor_sync(0xffffffffu, in_sum, off);
    }
    if (lane == 0) {
        s_dd[w] = dd_sum;
        s_in[w] = in_sum;
    }
    __syncthreads();
    if (CELLS) {
#pragma unroll
        for (int c = 0; c < KT; c++) {
            if (c < K) {
                if (lane < D && sum_lo[c] != 0.0) atomicAdd(packed + L.cells() + (long) c * D + lane, sum_lo[c]);
                if (HI && lane + 32 < D && sum_hi[c] != 0.0)
                    atomicAdd(packed + L.cells() + (long) c * D + lane + 32, sum_hi[c]);
            }
        }
        if (threadIdx.x < K && s_cnt[threadIdx.x] > 0)
            atomicAdd(packed + L.counts() + threadIdx.x, (double) s_cnt[threadIdx.x]);
    }
    if (threadIdx.x == 0) {
        double a = 0.0, b2 = 0.0;
#pragma unroll
        for (int ww = 0; ww < NW; ww++) {
            a += s_dd[ww];
            b2 += s_in[ww];
        }
        if (!CELLS) atomicAdd(packed + L.dd(), a);
        atomicAdd(packed + L.inertia(), b2);
    }
}

// per-cell distortion = cellsum . w - count, DD = their sum (same identity as vq_cell_finish)
__global__ void __launch_bounds__(256) vq_small_finish(const double *__restrict__ raas, PackedLayout L,
                                                       double *__restrict__ packed) {
    const int K = L.K, D = L.D;
    const int lane = threadIdx.x & 31, c = threadIdx.x >> 5;   // one CTA, warp c handles cells c, c + 8, ..
    double dd = 0.0;
    for (int cc = c; cc < K; cc += 8) {
        double part = 0.0;
        for (int m = lane; m < D; m += 32) {
            const double wv = m > 0 ? 2.0 * raas[(long) cc * D + m] : raas[(long) cc * D];
            part = fma(packed[L.cells() + (long) cc * D + m], wv, part);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
        const double cnt = packed[L.counts() + cc];
        const double dis = cnt > 0.0 ? part - cnt : 0.0;
        if (lane == 0) packed[L.discel() + cc] = dis;
        dd += dis;
    }
    __shared__ double s_part[8];
    if (lane == 0) s_part[c] = dd;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        for (int i = 0; i < 8; i++) a += s_part[i];
        packed[L.dd()] += a;
    }
}

static bool small2_enabled() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("ECOZ2_VQ_SMALL2"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}

template <int DT, int KT>
static void launch_pass_small2_k(const double *x64, long n, const double *raas, int K, double *packed,
                                 int32_t *idx, double *mindist, uint16_t *seq, bool with_cells, int num_sms,
                                 cudaStream_t s) {
    PackedLayout L{K, DT};
    const size_t smem = sizeof(double) * ((size_t) kSmallTile * DT + (size_t) DT * KT);
    const int bulk_ok = (reinterpret_cast<uintptr_t>(x64) & 15) == 0;
    long blocks = (long) num_sms * 5;
    const long tiles = (n + kSmallTile - 1) / kSmallTile;
    if (blocks > tiles) blocks = tiles;
    if (with_cells) {
        cudaFuncSetAttribute(vq_pass_small2<DT, KT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        vq_pass_small2<DT, KT, true><<<(unsigned) blocks, kSmallThreads, smem, s>>>(x64, n, raas, K, L, packed, idx,
                                                                                  mindist, seq, bulk_ok);
        vq_small_finish<<<1, 256, 0, s>>>(raas, L, packed);
    } else {
        cudaFuncSetAttribute(vq_pass_small2<DT, KT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        vq_pass_small2<DT, KT, false><<<(unsigned) blocks, kSmallThreads, smem, s>>>(x64, n, raas, K, L, packed, idx,
                                                                                   mindist, seq, bulk_ok);
    }
}
template <int DT>
static void launch_pass_small2_d(const double *x64, long n, const double *raas, int K, double *packed,
                                 int32_t *idx, double *mindist, uint16_t *seq, bool with_cells, int num_sms,
                                 cudaStream_t s) {
    if (K <= 2) launch_pass_small2_k<DT, 2>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
    else if (K <= 4) launch_pass_small2_k<DT, 4>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
    else if (K <= 8) launch_pass_small2_k<DT, 8>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
    else launch_pass_small2_k<DT, 16>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
}

// largest codebook the thread-per-frame kernel takes (ECOZ2_VQ_SMALL2_MAXK, up to 16).  Measured on
// B200, ms per pass.  10M frames, P = 36, against 1.47-1.50 of tensor-core search + sorted cell sums:
// K = 2: 0.56, 4: 0.75, 8: 1.01, 16: 1.58 (jump per row; with add + select: 0.54, 0.79, 1.42, 2.85).
// 20M frames, P = 12, against 1.31-1.42: add + select 0.70, 0.99, 1.64, 2.91 (jump: 0.87, 1.25, 1.65,
// 2.20).  Defaults: K <= 8 for rows longer than 32, K <= 4 for short rows.
static int small2_max_k(int D) {
    static int v = -2;
    if (v == -2) {
        const char *e = getenv("ECOZ2_VQ_SMALL2_MAXK");
        v = e ? atoi(e) : -1;
        if (v > kSmallMaxK) v = kSmallMaxK;
    }
    if (v >= 0) return v;
    return D > 32 ? 8 : 4;
}
static bool small2_takes(int K, int D) { return small2_enabled() && K <= small2_max_k(D) && (D == 37 || D == 21 || D == 13); }

bool pass_small_fast(int K, int D) { return small2_takes(K, D); }
bool pass_small_supported(int K, int D) { return (K <= 32 && (D == 37 || D == 13)) || small2_takes(K, D); }

template <int DT>
static void launch_pass_small_d(const double *x64, long n, const double *raas, int K, double *packed,
                                int32_t *idx, double *mindist, uint16_t *seq, bool with_cells,
                                int num_sms, cudaStream_t s) {
    constexpr int XROW = (DT + 1) & ~1;
    PackedLayout L{K, DT};
    constexpr int kWarps = 4;   // 128 threads: ~160 registers/thread still leaves 3 CTAs per SM
    const size_t smem = sizeof(double) * (kWarps * (size_t) 4 * XROW + kWarps * (size_t) K * DT);
    int per_sm = (int) ((200 * 1024) / (smem + 1024));
    if (per_sm > 3) per_sm = 3;
    if (per_sm < 1) per_sm = 1;
    long blocks = (long) num_sms * per_sm;
    const long need = (n + 127) / 128;
    if (blocks > need) blocks = need < 1 ? 1 : need;
    if (with_cells) {
        cudaFuncSetAttribute(vq_pass_small<DT, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        vq_pass_small<DT, true><<<(unsigned) blocks, 128, smem, s>>>(x64, n, raas, L, packed, idx, mindist, seq);
    } else {
        cudaFuncSetAttribute(vq_pass_small<DT, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem);
        vq_pass_small<DT, false><<<(unsigned) blocks, 128, smem, s>>>(x64, n, raas, L, packed, idx, mindist, seq);
    }
}
void launch_pass_small(const double *x64, long n, int D, const double *raas, int K, double *packed,
                       int32_t *idx, double *mindist, uint16_t *seq, bool with_cells, int num_sms,
                       cudaStream_t s) {
    if (n <= 0) return;
    if (small2_takes(K, D)) {
        if (D == 37) launch_pass_small2_d<37>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
        else if (D == 21) launch_pass_small2_d<21>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
        else launch_pass_small2_d<13>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
        return;
    }
    if (D == 37) launch_pass_small_d<37>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
    else launch_pass_small_d<13>(x64, n, raas, K, packed, idx, mindist, seq, with_cells, num_sms, s);
}

// --------------------------------------------------------------------------------
// (c) centroid / split: one warp per codeword, FP64, shared-memory scratch
// --------------------------------------------------------------------------------
// step-up recursion rc -> predictor a (lpca_rc, lpca.c:106-125); a[] in shared
__device__ __forceinline__ void warp_step_up(int P, const double *rc, double *a, int lane) {
    if (lane == 0) a[0] = 1.0;
    __syncwarp();
    for (int k = 1; k <= P; k++) {
        const double kk = rc[k];
        const int half = k >> 1;
        double lo[4], hi[4];  // P <= 207 -> half <= 103 -> <= 4 pairs per lane
        int cnt = 0;
        for (int i = 1 + lane; i <= half; i += 32, cnt++) { lo[cnt] = a[i]; hi[cnt] = a[k - i]; }
        __syncwarp();
        cnt = 0;
        for (int i = 1 + lane; i <= half; i += 32, cnt++) {
            a[i] = lo[cnt] + kk * hi[cnt];
            a[k - i] = hi[cnt] + kk * lo[cnt];   // i == k-i: both writes give the same value
        }
        if (lane == 0) a[k] = kk;
        __syncwarp();
    }
}
// raa[n] = sum_{k=0}^{P-n} a[k] a[k+n]  (ref2raas.c:17-23, vq_learn.i:138-144)
__device__ __forceinline__ void warp_autocorr(int P, const double *a, double *raa, int lane) {
    for (int n = lane; n <= P; n += 32) {
        double s = 0.0;
        for (int k = 0; k <= P - n; k++) s += a[k] * a[k + n];
        raa[n] = s;
    }
}

__global__ void __launch_bounds__(256)
vq_centroids(double *__restrict__ packed, PackedLayout L, double *__restrict__ refl,
             double *__restrict__ raas) {
    __shared__ double s_a[8][kMaxD];
    __shared__ double s_r[8][kMaxD];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i = blockIdx.x * 8 + w;
    const int D = L.D, P = D - 1;
    if (i >= L.K) return;
    if (!(packed[L.counts() + i] > 0.0)) return;  // empty cell: codeword unchanged (vq_learn.i:133)
    double *cel = packed + L.cells() + (long) i * D;
    double *a = s_a[w], *r = s_r[w];
    double *rc = refl + (long) i * D;
    for (int n = lane; n < D; n += 32) { r[n] = cel[n]; a[n] = 0.0; }
    __syncwarp();
    // lpca_r (lpca.c:64-101)
    double pe = r[0];
    if (pe == 0.0) {
        for (int n = lane; n < D; n += 32) { rc[n] = 0.0; raas[(long) i * D + n] = 0.0; }
        return;
    }
    if (lane == 0) a[0] = 1.0;
    __syncwarp();
    for (int k = 1; k <= P; k++) {
        double part = 0.0;
        for (int m = 1 + lane; m <= k; m += 32) part -= a[k - m] * r[m];
        const double sum = warp_sum(part);
        const double kk = sum / pe;
        const int half = k >> 1;
        double lo[4], hi[4];
        int cnt = 0;
        for (int m = 1 + lane; m <= half; m += 32, cnt++) { lo[cnt] = a[m]; hi[cnt] = a[k - m]; }
        __syncwarp();
        cnt = 0;
        for (int m = 1 + lane; m <= half; m += 32, cnt++) {
            a[m] = lo[cnt] + kk * hi[cnt];
            a[k - m] = hi[cnt] + kk * lo[cnt];
        }
        if (lane == 0) { a[k] = kk; rc[k] = kk; }
        __syncwarp();
        pe *= 1.0 - kk * kk;
        if (pe <= 0.0) break;  // lpca.c:97-98 (remaining predictor entries stay 0 here)
    }
    warp_autocorr(P, a, raas + (long) i * D, lane);
    if (pe != 0.0)
        for (int n = lane; n < D; n += 32) cel[n] = r[n] / pe;
}
void launch_centroids(double *packed, int K, int D, double *refl, double *raas, cudaStream_t s) {
    PackedLayout L{K, D};
    vq_centroids<<<(K + 7) / 8, 256, 0, s>>>(packed, L, refl, raas);
}

__global__ void __launch_bounds__(256)
vq_refl_to_raas(const double *__restrict__ refl, double *__restrict__ raas, int K, int D) {
    __shared__ double s_a[8][kMaxD];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int i = blockIdx.x * 8 + w;
    if (i >= K) return;
    warp_step_up(D - 1, refl + (long) i * D, s_a[w], lane);
    warp_autocorr(D - 1, s_a[w], raas + (long) i * D, lane);
}
void launch_refl_to_raas(const double *refl, double *raas, int K, int D, cudaStream_t s) {
    vq_refl_to_raas<<<(K + 7) / 8, 256, 0, s>>>(refl, raas, K, D);
}

// children 2j (x 0.99f) and 2j+1 (x 1.01f) of every row j, on ALL D entries; the
// factors are float literals widened to double (vq_learn.i:54-55).  Written to a
// scratch copy first so the in-place doubling has no ordering hazard.
__global__ void vq_grow_copy(const double *__restrict__ refl_in, double *__restrict__ refl_out,
                             int K, int D) {
    const long e = (long) blockIdx.x * blockDim.x + threadIdx.x;
    if (e >= (long) K * D) return;
    const int j = (int) (e / D), n = (int) (e - (long) j * D);
    const double v = refl_in[e];
    refl_out[(long) (2 * j) * D + n] = v * (double) 0.99f;
    refl_out[(long) (2 * j + 1) * D + n] = v * (double) 1.01f;
}
void launch_grow(double *refl, double *raas, int K, int D, cudaStream_t s) {
    // raas is dead until refl_to_raas below: use it as the scratch for the old rows
    const long total = (long) K * D;
    cudaMemcpyAsync(raas, refl, sizeof(double) * total, cudaMemcpyDeviceToDevice, s);
    vq_grow_copy<<<(unsigned) ((total + 255) / 256), 256, 0, s>>>(raas, refl, K, D);
    launch_refl_to_raas(refl, raas, 2 * K, D, s);
}

// --------------------------------------------------------------------------------
// sigma: sum_{i != j} [ distortion(cells_j, raas_i) - 1 ]  (sigma.c:14-25)
// CTA per codeword i, threads over cells j.
// --------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
vq_sigma_sum(const double *__restrict__ raas, const double *__restrict__ cells, int K, int D,
             double *__restrict__ out) {
    __shared__ double s_w[kMaxD];
    __shared__ double s_part[8];
    const int i = blockIdx.x;
    for (int n = threadIdx.x; n < D; n += blockDim.x) {
        const double v = raas[(long) i * D + n];
        s_w[n] = n > 0 ? 2.0 * v : v;
    }
    __syncthreads();
    double acc = 0.0;
    for (int j = threadIdx.x; j < K; j += blockDim.x) {
        if (j == i) continue;
        const double *c = cells + (long) j * D;
        double d = 0.0;
        for (int n = 0; n < D; n++) d = fma(c[n], s_w[n], d);
        acc += d - 1.0;
    }
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
    __syncthreads();
    if (threadIdx.x == 0) {
        double t = 0.0;
        for (int k = 0; k < 8; k++) t += s_part[k];
        atomicAdd(out, t);
    }
}
void launch_sigma_sum(const double *raas, const double *cells, int K, int D, double *out,
                      cudaStream_t s) {
    vq_sigma_sum<<<K, 256, 0, s>>>(raas, cells, K, D, out);
}

// --------------------------------------------------------------------------------
// FP32 FFMA peak probe: 8 independent FMA chains per thread, nothing else.  Gives
// the measured denominator of the FP32 roofline on the box the bench runs on.
// --------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) vq_ffma_probe(float *out, int iters, float a, float b) {
    float c0 = threadIdx.x, c1 = c0 + 1.f, c2 = c0 + 2.f, c3 = c0 + 3.f;
    float c4 = c0 + 4.f, c5 = c0 + 5.f, c6 = c0 + 6.f, c7 = c0 + 7.f;
    for (int i = 0; i < iters; i++) {
#pragma unroll
        for (int u = 0; u < 16; u++) {
            c0 = fmaf(c0, a, b); c1 = fmaf(c1, a, b); c2 = fmaf(c2, a, b); c3 = fmaf(c3, a, b);
            c4 = fmaf(c4, a, b); c5 = fmaf(c5, a, b); c6 = fmaf(c6, a, b); c7 = fmaf(c7, a, b);
        }
    }
    const float r = ((c0 + c1) + (c2 + c3)) + ((c4 + c5) + (c6 + c7));
    if (r == 123.456f) out[0] = r;  // keep the chains alive
}
void launch_ffma_probe(float *out, int blocks, int iters, cudaStream_t s) {
    vq_ffma_probe<<<blocks, 256, 0, s>>>(out, iters, 0.999f, 0.001f);
}

}  // namespace vqb200
