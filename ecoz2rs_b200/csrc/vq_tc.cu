// vq_tc.cu -- tensor-core (tcgen05 / TMEM) variant of kernel (a), the fused
// distortion + argmin search, for sm_100a.
//
// The distortion of frame i against codeword j is the inner product x_i . w_j
// (w = (raa0, 2 raa1 .. 2 raaP), reference distortion.c:6-13).  Here the N x K
// products run on the 5th-generation tensor cores as split-TF32 ("3xTF32"):
//     x = xh + xl,  w = wh + wl   (each part representable in TF32)
//     x . w  ~=  xl.wh + xh.wl + xh.wh          (xl.wl ~ 2^-22 relative, dropped)
// accumulated in FP32 in tensor memory.  One CTA per SM walks frame tiles of 128
// frames (UMMA M = 128, one TMEM lane per frame); for each frame tile all codeword
// tiles (UMMA N = BN <= 256) stream past it:
//
//   warp 0      B producer: one cp.async.bulk (TMA engine) per (codeword tile, k-step)
//               stage; the codebook is pre-laid-out in global memory in the UMMA
//               canonical K-major core-matrix order, so a stage is one contiguous run
//   warp 1      MMA issuer: one thread issues 3 tcgen05.mma.kind::tf32 per k-step into
//               one of two TMEM accumulators, tcgen05.commit releases stages / signals
//               the epilogue
//   warps 2-3   A loaders: FP32 frames -> (hi, lo) split -> canonical smem layout,
//               double buffered (next frame tile is staged while this one is searched)
//   warps 4-7   epilogue: tcgen05.ld the accumulator (thread = frame, registers =
//               codewords), running (best, second, argbest) per frame; the N x K
//               matrix never leaves the SM.
//
// The values are only as good as 3xTF32, so -- exactly like the FP32 FFMA search --
// a frame whose best-two gap is inside the error bound is appended to flag_list and
// re-evaluated in FP64 by launch_recheck (vq_kernels.cu).  The reference semantics
// (strict '<' scan in index order, vq_learn_par.c:66-75, quantize.c:19-26) are kept:
// first index among equal values wins.
#include "vq_kernels.cuh"

#include <cuda_fp16.h>
#include <cfloat>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "vq_tc_common.cuh"

namespace vqb200 {


struct TcParams {
    const float *x;             // frames [n][D] FP32
    long n_frames;
    int D;
    int ksteps;                 // ceil((D + 1) / 8): D elements + the bound slot
    const int32_t *list;        // optional gather list (+ device-side count)
    const int32_t *list_count;
    const float *wtc;           // codeword stages, see tc_stage layout below
    int K;
    int BN;                     // codewords per tile (UMMA N): 64, 128 or 256
    const float *w1;            // sum_n |w_i[n]| per codeword
    const float *w1max;         // max_i w1[i]
    const float *wscale;        // FP16 format: the codebook's power-of-two scale G (device scalar)
    float err_scale;            // kappa * 2^-24
    int32_t *idx_out;
    int32_t *flag_list;
    int32_t *flag_count;
    float *flag_ub;
    unsigned long long *min_key;
    float *dump;                // debug: all distortions [n][dump_ld]
    int dump_ld;
    // filter mode (re-evaluation of flagged frames over the gather list): no argmin; every
    // (list position, codeword) whose lower bound is <= the frame's upper bound flag_ub[pos]
    // of the true minimum becomes a candidate for the FP64 evaluation
    int2 *cand;
    int32_t *cand_count;
    int cand_cap;
    int32_t *exact_list;        // frames whose candidates did not fit: all-FP64 search
    int32_t *exact_count;
    int mc;                     // 1: launched as clusters of 2 CTAs that share the codeword stream
    int raw_bufs;               // variant 2: staging buffers for raw frame halves (1 .. 8)
    // variant 2, FP16 format: the frames ALREADY as tensor-core operands (launch_prepare_frames_tc:
    // per 128-frame half-tile the scaled (hi, lo) parts in canonical K-major order + the 128 norms),
    // built once per frame set.  A half-tile is then two bulk copies straight into the operand
    // buffer: no conversion work per search at all (the loader warps idle).
    const unsigned char *xop;
    // fused cell sums (CELLS instance): the FP64 row of every frame OUTSIDE the error bound is added
    // to its cell with FP64 reductions (RED) by the loader warps while the tensor core is busy with
    // later tiles; flagged frames are added after their re-evaluation (vq_cell_fixup)
    const double *x64;
    double *cells;              // packed cell sums [K][D]
    int32_t *hist;              // frames per cell
    // several codebooks in one sweep (SEG instance; cbook_quantize of R models, vq/vq.c:154-157,
    // 285-315): the K columns are R codebooks of seg_cols codewords each (seg_cols = 32 .. BN dividing
    // BN, or a multiple of BN); the argmin restarts at every boundary, idx_out is [R][idx_stride] with
    // model-local indices, frames inside the bound are reported as keys (model << 40 | frame)
    int seg_cols;
    long idx_stride;
    unsigned long long *seg_key;
    int seg_cap;
    unsigned long long *prof;   // dev (exp bit 256): per CTA {SM clocks, ns} of the kernel
    int exp;                    // dev experiments (ECOZ2_VQ_TC_EXP): bit mask 1 no record path, 2 trivial reduction, 4 no tcgen05.ld, 8 no MMA, 16 no tcgen05.st, 32 no B copy
};

// Global layout of the search operand built by vq_prepare_codebook_tc:
//   stage (t, s) = codeword tile t, k-step s, contiguous 4 * BN * 16 bytes:
//     [part: hi, lo][k-chunk c: 0, 1][row r: 0..BN)[e: 0..4)  = element k = 8 s + 4 c + e
//   of codeword t * BN + r (zero beyond K / beyond D).
// Shared memory (dynamic):
//   [barriers, tmem slot | 1 KB][x0: 2 x 128 f32 | 1 KB][epilogue merge slots: 2 x 128 float4 | 4 KB]
//   [A: 2 buffers x (2 parts x 2 ksteps chunks x 128 rows x 16 B)][B: kStages x 4 BN x 16 B]
struct TcSmem {
    uint64_t full[kStages], empty[kStages];
    uint64_t tmem_full[2], tmem_empty[2];
    uint64_t a_full[2], a_empty[2];
    uint64_t raw_full;
    uint32_t tmem_slot;
};

// filter mode (see TcParams::cand): a frame's candidates are collected in registers (a
// flagged frame typically has 2-3) and published once per frame tile with one atomic per
// warp; a frame with more than kFilterSlots candidates goes to the all-FP64 search instead.
constexpr int kFilterSlots = 12;
struct FilterState {
    int cnt;
    int c[kFilterSlots];
    __device__ __forceinline__ void init() {
        cnt = 0;
#pragma unroll
        for (int k = 0; k < kFilterSlots; k++) c[k] = 0;
    }
    __device__ __forceinline__ void push(int col) {
#pragma unroll
        for (int k = kFilterSlots - 1; k > 0; k--) c[k] = c[k - 1];
        c[0] = col;
        cnt++;
    }
};

__device__ __forceinline__ void epi_filter32(const uint32_t (&r)[32], int col0, int nv_rel, float thr,
                                             FilterState &fs) {
#define F(i) __uint_as_float(r[i])
    float m = INFINITY;
#pragma unroll
    for (int j = 0; j < 32; j += 2) m = min3f(m, F(j), F(j + 1));
    if (nv_rel < 32) m = -INFINITY;   // ragged end: take the slow path, it masks per column
    if (__any_sync(0xffffffffu, m <= thr)) {
        uint32_t mask = 0;
#pragma unroll
        for (int j = 0; j < 32; j++) mask |= (j < nv_rel && F(j) <= thr) ? (1u << j) : 0u;
        while (mask) {
            const int j = __ffs(mask) - 1;
            mask &= mask - 1;
            fs.push(col0 + j);
        }
    }
#undef F
}

// publish the candidates of one warp's 32 frames; pos = list position of this lane's frame
__device__ __forceinline__ void epi_filter_flush(const FilterState &fs, int pos, long row,
                                                 const TcParams &p) {
    const int lane = threadIdx.x & 31;
    const bool over = fs.cnt > kFilterSlots;
    const int mine = (row >= 0 && !over) ? fs.cnt : 0;
    int incl = mine;
#pragma unroll
    for (int off = 1; off < 32; off <<= 1) {
        const int v = __shfl_up_sync(0xffffffffu, incl, off);
        if (lane >= off) incl += v;
    }
    const int total = __shfl_sync(0xffffffffu, incl, 31);
    int base = 0;
    if (lane == 31 && total > 0) base = atomicAdd(p.cand_count, total);
    base = __shfl_sync(0xffffffffu, base, 31);
    const int first = base + incl - mine;
    // a flagged frame always has its own best among the candidates; none at all means the
    // tensor-core values were unusable (NaN input ...): let the FP64 search decide
    bool spill = row >= 0 && (over || fs.cnt == 0);
#pragma unroll
    for (int k = 0; k < kFilterSlots; k++) {
        if (k < mine) {
            // c[0] is the newest; order inside the buffer is irrelevant
            if (first + k < p.cand_cap) p.cand[first + k] = make_int2(pos, fs.c[k]);
            else spill = true;
        }
    }
    if (spill) p.exact_list[atomicAdd(p.exact_count, 1)] = (int32_t) row;
}

// Columns [c_lo, c_hi) of one accumulator (thread = its TMEM lane), 32 at a time; the next
// tcgen05.ld is in flight while the current 32 values are reduced.  nv = valid codewords in
// this tile, tcol0 = codeword index of the tile's column 0.  Generic path: ragged tiles,
// filter mode, and (DBG) the distortion dump / experiment switches.
template <bool FILTER, bool DBG>
__device__ __forceinline__ void epi_tile(uint32_t taddr, int c_lo, int c_hi, int tcol0, int nv,
                                         EpiState &est, const TcParams &p, long row, float thr = 0.f,
                                         FilterState *fs = nullptr, float dump_mul = 1.f) {
    const int exp = DBG ? p.exp : 0;
    if (exp & 4) return;
    const int nv32 = (nv + 31) & ~31;
    const int c_end = nv32 < c_hi ? nv32 : c_hi;
    uint32_t ra[32], rb[32];
    int c = c_lo;
    if (c < c_end) tmem_ld32(taddr + c, ra);
#define VQ_TC_STEP(R)                                                                     \
    {                                                                                     \
        if (DBG && p.dump != nullptr && row >= 0) {                                       \
            _Pragma("unroll") for (int j = 0; j < 32; j++)                                \
                if (c + j < nv) p.dump[row * p.dump_ld + tcol0 + c + j] = dump_mul * __uint_as_float(R[j]); \
        }                                                                                 \
        if (FILTER) epi_filter32(R, tcol0 + c, nv - c, thr, *fs);                         \
        else epi_step32(R, tcol0 + c, nv - c, est, exp);                                  \
    }
    while (c < c_end) {
        tmem_ld_wait();
        if (c + 32 < c_end) tmem_ld32(taddr + c + 32, rb);
        VQ_TC_STEP(ra);
        c += 32;
        if (c >= c_end) break;
        tmem_ld_wait();
        if (c + 32 < c_end) tmem_ld32(taddr + c + 32, ra);
        VQ_TC_STEP(rb);
        c += 32;
    }
#undef VQ_TC_STEP
}

__global__ void __launch_bounds__(kThreads, 1) vq_assign_tc(const TcParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    TcSmem *sm = reinterpret_cast<TcSmem *>(smem_raw);
    float *s_x0 = reinterpret_cast<float *>(smem_raw + 1024);
    float4 *s_merge = reinterpret_cast<float4 *>(smem_raw + 2048);
    const int ksteps = p.ksteps;
    const int KC = 2 * ksteps;                              // 16-byte k-chunks per row
    const uint32_t a_part_bytes = (uint32_t) KC * kBM * 16;  // one part (hi or lo) of a frame tile
    const uint32_t a_buf_bytes = 2 * a_part_bytes;
    const int BN = p.BN;
    const uint32_t stage_bytes = (uint32_t) 4 * BN * 16;
    unsigned char *a_base = smem_raw + 6144;
    unsigned char *b_base = a_base + 2 * a_buf_bytes;
    unsigned char *raw_base = b_base + (size_t) kStages * stage_bytes;   // one tile of raw FP32 rows

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int K = p.K;
    const int n_wtiles = (K + BN - 1) / BN;
    const long total = p.list ? (long) *p.list_count : p.n_frames;
    const long n_ftiles = (total + kBM - 1) / kBM;
    // cluster mode: the two CTAs of a pair run the same number of frame tiles (the odd CTA's
    // last one may be empty) because they consume the shared codeword stream in lock step
    const long lead_off = p.mc ? (long) (blockIdx.x & 1) : 0;
    const uint32_t cta_rank = p.mc ? cluster_ctarank() : 0;

    if (tid == 0) {
        for (int i = 0; i < kStages; i++) {
            mbar_init(&sm->full[i], 1);
            mbar_init(&sm->empty[i], p.mc ? 2 : 1);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&sm->tmem_full[i], 1);
            mbar_init(&sm->tmem_empty[i], kEpiWarps);
            mbar_init(&sm->a_full[i], kLoaderThreads);
            mbar_init(&sm->a_empty[i], 1 + kEpiWarps);
        }
        mbar_init(&sm->raw_full, 1);
        fence_barrier_init();
    }
    if (warp == 1) tmem_alloc(&sm->tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    if (p.mc) cluster_sync_all();   // the peer's barriers exist before anything is sent to them
    tc_fence_after();
    const uint32_t tmem_base = sm->tmem_slot;

    if (warp == 0) {
        // ===== B producer =====
        if (lane == 0) {
            uint32_t st = 0, ph = 0;
            for (long ft = blockIdx.x; ft - lead_off < n_ftiles; ft += gridDim.x) {
                const float *src = p.wtc;
                for (int t = 0; t < n_wtiles; t++) {
                    for (int s = 0; s < ksteps; s++) {
                        mbar_wait(&sm->empty[st], ph ^ 1);
                        mbar_expect_tx(&sm->full[st], stage_bytes);
                        if (p.mc) {
                            // each CTA of the pair fetches half of the stage for both
                            const uint32_t half = stage_bytes / 2;
                            bulk_g2s_mc(b_base + (size_t) st * stage_bytes + cta_rank * half,
                                        reinterpret_cast<const unsigned char *>(src) + cta_rank * half,
                                        half, &sm->full[st], (uint16_t) 3);
                        } else {
                            bulk_g2s(b_base + (size_t) st * stage_bytes, src, stage_bytes, &sm->full[st]);
                        }
                        src += (size_t) BN * 16;
                        if (++st == kStages) { st = 0; ph ^= 1; }
                    }
                }
            }
        }
    } else if (warp == 1) {
        // ===== MMA issuer =====
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_tf32(kBM, BN);
            uint32_t st = 0, ph = 0;
            uint32_t tc = 0;  // accumulator tiles issued so far
            uint32_t it = 0;
            for (long ft = blockIdx.x; ft - lead_off < n_ftiles; ft += gridDim.x, it++) {
                const uint32_t ab = it & 1;
                mbar_wait(&sm->a_full[ab], (it >> 1) & 1);
                tc_fence_after();
                const uint32_t a_hi = s32(a_base + (size_t) ab * a_buf_bytes);
                const uint32_t a_lo = a_hi + a_part_bytes;
                for (int t = 0; t < n_wtiles; t++, tc++) {
                    const uint32_t acc = tc & 1;
                    mbar_wait(&sm->tmem_empty[acc], ((tc >> 1) & 1) ^ 1);
                    tc_fence_after();
                    const uint32_t d_tmem = tmem_base + acc * 256;
                    for (int s = 0; s < ksteps; s++) {
                        mbar_wait(&sm->full[st], ph);
                        tc_fence_after();
                        const uint32_t b_hi = s32(b_base + (size_t) st * stage_bytes);
                        const uint32_t b_lo = b_hi + 2 * BN * 16;
                        const uint32_t koff = (uint32_t) (2 * s) * kBM * 16;
                        const uint64_t dah = umma_desc(a_hi + koff, kBM * 16, 128);
                        const uint64_t dal = umma_desc(a_lo + koff, kBM * 16, 128);
                        const uint64_t dbh = umma_desc(b_hi, BN * 16, 128);
                        const uint64_t dbl = umma_desc(b_lo, BN * 16, 128);
                        // small terms first
                        if (!(p.exp & 8)) {
                            umma_tf32(d_tmem, dal, dbh, idesc, s > 0);
                            umma_tf32(d_tmem, dah, dbl, idesc, 1);
                            umma_tf32(d_tmem, dah, dbh, idesc, 1);
                        }
                        if (p.mc) umma_commit_mc(&sm->empty[st], (uint16_t) 3);
                        else umma_commit(&sm->empty[st]);
                        if (++st == kStages) { st = 0; ph ^= 1; }
                    }
                    umma_commit(&sm->tmem_full[acc]);
                }
                umma_commit(&sm->a_empty[ab]);
            }
        }
    } else if (warp < kEpiWarp0) {
        // ===== A loaders: FP32 rows -> TF32 (hi, lo) -> canonical K-major layout =====
        // Full contiguous tiles arrive as ONE bulk copy into a raw staging buffer (issued a
        // tile ahead by loader thread 0); each loader thread then converts two rows.  Gathered
        // (list) or partial tiles are read straight from global memory.
        const int lt = tid - 64;
        const int D = p.D;
        const uint32_t raw_bytes = (uint32_t) kBM * D * 4;
        const bool contiguous = p.list == nullptr;
        if (lt == 0 && contiguous && (long) blockIdx.x * kBM + kBM <= total) {
            mbar_expect_tx(&sm->raw_full, raw_bytes);
            bulk_g2s(raw_base, p.x + (long) blockIdx.x * kBM * D, raw_bytes, &sm->raw_full);
        }
        uint32_t it = 0, rs = 0;
        for (long ft = blockIdx.x; ft - lead_off < n_ftiles; ft += gridDim.x, it++) {
            const uint32_t ab = it & 1;
            const long base = ft * kBM;
            const int nvalid = (int) (total - base < kBM ? (total - base > 0 ? total - base : 0) : kBM);
            const bool staged = contiguous && nvalid == kBM;
            if (staged) mbar_wait(&sm->raw_full, rs & 1);
            mbar_wait(&sm->a_empty[ab], ((it >> 1) & 1) ^ 1);
            unsigned char *hi = a_base + (size_t) ab * a_buf_bytes;
            unsigned char *lo = hi + a_part_bytes;
            for (int f = lt; f < kBM; f += kLoaderThreads) {
                const float *src;
                if (staged) {
                    src = reinterpret_cast<const float *>(raw_base) + f * D;
                } else {
                    const long row = f < nvalid ? (p.list ? (long) p.list[base + f] : base + f) : 0;
                    src = p.x + row * D;
                }
                const bool valid = f < nvalid;
                // norm of the error bound: max_k |x[k]| (= |rx[0]| for an autocorrelation row)
                float x0 = 0.f;
                if (valid)
                    for (int k = 0; k < D; k++) x0 = fmaxf(x0, fabsf(src[k]));
                for (int c = 0; c < KC; c++) {
                    float v[4];
#pragma unroll
                    for (int e = 0; e < 4; e++) {
                        const int k = 4 * c + e;
                        v[e] = (valid && k < D) ? src[k] : 0.f;
                        // slot D carries -e |x0|: with W1_j in the codeword's slot D the
                        // accumulator becomes the lower bound d~_ij - e |x0| W1_j
                        if (k == D) v[e] = -1.001f * p.err_scale * x0;
                    }
                    float4 h, l;
                    h.x = tf32_rna(v[0]); h.y = tf32_rna(v[1]); h.z = tf32_rna(v[2]); h.w = tf32_rna(v[3]);
                    l.x = tf32_rna(v[0] - h.x); l.y = tf32_rna(v[1] - h.y);
                    l.z = tf32_rna(v[2] - h.z); l.w = tf32_rna(v[3] - h.w);
                    *reinterpret_cast<float4 *>(hi + ((size_t) c * kBM + f) * 16) = h;
                    *reinterpret_cast<float4 *>(lo + ((size_t) c * kBM + f) * 16) = l;
                }
                s_x0[ab * kBM + f] = x0;
            }
            fence_proxy_async();
            mbar_arrive(&sm->a_full[ab]);
            if (staged) {
                rs++;
                // every loader thread is done with the staging buffer: fetch the next tile
                asm volatile("bar.sync 5, 64;" ::: "memory");
                const long nbase = (ft + gridDim.x) * kBM;
                if (lt == 0 && nbase + kBM <= total) {
                    mbar_expect_tx(&sm->raw_full, raw_bytes);
                    bulk_g2s(raw_base, p.x + nbase * D, raw_bytes, &sm->raw_full);
                }
            }
        }
    } else {
        // ===== epilogue: thread = frame (TMEM lane), registers = codewords =====
        const int q = warp & 3;                 // TMEM lane quarter this warp may read
        const int h = (warp - kEpiWarp0) >> 2;  // which half of every tile's columns
        const int f = q * 32 + lane;            // frame within the tile
        const uint32_t lane_addr = tmem_base + ((uint32_t) (q * 32) << 16);
        const int n_chunks = BN / 16;
        const int c_lo = h * (n_chunks / 2), c_hi = c_lo + n_chunks / 2;
        const float e2 = 2.f * 1.001f * p.err_scale;
        uint32_t tc = 0, it = 0;
        for (long ft = blockIdx.x; ft - lead_off < n_ftiles; ft += gridDim.x, it++) {
            const uint32_t ab = it & 1;
            const long base = ft * kBM;
            const int nvalid = (int) (total - base < kBM ? (total - base > 0 ? total - base : 0) : kBM);
            mbar_wait(&sm->a_full[ab], (it >> 1) & 1);
            const float x0 = s_x0[ab * kBM + f];
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm->a_empty[ab]);
            const long row = f < nvalid ? (p.list ? (long) p.list[base + f] : base + f) : -1;

            EpiState est;
            est.init();
            for (int t = 0; t < n_wtiles; t++, tc++) {
                const uint32_t acc = tc & 1;
                mbar_wait(&sm->tmem_full[acc], (tc >> 1) & 1);
                tc_fence_after();
                const int nv = K - t * BN < BN ? K - t * BN : BN;   // valid codewords in this tile
                epi_tile<false, true>(lane_addr + acc * 256, c_lo * 16, c_hi * 16, t * BN, nv, est, p, row);
                tc_fence_before();
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm->tmem_empty[acc]);
            }
            float best = est.best, second;
            int bidx;
            epi_finish(est, second, bidx);
            // merge the two column halves of each frame (upper half hands over via smem)
            float4 *slot = s_merge + (it & 1) * kBM + f;
            if (h == 1) *slot = make_float4(best, second, __int_as_float(bidx), 0.f);
            asm volatile("bar.sync %0, 64;" ::"r"(1 + q) : "memory");
            if (h == 0) {
                const float4 o = *slot;
                const int oi = __float_as_int(o.z);
                second = min3f(second, o.y, fmaxf(best, o.x));
                if (o.x < best || (o.x == best && oi < bidx)) { best = o.x; bidx = oi; }
                if (row >= 0) {
                    p.idx_out[row] = bidx;
                    if (p.flag_list != nullptr) {
                        // accumulator = lower bound L_ij = d~_ij - e |x0| W1_j of the true
                        // distortion; bidx is certainly the true argmin when every other lower
                        // bound exceeds its upper bound L + 2 e |x0| W1.  NaN-safe.
                        const float ub = fmaf(e2 * x0, p.w1[bidx], best);
                        if (!(second > ub)) {
                            const int pos = atomicAdd(p.flag_count, 1);
                            p.flag_list[pos] = (int32_t) row;
                            p.flag_ub[pos] = ub;
                            p.min_key[pos] = ~0ull;
                        }
                    }
                }
            }
        }
    }

    tc_fence_before();
    __syncthreads();
    if (p.mc) cluster_sync_all();   // nothing of the peer is still in flight towards this CTA
    if (warp == 1) tmem_dealloc(tmem_base, 512);
}

// =====================================================================================
// Variant 2: 256-frame tiles.
//
// Measured on B200 (scripts/exp): one SM ingests at most ~45 B/clk of bulk copies, and
// variant 1 needs 43 B/clk of codeword stages at the MMA floor -- it is bound by that stream.
// Here a frame tile is 256 frames = two UMMA M=128 halves that share every codeword tile, so
// the stream is halved.  Both operands come from shared memory (a tcgen05.mma.kind::tf32 with
// the A operand in TMEM was measured at 76-81 clk per 128x128x8 instead of 64), the codeword
// tile is whole (all k-steps), so the 3 KS MMAs of a tile run small cross terms first, and
// each half owns two 128-column accumulators: the tensor core never waits for an epilogue
// that keeps up on average.
//
//   warps 0-3    frame loaders, thread = frame row: raw FP32 row (TMA-staged in smem, or
//                gathered from global) -> TF32 hi/lo -> canonical K-major A[half] in smem
//   warps 4-7    epilogue of half 0,  warps 8-11 epilogue of half 1 (warp = TMEM lane quarter)
//   warp 12      codeword-tile producer (cp.async.bulk, whole tiles, kB2Stages ring)
//   warp 13      MMA issuer
//   warp 14      raw-frame producer (cp.async.bulk of 128 contiguous rows)
//
// TMEM columns: accumulator[half h][buffer b] at (2 b + h) * 128.
// Global codeword layout (vq_prepare_codebook_tc2): [tile][hi, lo][k-chunk][row][4 floats].
// =====================================================================================
namespace {
constexpr int k2Threads = 512;      // 4 warpgroups (setmaxnreg is per warpgroup); warp 15 idles
constexpr int kEpiRegs = 168, kLoaderRegs = 96, kRoleRegs = 80;   // 256*168 + 128*96 + 128*80 = 65536
constexpr int kB2Stages = 3;

struct Tc2Smem {
    uint64_t cell_full[8];                // [2 * (tile & 3) + half]: the half's cell indices are in s_cell
    uint64_t cell_free[4];                // [tile & 3]: the loader warps have added that tile's rows (pre-staged operands)
    uint64_t b_full[kB2Stages], b_empty[kB2Stages];
    uint64_t raw_full[8], raw_empty[8];   // raw-frame staging buffers (TcParams::raw_bufs of them)
    uint64_t a_full[4], a_free[4];        // [2 * frame-tile buffer + half]
    uint64_t acc_full[4], acc_empty[4];   // [2 * buffer + half]
    uint32_t tmem_slot;
};
}  // namespace

template <int KS, bool FILTER, bool DBG, bool H16, bool CELLS, bool SEG = false>
__global__ void __launch_bounds__(k2Threads, 1) vq_assign_tc2(const TcParams p) {
    const int exp = DBG ? p.exp : 0;
    unsigned long long prof_c0 = 0, prof_t0 = 0;
    if (DBG && p.prof != nullptr && threadIdx.x == 0) {
        prof_c0 = clock64();
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(prof_t0));
    }
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    Tc2Smem *sm = reinterpret_cast<Tc2Smem *>(smem_raw);
    float *s_x0 = reinterpret_cast<float *>(smem_raw + 1024);      // [4 slots][2 halves][128]
    const int D = p.D;
    const int BN = p.BN;
    const bool wide = BN == 256;                  // UMMA N = 256: see tc_tile_cols
    const uint32_t n_bstages = wide ? 2 : kB2Stages;
    const bool pre = H16 && !DBG && !FILTER && p.xop != nullptr;   // pre-staged frame operands
    // norms of the frame tiles in flight: [slot][half][128]; 8 slots when the producer can run far
    // ahead of the epilogue's once-per-tile finish (pre-staged operands), else 4
    const uint32_t x0_mask = pre ? 7u : 3u;
    // KS = MMAs per product term (TF32 format, 8 elements each) or 8-element chunks per split
    // part (FP16 format); either way a part of a row is KC 16-byte chunks
    constexpr int KC = H16 ? KS : 2 * KS;
    // FP16 format: a row is [hi: KC chunks][lo: KC chunks], a codeword row [lo][hi][zero chunk if
    // KC is odd].  The two cross terms are ONE product over the concatenated 2 KC chunks
    // ([xh|xl] . [wl|wh]), the main term pairs chunks (2s, 2s+1) of [xh|xl..] with [wh|0]:
    // KC + ceil(KC/2) MMAs of 16 elements instead of 3 ceil(KC/2).
    constexpr int BC = H16 ? 2 * KC + (KC & 1) : 2 * KC;             // chunks per codeword row
    constexpr int EPC = H16 ? 8 : 4;                                 // elements per chunk
    constexpr int NE = KC * EPC;                                     // padded row length
    constexpr uint32_t hdr_bytes = H16 ? 13312 : 5120;               // + per-frame scales, cell indices (FP16 format)
    float *s_scale = reinterpret_cast<float *>(smem_raw + 5120);     // [4 slots][2 halves][128], H16 only
    int32_t *s_cell = reinterpret_cast<int32_t *>(smem_raw + 9216);  // [4 slots][2 halves][128], CELLS only
    constexpr uint32_t a_part_bytes = (uint32_t) KC * kBM * 16;      // hi or lo of one half
    constexpr uint32_t a_half_bytes = 2 * a_part_bytes;
    const uint32_t raw_bytes = (uint32_t) kBM * D * 4;               // 128 rows
    const uint32_t raw_stride = (raw_bytes + 127u) & ~127u;
    unsigned char *raw_base = smem_raw + hdr_bytes;
    const uint32_t RB = (uint32_t) p.raw_bufs;                       // 2 when shared memory allows
    unsigned char *a_base = smem_raw + hdr_bytes + RB * raw_stride;
    const uint32_t tile_bytes = (uint32_t) BC * BN * 16;
    // the FP16 format leaves room for two frame tiles: the loaders run a whole tile ahead
    constexpr int ABUF = H16 ? 2 : 1;
    unsigned char *b_base = a_base + ABUF * 2 * a_half_bytes;

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const int K = p.K;
    const int n_wtiles = (K + BN - 1) / BN;
    const long total = p.list ? (long) *p.list_count : p.n_frames;
    const long n_ftiles = (total + 2 * kBM - 1) / (2 * kBM);

    if (tid == 0) {
        for (int i = 0; i < kB2Stages; i++) {
            mbar_init(&sm->b_full[i], 1);
            mbar_init(&sm->b_empty[i], 1);
        }
        for (int i = 0; i < 8; i++) {
            mbar_init(&sm->raw_full[i], 1);
            mbar_init(&sm->raw_empty[i], 4);
        }
        for (int i = 0; i < 4; i++) {
            mbar_init(&sm->a_full[i], pre ? 1 : 4);
            mbar_init(&sm->a_free[i], 1);
        }
        for (int i = 0; i < 4; i++) {
            mbar_init(&sm->acc_full[i], 1);
            mbar_init(&sm->acc_empty[i], 4);
        }
        for (int i = 0; i < 8; i++) mbar_init(&sm->cell_full[i], 4);
        for (int i = 0; i < 4; i++) mbar_init(&sm->cell_free[i], 4);
        fence_barrier_init();
    }
    if (warp == 13) tmem_alloc(&sm->tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = sm->tmem_slot;

    // register budget by role (warpgroups of 4 warps): the epilogue holds two 32-column
    // TMEM loads, the snapshot chunk and its state in registers
    if (warp >= 12) {
        setmaxnreg_dec<kRoleRegs>();
    }
    if (warp == 12) {
        // ===== codeword tile producer =====
        if (lane == 0) {
            uint32_t st = 0, ph = 0;
            for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x) {
                const unsigned char *src = reinterpret_cast<const unsigned char *>(p.wtc);
                for (int t = 0; t < n_wtiles; t++) {
                    mbar_wait_relaxed(&sm->b_empty[st], ph ^ 1);
                    if (exp & 32) {
                        mbar_arrive(&sm->b_full[st]);
                    } else {
                        mbar_expect_tx(&sm->b_full[st], tile_bytes);
                        bulk_g2s(b_base + (size_t) st * tile_bytes, src, tile_bytes, &sm->b_full[st]);
                    }
                    src += tile_bytes;
                    if (++st == n_bstages) { st = 0; ph ^= 1; }
                }
            }
        }
    } else if (warp == 14) {
        // ===== raw frame producer (full, contiguous halves only) =====
        if (pre) {
            // pre-staged operands: the half-tile record [hi | lo parts, a_half_bytes][128 norms] goes
            // straight into the operand buffer and the norm slot
            if (lane == 0) {
                const size_t rec = (size_t) a_half_bytes + kBM * sizeof(float);
                uint32_t it = 0;
                for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x, it++) {
                    // fused cell sums: the loader warps no longer pace the pipeline (they only add
                    // rows), so the ring of four cell-index slots needs its own back-pressure -- tile
                    // `it` may start only when the rows of tile it - 4 have been added
                    if (CELLS && it >= 4) mbar_wait(&sm->cell_free[it & 3], ((it >> 2) + 1) & 1);
                    for (int h = 0; h < 2; h++) {
                        const uint32_t ab = it & 1, aph = (it >> 1) & 1;
                        mbar_wait(&sm->a_free[2 * ab + h], aph ^ 1);
                        const unsigned char *src = p.xop + (size_t) (2 * ft + h) * rec;
                        mbar_expect_tx(&sm->a_full[2 * ab + h], (uint32_t) rec);
                        bulk_g2s(a_base + (size_t) (2 * ab + h) * a_half_bytes, src, a_half_bytes, &sm->a_full[2 * ab + h]);
                        bulk_g2s(s_x0 + ((it & x0_mask) * 2 + h) * kBM, src + a_half_bytes, kBM * sizeof(float),
                                 &sm->a_full[2 * ab + h]);
                    }
                }
            }
        } else if (lane == 0 && p.list == nullptr && !(exp & 64)) {
            uint32_t rs = 0;   // staged halves so far
            for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x) {
                for (int h = 0; h < 2; h++) {
                    const long base = ft * (2 * kBM) + h * kBM;
                    if (base + kBM > total) continue;
                    const uint32_t rb = rs % RB, rph = (rs / RB) & 1;
                    mbar_wait_relaxed(&sm->raw_empty[rb], rph ^ 1);
                    mbar_expect_tx(&sm->raw_full[rb], raw_bytes);
                    bulk_g2s(raw_base + rb * raw_stride, p.x + base * D, raw_bytes, &sm->raw_full[rb]);
                    rs++;
                }
            }
        }
    } else if (warp == 13) {
        // ===== MMA issuer =====
        // The whole warp runs the loop (uniform control flow keeps descriptors in uniform
        // registers); one elected lane issues the MMAs and their commits.
        const uint32_t idesc = H16 ? umma_idesc_f16(kBM, BN) : umma_idesc_tf32(kBM, BN);
        // descriptor of (half 0, hi, k-step 0) / (stage 0, hi, k-step 0); the start-address
        // field counts 16-byte units, so other operands are reached by adding bytes >> 4
        const uint64_t dA0 = umma_desc(s32(a_base), kBM * 16, 128);
        const uint64_t dB0 = umma_desc(s32(b_base), BN * 16, 128);
        constexpr uint32_t a_half16 = a_half_bytes >> 4, a_part16 = a_part_bytes >> 4;
        constexpr uint32_t a_k16 = (2 * kBM * 16) >> 4;
        const uint32_t b_tile16 = tile_bytes >> 4, b_part16 = (uint32_t) (KC * BN * 16) >> 4;
        const uint32_t b_k16 = (uint32_t) (2 * BN * 16) >> 4;
        uint32_t st = 0, ph = 0, u = 0, it = 0;
        long long pw_a = 0, pw_e = 0, pw_b = 0;   // dev (exp bit 256): clocks spent waiting
        for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x, it++) {
            for (int t = 0; t < n_wtiles; t++, u++) {
                // 128-column tiles: two accumulators per half (ping-pong); 256-column tiles: one per
                // half, the halves themselves alternate between the tensor core and the epilogue
                const uint32_t buf = wide ? 0u : (u & 1), acc_ph = wide ? (u & 1) : ((u >> 1) & 1);
                for (int h = 0; h < 2; h++) {
                    const uint32_t ab = ABUF == 2 ? (it & 1) : 0, aph = ABUF == 2 ? (it >> 1) & 1 : it & 1;
                    const long long pc0 = DBG ? clock64() : 0;
                    if (t == 0) mbar_wait(&sm->a_full[2 * ab + h], aph);
                    const long long pc1 = DBG ? clock64() : 0;
                    mbar_wait(&sm->acc_empty[2 * buf + h], acc_ph ^ 1);
                    const long long pc2 = DBG ? clock64() : 0;
                    if (h == 0) mbar_wait(&sm->b_full[st], ph);
                    if (DBG) {
                        const long long pc3 = clock64();
                        pw_a += pc1 - pc0; pw_e += pc2 - pc1; pw_b += pc3 - pc2;
                    }
                    tc_fence_after();
                    const uint64_t dah = dA0 + (uint32_t) (2 * ab + h) * a_half16, dal = dah + a_part16;
                    const uint64_t dbh = dB0 + st * b_tile16, dbl = dbh + b_part16;
                    const uint32_t d_tmem = tmem_base + (wide ? (uint32_t) h * 256u : (2 * buf + h) * 128u);
                    if (elect_one()) {
                        // small cross terms first, the dominant hi*hi products last
                        if (!(exp & 8)) {
#define VQ_UMMA(D_, A_, B_, ACC_)                                  \
    do {                                                           \
        if (H16) umma_f16(D_, A_, B_, idesc, ACC_);                \
        else umma_tf32(D_, A_, B_, idesc, ACC_);                   \
    } while (0)
                            if (H16) {
                                // dah = [xh|xl] of this half, dbh = [wl|wh|0] of this tile
#pragma unroll
                                for (int s = 0; s < KC; s++)
                                    VQ_UMMA(d_tmem, dah + s * a_k16, dbh + s * b_k16, s > 0);
#pragma unroll
                                for (int s = 0; s < (KC + 1) / 2; s++)
                                    VQ_UMMA(d_tmem, dah + s * a_k16, dbl + s * b_k16, 1);
                            } else {
#pragma unroll
                                for (int s = 0; s < KS; s++)
                                    VQ_UMMA(d_tmem, dal + s * a_k16, dbh + s * b_k16, s > 0);
#pragma unroll
                                for (int s = 0; s < KS; s++)
                                    VQ_UMMA(d_tmem, dah + s * a_k16, dbl + s * b_k16, 1);
#pragma unroll
                                for (int s = 0; s < KS; s++)
                                    VQ_UMMA(d_tmem, dah + s * a_k16, dbh + s * b_k16, 1);
                            }
#undef VQ_UMMA
                        }
                        umma_commit(&sm->acc_full[2 * buf + h]);
                        if (t == n_wtiles - 1) umma_commit(&sm->a_free[2 * ab + h]);
                        if (h == 1) umma_commit(&sm->b_empty[st]);
                    }
                    __syncwarp();
                    if (h == 1 && ++st == n_bstages) { st = 0; ph ^= 1; }
                }
            }
        }
        if (DBG && p.prof != nullptr && lane == 0) {
            unsigned long long *o = p.prof + 2048 + 8 * blockIdx.x;
            o[0] = pw_a; o[1] = pw_e; o[2] = pw_b;
        }
    } else if (warp < 4) {
        // ===== frame loaders: thread = frame row =====
        setmaxnreg_dec<kLoaderRegs>();
        const int r = warp * 32 + lane;
        uint32_t it = 0, rs = 0;
        // CELLS: the rows of tile `t` (this CTA's t-th tile) that fell OUTSIDE the error bound are added
        // to their cells two tiles later, when the epilogue has long finished with it; warp w takes rows
        // 32 w .. 32 w + 31 of both halves, lanes = components, 8 rows in flight
        auto accumulate_tile = [&](uint32_t t) {
            const long ft = (long) blockIdx.x + (long) t * gridDim.x;
            for (int h = 0; h < 2; h++) {
                const uint32_t slot = (t & 3) * 2 + h;
                mbar_wait_relaxed(&sm->cell_full[slot], (t >> 2) & 1);
                const int cmine = s_cell[slot * kBM + r];
                const long row0 = ft * (2 * kBM) + h * kBM + warp * 32;
                // the rows this warp adds NEXT time (tile t + 1) start their way from HBM to L2 now
                {
                    const long nrow0 = row0 + (long) gridDim.x * (2 * kBM);
                    if (nrow0 + 32 <= total) {
                        const char *nb = reinterpret_cast<const char *>(p.x64 + nrow0 * D);
                        for (int o = lane * 128; o < 32 * D * 8; o += 32 * 128)
                            asm volatile("prefetch.global.L2 [%0];" ::"l"(nb + o));
                    }
                }
                if (__ballot_sync(0xffffffffu, cmine >= 0) == 0u) continue;
#pragma unroll 1
                for (int u = 0; u < 32; u += 4) {
                    int cu[4];
                    double xa[4], xb[4];
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        cu[k] = __shfl_sync(0xffffffffu, cmine, u + k);
                        const double *xr = p.x64 + (row0 + u + k) * D;
                        xa[k] = (cu[k] >= 0 && lane < D) ? xr[lane] : 0.0;
                        xb[k] = (cu[k] >= 0 && lane + 32 < D) ? xr[lane + 32] : 0.0;
                    }
#pragma unroll
                    for (int k = 0; k < 4; k++) {
                        if (cu[k] >= 0) {
                            double *dst = p.cells + (long) cu[k] * D;
                            if (lane < D) atomicAdd(dst + lane, xa[k]);
                            if (lane + 32 < D) atomicAdd(dst + lane + 32, xb[k]);
                        }
                    }
                }
                if (cmine >= 0) atomicAdd(p.hist + cmine, 1);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm->cell_free[t & 3]);
        };
        for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x, it++) {
            if (CELLS && it >= 2) accumulate_tile(it - 2);
            if (pre) continue;   // the operands come ready-made from the producer warp
            for (int h = 0; h < 2; h++) {
                const long base = ft * (2 * kBM) + h * kBM;
                const bool staged = p.list == nullptr && base + kBM <= total && !(exp & 64);
                float v[NE];
                if (staged) {
                    const uint32_t rb = rs % RB, rph = (rs / RB) & 1;
                    mbar_wait_relaxed(&sm->raw_full[rb], rph);
                    rs++;
                    const float *src = reinterpret_cast<const float *>(raw_base + rb * raw_stride) + r * D;
#pragma unroll
                    for (int k = 0; k < NE; k++) v[k] = k < D ? src[k] : 0.f;
                    // the staging buffer may be overwritten as soon as raw_empty completes: the
                    // arrive must not be issued before every lane's loads have RETURNED, so it is
                    // made data dependent on all of them (a vote over a value built from them)
                    uint32_t dep = 0;
#pragma unroll
                    for (int k = 0; k < NE; k++) dep ^= __float_as_uint(v[k]);
                    fence_proxy_async();   // generic reads before the async-proxy overwrite
                    const uint32_t all_in = __all_sync(0xffffffffu, dep != 0x7fc12345u);
                    if (lane == 0) mbar_arrive_dep(&sm->raw_empty[rb], all_in);
                } else {
                    const bool valid = base + r < total;
                    const long row = valid ? (p.list ? (long) p.list[base + r] : base + r) : 0;
                    const float *src = p.x + row * D;
#pragma unroll
                    for (int k = 0; k < NE; k++) v[k] = (valid && k < D) ? __ldg(src + k) : 0.f;
                }
                // reference magnitude of the error bound: max_k |x[k]| (TF32 format, values unscaled),
                // or the same after the frame was scaled by a power of two into [2^14, 2^15)
                // (FP16 format: both split parts then stay inside the FP16 range, and the
                // argmin of a frame does not depend on its own positive scale)
                float xm = 0.f;
#pragma unroll
                for (int k = 0; k < NE; k++) xm = fmaxf(xm, fabsf(v[k]));
                float x0 = xm;
                if (H16) {
                    uint32_t ef = (__float_as_uint(xm) >> 23) & 0xffu;
                    ef = ef < 16u ? 16u : ef;
                    const float scale = __uint_as_float((268u - ef) << 23);   // 2^(14 - floor(log2 xm))
#pragma unroll
                    for (int k = 0; k < NE; k++) v[k] *= scale;
                    x0 = xm * scale;
                    if (DBG) s_scale[((it & 3) * 2 + h) * kBM + r] = scale;
                }
                // slot D carries -e x0; with W1_j in the codeword's slot D the accumulator
                // becomes the lower bound d~_ij - e x0 W1_j
#pragma unroll
                for (int k = 0; k < NE; k++)
                    if (k == D) v[k] = -1.001f * p.err_scale * x0;
                const uint32_t ab = ABUF == 2 ? (it & 1) : 0, aph = ABUF == 2 ? (it >> 1) & 1 : it & 1;
                mbar_wait_relaxed(&sm->a_free[2 * ab + h], aph ^ 1);
                unsigned char *hi = a_base + (size_t) (2 * ab + h) * a_half_bytes;
                unsigned char *lo = hi + a_part_bytes;
                if (H16) {
#pragma unroll
                    for (int c = 0; c < KC; c++) {
                        uint32_t hw[4], lw[4];
#pragma unroll
                        for (int e = 0; e < 4; e++) {
                            const float a = v[8 * c + 2 * e], b2 = v[8 * c + 2 * e + 1];
                            const __half2 hh = __floats2half2_rn(a, b2);
                            const float2 hf = __half22float2(hh);
                            const __half2 ll = __floats2half2_rn(a - hf.x, b2 - hf.y);
                            hw[e] = *reinterpret_cast<const uint32_t *>(&hh);
                            lw[e] = *reinterpret_cast<const uint32_t *>(&ll);
                        }
                        *reinterpret_cast<uint4 *>(hi + ((size_t) c * kBM + r) * 16) =
                            make_uint4(hw[0], hw[1], hw[2], hw[3]);
                        *reinterpret_cast<uint4 *>(lo + ((size_t) c * kBM + r) * 16) =
                            make_uint4(lw[0], lw[1], lw[2], lw[3]);
                    }
                } else {
#pragma unroll
                    for (int c = 0; c < KC; c++) {
                        float4 hv, lv;
                        hv.x = tf32_rna(v[4 * c + 0]); hv.y = tf32_rna(v[4 * c + 1]);
                        hv.z = tf32_rna(v[4 * c + 2]); hv.w = tf32_rna(v[4 * c + 3]);
                        lv.x = tf32_rna(v[4 * c + 0] - hv.x); lv.y = tf32_rna(v[4 * c + 1] - hv.y);
                        lv.z = tf32_rna(v[4 * c + 2] - hv.z); lv.w = tf32_rna(v[4 * c + 3] - hv.w);
                        *reinterpret_cast<float4 *>(hi + ((size_t) c * kBM + r) * 16) = hv;
                        *reinterpret_cast<float4 *>(lo + ((size_t) c * kBM + r) * 16) = lv;
                    }
                }
                s_x0[((it & x0_mask) * 2 + h) * kBM + r] = x0;
                fence_proxy_async();
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm->a_full[2 * ab + h]);
            }
        }
        if (CELLS) {   // drain: the last two tiles
            if (it >= 2) accumulate_tile(it - 2);
            if (it >= 1) accumulate_tile(it - 1);
        }
    } else if (warp < 12) {
        // ===== epilogue: thread = frame (TMEM lane), registers = codewords =====
        setmaxnreg_inc<kEpiRegs>();
        const int q = warp & 3;
        const int h = (warp - 4) >> 2;
        const int r = q * 32 + lane;
        const uint32_t lane_addr = tmem_base + ((uint32_t) (q * 32) << 16);
        const float e2 = 2.f * 1.001f * p.err_scale;
        uint32_t u = 0, it = 0;
        long long pw_f = 0;   // dev (exp bit 256): clocks spent waiting for accumulators
        const long long pw_t0 = DBG ? clock64() : 0;
        for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x, it++) {
            const long base = ft * (2 * kBM) + h * kBM;
            const long row = base + r < total ? (p.list ? (long) p.list[base + r] : base + r) : -1;
            EpiState est;
            est.init();
            float thr = -INFINITY;
            FilterState fst;
            fst.init();
            if (FILTER && row >= 0) thr = p.flag_ub[base + r];
            // SEG: a codebook of the concatenation is complete -> its argmin, bound test, restart
            auto seg_done = [&](int col0) {
                const float sbest = est.best;
                float ssecond;
                int sidx;
                epi_finish(est, ssecond, sidx);
                if (row >= 0) {
                    const int model = col0 / p.seg_cols;
                    // (a frame of NaNs never records a minimum: keep the index inside the model)
                    const int local = sidx - col0;
                    p.idx_out[(long) model * p.idx_stride + row] = (unsigned) local < (unsigned) p.seg_cols ? local : 0;
                    const float x0 = s_x0[((it & x0_mask) * 2 + h) * kBM + r];
                    if (pre) fence_proxy_async();   // generic read before a later bulk copy reuses the slot
                    const float ub = fmaf(e2 * x0, p.w1[(unsigned) sidx < (unsigned) K ? sidx : 0], sbest);
                    if (!(ssecond > ub)) {
                        const int pos = atomicAdd(p.flag_count, 1);
                        if (pos < p.seg_cap)
                            p.seg_key[pos] = ((unsigned long long) model << 40) | (unsigned long long) row;
                    }
                }
                est.init();
            };
            for (int t = 0; t < n_wtiles; t++, u++) {
                const uint32_t buf = wide ? 0u : (u & 1), acc_ph = wide ? (u & 1) : ((u >> 1) & 1);
                const long long pc0 = DBG ? clock64() : 0;
                mbar_wait(&sm->acc_full[2 * buf + h], acc_ph);
                if (DBG) pw_f += clock64() - pc0;
                tc_fence_after();
                const int nv = K - t * BN < BN ? K - t * BN : BN;
                const uint32_t taddr = lane_addr + (wide ? (uint32_t) h * 256u : (2 * buf + h) * 128u);
                const bool fast = !FILTER && (!DBG || (!(exp & 7) && p.dump == nullptr));
                if (SEG) {   // whole tiles only (the host pads the concatenation)
                    if (p.seg_cols < BN) {
                        if (BN == 256) epi_tile_seg<256>(taddr, t * BN, p.seg_cols, est, &sm->acc_empty[2 * buf + h], seg_done);
                        else if (BN == 128) epi_tile_seg<128>(taddr, t * BN, p.seg_cols, est, &sm->acc_empty[2 * buf + h], seg_done);
                        else epi_tile_seg<64>(taddr, t * BN, p.seg_cols, est, &sm->acc_empty[2 * buf + h], seg_done);
                    } else {
                        if (BN == 256) epi_tile_full<256>(taddr, t * BN, est, &sm->acc_empty[2 * buf + h]);
                        else if (BN == 128) epi_tile_full<128>(taddr, t * BN, est, &sm->acc_empty[2 * buf + h]);
                        else epi_tile_full<64>(taddr, t * BN, est, &sm->acc_empty[2 * buf + h]);
                        if (((t + 1) * BN) % p.seg_cols == 0) seg_done((t + 1) * BN - p.seg_cols);
                    }
                } else if (fast && nv == 256) {
                    epi_tile_full<256>(taddr, t * BN, est, &sm->acc_empty[2 * buf + h]);
                } else if (fast && nv == 128) {
                    epi_tile_full<128>(taddr, t * BN, est, &sm->acc_empty[2 * buf + h]);
                } else if (fast && nv == 64) {
                    epi_tile_full<64>(taddr, t * BN, est, &sm->acc_empty[2 * buf + h]);
                } else {
                    float dump_mul = 1.f;   // debug dump in the caller's units
                    if (DBG && H16) dump_mul = 1.f / (s_scale[((it & 3) * 2 + h) * kBM + r] * *p.wscale);
                    epi_tile<FILTER, DBG>(taddr, 0, BN, t * BN, nv, est, p, row, thr, &fst, dump_mul);
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm->acc_empty[2 * buf + h]);
                }
            }
            if (FILTER) {
                epi_filter_flush(fst, (int) (base + r), row, p);
                continue;
            }
            if (SEG) continue;
            float best = est.best, second;
            int bidx;
            epi_finish(est, second, bidx);
            if (DBG && exp) bidx = (unsigned) bidx < (unsigned) K ? bidx : 0;   // experiments leave it undefined
            int cell = -1;
            if (row >= 0) {
                p.idx_out[row] = bidx;
                cell = bidx;
                if (p.flag_list != nullptr && !(exp & 128)) {
                    // accumulator = lower bound L_ij = d~_ij - e |x0| W1_j of the true distortion;
                    // bidx is certainly the true argmin when every other lower bound exceeds its
                    // upper bound L + 2 e |x0| W1.  NaN-safe.
                    const float x0 = s_x0[((it & x0_mask) * 2 + h) * kBM + r];
                    if (pre) fence_proxy_async();   // generic read before a later bulk copy reuses the slot
                    const float ub = fmaf(e2 * x0, p.w1[bidx], best);
                    if (!(second > ub)) {
                        const int pos = atomicAdd(p.flag_count, 1);
                        p.flag_list[pos] = (int32_t) row;
                        p.flag_ub[pos] = ub;
                        p.min_key[pos] = ~0ull;
                        cell = -1;   // its row is added after the FP64 re-evaluation (vq_cell_fixup)
                    }
                }
            }
            if (CELLS) {
                s_cell[((it & 3) * 2 + h) * kBM + r] = cell;
                __syncwarp();
                if (lane == 0) mbar_arrive(&sm->cell_full[(it & 3) * 2 + h]);
            }
        }
        if (DBG && p.prof != nullptr && lane == 0 && q == 0) {
            unsigned long long *o = p.prof + 2048 + 8 * blockIdx.x + 3 + 2 * h;
            o[0] = pw_f; o[1] = clock64() - pw_t0;
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 13) tmem_dealloc(tmem_base, 512);
    if (DBG && p.prof != nullptr && threadIdx.x == 0) {
        unsigned long long t1;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
        p.prof[2 * blockIdx.x] = clock64() - prof_c0;
        p.prof[2 * blockIdx.x + 1] = t1 - prof_t0;
    }
}

// raas f64 [K][D] -> tiles [tile][hi, lo][k-chunk][row][4]  (zero padded; slot D = W1)
__global__ void vq_prepare_codebook_tc2(const double *__restrict__ raas, int K, int D, int BN,
                                        int ksteps, float *__restrict__ wtc) {
    const int KC = 2 * ksteps;
    const int n_wtiles = (K + BN - 1) / BN;
    const long items = (long) n_wtiles * KC * BN;   // (t, c, r)
    for (long it = (long) blockIdx.x * blockDim.x + threadIdx.x; it < items;
         it += (long) gridDim.x * blockDim.x) {
        const int r = (int) (it % BN);
        const int c = (int) ((it / BN) % KC);
        const int t = (int) (it / ((long) BN * KC));
        const int cw = t * BN + r;
        float4 h = make_float4(0.f, 0.f, 0.f, 0.f), l = h;
        float *hp = &h.x, *lp = &l.x;
        if (cw < K) {
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int k = 4 * c + e;
                if (k <= D) {
                    double w;
                    if (k < D) {
                        w = k > 0 ? 2.0 * raas[(long) cw * D + k] : raas[(long) cw * D];
                    } else {   // slot D: W1 = sum_n |w[n]|
                        w = fabs(raas[(long) cw * D]);
                        for (int n = 1; n < D; n++) w += 2.0 * fabs(raas[(long) cw * D + n]);
                        w *= 1.0 + 1e-6;
                    }
                    const float hi = tf32_rna((float) w);
                    hp[e] = hi;
                    lp[e] = tf32_rna((float) (w - (double) hi));
                }
            }
        }
        float *tile = wtc + (size_t) t * (2 * KC * BN * 4);
        *reinterpret_cast<float4 *>(tile + ((size_t) c * BN + r) * 4) = h;
        *reinterpret_cast<float4 *>(tile + ((size_t) (KC + c) * BN + r) * 4) = l;
    }
}

// FP16 format: raas f64 [K][D] -> tiles [tile][lo: KC chunks | hi: KC chunks | zero chunk if KC
// is odd][row][8 halves] (KC = `ksteps` = 8-element chunks per split part) of the codebook
// scaled by G = 2^g, g chosen from *w1max so that every scaled W1_j is below 2^14 (both split
// parts of every element then stay inside the FP16 range).  Slot D and w1h[j] carry
//     W1'_j = G W1_j + D 2^-25 / err_scale
// -- the second term covers the absolute error 2^-25 of a scaled codeword element whose low
// part is an FP16 subnormal (sum over k of |x_k| 2^-25 <= x_ref D 2^-25 = err_scale x_ref times
// that term).  w1h[n_wtiles BN] = G (debug dump, to return values in the caller's units).
__global__ void vq_prepare_codebook_tc2h(const double *__restrict__ raas, int K, int D, int BN,
                                         int ksteps, const float *__restrict__ w1max,
                                         float err_scale, float *__restrict__ wtc,
                                         float *__restrict__ w1h) {
    const int KC = ksteps;
    const int BC = 2 * KC + (KC & 1);
    const int n_wtiles = (K + BN - 1) / BN;
    const long items = (long) n_wtiles * KC * BN;   // (t, c, r)
    uint32_t ef = (__float_as_uint(*w1max) >> 23) & 0xffu;
    ef = ef < 14u ? 14u : (ef > 253u ? 253u : ef);
    const double G = (double) __uint_as_float((267u - ef) << 23);   // G w1max in [2^13, 2^14)
    const double extra = (double) D * 2.98023223876953125e-8 / (double) err_scale;
    if (blockIdx.x == 0 && threadIdx.x == 0) w1h[(long) n_wtiles * BN] = (float) G;
    for (long it = (long) blockIdx.x * blockDim.x + threadIdx.x; it < items;
         it += (long) gridDim.x * blockDim.x) {
        const int r = (int) (it % BN);
        const int c = (int) ((it / BN) % KC);
        const int t = (int) (it / ((long) BN * KC));
        const int cw = t * BN + r;
        __half hp[8], lp[8];
#pragma unroll
        for (int e = 0; e < 8; e++) { hp[e] = __float2half_rn(0.f); lp[e] = hp[e]; }
        if (cw < K) {
#pragma unroll
            for (int e = 0; e < 8; e++) {
                const int k = 8 * c + e;
                if (k <= D) {
                    double w;
                    if (k < D) {
                        w = G * (k > 0 ? 2.0 * raas[(long) cw * D + k] : raas[(long) cw * D]);
                    } else {   // slot D: W1'
                        w = fabs(raas[(long) cw * D]);
                        for (int n = 1; n < D; n++) w += 2.0 * fabs(raas[(long) cw * D + n]);
                        w = (G * w + extra) * (1.0 + 1e-6);
                        if (c == D / 8) w1h[cw] = (float) (w * (1.0 + 1e-6));
                    }
                    const __half hi = __float2half_rn((float) w);
                    hp[e] = hi;
                    lp[e] = __float2half_rn((float) (w - (double) __half2float(hi)));
                }
            }
        } else if (c == D / 8) {
            w1h[cw] = 0.f;
        }
        unsigned char *tile = reinterpret_cast<unsigned char *>(wtc) + (size_t) t * (BC * BN * 16);
        *reinterpret_cast<uint4 *>(tile + ((size_t) c * BN + r) * 16) = *reinterpret_cast<const uint4 *>(lp);
        *reinterpret_cast<uint4 *>(tile + ((size_t) (KC + c) * BN + r) * 16) = *reinterpret_cast<const uint4 *>(hp);
        if ((KC & 1) && c == 0)
            *reinterpret_cast<uint4 *>(tile + ((size_t) (2 * KC) * BN + r) * 16) = make_uint4(0u, 0u, 0u, 0u);
    }
}

// raas f64 [K][D] -> stages [tile][k-step][hi, lo][k-chunk][row][4]  (zero padded)
__global__ void vq_prepare_codebook_tc(const double *__restrict__ raas, int K, int D, int BN,
                                       int ksteps, float *__restrict__ wtc) {
    const int n_wtiles = (K + BN - 1) / BN;
    const long items = (long) n_wtiles * ksteps * 2 * BN;   // (t, s, c, r)
    for (long it = (long) blockIdx.x * blockDim.x + threadIdx.x; it < items;
         it += (long) gridDim.x * blockDim.x) {
        const int r = (int) (it % BN);
        const int c = (int) ((it / BN) & 1);
        const int s = (int) ((it / (2 * BN)) % ksteps);
        const int t = (int) (it / ((long) 2 * BN * ksteps));
        const int cw = t * BN + r;
        float4 h = make_float4(0.f, 0.f, 0.f, 0.f), l = h;
        float *hp = &h.x, *lp = &l.x;
        if (cw < K) {
#pragma unroll
            for (int e = 0; e < 4; e++) {
                const int k = 8 * s + 4 * c + e;
                if (k <= D) {
                    double w;
                    if (k < D) {
                        w = k > 0 ? 2.0 * raas[(long) cw * D + k] : raas[(long) cw * D];
                    } else {   // slot D: W1 = sum_n |w[n]|
                        w = fabs(raas[(long) cw * D]);
                        for (int n = 1; n < D; n++) w += 2.0 * fabs(raas[(long) cw * D + n]);
                        w *= 1.0 + 1e-6;
                    }
                    const float hi = tf32_rna((float) w);
                    hp[e] = hi;
                    lp[e] = tf32_rna((float) (w - (double) hi));
                }
            }
        }
        float *stage = wtc + ((size_t) t * ksteps + s) * (size_t) (4 * BN * 4);
        *reinterpret_cast<float4 *>(stage + ((size_t) (0 * 2 + c) * BN + r) * 4) = h;
        *reinterpret_cast<float4 *>(stage + ((size_t) (1 * 2 + c) * BN + r) * 4) = l;
    }
}

// ECOZ2_VQ_TC_VARIANT=1: both operands from shared memory (first version, kept for A/B runs);
// default 2: frame operand in tensor memory.
static int tc_variant() {
    static int v = 0;
    if (v == 0) {
        const char *e = getenv("ECOZ2_VQ_TC_VARIANT");
        v = (e && e[0] == '1') ? 1 : 2;
    }
    return v;
}

// Operand format of variant 2 (ECOZ2_VQ_TC_FMT=tf32 | f16, default f16): the same 22-bit
// two-way split either as TF32 parts (8 elements per MMA) or as FP16 parts of power-of-two
// scaled operands (16 elements per MMA at the same shared-memory bytes: fewer MMAs per tile).
static bool tc_f16() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("ECOZ2_VQ_TC_FMT");
        v = (tc_variant() == 2 && !(e && e[0] == 't')) ? 1 : 0;
    }
    return v == 1;
}
// D elements + the bound slot in groups of 8: MMAs per product term (TF32 format) or 16-byte
// chunks per split part (FP16 format)
static int tc_ksteps(int D) { return (D + 8) / 8; }
static int tc_max_ksteps() { return tc_f16() ? 6 : kMaxKSteps; }
// 16-byte chunks of one codeword row in a tile
static int tc_row_chunks(int D) {
    const int ks = tc_ksteps(D);
    return tc_f16() ? 2 * ks + (ks & 1) : 4 * ks;
}

// smallest codebook the tensor-core search takes (ECOZ2_VQ_TC_MIN_K, default 2: a 64-column
// tile that is mostly padding still beats the FP64 sweep of vq_pass_small)
static int tc_min_k() {
    static int v = 0;
    if (v == 0) {
        const char *e = getenv("ECOZ2_VQ_TC_MIN_K");
        v = e ? atoi(e) : 2;
        if (v < 1) v = 1;
    }
    return v;
}

// Variant 2, FP16 format: 256-column tiles for K >= 256.  A 128x256x16 MMA fetches 96 operand
// wavefronts in its 128 clocks and runs at the compute floor; a 128x128x16 one fetches 64 in 64
// and was measured at 73 clk alone, 84 with the epilogue's tcgen05.ld traffic, ~100 in the kernel
// (scripts/exp/ldtm_bench).  TMEM then holds ONE accumulator per 128-frame half: the two halves
// alternate between the tensor core and the epilogue.  ECOZ2_VQ_TC_WIDE=0 keeps 128.
static bool tc_wide() {
    static int v = -1;
    if (v < 0) {
        const char *e = getenv("ECOZ2_VQ_TC_WIDE");
        v = (tc_variant() == 2 && !(e && e[0] == '0')) ? 1 : 0;
    }
    return v == 1;
}

bool tc_operands_f16() { return tc_f16(); }

// Frames -> pre-staged operands of vq_assign_tc2 (FP16 format): per half-tile of 128 frames one
// record [hi part: KS chunks x 128 rows x 8 halves][lo part: same][128 norms f32], i.e. exactly what
// the loader warps write into shared memory (same operations in the same order: bit-identical
// operands), built ONCE per frame set.  Rows beyond n are zero.  thread = frame row.
template <int KS>
__global__ void __launch_bounds__(kBM)
vq_prepare_frames_tc2h(const float *__restrict__ x, long n, int D, float err_scale,
                       unsigned char *__restrict__ xop) {
    constexpr int NE = KS * 8;
    constexpr size_t part_bytes = (size_t) KS * kBM * 16;
    constexpr size_t rec = 2 * part_bytes + kBM * sizeof(float);
    const int r = threadIdx.x;
    const long g = blockIdx.x;                 // half-tile
    const long row = g * kBM + r;
    float v[NE];
    const float *src = x + row * D;
#pragma unroll
    for (int k = 0; k < NE; k++) v[k] = (row < n && k < D) ? __ldg(src + k) : 0.f;
    float xm = 0.f;
#pragma unroll
    for (int k = 0; k < NE; k++) xm = fmaxf(xm, fabsf(v[k]));
    uint32_t ef = (__float_as_uint(xm) >> 23) & 0xffu;
    ef = ef < 16u ? 16u : ef;
    const float scale = __uint_as_float((268u - ef) << 23);   // 2^(14 - floor(log2 xm))
#pragma unroll
    for (int k = 0; k < NE; k++) v[k] *= scale;
    const float x0 = xm * scale;
#pragma unroll
    for (int k = 0; k < NE; k++)
        if (k == D) v[k] = -1.001f * err_scale * x0;
    unsigned char *hi = xop + (size_t) g * rec;
    unsigned char *lo = hi + part_bytes;
#pragma unroll
    for (int c = 0; c < KS; c++) {
        uint32_t hw[4], lw[4];
#pragma unroll
        for (int e = 0; e < 4; e++) {
            const float a = v[8 * c + 2 * e], b2 = v[8 * c + 2 * e + 1];
            const __half2 hh = __floats2half2_rn(a, b2);
            const float2 hf = __half22float2(hh);
            const __half2 ll = __floats2half2_rn(a - hf.x, b2 - hf.y);
            hw[e] = *reinterpret_cast<const uint32_t *>(&hh);
            lw[e] = *reinterpret_cast<const uint32_t *>(&ll);
        }
        *reinterpret_cast<uint4 *>(hi + ((size_t) c * kBM + r) * 16) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
        *reinterpret_cast<uint4 *>(lo + ((size_t) c * kBM + r) * 16) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
    }
    reinterpret_cast<float *>(hi + 2 * part_bytes)[r] = x0;
}

int tc_tile_cols(int K, int D) {
    // (wide tiles need 2 x 13 chunks x 4 KB for D + 1 > 40: no room next to the frame tiles)
    if (tc_variant() == 2)
        return (K >= 256 && tc_f16() && tc_wide() && tc_ksteps(D) <= 5) ? 256 : (K >= 128 ? 128 : 64);
    return K >= 256 ? 256 : (K > 64 ? 128 : 64);
}

int tc_issued_flops_per_eval(int D, int *is_f16) {
    if (is_f16) *is_f16 = tc_f16() ? 1 : 0;
    const int ks = tc_ksteps(D);
    if (ks > tc_max_ksteps()) return 0;
    if (tc_f16()) return 2 * 16 * (ks + (ks + 1) / 2);   // cross product + main product, 16 per MMA
    return 2 * 8 * 3 * ks;
}

bool assign_tc_supported(int K, int D) { return K >= tc_min_k() && tc_ksteps(D) <= tc_max_ksteps(); }

// floats of the tiles themselves: n_wtiles x [hi, lo] x (2 ksteps chunks) x BN rows x 16 bytes
static size_t tc_tiles_floats(int K, int D) {
    const int BN = tc_tile_cols(K, D);
    const int n_wtiles = (K + BN - 1) / BN;
    return (size_t) n_wtiles * tc_row_chunks(D) * BN * 4;
}
// + (FP16 format) the scaled bound weights W1'_j per padded codeword and the scale G
size_t tc_codebook_floats(int K, int D) {
    const int BN = tc_tile_cols(K, D);
    const int n_wtiles = (K + BN - 1) / BN;
    return tc_tiles_floats(K, D) + (size_t) n_wtiles * BN + 4;
}

const float *tc_bound_weights(const float *wtc, int K, int D) { return wtc + tc_tiles_floats(K, D); }

void launch_prepare_codebook_tc(const double *raas, int K, int D, float *wtc, const float *w1max,
                                float err_scale, cudaStream_t s) {
    const int BN = tc_tile_cols(K, D), ksteps = tc_ksteps(D);
    const long items = (long) ((K + BN - 1) / BN) * ksteps * 2 * BN;
    const int blocks = (int) ((items + 255) / 256);
    if (tc_f16())
        vq_prepare_codebook_tc2h<<<blocks, 256, 0, s>>>(raas, K, D, BN, ksteps, w1max, err_scale, wtc,
                                                       wtc + tc_tiles_floats(K, D));
    else if (tc_variant() == 2)
        vq_prepare_codebook_tc2<<<blocks, 256, 0, s>>>(raas, K, D, BN, ksteps, wtc);
    else
        vq_prepare_codebook_tc<<<blocks, 256, 0, s>>>(raas, K, D, BN, ksteps, wtc);
}

static size_t tc_smem_bytes(int D, int BN) {
    const int ksteps = tc_ksteps(D);
    return 6144 + (size_t) 2 * 2 * (2 * ksteps) * kBM * 16 + (size_t) kStages * 4 * BN * 16 +
           (((size_t) kBM * D * 4 + 127) & ~(size_t) 127);
}
static size_t tc2_smem_bytes(int D, int BN, int raw_bufs) {
    const size_t raw_stride = ((size_t) kBM * D * 4 + 127) & ~(size_t) 127;
    const size_t KC = (tc_f16() ? 1 : 2) * (size_t) tc_ksteps(D);   // chunks per split part
    return (tc_f16() ? 13312 : 5120) + raw_bufs * raw_stride + (tc_f16() ? 2 : 1) * 2 * (2 * KC * kBM * 16) +
           (size_t) (BN == 256 ? 2 : kB2Stages) * tc_row_chunks(D) * BN * 16;
}
// two staging buffers for the raw frame halves when they fit (the second hides the bulk-copy
// latency behind the loaders' conversion of the first: it matters for small codebooks, where
// the loaders are the critical path)
// Small codebooks are bound by the frames in flight, not by the tensor core: a half-tile's bulk copy
// takes ~4000 clocks from HBM under load, so an SM needs ~100 KB in flight to take its share of the
// bandwidth (measured on B200: two 19 KB buffers = 2.6 TB/s, 0.56 ms per 10M-frame search whatever K).
// The narrow-tile configurations leave that much shared memory free: up to 8 staging buffers.
static int tc2_raw_bufs(int D, int BN) {
    static int cap = -1;
    if (cap < 0) { const char *e = getenv("ECOZ2_VQ_TC_RAW"); cap = e ? atoi(e) : 2; if (cap < 1) cap = 1; if (cap > 8) cap = 8; }
    int n = cap;
    while (n > 1 && tc2_smem_bytes(D, BN, n) > 227 * 1024) n--;
    return n;
}

template <int KS, bool FILTER, bool DBG, bool H16, bool CELLS, bool SEG = false>
static void launch_tc2_inst2(const TcParams &p, long grid, cudaStream_t s) {
    // per device (function attributes belong to the context): set on every launch, ~1 us
    cudaFuncSetAttribute(vq_assign_tc2<KS, FILTER, DBG, H16, CELLS, SEG>,
                         cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    vq_assign_tc2<KS, FILTER, DBG, H16, CELLS, SEG><<<(unsigned) grid, k2Threads, tc2_smem_bytes(p.D, p.BN, p.raw_bufs), s>>>(p);
}
template <int KS, bool H16>
static void launch_tc2_inst(const TcParams &p, long grid, cudaStream_t s) {
    if (H16 && p.seg_cols != 0) launch_tc2_inst2<KS, false, false, H16, false, H16>(p, grid, s);
    else if (p.cand != nullptr) launch_tc2_inst2<KS, true, false, H16, false>(p, grid, s);
    else if (p.exp != 0 || p.dump != nullptr) launch_tc2_inst2<KS, false, true, H16, false>(p, grid, s);
    else if (H16 && p.cells != nullptr) launch_tc2_inst2<KS, false, false, H16, H16>(p, grid, s);
    else launch_tc2_inst2<KS, false, false, H16, false>(p, grid, s);
}
// variant 2, any supported (format, k-steps)
static void launch_tc2(TcParams &p, long grid, cudaStream_t s) {
    p.raw_bufs = tc2_raw_bufs(p.D, p.BN);
    if (tc_f16()) {
        // the bound weights and the scale live behind the tiles (vq_prepare_codebook_tc2h)
        const float *w1h = p.wtc + tc_tiles_floats(p.K, p.D);
        p.w1 = w1h;
        p.wscale = w1h + (size_t) ((p.K + p.BN - 1) / p.BN) * p.BN;
        switch (p.ksteps) {
            case 1: launch_tc2_inst<1, true>(p, grid, s); break;
            case 2: launch_tc2_inst<2, true>(p, grid, s); break;
            case 3: launch_tc2_inst<3, true>(p, grid, s); break;
            case 4: launch_tc2_inst<4, true>(p, grid, s); break;
            case 5: launch_tc2_inst<5, true>(p, grid, s); break;
            default: launch_tc2_inst<6, true>(p, grid, s); break;
        }
        return;
    }
    switch (p.ksteps) {
        case 1: launch_tc2_inst<1, false>(p, grid, s); break;
        case 2: launch_tc2_inst<2, false>(p, grid, s); break;
        case 3: launch_tc2_inst<3, false>(p, grid, s); break;
        case 4: launch_tc2_inst<4, false>(p, grid, s); break;
        default: launch_tc2_inst<5, false>(p, grid, s); break;
    }
}

// pre-staged frame operands (see vq_prepare_frames_tc2h): bytes for n frames, 0 = not available
static bool tc_pre_enabled() {
    static int on = -1;
    if (on < 0) { const char *e = getenv("ECOZ2_VQ_TC_PRE"); on = (e && e[0] == '0') ? 0 : 1; }
    return on == 1;
}
size_t tc_frames_bytes(long n, int D) {
    if (!(tc_pre_enabled() && tc_variant() == 2 && tc_f16() && n > 0 && tc_ksteps(D) <= 6)) return 0;
    const size_t halves = 2 * (size_t) ((n + 2 * kBM - 1) / (2 * kBM));
    return halves * ((size_t) 2 * tc_ksteps(D) * kBM * 16 + kBM * sizeof(float));
}
void launch_prepare_frames_tc(const float *x32, long n, int D, float err_scale, void *xop, cudaStream_t s) {
    if (tc_frames_bytes(n, D) == 0) return;
    const unsigned halves = (unsigned) (2 * ((n + 2 * kBM - 1) / (2 * kBM)));
    unsigned char *o = static_cast<unsigned char *>(xop);
    switch (tc_ksteps(D)) {
        case 1: vq_prepare_frames_tc2h<1><<<halves, kBM, 0, s>>>(x32, n, D, err_scale, o); break;
        case 2: vq_prepare_frames_tc2h<2><<<halves, kBM, 0, s>>>(x32, n, D, err_scale, o); break;
        case 3: vq_prepare_frames_tc2h<3><<<halves, kBM, 0, s>>>(x32, n, D, err_scale, o); break;
        case 4: vq_prepare_frames_tc2h<4><<<halves, kBM, 0, s>>>(x32, n, D, err_scale, o); break;
        case 5: vq_prepare_frames_tc2h<5><<<halves, kBM, 0, s>>>(x32, n, D, err_scale, o); break;
        default: vq_prepare_frames_tc2h<6><<<halves, kBM, 0, s>>>(x32, n, D, err_scale, o); break;
    }
}

bool assign_tc_fuses_cells(int K, int D) {
    static int on = -1;
    if (on < 0) { const char *e = getenv("ECOZ2_VQ_TC_CELLS"); on = (e && e[0] == '0') ? 0 : 1; }
    return on && tc_variant() == 2 && tc_f16() && D <= 64 && assign_tc_supported(K, D);
}

void launch_assign_tc(const float *x32, long n, int D, const float *wtc, int K,
                      const int32_t *list, const int32_t *list_count, const SearchScratch &sc,
                      bool flag, const float *w1, const float *w1max, float err_scale, int32_t *idx,
                      float *dump, int dump_ld, int num_sms, cudaStream_t s, const double *x64,
                      double *cells, int32_t *hist, const void *xop) {
    if (n <= 0) return;
    TcParams p{};
    if (xop != nullptr && list == nullptr && dump == nullptr) p.xop = static_cast<const unsigned char *>(xop);
    if (cells != nullptr && list == nullptr && dump == nullptr && flag && assign_tc_fuses_cells(K, D)) {
        p.x64 = x64; p.cells = cells; p.hist = hist;
    }
    p.x = x32; p.n_frames = n; p.D = D; p.ksteps = tc_ksteps(D);
    p.list = list; p.list_count = list_count;
    p.wtc = wtc; p.K = K; p.BN = tc_tile_cols(K, D);
    p.w1 = w1; p.w1max = w1max; p.err_scale = err_scale; p.idx_out = idx;
    if (flag) {
        p.flag_list = sc.flag_list; p.flag_count = sc.counters + 0; p.flag_ub = sc.flag_ub;
        p.min_key = sc.min_key;
    }
    p.dump = dump; p.dump_ld = dump_ld;
    {
        static int exp = -1;
        if (exp < 0) { const char *e = getenv("ECOZ2_VQ_TC_EXP"); exp = e ? atoi(e) : 0; }
        p.exp = exp;
    }
    if (tc_variant() == 2) {
        long tiles = (n + 2 * kBM - 1) / (2 * kBM);
        long grid = tiles < num_sms ? tiles : num_sms;
        if (p.exp & 256) {   // dev: SM clock seen by the kernel itself (clock64 vs globaltimer)
            static unsigned long long *prof = nullptr;
            if (!prof) cudaMalloc(&prof, sizeof(unsigned long long) * 10 * 1024);
            cudaMemsetAsync(prof, 0, sizeof(unsigned long long) * 10 * 1024, s);
            p.prof = prof;
            launch_tc2(p, grid, s);
            std::vector<unsigned long long> h(10 * 1024);
            cudaStreamSynchronize(s);
            cudaMemcpy(h.data(), prof, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost);
            double clk = 0, ns = 0;
            for (long i = 0; i < grid; i++) { clk += (double) h[2 * i]; ns += (double) h[2 * i + 1]; }
            fprintf(stderr, "[tc prof] n=%ld K=%d: %.3f ms per CTA, SM clock %.0f MHz, %.0f clk per CTA\n", n, K,
                    ns / grid * 1e-6, clk / ns * 1e3, clk / grid);
            // where the roles of CTA 0.. wait (clocks, mean over CTAs): MMA issuer for frames /
            // free accumulator / codeword tile; epilogue of half 0 and 1 for a full accumulator
            double w[7] = {0, 0, 0, 0, 0, 0, 0};
            for (long i = 0; i < grid; i++)
                for (int k = 0; k < 7; k++) w[k] += (double) h[2048 + 8 * i + k] / grid;
            fprintf(stderr, "[tc prof]   MMA issuer waits: frames %.0f, accumulator %.0f, codewords %.0f | epilogue 0: "
                            "waits %.0f of %.0f | epilogue 1: waits %.0f of %.0f\n",
                    w[0], w[1], w[2], w[3], w[4], w[5], w[6]);
            return;
        }
        launch_tc2(p, grid, s);
        return;
    }
    const size_t smem = tc_smem_bytes(D, p.BN);
    // per device (function attributes belong to the context): set on every launch, ~1 us
    cudaFuncSetAttribute(vq_assign_tc, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024);
    long tiles = (n + kBM - 1) / kBM;
    long grid = tiles < num_sms ? tiles : num_sms;
    if (grid < 1) grid = 1;
    static int use_mc = -1;
    if (use_mc < 0) { const char *e = getenv("ECOZ2_VQ_TC_MULTICAST"); use_mc = (e && e[0] == '0') ? 0 : 1; }
    if (use_mc && grid >= 2) {
        // clusters of 2 CTAs share the codeword stream through TMA multicast
        grid &= ~1L;
        p.mc = 1;
        cudaLaunchConfig_t cfg = {};
        cfg.gridDim = dim3((unsigned) grid);
        cfg.blockDim = dim3(kThreads);
        cfg.dynamicSmemBytes = smem;
        cfg.stream = s;
        cudaLaunchAttribute attr[1];
        attr[0].id = cudaLaunchAttributeClusterDimension;
        attr[0].val.clusterDim.x = 2;
        attr[0].val.clusterDim.y = 1;
        attr[0].val.clusterDim.z = 1;
        cfg.attrs = attr;
        cfg.numAttrs = 1;
        cudaLaunchKernelEx(&cfg, vq_assign_tc, p);
        return;
    }
    vq_assign_tc<<<(unsigned) grid, kThreads, smem, s>>>(p);
}


// R codebooks of K codewords each, concatenated (wtc built from the R * K raas rows, padded by the
// caller to whole tiles): one sweep over the frames, idx_out[R][idx_stride] model-local argmins,
// (model, frame) pairs inside the error bound as keys in seg_key (count in sc.counters[0]; pairs
// beyond seg_cap are dropped -- the caller checks the count).
bool assign_tc_models_supported(int K, int D, long total_cols) {
    if (!(tc_variant() == 2 && tc_f16() && assign_tc_supported((int) total_cols, D))) return false;
    if (total_cols < 256 || total_cols % 256 != 0) return false;
    // (larger codebooks: the search is compute-bound anyway and the FP64 scan of the ambiguous pairs,
    // K evaluations each, would cost more than the R - 1 frame sweeps saved)
    return K == 32 || K == 64 || K == 128 || K == 256;
}
void launch_assign_tc_models(const float *x32, long n, int D, const float *wtc, int K, long total_cols,
                             const SearchScratch &sc, float err_scale, int32_t *idx_out, long idx_stride,
                             unsigned long long *seg_key, int seg_cap, int num_sms, cudaStream_t s,
                             const void *xop) {
    if (n <= 0) return;
    TcParams p{};
    p.xop = static_cast<const unsigned char *>(xop);
    p.x = x32; p.n_frames = n; p.D = D; p.ksteps = tc_ksteps(D);
    p.wtc = wtc; p.K = (int) total_cols; p.BN = tc_tile_cols((int) total_cols, D);
    p.err_scale = err_scale; p.idx_out = idx_out;
    p.flag_count = sc.counters + 0;
    p.seg_cols = K; p.idx_stride = idx_stride; p.seg_key = seg_key; p.seg_cap = seg_cap;
    long tiles = (n + 2 * kBM - 1) / (2 * kBM);
    long grid = tiles < num_sms ? tiles : num_sms;
    launch_tc2(p, grid, s);
}

bool filter_tc_supported() { return tc_variant() == 2; }

void launch_filter_tc(const float *x32, long n_cap, int D, const float *wtc, int K,
                      const SearchScratch &sc, float err_scale, int num_sms, cudaStream_t s) {
    if (n_cap <= 0) return;
    TcParams p{};
    p.x = x32; p.n_frames = n_cap; p.D = D; p.ksteps = tc_ksteps(D);
    p.list = sc.flag_list; p.list_count = sc.counters + 0;
    p.wtc = wtc; p.K = K; p.BN = tc_tile_cols(K, D);
    p.err_scale = err_scale; p.flag_ub = sc.flag_ub;
    p.cand = sc.cand; p.cand_count = sc.counters + 1; p.cand_cap = (int) sc.cand_cap;
    p.exact_list = sc.exact_list; p.exact_count = sc.counters + 2;
    // device-side count: persistent grid sized for the worst case
    long tiles = (n_cap + 2 * kBM - 1) / (2 * kBM);
    long grid = tiles < num_sms ? tiles : num_sms;
    launch_tc2(p, grid, s);
}

}  // namespace vqb200
