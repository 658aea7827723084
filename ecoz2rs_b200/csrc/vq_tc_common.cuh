// vq_tc_common.cuh -- device helpers shared by the tensor-core kernels (vq_tc.cu, vq_mid.cu):
// PTX wrappers (mbarrier, bulk copies, tcgen05 MMA / TMEM), UMMA descriptors and the fused
// argmin epilogue over TMEM accumulators.  Internal linkage (anonymous namespaces).
#pragma once
#include <cuda_fp16.h>
#include <cfloat>
#include <cmath>
#include <stdint.h>

namespace vqb200 {

namespace {

constexpr int kBM = 128;        // frames per tile = UMMA M = TMEM lanes
constexpr int kStages = 6;      // B stage ring
constexpr int kMaxKSteps = 5;   // D <= 40
constexpr int kThreads = 384;    // 4 role warps + 8 epilogue warps
constexpr int kLoaderThreads = 64;
constexpr int kEpiWarp0 = 4;
constexpr int kEpiWarps = 8;     // two per TMEM lane quarter, each takes half of the columns

// ---- PTX wrappers -----------------------------------------------------------------
__device__ __forceinline__ uint32_t s32(const void *p) {
    return (uint32_t) __cvta_generic_to_shared(p);
}
__device__ __forceinline__ void mbar_init(uint64_t *bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(s32(bar)), "r"(count));
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t *bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(s32(bar)),
                 "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t *bar) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(s32(bar)) : "memory");
}
// arrive whose issue depends on `dep` having been computed
__device__ __forceinline__ void mbar_arrive_dep(uint64_t *bar, uint32_t dep) {
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0]; // %1" ::"r"(s32(bar)), "r"(dep) : "memory");
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
        : "r"(s32(bar)), "r"(parity)
        : "memory");
    return ok != 0;
}
__device__ __forceinline__ void mbar_wait(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) {
    }
}
// wait of a warp that is far ahead of its consumer (producers, frame loaders): sleeping
// between polls leaves the issue slots of its SM sub-partition to the epilogue warps
__device__ __forceinline__ void mbar_wait_relaxed(uint64_t *bar, uint32_t parity) {
    while (!mbar_try_wait(bar, parity)) __nanosleep(200);
}
template <int REGS>
__device__ __forceinline__ void setmaxnreg_inc() {
    asm volatile("setmaxnreg.inc.sync.aligned.u32 %0;" ::"n"(REGS));
}
template <int REGS>
__device__ __forceinline__ void setmaxnreg_dec() {
    asm volatile("setmaxnreg.dec.sync.aligned.u32 %0;" ::"n"(REGS));
}
__device__ __forceinline__ void bulk_g2s(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                         uint64_t *bar) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::
            "r"(s32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(s32(bar))
        : "memory");
}
__device__ __forceinline__ void fence_barrier_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_before() {
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tc_fence_after() {
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
}
__device__ __forceinline__ void tmem_alloc(uint32_t *slot_smem, uint32_t cols) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(
                     s32(slot_smem)),
                 "r"(cols)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ void tmem_dealloc(uint32_t taddr, uint32_t cols) {
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(taddr), "r"(cols)
                 : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], TF32 inputs, FP32 accumulate
__device__ __forceinline__ void umma_tf32(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc,
                                          uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], FP16 inputs (K = 16 per instruction), FP32 accumulate
__device__ __forceinline__ void umma_f16(uint32_t d_tmem, uint64_t adesc, uint64_t bdesc,
                                         uint32_t idesc, uint32_t accumulate) {
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n" ::"r"(d_tmem),
        "l"(adesc), "l"(bdesc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// mbarrier arrives when every tcgen05.mma issued so far by this thread has completed
__device__ __forceinline__ void umma_commit(uint64_t *bar) {
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::
                     "r"(s32(bar))
                 : "memory");
}
// multicast variants (thread-block cluster of 2): the copy lands at the same CTA-relative
// offset in every CTA of cta_mask and signals each one's mbarrier at that offset
__device__ __forceinline__ void bulk_g2s_mc(void *dst_smem, const void *src_gmem, uint32_t bytes,
                                            uint64_t *bar, uint16_t cta_mask) {
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.multicast::cluster "
        "[%0], [%1], %2, [%3], %4;" ::"r"(s32(dst_smem)),
        "l"(src_gmem), "r"(bytes), "r"(s32(bar)), "h"(cta_mask)
        : "memory");
}
__device__ __forceinline__ void umma_commit_mc(uint64_t *bar, uint16_t cta_mask) {
    asm volatile(
        "tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.multicast::cluster.b64 "
        "[%0], %1;" ::"r"(s32(bar)),
        "h"(cta_mask)
        : "memory");
}
__device__ __forceinline__ void cluster_sync_all() {
    asm volatile("barrier.cluster.arrive.release.aligned;" ::: "memory");
    asm volatile("barrier.cluster.wait.acquire.aligned;" ::: "memory");
}
__device__ __forceinline__ uint32_t cluster_ctarank() {
    uint32_t r;
    asm volatile("mov.u32 %0, %%cluster_ctarank;" : "=r"(r));
    return r;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t (&v)[32]) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]),
          "=r"(v[7]), "=r"(v[8]), "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]),
          "=r"(v[14]), "=r"(v[15]), "=r"(v[16]), "=r"(v[17]), "=r"(v[18]), "=r"(v[19]), "=r"(v[20]),
          "=r"(v[21]), "=r"(v[22]), "=r"(v[23]), "=r"(v[24]), "=r"(v[25]), "=r"(v[26]), "=r"(v[27]),
          "=r"(v[28]), "=r"(v[29]), "=r"(v[30]), "=r"(v[31])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ bool elect_one() {
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "elect.sync _|p, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0;
}
__device__ __forceinline__ void tmem_ld_wait() {
    asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}
__device__ __forceinline__ float tf32_rna(float v) {
    uint32_t r;
    asm("cvt.rna.tf32.f32 %0, %1;" : "=r"(r) : "f"(v));
    return __uint_as_float(r);
}

// UMMA shared-memory matrix descriptor, K-major, no swizzle: core matrix = 8 rows x 16 B
// stored as 128 contiguous bytes; lbo = byte distance between the two core matrices of a
// k-step (K direction), sbo = byte distance between 8-row groups (M/N direction).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t lbo, uint32_t sbo) {
    return (uint64_t) ((saddr >> 4) & 0x3FFFu) | ((uint64_t) ((lbo >> 4) & 0x3FFFu) << 16) |
           ((uint64_t) ((sbo >> 4) & 0x3FFFu) << 32) | (1ull << 46);
}
// instruction descriptor: D = F32, A = B = TF32, both K-major, M x N
__host__ __device__ constexpr uint32_t umma_idesc_tf32(int M, int N) {
    return (1u << 4) | (2u << 7) | (2u << 10) | ((uint32_t) (N >> 3) << 17) |
           ((uint32_t) (M >> 4) << 24);
}
// same for kind::f16 with A = B = FP16 (format code 0)
__host__ __device__ constexpr uint32_t umma_idesc_f16(int M, int N) {
    return (1u << 4) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (M >> 4) << 24);
}

}  // namespace

namespace {
__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }


// Running state of one frame (one epilogue thread): minimum, runner-up, and a snapshot of
// the 16-codeword chunk that holds the minimum.  Only the FINAL minimum's chunk matters for
// the runner-up inside a chunk (a superseded minimum is itself a smaller-or-equal contender
// than anything else in its chunk), so the position inside the chunk is resolved once per
// frame (epi_finish), not per record.
struct EpiState {
    float best, second;
    int col;          // first codeword of the snapshot chunk
    float snap[16];
    __device__ __forceinline__ void init() {
        best = INFINITY; second = INFINITY; col = 0;
#pragma unroll
        for (int j = 0; j < 16; j++) snap[j] = INFINITY;
    }
};

// One chunk of 16 consecutive codewords: 10 FMNMX(3) for the chunk minimum + 4 ops; when some
// lane of the warp sees a new minimum those lanes copy the chunk (predicated moves).
__device__ __forceinline__ void epi_chunk(const float (&v)[16], int col0, EpiState &st, int exp = 0) {
    if (exp & 2) { st.best = fminf(st.best, v[0]); return; }
    const float q0 = fminf(min3f(v[0], v[1], v[2]), v[3]);
    const float q1 = fminf(min3f(v[4], v[5], v[6]), v[7]);
    const float q2 = fminf(min3f(v[8], v[9], v[10]), v[11]);
    const float q3 = fminf(min3f(v[12], v[13], v[14]), v[15]);
    const float m = fminf(min3f(q0, q1, q2), q3);
    const bool rec = m < st.best;
    // not a record: every value of the chunk is >= m >= best, so m is the chunk's only
    // contender for runner-up.  record: the old best is.
    st.second = fminf(st.second, fmaxf(m, st.best));
    st.best = fminf(st.best, m);
    if (!(exp & 1) && __any_sync(0xffffffffu, rec)) {
        if (rec) {
#pragma unroll
            for (int j = 0; j < 16; j++) st.snap[j] = v[j];
            st.col = col0;
        }
    }
}

// 32 consecutive codewords (two independent 16-chunks -> two min trees in flight).
// nv_rel = number of valid codewords from col0 on (>= 32 except at the end of the codebook).
__device__ __forceinline__ void epi_step32(const uint32_t (&r)[32], int col0, int nv_rel,
                                           EpiState &st, int exp) {
#define F(i) __uint_as_float(r[i])
    if (exp & 2) { st.best = fminf(st.best, F(0)); return; }
    if (nv_rel < 32) {   // ragged end of the codebook: generic path
        float v[16];
#pragma unroll
        for (int j = 0; j < 16; j++) v[j] = j < nv_rel ? F(j) : INFINITY;
        epi_chunk(v, col0, st, exp);
        if (nv_rel > 16) {
#pragma unroll
            for (int j = 0; j < 16; j++) v[j] = 16 + j < nv_rel ? F(16 + j) : INFINITY;
            epi_chunk(v, col0 + 16, st, exp);
        }
        return;
    }
    // minimum of each 16-codeword chunk in 8 three-input ops (16 -> 6 -> 2 -> 1), the two
    // chunks interleaved: two independent dependency chains
    const float a0 = min3f(F(0), F(1), F(2)), b0 = min3f(F(16), F(17), F(18));
    const float a1 = min3f(F(3), F(4), F(5)), b1 = min3f(F(19), F(20), F(21));
    const float a2 = min3f(F(6), F(7), F(8)), b2 = min3f(F(22), F(23), F(24));
    const float a3 = min3f(F(9), F(10), F(11)), b3 = min3f(F(25), F(26), F(27));
    const float a4 = min3f(F(12), F(13), F(14)), b4 = min3f(F(28), F(29), F(30));
    const float a5 = min3f(a0, a1, F(15)), b5 = min3f(b0, b1, F(31));
    const float a6 = min3f(a2, a3, a4), b6 = min3f(b2, b3, b4);
    const float mA = fminf(a5, a6), mB = fminf(b5, b6);
    const float m = fminf(mA, mB);
    const bool rec = m < st.best;
    // contenders for runner-up: the larger chunk minimum always; m itself unless it is the new
    // best, in which case the old best is
    st.second = min3f(st.second, fmaxf(m, st.best), fmaxf(mA, mB));
    st.best = fminf(st.best, m);
    if (!(exp & 1)) {
        // snapshot of the record chunk; the two chunks vote separately -- late in the codebook a
        // step sees records in one or two lanes, i.e. usually in one chunk only
        const bool ra_ = rec && mA <= mB, rb_ = rec && !(mA <= mB);
        if (__any_sync(0xffffffffu, ra_)) {
            if (ra_) {
#pragma unroll
                for (int j = 0; j < 16; j++) st.snap[j] = F(j);
                st.col = col0;
            }
        }
        if (__any_sync(0xffffffffu, rb_)) {
            if (rb_) {
#pragma unroll
                for (int j = 0; j < 16; j++) st.snap[j] = F(16 + j);
                st.col = col0 + 16;
            }
        }
    }
#undef F
}

// -> codeword of the minimum (first position among equal values) and the overall runner-up
__device__ __forceinline__ void epi_finish(const EpiState &st, float &second, int &bidx) {
    int jm = 15;
#pragma unroll
    for (int j = 14; j >= 0; j--) jm = st.snap[j] == st.best ? j : jm;
    float s2 = INFINITY;
#pragma unroll
    for (int j = 0; j < 16; j++) s2 = fminf(s2, j == jm ? INFINITY : st.snap[j]);
    second = fminf(st.second, s2);
    bidx = st.col + jm;
}

// Full tile of NCOLS codewords, search mode: fully unrolled (constant TMEM offsets, no
// per-step checks).  The accumulator is handed back to the MMA warp (release_bar) as soon as
// its last columns are in registers, before they are reduced.
template <int NCOLS>
__device__ __forceinline__ void epi_tile_full(uint32_t taddr, int tcol0, EpiState &est,
                                              uint64_t *release_bar) {
    constexpr int NS = NCOLS / 32;
    uint32_t ra[32], rb[32];
    tmem_ld32(taddr, ra);
#pragma unroll
    for (int s = 0; s < NS; s += 2) {
        tmem_ld_wait();
        if (s + 1 < NS) {
            tmem_ld32(taddr + 32 * (s + 1), rb);
        } else {
            tc_fence_before();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(release_bar);
        }
        epi_step32(ra, tcol0 + 32 * s, 32, est, 0);
        if (s + 1 < NS) {
            tmem_ld_wait();
            if (s + 2 < NS) {
                tmem_ld32(taddr + 32 * (s + 2), ra);
            } else {
                tc_fence_before();
                __syncwarp();
                if ((threadIdx.x & 31) == 0) mbar_arrive(release_bar);
            }
            epi_step32(rb, tcol0 + 32 * (s + 1), 32, est, 0);
        }
    }
}

// Same sweep over a tile that holds SEVERAL codebooks of seg_cols columns each (seg_cols a multiple
// of 32 that divides NCOLS): seg_done(first global column of the finished codebook) is called after
// the last 32 columns of each one (it finishes the argmin state and resets it).
template <int NCOLS, class F>
__device__ __forceinline__ void epi_tile_seg(uint32_t taddr, int tcol0, int seg_cols, EpiState &est,
                                             uint64_t *release_bar, F &&seg_done) {
    constexpr int NS = NCOLS / 32;
    uint32_t ra[32], rb[32];
    tmem_ld32(taddr, ra);
#pragma unroll
    for (int s = 0; s < NS; s += 2) {
        tmem_ld_wait();
        if (s + 1 < NS) {
            tmem_ld32(taddr + 32 * (s + 1), rb);
        } else {
            tc_fence_before();
            __syncwarp();
            if ((threadIdx.x & 31) == 0) mbar_arrive(release_bar);
        }
        epi_step32(ra, tcol0 + 32 * s, 32, est, 0);
        if ((32 * (s + 1)) % seg_cols == 0) seg_done(tcol0 + 32 * (s + 1) - seg_cols);
        if (s + 1 < NS) {
            tmem_ld_wait();
            if (s + 2 < NS) {
                tmem_ld32(taddr + 32 * (s + 2), ra);
            } else {
                tc_fence_before();
                __syncwarp();
                if ((threadIdx.x & 31) == 0) mbar_arrive(release_bar);
            }
            epi_step32(rb, tcol0 + 32 * (s + 1), 32, est, 0);
            if ((32 * (s + 2)) % seg_cols == 0) seg_done(tcol0 + 32 * (s + 2) - seg_cols);
        }
    }
}
}  // namespace

}  // namespace vqb200
