// Dev microbenchmark (not part of the product): how well do TMEM loads (tcgen05.ld 32x32b)
// overlap with the min-tree arithmetic of the search epilogue inside one SM sub-partition?
// 8 warps (2 per sub-partition) as in vq_assign_tc2; no MMAs.
//   mode 0: loads only    mode 1: arithmetic only (registers)    mode 2: both, double-buffered
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o epi_bench epi_bench.cu
#include <cstdio>
#include <cstdint>
#include <cstdlib>
#include <cuda_runtime.h>

__device__ __forceinline__ uint32_t s32(const void *p) { return (uint32_t) __cvta_generic_to_shared(p); }

template <int N> __device__ __forceinline__ void tmem_ld(uint32_t taddr, uint32_t *v);
template <> __device__ __forceinline__ void tmem_ld<32>(uint32_t taddr, uint32_t *r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, "
        "%16, %17, %18, %19, %20, %21, %22, %23, %24, %25, %26, %27, %28, %29, %30, %31}, [%32];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]),
          "=r"(r[16]), "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]),
          "=r"(r[24]), "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
        : "r"(taddr) : "memory");
}
template <> __device__ __forceinline__ void tmem_ld<16>(uint32_t taddr, uint32_t *r) {
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr) : "memory");
}
__device__ __forceinline__ void ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ float min3f(float a, float b, float c) { return fminf(fminf(a, b), c); }

struct St { float best, second; int col; float snap[16]; };

// the product's arithmetic per 32 columns, snapshot votes included
template <bool SNAP>
__device__ __forceinline__ void step32(const uint32_t (&r)[32], int col0, St &st) {
#define F(i) __uint_as_float(r[i])
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
    st.second = min3f(st.second, fmaxf(m, st.best), fmaxf(mA, mB));
    st.best = fminf(st.best, m);
    if (SNAP) {
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

// MODE 0 loads only, 1 arithmetic only, 2 both (double-buffered like epi_tile_full), 3 both with
// x16 loads (4 in flight)
template <int MODE, bool SNAP>
__global__ void __launch_bounds__(256, 1) bench(int iters, unsigned long long *out_clk, float *sink, int warps_active) {
    __shared__ uint32_t slot;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(s32(&slot)), "r"(512) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    const uint32_t tb = slot;
    // fill TMEM with decreasing-ish pseudo-random positive floats so that records happen at a
    // realistic (1/position) rate
    {
        const uint32_t taddr = tb + ((uint32_t) ((warp & 3) * 32) << 16);
        if (warp < 4) {
            for (int c = 0; c < 512; c++) {
                uint32_t h = (uint32_t) (c * 2654435761u) ^ (uint32_t) (threadIdx.x * 40503u);
                h ^= h >> 13; h *= 0x5bd1e995u; h ^= h >> 15;
                const float v = 1.0f + (float) (h & 0xffffff) * (1.0f / 16777216.0f);
                asm volatile("tcgen05.st.sync.aligned.32x32b.x1.b32 [%0], {%1};" ::"r"(taddr + c), "r"(__float_as_uint(v)) : "memory");
            }
            asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory");
        }
    }
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory");
    unsigned long long t0 = clock64();
    if (warp < warps_active) {
        const uint32_t taddr = tb + ((uint32_t) ((warp & 3) * 32) << 16) + (uint32_t) ((warp >> 2) * 256);
        St st;
        st.best = INFINITY; st.second = INFINITY; st.col = 0;
#pragma unroll
        for (int j = 0; j < 16; j++) st.snap[j] = INFINITY;
        uint32_t ra[32], rb[32];
        if (MODE == 1) {
            tmem_ld<32>(taddr, ra);
            tmem_ld<32>(taddr + 32, rb);
            ld_wait();
        }
        for (int i = 0; i < iters; i++) {
            // one "frame tile": 8 steps of 32 columns = 256 columns; state reset per tile
            if ((i & 7) == 0) { st.best = INFINITY; st.second = INFINITY; }
            if (MODE == 0) {
#pragma unroll
                for (int s = 0; s < 8; s += 2) {
                    tmem_ld<32>(taddr + 32 * s, ra);
                    tmem_ld<32>(taddr + 32 * s + 32, rb);
                    ld_wait();
                    st.best = fminf(st.best, __uint_as_float(ra[i & 31]) + __uint_as_float(rb[i & 31]));
                }
            } else if (MODE == 1) {
#pragma unroll
                for (int s = 0; s < 8; s += 2) {
                    step32<SNAP>(ra, 32 * s + 256 * i, st);
                    step32<SNAP>(rb, 32 * s + 32 + 256 * i, st);
                    // keep the values changing so nothing is hoisted
                    ra[(s + i) & 31] ^= 1u; rb[(s + i) & 31] ^= 1u;
                }
            } else if (MODE == 2) {
                tmem_ld<32>(taddr, ra);
#pragma unroll
                for (int s = 0; s < 8; s += 2) {
                    ld_wait();
                    tmem_ld<32>(taddr + 32 * (s + 1), rb);
                    step32<SNAP>(ra, 32 * s + 256 * i, st);
                    ld_wait();
                    if (s + 2 < 8) tmem_ld<32>(taddr + 32 * (s + 2), ra);
                    step32<SNAP>(rb, 32 * (s + 1) + 256 * i, st);
                }
            }
        }
        float o = st.best + st.second + (float) st.col;
#pragma unroll
        for (int j = 0; j < 16; j++) o += st.snap[j];
        if (o == 12345.f) sink[threadIdx.x] = o;
    }
    unsigned long long t1 = clock64();
    if (lane == 0) out_clk[blockIdx.x * 8 + warp] = t1 - t0;
    asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory");
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tb), "r"(512) : "memory");
}

template <int MODE, bool SNAP>
void run(const char *name, int warps_active, int iters) {
    unsigned long long *clk; float *sink;
    cudaMalloc(&clk, 148 * 8 * 8); cudaMalloc(&sink, 4096);
    cudaMemset(clk, 0, 148 * 8 * 8);
    for (int rep = 0; rep < 2; rep++) bench<MODE, SNAP><<<148, 256>>>(iters, clk, sink, warps_active);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("CUDA error %s\n", cudaGetErrorString(e)); exit(1); }
    unsigned long long h[8];
    cudaMemcpy(h, clk, sizeof(h), cudaMemcpyDeviceToHost);
    printf("%-34s warps %d: %7.1f clk per 32-column step per warp (warp 0), %7.1f (warp %d)\n", name, warps_active,
           (double) h[0] / (iters * 8.0), (double) h[warps_active - 1] / (iters * 8.0), warps_active - 1);
    cudaFree(clk); cudaFree(sink);
}

int main() {
    const int it = 4000;
    for (int w : {4, 8}) {
        run<0, false>("loads only (2 in flight)", w, it);
        run<1, false>("arithmetic only, no snapshot", w, it);
        run<1, true>("arithmetic only, snapshot", w, it);
        run<2, false>("loads + arithmetic, no snapshot", w, it);
        run<2, true>("loads + arithmetic, snapshot", w, it);
    }
    return 0;
}
