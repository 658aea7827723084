// vq_cells.cu -- kernel (b'), cell accumulation without per-frame atomics.
//
// reference: add_to_cell / init_cells (vq_learn.i:84-106), thread merge (vq_learn_par.c:88-111),
// DD += ddmin - 1 (vq_learn_par.c:77-79).
//
// vq_accumulate (vq_kernels.cu) issues D FP64 atomics per frame (consecutive frames rarely
// share a cell); at 10M frames that is 370M atomics and 2.2 ms per pass -- and far worse
// for small codebooks where the atomics collide.  Here the frames are first grouped by cell
// with a counting sort of their indices (histogram in shared memory, exclusive scan, scatter
// with shared-memory ranks), then one warp per (cell, 256-frame chunk) gathers its rows and
// sums them in registers: HBM traffic is the 8 D bytes per frame that must be read anyway,
// and only D atomics per 256 frames remain.
//
// The per-cell distortion needs no per-frame value either: the distortion is linear in the
// frame, so  sum_v (x_v . w_c - 1) = cellsum_c . w_c - count_c.  Per-frame minDist values
// are produced on demand (stage reports, ecoz2b200_last_assignment) by vq_accumulate's
// distortion-only instance.
#include "vq_kernels.cuh"

namespace vqb200 {

namespace {

constexpr int kCellChunk = 256;      // frames per work item of vq_cell_sums
constexpr int kScatterFrames = 16;   // frames per thread and block iteration of vq_cell_scatter

__global__ void __launch_bounds__(256)
vq_cell_hist(const int32_t *__restrict__ idx, long n, int K, int32_t *__restrict__ hist,
             uint16_t *__restrict__ seq) {
    extern __shared__ int32_t sh[];
    for (int c = threadIdx.x; c < K; c += blockDim.x) sh[c] = 0;
    __syncthreads();
    const long stride = (long) gridDim.x * blockDim.x;
    for (long v = (long) blockIdx.x * blockDim.x + threadIdx.x; v < n; v += stride) {
        const int c = idx[v];
        atomicAdd(&sh[c], 1);
        if (seq) seq[v] = (uint16_t) c;
    }
    __syncthreads();
    for (int c = threadIdx.x; c < K; c += blockDim.x)
        if (sh[c]) atomicAdd(&hist[c], sh[c]);
}

// one CTA of 1024 threads: offsets = exclusive scan of hist, cursor = offsets,
// item_start = exclusive scan of ceil(hist / kCellChunk)
__global__ void __launch_bounds__(1024)
vq_cell_scan(const int32_t *__restrict__ hist, int K, int32_t *__restrict__ offsets,
             int32_t *__restrict__ cursor, int32_t *__restrict__ item_start) {
    __shared__ int s_a[1024], s_b[1024];
    const int per = (K + 1023) / 1024;
    const int c0 = threadIdx.x * per;
    int a = 0, b = 0;
    for (int i = 0; i < per; i++) {
        const int c = c0 + i;
        if (c < K) {
            a += hist[c];
            b += (hist[c] + kCellChunk - 1) / kCellChunk;
        }
    }
    s_a[threadIdx.x] = a;
    s_b[threadIdx.x] = b;
    __syncthreads();
    // Hillis-Steele inclusive scan over the 1024 partials
    for (int off = 1; off < 1024; off <<= 1) {
        int va = 0, vb = 0;
        if ((int) threadIdx.x >= off) { va = s_a[threadIdx.x - off]; vb = s_b[threadIdx.x - off]; }
        __syncthreads();
        s_a[threadIdx.x] += va;
        s_b[threadIdx.x] += vb;
        __syncthreads();
    }
    int ra = s_a[threadIdx.x] - a, rb = s_b[threadIdx.x] - b;   // exclusive prefix of this thread
    for (int i = 0; i < per; i++) {
        const int c = c0 + i;
        if (c < K) {
            offsets[c] = ra;
            cursor[c] = ra;
            item_start[c] = rb;
            ra += hist[c];
            rb += (hist[c] + kCellChunk - 1) / kCellChunk;
        }
    }
    if (threadIdx.x == 1023) {
        offsets[K] = s_a[1023];
        item_start[K] = s_b[1023];
    }
}

__global__ void __launch_bounds__(256)
vq_cell_scatter(const int32_t *__restrict__ idx, long n, int K, int32_t *__restrict__ cursor,
                int32_t *__restrict__ perm) {
    extern __shared__ int32_t slot[];
    constexpr int FB = 256 * kScatterFrames;
    const long n_blocks = (n + FB - 1) / FB;
    for (long b = blockIdx.x; b < n_blocks; b += gridDim.x) {
        for (int c = threadIdx.x; c < K; c += 256) slot[c] = 0;
        __syncthreads();
        const long base = b * FB;
        int cell[kScatterFrames], rank[kScatterFrames];
#pragma unroll
        for (int i = 0; i < kScatterFrames; i++) {
            const long v = base + i * 256 + threadIdx.x;
            cell[i] = v < n ? idx[v] : -1;
        }
#pragma unroll
        for (int i = 0; i < kScatterFrames; i++)
            rank[i] = cell[i] >= 0 ? atomicAdd(&slot[cell[i]], 1) : 0;
        __syncthreads();
        // reserve this block's range of every cell it touches
        for (int c = threadIdx.x; c < K; c += 256) {
            const int cnt = slot[c];
            if (cnt) slot[c] = atomicAdd(&cursor[c], cnt);
        }
        __syncthreads();
#pragma unroll
        for (int i = 0; i < kScatterFrames; i++) {
            const long v = base + i * 256 + threadIdx.x;
            if (cell[i] >= 0) perm[slot[cell[i]] + rank[i]] = (int32_t) v;
        }
        __syncthreads();
    }
}

// one warp per (cell, chunk of kCellChunk frames): lane l owns components l, l + 32, ...
template <int NR>
__global__ void __launch_bounds__(256)
vq_cell_sums(const double *__restrict__ x64, const int32_t *__restrict__ perm,
             const int32_t *__restrict__ offsets, const int32_t *__restrict__ item_start, int K,
             int D, PackedLayout L, double *__restrict__ packed) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    const int n_items = item_start[K];
    for (int item = warp; item < n_items; item += n_warps) {
        // largest c with item_start[c] <= item
        int lo = 0, hi = K;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (item_start[mid] <= item) lo = mid; else hi = mid;
        }
        const int c = lo;
        const int first = offsets[c] + (item - item_start[c]) * kCellChunk;
        const int last = min(first + kCellChunk, offsets[c + 1]);
        double s[NR];
#pragma unroll
        for (int r = 0; r < NR; r++) s[r] = 0.0;
        for (int g = first; g < last; g += 32) {
            const int mine = g + lane < last ? perm[g + lane] : -1;
            const int cnt = min(32, last - g);
            int u = 0;
            for (; u + 8 <= cnt; u += 8) {
                double xv[8][NR];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const long row = __shfl_sync(0xffffffffu, mine, u + k);
#pragma unroll
                    for (int r = 0; r < NR; r++) {
                        const int m = lane + 32 * r;
                        xv[k][r] = m < D ? x64[row * D + m] : 0.0;
                    }
                }
#pragma unroll
                for (int k = 0; k < 8; k++)
#pragma unroll
                    for (int r = 0; r < NR; r++) s[r] += xv[k][r];
            }
            for (; u < cnt; u++) {
                const long row = __shfl_sync(0xffffffffu, mine, u);
#pragma unroll
                for (int r = 0; r < NR; r++) {
                    const int m = lane + 32 * r;
                    if (m < D) s[r] += x64[row * D + m];
                }
            }
        }
#pragma unroll
        for (int r = 0; r < NR; r++) {
            const int m = lane + 32 * r;
            if (m < D) atomicAdd(packed + L.cells() + (long) c * D + m, s[r]);
        }
    }
}

// counts, per-cell distortion  cellsum . w - count,  DD = sum of them; one warp per cell
__global__ void __launch_bounds__(256)
vq_cell_finish(const double *__restrict__ raas, const int32_t *__restrict__ hist, int K, int D,
               PackedLayout L, double *__restrict__ packed) {
    const int lane = threadIdx.x & 31;
    const int c = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    double dis = 0.0;
    if (c < K) {
        double part = 0.0;
        for (int m = lane; m < D; m += 32) {
            const double w = m > 0 ? 2.0 * raas[(long) c * D + m] : raas[(long) c * D];
            part = fma(packed[L.cells() + (long) c * D + m], w, part);
        }
#pragma unroll
        for (int off = 16; off >= 1; off >>= 1) part += __shfl_xor_sync(0xffffffffu, part, off);
        const int cnt = hist[c];
        dis = cnt > 0 ? part - (double) cnt : 0.0;
        if (lane == 0) {
            packed[L.counts() + c] = (double) cnt;
            packed[L.discel() + c] = dis;
        }
    }
    __shared__ double s_dd[8];
    if (lane == 0) s_dd[threadIdx.x >> 5] = dis;
    __syncthreads();
    if (threadIdx.x == 0) {
        double a = 0.0;
        for (int i = 0; i < 8; i++) a += s_dd[i];
        atomicAdd(packed + L.dd(), a);
    }
}

size_t align256(size_t v) { return (v + 255) & ~(size_t) 255; }

}  // namespace

bool cell_sort_supported(int K) { return K >= 2 && K <= 8192; }

size_t cell_scratch_bytes(long cap, int max_K) {
    return align256(sizeof(int32_t) * (size_t) cap) + 4 * align256(sizeof(int32_t) * (size_t) (max_K + 1));
}

void cell_scratch_carve(void *base, long cap, int max_K, CellScratch *out) {
    char *p = static_cast<char *>(base);
    out->perm = reinterpret_cast<int32_t *>(p); p += align256(sizeof(int32_t) * (size_t) cap);
    out->hist = reinterpret_cast<int32_t *>(p); p += align256(sizeof(int32_t) * (size_t) (max_K + 1));
    out->offsets = reinterpret_cast<int32_t *>(p); p += align256(sizeof(int32_t) * (size_t) (max_K + 1));
    out->cursor = reinterpret_cast<int32_t *>(p); p += align256(sizeof(int32_t) * (size_t) (max_K + 1));
    out->item_start = reinterpret_cast<int32_t *>(p);
    out->cap = cap;
}

void launch_cell_sums(const double *x64, const int32_t *idx, long n, int D, const double *raas,
                      int K, double *packed, uint16_t *seq, const CellScratch &cs, int num_sms,
                      cudaStream_t s) {
    const PackedLayout L{K, D};
    cudaMemsetAsync(cs.hist, 0, sizeof(int32_t) * (size_t) (K + 1), s);
    if (n > 0) {
        const size_t smem = sizeof(int32_t) * (size_t) K;
        long blocks = (n + 256 * 32 - 1) / (256 * 32);
        if (blocks > (long) num_sms * 8) blocks = (long) num_sms * 8;
        vq_cell_hist<<<(unsigned) blocks, 256, smem, s>>>(idx, n, K, cs.hist, seq);
    }
    vq_cell_scan<<<1, 1024, 0, s>>>(cs.hist, K, cs.offsets, cs.cursor, cs.item_start);
    if (n > 0) {
        const size_t smem = sizeof(int32_t) * (size_t) K;
        long blocks = (n + 256 * kScatterFrames - 1) / (256 * kScatterFrames);
        if (blocks > (long) num_sms * 4) blocks = (long) num_sms * 4;
        vq_cell_scatter<<<(unsigned) blocks, 256, smem, s>>>(idx, n, K, cs.cursor, cs.perm);
        const int grid = num_sms * 8;
        switch ((D + 31) / 32) {
            case 1: vq_cell_sums<1><<<grid, 256, 0, s>>>(x64, cs.perm, cs.offsets, cs.item_start, K, D, L, packed); break;
            case 2: vq_cell_sums<2><<<grid, 256, 0, s>>>(x64, cs.perm, cs.offsets, cs.item_start, K, D, L, packed); break;
            case 3: vq_cell_sums<3><<<grid, 256, 0, s>>>(x64, cs.perm, cs.offsets, cs.item_start, K, D, L, packed); break;
            case 4: vq_cell_sums<4><<<grid, 256, 0, s>>>(x64, cs.perm, cs.offsets, cs.item_start, K, D, L, packed); break;
            default: vq_cell_sums<7><<<grid, 256, 0, s>>>(x64, cs.perm, cs.offsets, cs.item_start, K, D, L, packed); break;
        }
    }
    vq_cell_finish<<<(K + 7) / 8, 256, 0, s>>>(raas, cs.hist, K, D, L, packed);
}

}  // namespace vqb200
