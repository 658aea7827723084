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
#include "vq_tc_common.cuh"

#include <cstdint>
#include <cstdlib>

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

// The same sums from a PADDED copy of the frames (rows of ld doubles, ld a multiple of 8 = whole
// 64-byte DRAM atoms, pad columns zero): a gathered 296-byte row of the dense array straddles
// atoms (ncu: 789 MB read for 608 MB of rows, 1.30x), a 320-byte row of the copy does not.  Lane l
// < ld / 2 loads components (2 l, 2 l + 1) as one 16-byte access: one load instruction per row.
__global__ void __launch_bounds__(256)
vq_cell_sums_padded(const double *__restrict__ xp, int ld, const int32_t *__restrict__ perm,
                    const int32_t *__restrict__ offsets, const int32_t *__restrict__ item_start, int K,
                    int D, PackedLayout L, double *__restrict__ packed) {
    const int lane = threadIdx.x & 31;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    const int n_items = item_start[K];
    const bool active = 2 * lane < ld;
    for (int item = warp; item < n_items; item += n_warps) {
        int lo = 0, hi = K;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (item_start[mid] <= item) lo = mid; else hi = mid;
        }
        const int c = lo;
        const int first = offsets[c] + (item - item_start[c]) * kCellChunk;
        const int last = min(first + kCellChunk, offsets[c + 1]);
        double s0 = 0.0, s1 = 0.0;
        for (int g = first; g < last; g += 32) {
            const int mine = g + lane < last ? perm[g + lane] : -1;
            const int cnt = min(32, last - g);
            int u = 0;
            for (; u + 8 <= cnt; u += 8) {
                double2 xv[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const long row = __shfl_sync(0xffffffffu, mine, u + k);
                    xv[k] = active ? reinterpret_cast<const double2 *>(xp + row * ld)[lane] : make_double2(0.0, 0.0);
                }
#pragma unroll
                for (int k = 0; k < 8; k++) { s0 += xv[k].x; s1 += xv[k].y; }
            }
            for (; u < cnt; u++) {
                const long row = __shfl_sync(0xffffffffu, mine, u);
                if (active) {
                    const double2 v = reinterpret_cast<const double2 *>(xp + row * ld)[lane];
                    s0 += v.x; s1 += v.y;
                }
            }
        }
        if (2 * lane < D) atomicAdd(packed + L.cells() + (long) c * D + 2 * lane, s0);
        if (2 * lane + 1 < D) atomicAdd(packed + L.cells() + (long) c * D + 2 * lane + 1, s1);
    }
}

// dense rows [n][D] -> padded rows [n][ld] (pad columns zero)
__global__ void vq_pad_rows(const double *__restrict__ x, double *__restrict__ xp, long n, int D, int ld) {
    const long total = n * ld;
    for (long i = (long) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (long) gridDim.x * blockDim.x) {
        const long row = i / ld;
        const int k = (int) (i - row * ld);
        xp[i] = k < D ? x[row * D + k] : 0.0;
    }
}

// D <= 16 (P <= 15): a row is at most 128 bytes, so one load instruction of the kernel above
// would keep 16+ lanes idle.  Here a warp load covers TWO rows (lanes 0-15: row u, lanes 16-31:
// row u + 1), 8 such loads in flight; the two half-warp sums meet in one shuffle at the end.
__global__ void __launch_bounds__(256)
vq_cell_sums_pair(const double *__restrict__ x64, const int32_t *__restrict__ perm,
                  const int32_t *__restrict__ offsets, const int32_t *__restrict__ item_start, int K,
                  int D, PackedLayout L, double *__restrict__ packed) {
    const int lane = threadIdx.x & 31;
    const int half = lane >> 4, m = lane & 15;
    const int warp = (blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const int n_warps = (gridDim.x * blockDim.x) >> 5;
    const int n_items = item_start[K];
    for (int item = warp; item < n_items; item += n_warps) {
        int lo = 0, hi = K;
        while (hi - lo > 1) {
            const int mid = (lo + hi) >> 1;
            if (item_start[mid] <= item) lo = mid; else hi = mid;
        }
        const int c = lo;
        const int first = offsets[c] + (item - item_start[c]) * kCellChunk;
        const int last = min(first + kCellChunk, offsets[c + 1]);
        double s = 0.0;
        for (int g = first; g < last; g += 32) {
            const int mine = g + lane < last ? perm[g + lane] : -1;
            const int cnt = min(32, last - g);
            for (int u = 0; u < cnt; u += 16) {
                double xv[8];
#pragma unroll
                for (int k = 0; k < 8; k++) {
                    const int src = u + 2 * k + half;             // this half-warp's row of pair k
                    const long row = __shfl_sync(0xffffffffu, mine, src & 31);
                    xv[k] = (src < cnt && m < D) ? x64[row * D + m] : 0.0;
                }
#pragma unroll
                for (int k = 0; k < 8; k++) s += xv[k];
            }
        }
        s += __shfl_down_sync(0xffffffffu, s, 16);
        if (lane < D) atomicAdd(packed + L.cells() + (long) c * D + lane, s);
    }
}

// ---- streaming variant for cell tables that fit in shared memory (K D 8 bytes <= ~200 KB:
// every stage up to K = 512 at P = 36, every stage of the P = 12 configuration).
// The frames are read ONCE in natural order, fully coalesced (a warp load = 32 consecutive doubles
// of the frame array, whatever D is), and added into a shared-memory table [replica][K][D] with
// FP64 compare-and-swap atomics (ATOMS.CAST.SPIN; shared memory has no native FP64 add); warps are
// spread over R replicas of the table so that small codebooks do not serialise on a few addresses.
// Reads 8 D bytes per frame instead of the gather of the sorted variant, whose 104..296-byte rows in
// cell order cost 1.3-1.6x their size in DRAM atoms; pays ~10 instructions per element for the CAS
// loop.  Used for short rows (D <= 20), where the gather is at its worst.  Counts go through a
// shared int table.
__global__ void __launch_bounds__(1024, 1)
vq_cell_sums_stream(const double *__restrict__ x64, const int32_t *__restrict__ idx, long n, int K, int D,
                    int R, PackedLayout L, double *__restrict__ packed, int32_t *__restrict__ hist,
                    uint16_t *__restrict__ seq) {
    extern __shared__ __align__(16) unsigned char cs_smem[];
    double *tab = reinterpret_cast<double *>(cs_smem);                 // [R][K * D]
    int32_t *cnt = reinterpret_cast<int32_t *>(tab + (size_t) R * K * D);   // [K]
    const int KD = K * D;
    for (int i = threadIdx.x; i < R * KD; i += blockDim.x) tab[i] = 0.0;
    for (int i = threadIdx.x; i < K; i += blockDim.x) cnt[i] = 0;
    __syncthreads();
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, n_warps = blockDim.x >> 5;
    double *mytab = tab + (size_t) (warp % R) * KD;
    // this CTA's contiguous range of frames, split into per-warp rounds of 32 doubles
    const long per_cta = (n + gridDim.x - 1) / gridDim.x;
    const long f_lo = (long) blockIdx.x * per_cta;
    const long f_hi = f_lo + per_cta < n ? f_lo + per_cta : n;
    if (f_lo < f_hi) {
        const long e_lo = f_lo * D, e_hi = f_hi * D;                    // element range
        // element offset -> (frame, component) by multiplication: q = floor(off / D) for every
        // off with off * D < 2^32 (checked by the launcher)
        const unsigned magic = (unsigned) ((0x100000000ull + (unsigned) D - 1) / (unsigned) D);
        constexpr int U = 4;                                            // loads in flight per lane
        for (long e0 = e_lo + (long) warp * 32 * U; e0 < e_hi; e0 += (long) n_warps * 32 * U) {
            double v[U];
            long e[U];
#pragma unroll
            for (int u = 0; u < U; u++) {
                e[u] = e0 + u * 32 + lane;
                v[u] = e[u] < e_hi ? x64[e[u]] : 0.0;
            }
#pragma unroll
            for (int u = 0; u < U; u++) {
                if (e[u] < e_hi) {
                    const unsigned off = (unsigned) (e[u] - e_lo);
                    const unsigned fr = __umulhi(off, magic);
                    const int m = (int) (off - fr * (unsigned) D);
                    const long f = f_lo + fr;
                    const int c = idx[f];
                    atomicAdd(mytab + c * D + m, v[u]);
                    if (m == 0) {
                        atomicAdd(cnt + c, 1);
                        if (seq) seq[f] = (uint16_t) c;
                    }
                }
            }
        }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < KD; i += blockDim.x) {
        double s = tab[i];
        for (int r = 1; r < R; r++) s += tab[(size_t) r * KD + i];
        if (s != 0.0) atomicAdd(packed + L.cells() + i, s);
    }
    for (int i = threadIdx.x; i < K; i += blockDim.x)
        if (cnt[i]) atomicAdd(hist + i, cnt[i]);
}

// ---- streaming variant without atomics: one PRIVATE table per warp --------------------------------
// For tables small enough that every consumer warp of the CTA holds its own copy ([K][D + 1] doubles,
// column D counts the frames; K <= 32 with 16 warps, K <= 64 with 8 at P = 36).  The frames are read
// ONCE, in natural order: a producer warp streams tiles of 64 consecutive frames and their cell
// indices (two cp.async.bulk per tile, S stages in flight) into a shared-memory ring; each of the W
// consumer warps takes 64 / W rows of a tile and adds them into ITS table with plain shared-memory
// read-modify-writes -- element (c, m) of a warp's table is only ever touched by lane m (mod 32) of
// that warp, so program order is all the ordering needed.  One flush per CTA at the end: K D FP64
// atomics.  A last partial tile is read from global memory.
// Measured on B200, 10M frames at P = 36: 0.63 ms (K = 16, 32) and 0.70 ms (K = 64) against 0.79 ms of
// the sorted gather; ncu: DRAM 58 %, shared-memory pipe 63 % (row in by bulk copy, row out, table row
// out and back in: ~1.2 KB per frame at 128 B/clk), issue slots 58 %.  Rows of 13 doubles (two frames per
// instruction were tried) gain nothing over vq_cell_sums_stream and keep that kernel.
constexpr int kPrivTile = 64;

template <int NR, int W>
__global__ void __launch_bounds__(32 * (W + 1), 1)
vq_cell_sums_private(const double *__restrict__ x64, const int32_t *__restrict__ idx, long n, int K, int D,
                     int S, PackedLayout L, double *__restrict__ packed, int32_t *__restrict__ hist,
                     uint16_t *__restrict__ seq) {
    constexpr int TF = kPrivTile, FPW = TF / W;
    static_assert(FPW % 4 == 0, "rows per warp: whole batches of 4");
    extern __shared__ __align__(128) unsigned char cp_smem[];
    const int RS = D + 1;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    double *tab_all = reinterpret_cast<double *>(cp_smem);
    const size_t tab_bytes = (sizeof(double) * (size_t) W * K * RS + 127) & ~(size_t) 127;
    unsigned char *ring = cp_smem + tab_bytes;
    // a stage = the tile's rows (a multiple of 16 bytes: TF is even) + its TF cell indices
    const uint32_t stage_bytes = (uint32_t) (TF * D * sizeof(double));
    const uint32_t idx_off = (stage_bytes + 127u) & ~127u;
    const uint32_t stage_stride = idx_off + TF * (uint32_t) sizeof(int32_t);
    uint64_t *full = reinterpret_cast<uint64_t *>(ring + (size_t) S * stage_stride);
    uint64_t *empty = full + S;
    if (threadIdx.x == 0) {
        for (int i = 0; i < S; i++) { mbar_init(&full[i], 1); mbar_init(&empty[i], W); }
        fence_barrier_init();
    }
    for (int e = threadIdx.x; e < W * K * RS; e += blockDim.x) tab_all[e] = 0.0;
    __syncthreads();

    const long n_full = n / TF;
    if (warp == W) {
        // producer: two bulk copies per tile
        if (lane == 0) {
            long i = 0;
            for (long t = blockIdx.x; t < n_full; t += gridDim.x, i++) {
                const int st = (int) (i % S);
                if (i >= S) mbar_wait(&empty[st], (uint32_t) (((i / S) & 1) ^ 1));   // (one lane spinning)
                mbar_expect_tx(&full[st], stage_bytes + TF * (uint32_t) sizeof(int32_t));
                bulk_g2s(ring + (size_t) st * stage_stride, x64 + t * TF * D, stage_bytes, &full[st]);
                bulk_g2s(ring + (size_t) st * stage_stride + idx_off, idx + t * TF, TF * (uint32_t) sizeof(int32_t),
                         &full[st]);
            }
        }
    } else {
        double *tab = tab_all + (size_t) warp * K * RS;
        // rows[0 .. cnt) of this warp, cc = cell of row `lane` (-1: none)
        auto process = [&](const double *rows, int cc, int cnt) {
            double xv[FPW][NR];
#pragma unroll
            for (int k = 0; k < FPW; k++)
#pragma unroll
                for (int r = 0; r < NR; r++) {
                    const int m = lane + 32 * r;
                    double v = 0.0;
                    if (k < cnt) { if (m < D) v = rows[k * D + m]; else if (m == D) v = 1.0; }
                    xv[k][r] = v;
                }
            // Batches of 4 frames.  The read-modify-writes of one frame after the other are a chain of
            // dependent LDS -> DADD -> STS (~100 clocks each); when the 4 cells of a batch differ
            // (warp-uniform test: one match + one ballot per call) the 4 chains are independent and
            // are issued together.
            const unsigned peers = __match_any_sync(0xffffffffu, cc);
            const unsigned batch = 0xFu << (lane & ~3);
            const unsigned uniq = __ballot_sync(0xffffffffu, cc >= 0 && (peers & batch) == (1u << lane));
#pragma unroll
            for (int b = 0; b < FPW / 4; b++) {
                if (((uniq >> (4 * b)) & 0xFu) == 0xFu) {
                    int cs[4];
                    double t[4][NR];
#pragma unroll
                    for (int j = 0; j < 4; j++) cs[j] = __shfl_sync(0xffffffffu, cc, 4 * b + j);
#pragma unroll
                    for (int j = 0; j < 4; j++)
#pragma unroll
                        for (int r = 0; r < NR; r++) {
                            const int m = lane + 32 * r;
                            t[j][r] = m <= D ? tab[cs[j] * RS + m] : 0.0;
                        }
#pragma unroll
                    for (int j = 0; j < 4; j++)
#pragma unroll
                        for (int r = 0; r < NR; r++) {
                            const int m = lane + 32 * r;
                            if (m <= D) tab[cs[j] * RS + m] = t[j][r] + xv[4 * b + j][r];
                        }
                    continue;
                }
#pragma unroll
                for (int k = 4 * b; k < 4 * b + 4; k++) {
                    const int c = __shfl_sync(0xffffffffu, cc, k);
                    if (c >= 0) {                              // warp-uniform
#pragma unroll
                        for (int r = 0; r < NR; r++) {
                            const int m = lane + 32 * r;
                            if (m <= D) tab[c * RS + m] += xv[k][r];
                        }
                    }
                }
            }
        };
        long i = 0;
        for (long t = blockIdx.x; t < n_full; t += gridDim.x, i++) {
            const int st = (int) (i % S);
            const long v0 = t * TF + warp * FPW;
            const unsigned char *stage = ring + (size_t) st * stage_stride;
            mbar_wait(&full[st], (uint32_t) ((i / S) & 1));
            const int cc = lane < FPW ? reinterpret_cast<const int32_t *>(stage + idx_off)[warp * FPW + lane] : -1;
            if (seq != nullptr && lane < FPW) seq[v0 + lane] = (uint16_t) cc;
            process(reinterpret_cast<const double *>(stage) + warp * FPW * D, cc, FPW);
            __syncwarp();
            if (lane == 0) mbar_arrive(&empty[st]);
        }
        if (n_full * TF < n && (long) blockIdx.x == n_full % gridDim.x) {
            const long v0 = n_full * TF + warp * FPW;
            long mine = n - v0;
            if (mine > FPW) mine = FPW;
            if (mine > 0) {
                const int cc = lane < mine ? idx[v0 + lane] : -1;
                if (seq != nullptr && lane < mine) seq[v0 + lane] = (uint16_t) cc;
                process(x64 + v0 * D, cc, (int) mine);
            }
        }
    }
    __syncthreads();
    for (int e = threadIdx.x; e < K * RS; e += blockDim.x) {
        double sum = tab_all[e];
        for (int i = 1; i < W; i++) sum += tab_all[(size_t) i * K * RS + e];
        if (sum != 0.0) {
            const int c = e / RS, m = e - c * RS;
            if (m < D) atomicAdd(packed + L.cells() + (long) c * D + m, sum);
            else atomicAdd(hist + c, (int32_t) (sum + 0.5));
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

void launch_cell_finish(const double *raas, const int32_t *hist, int K, int D, double *packed, cudaStream_t s) {
    const PackedLayout L{K, D};
    vq_cell_finish<<<(K + 7) / 8, 256, 0, s>>>(raas, hist, K, D, L, packed);
}

// shared memory of the streaming variant with R table replicas
static size_t stream_smem_bytes(int K, int D, int R) {
    return sizeof(double) * (size_t) R * K * D + sizeof(int32_t) * (size_t) K;
}
static bool cell_stream_enabled() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("ECOZ2_VQ_CELL_STREAM"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}

// ECOZ2_VQ_CELL_PRIV=0 switches the private-table variant off
static bool cell_private_enabled() {
    static int v = -1;
    if (v < 0) { const char *e = getenv("ECOZ2_VQ_CELL_PRIV"); v = (e && e[0] == '0') ? 0 : 1; }
    return v == 1;
}
int cell_pad_ld(int D) {
    static int on = -1;
    // OFF by default: measured on B200 (10M frames, K = 16 .. 2048) the sums take 0.79-0.82 ms from the
    // padded copy and 0.78-0.84 ms from the dense rows -- the gather is bound by DRAM random-access
    // efficiency (63-67 % of peak at 320-byte granularity), not by the 1.3x over-fetch, and the copy
    // costs 3.2 GB of HBM per 10M frames.  ECOZ2_VQ_CELL_PAD=1 switches it on for A/B runs.
    if (on < 0) { const char *e = getenv("ECOZ2_VQ_CELL_PAD"); on = (e && e[0] == '1') ? 1 : 0; }
    const int ld = (D + 7) / 8 * 8;
    // (short rows take the streaming kernel; rows that are whole atoms already need no copy)
    return (on && D > 20 && ld != D && ld <= 64) ? ld : 0;
}
void launch_pad_rows(const double *x64, double *xp, long n, int D, int ld, int num_sms, cudaStream_t s) {
    if (n <= 0) return;
    vq_pad_rows<<<num_sms * 16, 256, 0, s>>>(x64, xp, n, D, ld);
}

void launch_cell_sums(const double *x64, const int32_t *idx, long n, int D, const double *raas,
                      int K, double *packed, uint16_t *seq, const CellScratch &cs, int num_sms,
                      cudaStream_t s, const double *x64_padded, int ld) {
    const PackedLayout L{K, D};
    cudaMemsetAsync(cs.hist, 0, sizeof(int32_t) * (size_t) (K + 1), s);
    constexpr size_t kStreamSmemMax = 200 * 1024;
    // private tables (no atomics): rows of 16 .. 62 doubles, when 16 or 8 copies of the [K][D + 1] table
    // and a ring of at least 3 tiles fit into the shared memory of an SM
    if (n > 0 && cell_private_enabled() && D > 15 && D <= 62 && x64_padded == nullptr &&
        ((reinterpret_cast<uintptr_t>(x64) | reinterpret_cast<uintptr_t>(idx)) & 15) == 0) {
        constexpr size_t kPrivSmemMax = 224 * 1024;
        const size_t tab1 = sizeof(double) * (size_t) K * (D + 1);
        const size_t stage = ((sizeof(double) * (size_t) kPrivTile * D + 127) & ~(size_t) 127) +
                             sizeof(int32_t) * (size_t) kPrivTile;      // rows + cell indices of a tile
        int W = 0, S = 0;
        for (int w : {16, 8}) {
            const size_t tabs = (tab1 * w + 127) & ~(size_t) 127;
            if (tabs + 3 * (stage + 16) > kPrivSmemMax) continue;
            W = w;
            S = (int) ((kPrivSmemMax - tabs) / (stage + 16));
            if (S > 8) S = 8;
            break;
        }
        if (W > 0) {
            const size_t smem = ((tab1 * W + 127) & ~(size_t) 127) + (size_t) S * (stage + 16);
            const long tiles = (n + kPrivTile - 1) / kPrivTile;
            const long grid = tiles < num_sms ? tiles : num_sms;
#define PRIV_LAUNCH(NRV, WV)                                                                               \
    do {                                                                                                   \
        cudaFuncSetAttribute(vq_cell_sums_private<NRV, WV>, cudaFuncAttributeMaxDynamicSharedMemorySize,   \
                             (int) kPrivSmemMax);                                                          \
        vq_cell_sums_private<NRV, WV><<<(unsigned) grid, 32 * (WV + 1), smem, s>>>(                        \
            x64, idx, n, K, D, S, L, packed, cs.hist, seq);                                                \
    } while (0)
            if (D <= 31) { if (W == 16) PRIV_LAUNCH(1, 16); else PRIV_LAUNCH(1, 8); }
            else { if (W == 16) PRIV_LAUNCH(2, 16); else PRIV_LAUNCH(2, 8); }
#undef PRIV_LAUNCH
            vq_cell_finish<<<(K + 7) / 8, 256, 0, s>>>(raas, cs.hist, K, D, L, packed);
            return;
        }
    }
    const long stream_grid = n > 0 ? ((n + 4095) / 4096 < num_sms ? (n + 4095) / 4096 : num_sms) : 1;
    const long per_cta = (n + stream_grid - 1) / stream_grid;
    // measured on B200 (10M x P=36: 0.75 ms either way -- the CAS atomics cost what the gather
    // loses; 20M x P=12: 0.57-0.67 ms against 0.88-1.02): worth it for short rows only
    if (n > 0 && cell_stream_enabled() && D <= 20 && stream_smem_bytes(K, D, 1) <= kStreamSmemMax &&
        (double) per_cta * D * D < 4.0e9) {
        int R = (int) ((kStreamSmemMax - sizeof(int32_t) * (size_t) K) / (sizeof(double) * (size_t) K * D));
        if (R > 32) R = 32;
        // per device (function attributes belong to the context): set on every launch, ~1 us
        cudaFuncSetAttribute(vq_cell_sums_stream, cudaFuncAttributeMaxDynamicSharedMemorySize,
                             (int) kStreamSmemMax);
        const long grid = stream_grid;
        vq_cell_sums_stream<<<(unsigned) grid, 1024, stream_smem_bytes(K, D, R), s>>>(
            x64, idx, n, K, D, R, L, packed, cs.hist, seq);
        vq_cell_finish<<<(K * 32 + 255) / 256, 256, 0, s>>>(raas, cs.hist, K, D, L, packed);
        return;
    }
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
        if (x64_padded != nullptr && ld > 0) {
            vq_cell_sums_padded<<<grid, 256, 0, s>>>(x64_padded, ld, cs.perm, cs.offsets, cs.item_start, K, D, L, packed);
        } else if (D <= 16) {
            vq_cell_sums_pair<<<grid, 256, 0, s>>>(x64, cs.perm, cs.offsets, cs.item_start, K, D, L, packed);
        } else switch ((D + 31) / 32) {
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
