// vq_mid.cu -- the whole k-means pass in ONE sweep for codebooks that fit one tensor-core tile
// (K <= 128 at P = 36): search + cell sums, the FP64 frames read once.
//
// reference: assign loop vq_learn_par.c:62-81, add_to_cell vq_learn.i:99-106.
//
// Stages K = 16 .. 128 of `vq learn` are HBM-bound (K/2 flop per byte), and the two-kernel pass read
// the frames twice: FP32 copies for the tensor-core search, FP64 originals gathered in cell order
// for the sums (4.44 GB per 10M frames against 2.96 GB that must be read).  Here a persistent CTA
// per SM streams half-tiles of 128 FP64 frames through a ring of bulk copies and does everything
// while a half-tile is in shared memory:
//
//   warp 16      raw producer: ONE cp.async.bulk per half-tile (128 rows x D doubles, contiguous)
//   warps 0-7    loaders (one warpgroup per half of a tile pair), thread = frame row: FP64 row ->
//                FP32 -> power-of-two scaling -> FP16 (hi, lo) split -> UMMA canonical K-major
//                operand (same numerics as vq_assign_tc2)
//   warp 17      MMA issuer: the codebook tile is resident in shared memory for the whole kernel
//                (loaded once by warp 18); 8 tcgen05.mma per half-tile into a TMEM accumulator
//   warps 8-15   epilogue + cell sums.  Epilogue (one warpgroup per half, thread = frame):
//                tcgen05.ld, (best, runner-up, argbest), error-bound test; writes idx, appends frames
//                inside the bound to flag_list and every other frame's row number to the list of
//                its cell (s_list[ring slot][cell][CAP], one shared-memory atomic per frame; a row
//                that finds its list full goes to late_list).  Cell sums (all 8 warps, every
//                half-tile): lane = cell, warp = a slice of NCW components kept in REGISTERS for the
//                whole kernel; a lane walks its cell's row list and adds the FP64 row -- still in
//                the raw buffer -- with plain loads and DADDs: no atomics on the sums, no table in
//                shared memory.  The raw buffer returns to the ring when the loaders and all 8
//                warps are done with it.
//
// Frames inside the error bound are NOT accumulated here: they go through the usual FP64
// re-evaluation (launch_recheck) and vq_cell_fixup adds their rows (and the late_list rows)
// afterwards with the final index.  At the end every warp flushes its registers with one FP64
// atomic per (cell, component); counts, per-cell distortions and DD follow from the sums exactly as
// in vq_cells.cu (launch_cell_finish).
//
// STATUS: experimental, OFF by default (ECOZ2_VQ_MID_MAXK=0).  Measured on B200, 10M frames, per pass
// at K = 16 / 32 / 64 / 128: 2.0 / 2.1 / 1.34 / 2.0 ms against 1.5 ms of search + sorted sums.  Earlier
// designs: shared-memory cell table by ownership 2.5-2.9 ms; ballots per cell 1.56 / 1.59 / 1.79 /
// 2.55; register switch 1.8-4.4.  Without the cell-sum loop (ECOZ2_VQ_MID_EXP=1) the sweep takes
// 1.06-1.10 ms: the loaders' FP64 -> FP16 conversion bounds it.  DESIGN.md 10.3.
#include "vq_kernels.cuh"
#include "vq_tc_common.cuh"

#include <cstdio>
#include <cstdlib>
#include <vector>

namespace vqb200 {

namespace {

constexpr int kMidThreads = 640;          // 5 warpgroups: 2 x loaders, 2 x epilogue / cell sums, roles
// setmaxnreg moves registers inside the CTA's LAUNCH allocation (640 threads x 96 = 61440), never
// beyond it: 128 x (2 x 96 loaders + 2 x 112 epilogue + 64 roles) = 61440
constexpr int kMidEpiRegs = 112, kMidSmallRegs = 64;   // loaders stay at the launch value 96
constexpr int kMidMaxRaw = 4;
#ifndef MID_SLEEP_PRODUCER
#define MID_SLEEP_PRODUCER 0
#define MID_SLEEP_MMA 0
#define MID_SLEEP_LOADER 0
#endif

struct MidParams {
    const double *x;            // frames [n][D] FP64
    long n_frames;
    int D;
    const float *wtc;           // the codebook's ONE tile, vq_prepare_codebook_tc2h layout
    int K;
    int BN;                     // tile columns (UMMA N): 64, 128 or 256, >= K
    const float *w1;            // W1'_j (bound weights, behind the tile)
    float err_scale;            // kappa * 2^-24
    int32_t *idx_out;
    int32_t *flag_list;
    int32_t *flag_count;
    float *flag_ub;
    unsigned long long *min_key;
    double *cells;              // packed cell sums [K][D]
    int32_t *hist;              // frames per cell [K]
    int32_t *late_list;         // rows whose cell list was full: added by vq_cell_fixup
    int32_t *late_count;
    int list_cap;               // rows per cell list and half-tile (1024 / cells, at most 128)
    int n_raw;                  // depth of the raw half-tile ring
    int staged;                 // 1: x is 16-byte aligned, half-tiles arrive as bulk copies
    unsigned long long *prof;   // dev: per CTA wait clocks of every role
    int exp;                    // dev experiments (ECOZ2_VQ_MID_EXP): 1 no cell-sum loop, 2 no lists, 4 no tmem epilogue
};

struct MidSmem {
    uint64_t raw_full[8], raw_empty[8], cell_full[8];
    uint64_t a_full[2], a_free[2];
    uint64_t acc_full[4], acc_empty[4];
    uint64_t b_full;
    uint32_t tmem_slot;
};

__device__ __forceinline__ void mbar_wait_prof(uint64_t *bar, uint32_t parity, long long &acc, bool prof) {
    if (!prof) { mbar_wait(bar, parity); return; }
    const long long t0 = clock64();
    mbar_wait(bar, parity);
    acc += clock64() - t0;
}
// wait of a role that is normally AHEAD of its producer (loaders for data, the raw producer for a free
// slot, the MMA warp for operands): sleeping between polls keeps its polling out of the issue slots and
// the shared-memory pipe of the epilogue / cell-sum warps, which are the critical path
template <int NS>
__device__ __forceinline__ void mbar_wait_sleep(uint64_t *bar, uint32_t parity, long long &acc, bool prof) {
    const long long t0 = prof ? clock64() : 0;
    while (!mbar_try_wait(bar, parity)) {
        if (NS > 0) __nanosleep(NS);
    }
    if (prof) acc += clock64() - t0;
}

}  // namespace

// KS: 16-byte chunks per split part of a row (D + 1 <= 8 KS); NCW: components per cell-sum warp
template <int KS, int NCW>
__global__ void __launch_bounds__(kMidThreads, 1) vq_pass_mid(const MidParams p) {
    extern __shared__ __align__(1024) unsigned char smem_raw[];
    MidSmem *sm = reinterpret_cast<MidSmem *>(smem_raw);
    const int D = p.D, K = p.K, BN = p.BN;
    const uint32_t NRAW = (uint32_t) p.n_raw;
    const bool PROF = p.prof != nullptr;
    constexpr int KC = KS;                       // 16-byte chunks per split part (8 halves each)
    constexpr int BC = 2 * KC + (KC & 1);        // chunks per codeword row: [lo | hi | zero pad]
    constexpr int NE = KC * 8;                   // padded row length
    constexpr uint32_t a_part_bytes = (uint32_t) KC * kBM * 16;
    constexpr uint32_t a_half_bytes = 2 * a_part_bytes;
    float *s_x0 = reinterpret_cast<float *>(smem_raw + 1024);                     // [NRAW][128]
    int32_t *s_cell = reinterpret_cast<int32_t *>(smem_raw + 1024 + NRAW * 512);  // [NRAW][128]
    int32_t *s_lcnt = reinterpret_cast<int32_t *>(smem_raw + 1024 + NRAW * 1024); // [NRAW][128] rows per cell
    uint8_t *s_list = smem_raw + 1024 + NRAW * 1536;                              // [NRAW][K][CAP] row numbers
    float *s_w1 = reinterpret_cast<float *>(smem_raw + 1024 + NRAW * 2560);       // [128] bound weights
    const double *s_zero = reinterpret_cast<const double *>(smem_raw + 1536 + NRAW * 2560);   // [64] zeros
    const uint32_t hdr_bytes = 2048 + NRAW * 2560;
    const uint32_t raw_bytes = (uint32_t) kBM * D * 8;
    const uint32_t raw_stride = (raw_bytes + 127u) & ~127u;
    unsigned char *raw_base = smem_raw + hdr_bytes;
    unsigned char *a_base = raw_base + NRAW * raw_stride;
    unsigned char *b_base = a_base + 2 * a_half_bytes;
    const uint32_t tile_bytes = (uint32_t) BC * BN * 16;

    const int tid = threadIdx.x;
    const int warp = tid >> 5, lane = tid & 31;
    const long total = p.n_frames;
    const long n_ftiles = (total + 2 * kBM - 1) / (2 * kBM);
    const bool wide = BN == 256;

    if (tid == 0) {
        for (int i = 0; i < 8; i++) {
            mbar_init(&sm->raw_full[i], 1);
            mbar_init(&sm->raw_empty[i], 12);    // 4 loader warps of the half + 8 epilogue / cell-sum warps
            mbar_init(&sm->cell_full[i], 4);
        }
        for (int i = 0; i < 2; i++) {
            mbar_init(&sm->a_full[i], 4);
            mbar_init(&sm->a_free[i], 1);
        }
        for (int i = 0; i < 4; i++) {
            mbar_init(&sm->acc_full[i], 1);
            mbar_init(&sm->acc_empty[i], 4);
        }
        mbar_init(&sm->b_full, 1);
        fence_barrier_init();
    }
    if (warp == 17) tmem_alloc(&sm->tmem_slot, 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem_base = sm->tmem_slot;
    long long pw0 = 0, pw1 = 0, pw2 = 0;   // dev: clocks spent in the role's waits
    const long long pt0 = PROF ? clock64() : 0;

    if (warp >= 16) {
        setmaxnreg_dec<kMidSmallRegs>();
        if (warp == 18) {
            // ===== codebook: one tile, resident for the whole kernel =====
            if (lane == 0) {
                mbar_expect_tx(&sm->b_full, tile_bytes);
                bulk_g2s(b_base, p.wtc, tile_bytes, &sm->b_full);
            }
        } else if (warp == 16) {
            // ===== raw producer =====
            if (lane == 0) {
                uint32_t j = 0;
                for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x) {
                    for (int h = 0; h < 2; h++, j++) {
                        const long base = ft * (2 * kBM) + h * kBM;
                        const uint32_t rb = j % NRAW, rph = (j / NRAW) & 1;
                        mbar_wait_sleep<MID_SLEEP_PRODUCER>(&sm->raw_empty[rb], rph ^ 1, pw0, PROF);
                        if (p.staged && base + kBM <= total) {
                            mbar_expect_tx(&sm->raw_full[rb], raw_bytes);
                            bulk_g2s(raw_base + rb * raw_stride, p.x + base * D, raw_bytes, &sm->raw_full[rb]);
                        } else {
                            mbar_arrive(&sm->raw_full[rb]);   // the loaders fetch the rows themselves
                        }
                    }
                }
            }
        } else if (warp == 17) {
            // ===== MMA issuer (whole warp in the loop, one elected lane issues) =====
            const uint32_t idesc = umma_idesc_f16(kBM, BN);
            const uint64_t dA0 = umma_desc(s32(a_base), kBM * 16, 128);
            const uint64_t dB0 = umma_desc(s32(b_base), BN * 16, 128);
            constexpr uint32_t a_half16 = a_half_bytes >> 4;
            constexpr uint32_t a_k16 = (2 * kBM * 16) >> 4;
            const uint32_t b_part16 = (uint32_t) (KC * BN * 16) >> 4;
            const uint32_t b_k16 = (uint32_t) (2 * BN * 16) >> 4;
            mbar_wait(&sm->b_full, 0);
            uint32_t j = 0;
            for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x) {
                for (int h = 0; h < 2; h++, j++) {
                    const uint32_t aph = (j >> 1) & 1;
                    const uint32_t buf = wide ? 0u : ((j >> 1) & 1);
                    const uint32_t acc_ph = wide ? ((j >> 1) & 1) : ((j >> 2) & 1);
                    mbar_wait_sleep<MID_SLEEP_MMA>(&sm->a_full[h], aph, pw0, PROF);
                    mbar_wait_prof(&sm->acc_empty[2 * buf + h], acc_ph ^ 1, pw1, PROF);
                    tc_fence_after();
                    const uint64_t dah = dA0 + (uint32_t) h * a_half16;          // [xh | xl] of this half
                    const uint64_t dbh = dB0, dbl = dB0 + b_part16;              // [wl | wh | 0]
                    const uint32_t d_tmem = tmem_base + (wide ? (uint32_t) h * 256u : (2 * buf + h) * 128u);
                    if (elect_one()) {
                        // the two cross terms as ONE product over [xh|xl].[wl|wh], then xh.wh
#pragma unroll
                        for (int s = 0; s < KC; s++)
                            umma_f16(d_tmem, dah + s * a_k16, dbh + s * b_k16, idesc, s > 0);
#pragma unroll
                        for (int s = 0; s < (KC + 1) / 2; s++)
                            umma_f16(d_tmem, dah + s * a_k16, dbl + s * b_k16, idesc, 1);
                        umma_commit(&sm->acc_full[2 * buf + h]);
                        umma_commit(&sm->a_free[h]);
                    }
                    __syncwarp();
                }
            }
        }
    } else if (warp < 8) {
        // ===== loaders: one warpgroup per half of every tile, thread = frame row =====
        const int h = warp >> 2;
        const int r = (warp & 3) * 32 + lane;
        uint32_t j = (uint32_t) h;
        for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x, j += 2) {
            const long base = ft * (2 * kBM) + h * kBM;
            const uint32_t rb = j % NRAW, rph = (j / NRAW) & 1;
            const bool staged = p.staged && base + kBM <= total;
            mbar_wait_sleep<MID_SLEEP_LOADER>(&sm->raw_full[rb], rph, pw0, PROF);
            double *rows = reinterpret_cast<double *>(raw_base + rb * raw_stride);
            float v[NE];
            if (staged) {
                const double *src = rows + r * D;
#pragma unroll
                for (int k = 0; k < NE; k++) v[k] = k < D ? (float) src[k] : 0.f;
            } else {
                // partial half-tile / unaligned frames: global loads, and the FP64 row is parked in
                // the raw buffer for the cell sums
                const bool valid = base + r < total;
                const double *src = p.x + (valid ? base + r : 0) * D;
#pragma unroll
                for (int k = 0; k < NE; k++) {
                    const double xv = (valid && k < D) ? src[k] : 0.0;
                    if (k < D) rows[r * D + k] = xv;
                    v[k] = (float) xv;
                }
            }
            // the staging buffer goes back to the ring only when every lane's loads have RETURNED:
            // the arrive is made data dependent on them
            uint32_t dep = 0;
#pragma unroll
            for (int k = 0; k < NE; k++) dep ^= __float_as_uint(v[k]);
            fence_proxy_async();
            const uint32_t all_in = __all_sync(0xffffffffu, dep != 0x7fc12345u);
            if (lane == 0) mbar_arrive_dep(&sm->raw_empty[rb], all_in);
            // power-of-two scaling into max_k |x[k]| in [2^14, 2^15): both split parts stay inside
            // the FP16 range (see vq_assign_tc2)
            float xm = 0.f;
#pragma unroll
            for (int k = 0; k < NE; k++) xm = fmaxf(xm, fabsf(v[k]));
            uint32_t ef = (__float_as_uint(xm) >> 23) & 0xffu;
            ef = ef < 16u ? 16u : ef;
            const float scale = __uint_as_float((268u - ef) << 23);
#pragma unroll
            for (int k = 0; k < NE; k++) v[k] *= scale;
            const float x0 = xm * scale;
            // slot D carries -e x0: the accumulator becomes the lower bound d~ - e x0 W1'_j
#pragma unroll
            for (int k = 0; k < NE; k++)
                if (k == D) v[k] = -1.001f * p.err_scale * x0;
            s_x0[rb * kBM + r] = x0;
            mbar_wait_prof(&sm->a_free[h], ((j >> 1) & 1) ^ 1, pw1, PROF);
            unsigned char *hi = a_base + (size_t) h * a_half_bytes;
            unsigned char *lo = hi + a_part_bytes;
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
                *reinterpret_cast<uint4 *>(hi + ((size_t) c * kBM + r) * 16) = make_uint4(hw[0], hw[1], hw[2], hw[3]);
                *reinterpret_cast<uint4 *>(lo + ((size_t) c * kBM + r) * 16) = make_uint4(lw[0], lw[1], lw[2], lw[3]);
            }
            fence_proxy_async();
            __syncwarp();
            if (lane == 0) mbar_arrive(&sm->a_full[h]);
        }
    } else {
        // ===== epilogue (warpgroup per half, thread = frame) + cell sums (all 8 warps, every half-tile) =====
        setmaxnreg_inc<kMidEpiRegs>();
        const int q = warp & 3;
        const int h = (warp - 8) >> 2;
        const int w8 = warp - 8;
        const int r = q * 32 + lane;
        const int t8 = tid - 8 * 32;             // 0 .. 255
        const uint32_t lane_addr = tmem_base + ((uint32_t) (q * 32) << 16);
        const float e2 = 2.f * 1.001f * p.err_scale;
        // cell sums: LANE = cell, REGISTERS = components.  G groups of 32 cells, each group's D components
        // split over NP = 8 / G warps (NCW components per warp, statically indexed registers).  Codebooks
        // of at most 16 cells use two lanes per cell (alternate rows of its list).
        const int G = K > 64 ? 4 : (K > 32 ? 2 : 1);
        const int gi = w8 % G, pi = w8 / G;
        const int LPC = K <= 16 ? 2 : 1;
        const int mc = K <= 16 ? (lane & 15) : gi * 32 + lane;      // this lane's cell (>= K: never filled)
        const int sub = K <= 16 ? (lane >> 4) : 0;
        const int m_lo = pi * NCW;                                  // first component of this warp
        const int CAP = p.list_cap;                                 // rows a cell's list takes per half-tile
        double acc[NCW];
#pragma unroll
        for (int m = 0; m < NCW; m++) acc[m] = 0.0;
        int my_count = 0;
        for (int e = t8; e < NRAW * kBM; e += 256) s_lcnt[e] = 0;
        for (int e = t8; e < K; e += 256) s_w1[e] = p.w1[e];
        if (t8 < 64) reinterpret_cast<double *>(smem_raw + 1536 + NRAW * 2560)[t8] = 0.0;
        asm volatile("bar.sync 1, 256;" ::: "memory");
        uint32_t j = 0;
        for (long ft = blockIdx.x; ft < n_ftiles; ft += gridDim.x) {
            for (int hh = 0; hh < 2; hh++, j++) {
                const uint32_t rb = j % NRAW;
                int32_t *lcnt = s_lcnt + rb * kBM;
                uint8_t *list = s_list + rb * 1024;
                if (hh == h) {
                    // ---- epilogue of this warpgroup's half
                    const long base = ft * (2 * kBM) + h * kBM;
                    const long row = base + r < total ? base + r : -1;
                    const uint32_t buf = wide ? 0u : ((j >> 1) & 1);
                    const uint32_t acc_ph = wide ? ((j >> 1) & 1) : ((j >> 2) & 1);
                    EpiState est;
                    est.init();
                    mbar_wait_prof(&sm->acc_full[2 * buf + h], acc_ph, pw0, PROF);
                    tc_fence_after();
                    // the accumulator is full, so the ring slot was refilled, so ALL 8 warps are done with
                    // the slot's previous half-tile: its list counters can be cleared for this one
                    lcnt[r] = 0;
                    const uint32_t taddr = lane_addr + (wide ? (uint32_t) h * 256u : (2 * buf + h) * 128u);
                    if (p.exp & 4) {
                        est.best = 1.f; est.second = 2.f;
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&sm->acc_empty[2 * buf + h]);
                    } else if (K == 128) {
                        epi_tile_full<128>(taddr, 0, est, &sm->acc_empty[2 * buf + h]);
                    } else if (K == 64) {
                        epi_tile_full<64>(taddr, 0, est, &sm->acc_empty[2 * buf + h]);
                    } else {
                        // ragged codebook: columns >= K are masked inside epi_step32
                        uint32_t ra[32];
                        for (int c = 0; c < K; c += 32) {
                            tmem_ld32(taddr + c, ra);
                            tmem_ld_wait();
                            epi_step32(ra, c, K - c, est, 0);
                        }
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) mbar_arrive(&sm->acc_empty[2 * buf + h]);
                    }
                    float second;
                    int bidx;
                    epi_finish(est, second, bidx);
                    asm volatile("bar.sync %0, 128;" ::"r"(2 + h) : "memory");   // counters cleared by all 128 threads
                    if (row >= 0) {
                        p.idx_out[row] = bidx;
                        // accumulator = lower bound L_ij of the true distortion; bidx is certainly the true
                        // argmin when every other lower bound exceeds its upper bound L + 2 e x0 W1'.  NaN-safe.
                        const float x0 = s_x0[rb * kBM + r];
                        const float ub = fmaf(e2 * x0, s_w1[bidx], est.best);
                        if (!(second > ub)) {
                            const int pos = atomicAdd(p.flag_count, 1);
                            p.flag_list[pos] = (int32_t) row;
                            p.flag_ub[pos] = ub;
                            p.min_key[pos] = ~0ull;
                        } else {
                            // the row joins its cell's list of this half-tile; a full list (one cell taking
                            // more than CAP of the 128 rows) sends it to the late list: vq_cell_fixup adds it
                            const int pos = (p.exp & 2) ? 0 : atomicAdd(&lcnt[bidx], 1);
                            if (p.exp & 2) {
                            } else if (pos < CAP) {
                                list[bidx * CAP + pos] = (uint8_t) r;
                            } else {
                                const int lp = atomicAdd(p.late_count, 1);
                                p.late_list[lp] = (int32_t) row;
                            }
                        }
                    }
                    __syncwarp();
                    if (lane == 0) mbar_arrive(&sm->cell_full[rb]);
                }
                // ---- cell sums of half-tile j: every lane walks the row list of ITS cell and adds the
                // components this warp owns; all loads of a step are independent
                mbar_wait_prof(&sm->cell_full[rb], (j / NRAW) & 1, pw1, PROF);
                const double *rows = reinterpret_cast<const double *>(raw_base + rb * raw_stride);
                int n = mc < K ? lcnt[mc] : 0;
                n = n < CAP ? n : CAP;
                if (pi == 0 && sub == 0) my_count += n;
                const int steps = (n - sub + LPC - 1) / LPC;          // rows sub, sub + LPC, ...
                const int max_steps = __reduce_max_sync(0xffffffffu, steps > 0 ? steps : 0);
                // dead lanes of a step read a row of zeros; components past D (last part only) read into the
                // next row and land in registers that are never flushed: no selects, no bounds checks
                const uint8_t *my_list = list + mc * CAP + sub;
                for (int st = 0; st < ((p.exp & 1) ? 0 : max_steps); st++) {
                    const double *xr = st < steps ? rows + (int) my_list[st * LPC] * D + m_lo : s_zero;
#pragma unroll
                    for (int m = 0; m < NCW; m++) acc[m] += xr[m];
                }
                // every row this warp needs has been read (the arrive depends on the sums): the raw buffer
                // and the slot's lists may be reused
                uint32_t dep = (uint32_t) max_steps;
#pragma unroll
                for (int m = 0; m < NCW; m++) dep ^= (uint32_t) __double2hiint(acc[m]);
                const uint32_t all_in = __all_sync(0xffffffffu, dep != 0x7ff12345u);
                if (lane == 0) mbar_arrive_dep(&sm->raw_empty[rb], all_in);
            }
        }
        // one FP64 atomic per (CTA, cell, component)
        if (mc < K) {
#pragma unroll
            for (int m = 0; m < NCW; m++)
                if (m_lo + m < D && acc[m] != 0.0) atomicAdd(p.cells + (long) mc * D + m_lo + m, acc[m]);
            if (my_count > 0) atomicAdd(p.hist + mc, my_count);
        }
    }

    if (PROF && lane == 0) {
        unsigned long long *o = p.prof + ((size_t) blockIdx.x * 20 + warp) * 4;
        o[0] = (unsigned long long) pw0; o[1] = (unsigned long long) pw1; o[2] = (unsigned long long) pw2;
        o[3] = (unsigned long long) (clock64() - pt0);
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 17) tmem_dealloc(tmem_base, 512);
}

// rows the fused sweep did not add -- the flagged frames after their FP64 re-evaluation, and the rows
// whose cell list was full: add the row to its final cell (FP64 global atomics; normally a fraction of
// a percent of the frames); one warp per frame
__global__ void __launch_bounds__(256)
vq_cell_fixup(const double *__restrict__ x64, const int32_t *__restrict__ idx, int D,
              const int32_t *__restrict__ list, const int32_t *__restrict__ count,
              double *__restrict__ cells, int32_t *__restrict__ hist) {
    const int lane = threadIdx.x & 31;
    const long warp = ((long) blockIdx.x * blockDim.x + threadIdx.x) >> 5;
    const long n_warps = ((long) gridDim.x * blockDim.x) >> 5;
    const long total = *count;
    for (long k = warp; k < total; k += n_warps) {
        const long row = list[k];
        const int c = idx[row];
        for (int m = lane; m < D; m += 32) atomicAdd(cells + (long) c * D + m, x64[row * D + m]);
        if (lane == 0) atomicAdd(hist + c, 1);
    }
}

static size_t mid_smem_bytes(int D, int BN, int n_raw) {
    const int KC = (D + 8) / 8;
    const int BC = 2 * KC + (KC & 1);
    const size_t raw_stride = ((size_t) kBM * D * 8 + 127) & ~(size_t) 127;
    return (size_t) 2048 + (size_t) n_raw * 2560 + n_raw * raw_stride + (size_t) 2 * 2 * KC * kBM * 16 +
           (size_t) BC * BN * 16;
}
static int mid_raw_depth(int D, int BN) {
    int n = kMidMaxRaw;
    while (n > 0 && mid_smem_bytes(D, BN, n) > 227 * 1024) n--;
    return n;
}
static int mid_max_k() {
    // experimental, off by default: on B200 the fused sweep does not beat search + sorted cell sums yet
    // (see DESIGN.md); ECOZ2_VQ_MID_MAXK=128 switches it on for 9 <= K <= 128 (read at every pass)
    const char *e = getenv("ECOZ2_VQ_MID_MAXK");
    int v = e ? atoi(e) : 0;
    return v > 128 ? 128 : v;
}

bool pass_mid_supported(int K, int D) {
    if (!tc_operands_f16() || K < 9 || K > mid_max_k()) return false;
    const int KC = (D + 8) / 8;
    if (!(KC == 2 || KC == 3 || KC == 5)) return false;   // instantiated: P = 12 (8..14), 20 (16..22), 36 (32..38)
    const int BN = tc_tile_cols(K, D);
    if (K > BN) return false;                       // the codebook must be ONE tile
    return mid_raw_depth(D, BN) >= 2;
}

template <int KS, int NCW>
static void launch_mid_inst(const MidParams &p, long grid, size_t smem, cudaStream_t s) {
    cudaFuncSetAttribute(vq_pass_mid<KS, NCW>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024);
    vq_pass_mid<KS, NCW><<<(unsigned) grid, kMidThreads, smem, s>>>(p);
    const cudaError_t e = cudaPeekAtLastError();
    if (e != cudaSuccess) fprintf(stderr, "ecoz2_vq_b200: vq_pass_mid launch: %s\n", cudaGetErrorString(e));
}
// NP = 8 / G warps share the D components of a group of 32 cells: NCW = ceil(Dmax / NP) each
template <int KS, int DMAX>
static void launch_mid_k(const MidParams &p, long grid, size_t smem, cudaStream_t s) {
    if (p.K > 64) launch_mid_inst<KS, (DMAX + 1) / 2>(p, grid, smem, s);
    else if (p.K > 32) launch_mid_inst<KS, (DMAX + 3) / 4>(p, grid, smem, s);
    else launch_mid_inst<KS, (DMAX + 7) / 8>(p, grid, smem, s);
}

void launch_pass_mid(const double *x64, long n, int D, const float *wtc, int K, const SearchScratch &sc,
                     float err_scale, int32_t *idx, double *packed, int32_t *hist, int32_t *late_list,
                     int32_t *late_count, bool zero_hist, int num_sms, cudaStream_t s) {
    if (zero_hist) cudaMemsetAsync(hist, 0, sizeof(int32_t) * (size_t) (K + 1), s);
    cudaMemsetAsync(late_count, 0, sizeof(int32_t), s);
    if (n <= 0) return;
    MidParams p{};
    p.x = x64; p.n_frames = n; p.D = D; p.wtc = wtc; p.K = K; p.BN = tc_tile_cols(K, D);
    p.w1 = tc_bound_weights(wtc, K, D);
    p.err_scale = err_scale; p.idx_out = idx;
    p.flag_list = sc.flag_list; p.flag_count = sc.counters + 0; p.flag_ub = sc.flag_ub; p.min_key = sc.min_key;
    p.cells = packed; p.hist = hist;
    p.late_list = late_list; p.late_count = late_count;
    p.list_cap = K <= 16 ? 64 : (K <= 32 ? 32 : (K <= 64 ? 16 : 8));
    p.n_raw = mid_raw_depth(D, p.BN);
    {
        static int cap = -1;
        if (cap < 0) { const char *e = getenv("ECOZ2_VQ_MID_RAW"); cap = e ? atoi(e) : 0; }
        if (cap >= 2 && cap < p.n_raw) p.n_raw = cap;
    }
    { static int ex = -1; if (ex < 0) { const char *e = getenv("ECOZ2_VQ_MID_EXP"); ex = e ? atoi(e) : 0; } p.exp = ex; }
    p.staged = (reinterpret_cast<uintptr_t>(x64) & 15) == 0 && ((size_t) kBM * D * 8) % 16 == 0;
    const long tiles = (n + 2 * kBM - 1) / (2 * kBM);
    const long grid = tiles < num_sms ? tiles : num_sms;
    const size_t smem = mid_smem_bytes(D, p.BN, p.n_raw);
    static int prof_on = -1;
    if (prof_on < 0) prof_on = getenv("ECOZ2_VQ_MID_PROF") ? 1 : 0;
    static unsigned long long *prof = nullptr;
    if (prof_on) {
        if (!prof) cudaMalloc(&prof, sizeof(unsigned long long) * 160 * 20 * 4);
        cudaMemsetAsync(prof, 0, sizeof(unsigned long long) * 160 * 20 * 4, s);
        p.prof = prof;
    }
    switch ((D + 8) / 8) {
        case 2: launch_mid_k<2, 15>(p, grid, smem, s); break;
        case 3: launch_mid_k<3, 23>(p, grid, smem, s); break;
        default: launch_mid_k<5, 39>(p, grid, smem, s); break;
    }
    if (prof_on) {
        std::vector<unsigned long long> h((size_t) 160 * 20 * 4);
        cudaStreamSynchronize(s);
        cudaMemcpy(h.data(), prof, sizeof(unsigned long long) * h.size(), cudaMemcpyDeviceToHost);
        double w[20][4] = {};
        for (long b = 0; b < grid; b++)
            for (int wp = 0; wp < 20; wp++)
                for (int k = 0; k < 4; k++) w[wp][k] += (double) h[((size_t) b * 20 + wp) * 4 + k] / grid;
        const double halves = 2.0 * ((double) tiles / grid);
        fprintf(stderr,
                "[mid prof] n=%ld K=%d raw=%d: %.0f clk per half-tile | waits per half-tile: producer ring %.0f | "
                "loader (own halves) raw %.0f a_free %.0f | mma a_full %.0f acc_empty %.0f | epilogue acc_full %.0f "
                "cell_full %.0f\n",
                n, K, p.n_raw, w[16][3] / halves, w[16][0] / halves, 2.0 * w[0][0] / halves, 2.0 * w[0][1] / halves,
                w[17][0] / halves, w[17][1] / halves, w[8][0] / halves, w[8][1] / halves);
    }
}

void launch_cell_fixup(const double *x64, const int32_t *idx, int D, const int32_t *list, const int32_t *count,
                       double *packed, int32_t *hist, int num_sms, cudaStream_t s) {
    vq_cell_fixup<<<num_sms * 2, 256, 0, s>>>(x64, idx, D, list, count, packed, hist);
}

}  // namespace vqb200
