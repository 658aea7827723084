// vq_kernels.cuh -- launch interface of the sm_100a kernels of the VQ path.
//
// Notation (reference ecoz2/src/vq): N frames, K codewords, P prediction order,
// D = P + 1.  A frame is the gain-normalised autocorrelation rx[0..P]; a
// codeword is the predictor autocorrelation raa[0..P]; the distortion is
//     rx[0]*raa[0] + 2 * sum_{n=1..P} rx[n]*raa[n]        (distortion.c:6-13)
// which the kernels evaluate as the plain inner product rx . w with
//     w = (raa[0], 2 raa[1], ..., 2 raa[P]).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace vqb200 {

// codewords per shared-memory tile of the search kernels
constexpr int kTileF32 = 128;   // FP32 search:  128 frames x 128 codewords per CTA step
constexpr int kTileF64 = 64;    // FP64 search:   64 frames x  64 codewords per CTA step
constexpr int kMaxD = 208;      // >= 1 + MAX_PREDICTION_ORDER (reference include/vq.h:13)

// Packed accumulation buffer of one k-means pass (what the NCCL allreduce sums):
//   [K*D cell sums | K counts | K cell distortions | DD | inertia | pad]
struct PackedLayout {
    int K, D;
    __host__ __device__ long cells() const { return 0; }
    __host__ __device__ long counts() const { return (long) K * D; }
    __host__ __device__ long discel() const { return (long) K * D + K; }
    __host__ __device__ long dd() const { return (long) K * D + 2L * K; }
    __host__ __device__ long inertia() const { return dd() + 1; }
    __host__ __device__ long size() const { return dd() + 2; }
};

// ---- data staging -----------------------------------------------------------
// x64 [n][D] f64 -> x32 [n][D] f32 (dense rows, round-to-nearest)
void launch_convert_frames(const double *x64, float *x32, long n, int D, cudaStream_t s);

// raas f64 [K][D] -> search operands:
//   w32 tiles [ceil(K/128)][D][128] f32, w64 tiles [ceil(K/64)][D][64] f64 (zero padded)
//   *w1max (f32) = max_i sum_n |w_i[n]|   (for the FP32 ambiguity bound)
//   w1[i] (f32) = sum_n |w_i[n]|, *w1max = max_i w1[i]
void launch_prepare_codebook(const double *raas, int K, int D, float *w32, double *w64,
                             float *w1, float *w1max, cudaStream_t s);

// ---- (a) fused distortion + argmin ------------------------------------------
// Device scratch of the FP32 search + FP64 re-evaluation of one frame block of
// capacity `cap` frames (one cudaMalloc, carved by scratch_carve).
struct SearchScratch {
    int32_t *flag_list = nullptr;          // [cap]  frames whose FP32 best-two gap is ambiguous
    float *flag_ub = nullptr;              // [cap]  upper bound of their true minimum
    unsigned long long *min_key = nullptr; // [cap]  ordered bits of their FP64 minimum
    int2 *cand = nullptr;                  // [cand_cap] (list position, codeword) candidates
    double *cand_d = nullptr;              // [cand_cap] their FP64 distortions
    int32_t *exact_list = nullptr;         // [cap]  frames sent to the all-FP64 search
    int32_t *counters = nullptr;           // [4]    flag_count, cand_count, exact_count, -
    long cap = 0, cand_cap = 0;
};
size_t scratch_bytes(long cap);
void scratch_carve(void *base, long cap, SearchScratch *out);

// FP32 search over all n frames.  idx[v] = argmin (first index on ties among the
// FP32 values); frames whose best-two gap is below
//     2 * err_scale * |rx[0]| * (*w1max)          (err_scale = kappa * 2^-24)
// are appended to sc.flag_list (sc.counters must be zeroed by the caller).
void launch_assign_f32(const float *x32, long n, int D, const float *w32, int K,
                       const SearchScratch &sc, const float *w1, const float *w1max,
                       float err_scale, int32_t *idx, int num_sms, cudaStream_t s);

// FP64 search.  list == nullptr: all n frames.  Otherwise only the frames
// list[0 .. *list_count) (device-side count; no host sync).  Exact reference
// semantics: strict '<' in index order, so the first of equal minima wins.
void launch_assign_f64(const double *x64, long n, int D, const double *w64, int K,
                       const int32_t *list, const int32_t *list_count, int32_t *idx,
                       int num_sms, cudaStream_t s);

// Re-evaluation of the flagged frames (4 launches, no host sync):
//  1. the FP32 tile search repeated over the flagged frames only SELECTS candidates
//     (d32_j - c W1_j <= ub, c = 2 err_scale |rx0|);
//  2. vq_candidates_eval: every candidate's distortion in FP64 + per-frame minimum;
//  3. vq_candidates_select: lowest codeword index among the FP64-minimal candidates;
//  4. all-FP64 search of the (normally zero) frames whose candidates overflowed.
void launch_recheck(const float *x32, const double *x64, long n, int D, const float *w32,
                    const double *w64, const float *w1, const float *w1max, const double *raas,
                    int K, float err_scale, const SearchScratch &sc, int32_t *idx, int num_sms,
                    cudaStream_t s, const float *wtc = nullptr, float err_tc = 0.f);

// ---- (a') tensor-core search (vq_tc.cu): tcgen05 split-TF32, FP32 accumulate in TMEM --------
// Supported for K >= 64 and D <= 40.  wtc = codebook in the UMMA canonical K-major layout,
// split into TF32 (hi, lo) parts (tc_codebook_floats(K, D) floats, built from raas).
// Same outputs as launch_assign_f32 (idx + flag_list of frames inside the error bound when
// `flag`); list / list_count optionally restrict the search to a gathered subset.
// dump (debug): every distortion value, [n][dump_ld] floats.
bool assign_tc_supported(int K, int D);
// step 1 of launch_recheck on the tensor cores: candidates (lower bound <= flag_ub) of the
// frames in sc.flag_list, same outputs as the FP32 filter
bool filter_tc_supported();
void launch_filter_tc(const float *x32, long n_cap, int D, const float *wtc, int K,
                      const SearchScratch &sc, float err_scale, int num_sms, cudaStream_t s);
int tc_tile_cols(int K, int D);
bool tc_operands_f16();                                  // FP16 split parts (default) or TF32 parts
const float *tc_bound_weights(const float *wtc, int K, int D);   // W1'_j behind the tiles (FP16 format)
size_t tc_codebook_floats(int K, int D);
int tc_issued_flops_per_eval(int D, int *is_f16);
void launch_prepare_codebook_tc(const double *raas, int K, int D, float *wtc, const float *w1max,
                                float err_scale, cudaStream_t s);
// cells / hist (optional, with x64 = the FP64 originals of x32's rows; needs assign_tc_fuses_cells):
// the search also adds the FP64 row of every frame OUTSIDE the error bound to packed cell sums and
// counts (FP64 reductions issued by otherwise idle warps while the tensor core works); the flagged
// frames follow through launch_cell_fixup after launch_recheck, then launch_cell_finish.
bool assign_tc_fuses_cells(int K, int D);
// R codebooks of K codewords in ONE sweep (wtc = launch_prepare_codebook_tc of the R * K concatenated
// raas rows, total_cols = R * K padded to a multiple of 256 with copies of real codewords):
// idx_out[model][idx_stride] model-local argmins; (model << 40 | frame) keys of the pairs inside the
// error bound in seg_key (count in sc.counters[0], may exceed seg_cap: then keys were dropped).
bool assign_tc_models_supported(int K, int D, long total_cols);
void launch_assign_tc_models(const float *x32, long n, int D, const float *wtc, int K, long total_cols,
                             const SearchScratch &sc, float err_scale, int32_t *idx_out, long idx_stride,
                             unsigned long long *seg_key, int seg_cap, int num_sms, cudaStream_t s,
                             const void *xop = nullptr);
// The frames as ready-made operands of the tensor-core search (FP16 format; built once per frame set
// from the FP32 copy with the operations of the search's own loaders): tc_frames_bytes(n, D) bytes
// (0 = not available); pass the buffer as `xop` to launch_assign_tc / launch_assign_tc_models when
// the search runs over ALL n frames in order (no gather list).
size_t tc_frames_bytes(long n, int D);
void launch_prepare_frames_tc(const float *x32, long n, int D, float err_scale, void *xop, cudaStream_t s);
// all-FP64 argmin (strict <, first index wins: quantize.c:14-30) of every (model, frame) key over the
// model's K codewords; raas_all [R * K][D].  One warp per key.
void launch_models_recheck(const double *x64, int D, const double *raas_all, int K, const unsigned long long *keys,
                           const int32_t *count, int cap, int32_t *idx_out, long idx_stride, int num_sms,
                           cudaStream_t s);
// FP64 distortion - 1 of every frame's winner in every model, summed per chunk of frames:
// chunks[c] = {first frame, frames} (never across a segment boundary), partial[c][R].  Deterministic.
void launch_models_distortion(const double *x64, int D, const double *raas_all, int K, int R, const int32_t *idx_all,
                              long idx_stride, const long *chunks, int num_chunks, double *partial, int num_sms,
                              cudaStream_t s);
void launch_assign_tc(const float *x32, long n, int D, const float *wtc, int K,
                      const int32_t *list, const int32_t *list_count, const SearchScratch &sc,
                      bool flag, const float *w1, const float *w1max, float err_scale, int32_t *idx,
                      float *dump, int dump_ld, int num_sms, cudaStream_t s, const double *x64 = nullptr,
                      double *cells = nullptr, int32_t *hist = nullptr, const void *xop = nullptr);

// ---- (b) cell accumulation ----------------------------------------------------
// For every frame: dd = FP64 distortion against its codeword raas[idx];
//   mindist[v] = dd - 1 (if mindist != nullptr); seq[v] = (u16) idx (if seq != nullptr)
//   with_cells: packed.cells[idx] += rx, counts[idx] += 1, discel[idx] += dd - 1
//   packed.dd += dd - 1 ; packed.inertia += dd * dd
//   neg_list / neg_count (optional): frames whose winner distortion is NEGATIVE are appended there
//   instead of adding dd * dd (inertia.c:21-25 wants min_i dd_i^2: launch_inertia_neg scans them)
void launch_accumulate(const double *x64, const int32_t *idx, long n, int D,
                       const double *raas, int K, double *packed, double *mindist,
                       uint16_t *seq, bool with_cells, int num_sms, cudaStream_t s,
                       int32_t *neg_list = nullptr, int32_t *neg_count = nullptr);
// packed.inertia += min_i distortion(x_v, raas_i)^2 for the listed frames (exact FP64 scan)
void launch_inertia_neg(const double *x64, int D, const double *raas, int K, const int32_t *neg_list,
                        const int32_t *neg_count, double *packed, int num_sms, cudaStream_t s);
// run-time audit of the search (ECOZ2_VQ_AUDIT=n): all-FP64 scan of every n-th frame against idx;
// out[3] += {audited, index differs, chosen codeword worse than the minimum by > 1e-6 relative}
void launch_audit(const double *x64, long n, int D, const double *raas, int K, const int32_t *idx,
                  int every, unsigned long long *out, int num_sms, cudaStream_t s);

// ---- (b') cell accumulation by counting sort + gather (vq_cells.cu) ---------------------
// Same packed outputs as launch_accumulate(with_cells = true) -- cell sums, counts, cell
// distortions, DD -- without per-frame atomics and without per-frame distortions (the cell
// distortion is cellsum . w - count).  mindist is NOT produced: call launch_accumulate's
// distortion-only form when per-frame values are wanted.  2 <= K <= 8192.
struct CellScratch {
    int32_t *perm = nullptr;        // [cap]  frame indices grouped by cell
    int32_t *hist = nullptr;        // [K+1]  frames per cell
    int32_t *offsets = nullptr;     // [K+1]  first position of each cell in perm
    int32_t *cursor = nullptr;      // [K+1]  scatter cursors
    int32_t *item_start = nullptr;  // [K+1]  first 256-frame work item of each cell
    long cap = 0;
};
bool cell_sort_supported(int K);
size_t cell_scratch_bytes(long cap, int max_K);
void cell_scratch_carve(void *base, long cap, int max_K, CellScratch *out);
// A padded copy of the frames for the gather of the sorted path: rows of cell_pad_ld(D) doubles (whole
// 64-byte atoms; 0 = not worth it for this D), made once per frame set by launch_pad_rows.
int cell_pad_ld(int D);
void launch_pad_rows(const double *x64, double *xp, long n, int D, int ld, int num_sms, cudaStream_t s);
void launch_cell_sums(const double *x64, const int32_t *idx, long n, int D, const double *raas,
                      int K, double *packed, uint16_t *seq, const CellScratch &cs, int num_sms,
                      cudaStream_t s, const double *x64_padded = nullptr, int ld = 0);

// ---- (a+b) fused pass for codebooks of ONE tensor-core tile (vq_mid.cu), 8 < K <= 128 ------------
// One sweep over the FP64 frames: tensor-core search (same operands, same error bound and flag
// protocol as launch_assign_tc) + cell sums of every frame OUTSIDE the bound into packed (cell sums)
// and hist (counts).  The flagged frames (sc.flag_list) are re-evaluated by launch_recheck as usual;
// launch_cell_fixup then adds their rows, launch_cell_finish derives counts / distortions / DD.
// zero_hist = false accumulates into hist across calls (chunked host pass).
bool pass_mid_supported(int K, int D);
// late_list [n] / late_count [1]: rows the sweep could not take (a cell's per-tile list was full).
void launch_pass_mid(const double *x64, long n, int D, const float *wtc, int K, const SearchScratch &sc,
                     float err_scale, int32_t *idx, double *packed, int32_t *hist, int32_t *late_list,
                     int32_t *late_count, bool zero_hist, int num_sms, cudaStream_t s);
// packed cells[idx[row]] += x64[row], hist[idx[row]] += 1 for the rows list[0 .. *count)
void launch_cell_fixup(const double *x64, const int32_t *idx, int D, const int32_t *list, const int32_t *count,
                       double *packed, int32_t *hist, int num_sms, cudaStream_t s);
// counts, per-cell distortion (cellsum . w - count) and DD from the cell sums and hist (vq_cells.cu)
void launch_cell_finish(const double *raas, const int32_t *hist, int K, int D, double *packed, cudaStream_t s);

// ---- (a+b) fused small-codebook pass (K <= 32, D = 37 or 13), all FP64 ----------
// One sweep over the FP64 frames: exact argmin (first index on ties), idx / mindist / seq
// outputs, and (with_cells) the packed cell sums, counts, distortions; DD and inertia always.
bool pass_small_supported(int K, int D);
// the thread-per-frame version (vq_pass_small2) takes this codebook: faster than search + cell sums,
// so the tensor-core mode uses it too (ECOZ2_VQ_SMALL2=0: first version, FP32/FP64 modes only)
bool pass_small_fast(int K, int D);
void launch_pass_small(const double *x64, long n, int D, const double *raas, int K, double *packed,
                       int32_t *idx, double *mindist, uint16_t *seq, bool with_cells, int num_sms,
                       cudaStream_t s);

// ---- (c) centroid update / split ----------------------------------------------
// calculate_reflections (vq_learn.i:125-155): for every cell with count > 0,
// Levinson-Durbin on the raw cell sum -> refl, raas; cell sum /= errPred.
void launch_chunk_sums(const double *md, const long *chunks, int num_chunks, int R, int m, double *partial,
                       cudaStream_t s);
void launch_models_segments(const double *partial, const int *seg_first_chunk, int num_segments, int R, double *sums,
                            cudaStream_t s);
// idx (model-local, < 65536) -> u16 symbols
void launch_idx_to_u16(const int32_t *idx, uint16_t *seq, size_t n, cudaStream_t s);
void launch_centroids(double *packed, int K, int D, double *refl, double *raas,
                      cudaStream_t s);
// grow_codebook (vq_learn.i:47-72): refl K -> 2K rows (x 0.99f / x 1.01f), then
// reflections_to_raas over all 2K rows.
void launch_grow(double *refl, double *raas, int K, int D, cudaStream_t s);
// reflections_to_raas (ref2raas.c:8-26)
void launch_refl_to_raas(const double *refl, double *raas, int K, int D, cudaStream_t s);

// ---- stage evaluation -----------------------------------------------------------
// *out += sum_{i != j} [ distortion(cells_j, raas_i) - 1 ]      (sigma.c:14-25)
void launch_sigma_sum(const double *raas, const double *cells, int K, int D, double *out,
                      cudaStream_t s);

// ---- stage report on the device (vq_report.cu): "<cbook>.min_dists.csv", report.c:109-123 -------
// Lines "%d,%f\n" (codeword, minDistortion) of n frames as text, byte for byte what fprintf prints.
// launch_format_lengths: line lengths + offsets into `scratch` (format_scratch_bytes(n)); then
// *total (bytes of text) and *bad (1: a value was inf / nan / >= 2^63 -- format on the host instead)
// can be read back (format_result_ptrs); launch_format_write writes the text (capacity >= *total).
size_t format_scratch_bytes(long n);
size_t format_text_capacity(long n);
void launch_format_lengths(const int32_t *cw, const double *md, long n, void *scratch, cudaStream_t s);
void format_result_ptrs(void *scratch, long n, const unsigned long long **total, const int **bad);
void launch_format_write(const int32_t *cw, const double *md, long n, const void *scratch, char *text,
                         cudaStream_t s);

// FP32 FFMA probe: blocks x 256 threads x iters x 128 FMAs each
void launch_ffma_probe(float *out, int blocks, int iters, cudaStream_t s);

}  // namespace vqb200
