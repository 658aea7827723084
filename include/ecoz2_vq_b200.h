/* ecoz2_vq_b200.h -- C ABI of the B200-native ECOZ2 vector-quantization path.
 *
 * Shared library: ecoz2rs_b200/csrc/libecoz2_vq_b200.so (built by
 * ecoz2rs_b200/csrc/Makefile, CUDA sm_100a).  No torch / C++ types cross this
 * boundary: plain pointers, sizes and opaque handles only.
 *
 * PART 1 is the drop-in boundary: the three `ecoz2_vq_*` entry points of the
 * reference's public C API with identical signatures and file/callback
 * behaviour, i.e. exactly what the Rust crate's `extern "C"` block
 * (reference src/ecoz2_lib/mod.rs:96-122) and the C mains
 * (reference ecoz2/src/x/vq.learn.c:107-128, vq.quantize.c:41) bind.
 *
 * PART 2 is the in-memory API the host mirror (ecoz2rs_b200/vq.py), the parity
 * tests and bench.py call: the same algorithm steps with the frames / codebooks
 * passed as arrays instead of `.prd` / `.cbook` files.
 *
 * There is NO CPU fallback: every compute entry point needs a CUDA device and
 * returns ECOZ2B200_ERR_CUDA (with a message on stderr) when there is none.
 */
#ifndef ECOZ2_VQ_B200_H
#define ECOZ2_VQ_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

/* ===================================================================== PART 1
 * Drop-in replacements (reference ecoz2/src/include/ecoz2.h).               */

/* reference ecoz2/src/include/ecoz2.h:56-59.  Invoked synchronously on the
 * calling thread once per finished codebook size, after the `.cbook` is saved
 * and before the stage report is written (ecoz2/src/vq/vq_learn_par.c:133-143). */
typedef void (*vq_learn_callback_t)(void *callback_target, int M,
                                    double avgDistortion, double sigma,
                                    double inertia);

/* reference ecoz2/src/include/ecoz2.h:62-70 -> ecoz2/src/ecoz2/ecoz2.c:109-129
 * -> ecoz2/src/vq/vq_learn.c:7-32.  LBG from 2 codewords up to
 * ECOZ2_VQ_MAX_M (env, default 4096 = the reference's MAX_CODEBOOK_SIZE).
 * Writes data/codebooks/<class>/eps_<e>_M_<KKKK>.cbook (+ .rpt, .rpt.csv,
 * .cards_dists.csv, .min_dists.csv) relative to the CWD.
 * Returns 0 ok, 2 predictor load failure, 1 memory/codebook, 3 CUDA failure. */
int ecoz2_vq_learn(int prediction_order, double epsilon,
                   const char *codebook_class_name,
                   const char *predictor_filenames[], int num_predictors,
                   void *callback_target, vq_learn_callback_t callback);

/* reference ecoz2/src/include/ecoz2.h:73-80 -> ecoz2/src/vq/vq_learn.c:34-77.
 * Resume: loads reflections from `base_codebook`, starts at twice its size. */
int ecoz2_vq_learn_using_base_codebook(char *base_codebook, double epsilon,
                                       const char *predictor_filenames[],
                                       int num_predictors,
                                       void *callback_target,
                                       vq_learn_callback_t callback);

/* reference ecoz2/src/include/ecoz2.h:82-87 -> ecoz2/src/vq/vq_quantize.c:13-111.
 * One `.seq` per `.prd` under data/sequences[/TRAIN|/TEST]/M<K>/<class>/. */
int ecoz2_vq_quantize(const char *nom_raas, const char *predictor_filenames[],
                      int num_predictors, int show_filenames);

/* `ecoz2 vq classify`: replaces ecoz2_vq_classify, reference
 * ecoz2/src/include/ecoz2.h:95-101 -> vq_classify, ecoz2/src/vq/vq_classify.c:185-327.
 * Every predictor is quantized against every codebook (all of one size); stdout gets the
 * reference's model list, optional ranking of misclassified files, confusion matrix and
 * candidate-order table; stderr one green '*' / red '!' per file.  Returns 0, or 1 on a
 * codebook load / conformity error.  All frames stay resident in HBM: one search pass per
 * model over all files.                                                                  */
int ecoz2_vq_classify(char **cb_filenames, int num_codebooks, char **prd_filenames,
                      int num_predictors, int show_ranked);

/* The reference's INNER C API (ecoz2/src/include/vq.h:82-117) with identical signatures: what its
 * public forwards (ecoz2/src/ecoz2/ecoz2.c:109-193) and its C mains (ecoz2/src/x/vq.learn.c:107-128,
 * x/vq.quantize.c:41, x/vq.classify.c) call.  With these exported, ecoz2.c and the mains link against
 * this library unmodified (integration/build.rs.patch; oracle/Makefile target `dropin` builds the
 * reference's vq.learn / vq.quantize mains that way).  `use_par` (serial / OpenMP loop of the
 * reference) is accepted and ignored: both are the CUDA path.                                  */
int vq_learn(int prediction_order, double epsilon, const char *codebook_class_name,
             const char *predictor_filenames[], int num_predictors, int use_par,
             void *callback_target, vq_learn_callback_t callback);
int vq_learn_using_base_codebook(char *base_codebook, double epsilon,
                                 const char *predictor_filenames[], int num_predictors, int use_par,
                                 void *callback_target, vq_learn_callback_t callback);
int vq_quantize(const char *nom_raas, const char *predictor_filenames[], int num_predictors,
                int show_filenames);
int vq_classify(char **cb_filenames, int num_codebooks, char **prd_filenames, int num_predictors,
                int show_ranked);
/* quantize (ecoz2/src/include/vq.h:74-78, ecoz2/src/vq/quantize.c:8-34): the in-memory quantizer that the
 * reference's cbook_quantize (vq/vq.c:154-157) and through it its SeqProvider (vq/vq.c:285-315, used by
 * hmm_classify.c:396) call.  `prd` is the reference's Predictor (include/lpc.h:29-34: className[96],
 * sample_t **vectors, int T, int P), declared here as ecoz2_ref_predictor with the same layout; seq
 * receives T symbols (may be NULL); returns the average distortion sum(ddmin - 1) / T.  With this
 * exported, quantize.c leaves the reference's build and vq.c links unmodified.  The engine behind it
 * is created on first use and kept for the process (devices: ECOZ2_VQ_GPUS); the codebook is uploaded
 * again only when raas changes.  On a CUDA failure the process exits with status 1, as the
 * reference does on its own fatal errors (the signature has no error channel).                  */
typedef struct {
    char className[96];
    double **vectors;
    int T;
    int P;
} ecoz2_ref_predictor;
double quantize(double *raas, int num_raas, ecoz2_ref_predictor *prd, uint16_t *seq);

/* ===================================================================== PART 2
 * In-memory API.                                                            */

#define ECOZ2B200_OK 0
#define ECOZ2B200_ERR_ARG 1
#define ECOZ2B200_ERR_IO 2
#define ECOZ2B200_ERR_CUDA 3
#define ECOZ2B200_ERR_NCCL 4

/* Arithmetic mode of the nearest-codeword search (reference
 * ecoz2/src/vq/vq_learn_par.c:62-81, quantize.c:14-30 are FP64).
 *   TENSOR: tcgen05 tensor-core search with two-way split operands (22-bit
 *          products, FP32 accumulate in tensor memory) + FP64 re-evaluation of
 *          every frame whose best-two gap is inside the error bound (default;
 *          index-exact vs FP64).  The parts are FP16 values of power-of-two
 *          scaled operands (ECOZ2_VQ_TC_FMT=f16, default, P <= 46) or TF32 values
 *          (ECOZ2_VQ_TC_FMT=tf32, P <= 38); wider frames take the FP32 path.
 *   FP32:  FP32 FFMA search + the same FP64 re-evaluation.
 *   FP64:  the whole search in FP64 (verification mode).                      */
#define ECOZ2B200_MODE_FP32 0
#define ECOZ2B200_MODE_FP64 1
#define ECOZ2B200_MODE_TENSOR 2
#define ECOZ2B200_MODE_TF32X3 ECOZ2B200_MODE_TENSOR /* first name of the mode (round 1), kept as an alias */

typedef struct ecoz2b200_engine ecoz2b200_engine;

const char *ecoz2b200_version(void);
/* last error message of the calling thread ("" if none) */
const char *ecoz2b200_last_error(void);
/* number of visible CUDA devices (0 if none / no driver) */
int ecoz2b200_device_count(void);

/* One engine = the frames of ONE process, sharded over `num_devices` local
 * GPUs (device ordinals in `devices`), plus the replicated codebook state.
 * `max_M` bounds the codebook size (<= 65536, Symbol is u16:
 * reference ecoz2/src/include/symbol.h:7).                                   */
ecoz2b200_engine *ecoz2b200_create(const int *devices, int num_devices, int P,
                                   int max_M);
void ecoz2b200_destroy(ecoz2b200_engine *e);

int ecoz2b200_set_mode(ecoz2b200_engine *e, int mode);
/* kappa of the FP32 ambiguity bound: a frame is re-evaluated in FP64 when
 * d2 - d1 < 2 * kappa * 2^-24 * rx[0] * max_i sum_n |w_i[n]|.                 */
int ecoz2b200_set_recheck_kappa(ecoz2b200_engine *e, double kappa);
/* same for the TF32X3 tensor-core search (its own error constant)             */
int ecoz2b200_set_recheck_kappa_tc(ecoz2b200_engine *e, double kappa);
/* run-time audit of the search (same as ECOZ2_VQ_AUDIT=n at creation): after every search, every
 * `every`-th frame also gets the all-FP64 scan of vq_learn_par.c:62-81 and the device counts audited
 * frames, differing indices and differences that are not FP64-level ties (ecoz2b200_counters);
 * 0 switches it off.                                                                              */
int ecoz2b200_set_audit(ecoz2b200_engine *e, int every);
/* test aid: the raw tensor-core distortion values of the first n_rows resident
 * frames of shard 0 against the current codebook, out[n_rows][K] (TF32X3 mode) */
int ecoz2b200_debug_tc_distortions(ecoz2b200_engine *e, float *out, long n_rows);
/* The lines "<codeword>,<minDistortion>\n" of the LAST pass's assignment of all local frames, in
 * append order, formatted on the device: byte for byte the "%d,%f\n" lines of the reference's
 * "<cbook>.min_dists.csv" (ecoz2/src/vq/report.c:118-121).  *text points into a pinned host buffer
 * owned by the engine, valid until the call after the next.  ECOZ2B200_ERR_ARG when a value is
 * inf / nan / >= 2^62 (the file writer then formats on the host).                             */
int ecoz2b200_min_dists_text(ecoz2b200_engine *e, const char **text, size_t *bytes);
/* test aid: the device formatter alone on host arrays (device 0): n lines into out[capacity];
 * *bytes = size of the text.  ECOZ2B200_ERR_ARG if a value cannot be formatted or it does not fit. */
int ecoz2b200_debug_format_lines(const int32_t *codeword, const double *min_dist, long n, char *out,
                                 size_t capacity, size_t *bytes);
/* The same device formatter for callers that must NOT allocate (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
 * Calls no cudaMalloc / cudaFree or any allocating
 * routine: the kernels' scratch (per-block sums, the "cannot format" flag) is a fixed __device__
 * array of the library, 128.3 KB held by every GPU that loads it until the process exits, which
 * limits one call to ECOZ2B200_NOALLOC_MAX_LINES lines (more -> ECOZ2B200_ERR_ARG and a message in
 * ecoz2b200_last_error).  ALL pointers are device-accessible memory of the CURRENT device, owned
 * by the caller: d_codeword[n], d_min_dist[n] in; d_text[text_capacity] out, text_capacity >=
 * 40 * n (else ERR_ARG); *d_bytes out = size of the text, or ~0 when a value is inf / nan /
 * >= 2^62 (d_text is then incomplete).  `stream` is a cudaStream_t or NULL; the work is queued
 * on it, nothing is synchronised.  Every call on a device shares the one scratch array: calls
 * whose kernels can overlap on the GPU (different streams without an ordering) race on it.     */
#define ECOZ2B200_NOALLOC_MAX_LINES 4194304L
int ecoz2b200_format_lines_noalloc(const int32_t *d_codeword, const double *d_min_dist, long n,
                                   char *d_text, size_t text_capacity,
                                   unsigned long long *d_bytes, void *stream);
/* measurement aid: tensor-core FLOPs the TF32X3 search issues per (frame,
 * codeword) pair at prediction order P (padding and all three product terms
 * included); 0 if that order takes the FP32 path.  *is_f16 (optional) = 1 when
 * the operands are FP16 parts, 0 for TF32 parts.                              */
int ecoz2b200_tc_issued_flops_per_eval(int P, int *is_f16);
/* test aid (host only, no device needed): the "<cbook>.min_dists.csv" writer of the learn report
 * (reference report.c:109-123: N lines "%d,%f"), formatted by all host threads with an exact
 * integer "%f".  Returns 0, or ECOZ2B200_ERR_IO.                                          */
int ecoz2b200_write_min_dists_csv(const char *path, const int32_t *codeword, const double *min_dist,
                                  long n);

/* Multi-process sharding (one process per GPU, e.g. under torchrun): all ranks
 * call comm_init with the same 128-byte id obtained once from
 * ecoz2b200_nccl_unique_id().  Afterwards every k-means pass ends with ONE
 * ncclAllReduce(sum, f64) of the packed [K*(P+1) cell sums | K counts |
 * K cell distortions | DD] buffer.  Engines with several local devices and no
 * comm_init use ncclCommInitAll internally.                                   */
int ecoz2b200_nccl_unique_id(unsigned char id_out[128]);
/* The sharding rule (host only, no device needed): rank `rank` of `world_size` owns the frames
 * [*lo, *hi) of `total` in append order -- contiguous blocks of ceil(total / world_size), the
 * reference's thread partition vq_learn_par.c:62 made contiguous.  The same function splits an
 * engine's frames over its local devices.                                                  */
int ecoz2b200_shard_range(long total, int rank, int world_size, long *lo, long *hi);
/* Offsets (in doubles) of the packed per-pass buffer the allreduce sums:
 * offsets = {cell sums [K][P+1], counts [K], cell distortions [K], DD, inertia, total size} */
int ecoz2b200_packed_layout(int K, int P, long offsets[6]);
int ecoz2b200_comm_init(ecoz2b200_engine *e, const unsigned char id[128],
                        int rank, int world_size);

/* ---- frames (the Predictor vectors, reference ecoz2/src/include/lpc.h:29-34:
 * N rows of D = P+1 gain-normalised autocorrelations, FP64).
 * Frames are appended in call order and split into contiguous blocks over the
 * engine's devices when `frames_commit` is called.                            */
int ecoz2b200_frames_clear(ecoz2b200_engine *e);
/* Announce the total number of frames that will be appended: fixes the shard
 * boundaries so every append is copied straight to its GPU.  Required before
 * appending when the engine has more than one device; optional otherwise.    */
int ecoz2b200_frames_reserve(ecoz2b200_engine *e, long total);
/* host buffer, row-major [n][P+1] f64 (copied) */
int ecoz2b200_frames_append_host(ecoz2b200_engine *e, const double *x, long n);
/* device buffer on devices[0], row-major [n][P+1] f64 (copied device->device) */
int ecoz2b200_frames_append_device(ecoz2b200_engine *e, const double *x_dev,
                                   long n);
int ecoz2b200_frames_commit(ecoz2b200_engine *e);
long ecoz2b200_frames_count(const ecoz2b200_engine *e);

/* ---- codebook state (replicated on every device) -------------------------- */
/* reference ecoz2/src/vq/vq_learn.i:31-44 : 2 codewords, k1 = -/+0.5 */
int ecoz2b200_codebook_initial(ecoz2b200_engine *e);
/* set from reflections [K][P+1] (reference cb_load + reflections_to_raas,
 * ecoz2/src/vq/vq_quantize.c:25-40, ecoz2/src/lpc/ref2raas.c:8-26) */
int ecoz2b200_codebook_set_reflections(ecoz2b200_engine *e, const double *refl,
                                       int K);
/* set raas directly [K][P+1] (teacher forcing in tests; reflections zeroed) */
int ecoz2b200_codebook_set_raas(ecoz2b200_engine *e, const double *raas, int K);
/* reference grow_codebook, ecoz2/src/vq/vq_learn.i:47-72 : K -> 2K */
int ecoz2b200_codebook_grow(ecoz2b200_engine *e);
int ecoz2b200_codebook_size(const ecoz2b200_engine *e);
/* copy out [K][P+1] f64; either pointer may be NULL */
int ecoz2b200_codebook_get(ecoz2b200_engine *e, double *refl_out,
                           double *raas_out);

/* ---- steps ------------------------------------------------------------------ */
/* Nearest codeword for every frame (strict '<', first index wins).
 * Outputs are host buffers over ALL local frames in append order; any may be
 * NULL.  idx: int32 [N]; ddmin: f64 [N] raw distortion of the winner (the
 * reference's ddmin, before the -1).  *num_rechecked: frames re-evaluated in
 * FP64.  reference ecoz2/src/vq/vq_learn_par.c:62-81                           */
int ecoz2b200_assign(ecoz2b200_engine *e, int32_t *idx_out, double *ddmin_out,
                     long *num_rechecked);

/* One k-means pass: assign + cell accumulation (+ allreduce) + centroid update.
 * reference ecoz2/src/vq/vq_learn_par.c:50-125, vq_learn.i:84-155.
 * Host outputs (any may be NULL): cells [K][P+1] raw sums (before the errPred
 * normalisation), cardd int32 [K], discel f64 [K], *DD.
 * After the call the engine's codebook is the post-update one and its cell
 * buffer holds the errPred-normalised sums (for sigma).                        */
int ecoz2b200_kmeans_pass(ecoz2b200_engine *e, double *cells_out,
                          int32_t *cardd_out, double *discel_out,
                          double *DD_out);
/* The same pass over n frames x[n][P+1] that are still in HOST memory (pinned memory for
 * full speed): they replace the resident frame set and are uploaded in chunks on a copy
 * stream while the search of the previous chunk runs, so the pass costs the PCIe transfer
 * and little else.  Equivalent to frames_clear + frames_reserve + frames_append_host +
 * frames_commit + kmeans_pass (same results; the frames stay resident afterwards).  With
 * comm_init every rank passes ITS frames and gets the global sums.                     */
int ecoz2b200_kmeans_pass_host(ecoz2b200_engine *e, const double *x, long n, double *cells_out,
                               int32_t *cardd_out, double *discel_out, double *DD_out);

/* reference calculateSigma, ecoz2/src/vq/sigma.c:12-29, on the engine's current
 * codebook and normalised cells (i.e. right after kmeans_pass). */
int ecoz2b200_sigma(ecoz2b200_engine *e, double avgDistortion, double *sigma_out);
/* reference calculateInertia, ecoz2/src/vq/inertia.c:10-31 (global over ranks) */
int ecoz2b200_inertia(ecoz2b200_engine *e, double *inertia_out);

/* Per-frame assignment of the LAST pass (local frames): idx int32 [N] and
 * minDist f64 [N] (= ddmin - 1, reference CodewordAndMinDist,
 * ecoz2/src/include/vq.h:60-63).                                               */
int ecoz2b200_last_assignment(ecoz2b200_engine *e, int32_t *idx_out,
                              double *mindist_out);

/* ---- whole LBG run (reference learn_par, ecoz2/src/vq/vq_learn_par.c:3-167) -- */
typedef struct {
    int M;                 /* codebook size of the finished stage */
    int passes;            /* pass + 1 */
    double avgDistortion;  /* DD / N of the last pass (pre-update codebook) */
    double sigma;
    double inertia;
    const double *reflections; /* [M][P+1] post-update (what cb_save writes) */
    const double *raas;        /* [M][P+1] post-update */
    const int32_t *cardd;      /* [M] */
    const double *discel;      /* [M] */
    double stage_seconds;      /* wall time of this stage (host clock) */
    double assign_ms;          /* device time in the search kernels, this stage */
    double evals;              /* frame x codeword evaluations, this stage (global) */
} ecoz2b200_stage_info;

typedef void (*ecoz2b200_stage_cb)(void *ctx, const ecoz2b200_stage_info *info);
typedef void (*ecoz2b200_pass_cb)(void *ctx, int M, int pass, double DD,
                                  double pass_seconds, int empty_cells);

/* Starts from the engine's current codebook (initial / set_reflections+grow),
 * runs until the codebook size reaches `stop_M` (the reference's
 * MAX_CODEBOOK_SIZE, ecoz2/src/vq/vq_learn_par.c:147).  Both callbacks run on
 * the calling thread.                                                          */
int ecoz2b200_learn(ecoz2b200_engine *e, double epsilon, int stop_M,
                    ecoz2b200_stage_cb on_stage, ecoz2b200_pass_cb on_pass,
                    void *ctx);

/* ---- quantize (reference quantize, ecoz2/src/vq/quantize.c:8-34) ------------ */
/* Host in, host out: frames [T][P+1] f64 against the engine's current codebook;
 * seq u16 [T] (may be NULL), *avg_distortion = sum(ddmin - 1) / T.
 * Includes the host->device and device->host copies (the e2e path); uses the
 * engine's devices, T split in contiguous blocks.  Does not disturb the
 * engine's resident training frames.                                           */
int ecoz2b200_quantize_host(ecoz2b200_engine *e, const double *x, long T,
                            uint16_t *seq_out, double *avg_distortion);
/* Same on the engine's resident frames (inputs already in HBM); seq_out is a
 * host buffer of frames_count() entries or NULL.                               */
int ecoz2b200_quantize_resident(ecoz2b200_engine *e, uint16_t *seq_out,
                                double *avg_distortion);

/* cbook_quantize (ecoz2/src/vq/vq.c:154-157) of ALL resident frames against R codebooks of K codewords
 * at once: the work a SeqProvider does per predictor file for `hmm classify` (vq/vq.c:285-315:
 * next_seq.sequences[r], next_seq.dists[r]) and vq_classify.c:265-280 per (model, file), here for every
 * file in one call.  reflections [R][K][1+P] (what the R `.cbook` files hold); segment f = frames
 * [seg_offsets[f], seg_offsets[f+1]) of the append order = predictor file f (seg_offsets[0] = 0,
 * seg_offsets[num_segments] = frames_count()).  seq_out [R][frames_count()] symbols (or NULL),
 * dists_out [R][num_segments] = the `dist` cbook_quantize returns for (model r, file f) (or NULL).
 * In tensor mode with K = 32, 64, 128 or 256 all R searches are ONE sweep over the
 * frames (the codebooks concatenated, the argmin restarted at every codebook; *one_sweep = 1) and
 * the engine's own codebook is untouched; otherwise R sweeps, and the engine's codebook ends as
 * model R-1.  Not available with one process per GPU (comm_init).                              */
int ecoz2b200_quantize_models(ecoz2b200_engine *e, const double *reflections, int R, int K,
                              const long *seg_offsets, int num_segments, uint16_t *seq_out,
                              double *dists_out, int *one_sweep);

/* ---- timing / counters (CUDA events on the engine's own streams) ------------ */
typedef struct {
    double assign_ms;      /* search kernels (incl. FP64 re-evaluation) */
    double accumulate_ms;  /* cell accumulation */
    double update_ms;      /* centroid / split / codebook staging */
    double allreduce_ms;   /* NCCL */
    long kernel_launches;  /* kernels launched by this library since reset */
    double evals;          /* frame x codeword evaluations since reset (local) */
    long rechecked;        /* frames re-evaluated in FP64 since reset */
    double search_main_ms; /* the dominant kernel alone: the full search over all frames
                              (FP32 pass in FP32 mode), CUDA events on its stream */
    long search_main_launches;
    /* run-time audit (environment ECOZ2_VQ_AUDIT=n): after every search every n-th frame is searched
     * again over ALL codewords in FP64 on the device and compared with the index the tensor-core /
     * FP32 search + FP64 re-evaluation produced                                                 */
    long audited;          /* frames audited since reset */
    long audit_idx_diff;   /* of those: index differs from the all-FP64 scan (includes FP64-level ties) */
    long audit_mismatches; /* of those: the chosen codeword is worse than the FP64 minimum by more than
                              1e-6 relative, i.e. outside north_star's exemption: must stay 0 */
} ecoz2b200_counters;
int ecoz2b200_counters_reset(ecoz2b200_engine *e);
int ecoz2b200_counters_get(ecoz2b200_engine *e, ecoz2b200_counters *out);

/* Measured FP32 FFMA throughput of `device` in TFLOP/s (a pure FMA-chain kernel
 * filling every SM for ~`millis` ms): the denominator of the FP32 roofline.    */
double ecoz2b200_ffma_peak_tflops(int device, double millis);

/* ---- file formats (byte-compatible with the reference) ---------------------- */
/* `.prd` header probe, reference ecoz2/src/lpc/prd.c:34-74 */
int ecoz2b200_prd_info(const char *path, int *P, int *T, char class_name[96]);
/* reads the T x (P+1) f64 payload into `out` */
int ecoz2b200_prd_read(const char *path, double *out, long capacity_values);
/* reference cb_save / cb_load, ecoz2/src/vq/vq.c:22-116 */
int ecoz2b200_cbook_save(const char *path, const char *class_name,
                         const double *reflections, int M, int P);
int ecoz2b200_cbook_info(const char *path, int *P, int *M, char class_name[96]);
int ecoz2b200_cbook_read(const char *path, double *refl_out,
                         long capacity_values);
/* reference seq_save, ecoz2/src/hmm/symbol.c:17-36 */
int ecoz2b200_seq_save(const char *path, const char *class_name,
                       const uint16_t *seq, int T, int M);

/* ---- path helpers (host only; what decides where `.cbook` / `.seq` files land) ------------- */
/* class of a file = name of its parent directory, "" without one; reference get_class_name,
 * ecoz2/src/utl/fileutil.c:14-30 (truncated to 95 characters like the 96-byte header field) */
int ecoz2b200_class_name_of(const char *path, char class_name[96]);
/* `data/<base_dir>/<class or "_">/<stem><ext>`, directories created under the current working
 * directory; reference get_output_filename + camext, ecoz2/src/utl/fileutil.c:32-63, 97-119.
 * Returns ECOZ2B200_ERR_ARG when `out` (capacity bytes) is too small. */
int ecoz2b200_output_filename(const char *from, const char *base_dir, const char *ext,
                              char *out, long capacity);

#ifdef __cplusplus
}
#endif
#endif /* ECOZ2_VQ_B200_H */
