// vq_engine.h -- host side of the B200 VQ path: device shards, the k-means pass,
// the LBG driver.  Internal C++ interface behind include/ecoz2_vq_b200.h.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#include <condition_variable>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/ecoz2_vq_b200.h"
#include "vq_kernels.cuh"

namespace vqb200 {

void set_last_error(const std::string &msg);

// Every extern "C" entry point leaves the calling thread's current CUDA context as it found it
// (the engine switches devices with cudaSetDevice).  Saved and restored through the driver's
// cuCtxGetCurrent / cuCtxSetCurrent (looked up with cudaGetDriverEntryPoint, so libcuda is not a
// link dependency): unlike cudaSetDevice(previous) this does not create a context on a device the
// caller had only selected implicitly.
class DeviceGuard {
  public:
    DeviceGuard();
    ~DeviceGuard();
    DeviceGuard(const DeviceGuard &) = delete;
    DeviceGuard &operator=(const DeviceGuard &) = delete;

  private:
    void *ctx_ = nullptr;
    bool have_ = false;
};
const char *get_last_error();

struct NcclApi;  // dlopen'ed libnccl (vq_nccl.cpp)

// The sharding rule of the path: `total` frames in append order are split into
// `parts` contiguous blocks of ceil(total / parts) frames; the last blocks may be shorter or empty.
// Used for the devices of one engine (frames_reserve) and, by the callers, for the ranks of a
// multi-process run (ecoz2b200_shard_range).
inline long shard_block(long total, int parts) { return parts > 0 ? (total + parts - 1) / parts : total; }
inline void shard_range(long total, int part, int parts, long *lo, long *hi) {
    const long per = shard_block(total, parts);
    long a = (long) part * per;
    if (a > total) a = total;
    long b = a + per;
    if (b > total) b = total;
    *lo = a;
    *hi = b;
}

// One GPU's share of the frames + a replica of the codebook state.
struct Shard {
    int device = 0;
    int num_sms = 148;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_stream = nullptr;   // uploads that overlap the search (kmeans_pass_host)
    cudaEvent_t ev_chunk[2] = {};
    cudaEvent_t ev_qfree[2] = {};         // quantize_host: chunk buffer b may be overwritten
    cudaEvent_t ev[8] = {};   // phase boundaries of the current pass (ev[6], ev[7]: FP32 search)
    void *comm = nullptr;     // ncclComm_t

    long n = 0;               // frames resident on this shard
    long frame_offset = 0;    // index of its first frame in the engine's append order
    double *x64 = nullptr;    // [n][D]   original FP64 frames (exact re-evaluation, cell sums)
    float *x32 = nullptr;     // [n][D]   FP32 copy for the search
    int32_t *idx = nullptr;   // [n]   assignment of the last k-means pass
    int32_t *idx_tmp = nullptr;  // [n] scratch assignment (inertia / quantize)
    uint16_t *seq = nullptr;  // [n]   symbols (quantize)
    double *mindist = nullptr;  // [n]   ddmin - 1 of the last pass
    void *sc_mem = nullptr;   // scratch of the FP32 search / FP64 re-evaluation
    SearchScratch sc;
    void *cs_mem = nullptr;   // scratch of the sorted cell accumulation
    CellScratch cs;
    long cap = 0;             // capacity (frames) of the per-frame buffers
    // padded copy of x64 for the gather of the sorted cell sums (rows of whole 64-byte atoms,
    // vq_cells.cu); made lazily by the first pass after the frame set changed
    double *x64p = nullptr;
    long x64p_cap = 0;
    bool x64p_valid = false;
    // the frames as ready-made operands of the tensor-core search (launch_prepare_frames_tc): built by
    // the first search after the frame set (or the bound constant) changed
    void *xop = nullptr;
    size_t xop_cap = 0;
    bool xop_valid = false;
    float xop_err = 0.f;

    // codebook replica
    double *refl = nullptr;   // [max_M][D]
    double *raas = nullptr;   // [max_M][D]
    float *w32 = nullptr;     // search tiles
    double *w64 = nullptr;
    float *w1max = nullptr;
    float *w1 = nullptr;      // [K] sum_n |w_i[n]|
    float *wtc = nullptr;     // tensor-core search operand (UMMA canonical layout, TF32 hi/lo)
    double *packed = nullptr; // PackedLayout(max_M, D)
    double *packed_scratch = nullptr;  // same size: sink of the on-demand minDist pass
    double *raas_prev = nullptr;       // [max_M][D] codebook the last assignment was made against
    double *scalars = nullptr;  // [8] misc device scalars (sigma sum, ...)
    void *fmt_scratch = nullptr;   // report formatter: line offsets
    char *fmt_text = nullptr;      // report formatter: text of this shard's frames
    long fmt_cap = 0;              // frames fmt_text is sized for
    long fmt_scratch_cap = 0;      // frames fmt_scratch is sized for
    // report texts that stay in HBM until a writer thread has copied them out (min_dists_text_device)
    struct TextBuf { char *d = nullptr; size_t cap = 0; cudaEvent_t ready = nullptr; bool busy = false; };
    std::vector<TextBuf> text_pool;
    // device-side statistics, read only when the counters are asked for (no per-pass host sync):
    // [0] frames audited, [1] audit: index differs, [2] audit: real mis-assignment, [3] frames
    // re-evaluated in FP64
    unsigned long long *dstat = nullptr;
};

struct Counters {
    double assign_ms = 0, accumulate_ms = 0, update_ms = 0, allreduce_ms = 0;
    double search_main_ms = 0;  // the dominant kernel alone (FP32 / FP64 full search)
    long search_main_launches = 0;
    long kernel_launches = 0;
    double evals = 0;
    long rechecked = 0;
    long audited = 0, audit_idx_diff = 0, audit_mismatches = 0;
};

class Engine {
  public:
    Engine(const int *devices, int num_devices, int P, int max_M);
    ~Engine();
    bool ok() const { return ok_; }

    int set_mode(int mode);
    int set_kappa(double kappa);
    int set_kappa_tc(double kappa);
    int set_audit(int every) { if (every < 0) return ECOZ2B200_ERR_ARG; audit_every_ = every; return ECOZ2B200_OK; }
    // debug: the tensor-core search's raw distortion values of the first n_rows frames of
    // shard 0 against the current codebook, [n_rows][K] floats
    int debug_tc_distortions(float *out, long n_rows);
    int comm_init(const unsigned char id[128], int rank, int world);

    int frames_reserve(long total);
    int frames_clear();
    int frames_append_host(const double *x, long n);
    int frames_append_device(const double *x_dev, long n);
    int frames_commit();
    long frames_count() const { return n_total_; }

    int codebook_initial();
    int codebook_set_reflections(const double *refl, int K);
    int codebook_set_raas(const double *raas, int K);
    int codebook_grow();
    int codebook_size() const { return K_; }
    int codebook_get(double *refl_out, double *raas_out);

    // ddmin_out: the winner's distortion per frame (raw), or distortion - 1 (the reference's
    // minDist / quantize summand, computed on the device) when raw = false
    int assign(int32_t *idx_out, double *ddmin_out, long *num_rechecked, bool raw = true);
    int kmeans_pass(double *cells_out, int32_t *cardd_out, double *discel_out, double *DD_out);
    // the same pass over n frames that are still in HOST memory (pinned for full speed): they
    // replace the resident frame set, uploaded in chunks on a copy stream while the search of
    // the previous chunk runs; cell sums, allreduce and update follow as in kmeans_pass.
    int kmeans_pass_host(const double *x, long n, double *cells_out, int32_t *cardd_out,
                         double *discel_out, double *DD_out);
    int sigma(double avgDistortion, double *sigma_out);
    int inertia(double *inertia_out);
    // Random-access fill of a RESERVED frame set, for loaders that read several files / ranges at
    // once: frames [pos, pos + n) from host memory (pinned for full speed), copied on
    // stream_by_device[device ordinal of the owning shard] and complete on return.  Thread-safe
    // (touches no engine state); frames_mark_filled(total) then declares the set complete.
    int frames_put_host(long pos, const double *x, long n, const cudaStream_t *stream_by_device);
    int frames_mark_filled(long total);
    int last_assignment(int32_t *idx_out, double *mindist_out);
    // the lines "codeword,minDistortion\n" of the last pass's assignment (all local frames in append
    // order) formatted ON THE DEVICE, byte for byte the reference's "%d,%f\n" (report.c:118-121); the
    // text lands in one of two pinned host buffers owned by the engine (valid until the call after the
    // next).  ECOZ2B200_ERR_ARG with *bytes = 0 when a value cannot be formatted there (inf / nan).
    int min_dists_text(const char **text, size_t *bytes);
    std::vector<int> devices() const {
        std::vector<int> d;
        for (const Shard &s : shards_) d.push_back(s.device);
        return d;
    }
    // Same lines, but the text STAYS ON THE DEVICE: one segment per shard (append order), each in a
    // buffer of a per-shard pool, complete when `ready` has fired.  The caller (the report's writer
    // threads, vq_io.cpp) copies the segments out on streams of its own while the next stages run and
    // hands each one back with release_text (thread-safe).  When all kTextPoolMax buffers of a shard
    // are with writers the call waits for one to come back.
    struct TextSegment { int shard; int slot; int device; const char *d_text; size_t bytes; cudaEvent_t ready; };
    int min_dists_text_device(std::vector<TextSegment> &segs);
    int min_dists_text_device_impl(std::vector<TextSegment> &segs);
    int ensure_fmt_scratch(Shard &s);
    const void *frame_operands(Shard &s, float err_tc);   // pre-staged operands of s.x32, or nullptr
    void release_text(const TextSegment &seg);
    int learn(double eps, int stop_M, ecoz2b200_stage_cb on_stage, ecoz2b200_pass_cb on_pass,
              void *ctx);
    int quantize_host(const double *x, long T, uint16_t *seq_out, double *avg);
    // cbook_quantize (vq/vq.c:154-157) of ALL resident frames against R codebooks of K codewords
    // (reflections [R][K][D]) -- the work of a SeqProvider (vq/vq.c:285-315) / of vq_classify.c:265-280
    // for every predictor file at once.  Segment f = frames [seg_offsets[f], seg_offsets[f + 1]) of the
    // append order (a predictor file).  seq_out [R][n] symbols (optional), dists_out [R][num_segments]
    // average distortion (the `dist` of cbook_quantize) (optional).  The engine's own codebook is
    // left alone on the one-sweep route (tensor mode, K = 32, 64, 128 or 256) and is
    // overwritten by the last model otherwise.  *one_sweep (optional) says which route ran.
    int quantize_models(const double *reflections, int R, int K, const long *seg_offsets, int num_segments,
                        uint16_t *seq_out, double *dists_out, int *one_sweep);
    int quantize_resident(uint16_t *seq_out, double *avg);

    void counters_reset();
    const Counters &counters();   // fetches the device-side statistics (synchronises the streams)

    int P() const { return P_; }
    int D() const { return D_; }
    long global_frames() const { return n_global_; }
    int world() const { return world_; }
    int rank() const { return rank_; }

  private:
    int ensure_capacity(Shard &s, long n);
    int ensure_local_comms();
    // fuse_cells: the tensor-core search also adds every frame outside the error bound to the packed
    // cell sums / s.cs.hist (the flagged ones follow after their re-evaluation)
    void search(Shard &s, const double *x64, const float *x32, long n, int32_t *idx,
                const SearchScratch &sc, bool fuse_cells = false);
    // search + winner distortion (+ cell sums) of one block of frames; records ev[1]
    // between the two phases.  K <= 32: one fused FP64 kernel.
    void pass_kernels(Shard &s, const double *x64, const float *x32, long n, int32_t *idx,
                      const SearchScratch &sc, double *mindist, uint16_t *seq, bool with_cells);
    // fused mid-size pass of one block of frames: tensor-core search + cell sums of the frames outside
    // the error bound (vq_pass_mid), FP64 re-evaluation of the others, their rows added afterwards
    void mid_pass(Shard &s, const double *x64, const float *x32, long n, int32_t *idx,
                  const SearchScratch &sc, bool first_block);
    int stage_codebook();  // raas -> search operands on every shard
    int pass_tail(double *cells_out, int32_t *cardd_out, double *discel_out, double *DD_out);
    int sync_all();
    void harvest_times(bool with_cells);

    bool small_path() const {
        return use_small_ && pass_small_supported(K_, D_) && (!tc_path() || pass_small_fast(K_, D_));
    }
    bool tc_path() const { return mode_ == ECOZ2B200_MODE_TENSOR && assign_tc_supported(K_, D_); }
    bool mid_path() const { return tc_path() && !small_path() && pass_mid_supported(K_, D_); }
    bool fuse_path() const {
        return tc_path() && !small_path() && !mid_path() && K_ >= fuse_min_k_ && assign_tc_fuses_cells(K_, D_);
    }
    int fuse_min_k_ = 4096;      // ECOZ2_VQ_TC_CELLS_MINK: smallest codebook whose search carries the cell sums (K=2048: 4.2 against 4.4 ms per pass, but the search kernel itself 4.06 against 3.3 ms; K=1024: 3.2 against 2.75)
    bool use_small_ = true;   // ECOZ2_VQ_FUSED_SMALL=0 forces the tiled path at every K
    bool use_cell_sort_ = true;  // ECOZ2_VQ_CELL_SORT=0: per-frame atomics (vq_accumulate) instead
    bool mindist_valid_ = true;  // s.mindist holds the last pass's values (else: compute on demand)
    bool inertia_exact_ = false; // inside inertia(): frames with a negative minimum take the exact scan
    int audit_every_ = 0;        // ECOZ2_VQ_AUDIT=n: all-FP64 search of every n-th frame after each search
    int K_prev_ = 0;             // rows of raas_prev
    int ensure_mindist();
    bool ok_ = false;
    int P_, D_, max_M_;
    int K_ = 0;
    int mode_ = ECOZ2B200_MODE_TENSOR;   // falls back to the FFMA search where unsupported
    double kappa_ = 8.0;
    double kappa_tc_ = 40.0;
    std::vector<Shard> shards_;
    long n_total_ = 0;   // local frames (all shards)
    long n_global_ = 0;  // all ranks
    long reserved_ = 0;  // frames_reserve() total (fixes the shard boundaries)
    long per_shard_ = 0;
    long appended_ = 0;
    bool committed_ = false;
    // scratch of quantize_host (per shard, lazily allocated)
    struct QBuf { double *x64 = nullptr; float *x32 = nullptr; int32_t *idx = nullptr;
                  void *sc_mem = nullptr; SearchScratch sc; uint16_t *seq = nullptr; long cap = 0; };
    std::vector<QBuf> qbuf_;
    // scratch of quantize_models (per shard, lazily allocated, grows)
    struct MBuf {
        double *refl = nullptr, *raas = nullptr, *w64 = nullptr; float *w32 = nullptr, *w1 = nullptr, *w1max = nullptr, *wtc = nullptr;
        long cols = 0;                       // codewords the operand buffers are sized for
        int32_t *idx = nullptr; size_t idx_items = 0;
        uint16_t *seq = nullptr; size_t seq_items = 0;
        long *chunks = nullptr; int *seg_first = nullptr; double *partial = nullptr, *sums = nullptr;
        size_t chunk_cap = 0, seg_cap = 0, partial_cap = 0, sums_cap = 0;
    };
    std::vector<MBuf> mbuf_;
    int append_block(const double *x, long n, bool device_src);
    int rank_ = 0, world_ = 1;
    bool multi_process_ = false;
    bool local_comms_ready_ = false;
    Counters ctr_;
    char *h_text_[2] = {nullptr, nullptr};   // pinned: report text, double buffered
    size_t h_text_cap_[2] = {0, 0};
    int h_text_slot_ = 0;
    static constexpr int kTextPoolMax = 12;   // report texts in flight per shard (one LBG run 2 -> 4096)
    std::mutex text_mu_;
    std::condition_variable text_cv_;
    std::vector<double> h_packed_;  // host mirror of the packed buffer tail / cells
    double h_count_ = 0;            // kmeans_pass_host: frames of all ranks (read back with the pass)
};

}  // namespace vqb200
