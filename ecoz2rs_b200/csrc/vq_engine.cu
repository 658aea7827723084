// vq_engine.cu -- device shards, the k-means pass and the LBG driver.
//
// Control flow restated from the reference (ecoz2/src/vq/vq_learn_par.c:42-163);
// the arithmetic lives in vq_kernels.cu.  One host thread drives every local GPU
// through its own stream; all cross-GPU / cross-rank traffic is ONE
// ncclAllReduce(sum, f64) of the packed cell buffer per pass.
#include "vq_engine.h"

#include <cfloat>
#include <chrono>
#include <mutex>
#include <cmath>
#include <cstdio>
#include <cstdlib>
#include <cstring>

#include "vq_nccl.h"

namespace vqb200 {

static thread_local std::string g_last_error;
void set_last_error(const std::string &msg) {
    g_last_error = msg;
    fprintf(stderr, "ecoz2_vq_b200: %s\n", msg.c_str());
}
const char *get_last_error() { return g_last_error.c_str(); }

namespace {
typedef int (*CtxGetFn)(void **);
typedef int (*CtxSetFn)(void *);
CtxGetFn g_ctx_get = nullptr;
CtxSetFn g_ctx_set = nullptr;
std::once_flag g_ctx_once;
void ctx_fns_init() {
    void *g = nullptr, *s = nullptr;
    cudaDriverEntryPointQueryResult q1, q2;
    if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &g, cudaEnableDefault, &q1) == cudaSuccess &&
        cudaGetDriverEntryPoint("cuCtxSetCurrent", &s, cudaEnableDefault, &q2) == cudaSuccess &&
        q1 == cudaDriverEntryPointSuccess && q2 == cudaDriverEntryPointSuccess && g && s) {
        g_ctx_get = reinterpret_cast<CtxGetFn>(g);
        g_ctx_set = reinterpret_cast<CtxSetFn>(s);
    } else {
        cudaGetLastError();
    }
}
}  // namespace

DeviceGuard::DeviceGuard() {
    std::call_once(g_ctx_once, ctx_fns_init);
    if (g_ctx_get && g_ctx_get(&ctx_) == 0) have_ = true;
}
DeviceGuard::~DeviceGuard() {
    if (!have_ || !g_ctx_set) return;
    void *now = nullptr;
    if (g_ctx_get(&now) == 0 && now == ctx_) return;
    g_ctx_set(ctx_);
}

#define CUDA_TRY(expr)                                                                     \
    do {                                                                                   \
        cudaError_t _e = (expr);                                                           \
        if (_e != cudaSuccess) {                                                           \
            set_last_error(std::string(#expr) + ": " + cudaGetErrorString(_e));            \
            return ECOZ2B200_ERR_CUDA;                                                     \
        }                                                                                  \
    } while (0)

#define NCCL_TRY(api, expr)                                                                \
    do {                                                                                   \
        int _r = (expr);                                                                   \
        if (_r != 0) {                                                                     \
            set_last_error(std::string(#expr) + ": " + (api)->GetErrorString(_r));         \
            return ECOZ2B200_ERR_NCCL;                                                     \
        }                                                                                  \
    } while (0)

static double now_s() {
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}

// ---------------------------------------------------------------------------------
Engine::Engine(const int *devices, int num_devices, int P, int max_M)
    : P_(P), D_(P + 1), max_M_(max_M) {
    if (P < 1 || D_ > kMaxD || max_M < 2 || max_M > 65536 || num_devices < 1) {
        set_last_error("ecoz2b200_create: bad argument");
        return;
    }
    int visible = 0;
    if (cudaGetDeviceCount(&visible) != cudaSuccess || visible == 0) {
        set_last_error("no CUDA device: the B200 VQ path has no CPU fallback");
        return;
    }
    shards_.resize(num_devices);
    qbuf_.resize(2 * (size_t) num_devices);   // two chunk buffers per device (quantize_host)
    const PackedLayout L{max_M_, D_};
    const size_t cb = sizeof(double) * (size_t) max_M_ * D_;
    const int K32 = (max_M_ + kTileF32 - 1) / kTileF32 * kTileF32;
    const int K64 = (max_M_ + kTileF64 - 1) / kTileF64 * kTileF64;
    for (int i = 0; i < num_devices; i++) {
        Shard &s = shards_[i];
        s.device = devices ? devices[i] : i;
        if (s.device < 0 || s.device >= visible) {
            set_last_error("ecoz2b200_create: device ordinal out of range");
            return;
        }
        auto chk = [&](cudaError_t e, const char *what) {
            if (e != cudaSuccess) {
                set_last_error(std::string(what) + ": " + cudaGetErrorString(e));
                return false;
            }
            return true;
        };
        if (!chk(cudaSetDevice(s.device), "cudaSetDevice")) return;
        cudaDeviceProp prop;
        if (!chk(cudaGetDeviceProperties(&prop, s.device), "cudaGetDeviceProperties")) return;
        s.num_sms = prop.multiProcessorCount;
        if (!chk(cudaStreamCreateWithFlags(&s.stream, cudaStreamNonBlocking), "cudaStreamCreate")) return;
        for (auto &e : s.ev)
            if (!chk(cudaEventCreate(&e), "cudaEventCreate")) return;
        if (!chk(cudaMalloc(&s.refl, cb), "cudaMalloc refl")) return;
        if (!chk(cudaMalloc(&s.raas, cb), "cudaMalloc raas")) return;
        if (!chk(cudaMalloc(&s.w32, sizeof(float) * (size_t) K32 * D_), "cudaMalloc w32")) return;
        if (!chk(cudaMalloc(&s.w64, sizeof(double) * (size_t) K64 * D_), "cudaMalloc w64")) return;
        if (!chk(cudaMalloc(&s.w1max, 16), "cudaMalloc w1max")) return;
        if (tc_codebook_floats(max_M_, D_) > 0 &&
            !chk(cudaMalloc(&s.wtc, sizeof(float) * tc_codebook_floats(max_M_ < 256 ? 256 : max_M_, D_)),
                 "cudaMalloc wtc")) return;
        if (!chk(cudaMalloc(&s.w1, sizeof(float) * (size_t) K32), "cudaMalloc w1")) return;
        if (!chk(cudaMalloc(&s.packed, sizeof(double) * L.size()), "cudaMalloc packed")) return;
        if (!chk(cudaMalloc(&s.packed_scratch, sizeof(double) * L.size()), "cudaMalloc packed_scratch")) return;
        if (!chk(cudaMalloc(&s.raas_prev, cb), "cudaMalloc raas_prev")) return;
        if (!chk(cudaMalloc(&s.scalars, sizeof(double) * 8), "cudaMalloc scalars")) return;
        if (!chk(cudaMalloc(&s.dstat, sizeof(unsigned long long) * 8), "cudaMalloc dstat")) return;
        // on s.stream, NOT cudaMemset: that one runs on the legacy default stream, returns before the
        // fill has run and is not ordered against the non-blocking s.stream -- a codebook uploaded
        // right after the constructor could be zeroed again (seen as a rare wrong pass on cold boxes)
        if (!chk(cudaMemsetAsync(s.dstat, 0, sizeof(unsigned long long) * 8, s.stream), "cudaMemsetAsync")) return;
        if (!chk(cudaMemsetAsync(s.refl, 0, cb, s.stream), "cudaMemsetAsync")) return;
        if (!chk(cudaMemsetAsync(s.raas, 0, cb, s.stream), "cudaMemsetAsync")) return;
        if (!chk(cudaStreamSynchronize(s.stream), "cudaStreamSynchronize")) return;
    }
    const char *fs = getenv("ECOZ2_VQ_FUSED_SMALL");
    if (fs && fs[0] == '0') use_small_ = false;
    const char *csort = getenv("ECOZ2_VQ_CELL_SORT");
    if (csort && csort[0] == '0') use_cell_sort_ = false;
    const char *au = getenv("ECOZ2_VQ_AUDIT");
    if (au && atoi(au) > 0) audit_every_ = atoi(au);
    const char *fk = getenv("ECOZ2_VQ_TC_CELLS_MINK");
    if (fk && atoi(fk) > 0) fuse_min_k_ = atoi(fk);
    ok_ = true;
}

Engine::~Engine() {
    for (char *p : h_text_)
        if (p) cudaFreeHost(p);
    const NcclApi *api = (local_comms_ready_ || multi_process_) ? nccl_api() : nullptr;
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        cudaSetDevice(s.device);
        if (s.stream) cudaStreamSynchronize(s.stream);
        if (s.comm && api) api->CommDestroy(s.comm);
        cudaFree(s.x64); cudaFree(s.x32); cudaFree(s.idx); cudaFree(s.idx_tmp); cudaFree(s.seq);
        cudaFree(s.mindist); cudaFree(s.sc_mem);
        cudaFree(s.refl); cudaFree(s.raas); cudaFree(s.w32); cudaFree(s.w64); cudaFree(s.w1max); cudaFree(s.w1);
        cudaFree(s.wtc); cudaFree(s.cs_mem); cudaFree(s.packed_scratch); cudaFree(s.raas_prev);
        cudaFree(s.packed); cudaFree(s.scalars); cudaFree(s.dstat);
        cudaFree(s.fmt_scratch); cudaFree(s.fmt_text); cudaFree(s.x64p); cudaFree(s.xop);
        for (auto &tb : s.text_pool) {
            cudaFree(tb.d);
            if (tb.ready) cudaEventDestroy(tb.ready);
        }
        if (i < mbuf_.size()) {
            MBuf &m = mbuf_[i];
            cudaFree(m.refl); cudaFree(m.raas); cudaFree(m.w64); cudaFree(m.w32); cudaFree(m.w1); cudaFree(m.w1max);
            cudaFree(m.wtc); cudaFree(m.idx); cudaFree(m.seq); cudaFree(m.chunks); cudaFree(m.seg_first);
            cudaFree(m.partial); cudaFree(m.sums);
        }
        for (int b = 0; b < 2; b++) {
            QBuf &q = qbuf_[2 * i + b];
            cudaFree(q.x64); cudaFree(q.x32); cudaFree(q.idx); cudaFree(q.sc_mem); cudaFree(q.seq);
        }
        for (auto &e : s.ev)
            if (e) cudaEventDestroy(e);
        if (s.stream) cudaStreamDestroy(s.stream);
        if (s.copy_stream) cudaStreamDestroy(s.copy_stream);
        for (cudaEvent_t e : s.ev_chunk)
            if (e) cudaEventDestroy(e);
        for (cudaEvent_t e : s.ev_qfree)
            if (e) cudaEventDestroy(e);
    }
}

int Engine::set_mode(int mode) {
    if (mode != ECOZ2B200_MODE_FP32 && mode != ECOZ2B200_MODE_FP64 && mode != ECOZ2B200_MODE_TENSOR) {
        set_last_error("set_mode: mode not available in this build");
        return ECOZ2B200_ERR_ARG;
    }
    mode_ = mode;
    if (K_ > 0) {   // the tensor-core operand exists only in its mode
        int rc = stage_codebook();
        if (rc) return rc;
        return sync_all();
    }
    return ECOZ2B200_OK;
}
int Engine::set_kappa(double kappa) {
    if (!(kappa >= 0)) return ECOZ2B200_ERR_ARG;
    kappa_ = kappa;
    return ECOZ2B200_OK;
}
int Engine::set_kappa_tc(double kappa) {
    if (!(kappa > 0)) return ECOZ2B200_ERR_ARG;
    kappa_tc_ = kappa;
    if (K_ > 0 && tc_path()) {   // the staged bound weights depend on it (FP16 operand format)
        int rc = stage_codebook();
        if (rc) return rc;
        return sync_all();
    }
    return ECOZ2B200_OK;
}

// ---------------------------------------------------------------------------------
// communicators
int Engine::comm_init(const unsigned char id[128], int rank, int world) {
    if (shards_.size() != 1) {
        set_last_error("comm_init: multi-process mode needs exactly one device per engine");
        return ECOZ2B200_ERR_ARG;
    }
    if (world < 1 || rank < 0 || rank >= world) return ECOZ2B200_ERR_ARG;
    rank_ = rank;
    world_ = world;
    if (world == 1) return ECOZ2B200_OK;
    const NcclApi *api = nccl_api();
    if (!api) return ECOZ2B200_ERR_NCCL;
    NcclUniqueId uid;
    memcpy(uid.internal, id, 128);
    CUDA_TRY(cudaSetDevice(shards_[0].device));
    NCCL_TRY(api, api->CommInitRank(&shards_[0].comm, world, uid, rank));
    multi_process_ = true;
    return ECOZ2B200_OK;
}

int Engine::ensure_local_comms() {
    if (shards_.size() == 1 || local_comms_ready_) return ECOZ2B200_OK;
    const NcclApi *api = nccl_api();
    if (!api) return ECOZ2B200_ERR_NCCL;
    std::vector<void *> comms(shards_.size());
    std::vector<int> devs(shards_.size());
    for (size_t i = 0; i < shards_.size(); i++) devs[i] = shards_[i].device;
    NCCL_TRY(api, api->CommInitAll(comms.data(), (int) shards_.size(), devs.data()));
    for (size_t i = 0; i < shards_.size(); i++) shards_[i].comm = comms[i];
    local_comms_ready_ = true;
    return ECOZ2B200_OK;
}

// in-place sum over every shard of every rank of packed[first .. first+count)
static int allreduce_range(std::vector<Shard> &shards, long first, long count, bool needed) {
    if (!needed) return ECOZ2B200_OK;
    const NcclApi *api = nccl_api();
    if (!api) return ECOZ2B200_ERR_NCCL;
    NCCL_TRY(api, api->GroupStart());
    for (Shard &s : shards) {
        int r = api->AllReduce(s.packed + first, s.packed + first, (size_t) count, kNcclFloat64,
                               kNcclSum, s.comm, s.stream);
        if (r != 0) {
            api->GroupEnd();
            set_last_error(std::string("ncclAllReduce: ") + api->GetErrorString(r));
            return ECOZ2B200_ERR_NCCL;
        }
    }
    NCCL_TRY(api, api->GroupEnd());
    return ECOZ2B200_OK;
}

int Engine::sync_all() {
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
    }
    return ECOZ2B200_OK;
}

// ---------------------------------------------------------------------------------
// frames
int Engine::ensure_capacity(Shard &s, long n) {
    if (n <= s.cap) return ECOZ2B200_OK;
    CUDA_TRY(cudaSetDevice(s.device));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    double *nx64 = nullptr;
    CUDA_TRY(cudaMalloc(&nx64, sizeof(double) * (size_t) n * D_));
    if (s.x64 && s.n > 0) {   // (a device-to-device cudaMemcpy would run on the legacy stream and return early)
        CUDA_TRY(cudaMemcpyAsync(nx64, s.x64, sizeof(double) * (size_t) s.n * D_, cudaMemcpyDeviceToDevice,
                                 s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
    }
    cudaFree(s.x64); s.x64 = nx64;
    cudaFree(s.x32); cudaFree(s.idx); cudaFree(s.idx_tmp); cudaFree(s.seq); cudaFree(s.mindist);
    cudaFree(s.sc_mem); cudaFree(s.cs_mem);
    s.x32 = nullptr; s.idx = s.idx_tmp = nullptr; s.sc_mem = nullptr; s.seq = nullptr; s.mindist = nullptr;
    s.cs_mem = nullptr;
    CUDA_TRY(cudaMalloc(&s.x32, sizeof(float) * (size_t) n * D_));
    CUDA_TRY(cudaMalloc(&s.idx, sizeof(int32_t) * (size_t) n));
    CUDA_TRY(cudaMalloc(&s.idx_tmp, sizeof(int32_t) * (size_t) n));
    CUDA_TRY(cudaMalloc(&s.seq, sizeof(uint16_t) * (size_t) n));
    CUDA_TRY(cudaMalloc(&s.mindist, sizeof(double) * (size_t) n));
    CUDA_TRY(cudaMalloc(&s.sc_mem, scratch_bytes(n)));
    scratch_carve(s.sc_mem, n, &s.sc);
    CUDA_TRY(cudaMalloc(&s.cs_mem, cell_scratch_bytes(n, max_M_)));
    cell_scratch_carve(s.cs_mem, n, max_M_, &s.cs);
    s.cap = n;
    return ECOZ2B200_OK;
}

int Engine::frames_clear() {
    for (Shard &s : shards_) { s.n = 0; s.x64p_valid = false; s.xop_valid = false; }
    n_total_ = n_global_ = 0;
    reserved_ = per_shard_ = appended_ = 0;
    committed_ = false;
    return ECOZ2B200_OK;
}

int Engine::frames_reserve(long total) {
    if (total < 0) return ECOZ2B200_ERR_ARG;
    if (appended_ > 0) {
        set_last_error("frames_reserve: call before the first append");
        return ECOZ2B200_ERR_ARG;
    }
    const long S = (long) shards_.size();
    reserved_ = total;
    per_shard_ = shard_block(total, (int) S);
    for (long i = 0; i < S; i++) {
        shards_[i].frame_offset = i * per_shard_;
        int rc = ensure_capacity(shards_[i], per_shard_ > 0 ? per_shard_ : 1);
        if (rc) return rc;
    }
    return ECOZ2B200_OK;
}

int Engine::append_block(const double *x, long n, bool device_src) {
    if (n < 0 || (n > 0 && !x)) return ECOZ2B200_ERR_ARG;
    if (committed_) {  // appending after a commit starts a new frame set
        set_last_error("frames_append: frames already committed; call frames_clear first");
        return ECOZ2B200_ERR_ARG;
    }
    const cudaMemcpyKind kind = device_src ? cudaMemcpyDeviceToDevice : cudaMemcpyHostToDevice;
    if (reserved_ == 0) {
        if (shards_.size() != 1) {
            set_last_error("frames_append: call frames_reserve first when using several devices");
            return ECOZ2B200_ERR_ARG;
        }
        Shard &s = shards_[0];
        if (s.n + n > s.cap) {
            long want = s.cap * 2 > s.n + n ? s.cap * 2 : s.n + n;
            int rc = ensure_capacity(s, want);
            if (rc) return rc;
        }
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaMemcpyAsync(s.x64 + (size_t) s.n * D_, x, sizeof(double) * (size_t) n * D_, kind, s.stream));
        s.n += n;
        appended_ += n;
        return ECOZ2B200_OK;
    }
    if (appended_ + n > reserved_) {
        set_last_error("frames_append: more frames than reserved");
        return ECOZ2B200_ERR_ARG;
    }
    long pos = appended_, left = n;
    const double *src = x;
    while (left > 0) {
        const long si = pos / per_shard_;
        Shard &s = shards_[si];
        const long room = s.frame_offset + per_shard_ - pos;
        const long take = left < room ? left : room;
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaMemcpyAsync(s.x64 + (size_t) (pos - s.frame_offset) * D_, src,
                                 sizeof(double) * (size_t) take * D_, kind, s.stream));
        s.n = pos - s.frame_offset + take;
        pos += take; left -= take; src += (size_t) take * D_;
    }
    appended_ += n;
    return ECOZ2B200_OK;
}

int Engine::frames_put_host(long pos, const double *x, long n, const cudaStream_t *stream_by_device) {
    if (n < 0 || pos < 0 || (n > 0 && !x) || committed_ || reserved_ == 0 || pos + n > reserved_) {
        set_last_error("frames_put_host: outside the reserved frame set");
        return ECOZ2B200_ERR_ARG;
    }
    long left = n;
    const double *src = x;
    cudaStream_t used[2] = {nullptr, nullptr};
    int used_dev[2] = {-1, -1}, n_used = 0;
    while (left > 0) {
        const long si = pos / per_shard_;
        const Shard &s = shards_[si];
        const long room = s.frame_offset + per_shard_ - pos;
        const long take = left < room ? left : room;
        cudaStream_t st = stream_by_device[s.device];
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaMemcpyAsync(s.x64 + (size_t) (pos - s.frame_offset) * D_, src,
                                 sizeof(double) * (size_t) take * D_, cudaMemcpyHostToDevice, st));
        if (n_used < 2) { used[n_used] = st; used_dev[n_used] = s.device; n_used++; }
        else CUDA_TRY(cudaStreamSynchronize(st));   // a chunk over more than two shards: rare, just wait
        pos += take; left -= take; src += (size_t) take * D_;
    }
    for (int u = 0; u < n_used; u++) {
        CUDA_TRY(cudaSetDevice(used_dev[u]));
        CUDA_TRY(cudaStreamSynchronize(used[u]));
    }
    return ECOZ2B200_OK;
}

int Engine::frames_mark_filled(long total) {
    if (committed_ || reserved_ == 0 || total < 0 || total > reserved_) {
        set_last_error("frames_mark_filled: outside the reserved frame set");
        return ECOZ2B200_ERR_ARG;
    }
    for (Shard &s : shards_) {
        long n = total - s.frame_offset;
        if (n < 0) n = 0;
        if (n > per_shard_) n = per_shard_;
        s.n = n;
    }
    appended_ = total;
    return ECOZ2B200_OK;
}

int Engine::frames_append_host(const double *x, long n) { return append_block(x, n, false); }
int Engine::frames_append_device(const double *x_dev, long n) {
    // the source lives on devices[0]; make sure work queued there by the caller's
    // (default / torch) stream is complete before our stream reads it
    cudaSetDevice(shards_[0].device);
    cudaDeviceSynchronize();
    return append_block(x_dev, n, true);
}

int Engine::frames_commit() {
    n_total_ = 0;
    long off = 0;
    for (Shard &s : shards_) {
        s.x64p_valid = false; s.xop_valid = false;
        if (reserved_ == 0) s.frame_offset = off;
        off += s.n;
        n_total_ += s.n;
        if (s.n > 0) {
            CUDA_TRY(cudaSetDevice(s.device));
            launch_convert_frames(s.x64, s.x32, s.n, D_, s.stream);
            ctr_.kernel_launches++;
            // tensor mode: the frames as ready-made search operands, once per frame set (rebuilt by
            // the search itself if the bound constant changes later)
            if (mode_ == ECOZ2B200_MODE_TENSOR && assign_tc_supported(max_M_ < 2 ? 2 : max_M_, D_))
                frame_operands(s, (float) (kappa_tc_ * ldexp(1.0, -24)));
        }
    }
    n_global_ = n_total_;
    if (multi_process_) {
        Shard &s = shards_[0];
        CUDA_TRY(cudaSetDevice(s.device));
        const double mine = (double) n_total_;
        CUDA_TRY(cudaMemcpyAsync(s.packed, &mine, sizeof(double), cudaMemcpyHostToDevice, s.stream));
        int rc = allreduce_range(shards_, 0, 1, true);
        if (rc) return rc;
        double all = 0;
        CUDA_TRY(cudaMemcpyAsync(&all, s.packed, sizeof(double), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        n_global_ = (long) (all + 0.5);
    } else if (shards_.size() > 1) {
        int rc = ensure_local_comms();
        if (rc) return rc;
    }
    int rc = sync_all();
    if (rc) return rc;
    CUDA_TRY(cudaGetLastError());
    committed_ = true;
    return ECOZ2B200_OK;
}

// ---------------------------------------------------------------------------------
// codebook
int Engine::stage_codebook() {
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        launch_prepare_codebook(s.raas, K_, D_, s.w32, s.w64, s.w1, s.w1max, s.stream);
        ctr_.kernel_launches++;
        if (tc_path()) {
            launch_prepare_codebook_tc(s.raas, K_, D_, s.wtc, s.w1max,
                                       (float) (kappa_tc_ * ldexp(1.0, -24)), s.stream);
            ctr_.kernel_launches++;
        }
    }
    return ECOZ2B200_OK;
}

int Engine::codebook_initial() {
    // vq_learn.i:31-44 : reflections [_, -0.5, 0...] and [_, +0.5, 0...]
    std::vector<double> r((size_t) 2 * D_, 0.0);
    r[1] = -0.5;
    r[D_ + 1] = 0.5;
    return codebook_set_reflections(r.data(), 2);
}

int Engine::codebook_set_reflections(const double *refl, int K) {
    if (!refl || K < 1 || K > max_M_) {
        set_last_error("codebook_set_reflections: bad size");
        return ECOZ2B200_ERR_ARG;
    }
    K_ = K;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaMemcpyAsync(s.refl, refl, sizeof(double) * (size_t) K * D_, cudaMemcpyHostToDevice, s.stream));
        launch_refl_to_raas(s.refl, s.raas, K, D_, s.stream);
        ctr_.kernel_launches++;
    }
    int rc = stage_codebook();
    if (rc) return rc;
    return sync_all();
}

int Engine::codebook_set_raas(const double *raas, int K) {
    if (!raas || K < 1 || K > max_M_) return ECOZ2B200_ERR_ARG;
    K_ = K;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaMemsetAsync(s.refl, 0, sizeof(double) * (size_t) K * D_, s.stream));
        CUDA_TRY(cudaMemcpyAsync(s.raas, raas, sizeof(double) * (size_t) K * D_, cudaMemcpyHostToDevice, s.stream));
    }
    int rc = stage_codebook();
    if (rc) return rc;
    return sync_all();
}

int Engine::codebook_grow() {
    if (K_ < 1 || 2 * K_ > max_M_) {
        set_last_error("codebook_grow: maximum codebook size reached");  // vq_learn.i:48-51
        return ECOZ2B200_ERR_ARG;
    }
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        launch_grow(s.refl, s.raas, K_, D_, s.stream);
        ctr_.kernel_launches += 2;
    }
    K_ *= 2;
    return stage_codebook();
}

int Engine::codebook_get(double *refl_out, double *raas_out) {
    Shard &s = shards_[0];
    CUDA_TRY(cudaSetDevice(s.device));
    const size_t bytes = sizeof(double) * (size_t) K_ * D_;
    if (refl_out) CUDA_TRY(cudaMemcpyAsync(refl_out, s.refl, bytes, cudaMemcpyDeviceToHost, s.stream));
    if (raas_out) CUDA_TRY(cudaMemcpyAsync(raas_out, s.raas, bytes, cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    return ECOZ2B200_OK;
}

// debug aid (ECOZ2_VQ_DEBUG_POISON=1): fill every SM's shared memory with NaN patterns
// before a search so any read of not-yet-delivered tile data becomes visible.
__global__ void vq_debug_poison_smem() {
    extern __shared__ unsigned int poison[];
    for (int i = threadIdx.x; i < 50 * 1024; i += blockDim.x) poison[i] = 0x7fc00000u + i;
    __syncthreads();
    if (poison[(threadIdx.x * 7919) % (50 * 1024)] == 1u) printf("x");
}
static void debug_poison(const Shard &s) {
    static int on = -1;
    if (on < 0) on = getenv("ECOZ2_VQ_DEBUG_POISON") ? 1 : 0;
    if (!on) return;
    cudaFuncSetAttribute(vq_debug_poison_smem, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    vq_debug_poison_smem<<<s.num_sms, 256, 200 * 1024, s.stream>>>();
}

// dstat[3] += *flag_count (frames the search handed to the FP64 re-evaluation)
__global__ void vq_stat_add(const int32_t *__restrict__ counter, unsigned long long *__restrict__ stat) {
    if (threadIdx.x == 0 && blockIdx.x == 0) *stat += (unsigned long long) *counter;
}

void Engine::counters_reset() {
    ctr_ = Counters();
    for (Shard &s : shards_) {
        cudaSetDevice(s.device);
        cudaMemsetAsync(s.dstat, 0, sizeof(unsigned long long) * 8, s.stream);
    }
}

const Counters &Engine::counters() {
    long st[4] = {0, 0, 0, 0};
    for (Shard &s : shards_) {
        unsigned long long h[4] = {0, 0, 0, 0};
        cudaSetDevice(s.device);
        cudaMemcpyAsync(h, s.dstat, sizeof h, cudaMemcpyDeviceToHost, s.stream);
        cudaStreamSynchronize(s.stream);
        for (int i = 0; i < 4; i++) st[i] += (long) h[i];
    }
    ctr_.audited = st[0];
    ctr_.audit_idx_diff = st[1];
    ctr_.audit_mismatches = st[2];
    ctr_.rechecked = st[3];
    return ctr_;
}

// ---------------------------------------------------------------------------------
// (a) the search on one shard: FP32 main pass + FP64 re-evaluation of the frames
// inside the FP32 ambiguity bound, or the all-FP64 verification mode.
const void *Engine::frame_operands(Shard &s, float err_tc) {
    const size_t bytes = tc_frames_bytes(s.n, D_);
    if (bytes == 0 || !tc_operands_f16()) return nullptr;
    if (s.xop_valid && s.xop_err == err_tc) return s.xop;
    if (s.xop_cap < bytes) {
        cudaFree(s.xop);
        s.xop = nullptr; s.xop_cap = 0;
        const size_t want = tc_frames_bytes(s.cap > s.n ? s.cap : s.n, D_);
        if (cudaMalloc(&s.xop, want) != cudaSuccess) {   // no room: the loaders convert per search
            cudaGetLastError();
            s.xop = nullptr;
            return nullptr;
        }
        s.xop_cap = want;
    }
    launch_prepare_frames_tc(s.x32, s.n, D_, err_tc, s.xop, s.stream);
    ctr_.kernel_launches++;
    s.xop_valid = true;
    s.xop_err = err_tc;
    return s.xop;
}

void Engine::search(Shard &s, const double *x64, const float *x32, long n, int32_t *idx,
                    const SearchScratch &sc, bool fuse_cells) {
    if (n <= 0) return;
    debug_poison(s);
    cudaEventRecord(s.ev[6], s.stream);
    if (mode_ == ECOZ2B200_MODE_FP64) {
        launch_assign_f64(x64, n, D_, s.w64, K_, nullptr, nullptr, idx, s.num_sms, s.stream);
        cudaEventRecord(s.ev[7], s.stream);
        ctr_.kernel_launches += 1;
    } else if (tc_path()) {
        // tcgen05 split-TF32 search; same flag / re-evaluation protocol with its own kappa.
        const float err_tc = (float) (kappa_tc_ * ldexp(1.0, -24));
        const float err_scale = (float) (kappa_ * ldexp(1.0, -24));
        cudaMemsetAsync(sc.counters, 0, 4 * sizeof(int32_t), s.stream);
        const void *xop = (x32 == s.x32 && n == s.n) ? frame_operands(s, err_tc) : nullptr;
        launch_assign_tc(x32, n, D_, s.wtc, K_, nullptr, nullptr, sc, true, s.w1, s.w1max, err_tc, idx,
                         nullptr, 0, s.num_sms, s.stream, x64, fuse_cells ? s.packed : nullptr,
                         fuse_cells ? s.cs.hist : nullptr, xop);
        cudaEventRecord(s.ev[7], s.stream);
        launch_recheck(x32, x64, n, D_, s.w32, s.w64, s.w1, s.w1max, s.raas, K_, 2.f * err_scale, sc,
                       idx, s.num_sms, s.stream, s.wtc, err_tc);
        vq_stat_add<<<1, 32, 0, s.stream>>>(sc.counters, s.dstat + 3);
        ctr_.kernel_launches += 6;
        if (fuse_cells) {
            launch_cell_fixup(x64, idx, D_, sc.flag_list, sc.counters + 0, s.packed, s.cs.hist, s.num_sms, s.stream);
            ctr_.kernel_launches++;
        }
    } else {
        // |error| of one FP32 distortion <= kappa * 2^-24 * |rx[0]| * sum_n |w[n]|.
        // Main pass: flag when the best-two gap is below twice that bound taken with the
        // largest codeword norm.  Re-evaluation: per-codeword norms, twice the kappa.
        const float err_scale = (float) (kappa_ * ldexp(1.0, -24));
        cudaMemsetAsync(sc.counters, 0, 4 * sizeof(int32_t), s.stream);
        launch_assign_f32(x32, n, D_, s.w32, K_, sc, s.w1, s.w1max, err_scale, idx, s.num_sms,
                          s.stream);
        cudaEventRecord(s.ev[7], s.stream);
        launch_recheck(x32, x64, n, D_, s.w32, s.w64, s.w1, s.w1max, s.raas, K_, 2.f * err_scale, sc,
                       idx, s.num_sms, s.stream);
        vq_stat_add<<<1, 32, 0, s.stream>>>(sc.counters, s.dstat + 3);
        ctr_.kernel_launches += 6;
    }
    if (audit_every_ > 0 && mode_ != ECOZ2B200_MODE_FP64) {
        launch_audit(x64, n, D_, s.raas, K_, idx, audit_every_, s.dstat, s.num_sms, s.stream);
        ctr_.kernel_launches++;
    }
    ctr_.evals += (double) n * K_;
}

void Engine::mid_pass(Shard &s, const double *x64, const float *x32, long n, int32_t *idx,
                      const SearchScratch &sc, bool first_block) {
    if (n <= 0) {
        if (first_block) cudaMemsetAsync(s.cs.hist, 0, sizeof(int32_t) * (size_t) (K_ + 1), s.stream);
        return;
    }
    debug_poison(s);
    const float err_tc = (float) (kappa_tc_ * ldexp(1.0, -24));
    const float err_scale = (float) (kappa_ * ldexp(1.0, -24));
    cudaMemsetAsync(sc.counters, 0, 4 * sizeof(int32_t), s.stream);
    // (the sorted path's scratch is idle here: perm takes the rows whose cell list was full)
    launch_pass_mid(x64, n, D_, s.wtc, K_, sc, err_tc, idx, s.packed, s.cs.hist, s.cs.perm, s.cs.cursor,
                    first_block, s.num_sms, s.stream);
    cudaEventRecord(s.ev[7], s.stream);
    launch_recheck(x32, x64, n, D_, s.w32, s.w64, s.w1, s.w1max, s.raas, K_, 2.f * err_scale, sc, idx,
                   s.num_sms, s.stream, s.wtc, err_tc);
    vq_stat_add<<<1, 32, 0, s.stream>>>(sc.counters, s.dstat + 3);
    launch_cell_fixup(x64, idx, D_, sc.flag_list, sc.counters + 0, s.packed, s.cs.hist, s.num_sms, s.stream);
    launch_cell_fixup(x64, idx, D_, s.cs.perm, s.cs.cursor, s.packed, s.cs.hist, s.num_sms, s.stream);
    ctr_.kernel_launches += 9;
    if (audit_every_ > 0) {
        launch_audit(x64, n, D_, s.raas, K_, idx, audit_every_, s.dstat, s.num_sms, s.stream);
        ctr_.kernel_launches++;
    }
    ctr_.evals += (double) n * K_;
}

void Engine::pass_kernels(Shard &s, const double *x64, const float *x32, long n, int32_t *idx,
                          const SearchScratch &sc, double *mindist, uint16_t *seq, bool with_cells) {
    if (with_cells && mid_path() && seq == nullptr) {
        // one sweep over the FP64 frames: search + cell sums (vq_mid.cu)
        cudaEventRecord(s.ev[6], s.stream);
        mid_pass(s, x64, x32, n, idx, sc, true);
        cudaEventRecord(s.ev[1], s.stream);
        cudaMemcpyAsync(s.raas_prev, s.raas, sizeof(double) * (size_t) K_ * D_, cudaMemcpyDeviceToDevice,
                        s.stream);
        launch_cell_finish(s.raas, s.cs.hist, K_, D_, s.packed, s.stream);
        ctr_.kernel_launches++;
        K_prev_ = K_;
        if (idx == s.idx) mindist_valid_ = false;
        return;
    }
    if (small_path()) {
        // small codebooks are HBM-bound: one all-FP64 sweep does search + sums
        cudaEventRecord(s.ev[6], s.stream);
        launch_pass_small(x64, n, D_, s.raas, K_, s.packed, idx, mindist, seq, with_cells, s.num_sms,
                          s.stream);
        cudaEventRecord(s.ev[7], s.stream);
        cudaEventRecord(s.ev[1], s.stream);
        ctr_.kernel_launches += n > 0;
        ctr_.evals += (double) n * K_;
        if (mindist == s.mindist) mindist_valid_ = true;
        return;
    }
    const bool fuse = with_cells && seq == nullptr && fuse_path();
    if (fuse) cudaMemsetAsync(s.cs.hist, 0, sizeof(int32_t) * (size_t) (K_ + 1), s.stream);
    search(s, x64, x32, n, idx, sc, fuse);
    cudaEventRecord(s.ev[1], s.stream);
    if (fuse) {
        cudaMemcpyAsync(s.raas_prev, s.raas, sizeof(double) * (size_t) K_ * D_, cudaMemcpyDeviceToDevice,
                        s.stream);
        launch_cell_finish(s.raas, s.cs.hist, K_, D_, s.packed, s.stream);
        ctr_.kernel_launches++;
        K_prev_ = K_;
        if (idx == s.idx) mindist_valid_ = false;
        return;
    }
    if (with_cells && use_cell_sort_ && cell_sort_supported(K_) && x64 == s.x64) {
        // grouped accumulation; per-frame minDist values are produced on demand from the
        // codebook this assignment was made against
        cudaMemcpyAsync(s.raas_prev, s.raas, sizeof(double) * (size_t) K_ * D_, cudaMemcpyDeviceToDevice,
                        s.stream);
        const int ld = cell_pad_ld(D_);
        if (ld > 0 && !s.x64p_valid && n == s.n) {
            if (s.x64p_cap < s.cap) {
                cudaFree(s.x64p);
                s.x64p = nullptr; s.x64p_cap = 0;
                if (cudaMalloc(&s.x64p, sizeof(double) * (size_t) s.cap * ld) == cudaSuccess) s.x64p_cap = s.cap;
                else cudaGetLastError();   // no room for the copy: the dense rows serve
            }
            if (s.x64p) {
                launch_pad_rows(s.x64, s.x64p, s.n, D_, ld, s.num_sms, s.stream);
                ctr_.kernel_launches++;
                s.x64p_valid = true;
            }
        }
        const bool padded = ld > 0 && s.x64p_valid && n == s.n;
        launch_cell_sums(x64, idx, n, D_, s.raas, K_, s.packed, seq, s.cs, s.num_sms, s.stream,
                         padded ? s.x64p : nullptr, padded ? ld : 0);
        K_prev_ = K_;
        mindist_valid_ = false;
        ctr_.kernel_launches += n > 0 ? 5 : 2;
        return;
    }
    if (inertia_exact_ && !with_cells) {
        // inertia.c:21-25 as written: frames whose minimum distortion is negative (impossible for
        // gain-normalised input) are listed and scanned exactly for min_i dd_i^2
        cudaMemsetAsync(sc.counters + 3, 0, sizeof(int32_t), s.stream);
        launch_accumulate(x64, idx, n, D_, s.raas, K_, s.packed, mindist, seq, false, s.num_sms, s.stream,
                          sc.exact_list, sc.counters + 3);
        launch_inertia_neg(x64, D_, s.raas, K_, sc.exact_list, sc.counters + 3, s.packed, s.num_sms, s.stream);
        ctr_.kernel_launches += 2 * (n > 0);
        return;
    }
    launch_accumulate(x64, idx, n, D_, s.raas, K_, s.packed, mindist, seq, with_cells, s.num_sms,
                      s.stream);
    if (mindist == s.mindist) mindist_valid_ = true;
    ctr_.kernel_launches += n > 0;
}

int Engine::ensure_mindist() {
    if (mindist_valid_) return ECOZ2B200_OK;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        launch_accumulate(s.x64, s.idx, s.n, D_, s.raas_prev, K_prev_, s.packed_scratch, s.mindist, nullptr,
                          false, s.num_sms, s.stream);
        ctr_.kernel_launches += s.n > 0;
    }
    mindist_valid_ = true;
    return ECOZ2B200_OK;
}

int Engine::debug_tc_distortions(float *out, long n_rows) {
    if (!committed_ || K_ < 1 || !tc_path()) {
        set_last_error("debug_tc_distortions: needs committed frames, a codebook and TF32X3 mode");
        return ECOZ2B200_ERR_ARG;
    }
    Shard &s = shards_[0];
    CUDA_TRY(cudaSetDevice(s.device));
    if (n_rows > s.n) n_rows = s.n;
    const int BN = tc_tile_cols(K_, D_);
    const int ld = (K_ + BN - 1) / BN * BN;
    float *dump = nullptr;
    CUDA_TRY(cudaMalloc(&dump, sizeof(float) * (size_t) n_rows * ld));
    launch_assign_tc(s.x32, n_rows, D_, s.wtc, K_, nullptr, nullptr, s.sc, false, s.w1, s.w1max, 0.f,
                     s.idx_tmp, dump, ld, s.num_sms, s.stream);
    cudaError_t e = cudaMemcpy2DAsync(out, sizeof(float) * K_, dump, sizeof(float) * ld,
                                      sizeof(float) * K_, (size_t) n_rows, cudaMemcpyDeviceToHost,
                                      s.stream);
    if (e == cudaSuccess) e = cudaStreamSynchronize(s.stream);
    cudaFree(dump);
    CUDA_TRY(e);
    return ECOZ2B200_OK;
}

int Engine::assign(int32_t *idx_out, double *ddmin_out, long *num_rechecked, bool raw) {
    if (!committed_ || K_ < 1) {
        set_last_error("assign: frames not committed or no codebook");
        return ECOZ2B200_ERR_ARG;
    }
    const PackedLayout L{K_, D_};
    long rechecked = 0;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaMemsetAsync(s.packed + L.dd(), 0, 2 * sizeof(double), s.stream));
        pass_kernels(s, s.x64, s.x32, s.n, s.idx, s.sc, s.mindist, nullptr, false);
    }
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        if (idx_out)
            CUDA_TRY(cudaMemcpyAsync(idx_out + s.frame_offset, s.idx, sizeof(int32_t) * (size_t) s.n,
                                     cudaMemcpyDeviceToHost, s.stream));
        if (ddmin_out)
            CUDA_TRY(cudaMemcpyAsync(ddmin_out + s.frame_offset, s.mindist, sizeof(double) * (size_t) s.n,
                                     cudaMemcpyDeviceToHost, s.stream));
        int32_t fc = 0;
        if (mode_ != ECOZ2B200_MODE_FP64 && !small_path())
            CUDA_TRY(cudaMemcpyAsync(&fc, s.sc.counters, sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        rechecked += fc;
    }
    CUDA_TRY(cudaGetLastError());
    if (ddmin_out && raw)
        for (long v = 0; v < n_total_; v++) ddmin_out[v] += 1.0;  // raw distortion (before the -1)
    if (num_rechecked) *num_rechecked = rechecked;
    return ECOZ2B200_OK;
}

void Engine::harvest_times(bool with_cells) {
    // max over shards of every phase (events were recorded on each shard's stream)
    double a = 0, b = 0, c = 0, d = 0;
    for (Shard &s : shards_) {
        cudaSetDevice(s.device);
        float ms = 0;
        if (cudaEventElapsedTime(&ms, s.ev[0], s.ev[1]) == cudaSuccess && ms > a) a = ms;
        if (cudaEventElapsedTime(&ms, s.ev[1], s.ev[2]) == cudaSuccess && ms > b) b = ms;
        if (with_cells) {
            if (cudaEventElapsedTime(&ms, s.ev[2], s.ev[3]) == cudaSuccess && ms > c) c = ms;
            if (cudaEventElapsedTime(&ms, s.ev[3], s.ev[4]) == cudaSuccess && ms > d) d = ms;
        }
    }
    double m = 0;
    for (Shard &s : shards_) {
        cudaSetDevice(s.device);
        float ms = 0;
        if (s.n > 0 && cudaEventElapsedTime(&ms, s.ev[6], s.ev[7]) == cudaSuccess && ms > m) m = ms;
    }
    ctr_.search_main_ms += m;
    ctr_.search_main_launches += 1;
    ctr_.assign_ms += a;
    ctr_.accumulate_ms += b;
    ctr_.allreduce_ms += c;
    ctr_.update_ms += d;
}

// One k-means pass (vq_learn_par.c:50-125).
int Engine::kmeans_pass(double *cells_out, int32_t *cardd_out, double *discel_out, double *DD_out) {
    if (!committed_ || K_ < 1) {
        set_last_error("kmeans_pass: frames not committed or no codebook");
        return ECOZ2B200_ERR_ARG;
    }
    const PackedLayout L{K_, D_};
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaEventRecord(s.ev[0], s.stream));
        CUDA_TRY(cudaMemsetAsync(s.packed, 0, sizeof(double) * L.size(), s.stream));
        pass_kernels(s, s.x64, s.x32, s.n, s.idx, s.sc, s.mindist, nullptr, true);
        CUDA_TRY(cudaEventRecord(s.ev[2], s.stream));
    }
    return pass_tail(cells_out, cardd_out, discel_out, DD_out);
}

int Engine::kmeans_pass_host(const double *x, long n, double *cells_out, int32_t *cardd_out,
                             double *discel_out, double *DD_out) {
    if (K_ < 1 || n <= 0 || !x) {
        set_last_error("kmeans_pass_host: no codebook or no frames");
        return ECOZ2B200_ERR_ARG;
    }
    const bool pipelined = !small_path() && use_cell_sort_ && cell_sort_supported(K_);
    if (!pipelined) {   // plain sequence: upload, commit, pass
        int rc = frames_clear();
        if (!rc) rc = frames_reserve(n);
        if (!rc) rc = frames_append_host(x, n);
        if (!rc) rc = frames_commit();
        if (rc) return rc;
        return kmeans_pass(cells_out, cardd_out, discel_out, DD_out);
    }
    const PackedLayout L{K_, D_};
    const long S = (long) shards_.size();
    frames_clear();
    reserved_ = n;
    per_shard_ = shard_block(n, (int) S);
    const long chunk = 1L << 20;   // frames per upload chunk (310 MB of FP64 at P = 36)
    for (long i = 0; i < S; i++) {
        Shard &s = shards_[i];
        s.frame_offset = i * per_shard_;
        long mine = n - s.frame_offset;
        if (mine > per_shard_) mine = per_shard_;
        if (mine < 0) mine = 0;
        int rc = ensure_capacity(s, mine > 0 ? mine : 1);
        if (rc) return rc;
        s.n = mine;
        s.x64p_valid = false; s.xop_valid = false;   // new frames: the padded copy is rebuilt by the next resident pass
        CUDA_TRY(cudaSetDevice(s.device));
        if (!s.copy_stream) {
            CUDA_TRY(cudaStreamCreateWithFlags(&s.copy_stream, cudaStreamNonBlocking));
            for (int k = 0; k < 2; k++) CUDA_TRY(cudaEventCreateWithFlags(&s.ev_chunk[k], cudaEventDisableTiming));
        }
        // the copy stream must not overwrite frames an earlier pass on s.stream still reads
        CUDA_TRY(cudaEventRecord(s.ev_chunk[0], s.stream));
        CUDA_TRY(cudaStreamWaitEvent(s.copy_stream, s.ev_chunk[0], 0));
        CUDA_TRY(cudaEventRecord(s.ev[0], s.stream));
        CUDA_TRY(cudaMemsetAsync(s.packed, 0, sizeof(double) * L.size(), s.stream));
    }
    // chunks of all shards interleaved: every device's copy engine and SMs stay busy
    long max_chunks = (per_shard_ + chunk - 1) / chunk;
    for (long c = 0; c < max_chunks; c++) {
        for (Shard &s : shards_) {
            const long off = c * chunk;
            if (off >= s.n) continue;
            const long m = s.n - off < chunk ? s.n - off : chunk;
            CUDA_TRY(cudaSetDevice(s.device));
            CUDA_TRY(cudaMemcpyAsync(s.x64 + (size_t) off * D_, x + (size_t) (s.frame_offset + off) * D_,
                                     sizeof(double) * (size_t) m * D_, cudaMemcpyHostToDevice, s.copy_stream));
            CUDA_TRY(cudaEventRecord(s.ev_chunk[c & 1], s.copy_stream));
            // the copy stream carries nothing but copies (a kernel queued there would wait for
            // the persistent search kernel to leave the SMs and hold the next copy back)
            CUDA_TRY(cudaStreamWaitEvent(s.stream, s.ev_chunk[c & 1], 0));
            launch_convert_frames(s.x64 + (size_t) off * D_, s.x32 + (size_t) off * D_, m, D_, s.stream);
            if (mid_path()) {
                mid_pass(s, s.x64 + (size_t) off * D_, s.x32 + (size_t) off * D_, m, s.idx + off, s.sc, c == 0);
            } else {
                if (fuse_path() && c == 0)
                    cudaMemsetAsync(s.cs.hist, 0, sizeof(int32_t) * (size_t) (K_ + 1), s.stream);
                search(s, s.x64 + (size_t) off * D_, s.x32 + (size_t) off * D_, m, s.idx + off, s.sc, fuse_path());
            }
            ctr_.kernel_launches++;
        }
    }
    appended_ = n;
    n_total_ = n_global_ = n;
    if (S > 1) {
        int rc = ensure_local_comms();
        if (rc) return rc;
    }
    h_count_ = (double) n;   // member: the asynchronous read-back must not outlive its target
    if (multi_process_) {
        // frames of all ranks (the divisor of the average distortion): a one-double allreduce queued
        // behind the chunks, read back with the pass results
        Shard &s = shards_[0];
        const NcclApi *api = nccl_api();
        if (!api) return ECOZ2B200_ERR_NCCL;
        CUDA_TRY(cudaMemcpyAsync(s.scalars + 4, &h_count_, sizeof(double), cudaMemcpyHostToDevice, s.stream));
        NCCL_TRY(api, api->AllReduce(s.scalars + 4, s.scalars + 4, 1, kNcclFloat64, kNcclSum, s.comm, s.stream));
        CUDA_TRY(cudaMemcpyAsync(&h_count_, s.scalars + 4, sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    }
    committed_ = true;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaEventRecord(s.ev[1], s.stream));
        if (s.n > 0) {
            cudaMemcpyAsync(s.raas_prev, s.raas, sizeof(double) * (size_t) K_ * D_, cudaMemcpyDeviceToDevice,
                            s.stream);
            if (mid_path() || fuse_path()) {
                launch_cell_finish(s.raas, s.cs.hist, K_, D_, s.packed, s.stream);
                ctr_.kernel_launches++;
            } else {
                launch_cell_sums(s.x64, s.idx, s.n, D_, s.raas, K_, s.packed, nullptr, s.cs, s.num_sms, s.stream);
                ctr_.kernel_launches += 5;
            }
        }
        CUDA_TRY(cudaEventRecord(s.ev[2], s.stream));
    }
    K_prev_ = K_;
    mindist_valid_ = false;
    const int rc = pass_tail(cells_out, cardd_out, discel_out, DD_out);   // synchronises the streams
    n_global_ = (long) (h_count_ + 0.5);
    return rc;
}

// allreduce of the packed sums, read-back, centroid update and re-staging of the codebook
int Engine::pass_tail(double *cells_out, int32_t *cardd_out, double *discel_out, double *DD_out) {
    const PackedLayout L{K_, D_};
    const bool reduce = shards_.size() > 1 || multi_process_;
    int rc = allreduce_range(shards_, 0, L.size(), reduce);
    if (rc) return rc;
    h_packed_.resize((size_t) L.size());
    {
        Shard &s = shards_[0];
        CUDA_TRY(cudaSetDevice(s.device));
        // tail [counts | discel | DD | inertia] always; raw cell sums only on request
        CUDA_TRY(cudaMemcpyAsync(h_packed_.data() + L.counts(), s.packed + L.counts(),
                                 sizeof(double) * (size_t) (L.size() - L.counts()),
                                 cudaMemcpyDeviceToHost, s.stream));
        if (cells_out)
            CUDA_TRY(cudaMemcpyAsync(cells_out, s.packed, sizeof(double) * (size_t) K_ * D_,
                                     cudaMemcpyDeviceToHost, s.stream));
    }
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaEventRecord(s.ev[3], s.stream));
        launch_centroids(s.packed, K_, D_, s.refl, s.raas, s.stream);
        launch_prepare_codebook(s.raas, K_, D_, s.w32, s.w64, s.w1, s.w1max, s.stream);
        ctr_.kernel_launches += 2;
        if (tc_path()) {
            launch_prepare_codebook_tc(s.raas, K_, D_, s.wtc, s.w1max,
                                       (float) (kappa_tc_ * ldexp(1.0, -24)), s.stream);
            ctr_.kernel_launches++;
        }
        CUDA_TRY(cudaEventRecord(s.ev[4], s.stream));
    }
    rc = sync_all();
    if (rc) return rc;
    CUDA_TRY(cudaGetLastError());
    harvest_times(true);
    if (cardd_out)
        for (int i = 0; i < K_; i++) cardd_out[i] = (int32_t) (h_packed_[L.counts() + i] + 0.5);
    if (discel_out) memcpy(discel_out, h_packed_.data() + L.discel(), sizeof(double) * K_);
    if (DD_out) *DD_out = h_packed_[L.dd()];
    return ECOZ2B200_OK;
}

int Engine::sigma(double avgDistortion, double *sigma_out) {
    // sigma.c:12-29 on the post-update raas and the errPred-normalised cells
    Shard &s = shards_[0];
    CUDA_TRY(cudaSetDevice(s.device));
    CUDA_TRY(cudaMemsetAsync(s.scalars, 0, sizeof(double), s.stream));
    launch_sigma_sum(s.raas, s.packed, K_, D_, s.scalars, s.stream);
    ctr_.kernel_launches++;
    double sum = 0;
    CUDA_TRY(cudaMemcpyAsync(&sum, s.scalars, sizeof(double), cudaMemcpyDeviceToHost, s.stream));
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    CUDA_TRY(cudaGetLastError());
    const double dpc = sum / ((double) K_ * (double) (K_ - 1));
    if (sigma_out) *sigma_out = dpc / avgDistortion;
    return ECOZ2B200_OK;
}

int Engine::inertia(double *inertia_out) {
    // inertia.c:10-31: sum_v min_i distortion^2 with the current (post-update) codebook.
    // min_i dd_i^2 == (min_i dd_i)^2 whenever the minimum is >= 0 (every gain-normalised frame);
    // the other frames take the exact scan (pass_kernels, vq_pass_small*).
    const PackedLayout L{K_, D_};
    const bool reduce = shards_.size() > 1 || multi_process_;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaEventRecord(s.ev[0], s.stream));
        CUDA_TRY(cudaMemsetAsync(s.packed + L.dd(), 0, 2 * sizeof(double), s.stream));
        inertia_exact_ = true;
        pass_kernels(s, s.x64, s.x32, s.n, s.idx_tmp, s.sc, nullptr, nullptr, false);
        inertia_exact_ = false;
        CUDA_TRY(cudaEventRecord(s.ev[2], s.stream));
    }
    int rc = allreduce_range(shards_, L.dd(), 2, reduce);
    if (rc) return rc;
    double tail[2] = {0, 0};
    Shard &s0 = shards_[0];
    CUDA_TRY(cudaSetDevice(s0.device));
    CUDA_TRY(cudaMemcpyAsync(tail, s0.packed + L.dd(), 2 * sizeof(double), cudaMemcpyDeviceToHost, s0.stream));
    rc = sync_all();
    if (rc) return rc;
    CUDA_TRY(cudaGetLastError());
    harvest_times(false);
    if (inertia_out) *inertia_out = tail[1];
    return ECOZ2B200_OK;
}

int Engine::last_assignment(int32_t *idx_out, double *mindist_out) {
    if (mindist_out) {
        int rc = ensure_mindist();
        if (rc) return rc;
    }
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        if (idx_out)
            CUDA_TRY(cudaMemcpyAsync(idx_out + s.frame_offset, s.idx, sizeof(int32_t) * (size_t) s.n,
                                     cudaMemcpyDeviceToHost, s.stream));
        if (mindist_out)
            CUDA_TRY(cudaMemcpyAsync(mindist_out + s.frame_offset, s.mindist, sizeof(double) * (size_t) s.n,
                                     cudaMemcpyDeviceToHost, s.stream));
    }
    return sync_all();
}

int Engine::min_dists_text(const char **text, size_t *bytes) {
    if (text) *text = nullptr;
    if (bytes) *bytes = 0;
    int rc = ensure_mindist();
    if (rc) return rc;
    // pass 1: line lengths and offsets on every shard
    std::vector<unsigned long long> totals(shards_.size(), 0);
    size_t all = 0;
    bool bad_any = false;
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        rc = ensure_fmt_scratch(s);
        if (rc) return rc;
        if (s.fmt_cap < s.n) {
            CUDA_TRY(cudaStreamSynchronize(s.stream));
            cudaFree(s.fmt_text);
            s.fmt_text = nullptr; s.fmt_cap = 0;
            CUDA_TRY(cudaMalloc(&s.fmt_text, format_text_capacity(s.n)));
            s.fmt_cap = s.n;
        }
        launch_format_lengths(s.idx, s.mindist, s.n, s.fmt_scratch, s.stream);
        ctr_.kernel_launches += 2;
    }
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        const unsigned long long *d_total;
        const int *d_bad;
        format_result_ptrs(s.fmt_scratch, s.n, &d_total, &d_bad);
        int bad = 0;
        CUDA_TRY(cudaMemcpyAsync(&totals[i], d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        bad_any = bad_any || bad != 0;
        all += (size_t) totals[i];
    }
    if (bad_any) return ECOZ2B200_ERR_ARG;   // caller formats on the host (inf / nan / huge values)
    // pass 2: the text, straight into a pinned host buffer
    const int slot = h_text_slot_;
    h_text_slot_ ^= 1;
    if (h_text_cap_[slot] < all) {
        if (h_text_[slot]) cudaFreeHost(h_text_[slot]);
        h_text_[slot] = nullptr;
        h_text_cap_[slot] = 0;
        const size_t want = all + all / 8 + 4096;
        CUDA_TRY(cudaMallocHost(&h_text_[slot], want));
        h_text_cap_[slot] = want;
    }
    size_t off = 0;
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        launch_format_write(s.idx, s.mindist, s.n, s.fmt_scratch, s.fmt_text, s.stream);
        ctr_.kernel_launches++;
        CUDA_TRY(cudaMemcpyAsync(h_text_[slot] + off, s.fmt_text, (size_t) totals[i], cudaMemcpyDeviceToHost, s.stream));
        off += (size_t) totals[i];
    }
    rc = sync_all();
    if (rc) return rc;
    CUDA_TRY(cudaGetLastError());
    if (text) *text = h_text_[slot];
    if (bytes) *bytes = all;
    return ECOZ2B200_OK;
}

int Engine::ensure_fmt_scratch(Shard &s) {
    if (s.fmt_scratch_cap >= s.n) return ECOZ2B200_OK;
    CUDA_TRY(cudaStreamSynchronize(s.stream));
    cudaFree(s.fmt_scratch);
    s.fmt_scratch = nullptr; s.fmt_scratch_cap = 0;
    CUDA_TRY(cudaMalloc(&s.fmt_scratch, format_scratch_bytes(s.n)));
    s.fmt_scratch_cap = s.n;
    return ECOZ2B200_OK;
}

int Engine::min_dists_text_device(std::vector<TextSegment> &segs) {
    segs.clear();
    const int rc = min_dists_text_device_impl(segs);
    if (rc) {   // hand back what was taken before the failure
        for (const TextSegment &g : segs) release_text(g);
        segs.clear();
    }
    return rc;
}

int Engine::min_dists_text_device_impl(std::vector<TextSegment> &segs) {
    static const bool timing = getenv("ECOZ2_VQ_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    const double t0 = now();
    int rc = ensure_mindist();
    if (rc) return rc;
    if (timing) { sync_all(); fprintf(stderr, "[timing]   text: minDist %.2f ms\n", (now() - t0) * 1e3); }
    const double t1 = now();
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        rc = ensure_fmt_scratch(s);
        if (rc) return rc;
        launch_format_lengths(s.idx, s.mindist, s.n, s.fmt_scratch, s.stream);
        ctr_.kernel_launches += 2;
    }
    std::vector<unsigned long long> totals(shards_.size(), 0);
    bool bad_any = false;
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        const unsigned long long *d_total;
        const int *d_bad;
        format_result_ptrs(s.fmt_scratch, s.n, &d_total, &d_bad);
        int bad = 0;
        CUDA_TRY(cudaMemcpyAsync(&totals[i], d_total, sizeof(unsigned long long), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaMemcpyAsync(&bad, d_bad, sizeof(int), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        bad_any = bad_any || bad != 0;
    }
    if (bad_any) return ECOZ2B200_ERR_ARG;   // caller formats on the host (inf / nan / huge values)
    if (timing) fprintf(stderr, "[timing]   text: lengths + scan + read-back %.2f ms\n", (now() - t1) * 1e3);
    const double t2 = now();
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        const size_t need = (size_t) totals[i];
        // First text of a run: allocate the pool for all the stages to come NOW, while no writer
        // thread is active -- cudaMalloc next to another thread's copies was measured to stall both
        // for tens to hundreds of milliseconds.  Lines grow from "0,0.123456" to "4095,0.123456":
        // half as much again covers it (a text that still does not fit takes the slow path below).
        if (s.text_pool.empty()) {
            int stages = 1;
            for (int k = K_ < 1 ? 1 : K_; 2 * k <= max_M_; k *= 2) stages++;
            if (stages > kTextPoolMax) stages = kTextPoolMax;
            const size_t cap = need + need / 2 + 4096;
            size_t free_b = 0, total_b = 0;
            if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) free_b = 0;
            while (stages > 2 && (size_t) stages * cap > free_b / 4) stages--;
            std::lock_guard<std::mutex> lk(text_mu_);
            for (int b = 0; b < stages; b++) {
                Shard::TextBuf tb;
                if (cudaMalloc(&tb.d, cap) != cudaSuccess) { cudaGetLastError(); break; }
                tb.cap = cap;
                if (cudaEventCreateWithFlags(&tb.ready, cudaEventDisableTiming) != cudaSuccess) {
                    cudaFree(tb.d);
                    break;
                }
                s.text_pool.push_back(tb);
            }
            if (timing)
                fprintf(stderr, "[timing]   text: pool of %zu x %.0f MB allocated, %.2f ms since the lengths\n",
                        s.text_pool.size(), cap / 1e6, (now() - t2) * 1e3);
        }
        // a free buffer of the pool that is large enough; else a new one; else (pool full or out of
        // memory) wait for a writer to hand one back
        int slot = -1;
        {
            std::unique_lock<std::mutex> lk(text_mu_);
            for (;;) {
                int small_free = -1, busy = 0;
                for (size_t b = 0; b < s.text_pool.size() && slot < 0; b++) {
                    Shard::TextBuf &tb = s.text_pool[b];
                    if (tb.busy) { busy++; continue; }
                    if (tb.cap >= need) slot = (int) b;
                    else small_free = (int) b;
                }
                if (slot >= 0) break;
                const size_t want = need + need / 8 + 4096;
                if (small_free >= 0 || (int) s.text_pool.size() < kTextPoolMax) {
                    if (small_free < 0) {
                        s.text_pool.emplace_back();
                        small_free = (int) s.text_pool.size() - 1;
                    }
                    Shard::TextBuf &tb = s.text_pool[small_free];
                    if (tb.d) { cudaFree(tb.d); tb.d = nullptr; tb.cap = 0; }
                    if (!tb.ready && cudaEventCreateWithFlags(&tb.ready, cudaEventDisableTiming) != cudaSuccess) {
                        set_last_error("min_dists_text_device: cudaEventCreate failed");
                        return ECOZ2B200_ERR_CUDA;
                    }
                    if (cudaMalloc(&tb.d, want) == cudaSuccess) {
                        tb.cap = want;
                        slot = small_free;
                        break;
                    }
                    cudaGetLastError();   // out of memory: fall through and wait for a busy buffer
                    tb.d = nullptr;
                    if (busy == 0) {
                        set_last_error("min_dists_text_device: out of device memory for the report text");
                        return ECOZ2B200_ERR_CUDA;
                    }
                }
                text_cv_.wait(lk);
            }
            s.text_pool[slot].busy = true;
        }
        Shard::TextBuf &tb = s.text_pool[slot];
        launch_format_write(s.idx, s.mindist, s.n, s.fmt_scratch, tb.d, s.stream);
        ctr_.kernel_launches++;
        CUDA_TRY(cudaEventRecord(tb.ready, s.stream));
        segs.push_back(TextSegment{(int) i, slot, s.device, tb.d, need, tb.ready});
    }
    if (timing) fprintf(stderr, "[timing]   text: buffer + write launch %.2f ms\n", (now() - t2) * 1e3);
    CUDA_TRY(cudaGetLastError());
    return ECOZ2B200_OK;
}

void Engine::release_text(const TextSegment &seg) {
    {
        std::lock_guard<std::mutex> lk(text_mu_);
        shards_[seg.shard].text_pool[seg.slot].busy = false;
    }
    text_cv_.notify_all();
}

// ---------------------------------------------------------------------------------
// LBG driver (vq_learn_par.c:42-163):
//   - DDprv starts at DBL_MAX and is NOT reset when the codebook grows;
//   - a size is done when pass > 0 && (DDprv - DD) / DD < eps;
//   - the centroid update runs before the convergence test: the saved codebook is
//     post-update, the reported average distortion pre-update;
//   - sigma / inertia use the post-update codebook.
int Engine::learn(double eps, int stop_M, ecoz2b200_stage_cb on_stage, ecoz2b200_pass_cb on_pass,
                  void *ctx) {
    if (!committed_ || K_ < 1 || n_global_ < 1) {
        set_last_error("learn: frames not committed or no codebook");
        return ECOZ2B200_ERR_ARG;
    }
    if (stop_M > max_M_) stop_M = max_M_;
    std::vector<double> h_refl, h_raas, h_discel;
    std::vector<int32_t> h_cardd;
    double DDprv = DBL_MAX;
    int pass = 0;
    static const bool timing = getenv("ECOZ2_VQ_TIMING") != nullptr;
    double tm_pass = 0;
    double t_stage = now_s();
    Counters at_stage_start = ctr_;
    for (;;) {
        const double t_pass = now_s();
        double DD = 0;
        h_cardd.resize(K_);
        h_discel.resize(K_);
        int rc = kmeans_pass(nullptr, h_cardd.data(), h_discel.data(), &DD);
        if (rc) return rc;
        tm_pass += now_s() - t_pass;
        const double avg = DD / (double) n_global_;
        if (on_pass) {
            int empty = 0;
            for (int i = 0; i < K_; i++) empty += h_cardd[i] == 0;
            on_pass(ctx, K_, pass, DD, now_s() - t_pass, empty);
        }

        if (pass > 0 && ((DDprv - DD) / DD) < eps) {
            const double tq0 = now_s();
            h_refl.resize((size_t) K_ * D_);
            h_raas.resize((size_t) K_ * D_);
            rc = codebook_get(h_refl.data(), h_raas.data());
            if (rc) return rc;
            const double tq1 = now_s();
            double sig = 0, inert = 0;
            rc = sigma(avg, &sig);
            if (rc) return rc;
            const double tq2 = now_s();
            rc = inertia(&inert);
            if (rc) return rc;
            const double tq3 = now_s();
            ecoz2b200_stage_info info;
            info.M = K_;
            info.passes = pass + 1;
            info.avgDistortion = avg;
            info.sigma = sig;
            info.inertia = inert;
            info.reflections = h_refl.data();
            info.raas = h_raas.data();
            info.cardd = h_cardd.data();
            info.discel = h_discel.data();
            info.stage_seconds = now_s() - t_stage;
            info.assign_ms = ctr_.assign_ms - at_stage_start.assign_ms;
            info.evals = (ctr_.evals - at_stage_start.evals);
            if (on_stage) on_stage(ctx, &info);
            const double tq4 = now_s();
            if (timing)
                fprintf(stderr, "[timing] learn M=%d: %d passes %.2f ms | codebook read %.2f, sigma %.2f, inertia %.2f, "
                                "stage callback %.2f ms\n", K_, pass + 1, tm_pass * 1e3, (tq1 - tq0) * 1e3,
                        (tq2 - tq1) * 1e3, (tq3 - tq2) * 1e3, (tq4 - tq3) * 1e3);
            tm_pass = 0;
            if (K_ >= stop_M) break;
            pass = 0;
            rc = codebook_grow();
            if (rc) return rc;
            t_stage = now_s();
            at_stage_start = ctr_;
        } else {
            pass++;
        }
        DDprv = DD;
    }
    return ECOZ2B200_OK;
}

// ---------------------------------------------------------------------------------
// quantize (quantize.c:8-34)
int Engine::quantize_resident(uint16_t *seq_out, double *avg) {
    if (!committed_ || K_ < 1) {
        set_last_error("quantize_resident: frames not committed or no codebook");
        return ECOZ2B200_ERR_ARG;
    }
    const PackedLayout L{K_, D_};
    const bool reduce = shards_.size() > 1 || multi_process_;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaEventRecord(s.ev[0], s.stream));
        CUDA_TRY(cudaMemsetAsync(s.packed + L.dd(), 0, 2 * sizeof(double), s.stream));
        pass_kernels(s, s.x64, s.x32, s.n, s.idx_tmp, s.sc, nullptr, s.seq, false);
        CUDA_TRY(cudaEventRecord(s.ev[2], s.stream));
        if (seq_out)
            CUDA_TRY(cudaMemcpyAsync(seq_out + s.frame_offset, s.seq, sizeof(uint16_t) * (size_t) s.n,
                                     cudaMemcpyDeviceToHost, s.stream));
    }
    int rc = allreduce_range(shards_, L.dd(), 2, reduce);
    if (rc) return rc;
    double tail[2] = {0, 0};
    Shard &s0 = shards_[0];
    CUDA_TRY(cudaSetDevice(s0.device));
    CUDA_TRY(cudaMemcpyAsync(tail, s0.packed + L.dd(), 2 * sizeof(double), cudaMemcpyDeviceToHost, s0.stream));
    rc = sync_all();
    if (rc) return rc;
    CUDA_TRY(cudaGetLastError());
    harvest_times(false);
    if (avg) *avg = n_global_ > 0 ? tail[0] / (double) n_global_ : 0.0;
    return ECOZ2B200_OK;
}

// ---------------------------------------------------------------------------------
// R codebooks at once (cbook_quantize for a SeqProvider / vq_classify): see vq_engine.h, vq_models.cu
int Engine::quantize_models(const double *reflections, int R, int K, const long *seg_offsets, int num_segments,
                            uint16_t *seq_out, double *dists_out, int *one_sweep) {
    if (one_sweep) *one_sweep = 0;
    if (!committed_ || multi_process_ || !reflections || R < 1 || K < 1 || K > 65536 || num_segments < 0 ||
        (num_segments > 0 && !seg_offsets)) {
        set_last_error("quantize_models: frames not committed, one process per GPU, or bad argument");
        return ECOZ2B200_ERR_ARG;
    }
    if (num_segments > 0 && (seg_offsets[0] != 0 || seg_offsets[num_segments] != n_total_)) {
        set_last_error("quantize_models: segments must cover the resident frames");
        return ECOZ2B200_ERR_ARG;
    }
    for (int f = 0; f < num_segments; f++)
        if (seg_offsets[f + 1] < seg_offsets[f]) {
            set_last_error("quantize_models: segment offsets must not decrease");
            return ECOZ2B200_ERR_ARG;
        }
    static const bool timing = getenv("ECOZ2_VQ_TIMING") != nullptr;
    auto now = [] { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); };
    double t_mark = now();
    auto mark = [&](const char *what) {
        if (!timing) return;
        sync_all();
        const double t = now();
        fprintf(stderr, "[timing]   quantize_models: %s %.2f ms\n", what, (t - t_mark) * 1e3);
        t_mark = t;
    };
    if (mbuf_.size() != shards_.size()) mbuf_.resize(shards_.size());
    // one sweep: the concatenation padded to whole 256-column tiles with copies of the first model
    long r_pad = R;
    while (((long) r_pad * K) % 256 != 0) r_pad++;
    const long cols = r_pad * K;
    const bool fused = mode_ == ECOZ2B200_MODE_TENSOR && assign_tc_models_supported(K, D_, cols) &&
                       cols <= (1L << 20) && !getenv("ECOZ2_VQ_MODELS_SERIAL");
    if (!fused && K > max_M_) {
        set_last_error("quantize_models: codebook larger than the engine's max_M");
        return ECOZ2B200_ERR_ARG;
    }
    const float err_tc = (float) (kappa_tc_ * ldexp(1.0, -24));
    const size_t cb_vals = (size_t) K * D_;
    const long kChunkFrames = 1024;
    std::vector<std::vector<long>> h_chunks(shards_.size());
    std::vector<std::vector<int>> h_first(shards_.size());
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        MBuf &b = mbuf_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        // ---- buffers
        const size_t idx_items = (size_t) r_pad * (size_t) s.n;
        if (b.idx_items < idx_items) {
            CUDA_TRY(cudaStreamSynchronize(s.stream));
            cudaFree(b.idx); b.idx = nullptr; b.idx_items = 0;
            CUDA_TRY(cudaMalloc(&b.idx, sizeof(int32_t) * idx_items));
            b.idx_items = idx_items;
        }
        if (b.cols < cols) {
            CUDA_TRY(cudaStreamSynchronize(s.stream));
            cudaFree(b.refl); cudaFree(b.raas); cudaFree(b.w64); cudaFree(b.w32); cudaFree(b.w1); cudaFree(b.w1max);
            cudaFree(b.wtc);
            b = MBuf{.idx = b.idx, .idx_items = b.idx_items, .seq = b.seq, .seq_items = b.seq_items,
                     .chunks = b.chunks, .seg_first = b.seg_first, .partial = b.partial, .sums = b.sums,
                     .chunk_cap = b.chunk_cap, .seg_cap = b.seg_cap, .partial_cap = b.partial_cap,
                     .sums_cap = b.sums_cap};
            CUDA_TRY(cudaMalloc(&b.refl, sizeof(double) * (size_t) cols * D_));
            CUDA_TRY(cudaMalloc(&b.raas, sizeof(double) * (size_t) cols * D_));
            CUDA_TRY(cudaMalloc(&b.w64, sizeof(double) * (size_t) ((cols + 63) / 64) * 64 * D_));
            CUDA_TRY(cudaMalloc(&b.w32, sizeof(float) * (size_t) ((cols + 127) / 128) * 128 * D_));
            CUDA_TRY(cudaMalloc(&b.w1, sizeof(float) * (size_t) ((cols + 127) / 128) * 128));
            CUDA_TRY(cudaMalloc(&b.w1max, sizeof(float) * 4));
            if (assign_tc_supported((int) cols, D_))
                CUDA_TRY(cudaMalloc(&b.wtc, sizeof(float) * tc_codebook_floats((int) cols, D_)));
            b.cols = cols;
        }
        // ---- codebooks: reflections -> raas for all models (padding = copies of model 0)
        CUDA_TRY(cudaMemcpyAsync(b.refl, reflections, sizeof(double) * cb_vals * R, cudaMemcpyHostToDevice, s.stream));
        for (long m = R; m < r_pad; m++)
            CUDA_TRY(cudaMemcpyAsync(b.refl + cb_vals * m, b.refl, sizeof(double) * cb_vals, cudaMemcpyDeviceToDevice,
                                     s.stream));
        launch_refl_to_raas(b.refl, b.raas, (int) cols, D_, s.stream);
        ctr_.kernel_launches++;
    }
    // ---- chunk table per shard for the per-(model, file) sums (chunks never cross a segment)
    std::vector<double> sums((size_t) num_segments * R, 0.0);
    std::vector<size_t> n_chunks_of(shards_.size(), 0);
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        MBuf &b = mbuf_[i];
        if (s.n <= 0 || num_segments == 0 || !dists_out) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        std::vector<long> &ch = h_chunks[i];
        std::vector<int> &first = h_first[i];
        const long lo = s.frame_offset, hi = s.frame_offset + s.n;
        for (int f = 0; f < num_segments; f++) {
            first.push_back((int) (ch.size() / 2));
            const long a = seg_offsets[f] > lo ? seg_offsets[f] : lo;
            const long e = seg_offsets[f + 1] < hi ? seg_offsets[f + 1] : hi;
            for (long c = a; c < e; c += kChunkFrames) {
                ch.push_back(c - lo);
                ch.push_back(e - c < kChunkFrames ? e - c : kChunkFrames);
            }
        }
        first.push_back((int) (ch.size() / 2));
        const size_t n_chunks = ch.size() / 2;
        n_chunks_of[i] = n_chunks;
        if (n_chunks == 0) continue;
        if (b.chunk_cap < n_chunks || b.seg_cap < (size_t) num_segments + 1 || b.partial_cap < n_chunks * R ||
            b.sums_cap < (size_t) num_segments * R) {
            CUDA_TRY(cudaStreamSynchronize(s.stream));
            cudaFree(b.chunks); cudaFree(b.seg_first); cudaFree(b.partial); cudaFree(b.sums);
            b.chunks = nullptr; b.seg_first = nullptr; b.partial = nullptr; b.sums = nullptr;
            CUDA_TRY(cudaMalloc(&b.chunks, sizeof(long) * 2 * n_chunks));
            CUDA_TRY(cudaMalloc(&b.seg_first, sizeof(int) * ((size_t) num_segments + 1)));
            CUDA_TRY(cudaMalloc(&b.partial, sizeof(double) * n_chunks * R));
            CUDA_TRY(cudaMalloc(&b.sums, sizeof(double) * (size_t) num_segments * R));
            b.chunk_cap = n_chunks; b.seg_cap = (size_t) num_segments + 1; b.partial_cap = n_chunks * R;
            b.sums_cap = (size_t) num_segments * R;
        }
        CUDA_TRY(cudaMemcpyAsync(b.chunks, ch.data(), sizeof(long) * ch.size(), cudaMemcpyHostToDevice, s.stream));
        CUDA_TRY(cudaMemcpyAsync(b.seg_first, first.data(), sizeof(int) * first.size(), cudaMemcpyHostToDevice, s.stream));
    }
    mark("buffers + codebooks");
    if (fused) {
        for (size_t i = 0; i < shards_.size(); i++) {
            Shard &s = shards_[i];
            MBuf &b = mbuf_[i];
            if (s.n <= 0) continue;
            CUDA_TRY(cudaSetDevice(s.device));
            launch_prepare_codebook(b.raas, (int) cols, D_, b.w32, b.w64, b.w1, b.w1max, s.stream);
            launch_prepare_codebook_tc(b.raas, (int) cols, D_, b.wtc, b.w1max, err_tc, s.stream);
            CUDA_TRY(cudaMemsetAsync(s.sc.counters, 0, 4 * sizeof(int32_t), s.stream));
            debug_poison(s);
            const long seg_cap = s.sc.cap < (1L << 30) ? s.sc.cap : (1L << 30);
            launch_assign_tc_models(s.x32, s.n, D_, b.wtc, K, cols, s.sc, err_tc, b.idx, s.n, s.sc.min_key,
                                    (int) seg_cap, s.num_sms, s.stream, frame_operands(s, err_tc));
            mark("search");
            launch_models_recheck(s.x64, D_, b.raas, K, s.sc.min_key, s.sc.counters, (int) seg_cap, b.idx, s.n,
                                  s.num_sms, s.stream);
            vq_stat_add<<<1, 32, 0, s.stream>>>(s.sc.counters, s.dstat + 3);
            ctr_.kernel_launches += 5;
            ctr_.evals += (double) s.n * (double) R * (double) K;
        }
        // more pairs inside the bound than the key list holds: cannot happen short of degenerate
        // codebooks (every frame of every model); then the serial route is the answer
        for (size_t i = 0; i < shards_.size(); i++) {
            Shard &s = shards_[i];
            if (s.n <= 0) continue;
            CUDA_TRY(cudaSetDevice(s.device));
            int32_t fc = 0;
            CUDA_TRY(cudaMemcpyAsync(&fc, s.sc.counters, sizeof(int32_t), cudaMemcpyDeviceToHost, s.stream));
            CUDA_TRY(cudaStreamSynchronize(s.stream));
            if ((long) fc > s.sc.cap) {
                set_last_error("quantize_models: too many ambiguous (model, frame) pairs for one sweep");
                return ECOZ2B200_ERR_ARG;
            }
        }
        if (one_sweep) *one_sweep = 1;
    } else {
        for (int m = 0; m < R; m++) {
            int rc = codebook_set_reflections(reflections + cb_vals * m, K);
            if (rc) return rc;
            const PackedLayout L{K_, D_};
            for (size_t i = 0; i < shards_.size(); i++) {
                Shard &s = shards_[i];
                if (s.n <= 0) continue;
                CUDA_TRY(cudaSetDevice(s.device));
                CUDA_TRY(cudaMemsetAsync(s.packed + L.dd(), 0, 2 * sizeof(double), s.stream));
                pass_kernels(s, s.x64, s.x32, s.n, mbuf_[i].idx + (size_t) m * s.n, s.sc, s.mindist, nullptr, false);
                // the pass has left ddmin - 1 of every frame in s.mindist: per-chunk sums for this model
                if (n_chunks_of[i] > 0) {
                    launch_chunk_sums(s.mindist, mbuf_[i].chunks, (int) n_chunks_of[i], R, m, mbuf_[i].partial, s.stream);
                    ctr_.kernel_launches++;
                }
            }
        }
    }
    mark("search + re-evaluation");
    // ---- distortions per (model, segment)
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        MBuf &b = mbuf_[i];
        if (n_chunks_of[i] == 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        if (fused) {
            launch_models_distortion(s.x64, D_, b.raas, K, R, b.idx, s.n, b.chunks, (int) n_chunks_of[i], b.partial,
                                     s.num_sms, s.stream);
            ctr_.kernel_launches++;
        }
        launch_models_segments(b.partial, b.seg_first, num_segments, R, b.sums, s.stream);
        ctr_.kernel_launches++;
    }
    mark("distortions");
    // ---- results to the host
    std::vector<std::vector<double>> h_sums(shards_.size());
    for (size_t i = 0; i < shards_.size(); i++) {
        Shard &s = shards_[i];
        MBuf &b = mbuf_[i];
        if (s.n <= 0) continue;
        CUDA_TRY(cudaSetDevice(s.device));
        if (dists_out && num_segments > 0 && !h_chunks[i].empty()) {
            h_sums[i].resize((size_t) num_segments * R);
            CUDA_TRY(cudaMemcpyAsync(h_sums[i].data(), b.sums, sizeof(double) * h_sums[i].size(), cudaMemcpyDeviceToHost,
                                     s.stream));
        }
        if (seq_out) {
            const size_t items = (size_t) R * (size_t) s.n;
            if (b.seq_items < items) {
                CUDA_TRY(cudaStreamSynchronize(s.stream));
                cudaFree(b.seq); b.seq = nullptr; b.seq_items = 0;
                CUDA_TRY(cudaMalloc(&b.seq, sizeof(uint16_t) * items));
                b.seq_items = items;
            }
            launch_idx_to_u16(b.idx, b.seq, items, s.stream);
            ctr_.kernel_launches++;
            for (int m = 0; m < R; m++)
                CUDA_TRY(cudaMemcpyAsync(seq_out + (size_t) m * n_total_ + s.frame_offset, b.seq + (size_t) m * s.n,
                                         sizeof(uint16_t) * (size_t) s.n, cudaMemcpyDeviceToHost, s.stream));
        }
    }
    int rc = sync_all();
    if (rc) return rc;
    CUDA_TRY(cudaGetLastError());
    if (dists_out) {
        for (size_t i = 0; i < shards_.size(); i++)
            for (size_t k = 0; k < h_sums[i].size(); k++) sums[k] += h_sums[i][k];
        for (int m = 0; m < R; m++)
            for (int f = 0; f < num_segments; f++) {
                const long T = seg_offsets[f + 1] - seg_offsets[f];
                dists_out[(size_t) m * num_segments + f] = T > 0 ? sums[(size_t) f * R + m] / (double) T : 0.0;
            }
    }
    return ECOZ2B200_OK;
}

int Engine::quantize_host(const double *x, long T, uint16_t *seq_out, double *avg) {
    if (K_ < 1 || T < 0 || (T > 0 && !x)) {
        set_last_error("quantize_host: no codebook or bad argument");
        return ECOZ2B200_ERR_ARG;
    }
    const PackedLayout L{K_, D_};
    const long S = (long) shards_.size();
    const long chunk_cap = 1L << 20;  // frames per device chunk
    const long want = T < chunk_cap ? (T > 0 ? T : 1) : chunk_cap;
    for (long i = 0; i < S; i++) {
        Shard &s = shards_[i];
        CUDA_TRY(cudaSetDevice(s.device));
        if (!s.copy_stream) {
            CUDA_TRY(cudaStreamCreateWithFlags(&s.copy_stream, cudaStreamNonBlocking));
            for (int k = 0; k < 2; k++) CUDA_TRY(cudaEventCreateWithFlags(&s.ev_chunk[k], cudaEventDisableTiming));
        }
        for (int b = 0; b < 2; b++) {
            if (!s.ev_qfree[b]) CUDA_TRY(cudaEventCreateWithFlags(&s.ev_qfree[b], cudaEventDisableTiming));
            QBuf &q = qbuf_[2 * i + b];
            if (q.cap < want) {
                CUDA_TRY(cudaStreamSynchronize(s.stream));
                cudaFree(q.x64); cudaFree(q.x32); cudaFree(q.idx); cudaFree(q.sc_mem); cudaFree(q.seq);
                q = QBuf();
                CUDA_TRY(cudaMalloc(&q.x64, sizeof(double) * (size_t) want * D_));
                CUDA_TRY(cudaMalloc(&q.x32, sizeof(float) * (size_t) want * D_));
                CUDA_TRY(cudaMalloc(&q.idx, sizeof(int32_t) * (size_t) want));
                CUDA_TRY(cudaMalloc(&q.sc_mem, scratch_bytes(want)));
                scratch_carve(q.sc_mem, want, &q.sc);
                CUDA_TRY(cudaMalloc(&q.seq, sizeof(uint16_t) * (size_t) want));
                q.cap = want;
            }
            // whatever s.stream still does with the buffer (an earlier call) comes first
            CUDA_TRY(cudaEventRecord(s.ev_qfree[b], s.stream));
        }
        CUDA_TRY(cudaMemsetAsync(s.packed + L.dd(), 0, 2 * sizeof(double), s.stream));
    }
    // contiguous chunks dealt round-robin to the devices; a chunk is uploaded on the device's
    // copy stream into one of its two buffers while the previous chunk is searched:
    // H2D | f32 copy -> search -> FP64 distortion + symbols -> D2H
    long pos = 0;
    int turn = 0;
    std::vector<int> used((size_t) S, 0);
    while (pos < T) {
        Shard &s = shards_[turn];
        const int b = used[(size_t) turn]++ & 1;
        QBuf &q = qbuf_[2 * (size_t) turn + b];
        const long n = T - pos < q.cap ? T - pos : q.cap;
        CUDA_TRY(cudaSetDevice(s.device));
        CUDA_TRY(cudaStreamWaitEvent(s.copy_stream, s.ev_qfree[b], 0));
        CUDA_TRY(cudaMemcpyAsync(q.x64, x + (size_t) pos * D_, sizeof(double) * (size_t) n * D_,
                                 cudaMemcpyHostToDevice, s.copy_stream));
        CUDA_TRY(cudaEventRecord(s.ev_chunk[b], s.copy_stream));
        CUDA_TRY(cudaStreamWaitEvent(s.stream, s.ev_chunk[b], 0));
        if (!small_path()) {
            launch_convert_frames(q.x64, q.x32, n, D_, s.stream);
            ctr_.kernel_launches++;
        }
        pass_kernels(s, q.x64, q.x32, n, q.idx, q.sc, nullptr, q.seq, false);
        if (seq_out)
            CUDA_TRY(cudaMemcpyAsync(seq_out + pos, q.seq, sizeof(uint16_t) * (size_t) n,
                                     cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaEventRecord(s.ev_qfree[b], s.stream));
        pos += n;
        turn = (turn + 1) % (int) S;
    }
    double total = 0;
    for (Shard &s : shards_) {
        CUDA_TRY(cudaSetDevice(s.device));
        double tail[2] = {0, 0};
        CUDA_TRY(cudaMemcpyAsync(tail, s.packed + L.dd(), 2 * sizeof(double), cudaMemcpyDeviceToHost, s.stream));
        CUDA_TRY(cudaStreamSynchronize(s.stream));
        total += tail[0];
    }
    CUDA_TRY(cudaGetLastError());
    if (avg) *avg = T > 0 ? total / (double) T : 0.0;
    return ECOZ2B200_OK;
}

}  // namespace vqb200
