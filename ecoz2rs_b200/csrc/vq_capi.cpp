// vq_capi.cpp -- the extern "C" surface declared in include/ecoz2_vq_b200.h.
//
// PART 1: drop-in `ecoz2_vq_learn`, `ecoz2_vq_learn_using_base_codebook`,
//         `ecoz2_vq_quantize` (reference ecoz2/src/ecoz2/ecoz2.c:109-164 ->
//         ecoz2/src/vq/vq_learn.c, vq_learn.i:161-247, vq_quantize.c:13-111).
// PART 2: thin forwards to vqb200::Engine.
#include <cfloat>
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "vq_engine.h"
#include "vq_io.h"
#include "vq_hostio.h"
#include "vq_nccl.h"

using vqb200::Engine;

struct ecoz2b200_engine {
    Engine impl;
    ecoz2b200_engine(const int *d, int n, int P, int M) : impl(d, n, P, M) {}
};

namespace {

double wall_s() {
    using namespace std::chrono;
    return duration<double>(steady_clock::now().time_since_epoch()).count();
}

// ECOZ2_VQ_MODE=tensor|fp32|fp64 overrides the search arithmetic (default tensor; "tf32x3" = its old name)
static void apply_mode_env(vqb200::Engine &eng) {
    const char *m = getenv("ECOZ2_VQ_MODE");
    if (!m) return;
    if (strcmp(m, "fp64") == 0) eng.set_mode(ECOZ2B200_MODE_FP64);
    else if (strcmp(m, "fp32") == 0) eng.set_mode(ECOZ2B200_MODE_FP32);
    else if (strcmp(m, "tf32x3") == 0 || strcmp(m, "tensor") == 0) eng.set_mode(ECOZ2B200_MODE_TENSOR);
}

// ECOZ2_VQ_GPUS: unset/"" -> device 0; "all" -> every visible device; else "0,2,3"
std::vector<int> devices_from_env() {
    std::vector<int> devs;
    const char *env = getenv("ECOZ2_VQ_GPUS");
    int visible = ecoz2b200_device_count();
    if (!env || !*env) {
        devs.push_back(0);
    } else if (strcmp(env, "all") == 0) {
        for (int i = 0; i < visible; i++) devs.push_back(i);
    } else {
        std::string s(env);
        size_t pos = 0;
        while (pos < s.size()) {
            size_t comma = s.find(',', pos);
            if (comma == std::string::npos) comma = s.size();
            if (comma > pos) devs.push_back(atoi(s.substr(pos, comma - pos).c_str()));
            pos = comma + 1;
        }
    }
    if (devs.empty()) devs.push_back(0);
    return devs;
}

int env_int(const char *name, int dflt) {
    const char *v = getenv(name);
    return (v && *v) ? atoi(v) : dflt;
}

struct LearnCtx {
    Engine *eng;
    std::string class_name, prefix, cb_filename;
    vqb200::Report *report;
    vqb200::DeviceTextWriter *text_writer;
    void *target;
    vq_learn_callback_t callback;
    double DDprv = DBL_MAX;
    long tot_vecs;
    bool min_dists;
    double t_stage;
    int P;
    int rc = 0;
};

void on_pass(void *c, int M, int pass, double DD, double secs, int empty_cells) {
    LearnCtx *ctx = (LearnCtx *) c;
    // same progress lines as the reference (vq_learn_par.c:44,114-123)
    printf("(%d)", pass);
    printf("\n      (pass=%d) min processing took %.2fs\n", pass, secs);
    printf("\tDP=%g\tDDprv=%g\tDD=%g\t(DDprv-DD)/DD=%g\r", DD / ctx->tot_vecs, ctx->DDprv, DD,
           (ctx->DDprv - DD) / DD);
    fflush(stdout);
    if (empty_cells > 0)  // review_cells, vq_learn.i:108-119
        printf("\nWARN: review_cells: %d empty cell(s) for codebook size %d)\n", empty_cells, M);
    ctx->DDprv = DD;
}

void on_stage(void *c, const ecoz2b200_stage_info *info) {
    LearnCtx *ctx = (LearnCtx *) c;
    printf("\n      (%d-cbook) processing took %.2fs\n", info->M, wall_s() - ctx->t_stage);
    // codebook saved with reflections (vq_learn_par.c:133)
    if (vqb200::cbook_save(ctx->cb_filename.c_str(), ctx->class_name.c_str(), info->reflections,
                           info->M, ctx->P))
        fprintf(stderr, "error saving codebook %s\n", ctx->cb_filename.c_str());
    if (ctx->callback)
        ctx->callback(ctx->target, info->M, info->avgDistortion, info->sigma, info->inertia);
    std::vector<int32_t> cw;
    std::vector<double> md;
    const double t_rep0 = wall_s();
    bool text_done = false;
    if (ctx->min_dists) {
        // "<cbook>.min_dists.csv": formatted on the device (vq_report.cu); the text stays in HBM and
        // a pool of writer threads copies it out and writes the file while the next stages run
        // (vq_hostio.h: one file does not fill faster with more threads, several files do).
        std::vector<Engine::TextSegment> segs;
        if (ctx->text_writer && ctx->eng->min_dists_text_device(segs) == ECOZ2B200_OK) {
            ctx->text_writer->submit(ctx->cb_filename + ".min_dists.csv", "codeword,minDistortion\n",
                                     std::move(segs));
            text_done = true;
        } else {   // values the device formatter does not take (inf / nan): the host formatter
            cw.resize((size_t) ctx->tot_vecs);
            md.resize((size_t) ctx->tot_vecs);
            if (ctx->eng->last_assignment(cw.data(), md.data())) ctx->rc = ECOZ2B200_ERR_CUDA;
        }
    }
    const double t_rep1 = wall_s();
    const bool host_lines = ctx->min_dists && !text_done;
    ctx->report->cbook(ctx->cb_filename, info->passes, info->avgDistortion, info->sigma,
                       info->inertia, info->M, info->cardd, info->discel,
                       host_lines ? cw.data() : nullptr, host_lines ? md.data() : nullptr);
    if (getenv("ECOZ2_VQ_TIMING"))
        fprintf(stderr, "[timing] M=%d report text / read-back %.1f ms, report files %.1f ms\n", info->M,
                (t_rep1 - t_rep0) * 1e3, (wall_s() - t_rep1) * 1e3);
    printf("\n");
    // name of the next size (vq_learn_par.c:152-154)
    char buf[64];
    snprintf(buf, sizeof buf, "_M_%04d.cbook", info->M * 2);
    ctx->cb_filename = ctx->prefix + buf;
    ctx->t_stage = wall_s();
}

// vq_learn_ (vq_learn.i:161-247)
int learn_files(const double *base_refl, int base_K, int P, double eps, const char *class_name,
                const char *files[], int num_files, void *target, vq_learn_callback_t callback) {
    const double t_call = wall_s();
    const std::string class_dir = std::string("data/codebooks/") + class_name;
    vqb200::mk_dirs(class_dir);
    char epsbuf[64];
    snprintf(epsbuf, sizeof epsbuf, "%g", eps);
    const std::string prefix = class_dir + "/eps_" + epsbuf;

    // predictor headers first: total size fixes the GPU shard boundaries
    std::vector<int> Ts(num_files);
    long tot_vecs = 0;
    for (int i = 0; i < num_files; i++) {
        int p = 0, t = 0;
        if (vqb200::prd_info(files[i], &p, &t, nullptr) || p != P) {
            fprintf(stderr, "error loading predictor %s\n", files[i]);
            return 2;
        }
        Ts[i] = t;
        tot_vecs += t;
    }
    const int stop_M = env_int("ECOZ2_VQ_MAX_M", 4096);
    std::vector<int> devs = devices_from_env();
    std::unique_ptr<ecoz2b200_engine> h(
        new ecoz2b200_engine(devs.data(), (int) devs.size(), P, stop_M < 2 ? 2 : stop_M));
    Engine &eng = h->impl;
    if (!eng.ok()) return ECOZ2B200_ERR_CUDA;
    apply_mode_env(eng);
    int rc = eng.frames_reserve(tot_vecs);
    if (rc) return ECOZ2B200_ERR_CUDA;
    // streams + pinned chunks of the loader threads and, later, of the report writers (vq_hostio.h)
    // (small inputs: fewer threads, pinning memory costs more than it saves)
    const size_t total_bytes = (size_t) tot_vecs * (size_t) (P + 1) * sizeof(double);
    const int small_threads = (int) (total_bytes >> 25) + 1;             // one per 32 MB
    vqb200::IoLanes lanes(eng.devices(), small_threads < 8 ? small_threads : 0);
    if (lanes.ready() && tot_vecs > 0 && !getenv("ECOZ2_VQ_LOAD_SERIAL")) {
        const double t_load = wall_s();
        const int lr = vqb200::load_prd_frames(eng, lanes, files, Ts.data(), num_files, P);
        if (lr > 0) {
            fprintf(stderr, "error loading predictor %s\n", files[lr - 1]);
            return 2;
        }
        if (lr < 0) return ECOZ2B200_ERR_CUDA;
        if (getenv("ECOZ2_VQ_TIMING"))
            fprintf(stderr, "[timing] %ld frames of %d file(s) loaded in %.1f ms by %d thread(s)\n", tot_vecs,
                    num_files, (wall_s() - t_load) * 1e3, lanes.threads());
    } else {
        std::vector<double> buf;
        for (int i = 0; i < num_files; i++) {
            const size_t vals = (size_t) Ts[i] * (P + 1);
            if (buf.size() < vals) buf.resize(vals);
            if (vqb200::prd_read(files[i], buf.data(), (long) vals)) {
                fprintf(stderr, "error loading predictor %s\n", files[i]);
                return 2;
            }
            rc = eng.frames_append_host(buf.data(), Ts[i]);
            if (rc) return ECOZ2B200_ERR_CUDA;
            // the staging vector is reused: the (pageable) copy above is complete on return
            if (eng.frames_count() < 0) return 1;
        }
    }
    rc = eng.frames_commit();
    if (rc) return ECOZ2B200_ERR_CUDA;
    printf("%ld training vectors (ε=%g)\n", tot_vecs, eps);

    if (base_refl) {  // initial_codebook_from_base, vq_learn.i:75-82
        if (2 * base_K > (stop_M < 2 ? 2 : stop_M)) {
            fprintf(stderr, "ERROR: maximum codebook size reached: %d", stop_M);
            return 1;
        }
        rc = eng.codebook_set_reflections(base_refl, base_K);
        if (!rc) rc = eng.codebook_grow();
    } else {
        rc = eng.codebook_initial();
    }
    if (rc) return rc == ECOZ2B200_ERR_ARG ? 1 : ECOZ2B200_ERR_CUDA;

    vqb200::Report report;
    report.prepare(prefix + ".rpt", tot_vecs, eps);

    LearnCtx ctx;
    ctx.eng = &eng;
    ctx.class_name = class_name;
    ctx.prefix = prefix;
    ctx.report = &report;
    ctx.min_dists = env_int("ECOZ2_VQ_MIN_DISTS", 1) != 0;
    // writer threads for the device-formatted min_dists.csv (streams and pinned chunks are set up
    // here, before the run); ECOZ2_VQ_REPORT_HOST=1 keeps the host formatter
    std::unique_ptr<vqb200::DeviceTextWriter> text_writer;
    if (ctx.min_dists && !getenv("ECOZ2_VQ_REPORT_HOST")) {
        if (lanes.ready()) text_writer.reset(new vqb200::DeviceTextWriter(&eng, &lanes));
    }
    ctx.text_writer = text_writer.get();
    ctx.target = target;
    ctx.callback = callback;
    ctx.tot_vecs = tot_vecs;
    ctx.min_dists = env_int("ECOZ2_VQ_MIN_DISTS", 1) != 0;
    ctx.P = P;
    char buf[64];
    snprintf(buf, sizeof buf, "_M_%04d.cbook", eng.codebook_size());
    ctx.cb_filename = prefix + buf;
    printf("%s\n", ctx.cb_filename.c_str());
    printf("\n(devices=%d, B200 CUDA path)\n", (int) devs.size());
    const double t0 = wall_s();
    ctx.t_stage = t0;
    if (getenv("ECOZ2_VQ_TIMING"))
        fprintf(stderr, "[timing] %.1f ms from entry to the first pass\n", (t0 - t_call) * 1e3);
    rc = eng.learn(eps, stop_M, on_stage, on_pass, &ctx);
    printf("\nprocessing took %.2fs\n", wall_s() - t0);
    const double t_fin = wall_s();
    if (text_writer) text_writer->finish();   // the last stages' files
    if (getenv("ECOZ2_VQ_TIMING"))
        fprintf(stderr, "[timing] waited %.1f ms for the report writers after the last stage\n",
                (wall_s() - t_fin) * 1e3);
    report.close();
    if (getenv("ECOZ2_VQ_TIMING")) {   // where the rest of the wall time goes: tear-down in the open
        const double t1 = wall_s();
        text_writer.reset();
        h.reset();
        fprintf(stderr, "[timing] whole call %.1f ms so far; engine tear-down %.1f ms\n", (t1 - t_call) * 1e3,
                (wall_s() - t1) * 1e3);
    }
    if (rc) return ECOZ2B200_ERR_CUDA;
    return ctx.rc;
}

}  // namespace

int ecoz2b200_classify_entry(char **cb_filenames, int num_codebooks, char **prd_filenames,
                             int num_predictors, int show_ranked);   // vq_classify.cpp

// shared with vq_classify.cpp
void ecoz2b200_apply_mode_env(vqb200::Engine &eng) { apply_mode_env(eng); }
std::vector<int> ecoz2b200_devices_from_env() { return devices_from_env(); }

// ============================================================================ PART 1
extern "C" {

// The exported names below forward to these internal bodies: when the library is linked next to the
// reference's own ecoz2.c (integration/build.rs.patch), ecoz2_vq_learn resolves to THAT file's forward,
// which calls vq_learn -- which must not call ecoz2_vq_learn back through the dynamic symbol table.
static int learn_entry(int prediction_order, double epsilon, const char *codebook_class_name,
                       const char *predictor_filenames[], int num_predictors, void *callback_target,
                       vq_learn_callback_t callback) {
    vqb200::DeviceGuard device_guard_;
    // vq_learn.c:17-19
    printf("\nCodebook generation:\n\n");
    printf("prediction_order=%d class='%s'  epsilon=%g\n\n", prediction_order, codebook_class_name,
           epsilon);
    return learn_files(nullptr, 0, prediction_order, epsilon, codebook_class_name,
                       predictor_filenames, num_predictors, callback_target, callback);
}

int ecoz2_vq_learn(int prediction_order, double epsilon, const char *codebook_class_name,
                   const char *predictor_filenames[], int num_predictors, void *callback_target,
                   vq_learn_callback_t callback) {
    return learn_entry(prediction_order, epsilon, codebook_class_name, predictor_filenames, num_predictors,
                       callback_target, callback);
}

static int learn_base_entry(char *base_codebook, double epsilon, const char *predictor_filenames[],
                            int num_predictors, void *callback_target, vq_learn_callback_t callback) {
    vqb200::DeviceGuard device_guard_;
    // vq_learn.c:44-60
    int P = 0, M = 0;
    char cls[vqb200::kMaxClassNameLen];
    if (vqb200::cbook_info(base_codebook, &P, &M, cls)) {
        fprintf(stderr, "error loading base codebook %s.\n", base_codebook);
        return 2;
    }
    std::vector<double> refl((size_t) M * (P + 1));
    if (vqb200::cbook_read(base_codebook, refl.data(), (long) refl.size())) {
        fprintf(stderr, "error loading base codebook %s.\n", base_codebook);
        return 2;
    }
    printf("\nCodebook generation:\n\n");
    printf("base_codebook: %s  num_vecs=%d prediction_order=%d class='%s'  epsilon=%g\n\n",
           base_codebook, M, P, cls, epsilon);
    return learn_files(refl.data(), M, P, epsilon, cls, predictor_filenames, num_predictors,
                       callback_target, callback);
}

int ecoz2_vq_learn_using_base_codebook(char *base_codebook, double epsilon,
                                       const char *predictor_filenames[], int num_predictors,
                                       void *callback_target, vq_learn_callback_t callback) {
    return learn_base_entry(base_codebook, epsilon, predictor_filenames, num_predictors, callback_target,
                            callback);
}

static int quantize_entry(const char *nom_raas, const char *predictor_filenames[], int num_predictors,
                          int show_filenames) {
    vqb200::DeviceGuard device_guard_;
    // vq_quantize.c:13-111
    int P = 0, M = 0;
    char cb_class[vqb200::kMaxClassNameLen];
    if (vqb200::cbook_info(nom_raas, &P, &M, cb_class)) {
        printf("error loading codebook (%d).\n", vqb200::cb_errnum_last);
        return 1;
    }
    std::vector<double> refl((size_t) M * (P + 1));
    if (vqb200::cbook_read(nom_raas, refl.data(), (long) refl.size())) {
        printf("error loading codebook (%d).\n", vqb200::cb_errnum_last);
        return 1;
    }
    static const bool timing = getenv("ECOZ2_VQ_TIMING") != nullptr;
    double t_mark = wall_s();
    auto mark = [&](const char *what) {
        if (!timing) return;
        const double t = wall_s();
        fprintf(stderr, "[timing] vq_quantize: %s %.1f ms\n", what, (t - t_mark) * 1e3);
        t_mark = t;
    };
    std::vector<int> devs = devices_from_env();
    std::unique_ptr<ecoz2b200_engine> h(
        new ecoz2b200_engine(devs.data(), (int) devs.size(), P, M < 2 ? 2 : M));
    Engine &eng = h->impl;
    if (!eng.ok()) return ECOZ2B200_ERR_CUDA;
    apply_mode_env(eng);
    if (eng.codebook_set_reflections(refl.data(), M)) return ECOZ2B200_ERR_CUDA;
    printf("%s: %d symbols\n", nom_raas, M);
    mark("engine + codebook");

    // Headers first.  The reference works file by file and stops at the first one it cannot load
    // (vq_quantize.c:49-53), after having written the sequences of the files before it: same here.
    std::vector<int> Ts;
    int first_bad = -1;
    for (int i = 0; i < num_predictors; i++) {
        int p = 0, T = 0;
        if (vqb200::prd_info(predictor_filenames[i], &p, &T, nullptr) || p != P) {
            first_bad = i;
            break;
        }
        Ts.push_back(T);
    }
    int num_ok = (int) Ts.size();   // files [0, num_ok) are processed; shrinks if a file turns out truncated
    // All frames of a batch of files are resident at once: loaded by the reader threads of
    // vq_hostio.cpp, quantized in one sweep, the per-file distortions summed on the device
    // (Engine::quantize_models with one model).  A batch is bounded by the free device memory.
    size_t free_b = 0, total_b = 0;
    {
        cudaSetDevice(devs[0]);
        if (cudaMemGetInfo(&free_b, &total_b) != cudaSuccess) free_b = (size_t) 8 << 30;
    }
    const size_t per_frame = (size_t) (P + 1) * 12 + 64;            // f64 + f32 rows, per-frame outputs
    long budget = (long) (free_b / 2 / per_frame) * (long) devs.size();
    if (budget < (1L << 20)) budget = 1L << 20;
    std::unique_ptr<vqb200::IoLanes> lanes;

    int num_seqs = 0;
    double ddprm_total = 0;
    std::vector<uint16_t> seq_all;
    std::vector<double> dd;
    std::vector<long> seg;
    int done = 0;
    while (done < num_ok) {
        int hi = done;
        long frames = 0;
        while (hi < num_ok && (hi == done || frames + Ts[hi] <= budget)) frames += Ts[hi++];
        if (frames > 0) {
            if (!lanes) {
                const size_t bytes = (size_t) frames * (size_t) (P + 1) * sizeof(double);
                const int small_threads = (int) (bytes >> 25) + 1;
                lanes.reset(new vqb200::IoLanes(eng.devices(), small_threads < 8 ? small_threads : 0));
                if (!lanes->ready()) return ECOZ2B200_ERR_CUDA;
            }
            if (eng.frames_clear() || eng.frames_reserve(frames)) return ECOZ2B200_ERR_CUDA;
            const int lr = vqb200::load_prd_frames(eng, *lanes, predictor_filenames + done, Ts.data() + done, hi - done, P);
            if (lr < 0) return ECOZ2B200_ERR_CUDA;
            if (lr > 0) {   // shorter than its header says: everything before it is still written
                first_bad = done + lr - 1;
                num_ok = first_bad;   // re-run the batch up to that file
                continue;
            }
            if (eng.frames_commit()) return ECOZ2B200_ERR_CUDA;
            mark("frames loaded");
            seg.assign((size_t) (hi - done) + 1, 0);
            for (int i = done; i < hi; i++) seg[(size_t) (i - done) + 1] = seg[(size_t) (i - done)] + Ts[i];
            seq_all.resize((size_t) frames);
            dd.assign((size_t) (hi - done), 0.0);
            if (eng.quantize_models(refl.data(), 1, M, seg.data(), hi - done, seq_all.data(), dd.data(), nullptr))
                return ECOZ2B200_ERR_CUDA;
            mark("quantized, symbols on the host");
        }
        for (int i = done; i < hi; i++) {
        const char *prd_filename = predictor_filenames[i];
        const int T = Ts[i];
        const uint16_t *seq = seq_all.data() + (frames > 0 ? seg[(size_t) (i - done)] : 0);
        const double ddprm = frames > 0 ? dd[(size_t) (i - done)] : 0.0;
        ddprm_total += ddprm;
        num_seqs++;

        const std::string className = vqb200::get_class_name(prd_filename);
        const char *base_dir = "sequences";
        if (strstr(prd_filename, "TRAIN/")) base_dir = "sequences/TRAIN";
        else if (strstr(prd_filename, "TEST/")) base_dir = "sequences/TEST";
        const std::string base_dir_with_m = std::string(base_dir) + "/M" + std::to_string(M);
        const std::string seq_filename =
            vqb200::get_output_filename(prd_filename, base_dir_with_m, ".seq");
        if (vqb200::seq_save(seq_filename.c_str(), className.c_str(), seq, T, M))
            printf("error saving sequence\n");
        if (show_filenames)
            printf("\n%s className='%s' (%1.4g) = ", seq_filename.c_str(), className.c_str(), ddprm);
        if (className.empty()) printf("\nWARN%s NO className", seq_filename.c_str());
        if (show_filenames) {  // seq_show, symbol.c:84-98
            const int lim = 40, mrg = 10;
            printf("<(%d): ", T);
            const char *comma = "";
            for (int t = 0; t < T; t++) {
                if (T > lim && t >= mrg && t < T - mrg) {
                    printf(", ...");
                    t = T - mrg;
                }
                printf("%s%d", comma, seq[t]);
                comma = ", ";
            }
            printf(">\n");
        }
        }
        done = hi;
        mark(".seq files written");
    }
    if (first_bad >= 0) {
        fprintf(stderr, "%s: error loading predictor.\n", predictor_filenames[first_bad]);
        return 2;
    }
    if (num_seqs > 0)
        printf("\ntotal: %d sequences; total  average distortion = %f\n", num_seqs,
               ddprm_total / num_seqs);
    return 0;
}

// ---- the inner C API of the reference (ecoz2/src/include/vq.h:82-117), i.e. what its public
// forwards (ecoz2/src/ecoz2/ecoz2.c:109-193) and its C mains (x/vq.learn.c:107-128,
// x/vq.quantize.c:41, x/vq.classify.c) call.  Exporting these as well lets ecoz2.c and the mains
// link against this library UNMODIFIED once vq/vq_learn.c, vq_quantize.c, vq_classify.c (+ report.c,
// sigma.c, inertia.c) are dropped from the build.  `use_par` selects between the reference's
// serial and OpenMP loops; both are the one CUDA path here.
int ecoz2_vq_quantize(const char *nom_raas, const char *predictor_filenames[], int num_predictors,
                      int show_filenames) {
    return quantize_entry(nom_raas, predictor_filenames, num_predictors, show_filenames);
}

int vq_learn(int prediction_order, double epsilon, const char *codebook_class_name,
             const char *predictor_filenames[], int num_predictors, int use_par, void *callback_target,
             vq_learn_callback_t callback) {
    (void) use_par;
    return learn_entry(prediction_order, epsilon, codebook_class_name, predictor_filenames,
                       num_predictors, callback_target, callback);
}
int vq_learn_using_base_codebook(char *base_codebook, double epsilon, const char *predictor_filenames[],
                                 int num_predictors, int use_par, void *callback_target,
                                 vq_learn_callback_t callback) {
    (void) use_par;
    return learn_base_entry(base_codebook, epsilon, predictor_filenames, num_predictors, callback_target,
                            callback);
}
int vq_quantize(const char *nom_raas, const char *predictor_filenames[], int num_predictors,
                int show_filenames) {
    return quantize_entry(nom_raas, predictor_filenames, num_predictors, show_filenames);
}
int ecoz2_vq_classify(char **cb_filenames, int num_codebooks, char **prd_filenames, int num_predictors,
                      int show_ranked) {
    return ecoz2b200_classify_entry(cb_filenames, num_codebooks, prd_filenames, num_predictors, show_ranked);
}
int vq_classify(char **cb_filenames, int num_codebooks, char **prd_filenames, int num_predictors,
                int show_ranked) {
    return ecoz2b200_classify_entry(cb_filenames, num_codebooks, prd_filenames, num_predictors, show_ranked);
}

// quantize.c:8-34 behind the reference's own cbook_quantize / SeqProvider (vq/vq.c:154-157, 285-315)
double quantize(double *raas, int num_raas, ecoz2_ref_predictor *prd, uint16_t *seq) {
    static std::mutex mu;
    static ecoz2b200_engine *cached = nullptr;   // kept for the life of the process (never torn down at exit)
    static int cached_P = -1, cached_max = 0;
    static std::vector<double> cached_raas;       // the codebook currently staged on the device
    static std::vector<double> rows;
    if (!raas || !prd || num_raas < 1 || prd->P < 0) {
        fprintf(stderr, "quantize: bad argument\n");
        exit(1);
    }
    const int P = prd->P, T = prd->T, D = P + 1;
    if (T <= 0) return 0.0 / 0.0;                  // the reference divides by T
    std::lock_guard<std::mutex> lock(mu);
    vqb200::DeviceGuard device_guard_;
    if (!cached || cached_P != P || cached_max < num_raas) {
        delete cached;
        cached = nullptr;
        cached_raas.clear();
        std::vector<int> devs = devices_from_env();
        int cap = 2;
        while (cap < num_raas) cap *= 2;
        cached = new ecoz2b200_engine(devs.data(), (int) devs.size(), P, cap);
        if (!cached->impl.ok()) {
            fprintf(stderr, "quantize: %s\n", vqb200::get_last_error());
            exit(1);
        }
        apply_mode_env(cached->impl);
        cached_P = P;
        cached_max = cap;
    }
    Engine &eng = cached->impl;
    const size_t vals = (size_t) num_raas * D;
    if (cached_raas.size() != vals || memcmp(cached_raas.data(), raas, vals * sizeof(double)) != 0) {
        if (eng.codebook_set_raas(raas, num_raas)) {
            fprintf(stderr, "quantize: %s\n", vqb200::get_last_error());
            exit(1);
        }
        cached_raas.assign(raas, raas + vals);
    }
    if (rows.size() < (size_t) T * D) rows.resize((size_t) T * D);
    for (int t = 0; t < T; t++) memcpy(rows.data() + (size_t) t * D, prd->vectors[t], sizeof(double) * D);
    double ddprm = 0;
    if (eng.quantize_host(rows.data(), T, seq, &ddprm)) {
        fprintf(stderr, "quantize: %s\n", vqb200::get_last_error());
        exit(1);
    }
    return ddprm;
}

// ============================================================================ PART 2
const char *ecoz2b200_version(void) { return "ecoz2_vq_b200 0.1 (sm_100a)"; }
const char *ecoz2b200_last_error(void) { return vqb200::get_last_error(); }

int ecoz2b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) {
        cudaGetLastError();
        return 0;
    }
    return n;
}

ecoz2b200_engine *ecoz2b200_create(const int *devices, int num_devices, int P, int max_M) {
    vqb200::DeviceGuard device_guard_;
    ecoz2b200_engine *e = new ecoz2b200_engine(devices, num_devices, P, max_M);
    if (!e->impl.ok()) {
        delete e;
        return nullptr;
    }
    return e;
}
void ecoz2b200_destroy(ecoz2b200_engine *e) {
    vqb200::DeviceGuard device_guard_;
    delete e;
}

#define ENG(e)                                  \
    if (!(e)) return ECOZ2B200_ERR_ARG;         \
    vqb200::DeviceGuard device_guard_;          \
    Engine &eng = (e)->impl

int ecoz2b200_set_mode(ecoz2b200_engine *e, int mode) { ENG(e); return eng.set_mode(mode); }
int ecoz2b200_set_recheck_kappa(ecoz2b200_engine *e, double k) { ENG(e); return eng.set_kappa(k); }
int ecoz2b200_set_recheck_kappa_tc(ecoz2b200_engine *e, double k) { ENG(e); return eng.set_kappa_tc(k); }
int ecoz2b200_set_audit(ecoz2b200_engine *e, int every) { ENG(e); return eng.set_audit(every); }
int ecoz2b200_debug_tc_distortions(ecoz2b200_engine *e, float *out, long n_rows) {
    ENG(e);
    return eng.debug_tc_distortions(out, n_rows);
}

int ecoz2b200_write_min_dists_csv(const char *path, const int32_t *codeword, const double *min_dist,
                                  long n) {
    if (!path || n < 0 || (n > 0 && (!codeword || !min_dist))) return ECOZ2B200_ERR_ARG;
    return vqb200::write_min_dists_csv(path, codeword, min_dist, n) ? ECOZ2B200_OK : ECOZ2B200_ERR_IO;
}

int ecoz2b200_debug_format_lines(const int32_t *codeword, const double *min_dist, long n, char *out,
                                 size_t capacity, size_t *bytes) {
    if (n < 0 || (n > 0 && (!codeword || !min_dist)) || !bytes) return ECOZ2B200_ERR_ARG;
    *bytes = 0;
    if (n == 0) return ECOZ2B200_OK;
    int32_t *d_cw = nullptr;
    double *d_md = nullptr;
    void *d_scr = nullptr;
    char *d_txt = nullptr;
    int rc = ECOZ2B200_ERR_CUDA;
    unsigned long long total = 0;
    int bad = 0;
    const unsigned long long *d_total;
    const int *d_bad;
    if (cudaMalloc(&d_cw, sizeof(int32_t) * (size_t) n) != cudaSuccess) goto done;
    if (cudaMalloc(&d_md, sizeof(double) * (size_t) n) != cudaSuccess) goto done;
    if (cudaMalloc(&d_scr, vqb200::format_scratch_bytes(n)) != cudaSuccess) goto done;
    if (cudaMalloc(&d_txt, vqb200::format_text_capacity(n)) != cudaSuccess) goto done;
    cudaMemcpy(d_cw, codeword, sizeof(int32_t) * (size_t) n, cudaMemcpyHostToDevice);
    cudaMemcpy(d_md, min_dist, sizeof(double) * (size_t) n, cudaMemcpyHostToDevice);
    vqb200::launch_format_lengths(d_cw, d_md, n, d_scr, 0);
    vqb200::format_result_ptrs(d_scr, n, &d_total, &d_bad);
    cudaMemcpy(&total, d_total, sizeof total, cudaMemcpyDeviceToHost);
    cudaMemcpy(&bad, d_bad, sizeof bad, cudaMemcpyDeviceToHost);
    if (bad) { rc = ECOZ2B200_ERR_ARG; goto done; }
    *bytes = (size_t) total;
    if (total > capacity || !out) { rc = ECOZ2B200_ERR_ARG; goto done; }
    vqb200::launch_format_write(d_cw, d_md, n, d_scr, d_txt, 0);
    if (cudaMemcpy(out, d_txt, (size_t) total, cudaMemcpyDeviceToHost) == cudaSuccess &&
        cudaGetLastError() == cudaSuccess)
        rc = ECOZ2B200_OK;
done:
    cudaFree(d_cw); cudaFree(d_md); cudaFree(d_scr); cudaFree(d_txt);
    return rc;
}

int ecoz2b200_min_dists_text(ecoz2b200_engine *e, const char **text, size_t *bytes) {
    ENG(e);
    return eng.min_dists_text(text, bytes);
}

int ecoz2b200_tc_issued_flops_per_eval(int P, int *is_f16) {
    return vqb200::tc_issued_flops_per_eval(P + 1, is_f16);
}

int ecoz2b200_shard_range(long total, int rank, int world_size, long *lo, long *hi) {
    if (total < 0 || world_size < 1 || rank < 0 || rank >= world_size || !lo || !hi) return ECOZ2B200_ERR_ARG;
    vqb200::shard_range(total, rank, world_size, lo, hi);
    return ECOZ2B200_OK;
}
int ecoz2b200_packed_layout(int K, int P, long offsets[6]) {
    if (K < 1 || P < 1 || !offsets) return ECOZ2B200_ERR_ARG;
    const vqb200::PackedLayout L{K, P + 1};
    offsets[0] = L.cells(); offsets[1] = L.counts(); offsets[2] = L.discel();
    offsets[3] = L.dd(); offsets[4] = L.inertia(); offsets[5] = L.size();
    return ECOZ2B200_OK;
}

int ecoz2b200_nccl_unique_id(unsigned char id_out[128]) {
    const vqb200::NcclApi *api = vqb200::nccl_api();
    if (!api) return ECOZ2B200_ERR_NCCL;
    vqb200::NcclUniqueId id;
    if (api->GetUniqueId(&id) != 0) return ECOZ2B200_ERR_NCCL;
    memcpy(id_out, id.internal, 128);
    return ECOZ2B200_OK;
}
int ecoz2b200_comm_init(ecoz2b200_engine *e, const unsigned char id[128], int rank, int world) {
    ENG(e);
    return eng.comm_init(id, rank, world);
}

int ecoz2b200_frames_clear(ecoz2b200_engine *e) { ENG(e); return eng.frames_clear(); }
int ecoz2b200_frames_reserve(ecoz2b200_engine *e, long total) { ENG(e); return eng.frames_reserve(total); }
int ecoz2b200_frames_append_host(ecoz2b200_engine *e, const double *x, long n) {
    ENG(e);
    return eng.frames_append_host(x, n);
}
int ecoz2b200_frames_append_device(ecoz2b200_engine *e, const double *x, long n) {
    ENG(e);
    return eng.frames_append_device(x, n);
}
int ecoz2b200_frames_commit(ecoz2b200_engine *e) { ENG(e); return eng.frames_commit(); }
long ecoz2b200_frames_count(const ecoz2b200_engine *e) { return e ? e->impl.frames_count() : -1; }

int ecoz2b200_codebook_initial(ecoz2b200_engine *e) { ENG(e); return eng.codebook_initial(); }
int ecoz2b200_codebook_set_reflections(ecoz2b200_engine *e, const double *r, int K) {
    ENG(e);
    return eng.codebook_set_reflections(r, K);
}
int ecoz2b200_codebook_set_raas(ecoz2b200_engine *e, const double *r, int K) {
    ENG(e);
    return eng.codebook_set_raas(r, K);
}
int ecoz2b200_codebook_grow(ecoz2b200_engine *e) {
    ENG(e);
    int rc = eng.codebook_grow();
    return rc;
}
int ecoz2b200_codebook_size(const ecoz2b200_engine *e) { return e ? e->impl.codebook_size() : -1; }
int ecoz2b200_codebook_get(ecoz2b200_engine *e, double *refl, double *raas) {
    ENG(e);
    return eng.codebook_get(refl, raas);
}

int ecoz2b200_assign(ecoz2b200_engine *e, int32_t *idx, double *ddmin, long *nre) {
    ENG(e);
    return eng.assign(idx, ddmin, nre);
}
int ecoz2b200_kmeans_pass(ecoz2b200_engine *e, double *cells, int32_t *cardd, double *discel,
                          double *DD) {
    ENG(e);
    return eng.kmeans_pass(cells, cardd, discel, DD);
}
int ecoz2b200_kmeans_pass_host(ecoz2b200_engine *e, const double *x, long n, double *cells,
                               int32_t *cardd, double *discel, double *DD) {
    ENG(e);
    return eng.kmeans_pass_host(x, n, cells, cardd, discel, DD);
}
int ecoz2b200_sigma(ecoz2b200_engine *e, double avg, double *out) { ENG(e); return eng.sigma(avg, out); }
int ecoz2b200_inertia(ecoz2b200_engine *e, double *out) { ENG(e); return eng.inertia(out); }
int ecoz2b200_last_assignment(ecoz2b200_engine *e, int32_t *idx, double *md) {
    ENG(e);
    return eng.last_assignment(idx, md);
}
int ecoz2b200_learn(ecoz2b200_engine *e, double eps, int stop_M, ecoz2b200_stage_cb on_stage,
                    ecoz2b200_pass_cb on_pass_cb, void *ctx) {
    ENG(e);
    return eng.learn(eps, stop_M, on_stage, on_pass_cb, ctx);
}
int ecoz2b200_quantize_host(ecoz2b200_engine *e, const double *x, long T, uint16_t *seq,
                            double *avg) {
    ENG(e);
    return eng.quantize_host(x, T, seq, avg);
}
int ecoz2b200_quantize_resident(ecoz2b200_engine *e, uint16_t *seq, double *avg) {
    ENG(e);
    return eng.quantize_resident(seq, avg);
}

int ecoz2b200_quantize_models(ecoz2b200_engine *e, const double *reflections, int R, int K,
                              const long *seg_offsets, int num_segments, uint16_t *seq_out, double *dists_out,
                              int *one_sweep) {
    ENG(e);
    return eng.quantize_models(reflections, R, K, seg_offsets, num_segments, seq_out, dists_out, one_sweep);
}

int ecoz2b200_counters_reset(ecoz2b200_engine *e) { ENG(e); eng.counters_reset(); return 0; }
int ecoz2b200_counters_get(ecoz2b200_engine *e, ecoz2b200_counters *out) {
    ENG(e);
    if (!out) return ECOZ2B200_ERR_ARG;
    const vqb200::Counters &c = eng.counters();
    out->assign_ms = c.assign_ms;
    out->accumulate_ms = c.accumulate_ms;
    out->update_ms = c.update_ms;
    out->allreduce_ms = c.allreduce_ms;
    out->kernel_launches = c.kernel_launches;
    out->evals = c.evals;
    out->rechecked = c.rechecked;
    out->search_main_ms = c.search_main_ms;
    out->search_main_launches = c.search_main_launches;
    out->audited = c.audited;
    out->audit_idx_diff = c.audit_idx_diff;
    out->audit_mismatches = c.audit_mismatches;
    return ECOZ2B200_OK;
}

double ecoz2b200_ffma_peak_tflops(int device, double millis) {
    vqb200::DeviceGuard device_guard_;
    if (cudaSetDevice(device) != cudaSuccess) return 0.0;
    cudaDeviceProp prop;
    if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) return 0.0;
    float *out = nullptr;
    if (cudaMalloc(&out, 16) != cudaSuccess) return 0.0;
    cudaEvent_t a, b;
    cudaEventCreate(&a);
    cudaEventCreate(&b);
    const int blocks = prop.multiProcessorCount * 8;
    int iters = 2000;
    double best = 0.0;
    for (int rep = 0; rep < 6; rep++) {
        cudaEventRecord(a, 0);
        vqb200::launch_ffma_probe(out, blocks, iters, 0);
        cudaEventRecord(b, 0);
        if (cudaEventSynchronize(b) != cudaSuccess) break;
        float ms = 0;
        cudaEventElapsedTime(&ms, a, b);
        const double flops = 2.0 * 128.0 * iters * 256.0 * blocks;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
        if (ms < millis && iters < (1 << 24)) iters = (int) (iters * (millis / (ms > 0.01 ? ms : 0.01)) * 1.05) + 1;
    }
    cudaEventDestroy(a);
    cudaEventDestroy(b);
    cudaFree(out);
    return best;
}

int ecoz2b200_prd_info(const char *path, int *P, int *T, char class_name[96]) {
    return vqb200::prd_info(path, P, T, class_name) ? ECOZ2B200_ERR_IO : ECOZ2B200_OK;
}
int ecoz2b200_prd_read(const char *path, double *out, long cap) {
    return vqb200::prd_read(path, out, cap) ? ECOZ2B200_ERR_IO : ECOZ2B200_OK;
}
int ecoz2b200_cbook_save(const char *path, const char *cls, const double *refl, int M, int P) {
    return vqb200::cbook_save(path, cls, refl, M, P) ? ECOZ2B200_ERR_IO : ECOZ2B200_OK;
}
int ecoz2b200_cbook_info(const char *path, int *P, int *M, char class_name[96]) {
    return vqb200::cbook_info(path, P, M, class_name) ? ECOZ2B200_ERR_IO : ECOZ2B200_OK;
}
int ecoz2b200_cbook_read(const char *path, double *out, long cap) {
    return vqb200::cbook_read(path, out, cap) ? ECOZ2B200_ERR_IO : ECOZ2B200_OK;
}
int ecoz2b200_seq_save(const char *path, const char *cls, const uint16_t *seq, int T, int M) {
    return vqb200::seq_save(path, cls, seq, T, M) ? ECOZ2B200_ERR_IO : ECOZ2B200_OK;
}

int ecoz2b200_class_name_of(const char *path, char class_name[96]) {
    if (!path || !class_name) return ECOZ2B200_ERR_ARG;
    const std::string cls = vqb200::get_class_name(path);
    memset(class_name, 0, 96);
    memcpy(class_name, cls.data(), cls.size() < 95 ? cls.size() : 95);
    return ECOZ2B200_OK;
}
int ecoz2b200_output_filename(const char *from, const char *base_dir, const char *ext, char *out, long capacity) {
    if (!from || !base_dir || !ext || !out || capacity < 1) return ECOZ2B200_ERR_ARG;
    const std::string name = vqb200::get_output_filename(from, base_dir, ext);
    if ((long) name.size() + 1 > capacity) return ECOZ2B200_ERR_ARG;
    memcpy(out, name.c_str(), name.size() + 1);
    return ECOZ2B200_OK;
}

}  // extern "C"
