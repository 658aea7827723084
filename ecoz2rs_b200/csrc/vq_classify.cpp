// vq_classify.cpp -- `ecoz2 vq classify` on the B200 path.
//
// Drop-in for the reference's ecoz2_vq_classify (ecoz2/src/include/ecoz2.h:95-101, which forwards
// to vq_classify, ecoz2/src/vq/vq_classify.c:185-327): every predictor file is quantized against
// every codebook model (cbook_quantize -> quantize, vq/vq.c:154-157, vq/quantize.c:8-34), the
// models are ranked by average distortion, and a confusion matrix + candidate-order table goes to
// stdout in the reference's layout.
//
// B200 shape of the work: the reference runs R x F serial quantize calls (R models, F files).
// Here ALL frames of ALL files are resident in HBM once; each model is one pass of the search
// kernel over them (launch_assign_tc + FP64 re-evaluation + FP64 distortion of the winner); the
// per-file averages are segment sums of the per-frame FP64 distortions, taken in frame order like
// the reference's accumulation (quantize.c:27, 32).
#include "../../include/ecoz2_vq_b200.h"
#include "vq_engine.h"
#include "vq_hostio.h"
#include "vq_io.h"

#include <cstdio>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

using vqb200::Engine;

std::vector<int> ecoz2b200_devices_from_env();
void ecoz2b200_apply_mode_env(Engine &eng);

namespace {

constexpr int kMaxModels = 256;    // MAX_CB_MODELS, reference include/vq.h:16
constexpr int kMaxClasses = 128;   // vq_classify.c:16
constexpr int kTotal = kMaxClasses - 1;

struct Model {
    std::string file, class_name;
    std::vector<double> refl;
};

struct Tables {
    // result[c][0] = tests of class c; result[c][i] = tests whose own model was the i-th candidate
    // (vq_classify.c:26-31); row kTotal = totals.  confusion[i][j] as vq_classify.c:33-37.
    std::vector<std::vector<int>> result, confusion;
    Tables(int num_models)
        : result(kMaxClasses, std::vector<int>(num_models + 1, 0)),
          confusion(kTotal, std::vector<int>(kTotal, 0)) {}
};

// vq_classify.c:64-161
void report_results(const Tables &tb, const std::vector<Model> &models) {
    const int num_models = (int) models.size();
    const auto &result = tb.result;
    if (result[kTotal][0] == 0) return;
    auto cname = [&](int i) { return i < num_models ? models[i].class_name.c_str() : ""; };
    int margin = 0;
    for (int i = 0; i < kTotal; i++)
        if (result[i][0]) {
            const int len = (int) strlen(cname(i));
            if (margin < len) margin = len;
        }
    margin += 2;

    printf("\n\n");
    printf("%*s ", margin, "");
    printf("Confusion matrix:\n");
    printf("%*s ", margin, "");
    printf("     ");
    for (int j = 0; j < kTotal; j++)
        if (result[j][0] > 0) printf("%3d ", j);
    printf("    tests   errors\n");
    for (int i = 0; i < kTotal; i++) {
        if (result[i][0] == 0) continue;
        printf("\n");
        printf("%*s ", margin, cname(i));
        printf("%3d  ", i);
        int num_errs = 0;
        for (int j = 0; j < kTotal; j++)
            if (result[j][0] > 0) {
                printf("%3d ", tb.confusion[i][j]);
                if (i != j) num_errs += tb.confusion[i][j];
            }
        printf("%8d%8d", result[i][0], num_errs);
    }

    printf("\n\n");
    printf("%*s ", margin, "");
    printf("class     accuracy    tests      candidate order\n");
    int num_classes = 0;
    float avg_accuracy = 0, error_rate = 0;
    for (int classId = 0; classId <= kTotal; classId++) {
        if (result[classId][0] == 0) continue;
        const int num_tests = result[classId][0];
        const int correct_tests = result[classId][1];
        const float acc = (float) correct_tests / num_tests;
        const float err = (float) (num_tests - correct_tests) / num_tests;
        if (classId < kTotal) {
            ++num_classes;
            avg_accuracy += acc;
            error_rate += err;
            printf("%*s ", margin, cname(classId));
            printf("  %3d    ", classId);
        } else {
            printf("\n");
            printf("%*s ", margin, "");
            printf("  TOTAL  ");
        }
        printf("  %6.2f%%   %3d        ", (float) (100.0 * acc), num_tests);
        for (int i = 1; i <= num_models; i++) printf("%3d ", result[classId][i]);
        printf("\n");
    }
    printf("  avg_accuracy  %6.2f%%\n", (float) (100.0 * avg_accuracy) / num_classes);
    printf("    error_rate  %6.2f%%\n", (float) (100.0 * error_rate) / num_classes);
    printf("\n");
}

// Order of the models by DEcreasing distortion, the reference's selection (vq_classify.c:164-183)
// including its quirk: once model 0 is taken, a remaining model is only picked while its
// distortion is > 0 (average distortions are positive, so this never bites in practice).
void sort_dists(const std::vector<double> &dists, std::vector<int> &ordp) {
    const int n = (int) dists.size();
    std::vector<char> marked(n, 0);
    for (int i = 0; i < n; i++) {
        int cmp_index = -1;
        double cmp_dist = 0;
        for (int j = 0; j < n; j++)
            if (!marked[j] && (j == 0 || dists[j] > cmp_dist)) {
                cmp_index = j;
                cmp_dist = dists[j];
            }
        if (cmp_index >= 0) {
            ordp[i] = cmp_index;
            marked[cmp_index] = 1;
        }
    }
}

}  // namespace

// internal body: the exported names ecoz2_vq_classify / vq_classify (vq_capi.cpp) forward here, never
// to each other through the dynamic symbol table (see learn_entry in vq_capi.cpp)
int ecoz2b200_classify_entry(char **cb_filenames, int num_codebooks, char **prd_filenames,
                                 int num_predictors, int show_ranked) {
    vqb200::DeviceGuard device_guard_;
    // the reference asserts num_codebooks < MAX_CB_MODELS (256) but its tables have 127 class rows
    // (vq_classify.c:16-37): more models than that corrupt its memory; here they are refused
    if (!(0 < num_codebooks && num_codebooks <= kTotal && num_codebooks < kMaxModels) || num_predictors <= 0) {
        fprintf(stderr, "vq_classify: 0 < num_codebooks <= %d and num_predictors > 0 required\n", kTotal);
        return 1;
    }
    const int num_models = num_codebooks;
    std::vector<Model> models(num_models);

    printf("\nLoading models:\n");
    int P = 0, num_vecs = 0;
    for (int i = 0; i < num_models; i++) {
        printf("%2d: %s\n", i, cb_filenames[i]);
        int p = 0, m = 0;
        char cls[vqb200::kMaxClassNameLen];
        Model &mo = models[i];
        mo.file = cb_filenames[i];
        bool ok = vqb200::cbook_info(cb_filenames[i], &p, &m, cls) == 0;
        if (ok) {
            mo.refl.resize((size_t) m * (p + 1));
            ok = vqb200::cbook_read(cb_filenames[i], mo.refl.data(), (long) mo.refl.size()) == 0;
        }
        if (!ok) {
            if (i == 0) fprintf(stderr, "%s: error loading model\n", cb_filenames[i]);
            else fprintf(stderr, "could not load model %s\n", cb_filenames[i]);
            return 1;
        }
        mo.class_name = cls;
        if (i == 0) {
            P = p;
            num_vecs = m;
        } else if (num_vecs != m || P != p) {   // the reference asserts equal P (vq.c:155)
            fprintf(stderr, ": conformity error.\n");
            return 1;
        }
    }
    auto class_id = [&](const char *name) {
        for (int i = 0; i < num_models; i++)
            if (models[i].class_name == name) return i;
        return -1;
    };

    // ---- load every classifiable predictor into one resident frame set
    struct Instance { int file, classId; long first, T; };
    std::vector<Instance> inst;
    long n_total = 0;
    for (int i = 0; i < num_predictors; i++) {
        int p = 0, T = 0;
        char cls[vqb200::kMaxClassNameLen];
        if (vqb200::prd_info(prd_filenames[i], &p, &T, cls) || p != P) {
            inst.push_back({i, -2, 0, 0});   // load error, reported in file order below
            continue;
        }
        const int cid = class_id(cls);
        inst.push_back({i, cid, n_total, cid >= 0 ? (long) T : 0L});
        if (cid >= 0) n_total += T;
    }

    std::unique_ptr<Engine> h;
    std::vector<std::vector<double>> dists(num_models);   // [model][instance]
    if (n_total > 0) {
        std::vector<int> devs = ecoz2b200_devices_from_env();
        h.reset(new Engine(devs.data(), (int) devs.size(), P, num_vecs < 2 ? 2 : num_vecs));
        Engine &eng = *h;
        if (!eng.ok()) return ECOZ2B200_ERR_CUDA;
        ecoz2b200_apply_mode_env(eng);
        if (eng.frames_reserve(n_total)) return ECOZ2B200_ERR_CUDA;
        // the classifiable files, in order: loaded by the reader threads of vq_hostio.cpp
        std::vector<const char *> lf;
        std::vector<int> lT;
        std::vector<size_t> lk;
        for (size_t k = 0; k < inst.size(); k++)
            if (inst[k].classId >= 0) {
                lf.push_back(prd_filenames[inst[k].file]);
                lT.push_back((int) inst[k].T);
                lk.push_back(k);
            }
        const size_t total_bytes = (size_t) n_total * (size_t) (P + 1) * sizeof(double);
        const int small_threads = (int) (total_bytes >> 25) + 1;
        vqb200::IoLanes lanes(eng.devices(), small_threads < 8 ? small_threads : 0);
        if (!lanes.ready()) return ECOZ2B200_ERR_CUDA;
        const int lr = vqb200::load_prd_frames(eng, lanes, lf.data(), lT.data(), (int) lf.size(), P);
        if (lr < 0) return ECOZ2B200_ERR_CUDA;
        if (lr > 0) {   // a file shorter than its header says (prd_info had succeeded)
            fprintf(stderr, "%s: error loading predictor.\n", lf[(size_t) lr - 1]);
            return 1;
        }
        if (eng.frames_commit()) return ECOZ2B200_ERR_CUDA;

        // ---- all models against all frames (one sweep when the codebook size allows it,
        // Engine::quantize_models), per-file average distortions from the device
        std::vector<double> refl_all;
        for (int r = 0; r < num_models; r++) refl_all.insert(refl_all.end(), models[r].refl.begin(), models[r].refl.end());
        std::vector<long> seg(lk.size() + 1, 0);
        for (size_t j = 0; j < lk.size(); j++) seg[j + 1] = seg[j] + inst[lk[j]].T;
        std::vector<double> dd((size_t) num_models * lk.size());
        if (eng.quantize_models(refl_all.data(), num_models, num_vecs, seg.data(), (int) lk.size(), nullptr, dd.data(),
                                nullptr))
            return ECOZ2B200_ERR_CUDA;
        for (int r = 0; r < num_models; r++) {
            dists[r].assign(inst.size(), 0.0);
            for (size_t j = 0; j < lk.size(); j++) dists[r][lk[j]] = dd[(size_t) r * lk.size() + j];
        }
    }

    // ---- ranking and tables, in file order (vq_classify.c:246-322)
    Tables tb(num_models);
    std::vector<double> d(num_models);
    std::vector<int> ordp(num_models, 0);
    printf("\n");
    for (size_t k = 0; k < inst.size(); k++) {
        const Instance &it = inst[k];
        const char *filename = prd_filenames[it.file];
        if (it.classId == -2) {
            fprintf(stderr, "%s: error loading predictor.\n", filename);
            continue;
        }
        if (it.classId < 0) {
            char cls[vqb200::kMaxClassNameLen] = "";
            int p = 0, T = 0;
            vqb200::prd_info(filename, &p, &T, cls);
            fprintf(stderr, "\n%s: no model loaded for className='%s'\n", filename, cls);
            continue;
        }
        const int classId = it.classId;
        tb.result[kTotal][0]++;
        tb.result[classId][0]++;
        for (int r = 0; r < num_models; r++) d[r] = dists[r][k];
        sort_dists(d, ordp);
        const int best = ordp[num_models - 1];
        const bool correct = classId == best;
        fprintf(stderr, correct ? "\x1b[32m*\x1b[0m" : "\x1b[31m!\x1b[0m");
        fflush(stderr);
        if (show_ranked && !correct) {
            printf("\n%s: '%s'\n", filename, models[classId].class_name.c_str());
            int index = 0;
            for (int r = num_models - 1; r >= 0; r--, index++) {
                const char *mark = classId == ordp[r] ? "*" : "";
                printf("  [%2d] %1s %-60s : %e  : '%s'\n", index, mark, cb_filenames[ordp[r]], d[ordp[r]],
                       models[ordp[r]].class_name.c_str());
                if (classId == ordp[r]) break;
            }
            printf("\n");
        }
        if (classId < kTotal && best < kTotal) tb.confusion[classId][best]++;
        if (correct) {
            tb.result[kTotal][1]++;
            tb.result[classId][1]++;
        } else {
            for (int i = 1; i < num_models; i++)
                if (ordp[num_models - 1 - i] == classId) {
                    tb.result[kTotal][i + 1]++;
                    tb.result[classId][i + 1]++;
                    break;
                }
        }
    }
    report_results(tb, models);
    fflush(stdout);
    return 0;
}
