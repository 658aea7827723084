// ecoz2_cli.cpp -- `ecoz2 vq learn|quantize|classify` and `ecoz2 lpc` command line on top of the C ABI.
//
// The reference's CLI is the Rust binary (src/main.rs:95-97 -> src/vq/mod.rs); there
// is no Rust toolchain in this environment, so this C++ driver keeps the same
// sub-commands, option names, defaults and file-resolution rules
// (src/vq/mod.rs:39-96,151-216; src/utl/mod.rs:66-260) and calls the very same
// three C entry points the Rust FFI shim binds (src/ecoz2_lib/mod.rs:96-122).
#include <dirent.h>
#include <sys/stat.h>

#include <algorithm>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fstream>
#include <sstream>
#include <string>
#include <vector>

#include "ecoz2_lpc_b200.h"
#include "ecoz2_vq_b200.h"

namespace {

bool ends_with(const std::string &s, const std::string &suf) {
    return s.size() >= suf.size() && s.compare(s.size() - suf.size(), suf.size(), suf) == 0;
}
bool is_dir(const std::string &p) {
    struct stat st;
    return stat(p.c_str(), &st) == 0 && S_ISDIR(st.st_mode);
}
bool is_file(const std::string &p) {
    struct stat st;
    return stat(p.c_str(), &st) == 0 && S_ISREG(st.st_mode);
}
// src/utl/mod.rs:246-260 (WalkDir)
void walk(const std::string &dir, const std::string &ext, std::vector<std::string> &out) {
    DIR *d = opendir(dir.c_str());
    if (!d) return;
    while (dirent *e = readdir(d)) {
        const std::string name = e->d_name;
        if (name == "." || name == "..") continue;
        const std::string path = dir + "/" + name;
        if (is_dir(path)) walk(path, ext, out);
        else if (is_file(path) && ends_with(path, ext)) out.push_back(path);
    }
    closedir(d);
}
// src/utl/mod.rs:146-196: rows `tt,class,selection`, '#' comments, header line
bool files_from_csv(const std::string &csv, const std::string &tt, const std::string &cls,
                    const std::string &subdir, const std::string &ext, const std::string *templ,
                    std::vector<std::string> &out) {
    std::ifstream in(csv);
    if (!in) return false;
    std::string line;
    bool header = true;
    while (std::getline(in, line)) {
        if (line.empty() || line[0] == '#') continue;
        if (header) { header = false; continue; }
        std::stringstream ss(line);
        std::string r_tt, r_class, r_sel;
        std::getline(ss, r_tt, ',');
        std::getline(ss, r_class, ',');
        std::getline(ss, r_sel, ',');
        while (!r_sel.empty() && (r_sel.back() == '\r' || r_sel.back() == ' ')) r_sel.pop_back();
        if (tt != r_tt) continue;
        if (!cls.empty() && cls != r_class) continue;
        std::string f;
        if (templ) {
            f = *templ;
            for (size_t p; (p = f.find("{class}")) != std::string::npos;) f.replace(p, 7, r_class);
            for (size_t p; (p = f.find("{selection}")) != std::string::npos;) f.replace(p, 11, r_sel);
        } else {
            f = "data/" + subdir + "/" + r_class + "/" + r_sel + ext;
        }
        out.push_back(f);
    }
    return !out.empty();
}
// Order of Rust's PathBuf (what `list.sort()` of src/utl/mod.rs:222,244 uses): component by
// component, not byte by byte -- "a/x.prd" sorts before "a-b/x.prd" although '-' < '/'.
// Components as std::path::Path::components(): repeated separators and inner "." dropped; kinds order
// RootDir < CurDir < ParentDir < Normal, normal components compare as bytes.
struct PathKey {
    std::vector<std::pair<int, std::string>> comps;
    explicit PathKey(const std::string &p) {
        size_t i = 0;
        if (!p.empty() && p[0] == '/') comps.emplace_back(0, "");
        bool first = true;
        while (i <= p.size()) {
            const size_t j = p.find('/', i);
            const std::string c = p.substr(i, j == std::string::npos ? std::string::npos : j - i);
            if (c == ".") { if (first && comps.empty()) comps.emplace_back(1, ""); }
            else if (c == "..") comps.emplace_back(2, "");
            else if (!c.empty()) comps.emplace_back(3, c);
            if (!c.empty()) first = false;
            if (j == std::string::npos) break;
            i = j + 1;
        }
    }
    bool operator<(const PathKey &o) const { return comps < o.comps; }
};
bool path_less(const std::string &a, const std::string &b) { return PathKey(a) < PathKey(b); }

// src/utl/mod.rs:66-83,201-243
bool resolve(const std::vector<std::string> &given, const std::string &tt, const std::string &cls,
             const std::string &subdir, const std::string &ext, const std::string *templ,
             std::vector<std::string> &out) {
    if (given.size() == 1 && ends_with(given[0], ".csv"))
        return files_from_csv(given[0], tt, cls, subdir, ext, templ, out);
    for (const std::string &g : given) {
        if (is_dir(g)) walk(g, ext, out);
        else if (is_file(g) && ends_with(g, ext)) out.push_back(g);
    }
    std::sort(out.begin(), out.end(), path_less);
    return !out.empty();
}

void step(void *, int M, double avg, double sigma, double inertia) {
    // src/ecoz2_lib/mod.rs:59-63
    printf("   Ecoz2ObserverRef.step: M=%d avg_distortion=%g sigma=%g inertia=%g\n", M, avg, sigma,
           inertia);
}

int usage() {
    fprintf(stderr,
            "usage:\n"
            "  ecoz2 vq learn [-B <codebook> | -P <P>] [-e <eps>=0.05] [--class-name <class>]\n"
            "                 --predictors <files|dirs|one.csv>...\n"
            "  ecoz2 vq quantize --codebook <cb> --predictors <files|dirs|one.csv>...\n"
            "                 [--predictors-dir-template <t>] [--tt TRAIN|TEST] [--class-name <c>] [-s]\n"
            "  ecoz2 vq classify [-r] --codebooks <files|dirs>... --tt TRAIN|TEST\n"
            "                 --predictors <files|dirs|one.csv>...\n"
            "  ecoz2 lpc [-P <P>=36] [-W <ms>=45] [-O <ms>=15] [-m <minpc>=0] [-s <split>=0] [-X <secs>=5]\n"
            "                 [--signals-dir-template <t>] [--tt TRAIN|TEST] [--class <c>] [--verbose]\n"
            "                 --signals <wav files|dirs|one.csv>...\n"
            "environment: ECOZ2_VQ_GPUS=all|0,1,..  ECOZ2_VQ_MAX_M=<stop size>  ECOZ2_VQ_MIN_DISTS=0|1\n");
    return 2;
}


// `ecoz2 lpc`, reference src/lpc/mod.rs:16-143 (the --zrs / --zrsp Rust-native variants are not
// part of the C path and are not offered)
int main_lpc(int argc, char **argv) {
    int P = 36, W = 45, O = 15, minpc = 0;
    float split = 0.f, mintrpt = 5.f;
    bool verbose = false, have_class = false;
    std::string templ = "data/signals", tt, cls, v;
    std::vector<std::string> signals, files;
    for (int i = 2; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&](std::string &dst) { if (i + 1 < argc) dst = argv[++i]; };
        if (a == "-P" || a == "--prediction-order") { next(v); P = atoi(v.c_str()); }
        else if (a == "-W" || a == "--window-length-ms") { next(v); W = atoi(v.c_str()); }
        else if (a == "-O" || a == "--offset-length-ms") { next(v); O = atoi(v.c_str()); }
        else if (a == "-m" || a == "--minpc") { next(v); minpc = atoi(v.c_str()); }
        else if (a == "-s" || a == "--split") { next(v); split = (float) atof(v.c_str()); }
        else if (a == "-X") { next(v); mintrpt = (float) atof(v.c_str()); }
        else if (a == "--signals-dir-template") next(templ);
        else if (a == "--tt") next(tt);
        else if (a == "--class") { next(cls); have_class = true; }
        else if (a == "--verbose") verbose = true;
        else if (a == "--signals") {
            while (i + 1 < argc && argv[i + 1][0] != '-') signals.push_back(argv[++i]);
        } else return usage();
    }
    if (signals.empty()) return usage();
    if (!resolve(signals, tt, have_class ? cls : "", "", ".wav", &templ, files)) {
        printf("No signals given\n");
        return 1;
    }
    std::vector<char *> ptrs;
    for (auto &f : files) ptrs.push_back(&f[0]);
    return ecoz2_lpc_signals(P, W, O, minpc, split, ptrs.data(), (int) ptrs.size(), mintrpt, verbose ? 1 : 0);
}

}  // namespace

int main(int argc, char **argv) {
    if (argc >= 2 && strcmp(argv[1], "lpc") == 0) return main_lpc(argc, argv);
    if (argc < 3 || strcmp(argv[1], "vq") != 0) return usage();
    const std::string cmd = argv[2];
    std::string base_cb, class_name, codebook, tt, templ = "data/predictors";
    bool have_class = false, show = false, have_templ = false, show_ranked = false;
    std::vector<std::string> codebooks;
    int P = -1;
    double eps = 0.05;
    std::vector<std::string> predictors;
    for (int i = 3; i < argc; i++) {
        const std::string a = argv[i];
        auto next = [&](std::string &dst) { if (i + 1 < argc) dst = argv[++i]; };
        std::string v;
        if (a == "-B" || a == "--base-codebook") next(base_cb);
        else if (a == "-P" || a == "--prediction-order") { next(v); P = atoi(v.c_str()); }
        else if (a == "-e" || a == "--epsilon") { next(v); eps = atof(v.c_str()); }
        else if (a == "--class-name") { next(class_name); have_class = true; }
        else if (a == "--codebook") next(codebook);
        else if (a == "--tt") next(tt);
        else if (a == "--predictors-dir-template") { next(templ); have_templ = true; }
        else if (a == "-s" || a == "--show-filenames") show = true;
        else if (a == "--exp-key") next(v);  // comet logging is out of scope (needs network)
        else if (a == "-r" || a == "--show-ranked") show_ranked = true;
        else if (a == "--codebooks") {
            while (i + 1 < argc && argv[i + 1][0] != '-') codebooks.push_back(argv[++i]);
        } else if (a == "--predictors") {
            while (i + 1 < argc && argv[i + 1][0] != '-') predictors.push_back(argv[++i]);
        } else return usage();
    }
    (void) have_templ;
    std::vector<std::string> files;
    if (cmd == "files") {   // (no counterpart: prints the resolved predictor list, one per line, in run order)
        if (!resolve(predictors, tt.empty() ? "TRAIN" : tt, have_class ? class_name : "", "predictors", ".prd",
                     nullptr, files))
            return 1;
        for (const std::string &f : files) printf("%s\n", f.c_str());
        return 0;
    }
    if (cmd == "learn") {
        if (!base_cb.empty() && P >= 0) {
            printf("Only one of base codebook or prediction order expected\n");
            return 1;
        }
        if (base_cb.empty() && P < 0) return usage();
        if (!resolve(predictors, "TRAIN", have_class ? class_name : "", "predictors", ".prd", nullptr,
                     files)) {
            printf("No predictors given\n");
            return 1;
        }
        const std::string cls = have_class ? class_name : "_";
        printf("vq_learn: base_codebook_opt=%s prediction_order=%d, epsilon=%g codebook_class_name=%s "
               "predictor_filenames: %zu\n",
               base_cb.empty() ? "None" : base_cb.c_str(), P, eps, cls.c_str(), files.size());
        std::vector<const char *> ptrs;
        for (auto &f : files) ptrs.push_back(f.c_str());
        if (!base_cb.empty())
            return ecoz2_vq_learn_using_base_codebook(&base_cb[0], eps, ptrs.data(), (int) ptrs.size(),
                                                      nullptr, step);
        return ecoz2_vq_learn(P, eps, cls.c_str(), ptrs.data(), (int) ptrs.size(), nullptr, step);
    }
    if (cmd == "quantize") {
        if (codebook.empty() || predictors.empty()) return usage();
        if (!resolve(predictors, tt, have_class ? class_name : "", "", ".prd", &templ, files)) {
            printf("No predictors given\n");
            return 1;
        }
        printf("number of predictor files: %zu\n", files.size());
        printf("nom_raas = %s\n", codebook.c_str());
        std::vector<const char *> ptrs;
        for (auto &f : files) ptrs.push_back(f.c_str());
        return ecoz2_vq_quantize(codebook.c_str(), ptrs.data(), (int) ptrs.size(), show ? 1 : 0);
    }
    if (cmd == "classify") {   // reference src/vq/mod.rs:218-246
        if (codebooks.empty() || predictors.empty() || tt.empty()) return usage();
        std::vector<std::string> cbs;
        for (const std::string &g : codebooks) {
            if (is_dir(g)) walk(g, ".cbook", cbs);
            else if (is_file(g) && ends_with(g, ".cbook")) cbs.push_back(g);
        }
        std::sort(cbs.begin(), cbs.end(), path_less);
        if (cbs.empty() || !resolve(predictors, tt, "", "predictors", ".prd", nullptr, files)) {
            printf("No codebooks or predictors given\n");
            return 1;
        }
        printf("number of codebooks: %zu  number of predictors: %zu\n", cbs.size(), files.size());
        printf("show_ranked = %s\n", show_ranked ? "true" : "false");
        std::vector<char *> cb_ptrs, prd_ptrs;
        for (auto &f : cbs) cb_ptrs.push_back(&f[0]);
        for (auto &f : files) prd_ptrs.push_back(&f[0]);
        return ecoz2_vq_classify(cb_ptrs.data(), (int) cb_ptrs.size(), prd_ptrs.data(), (int) prd_ptrs.size(),
                                 show_ranked ? 1 : 0);
    }
    return usage();
}
