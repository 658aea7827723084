// lpc_signals.cpp -- drop-in for ecoz2_lpc_signals (reference ecoz2/src/include/ecoz2.h:23-33,
// ecoz2/src/lpc/lpc_signals.c:38-153): wav files -> data/predictors[/TRAIN|/TEST]/<class>/<name>.prd.
// The per-signal analysis (lpa_on_signal) runs on the GPU (lpc_front.cu); this file is the
// host glue: class grouping, the rand()-driven shuffle and train/test split, wav decoding
// and the .prd writer.
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>

#include <string>
#include <vector>

#include "../../include/ecoz2_lpc_b200.h"
#include "vq_io.h"

namespace {

// Mono RIFF/WAVE -> samples as the reference sees them: dr_wav's float32 conversion
// (sgn.c:16 drwav_open_and_read_file_f32) widened to double (sgn.c:40-42).
bool wav_load(const char *path, std::vector<double> *out, unsigned *sample_rate, std::string *err) {
    FILE *f = fopen(path, "rb");
    if (!f) { *err = "cannot open"; return false; }
    std::vector<unsigned char> b;
    fseek(f, 0, SEEK_END);
    const long sz = ftell(f);
    fseek(f, 0, SEEK_SET);
    b.resize(sz > 0 ? (size_t) sz : 0);
    const bool rd = sz > 0 && fread(b.data(), 1, (size_t) sz, f) == (size_t) sz;
    fclose(f);
    if (!rd || sz < 12 || memcmp(b.data(), "RIFF", 4) != 0 || memcmp(b.data() + 8, "WAVE", 4) != 0) {
        *err = "not a RIFF/WAVE file";
        return false;
    }
    auto u16 = [&](size_t o) { return (unsigned) b[o] | ((unsigned) b[o + 1] << 8); };
    auto u32 = [&](size_t o) { return (uint32_t) u16(o) | ((uint32_t) u16(o + 2) << 16); };
    unsigned fmt = 0, channels = 0, bits = 0;
    size_t data_off = 0, data_len = 0;
    for (size_t o = 12; o + 8 <= b.size();) {
        const uint32_t len = u32(o + 4);
        if (memcmp(&b[o], "fmt ", 4) == 0 && len >= 16 && o + 8 + 16 <= b.size()) {
            fmt = u16(o + 8);
            channels = u16(o + 10);
            *sample_rate = u32(o + 12);
            bits = u16(o + 22);
            if (fmt == 0xFFFE && len >= 26) fmt = u16(o + 32);   // WAVE_FORMAT_EXTENSIBLE: sub-format
        } else if (memcmp(&b[o], "data", 4) == 0) {
            data_off = o + 8;
            data_len = len <= b.size() - data_off ? len : b.size() - data_off;
            break;
        }
        o += 8 + (size_t) len + (len & 1);
    }
    if (!fmt || !data_off) { *err = "missing fmt or data chunk"; return false; }
    if (channels != 1) { *err = "only one channel expected"; return false; }
    const size_t bps = bits / 8;
    if (bps == 0) { *err = "unsupported sample size"; return false; }
    const size_t n = data_len / bps;
    out->resize(n);
    const unsigned char *d = b.data() + data_off;
    for (size_t i = 0; i < n; i++) {
        const unsigned char *s = d + i * bps;
        float v;
        if (fmt == 1 && bits == 16) {
            v = (float) (int16_t) (s[0] | (s[1] << 8)) * 0.000030517578125f;
        } else if (fmt == 1 && bits == 8) {
            v = (float) s[0] * 0.00784313725490196078f - 1.f;
        } else if (fmt == 1 && bits == 24) {
            const int32_t x = (int32_t) ((uint32_t) s[0] << 8 | (uint32_t) s[1] << 16 | (uint32_t) s[2] << 24) >> 8;
            v = (float) ((double) x / 8388608.0);
        } else if (fmt == 1 && bits == 32) {
            int32_t x;
            memcpy(&x, s, 4);
            v = (float) ((double) x / 2147483648.0);
        } else if (fmt == 3 && bits == 32) {
            memcpy(&v, s, 4);
        } else if (fmt == 3 && bits == 64) {
            double x;
            memcpy(&x, s, 8);
            v = (float) x;
        } else {
            *err = "unsupported wav encoding";
            return false;
        }
        (*out)[i] = (double) v;
    }
    return true;
}

struct ByClass {
    std::string name;
    std::vector<char *> files;
};

double now_sec() {
    struct timespec ts;
    clock_gettime(CLOCK_MONOTONIC, &ts);
    return (double) ts.tv_sec + 1e-9 * (double) ts.tv_nsec;
}

}  // namespace

extern "C" int ecoz2_lpc_signals(int P, int windowLengthMs, int offsetLengthMs, int minpc, float split,
                                 char *sgn_filenames[], int num_signals, float mintrpt, int verbose) {
    std::vector<ByClass> classes;   // in order of first appearance (lpc_signals.c:21-36)
    for (int i = 0; i < num_signals; ++i) {
        const std::string cls = vqb200::get_class_name(sgn_filenames[i]);
        ByClass *e = nullptr;
        for (auto &c : classes)
            if (c.name == cls) { e = &c; break; }
        if (!e) {
            classes.push_back(ByClass{cls, {}});
            e = &classes.back();
        }
        e->files.push_back(sgn_filenames[i]);
    }
    printf("lpc_signals: number of classes: %ld\n", (long) classes.size());

    for (auto &e : classes) {
        // list_shuffle (ecoz2/src/utl/list.c:84-92): same libc rand() stream as the reference
        const long n = (long) e.files.size();
        for (long i = 0; i < n; i++) {
            const long j = i + rand() % (n - i);
            std::swap(e.files[i], e.files[j]);
        }
        if (minpc > 0 && n < minpc) {
            printf("class '%s': insufficient #signals=%ld\n", e.name.c_str(), n);
            continue;
        }
        int numTrain = -1;
        if (split > 0) numTrain = (int) (split * n);
        printf("class '%s': %ld\n", e.name.c_str(), n);

        for (long j = 0; j < n; j++) {
            char *sgn_filename = e.files[j];
            if (verbose) printf("  %s\n", sgn_filename);
            std::vector<double> samples;
            unsigned rate = 0;
            std::string err;
            if (!wav_load(sgn_filename, &samples, &rate, &err)) {
                fprintf(stderr, "Error loading wave file: %s (%s)\n", sgn_filename, err.c_str());
                fprintf(stderr, "%s: error loading signal\n", sgn_filename);
                continue;
            }
            const double t0 = now_sec();
            const long T = ecoz2b200_lpa_num_frames((long) samples.size(), (double) rate, windowLengthMs,
                                                    offsetLengthMs);
            std::vector<double> vectors;
            long got = -1;
            int n_err = 0;
            if (T >= 0) {
                vectors.resize((size_t) T * (P + 1));
                got = ecoz2b200_lpa_on_signal(samples.data(), (long) samples.size(), (double) rate, P,
                                              windowLengthMs, offsetLengthMs, vectors.data(), T, &n_err);
            }
            const double elapsed = now_sec() - t0;
            if (got < 0) {
                fprintf(stderr, "ERROR: %s\n", ecoz2b200_lpc_last_error());
                fprintf(stderr, "cannot create lpc predictor\n");
                continue;
            }
            if (verbose)
                printf("lpa_on_signal: P=%d numSamples=%ld sampleRate=%ld T=%ld\n", P, (long) samples.size(),
                       (long) rate, got);
            if (n_err) fprintf(stderr, "ERROR: lpa_on_signal: lpca error on %d frame(s)\n", n_err);
            if (elapsed >= mintrpt && verbose) printf("processing took %.2fs\n", elapsed);

            const char *base_dir = "predictors";
            if (numTrain > 0) base_dir = j <= numTrain ? "predictors/TRAIN" : "predictors/TEST";
            const std::string prd_filename = vqb200::get_output_filename(sgn_filename, base_dir, ".prd");
            if (vqb200::prd_save(prd_filename.c_str(), e.name.c_str(), vectors.data(), (int) got, P)) {
                fprintf(stderr, "%s: error saving predictor\n", prd_filename.c_str());
            } else if (verbose) {
                printf("%s: '%s': predictor saved\n", prd_filename.c_str(), e.name.c_str());
            }
            if (verbose) printf("\n");
        }
    }
    return 0;
}
