// vq_io.cpp -- ECOZ2 file formats, output paths and report writers (host only).
// Written against the format descriptions in SURVEY.md §8(a) rows 1, 11, 12, 16, 18;
// every function cites the reference code whose bytes it must reproduce.
#include "vq_io.h"

#include <fcntl.h>
#include <sys/mman.h>
#include <unistd.h>

#include <errno.h>
#include <string.h>
#include <sys/stat.h>

#include <math.h>
#include <stdint.h>

#include <memory>
#include <thread>
#include <vector>

namespace vqb200 {

int cb_errnum_last = 0;

namespace {

struct FileCloser {
    void operator()(FILE *f) const { if (f) fclose(f); }
};
using FilePtr = std::unique_ptr<FILE, FileCloser>;

// 16-byte ident, NUL padded (fileutil.c:121-139)
bool write_ident(FILE *f, const char *ident) {
    char chunk[kFileIdentLen];
    memset(chunk, 0, sizeof chunk);
    size_t len = strlen(ident);
    if (len >= (size_t) kFileIdentLen) len = kFileIdentLen - 1;
    memcpy(chunk, ident, len);
    return fwrite(chunk, sizeof chunk, 1, f) == 1;
}
// prefix match on strlen(ident) bytes (fileutil.c:141-160)
bool read_ident(FILE *f, const char *ident) {
    char chunk[kFileIdentLen];
    if (fread(chunk, sizeof chunk, 1, f) != 1) return false;
    return strncmp(ident, chunk, strlen(ident)) == 0;
}
// 96 bytes: name, NUL, then '_' padding (fileutil.c:162-182)
bool write_class(FILE *f, const char *name) {
    char chunk[kMaxClassNameLen];
    size_t len = strlen(name);
    if (len >= (size_t) kMaxClassNameLen) len = kMaxClassNameLen - 1;
    memcpy(chunk, name, len);
    chunk[len] = 0;
    for (size_t i = len + 1; i < (size_t) kMaxClassNameLen; i++) chunk[i] = '_';
    return fwrite(chunk, sizeof chunk, 1, f) == 1;
}
bool read_class(FILE *f, char *name) {
    if (fread(name, kMaxClassNameLen, 1, f) != 1) return false;
    name[kMaxClassNameLen - 1] = 0;
    return true;
}

int read_header(const char *path, const char *ident, int *a, int *b, char *class_name, FilePtr *keep) {
    FilePtr f(fopen(path, "rb"));
    if (!f) return 1;
    if (!read_ident(f.get(), ident)) return 2;
    char cls[kMaxClassNameLen];
    if (!read_class(f.get(), cls)) return 2;
    int32_t v[2];
    if (fread(v, sizeof(int32_t), 2, f.get()) != 2) return 3;
    if (a) *a = v[0];
    if (b) *b = v[1];
    if (class_name) memcpy(class_name, cls, kMaxClassNameLen);
    if (keep) *keep = std::move(f);
    return 0;
}

}  // namespace

// ---- .prd ------------------------------------------------------------------------
int prd_info(const char *path, int *P, int *T, char class_name[kMaxClassNameLen]) {
    return read_header(path, "<predictor>", P, T, class_name, nullptr);
}
int prd_read(const char *path, double *out, long capacity_values) {
    int P = 0, T = 0;
    FilePtr f;
    int rc = read_header(path, "<predictor>", &P, &T, nullptr, &f);
    if (rc) return rc;
    const long want = (long) T * (P + 1);
    if (P < 0 || T < 0 || want > capacity_values) return 3;
    if (want > 0 && fread(out, sizeof(double), (size_t) want, f.get()) != (size_t) want) return 4;
    return 0;
}

// reference ecoz2/src/lpc/prd.c:76-115
int prd_save(const char *path, const char *class_name, const double *vectors, int T, int P) {
    FilePtr f(fopen(path, "wb"));
    if (!f) return 1;
    const int32_t hdr[2] = {P, T};
    if (!write_ident(f.get(), "<predictor>") || !write_class(f.get(), class_name) ||
        fwrite(hdr, sizeof(int32_t), 2, f.get()) != 2)
        return 2;
    const size_t n = (size_t) T * (P + 1);
    if (n > 0 && fwrite(vectors, sizeof(double), n, f.get()) != n) return 2;
    return 0;
}

// ---- .cbook ------------------------------------------------------------------------
int cbook_save(const char *path, const char *class_name, const double *refl, int M, int P) {
    FilePtr f(fopen(path, "wb"));
    if (!f) return cb_errnum_last = 1;
    const int32_t hdr[2] = {P, M};
    if (!write_ident(f.get(), "<codebook>") || !write_class(f.get(), class_name) ||
        fwrite(hdr, sizeof(int32_t), 2, f.get()) != 2)
        return cb_errnum_last = 1;
    const size_t n = (size_t) M * (P + 1);
    if (n > 0 && fwrite(refl, sizeof(double), n, f.get()) != n) return cb_errnum_last = 1;
    return cb_errnum_last = 0;
}
int cbook_info(const char *path, int *P, int *M, char class_name[kMaxClassNameLen]) {
    return cb_errnum_last = read_header(path, "<codebook>", P, M, class_name, nullptr);
}
int cbook_read(const char *path, double *out, long capacity_values) {
    int P = 0, M = 0;
    FilePtr f;
    int rc = read_header(path, "<codebook>", &P, &M, nullptr, &f);
    if (rc) return cb_errnum_last = rc;
    const long want = (long) M * (P + 1);
    if (P < 0 || M < 0 || want > capacity_values) return cb_errnum_last = 3;
    if (want > 0 && fread(out, sizeof(double), (size_t) want, f.get()) != (size_t) want)
        return cb_errnum_last = 4;
    return cb_errnum_last = 0;
}

// ---- .seq ------------------------------------------------------------------------
int seq_save(const char *path, const char *class_name, const uint16_t *seq, int T, int M) {
    FilePtr f(fopen(path, "wb"));
    if (!f) return 1;
    if (T <= 0) printf("WARN: seq_save: T=%d <= 0: %s\n", T, path);
    const int32_t hdr[2] = {T, M};
    if (!write_ident(f.get(), "<sequence>") || !write_class(f.get(), class_name) ||
        fwrite(hdr, sizeof(int32_t), 2, f.get()) != 2)
        return 2;
    if (T > 0 && fwrite(seq, sizeof(uint16_t), (size_t) T, f.get()) != (size_t) T) return 2;
    return 0;
}

// ---- paths -------------------------------------------------------------------------
// class = name of the parent directory, "" if the path has fewer than two '/'
// (fileutil.c:14-30)
std::string get_class_name(const std::string &path) {
    const size_t last = path.rfind('/');
    if (last == std::string::npos) return "";
    const std::string dir = path.substr(0, last);
    const size_t prev = dir.rfind('/');
    if (prev == std::string::npos) return "";
    std::string cls = dir.substr(prev + 1);
    if (cls.size() >= (size_t) kMaxClassNameLen) cls.resize(kMaxClassNameLen - 1);
    return cls;
}

int mk_dirs(const std::string &path) {
    // every prefix up to each '/', then the whole path; EEXIST is fine (fileutil.c:65-95)
    for (size_t pos = path.find('/'); ; pos = path.find('/', pos + 1)) {
        const std::string part = pos == std::string::npos ? path : path.substr(0, pos);
        if (!part.empty() && mkdir(part.c_str(), S_IRWXU) != 0 && errno != EEXIST) {
            printf("ERROR creating directory %s: %s\n", part.c_str(), strerror(errno));
            return -1;
        }
        if (pos == std::string::npos) break;
    }
    return 0;
}

// extension swap as the reference's camext (fileutil.c:97-119): the last '.'
// anywhere in the path; if fewer than 2 characters follow it the new extension
// is appended instead.
static std::string change_ext(const std::string &name, const std::string &ext) {
    const size_t dot = name.rfind('.');
    if (dot == std::string::npos) return name;
    if (name.size() - dot < 3) return name + ext;
    if (ext.empty() || ext == ".") return name.substr(0, dot);
    if (ext[0] == '.') return name.substr(0, dot) + ext;
    return name.substr(0, dot) + "." + ext;
}

std::string get_output_filename(const std::string &from, const std::string &base_dir,
                                const std::string &ext) {
    // fileutil.c:32-63: class "_" unless the path has a parent directory name
    std::string tmp = change_ext(from, ext);
    std::string cls = "_", simple = tmp;
    const size_t last = tmp.rfind('/');
    if (last != std::string::npos) {
        simple = tmp.substr(last + 1);
        const std::string dir = tmp.substr(0, last);
        const size_t prev = dir.rfind('/');
        if (prev != std::string::npos) cls = dir.substr(prev + 1);
    }
    const std::string class_dir = "data/" + base_dir + "/" + cls;
    mk_dirs(class_dir);
    return class_dir + "/" + simple;
}

// ---- report ------------------------------------------------------------------------
void Report::prepare(const std::string &rpt_filename, long tot_vecs, double eps) {
    tot_vecs_ = tot_vecs;
    printf("Report: %s\n", rpt_filename.c_str());
    file_ = fopen(rpt_filename.c_str(), "w");
    if (!file_) printf("error creating report file %s\n", rpt_filename.c_str());
    if (file_)
        fprintf(file_, "Codebook generation\n\n%ld training vectors. (ε = %g)\n\n", tot_vecs, eps);
    const std::string csv = rpt_filename + ".csv";
    file_csv_ = fopen(csv.c_str(), "w");
    if (!file_csv_) {
        printf("error creating report file %s\n", csv.c_str());
    } else {
        fprintf(file_csv_, "# %ld training vectors. (ε = %g)\n", tot_vecs, eps);
        fprintf(file_csv_, "%s,%s,%s,%s,%s\n", "M", "passes", "DDprm", "σ", "inertia");
    }
}

void Report::cbook(const std::string &cb_filename, int num_pas, double DDprm, double sigma,
                   double inertia, int M, const int32_t *cardd, const double *discel,
                   const int32_t *codeword, const double *min_dist) {
    if (!file_) return;
    fprintf(file_, "### %s (%2d passes)  Dprm = %f  σ = %f  inertia = %f\n", cb_filename.c_str(),
            num_pas, DDprm, sigma, inertia);
    if (file_csv_) {
        fprintf(file_csv_, "%d,%d,%f,%f,%f\n", M, num_pas, DDprm, sigma, inertia);
        fflush(file_csv_);
    }
    int empty = 0;
    for (int i = 0; i < M; i++)
        if (!cardd[i]) empty++;
    fprintf(file_, "Cardinalities:");
    if (empty) fprintf(file_, "  (%d empty cells)", empty);
    fprintf(file_, "\n");
    for (int i = 0; i < M; i++) fprintf(file_, "%8d ", cardd[i]);
    fprintf(file_, "\n");
    fprintf(file_, "average cell distortion:\n");
    for (int i = 0; i < M; i++) {
        if (cardd[i]) fprintf(file_, "%8.3g ", discel[i] / cardd[i]);
        else fprintf(file_, "%8s ", "_");
    }
    fprintf(file_, "\n\n");
    fflush(file_);

    {
        const std::string name = cb_filename + ".cards_dists.csv";
        FILE *f = fopen(name.c_str(), "w");
        if (!f) {
            printf("error creating %s\n", name.c_str());
        } else {
            fprintf(f, "%s,%s,%s\n", "index", "cardinality", "distortion");
            for (int i = 0; i < M; i++) {
                fprintf(f, "%d,%d,", i, cardd[i]);
                if (cardd[i] > 0) fprintf(f, "%f\n", discel[i] / cardd[i]);
                else fprintf(f, "NaN\n");
            }
            fclose(f);
        }
    }
    if (codeword && min_dist) {
        const std::string name = cb_filename + ".min_dists.csv";
        if (!write_min_dists_csv(name.c_str(), codeword, min_dist, tot_vecs_))
            printf("error creating %s\n", name.c_str());
    }
}

// ---- "<cbook>.min_dists.csv" (report.c:109-123): N lines "%d,%f\n" per stage.  With passes of
// milliseconds, 10M glibc fprintf calls per stage (2 s) would BE the run time of `vq learn`
// (SURVEY.md 8f-2), so the lines are formatted by all host threads into per-thread buffers with an
// exact integer "%f" (below) and written in order.

// decimal digits of v, returns the end
static inline char *fmt_u64(char *p, uint64_t v) {
    char tmp[24];
    int n = 0;
    do { tmp[n++] = (char) ('0' + v % 10); v /= 10; } while (v);
    while (n) *p++ = tmp[--n];
    return p;
}

// printf("%f") of a finite |x| < 2^63: x = m 2^e exactly, so round(x 10^6) is an exact integer
// computation (m 10^6 < 2^73 in 128 bits, shifted by -e with round-half-to-even on the exact
// remainder -- what glibc does in the default rounding mode).  Anything else: snprintf.
char *format_f6(char *p, double x) {
    uint64_t bits;
    memcpy(&bits, &x, sizeof bits);
    const int ef = (int) ((bits >> 52) & 0x7ff);
    uint64_t m = bits & ((1ull << 52) - 1);
    if (ef == 0x7ff || ef >= 1023 + 62) {   // inf / nan / huge: at most kMaxF6Chars (DBL_MAX has 309 digits)
        const int w = snprintf(p, kMaxF6Chars + 1, "%f", x);
        return p + (w < 0 ? 0 : (w > kMaxF6Chars ? kMaxF6Chars : w));
    }
    if (bits >> 63) *p++ = '-';
    int e;
    if (ef == 0) e = -1074;            // subnormal (or zero)
    else { m |= 1ull << 52; e = ef - 1075; }
    unsigned __int128 v = (unsigned __int128) m * 1000000u;   // x 10^6 = v 2^e
    if (e >= 0) {
        v <<= e;                       // ef < 1023 + 62 keeps this below 2^(73+9)... guarded: e <= 9
    } else if (-e >= 127) {
        v = 0;                         // below 2^-54: rounds to 0.000000
    } else {
        const int sh = -e;
        const unsigned __int128 q = v >> sh, r = v - (q << sh), half = (unsigned __int128) 1 << (sh - 1);
        v = q + ((r > half || (r == half && (q & 1))) ? 1 : 0);
    }
    const uint64_t ip = (uint64_t) (v / 1000000u), fp = (uint64_t) (v % 1000000u);
    p = fmt_u64(p, ip);
    *p++ = '.';
    for (int k = 5, f = (int) fp; k >= 0; k--, f /= 10) p[k] = (char) ('0' + f % 10);
    return p + 6;
}

static inline char *format_i32(char *p, int32_t v) {
    if (v < 0) { *p++ = '-'; return fmt_u64(p, (uint64_t) (-(int64_t) v)); }
    return fmt_u64(p, (uint64_t) v);
}

bool write_min_dists_csv(const char *name, const int32_t *codeword, const double *min_dist, long n) {
    FILE *f = fopen(name, "w");
    if (!f) return false;
    fputs("codeword,minDistortion\n", f);
    unsigned hw = std::thread::hardware_concurrency();
    const int T = (int) (hw == 0 ? 1 : (hw > 32 ? 32 : hw));
    const long kLines = 1L << 17;                 // lines per thread and round; a line is < 64 chars
    std::vector<std::vector<char>> bufs((size_t) T);
    std::vector<size_t> used((size_t) T, 0);
    bool ok = true;
    for (long base = 0; base < n && ok; base += (long) T * kLines) {
        auto work = [&](int t) {
            const long lo = base + (long) t * kLines;
            long hi = lo + kLines;
            if (hi > n) hi = n;
            used[(size_t) t] = 0;
            if (lo >= hi) return;
            std::vector<char> &b = bufs[(size_t) t];
            if (b.size() < (size_t) kLines * 64) b.resize((size_t) kLines * 64);
            char *p = b.data();
            for (long v = lo; v < hi; v++) {
                // usual lines are < 64 chars; a huge distortion ("%f" of 1e300) needs up to kMaxF6Chars
                if ((size_t) (p - b.data()) + 16 + kMaxF6Chars > b.size()) {
                    const size_t at = (size_t) (p - b.data());
                    b.resize(b.size() * 2);
                    p = b.data() + at;
                }
                p = format_i32(p, codeword[v]);
                *p++ = ',';
                p = format_f6(p, min_dist[v]);
                *p++ = '\n';
            }
            used[(size_t) t] = (size_t) (p - b.data());
        };
        const long left = n - base;
        const int active = (int) ((left + kLines - 1) / kLines < T ? (left + kLines - 1) / kLines : T);
        std::vector<std::thread> th;
        for (int t = 1; t < active; t++) th.emplace_back(work, t);
        work(0);
        for (auto &x : th) x.join();
        for (int t = 0; t < active && ok; t++)
            if (used[(size_t) t]) ok = fwrite(bufs[(size_t) t].data(), 1, used[(size_t) t], f) == used[(size_t) t];
    }
    return fclose(f) == 0 && ok;
}

bool write_text_file(const char *name, const char *header, const char *text, size_t bytes) {
    const int fd = open(name, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return false;
    const size_t hl = strlen(header);
    const size_t total = hl + bytes;
    bool ok = true;
    // Large files: the copy into fresh page-cache pages is what costs, and write() / pwrite() to one
    // file are serialised by its inode lock.  Mapping the file lets several threads fault and fill
    // disjoint ranges at once.
    void *map = MAP_FAILED;
    // (posix_fallocate + MAP_POPULATE first was measured slower on tmpfs: 2.8 s against 1.6 s for the
    // eleven 120 MB files of configs[1])
    if (bytes >= (8u << 20) && ftruncate(fd, (off_t) total) == 0)
        map = mmap(nullptr, total, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    if (map != MAP_FAILED) {
        char *dst = static_cast<char *>(map);
        memcpy(dst, header, hl);
        unsigned hw = std::thread::hardware_concurrency();
        const int T = (int) (hw == 0 ? 1 : (hw > 16 ? 16 : hw));
        const size_t per = ((bytes + T - 1) / T + 4095) & ~(size_t) 4095;
        auto work = [&](int t) {
            const size_t lo = (size_t) t * per;
            if (lo >= bytes) return;
            const size_t hi = lo + per < bytes ? lo + per : bytes;
            memcpy(dst + hl + lo, text + lo, hi - lo);
        };
        std::vector<std::thread> th;
        for (int t = 1; t < T; t++) th.emplace_back(work, t);
        work(0);
        for (auto &x : th) x.join();
        ok = munmap(map, total) == 0;
    } else {
        ok = (size_t) write(fd, header, hl) == hl;
        size_t lo = 0;
        while (ok && lo < bytes) {
            const ssize_t w = write(fd, text + lo, bytes - lo);
            if (w <= 0) ok = false;
            else lo += (size_t) w;
        }
    }
    return ::close(fd) == 0 && ok;
}

void Report::writer_loop() {
    for (;;) {
        Job job;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || !jobs_.empty(); });
            if (jobs_.empty()) return;
            job = jobs_.front();
            jobs_.pop_front();
        }
        const std::string name = job.name + ".min_dists.csv";
        if (!write_text_file(name.c_str(), "codeword,minDistortion\n", job.text, job.bytes))
            printf("error creating %s\n", name.c_str());
        {
            std::lock_guard<std::mutex> lk(mu_);
            outstanding_--;
        }
        cv_.notify_all();
    }
}

void Report::min_dists_submit(const std::string &cb_filename, const char *text, size_t bytes) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        if (!writer_.joinable()) writer_ = std::thread(&Report::writer_loop, this);
        jobs_.push_back(Job{cb_filename, text, bytes});
        outstanding_++;
    }
    cv_.notify_all();
}

void Report::wait_pending(size_t max_outstanding) {
    std::unique_lock<std::mutex> lk(mu_);
    cv_.wait(lk, [&] { return outstanding_ <= max_outstanding; });
}

void Report::close() {
    if (writer_.joinable()) {
        wait_pending(0);
        {
            std::lock_guard<std::mutex> lk(mu_);
            stop_ = true;
        }
        cv_.notify_all();
        writer_.join();
        stop_ = false;
    }
    if (file_) { fclose(file_); file_ = nullptr; }
    if (file_csv_) { fclose(file_csv_); file_csv_ = nullptr; }
}

}  // namespace vqb200
