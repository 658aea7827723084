// vq_io.h -- ECOZ2 binary formats (.prd / .cbook / .seq), output-path rules and
// the stage report writers, byte/column compatible with the reference.
#pragma once
#include <stdint.h>
#include <stdio.h>

#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

namespace vqb200 {

constexpr int kFileIdentLen = 16;     // reference ecoz2/src/include/utl.h (FILE_IDENT_LEN)
constexpr int kMaxClassNameLen = 96;  // reference ecoz2/src/include/utl.h (MAX_CLASS_NAME_LEN)

// `.prd`  (reference ecoz2/src/lpc/prd.c:34-115):
//   "<predictor>" ident | class name | i32 P | i32 T | T x (1+P) f64
constexpr long kPrdDataOffset = kFileIdentLen + kMaxClassNameLen + 2 * 4;   // ident | class | P | T | data
int prd_info(const char *path, int *P, int *T, char class_name[kMaxClassNameLen]);
int prd_read(const char *path, double *out, long capacity_values);
int prd_save(const char *path, const char *class_name, const double *vectors, int T, int P);

// `.cbook` (reference ecoz2/src/vq/vq.c:22-116):
//   "<codebook>" ident | class name | i32 P | i32 M | M x (1+P) f64 reflections
int cbook_save(const char *path, const char *class_name, const double *refl, int M, int P);
int cbook_info(const char *path, int *P, int *M, char class_name[kMaxClassNameLen]);
int cbook_read(const char *path, double *out, long capacity_values);
extern int cb_errnum_last;  // reference cb_errnum (vq.c:20)

// `.seq` (reference ecoz2/src/hmm/symbol.c:17-36):
//   "<sequence>" ident | class name | i32 T | i32 M | T x u16
int seq_save(const char *path, const char *class_name, const uint16_t *seq, int T, int M);

// reference ecoz2/src/utl/fileutil.c:14-119
// "%f" of x appended at p (exact integer arithmetic, identical to printf); returns the end
constexpr int kMaxF6Chars = 320;   // "%f" of -DBL_MAX: sign + 309 digits + "." + 6 digits
char *format_f6(char *p, double x);   // writes at most kMaxF6Chars chars, no terminator
// "<cbook>.min_dists.csv" (report.c:109-123), formatted by all host threads
bool write_min_dists_csv(const char *name, const int32_t *codeword, const double *min_dist, long n);

std::string get_class_name(const std::string &path);
int mk_dirs(const std::string &path);
// data/<base_dir>/<class>/<stem><ext>; creates the directory
std::string get_output_filename(const std::string &from, const std::string &base_dir,
                                const std::string &ext);

// reference ecoz2/src/vq/report.c:14-135
class Report {
  public:
    void prepare(const std::string &rpt_filename, long tot_vecs, double eps);
    void cbook(const std::string &cb_filename, int num_pas, double DDprm, double sigma,
               double inertia, int M, const int32_t *cardd, const double *discel,
               const int32_t *codeword, const double *min_dist /* both may be null */);
    // "<cbook>.min_dists.csv" from text that is already formatted (Engine::min_dists_text): handed to
    // a writer thread, so that the next stage runs while the file is written.  `text` must stay valid
    // until wait_pending() says at most one LATER job is outstanding (the engine double-buffers it).
    void min_dists_submit(const std::string &cb_filename, const char *text, size_t bytes);
    void wait_pending(size_t max_outstanding);
    void close();   // joins the writer
    ~Report() { close(); }

  private:
    struct Job { std::string name; const char *text; size_t bytes; };
    void writer_loop();
    long tot_vecs_ = 0;
    FILE *file_ = nullptr;
    FILE *file_csv_ = nullptr;
    std::thread writer_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Job> jobs_;
    size_t outstanding_ = 0;
    bool stop_ = false;
};

// header line + `bytes` of ready-made lines -> file, the copy split over several threads (pwrite)
bool write_text_file(const char *name, const char *header, const char *text, size_t bytes);

}  // namespace vqb200
