// vq_hostio.h -- the host-side I/O pipelines of the file-level entry points (ecoz2_vq_learn ...):
// a parallel `.prd` loader and the writer pool of the per-frame stage report.
//
// reference: prd_load, ecoz2/src/lpc/prd.c:34-74 (one fread per file, called in a loop by
// vq_learn.i:190-207); report_cbook, ecoz2/src/vq/report.c:109-123 ("<cbook>.min_dists.csv", N lines
// per stage).
//
// Once a stage of `vq learn` takes milliseconds these two ARE the run time of the entry point
// (10M frames, 2 -> 2048: 0.85 s of loading, 0.1 s of learning, 0.5-1.3 s of report writing when
// done naively).  Both are host-memory-bandwidth work that one thread cannot saturate:
//
//   loading   N reader threads pread() chunks of the files (any file, any offset: the global frame
//             position of a chunk is known from the headers) into pinned memory and copy them to
//             their place in the reserved frame set on streams of their own;
//   reports   ONE file does not fill faster with more threads (the inode serialises write(), the
//             mapping's page tree serialises faults: 2.4 GB/s with 8 threads), DIFFERENT files do
//             (8.6 GB/s with 8 single-threaded writers, scripts/exp/pagecache_fill.cpp).  So the
//             texts of the stages stay in HBM (Engine::min_dists_text_device, a pool of buffers),
//             the LBG run goes on, and N writer threads drain them: one thread per file, copying
//             chunks to pinned memory on a stream of its own (double buffered) and write()-ing them.
//
// Streams and pinned chunks (IoLanes) are created ONCE, up front, and shared by both pipelines:
// cudaMallocHost / cudaStreamCreate / cudaMalloc issued next to another thread's copies and
// launches were measured to block for 100-250 ms.
#pragma once
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "vq_engine.h"

namespace vqb200 {

// per I/O thread: two pinned chunk buffers; per (thread, device): a stream
class IoLanes {
  public:
    // threads = 0: ECOZ2_VQ_IO_THREADS, else min(8, half the host threads)
    IoLanes(const std::vector<int> &devices, int threads = 0, size_t chunk_bytes = 2u << 20);
    ~IoLanes();
    IoLanes(const IoLanes &) = delete;
    IoLanes &operator=(const IoLanes &) = delete;
    bool ready() const { return ready_; }   // false: pinned memory / streams could not be set up
    int threads() const { return n_threads_; }
    size_t chunk() const { return chunk_; }
    int device_slots() const { return n_slots_; }                        // 1 + highest device ordinal
    char *buffer(int t, int b) const { return pinned_ + ((size_t) t * 2 + b) * chunk_; }
    cudaStream_t stream(int t, int device) const { return streams_[(size_t) t * n_slots_ + device]; }
    const cudaStream_t *streams(int t) const { return &streams_[(size_t) t * n_slots_]; }

  private:
    int n_threads_ = 1, n_slots_ = 0;
    size_t chunk_ = 0;
    char *pinned_ = nullptr;
    std::vector<cudaStream_t> streams_;   // [thread][device ordinal], null where the engine has no shard
    bool ready_ = true;
};

// The frames of the `.prd` files, in file order, into the engine's RESERVED frame set
// (frames_reserve(sum of T) first; frames_commit after).  T[i] = frames of file i (from prd_info).
// Returns 0, or 1 + index of the file that could not be read, or -1 on a CUDA error.
int load_prd_frames(Engine &eng, IoLanes &lanes, const char *const files[], const int *T, int num_files, int P);

class DeviceTextWriter {
  public:
    DeviceTextWriter(Engine *eng, IoLanes *lanes) : eng_(eng), lanes_(lanes) {}
    ~DeviceTextWriter() { finish(); }
    // header + the segments in order -> file `name`; the segments go back to the engine when copied
    void submit(const std::string &name, const std::string &header, std::vector<Engine::TextSegment> segs);
    // waits for every submitted file, joins the threads; false if some file could not be written
    bool finish();

  private:
    struct Job { std::string name, header; std::vector<Engine::TextSegment> segs; };
    void worker(int t);
    bool write_job(const Job &job, int t);
    Engine *eng_;
    IoLanes *lanes_;
    std::vector<std::thread> threads_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Job> jobs_;
    bool stop_ = false;
    bool ok_ = true;
};

}  // namespace vqb200
