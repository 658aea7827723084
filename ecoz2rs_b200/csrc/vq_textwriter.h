// vq_textwriter.h -- report files from text that was formatted on the device and is still there.
//
// reference: report_cbook, ecoz2/src/vq/report.c:109-123 ("<cbook>.min_dists.csv", N lines per stage).
// A stage of `vq learn` takes milliseconds, its 120 MB report file (10M frames) 50 ms or more of
// page-cache filling -- and ONE file does not fill faster with more threads (the inode serialises
// write(), the mapping's page tree serialises faults: 2.4 GB/s with 8 threads), while DIFFERENT
// files do (8.6 GB/s with 8 single-threaded writers, scripts/exp/pagecache_fill.cpp).  So the texts
// of the stages stay in HBM (Engine::min_dists_text_device, a pool of buffers), the LBG run goes
// on, and a pool of writer threads drains them: one thread per file, copying 4 MB chunks to pinned
// memory on a stream of its own (double buffered) and write()-ing them.  Streams and pinned chunks
// are created up front by the constructor: nothing is allocated while the LBG run is going.
#pragma once
#include <condition_variable>
#include <deque>
#include <mutex>
#include <string>
#include <thread>
#include <vector>

#include "vq_engine.h"

namespace vqb200 {

// per (writer thread, device): a stream and two pinned chunk buffers
struct WriterLane {
    cudaStream_t stream = nullptr;
    char *h[2] = {nullptr, nullptr};
};

class DeviceTextWriter {
  public:
    // threads = 0: ECOZ2_VQ_REPORT_THREADS, else min(8, half the host threads)
    explicit DeviceTextWriter(Engine *eng, int threads = 0);
    ~DeviceTextWriter();
    bool ready() const { return ready_; }   // false: pinned memory / streams could not be set up
    // header + the segments in order -> file `name`; the segments go back to the engine when copied
    void submit(const std::string &name, const std::string &header, std::vector<Engine::TextSegment> segs);
    // waits for every submitted file, joins the threads; false if some file could not be written
    bool finish();

  private:
    struct Job { std::string name, header; std::vector<Engine::TextSegment> segs; };
    void worker(int t);
    bool write_job(const Job &job, WriterLane *lanes);
    Engine *eng_;
    int n_threads_;
    int n_lanes_ = 0;                    // 1 + highest device ordinal of the engine
    std::vector<int> devices_;
    std::vector<WriterLane> lanes_;      // [thread][device]
    void *pinned_ = nullptr;             // all chunk buffers, one allocation
    bool ready_ = true;
    std::vector<std::thread> threads_;
    std::mutex mu_;
    std::condition_variable cv_;
    std::deque<Job> jobs_;
    bool stop_ = false;
    bool ok_ = true;
};

}  // namespace vqb200
