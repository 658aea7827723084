// vq_textwriter.cpp -- see vq_textwriter.h
#include "vq_textwriter.h"

#include <chrono>
#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

namespace vqb200 {

namespace {
constexpr size_t kChunk = 2u << 20;
}  // namespace

// per (writer thread, device): a stream and two pinned chunk buffers, all created by the
// constructor -- a writer thread makes no allocating CUDA call (cudaMallocHost / cudaStreamCreate
// next to another thread's copies and launches were measured to block for 100-250 ms)

DeviceTextWriter::DeviceTextWriter(Engine *eng, int threads) : eng_(eng) {
    if (threads <= 0) {
        const char *v = getenv("ECOZ2_VQ_REPORT_THREADS");
        threads = (v && *v) ? atoi(v) : 0;
    }
    if (threads <= 0) {
        const unsigned hw = std::thread::hardware_concurrency();
        threads = (int) (hw / 2);
        if (threads > 8) threads = 8;
    }
    n_threads_ = threads < 1 ? 1 : threads;
    devices_ = eng->devices();
    int max_dev = 0;
    for (int d : devices_) max_dev = d > max_dev ? d : max_dev;
    n_lanes_ = max_dev + 1;
    lanes_.assign((size_t) n_threads_ * n_lanes_, WriterLane());
    const size_t per_lane = 2 * kChunk;
    if (cudaHostAlloc(&pinned_, per_lane * n_threads_ * devices_.size(), cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        pinned_ = nullptr;
        ready_ = false;
        return;
    }
    int prev = 0;
    cudaGetDevice(&prev);
    size_t slice = 0;
    for (int t = 0; t < n_threads_; t++)
        for (int d : devices_) {
            WriterLane &l = lanes_[(size_t) t * n_lanes_ + d];
            cudaSetDevice(d);
            if (cudaStreamCreateWithFlags(&l.stream, cudaStreamNonBlocking) != cudaSuccess) ready_ = false;
            l.h[0] = static_cast<char *>(pinned_) + slice * per_lane;
            l.h[1] = l.h[0] + kChunk;
            slice++;
        }
    cudaSetDevice(prev);
}

DeviceTextWriter::~DeviceTextWriter() {
    finish();
    for (WriterLane &l : lanes_)
        if (l.stream) cudaStreamDestroy(l.stream);
    if (pinned_) cudaFreeHost(pinned_);
}

void DeviceTextWriter::submit(const std::string &name, const std::string &header,
                              std::vector<Engine::TextSegment> segs) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        jobs_.push_back(Job{name, header, std::move(segs)});
        if ((int) threads_.size() < n_threads_)   // one more writer per file until the pool is full
            threads_.emplace_back(&DeviceTextWriter::worker, this, (int) threads_.size());
    }
    cv_.notify_one();
}

bool DeviceTextWriter::finish() {
    {
        std::lock_guard<std::mutex> lk(mu_);
        stop_ = true;
    }
    cv_.notify_all();
    for (auto &t : threads_) t.join();
    threads_.clear();
    std::lock_guard<std::mutex> lk(mu_);
    stop_ = false;
    const bool ok = ok_;
    ok_ = true;
    return ok;
}

void DeviceTextWriter::worker(int t) {
    WriterLane *lanes = &lanes_[(size_t) t * n_lanes_];
    for (;;) {
        Job job;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || !jobs_.empty(); });
            if (jobs_.empty()) break;       // stop_ and nothing left
            job = std::move(jobs_.front());
            jobs_.pop_front();
        }
        const bool ok = write_job(job, lanes);
        for (const Engine::TextSegment &g : job.segs) eng_->release_text(g);
        if (!ok) {
            printf("error creating %s\n", job.name.c_str());
            std::lock_guard<std::mutex> lk(mu_);
            ok_ = false;
        }
    }
}

static bool write_all(int fd, const char *p, size_t n) {
    while (n > 0) {
        const ssize_t w = write(fd, p, n);
        if (w <= 0) return false;
        p += w;
        n -= (size_t) w;
    }
    return true;
}

static double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

bool DeviceTextWriter::write_job(const Job &job, WriterLane *lanes) {
    static const bool timing = getenv("ECOZ2_VQ_TIMING") != nullptr;
    const double t0 = now_s();
    double t_sync = 0, t_write = 0;
    const int fd = open(job.name.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return false;
    bool ok = write_all(fd, job.header.data(), job.header.size());
    for (const Engine::TextSegment &g : job.segs) {
        if (!ok) break;
        if (g.device < 0 || g.device >= n_lanes_ || !lanes[g.device].stream ||
            cudaSetDevice(g.device) != cudaSuccess) { ok = false; break; }
        WriterLane &l = lanes[g.device];
        // the text is complete when the engine's stream has passed `ready`
        if (cudaStreamWaitEvent(l.stream, g.ready, 0) != cudaSuccess) { ok = false; break; }
        const size_t n_chunks = (g.bytes + kChunk - 1) / kChunk;
        auto issue = [&](size_t k) {
            const size_t lo = k * kChunk;
            const size_t len = g.bytes - lo < kChunk ? g.bytes - lo : kChunk;
            return cudaMemcpyAsync(l.h[k & 1], g.d_text + lo, len, cudaMemcpyDeviceToHost, l.stream) == cudaSuccess;
        };
        if (n_chunks > 0 && !issue(0)) { ok = false; break; }
        for (size_t k = 0; k < n_chunks && ok; k++) {
            const double ts = now_s();
            if (cudaStreamSynchronize(l.stream) != cudaSuccess) { ok = false; break; }
            if (k + 1 < n_chunks && !issue(k + 1)) { ok = false; break; }   // copies while chunk k is written
            const size_t lo = k * kChunk;
            const size_t len = g.bytes - lo < kChunk ? g.bytes - lo : kChunk;
            const double tw = now_s();
            ok = write_all(fd, l.h[k & 1], len);
            t_sync += tw - ts;
            t_write += now_s() - tw;
        }
        if (!ok) cudaStreamSynchronize(l.stream);   // nothing of ours in flight when the buffer goes back
    }
    const bool closed = ::close(fd) == 0;
    if (timing)
        fprintf(stderr, "[timing]   writer: %s %.1f ms (waiting for copies %.1f, write() %.1f)\n",
                job.name.c_str(), (now_s() - t0) * 1e3, t_sync * 1e3, t_write * 1e3);
    return closed && ok;
}

}  // namespace vqb200
