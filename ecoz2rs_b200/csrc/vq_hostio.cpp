// vq_hostio.cpp -- see vq_hostio.h
#include "vq_hostio.h"

#include "vq_io.h"

#include <algorithm>
#include <atomic>
#include <chrono>
#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

namespace vqb200 {

IoLanes::IoLanes(const std::vector<int> &devices, int threads, size_t chunk_bytes) : chunk_(chunk_bytes) {
    if (threads <= 0) {
        const char *v = getenv("ECOZ2_VQ_IO_THREADS");
        threads = (v && *v) ? atoi(v) : 0;
    }
    if (threads <= 0) {
        const unsigned hw = std::thread::hardware_concurrency();
        threads = (int) (hw / 2);
        if (threads > 8) threads = 8;
    }
    n_threads_ = threads < 1 ? 1 : threads;
    int max_dev = 0;
    for (int d : devices) max_dev = d > max_dev ? d : max_dev;
    n_slots_ = max_dev + 1;
    streams_.assign((size_t) n_threads_ * n_slots_, nullptr);
    void *p = nullptr;
    if (cudaHostAlloc(&p, 2 * chunk_ * n_threads_, cudaHostAllocPortable) != cudaSuccess) {
        cudaGetLastError();
        ready_ = false;
        return;
    }
    pinned_ = static_cast<char *>(p);
    int prev = 0;
    cudaGetDevice(&prev);
    for (int t = 0; t < n_threads_; t++)
        for (int d : devices) {
            cudaStream_t &st = streams_[(size_t) t * n_slots_ + d];
            if (st) continue;
            cudaSetDevice(d);
            if (cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking) != cudaSuccess) {
                cudaGetLastError();
                st = nullptr;
                ready_ = false;
            }
        }
    cudaSetDevice(prev);
}

IoLanes::~IoLanes() {
    for (cudaStream_t st : streams_)
        if (st) cudaStreamDestroy(st);
    if (pinned_) cudaFreeHost(pinned_);
}

// ---------------------------------------------------------------------------------------------
// parallel .prd loader
int load_prd_frames(Engine &eng, IoLanes &lanes, const char *const files[], const int *T, int num_files, int P) {
    const size_t row = sizeof(double) * (size_t) (P + 1);
    const long per_chunk = (long) (lanes.chunk() / row);
    if (per_chunk < 1) return -1;
    struct Item { int file; long first, count, pos; };
    std::vector<Item> items;
    long pos = 0;
    for (int i = 0; i < num_files; i++) {
        for (long f = 0; f < T[i]; f += per_chunk) {
            const long c = T[i] - f < per_chunk ? T[i] - f : per_chunk;
            items.push_back(Item{i, f, c, pos});
            pos += c;
        }
    }
    std::atomic<size_t> next(0);
    std::atomic<int> failed(0);   // 0 ok, 1 + file index, -1 CUDA
    auto work = [&](int t) {
        int fd = -1, fd_file = -1;
        char *buf = lanes.buffer(t, 0);
        for (;;) {
            const size_t k = next.fetch_add(1);
            if (k >= items.size() || failed.load() != 0) break;
            const Item &it = items[k];
            if (fd_file != it.file) {
                if (fd >= 0) ::close(fd);
                fd = open(files[it.file], O_RDONLY);
                fd_file = it.file;
                if (fd < 0) { failed.store(1 + it.file); break; }
            }
            size_t want = (size_t) it.count * row, got = 0;
            const off_t off = (off_t) kPrdDataOffset + (off_t) it.first * (off_t) row;
            while (got < want) {
                const ssize_t r = pread(fd, buf + got, want - got, off + (off_t) got);
                if (r <= 0) break;
                got += (size_t) r;
            }
            if (got != want) { failed.store(1 + it.file); break; }
            // to its place in the frame set; the buffer is reused only when the copy is done
            if (eng.frames_put_host(it.pos, reinterpret_cast<const double *>(buf), it.count, lanes.streams(t))) {
                failed.store(-1);
                break;
            }
        }
        if (fd >= 0) ::close(fd);
    };
    std::vector<std::thread> th;
    const int n_thr = (int) std::min<size_t>((size_t) lanes.threads(), items.size());
    for (int t = 1; t < n_thr; t++) th.emplace_back(work, t);
    work(0);
    for (auto &x : th) x.join();
    if (failed.load() != 0) return failed.load();
    return eng.frames_mark_filled(pos) ? -1 : 0;
}

// ---------------------------------------------------------------------------------------------
// report writer pool
void DeviceTextWriter::submit(const std::string &name, const std::string &header,
                              std::vector<Engine::TextSegment> segs) {
    {
        std::lock_guard<std::mutex> lk(mu_);
        jobs_.push_back(Job{name, header, std::move(segs)});
        if ((int) threads_.size() < lanes_->threads())   // one more writer per file until the pool is full
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
    for (;;) {
        Job job;
        {
            std::unique_lock<std::mutex> lk(mu_);
            cv_.wait(lk, [&] { return stop_ || !jobs_.empty(); });
            if (jobs_.empty()) break;       // stop_ and nothing left
            job = std::move(jobs_.front());
            jobs_.pop_front();
        }
        const bool ok = write_job(job, t);
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

bool DeviceTextWriter::write_job(const Job &job, int t) {
    const size_t kChunk = lanes_->chunk();
    char *const h[2] = {lanes_->buffer(t, 0), lanes_->buffer(t, 1)};
    static const bool timing = getenv("ECOZ2_VQ_TIMING") != nullptr;
    const double t0 = now_s();
    double t_sync = 0, t_write = 0;
    const int fd = open(job.name.c_str(), O_WRONLY | O_CREAT | O_TRUNC, 0644);
    if (fd < 0) return false;
    bool ok = write_all(fd, job.header.data(), job.header.size());
    for (const Engine::TextSegment &g : job.segs) {
        if (!ok) break;
        if (g.device < 0 || g.device >= lanes_->device_slots() || !lanes_->stream(t, g.device) ||
            cudaSetDevice(g.device) != cudaSuccess) { ok = false; break; }
        const cudaStream_t st = lanes_->stream(t, g.device);
        // the text is complete when the engine's stream has passed `ready`
        if (cudaStreamWaitEvent(st, g.ready, 0) != cudaSuccess) { ok = false; break; }
        const size_t n_chunks = (g.bytes + kChunk - 1) / kChunk;
        auto issue = [&](size_t k) {
            const size_t lo = k * kChunk;
            const size_t len = g.bytes - lo < kChunk ? g.bytes - lo : kChunk;
            return cudaMemcpyAsync(h[k & 1], g.d_text + lo, len, cudaMemcpyDeviceToHost, st) == cudaSuccess;
        };
        if (n_chunks > 0 && !issue(0)) { ok = false; break; }
        for (size_t k = 0; k < n_chunks && ok; k++) {
            const double ts = now_s();
            if (cudaStreamSynchronize(st) != cudaSuccess) { ok = false; break; }
            if (k + 1 < n_chunks && !issue(k + 1)) { ok = false; break; }   // copies while chunk k is written
            const size_t lo = k * kChunk;
            const size_t len = g.bytes - lo < kChunk ? g.bytes - lo : kChunk;
            const double tw = now_s();
            ok = write_all(fd, h[k & 1], len);
            t_sync += tw - ts;
            t_write += now_s() - tw;
        }
        if (!ok) cudaStreamSynchronize(st);   // nothing of ours in flight when the buffer goes back
    }
    const bool closed = ::close(fd) == 0;
    if (timing)
        fprintf(stderr, "[timing]   writer: %s %.1f ms (waiting for copies %.1f, write() %.1f)\n",
                job.name.c_str(), (now_s() - t0) * 1e3, t_sync * 1e3, t_write * 1e3);
    return closed && ok;
}

}  // namespace vqb200
