// dev experiment: how fast can 120 MB text files be created on tmpfs?
// g++ -O2 -std=c++17 -pthread pagecache_fill.cpp -o pagecache_fill && ./pagecache_fill /dev/shm/pcf 11 120
#include <chrono>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <fcntl.h>
#include <string>
#include <sys/mman.h>
#include <thread>
#include <unistd.h>
#include <vector>
static double now() { return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static void mm_fill(const char *name, const char *src, size_t bytes, int T) {
    int fd = open(name, O_RDWR | O_CREAT | O_TRUNC, 0644);
    if (ftruncate(fd, bytes)) abort();
    char *dst = (char *) mmap(nullptr, bytes, PROT_READ | PROT_WRITE, MAP_SHARED, fd, 0);
    size_t per = ((bytes + T - 1) / T + 4095) & ~(size_t) 4095;
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back([=] { size_t lo = t * per; if (lo >= bytes) return; size_t hi = lo + per < bytes ? lo + per : bytes; memcpy(dst + lo, src + lo, hi - lo); });
    for (auto &x : th) x.join();
    munmap(dst, bytes); close(fd);
}
static void wr_fill(const char *name, const char *src, size_t bytes, size_t chunk) {
    int fd = open(name, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    for (size_t lo = 0; lo < bytes;) { size_t c = bytes - lo < chunk ? bytes - lo : chunk; ssize_t w = write(fd, src + lo, c); if (w <= 0) abort(); lo += w; }
    close(fd);
}
static void pw_fill(const char *name, const char *src, size_t bytes, int T) {
    int fd = open(name, O_WRONLY | O_CREAT | O_TRUNC, 0644);
    size_t per = ((bytes + T - 1) / T + 4095) & ~(size_t) 4095;
    std::vector<std::thread> th;
    for (int t = 0; t < T; t++) th.emplace_back([=] { size_t lo = t * per; if (lo >= bytes) return; size_t hi = lo + per < bytes ? lo + per : bytes; while (lo < hi) { ssize_t w = pwrite(fd, src + lo, hi - lo, lo); if (w <= 0) abort(); lo += w; } });
    for (auto &x : th) x.join();
    close(fd);
}
int main(int argc, char **argv) {
    std::string dir = argv[1]; int nf = atoi(argv[2]); size_t mb = atoi(argv[3]); size_t bytes = mb << 20;
    int T = argc > 4 ? atoi(argv[4]) : 8;
    char *src = (char *) malloc(bytes); memset(src, 'x', bytes);
    auto names = [&](int i) { return dir + "_" + std::to_string(i); };
    auto cleanup = [&] { for (int i = 0; i < nf; i++) unlink(names(i).c_str()); };
    double t0;
    cleanup(); t0 = now(); for (int i = 0; i < nf; i++) mm_fill(names(i).c_str(), src, bytes, T);
    printf("mmap %d thr, files in sequence: %.3f s  %.2f GB/s\n", T, now() - t0, nf * bytes / (now() - t0) / 1e9);
    cleanup(); t0 = now(); for (int i = 0; i < nf; i++) wr_fill(names(i).c_str(), src, bytes, 1 << 20);
    printf("write 1 thr: %.3f s  %.2f GB/s\n", now() - t0, nf * bytes / (now() - t0) / 1e9);
    cleanup(); t0 = now(); for (int i = 0; i < nf; i++) pw_fill(names(i).c_str(), src, bytes, T);
    printf("pwrite %d thr one file: %.3f s  %.2f GB/s\n", T, now() - t0, nf * bytes / (now() - t0) / 1e9);
    cleanup(); t0 = now();
    { std::vector<std::thread> th; for (int i = 0; i < nf; i++) th.emplace_back([&, i] { wr_fill(names(i).c_str(), src, bytes, 1 << 20); }); for (auto &x : th) x.join(); }
    printf("write, %d files in parallel (1 thr each): %.3f s  %.2f GB/s\n", nf, now() - t0, nf * bytes / (now() - t0) / 1e9);
    cleanup(); t0 = now();
    { std::vector<std::thread> th; for (int i = 0; i < nf; i++) th.emplace_back([&, i] { mm_fill(names(i).c_str(), src, bytes, 2); }); for (auto &x : th) x.join(); }
    printf("mmap 2 thr, %d files in parallel: %.3f s  %.2f GB/s\n", nf, now() - t0, nf * bytes / (now() - t0) / 1e9);
    // rewrite of existing files (pages already in the cache)
    t0 = now(); for (int i = 0; i < nf; i++) { int fd = open(names(i).c_str(), O_WRONLY); size_t lo = 0; while (lo < bytes) { ssize_t w = pwrite(fd, src + lo, bytes - lo, lo); lo += w; } close(fd); }
    printf("rewrite existing pages, 1 thr: %.3f s  %.2f GB/s\n", now() - t0, nf * bytes / (now() - t0) / 1e9);
    cleanup();
    return 0;
}
