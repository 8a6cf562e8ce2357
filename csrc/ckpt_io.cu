// Checkpoint snapshot engine (replaces K22/K23): streams device buffers to a shard file (and
// back) through two pinned staging buffers, so the device->host copy of chunk k+1 overlaps the file write of
// chunk k, with several pwrite/pread threads per chunk.  The reference's `torch.save` pickles 873 CUDA tensors
// one by one through pageable memory into one file: 33.6 s for 48 GB (utils.py:75-80, logs/output_444664.out:94-95).
// Here every rank writes only its own slice (48 GB / W), in parallel, and fsyncs before returning — that is
// the "durable" point of the SIGUSR1 path.
//
// Host code only (no kernels); the copy stream is private and non-blocking, so a snapshot taken between steps
// does not serialise against the compute stream.  The staging buffers are allocated once (ftl_ckpt_init, called
// at trainer start) so the exit path itself does not allocate.
#include <cuda_runtime.h>
#include <errno.h>
#include <stdio.h>
#include <fcntl.h>
#include <stdint.h>
#include <string.h>
#include <unistd.h>

#include <atomic>
#include <chrono>
#include <thread>
#include <vector>

#include "common/util.cuh"

namespace {

struct Staging {
    void* buf[2] = {nullptr, nullptr};
    int64_t bytes = 0;
    cudaStream_t stream = nullptr;
    cudaEvent_t done[2] = {nullptr, nullptr};
};
Staging g_stage;
char g_err[256] = "";

double now_s() {
    return std::chrono::duration<double>(std::chrono::steady_clock::now().time_since_epoch()).count();
}

int ensure_staging(int64_t bytes) {
    if (g_stage.buf[0] && g_stage.bytes >= bytes) return 0;
    for (int i = 0; i < 2; ++i) {
        if (g_stage.buf[i]) cudaFreeHost(g_stage.buf[i]);
        cudaError_t e = cudaHostAlloc(&g_stage.buf[i], (size_t)bytes, cudaHostAllocDefault);
        if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "cudaHostAlloc(%lld): %s", (long long)bytes, cudaGetErrorString(e)); return (int)e; }
        if (!g_stage.done[i]) cudaEventCreateWithFlags(&g_stage.done[i], cudaEventDisableTiming);
    }
    if (!g_stage.stream) cudaStreamCreateWithFlags(&g_stage.stream, cudaStreamNonBlocking);
    g_stage.bytes = bytes;
    return 0;
}

// n bytes at file offset `off`, split over `nthreads` threads.  write=true: pwrite from buf, else pread into buf.
bool file_io(int fd, char* buf, int64_t n, int64_t off, int nthreads, bool write) {
    if (nthreads < 1) nthreads = 1;
    std::vector<std::thread> th;
    std::vector<int> ok(nthreads, 1);
    const int64_t per = ((n + nthreads - 1) / nthreads + 4095) / 4096 * 4096;
    for (int t = 0; t < nthreads; ++t) {
        const int64_t lo = (int64_t)t * per, hi = lo + per < n ? lo + per : n;
        if (lo >= hi) break;
        th.emplace_back([=, &ok]() {
            int64_t done = 0;
            while (done < hi - lo) {
                ssize_t r = write ? pwrite(fd, buf + lo + done, (size_t)(hi - lo - done), off + lo + done)
                                  : pread(fd, buf + lo + done, (size_t)(hi - lo - done), off + lo + done);
                if (r < 0 && errno == EINTR) continue;   // SIGUSR1/SIGTERM during a save must not break it
                if (r <= 0) { ok[t] = 0; return; }
                done += r;
            }
        });
    }
    for (auto& t : th) t.join();
    for (int v : ok) if (!v) return false;
    return true;
}

}  // namespace

FTL_API const char* ftl_ckpt_last_error(void) { return g_err; }

FTL_API int ftl_ckpt_init(int64_t stage_bytes) { return ensure_staging(stage_bytes); }

// Write `nseg` device segments into `path` at the given file offsets (the caller lays out the file and writes
// its own header).  timings[0] = seconds in device->host copies (exposed), [1] = seconds in file writes,
// [2] = fsync, [3] = total.  Returns 0 or an error code (message in ftl_ckpt_last_error()).
FTL_API int ftl_ckpt_write(const char* path, int nseg, const void* const* dev_ptrs, const int64_t* sizes,
                           const int64_t* file_offsets, int nthreads, int do_fsync, double* timings) {
    g_err[0] = 0;
    const double t_start = now_s();
    if (int e = ensure_staging(g_stage.bytes > 0 ? g_stage.bytes : (int64_t)256 << 20)) return e;
    const int fd = open(path, O_WRONLY | O_CREAT, 0644);
    if (fd < 0) { snprintf(g_err, sizeof g_err, "open(%s): %s", path, strerror(errno)); return -1; }
    double t_copy = 0, t_write = 0;
    std::thread writer[2];
    std::atomic<bool> io_ok{true};
    int k = 0;
    for (int s = 0; s < nseg && io_ok; ++s) {
        const char* src = static_cast<const char*>(dev_ptrs[s]);
        for (int64_t done = 0; done < sizes[s] && io_ok; ) {
            const int b = k & 1;
            if (writer[b].joinable()) { const double t0 = now_s(); writer[b].join(); t_write += now_s() - t0; }
            const int64_t n = sizes[s] - done < g_stage.bytes ? sizes[s] - done : g_stage.bytes;
            const double t0 = now_s();
            cudaError_t e = cudaMemcpyAsync(g_stage.buf[b], src + done, (size_t)n, cudaMemcpyDeviceToHost, g_stage.stream);
            if (e == cudaSuccess) e = cudaStreamSynchronize(g_stage.stream);
            t_copy += now_s() - t0;
            if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "D2H copy: %s", cudaGetErrorString(e)); io_ok = false; break; }
            const int64_t off = file_offsets[s] + done;
            char* hb = static_cast<char*>(g_stage.buf[b]);
            writer[b] = std::thread([=, &io_ok]() { if (!file_io(fd, hb, n, off, nthreads, true)) io_ok = false; });
            done += n;
            ++k;
        }
    }
    for (int b = 0; b < 2; ++b)
        if (writer[b].joinable()) { const double t0 = now_s(); writer[b].join(); t_write += now_s() - t0; }
    double t_sync = 0;
    if (io_ok && do_fsync) { const double t0 = now_s(); if (fsync(fd) != 0) io_ok = false; t_sync = now_s() - t0; }
    close(fd);
    if (!io_ok && !g_err[0]) snprintf(g_err, sizeof g_err, "write(%s): %s", path, strerror(errno));
    if (timings) { timings[0] = t_copy; timings[1] = t_write; timings[2] = t_sync; timings[3] = now_s() - t_start; }
    return io_ok ? 0 : -2;
}

// Inverse: file -> device segments.
FTL_API int ftl_ckpt_read(const char* path, int nseg, void* const* dev_ptrs, const int64_t* sizes,
                          const int64_t* file_offsets, int nthreads, double* timings) {
    g_err[0] = 0;
    const double t_start = now_s();
    if (int e = ensure_staging(g_stage.bytes > 0 ? g_stage.bytes : (int64_t)256 << 20)) return e;
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { snprintf(g_err, sizeof g_err, "open(%s): %s", path, strerror(errno)); return -1; }
    bool ok = true;
    int k = 0;
    for (int s = 0; s < nseg && ok; ++s) {
        char* dst = static_cast<char*>(dev_ptrs[s]);
        for (int64_t done = 0; done < sizes[s] && ok; ) {
            const int b = k & 1;
            cudaEventSynchronize(g_stage.done[b]);   // the H2D copy that last used this buffer has finished
            const int64_t n = sizes[s] - done < g_stage.bytes ? sizes[s] - done : g_stage.bytes;
            if (!file_io(fd, static_cast<char*>(g_stage.buf[b]), n, file_offsets[s] + done, nthreads, false)) { ok = false; break; }
            cudaError_t e = cudaMemcpyAsync(dst + done, g_stage.buf[b], (size_t)n, cudaMemcpyHostToDevice, g_stage.stream);
            if (e != cudaSuccess) { snprintf(g_err, sizeof g_err, "H2D copy: %s", cudaGetErrorString(e)); ok = false; break; }
            cudaEventRecord(g_stage.done[b], g_stage.stream);
            done += n;
            ++k;
        }
    }
    cudaStreamSynchronize(g_stage.stream);
    close(fd);
    if (!ok && !g_err[0]) snprintf(g_err, sizeof g_err, "read(%s): %s", path, strerror(errno));
    if (timings) timings[3] = now_s() - t_start;
    return ok ? 0 : -2;
}
