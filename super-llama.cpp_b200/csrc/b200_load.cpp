// b200_load.cpp -- weight load path: file -> pinned staging ring -> device, repack overlapped, per-GPU sharded reads.
//
// What it replaces: the upload loop of llama_model_loader::load_all_data (reference src/llama-model-loader.cpp:1136-1186) -- ONE host thread
// that alternates between read() into one of four 1 MiB pinned buffers and ggml_backend_tensor_set_async, so the file read and the PCIe copy
// never overlap for longer than one chunk and the device-side re-layout of a weight (ggml-cuda has none; ours: b200_repack) would wait for
// the whole tensor -- and, for split-mode=row, ggml_backend_cuda_split_buffer_set_tensor (ggml/src/ggml-cuda/ggml-cuda.cu:900-937), which
// receives the WHOLE tensor on the host and copies each device's row slice from it.
// Here: reader threads pread chunk after chunk into a ring of pinned buffers while the calling thread issues the host-to-device copies on
// a copy stream of its own; a buffer goes back to the reader when its copy's event has completed.  When the last chunk of a tensor has been
// issued, its repack is queued on the context's compute stream behind an event of the copy stream, so it runs while the next tensor's
// bytes are still arriving.  An item names a byte range of the file and a byte offset inside the destination, so a rank (or a device of the
// split buffer type) reads only the rows it owns: b200_load_row_slice() turns (row_lo, row_hi) of b200_split_row_range into such an item.
// Host code only; no kernel of its own (b200_repack launches the permutation kernels of b200_repack.cu).
#include "b200_common.cuh"
#include <atomic>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>
#include <chrono>
#include <errno.h>
#include <fcntl.h>
#include <stdlib.h>
#include <string.h>
#include <unistd.h>

namespace {
struct chunk { int item; size_t off, n; };      // bytes [off, off + n) of item's range
struct ring {
    std::mutex m; std::condition_variable cv;
    std::vector<int> filled, freed;             // buffer indices: read and waiting for their copy / copied and free again
    bool failed = false; int err = 0;
};
}

extern "C" int b200_load_row_slice(int type, int64_t ne0, int64_t file_offset, int64_t row_lo, int64_t row_hi, b200_load_item * item) {
    const int bs = b200_blck_size(type);
    if (!item || bs <= 0 || ne0 <= 0 || ne0 % bs != 0 || row_lo < 0 || row_hi < row_lo) { b200_set_error("load_row_slice: bad arguments"); return B200_ERR_INVALID; }
    const size_t rb = b200_row_size(type, ne0);
    item->file_offset = file_offset + (int64_t) (rb * (size_t) row_lo);
    item->nbytes = rb * (size_t) (row_hi - row_lo);
    item->dst_offset = 0;                        // the slice is a tensor of its own on the device that owns it
    return B200_OK;
}

extern "C" int b200_load_file(b200_ctx * ctx, const char * path, const b200_load_item * items, int n_items, size_t chunk_bytes, int n_buffers,
                              b200_load_stats * stats) {
    if (!ctx || !path || (n_items > 0 && !items) || n_items < 0) { b200_set_error("load_file: null argument"); return B200_ERR_INVALID; }
    if (chunk_bytes == 0) chunk_bytes = 4u << 20;
    if (n_buffers <= 0) n_buffers = 8;
    if (chunk_bytes < 4096 || chunk_bytes > (1u << 30) || n_buffers < 2 || n_buffers > 64) { b200_set_error("load_file: chunk_bytes %zu (4 KiB .. 1 GiB), n_buffers %d (2 .. 64)", chunk_bytes, n_buffers); return B200_ERR_INVALID; }
    for (int i = 0; i < n_items; ++i) {
        const b200_load_item & it = items[i];
        if (!it.dst || !it.dst->data || it.file_offset < 0) { b200_set_error("load_file: item %d has no destination / a negative offset", i); return B200_ERR_INVALID; }
        if (it.dst_offset + it.nbytes > b200_tensor_span(it.dst)) { b200_set_error("load_file: item %d writes %zu bytes at %zu into a tensor of %zu", i, it.nbytes, it.dst_offset, b200_tensor_span(it.dst)); return B200_ERR_INVALID; }
        if (it.repack && (it.dst_offset != 0 || it.nbytes != b200_tensor_span(it.dst) || !b200_repack_supported(it.dst->type, it.dst->ne[0]))) {
            b200_set_error("load_file: item %d asks for a repack of a partial write or of a type / row length without a device layout", i); return B200_ERR_INVALID;
        }
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int fd = open(path, O_RDONLY);
    if (fd < 0) { b200_set_error("load_file: %s: %s", path, strerror(errno)); return B200_ERR_INVALID; }

    std::vector<chunk> plan;
    size_t total = 0;
    for (int i = 0; i < n_items; ++i) {
        for (size_t off = 0; off < items[i].nbytes; off += chunk_bytes) plan.push_back({ i, off, std::min(chunk_bytes, items[i].nbytes - off) });
        total += items[i].nbytes;
    }
    std::vector<void *> buf(n_buffers, nullptr);
    std::vector<cudaEvent_t> done(n_buffers, nullptr);
    cudaStream_t copy = nullptr;
    cudaEvent_t tensor_in = nullptr;
    int rc = B200_OK;
    auto cleanup = [&]() {
        for (void * b : buf) if (b) cudaFreeHost(b);
        for (cudaEvent_t e : done) if (e) cudaEventDestroy(e);
        if (tensor_in) cudaEventDestroy(tensor_in);
        if (copy) cudaStreamDestroy(copy);
        close(fd);
    };
#define LD_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { b200_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); rc = B200_ERR_CUDA; } } while (0)
    for (int b = 0; b < n_buffers && rc == B200_OK; ++b) { LD_TRY(cudaHostAlloc(&buf[b], chunk_bytes, cudaHostAllocDefault)); if (rc == B200_OK) LD_TRY(cudaEventCreateWithFlags(&done[b], cudaEventDisableTiming)); }
    if (rc == B200_OK) LD_TRY(cudaStreamCreateWithFlags(&copy, cudaStreamNonBlocking));
    if (rc == B200_OK) LD_TRY(cudaEventCreateWithFlags(&tensor_in, cudaEventDisableTiming));
    if (rc != B200_OK) { cleanup(); return rc; }

    const auto t0 = std::chrono::steady_clock::now();
    // GGML_B200_LOAD_SERIAL=1 (measurement only): the shape of the reference's loop -- ONE thread alternating between read() and the async copy,
    // everything on one stream, the repack behind the tensor's last chunk on that same stream (what the plugin's buffer hook does today)
    static const bool serial = getenv("GGML_B200_LOAD_SERIAL") && atoi(getenv("GGML_B200_LOAD_SERIAL"));
    if (serial) {
        double rs = 0.0; size_t sent = 0;
        for (size_t c = 0; c < plan.size() && rc == B200_OK; ++c) {
            const int b = (int) (c % (size_t) n_buffers);
            const b200_load_item & it = items[plan[c].item];
            if (c >= (size_t) n_buffers) LD_TRY(cudaEventSynchronize(done[b]));
            const auto r0 = std::chrono::steady_clock::now();
            size_t got = 0;
            while (got < plan[c].n && rc == B200_OK) {
                const ssize_t k = pread(fd, (char *) buf[b] + got, plan[c].n - got, (off_t) (it.file_offset + (int64_t) (plan[c].off + got)));
                if (k < 0 && errno == EINTR) continue;
                if (k <= 0) { b200_set_error("load_file: %s: read failed: %s", path, strerror(k < 0 ? errno : EIO)); rc = B200_ERR_INVALID; break; }
                got += (size_t) k;
            }
            rs += std::chrono::duration<double>(std::chrono::steady_clock::now() - r0).count();
            if (rc != B200_OK) break;
            LD_TRY(cudaMemcpyAsync((char *) it.dst->data + it.dst_offset + plan[c].off, buf[b], plan[c].n, cudaMemcpyHostToDevice, ctx->stream));
            if (rc == B200_OK) LD_TRY(cudaEventRecord(done[b], ctx->stream));
            sent += plan[c].n;
            if (rc == B200_OK && plan[c].off + plan[c].n == it.nbytes) {
                if (it.repack) { rc = b200_repack(ctx, it.dst->type, it.dst->data, it.dst->ne[0], it.dst->ne[1] * it.dst->ne[2] * it.dst->ne[3], it.dst->nb[1]); if (rc == B200_OK) it.dst->flags |= B200_TENSOR_REPACKED; }
                b200_note_write(ctx, it.dst->data, b200_tensor_span(it.dst));
            }
        }
        const cudaError_t es = cudaStreamSynchronize(ctx->stream);
        if (rc == B200_OK && es != cudaSuccess) { b200_set_error("load_file: %s", cudaGetErrorString(es)); rc = B200_ERR_CUDA; }
        if (stats) { stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count(); stats->read_seconds = rs; stats->bytes = (uint64_t) sent; stats->chunks = (uint64_t) plan.size(); }
        cleanup();
        return rc;
    }
    ring R;
    for (int b = 0; b < n_buffers; ++b) R.freed.push_back(b);
    // the readers: one pread from the page cache is a single core's memcpy (~11 GB/s measured on the B200 boxes' hosts), a fifth of what the
    // PCIe link takes, so several threads read different chunks at once (GGML_B200_LOAD_THREADS, default 4, at most n_buffers - 1).  Chunks are
    // claimed in plan order and may complete out of order; the copies go out in the order the chunks arrive
    static const int want_threads = getenv("GGML_B200_LOAD_THREADS") ? atoi(getenv("GGML_B200_LOAD_THREADS")) : 4;
    const int n_readers = std::max(1, std::min(std::min(want_threads, n_buffers - 1), (int) plan.size()));
    std::atomic<size_t> next(0);
    std::vector<double> read_t(n_readers, 0.0);
    std::vector<std::pair<int, size_t>> filled;                      // (buffer, chunk) read and waiting for its copy; guarded by R.m
    auto read_loop = [&](int me) {
        while (true) {
            const size_t c = next.fetch_add(1);
            if (c >= plan.size()) return;
            int b;
            {
                std::unique_lock<std::mutex> lk(R.m);
                R.cv.wait(lk, [&] { return !R.freed.empty() || R.failed; });
                if (R.failed) return;
                b = R.freed.front(); R.freed.erase(R.freed.begin());
            }
            const auto r0 = std::chrono::steady_clock::now();
            size_t got = 0; int err = 0;
            while (got < plan[c].n) {
                const ssize_t k = pread(fd, (char *) buf[b] + got, plan[c].n - got, (off_t) (items[plan[c].item].file_offset + (int64_t) (plan[c].off + got)));
                if (k < 0 && errno == EINTR) continue;
                if (k <= 0) { err = k < 0 ? errno : EIO; break; }     // (0: the file ends inside the range)
                got += (size_t) k;
            }
            read_t[me] += std::chrono::duration<double>(std::chrono::steady_clock::now() - r0).count();
            std::lock_guard<std::mutex> lk(R.m);
            if (err) { R.failed = true; R.err = err; R.cv.notify_all(); return; }
            filled.push_back({ b, c });
            R.cv.notify_all();
        }
    };
    std::vector<std::thread> readers;
    for (int t = 0; t < n_readers; ++t) readers.emplace_back(read_loop, t);
    // the uploader (this thread)
    std::vector<int> in_flight;                  // buffers whose copy has been issued, oldest first
    std::vector<size_t> left(n_items, 0);        // bytes of an item whose copy has not been issued yet
    for (int i = 0; i < n_items; ++i) left[i] = items[i].nbytes;
    size_t issued = 0;
    for (size_t k = 0; k < plan.size() && rc == B200_OK; ++k) {
        int b = -1; size_t c = 0;
        while (b < 0) {
            // hand back every buffer whose copy has finished; block on the oldest one only when the readers have nothing for us and no buffer is free
            while (!in_flight.empty() && cudaEventQuery(done[in_flight.front()]) == cudaSuccess) {
                std::lock_guard<std::mutex> lk(R.m);
                R.freed.push_back(in_flight.front()); in_flight.erase(in_flight.begin()); R.cv.notify_all();
            }
            std::unique_lock<std::mutex> lk(R.m);
            if (R.failed) { b200_set_error("load_file: %s: read failed: %s", path, strerror(R.err)); rc = B200_ERR_INVALID; break; }
            if (!filled.empty()) { b = filled.front().first; c = filled.front().second; filled.erase(filled.begin()); break; }
            if ((int) in_flight.size() == n_buffers) {        // every buffer is waiting for its copy: wait for the oldest
                lk.unlock();
                LD_TRY(cudaEventSynchronize(done[in_flight.front()]));
                if (rc != B200_OK) break;
                continue;
            }
            R.cv.wait_for(lk, std::chrono::microseconds(100));
        }
        if (rc != B200_OK) break;
        const b200_load_item & it = items[plan[c].item];
        LD_TRY(cudaMemcpyAsync((char *) it.dst->data + it.dst_offset + plan[c].off, buf[b], plan[c].n, cudaMemcpyHostToDevice, copy));
        if (rc == B200_OK) LD_TRY(cudaEventRecord(done[b], copy));
        in_flight.push_back(b);
        issued += plan[c].n;
        left[plan[c].item] -= plan[c].n;
        if (rc == B200_OK && left[plan[c].item] == 0) {
            // every chunk of the tensor is on the copy stream: its re-layout goes to the compute stream behind that point, and overlaps what follows
            if (it.repack) {
                LD_TRY(cudaEventRecord(tensor_in, copy));
                if (rc == B200_OK) LD_TRY(cudaStreamWaitEvent(ctx->stream, tensor_in, 0));
                if (rc == B200_OK) rc = b200_repack(ctx, it.dst->type, it.dst->data, it.dst->ne[0], it.dst->ne[1] * it.dst->ne[2] * it.dst->ne[3], it.dst->nb[1]);
                if (rc == B200_OK) it.dst->flags |= B200_TENSOR_REPACKED;
            }
            b200_note_write(ctx, it.dst->data, b200_tensor_span(it.dst));
        }
    }
    if (rc != B200_OK) { std::lock_guard<std::mutex> lk(R.m); R.failed = true; R.cv.notify_all(); }
    for (auto & t : readers) t.join();
    double read_s = 0.0;
    for (double v : read_t) read_s = std::max(read_s, v);
    if (rc == B200_OK && R.failed) { b200_set_error("load_file: %s: read failed: %s", path, strerror(R.err)); rc = B200_ERR_INVALID; }
    cudaError_t e1 = cudaStreamSynchronize(copy), e2 = cudaStreamSynchronize(ctx->stream);
    if (rc == B200_OK && (e1 != cudaSuccess || e2 != cudaSuccess)) { b200_set_error("load_file: %s", cudaGetErrorString(e1 != cudaSuccess ? e1 : e2)); rc = B200_ERR_CUDA; }
    if (stats) {
        stats->seconds = std::chrono::duration<double>(std::chrono::steady_clock::now() - t0).count();
        stats->read_seconds = read_s; stats->bytes = (uint64_t) issued; stats->chunks = (uint64_t) plan.size();
    }
    (void) total;
    cleanup();
    return rc;
#undef LD_TRY
}
