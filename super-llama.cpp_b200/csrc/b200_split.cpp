// b200_split.cpp -- the split-mode=row buffer type: what the registry answers to "ggml_backend_split_buffer_type"
// (asked by src/llama-model.cpp:393-411; the reference's CUDA answer is ggml/src/ggml-cuda/ggml-cuda.cu:777-1000).
//
// A tensor of this buffer type is a weight matrix whose rows are divided among ALL devices of the registry: device i keeps rows
// [lo_i, hi_i) -- the partition rule of ggml-cuda.cu:777-803, through b200_split_row_range -- in an allocation of its own.  The
// buffer object itself owns no memory (its base is a dummy address, tensor->data is never dereferenced); the per-device slices hang off
// tensor->extra.  Weights arrive whole through set_tensor, are cut by rows, copied to their devices and -- where the kernels have a
// device layout for the type (b200_repack) -- re-laid out per slice, exactly as plain weight buffers do; get_tensor undoes it.
// Only MUL_MAT reads such tensors (b200_backend.cpp: split_mul_mat): every device multiplies its rows with the activation, which it
// reads from the main device over NVLink peer access, and stores its rows of the result straight into the main device's dst.
#include "b200_split.h"
#include <cuda_runtime.h>
#include <mutex>
#include <string>
#include <vector>
#include <stdio.h>
#include <stdlib.h>

#define SPLIT_CHECK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
        fprintf(stderr, "ggml-b200: %s:%d: %s failed: %s\n", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); abort(); } } while (0)

static const int64_t k_row_rounding = 64;                  // slice boundaries fall on multiples of the GEMM's half tile

struct split_buft_ctx {
    int main_device; int n; float start[B200_SPLIT_MAX_DEVICES];      // start[i]: cumulative fraction of the rows at which device i begins
    std::string name;
};
struct split_buffer_ctx { std::vector<b200_split_extra *> extras; };

static void slice_range(const split_buft_ctx * bc, int64_t nrows, int i, int64_t * lo, int64_t * hi) {
    if (b200_split_row_range(nrows, bc->start, bc->n, i, k_row_rounding, lo, hi) != B200_OK) { fprintf(stderr, "ggml-b200: split_row_range: %s\n", b200_last_error()); abort(); }
}

// peers read activations from and write results into each other's memory: enable access between every pair of distinct GPUs once
static void enable_peer_access() {
    static std::once_flag once;
    std::call_once(once, [] {
        const int n = b200_logical_devices();
        for (int a = 0; a < n; ++a) for (int b = 0; b < n; ++b) {
            const int da = b200_logical_to_cuda(a), db = b200_logical_to_cuda(b);
            if (da == db) continue;
            int can = 0;
            if (cudaDeviceCanAccessPeer(&can, da, db) != cudaSuccess || !can) { fprintf(stderr, "ggml-b200: device %d cannot access device %d: split-mode row needs peer access\n", da, db); abort(); }
            SPLIT_CHECK(cudaSetDevice(da));
            const cudaError_t e = cudaDeviceEnablePeerAccess(db, 0);
            if (e != cudaSuccess && e != cudaErrorPeerAccessAlreadyEnabled) SPLIT_CHECK(e);
            cudaGetLastError();
        }
    });
}

// ---- buffer -------------------------------------------------------------------------------------------------------------------------
static void sbuf_free(ggml_backend_buffer_t b) {
    split_buffer_ctx * c = (split_buffer_ctx *) b->context;
    for (b200_split_extra * e : c->extras) {
        for (int i = 0; i < e->n; ++i) if (e->s[i].data) { cudaSetDevice(b200_logical_to_cuda(i)); cudaFree(e->s[i].data); }
        delete e;
    }
    delete c;
}
static void * sbuf_base(ggml_backend_buffer_t) { return (void *) 0x1000; }      // never dereferenced: the slices are in tensor->extra
static enum ggml_status sbuf_init_tensor(ggml_backend_buffer_t b, ggml_tensor * t) {
    if (t->view_src != nullptr) { fprintf(stderr, "ggml-b200: views of row-split tensors are not supported (%s)\n", t->name); abort(); }
    const split_buft_ctx * bc = (const split_buft_ctx *) b->buft->context;
    const size_t row = b200_row_size((int) t->type, t->ne[0]);
    if (t->nb[1] != row || t->ne[2] * t->ne[3] != 1) { fprintf(stderr, "ggml-b200: row split needs a contiguous matrix (%s)\n", t->name); abort(); }
    b200_split_extra * e = new b200_split_extra{};
    e->n = bc->n;
    ((split_buffer_ctx *) b->context)->extras.push_back(e);
    for (int i = 0; i < bc->n; ++i) {
        slice_range(bc, t->ne[1], i, &e->s[i].lo, &e->s[i].hi);
        const int64_t rows = e->s[i].hi - e->s[i].lo;
        if (rows <= 0) continue;
        // one row of 512 elements of padding behind the slice, zeroed: kernels may over-read a row tail (as MATRIX_ROW_PADDING, ggml-cuda.cu:871-885)
        const size_t size = row * rows, pad = b200_row_size((int) t->type, 512);
        SPLIT_CHECK(cudaSetDevice(b200_logical_to_cuda(i)));
        if (cudaMalloc(&e->s[i].data, size + pad) != cudaSuccess) { fprintf(stderr, "ggml-b200: out of memory for %s rows [%lld, %lld) on device %d\n", t->name, (long long) e->s[i].lo, (long long) e->s[i].hi, i); abort(); }
        SPLIT_CHECK(cudaMemset((char *) e->s[i].data + size, 0, pad));
    }
    t->extra = e;
    return GGML_STATUS_SUCCESS;
}
static void slice_layout(ggml_backend_buffer_t b, const ggml_tensor * t, b200_split_slice & s, int device, bool to_device_layout) {
    if (to_device_layout == (s.repacked != 0) || s.hi <= s.lo) return;
    if (to_device_layout && (!b200_repack_policy(b) || !b200_repack_supported((int) t->type, t->ne[0]))) return;
    b200_ctx * u = b200_util_ctx(b200_logical_to_cuda(device));
    const size_t row = b200_row_size((int) t->type, t->ne[0]);
    const int rc = to_device_layout ? b200_repack(u, (int) t->type, s.data, t->ne[0], s.hi - s.lo, row) : b200_unrepack(u, (int) t->type, s.data, t->ne[0], s.hi - s.lo, row);
    if (rc != B200_OK || b200_ctx_synchronize(u) != B200_OK) { fprintf(stderr, "ggml-b200: repack of a row slice failed: %s\n", b200_last_error()); abort(); }
    s.repacked = to_device_layout ? 1 : 0;
}
static void sbuf_set_tensor(ggml_backend_buffer_t b, ggml_tensor * t, const void * data, size_t off, size_t n) {
    b200_split_extra * e = (b200_split_extra *) t->extra;
    const size_t row = b200_row_size((int) t->type, t->ne[0]);
    if (off != 0 || n != row * (size_t) t->ne[1]) { fprintf(stderr, "ggml-b200: row-split tensors are written whole (%s)\n", t->name); abort(); }   // as ggml-cuda.cu:902-903
    for (int i = 0; i < e->n; ++i) {
        b200_split_slice & s = e->s[i];
        if (s.hi <= s.lo) continue;
        SPLIT_CHECK(cudaSetDevice(b200_logical_to_cuda(i)));
        s.repacked = 0;
        SPLIT_CHECK(cudaMemcpyAsync(s.data, (const char *) data + row * s.lo, row * (s.hi - s.lo), cudaMemcpyHostToDevice, cudaStreamPerThread));
    }
    for (int i = 0; i < e->n; ++i) {
        if (e->s[i].hi <= e->s[i].lo) continue;
        SPLIT_CHECK(cudaSetDevice(b200_logical_to_cuda(i)));
        SPLIT_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
        slice_layout(b, t, e->s[i], i, true);
    }
}
static void sbuf_get_tensor(ggml_backend_buffer_t b, const ggml_tensor * t, void * data, size_t off, size_t n) {
    b200_split_extra * e = (b200_split_extra *) t->extra;
    const size_t row = b200_row_size((int) t->type, t->ne[0]);
    if (off != 0 || n != row * (size_t) t->ne[1]) { fprintf(stderr, "ggml-b200: row-split tensors are read whole (%s)\n", t->name); abort(); }
    for (int i = 0; i < e->n; ++i) {
        b200_split_slice & s = e->s[i];
        if (s.hi <= s.lo) continue;
        const bool was = s.repacked != 0;
        slice_layout(b, t, s, i, false);                          // the host sees GGUF bytes
        SPLIT_CHECK(cudaSetDevice(b200_logical_to_cuda(i)));
        SPLIT_CHECK(cudaMemcpyAsync((char *) data + row * s.lo, s.data, row * (s.hi - s.lo), cudaMemcpyDeviceToHost, cudaStreamPerThread));
        SPLIT_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
        if (was) slice_layout(b, t, s, i, true);
    }
}
static void sbuf_clear(ggml_backend_buffer_t, uint8_t) {}
static const ggml_backend_buffer_i k_split_buffer_iface = { sbuf_free, sbuf_base, sbuf_init_tensor, nullptr, sbuf_set_tensor, sbuf_get_tensor, nullptr, sbuf_clear, nullptr };

// ---- buffer type -----------------------------------------------------------------------------------------------------------------
static const char * sbuft_name(ggml_backend_buffer_type_t buft) { return ((split_buft_ctx *) buft->context)->name.c_str(); }
static ggml_backend_buffer_t sbuft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    // memory is allocated per tensor and per device in init_tensor; `size` (the sum of get_alloc_size) is only what the buffer reports
    return new ggml_backend_buffer{ k_split_buffer_iface, buft, new split_buffer_ctx, size, GGML_BACKEND_BUFFER_USAGE_ANY };
}
static size_t sbuft_alignment(ggml_backend_buffer_type_t) { return 128; }
static size_t sbuft_alloc_size(ggml_backend_buffer_type_t buft, const ggml_tensor * t) {
    const split_buft_ctx * bc = (const split_buft_ctx *) buft->context;
    const size_t row = b200_row_size((int) t->type, t->ne[0]), pad = b200_row_size((int) t->type, 512);
    size_t total = 0;
    for (int i = 0; i < bc->n; ++i) { int64_t lo, hi; slice_range(bc, t->ne[1], i, &lo, &hi); if (hi > lo) total += row * (hi - lo) + pad; }
    return total;
}
static bool sbuft_is_host(ggml_backend_buffer_type_t) { return false; }
static const ggml_backend_buffer_type_i k_split_buft_iface = { sbuft_name, sbuft_alloc, sbuft_alignment, nullptr, sbuft_alloc_size, sbuft_is_host };
bool b200_buft_is_split(ggml_backend_buffer_type_t buft) { return buft && buft->iface.get_name == sbuft_name; }

// One buffer type per (main device, split) pair, kept for the life of the process (the caller compares buffer types by address).
// tensor_split: one weight per device, all zero = even (llama's default; the reference then weighs by free memory, ggml-cuda.cu:1003-1030,
// which on a box of identical GPUs is the same thing).
ggml_backend_buffer_type_t b200_split_buffer_type(int main_device, const float * tensor_split) {
    static std::mutex mu; static std::vector<ggml_backend_buffer_type *> made;
    const int n = b200_logical_devices();
    if (n < 1 || main_device < 0 || main_device >= n || n > B200_SPLIT_MAX_DEVICES) return nullptr;
    enable_peer_access();
    split_buft_ctx want; want.main_device = main_device; want.n = n;
    double sum = 0;
    for (int i = 0; i < n; ++i) sum += tensor_split ? (double) tensor_split[i] : 0.0;
    double acc = 0;
    for (int i = 0; i < n; ++i) { want.start[i] = sum > 0 ? (float) (acc / sum) : (float) i / n; acc += tensor_split ? tensor_split[i] : 0.0; }
    std::lock_guard<std::mutex> g(mu);
    for (ggml_backend_buffer_type * b : made) {
        const split_buft_ctx * c = (const split_buft_ctx *) b->context;
        bool same = c->main_device == main_device && c->n == n;
        for (int i = 0; i < n && same; ++i) same = c->start[i] == want.start[i];
        if (same) return b;
    }
    // the name carries the main device (as ggml-cuda.cu:1095 does): llama's loader keys its per-buffer-type contexts by NAME
    // (ggml_backend_buft_comparator), so two split types with one name would collapse into the one that was listed first and every
    // layer's matmuls would run on that device's backend, away from the rest of the layer
    want.name = "B200_" + std::to_string(main_device) + "_Split";
    ggml_backend_buffer_type * b = new ggml_backend_buffer_type{ k_split_buft_iface, b200_logical_dev(main_device), new split_buft_ctx(want) };
    made.push_back(b);
    return b;
}
