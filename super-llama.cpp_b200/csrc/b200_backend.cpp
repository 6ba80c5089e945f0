// b200_backend.cpp -- the ggml backend plugin ("B200"): registry, devices, buffer types, buffers,
// backend streams, supports_op and graph_compute.  It fills the reference's vtables
// (ggml/src/ggml-backend-impl.h:17-210) and translates each ggml_tensor node into one call of the
// flat operator API (include/ggml-b200.h).  No symbol of libggml is needed at link time, so the
// same .so loads into llama-bench / test-backend-ops via GGML_BACKEND_PATH and into Python via ctypes.
//
// Ownership follows ggml-backend.cpp: ggml deletes `ggml_backend_buffer` objects itself after calling
// free_buffer (ggml/src/ggml-backend.cpp, ggml_backend_buffer_free), so they are allocated with `new`
// and the same layout; backends are freed through iface.free.
#include "b200_common.cuh"
#include "b200_glue.cuh"   // non-hot-path ops so whole llama graphs stay on the device
#include "../../include/ggml-b200-abi.h"
#include "b200_split.h"
#include <chrono>
#include <string>
#include <vector>
#include <mutex>
#include <atomic>
#include <unordered_set>
#include <string.h>
#include <stdlib.h>

#define B200_MAX_DEVICES 16
#define B200_LOG(...) do { fprintf(stderr, "ggml-b200: " __VA_ARGS__); fputc('\n', stderr); } while (0)
#define B200_CHECK(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
        B200_LOG("%s:%d: %s failed: %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); abort(); } } while (0)

// ---- ggml helpers we cannot import (no link dependency on libggml-base) ------------------------
struct type_info { size_t size; int64_t blck; };
static const type_info k_types[GGML_TYPE_COUNT] = {            // ggml/src/ggml.c type_traits table
    {4,1},{2,1},{18,32},{20,32},{0,0},{0,0},{22,32},{24,32},{34,32},{36,32},
    {84,256},{110,256},{144,256},{176,256},{210,256},{292,256},{66,256},{74,256},{98,256},{50,256},
    {18,32},{110,256},{82,256},{136,256},{1,1},{2,1},{4,1},{8,1},{8,1},{56,256},
    {2,1},{0,0},{0,0},{0,0},{54,256},{66,256},{0,0},{0,0},{0,0},{17,32},
};
static size_t  t_size(int t) { return (t >= 0 && t < GGML_TYPE_COUNT) ? k_types[t].size : 0; }
static int64_t t_blck(int t) { return (t >= 0 && t < GGML_TYPE_COUNT) ? k_types[t].blck : 0; }

static size_t tensor_nbytes(const ggml_tensor * t) {           // ggml_nbytes, ggml/src/ggml.c
    for (int i = 0; i < GGML_MAX_DIMS; ++i) if (t->ne[i] <= 0) return 0;
    const int64_t blck = t_blck(t->type);
    size_t n;
    if (blck == 1) { n = t_size(t->type); for (int i = 0; i < GGML_MAX_DIMS; ++i) n += (t->ne[i] - 1) * t->nb[i]; }
    else           { n = t->ne[0] * t->nb[0] / blck; for (int i = 1; i < GGML_MAX_DIMS; ++i) n += (t->ne[i] - 1) * t->nb[i]; }
    return n;
}
static int64_t tensor_nelements(const ggml_tensor * t) { return t->ne[0] * t->ne[1] * t->ne[2] * t->ne[3]; }
static bool tensor_is_contiguous(const ggml_tensor * t) {     // ggml_is_contiguous
    size_t next = t_size(t->type);
    if (t->ne[0] != t_blck(t->type) && t->nb[0] != next) return false;
    next *= t->ne[0] / t_blck(t->type);
    for (int i = 1; i < GGML_MAX_DIMS; ++i) { if (t->ne[i] != 1 && t->nb[i] != next) return false; next *= t->ne[i]; }
    return true;
}
static bool tensor_rows_contiguous(const ggml_tensor * t) { return t->nb[0] == t_size(t->type); }
static bool same_shape(const ggml_tensor * a, const ggml_tensor * b) {
    return a->ne[0] == b->ne[0] && a->ne[1] == b->ne[1] && a->ne[2] == b->ne[2] && a->ne[3] == b->ne[3];
}

static bool tensor_is_repacked(const ggml_tensor * t);
static b200_tensor to_b200(const ggml_tensor * t) {
    b200_tensor r; r.data = t->data; r.type = (int32_t) t->type; r.flags = tensor_is_repacked(t) ? B200_TENSOR_REPACKED : 0;
    for (int i = 0; i < 4; ++i) { r.ne[i] = t->ne[i]; r.nb[i] = t->nb[i]; }
    return r;
}

// ---- registry state -----------------------------------------------------------------------------
struct b200_device_ctx {
    int device;                             // CUDA device
    int index;                              // position in the registry (with GGML_B200_FAKE_DEVICES several entries share one CUDA device)
    std::string name, description;
    ggml_backend_buffer_type buft;          // default device buffer type
    ggml_backend_device      dev;
};
struct b200_reg_ctx { std::vector<b200_device_ctx *> devices; };
// `repacked`: weight tensors of this buffer that set_tensor re-laid out for the kernels (b200_repack): the load-time repack hook
struct b200_buffer_ctx { int device; void * base; std::mutex mu; std::unordered_set<const void *> repacked; std::atomic<int> n_repacked{0}; };   // n_repacked: size of the set, readable without the lock
// one instantiated CUDA graph per ggml graph this backend has computed before (be_graph_compute)
struct b200_graph_entry { uint64_t h1, h2; int n_nodes; void * exec; int seen; int failed; uint64_t used; };
struct b200_backend_ctx {
    int device; b200_ctx * ctx; std::string name;
    int index = 0;                                     // registry position of this backend's device
    // split-mode row (split_mul_mat): one helper context (= stream) per other device, created on first use, and the events that order
    // them against this backend's stream
    b200_ctx * peer[B200_SPLIT_MAX_DEVICES] = {}; cudaEvent_t ev_main = nullptr, ev_peer[B200_SPLIT_MAX_DEVICES] = {}; uint64_t n_split = 0;
    std::vector<b200_graph_entry> graphs; uint64_t graphs_epoch = 0, tick = 0, n_replayed = 0, n_captured = 0, n_eager = 0;
    uint64_t host_ns_compute = 0, host_ns_identity = 0, host_calls = 0;   // GGML_B200_GRAPH_STATS: host time inside graph_compute / its identity hash
    // GGML_B200_GRAPH_STATS: device time of every replay (event pairs, read in be_free) and host time spent inside graph_compute
    std::vector<std::pair<cudaEvent_t, cudaEvent_t>> replay_events; double host_us = 0;
    uint64_t n_fuse[5] = { 0, 0, 0, 0, 0 };          // launches saved by: matmul groups, [1] = decode chains launched, ADD+RMS_NORM+MUL, ROPE/SET_ROWS, gate+up+GLU (counted while walking nodes)
};
// bumped whenever something a captured graph baked in may have changed under it: a weight tensor changed layout, a buffer was freed
static std::atomic<uint64_t> g_layout_epoch{1};

static ggml_backend_reg g_reg;
static b200_reg_ctx     g_reg_ctx;
static ggml_backend_buffer_type g_host_buft;

static bool buft_is_b200(ggml_backend_buffer_type_t buft);
static bool buft_is_b200_host(ggml_backend_buffer_type_t buft);

// ---- load-time weight repack -------------------------------------------------------------------------------
// Weights arrive through set_tensor in GGUF layout.  In a buffer marked USAGE_WEIGHTS (llama_model_loader does that before it
// uploads; GGML_B200_REPACK=all extends it to every buffer, GGML_B200_REPACK=0 turns it off) a whole-tensor write of a type with
// a device layout (b200_repack_supported) is permuted in place right after the copy; get_tensor and partial writes undo /
// redo it, so the host only ever sees GGUF bytes.  Only MUL_MAT / MUL_MAT_ID read repacked tensors (dev_supports_op).
static int repack_mode() {
    static int m = -1;
    if (m < 0) { const char * e = getenv("GGML_B200_REPACK"); m = !e ? 1 : !strcmp(e, "all") ? 2 : atoi(e) ? 1 : 0; }
    return m;
}
b200_ctx * b200_util_ctx(int device) {                      // one helper context per CUDA device for the repack kernels
    static std::mutex mu; static b200_ctx * ctxs[B200_MAX_DEVICES] = {};
    std::lock_guard<std::mutex> g(mu);
    if (!ctxs[device]) ctxs[device] = b200_ctx_create(device);
    return ctxs[device];
}
static b200_ctx * util_ctx(int device) { return b200_util_ctx(device); }
int b200_repack_policy(const ggml_backend_buffer * b) {
    const int mode = repack_mode();
    return mode == 2 || (mode == 1 && b->usage == GGML_BACKEND_BUFFER_USAGE_WEIGHTS);
}
int                b200_logical_devices()        { return (int) g_reg_ctx.devices.size(); }
int                b200_logical_to_cuda(int i)   { return g_reg_ctx.devices[i]->device; }
ggml_backend_dev_t b200_logical_dev(int i)       { return &g_reg_ctx.devices[i]->dev; }
static bool tensor_is_repacked(const ggml_tensor * t) {
    const ggml_tensor * base = t->view_src ? t->view_src : t;
    if (!base->buffer || !buft_is_b200(base->buffer->buft) || t_blck(base->type) <= 1) return false;
    b200_buffer_ctx * c = (b200_buffer_ctx *) base->buffer->context;
    std::lock_guard<std::mutex> g(c->mu);
    return c->repacked.count(base->data) != 0;
}
static void tensor_repack(ggml_backend_buffer_t b, const ggml_tensor * t, bool inverse) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) b->context;
    b200_ctx * u = util_ctx(c->device);
    if (!u) return;
    const int rc = inverse ? b200_unrepack(u, (int) t->type, t->data, t->ne[0], t->ne[1] * t->ne[2] * t->ne[3], t->nb[1])
                           : b200_repack(u, (int) t->type, t->data, t->ne[0], t->ne[1] * t->ne[2] * t->ne[3], t->nb[1]);
    if (rc != B200_OK || b200_ctx_synchronize(u) != B200_OK) { fprintf(stderr, "ggml-b200: repack failed: %s\n", b200_last_error()); abort(); }
    std::lock_guard<std::mutex> g(c->mu);
    if (inverse) c->repacked.erase(t->data); else c->repacked.insert(t->data);
    c->n_repacked.store((int) c->repacked.size());
    ++g_layout_epoch;
}
static bool tensor_wants_repack(ggml_backend_buffer_t b, const ggml_tensor * t) {
    const int mode = repack_mode();
    if (mode == 0 || (mode == 1 && b->usage != GGML_BACKEND_BUFFER_USAGE_WEIGHTS)) return false;
    if (t->view_src || t_blck(t->type) <= 1 || !b200_repack_supported((int) t->type, t->ne[0])) return false;
    // rows densely packed one after the other (what the in-place group permutation assumes)
    return t->nb[0] == t_size(t->type) && t->nb[1] == (size_t) (t->ne[0] / t_blck(t->type)) * t_size(t->type) &&
           t->nb[2] == t->nb[1] * (size_t) t->ne[1] && t->nb[3] == t->nb[2] * (size_t) t->ne[2] && ((uintptr_t) t->data & 15) == 0;
}

// ---- buffer ---------------------------------------------------------------------------------------
static void buf_free(ggml_backend_buffer_t b) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) b->context;
    cudaSetDevice(c->device); cudaFree(c->base); delete c;
    ++g_layout_epoch;
}
static void * buf_get_base(ggml_backend_buffer_t b) { return ((b200_buffer_ctx *) b->context)->base; }
static enum ggml_status buf_init_tensor(ggml_backend_buffer_t b, ggml_tensor * t) {
    // a tensor that starts its life at an address forgets what an earlier tenant of that address was (compute buffers under
    // GGML_B200_REPACK=all hand one block to tensors of different shapes: the layout mark must not outlive the tensor it was made for)
    if (t->view_src == nullptr && ((b200_buffer_ctx *) b->context)->n_repacked.load(std::memory_order_relaxed) != 0) {   // (the graph allocator calls this for every node of every graph)
        b200_buffer_ctx * c = (b200_buffer_ctx *) b->context;
        std::lock_guard<std::mutex> g(c->mu);
        if (c->repacked.erase(t->data)) { c->n_repacked.store((int) c->repacked.size()); ++g_layout_epoch; }
    }
    // quantized rows may be padded by get_alloc_size: zero the tail so padded reads are defined (as ggml-cuda.cu:640-656)
    if (t->view_src == nullptr && t_blck(t->type) > 1) {
        const size_t orig = tensor_nbytes(t), padded = b->buft->iface.get_alloc_size(b->buft, t);
        if (padded > orig) { b200_buffer_ctx * c = (b200_buffer_ctx *) b->context; cudaSetDevice(c->device); B200_CHECK(cudaMemset((char *) t->data + orig, 0, padded - orig)); }
    }
    return GGML_STATUS_SUCCESS;
}
static void buf_memset_tensor(ggml_backend_buffer_t b, ggml_tensor * t, uint8_t v, size_t off, size_t n) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) b->context; B200_CHECK(cudaSetDevice(c->device));
    const ggml_tensor * base = t->view_src ? t->view_src : t;
    const bool was = tensor_is_repacked(t);                  // (offset, length) name GGUF bytes: a re-laid-out weight goes back first, as in set_tensor
    if (was && !(off == 0 && n == tensor_nbytes(t))) tensor_repack(b, base, true);
    B200_CHECK(cudaMemsetAsync((char *) t->data + off, v, n, cudaStreamPerThread)); B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    if (was && !(off == 0 && n == tensor_nbytes(t))) tensor_repack(b, base, false);
}
static void buf_set_tensor(ggml_backend_buffer_t b, ggml_tensor * t, const void * data, size_t off, size_t n) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) b->context; B200_CHECK(cudaSetDevice(c->device));
    const ggml_tensor * base = t->view_src ? t->view_src : t;
    const bool was = tensor_is_repacked(t);
    if (was) tensor_repack(b, base, true);                   // the host writes GGUF bytes: go back to GGUF layout first
    B200_CHECK(cudaMemcpyAsync((char *) t->data + off, data, n, cudaMemcpyHostToDevice, cudaStreamPerThread)); B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    // llama_model_loader without mmap uploads a tensor in ascending chunks (src/llama-model-loader.cpp:1136-1186): the write that
    // reaches the last byte completes the tensor
    if (was || (off + n == tensor_nbytes(t) && tensor_wants_repack(b, t))) tensor_repack(b, base, false);
}
static void buf_get_tensor(ggml_backend_buffer_t b, const ggml_tensor * t, void * data, size_t off, size_t n) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) b->context; B200_CHECK(cudaSetDevice(c->device));
    const ggml_tensor * base = t->view_src ? t->view_src : t;
    const bool was = tensor_is_repacked(t);
    if (was) tensor_repack(b, base, true);
    B200_CHECK(cudaMemcpyAsync(data, (const char *) t->data + off, n, cudaMemcpyDeviceToHost, cudaStreamPerThread)); B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    if (was) tensor_repack(b, base, false);
}
static bool buf_cpy_tensor(ggml_backend_buffer_t b, const ggml_tensor * src, ggml_tensor * dst) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) b->context;
    if (!src->buffer || !buft_is_b200(src->buffer->buft)) return false;
    if (tensor_is_repacked(src) || tensor_is_repacked(dst)) return false;     // go through get_tensor / set_tensor, which translate the layout
    b200_buffer_ctx * sc = (b200_buffer_ctx *) src->buffer->context;
    B200_CHECK(cudaSetDevice(c->device));
    if (sc->device == c->device) B200_CHECK(cudaMemcpyAsync(dst->data, src->data, tensor_nbytes(src), cudaMemcpyDeviceToDevice, cudaStreamPerThread));
    else                         B200_CHECK(cudaMemcpyPeerAsync(dst->data, c->device, src->data, sc->device, tensor_nbytes(src), cudaStreamPerThread));
    B200_CHECK(cudaStreamSynchronize(cudaStreamPerThread));
    return true;
}
static void buf_clear(ggml_backend_buffer_t b, uint8_t v) {
    b200_buffer_ctx * c = (b200_buffer_ctx *) b->context; B200_CHECK(cudaSetDevice(c->device));
    B200_CHECK(cudaMemset(c->base, v, b->size));
}
static const ggml_backend_buffer_i k_buffer_iface = {
    buf_free, buf_get_base, buf_init_tensor, buf_memset_tensor, buf_set_tensor, buf_get_tensor, buf_cpy_tensor, buf_clear, nullptr,
};

// ---- device buffer type ---------------------------------------------------------------------------
static const char * buft_name(ggml_backend_buffer_type_t buft) { return ((b200_device_ctx *) buft->context)->name.c_str(); }
static ggml_backend_buffer_t buft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    b200_device_ctx * d = (b200_device_ctx *) buft->context;
    if (cudaSetDevice(d->device) != cudaSuccess) return nullptr;
    void * p = nullptr;
    if (cudaMalloc(&p, size ? size : 1) != cudaSuccess) { cudaGetLastError(); B200_LOG("alloc of %.2f MiB on device %d failed", size / 1048576.0, d->device); return nullptr; }
    return new ggml_backend_buffer{ k_buffer_iface, buft, new b200_buffer_ctx{ d->device, p }, size, GGML_BACKEND_BUFFER_USAGE_ANY };
}
static size_t buft_alignment(ggml_backend_buffer_type_t) { return 256; }
static size_t buft_alloc_size(ggml_backend_buffer_type_t, const ggml_tensor * t) {
    // pad quantized rows to a multiple of 512 elements so kernels may over-read a row tail (as MATRIX_ROW_PADDING, ggml-cuda common.cuh:148)
    size_t n = tensor_nbytes(t);
    if (t_blck(t->type) > 1 && t->view_src == nullptr && t->ne[0] % 512 != 0) n += (512 - t->ne[0] % 512) / t_blck(t->type) * t_size(t->type);
    return n;
}
static bool buft_is_host(ggml_backend_buffer_type_t) { return false; }
static const ggml_backend_buffer_type_i k_buft_iface = { buft_name, buft_alloc, buft_alignment, nullptr, buft_alloc_size, buft_is_host };
static bool buft_is_b200(ggml_backend_buffer_type_t buft) { return buft->iface.get_name == buft_name; }

// ---- pinned host buffer type ------------------------------------------------------------------------
static void hbuf_free(ggml_backend_buffer_t b) { cudaFreeHost(b->context); }
static void * hbuf_base(ggml_backend_buffer_t b) { return b->context; }
static void hbuf_memset_tensor(ggml_backend_buffer_t, ggml_tensor * t, uint8_t v, size_t off, size_t n) { memset((char *) t->data + off, v, n); }
static void hbuf_set_tensor(ggml_backend_buffer_t, ggml_tensor * t, const void * d, size_t off, size_t n) { memcpy((char *) t->data + off, d, n); }
static void hbuf_get_tensor(ggml_backend_buffer_t, const ggml_tensor * t, void * d, size_t off, size_t n) { memcpy(d, (const char *) t->data + off, n); }
static bool hbuf_cpy_tensor(ggml_backend_buffer_t, const ggml_tensor * src, ggml_tensor * dst) {
    if (src->buffer && src->buffer->buft->iface.is_host && src->buffer->buft->iface.is_host(src->buffer->buft)) { memcpy(dst->data, src->data, tensor_nbytes(src)); return true; }
    return false;
}
static void hbuf_clear(ggml_backend_buffer_t b, uint8_t v) { memset(b->context, v, b->size); }
static const ggml_backend_buffer_i k_host_buffer_iface = { hbuf_free, hbuf_base, nullptr, hbuf_memset_tensor, hbuf_set_tensor, hbuf_get_tensor, hbuf_cpy_tensor, hbuf_clear, nullptr };
static const char * hbuft_name(ggml_backend_buffer_type_t) { return "B200_Host"; }
static ggml_backend_buffer_t hbuft_alloc(ggml_backend_buffer_type_t buft, size_t size) {
    void * p = nullptr;
    if (cudaMallocHost(&p, size ? size : 1) != cudaSuccess) { cudaGetLastError(); return nullptr; }
    return new ggml_backend_buffer{ k_host_buffer_iface, buft, p, size, GGML_BACKEND_BUFFER_USAGE_ANY };
}
static size_t hbuft_alignment(ggml_backend_buffer_type_t) { return 64; }
static bool hbuft_is_host(ggml_backend_buffer_type_t) { return true; }
static const ggml_backend_buffer_type_i k_hbuft_iface = { hbuft_name, hbuft_alloc, hbuft_alignment, nullptr, nullptr, hbuft_is_host };
static bool buft_is_b200_host(ggml_backend_buffer_type_t buft) { return buft->iface.get_name == hbuft_name; }

// ---- backend (stream) ---------------------------------------------------------------------------------
static ggml_guid g_guid = { 0xb2, 0x00, 0x5a, 0x10, 0x0a, 0x67, 0x67, 0x6d, 0x6c, 0x2d, 0x62, 0x32, 0x30, 0x30, 0x00, 0x01 };

static uint64_t g_optimize_ns = 0, g_optimize_calls = 0;
static const char * be_name(ggml_backend_t be) { return ((b200_backend_ctx *) be->context)->name.c_str(); }
static void be_free(ggml_backend_t be) {
    b200_backend_ctx * c = (b200_backend_ctx *) be->context;
    for (auto & e : c->graphs) if (e.exec) b200_graph_destroy(e.exec);
    if (!c->replay_events.empty()) {
        cudaStreamSynchronize(c->ctx->stream);
        double ms = 0, mn = 1e30; for (auto & p : c->replay_events) { float t = 0; cudaEventElapsedTime(&t, p.first, p.second); ms += t; if (t < mn) mn = t; cudaEventDestroy(p.first); cudaEventDestroy(p.second); }
        fprintf(stderr, "ggml-b200 graphs: %zu replays timed: %.1f us each on the device (fastest %.1f us)\n", c->replay_events.size(), ms * 1e3 / c->replay_events.size(), mn * 1e3);
    }
    if (getenv("GGML_B200_GRAPH_STATS"))
        fprintf(stderr, "ggml-b200 fusions (launches saved while walking nodes): matmul groups %llu, add+norm+mul %llu, rope/set_rows %llu, gate+up+glu %llu, decode chains %llu\n",
                (unsigned long long) c->n_fuse[0], (unsigned long long) c->n_fuse[2], (unsigned long long) c->n_fuse[3], (unsigned long long) c->n_fuse[4], (unsigned long long) c->n_fuse[1]);
    if (getenv("GGML_B200_GRAPH_STATS") && c->host_calls)
        fprintf(stderr, "ggml-b200 host: %llu graph_compute calls, %.1f us each on the host (%.1f us of it the identity hash); graph_optimize %.1f us per call\n", (unsigned long long) c->host_calls,
                c->host_ns_compute * 1e-3 / c->host_calls, c->host_ns_identity * 1e-3 / c->host_calls, g_optimize_calls ? g_optimize_ns * 1e-3 / g_optimize_calls : 0.0);
    if (getenv("GGML_B200_GRAPH_STATS") && c->n_split)
        fprintf(stderr, "ggml-b200 split: %s ran %llu row-split MUL_MATs over %d devices\n", c->name.c_str(), (unsigned long long) c->n_split, b200_logical_devices());
    if (getenv("GGML_B200_PROFILE") || getenv("GGML_B200_GRAPH_STATS"))
        fprintf(stderr, "ggml-b200 graphs: %llu computed launch by launch, %llu recorded, %llu replayed\n", (unsigned long long) c->n_eager, (unsigned long long) c->n_captured, (unsigned long long) c->n_replayed);
    if (getenv("GGML_B200_PROFILE")) {                           // per-kernel-kind device time of everything this backend launched
        static const char * kinds[] = { "quantize_q8_1", "gemv", "gemv_id", "gemm", "fattn", "fattn_combine", "matmul_float", "glue", "mmid_sort", "convert" };
        for (int k = 0; k < 10; ++k) {
            double ms = 0; uint64_t n = 0;
            if (b200_ctx_profile_get(c->ctx, k, &ms, &n) == B200_OK && n) fprintf(stderr, "ggml-b200 profile: %-14s %8llu launches %10.3f ms  %8.2f us each\n", kinds[k], (unsigned long long) n, ms, ms * 1e3 / n);
        }
    }
    for (int i = 0; i < B200_SPLIT_MAX_DEVICES; ++i) {
        if (c->peer[i]) { b200_ctx_synchronize(c->peer[i]); b200_ctx_destroy(c->peer[i]); }
        if (c->ev_peer[i]) { cudaSetDevice(b200_logical_to_cuda(i)); cudaEventDestroy(c->ev_peer[i]); }
    }
    if (c->ev_main) { cudaSetDevice(c->device); cudaEventDestroy(c->ev_main); }
    b200_ctx_destroy(c->ctx); delete c; delete be;
}
static void be_set_tensor_async(ggml_backend_t be, ggml_tensor * t, const void * data, size_t off, size_t n) {
    b200_backend_ctx * c = (b200_backend_ctx *) be->context; B200_CHECK(cudaSetDevice(c->device));
    b200_note_write(c->ctx, (char *) t->data + off, n);
    if (t->buffer && buft_is_b200(t->buffer->buft) && (tensor_is_repacked(t) || (off + n == tensor_nbytes(t) && tensor_wants_repack(t->buffer, t)))) {
        B200_CHECK(cudaStreamSynchronize(c->ctx->stream));       // layout translation is synchronous: weights are uploaded once
        buf_set_tensor(t->buffer, t, data, off, n);
        return;
    }
    B200_CHECK(cudaMemcpyAsync((char *) t->data + off, data, n, cudaMemcpyHostToDevice, c->ctx->stream));
}
static void be_get_tensor_async(ggml_backend_t be, const ggml_tensor * t, void * data, size_t off, size_t n) {
    b200_backend_ctx * c = (b200_backend_ctx *) be->context; B200_CHECK(cudaSetDevice(c->device));
    if (t->buffer && buft_is_b200(t->buffer->buft) && tensor_is_repacked(t)) {
        B200_CHECK(cudaStreamSynchronize(c->ctx->stream));
        buf_get_tensor(t->buffer, t, data, off, n);
        return;
    }
    B200_CHECK(cudaMemcpyAsync(data, (const char *) t->data + off, n, cudaMemcpyDeviceToHost, c->ctx->stream));
}
static bool be_is_b200(ggml_backend_t be) { return be && be->iface.get_name == be_name; }
static bool be_cpy_tensor_async(ggml_backend_t be_src, ggml_backend_t be_dst, const ggml_tensor * src, ggml_tensor * dst) {
    if (!be_is_b200(be_src) || !be_is_b200(be_dst)) return false;
    if (tensor_is_repacked(src) || tensor_is_repacked(dst)) return false;
    if (!src->buffer || !dst->buffer || !buft_is_b200(src->buffer->buft) || !buft_is_b200(dst->buffer->buft)) return false;
    b200_backend_ctx * cs = (b200_backend_ctx *) be_src->context, * cd = (b200_backend_ctx *) be_dst->context;
    b200_buffer_ctx * bs = (b200_buffer_ctx *) src->buffer->context, * bd = (b200_buffer_ctx *) dst->buffer->context;
    if (cs->device != bs->device || cd->device != bd->device) return false;
    b200_note_write(cd->ctx, dst->data, tensor_nbytes(dst));
    // copy on the source stream, then make the destination stream wait for it (as ggml-cuda.cu:2700-2760)
    B200_CHECK(cudaSetDevice(cs->device));
    if (cs->device == cd->device) B200_CHECK(cudaMemcpyAsync(dst->data, src->data, tensor_nbytes(dst), cudaMemcpyDeviceToDevice, cs->ctx->stream));
    else                          B200_CHECK(cudaMemcpyPeerAsync(dst->data, cd->device, src->data, cs->device, tensor_nbytes(dst), cs->ctx->stream));
    if (be_src != be_dst) {
        cudaEvent_t ev; B200_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
        B200_CHECK(cudaEventRecord(ev, cs->ctx->stream));
        B200_CHECK(cudaSetDevice(cd->device));
        B200_CHECK(cudaStreamWaitEvent(cd->ctx->stream, ev, 0));
        B200_CHECK(cudaEventDestroy(ev));
    }
    return true;
}
static void be_synchronize(ggml_backend_t be) {
    b200_backend_ctx * c = (b200_backend_ctx *) be->context; B200_CHECK(cudaSetDevice(c->device));
    B200_CHECK(cudaStreamSynchronize(c->ctx->stream));
}

static bool op_is_view(int op) { return op == GGML_OP_NONE || op == GGML_OP_RESHAPE || op == GGML_OP_VIEW || op == GGML_OP_PERMUTE || op == GGML_OP_TRANSPOSE; }

static int compute_node(b200_ctx * ctx, ggml_tensor * node) {
    switch ((int) node->op) {
        case GGML_OP_MUL_MAT: {
            b200_tensor w = to_b200(node->src[0]), x = to_b200(node->src[1]), d = to_b200(node);
            return b200_mul_mat(ctx, &w, &x, &d);
        }
        case GGML_OP_MUL_MAT_ID: {
            b200_tensor as = to_b200(node->src[0]), b = to_b200(node->src[1]), ids = to_b200(node->src[2]), d = to_b200(node);
            return b200_mul_mat_id(ctx, &as, &b, &ids, &d);
        }
        case GGML_OP_FLASH_ATTN_EXT: {
            b200_tensor q = to_b200(node->src[0]), k = to_b200(node->src[1]), v = to_b200(node->src[2]), d = to_b200(node), m, s;
            if (node->src[3]) m = to_b200(node->src[3]);
            if (node->src[4]) s = to_b200(node->src[4]);
            float scale, max_bias, softcap;
            memcpy(&scale, node->op_params + 0, 4); memcpy(&max_bias, node->op_params + 1, 4); memcpy(&softcap, node->op_params + 2, 4);
            return b200_flash_attn_ext(ctx, &q, &k, &v, node->src[3] ? &m : nullptr, node->src[4] ? &s : nullptr, &d, scale, max_bias, softcap);
        }
        default: {
            const int rc = b200_glue_compute(ctx, node);
            b200_note_write(ctx, node->data, tensor_nbytes(node));
            return rc;
        }
    }
}

static bool fuse_off() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_NO_FUSE"); v = (e && atoi(e)) ? 1 : 0; } return v != 0; }
static bool ranges_overlap(const ggml_tensor * a, const ggml_tensor * b) {
    if (!a || !b || !a->data || !b->data) return false;
    const char * pa = (const char *) a->data, * pb = (const char *) b->data;
    return pa < pb + tensor_nbytes(b) && pb < pa + tensor_nbytes(a);
}
static int next_work(const ggml_cgraph * g, int i) {            // next node after i that launches something, or -1
    for (int j = i + 1; j < g->n_nodes; ++j) if (!op_is_view((int) g->nodes[j]->op) && tensor_nelements(g->nodes[j]) != 0) return j;
    return -1;
}
// node i is a ROPE: if it opens the KV-cache update of a llama layer -- [ROPE(Q)] ROPE(K) SET_ROWS(K) SET_ROWS(V), only views in between --
// run the whole run as one launch and return the index of its last node; -1 when it is not that pattern (b200_glue_rope_kv checks the rest)
static bool ranges_overlap(const ggml_tensor * a, const ggml_tensor * b);
static int try_rope_kv(b200_backend_ctx * c, ggml_cgraph * g, int i) {
    int j1 = next_work(g, i), j2 = j1 < 0 ? -1 : next_work(g, j1), j3 = -1;
    if (j2 < 0) return -1;
    ggml_tensor * rq = nullptr, * rk, * sk, * sv;
    if (g->nodes[j1]->op == GGML_OP_ROPE) {
        j3 = next_work(g, j2);
        if (j3 < 0) return -1;
        sk = g->nodes[j2]; sv = g->nodes[j3];
        // either order: the rope that SET_ROWS(K) stores is the K one (graph_optimize puts it first, see there)
        const bool first_is_k = sk->op == GGML_OP_SET_ROWS && sk->src[0] && sk->src[0]->data == g->nodes[i]->data;
        rk = first_is_k ? g->nodes[i] : g->nodes[j1]; rq = first_is_k ? g->nodes[j1] : g->nodes[i];
    } else { rk = g->nodes[i]; sk = g->nodes[j1]; sv = g->nodes[j2]; j3 = j2; }
    if (sk->op != GGML_OP_SET_ROWS || sv->op != GGML_OP_SET_ROWS) return -1;
    if (b200_glue_rope_kv(c->ctx, rq, rk, sk, sv) != B200_OK) {
        static int said = 0;
        if (!said++ && getenv("GGML_B200_DEBUG")) B200_LOG("rope/KV-store fusion declined at node %d (%s): %s", i, g->nodes[i]->name, b200_last_error());
        // ROPE(K) ROPE(Q) SET_ROWS SET_ROWS where the allocator put Q's result over K's input (free blocks coalesce): K + the two cache
        // writes still share a launch, ROPE(Q) follows.  It moves behind the SET_ROWS nodes, which it must not disturb.
        if (!rq || rk != g->nodes[i]) return -1;
        const ggml_tensor * keep[5] = { rk, sk->src[0], sv->src[0], sk, sv };
        for (const ggml_tensor * t : keep) if (ranges_overlap(rq, t)) return -1;
        if (b200_glue_rope_kv(c->ctx, nullptr, rk, sk, sv) != B200_OK || compute_node(c->ctx, rq) != B200_OK) return -1;
    }
    ggml_tensor * outs[4] = { rq, rk, sk, sv };
    for (ggml_tensor * t : outs) if (t) b200_note_write(c->ctx, t->data, tensor_nbytes(t));
    return j3;
}
// nodes a, b: two adjacent MUL_MATs of one activation.  If the next node is GLU(SWIGLU) of exactly their results (ffn_gate, ffn_up, in
// either order) the tensor-core GEMV finishes it in its epilogue (b200_mul_mat_gate_up_swiglu).  Returns the GLU's index or -1.
static int try_gate_up_glu(b200_backend_ctx * c, ggml_cgraph * g, int a, int b) {
    const int j = next_work(g, b);
    if (j < 0) return -1;
    ggml_tensor * glu = g->nodes[j], * na = g->nodes[a], * nb = g->nodes[b];
    if (glu->op != GGML_OP_GLU || glu->op_params[0] != GGML_GLU_OP_SWIGLU || glu->op_params[1] != 0 || !glu->src[1]) return -1;
    ggml_tensor * gate = glu->src[0] == na ? na : glu->src[0] == nb ? nb : nullptr, * up = glu->src[1] == na ? na : glu->src[1] == nb ? nb : nullptr;
    if (!gate || !up || gate == up || glu->type != GGML_TYPE_F32 || !tensor_rows_contiguous(glu)) return -1;
    // the GLU column is written while other CTAs may still be reading the activation / weights, and beside gate and up
    const ggml_tensor * others[5] = { gate, up, gate->src[0], up->src[0], gate->src[1] };
    for (const ggml_tensor * o : others) if (ranges_overlap(glu, o)) return -1;
    b200_tensor wg = to_b200(gate->src[0]), wu = to_b200(up->src[0]), x = to_b200(gate->src[1]), dg = to_b200(gate), du = to_b200(up), dl = to_b200(glu);
    if (b200_mul_mat_gate_up_swiglu(c->ctx, &wg, &wu, &x, &dg, &du, &dl) != B200_OK) return -1;
    return j;
}

// ---- the persistent decode chain (b200_decode_chain, b200_chain.cu) -----------------------------------------------------------------
// Node i opens a run of a llama layer's decode nodes between two attention operators:
//     MUL_MAT(s) of one column  [GLU(SWIGLU)]  { ADD, RMS_NORM, MUL, MUL_MAT(s) [GLU(SWIGLU)]  |  MUL_MAT of the GLU column } ...
// (src/models/llama.cpp:40-130: wo, +inpSA, ffn_norm, ffn_gate / ffn_up, swiglu, ffn_down, +ffn_inp, attn_norm of the next layer, wq / wk / wv;
// in the last layer the run ends with output_norm and the output head).  A run may also open at the ADD.  Two or more MUL_MAT groups go out
// as ONE kernel.  Returns the index of the last node taken, or -1 (then the nodes run one by one as before).
// The graph allocator computes ADD and MUL in place where it can (ffn_inp over wo's result, the scaled norm over the norm): inside the chain
// every CTA still reads the addend while one of them would overwrite it.  So a result that the next ADD overwrites is kept in a private column
// instead (nothing else reads it -- that is why the allocator could reuse it), and a norm value that the MUL overwrites is not stored.
static bool chain_on() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_NO_CHAIN"); v = (e && atoi(e)) ? 0 : 1; } return v != 0; }
static bool chain_col(const ggml_tensor * t) {
    return t && t->type == GGML_TYPE_F32 && t->ne[1] == 1 && t->ne[2] == 1 && t->ne[3] == 1 && t->nb[0] == 4 && t->data && ((uintptr_t) t->data & 15) == 0;
}
static bool chain_weight(const ggml_tensor * w) {
    return w && (w->type == GGML_TYPE_Q4_K || w->type == GGML_TYPE_Q6_K) && w->ne[2] == 1 && w->ne[3] == 1 && w->ne[0] % 2048 == 0 && w->ne[1] % 4 == 0 && w->data &&
           !b200_split_of(w);
}
struct chain_build {
    b200_chain_rope tail[3];                          // ROPE / KV-cache write behind the last phase's wq / wk / wv (chain_tail)
    b200_chain_phase ph[B200_CHAIN_MAX_PHASES];
    b200_tensor t[B200_CHAIN_MAX_PHASES][16];         // storage behind the pointers: w[3], dst[3], in0, in1, normw, out_add, out_norm, out_mul, glu
    int n = 0;
};
// every reader of t in the graph is one of `ok[0 .. nok)` (views of t count as readers of t), and t is not a graph output
static bool chain_only_read_by(const ggml_cgraph * g, const ggml_tensor * t, ggml_tensor * const * ok, int nok) {
    if (t->flags & 2) return false;                   // GGML_TENSOR_FLAG_OUTPUT (ggml/include/ggml.h)
    for (int k = 0; k < g->n_nodes; ++k) {
        const ggml_tensor * n = g->nodes[k];
        bool reads = n->view_src == t;
        for (int sidx = 0; sidx < GGML_MAX_SRC && !reads; ++sidx) reads = n->src[sidx] == t;
        if (!reads) continue;
        bool allowed = false;
        for (int q = 0; q < nok; ++q) allowed |= ok[q] == n;
        if (!allowed) return false;
    }
    return true;
}
// MUL_MAT group at node j (adjacent matmuls of one column, views in between) [+ GLU]: fills phase `ph`, returns the last node index or -1
static int chain_group(ggml_cgraph * g, int j, const ggml_tensor * x, chain_build & cb, int ph) {
    ggml_tensor * mm[3]; int nm = 0, last = -1;
    for (int k = j; k < g->n_nodes && nm < 3; ++k) {
        ggml_tensor * n = g->nodes[k];
        if (op_is_view((int) n->op) || tensor_nelements(n) == 0) continue;
        if (n->op != GGML_OP_MUL_MAT || n->src[1] != x || !chain_weight(n->src[0]) || !chain_col(n) || n->src[0]->ne[0] != x->ne[0]) break;
        mm[nm++] = n; last = k;
    }
    if (nm == 0) return -1;
    b200_chain_phase & P = cb.ph[ph];
    P.epilogue = B200_CHAIN_Y_PLAIN; P.n_w = nm; P.glu = nullptr;
    if (nm == 2) {
        const int k = next_work(g, last);
        if (k >= 0) {
            ggml_tensor * glu = g->nodes[k];
            if (glu->op == GGML_OP_GLU && glu->op_params[0] == GGML_GLU_OP_SWIGLU && glu->op_params[1] == 0 && glu->src[1] && chain_col(glu) &&
                ((glu->src[0] == mm[0] && glu->src[1] == mm[1]) || (glu->src[0] == mm[1] && glu->src[1] == mm[0]))) {
                if (glu->src[0] == mm[1]) { ggml_tensor * t = mm[0]; mm[0] = mm[1]; mm[1] = t; }       // gate first
                P.epilogue = B200_CHAIN_Y_SWIGLU;
                cb.t[ph][12] = to_b200(glu); P.glu = &cb.t[ph][12];
                last = k;
            }
        }
    }
    for (int i = 0; i < nm; ++i) {
        cb.t[ph][i] = to_b200(mm[i]->src[0]); cb.t[ph][3 + i] = to_b200(mm[i]);
        P.w[i] = &cb.t[ph][i]; P.dst[i] = &cb.t[ph][3 + i];
    }
    return last;
}
static bool chain_tail_on() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_NO_CHAIN_TAIL"); v = (e && atoi(e)) ? 0 : 1; } return v != 0; }
static bool bytes_overlap(const void * a, size_t na, const void * b, size_t nb) {
    return a && b && (const char *) a < (const char *) b + nb && (const char *) b < (const char *) a + na;
}
// The chain's last phase is wq / wk / wv and node `at` opens [ROPE(Q)] ROPE(K) SET_ROWS(K) SET_ROWS(V) on exactly their results (one token):
// the row owners of the chain kernel rotate the Q / K rows and write the K / V cache rows themselves.  Returns the SET_ROWS(V) index, or -1
static int chain_tail(ggml_cgraph * g, int at, chain_build & cb) {
    if (!chain_tail_on() || at < 0 || g->nodes[at]->op != GGML_OP_ROPE) return -1;
    b200_chain_phase & P = cb.ph[cb.n - 1];
    if (P.n_w != 3 || P.epilogue != B200_CHAIN_Y_PLAIN) return -1;
    const int j1 = next_work(g, at), j2 = j1 < 0 ? -1 : next_work(g, j1), j3 = j2 < 0 ? -1 : next_work(g, j2);
    if (j3 < 0 || g->nodes[j1]->op != GGML_OP_ROPE) return -1;
    ggml_tensor * sk = g->nodes[j2], * sv = g->nodes[j3];
    if (sk->op != GGML_OP_SET_ROWS || sv->op != GGML_OP_SET_ROWS || !sk->src[0] || !sv->src[0]) return -1;
    const bool first_is_k = sk->src[0]->data == g->nodes[at]->data;
    ggml_tensor * rk = first_is_k ? g->nodes[at] : g->nodes[j1], * rq = first_is_k ? g->nodes[j1] : g->nodes[at];
    b200_chain_rope plan[3]; bool dead_in[3];
    if (b200_glue_rope_kv_plan(rq, rk, sk, sv, plan, dead_in) != B200_OK) {
        static int said = 0;
        if (!said++ && getenv("GGML_B200_DEBUG")) B200_LOG("decode chain: rope / KV-store tail declined at node %d (%s): %s", at, g->nodes[at]->name, b200_last_error());
        return -1;
    }
    const void * in[3] = { rq->src[0]->data, rk->src[0]->data, sv->src[0]->data };
    const int64_t n_in[3] = { tensor_nelements(rq->src[0]), tensor_nelements(rk->src[0]), tensor_nelements(sv->src[0]) };
    int which[3] = { -1, -1, -1 };
    for (int i = 0; i < 3; ++i) for (int k = 0; k < 3; ++k) if (P.dst[i]->data == in[k] && P.dst[i]->ne[0] == n_in[k]) which[i] = k;
    if (which[0] < 0 || which[1] < 0 || which[2] < 0 || which[0] == which[1] || which[0] == which[2] || which[1] == which[2]) return -1;
    // the rotated rows are stored while other CTAs may still be reading the phase's own inputs
    const ggml_tensor * outs[2] = { rq, rk };
    const b200_tensor * ins[4] = { P.in0, P.in1, P.normw, P.out_add };
    for (const ggml_tensor * o : outs) for (const b200_tensor * t : ins) if (t && bytes_overlap(o->data, tensor_nbytes(o), t->data, (size_t) t->ne[0] * 4)) return -1;
    for (int k = 0; k < 3; ++k) { cb.tail[k] = plan[k]; cb.tail[k].skip_dst = dead_in[k] ? 1 : 0; }
    for (int i = 0; i < 3; ++i) P.tail[i] = &cb.tail[which[i]];
    return j3;
}
static int try_chain(b200_backend_ctx * c, ggml_cgraph * g, int i) {
    if (!chain_on() || fuse_off()) return -1;
    static chain_build cb_store[B200_MAX_DEVICES];
    chain_build & cb = cb_store[c->device % B200_MAX_DEVICES];
    cb.n = 0;
    memset(cb.ph, 0, sizeof(cb.ph));
    ggml_tensor * res[B200_CHAIN_MAX_PHASES] = {};       // the single result column of a phase (what an ADD may take), or its glu
    int last = -1, at = i, why = 0, n_add = 0;
    const void * last_sum = nullptr;                     // the out_add column of the chain's last ADD so far
    while (cb.n < B200_CHAIN_MAX_PHASES) {
        ggml_tensor * n0 = at >= 0 && at < g->n_nodes ? g->nodes[at] : nullptr;
        if (!n0) { why = 1; break; }
        const int ph = cb.n;
        b200_chain_phase & P = cb.ph[ph];
        int end = -1;
        bool b_is_last_sum = false;
        void * saved = ph > 0 ? cb.t[ph - 1][3].data : nullptr;      // (the phase before may get a private result column: undone if this phase is dropped)
        if (n0->op == GGML_OP_MUL_MAT) {
            ggml_tensor * x = n0->src[1];
            if (!chain_col(x)) { why = 2; break; }
            if (ph > 0 && x != res[ph - 1]) { why = 3; break; }        // inside a run a bare matmul continues from the GLU column (or the single result) before it
            P.prologue = B200_CHAIN_X_PLAIN;
            cb.t[ph][6] = to_b200(x); P.in0 = &cb.t[ph][6];
            end = chain_group(g, at, x, cb, ph);
        } else if (n0->op == GGML_OP_ADD) {
            const int j1 = next_work(g, at), j2 = j1 < 0 ? -1 : next_work(g, j1), j3 = j2 < 0 ? -1 : next_work(g, j2);
            if (j3 < 0) { why = 4; break; }
            ggml_tensor * add = n0, * norm = g->nodes[j1], * mul = g->nodes[j2];
            if (norm->op != GGML_OP_RMS_NORM || norm->src[0] != add || mul->op != GGML_OP_MUL) { why = 5; break; }
            ggml_tensor * nw = mul->src[0] == norm ? mul->src[1] : mul->src[1] == norm ? mul->src[0] : nullptr;
            if (!nw || nw->type != GGML_TYPE_F32 || nw->ne[0] != add->ne[0] || nw->nb[0] != 4 || tensor_nelements(nw) != add->ne[0] || ((uintptr_t) nw->data & 15)) { why = 6; break; }
            if (!chain_col(add) || !chain_col(norm) || !chain_col(mul) || !chain_col(add->src[0]) || !chain_col(add->src[1])) { why = 7; break; }
            ggml_tensor * a = add->src[0], * b = add->src[1];
            if (ph > 0) { if (b == res[ph - 1]) { b = a; a = res[ph - 1]; } else if (a != res[ph - 1]) { why = 8; break; } }    // in0 = the result of the phase before
            P.prologue = B200_CHAIN_X_ADD_NORM_MUL;
            b_is_last_sum = last_sum && b->data == last_sum;
            memcpy(&P.eps, norm->op_params, 4);
            cb.t[ph][6] = to_b200(a); cb.t[ph][7] = to_b200(b); cb.t[ph][8] = to_b200(nw);
            cb.t[ph][9] = to_b200(add); cb.t[ph][10] = to_b200(norm); cb.t[ph][11] = to_b200(mul);
            P.in0 = &cb.t[ph][6]; P.in1 = &cb.t[ph][7]; P.normw = &cb.t[ph][8];
            P.out_add = &cb.t[ph][9]; P.out_norm = &cb.t[ph][10]; P.out_mul = &cb.t[ph][11];
            // the norm and its scaled form are read by the MUL / by this phase's matmuls only: then they need no store (the allocator usually
            // gives them the memory of a column that has just died -- the addend, or each other)
            end = chain_group(g, j3, mul, cb, ph);
            if (end < 0) { why = 9; break; }
            {
                ggml_tensor * mm[3] = {}; int nm = 0;
                for (int k = j3; k <= end; ++k) if (g->nodes[k]->op == GGML_OP_MUL_MAT && g->nodes[k]->src[1] == mul) mm[nm++] = g->nodes[k];
                ggml_tensor * only_mul[1] = { mul };
                if (chain_only_read_by(g, norm, only_mul, 1)) P.out_norm = nullptr;
                if (chain_only_read_by(g, mul, mm, nm)) P.out_mul = nullptr;
            }
            // what IS stored must not land on anything the phase still reads; an ADD in place over the result of the phase before is the one
            // case with a remedy: that result goes to a private column instead (see above)
            const ggml_tensor * live_out[2] = { P.out_norm ? norm : nullptr, P.out_mul ? mul : nullptr };
            bool bad = ranges_overlap(add, b) || (P.out_norm && P.out_mul && ranges_overlap(norm, mul));
            for (const ggml_tensor * o : live_out) if (o) bad = bad || ranges_overlap(o, a) || ranges_overlap(o, b) || ranges_overlap(o, add) || ranges_overlap(o, nw);
            if (bad) { end = -1; }
            if (ranges_overlap(add, a)) {
                if (ph == 0 || add->data != a->data) { why = 10; break; }
                float * priv = (float *) b200_scratch(c->ctx, 4, (size_t) B200_CHAIN_MAX_PHASES * 32768 * 4);
                if (!priv) { why = 11; break; }
                priv += (size_t) ph * 32768;
                cb.t[ph - 1][3].data = priv; cb.t[ph][6].data = priv;                  // phase before: dst[0]; this phase: in0
            }
        } else { why = 12; break; }
        if (end < 0) { if (ph > 0) cb.t[ph - 1][3].data = saved; { why = 13; break; } }
        res[ph] = P.epilogue == B200_CHAIN_Y_SWIGLU ? g->nodes[end] : (P.n_w == 1 ? g->nodes[end] : nullptr);
        // a phase's results must not land on what the same phase still reads.  (Not the addend of a later phase when it is the chain's first
        // from outside or the sum of the ADD before: the kernel holds those in registers from the start / from that ADD on, so a result may
        // take the addend's memory -- the allocator often gives wq's result the block ffn_inp just left.)
        bool clash = false;
        for (int k = 0; k <= P.n_w && !clash; ++k) for (int q = 6; q <= 8 && !clash; ++q) {
            if (q == 7 && ph > 0 && (n_add == 0 || b_is_last_sum)) continue;
            const b200_tensor * in = q == 6 ? P.in0 : q == 7 ? P.in1 : P.normw, * o = k < P.n_w ? P.dst[k] : P.glu;
            if (in && in->data && o && o->data && (const char *) in->data < (const char *) o->data + o->ne[0] * 4 && (const char *) o->data < (const char *) in->data + in->ne[0] * 4) clash = true;
        }
        if (clash) { if (ph > 0) cb.t[ph - 1][3].data = saved; { why = 14; break; } }
        if (P.prologue == B200_CHAIN_X_ADD_NORM_MUL) { ++n_add; last_sum = P.out_add->data; }
        ++cb.n; last = end;
        at = next_work(g, end);
        if (!res[ph]) { why = 15; break; }                              // several result columns (wq / wk / wv): nothing in a chain continues from them
    }
    { static int said = 0; if (said < 12 && getenv("GGML_B200_DEBUG")) { ++said; B200_LOG("decode chain at node %d (%s): %d phases, stopped at node %d (%s) for reason %d", i, g->nodes[i]->name, cb.n, at, at >= 0 && at < g->n_nodes ? g->nodes[at]->name : "-", why); } }
    if (cb.n < 2) return -1;
    const int tail_end = why == 15 ? chain_tail(g, at, cb) : -1;
    if (tail_end >= 0 && !b200_decode_chain_supported(c->ctx, cb.ph, cb.n)) { for (int k = 0; k < 3; ++k) cb.ph[cb.n - 1].tail[k] = nullptr; }
    else if (tail_end >= 0) { last = tail_end; c->n_fuse[3] += 3; }
    if (!b200_decode_chain_supported(c->ctx, cb.ph, cb.n)) {
        static int said = 0;
        if (!said++ && getenv("GGML_B200_DEBUG")) B200_LOG("decode chain declined at node %d (%s): %s", i, g->nodes[i]->name, b200_last_error());
        return -1;
    }
    if (b200_decode_chain(c->ctx, cb.ph, cb.n) != B200_OK) return -1;
    for (int k = i; k <= last; ++k) if (!op_is_view((int) g->nodes[k]->op)) b200_note_write(c->ctx, g->nodes[k]->data, tensor_nbytes(g->nodes[k]));
    return last;
}
// MUL_MAT whose weight lives in the row-split buffer type (b200_split.cpp): every device of the registry multiplies ITS rows of the weight
// with the activation -- read in place from this backend's device over NVLink peer access -- and stores its rows of the result straight
// into this backend's dst (peer stores): the all-gather of the reference's split path (ggml-cuda.cu:1451-1769: copy src1 to every device,
// compute, copy the dst slices back) happens inside the kernels' own loads and stores.  Ordering: the other devices' streams wait for an
// event recorded on this stream (the activation is ready), this stream waits for one event per device (its rows have landed).
// Up to three MUL_MATs of the same activation go out behind ONE hand-over (one event each way per device instead of one per node); each
// device takes its slices of all of them in one b200_mul_mat_multi call.  Everything here is stream-ordered (events, no host wait), so
// the whole exchange records into a CUDA graph: a wait on an event of a capturing stream pulls the peer's stream into the capture, and the
// wait on the peer's event joins it again before the capture ends.
static inline void pdl_hold(b200_ctx * ctx) { if (ctx->pdl) { ctx->pdl_hold = ctx->pdl; ctx->pdl = 0; } }   // the next launch waits for ALL it depends on
static int split_mul_mat(b200_backend_ctx * c, ggml_tensor * const * nodes, int n) {
    const ggml_tensor * x = nodes[0]->src[1];
    const b200_split_extra * e[3];
    for (int k = 0; k < n; ++k) {
        const ggml_tensor * w = nodes[k]->src[0];
        e[k] = (const b200_split_extra *) w->extra;
        if (!e[k] || w->ne[2] * w->ne[3] != 1) { b200_set_error("row-split MUL_MAT: weight without slices or with batch dimensions"); return B200_ERR_UNSUPPORTED; }
    }
    if (!c->ev_main) {
        if (c->ctx->capturing) { b200_set_error("row-split MUL_MAT: events must exist before a graph is recorded"); return B200_ERR_UNSUPPORTED; }
        B200_CHECK(cudaEventCreateWithFlags(&c->ev_main, cudaEventDisableTiming));
    }
    B200_CHECK(cudaEventRecord(c->ev_main, c->ctx->stream));
    const b200_tensor bx = to_b200(x);
    int rc = B200_OK;
    for (int i = 0; i < e[0]->n && rc == B200_OK; ++i) {
        b200_tensor bw[3], bd[3]; const b200_tensor * wp[3]; b200_tensor * dp[3]; int cnt = 0;
        for (int k = 0; k < n; ++k) {
            if (i >= e[k]->n) continue;
            const b200_split_slice & s = e[k]->s[i];
            if (s.hi <= s.lo) continue;
            const ggml_tensor * w = nodes[k]->src[0];
            b200_tensor & W = bw[cnt]; b200_tensor & D = bd[cnt];
            W.data = s.data; W.type = (int32_t) w->type; W.flags = s.repacked ? B200_TENSOR_REPACKED : 0;
            W.ne[0] = w->ne[0]; W.ne[1] = s.hi - s.lo; W.ne[2] = W.ne[3] = 1;
            W.nb[0] = w->nb[0]; W.nb[1] = w->nb[1]; W.nb[2] = W.nb[3] = w->nb[1] * (size_t) W.ne[1];
            D = to_b200(nodes[k]);
            D.data = (char *) nodes[k]->data + s.lo * sizeof(float); D.ne[0] = s.hi - s.lo;
            wp[cnt] = &W; dp[cnt] = &D; ++cnt;
        }
        if (!cnt) continue;
        b200_ctx * ctx = c->ctx;
        if (i != c->index) {
            if (!c->peer[i]) {
                if (c->ctx->capturing) { b200_set_error("row-split MUL_MAT: a peer context must exist before a graph is recorded"); return B200_ERR_UNSUPPORTED; }
                c->peer[i] = b200_ctx_create(b200_logical_to_cuda(i));
                if (!c->peer[i]) return B200_ERR_ALLOC;
                B200_CHECK(cudaSetDevice(b200_logical_to_cuda(i)));
                B200_CHECK(cudaEventCreateWithFlags(&c->ev_peer[i], cudaEventDisableTiming));
            }
            ctx = c->peer[i];
            B200_CHECK(cudaSetDevice(ctx->device));
            B200_CHECK(cudaStreamWaitEvent(ctx->stream, c->ev_main, 0));
            b200_ctx_invalidate(ctx);                         // its cached activation images were made from older contents of these addresses
            pdl_hold(ctx);
        }
        rc = cnt == 1 ? b200_mul_mat(ctx, wp[0], &bx, dp[0]) : b200_mul_mat_multi(ctx, wp, cnt, &bx, dp);
        if (i != c->index) {
            B200_CHECK(cudaEventRecord(c->ev_peer[i], ctx->stream));
            B200_CHECK(cudaSetDevice(c->device));
            B200_CHECK(cudaStreamWaitEvent(c->ctx->stream, c->ev_peer[i], 0));
            pdl_hold(c->ctx);
        }
    }
    B200_CHECK(cudaSetDevice(c->device));
    for (int k = 0; k < n; ++k) b200_note_write(c->ctx, nodes[k]->data, tensor_nbytes(nodes[k]));
    c->n_split += n;
    return rc;
}

static enum ggml_status run_nodes(b200_backend_ctx * c, ggml_cgraph * g) {
    c->ctx->qcache.valid = 0; c->ctx->mcache.valid = 0;      // tensors may have been rewritten through buffer set_tensor since the last graph
    for (int i = 0; i < g->n_nodes; ++i) {
        ggml_tensor * node = g->nodes[i];
        if (op_is_view((int) node->op) || tensor_nelements(node) == 0) continue;
        int rc;
        if (node->op == GGML_OP_MUL_MAT && b200_split_of(node->src[0])) {
            // the row-split matmuls that follow and read the same activation (wq / wk / wv, ffn_gate / ffn_up after be_graph_optimize) share
            // the hand-over; same rule as below: a member's dst may not overlap what the members before it read or write
            ggml_tensor * grp[3] = { node, nullptr, nullptr }; int ngr = 1, last = i;
            for (int j = i + 1; j < g->n_nodes && ngr < 3 && !fuse_off(); ++j) {
                ggml_tensor * nj = g->nodes[j];
                if (op_is_view((int) nj->op) || tensor_nelements(nj) == 0) continue;
                if (nj->op != GGML_OP_MUL_MAT || nj->src[1] != node->src[1] || !b200_split_of(nj->src[0])) break;
                bool clash = false;
                for (int k = 0; k < ngr && !clash; ++k) clash = ranges_overlap(nj, grp[k]) || ranges_overlap(nj, grp[k]->src[1]);
                if (clash) break;
                grp[ngr++] = nj; last = j;
            }
            if ((rc = split_mul_mat(c, grp, ngr)) != B200_OK) { B200_LOG("graph_compute: row-split node %d (%s) failed: %s", i, node->name, b200_last_error()); return GGML_STATUS_FAILED; }
            i = last;
            continue;
        }
        // decode: the whole run of a layer's nodes between two attention operators as one persistent kernel
        if ((node->op == GGML_OP_MUL_MAT || node->op == GGML_OP_ADD) && node->ne[1] == 1 && (rc = try_chain(c, g, i)) >= 0) { c->n_fuse[1] += 1; i = rc; continue; }
        // MUL_MAT nodes that follow each other (views in between do no work) and read the same src1 -- ffn_gate / ffn_up, attn_v / attn_k
        // of a llama layer -- go out as ONE launch (b200_mul_mat_multi; it falls back to one launch each when the kernels cannot take them
        // together).  Only adjacent nodes: running a later node early could write into memory the graph allocator still lends to a
        // tensor in between, so a candidate's dst must also not overlap anything the nodes before it in the group read or write.
        int group[3] = { i, -1, -1 }, ng = 1;
        if (node->op == GGML_OP_MUL_MAT && !fuse_off()) {
            for (int j = i + 1; j < g->n_nodes && ng < 3; ++j) {
                ggml_tensor * nj = g->nodes[j];
                if (op_is_view((int) nj->op) || tensor_nelements(nj) == 0) continue;
                if (nj->op != GGML_OP_MUL_MAT || nj->src[1] != node->src[1] || nj->src[0]->type != node->src[0]->type || b200_split_of(nj->src[0])) break;
                bool clash = false;
                for (int k = 0; k < ng && !clash; ++k) {
                    const ggml_tensor * nk = g->nodes[group[k]];
                    clash = ranges_overlap(nj, nk) || ranges_overlap(nj, nk->src[0]) || ranges_overlap(nj, nk->src[1]);
                }
                if (clash) break;
                group[ng++] = j;
            }
        }
        if (ng == 2 && (rc = try_gate_up_glu(c, g, group[0], group[1])) >= 0) {
            i = rc; rc = B200_OK; c->n_fuse[0] += 1; c->n_fuse[4] += 1;       // ffn_gate + ffn_up + SwiGLU: one launch
        } else if (ng > 1) {
            b200_tensor w[3], d[3], x = to_b200(node->src[1]);
            const b200_tensor * wp[3]; b200_tensor * dp[3];
            for (int k = 0; k < ng; ++k) { w[k] = to_b200(g->nodes[group[k]]->src[0]); d[k] = to_b200(g->nodes[group[k]]); wp[k] = &w[k]; dp[k] = &d[k]; }
            rc = b200_mul_mat_multi(c->ctx, wp, ng, &x, dp);
            if (rc == B200_OK) { i = group[ng - 1]; c->n_fuse[0] += ng - 1; }   // (views between the members need no work)
        } else if (node->op == GGML_OP_ADD && !fuse_off() && i + 2 < g->n_nodes && g->nodes[i + 1]->op == GGML_OP_RMS_NORM && g->nodes[i + 2]->op == GGML_OP_MUL &&
                   (rc = b200_glue_add_rms_norm_mul(c->ctx, node, g->nodes[i + 1], g->nodes[i + 2])) == B200_OK) {
            // residual add + the norm behind it + the norm weight: one launch
            for (int k = 0; k < 3; ++k) b200_note_write(c->ctx, g->nodes[i + k]->data, tensor_nbytes(g->nodes[i + k]));
            i += 2; c->n_fuse[2] += 2;
        } else if (node->op == GGML_OP_ROPE && !fuse_off() && (rc = try_rope_kv(c, g, i)) >= 0) {
            i = rc; rc = B200_OK; c->n_fuse[3] += 2;                   // [ROPE(Q)] ROPE(K) SET_ROWS(K) SET_ROWS(V): one launch
        } else if (node->op == GGML_OP_RMS_NORM && !fuse_off() && i + 1 < g->n_nodes && g->nodes[i + 1]->op == GGML_OP_MUL &&
                   (rc = b200_glue_rms_norm_mul(c->ctx, node, g->nodes[i + 1])) == B200_OK) {
            b200_note_write(c->ctx, node->data, tensor_nbytes(node));
            b200_note_write(c->ctx, g->nodes[i + 1]->data, tensor_nbytes(g->nodes[i + 1]));
            ++i;
        } else {
            rc = compute_node(c->ctx, node);
        }
        if (rc != B200_OK) {
            B200_LOG("graph_compute: node %d (%s, op %d) failed: %s", i, node->name, (int) node->op, b200_last_error());
            return GGML_STATUS_FAILED;
        }
    }
    return GGML_STATUS_SUCCESS;
}

// ---- CUDA graphs for repeated ggml graphs ----------------------------------------------------------------------------------------
// A decode step is ~500 launches of 2-15 us kernels: issued one by one the host cannot keep the device fed (llama-bench tg128 on
// Llama-3-8B Q4_K_M: 4.2 ms per token against ~2 ms of kernel time).  So the second time a graph with the same identity arrives -- same
// ops, op_params, shapes, strides and data pointers on every node and source, what ggml-cuda.cu's graph-properties check compares -- the node
// loop is recorded into a CUDA graph (PDL edges included) and from then on one cudaGraphLaunch replays it.  Tensor CONTENTS may differ
// freely between replays (token ids, positions, KV indices and the mask are tensors); everything the host decided while recording
// depends on identity only.  Recorded graphs are dropped when a scratch block moved (b200_ctx_scratch_generation) or a weight
// changed layout / a buffer was freed (g_layout_epoch).  Only small-batch graphs on plain device buffers (no row-split peers) are
// recorded; GGML_B200_GRAPHS=0 turns it off.
static bool split_graphs_on() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_SPLIT_GRAPHS"); v = (e && !atoi(e)) ? 0 : 1; } return v != 0; }
static bool graphs_on() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_GRAPHS"); v = (e && !atoi(e)) ? 0 : 1; } return v != 0; }
// The walk runs once per token in front of the replay, with the GPU idle (llama synchronizes after every token): 60 us for the ~1100
// nodes of an 8B decode graph when every value went through one dependent multiply chain and every source was hashed in full again
// (GGML_B200_GRAPH_STATS, profiles/r2_run13).  So: four independent accumulators (the multiplies overlap), and a tensor that was
// is one of the eight nodes in front -- most sources are -- only contributes WHICH of them it is (its own walk hashed it in full).  Nothing
// here depends on where the tensor structs themselves live: llama rebuilds them for every token.
struct id_hasher {
    uint64_t a[4] = { 0xcbf29ce484222325ull, 0x84222325cbf29ce4ull, 0x9e3779b97f4a7c15ull, 0xc2b2ae3d27d4eb4full };
    unsigned n = 0;
    inline void mix(uint64_t v) { uint64_t & h = a[n++ & 3]; h = (h ^ v) * 0x100000001b3ull; h ^= h >> 29; }
    inline void tensor(const ggml_tensor * t) {
        mix((uint64_t) (uintptr_t) t->data); mix(((uint64_t) t->type << 32) | (uint32_t) t->op);
        for (int i = 0; i < 4; ++i) { mix((uint64_t) t->ne[i]); mix((uint64_t) t->nb[i]); }
    }
    inline void finish(uint64_t & h1, uint64_t & h2) const {
        h1 = a[0] ^ (a[1] << 13 | a[1] >> 51) ^ (a[2] << 29 | a[2] >> 35) ^ (a[3] << 47 | a[3] >> 17);
        h2 = a[0] * 0x9e3779b97f4a7c15ull + a[1] * 0xc2b2ae3d27d4eb4full + a[2] * 0x165667b19e3779f9ull + a[3] * 0x27d4eb2f165667c5ull + n;
    }
};
static bool graph_identity(const ggml_cgraph * g, uint64_t & h1, uint64_t & h2) {
    id_hasher H;
    int work = 0;
    for (int i = 0; i < g->n_nodes; ++i) {
        const ggml_tensor * n = g->nodes[i];
        H.tensor(n);
        if (op_is_view((int) n->op)) continue;
        ++work;
        if (n->buffer && !buft_is_b200(n->buffer->buft)) return false;
        for (int k = 0; k < (int) (GGML_MAX_OP_PARAMS / sizeof(int32_t)); k += 2) H.mix(((uint64_t) (uint32_t) n->op_params[k] << 32) | (uint32_t) n->op_params[k + 1]);
        unsigned present = 0;
        for (int k = 0; k < GGML_MAX_SRC; ++k) {
            const ggml_tensor * s = n->src[k];
            if (!s) continue;
            present |= 1u << k;
            if (s->buffer && !buft_is_b200(s->buffer->buft)) {
                // a row-split weight of a MUL_MAT records too (its slices hang off ->extra; events order the peers' streams); anything else does not
                if (!(k == 0 && n->op == GGML_OP_MUL_MAT && b200_split_of(s) && split_graphs_on())) return false;
                H.mix((uint64_t) (uintptr_t) s->extra);
            }
            int back = 0;
            for (int j = 1; j <= 8 && j <= i && !back; ++j) if (g->nodes[i - j] == s) back = j;
            if (back) H.mix(0xbac0ull | (unsigned) back); else H.tensor(s);
        }
        H.mix(0x5eed0000ull | present);
        // prefill-sized work is not launch-bound and its paths may size scratch by content
        const int64_t tokens = n->op == GGML_OP_MUL_MAT ? n->ne[1] * n->ne[2] * n->ne[3] : n->op == GGML_OP_MUL_MAT_ID ? n->ne[2] * n->ne[3] :
                               n->op == GGML_OP_FLASH_ATTN_EXT ? n->ne[2] * n->ne[3] : 0;         // (attention result: [head size, heads, tokens])
        if (tokens > 64) return false;
    }
    H.finish(h1, h2);
    return work >= 2;
}
static enum ggml_status be_graph_compute_(ggml_backend_t be, ggml_cgraph * g);
static inline uint64_t host_now_ns() { return (uint64_t) std::chrono::duration_cast<std::chrono::nanoseconds>(std::chrono::steady_clock::now().time_since_epoch()).count(); }
static enum ggml_status be_graph_compute(ggml_backend_t be, ggml_cgraph * g) {
    static const bool stats = getenv("GGML_B200_GRAPH_STATS") != nullptr;
    if (!stats) return be_graph_compute_(be, g);
    b200_backend_ctx * c = (b200_backend_ctx *) be->context;
    const uint64_t t0 = host_now_ns();
    const enum ggml_status st = be_graph_compute_(be, g);
    c->host_ns_compute += host_now_ns() - t0; ++c->host_calls;
    return st;
}
static enum ggml_status be_graph_compute_(ggml_backend_t be, ggml_cgraph * g) {
    b200_backend_ctx * c = (b200_backend_ctx *) be->context;
    if (cudaSetDevice(c->device) != cudaSuccess) return GGML_STATUS_FAILED;
    uint64_t h1, h2;
    static const bool stats = getenv("GGML_B200_GRAPH_STATS") != nullptr;
    const uint64_t ti = stats ? host_now_ns() : 0;
    const bool same = graphs_on() && !c->ctx->profile && !c->ctx->trace && !c->ctx->capturing && graph_identity(g, h1, h2);
    if (stats) c->host_ns_identity += host_now_ns() - ti;
    if (!same) { ++c->n_eager; return run_nodes(c, g); }
    uint64_t epoch = g_layout_epoch.load() * 0x10001ull + b200_ctx_scratch_generation(c->ctx);
    for (b200_ctx * p : c->peer) if (p) epoch += 0x101ull * b200_ctx_scratch_generation(p) + 1;       // (a peer appearing or growing a block is a new epoch)
    if (epoch != c->graphs_epoch) {                            // recordings are stale; what was seen before is recorded again right away
        B200_CHECK(cudaStreamSynchronize(c->ctx->stream));
        for (auto & x : c->graphs) { if (x.exec) b200_graph_destroy(x.exec); x.exec = nullptr; x.failed = 0; }
        c->graphs_epoch = epoch;
    }
    b200_graph_entry * e = nullptr;
    for (auto & x : c->graphs) if (x.h1 == h1 && x.h2 == h2 && x.n_nodes == g->n_nodes) { e = &x; break; }
    if (!e) {                                                   // first sight: run eagerly (sizes the scratch blocks), remember
        if (c->graphs.size() >= 16) {
            size_t lru = 0; for (size_t i = 1; i < c->graphs.size(); ++i) if (c->graphs[i].used < c->graphs[lru].used) lru = i;
            if (c->graphs[lru].exec) { B200_CHECK(cudaStreamSynchronize(c->ctx->stream)); b200_graph_destroy(c->graphs[lru].exec); }
            c->graphs.erase(c->graphs.begin() + lru);
        }
        c->graphs.push_back({ h1, h2, g->n_nodes, nullptr, 1, 0, ++c->tick });
        ++c->n_eager;
        return run_nodes(c, g);
    }
    e->used = ++c->tick;
    if (!e->exec && !e->failed) {
        if (b200_graph_begin(c->ctx) != B200_OK) { e->failed = 1; cudaGetLastError(); }
        else {
            for (b200_ctx * p : c->peer) if (p) p->capturing = 1;                  // their streams join the capture through the hand-over events
            const enum ggml_status st = run_nodes(c, g);
            for (b200_ctx * p : c->peer) if (p) { p->capturing = 0; b200_ctx_invalidate(p); }
            void * exec = b200_graph_end(c->ctx);
            c->ctx->qcache.valid = 0; c->ctx->mcache.valid = 0;
            if (st == GGML_STATUS_SUCCESS && exec) { e->exec = exec; ++c->n_captured; }
            else {
                if (exec) b200_graph_destroy(exec);
                cudaGetLastError(); e->failed = 1;
                B200_LOG("graph_compute: recording a %d-node graph failed (%s): it keeps running launch by launch", g->n_nodes, b200_last_error());
            }
        }
    }
    if (e->exec) {
        cudaEvent_t e0 = nullptr, e1 = nullptr;
        if (stats && c->replay_events.size() < 8192) { cudaEventCreate(&e0); cudaEventCreate(&e1); cudaEventRecord(e0, c->ctx->stream); }
        const int rc = b200_graph_launch(c->ctx, e->exec);
        if (e0) { cudaEventRecord(e1, c->ctx->stream); c->replay_events.push_back({ e0, e1 }); }
        if (rc == B200_OK) { ++c->n_replayed; return GGML_STATUS_SUCCESS; }
        B200_LOG("graph_compute: graph launch failed: %s", b200_last_error());
        return GGML_STATUS_FAILED;
    }
    ++c->n_eager;
    return run_nodes(c, g);
}

static void be_event_record(ggml_backend_t be, ggml_backend_event_t ev) {
    b200_backend_ctx * c = (b200_backend_ctx *) be->context; B200_CHECK(cudaSetDevice(c->device));
    B200_CHECK(cudaEventRecord((cudaEvent_t) ev->context, c->ctx->stream));
}
static void be_event_wait(ggml_backend_t be, ggml_backend_event_t ev) {
    b200_backend_ctx * c = (b200_backend_ctx *) be->context; B200_CHECK(cudaSetDevice(c->device));
    B200_CHECK(cudaStreamWaitEvent(c->ctx->stream, (cudaEvent_t) ev->context, 0));
}
// graph_optimize (ggml_backend_sched calls it on every split BEFORE the graph allocator runs, ggml-backend.cpp:1321-1324, so moving a node
// cannot break buffer reuse): MUL_MATs of plain weight tensors that read the same activation are pulled up next to the first of them,
// same weight types side by side.  llama's attention block arrives as  Q = wq x, ROPE(Q), V = wv x, K = wk x, ROPE(K), SET_ROWS x 2:
// afterwards the three matmuls are adjacent (one launch for equal types, be_graph_compute) and so are the four glue nodes
// (b200_glue_rope_kv).  A weight matmul depends on its activation and a leaf only, so it may run anywhere after the activation exists.
static bool is_weight_leaf(const ggml_tensor * w) {
    return w && w->op == GGML_OP_NONE && w->view_src == nullptr && w->buffer && w->buffer->usage == GGML_BACKEND_BUFFER_USAGE_WEIGHTS;
}
static void graph_optimize_(ggml_cgraph * g);
static void be_graph_optimize(ggml_backend_t, ggml_cgraph * g) {
    static const bool off = [] { const char * e = getenv("GGML_B200_NO_REORDER"); return e && atoi(e); }();
    if (off || fuse_off()) return;
    static const bool stats = getenv("GGML_B200_GRAPH_STATS") != nullptr;
    const uint64_t t0 = stats ? host_now_ns() : 0;
    graph_optimize_(g);
    if (stats) { g_optimize_ns += host_now_ns() - t0; ++g_optimize_calls; }
}
static void graph_optimize_(ggml_cgraph * g) {
    for (int i = 0; i < g->n_nodes; ++i) {
        ggml_tensor * a = g->nodes[i];
        if (a->op != GGML_OP_MUL_MAT || !is_weight_leaf(a->src[0])) continue;
        int at = i + 1;                                             // where the next member of the group goes
        int members[3] = { i, -1, -1 }, nm = 1;
        for (int j = i + 1; j < g->n_nodes && j <= i + 16 && nm < 3; ++j) {
            ggml_tensor * b = g->nodes[j];
            if (b->op != GGML_OP_MUL_MAT || b->src[1] != a->src[1] || !is_weight_leaf(b->src[0])) continue;
            if (j != at) { for (int k = j; k > at; --k) g->nodes[k] = g->nodes[k - 1]; g->nodes[at] = b; }
            members[nm++] = at; ++at;
        }
        if (nm == 3) {                                              // equal weight types next to each other, first-come order otherwise
            ggml_tensor * m[3] = { g->nodes[members[0]], g->nodes[members[1]], g->nodes[members[2]] };
            if (m[0]->src[0]->type != m[1]->src[0]->type && m[0]->src[0]->type == m[2]->src[0]->type) { g->nodes[members[1]] = m[2]; g->nodes[members[2]] = m[1]; }
        }
        i = at - 1;
    }
    // ROPE(Q) [views] ROPE(K) [views] SET_ROWS(K result): run ROPE(K) first.  The allocator gives neither rope its input's memory (their
    // parents are reshapes), so in the original order ROPE(K)'s result lands in the block Q's matmul result has just vacated, and the two
    // ropes could not share a launch (b200_glue_rope_kv declines when one section writes what another reads).  With K first its result is
    // allocated while Q's input is still live, and Q's (larger) result cannot fit the block K's input vacates.
    for (int i = 0; i + 2 < g->n_nodes; ++i) {
        if (g->nodes[i]->op != GGML_OP_ROPE) continue;
        int j = i + 1; while (j < g->n_nodes && op_is_view((int) g->nodes[j]->op)) ++j;
        if (j >= g->n_nodes || g->nodes[j]->op != GGML_OP_ROPE) continue;
        int k = j + 1; while (k < g->n_nodes && op_is_view((int) g->nodes[k]->op)) ++k;
        if (k >= g->n_nodes || g->nodes[k]->op != GGML_OP_SET_ROWS) continue;
        const ggml_tensor * stored = g->nodes[k]->src[0];
        if (!stored || (stored != g->nodes[j] && stored->view_src != g->nodes[j])) continue;
        bool used = false;                                          // nothing in between may read ROPE(Q)'s result (only views are in between)
        for (int m = i + 1; m <= j && !used; ++m) for (int sidx = 0; sidx < GGML_MAX_SRC; ++sidx) used |= g->nodes[m]->src[sidx] == g->nodes[i];
        if (used) continue;
        ggml_tensor * rq = g->nodes[i];
        for (int m = i; m < j; ++m) g->nodes[m] = g->nodes[m + 1];
        g->nodes[j] = rq;
        i = k;
    }
}
static const ggml_backend_i k_backend_iface = {
    be_name, be_free, be_set_tensor_async, be_get_tensor_async, be_cpy_tensor_async, be_synchronize,
    nullptr, nullptr, nullptr, nullptr, be_graph_compute, be_event_record, be_event_wait, be_graph_optimize,
};

// ---- device -------------------------------------------------------------------------------------------
static const char * dev_name(ggml_backend_dev_t d) { return ((b200_device_ctx *) d->context)->name.c_str(); }
static const char * dev_desc(ggml_backend_dev_t d) { return ((b200_device_ctx *) d->context)->description.c_str(); }
static void dev_memory(ggml_backend_dev_t d, size_t * free_b, size_t * total_b) {
    b200_device_ctx * c = (b200_device_ctx *) d->context;
    *free_b = *total_b = 0;
    if (cudaSetDevice(c->device) == cudaSuccess) cudaMemGetInfo(free_b, total_b);
}
static enum ggml_backend_dev_type dev_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }
static void dev_props(ggml_backend_dev_t d, ggml_backend_dev_props * p) {
    p->name = dev_name(d); p->description = dev_desc(d); p->type = GGML_BACKEND_DEVICE_TYPE_GPU; p->device_id = nullptr;
    dev_memory(d, &p->memory_free, &p->memory_total);
    p->caps.async = true; p->caps.host_buffer = true; p->caps.buffer_from_host_ptr = false; p->caps.events = true;
}
static ggml_backend_t dev_init_backend(ggml_backend_dev_t d, const char *) {
    b200_device_ctx * c = (b200_device_ctx *) d->context;
    b200_ctx * ctx = b200_ctx_create(c->device);
    if (ctx && getenv("GGML_B200_PROFILE")) b200_ctx_profile(ctx, 1);
    if (!ctx) { B200_LOG("init_backend: %s", b200_last_error()); return nullptr; }
    b200_backend_ctx * bc = new b200_backend_ctx{ c->device, ctx, c->name };
    bc->index = c->index;
    return new ggml_backend{ &g_guid, k_backend_iface, d, bc };
}
static ggml_backend_buffer_type_t dev_buft(ggml_backend_dev_t d) { return &((b200_device_ctx *) d->context)->buft; }
static ggml_backend_buffer_type_t dev_host_buft(ggml_backend_dev_t) { return &g_host_buft; }

static bool mm_weight_ok(const ggml_tensor * w) {
    switch ((int) w->type) {
        case GGML_TYPE_Q4_0: case GGML_TYPE_Q8_0: case GGML_TYPE_Q4_K: case GGML_TYPE_Q5_K: case GGML_TYPE_Q6_K:
            return tensor_rows_contiguous(w) && w->ne[0] % t_blck(w->type) == 0 &&
                   (w->type == GGML_TYPE_Q4_K || w->type == GGML_TYPE_Q5_K ? (w->nb[1] % 16 == 0 && w->nb[2] % 16 == 0 && w->nb[3] % 16 == 0) : true);
        case GGML_TYPE_F32: case GGML_TYPE_F16: case GGML_TYPE_BF16:
            return tensor_rows_contiguous(w);
    }
    return false;
}

// GGML_B200_DECLINE_OPS=ROPE,GLU,... (debugging): leave these ops to the CPU backend, to bisect a wrong result in a whole graph
static bool op_declined(const ggml_tensor * op) {
    static const char * env = getenv("GGML_B200_DECLINE_OPS");
    if (!env) return false;
    static const struct { int op; const char * name; } names[] = {
        { GGML_OP_MUL_MAT, "MUL_MAT" }, { GGML_OP_MUL_MAT_ID, "MUL_MAT_ID" }, { GGML_OP_FLASH_ATTN_EXT, "FLASH_ATTN_EXT" }, { GGML_OP_RMS_NORM, "RMS_NORM" },
        { GGML_OP_ADD, "ADD" }, { GGML_OP_MUL, "MUL" }, { GGML_OP_ROPE, "ROPE" }, { GGML_OP_SET_ROWS, "SET_ROWS" }, { GGML_OP_GET_ROWS, "GET_ROWS" },
        { GGML_OP_CPY, "CPY" }, { GGML_OP_CONT, "CONT" }, { GGML_OP_GLU, "GLU" }, { GGML_OP_SCALE, "SCALE" } };
    for (const auto & n : names) {
        if (n.op != (int) op->op) continue;
        const char * f = strstr(env, n.name);
        while (f) {                                                  // whole-word match (MUL vs MUL_MAT)
            const char after = f[strlen(n.name)]; const bool start_ok = f == env || f[-1] == ',';
            if (start_ok && (after == 0 || after == ',')) {
                if (!strcmp(n.name, "MUL_MAT") && getenv("GGML_B200_DECLINE_DECODE_ONLY") && op->ne[1] > 1) return false;
                return true;
            }
            f = strstr(f + 1, n.name);
        }
    }
    return false;
}
static bool dev_supports_op(ggml_backend_dev_t, const ggml_tensor * op) {
    if (op_declined(op)) return false;
    // row-split tensors are weights of plain MUL_MATs and nothing else (as ggml-cuda.cu:4304-4311, :4371-4374)
    for (int i = 0; i < GGML_MAX_SRC; ++i) {
        if (!op->src[i] || !op->src[i]->buffer || !b200_buft_is_split(op->src[i]->buffer->buft)) continue;
        if (op->op != GGML_OP_MUL_MAT || i != 0 || op->src[0]->ne[2] * op->src[0]->ne[3] != 1 || op->src[0]->view_src) return false;
    }
    switch ((int) op->op) {
        case GGML_OP_NONE: case GGML_OP_RESHAPE: case GGML_OP_VIEW: case GGML_OP_PERMUTE: case GGML_OP_TRANSPOSE:
            return true;
        case GGML_OP_MUL_MAT: {
            const ggml_tensor * w = op->src[0], * x = op->src[1];
            return mm_weight_ok(w) && x->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && x->nb[0] == 4 && op->nb[0] == 4 &&
                   x->ne[2] * x->ne[3] <= 65535;
        }
        case GGML_OP_MUL_MAT_ID: {
            const ggml_tensor * w = op->src[0], * x = op->src[1], * ids = op->src[2];
            return mm_weight_ok(w) && t_blck(w->type) > 1 && x->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && ids->type == GGML_TYPE_I32 &&
                   x->nb[0] == 4 && op->nb[0] == 4 && ids->nb[0] == 4 && ids->ne[0] * ids->ne[1] <= 65535;
        }
        case GGML_OP_FLASH_ATTN_EXT: {
            const ggml_tensor * q = op->src[0], * k = op->src[1], * v = op->src[2], * m = op->src[3];
            const int64_t DK = k->ne[0], DV = v->ne[0];
            if (q->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32) return false;
            if (m && m->type != GGML_TYPE_F16) return false;
            if (k->type != GGML_TYPE_F16 || v->type != GGML_TYPE_F16) {          // quantized KV cache: q8_0 / q4_0 at head size 64 / 128 (b200_fattn_q.cu)
                if (q->nb[0] != 4 || q->ne[2] > 65535 || q->ne[3] > 65535) return false;
                const b200_tensor bq = to_b200(q), bk = to_b200(k), bv = to_b200(v);
                return b200_fattn_kvq_supported(&bq, &bk, &bv);
            }
            if (DK % 64 != 0 || DK > 256 || (DV != 64 && DV != 128 && DV != 256)) return false;
            if (q->nb[0] != 4 || k->nb[0] != 2 || v->nb[0] != 2) return false;
            if (k->nb[1] % 16 || k->nb[2] % 16 || k->nb[3] % 16 || v->nb[1] % 4 || v->nb[2] % 4 || v->nb[3] % 4) return false;
            if (q->ne[2] > 65535 || q->ne[3] > 65535) return false;
            return true;
        }
        default:
            for (int i = 0; i < GGML_MAX_SRC; ++i) if (op->src[i] && tensor_is_repacked(op->src[i])) return false;   // only the matmuls read device layouts
            return b200_glue_supports(op);
    }
}
static bool dev_supports_buft(ggml_backend_dev_t d, ggml_backend_buffer_type_t buft) {
    if (buft_is_b200_host(buft)) return false;   // host tensors are copied in by the scheduler
    return (buft_is_b200(buft) || b200_buft_is_split(buft)) && buft->device == d;
}
static bool dev_offload_op(ggml_backend_dev_t, const ggml_tensor * op) {
    // worth shipping host-resident weights over only for real batches; the batch size of an op as the reference counts it
    // (get_op_batch_size, ggml/src/ggml-cuda/ggml-cuda.cu:4675-4694)
    const int64_t min_batch = 32;
    int64_t batch;
    switch ((int) op->op) {
        case GGML_OP_GET_ROWS:   batch = 0; break;
        case GGML_OP_MUL_MAT:    batch = op->ne[1]; break;
        case GGML_OP_MUL_MAT_ID: case GGML_OP_ROPE: batch = op->ne[2]; break;
        default:                 batch = op->ne[1] * op->ne[2] * op->ne[3]; break;
    }
    return batch >= min_batch;
}
static ggml_backend_event_t dev_event_new(ggml_backend_dev_t d) {
    b200_device_ctx * c = (b200_device_ctx *) d->context; B200_CHECK(cudaSetDevice(c->device));
    cudaEvent_t ev; B200_CHECK(cudaEventCreateWithFlags(&ev, cudaEventDisableTiming));
    return new ggml_backend_event{ d, ev };
}
static void dev_event_free(ggml_backend_dev_t, ggml_backend_event_t ev) { cudaEventDestroy((cudaEvent_t) ev->context); delete ev; }
static void dev_event_sync(ggml_backend_dev_t, ggml_backend_event_t ev) { B200_CHECK(cudaEventSynchronize((cudaEvent_t) ev->context)); }
static const ggml_backend_device_i k_device_iface = {
    dev_name, dev_desc, dev_memory, dev_type, dev_props, dev_init_backend, dev_buft, dev_host_buft, nullptr,
    dev_supports_op, dev_supports_buft, dev_offload_op, dev_event_new, dev_event_free, dev_event_sync,
};

// ---- registry -------------------------------------------------------------------------------------------
static const char * reg_name(ggml_backend_reg_t) { return "B200"; }
static size_t reg_dev_count(ggml_backend_reg_t) { return g_reg_ctx.devices.size(); }
static ggml_backend_dev_t reg_get_dev(ggml_backend_reg_t, size_t i) { return i < g_reg_ctx.devices.size() ? &g_reg_ctx.devices[i]->dev : nullptr; }

static ggml_backend_feature g_features[] = { { "ARCH", "sm_100a" }, { "TCGEN05", "1" }, { "TMA", "1" }, { nullptr, nullptr } };
static ggml_backend_feature * reg_get_features(ggml_backend_reg_t) { return g_features; }

ggml_backend_buffer_type_t b200_split_buffer_type(int main_device, const float * tensor_split);   // b200_split.cpp

static void * reg_proc_address(ggml_backend_reg_t, const char * name) {
    if (!strcmp(name, "ggml_backend_get_features"))      return (void *) reg_get_features;
    if (!strcmp(name, "ggml_backend_split_buffer_type")) return (void *) b200_split_buffer_type;
    if (!strcmp(name, "ggml_backend_b200_graph_identity"))      // host-only: the key of the CUDA-graph cache (tests/reorder_probe.cpp)
        return (void *) +[](const ggml_cgraph * g, uint64_t * h1, uint64_t * h2) -> int { return graph_identity(g, *h1, *h2) ? 1 : 0; };
    if (!strcmp(name, "ggml_backend_b200_graph_optimize")) return (void *) +[](ggml_cgraph * g) { be_graph_optimize(nullptr, g); };   // host-only: tests/reorder_probe.cpp
    if (!strcmp(name, "ggml_backend_b200_tensor_is_repacked")) return (void *) +[](const ggml_tensor * t) -> int { return tensor_is_repacked(t) ? 1 : 0; };   // tests/upload_probe.cpp
    return nullptr;    // "ggml_backend_set_n_threads", "ggml_backend_dev_get_extra_bufts": not applicable
}
static const ggml_backend_reg_i k_reg_iface = { reg_name, reg_dev_count, reg_get_dev, reg_proc_address };

extern "C" B200_API ggml_backend_reg_t ggml_backend_b200_reg(void) {
    static std::once_flag once;
    std::call_once(once, [] {
        g_reg.api_version = GGML_BACKEND_API_VERSION; g_reg.iface = k_reg_iface; g_reg.context = &g_reg_ctx;
        g_host_buft.iface = k_hbuft_iface; g_host_buft.device = nullptr; g_host_buft.context = nullptr;
        const int n_cuda = b200_device_count();
        // GGML_B200_FAKE_DEVICES=N: list N devices, dealt round-robin onto the GPUs that exist, so that a one-GPU box can run the
        // multi-device paths (split-mode row, layer split) end to end; each keeps its own stream, buffers and row slices
        const char * fake = getenv("GGML_B200_FAKE_DEVICES");
        const int n = (fake && atoi(fake) > 0 && n_cuda > 0) ? atoi(fake) : n_cuda;
        for (int i = 0; i < n && i < B200_MAX_DEVICES; ++i) {
            cudaDeviceProp prop; if (cudaGetDeviceProperties(&prop, i % n_cuda) != cudaSuccess) continue;
            b200_device_ctx * d = new b200_device_ctx;
            d->device = i % n_cuda; d->index = (int) g_reg_ctx.devices.size(); d->name = "B200_" + std::to_string(i); d->description = prop.name;
            d->dev.iface = k_device_iface; d->dev.reg = &g_reg; d->dev.context = d;
            d->buft.iface = k_buft_iface; d->buft.device = &d->dev; d->buft.context = d;
            g_reg_ctx.devices.push_back(d);
        }
        if (!g_reg_ctx.devices.empty()) g_host_buft.device = &g_reg_ctx.devices[0]->dev;
    });
    return &g_reg;
}
extern "C" B200_API ggml_backend_reg_t ggml_backend_init(void) { return ggml_backend_b200_reg(); }
extern "C" B200_API int ggml_backend_score(void) { return b200_device_count() > 0 ? 100 : 0; }
