// b200_api.cu -- the flat C-ABI (include/ggml-b200.h): context, memory, operator dispatch.
// The ggml plugin (b200_backend.cpp) and the ctypes bindings (tests/, bench.py) both land here.
#include "b200_common.cuh"
#include <stdarg.h>
#include <string.h>
#include <stdlib.h>

static thread_local char g_err[512] = "";

void b200_set_error(const char * fmt, ...) {
    va_list ap; va_start(ap, fmt); vsnprintf(g_err, sizeof(g_err), fmt, ap); va_end(ap);
    if (getenv("GGML_B200_DEBUG")) fprintf(stderr, "ggml-b200: %s\n", g_err);
}
extern "C" const char * b200_last_error(void) { return g_err; }
static void prof_clear(b200_ctx * ctx);

extern "C" int b200_device_count(void) {
    int n = 0;
    if (cudaGetDeviceCount(&n) != cudaSuccess) { cudaGetLastError(); return 0; }
    int ok = 0;
    for (int i = 0; i < n; ++i) {
        cudaDeviceProp prop;
        if (cudaGetDeviceProperties(&prop, i) == cudaSuccess && prop.major == 10) ok++;   // sm_100a code only runs on cc 10.x
    }
    return ok == n ? n : 0;
}

extern "C" b200_ctx * b200_ctx_create(int device) {
    const int n = b200_device_count();
    if (device < 0 || device >= n) { b200_set_error("b200_ctx_create: device %d not available (%d sm_100 devices)", device, n); return nullptr; }
    if (cudaSetDevice(device) != cudaSuccess) { b200_set_error("cudaSetDevice(%d) failed", device); return nullptr; }
    b200_ctx * ctx = (b200_ctx *) calloc(1, sizeof(b200_ctx));
    ctx->device = device; ctx->prof_open = -1;
    { const char * e = getenv("GGML_B200_NO_PDL"); ctx->pdl = (e && atoi(e)) ? 0 : 1; }
    if (cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking) != cudaSuccess) { b200_set_error("stream create failed"); free(ctx); return nullptr; }
    cudaDeviceGetAttribute(&ctx->sm_count, cudaDevAttrMultiProcessorCount, device);
    return ctx;
}
extern "C" void b200_ctx_destroy(b200_ctx * ctx) {
    if (!ctx) return;
    cudaSetDevice(ctx->device);
    cudaStreamSynchronize(ctx->stream);
    prof_clear(ctx); free(ctx->prof);
    for (int i = 0; i < 8; ++i) if (ctx->scratch[i]) cudaFree(ctx->scratch[i]);
    cudaStreamDestroy(ctx->stream);
    free(ctx);
}
// ---- per-kernel event profile ------------------------------------------------------------------
void b200_prof_begin(b200_ctx * ctx, int kind) {
    if (ctx->capturing) return;
    if (ctx->prof_n == ctx->prof_cap) {
        ctx->prof_cap = ctx->prof_cap ? ctx->prof_cap * 2 : 4096;
        ctx->prof = (b200_prof_rec *) realloc(ctx->prof, sizeof(b200_prof_rec) * ctx->prof_cap);
    }
    b200_prof_rec * r = &ctx->prof[ctx->prof_n];
    r->kind = kind;
    if (cudaEventCreate(&r->e0) != cudaSuccess || cudaEventCreate(&r->e1) != cudaSuccess) { ctx->prof_open = -1; return; }
    cudaEventRecord(r->e0, ctx->stream);
    ctx->prof_open = ctx->prof_n++;
}
void b200_prof_end(b200_ctx * ctx) {
    if (ctx->capturing || ctx->prof_open < 0) return;
    cudaEventRecord(ctx->prof[ctx->prof_open].e1, ctx->stream);
    ctx->prof_open = -1;
}
static void prof_clear(b200_ctx * ctx) {
    for (int i = 0; i < ctx->prof_n; ++i) { cudaEventDestroy(ctx->prof[i].e0); cudaEventDestroy(ctx->prof[i].e1); }
    ctx->prof_n = 0; ctx->prof_open = -1;
}
extern "C" int b200_ctx_profile(b200_ctx * ctx, int enable) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    prof_clear(ctx);
    ctx->profile = enable ? 1 : 0;
    return B200_OK;
}
extern "C" int b200_ctx_profile_get(b200_ctx * ctx, int kind, double * ms_total, uint64_t * launches) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    double t = 0.0; uint64_t n = 0;
    for (int i = 0; i < ctx->prof_n; ++i) if (ctx->prof[i].kind == kind) {
        float ms = 0.0f;
        B200_CUDA_TRY(cudaEventElapsedTime(&ms, ctx->prof[i].e0, ctx->prof[i].e1));
        t += ms; n++;
    }
    if (ms_total) *ms_total = t;
    if (launches) *launches = n;
    return B200_OK;
}
extern "C" uint64_t b200_ctx_launch_count_kind(const b200_ctx * ctx, int kind) { return kind >= 0 && kind < B200_K_COUNT ? ctx->launches_kind[kind] : 0; }

// ---- events and graphs ---------------------------------------------------------------------------
extern "C" void * b200_event_create(void) { cudaEvent_t e; if (cudaEventCreate(&e) != cudaSuccess) { cudaGetLastError(); return nullptr; } return (void *) e; }
extern "C" int b200_event_destroy(void * ev) { B200_CUDA_TRY(cudaEventDestroy((cudaEvent_t) ev)); return B200_OK; }
extern "C" int b200_event_record(b200_ctx * ctx, void * ev) { B200_CUDA_TRY(cudaEventRecord((cudaEvent_t) ev, ctx->stream)); return B200_OK; }
extern "C" int b200_event_elapsed_ms(void * ev0, void * ev1, float * ms) {
    B200_CUDA_TRY(cudaEventSynchronize((cudaEvent_t) ev1));
    B200_CUDA_TRY(cudaEventElapsedTime(ms, (cudaEvent_t) ev0, (cudaEvent_t) ev1));
    return B200_OK;
}
extern "C" int b200_debug_trace(b200_ctx * ctx, void * buf, int cap) { ctx->trace = (unsigned long long *) buf; ctx->trace_cap = buf ? cap : 0; return B200_OK; }
extern "C" int b200_graph_begin(b200_ctx * ctx) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamBeginCapture(ctx->stream, cudaStreamCaptureModeThreadLocal));
    ctx->capturing = 1;
    return B200_OK;
}
extern "C" void * b200_graph_end(b200_ctx * ctx) {
    cudaGraph_t g = nullptr; cudaGraphExec_t ge = nullptr;
    ctx->capturing = 0;
    if (cudaStreamEndCapture(ctx->stream, &g) != cudaSuccess || !g) { b200_set_error("graph capture failed: %s", cudaGetErrorString(cudaGetLastError())); return nullptr; }
    if (cudaGraphInstantiate(&ge, g, 0) != cudaSuccess) { b200_set_error("graph instantiate failed: %s", cudaGetErrorString(cudaGetLastError())); ge = nullptr; }
    cudaGraphDestroy(g);
    return (void *) ge;
}
extern "C" int b200_graph_launch(b200_ctx * ctx, void * ge) { B200_CUDA_TRY(cudaGraphLaunch((cudaGraphExec_t) ge, ctx->stream)); return B200_OK; }
extern "C" int b200_graph_destroy(void * ge) { B200_CUDA_TRY(cudaGraphExecDestroy((cudaGraphExec_t) ge)); return B200_OK; }

extern "C" uint64_t b200_ctx_scratch_generation(const b200_ctx * ctx) { return ctx->scratch_gen; }
extern "C" int      b200_ctx_device(const b200_ctx * ctx) { return ctx->device; }
extern "C" void *   b200_ctx_stream(const b200_ctx * ctx) { return (void *) ctx->stream; }
extern "C" uint64_t b200_ctx_launch_count(const b200_ctx * ctx) { return ctx->launches; }
extern "C" int      b200_ctx_synchronize(b200_ctx * ctx) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return B200_OK;
}

void * b200_scratch(b200_ctx * ctx, int slot, size_t nbytes) {
    if (ctx->scratch_size[slot] >= nbytes && ctx->scratch[slot]) return ctx->scratch[slot];
    if (ctx->capturing) { b200_set_error("scratch slot %d must grow to %zu bytes during graph capture: run the step once eagerly first", slot, nbytes); return nullptr; }
    // growing: earlier kernels on the stream may still read the old block -> drain first (rare, sizes converge fast)
    cudaStreamSynchronize(ctx->stream);
    if (ctx->scratch[slot]) cudaFree(ctx->scratch[slot]);
    size_t want = nbytes + nbytes / 2; if (want < (1u << 20)) want = 1u << 20;
    ctx->scratch[slot] = nullptr; ctx->scratch_size[slot] = 0; ++ctx->scratch_gen;
    if (cudaMalloc(&ctx->scratch[slot], want) != cudaSuccess) { cudaGetLastError(); b200_set_error("scratch alloc of %zu bytes failed", want); return nullptr; }
    ctx->scratch_size[slot] = want;
    return ctx->scratch[slot];
}

extern "C" void * b200_malloc(b200_ctx * ctx, size_t nbytes) {
    void * p = nullptr;
    if (cudaSetDevice(ctx->device) != cudaSuccess || cudaMalloc(&p, nbytes ? nbytes : 1) != cudaSuccess) {
        cudaGetLastError(); b200_set_error("cudaMalloc(%zu) failed", nbytes); return nullptr;
    }
    return p;
}
extern "C" int b200_free(b200_ctx * ctx, void * dptr) { ctx->qcache.valid = 0; B200_CUDA_TRY(cudaSetDevice(ctx->device)); B200_CUDA_TRY(cudaFree(dptr)); return B200_OK; }
extern "C" void * b200_host_alloc(size_t nbytes) {
    void * p = nullptr;
    if (cudaMallocHost(&p, nbytes ? nbytes : 1) != cudaSuccess) { cudaGetLastError(); b200_set_error("cudaMallocHost(%zu) failed", nbytes); return nullptr; }
    return p;
}
extern "C" int b200_host_free(void * hptr) { B200_CUDA_TRY(cudaFreeHost(hptr)); return B200_OK; }
extern "C" int b200_memcpy_h2d(b200_ctx * ctx, void * dst, const void * src, size_t n) {
    b200_note_write(ctx, dst, n);
    B200_CUDA_TRY(cudaMemcpyAsync(dst, src, n, cudaMemcpyHostToDevice, ctx->stream)); return B200_OK;
}
extern "C" int b200_memcpy_d2h(b200_ctx * ctx, void * dst, const void * src, size_t n) {
    B200_CUDA_TRY(cudaMemcpyAsync(dst, src, n, cudaMemcpyDeviceToHost, ctx->stream)); return B200_OK;
}
extern "C" int b200_memset(b200_ctx * ctx, void * dst, int value, size_t n) {
    b200_note_write(ctx, dst, n);
    B200_CUDA_TRY(cudaMemsetAsync(dst, value, n, ctx->stream)); return B200_OK;
}

// ---- operators -----------------------------------------------------------------------------
extern "C" int b200_quantize_q8_1(b200_ctx * ctx, const float * x, size_t x_row_stride, void * y, int64_t k, int64_t nrows, int variant) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    return b200_launch_quantize_q8_1(ctx, x, x_row_stride, y, k, nrows, variant);
}

size_t b200_tensor_span(const b200_tensor * t) {                 // ggml_nbytes
    for (int i = 0; i < 4; ++i) if (t->ne[i] <= 0) return 0;
    const int64_t blck = b200_blck_size(t->type);
    size_t n = blck > 1 ? (size_t) t->ne[0] * t->nb[0] / blck : b200_type_size(t->type) + (size_t) (t->ne[0] - 1) * t->nb[0];
    for (int i = 1; i < 4; ++i) n += (size_t) (t->ne[i] - 1) * t->nb[i];
    return n;
}
void b200_note_write(b200_ctx * ctx, const void * ptr, size_t nbytes) {
    const char * b = (const char *) ptr;
    if (ctx->qcache.valid) { const char * a = (const char *) ctx->qcache.data; if (b < a + ctx->qcache.span && a < b + nbytes) ctx->qcache.valid = 0; }
    if (ctx->mcache.valid) { const char * a = (const char *) ctx->mcache.data; if (b < a + ctx->mcache.span && a < b + nbytes) ctx->mcache.valid = 0; }
}
extern "C" int b200_ctx_note_write(b200_ctx * ctx, const void * ptr, size_t nbytes) { b200_note_write(ctx, ptr, nbytes); return B200_OK; }
extern "C" int b200_ctx_invalidate(b200_ctx * ctx) { ctx->qcache.valid = 0; ctx->mcache.valid = 0; return B200_OK; }

// largest K whose activation column every GEMV CTA quantizes itself; above it one quantize launch writes an image (GGML_B200_FUSED_QUANT_MAX_K)
static int64_t fused_quant_max_k() {
    static int64_t v = -1;
    if (v < 0) { const char * e = getenv("GGML_B200_FUSED_QUANT_MAX_K"); v = e ? atoll(e) : B200_GEMV_FUSED_QUANT_MAX_K; }
    return v;
}
// q8_1 image of x in scratch slot 0, reused when the same activation tensor feeds consecutive matmuls
static const void * quantized_activations(b200_ctx * ctx, const b200_tensor * x, int variant, int * rc) {
    const int64_t K = x->ne[0], ncol = x->ne[1] * x->ne[2] * x->ne[3];
    auto & qc = ctx->qcache;
    *rc = B200_OK;
    if (qc.valid && qc.data == x->data && qc.variant == variant && ctx->scratch[0] &&
        !memcmp(qc.ne, x->ne, sizeof(qc.ne)) && !memcmp(qc.nb, x->nb, sizeof(qc.nb))) return ctx->scratch[0];
    qc.valid = 0;
    void * image = b200_scratch(ctx, 0, variant == B200_ACT_BF16 ? (size_t) ncol * K * 2 : b200_q8_1_soa_size(K, ncol));
    if (!image) { *rc = B200_ERR_ALLOC; return nullptr; }
    if (variant == B200_ACT_BF16) *rc = b200_launch_act_bf16(ctx, (const float *) x->data, x->nb[1], K, x->ne[1], image);   // 2-D only (b200_gemm_supported)
    else *rc = b200_launch_quantize_q8_1_soa(ctx, (const float *) x->data, x->nb[1], x->nb[2], x->nb[3], K, x->ne[1], x->ne[2], x->ne[3], image, variant);
    if (*rc) return nullptr;
    qc.data = x->data; memcpy(qc.ne, x->ne, sizeof(qc.ne)); memcpy(qc.nb, x->nb, sizeof(qc.nb));
    qc.span = b200_tensor_span(x); qc.variant = variant; qc.valid = 1;
    return image;
}

static bool is_quant(int t) { return t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K; }
static bool is_float_w(int t) { return t == B200_TYPE_F32 || t == B200_TYPE_F16 || t == B200_TYPE_BF16; }

static int q8_1_variant_env() {
    static int v = -1;
    if (v < 0) { const char * e = getenv("GGML_B200_Q8_1_REF"); v = (e && atoi(e)) ? B200_Q8_1_REF : B200_Q8_1_X86; }
    return v;
}

extern "C" int b200_mul_mat(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst) {
    if (!ctx || !w || !x || !dst) { b200_set_error("mul_mat: null argument"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int64_t K = w->ne[0], M = w->ne[1];
    if (x->ne[0] != K || dst->ne[0] != M || dst->ne[1] != x->ne[1] || dst->ne[2] != x->ne[2] || dst->ne[3] != x->ne[3] ||
        w->ne[2] <= 0 || w->ne[3] <= 0 || x->ne[2] % w->ne[2] != 0 || x->ne[3] % w->ne[3] != 0) {
        b200_set_error("mul_mat: shape mismatch"); return B200_ERR_INVALID;
    }
    if (x->type != B200_TYPE_F32 || dst->type != B200_TYPE_F32 || x->nb[0] != 4 || dst->nb[0] != 4) {
        b200_set_error("mul_mat: activations and result must be f32 with contiguous rows"); return B200_ERR_UNSUPPORTED;
    }
    if (M == 0 || x->ne[1] * x->ne[2] * x->ne[3] == 0) return B200_OK;
    if (is_quant(w->type)) {
        if (K % b200_blck_size(w->type) != 0) { b200_set_error("mul_mat: K=%lld not a multiple of the block size", (long long) K); return B200_ERR_INVALID; }
        if (w->nb[0] != b200_type_size(w->type)) { b200_set_error("mul_mat: quantized weight rows must be contiguous"); return B200_ERR_UNSUPPORTED; }
        int rc;
        if (x->ne[1] > B200_GEMV_MAX_COLS && b200_gemm_supported(w, x)) {
            const void * image = quantized_activations(ctx, x, B200_ACT_BF16, &rc);   // bf16 [N][K], shared by wq/wk/wv and gate/up
            if (!image) return rc;
            rc = b200_launch_gemm_q(ctx, w, image, x->ne[1], dst);
        }
        else if (q8_1_variant_env() == B200_Q8_1_X86 && b200_gemv_stream_supported(w, x)) {
            float * d = (float *) dst->data;
            const void * image = nullptr;
            if (w->ne[0] > fused_quant_max_k()) { image = quantized_activations(ctx, x, B200_Q8_1_X86, &rc); if (!image) return rc; }
            rc = b200_launch_gemv_stream(ctx, &w, 1, x, image, &d);      // TMA-streamed GEMV, q8_1 quantization fused for small K
        } else {
            const void * image = quantized_activations(ctx, x, q8_1_variant_env(), &rc);
            if (!image) return rc;
            rc = b200_launch_gemv_q(ctx, w, image, x->ne[1], x->ne[2], x->ne[3], dst);
        }
        b200_note_write(ctx, dst->data, b200_tensor_span(dst));
        return rc;
    }
    if (is_float_w(w->type)) { const int rc = b200_launch_matmul_float(ctx, w, x, dst); b200_note_write(ctx, dst->data, b200_tensor_span(dst)); return rc; }
    b200_set_error("mul_mat: weight type %d is outside the hot path", w->type);
    return B200_ERR_UNSUPPORTED;
}

// gate = wg x, up = wu x, glu = swiglu(gate, up) for one activation column in ONE launch: the tensor-core q4_K GEMV hands every CTA the
// same rows of both weights, so its epilogue can finish the GLU node as well (no launch, no re-read of the two n_ff columns).
// B200_ERR_UNSUPPORTED when that kernel does not serve the shape: the caller runs b200_mul_mat_multi and the GLU op then.
extern "C" int b200_mul_mat_gate_up_swiglu(b200_ctx * ctx, const b200_tensor * wg, const b200_tensor * wu, const b200_tensor * x,
                                           b200_tensor * gate, b200_tensor * up, b200_tensor * glu) {
    if (!ctx || !wg || !wu || !x || !gate || !up || !glu) { b200_set_error("mul_mat_gate_up_swiglu: null argument"); return B200_ERR_INVALID; }
    const b200_tensor * w[2] = { wg, wu };
    const int64_t K = wg->ne[0], M = wg->ne[1];
    b200_tensor * cols[3] = { gate, up, glu };
    for (b200_tensor * t : cols)
        if (t->type != B200_TYPE_F32 || t->nb[0] != 4 || t->ne[0] != M || t->ne[1] != 1 || t->ne[2] != 1 || t->ne[3] != 1) { b200_set_error("mul_mat_gate_up_swiglu: outputs are f32 columns of M values"); return B200_ERR_UNSUPPORTED; }
    if (wu->ne[0] != K || wu->ne[1] != M || x->ne[0] != K || x->ne[1] != 1 || x->ne[2] != 1 || x->ne[3] != 1 || x->type != B200_TYPE_F32 || x->nb[0] != 4) { b200_set_error("mul_mat_gate_up_swiglu: shape mismatch"); return B200_ERR_UNSUPPORTED; }
    if (q8_1_variant_env() != B200_Q8_1_X86 || !b200_gemv_stream_supported(wg, x) || !b200_gemv_stream_supported(wu, x) || !b200_gemv_mma_supported(ctx, w, 2, x)) {
        b200_set_error("mul_mat_gate_up_swiglu: not a tensor-core GEMV shape"); return B200_ERR_UNSUPPORTED;
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    int rc;
    const void * image = nullptr;
    if (K > fused_quant_max_k()) { image = quantized_activations(ctx, x, B200_Q8_1_X86, &rc); if (!image) return rc; }
    float * d[2] = { (float *) gate->data, (float *) up->data };
    rc = b200_launch_gemv_mma(ctx, w, 2, x, image, d, nullptr, (float *) glu->data);
    for (b200_tensor * t : cols) b200_note_write(ctx, t->data, b200_tensor_span(t));
    return rc;
}

// n MUL_MATs that share src1 (wq/wk/wv, ffn_gate/ffn_up): one launch when the streaming GEMV can take them all
extern "C" int b200_mul_mat_multi(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, b200_tensor * const * dst) {
    if (!ctx || !w || !x || !dst || n <= 0) { b200_set_error("mul_mat_multi: bad arguments"); return B200_ERR_INVALID; }
    bool fuse = n > 1 && n <= 3 && q8_1_variant_env() == B200_Q8_1_X86;
    for (int i = 0; i < n && fuse; ++i) {
        fuse = b200_gemv_stream_supported(w[i], x) && w[i]->type == w[0]->type && w[i]->ne[0] == w[0]->ne[0] && x->ne[0] == w[i]->ne[0] &&
               dst[i]->type == B200_TYPE_F32 && dst[i]->nb[0] == 4 && dst[i]->ne[0] == w[i]->ne[1] && dst[i]->ne[1] == 1 && dst[i]->ne[2] == 1 && dst[i]->ne[3] == 1;
    }
    // prefill: the same fusion for the tcgen05 GEMM (grid.y walks the row tiles of all weights)
    bool fuse_gemm = !fuse && n > 1 && n <= 3 && x->ne[1] > B200_GEMV_MAX_COLS;
    for (int i = 0; i < n && fuse_gemm; ++i)
        fuse_gemm = is_quant(w[i]->type) && b200_gemm_supported(w[i], x) && w[i]->type == w[0]->type && w[i]->ne[0] == w[0]->ne[0] && x->ne[0] == w[i]->ne[0] &&
                    dst[i]->type == B200_TYPE_F32 && dst[i]->nb[0] == 4 && dst[i]->ne[0] == w[i]->ne[1] && dst[i]->ne[1] == x->ne[1] && dst[i]->ne[2] == 1 && dst[i]->ne[3] == 1;
    if (fuse_gemm) {
        B200_CUDA_TRY(cudaSetDevice(ctx->device));
        int rc;
        const void * image = quantized_activations(ctx, x, B200_ACT_BF16, &rc);
        if (!image) return rc;
        rc = b200_launch_gemm_multi(ctx, w, n, image, x->ne[1], dst);
        for (int i = 0; i < n; ++i) b200_note_write(ctx, dst[i]->data, b200_tensor_span(dst[i]));
        return rc;
    }
    if (!fuse) {
        // three weights of two types (Q4_K_M promotes attn_v to q6_K on half the layers): the two of equal type still share a launch
        if (n == 3 && !(w[0]->type == w[1]->type && w[1]->type == w[2]->type)) {
            const int a = w[0]->type == w[1]->type ? 0 : w[0]->type == w[2]->type ? 0 : w[1]->type == w[2]->type ? 1 : -1;
            if (a >= 0) {
                const int b = (a == 0 && w[0]->type == w[1]->type) ? 1 : 2, o = 3 - a - b;
                const b200_tensor * wp[2] = { w[a], w[b] }; b200_tensor * dp[2] = { dst[a], dst[b] };
                int rc = b200_mul_mat_multi(ctx, wp, 2, x, dp);
                if (rc) return rc;
                return b200_mul_mat(ctx, w[o], x, dst[o]);
            }
        }
        for (int i = 0; i < n; ++i) { const int rc = b200_mul_mat(ctx, w[i], x, dst[i]); if (rc) return rc; }
        return B200_OK;
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    float * d[3];
    for (int i = 0; i < n; ++i) d[i] = (float *) dst[i]->data;
    int rc;
    const void * image = nullptr;
    if (w[0]->ne[0] > fused_quant_max_k()) { image = quantized_activations(ctx, x, B200_Q8_1_X86, &rc); if (!image) return rc; }
    rc = b200_launch_gemv_stream(ctx, w, n, x, image, d);
    for (int i = 0; i < n; ++i) b200_note_write(ctx, dst[i]->data, b200_tensor_span(dst[i]));
    return rc;
}

// ---- split-mode=row over peer memory ----------------------------------------------------------------------------------------
extern "C" int b200_peer_handle(b200_ctx * ctx, void * dptr, void * handle64) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    static_assert(sizeof(cudaIpcMemHandle_t) == 64, "ipc handle size");
    B200_CUDA_TRY(cudaIpcGetMemHandle((cudaIpcMemHandle_t *) handle64, dptr));
    return B200_OK;
}
extern "C" int b200_peer_open(b200_ctx * ctx, const void * handle64, void ** dptr) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    cudaIpcMemHandle_t h; memcpy(&h, handle64, sizeof(h));
    B200_CUDA_TRY(cudaIpcOpenMemHandle(dptr, h, cudaIpcMemLazyEnablePeerAccess));
    return B200_OK;
}
extern "C" int b200_peer_close(b200_ctx * ctx, void * dptr) {
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    B200_CUDA_TRY(cudaIpcCloseMemHandle(dptr));
    return B200_OK;
}
static bool exchange_ok(const b200_exchange * ex) { return ex && ex->world >= 1 && ex->world <= B200_MAX_PEERS && ex->rank >= 0 && ex->rank < ex->world && ex->base[ex->rank]; }

// n <= 3 row-split MUL_MATs of one activation column: weight i holds rows [row_lo[i], row_lo[i] + w[i]->ne[1]) of dst[i].
// Runs of equal weight type share one launch (compute + peer stores + receive in ONE kernel); a rank with no rows only receives.
extern "C" int b200_mul_mat_rowsplit_multi(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, b200_tensor * const * dst,
                                           const int64_t * row_lo, const b200_exchange * ex) {
    if (!ctx || !w || !x || !dst || !row_lo || n < 1 || n > 3 || !exchange_ok(ex)) { b200_set_error("mul_mat_rowsplit: bad arguments"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    char * own = (char *) ex->base[ex->rank];
    unsigned * ctrl = (unsigned *) (own + ex->ctrl_off);
    for (int i = 0; i < n; ++i) {
        if (!w[i] || !dst[i]) { b200_set_error("mul_mat_rowsplit: null tensor"); return B200_ERR_INVALID; }
        if (x->ne[0] != w[i]->ne[0] || row_lo[i] < 0 || row_lo[i] + w[i]->ne[1] > dst[i]->ne[0] || dst[i]->ne[1] != x->ne[1] || dst[i]->type != B200_TYPE_F32 || dst[i]->nb[0] != 4) {
            b200_set_error("mul_mat_rowsplit: shape mismatch"); return B200_ERR_INVALID;
        }
        if (x->ne[1] != 1 || dst[i]->ne[1] != 1) { b200_set_error("mul_mat_rowsplit: one activation column (decode) only"); return B200_ERR_UNSUPPORTED; }
        const size_t doff = (size_t) ((char *) dst[i]->data - own);
        if ((char *) dst[i]->data < own || doff + (size_t) dst[i]->ne[0] * 4 > ex->ctrl_off) { b200_set_error("mul_mat_rowsplit: dst is not inside the exchange arena"); return B200_ERR_INVALID; }
        if (w[i]->ne[1] > 0 && !(is_quant(w[i]->type) && q8_1_variant_env() == B200_Q8_1_X86 && b200_gemv_stream_supported(w[i], x))) {
            b200_set_error("mul_mat_rowsplit: only the streaming GEMV is fused with the exchange (weight type %d, K=%lld)", w[i]->type, (long long) w[i]->ne[0]);
            return B200_ERR_UNSUPPORTED;
        }
    }
    b200_rowsplit_peers pe; pe.nll = 0; pe.arena = own; pe.ctrl = ctrl;
    for (int r = 0; r < ex->world; ++r) pe.ll_base[pe.nll++] = (char *) ex->base[r] + ex->ctrl_off + 256;   // mirror: 8 bytes per f32 of the arena
    pe.ll_self = pe.ll_base[ex->rank];
    int rc = B200_OK;
    const void * image = nullptr;
    if (w[0]->ne[0] > fused_quant_max_k()) { image = quantized_activations(ctx, x, B200_Q8_1_X86, &rc); if (!image) return rc; }
    for (int i = 0; i < n && rc == B200_OK; ) {
        int j = i + 1;
        while (j < n && w[j]->type == w[i]->type && (w[j]->flags & B200_TENSOR_REPACKED) == (w[i]->flags & B200_TENSOR_REPACKED)) ++j;
        int64_t rows = 0;
        float * d[3];
        for (int t = i; t < j; ++t) {
            rows += w[t]->ne[1];
            d[t - i] = (float *) dst[t]->data + row_lo[t];
            pe.full[t - i] = (float *) dst[t]->data; pe.full_m[t - i] = (int) dst[t]->ne[0];
        }
        if (rows > 0) rc = b200_launch_gemv_stream(ctx, w + i, j - i, x, image, d, &pe);
        else for (int t = i; t < j && rc == B200_OK; ++t)                  // no local rows: receive only
            rc = b200_launch_exchange_unpack(ctx, (const char *) pe.ll_self + 2 * (size_t) ((char *) dst[t]->data - own), (float *) dst[t]->data, dst[t]->ne[0], ctrl);
        i = j;
    }
    ctx->qcache.valid = 0; ctx->mcache.valid = 0;
    for (int i = 0; i < n; ++i) b200_note_write(ctx, dst[i]->data, b200_tensor_span(dst[i]));
    return rc;
}
extern "C" int b200_mul_mat_rowsplit(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst, int64_t row_lo, const b200_exchange * ex) {
    return b200_mul_mat_rowsplit_multi(ctx, &w, 1, x, &dst, &row_lo, ex);
}
extern "C" int b200_exchange_error(b200_ctx * ctx, const b200_exchange * ex) {
    if (!ctx || !exchange_ok(ex)) return -1;
    unsigned e = 0;
    if (cudaMemcpy(&e, (unsigned *) ((char *) ex->base[ex->rank] + ex->ctrl_off) + 10, 4, cudaMemcpyDeviceToHost) != cudaSuccess) return -1;
    return (int) e;
}

extern "C" int b200_mul_mat_id(b200_ctx * ctx, const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, b200_tensor * dst) {
    if (!ctx || !as || !b || !ids || !dst) { b200_set_error("mul_mat_id: null argument"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int64_t K = as->ne[0], M = as->ne[1], n_used = ids->ne[0], n_tokens = ids->ne[1];
    if (as->ne[3] != 1 || b->ne[3] != 1 || dst->ne[3] != 1 || b->ne[0] != K || b->ne[2] != n_tokens || dst->ne[0] != M ||
        dst->ne[1] != n_used || dst->ne[2] != n_tokens || b->ne[1] <= 0 || n_used % b->ne[1] != 0 || ids->type != B200_TYPE_I32) {
        b200_set_error("mul_mat_id: shape mismatch"); return B200_ERR_INVALID;
    }
    if (b->type != B200_TYPE_F32 || dst->type != B200_TYPE_F32 || b->nb[0] != 4 || dst->nb[0] != 4 || !is_quant(as->type) ||
        as->nb[0] != b200_type_size(as->type) || K % b200_blck_size(as->type) != 0) {
        b200_set_error("mul_mat_id: unsupported type/layout"); return B200_ERR_UNSUPPORTED;
    }
    if (M == 0 || n_used * n_tokens == 0) return B200_OK;
    if (n_tokens > B200_GEMV_MAX_COLS && b200_gemm_id_supported(as, b, ids)) {
        const int rc = b200_launch_gemm_id_q(ctx, as, b, ids, dst);        // grouped tcgen05 GEMM (prefill)
        b200_note_write(ctx, dst->data, b200_tensor_span(dst));
        return rc;
    }
    int rc;
    // decode (one token, <= 3 experts used): the experts are segments of the streaming GEMV (TMA ring, fused q8_1 quantization), which
    // picks each segment's expert from the device-side ids; shared activations (gate/up) = one launch, per-slot activations (down) = one
    // launch per slot
    static const int no_stream_id = getenv("GGML_B200_NO_GEMV_STREAM_ID") != nullptr;
    if (!no_stream_id && n_tokens == 1 && n_used <= 3 && q8_1_variant_env() == B200_Q8_1_X86 && ids->nb[0] == 4 && dst->nb[1] % 4 == 0 && b->nb[1] % 16 == 0) {
        b200_tensor w2 = *as; w2.ne[2] = 1; w2.ne[3] = 1;                   // expert 0 as a plain [K, M] matrix
        b200_tensor x2 = *b;  x2.ne[1] = 1; x2.ne[2] = 1; x2.ne[3] = 1;
        if (b200_gemv_stream_supported(&w2, &x2)) {
            const b200_tensor * ws[3] = { &w2, &w2, &w2 };
            float * d[3];
            for (int64_t sl = 0; sl < n_used; ++sl) d[sl] = (float *) ((char *) dst->data + sl * dst->nb[1]);
            b200_gemv_ids idm; idm.ids = (const int *) ids->data; idm.stride = as->nb[2]; idm.ads = nullptr;
            rc = B200_OK;
            if (b->ne[1] == 1) {
                const void * image = nullptr;
                if (K > fused_quant_max_k()) { image = quantized_activations(ctx, &x2, B200_Q8_1_X86, &rc); if (!image) return rc; }
                idm.base = 0;
                rc = b200_launch_gemv_stream(ctx, ws, (int) n_used, &x2, image, d, nullptr, &idm);
            } else {
                const char * image = nullptr;                               // all slots' columns quantized by one launch
                if (K > fused_quant_max_k()) { image = (const char *) quantized_activations(ctx, b, B200_Q8_1_X86, &rc); if (!image) return rc; }
                for (int64_t sl = 0; sl < n_used && rc == B200_OK; ++sl) {
                    x2.data = (char *) b->data + sl * b->nb[1];
                    idm.base = (int) sl;
                    idm.ads = image ? image + (size_t) n_used * K + (size_t) sl * (K / 32) * 4 : nullptr;
                    rc = b200_launch_gemv_stream(ctx, ws, 1, &x2, image ? image + (size_t) sl * K : nullptr, d + sl, nullptr, &idm);
                }
            }
            b200_note_write(ctx, dst->data, b200_tensor_span(dst));
            return rc;
        }
    }
    const void * image = quantized_activations(ctx, b, q8_1_variant_env(), &rc);
    if (!image) return rc;
    rc = b200_launch_gemv_id_q(ctx, as, image, b->ne[1], n_tokens, ids, dst);
    b200_note_write(ctx, dst->data, b200_tensor_span(dst));
    return rc;
}

extern "C" int b200_flash_attn_ext(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                                   const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                                   float scale, float max_bias, float logit_softcap) {
    if (!ctx || !q || !k || !v || !dst) { b200_set_error("flash_attn_ext: null argument"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int64_t DK = k->ne[0], DV = v->ne[0];
    if (q->ne[0] != DK || v->ne[1] != k->ne[1] || dst->ne[0] != DV || dst->ne[1] != q->ne[2] || dst->ne[2] != q->ne[1] || dst->ne[3] != q->ne[3] ||
        k->ne[2] <= 0 || q->ne[2] % k->ne[2] != 0 || q->ne[2] % v->ne[2] != 0 || q->ne[3] % k->ne[3] != 0 || q->ne[3] % v->ne[3] != 0) {
        b200_set_error("flash_attn_ext: shape mismatch"); return B200_ERR_INVALID;
    }
    if (k->type != B200_TYPE_F16 || v->type != B200_TYPE_F16) {              // quantized KV cache (b200_fattn_q.cu)
        if (q->type != B200_TYPE_F32 || dst->type != B200_TYPE_F32 || (mask && mask->type != B200_TYPE_F16) || q->nb[0] != 4 || dst->nb[0] != 4 ||
            !b200_fattn_kvq_supported(q, k, v)) {
            b200_set_error("flash_attn_ext: K/V types outside f16 and q8_0 / q4_0 at head size 64 / 128 (DK=%lld DV=%lld)", (long long) DK, (long long) DV);
            return B200_ERR_UNSUPPORTED;
        }
        if (mask && (mask->ne[0] < k->ne[1] || mask->ne[1] < q->ne[1])) { b200_set_error("flash_attn_ext: mask too small"); return B200_ERR_INVALID; }
        const int rcq = b200_launch_fattn_kvq(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap);
        b200_note_write(ctx, dst->data, b200_tensor_span(dst));
        return rcq;
    }
    if (q->type != B200_TYPE_F32 || k->type != B200_TYPE_F16 || v->type != B200_TYPE_F16 || dst->type != B200_TYPE_F32 ||
        (mask && mask->type != B200_TYPE_F16) || q->nb[0] != 4 || k->nb[0] != 2 || v->nb[0] != 2 || dst->nb[0] != 4 ||
        DK % 64 != 0 || DV % 64 != 0 || DK > 256 || (DV != 64 && DV != 128 && DV != 256) ||
        (k->nb[1] % 16) || (k->nb[2] % 16) || (k->nb[3] % 16) || ((uintptr_t) k->data % 16) ||
        (v->nb[1] % 4) || (v->nb[2] % 4) || (v->nb[3] % 4) || ((uintptr_t) v->data % 4)) {
        b200_set_error("flash_attn_ext: type/head-size/alignment outside the hot path (DK=%lld DV=%lld)", (long long) DK, (long long) DV);
        return B200_ERR_UNSUPPORTED;
    }
    if (mask && (mask->ne[0] < k->ne[1] || mask->ne[1] < q->ne[1])) { b200_set_error("flash_attn_ext: mask too small"); return B200_ERR_INVALID; }
    static const int no_dec = getenv("GGML_B200_FATTN_VEC") != nullptr;
    int rc = B200_ERR_UNSUPPORTED;
    if (!no_dec && b200_fattn_gqa_supported(q, k, v, mask, dst)) rc = b200_launch_fattn_gqa(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap);
    if (rc != B200_ERR_UNSUPPORTED) {}
    else if (!no_dec && b200_fattn_dec_supported(q, k, v)) rc = b200_launch_fattn_dec(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap);
    else {
        if (!no_dec && b200_fattn_tc_supported(q, k, v, mask, dst)) rc = b200_launch_fattn_tc(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap);
        if (rc == B200_ERR_UNSUPPORTED) rc = b200_launch_fattn_vec(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap);
    }
    b200_note_write(ctx, dst->data, b200_tensor_span(dst));
    return rc;
}

extern "C" int b200_flash_attn_ext_paged(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_kv_pages * pages,
                                         const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst, float scale, float max_bias, float logit_softcap) {
    if (!ctx || !q || !k || !v || !dst || !pages || !pages->table) { b200_set_error("flash_attn_ext_paged: null argument"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    const int64_t DK = k->ne[0], DV = v->ne[0], KV = k->ne[1];
    const int pt = pages->page_tokens;
    if (pt < 16 || (pt & (pt - 1)) || pages->n_pages < 0 || (int64_t) pages->n_pages * pt < KV || pages->pool_tokens < pt) {
        b200_set_error("flash_attn_ext_paged: page_tokens %d (power of two >= 16), n_pages %d, pool_tokens %lld do not cover %lld keys", pt, pages->n_pages,
                       (long long) pages->pool_tokens, (long long) KV);
        return B200_ERR_INVALID;
    }
    if (q->ne[0] != DK || v->ne[1] != KV || dst->ne[0] != DV || dst->ne[1] != q->ne[2] || dst->ne[2] != q->ne[1] || dst->ne[3] != q->ne[3] ||
        k->ne[2] <= 0 || q->ne[2] % k->ne[2] != 0 || q->ne[2] % v->ne[2] != 0 || k->ne[3] != 1 || v->ne[3] != 1) {
        b200_set_error("flash_attn_ext_paged: shape mismatch"); return B200_ERR_INVALID;
    }
    if (q->type != B200_TYPE_F32 || k->type != B200_TYPE_F16 || v->type != B200_TYPE_F16 || dst->type != B200_TYPE_F32 || (mask && mask->type != B200_TYPE_F16) ||
        q->nb[0] != 4 || k->nb[0] != 2 || v->nb[0] != 2 || dst->nb[0] != 4) {
        b200_set_error("flash_attn_ext_paged: types outside the hot path"); return B200_ERR_UNSUPPORTED;
    }
    if (mask && (mask->ne[0] < KV || mask->ne[1] < q->ne[1])) { b200_set_error("flash_attn_ext_paged: mask too small"); return B200_ERR_INVALID; }
    int rc = B200_ERR_UNSUPPORTED;
    if (b200_fattn_gqa_supported(q, k, v, mask, dst)) rc = b200_launch_fattn_gqa(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap, pages);
    if (rc == B200_ERR_UNSUPPORTED && b200_fattn_dec_supported(q, k, v)) rc = b200_launch_fattn_dec(ctx, q, k, v, mask, sinks, dst, scale, max_bias, logit_softcap, pages);
    if (rc == B200_ERR_UNSUPPORTED) b200_set_error("flash_attn_ext_paged: only query batches of <= 8 tokens with f16 K/V of head size 64 / 128 / 256");
    b200_note_write(ctx, dst->data, b200_tensor_span(dst));
    return rc;
}

// ---- row split helper: ggml-cuda.cu:777-803 semantics -----------------------------------------
extern "C" int b200_split_row_range(int64_t nrows, const float * tensor_split, int world, int rank, int64_t rounding,
                                    int64_t * row_low, int64_t * row_high) {
    if (world <= 0 || rank < 0 || rank >= world || rounding <= 0 || !row_low || !row_high) { b200_set_error("split_row_range: bad arguments"); return B200_ERR_INVALID; }
    // tensor_split[i] = cumulative fraction at which device i starts (as the reference normalises it); NULL = even
    auto start_frac = [&](int r) -> double {
        if (r >= world) return 1.0;
        if (!tensor_split) return (double) r / world;
        return (double) tensor_split[r];
    };
    int64_t lo = rank == 0 ? 0 : (int64_t) ((float) nrows * (float) start_frac(rank));   // float product, as the reference
    lo -= lo % rounding;
    int64_t hi;
    if (rank == world - 1) hi = nrows;
    else { hi = (int64_t) ((float) nrows * (float) start_frac(rank + 1)); hi -= hi % rounding; }
    *row_low = lo; *row_high = hi;
    return B200_OK;
}
