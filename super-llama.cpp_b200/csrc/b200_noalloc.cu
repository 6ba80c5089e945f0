// b200_noalloc.cu -- b200_mul_mat_noalloc: a quantized MUL_MAT for callers that may not allocate (code under CUDA graph capture,
// tight serving loops, harnesses that forbid cudaMalloc).  No counterpart in the reference.
//
// b200_mul_mat's general small-batch path (quantize_q8_1 image -> k_gemv, b200_gemv.cu) keeps the q8_1 image of the activations in a
// scratch block that the context cudaMallocs on demand (b200_scratch slot 0).  Here the same two kernels run with the image in a
// file-scope device array instead: B200_NOALLOC_SCRATCH_BYTES per GPU that loads the library, held until the process exits.
//   * the array's NAME is a host-side stand-in in host code, not a device address: the address comes from cudaGetSymbolAddress, once per
//     device with that device current (noalloc_scratch()), and only that pointer is handed to the kernels;
//   * nothing has to start at zero: the quantizer overwrites every byte of the image the GEMV then reads, so the previous call's
//     contents are dead by construction;
//   * every call on a device shares the array: calls that overlap on the GPU (two contexts / streams on one device) race.  Calls on one
//     context are ordered by its stream.
// Calls no cudaMalloc / cudaFree / event or stream creation; rejects anything that would not fit the array.
#include "b200_common.cuh"
#include <mutex>

static __device__ __align__(256) unsigned char g_noalloc_scratch[B200_NOALLOC_SCRATCH_BYTES];

#define B200_NOALLOC_MAX_DEVICES 64
static void *     g_scratch_addr[B200_NOALLOC_MAX_DEVICES];
static std::mutex g_scratch_mu;

// device address of g_noalloc_scratch on `device` (which must be current); resolved once per device
static void * noalloc_scratch(int device) {
    if (device < 0 || device >= B200_NOALLOC_MAX_DEVICES) { b200_set_error("mul_mat_noalloc: device %d out of range", device); return nullptr; }
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (!g_scratch_addr[device]) {
        void * p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) { cudaGetLastError(); b200_set_error("mul_mat_noalloc: cudaGetSymbolAddress -> %s", cudaGetErrorString(e)); return nullptr; }
        g_scratch_addr[device] = p;
    }
    return g_scratch_addr[device];
}

extern "C" size_t b200_noalloc_scratch_bytes(void) { return B200_NOALLOC_SCRATCH_BYTES; }

extern "C" int b200_mul_mat_noalloc(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst) {
    if (!ctx || !w || !x || !dst) { b200_set_error("mul_mat_noalloc: null argument"); return B200_ERR_INVALID; }
    const int64_t K = w->ne[0], M = w->ne[1];
    if (x->ne[0] != K || dst->ne[0] != M || dst->ne[1] != x->ne[1] || dst->ne[2] != x->ne[2] || dst->ne[3] != x->ne[3] ||
        w->ne[2] <= 0 || w->ne[3] <= 0 || x->ne[2] % w->ne[2] != 0 || x->ne[3] % w->ne[3] != 0) {
        b200_set_error("mul_mat_noalloc: shape mismatch"); return B200_ERR_INVALID;
    }
    const int t = w->type;
    if (!(t == B200_TYPE_Q4_0 || t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_K || t == B200_TYPE_Q5_K || t == B200_TYPE_Q6_K) ||
        x->type != B200_TYPE_F32 || dst->type != B200_TYPE_F32 || x->nb[0] != 4 || dst->nb[0] != 4 || w->nb[0] != b200_type_size(t)) {
        b200_set_error("mul_mat_noalloc: quantized weights with contiguous rows, f32 activations and result only"); return B200_ERR_UNSUPPORTED;
    }
    if (K <= 0 || K % b200_blck_size(t) != 0) { b200_set_error("mul_mat_noalloc: K=%lld not a multiple of the block size", (long long) K); return B200_ERR_INVALID; }
    const int64_t ncol = x->ne[1] * x->ne[2] * x->ne[3];
    if (M == 0 || ncol == 0) return B200_OK;
    if (x->ne[1] > B200_GEMV_MAX_COLS) {
        b200_set_error("mul_mat_noalloc: %lld activation columns, at most %d (small-batch GEMV only)", (long long) x->ne[1], B200_GEMV_MAX_COLS); return B200_ERR_UNSUPPORTED;
    }
    const size_t need = b200_q8_1_soa_size(K, ncol);
    if (need > (size_t) B200_NOALLOC_SCRATCH_BYTES) {
        b200_set_error("mul_mat_noalloc: the q8_1 image of %lld columns x K=%lld needs %zu bytes, the static scratch holds %zu (B200_NOALLOC_SCRATCH_BYTES)",
                       (long long) ncol, (long long) K, need, (size_t) B200_NOALLOC_SCRATCH_BYTES);
        return B200_ERR_INVALID;
    }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    void * image = noalloc_scratch(ctx->device);
    if (!image) return B200_ERR_CUDA;
    const int prof = ctx->profile;                 // the per-launch event profile creates CUDA events: off for this call
    ctx->profile = 0;
    int rc = b200_launch_quantize_q8_1_soa(ctx, (const float *) x->data, x->nb[1], x->nb[2], x->nb[3], K, x->ne[1], x->ne[2], x->ne[3], image, B200_Q8_1_X86);
    if (rc == B200_OK) rc = b200_launch_gemv_q(ctx, w, image, x->ne[1], x->ne[2], x->ne[3], dst);
    ctx->profile = prof;
    b200_note_write(ctx, dst->data, b200_tensor_span(dst));
    return rc;
}
