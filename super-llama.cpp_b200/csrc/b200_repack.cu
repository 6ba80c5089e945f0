// b200_repack.cu -- load-time weight repacking ("weight repacking at GGUF load time into TMA-friendly tiles").
//
// q4_K / q5_K blocks (144 / 176 B) already have every field 16-byte aligned and are streamed raw.  q6_K (210 B), q8_0
// (34 B) and q4_0 (18 B) blocks are only 2-byte aligned, which rules out 128-bit shared loads and bulk-copy alignment.
// For those, every aligned group of 8 consecutive blocks of a row is permuted IN PLACE into a field-major, lane-major
// layout (see b200_gemv_stream.cu for the q6_K group map).  The permutation stays inside the group, so row strides,
// tensor sizes, row views and expert slices are unchanged; un-repacking (get_tensor) is the exact inverse.
// The format on disk / in host memory is always the reference's (ggml/src/ggml-common.h:170-344).
#include "b200_common.cuh"

// raw byte offset inside a group of 8 q6_K blocks (0..1679) -> offset in the repacked group
static __device__ __forceinline__ int q6k_fwd(int s) {
    const int b = s / 210, o = s - b * 210;
    if (o < 128) { const int n = o >> 6, ab = (o >> 5) & 1, j = (o >> 4) & 1; return ((((n * 2 + j) * 2 + ab) * 8 + b) << 4) + (o & 15); }
    if (o < 192) { const int i = o - 128, n = i >> 5, j = (i >> 4) & 1;       return 1024 + (((n * 2 + j) * 8 + b) << 4) + (i & 15); }
    if (o < 208) return 1536 + b * 16 + (o - 192);
    return 1664 + b * 2 + (o - 208);
}

// q8_0 / q4_0: every aligned run of 8 blocks (256 weights: 272 / 144 B, both multiples of 16) becomes [8 x fp16 d][8 x qs],
// so the quants sit at 16-byte aligned addresses and a run is one bulk-copy / TMA-box unit
static __device__ __forceinline__ int q80_fwd(int s) { const int b = s / 34, o = s - b * 34; return o < 2 ? b * 2 + o : 16 + b * 32 + (o - 2); }
static __device__ __forceinline__ int q40_fwd(int s) { const int b = s / 18, o = s - b * 18; return o < 2 ? b * 2 + o : 16 + b * 16 + (o - 2); }

// one thread block per row chunk of 8 runs; TYPE selects the byte map
template <int TYPE, bool INVERSE>
__global__ void __launch_bounds__(256) k_repack_run8(uint8_t * data, size_t row_stride, int runs_per_row, int64_t nruns) {
    constexpr int RUN = TYPE == B200_TYPE_Q8_0 ? 272 : 144;
    __shared__ uint8_t buf[RUN];
    for (int64_t g = blockIdx.x; g < nruns; g += gridDim.x) {
        uint8_t * grp = data + (size_t) (g / runs_per_row) * row_stride + (size_t) (g % runs_per_row) * RUN;
        for (int i = threadIdx.x; i < RUN / 4; i += blockDim.x) ((uint32_t *) buf)[i] = ((const uint32_t *) grp)[i];
        __syncthreads();
        for (int s = threadIdx.x; s < RUN; s += blockDim.x) {
            const int d = TYPE == B200_TYPE_Q8_0 ? q80_fwd(s) : q40_fwd(s);
            if (INVERSE) grp[s] = buf[d];
            else         grp[d] = buf[s];
        }
        __syncthreads();
    }
}

template <bool INVERSE>
__global__ void __launch_bounds__(256) k_repack_q6k(uint8_t * data, size_t row_stride, int groups_per_row, int64_t ngroups) {
    __shared__ uint8_t buf[1680];
    for (int64_t g = blockIdx.x; g < ngroups; g += gridDim.x) {
        uint8_t * grp = data + (size_t) (g / groups_per_row) * row_stride + (size_t) (g % groups_per_row) * 1680;
        for (int i = threadIdx.x; i < 1680 / 4; i += blockDim.x) ((uint32_t *) buf)[i] = ((const uint32_t *) grp)[i];
        __syncthreads();
        for (int s = threadIdx.x; s < 1680; s += blockDim.x) {
            if (INVERSE) grp[s] = buf[q6k_fwd(s)];
            else         grp[q6k_fwd(s)] = buf[s];
        }
        __syncthreads();
    }
}

extern "C" int b200_repack_supported(int type, int64_t k) {
    if (type == B200_TYPE_Q6_K) return k > 0 && k % 2048 == 0;
    if (type == B200_TYPE_Q8_0 || type == B200_TYPE_Q4_0) return k > 0 && k % 256 == 0;
    return 0;
}

static int repack_impl(b200_ctx * ctx, int type, void * data, int64_t k, int64_t nrows, size_t row_stride, bool inverse) {
    if (!b200_repack_supported(type, k)) { b200_set_error("repack: type %d with k=%lld has no repacked layout", type, (long long) k); return B200_ERR_UNSUPPORTED; }
    if (((uintptr_t) data & 3) || (row_stride & 3)) { b200_set_error("repack: rows must be 4-byte aligned"); return B200_ERR_INVALID; }
    if (nrows <= 0) return B200_OK;
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    if (type != B200_TYPE_Q6_K) {
        const int rpr = (int) (k / 256);
        const int64_t nruns = nrows * rpr;
        const unsigned grid = (unsigned) (nruns < 148 * 32 ? nruns : 148 * 32);
        uint8_t * d = (uint8_t *) data;
        if (type == B200_TYPE_Q8_0) {
            if (inverse) B200_LAUNCH(ctx, B200_K_CONVERT, k_repack_run8<B200_TYPE_Q8_0, true><<<grid, 256, 0, ctx->stream>>>(d, row_stride, rpr, nruns));
            else         B200_LAUNCH(ctx, B200_K_CONVERT, k_repack_run8<B200_TYPE_Q8_0, false><<<grid, 256, 0, ctx->stream>>>(d, row_stride, rpr, nruns));
        } else {
            if (inverse) B200_LAUNCH(ctx, B200_K_CONVERT, k_repack_run8<B200_TYPE_Q4_0, true><<<grid, 256, 0, ctx->stream>>>(d, row_stride, rpr, nruns));
            else         B200_LAUNCH(ctx, B200_K_CONVERT, k_repack_run8<B200_TYPE_Q4_0, false><<<grid, 256, 0, ctx->stream>>>(d, row_stride, rpr, nruns));
        }
        return B200_OK;
    }
    const int gpr = (int) (k / 2048);
    const int64_t ngroups = nrows * gpr;
    const unsigned grid = (unsigned) (ngroups < 148 * 16 ? ngroups : 148 * 16);
    if (inverse) B200_LAUNCH(ctx, B200_K_CONVERT, k_repack_q6k<true><<<grid, 256, 0, ctx->stream>>>((uint8_t *) data, row_stride, gpr, ngroups));
    else         B200_LAUNCH(ctx, B200_K_CONVERT, k_repack_q6k<false><<<grid, 256, 0, ctx->stream>>>((uint8_t *) data, row_stride, gpr, ngroups));
    return B200_OK;
}
extern "C" int b200_repack(b200_ctx * ctx, int type, void * data, int64_t k, int64_t nrows, size_t row_stride) {
    return repack_impl(ctx, type, data, k, nrows, row_stride, false);
}
extern "C" int b200_unrepack(b200_ctx * ctx, int type, void * data, int64_t k, int64_t nrows, size_t row_stride) {
    return repack_impl(ctx, type, data, k, nrows, row_stride, true);
}
