// b200_matmul_float.cu -- MUL_MAT with f32 / f16 / bf16 weights (not a north-star type; kept so that
// graphs with a float matmul, e.g. -fa 0 attention or an unquantized output head, do not bounce
// to the CPU).  Plain warp-per-row dot products, <= 8 activation columns per pass.
// Reference semantics: ggml_compute_forward_mul_mat, ggml/src/ggml-cpu/ggml-cpu.c:1226-1418.
#include "b200_common.cuh"

struct mmf_args {
    const char * w; size_t w_nb1, w_nb2, w_nb3;
    const char * x; size_t x_nb1, x_nb2, x_nb3;
    float * dst; size_t d_nb1, d_nb2, d_nb3;
    int64_t K, M, ne11, ne12, ne13, r2, r3, col_base;
};

template <typename T> static __device__ __forceinline__ float to_f32(T v);
template <> __device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <> __device__ __forceinline__ float to_f32<__half>(__half v) { return __half2float(v); }
template <> __device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }

template <typename T, int NC>
__global__ void __launch_bounds__(128) k_matmul_float(const mmf_args p) {
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row = (int64_t) blockIdx.x * 4 + warp;
    if (row >= p.M) return;
    const int64_t i12 = blockIdx.y % p.ne12, i13 = blockIdx.y / p.ne12;
    const T * wr = (const T *) (p.w + row * p.w_nb1 + (i12 / p.r2) * p.w_nb2 + (i13 / p.r3) * p.w_nb3);
    const char * xb = p.x + i12 * p.x_nb2 + i13 * p.x_nb3;
    float acc[NC];
#pragma unroll
    for (int n = 0; n < NC; ++n) acc[n] = 0.0f;
    for (int64_t i = lane; i < p.K; i += 32) {
        const float wv = to_f32<T>(wr[i]);
#pragma unroll
        for (int n = 0; n < NC; ++n) acc[n] += wv * ((const float *) (xb + (p.col_base + n) * p.x_nb1))[i];
    }
    float * drow = (float *) ((char *) p.dst + i12 * p.d_nb2 + i13 * p.d_nb3) + row;
#pragma unroll
    for (int n = 0; n < NC; ++n) {
        const float s = warp_sum(acc[n]);
        if (lane == 0) *(float *) ((char *) drow + (p.col_base + n) * p.d_nb1) = s;
    }
}

template <typename T>
static int launch_mmf(b200_ctx * ctx, mmf_args p) {
    dim3 grid((unsigned) ((p.M + 3) / 4), (unsigned) (p.ne12 * p.ne13));
    for (int64_t c0 = 0; c0 < p.ne11; c0 += 8) {
        const int nc = (int) (p.ne11 - c0 < 8 ? p.ne11 - c0 : 8);
        p.col_base = c0;
        switch (nc) {
            case 1: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 1><<<grid, 128, 0, ctx->stream>>>(p)); break;
            case 2: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 2><<<grid, 128, 0, ctx->stream>>>(p)); break;
            case 3: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 3><<<grid, 128, 0, ctx->stream>>>(p)); break;
            case 4: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 4><<<grid, 128, 0, ctx->stream>>>(p)); break;
            case 5: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 5><<<grid, 128, 0, ctx->stream>>>(p)); break;
            case 6: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 6><<<grid, 128, 0, ctx->stream>>>(p)); break;
            case 7: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 7><<<grid, 128, 0, ctx->stream>>>(p)); break;
            default: B200_LAUNCH(ctx, B200_K_MATMUL_FLOAT, k_matmul_float<T, 8><<<grid, 128, 0, ctx->stream>>>(p)); break;
        }
    }
    return B200_OK;
}

int b200_launch_matmul_float(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst) {
    if (w->nb[0] != b200_type_size(w->type)) { b200_set_error("matmul_float: transposed weights unsupported"); return B200_ERR_UNSUPPORTED; }
    mmf_args p;
    p.w = (const char *) w->data; p.w_nb1 = w->nb[1]; p.w_nb2 = w->nb[2]; p.w_nb3 = w->nb[3];
    p.x = (const char *) x->data; p.x_nb1 = x->nb[1]; p.x_nb2 = x->nb[2]; p.x_nb3 = x->nb[3];
    p.dst = (float *) dst->data; p.d_nb1 = dst->nb[1]; p.d_nb2 = dst->nb[2]; p.d_nb3 = dst->nb[3];
    p.K = w->ne[0]; p.M = w->ne[1]; p.ne11 = x->ne[1]; p.ne12 = x->ne[2]; p.ne13 = x->ne[3];
    p.r2 = x->ne[2] / w->ne[2]; p.r3 = x->ne[3] / w->ne[3]; p.col_base = 0;
    if (p.ne12 * p.ne13 > 65535) { b200_set_error("matmul_float: batch too large"); return B200_ERR_UNSUPPORTED; }
    switch (w->type) {
        case B200_TYPE_F32:  return launch_mmf<float>(ctx, p);
        case B200_TYPE_F16:  return launch_mmf<__half>(ctx, p);
        case B200_TYPE_BF16: return launch_mmf<__nv_bfloat16>(ctx, p);
    }
    return B200_ERR_UNSUPPORTED;
}
