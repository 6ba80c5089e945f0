// b200_quantize.cu -- f32 activations -> block_q8_1, bit-exact with the reference's ggml-cpu symbol.
//
// Reference: quantize_row_q8_1, x86 AVX2 body ggml/src/ggml-cpu/arch/x86/quants.c:388-449 (variant X86,
// what libggml-cpu.so executes on an AVX2 box) and the scalar quantize_row_q8_1_ref
// ggml/src/ggml-quants.c:225-258 (variant REF).  They differ in the reciprocal (127/amax vs
// 1/(amax/127)) and in the rounding of exact .5 ties (half-to-even vs half-away-from-zero).
// Note the reference's CUDA kernel (ggml-cuda/quantize.cu:31-47) matches neither; we follow the CPU.
//
// One warp owns one 32-element block: lane i holds x[i]; amax via shuffles; every arithmetic step
// uses an explicitly rounded intrinsic so nvcc cannot contract or reassociate it.
// Bytes: reads 128 B, writes 36 B per block -> HBM/L2 bound, negligible next to the GEMV it feeds.
#include "b200_common.cuh"

template <int VARIANT>
__global__ void __launch_bounds__(256) k_quantize_q8_1(const float * __restrict__ x, size_t x_row_stride,
                                                      blk_q8_1 * __restrict__ y, int64_t nblk_row, int64_t nblk_total) {
    const int     lane = threadIdx.x & 31;
    const int64_t ib   = (int64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);   // global block index
    if (ib >= nblk_total) return;                                                          // warp-uniform
    const int64_t row = ib / nblk_row;
    const int64_t bi  = ib - row * nblk_row;

    const float * xr = (const float *) ((const char *) x + row * x_row_stride);
    const float   v  = xr[bi * 32 + lane];

    int q, isum; float d;
    q8_1_block<VARIANT>(v, q, d, isum);

    // pack 4 int8 per lane%4==0 lane -> 8 aligned 32-bit stores per block
    unsigned t = (unsigned) q & 0xffu;
    t |= __shfl_down_sync(0xffffffffu, t, 1) << 8;
    t |= __shfl_down_sync(0xffffffffu, t, 2) << 16;

    blk_q8_1 * yb = y + ib;
    if ((lane & 3) == 0) *(unsigned *) (yb->qs + lane) = t;
    if (lane == 1) {
        *(__half2 *) &yb->d = q8_1_ds(d, isum);
    }
}

int b200_launch_quantize_q8_1(b200_ctx * ctx, const float * x, size_t x_row_stride, void * y,
                              int64_t k, int64_t nrows, int variant) {
    if (k % 32 != 0 || k <= 0 || nrows < 0) { b200_set_error("quantize_q8_1: k=%lld not a multiple of 32", (long long) k); return B200_ERR_INVALID; }
    if (nrows == 0) return B200_OK;
    const int64_t nblk_row = k / 32, total = nblk_row * nrows;
    const unsigned grid = (unsigned) ((total + 7) / 8);
    if (variant == B200_Q8_1_REF)
        B200_LAUNCH(ctx, B200_K_QUANTIZE, k_quantize_q8_1<B200_Q8_1_REF><<<grid, 256, 0, ctx->stream>>>(x, x_row_stride, (blk_q8_1 *) y, nblk_row, total));
    else
        B200_LAUNCH(ctx, B200_K_QUANTIZE, k_quantize_q8_1<B200_Q8_1_X86><<<grid, 256, 0, ctx->stream>>>(x, x_row_stride, (blk_q8_1 *) y, nblk_row, total));
    return B200_OK;
}

// ---- internal SoA activation image -----------------------------------------------------------
// Same arithmetic, different placement: the GEMV wants 16-byte aligned quant streams, which the
// 36-byte AoS block_q8_1 cannot give.  For `ncol_total` activation columns of length k the image is
//     qs : int8  [ncol_total][k]        (16-byte aligned per column because k % 32 == 0)
//     ds : half2 [ncol_total][k/32]     (d, s = d*sum q)  -- starts at byte offset ncol_total*k
// Column index = (i3*ne2 + i2)*ne1 + i1 for a strided f32 source x[k, ne1, ne2, ne3].
template <int VARIANT>
__global__ void __launch_bounds__(256) k_quantize_q8_1_soa(const float * __restrict__ x, size_t nb1, size_t nb2, size_t nb3,
                                                          int64_t ne1, int64_t ne2,
                                                          int8_t * __restrict__ qs, __half2 * __restrict__ ds,
                                                          int64_t nblk_col, int64_t nblk_total) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");    // lets a PDL-launched consumer prefetch its weights meanwhile
    const int     lane = threadIdx.x & 31;
    const int64_t ib   = (int64_t) blockIdx.x * (blockDim.x >> 5) + (threadIdx.x >> 5);
    if (ib >= nblk_total) return;
    const int64_t col = ib / nblk_col;
    const int64_t bi  = ib - col * nblk_col;
    const int64_t i1 = col % ne1, i2 = (col / ne1) % ne2, i3 = col / (ne1 * ne2);
    const float * xr = (const float *) ((const char *) x + i1 * nb1 + i2 * nb2 + i3 * nb3);
    const float   v  = xr[bi * 32 + lane];
    int q, isum; float d;
    q8_1_block<VARIANT>(v, q, d, isum);
    unsigned t = (unsigned) q & 0xffu;
    t |= __shfl_down_sync(0xffffffffu, t, 1) << 8;
    t |= __shfl_down_sync(0xffffffffu, t, 2) << 16;
    if ((lane & 3) == 0) *(unsigned *) (qs + ib * 32 + lane) = t;
    if (lane == 1) ds[ib] = q8_1_ds(d, isum);
}

int b200_launch_quantize_q8_1_soa(b200_ctx * ctx, const float * x, size_t nb1, size_t nb2, size_t nb3,
                                  int64_t k, int64_t ne1, int64_t ne2, int64_t ne3, void * image, int variant) {
    const int64_t ncol = ne1 * ne2 * ne3;
    if (k % 32 != 0 || k <= 0) { b200_set_error("quantize_q8_1_soa: k=%lld not a multiple of 32", (long long) k); return B200_ERR_INVALID; }
    if (ncol == 0) return B200_OK;
    const int64_t nblk_col = k / 32, total = nblk_col * ncol;
    int8_t *  qs = (int8_t *) image;
    __half2 * ds = (__half2 *) (qs + ncol * k);
    const unsigned grid = (unsigned) ((total + 7) / 8);
    if (variant == B200_Q8_1_REF)
        B200_LAUNCH(ctx, B200_K_QUANTIZE, k_quantize_q8_1_soa<B200_Q8_1_REF><<<grid, 256, 0, ctx->stream>>>(x, nb1, nb2, nb3, ne1, ne2, qs, ds, nblk_col, total));
    else
        B200_LAUNCH(ctx, B200_K_QUANTIZE, k_quantize_q8_1_soa<B200_Q8_1_X86><<<grid, 256, 0, ctx->stream>>>(x, nb1, nb2, nb3, ne1, ne2, qs, ds, nblk_col, total));
    return B200_OK;
}
