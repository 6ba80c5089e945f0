// b200_gemv.cu -- decode-time dequant-GEMV (MUL_MAT with <= 8 activation columns).
//
// Replaces the reference CPU inner loops ggml_vec_dot_q4_0_q8_0 / q8_0_q8_0 / q4_K_q8_K / q6_K_q8_K
// (ggml/src/ggml-cpu/quants.c:115,305,550; arch/x86/quants.c:543,1011,1741) driven by
// ggml_compute_forward_mul_mat (ggml/src/ggml-cpu/ggml-cpu.c:1226-1418).
//
// Bound: HBM.  Algorithmic bytes per launch = M * K * bytes_per_block / block_size (weights, read once)
// + the q8_1 activation image + 4*M*ncols output; weights dominate (DESIGN.md "GEMV").
//
// Shape of the kernel: one warp per weight row, lanes stride over "units" -- the largest piece of a
// GGUF block a lane can fetch with one 128-bit (or, for the 2-byte aligned legacy blocks, a few
// 32-bit) streaming loads -- integer dot products with dp4a against the SoA q8_1 image (L1/L2
// resident, shared by every row), fp32 scale algebra per unit, one warp-shuffle reduction per row.
// Weights are touched exactly once with ld.global.nc.L1::no_allocate so they never evict the
// activations from L1.
#include "b200_common.cuh"

struct act_image {            // SoA q8_1 image, see b200_quantize.cu
    const int8_t *  qs;       // [ncol][K]
    const __half2 * ds;       // [ncol][K/32]
    int64_t         K;
};

template <int TYPE> struct gemv_traits;

// ---- q4_K: unit = 16 B of qs = 16 weights of sub-block 2j (low nibbles) + 16 of 2j+1 (high) ----
template <> struct gemv_traits<B200_TYPE_Q4_K> {
    static constexpr int units_per_block = 8, block_elems = 256, block_bytes = 144;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const int64_t b = u >> 3; const int c = (int) (u & 7);
        const uint8_t * blk = row + b * 144;
        const int4 hdr = ld_stream_16(blk);                    // d, dmin, 12 scale bytes
        const int4 q   = ld_stream_16(blk + 16 + c * 16);
        const float2 dm = __half22float2(*(const __half2 *) &hdr.x);
        const int p  = c >> 1;                                 // sub-block pair
        const int j0 = 2 * p;
        const unsigned s0 = (unsigned) hdr.y, s1 = (unsigned) hdr.z, s2 = (unsigned) hdr.w;
        int sc0, m0, sc1, m1;
        {
            const int sh0 = 8 * (j0 & 3), sh1 = sh0 + 8;
            if (p < 2) {
                sc0 = (s0 >> sh0) & 63; m0 = (s1 >> sh0) & 63;
                sc1 = (s0 >> sh1) & 63; m1 = (s1 >> sh1) & 63;
            } else {
                sc0 = ((s2 >> sh0) & 0xF) | (((s0 >> (sh0 + 6)) & 3) << 4);
                m0  = ((s2 >> (sh0 + 4)) & 0xF) | (((s1 >> (sh0 + 6)) & 3) << 4);
                sc1 = ((s2 >> sh1) & 0xF) | (((s0 >> (sh1 + 6)) & 3) << 4);
                m1  = ((s2 >> (sh1 + 4)) & 0xF) | (((s1 >> (sh1 + 6)) & 3) << 4);
            }
        }
        const int wl[4] = { q.x & 0x0F0F0F0F, q.y & 0x0F0F0F0F, q.z & 0x0F0F0F0F, q.w & 0x0F0F0F0F };
        const int wh[4] = { (q.x >> 4) & 0x0F0F0F0F, (q.y >> 4) & 0x0F0F0F0F, (q.z >> 4) & 0x0F0F0F0F, (q.w >> 4) & 0x0F0F0F0F };
        const float dsc0 = dm.x * (float) sc0, dsc1 = dm.x * (float) sc1;
        const float mm0  = (c & 1) ? 0.0f : dm.y * (float) m0;  // the min term uses the whole-block sum s: added by the even lane only
        const float mm1  = (c & 1) ? 0.0f : dm.y * (float) m1;
        const int64_t ab = b * 8 + j0;                          // activation block of sub-block j0
        const int     ao = 16 * (c & 1);
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t *  aq = a.qs + (col0 + n) * a.K;
            const __half2 * ad = a.ds + (col0 + n) * (a.K >> 5);
            const int4 a0 = *(const int4 *) (aq + ab * 32 + ao);
            const int4 a1 = *(const int4 *) (aq + ab * 32 + 32 + ao);
            const float2 ds0 = __half22float2(ad[ab]), ds1 = __half22float2(ad[ab + 1]);
            int i0 = __dp4a(wl[0], a0.x, 0); i0 = __dp4a(wl[1], a0.y, i0); i0 = __dp4a(wl[2], a0.z, i0); i0 = __dp4a(wl[3], a0.w, i0);
            int i1 = __dp4a(wh[0], a1.x, 0); i1 = __dp4a(wh[1], a1.y, i1); i1 = __dp4a(wh[2], a1.z, i1); i1 = __dp4a(wh[3], a1.w, i1);
            acc[n] += dsc0 * ds0.x * (float) i0 + dsc1 * ds1.x * (float) i1 - (mm0 * ds0.y + mm1 * ds1.y);
        }
    }
};

// ---- q5_K: q4_K plus one high bit per weight from qh ------------------------------------------
template <> struct gemv_traits<B200_TYPE_Q5_K> {
    static constexpr int units_per_block = 8, block_elems = 256, block_bytes = 176;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const int64_t b = u >> 3; const int c = (int) (u & 7);
        const uint8_t * blk = row + b * 176;
        const int4 hdr = ld_stream_16(blk);
        const int4 qh  = ld_stream_16(blk + 16 + 16 * (c & 1));     // high bits of elements l = 16*(c&1)..+15, all 8 sub-blocks
        const int4 q   = ld_stream_16(blk + 48 + c * 16);
        const float2 dm = __half22float2(*(const __half2 *) &hdr.x);
        const int p  = c >> 1;
        const int j0 = 2 * p;
        const unsigned s0 = (unsigned) hdr.y, s1 = (unsigned) hdr.z, s2 = (unsigned) hdr.w;
        int sc0, m0, sc1, m1;
        {
            const int sh0 = 8 * (j0 & 3), sh1 = sh0 + 8;
            if (p < 2) {
                sc0 = (s0 >> sh0) & 63; m0 = (s1 >> sh0) & 63;
                sc1 = (s0 >> sh1) & 63; m1 = (s1 >> sh1) & 63;
            } else {
                sc0 = ((s2 >> sh0) & 0xF) | (((s0 >> (sh0 + 6)) & 3) << 4);
                m0  = ((s2 >> (sh0 + 4)) & 0xF) | (((s1 >> (sh0 + 6)) & 3) << 4);
                sc1 = ((s2 >> sh1) & 0xF) | (((s0 >> (sh1 + 6)) & 3) << 4);
                m1  = ((s2 >> (sh1 + 4)) & 0xF) | (((s1 >> (sh1 + 6)) & 3) << 4);
            }
        }
        const int qv[4] = { q.x, q.y, q.z, q.w }, hv[4] = { qh.x, qh.y, qh.z, qh.w };
        int wl[4], wh[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            wl[i] = (qv[i] & 0x0F0F0F0F)        | (((hv[i] >> j0)       & 0x01010101) << 4);
            wh[i] = ((qv[i] >> 4) & 0x0F0F0F0F) | (((hv[i] >> (j0 + 1)) & 0x01010101) << 4);
        }
        const float dsc0 = dm.x * (float) sc0, dsc1 = dm.x * (float) sc1;
        const float mm0  = (c & 1) ? 0.0f : dm.y * (float) m0;
        const float mm1  = (c & 1) ? 0.0f : dm.y * (float) m1;
        const int64_t ab = b * 8 + j0;
        const int     ao = 16 * (c & 1);
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t *  aq = a.qs + (col0 + n) * a.K;
            const __half2 * ad = a.ds + (col0 + n) * (a.K >> 5);
            const int4 a0 = *(const int4 *) (aq + ab * 32 + ao);
            const int4 a1 = *(const int4 *) (aq + ab * 32 + 32 + ao);
            const float2 ds0 = __half22float2(ad[ab]), ds1 = __half22float2(ad[ab + 1]);
            int i0 = __dp4a(wl[0], a0.x, 0); i0 = __dp4a(wl[1], a0.y, i0); i0 = __dp4a(wl[2], a0.z, i0); i0 = __dp4a(wl[3], a0.w, i0);
            int i1 = __dp4a(wh[0], a1.x, 0); i1 = __dp4a(wh[1], a1.y, i1); i1 = __dp4a(wh[2], a1.z, i1); i1 = __dp4a(wh[3], a1.w, i1);
            acc[n] += dsc0 * ds0.x * (float) i0 + dsc1 * ds1.x * (float) i1 - (mm0 * ds0.y + mm1 * ds1.y);
        }
    }
};

// ---- q6_K (raw GGUF, 210-byte blocks, 2-byte aligned): unit = 4 consecutive l of one 128-half ----
// -> 4 weights from each of the 4 groups (y[l], y[l+32], y[l+64], y[l+96]; ggml-quants.c:1762-1791)
template <> struct gemv_traits<B200_TYPE_Q6_K> {
    static constexpr int units_per_block = 16, block_elems = 256, block_bytes = 210;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const int64_t b = u >> 4; const int r = (int) (u & 15);
        const int h = r >> 3, l = 4 * (r & 7);
        const uint8_t * blk = row + b * 210;
        const int qa = ld_u16x2(blk + 64 * h + l);
        const int qb = ld_u16x2(blk + 64 * h + 32 + l);
        const int qh = ld_u16x2(blk + 128 + 32 * h + l);
        const int8_t * sc = (const int8_t *) (blk + 192) + 8 * h + (l >> 4);
        const float d = __half2float(*(const __half *) (blk + 208));
        int w[4];
        w[0] = __vsub4((qa & 0x0F0F0F0F)        | ((qh & 0x03030303) << 4),        0x20202020);
        w[1] = __vsub4((qb & 0x0F0F0F0F)        | (((qh >> 2) & 0x03030303) << 4), 0x20202020);
        w[2] = __vsub4(((qa >> 4) & 0x0F0F0F0F) | (((qh >> 4) & 0x03030303) << 4), 0x20202020);
        w[3] = __vsub4(((qb >> 4) & 0x0F0F0F0F) | (((qh >> 6) & 0x03030303) << 4), 0x20202020);
        float dsc[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) dsc[g] = d * (float) sc[2 * g];
        const int64_t ab = b * 8 + 4 * h;
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t *  aq = a.qs + (col0 + n) * a.K;
            const __half2 * ad = a.ds + (col0 + n) * (a.K >> 5);
            float s = 0.0f;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const int av = *(const int *) (aq + (ab + g) * 32 + l);
                s += dsc[g] * __low2float(ad[ab + g]) * (float) __dp4a(w[g], av, 0);
            }
            acc[n] += s;
        }
    }
};

// ---- q6_K in the group-SoA repack (b200_repack.cu): same unit, fields found through the group map ----------------
#define B200_TYPE_Q6_K_REPACKED 1014
template <> struct gemv_traits<B200_TYPE_Q6_K_REPACKED> {
    static constexpr int units_per_block = 16, block_elems = 256, block_bytes = 210;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const int64_t b = u >> 4; const int r = (int) (u & 15);
        const int h = r >> 3, l = 4 * (r & 7), j = l >> 4, ii = l & 15, bb = (int) (b & 7);
        const uint8_t * G = row + (b >> 3) * 1680;
        const int qa = *(const int *) (G + ((((h * 2 + j) * 2 + 0) * 8 + bb) << 4) + ii);
        const int qb = *(const int *) (G + ((((h * 2 + j) * 2 + 1) * 8 + bb) << 4) + ii);
        const int qh = *(const int *) (G + 1024 + (((h * 2 + j) * 8 + bb) << 4) + ii);
        const int8_t * sc = (const int8_t *) (G + 1536 + bb * 16) + 8 * h + j;
        const float d = __half2float(*(const __half *) (G + 1664 + bb * 2));
        int w[4];
        w[0] = __vsub4((qa & 0x0F0F0F0F)        | ((qh & 0x03030303) << 4),        0x20202020);
        w[1] = __vsub4((qb & 0x0F0F0F0F)        | (((qh >> 2) & 0x03030303) << 4), 0x20202020);
        w[2] = __vsub4(((qa >> 4) & 0x0F0F0F0F) | (((qh >> 4) & 0x03030303) << 4), 0x20202020);
        w[3] = __vsub4(((qb >> 4) & 0x0F0F0F0F) | (((qh >> 6) & 0x03030303) << 4), 0x20202020);
        float dsc[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) dsc[g] = d * (float) sc[2 * g];
        const int64_t ab = b * 8 + 4 * h;
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t *  aq = a.qs + (col0 + n) * a.K;
            const __half2 * ad = a.ds + (col0 + n) * (a.K >> 5);
            float s = 0.0f;
#pragma unroll
            for (int g = 0; g < 4; ++g) {
                const int av = *(const int *) (aq + (ab + g) * 32 + l);
                s += dsc[g] * __low2float(ad[ab + g]) * (float) __dp4a(w[g], av, 0);
            }
            acc[n] += s;
        }
    }
};

// ---- q4_0 (raw GGUF, 18-byte blocks): unit = one block --------------------------------------
template <> struct gemv_traits<B200_TYPE_Q4_0> {
    static constexpr int units_per_block = 1, block_elems = 32, block_bytes = 18;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const uint8_t * blk = row + u * 18;
        const float d = __half2float(*(const __half *) blk);
        int q[4];
#pragma unroll
        for (int i = 0; i < 4; ++i) q[i] = ld_u16x2(blk + 2 + 4 * i);
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t *  aq = a.qs + (col0 + n) * a.K + u * 32;
            const float2    ds = __half22float2(a.ds[(col0 + n) * (a.K >> 5) + u]);
            const int4 a0 = *(const int4 *) aq, a1 = *(const int4 *) (aq + 16);
            int s = 0;
            s = __dp4a(q[0] & 0x0F0F0F0F, a0.x, s); s = __dp4a(q[1] & 0x0F0F0F0F, a0.y, s);
            s = __dp4a(q[2] & 0x0F0F0F0F, a0.z, s); s = __dp4a(q[3] & 0x0F0F0F0F, a0.w, s);
            s = __dp4a((q[0] >> 4) & 0x0F0F0F0F, a1.x, s); s = __dp4a((q[1] >> 4) & 0x0F0F0F0F, a1.y, s);
            s = __dp4a((q[2] >> 4) & 0x0F0F0F0F, a1.z, s); s = __dp4a((q[3] >> 4) & 0x0F0F0F0F, a1.w, s);
            acc[n] += d * (ds.x * (float) s - 8.0f * ds.y);      // sum (q-8)*a = sum q*a - 8*sum a
        }
    }
};

// ---- q8_0 (raw GGUF, 34-byte blocks): unit = one block --------------------------------------
template <> struct gemv_traits<B200_TYPE_Q8_0> {
    static constexpr int units_per_block = 1, block_elems = 32, block_bytes = 34;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const uint8_t * blk = row + u * 34;
        const float d = __half2float(*(const __half *) blk);
        int q[8];
#pragma unroll
        for (int i = 0; i < 8; ++i) q[i] = ld_u16x2(blk + 2 + 4 * i);
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t * aq = a.qs + (col0 + n) * a.K + u * 32;
            const float    da = __low2float(a.ds[(col0 + n) * (a.K >> 5) + u]);
            const int4 a0 = *(const int4 *) aq, a1 = *(const int4 *) (aq + 16);
            int s = 0;
            s = __dp4a(q[0], a0.x, s); s = __dp4a(q[1], a0.y, s); s = __dp4a(q[2], a0.z, s); s = __dp4a(q[3], a0.w, s);
            s = __dp4a(q[4], a1.x, s); s = __dp4a(q[5], a1.y, s); s = __dp4a(q[6], a1.z, s); s = __dp4a(q[7], a1.w, s);
            acc[n] += d * da * (float) s;
        }
    }
};

// ---- q4_0 / q8_0 in the load-time run-of-8 layout (b200_repack.cu): [8 x d][8 x qs] per 256 weights, 16-byte aligned quants ----
#define B200_TYPE_Q4_0_REPACKED 1002
#define B200_TYPE_Q8_0_REPACKED 1008
template <> struct gemv_traits<B200_TYPE_Q4_0_REPACKED> {
    static constexpr int units_per_block = 1, block_elems = 32, block_bytes = 18;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const uint8_t * run = row + (u >> 3) * 144; const int c = (int) (u & 7);
        const float d = __half2float(*(const __half *) (run + 2 * c));
        const int4 qv = ld_stream_16(run + 16 + 16 * c);
        const int q[4] = { qv.x, qv.y, qv.z, qv.w };
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t *  aq = a.qs + (col0 + n) * a.K + u * 32;
            const float2    ds = __half22float2(a.ds[(col0 + n) * (a.K >> 5) + u]);
            const int4 a0 = *(const int4 *) aq, a1 = *(const int4 *) (aq + 16);
            int s = 0;
            s = __dp4a(q[0] & 0x0F0F0F0F, a0.x, s); s = __dp4a(q[1] & 0x0F0F0F0F, a0.y, s);
            s = __dp4a(q[2] & 0x0F0F0F0F, a0.z, s); s = __dp4a(q[3] & 0x0F0F0F0F, a0.w, s);
            s = __dp4a((q[0] >> 4) & 0x0F0F0F0F, a1.x, s); s = __dp4a((q[1] >> 4) & 0x0F0F0F0F, a1.y, s);
            s = __dp4a((q[2] >> 4) & 0x0F0F0F0F, a1.z, s); s = __dp4a((q[3] >> 4) & 0x0F0F0F0F, a1.w, s);
            acc[n] += d * (ds.x * (float) s - 8.0f * ds.y);
        }
    }
};
template <> struct gemv_traits<B200_TYPE_Q8_0_REPACKED> {
    static constexpr int units_per_block = 1, block_elems = 32, block_bytes = 34;
    template <int NC>
    static __device__ __forceinline__ void unit(const uint8_t * row, int64_t u, const act_image & a, int64_t col0, float * acc) {
        const uint8_t * run = row + (u >> 3) * 272; const int c = (int) (u & 7);
        const float d = __half2float(*(const __half *) (run + 2 * c));
        const int4 q0 = ld_stream_16(run + 16 + 32 * c), q1 = ld_stream_16(run + 32 + 32 * c);
#pragma unroll
        for (int n = 0; n < NC; ++n) {
            const int8_t * aq = a.qs + (col0 + n) * a.K + u * 32;
            const float    da = __low2float(a.ds[(col0 + n) * (a.K >> 5) + u]);
            const int4 a0 = *(const int4 *) aq, a1 = *(const int4 *) (aq + 16);
            int s = 0;
            s = __dp4a(q0.x, a0.x, s); s = __dp4a(q0.y, a0.y, s); s = __dp4a(q0.z, a0.z, s); s = __dp4a(q0.w, a0.w, s);
            s = __dp4a(q1.x, a1.x, s); s = __dp4a(q1.y, a1.y, s); s = __dp4a(q1.z, a1.z, s); s = __dp4a(q1.w, a1.w, s);
            acc[n] += d * da * (float) s;
        }
    }
};

// ------------------------------------------------------------------------------------------
struct gemv_args {
    const uint8_t * w; size_t w_nb1, w_nb2, w_nb3;     // weight rows
    act_image a;
    float * dst; size_t d_nb1, d_nb2, d_nb3;           // dst[row + col*nb1/4 ...]
    int64_t M;                                          // rows
    int64_t ne11, ne12, ne13;                           // activation cols and batch dims
    int64_t r2, r3;                                     // broadcast ratios ne12/ne02, ne13/ne03
    int64_t col_base;                                   // first column of this launch (for > 8 columns)
};

template <int TYPE, int NC, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_gemv(const gemv_args p) {
    using T = gemv_traits<TYPE>;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row = (int64_t) blockIdx.x * WARPS + warp;
    if (row >= p.M) return;
    const int64_t i12 = blockIdx.y % p.ne12, i13 = blockIdx.y / p.ne12;
    const uint8_t * wrow = p.w + row * p.w_nb1 + (i12 / p.r2) * p.w_nb2 + (i13 / p.r3) * p.w_nb3;
    const int64_t col0 = (i13 * p.ne12 + i12) * p.ne11 + p.col_base;
    const int64_t nunits = p.a.K / T::block_elems * T::units_per_block;

    float acc[NC];
#pragma unroll
    for (int n = 0; n < NC; ++n) acc[n] = 0.0f;

    int64_t u = lane;
    // 2-way unrolled so two units' weight loads are in flight per lane
    for (; u + 32 < nunits; u += 64) {
        T::template unit<NC>(wrow, u,      p.a, col0, acc);
        T::template unit<NC>(wrow, u + 32, p.a, col0, acc);
    }
    if (u < nunits) T::template unit<NC>(wrow, u, p.a, col0, acc);

    float * drow = (float *) ((char *) p.dst + i12 * p.d_nb2 + i13 * p.d_nb3) + row;
#pragma unroll
    for (int n = 0; n < NC; ++n) {
        const float s = warp_sum(acc[n]);
        if (lane == 0) *(float *) ((char *) drow + (p.col_base + n) * p.d_nb1) = s;
    }
}

template <int TYPE>
static int launch_gemv_type(b200_ctx * ctx, gemv_args p, int ncols) {
    constexpr int WARPS = 4;
    dim3 grid((unsigned) ((p.M + WARPS - 1) / WARPS), (unsigned) (p.ne12 * p.ne13));
    for (int c0 = 0; c0 < ncols; c0 += 8) {
        const int nc = ncols - c0 < 8 ? ncols - c0 : 8;
        p.col_base = c0;
        switch (nc) {
            case 1: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 1, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
            case 2: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 2, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
            case 3: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 3, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
            case 4: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 4, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
            case 5: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 5, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
            case 6: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 6, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
            case 7: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 7, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
            default: B200_LAUNCH(ctx, B200_K_GEMV, k_gemv<TYPE, 8, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
        }
    }
    return B200_OK;
}

// w: quantized [K, M, ne02, ne03]; image: SoA q8_1 of x[K, ne11, ne12, ne13]; dst f32
int b200_launch_gemv_q(b200_ctx * ctx, const b200_tensor * w, const void * image, int64_t ne11, int64_t ne12, int64_t ne13,
                       b200_tensor * dst) {
    gemv_args p;
    const int64_t K = w->ne[0], ncol = ne11 * ne12 * ne13;
    p.w = (const uint8_t *) w->data; p.w_nb1 = w->nb[1]; p.w_nb2 = w->nb[2]; p.w_nb3 = w->nb[3];
    p.a.qs = (const int8_t *) image; p.a.ds = (const __half2 *) ((const int8_t *) image + ncol * K); p.a.K = K;
    p.dst = (float *) dst->data; p.d_nb1 = dst->nb[1]; p.d_nb2 = dst->nb[2]; p.d_nb3 = dst->nb[3];
    p.M = w->ne[1]; p.ne11 = ne11; p.ne12 = ne12; p.ne13 = ne13;
    p.r2 = ne12 / w->ne[2]; p.r3 = ne13 / w->ne[3]; p.col_base = 0;
    if (p.ne12 * p.ne13 > 65535) { b200_set_error("gemv: batch %lld too large", (long long) (ne12 * ne13)); return B200_ERR_UNSUPPORTED; }
    switch (w->type) {
        case B200_TYPE_Q4_K: return launch_gemv_type<B200_TYPE_Q4_K>(ctx, p, (int) ne11);
        case B200_TYPE_Q5_K: return launch_gemv_type<B200_TYPE_Q5_K>(ctx, p, (int) ne11);
        case B200_TYPE_Q6_K: return (w->flags & B200_TENSOR_REPACKED) ? launch_gemv_type<B200_TYPE_Q6_K_REPACKED>(ctx, p, (int) ne11)
                                                                        : launch_gemv_type<B200_TYPE_Q6_K>(ctx, p, (int) ne11);
        case B200_TYPE_Q4_0: return (w->flags & B200_TENSOR_REPACKED) ? launch_gemv_type<B200_TYPE_Q4_0_REPACKED>(ctx, p, (int) ne11)
                                                                        : launch_gemv_type<B200_TYPE_Q4_0>(ctx, p, (int) ne11);
        case B200_TYPE_Q8_0: return (w->flags & B200_TENSOR_REPACKED) ? launch_gemv_type<B200_TYPE_Q8_0_REPACKED>(ctx, p, (int) ne11)
                                                                        : launch_gemv_type<B200_TYPE_Q8_0>(ctx, p, (int) ne11);
    }
    b200_set_error("gemv: weight type %d unsupported", w->type);
    return B200_ERR_UNSUPPORTED;
}

// ---- MUL_MAT_ID at decode batch: one GEMV per (token, used-expert slot), expert picked on device ----
// Reference semantics: ggml/src/ggml.c:3205-3241, CPU ggml/src/ggml-cpu/ggml-cpu.c:1500-1700.
//   dst[:, e, t] = as[:, :, ids[e, t]]^T * b[:, e % ne11, t]
struct gemv_id_args {
    const uint8_t * w; size_t w_nb1, w_nb2;            // as[K, M, n_expert]
    act_image a;                                        // image of b[K, ne11, n_tokens]: column = t*ne11 + i11
    const int32_t * ids; size_t ids_nb1;                // ids[n_used, n_tokens]
    float * dst; size_t d_nb1, d_nb2;                   // dst[M, n_used, n_tokens]
    int64_t M, ne11, n_used, n_expert;
};

template <int TYPE, int WARPS>
__global__ void __launch_bounds__(WARPS * 32) k_gemv_id(const gemv_id_args p) {
    using T = gemv_traits<TYPE>;
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t row = (int64_t) blockIdx.x * WARPS + warp;
    if (row >= p.M) return;
    const int64_t t = blockIdx.y / p.n_used, e = blockIdx.y % p.n_used;
    const int32_t expert = *(const int32_t *) ((const char *) p.ids + t * p.ids_nb1 + e * 4);
    if (expert < 0 || expert >= p.n_expert) return;     // reference asserts; we leave dst untouched
    const uint8_t * wrow = p.w + (int64_t) expert * p.w_nb2 + row * p.w_nb1;
    const int64_t col0 = t * p.ne11 + (e % p.ne11);
    const int64_t nunits = p.a.K / T::block_elems * T::units_per_block;
    float acc[1] = { 0.0f };
    int64_t u = lane;
    for (; u + 32 < nunits; u += 64) {
        T::template unit<1>(wrow, u,      p.a, col0, acc);
        T::template unit<1>(wrow, u + 32, p.a, col0, acc);
    }
    if (u < nunits) T::template unit<1>(wrow, u, p.a, col0, acc);
    const float s = warp_sum(acc[0]);
    if (lane == 0) *((float *) ((char *) p.dst + e * p.d_nb1 + t * p.d_nb2) + row) = s;
}

int b200_launch_gemv_id_q(b200_ctx * ctx, const b200_tensor * as, const void * image, int64_t ne11, int64_t n_tokens,
                          const b200_tensor * ids, b200_tensor * dst) {
    constexpr int WARPS = 4;
    gemv_id_args p;
    const int64_t K = as->ne[0], ncol = ne11 * n_tokens;
    p.w = (const uint8_t *) as->data; p.w_nb1 = as->nb[1]; p.w_nb2 = as->nb[2];
    p.a.qs = (const int8_t *) image; p.a.ds = (const __half2 *) ((const int8_t *) image + ncol * K); p.a.K = K;
    p.ids = (const int32_t *) ids->data; p.ids_nb1 = ids->nb[1];
    p.dst = (float *) dst->data; p.d_nb1 = dst->nb[1]; p.d_nb2 = dst->nb[2];
    p.M = as->ne[1]; p.ne11 = ne11; p.n_used = ids->ne[0]; p.n_expert = as->ne[2];
    const int64_t npairs = p.n_used * n_tokens;
    if (npairs > 65535) { b200_set_error("gemv_id: %lld (token, expert) pairs exceed grid.y", (long long) npairs); return B200_ERR_UNSUPPORTED; }
    dim3 grid((unsigned) ((p.M + WARPS - 1) / WARPS), (unsigned) npairs);
    switch (as->type) {
        case B200_TYPE_Q4_K: B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q4_K, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
        case B200_TYPE_Q5_K: B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q5_K, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p)); break;
        case B200_TYPE_Q6_K:
            if (as->flags & B200_TENSOR_REPACKED) B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q6_K_REPACKED, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p));
            else                                  B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q6_K, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p));
            break;
        case B200_TYPE_Q4_0:
            if (as->flags & B200_TENSOR_REPACKED) B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q4_0_REPACKED, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p));
            else                                  B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q4_0, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p));
            break;
        case B200_TYPE_Q8_0:
            if (as->flags & B200_TENSOR_REPACKED) B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q8_0_REPACKED, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p));
            else                                  B200_LAUNCH(ctx, B200_K_GEMV_ID, k_gemv_id<B200_TYPE_Q8_0, WARPS><<<grid, WARPS * 32, 0, ctx->stream>>>(p));
            break;
        default: b200_set_error("gemv_id: weight type %d unsupported", as->type); return B200_ERR_UNSUPPORTED;
    }
    return B200_OK;
}
