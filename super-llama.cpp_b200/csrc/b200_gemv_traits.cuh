// b200_gemv_traits.cuh -- per-type pieces of the streamed decode GEMVs, shared by k_gemv_stream (b200_gemv_stream.cu) and the
// persistent decode chain (b200_chain.cu): the shared-memory layout of the q8_1 activation image (store_unit) and the dot products of one
// warp over one stage of NR weight-row chunks (consume).  Arithmetic of ggml_vec_dot_q4_K_q8_K / q6_K_q8_K / q4_0_q8_0 / q8_0_q8_0
// (reference ggml/src/ggml-cpu/quants.c:115,305,550,708) over q8_1 activations.
#pragma once
#include "b200_common.cuh"
#include "b200_ptx.cuh"

#define GS_RW     2       // rows per warp per stage (k_gemv_stream)
#define GS_CB     8       // 256-weight blocks per row-chunk

template <int TYPE> struct gs_traits;

// ================================================ q4_K (raw GGUF blocks) ================================================
// 144-byte blocks, every field 16-byte aligned, so the raw row is streamed as is; chunk = 8 blocks = 1152 B.
// Lane -> work inside one stage: block b = lane & 7; it in {0, 1}; c = (lane >> 3) + 4 it is the 16-byte unit of the
// block's qs, i.e. sub-block pair p = c >> 1 = (lane >> 4) + 2 it, half = (lane >> 3) & 1.  All lanes are in the same
// branch of get_scale_min_k4 (ggml-quants.c:703-710) for a given `it`, so the 6-bit decode is straight-line code.
// Shared reads are conflict-free: a quarter-warp reads 8 blocks at stride 144 B = 36 words (banks 4i..4i+3).
// Activations in shared memory: quant planes plane0[c][gb] / plane1[c][gb] (16 B: the int8 that meet the low / high
// nibbles of unit (gb, c)) and ds4[p][gb] = (dA(2p), s(2p), dA(2p+1)/16, s(2p+1)); the /16 undoes the unshifted high
// nibbles (dp4a on q & 0xF0F0F0F0 gives 16 x the sum).
template <> struct gs_traits<B200_TYPE_Q4_K> {
    static constexpr int BLOCK_BYTES = 144;
    static __host__ __device__ constexpr int stages(int shape) { return shape == 1 ? 4 : 3; }
    static __host__ __device__ size_t act_bytes(int K) { return (size_t) K + (size_t) (K >> 8) * 64; }
    // the 16 quants of half hf of q8_1 block blk32 (one 16-byte unit); the first half also stores the block's (d, s)
    static __device__ __forceinline__ void store_unit(uint8_t * s_act, int K, int nblocks, int blk32, int hf, int4 w, float d, float s) {
        const int gb = blk32 >> 3, sub = blk32 & 7;
        *(int4 *) (s_act + (sub & 1) * (K >> 1) + (((sub >> 1) * 2 + hf) * nblocks + gb) * 16) = w;
        if (hf == 0) {
            float * ds = (float *) (s_act + K) + ((sub >> 1) * nblocks + gb) * 4 + (sub & 1) * 2;
            *(float2 *) ds = make_float2((sub & 1) ? d * 0.0625f : d, s);
        }
    }
    template <int NR, int NA>
    static __device__ __forceinline__ void consume(uint32_t st, int row_stride, int bl, int gb, int lane, uint32_t act, int K, int nblocks,
                                                   float (&acc)[NA]) {
        const int cq = lane >> 3, half = cq & 1, sh = 16 * (lane >> 4), shm = sh + 8 * half;
        const uint32_t blk = st + bl * 144;
#pragma unroll
        for (int it = 0; it < 2; ++it) {
            const int c = cq + 4 * it, pp = (lane >> 4) + 2 * it;
            const int4 a0 = lds128(act + (c * nblocks + gb) * 16), a1 = lds128(act + (K >> 1) + (c * nblocks + gb) * 16);
            const int4 dsr = lds128(act + K + (pp * nblocks + gb) * 16);
            const float dA0 = __int_as_float(dsr.x), dA1 = __int_as_float(dsr.z), ssel = __int_as_float(half ? dsr.w : dsr.y);
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                const uint32_t base = blk + r * row_stride;
                const int4 hdr = lds128(base);
                const int4 q   = lds128(base + 16 + c * 16);
                const unsigned s0 = (unsigned) hdr.y, s1 = (unsigned) hdr.z, s2 = (unsigned) hdr.w;
                int sc0, sc1, msel;
                if (it == 0) {
                    sc0  = (s0 >> sh) & 63;
                    sc1  = (s0 >> (sh + 8)) & 63;
                    msel = (s1 >> shm) & 63;
                } else {
                    sc0  = ((s2 >> sh) & 0xF)        | (((s0 >> (sh + 6)) & 3) << 4);
                    sc1  = ((s2 >> (sh + 8)) & 0xF)  | (((s0 >> (sh + 14)) & 3) << 4);
                    msel = ((s2 >> (shm + 4)) & 0xF) | (((s1 >> (shm + 6)) & 3) << 4);
                }
                int i0 = __dp4a(q.x & 0x0F0F0F0F, a0.x, 0); i0 = __dp4a(q.y & 0x0F0F0F0F, a0.y, i0);
                i0 = __dp4a(q.z & 0x0F0F0F0F, a0.z, i0);    i0 = __dp4a(q.w & 0x0F0F0F0F, a0.w, i0);
                int i1 = dp4a_us((unsigned) q.x & 0xF0F0F0F0u, a1.x, 0); i1 = dp4a_us((unsigned) q.y & 0xF0F0F0F0u, a1.y, i1);
                i1 = dp4a_us((unsigned) q.z & 0xF0F0F0F0u, a1.z, i1);    i1 = dp4a_us((unsigned) q.w & 0xF0F0F0F0u, a1.w, i1);
                const float2 dm = __half22float2(*(const __half2 *) &hdr.x);
                const float t = (float) (sc0 * i0) * dA0 + (float) (sc1 * i1) * dA1;
                acc[r] = fmaf(dm.x, t, acc[r]);
                acc[r] = fmaf(-dm.y, (float) msel * ssel, acc[r]);
            }
        }
    }
};

// ================================================ q6_K (group-SoA repack) ================================================
// Raw q6_K blocks are 210 bytes with 2-byte aligned fields: unusable for 128-bit shared loads.  At load time
// (b200_repack, b200_repack.cu) every aligned group of 8 blocks (1680 B) of a row is permuted in place into
//     QL [t = (n*2+j)*2+AB][b] 16 B   ql[64 n + 32 AB + 16 j + 0..15] of block b          offset 0
//     QH [nj = n*2+j][b]       16 B   qh[32 n + 16 j + 0..15]                              offset 1024
//     SC [b]                   16 B   scales[0..15]                                        offset 1536
//     D  [b]                    2 B   d                                                    offset 1664
// so a lane (b = lane & 7, nj = lane >> 3) reads 16-byte words that are contiguous across the 8 lanes of its quarter-warp.
// Unit (b, n, j) = 64 weights y[128 n + 32 g + 16 j + 0..15], g = 0..3 (ggml-quants.c:1762-1791), one int8 scale each:
// sc[8 n + j + 2 g].  Activations: planes[g][nj][gb] (16 int8) and dA4[n][gb] = (dA(4n), .., dA(4n+3)).
template <> struct gs_traits<B200_TYPE_Q6_K> {
    static constexpr int BLOCK_BYTES = 210;
    static __host__ __device__ constexpr int stages(int shape) { return shape == 1 ? 3 : 2; }
    static __host__ __device__ size_t act_bytes(int K) { return (size_t) K + (size_t) (K >> 8) * 32; }
    static __device__ __forceinline__ void store_unit(uint8_t * s_act, int K, int nblocks, int blk32, int hf, int4 w, float d, float s) {
        const int gb = blk32 >> 3, sub = blk32 & 7, n = sub >> 2, g = sub & 3;
        *(int4 *) (s_act + ((g * 4 + n * 2 + hf) * nblocks + gb) * 16) = w;
        if (hf == 0) ((float *) (s_act + K))[(n * nblocks + gb) * 4 + g] = d;
        (void) s;
    }
    template <int NR, int NA>
    static __device__ __forceinline__ void consume(uint32_t st, int row_stride, int bl, int gb, int lane, uint32_t act, int K, int nblocks,
                                                   float (&acc)[NA]) {
        const int nj = lane >> 3, n = nj >> 1, j = nj & 1;
        int4 a[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) a[g] = lds128(act + ((g * 4 + nj) * nblocks + gb) * 16);
        const int4 dar = lds128(act + K + (n * nblocks + gb) * 16);
        const float dA[4] = { __int_as_float(dar.x), __int_as_float(dar.y), __int_as_float(dar.z), __int_as_float(dar.w) };
        // the -32 of every 6-bit weight: sum((v - 32) a) = sum(v a) - 32 sum(a), exact in int32.  sum(a) of this lane's four 16-value units is
        // formed once per stage instead of subtracting 32 from every weight byte of every row (vsub4 is two instructions per word: a fifth of
        // this loop, cuobjdump of k_decode_chain)
        int base[4];
#pragma unroll
        for (int g = 0; g < 4; ++g) {
            int sa = __dp4a(0x01010101, a[g].x, 0); sa = __dp4a(0x01010101, a[g].y, sa); sa = __dp4a(0x01010101, a[g].z, sa); sa = __dp4a(0x01010101, a[g].w, sa);
            base[g] = -32 * sa;
        }
#pragma unroll
        for (int r = 0; r < NR; ++r) {
            const uint32_t G = st + r * row_stride;
            const int4 qa = lds128(G + ((nj * 2 + 0) * 8 + bl) * 16);
            const int4 qb = lds128(G + ((nj * 2 + 1) * 8 + bl) * 16);
            const int4 qh = lds128(G + 1024 + (nj * 8 + bl) * 16);
            const int2 scr = lds64(G + 1536 + bl * 16 + 8 * n);                 // scales[8 n .. 8 n + 7]
            unsigned short dh;
            asm volatile("ld.shared.u16 %0, [%1];" : "=h"(dh) : "r"(G + 1664 + bl * 2));
            const float d = __half2float(__ushort_as_half(dh));
            const int qav[4] = { qa.x, qa.y, qa.z, qa.w }, qbv[4] = { qb.x, qb.y, qb.z, qb.w }, qhv[4] = { qh.x, qh.y, qh.z, qh.w };
            const int av[4][4] = { { a[0].x, a[0].y, a[0].z, a[0].w }, { a[1].x, a[1].y, a[1].z, a[1].w },
                                   { a[2].x, a[2].y, a[2].z, a[2].w }, { a[3].x, a[3].y, a[3].z, a[3].w } };
            int I[4] = { base[0], base[1], base[2], base[3] };
#pragma unroll
            for (int w = 0; w < 4; ++w) {
                const int v0 = (qav[w] & 0x0F0F0F0F)        | ((qhv[w] << 4) & 0x30303030);
                const int v1 = (qbv[w] & 0x0F0F0F0F)        | ((qhv[w] << 2) & 0x30303030);
                const int v2 = ((qav[w] >> 4) & 0x0F0F0F0F) | (qhv[w] & 0x30303030);
                const int v3 = ((qbv[w] >> 4) & 0x0F0F0F0F) | ((qhv[w] >> 2) & 0x30303030);
                I[0] = __dp4a(v0, av[0][w], I[0]);          // (bytes 0 .. 63: the same as unsigned or signed)
                I[1] = __dp4a(v1, av[1][w], I[1]);
                I[2] = __dp4a(v2, av[2][w], I[2]);
                I[3] = __dp4a(v3, av[3][w], I[3]);
            }
            // sc[8 n + j + 2 g]: byte j + 2 g of the 8 loaded bytes, sign-extended
            const unsigned lo = (unsigned) scr.x >> (8 * j), hi = (unsigned) scr.y >> (8 * j);
            const int sc0 = (int) (signed char) (lo & 0xFF), sc1 = (int) (signed char) ((lo >> 16) & 0xFF);
            const int sc2 = (int) (signed char) (hi & 0xFF), sc3 = (int) (signed char) ((hi >> 16) & 0xFF);
            const float t = (float) (sc0 * I[0]) * dA[0] + (float) (sc1 * I[1]) * dA[1] + (float) (sc2 * I[2]) * dA[2] + (float) (sc3 * I[3]) * dA[3];
            acc[r] = fmaf(d, t, acc[r]);
        }
    }
};

// ============================================ q4_0 / q8_0 (run-of-8 repack) ============================================
// b200_repack turns every 8 consecutive blocks (256 weights) into [8 x fp16 d][8 x qs]: 144 B (q4_0) / 272 B (q8_0), both
// multiples of 16 with a bank stride of 4 words, so the same lane map as q4_K applies (b = lane & 7 picks the run, lane >> 3
// the 16-byte unit).  q4_0: unit c = sub-block c, low nibbles meet activations 32c..32c+15 (plane 0), high nibbles
// 32c+16..32c+31 (plane 1); value = d (q - 8), the -8 goes through q8_1's s = d_a * sum(a).  q8_0: unit c16 = 16 quants.
template <> struct gs_traits<B200_TYPE_Q4_0> {
    static constexpr int BLOCK_BYTES = 144;
    static __host__ __device__ constexpr int stages(int shape) { return shape == 1 ? 4 : 3; }
    static __host__ __device__ size_t act_bytes(int K) { return (size_t) K + (size_t) (K >> 5) * 8; }
    static __device__ __forceinline__ void store_unit(uint8_t * s_act, int K, int nblocks, int blk32, int hf, int4 w, float d, float s) {
        const int gb = blk32 >> 3, c = blk32 & 7;
        *(int4 *) (s_act + hf * (K >> 1) + (c * nblocks + gb) * 16) = w;
        if (hf == 0) *(float2 *) ((float *) (s_act + K) + (c * nblocks + gb) * 2) = make_float2(d, s);
    }
    template <int NR, int NA>
    static __device__ __forceinline__ void consume(uint32_t st, int row_stride, int bl, int gb, int lane, uint32_t act, int K, int nblocks,
                                                   float (&acc)[NA]) {
        const int cq = lane >> 3;
#pragma unroll
        for (int it = 0; it < 2; ++it) {
            const int c = cq + 4 * it;
            const int4 a0 = lds128(act + (c * nblocks + gb) * 16), a1 = lds128(act + (K >> 1) + (c * nblocks + gb) * 16);
            const int2 dsr = lds64(act + K + (c * nblocks + gb) * 8);
            const float dA = __int_as_float(dsr.x), sA = __int_as_float(dsr.y);
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                const uint32_t base = st + bl * 144 + r * row_stride;
                const int4 q = lds128(base + 16 + c * 16);
                unsigned short dh;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(dh) : "r"(base + 2 * c));
                int i0 = __dp4a(q.x & 0x0F0F0F0F, a0.x, 0); i0 = __dp4a(q.y & 0x0F0F0F0F, a0.y, i0);
                i0 = __dp4a(q.z & 0x0F0F0F0F, a0.z, i0);    i0 = __dp4a(q.w & 0x0F0F0F0F, a0.w, i0);
                int i1 = dp4a_us((unsigned) q.x & 0xF0F0F0F0u, a1.x, 0); i1 = dp4a_us((unsigned) q.y & 0xF0F0F0F0u, a1.y, i1);
                i1 = dp4a_us((unsigned) q.z & 0xF0F0F0F0u, a1.z, i1);    i1 = dp4a_us((unsigned) q.w & 0xF0F0F0F0u, a1.w, i1);
                acc[r] = fmaf(__half2float(__ushort_as_half(dh)), dA * (float) (i0 + (i1 >> 4)) - 8.0f * sA, acc[r]);
            }
        }
    }
};
template <> struct gs_traits<B200_TYPE_Q8_0> {
    static constexpr int BLOCK_BYTES = 272;
    static __host__ __device__ constexpr int stages(int shape) { return shape == 1 ? 3 : 2; }
    static __host__ __device__ size_t act_bytes(int K) { return (size_t) K + (size_t) (K >> 5) * 4; }
    static __device__ __forceinline__ void store_unit(uint8_t * s_act, int K, int nblocks, int blk32, int hf, int4 w, float d, float s) {
        const int gb = blk32 >> 3, sub = blk32 & 7;
        *(int4 *) (s_act + ((sub * 2 + hf) * nblocks + gb) * 16) = w;
        if (hf == 0) ((float *) (s_act + K))[sub * nblocks + gb] = d;
        (void) s;
    }
    template <int NR, int NA>
    static __device__ __forceinline__ void consume(uint32_t st, int row_stride, int bl, int gb, int lane, uint32_t act, int K, int nblocks,
                                                   float (&acc)[NA]) {
        const int cq = lane >> 3;
#pragma unroll
        for (int it = 0; it < 4; ++it) {
            const int c16 = cq + 4 * it, sub = c16 >> 1;
            const int4 a = lds128(act + (c16 * nblocks + gb) * 16);
            const float dA = __int_as_float(lds32(act + K + (sub * nblocks + gb) * 4));
#pragma unroll
            for (int r = 0; r < NR; ++r) {
                const uint32_t base = st + bl * 272 + r * row_stride;
                const int4 q = lds128(base + 16 + c16 * 16);
                unsigned short dh;
                asm volatile("ld.shared.u16 %0, [%1];" : "=h"(dh) : "r"(base + 2 * sub));
                int i = __dp4a(q.x, a.x, 0); i = __dp4a(q.y, a.y, i); i = __dp4a(q.z, a.z, i); i = __dp4a(q.w, a.w, i);
                acc[r] = fmaf(__half2float(__ushort_as_half(dh)) * dA, (float) i, acc[r]);
            }
        }
    }
};

