// b200_gemm.cu -- prefill MUL_MAT (more than 8 activation columns): dequantize-to-bf16 + tcgen05.mma with TMEM accumulators.
//
// Semantics: ggml_compute_forward_mul_mat (reference ggml/src/ggml-cpu/ggml-cpu.c:1226-1418): dst[M,N] = W[M,K] . X[K,N].
// Bound: tensor pipe (2*M*N*K flops); the weights are read once per 128-row tile and N-tile.
//
// Why bf16 and not the int8 MMA the north star sketches: GGUF scales change every 32 (q4_0/q8_0) or 16/32 (K-quants) values
// along K, so an int8 accumulator would have to be drained from TMEM and rescaled every 32 K -- 1/8th of a pipeline stage.
// Dequantizing to bf16 in shared memory keeps ONE fp32 accumulator in TMEM for the whole K loop; bf16 rounding of weights and
// activations costs ~1e-5 NMSE against the 5e-4 bar (tests/test-backend-ops.cpp:6080-6082).
//
// CTA = MT*128 weight rows x BN activation columns.  The limiter is L2->SM traffic (~6.3 KB/clk chip-wide): per 64-K step a CTA
// pulls BN*128 B of activations but only MT*128*36 B of 4.5-bit weights, so tall tiles (MT=2, BN=128) are preferred.
// The A operand (weights) never touches shared memory as bf16: an SS-mode
// M=128 MMA reads A and B from shared memory at the full 128 B/clk the SM has, which measured as the limiter of the first
// version of this kernel (profiles/r1_gemm_v1_*).  Here A is fed from TMEM.  Warp roles:
//   warps 0-7  A producers: thread = weight row = TMEM lane (MT=2: warps 4-7 own the second 128 rows; MT=1: the two groups
//              alternate 64-K steps).  Raw GGUF blocks of the 128-row tile arrive in a shared ring by one
//              TMA tiled load per 256-K block (a 2-D byte tensor map over the weight rows); the thread expands 64 K-values at a
//              time to bf16 (value = d*sc*q - dmin*m, two values per HFMA2.BF16) and writes them with one tcgen05.st (32 columns
//              of its lane).  After the K loop the same warps run the epilogue: tcgen05.ld the accumulator lanes and store fp32
//              columns (coalesced along M).
//   warp 8     B producer: one thread, TMA tiled loads (SWIZZLE_128B) of the bf16 activation image [N][K].
//   warp 9     TMEM owner + MMA issuer: one thread, four tcgen05.mma (A in TMEM, K=16 each) per stage; tcgen05.commit frees it.
//   warp 10    raw-weight producer: one thread, the TMA loads of the GGUF block ring.
#include "b200_common.cuh"
#include "b200_ptx.cuh"
#include <cuda.h>
#include <stdlib.h>

#define GM_BM      128
#define GM_BK      64
#define GM_PRODUCERS 8       // A-producer warps (two per TMEM lane quarter)
#define GM_THREADS ((GM_PRODUCERS + 3) * 32)

struct gm_args {
    // up to three weight tensors of one type that share the activations (wq/wk/wv, ffn_gate/ffn_up): grid.y walks their row tiles
    struct { float * dst; size_t d_ld; int M; int mtiles; } seg[3];   // dst[m + n * d_ld]
    int nseg;
    int N, K;
    int BN;                              // activation columns per CTA (UMMA N)
    int MT;                              // 128-row weight tiles per CTA (1 or 2): 2 halves the activation traffic per flop
    int SA, SB, RS;                      // ring depths: A stages in TMEM (power of two), B stages in smem (64-K steps), raw 256-K weight blocks
    // MUL_MAT_ID (grouped GEMM): activation rows sorted by expert; expert e owns rows [offsets[e], offsets[e] + counts[e]) of the
    // image, col_off[row] is where that (token, slot) pair's result column starts in dst (in floats).  NULL for plain MUL_MAT.
    const int * counts; const int * offsets; const long long * col_off;
    int dbg;                             // timing experiments only (GGML_B200_GEMM_DBG): 1 = no weight fetch, 2 = no expansion math, 4 = no activation fetch
    // plain MUL_MAT runs on a 1-D grid: CTA id < n_unsplit is the whole tile `id` (N tile fastest); the tiles after those -- the
    // partial last wave -- are cut in two along K, CTA n_unsplit + 2 j + h doing half h of tile n_unsplit + j, so that the tail
    // of the grid takes half a tile time.  Half 0 stores and raises flags[j]; half 1 waits for it and adds (fixed order: results
    // do not depend on timing); half 0 always has the smaller CTA id, so it is resident whenever half 1 is.
    int lin, ntn, n_unsplit; unsigned * flags;
};

// ---- activations f32 [K, N] -> bf16 image [N][K] ---------------------------------------------------------------------
// one CTA per (chunk of 4096 values, row): every thread issues its four 16-byte loads before the first store (a grid-stride loop with one load
// in flight and a 64-bit division per element ran at half the HBM rate: 0.90 ms of a 13.5 ms pp512 ubatch, ncu launch list of llama-bench)
__global__ void __launch_bounds__(256) k_act_bf16(const float * __restrict__ x, size_t nb1, int K, int N, __nv_bfloat16 * __restrict__ out) {
    // PDL: the GEMM behind us may start fetching and expanding its weights now; x comes from the kernel before us, and the image we
    // overwrite may still be read by the GEMM before us, so everything we touch waits for the whole previous grid
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const int n = blockIdx.y, k0 = blockIdx.x * 4096 + threadIdx.x * 4;
    const float * xr = (const float *) ((const char *) x + (size_t) n * nb1);
    float4 v[4];
#pragma unroll
    for (int j = 0; j < 4; ++j) { const int k = k0 + j * 1024; if (k < K) v[j] = *(const float4 *) (xr + k); }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const int k = k0 + j * 1024;
        if (k >= K) continue;
        __nv_bfloat162 a = __floats2bfloat162_rn(v[j].x, v[j].y), b = __floats2bfloat162_rn(v[j].z, v[j].w);
        uint2 o; o.x = *(uint32_t *) &a; o.y = *(uint32_t *) &b;
        *(uint2 *) (out + (size_t) n * K + k) = o;
    }
}
int b200_launch_act_bf16(b200_ctx * ctx, const float * x, size_t nb1, int64_t K, int64_t N, void * image) {
    if (K <= 0 || N <= 0) return B200_OK;
    if (K % 4 != 0 || K > (1 << 30) || N > 65535) { b200_set_error("act_bf16: K %lld, N %lld outside the launch shape", (long long) K, (long long) N); return B200_ERR_UNSUPPORTED; }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) ((K + 4095) / 4096), (unsigned) N); cfg.blockDim = dim3(256); cfg.dynamicSmemBytes = 0; cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    B200_LAUNCH(ctx, B200_K_CONVERT, cudaLaunchKernelEx(&cfg, k_act_bf16, x, nb1, (int) K, (int) N, (__nv_bfloat16 *) image));
    return B200_OK;
}

// ---- per-type expansion of one row's 64 K-values to bf16 -----------------------------------------------------------------
static __device__ __forceinline__ uint32_t bf2(float a, float b) { __nv_bfloat162 t = __floats2bfloat162_rn(a, b); return *(uint32_t *) &t; }
// (t - 128) * s + o on both halves; t holds 128+q as exact bf16 values
static __device__ __forceinline__ uint32_t fma_magic(uint32_t t, uint32_t s, uint32_t o) {
    const uint32_t c128 = 0x43004300u;
    __nv_bfloat162 r = __hfma2(__hsub2(*(__nv_bfloat162 *) &t, *(const __nv_bfloat162 *) &c128), *(__nv_bfloat162 *) &s, *(__nv_bfloat162 *) &o);
    return *(uint32_t *) &r;
}

template <int TYPE> struct gm_traits;

template <> struct gm_traits<B200_TYPE_Q4_K> {
    static constexpr int RAW_K = 256, ROW_BYTES = 144, TILE_BYTES = GM_BM * 144;   // raw slot of one 128-row tile: [128][144]
    static constexpr int A_FMT = 1;                                                // A operand written as bf16
    struct unit { float d, dmin; uint32_t sc_lo, sc_hi, m_lo, m_hi; };   // 8 x 6-bit scales / mins, one per byte
    static __device__ __forceinline__ void load_unit(uint32_t tile, int rt, int, unit & u) {
        const uint32_t raw = tile + rt * ROW_BYTES;
        const int4 h = lds128(raw);                                          // d, dmin, scales[12]
        const float2 dm = __half22float2(*(const __half2 *) &h.x);
        u.d = dm.x; u.dmin = dm.y;
        const uint32_t u0 = h.y, u1 = h.z, u2 = h.w;                         // get_scale_min_k4, ggml-quants.c:703-710
        u.sc_lo = u0 & 0x3f3f3f3fu; u.m_lo = u1 & 0x3f3f3f3fu;
        u.sc_hi = (u2 & 0x0f0f0f0fu) | (((u0 >> 6) & 0x03030303u) << 4);
        u.m_hi  = ((u2 >> 4) & 0x0f0f0f0fu) | (((u1 >> 6) & 0x03030303u) << 4);
    }
    // K-chunk g (0..3) of the block: sub-block 2g = low nibbles of qs[32g..32g+31], sub-block 2g+1 = high nibbles (ggml-quants.c:1352-1374)
    template <int G> static __device__ __forceinline__ void expand(uint32_t tile, int rt, const unit & u, uint32_t (&out)[32]) {
        const uint32_t raw = tile + rt * ROW_BYTES;
        const uint32_t scs = G < 2 ? u.sc_lo : u.sc_hi, ms = G < 2 ? u.m_lo : u.m_hi;
        const int sh = (G & 1) * 16;
        const float d0 = u.d * (float) ((scs >> sh) & 0xff), d1 = u.d * (float) ((scs >> (sh + 8)) & 0xff);
        const float o0 = -u.dmin * (float) ((ms >> sh) & 0xff), o1 = -u.dmin * (float) ((ms >> (sh + 8)) & 0xff);
        const uint32_t s_lo = bf2(d0, d0), s_hi = bf2(d1, d1), o_lo = bf2(o0, o0), o_hi = bf2(o1, o1);
        uint32_t w[8];
        { const int4 a = lds128(raw + 16 + 32 * G), b = lds128(raw + 32 + 32 * G);
          w[0] = a.x; w[1] = a.y; w[2] = a.z; w[3] = a.w; w[4] = b.x; w[5] = b.y; w[6] = b.z; w[7] = b.w; }
        uint32_t * lo = out, * hi = out + 16;                                // natural K order: 32 low-nibble values, then 32 high
#pragma unroll
        for (int i = 0; i < 8; ++i) {
            // nibble q -> bf16 bits 0x4300|q = 128+q (exact); PRMT drops the 0x43 bytes in while pairing neighbours
            const uint32_t l4 = w[i] & 0x0F0F0F0Fu, h4 = (w[i] >> 4) & 0x0F0F0F0Fu;
            lo[2 * i]     = fma_magic(__byte_perm(l4, 0x43434343u, 0x4140), s_lo, o_lo);
            lo[2 * i + 1] = fma_magic(__byte_perm(l4, 0x43434343u, 0x4342), s_lo, o_lo);
            hi[2 * i]     = fma_magic(__byte_perm(h4, 0x43434343u, 0x4140), s_hi, o_hi);
            hi[2 * i + 1] = fma_magic(__byte_perm(h4, 0x43434343u, 0x4342), s_hi, o_hi);
        }
    }
};

// q6_K in the load-time repacked group layout (b200_repack.cu): per row, groups of 8 blocks = 1680 B =
//   QL[8 units][8 blocks][16 B] | QH[4 units][8 blocks][16 B] | SC[8 blocks][16 B] | D[8 blocks][2 B]
// Block b of a group is column b of a [13 x 128 B] byte matrix: one 4-D TMA box {16 B, 13, 1 group, 128 rows} fetches it for a
// whole tile ([128 rows][13][16 B] in the slot); the 16 B of fp16 d's come through a second 3-D map ([128 rows][16 B]).
template <> struct gm_traits<B200_TYPE_Q6_K> {
    static constexpr int RAW_K = 256, ROW_BYTES = 208, TILE_BYTES = GM_BM * (208 + 16);
    static constexpr int A_FMT = 1;
    struct unit { float d; int4 sc; };
    static __device__ __forceinline__ void load_unit(uint32_t tile, int rt, int blk, unit & u) {
        u.sc = lds128(tile + rt * ROW_BYTES + 12 * 16);
        const uint32_t dd = (uint32_t) lds32(tile + GM_BM * ROW_BYTES + rt * 16 + ((blk & 7) >> 1) * 4);
        u.d = __half2float(__ushort_as_half((unsigned short) ((blk & 1) ? dd >> 16 : dd & 0xffffu)));
    }
    // K-step G of the block: half n = G>>1, nibble hi = G&1 (ggml-quants.c:1762-1791).  Element e = ab*32 + j*16 + byte comes from QL unit
    // (n*2+j)*2+ab (nibble hi) and QH unit n*2+j (bits 4hi+2ab, +1), scale sc[8n + 4hi + 2ab + j]; value = d * sc * (q - 32).
    template <int G> static __device__ __forceinline__ void expand(uint32_t tile, int rt, const unit & u, uint32_t (&out)[32]) {
        constexpr int n = G >> 1, hi = G & 1;
        const uint32_t raw = tile + rt * ROW_BYTES;
        const uint32_t scw[4] = { (uint32_t) u.sc.x, (uint32_t) u.sc.y, (uint32_t) u.sc.z, (uint32_t) u.sc.w };
        const uint32_t sc4 = scw[2 * n + hi];                                // the four int8 scales of this K-step: (ab, j) = (0,0) (0,1) (1,0) (1,1)
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int4 qh = lds128(raw + (8 + n * 2 + j) * 16);
#pragma unroll
            for (int ab = 0; ab < 2; ++ab) {
                const int4 ql = lds128(raw + (((n * 2 + j) * 2 + ab) * 16));
                const float sf = u.d * (float) (int) (int8_t) ((sc4 >> (8 * (2 * ab + j))) & 0xffu);
                const uint32_t s2 = bf2(sf, sf);
                const uint32_t qlw[4] = { (uint32_t) ql.x, (uint32_t) ql.y, (uint32_t) ql.z, (uint32_t) ql.w };
                const uint32_t qhw[4] = { (uint32_t) qh.x, (uint32_t) qh.y, (uint32_t) qh.z, (uint32_t) qh.w };
#pragma unroll
                for (int w = 0; w < 4; ++w) {
                    constexpr int hs = 4 * hi;                               // bit offset of the two qh bits is hs + 2ab; they go to bits 4..5
                    const int sh = hs + 2 * ab;
                    const uint32_t lo4 = (qlw[w] >> hs) & 0x0F0F0F0Fu;
                    const uint32_t h2 = sh >= 4 ? (qhw[w] >> (sh - 4)) : (qhw[w] << (4 - sh));
                    const uint32_t q = (h2 & 0x30303030u) | lo4;             // four 6-bit values, one per byte
                    const uint32_t c160 = 0x43204320u;                       // bf16 160 = 128 + 32
                    uint32_t t0 = __byte_perm(q, 0x43434343u, 0x4140), t1 = __byte_perm(q, 0x43434343u, 0x4342);   // 128+q, exact
                    __nv_bfloat162 r0 = __hmul2(__hsub2(*(__nv_bfloat162 *) &t0, *(const __nv_bfloat162 *) &c160), *(const __nv_bfloat162 *) &s2);
                    __nv_bfloat162 r1 = __hmul2(__hsub2(*(__nv_bfloat162 *) &t1, *(const __nv_bfloat162 *) &c160), *(const __nv_bfloat162 *) &s2);
                    out[ab * 16 + j * 8 + w * 2]     = *(uint32_t *) &r0;
                    out[ab * 16 + j * 8 + w * 2 + 1] = *(uint32_t *) &r1;
                }
            }
        }
    }
};

// q4_0 / q8_0 in the run-of-8 layout of b200_repack.cu: per 256 weights [8 x fp16 d][8 x qs].
// q4_0: the same 144-byte rows as q4_K (one 2-D TMA box); value = d (q - 8).
template <> struct gm_traits<B200_TYPE_Q4_0> {
    static constexpr int RAW_K = 256, ROW_BYTES = 144, TILE_BYTES = GM_BM * 144;
    static constexpr int A_FMT = 1;
    struct unit { int4 d8; };
    static __device__ __forceinline__ void load_unit(uint32_t tile, int rt, int, unit & u) { u.d8 = lds128(tile + rt * ROW_BYTES); }
    template <int G> static __device__ __forceinline__ void expand(uint32_t tile, int rt, const unit & u, uint32_t (&out)[32]) {
        const uint32_t raw = tile + rt * ROW_BYTES;
        const uint32_t dw = G == 0 ? (uint32_t) u.d8.x : G == 1 ? (uint32_t) u.d8.y : G == 2 ? (uint32_t) u.d8.z : (uint32_t) u.d8.w;   // d of sub-blocks 2G, 2G+1
        const float2 df = __half22float2(*(const __half2 *) &dw);
#pragma unroll
        for (int h = 0; h < 2; ++h) {                                        // sub-block 2G + h: low nibbles = elements 0..15, high = 16..31
            const int4 qv = lds128(raw + 16 + 16 * (2 * G + h));
            const uint32_t w[4] = { (uint32_t) qv.x, (uint32_t) qv.y, (uint32_t) qv.z, (uint32_t) qv.w };
            const float d = h ? df.y : df.x;
            const uint32_t s2 = bf2(d, d), c136 = 0x43084308u;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const uint32_t l4 = w[i] & 0x0F0F0F0Fu, h4 = (w[i] >> 4) & 0x0F0F0F0Fu;
                uint32_t t[4] = { __byte_perm(l4, 0x43434343u, 0x4140), __byte_perm(l4, 0x43434343u, 0x4342),
                                  __byte_perm(h4, 0x43434343u, 0x4140), __byte_perm(h4, 0x43434343u, 0x4342) };
#pragma unroll
                for (int k = 0; k < 4; ++k) {
                    __nv_bfloat162 r = __hmul2(__hsub2(*(__nv_bfloat162 *) &t[k], *(const __nv_bfloat162 *) &c136), *(const __nv_bfloat162 *) &s2);
                    out[h * 16 + (k >> 1) * 8 + 2 * i + (k & 1)] = *(uint32_t *) &r;
                }
            }
        }
    }
};
// q8_0: 272-byte rows ([16 B of d][256 quants]) fetched as a 3-D box {16 B, 17, 128 rows}.  int8 -> float without a convert:
// u = q + 128 dropped into mantissa bits 8..15 of 0x47000000 is exactly 32768 + u, so d*q = fma(f, d, -32896 d) (the rounding of
// the constant is 2^-24 * 32896 d = 0.002 d, far below the bf16 rounding that follows); pairs are packed to bf16 by F2FP.
template <> struct gm_traits<B200_TYPE_Q8_0> {
    static constexpr int RAW_K = 256, ROW_BYTES = 272, TILE_BYTES = GM_BM * 272;
    static constexpr int A_FMT = 1;
    struct unit { int4 d8; };
    static __device__ __forceinline__ void load_unit(uint32_t tile, int rt, int, unit & u) { u.d8 = lds128(tile + rt * ROW_BYTES); }
    template <int G> static __device__ __forceinline__ void expand(uint32_t tile, int rt, const unit & u, uint32_t (&out)[32]) {
        const uint32_t raw = tile + rt * ROW_BYTES;
        const uint32_t dw = G == 0 ? (uint32_t) u.d8.x : G == 1 ? (uint32_t) u.d8.y : G == 2 ? (uint32_t) u.d8.z : (uint32_t) u.d8.w;
        const float2 df = __half22float2(*(const __half2 *) &dw);
#pragma unroll
        for (int h = 0; h < 2; ++h) {
            const float d = h ? df.y : df.x, c = -32896.0f * d;
#pragma unroll
            for (int v = 0; v < 2; ++v) {
                const int4 qv = lds128(raw + 16 + 32 * (2 * G + h) + 16 * v);
                const uint32_t w[4] = { (uint32_t) qv.x, (uint32_t) qv.y, (uint32_t) qv.z, (uint32_t) qv.w };
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const uint32_t wx = w[i] ^ 0x80808080u;
                    const float f0 = __uint_as_float(__byte_perm(wx, 0x47000000u, 0x7604)), f1 = __uint_as_float(__byte_perm(wx, 0x47000000u, 0x7614));
                    const float f2 = __uint_as_float(__byte_perm(wx, 0x47000000u, 0x7624)), f3 = __uint_as_float(__byte_perm(wx, 0x47000000u, 0x7634));
                    out[h * 16 + v * 8 + 2 * i]     = bf2(fmaf(f0, d, c), fmaf(f1, d, c));
                    out[h * 16 + v * 8 + 2 * i + 1] = bf2(fmaf(f2, d, c), fmaf(f3, d, c));
                }
            }
        }
    }
};

struct gm_maps { CUtensorMap w[3]; CUtensorMap w2[3]; };

template <int TYPE>
__global__ void __launch_bounds__(GM_THREADS, 2) k_gemm_q(const gm_args p, const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ gm_maps maps) {
    using T = gm_traits<TYPE>;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const int MT = p.MT, BN = p.BN, SA = p.SA, SB = p.SB, RS = p.RS;
    const uint32_t base  = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t b_sm  = base;                                             // [SB][BN rows][128 B], 128B-swizzled by TMA
    const uint32_t r_sm  = b_sm + SB * (BN * 128);                    // [RS][MT tiles][TILE_BYTES]
    const uint32_t raw_slot = (uint32_t) (MT * T::TILE_BYTES);
    const uint32_t bars  = r_sm + RS * raw_slot;
    const uint32_t fa0 = bars, ea0 = fa0 + 8 * SA, fb0 = ea0 + 8 * SA, eb0 = fb0 + 8 * SB;   // A written / A consumed / B landed / B consumed
    const uint32_t rf0 = eb0 + 8 * SB, re0 = rf0 + 8 * RS;           // raw block landed / raw block consumed
    const uint32_t tfull = re0 + 8 * RS, tslot = tfull + 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    int sg = 0, mtile = blockIdx.y, ntile = blockIdx.x, kpart = 0, ksplit = 1, split_j = 0;
    if (p.lin) {
        int id = blockIdx.x;
        if (id >= p.n_unsplit) { split_j = (id - p.n_unsplit) >> 1; kpart = (id - p.n_unsplit) & 1; ksplit = 2; id = p.n_unsplit + split_j; }
        mtile = id / p.ntn; ntile = id - mtile * p.ntn;
    }
    if (p.nseg > 1 && mtile >= p.seg[0].mtiles) { mtile -= p.seg[0].mtiles; sg = 1; if (p.nseg > 2 && mtile >= p.seg[1].mtiles) { mtile -= p.seg[1].mtiles; sg = 2; } }
    const CUtensorMap & tmap_w = maps.w[sg], & tmap_w2 = maps.w2[sg];
    const int pM = p.seg[sg].M;
    const int m0 = mtile * (MT * GM_BM);
    int n0 = ntile * BN, n_end = p.N, wrow0 = m0;
    if (p.counts) {                                                          // grouped: this CTA serves expert blockIdx.z
        const int e = blockIdx.z, cnt = p.counts[e];
        if (n0 >= cnt) return;                                               // uniform per CTA, before any barrier or TMEM allocation
        n0 += p.offsets[e]; n_end = p.offsets[e] + cnt; wrow0 = e * pM + m0; // experts are stacked rows of one tall weight matrix
    }
    const int units_all = p.K / T::RAW_K;
    const int u0 = ksplit == 2 ? kpart * (units_all / 2) : 0, u1 = ksplit == 2 && kpart == 0 ? units_all / 2 : units_all;   // this CTA's 256-K blocks
    const int nunits = u1 - u0, nks = nunits * (T::RAW_K / GM_BK);
    // TMEM columns: accumulators [0, MT*BN), then A stages: stage s, tile t at MT*BN + (s*MT + t)*32
    const uint32_t tneed = (uint32_t) (MT * BN + MT * 32 * SA);
    const uint32_t tcols = tneed <= 128 ? 128u : tneed <= 256 ? 256u : 512u;

    if (threadIdx.x == 0) {
        for (int s = 0; s < SA; ++s) { mbar_init(fa0 + 8 * s, 4 * MT); mbar_init(ea0 + 8 * s, 1); }
        for (int s = 0; s < SB; ++s) { mbar_init(fb0 + 8 * s, 1); mbar_init(eb0 + 8 * s, 1); }
        for (int s = 0; s < RS; ++s) { mbar_init(rf0 + 8 * s, 1); mbar_init(re0 + 8 * s, GM_PRODUCERS); }
        mbar_init(tfull, 1);
        mbar_fence_init();
    }
    if (warp == GM_PRODUCERS + 1) tmem_alloc(tslot, tcols);
    if (warp == GM_PRODUCERS && lane == 0) tma_prefetch_desc(&tmap_x);
    if (warp == GM_PRODUCERS + 2 && lane == 0) tma_prefetch_desc(&tmap_w);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tslot));
    const uint32_t tmem_a = tmem + (uint32_t) (MT * BN);
    // PDL: weights are static, so the raw-weight producer and the expanding warps run ahead under the previous kernel's tail (on the
    // SMs its last wave leaves idle); only the activation image (B producer) and dst (epilogue) wait for the previous grid
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    if (warp < GM_PRODUCERS) {
        // ===== A producers =====
        const int grp = warp >> 2, q = warp & 3;
        const int tile = MT == 2 ? grp : 0;                                  // which 128-row tile this thread's row is in
        const int rt = q * 32 + lane;                                        // row inside the tile = TMEM lane
        const uint32_t lane_base = (uint32_t) (q * 32) << 16;
        const int sshift = SA == 8 ? 3 : SA == 4 ? 2 : 1;
        int pending = -1;                                                    // stage whose tcgen05.st is still in flight
        auto flush = [&]() {
            if (pending >= 0) {
                tmem_st_wait(); tc_fence_before(); __syncwarp();
                if (lane == 0) mbar_arrive(fa0 + 8 * pending);
                pending = -1;
            }
        };
        auto emit = [&](int ks, const uint32_t (&o)[32]) {
            flush();                                                         // the expansion of this step overlapped the previous store
            const int stage = ks & (SA - 1);
            mbar_wait(ea0 + 8 * stage, (((uint32_t) ks >> sshift) & 1u) ^ 1u);
            tc_fence_after();
            tmem_st32(tmem_a + lane_base + (uint32_t) ((stage * MT + tile) * 32), o);
            pending = stage;
        };
        int slot = 0; uint32_t rph = 0;
        for (int unit = 0; unit < nunits; ++unit) {
            if (!(p.dbg & 1)) mbar_wait(rf0 + 8 * slot, rph);
            const uint32_t raw = r_sm + slot * raw_slot + tile * T::TILE_BYTES;
            typename T::unit u; T::load_unit(raw, rt, u0 + unit, u);
            uint32_t o[32];
            const int ks = unit * 4;
            if (p.dbg & 2) { for (int c = 0; c < 32; ++c) o[c] = 0x3c003c00u + (uint32_t) c; for (int g = 0; g < 4; ++g) if (MT == 2 || (g & 1) == grp) emit(ks + g, o); }
            else {
                if (MT == 2 || grp == 0) { T::template expand<0>(raw, rt, u, o); emit(ks + 0, o); }
                if (MT == 2 || grp == 1) { T::template expand<1>(raw, rt, u, o); emit(ks + 1, o); }
                if (MT == 2 || grp == 0) { T::template expand<2>(raw, rt, u, o); emit(ks + 2, o); }
                if (MT == 2 || grp == 1) { T::template expand<3>(raw, rt, u, o); emit(ks + 3, o); }
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(re0 + 8 * slot);
            if (++slot == RS) { slot = 0; rph ^= 1u; }
        }
        flush();
        // ===== epilogue: TMEM -> registers -> dst (columns of dst are contiguous along M) =====
        mbar_wait(tfull, 0);
        tc_fence_after();
        asm volatile("griddepcontrol.wait;" ::: "memory");                    // long satisfied; orders our stores after the previous grid
        const int mm = m0 + tile * GM_BM + q * 32 + lane;
        const int c_begin = MT == 2 ? 0 : grp * (BN / 2), c_end = MT == 2 ? BN : c_begin + BN / 2;   // MT=1: the groups split the columns
        float * drow = p.seg[sg].dst + mm;
        if (ksplit == 2 && kpart == 1) {                                     // the other half's stores must be visible before we add
            // bounded (2 s): the other half's CTA has a lower block index and is resident or done when this one runs, but a launch that was
            // cut short must not leave this one spinning for ever; a wait that ran out is recorded in flags[1024] and the tile's sum is wrong
            unsigned long long t0 = 0; unsigned polls = 0;
            while (*(volatile unsigned *) (p.flags + split_j) == 0u) {
                if ((++polls & 0x3ffu) != 0) continue;
                unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
                if (!t0) t0 = t;
                else if (t - t0 > 2000000000ull) { if (threadIdx.x % 32 == 0) atomicExch(p.flags + 1024, 1u); break; }
            }
            __threadfence();
        }
        for (int c0 = c_begin; c0 < c_end; c0 += 16) {
            if (n0 + c0 >= n_end) break;
            uint32_t r[16];
            tmem_ld16(tmem + lane_base + (uint32_t) (tile * BN + c0), r);
            tmem_ld_wait();
            if (mm < pM) {
                if (ksplit == 2 && kpart == 1) {                             // all 16 loads in flight before the first store (they may not be reordered past it)
                    float prev[16];
#pragma unroll
                    for (int j = 0; j < 16; ++j) prev[j] = n0 + c0 + j < n_end ? __ldcg(drow + (size_t) (n0 + c0 + j) * p.seg[sg].d_ld) : 0.0f;
#pragma unroll
                    for (int j = 0; j < 16; ++j) r[j] = __float_as_uint(prev[j] + __uint_as_float(r[j]));
                }
#pragma unroll
                for (int j = 0; j < 16; ++j)
                    if (n0 + c0 + j < n_end) drow[p.col_off ? (size_t) p.col_off[n0 + c0 + j] : (size_t) (n0 + c0 + j) * p.seg[sg].d_ld] = __uint_as_float(r[j]);
            }
        }
        if (ksplit == 2) {
            __threadfence();
            asm volatile("bar.sync 1, %0;" :: "n"(GM_PRODUCERS * 32));       // the eight epilogue warps
            if (threadIdx.x == 0) *(volatile unsigned *) (p.flags + split_j) = kpart == 0 ? 1u : 0u;   // half 1 re-arms the flag for the next launch
        }
        tc_fence_before();
    } else if (warp == GM_PRODUCERS) {
        // ===== B producer (control flow is warp-uniform; one elected lane issues) =====
        asm volatile("griddepcontrol.wait;" ::: "memory");                    // the image is written by the kernel before us
        int stage = 0; uint32_t eph = 1;
        for (int ks = 0; ks < nks; ++ks) {
            mbar_wait(eb0 + 8 * stage, eph);
            if (elect_one()) {
                if (p.dbg & 4) mbar_arrive(fb0 + 8 * stage);
                else {
                    mbar_arrive_expect_tx(fb0 + 8 * stage, (uint32_t) BN * 128u);
                    tma_load_2d(b_sm + stage * (BN * 128), &tmap_x, (u0 * (T::RAW_K / GM_BK) + ks) * GM_BK, n0, fb0 + 8 * stage);
                }
            }
            __syncwarp();
            if (++stage == SB) { stage = 0; eph ^= 1u; }
        }
    } else if (warp == GM_PRODUCERS + 1) {
        // ===== MMA issuer =====
        const uint32_t idesc = umma_idesc_bf16(GM_BM, BN) & ~((T::A_FMT ? 0u : 1u) << 7);     // A format: bf16 or fp16; B is always bf16
        int sa = 0, sb = 0; uint32_t pa = 0, pb = 0;
        for (int ks = 0; ks < nks; ++ks) {
            mbar_wait(fa0 + 8 * sa, pa);
            mbar_wait(fb0 + 8 * sb, pb);
            tc_fence_after();
            const uint64_t db = umma_desc_sw128(b_sm + sb * (BN * 128));
            const uint32_t ta = tmem_a + (uint32_t) (sa * MT * 32);
            if (elect_one()) {
#pragma unroll
                for (int k = 0; k < GM_BK / 16; ++k) umma_f16_ts(tmem, ta + 8u * k, db + 2 * k, idesc, (uint32_t) ((ks | k) != 0));
                if (MT == 2) {
#pragma unroll
                    for (int k = 0; k < GM_BK / 16; ++k) umma_f16_ts(tmem + (uint32_t) BN, ta + 32u + 8u * k, db + 2 * k, idesc, (uint32_t) ((ks | k) != 0));
                }
                umma_commit(ea0 + 8 * sa);
                umma_commit(eb0 + 8 * sb);
            }
            __syncwarp();
            if (++sa == SA) { sa = 0; pa ^= 1u; }
            if (++sb == SB) { sb = 0; pb ^= 1u; }
        }
        if (elect_one()) umma_commit(tfull);
        __syncwarp();
    } else {
        // ===== raw weight producer: one TMA box per 128-row tile and 256-K block (rows past M are zero-filled) =====
        if (!(p.dbg & 1)) {
            int slot = 0; uint32_t eph = 1;
            for (int unit = 0; unit < nunits; ++unit) {
                mbar_wait(re0 + 8 * slot, eph);
                if (elect_one()) {
                    mbar_arrive_expect_tx(rf0 + 8 * slot, raw_slot);
                    for (int t = 0; t < MT; ++t) {
                        const uint32_t dstp = r_sm + slot * raw_slot + t * T::TILE_BYTES;
                        if (TYPE == B200_TYPE_Q8_0) {
                            tma_load_3d(dstp, &tmap_w, 0, (u0 + unit) * 17, wrow0 + t * GM_BM, rf0 + 8 * slot);
                        } else if (TYPE == B200_TYPE_Q6_K) {
                            tma_load_4d(dstp, &tmap_w, ((u0 + unit) & 7) * 16, 0, (u0 + unit) >> 3, wrow0 + t * GM_BM, rf0 + 8 * slot);
                            tma_load_3d(dstp + GM_BM * T::ROW_BYTES, &tmap_w2, 0, (u0 + unit) >> 3, wrow0 + t * GM_BM, rf0 + 8 * slot);
                        } else {
                            tma_load_2d(dstp, &tmap_w, (u0 + unit) * T::ROW_BYTES, wrow0 + t * GM_BM, rf0 + 8 * slot);
                        }
                    }
                }
                __syncwarp();
                if (++slot == RS) { slot = 0; eph ^= 1u; }
            }
        }
    }
    __syncthreads();
    if (warp == GM_PRODUCERS + 1) { tc_fence_after(); tmem_dealloc(tmem, tcols); }
}

// ---- host side -------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn b200_get_encode_tiled() {
    static encode_tiled_fn fn = nullptr; static int tried = 0;
    if (!tried) {
        tried = 1;
        void * ptr = nullptr; cudaDriverEntryPointQueryResult qres;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &ptr, cudaEnableDefault, &qres) == cudaSuccess && qres == cudaDriverEntryPointSuccess)
            fn = (encode_tiled_fn) ptr;
    }
    return fn;
}

bool b200_gemm_supported(const b200_tensor * w, const b200_tensor * x) {
    static const int off = getenv("GGML_B200_NO_GEMM") != nullptr;
    if (off) return false;
    const bool q4k = w->type == B200_TYPE_Q4_K && !(w->flags & B200_TENSOR_REPACKED);
    const bool q6k = w->type == B200_TYPE_Q6_K && (w->flags & B200_TENSOR_REPACKED) && w->ne[0] % 2048 == 0;   // group layout of b200_repack.cu
    const bool q40 = w->type == B200_TYPE_Q4_0 && (w->flags & B200_TENSOR_REPACKED), q80 = w->type == B200_TYPE_Q8_0 && (w->flags & B200_TENSOR_REPACKED);
    if (!q4k && !q6k && !q40 && !q80) return false;
    if (w->ne[2] != 1 || w->ne[3] != 1 || x->ne[2] != 1 || x->ne[3] != 1) return false;
    if (((uintptr_t) w->data & 15) || (w->nb[1] & 15) || ((uintptr_t) x->data & 15) || (x->nb[1] & 15)) return false;
    if (w->ne[0] % 256 != 0 || w->ne[0] > (1 << 20) || w->ne[1] > (1 << 30) || x->ne[1] > (1 << 30)) return false;
    return x->ne[1] > B200_GEMV_MAX_COLS;
}

// tile shape.  Measured on B200 (profiles/r1_gemm_sweep.txt): 128x256 tiles beat two stacked 128x128 tiles (MT=2) whenever they
// give at least ~0.7 waves; below that, narrower tiles that fill more SMs win.
// Cost model of one launch in microseconds, fitted to scripts/gemv_bench.py --n 512 (profiles/r1_gemm_split.txt): a 128 x 256 tile
// takes 24.4 us per 4096 of K at the MMA rate one SM sustains, a 128 x 128 tile half of that / 0.85; every CTA adds ~10 us of fill,
// drain and epilogue; a tile of the partial last wave is cut in two along K (gm_args::lin) when both halves fit into that wave,
// which halves its MMA time and adds ~10 us for the second half's read-add-store epilogue behind the first's.
static double gemm_cost(int sms, int64_t tiles, int BN, int64_t K, bool can_split) {
    const double work = 24.4 * (double) K / 4096.0 * (BN == 256 ? 1.0 : BN / 256.0 / 0.85), c = 10.0, e = 10.0;
    const int64_t full = tiles / sms, left = tiles - full * sms;
    double t = (double) full * (work + c);
    if (left) t += (can_split && 2 * left <= sms && work / 2 + e < work) ? work / 2 + c + e : work + c;
    return t;
}
static void gemm_pick_tile(const b200_ctx * ctx, int64_t mt, int64_t N, int64_t K, bool can_split, int & MT, int & BN) {
    MT = 1;
    BN = N > 64 ? 128 : N > 32 ? 64 : 32;
    if (N > 128 && gemm_cost(ctx->sm_count, mt * ((N + 255) / 256), 256, K, can_split) <= gemm_cost(ctx->sm_count, mt * ((N + 127) / 128), 128, K, can_split)) BN = 256;
}

// tensor maps over one weight tensor (rows_total rows; for MUL_MAT_ID the experts stacked)
static bool gemm_weight_maps(encode_tiled_fn encode, const b200_tensor * w, int64_t rows_total, CUtensorMap * tmap_w, CUtensorMap * tmap_w2) {
    const cuuint32_t e4[4] = { 1, 1, 1, 1 };
    const int64_t K = w->ne[0];
    CUresult cw;
    if (w->type == B200_TYPE_Q8_0) {
        // repacked q8_0: rows of 17 x 16 B per 256 weights
        const cuuint64_t wdim[3] = { 16, (cuuint64_t) (K / 256) * 17, (cuuint64_t) rows_total };
        const cuuint64_t wstr[2] = { 16, (cuuint64_t) w->nb[1] };
        const cuuint32_t wbox[3] = { 16, 17, GM_BM };
        cw = encode(tmap_w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, w->data, wdim, wstr, wbox, e4, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        *tmap_w2 = *tmap_w;
    } else if (w->type != B200_TYPE_Q6_K) {
        // q4_K and repacked q4_0: the weight rows as a 2-D byte tensor; box = 144 bytes (256 weights) of 128 consecutive rows
        const cuuint64_t wdim[2] = { (cuuint64_t) (K / 256) * 144, (cuuint64_t) rows_total };
        const cuuint64_t wstr[1] = { (cuuint64_t) w->nb[1] };
        const cuuint32_t wbox[2] = { 144, GM_BM };
        cw = encode(tmap_w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, w->data, wdim, wstr, wbox, e4, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        *tmap_w2 = *tmap_w;
    } else {
        // repacked q6_K: {byte in unit-row (128), unit-row (13), group (K/2048), weight row}; box = column b (16 B) of the 13 unit-rows
        const cuuint64_t wdim[4] = { 128, 13, (cuuint64_t) (K / 2048), (cuuint64_t) rows_total };
        const cuuint64_t wstr[3] = { 128, 1680, (cuuint64_t) w->nb[1] };
        const cuuint32_t wbox[4] = { 16, 13, 1, GM_BM };
        cw = encode(tmap_w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 4, w->data, wdim, wstr, wbox, e4, CU_TENSOR_MAP_INTERLEAVE_NONE,
                    CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
        // the eight fp16 d's of a group: 16 B at byte 1664 of every group
        const cuuint64_t ddim[3] = { 16, (cuuint64_t) (K / 2048), (cuuint64_t) rows_total };
        const cuuint64_t dstr[2] = { 1680, (cuuint64_t) w->nb[1] };
        const cuuint32_t dbox[3] = { 16, 1, GM_BM };
        if (cw == CUDA_SUCCESS)
            cw = encode(tmap_w2, CU_TENSOR_MAP_DATA_TYPE_UINT8, 3, (char *) w->data + 1664, ddim, dstr, dbox, e4, CU_TENSOR_MAP_INTERLEAVE_NONE,
                        CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_NONE, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    return cw == CUDA_SUCCESS;
}

// image = bf16 [N][K] activations (b200_launch_act_bf16); nseg weights of one type sharing them; grouped (MUL_MAT_ID) when gi != NULL
struct gm_group { const int * counts; const int * offsets; const long long * col_off; int n_expert; int bn; };
static int gemm_launch(b200_ctx * ctx, const b200_tensor * const * ws, int nseg, const void * image, int64_t N, float * const * dst, const size_t * d_ld, const gm_group * gi) {
    encode_tiled_fn encode = b200_get_encode_tiled();
    if (!encode) { b200_set_error("gemm: cuTensorMapEncodeTiled not available from the driver"); return B200_ERR_CUDA; }
    const b200_tensor * w = ws[0];
    gm_args p;
    p.counts = gi ? gi->counts : nullptr; p.offsets = gi ? gi->offsets : nullptr; p.col_off = gi ? gi->col_off : nullptr;
    p.N = (int) N; p.K = (int) w->ne[0]; p.nseg = nseg;
    int64_t Mtot = 0;
    for (int i = 0; i < nseg; ++i) Mtot += ws[i]->ne[1];
    static const int force_bn = getenv("GGML_B200_GEMM_BN") ? atoi(getenv("GGML_B200_GEMM_BN")) : 0;
    static const int force_mt = getenv("GGML_B200_GEMM_MT") ? atoi(getenv("GGML_B200_GEMM_MT")) : 0;
    // measured in the pp512 chain: 46.2 k tok/s with the split, 46.8 k without (the second half's epilogue eats the gain): opt-in
    static const int want_split = getenv("GGML_B200_GEMM_KSPLIT") && atoi(getenv("GGML_B200_GEMM_KSPLIT")) != 0;
    const bool can_split = !gi && want_split && (p.K / 256) >= 2;
    int64_t mt128 = 0;
    for (int i = 0; i < nseg; ++i) mt128 += (ws[i]->ne[1] + GM_BM - 1) / GM_BM;
    gemm_pick_tile(ctx, mt128, p.N, p.K, can_split, p.MT, p.BN);
    if (force_bn) p.BN = force_bn;
    if (force_mt) p.MT = force_mt;
    if (gi) { p.MT = 1; p.BN = gi->bn; }
    if (p.MT == 2 && p.BN > 128) p.BN = 128;                              // TMEM: 2*BN accumulator columns + 256 columns of A stages
    static const int force_s = getenv("GGML_B200_GEMM_S") ? atoi(getenv("GGML_B200_GEMM_S")) : 0, force_rs = getenv("GGML_B200_GEMM_RS") ? atoi(getenv("GGML_B200_GEMM_RS")) : 0;
    static const int force_sb = getenv("GGML_B200_GEMM_SB") ? atoi(getenv("GGML_B200_GEMM_SB")) : 0;
    // depth in TIME is what hides the TMA and producer latencies (~1 us): as many 64-K steps in flight as TMEM and smem allow
    p.SA = force_s ? force_s : (p.MT * p.BN + p.MT * 256 <= 512 ? 8 : 4);
    p.RS = force_rs ? force_rs : (p.MT == 2 ? 2 : 3);
    p.SB = force_sb;
    static const int dbg = getenv("GGML_B200_GEMM_DBG") ? atoi(getenv("GGML_B200_GEMM_DBG")) : 0;
    p.dbg = dbg;
    CUtensorMap tmap;
    const cuuint64_t gdim[2] = { (cuuint64_t) p.K, (cuuint64_t) p.N };
    const cuuint64_t gstr[1] = { (cuuint64_t) p.K * 2 };
    const cuuint32_t box[2]  = { GM_BK, (cuuint32_t) p.BN };
    const cuuint32_t estr[2] = { 1, 1 };
    const CUresult cr = encode(&tmap, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, (void *) image, gdim, gstr, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cr != CUDA_SUCCESS) { b200_set_error("gemm: cuTensorMapEncodeTiled failed (%d)", (int) cr); return B200_ERR_CUDA; }
    gm_maps maps;
    unsigned mtiles_total = 0;
    for (int i = 0; i < 3; ++i) {
        const b200_tensor * wi = ws[i < nseg ? i : 0];
        if (!gemm_weight_maps(encode, wi, wi->ne[1] * (gi ? gi->n_expert : 1), &maps.w[i], &maps.w2[i])) {
            b200_set_error("gemm: cuTensorMapEncodeTiled (weights) failed"); return B200_ERR_CUDA;
        }
        p.seg[i].dst = i < nseg ? dst[i] : nullptr; p.seg[i].d_ld = i < nseg ? d_ld[i] : 0;
        p.seg[i].M = i < nseg ? (int) wi->ne[1] : 0;
        p.seg[i].mtiles = i < nseg ? (int) ((wi->ne[1] + p.MT * GM_BM - 1) / (p.MT * GM_BM)) : 0;
        mtiles_total += (unsigned) p.seg[i].mtiles;
    }
    const bool q6 = w->type == B200_TYPE_Q6_K, q8 = w->type == B200_TYPE_Q8_0;
    const int tile_bytes = q6 ? gm_traits<B200_TYPE_Q6_K>::TILE_BYTES : q8 ? gm_traits<B200_TYPE_Q8_0>::TILE_BYTES : gm_traits<B200_TYPE_Q4_K>::TILE_BYTES;
    const size_t raw_bytes = (size_t) p.RS * p.MT * tile_bytes;
    if (!force_sb) { p.SB = (int) ((220 * 1024 - 2048 - raw_bytes) / ((size_t) p.BN * 128)); if (p.SB > 8) p.SB = 8; }
    if (p.SB < 2) { b200_set_error("gemm: tile does not fit in shared memory"); return B200_ERR_UNSUPPORTED; }
    const size_t smem = 1024 + (size_t) p.SB * (p.BN * 128) + raw_bytes + 8 * (2 * p.SA + 2 * p.SB + 2 * p.RS + 1) + 16;
    if (!ctx->attr_done[5]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemm_q<B200_TYPE_Q4_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemm_q<B200_TYPE_Q6_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemm_q<B200_TYPE_Q4_0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemm_q<B200_TYPE_Q8_0>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
        ctx->attr_done[5] = 1;
    }
    dim3 grid((unsigned) ((p.N + p.BN - 1) / p.BN), mtiles_total, (unsigned) (gi ? gi->n_expert : 1));
    p.lin = 0; p.ntn = (int) grid.x; p.n_unsplit = 0; p.flags = nullptr;
    if (!gi) {
        const int64_t tiles = (int64_t) grid.x * mtiles_total;
        const int64_t left = tiles % ctx->sm_count;
        const double work = 24.4 * (double) p.K / 4096.0 * (p.BN == 256 ? 1.0 : p.BN / 256.0 / 0.85);
        const int64_t nsplit = (can_split && left && 2 * left <= ctx->sm_count && left <= 1024 && work / 2 + 10.0 < work) ? left : 0;
        if (nsplit) {
            const bool fresh = ctx->scratch[6] == nullptr;
            p.flags = (unsigned *) b200_scratch(ctx, 6, 1025 * sizeof(unsigned));          // fixed size, zeroed once, re-armed by the kernel; [1024]: a wait ran out
            if (!p.flags) return B200_ERR_ALLOC;
            if (fresh) B200_CUDA_TRY(cudaMemsetAsync(p.flags, 0, 1025 * sizeof(unsigned), ctx->stream));
        }
        p.lin = 1; p.n_unsplit = (int) (tiles - nsplit);
        grid = dim3((unsigned) (tiles + nsplit), 1, 1);
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(GM_THREADS); cfg.dynamicSmemBytes = smem; cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    static const int no_pdl = getenv("GGML_B200_GEMM_NO_PDL") != nullptr;
    cfg.attrs = attr; cfg.numAttrs = (ctx->pdl && !gi && !no_pdl) ? 1 : 0;    // grouped: the bucket tables are read before anything else
    switch (w->type) {
        case B200_TYPE_Q6_K: B200_LAUNCH(ctx, B200_K_GEMM, cudaLaunchKernelEx(&cfg, k_gemm_q<B200_TYPE_Q6_K>, p, tmap, maps)); break;
        case B200_TYPE_Q8_0: B200_LAUNCH(ctx, B200_K_GEMM, cudaLaunchKernelEx(&cfg, k_gemm_q<B200_TYPE_Q8_0>, p, tmap, maps)); break;
        case B200_TYPE_Q4_0: B200_LAUNCH(ctx, B200_K_GEMM, cudaLaunchKernelEx(&cfg, k_gemm_q<B200_TYPE_Q4_0>, p, tmap, maps)); break;
        default:             B200_LAUNCH(ctx, B200_K_GEMM, cudaLaunchKernelEx(&cfg, k_gemm_q<B200_TYPE_Q4_K>, p, tmap, maps)); break;
    }
    return B200_OK;
}

int b200_launch_gemm_q(b200_ctx * ctx, const b200_tensor * w, const void * image, int64_t N, b200_tensor * dst) {
    float * d = (float *) dst->data; const size_t ld = dst->nb[1] / 4;
    return gemm_launch(ctx, &w, 1, image, N, &d, &ld, nullptr);
}
// n <= 3 weights of one type and K sharing the activation image: one launch
int b200_launch_gemm_multi(b200_ctx * ctx, const b200_tensor * const * w, int n, const void * image, int64_t N, b200_tensor * const * dst) {
    float * d[3]; size_t ld[3];
    for (int i = 0; i < n; ++i) { d[i] = (float *) dst[i]->data; ld[i] = dst[i]->nb[1] / 4; }
    return gemm_launch(ctx, w, n, image, N, d, ld, nullptr);
}

// ---- MUL_MAT_ID as a grouped GEMM -----------------------------------------------------------------------------------------------
// Reference: ggml_compute_forward_mul_mat_id (ggml/src/ggml-cpu/ggml-cpu.c:1500-1700) buckets the (slot, token) pairs by expert
// on thread 0 and runs one matmul per bucket.  Here: one small kernel buckets the pairs on the device, a gather kernel writes
// the bf16 activation rows in bucket order, and ONE launch of the GEMM serves every expert (grid.z = expert; CTAs beyond an
// expert's row count exit at once); the epilogue scatters result columns to dst[:, slot, token].
__global__ void __launch_bounds__(1024) k_mmid_bucket(const char * ids, size_t ids_nb1, int n_used, int n_tokens, int n_expert, int ne11,
                                                       size_t b_nb1, size_t b_nb2, long long d_ld1, long long d_ld2,
                                                       int * counts, int * offsets, long long * col_off, long long * src_off) {
    extern __shared__ int sm[];                                             // cnt[n_expert], cur[n_expert]
    int * cnt = sm, * cur = sm + n_expert;
    for (int i = threadIdx.x; i < 2 * n_expert; i += blockDim.x) sm[i] = 0;
    __syncthreads();
    const int npairs = n_used * n_tokens;
    for (int i = threadIdx.x; i < npairs; i += blockDim.x) {
        const int t = i / n_used, s = i - t * n_used;
        const int e = ((const int *) (ids + (size_t) t * ids_nb1))[s];
        if (e >= 0 && e < n_expert) atomicAdd(&cnt[e], 1);
    }
    __syncthreads();
    if (threadIdx.x == 0) { int acc = 0; for (int e = 0; e < n_expert; ++e) { offsets[e] = acc; counts[e] = cnt[e]; cur[e] = acc; acc += cnt[e]; } }
    __syncthreads();
    for (int i = threadIdx.x; i < npairs; i += blockDim.x) {
        const int t = i / n_used, s = i - t * n_used;
        const int e = ((const int *) (ids + (size_t) t * ids_nb1))[s];
        if (e < 0 || e >= n_expert) continue;
        const int pos = atomicAdd(&cur[e], 1);
        col_off[pos] = (long long) t * d_ld2 + (long long) s * d_ld1;
        src_off[pos] = (long long) t * (long long) b_nb2 + (long long) (s % ne11) * (long long) b_nb1;
    }
}
__global__ void __launch_bounds__(256) k_act_bf16_gather(const char * __restrict__ b, const long long * __restrict__ src_off, int64_t K, int64_t nrows,
                                                          __nv_bfloat16 * __restrict__ out) {
    const int64_t per_row = K / 4, total = per_row * nrows;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t n = i / per_row, k = (i - n * per_row) * 4;
        const float4 v = *(const float4 *) (b + src_off[n] + k * 4);
        __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), c = __floats2bfloat162_rn(v.z, v.w);
        uint2 o; o.x = *(uint32_t *) &a; o.y = *(uint32_t *) &c;
        *(uint2 *) (out + n * K + k) = o;
    }
}

bool b200_gemm_id_supported(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids) {
    b200_tensor w2 = *as; w2.ne[2] = 1;                                      // one expert's matrix, for the type / alignment rules
    b200_tensor x2 = *b;  x2.ne[1] = 64; x2.ne[2] = 1; x2.nb[1] = (size_t) b->ne[0] * 4;
    if (!b200_gemm_supported(&w2, &x2)) return false;
    if (as->nb[2] != as->nb[1] * (size_t) as->ne[1] || as->ne[2] > 256 || as->ne[1] * as->ne[2] > (1ll << 30)) return false;   // experts densely stacked
    if (((uintptr_t) b->data & 15) || (b->nb[1] & 15) || (b->nb[2] & 15) || (ids->nb[1] & 3)) return false;
    return ids->ne[0] * ids->ne[1] > 32 && ids->ne[0] * ids->ne[1] <= (1 << 22);
}

int b200_launch_gemm_id_q(b200_ctx * ctx, const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, b200_tensor * dst) {
    const int n_used = (int) ids->ne[0], n_tokens = (int) ids->ne[1], n_expert = (int) as->ne[2], npairs = n_used * n_tokens;
    const int64_t K = as->ne[0];
    // scratch slot 3: counts | offsets | col_off | src_off ; slot 0: the bucket-ordered bf16 activation image
    const size_t ints = ((size_t) 2 * n_expert * sizeof(int) + 15) & ~(size_t) 15;
    char * meta = (char *) b200_scratch(ctx, 3, ints + (size_t) 2 * npairs * sizeof(long long));
    ctx->qcache.valid = 0;
    void * image = b200_scratch(ctx, 0, (size_t) npairs * K * 2);
    if (!meta || !image) return B200_ERR_ALLOC;
    int * counts = (int *) meta, * offsets = counts + n_expert;
    long long * col_off = (long long *) (meta + ints), * src_off = col_off + npairs;
    B200_LAUNCH(ctx, B200_K_MMID_SORT, k_mmid_bucket<<<1, 1024, 2 * n_expert * sizeof(int), ctx->stream>>>((const char *) ids->data, ids->nb[1], n_used, n_tokens,
        n_expert, (int) b->ne[1], b->nb[1], b->nb[2], (long long) (dst->nb[1] / 4), (long long) (dst->nb[2] / 4), counts, offsets, col_off, src_off));
    const int64_t total = K / 4 * npairs;
    const unsigned ggrid = (unsigned) ((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    B200_LAUNCH(ctx, B200_K_CONVERT, k_act_bf16_gather<<<ggrid, 256, 0, ctx->stream>>>((const char *) b->data, src_off, K, npairs, (__nv_bfloat16 *) image));
    gm_group gi;
    gi.counts = counts; gi.offsets = offsets; gi.col_off = col_off; gi.n_expert = n_expert;
    const int avg = npairs / (n_expert > 0 ? n_expert : 1);
    gi.bn = avg > 96 ? 128 : avg > 48 ? 64 : 32;
    b200_tensor w2 = *as; w2.ne[2] = 1;
    const b200_tensor * wp = &w2; float * dd = (float *) dst->data; const size_t ld0 = 0;
    return gemm_launch(ctx, &wp, 1, image, npairs, &dd, &ld0, &gi);
}
