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
// CTA = 128 weight rows x BN activation columns.  The A operand (weights) never touches shared memory as bf16: an SS-mode
// M=128 MMA reads A and B from shared memory at the full 128 B/clk the SM has, which measured as the limiter of the first
// version of this kernel (profiles/r1_gemm_v1_*).  Here A is fed from TMEM.  Warp roles:
//   warps 0-3  A producers: thread = weight row = TMEM lane.  Raw GGUF blocks of the 128-row tile arrive in a shared ring by one
//              TMA tiled load per 256-K block (a 2-D byte tensor map over the weight rows); the thread expands 64 K-values at a
//              time to bf16 (value = d*sc*q - dmin*m, two values per HFMA2.BF16) and writes them with one tcgen05.st (32 columns
//              of its lane).  After the K loop the same warps run the epilogue: tcgen05.ld the accumulator lanes and store fp32
//              columns (coalesced along M).
//   warp 4     B producer: one thread, TMA tiled loads (SWIZZLE_128B) of the bf16 activation image [N][K].
//   warp 5     TMEM owner + MMA issuer: one thread, four tcgen05.mma (A in TMEM, K=16 each) per stage; tcgen05.commit frees it.
//   warp 6     raw-weight producer: one thread, the TMA loads of the GGUF block ring.
#include "b200_common.cuh"
#include "b200_ptx.cuh"
#include <cuda.h>
#include <stdlib.h>

#define GM_BM      128
#define GM_BK      64
#define GM_THREADS 224
#define GM_STAGES  4          // A (TMEM) / B (smem) pipeline depth in 64-K steps
#define GM_RAW_STAGES 4      // ring of raw 256-K weight blocks

struct gm_args {
    const uint8_t * w; size_t w_nb1;
    float * dst; size_t d_ld;            // dst[m + n * d_ld]
    int M, N, K;
    int BN;
    int dbg;                             // timing experiments only (GGML_B200_GEMM_DBG): 1 = no weight fetch, 2 = no expansion math
};

// ---- activations f32 [K, N] -> bf16 image [N][K] ---------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_act_bf16(const float * __restrict__ x, size_t nb1, int64_t K, int64_t N, __nv_bfloat16 * __restrict__ out) {
    const int64_t per_row = K / 4, total = per_row * N;
    for (int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x; i < total; i += (int64_t) gridDim.x * blockDim.x) {
        const int64_t n = i / per_row, k = (i - n * per_row) * 4;
        const float4 v = *(const float4 *) ((const char *) x + n * nb1 + k * 4);
        __nv_bfloat162 a = __floats2bfloat162_rn(v.x, v.y), b = __floats2bfloat162_rn(v.z, v.w);
        uint2 o; o.x = *(uint32_t *) &a; o.y = *(uint32_t *) &b;
        *(uint2 *) (out + n * K + k) = o;
    }
}
int b200_launch_act_bf16(b200_ctx * ctx, const float * x, size_t nb1, int64_t K, int64_t N, void * image) {
    const int64_t total = K / 4 * N;
    const unsigned grid = (unsigned) ((total + 255) / 256 < 148 * 8 ? (total + 255) / 256 : 148 * 8);
    B200_LAUNCH(ctx, B200_K_CONVERT, k_act_bf16<<<grid, 256, 0, ctx->stream>>>(x, nb1, K, N, (__nv_bfloat16 *) image));
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
    static constexpr int RAW_K = 256, RAW_BYTES = 144;
    struct unit { float d, dmin; uint32_t sc_lo, sc_hi, m_lo, m_hi; };   // 8 x 6-bit scales / mins, one per byte
    static __device__ __forceinline__ void load_unit(uint32_t raw, unit & u) {
        const int4 h = lds128(raw);                                          // d, dmin, scales[12]
        const float2 dm = __half22float2(*(const __half2 *) &h.x);
        u.d = dm.x; u.dmin = dm.y;
        const uint32_t u0 = h.y, u1 = h.z, u2 = h.w;                         // get_scale_min_k4, ggml-quants.c:703-710
        u.sc_lo = u0 & 0x3f3f3f3fu; u.m_lo = u1 & 0x3f3f3f3fu;
        u.sc_hi = (u2 & 0x0f0f0f0fu) | (((u0 >> 6) & 0x03030303u) << 4);
        u.m_hi  = ((u2 >> 4) & 0x0f0f0f0fu) | (((u1 >> 6) & 0x03030303u) << 4);
    }
    // K-chunk g (0..3) of the block: sub-block 2g = low nibbles of qs[32g..32g+31], sub-block 2g+1 = high nibbles (ggml-quants.c:1352-1374)
    template <int G> static __device__ __forceinline__ void expand(uint32_t raw, const unit & u, uint32_t (&out)[32]) {
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

template <int TYPE>
__global__ void __launch_bounds__(GM_THREADS, 1) k_gemm_q(const gm_args p, const __grid_constant__ CUtensorMap tmap_x, const __grid_constant__ CUtensorMap tmap_w) {
    using T = gm_traits<TYPE>;
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base  = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t b_sm  = base;                                             // [GM_STAGES][BN rows][128 B], 128B-swizzled by TMA
    const uint32_t r_sm  = b_sm + GM_STAGES * (p.BN * 128);                  // [GM_RAW_STAGES][128 rows][RAW_BYTES]
    const uint32_t bars  = r_sm + GM_RAW_STAGES * GM_BM * T::RAW_BYTES;
    const uint32_t fa0 = bars, fb0 = fa0 + 8 * GM_STAGES, empty0 = fb0 + 8 * GM_STAGES;   // A written / B landed / stage consumed
    const uint32_t rf0 = empty0 + 8 * GM_STAGES, re0 = rf0 + 8 * GM_RAW_STAGES;           // raw block landed / raw block consumed
    const uint32_t tfull = re0 + 8 * GM_RAW_STAGES, tslot = tfull + 8;
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int n0 = blockIdx.x * p.BN, m0 = blockIdx.y * GM_BM;
    const int nks = p.K / GM_BK, nunits = p.K / T::RAW_K;
    const uint32_t tcols = p.BN + 32 * GM_STAGES <= 256 ? 256u : 512u;       // accumulator [0, BN) + A stages [BN, BN + 32*GM_STAGES)

    if (threadIdx.x == 0) {
        for (int s = 0; s < GM_STAGES; ++s) { mbar_init(fa0 + 8 * s, 4); mbar_init(fb0 + 8 * s, 1); mbar_init(empty0 + 8 * s, 1); }
        for (int s = 0; s < GM_RAW_STAGES; ++s) { mbar_init(rf0 + 8 * s, 1); mbar_init(re0 + 8 * s, 4); }
        mbar_init(tfull, 1);
        mbar_fence_init();
    }
    if (warp == 5) tmem_alloc(tslot, tcols);
    if (warp == 4 && lane == 0) tma_prefetch_desc(&tmap_x);
    if (warp == 6 && lane == 0) tma_prefetch_desc(&tmap_w);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tslot));
    const uint32_t tmem_a = tmem + (uint32_t) p.BN;

    if (warp < 4) {
        // ===== A producers =====
        const int row = warp * 32 + lane;
        const uint32_t lane_base = (uint32_t) (warp * 32) << 16;
        int stage = 0; uint32_t eph = 1;                                     // parity to wait for on empty[stage] (1: a fresh barrier passes)
        int slot = 0; uint32_t rph = 0;
        auto emit = [&](const uint32_t (&o)[32]) {
            mbar_wait(empty0 + 8 * stage, eph);
            tc_fence_after();
            tmem_st32(tmem_a + lane_base + 32u * stage, o);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(fa0 + 8 * stage);
            if (++stage == GM_STAGES) { stage = 0; eph ^= 1u; }
        };
        for (int unit = 0; unit < nunits; ++unit) {
            if (!(p.dbg & 1)) mbar_wait(rf0 + 8 * slot, rph);
            const uint32_t raw = r_sm + (slot * GM_BM + row) * T::RAW_BYTES;
            typename T::unit u; T::load_unit(raw, u);
            uint32_t o[32];
            if (p.dbg & 2) { for (int c = 0; c < 32; ++c) o[c] = u.sc_lo; emit(o); emit(o); emit(o); emit(o); }
            else {
            T::template expand<0>(raw, u, o); emit(o);
            T::template expand<1>(raw, u, o); emit(o);
            T::template expand<2>(raw, u, o); emit(o);
            T::template expand<3>(raw, u, o); emit(o);
            }
            __syncwarp();
            if (lane == 0) mbar_arrive(re0 + 8 * slot);
            if (++slot == GM_RAW_STAGES) { slot = 0; rph ^= 1u; }
        }
        // ===== epilogue: TMEM -> registers -> dst (columns of dst are contiguous along M) =====
        mbar_wait(tfull, 0);
        tc_fence_after();
        const int mm = m0 + row;
        float * drow = p.dst + mm;
        for (int c0 = 0; c0 < p.BN; c0 += 16) {
            if (n0 + c0 >= p.N) break;
            uint32_t r[16];
            tmem_ld16(tmem + lane_base + (uint32_t) c0, r);
            tmem_ld_wait();
            if (mm < p.M) {
#pragma unroll
                for (int j = 0; j < 16; ++j) if (n0 + c0 + j < p.N) drow[(size_t) (n0 + c0 + j) * p.d_ld] = __uint_as_float(r[j]);
            }
        }
        tc_fence_before();
    } else if (warp == 4) {
        // ===== B producer =====
        if (lane == 0) {
            int stage = 0; uint32_t eph = 1;
            for (int ks = 0; ks < nks; ++ks) {
                mbar_wait(empty0 + 8 * stage, eph);
                mbar_arrive_expect_tx(fb0 + 8 * stage, (uint32_t) p.BN * 128u);
                tma_load_2d(b_sm + stage * (p.BN * 128), &tmap_x, ks * GM_BK, n0, fb0 + 8 * stage);
                if (++stage == GM_STAGES) { stage = 0; eph ^= 1u; }
            }
        }
    } else if (warp == 5) {
        // ===== MMA issuer =====
        if (lane == 0) {
            const uint32_t idesc = umma_idesc_bf16(GM_BM, p.BN);
            int stage = 0; uint32_t fph = 0;
            for (int ks = 0; ks < nks; ++ks) {
                mbar_wait(fa0 + 8 * stage, fph);
                mbar_wait(fb0 + 8 * stage, fph);
                tc_fence_after();
                const uint64_t db = umma_desc_sw128(b_sm + stage * (p.BN * 128));
#pragma unroll
                for (int k = 0; k < GM_BK / 16; ++k) umma_f16_ts(tmem, tmem_a + 32u * stage + 8u * k, db + 2 * k, idesc, (uint32_t) ((ks | k) != 0));
                umma_commit(empty0 + 8 * stage);
                if (++stage == GM_STAGES) { stage = 0; fph ^= 1u; }
            }
            umma_commit(tfull);
        }
        __syncwarp();
    } else {
        // ===== raw weight producer: one 128-row x RAW_BYTES box per 256-K block (rows past M are zero-filled) =====
        if (lane == 0 && !(p.dbg & 1)) {
            int slot = 0; uint32_t eph = 1;
            for (int unit = 0; unit < nunits; ++unit) {
                mbar_wait(re0 + 8 * slot, eph);
                mbar_arrive_expect_tx(rf0 + 8 * slot, (uint32_t) (GM_BM * T::RAW_BYTES));
                tma_load_2d(r_sm + slot * (GM_BM * T::RAW_BYTES), &tmap_w, unit * T::RAW_BYTES, m0, rf0 + 8 * slot);
                if (++slot == GM_RAW_STAGES) { slot = 0; eph ^= 1u; }
            }
        }
    }
    __syncthreads();
    if (warp == 5) { tc_fence_after(); tmem_dealloc(tmem, tcols); }
}

// ---- host side -------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
static encode_tiled_fn get_encode_tiled() {
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
    if (w->type != B200_TYPE_Q4_K || (w->flags & B200_TENSOR_REPACKED)) return false;
    if (w->ne[2] != 1 || w->ne[3] != 1 || x->ne[2] != 1 || x->ne[3] != 1) return false;
    if (((uintptr_t) w->data & 15) || (w->nb[1] & 15) || ((uintptr_t) x->data & 15) || (x->nb[1] & 15)) return false;
    if (w->ne[0] % 256 != 0 || w->ne[0] > (1 << 20) || w->ne[1] > (1 << 30) || x->ne[1] > (1 << 30)) return false;
    return x->ne[1] > B200_GEMV_MAX_COLS;
}

static int gemm_pick_bn(const b200_ctx * ctx, int64_t M, int64_t N) {
    const int64_t mt = (M + GM_BM - 1) / GM_BM;
    if (N > 128 && mt * ((N + 255) / 256) >= (int64_t) ctx->sm_count * 3 / 4) return 256;
    if (N > 64) return 128;
    if (N > 32) return 64;
    return 32;
}

// image = bf16 [N][K] activations (b200_launch_act_bf16)
int b200_launch_gemm_q(b200_ctx * ctx, const b200_tensor * w, const void * image, int64_t N, b200_tensor * dst) {
    encode_tiled_fn encode = get_encode_tiled();
    if (!encode) { b200_set_error("gemm: cuTensorMapEncodeTiled not available from the driver"); return B200_ERR_CUDA; }
    gm_args p;
    p.w = (const uint8_t *) w->data; p.w_nb1 = w->nb[1];
    p.dst = (float *) dst->data; p.d_ld = dst->nb[1] / 4;
    p.M = (int) w->ne[1]; p.N = (int) N; p.K = (int) w->ne[0];
    static const int force_bn = getenv("GGML_B200_GEMM_BN") ? atoi(getenv("GGML_B200_GEMM_BN")) : 0;
    p.BN = force_bn ? force_bn : gemm_pick_bn(ctx, p.M, p.N);
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
    using T = gm_traits<B200_TYPE_Q4_K>;
    // the weight rows as a 2-D byte tensor: box = one 256-K block of 128 consecutive rows
    CUtensorMap tmap_w;
    const cuuint64_t wdim[2] = { (cuuint64_t) (p.K / T::RAW_K) * T::RAW_BYTES, (cuuint64_t) p.M };
    const cuuint64_t wstr[1] = { (cuuint64_t) p.w_nb1 };
    const cuuint32_t wbox[2] = { T::RAW_BYTES, GM_BM };
    const CUresult cw = encode(&tmap_w, CU_TENSOR_MAP_DATA_TYPE_UINT8, 2, (void *) p.w, wdim, wstr, wbox, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (cw != CUDA_SUCCESS) { b200_set_error("gemm: cuTensorMapEncodeTiled (weights) failed (%d)", (int) cw); return B200_ERR_CUDA; }
    const size_t smem = 1024 + (size_t) GM_STAGES * (p.BN * 128) + GM_RAW_STAGES * GM_BM * T::RAW_BYTES + 8 * (3 * GM_STAGES + 2 * GM_RAW_STAGES + 1) + 16;
    if (!ctx->attr_done[5]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemm_q<B200_TYPE_Q4_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 225 * 1024));
        ctx->attr_done[5] = 1;
    }
    dim3 grid((unsigned) ((p.N + p.BN - 1) / p.BN), (unsigned) ((p.M + GM_BM - 1) / GM_BM));
    B200_LAUNCH(ctx, B200_K_GEMM, k_gemm_q<B200_TYPE_Q4_K><<<grid, GM_THREADS, smem, ctx->stream>>>(p, tmap, tmap_w));
    return B200_OK;
}

bool b200_gemm_id_supported(const b200_tensor *, const b200_tensor *, const b200_tensor *) { return false; }
int  b200_launch_gemm_id_q(b200_ctx *, const b200_tensor *, const b200_tensor *, const b200_tensor *, b200_tensor *) { return B200_ERR_UNSUPPORTED; }
