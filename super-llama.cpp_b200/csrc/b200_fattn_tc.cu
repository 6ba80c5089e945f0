// b200_fattn_tc.cu -- GGML_OP_FLASH_ATTN_EXT for query batches (prefill): tcgen05 tensor cores, TMEM accumulators, TMA.
//
// Semantics: ggml_compute_forward_flash_attn_ext_f16_one_chunk (reference ggml/src/ggml-cpu/ops.cpp:8047-8272); like the
// reference the query row is rounded to fp16 (:8117-8118,8162) and K.q accumulates in fp32; P is rounded to fp16 before P.V
// (the reference accumulates V in fp16, :8193-8207; we keep the fp32 accumulator).
// Bound: tensor pipe, 2*n_head*n_batch*(DK+DV)*n_kv flops per layer (tests/test-backend-ops.cpp:6084-6089).
//
// CTA = 128 query rows of one head.  KV is walked in tiles of 128 keys; tiles whose mask is entirely -inf (causal future,
// unused cache cells) are skipped by every role, tiles whose mask is entirely zero skip the mask loads (k_mask_flags scans the
// mask once per mask tensor; the result is cached in the context until the mask is written again).
//   warp 4   TMA: K and V tiles straight from the strided KV-cache view (4-D tensor maps, SWIZZLE_128B), two stages each.
//   warp 5   MMA: S = Q.K^T (both K-major in smem) into a double-buffered TMEM tile; O += P.V with P read from TMEM and V
//            as an MN-major B operand -- the cache row layout [token][head_dim] is used as is, no transpose anywhere.
//   warps 0-3  softmax: thread = query row = TMEM lane.  tcgen05.ld the 128 scores, scale/softcap/mask, online max and sum,
//            rescale O in TMEM when the max moved, write P (fp16) back to TMEM with tcgen05.st.  Finally O/l -> dst.
#include "b200_fattn.cuh"
#include "b200_ptx.cuh"
#include <cuda.h>
#include <stdlib.h>

#define FT_BQ      128
#define FT_BKV     128
#define FT_THREADS 192

struct ft_extra {
    const uint8_t * flags;      // [m_ne3][m_ne2][nqt][nkt]: 0 = skip tile, 1 = mask all zero, 2 = mixed
    int nqt, nkt;
    int k_swap, v_swap;         // tensor-map dimension order: 0 = (d, token, head, s), 1 = (d, head, token, s)
};

// ---- mask scan ---------------------------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(128) k_mask_flags(const char * mask, size_t nb1, size_t nb2, size_t nb3, int64_t KV, int64_t NQ,
                                                     int nqt, int nkt, uint8_t * flags) {
    const int kt = blockIdx.x, qt = blockIdx.y, z = blockIdx.z;            // z = i3 * m_ne2 + i2 (nb2/nb3 applied by the caller's grid)
    const int64_t q = (int64_t) qt * FT_BQ + threadIdx.x;
    int fin = 0, nz = 0;
    if (q < NQ) {
        const __half * row = (const __half *) (mask + q * nb1 + (size_t) z * nb2);
        (void) nb3;
        const int64_t k0 = (int64_t) kt * FT_BKV, k1 = k0 + FT_BKV < KV ? k0 + FT_BKV : KV;
        for (int64_t k = k0; k < k1; ++k) {
            const unsigned short b = __half_as_ushort(row[k]);
            fin |= b != 0xFC00u;                                           // anything but -inf keeps the key alive
            nz  |= (b & 0x7FFFu) != 0;                                     // anything but +-0 needs the add
        }
    }
    fin = __syncthreads_or(fin); nz = __syncthreads_or(nz);
    if (threadIdx.x == 0) flags[((size_t) z * nqt + qt) * nkt + kt] = fin ? (nz || (int64_t) (kt + 1) * FT_BKV > KV ? 2 : 1) : 0;
}

// ---- the attention kernel ------------------------------------------------------------------------------------------------------
template <int D>
__global__ void __launch_bounds__(FT_THREADS, 1) k_fattn_tc(const fattn_args p, const ft_extra x,
                                                             const __grid_constant__ CUtensorMap tmap_k, const __grid_constant__ CUtensorMap tmap_v) {
    constexpr int HALVES = D / 64, TILE = HALVES * 16384;                    // one [128 rows][D] fp16 operand tile
    extern __shared__ __align__(1024) uint8_t smem_raw[];
    const uint32_t base = (smem_u32(smem_raw) + 1023u) & ~1023u;
    const uint32_t q_sm = base, k_sm = q_sm + TILE, v_sm = k_sm + 2 * TILE;
    const uint32_t bars = v_sm + 2 * TILE;
    const uint32_t k_full = bars, k_empty = bars + 16, v_full = bars + 32, v_empty = bars + 48;   // [2] each
    const uint32_t s_full = bars + 64, s_empty = bars + 80, p_full = bars + 96, o_done = bars + 104, tslot = bars + 112;
    uint8_t * fl_sm = smem_raw + (bars + 128 - smem_u32(smem_raw));          // live-tile flags of this query tile
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int qt = blockIdx.x; const int64_t iq2 = blockIdx.y, iq3 = blockIdx.z;
    const int64_t q0 = (int64_t) qt * FT_BQ;

    if (threadIdx.x == 0) {
        for (int s = 0; s < 2; ++s) {
            mbar_init(k_full + 8 * s, 1); mbar_init(k_empty + 8 * s, 1); mbar_init(v_full + 8 * s, 1); mbar_init(v_empty + 8 * s, 1);
            mbar_init(s_full + 8 * s, 1); mbar_init(s_empty + 8 * s, 4);
        }
        mbar_init(p_full, 4); mbar_init(o_done, 1);
        mbar_fence_init();
    }
    if (warp == 5) tmem_alloc(tslot, 512);
    if (warp == 4 && lane == 0) { tma_prefetch_desc(&tmap_k); tma_prefetch_desc(&tmap_v); }
    // tile flags -> smem
    {
        const uint8_t * fl = x.flags ? x.flags + (((size_t) (iq3 % p.m_ne3) * p.m_ne2 + (iq2 % p.m_ne2)) * x.nqt + qt) * x.nkt : nullptr;
        for (int i = threadIdx.x; i < x.nkt; i += FT_THREADS) fl_sm[i] = fl ? fl[i] : ((int64_t) (i + 1) * FT_BKV > p.KV ? 2 : 1);
    }
    // Q tile: f32 -> fp16, K-major 128B-swizzled (rows past N are zero)
    {
        const char * qb = p.q + iq2 * p.q_nb2 + iq3 * p.q_nb3;
        for (int idx = threadIdx.x; idx < FT_BQ * (D / 8); idx += FT_THREADS) {
            const int r = idx / (D / 8), c = idx % (D / 8);
            uint4 o = make_uint4(0, 0, 0, 0);
            if (q0 + r < p.N) {
                const float4 a = *(const float4 *) (qb + (q0 + r) * p.q_nb1 + c * 32), b = *(const float4 *) (qb + (q0 + r) * p.q_nb1 + c * 32 + 16);
                __half2 h0 = __floats2half2_rn(a.x, a.y), h1 = __floats2half2_rn(a.z, a.w), h2 = __floats2half2_rn(b.x, b.y), h3 = __floats2half2_rn(b.z, b.w);
                o = make_uint4(*(uint32_t *) &h0, *(uint32_t *) &h1, *(uint32_t *) &h2, *(uint32_t *) &h3);
            }
            sts128(q_sm + (c >> 3) * 16384 + (r >> 3) * 1024 + (r & 7) * 128 + (((c & 7) ^ (r & 7)) << 4), make_int4((int) o.x, (int) o.y, (int) o.z, (int) o.w));
        }
        fence_proxy_async();
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    // the kernels behind us (activation conversion, then the wo GEMM) are programmatic dependents: the GEMM's weight side may start on SMs
    // our last wave leaves idle; everything that reads our output waits for this whole grid first
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    uint32_t tmem;
    asm volatile("ld.shared.u32 %0, [%1];" : "=r"(tmem) : "r"(tslot));
    const uint32_t t_s = tmem, t_o = tmem + 256, t_p = tmem + 384;          // S[2][128] | O[<=128] | P[64]

    if (warp == 4) {
        // ===== TMA producer =====
        const int hk = (int) (iq2 / p.rk2), sk = (int) (iq3 / p.rk3), hv = (int) (iq2 / p.rv2), sv = (int) (iq3 / p.rv3);
        int cnt = 0;
        for (int kt = 0; kt < x.nkt; ++kt) {
            if (fl_sm[kt] == 0) continue;
            const int st = cnt & 1; const uint32_t ph = (((uint32_t) cnt >> 1) & 1u) ^ 1u;
            mbar_wait(k_empty + 8 * st, ph);
            if (elect_one()) {
                mbar_arrive_expect_tx(k_full + 8 * st, TILE);
                for (int h = 0; h < HALVES; ++h)
                    tma_load_4d(k_sm + st * TILE + h * 16384, &tmap_k, h * 64, x.k_swap ? hk : kt * FT_BKV, x.k_swap ? kt * FT_BKV : hk, sk, k_full + 8 * st);
            }
            __syncwarp();
            mbar_wait(v_empty + 8 * st, ph);
            if (elect_one()) {
                mbar_arrive_expect_tx(v_full + 8 * st, TILE);
                for (int h = 0; h < HALVES; ++h)
                    tma_load_4d(v_sm + st * TILE + h * 16384, &tmap_v, h * 64, x.v_swap ? hv : kt * FT_BKV, x.v_swap ? kt * FT_BKV : hv, sv, v_full + 8 * st);
            }
            __syncwarp();
            ++cnt;
        }
    } else if (warp == 5) {
        // ===== MMA issuer: QK(j) is issued before PV(j-1) so the softmax of tile j overlaps PV(j-1) =====
        const uint32_t id_qk = umma_idesc_16(FT_BQ, FT_BKV, 0, 0), id_pv = umma_idesc_16(FT_BQ, D, 0, 1);
        int cnt = 0;
        auto pv = [&](int c) {                                             // O (+)= P(c) . V(c)
            const int st = c & 1;
            mbar_wait(p_full, (uint32_t) c & 1u);
            mbar_wait(v_full + 8 * st, ((uint32_t) c >> 1) & 1u);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < FT_BKV / 16; ++kk)
                    umma_f16_ts(t_o, t_p + 8u * kk, umma_desc_sw128_mn(v_sm + st * TILE + kk * 2048, 16384, 1024), id_pv, (uint32_t) (c > 0 || kk > 0));
                umma_commit(o_done);
                umma_commit(v_empty + 8 * st);
            }
            __syncwarp();
        };
        for (int kt = 0; kt < x.nkt; ++kt) {
            if (fl_sm[kt] == 0) continue;
            const int st = cnt & 1;
            mbar_wait(s_empty + 8 * st, (((uint32_t) cnt >> 1) & 1u) ^ 1u);
            mbar_wait(k_full + 8 * st, ((uint32_t) cnt >> 1) & 1u);
            tc_fence_after();
            if (elect_one()) {
#pragma unroll
                for (int kk = 0; kk < D / 16; ++kk)
                    umma_f16(t_s + (uint32_t) (st * FT_BKV), umma_desc_sw128(q_sm + (kk >> 2) * 16384) + 2 * (kk & 3),
                             umma_desc_sw128(k_sm + st * TILE + (kk >> 2) * 16384) + 2 * (kk & 3), id_qk, (uint32_t) (kk > 0));
                umma_commit(s_full + 8 * st);
                umma_commit(k_empty + 8 * st);
            }
            __syncwarp();
            if (cnt > 0) pv(cnt - 1);
            ++cnt;
        }
        if (cnt > 0) pv(cnt - 1);
    } else {
        // ===== softmax / correction / epilogue: thread = query row =====
        const int r = warp * 32 + lane;
        const int64_t iq1 = q0 + r;
        const bool valid = iq1 < p.N;
        const uint32_t lane_base = (uint32_t) (warp * 32) << 16;
        const float slope = fattn_slope(p, (uint32_t) iq2);
        const __half * mrow = (p.mask && valid) ? (const __half *) (p.mask + iq1 * p.m_nb1 + (iq2 % p.m_ne2) * p.m_nb2 + (iq3 % p.m_ne3) * p.m_nb3) : nullptr;
        float m = -INFINITY, l = 0.0f;
        int cnt = 0;
        for (int kt = 0; kt < x.nkt; ++kt) {
            const int flag = fl_sm[kt];
            if (flag == 0) continue;
            const int st = cnt & 1;
            mbar_wait(s_full + 8 * st, ((uint32_t) cnt >> 1) & 1u);
            tc_fence_after();
            float s[FT_BKV];
#pragma unroll
            for (int c = 0; c < FT_BKV / 32; ++c) tmem_ld32(t_s + lane_base + (uint32_t) (st * FT_BKV + c * 32), *(uint32_t (*)[32]) &s[c * 32]);
            tmem_ld_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(s_empty + 8 * st);
            const int64_t k0 = (int64_t) kt * FT_BKV;
            float mx = -INFINITY;
#pragma unroll
            for (int c = 0; c < FT_BKV / 8; ++c) {
                uint4 mk = make_uint4(0, 0, 0, 0);
                if (flag == 2 && mrow) mk = *(const uint4 *) (mrow + k0 + c * 8);
                const __half * mh = (const __half *) &mk;
#pragma unroll
                for (int i = 0; i < 8; ++i) {
                    float v = s[c * 8 + i] * p.scale;
                    if (p.softcap != 0.0f) v = p.softcap * tanhf(v);
                    if (flag == 2) { v += slope * __half2float(mh[i]); if (k0 + c * 8 + i >= p.KV) v = -INFINITY; }
                    if (!valid) v = -INFINITY;
                    s[c * 8 + i] = v; mx = fmaxf(mx, v);
                }
            }
            const float m_new = fmaxf(m, mx);
            const float alpha = (m_new == -INFINITY || m == m_new) ? 1.0f : __expf(m - m_new);      // m = -inf -> 0
            const float mref = m_new == -INFINITY ? 0.0f : m_new;
            if (cnt > 0) {
                mbar_wait(o_done, (uint32_t) (cnt - 1) & 1u);              // PV(cnt-1) complete: O may be rescaled, P may be overwritten
                tc_fence_after();
                if (__any_sync(0xffffffffu, alpha != 1.0f)) {
#pragma unroll
                    for (int c = 0; c < D / 32; ++c) {
                        uint32_t o[32];
                        tmem_ld32(t_o + lane_base + (uint32_t) (c * 32), o);
                        tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                        tmem_st32(t_o + lane_base + (uint32_t) (c * 32), o);
                    }
                }
            }
            float sum = 0.0f;
            uint32_t pk[FT_BKV / 2];
#pragma unroll
            for (int i = 0; i < FT_BKV / 2; ++i) {
                const float p0 = __expf(s[2 * i] - mref), p1 = __expf(s[2 * i + 1] - mref);     // -inf -> 0
                sum += p0 + p1;
                __half2 h = __floats2half2_rn(p0, p1);
                pk[i] = *(uint32_t *) &h;
            }
            l = l * alpha + sum; m = m_new;
            tmem_st32(t_p + lane_base, *(uint32_t (*)[32]) &pk[0]);
            tmem_st32(t_p + lane_base + 32u, *(uint32_t (*)[32]) &pk[32]);
            tmem_st_wait();
            tc_fence_before();
            __syncwarp();
            if (lane == 0) mbar_arrive(p_full);
            ++cnt;
        }
        // ===== epilogue =====
        float sink_scale = 1.0f;
        if (p.sinks) {                                                     // ops.cpp:8241-8255
            const float sk = p.sinks[iq2];
            if (sk > m) { sink_scale = __expf(m - sk); l = l * sink_scale + 1.0f; }
            else        { l += __expf(sk - m); }
        }
        const float inv = (l == 0.0f || cnt == 0) ? 0.0f : sink_scale / l;
        if (cnt > 0) { mbar_wait(o_done, (uint32_t) (cnt - 1) & 1u); tc_fence_after(); }
        float * drow = (float *) ((char *) p.dst + iq2 * p.d_nb1 + iq1 * p.d_nb2 + iq3 * p.d_nb3);
#pragma unroll
        for (int c = 0; c < D / 32; ++c) {
            uint32_t o[32];
            if (cnt > 0) { tmem_ld32(t_o + lane_base + (uint32_t) (c * 32), o); tmem_ld_wait(); }
            if (valid) {
#pragma unroll
                for (int i = 0; i < 32; i += 4)
                    *(float4 *) (drow + c * 32 + i) = cnt > 0 ? make_float4(__uint_as_float(o[i]) * inv, __uint_as_float(o[i + 1]) * inv, __uint_as_float(o[i + 2]) * inv, __uint_as_float(o[i + 3]) * inv)
                                                              : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
            }
        }
        tc_fence_before();
    }
    __syncthreads();
    if (warp == 5) { tc_fence_after(); tmem_dealloc(tmem, 512); }
}

// ---- host side -------------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn b200_get_encode_tiled();

bool b200_fattn_tc_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst) {
    static const int off = getenv("GGML_B200_FATTN_NO_TC") != nullptr;
    if (off) return false;
    const int64_t D = k->ne[0];
    if ((D != 64 && D != 128) || v->ne[0] != D || q->ne[1] < 16) return false;
    if (((uintptr_t) q->data & 15) || (q->nb[1] & 15) || (q->nb[2] & 15) || (q->nb[3] & 15)) return false;
    if (((uintptr_t) k->data & 15) || (k->nb[1] & 15) || (k->nb[2] & 15) || (k->nb[3] & 15)) return false;
    if (((uintptr_t) v->data & 15) || (v->nb[1] & 15) || (v->nb[2] & 15) || (v->nb[3] & 15)) return false;
    if (((uintptr_t) dst->data & 15) || (dst->nb[1] & 15) || (dst->nb[2] & 15) || (dst->nb[3] & 15)) return false;
    if (mask && (((uintptr_t) mask->data & 15) || (mask->nb[1] & 15) || (mask->nb[2] & 15) || (mask->nb[3] & 15))) return false;
    if (k->ne[1] > (int64_t) 2048 * FT_BKV || q->ne[2] > 65535 || q->ne[3] > 65535) return false;
    return true;
}

// 4-D fp16 map over a K or V view [D, n_kv, n_head_kv, s]; the two middle dimensions are ordered by stride (the KV cache
// interleaves heads inside a token row, plain tensors do not); box = 64 values x 128 tokens of one head
static bool ft_make_map(encode_tiled_fn encode, CUtensorMap * map, const b200_tensor * t, int * swap) {
    *swap = (t->ne[2] > 1 && t->nb[2] < t->nb[1]) ? 1 : 0;
    const int a = *swap ? 2 : 1, b = *swap ? 1 : 2;
    cuuint64_t dim[4] = { (cuuint64_t) t->ne[0], (cuuint64_t) t->ne[a], (cuuint64_t) t->ne[b], (cuuint64_t) t->ne[3] };
    cuuint64_t str[3] = { (cuuint64_t) t->nb[a], (cuuint64_t) t->nb[b], (cuuint64_t) t->nb[3] };
    if (dim[2] == 1) str[1] = str[0] * dim[1];                             // strides of extent-1 dimensions only have to be well formed
    if (dim[3] == 1) str[2] = str[1] * dim[2];
    cuuint32_t box[4] = { 64, 1, 1, 1 }; box[*swap ? 2 : 1] = FT_BKV;
    const cuuint32_t es[4] = { 1, 1, 1, 1 };
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, t->data, dim, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// returns B200_ERR_UNSUPPORTED (without setting the error text) when the tensor maps cannot be built: the caller falls back
int b200_launch_fattn_tc(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                         const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                         float scale, float max_bias, float softcap) {
    encode_tiled_fn encode = b200_get_encode_tiled();
    if (!encode) return B200_ERR_UNSUPPORTED;
    fattn_args p;
    fattn_fill_args(p, q, k, v, mask, sinks, dst, scale, max_bias, softcap);
    if (p.N == 0 || p.H == 0 || q->ne[3] == 0) return B200_OK;
    ft_extra x;
    x.nqt = (int) ((p.N + FT_BQ - 1) / FT_BQ); x.nkt = (int) ((p.KV + FT_BKV - 1) / FT_BKV);
    CUtensorMap tk, tv;
    if (!ft_make_map(encode, &tk, k, &x.k_swap) || !ft_make_map(encode, &tv, v, &x.v_swap)) return B200_ERR_UNSUPPORTED;
    x.flags = nullptr;
    if (mask) {
        const size_t nflags = (size_t) mask->ne[3] * mask->ne[2] * x.nqt * x.nkt;
        auto & mc = ctx->mcache;
        const bool hit = mc.valid && mc.data == mask->data && ctx->scratch[2] && mc.n == p.N &&
                         !memcmp(mc.ne, mask->ne, sizeof(mc.ne)) && !memcmp(mc.nb, mask->nb, sizeof(mc.nb));
        uint8_t * flags = (uint8_t *) b200_scratch(ctx, 2, nflags);
        if (!flags) return B200_ERR_ALLOC;
        if (!hit) {
            mc.valid = 0;
            if (mask->ne[2] * mask->ne[3] > 1 && mask->nb[3] != mask->nb[2] * (size_t) mask->ne[2]) return B200_ERR_UNSUPPORTED;
            dim3 g((unsigned) x.nkt, (unsigned) x.nqt, (unsigned) (mask->ne[2] * mask->ne[3]));
            B200_LAUNCH(ctx, B200_K_CONVERT, k_mask_flags<<<g, 128, 0, ctx->stream>>>((const char *) mask->data, mask->nb[1], mask->nb[2], mask->nb[3],
                                                                                       p.KV, p.N, x.nqt, x.nkt, flags));
            mc.data = mask->data; memcpy(mc.ne, mask->ne, sizeof(mc.ne)); memcpy(mc.nb, mask->nb, sizeof(mc.nb));
            mc.span = b200_tensor_span(mask); mc.n = p.N; mc.valid = 1;
        }
        x.flags = flags;
    }
    const int D = p.DK;
    const size_t smem = 1024 + (size_t) 5 * (D / 64) * 16384 + 128 + ((x.nkt + 15) & ~15);
    if (!ctx->attr_done[6]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_tc<64>,  cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_tc<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        ctx->attr_done[6] = 1;
    }
    dim3 grid((unsigned) x.nqt, (unsigned) p.H, (unsigned) q->ne[3]);
    if (D == 64) B200_LAUNCH(ctx, B200_K_FATTN, k_fattn_tc<64><<<grid, FT_THREADS, smem, ctx->stream>>>(p, x, tk, tv));
    else         B200_LAUNCH(ctx, B200_K_FATTN, k_fattn_tc<128><<<grid, FT_THREADS, smem, ctx->stream>>>(p, x, tk, tv));
    return B200_OK;
}
