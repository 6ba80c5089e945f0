// b200_fattn.cu -- GGML_OP_FLASH_ATTN_EXT over the strided KV-cache views the fork hands to attention.
//
// Semantics restated from ggml_compute_forward_flash_attn_ext_f16_one_chunk
// (reference ggml/src/ggml-cpu/ops.cpp:8047-8272): per (query row, head): s = scale * K.q,
// optional softcap * tanh(s), + slope * mask, rows with mask == -inf skipped, online softmax with
// running max/sum, optional attention sinks, dst = VKQ / S stored permuted as [DV, H, N, S].
// K and V are f16 views of one contiguous cache buffer, heads interleaved inside each token row
// (src/llama-kv-cache.cpp:1008-1058): only nb[1..3] tell us where a (token, head) row lives, so
// the kernel is written against byte strides -- the identity page table of SURVEY 5.7.
//
// This file holds the "vec" kernel (any N; the bandwidth-bound path for decode, N small): one CTA
// per (query row, head), warps split the KV range, each lane scores one key (coalescing comes from
// 32 lanes touching 32 consecutive token rows), V is accumulated with the probabilities broadcast
// by shuffle, warps merge their (max, sum, acc) through shared memory.
#include "b200_common.cuh"
#include <math.h>

#include "b200_fattn.cuh"

#define FA_WARPS 4

template <int DV64>   // DV / 64
__global__ void __launch_bounds__(FA_WARPS * 32) k_fattn_vec(const fattn_args p) {
    extern __shared__ float smem[];
    float * sq   = smem;                                   // [DK] query row
    float * sacc = sq + p.DK;                              // [FA_WARPS][DV]
    float * sms  = sacc + FA_WARPS * p.DV;                 // [FA_WARPS][2] (M, S)

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int64_t iq1 = blockIdx.x, iq2 = blockIdx.y, iq3 = blockIdx.z;

    const float * qrow = (const float *) (p.q + iq1 * p.q_nb1 + iq2 * p.q_nb2 + iq3 * p.q_nb3);
    for (int i = threadIdx.x; i < p.DK; i += blockDim.x) sq[i] = qrow[i];
    __syncthreads();

    const uint32_t h = (uint32_t) iq2;
    const float slope = p.max_bias > 0.0f ? (h < p.n_head_log2 ? powf(p.m0, (float) (h + 1)) : powf(p.m1, (float) (2 * (h - p.n_head_log2) + 1))) : 1.0f;

    const char * kbase = p.k + (iq2 / p.rk2) * p.k_nb2 + (iq3 / p.rk3) * p.k_nb3;
    const char * vbase = p.v + (iq2 / p.rv2) * p.v_nb2 + (iq3 / p.rv3) * p.v_nb3;
    const __half * mrow = p.mask ? (const __half *) (p.mask + iq1 * p.m_nb1 + (iq2 % p.m_ne2) * p.m_nb2 + (iq3 % p.m_ne3) * p.m_nb3) : nullptr;

    float  M = -INFINITY, S = 0.0f;
    float2 acc[DV64];
#pragma unroll
    for (int i = 0; i < DV64; ++i) acc[i] = make_float2(0.0f, 0.0f);

    for (int64_t ic0 = (int64_t) warp * 32; ic0 < p.KV; ic0 += FA_WARPS * 32) {
        const int64_t ic = ic0 + lane;
        float s = -INFINITY;
        if (ic < p.KV) {
            const float mv = mrow ? slope * __half2float(mrow[ic]) : 0.0f;
            if (mv != -INFINITY) {
                const int4 * kr = (const int4 *) (kbase + ic * p.k_nb1);
                float dot = 0.0f;
                for (int i = 0; i < p.DK / 8; ++i) {
                    const int4 kk = kr[i];
                    const __half2 * kh = (const __half2 *) &kk;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {
                        const float2 kf = __half22float2(kh[j]);
                        dot += kf.x * sq[8 * i + 2 * j] + kf.y * sq[8 * i + 2 * j + 1];
                    }
                }
                s = dot * p.scale;
                if (p.softcap != 0.0f) s = p.softcap * tanhf(s);
                s += mv;
            }
        }
        const float mt = warp_max(s);
        if (mt == -INFINITY) continue;                      // whole tile masked (warp-uniform)
        const float Mn = fmaxf(M, mt);
        const float ms = expf(M - Mn);                      // M == -inf -> 0
        const float pr = s == -INFINITY ? 0.0f : expf(s - Mn);
        S = S * ms + warp_sum(pr);
        M = Mn;
#pragma unroll
        for (int i = 0; i < DV64; ++i) { acc[i].x *= ms; acc[i].y *= ms; }
        const unsigned live = __ballot_sync(0xffffffffu, pr != 0.0f);
        for (unsigned bits = live; bits; bits &= bits - 1) {
            const int j = __ffs(bits) - 1;
            const float pj = __shfl_sync(0xffffffffu, pr, j);
            const __half2 * vr = (const __half2 *) (vbase + (ic0 + j) * p.v_nb1);
#pragma unroll
            for (int i = 0; i < DV64; ++i) {
                const float2 vf = __half22float2(vr[32 * i + lane]);
                acc[i].x += pj * vf.x; acc[i].y += pj * vf.y;
            }
        }
    }

    // merge the warps
#pragma unroll
    for (int i = 0; i < DV64; ++i) {
        sacc[warp * p.DV + 64 * i + 2 * lane]     = acc[i].x;
        sacc[warp * p.DV + 64 * i + 2 * lane + 1] = acc[i].y;
    }
    if (lane == 0) { sms[2 * warp] = M; sms[2 * warp + 1] = S; }
    __syncthreads();

    float Mall = -INFINITY;
#pragma unroll
    for (int w = 0; w < FA_WARPS; ++w) Mall = fmaxf(Mall, sms[2 * w]);
    float Sall = 0.0f, wgt[FA_WARPS];
#pragma unroll
    for (int w = 0; w < FA_WARPS; ++w) {
        wgt[w] = sms[2 * w] == -INFINITY ? 0.0f : expf(sms[2 * w] - Mall);
        Sall += sms[2 * w + 1] * wgt[w];
    }
    float sink_scale = 1.0f;
    if (p.sinks) {                                           // ops.cpp:8241-8255
        const float sk = p.sinks[h];
        if (sk > Mall) { sink_scale = expf(Mall - sk); Sall = Sall * sink_scale + 1.0f; }
        else           { Sall += expf(sk - Mall); }
    }
    const float Sinv = Sall == 0.0f ? 0.0f : 1.0f / Sall;
    float * drow = (float *) ((char *) p.dst + iq2 * p.d_nb1 + iq1 * p.d_nb2 + iq3 * p.d_nb3);
    for (int d = threadIdx.x; d < p.DV; d += blockDim.x) {
        float o = 0.0f;
#pragma unroll
        for (int w = 0; w < FA_WARPS; ++w) o += sacc[w * p.DV + d] * wgt[w];
        drow[d] = o * sink_scale * Sinv;
    }
}

int b200_launch_fattn_vec(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                          const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                          float scale, float max_bias, float softcap) {
    fattn_args p;
    fattn_fill_args(p, q, k, v, mask, sinks, dst, scale, max_bias, softcap);
    if (p.N == 0 || p.H == 0 || q->ne[3] == 0) return B200_OK;
    if (p.H > 65535 || q->ne[3] > 65535) { b200_set_error("fattn: H or S exceeds grid limits"); return B200_ERR_UNSUPPORTED; }
    dim3 grid((unsigned) p.N, (unsigned) p.H, (unsigned) q->ne[3]);
    const size_t smem = (size_t) (p.DK + FA_WARPS * p.DV + 2 * FA_WARPS) * sizeof(float);
    switch (p.DV / 64) {
        case 1: B200_LAUNCH(ctx, B200_K_FATTN, k_fattn_vec<1><<<grid, FA_WARPS * 32, smem, ctx->stream>>>(p)); break;
        case 2: B200_LAUNCH(ctx, B200_K_FATTN, k_fattn_vec<2><<<grid, FA_WARPS * 32, smem, ctx->stream>>>(p)); break;
        case 4: B200_LAUNCH(ctx, B200_K_FATTN, k_fattn_vec<4><<<grid, FA_WARPS * 32, smem, ctx->stream>>>(p)); break;
        default: b200_set_error("fattn: DV=%d unsupported", p.DV); return B200_ERR_UNSUPPORTED;
    }
    return B200_OK;
}
