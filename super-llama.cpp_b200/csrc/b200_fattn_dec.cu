// b200_fattn_dec.cu -- GGML_OP_FLASH_ATTN_EXT for small query batches (decode): TMA row-gather, split-KV.
//
// Semantics: ggml_compute_forward_flash_attn_ext_f16_one_chunk (reference ggml/src/ggml-cpu/ops.cpp:8047-8272).
// Bound: HBM/L2 on K and V bytes (2 * n_kv * head_dim * 2 B per kv head and layer); at n_kv = 256 it is pure latency.
//
// K and V arrive as strided views of the KV cache with heads interleaved inside each token row
// (src/llama-kv-cache.cpp:1008-1058), so a (token, head) row is DK*2 contiguous bytes at an address only the strides
// (the identity page table of SURVEY 5.7) give.  Each warp owns tiles of 32 keys: every lane issues ONE cp.async.bulk for
// "its" K row and one for its V row (a per-row gather through the TMA engine -- no registers, no LSU slots, and it would
// follow a real page table unchanged), double-buffered on mbarriers.  K rows land at a padded stride (DK*2+16 B) so the
// lane-per-key dot product reads 128-bit words conflict-free; V rows land densely and are read lane-per-dim.
// Fully masked keys are never fetched; a tile whose 32 keys are all masked is skipped.
// Long contexts are split over CTAs (split-KV); partial (max, sum, acc) triples are merged by k_fattn_combine.
#include "b200_fattn.cuh"
#include "b200_ptx.cuh"
#include <stdlib.h>

#define FD_WARPS 4
#define FD_TILE  32

struct fd_extra {
    int nsplit, ntiles, tiles_per_split, nstage;
    int trigger;                // GGML_B200_FATTN_DEC_TRIGGER: 1 = PDL trigger once Q is in shared memory, 2 = at the top of the kernel
    float * partial;            // [S][H][N][nsplit][DV + 2] when nsplit > 1
};

template <int DV64>
__global__ void __launch_bounds__(FD_WARPS * 32, 1) k_fattn_dec(const fattn_args p, const fd_extra x) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int DKb = p.DK * 2, DVb = p.DV * 2, KROW = DKb + 16;
    const int stage_bytes = FD_TILE * (KROW + DVb);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    float *    sq   = (float *) (smem + FD_WARPS * x.nstage * stage_bytes);       // [DK] query row
    float *    sacc = sq + p.DK;                                                   // [FD_WARPS][DV]
    float *    sms  = sacc + FD_WARPS * p.DV;                                      // [FD_WARPS][2]
    uint64_t * sbar = (uint64_t *) (sms + 2 * FD_WARPS);                           // [FD_WARPS][2]
    const uint32_t ring = smem_u32(smem + warp * x.nstage * stage_bytes);
    const uint32_t bar0 = smem_u32(sbar + warp * 2);

    const int iq1 = blockIdx.x / x.nsplit, split = blockIdx.x % x.nsplit;
    const int64_t iq2 = blockIdx.y, iq3 = blockIdx.z;

    if (lane == 0) { mbar_init(bar0, 1); mbar_init(bar0 + 8, 1); mbar_fence_init(); }
    // no griddepcontrol.launch_dependents here: letting the PDL-launched GEMV behind us start on the SMs this small grid leaves free
    // measured 2 % SLOWER for the whole token (662 -> 649 tok/s): its CTAs on our SMs start late and become the stragglers
    // launched with the PDL attribute (GGML_B200_FATTN_DEC_PDL != 0): our launch latency hides under the previous kernel's tail; Q, the
    // mask and the newest K/V rows come from that kernel, so everything below waits for it.  (A no-op without the attribute.)
    if (x.trigger == 2) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");
    const float * qrow = (const float *) (p.q + iq1 * p.q_nb1 + iq2 * p.q_nb2 + iq3 * p.q_nb3);
    for (int i = threadIdx.x; i < p.DK; i += blockDim.x) sq[i] = qrow[i];
    __syncthreads();
    if (x.trigger == 1) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    const uint32_t h = (uint32_t) iq2;
    const float slope = fattn_slope(p, h);
    const char * kbase = p.k + (iq2 / p.rk2) * p.k_nb2 + (iq3 / p.rk3) * p.k_nb3;
    const char * vbase = p.v + (iq2 / p.rv2) * p.v_nb2 + (iq3 / p.rv3) * p.v_nb3;
    const __half * mrow = p.mask ? (const __half *) (p.mask + iq1 * p.m_nb1 + (iq2 % p.m_ne2) * p.m_nb2 + (iq3 % p.m_ne3) * p.m_nb3) : nullptr;

    const int t_begin = split * x.tiles_per_split, t_end = min(x.ntiles, t_begin + x.tiles_per_split);

    // per-stage bookkeeping of the tile in flight: which keys are live, and this lane's mask term
    unsigned live0 = 0u, live1 = 0u; float mv0 = 0.0f, mv1 = 0.0f;       // (scalars: no local-memory indexing)
    auto issue = [&](int tile, int stage) {                    // whole warp; every live lane gathers its own K and V row
        const int64_t ic = (int64_t) tile * FD_TILE + lane;
        float mv = 0.0f; bool ok = ic < p.KV;
        if (ok && mrow) { mv = slope * __half2float(mrow[ic]); ok = mv != -INFINITY; }
        const unsigned live = __ballot_sync(0xffffffffu, ok);
        if (stage) { live1 = live; mv1 = mv; } else { live0 = live; mv0 = mv; }
        if (live == 0u) return;
        const uint32_t bar = bar0 + 8 * stage, st = ring + stage * stage_bytes;
        if (lane == 0) mbar_arrive_expect_tx(bar, (uint32_t) __popc(live) * (uint32_t) (DKb + DVb));
        __syncwarp();
        if (ok) {
            const int64_t row = fattn_row(p, ic);                  // (paged cache: the page table is followed per row)
            bulk_g2s(st + lane * KROW, kbase + row * p.k_nb1, (uint32_t) DKb, bar);
            bulk_g2s(st + FD_TILE * KROW + lane * DVb, vbase + row * p.v_nb1, (uint32_t) DVb, bar);
        }
    };

    int tile_i = t_begin + warp;                               // next tile to issue
    for (int s = 0; s < x.nstage && tile_i < t_end; ++s, tile_i += FD_WARPS) issue(tile_i, s);

    float  M = -INFINITY, S = 0.0f;
    float2 acc[DV64];
#pragma unroll
    for (int i = 0; i < DV64; ++i) acc[i] = make_float2(0.0f, 0.0f);

    int stage = 0; unsigned phase = 0u;                        // bit s = parity of stage s
    for (int tile = t_begin + warp; tile < t_end; tile += FD_WARPS) {
        const unsigned live = stage ? live1 : live0;
        const float mv = stage ? mv1 : mv0;
        if (live != 0u) {
            mbar_wait(bar0 + 8 * stage, (phase >> stage) & 1u); phase ^= 1u << stage;
            const uint32_t st = ring + stage * stage_bytes;
            float s = -INFINITY;
            if ((live >> lane) & 1u) {
                const uint32_t kr = st + lane * KROW;
                float dot = 0.0f;
                for (int i = 0; i < p.DK / 8; ++i) {
                    const int4 kk = lds128(kr + 16 * i);
                    const float4 q0 = *(const float4 *) (sq + 8 * i), q1 = *(const float4 *) (sq + 8 * i + 4);
                    const float2 k0 = __half22float2(*(const __half2 *) &kk.x), k1 = __half22float2(*(const __half2 *) &kk.y);
                    const float2 k2 = __half22float2(*(const __half2 *) &kk.z), k3 = __half22float2(*(const __half2 *) &kk.w);
                    dot += k0.x * q0.x + k0.y * q0.y + k1.x * q0.z + k1.y * q0.w + k2.x * q1.x + k2.y * q1.y + k3.x * q1.z + k3.y * q1.w;
                }
                s = dot * p.scale;
                if (p.softcap != 0.0f) s = p.softcap * tanhf(s);
                s += mv;
            }
            const float mt = warp_max(s);
            if (mt != -INFINITY) {
                const float Mn = fmaxf(M, mt);
                const float ms = expf(M - Mn);                     // M == -inf -> 0
                const float pr = s == -INFINITY ? 0.0f : expf(s - Mn);
                S = S * ms + warp_sum(pr);
                M = Mn;
#pragma unroll
                for (int i = 0; i < DV64; ++i) { acc[i].x *= ms; acc[i].y *= ms; }
                const uint32_t vr = st + FD_TILE * KROW + lane * 4;
#pragma unroll 8
                for (int j = 0; j < FD_TILE; ++j) {
                    const float pj = __shfl_sync(0xffffffffu, pr, j);
                    if ((live >> j) & 1u) {                        // warp-uniform; dead rows hold stale bytes
#pragma unroll
                        for (int i = 0; i < DV64; ++i) {
                            const int raw = lds32(vr + j * DVb + 128 * i);
                            const float2 vf = __half22float2(*(const __half2 *) &raw);
                            acc[i].x += pj * vf.x; acc[i].y += pj * vf.y;
                        }
                    }
                }
            }
        }
        __syncwarp();
        if (tile_i < t_end) { issue(tile_i, stage); tile_i += FD_WARPS; } else if (stage) live1 = 0u; else live0 = 0u;
        if (x.nstage == 2) stage ^= 1;
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
    for (int w = 0; w < FD_WARPS; ++w) Mall = fmaxf(Mall, sms[2 * w]);
    float Sall = 0.0f, wgt[FD_WARPS];
#pragma unroll
    for (int w = 0; w < FD_WARPS; ++w) {
        wgt[w] = sms[2 * w] == -INFINITY ? 0.0f : expf(sms[2 * w] - Mall);
        Sall += sms[2 * w + 1] * wgt[w];
    }
    if (x.nsplit > 1) {
        float * part = x.partial + ((((int64_t) iq3 * p.H + iq2) * p.N + iq1) * x.nsplit + split) * (p.DV + 2);
        for (int d = threadIdx.x; d < p.DV; d += blockDim.x) {
            float o = 0.0f;
#pragma unroll
            for (int w = 0; w < FD_WARPS; ++w) o += sacc[w * p.DV + d] * wgt[w];
            part[d] = o;
        }
        if (threadIdx.x == 0) { part[p.DV] = Mall; part[p.DV + 1] = Sall; }
        return;
    }
    float sink_scale = 1.0f;
    if (p.sinks) {                                               // ops.cpp:8241-8255
        const float sk = p.sinks[h];
        if (sk > Mall) { sink_scale = expf(Mall - sk); Sall = Sall * sink_scale + 1.0f; }
        else           { Sall += expf(sk - Mall); }
    }
    const float Sinv = Sall == 0.0f ? 0.0f : 1.0f / Sall;
    float * drow = (float *) ((char *) p.dst + iq2 * p.d_nb1 + iq1 * p.d_nb2 + iq3 * p.d_nb3);
    for (int d = threadIdx.x; d < p.DV; d += blockDim.x) {
        float o = 0.0f;
#pragma unroll
        for (int w = 0; w < FD_WARPS; ++w) o += sacc[w * p.DV + d] * wgt[w];
        drow[d] = o * sink_scale * Sinv;
    }
}

// merges the nsplit partial (acc, max, sum) triples of one (query row, head): one CTA each
__global__ void __launch_bounds__(128) k_fattn_combine(const fattn_args p, const fd_extra x) {
    const int iq1 = blockIdx.x; const int64_t iq2 = blockIdx.y, iq3 = blockIdx.z;
    const float * part = x.partial + (((int64_t) iq3 * p.H + iq2) * p.N + iq1) * x.nsplit * (p.DV + 2);
    float Mall = -INFINITY;
    for (int s = 0; s < x.nsplit; ++s) Mall = fmaxf(Mall, part[s * (p.DV + 2) + p.DV]);
    float Sall = 0.0f;
    for (int s = 0; s < x.nsplit; ++s) {
        const float m = part[s * (p.DV + 2) + p.DV];
        if (m != -INFINITY) Sall += part[s * (p.DV + 2) + p.DV + 1] * expf(m - Mall);
    }
    float sink_scale = 1.0f;
    if (p.sinks) {
        const float sk = p.sinks[iq2];
        if (sk > Mall) { sink_scale = expf(Mall - sk); Sall = Sall * sink_scale + 1.0f; }
        else           { Sall += expf(sk - Mall); }
    }
    const float Sinv = Sall == 0.0f ? 0.0f : 1.0f / Sall;
    float * drow = (float *) ((char *) p.dst + iq2 * p.d_nb1 + iq1 * p.d_nb2 + iq3 * p.d_nb3);
    for (int d = threadIdx.x; d < p.DV; d += blockDim.x) {
        float o = 0.0f;
        for (int s = 0; s < x.nsplit; ++s) {
            const float m = part[s * (p.DV + 2) + p.DV];
            if (m != -INFINITY) o += part[s * (p.DV + 2) + d] * expf(m - Mall);
        }
        drow[d] = o * sink_scale * Sinv;
    }
}

// shared with b200_fattn_gqa.cu: partial layout [S][H][N][nsplit][DV + 2]
int b200_launch_fattn_combine(b200_ctx * ctx, const fattn_args & p, int64_t S, int nsplit, float * partial) {
    fd_extra x; x.nsplit = nsplit; x.ntiles = 0; x.tiles_per_split = 0; x.nstage = 0; x.partial = partial;
    dim3 cgrid((unsigned) p.N, (unsigned) p.H, (unsigned) S);
    B200_LAUNCH(ctx, B200_K_FATTN_COMBINE, k_fattn_combine<<<cgrid, 128, 0, ctx->stream>>>(p, x));
    return B200_OK;
}

bool b200_fattn_dec_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v) {
    if (q->ne[1] > 8) return false;                              // larger batches: the prefill kernels
    if ((v->nb[1] & 15) || (v->nb[2] & 15) || (v->nb[3] & 15) || ((uintptr_t) v->data & 15)) return false;
    if ((k->nb[1] & 15) || (k->nb[2] & 15) || (k->nb[3] & 15) || ((uintptr_t) k->data & 15)) return false;
    return k->ne[0] % 64 == 0 && k->ne[0] <= 256 && (v->ne[0] == 64 || v->ne[0] == 128 || v->ne[0] == 256);
}

int b200_launch_fattn_dec(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                          const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                          float scale, float max_bias, float softcap, const b200_kv_pages * pages) {
    fattn_args p;
    fattn_fill_args(p, q, k, v, mask, sinks, dst, scale, max_bias, softcap);
    fattn_set_pages(p, pages);
    const int64_t S = q->ne[3];
    if (p.N == 0 || p.H == 0 || S == 0) return B200_OK;
    if (p.H > 65535 || S > 65535) { b200_set_error("fattn: H or S exceeds grid limits"); return B200_ERR_UNSUPPORTED; }
    fd_extra x;
    x.ntiles = (int) ((p.KV + FD_TILE - 1) / FD_TILE);
    x.nstage = (p.DK + p.DV <= 256) ? 2 : 1;
    // split the KV range over CTAs until the grid covers the SMs, keeping >= 16 tiles (512 keys) per split: a split costs a second
    // launch (k_fattn_combine), ~9 us inside a token's kernel chain, which 4 tiles per warp do not
    const int64_t rows = p.N * p.H * S;
    int nsplit = (int) (ctx->sm_count / (rows > 0 ? rows : 1));
    nsplit = nsplit < 1 ? 1 : nsplit;
    if (nsplit > x.ntiles / 16) nsplit = x.ntiles / 16 > 0 ? x.ntiles / 16 : 1;
    x.tiles_per_split = (x.ntiles + nsplit - 1) / nsplit;
    x.nsplit = (x.ntiles + x.tiles_per_split - 1) / x.tiles_per_split;
    { static int tr = -1; if (tr < 0) { const char * e = getenv("GGML_B200_FATTN_DEC_TRIGGER"); tr = e ? atoi(e) : 2; } x.trigger = tr; }   // measured: 677 (none) -> 687 tok/s (top of the kernel)
    x.partial = nullptr;
    if (x.nsplit > 1) {
        x.partial = (float *) b200_scratch(ctx, 1, (size_t) rows * x.nsplit * (p.DV + 2) * sizeof(float));
        if (!x.partial) return B200_ERR_ALLOC;
    }
    const size_t smem = (size_t) FD_WARPS * x.nstage * FD_TILE * (p.DK * 2 + 16 + p.DV * 2) + (size_t) (p.DK + FD_WARPS * p.DV + 2 * FD_WARPS) * 4 + FD_WARPS * 16;
    if (!ctx->attr_done[4]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_dec<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_dec<2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_dec<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        ctx->attr_done[4] = 1;
    }
    static const int use_pdl = getenv("GGML_B200_FATTN_DEC_PDL") ? atoi(getenv("GGML_B200_FATTN_DEC_PDL")) : 1;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) (p.N * x.nsplit), (unsigned) p.H, (unsigned) S);
    cfg.blockDim = dim3(FD_WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = (ctx->pdl && use_pdl) ? 1 : 0;
    switch (p.DV / 64) {
        case 1: B200_LAUNCH(ctx, B200_K_FATTN, cudaLaunchKernelEx(&cfg, k_fattn_dec<1>, p, x)); break;
        case 2: B200_LAUNCH(ctx, B200_K_FATTN, cudaLaunchKernelEx(&cfg, k_fattn_dec<2>, p, x)); break;
        case 4: B200_LAUNCH(ctx, B200_K_FATTN, cudaLaunchKernelEx(&cfg, k_fattn_dec<4>, p, x)); break;
        default: b200_set_error("fattn: DV=%d unsupported", p.DV); return B200_ERR_UNSUPPORTED;
    }
    if (x.nsplit > 1) {
        dim3 cgrid((unsigned) p.N, (unsigned) p.H, (unsigned) S);
        B200_LAUNCH(ctx, B200_K_FATTN_COMBINE, k_fattn_combine<<<cgrid, 128, 0, ctx->stream>>>(p, x));
    }
    return B200_OK;
}
