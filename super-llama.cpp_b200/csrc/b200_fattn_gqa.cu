// b200_fattn_gqa.cu -- GGML_OP_FLASH_ATTN_EXT for decode (<= 16 query rows per kv head), f16 K/V, head size 64 / 128.
//
// Semantics: ggml_compute_forward_flash_attn_ext_f16_one_chunk (reference ggml/src/ggml-cpu/ops.cpp:8047-8272).
// Bound: HBM on the K and V bytes (2 * n_kv * D * 2 B per kv head and layer), each of which must be read ONCE:
//   * one CTA serves ALL query heads of a GQA group (and all <= 8 query tokens): its <= 16 (token, head) rows are the M dimension of
//     mma.sync.m16n8k16 (f16 in, f32 accumulate), so K/V tiles are fetched once per kv head, not once per query head;
//   * K/V tiles of 16 keys arrive by TMA (4-D tensor maps straight over the strided KV-cache views of src/llama-kv-cache.cpp:1008-1058,
//     SWIZZLE_128B, 64-value boxes) into per-warp rings of 3 stages with one mbarrier each; ldmatrix (.trans for V) reads them
//     conflict-free; tiles whose keys are all masked are never fetched;
//   * every warp runs its own online softmax over its tiles (no CTA-wide synchronisation in the loop); the 8 warps merge through shared
//     memory at the end, and long contexts are split over CTAs (split-KV) with the partial (acc, max, sum) merged by k_fattn_combine.
#include "b200_fattn.cuh"
#include "b200_ptx.cuh"
#include <cuda.h>
#include <cuda_fp16.h>

#define GA_WARPS  8
#define GA_TILE   16
#define GA_STAGES 3

struct ga_extra {
    int nsplit, ntiles, tiles_per_split;
    int flags_per_warp;         // bytes of per-tile mask classes each warp keeps in shared memory
    int G, nrows;               // query heads per kv head, live rows = N * G (<= 16); row r = token (r / G), head-in-group (r % G)
    int k_swap, v_swap;         // tensor-map dimension order: 0 = (d, token, head, s), 1 = (d, head, token, s)
    float * partial;            // [S][H][N][nsplit][D + 2] when nsplit > 1
    unsigned * counters;        // [S][Hkv] arrival counters (zero between launches) when the last CTA merges the splits; else NULL
};

static __device__ __forceinline__ void ldsm_x4(uint32_t addr, unsigned & r0, unsigned & r1, unsigned & r2, unsigned & r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
static __device__ __forceinline__ void ldsm_x4_t(uint32_t addr, unsigned & r0, unsigned & r1, unsigned & r2, unsigned & r3) {
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.trans.shared.b16 {%0, %1, %2, %3}, [%4];" : "=r"(r0), "=r"(r1), "=r"(r2), "=r"(r3) : "r"(addr));
}
static __device__ __forceinline__ void mma_f16(float (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.f16.f16.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
static __device__ __forceinline__ unsigned pack_h2(float a, float b) {
    const __half2 h = __floats2half2_rn(a, b);
    return *(const unsigned *) &h;
}

template <int D>
__global__ void __launch_bounds__(GA_WARPS * 32, 1) k_fattn_gqa(const fattn_args p, const ga_extra x, const __grid_constant__ CUtensorMap tmap_k,
                                                                 const __grid_constant__ CUtensorMap tmap_v) {
    constexpr int NB = D / 64;                              // 64-value boxes per row
    constexpr int BOX = GA_TILE * 128;                      // bytes of one box: 16 keys x 128 B
    constexpr int STAGE = 2 * NB * BOX;                     // K boxes then V boxes
    constexpr int KS = D / 16;                              // k-steps of Q K^T
    constexpr int NT = D / 8;                               // n-tiles of P V
    extern __shared__ uint8_t smem_raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) smem_raw + 1023) & ~(uintptr_t) 1023);      // swizzle atoms are 1 KB
    uint64_t * sbar = (uint64_t *) (smem + GA_WARPS * GA_STAGES * STAGE);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int g = lane >> 2, t = lane & 3;
    const uint32_t ring = smem_u32(smem + warp * GA_STAGES * STAGE);
    const uint32_t bar0 = smem_u32(sbar + warp * GA_STAGES);

    const int split = blockIdx.x;
    const int hkv = blockIdx.y; const int64_t iq3 = blockIdx.z;
    const int sk = (int) (iq3 / p.rk3), sv = (int) (iq3 / p.rv3);

    if (lane == 0) {
        for (int s = 0; s < GA_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
        if (warp == 0) { tma_prefetch_desc(&tmap_k); tma_prefetch_desc(&tmap_v); }
    }
    __syncwarp();
    // the matmul behind us is a programmatic dependent: let it start its prologue and weight prefetch now (it reads our output only
    // after its own griddepcontrol.wait).  Our own wait matters only with GGML_B200_FATTN_PDL: Q, the mask and the newest K/V rows
    // all come from kernels before us.
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    asm volatile("griddepcontrol.wait;" ::: "memory");

    // rows of this lane: g and g + 8
    const int ra = g, rb = g + 8;
    const bool va = ra < x.nrows, vb = rb < x.nrows;
    const int ta = va ? ra / x.G : 0, tb = vb ? rb / x.G : 0;                       // query token of each row
    const int ha = hkv * x.G + (va ? ra % x.G : 0), hb = hkv * x.G + (vb ? rb % x.G : 0);
    const __half * ma = p.mask ? (const __half *) (p.mask + ta * p.m_nb1 + (ha % p.m_ne2) * p.m_nb2 + (iq3 % p.m_ne3) * p.m_nb3) : nullptr;
    const __half * mb = p.mask ? (const __half *) (p.mask + tb * p.m_nb1 + (hb % p.m_ne2) * p.m_nb2 + (iq3 % p.m_ne3) * p.m_nb3) : nullptr;
    const float sla = fattn_slope(p, (uint32_t) ha), slb = fattn_slope(p, (uint32_t) hb);

    const int t_begin = split * x.tiles_per_split, t_end = min(x.ntiles, t_begin + x.tiles_per_split);

    auto fetch = [&](int tile, int stage) {
        if (lane == 0) {
            const uint32_t bar = bar0 + 8 * stage, st = ring + stage * STAGE;
            mbar_arrive_expect_tx(bar, STAGE);
            const int k0 = (int) fattn_row(p, (int64_t) tile * GA_TILE);      // (paged cache: a 16-key tile never straddles a page)
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                tma_load_4d(st + b * BOX, &tmap_k, b * 64, x.k_swap ? hkv : k0, x.k_swap ? k0 : hkv, sk, bar);
                tma_load_4d(st + (NB + b) * BOX, &tmap_v, b * 64, x.v_swap ? hkv : k0, x.v_swap ? k0 : hkv, sv, bar);
            }
        }
    };
    // the first tiles of the ring are fetched before anything is known about the mask (their latency overlaps the mask scan);
    // should one of them turn out fully masked it is simply processed as a mixed tile (all its probabilities come out 0)
    int tile_i = t_begin + warp;
    int nearly = 0;
    for (; nearly < GA_STAGES && tile_i < t_end; ++nearly, tile_i += GA_WARPS) fetch(tile_i, nearly);

    // Per-tile mask class, found once per warp for its own tiles (t_begin + warp + 8 i): 0 = every (token, key) pair masked: never
    // fetched; 2 = mask all zero (the bulk of a causal decode): no mask loads in the loop; 1 = mixed.  Lane l looks at key l & 15 for
    // the tokens l >> 4, +2, ...; eight tiles' loads are in flight at a time.
    uint8_t * sflag = (uint8_t *) (sbar + GA_WARPS * GA_STAGES) + warp * x.flags_per_warp;
    {
        const char * mbase = p.mask ? p.mask + ((hkv * x.G) % p.m_ne2) * p.m_nb2 + (iq3 % p.m_ne3) * p.m_nb3 : nullptr;
        int idx = 0;
        for (int tile0 = t_begin + warp; tile0 < t_end; tile0 += 8 * GA_WARPS, idx += 8) {   // 8 = the unroll below
            unsigned cls[8];                                 // bit 0: something visible, bit 1: something not plain zero
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const int tile = tile0 + u * GA_WARPS;
                const int64_t key = (int64_t) tile * GA_TILE + (lane & 15);
                unsigned c = 2u;                             // out of range: masked
                if (tile < t_end && key < p.KV) {
                    c = 1u;
                    if (mbase) {
                        c = 0u;
                        for (int n = lane >> 4; n < (int) p.N; n += 2) {
                            const float mv = __half2float(*(const __half *) (mbase + n * p.m_nb1 + key * 2));
                            c |= (mv != -INFINITY ? 1u : 0u) | (mv != 0.0f ? 2u : 0u);
                        }
                        if ((lane >> 4) >= (int) p.N) c = 0u;   // N == 1: the upper half-warp has no token
                    }
                }
                cls[u] = c;
            }
#pragma unroll
            for (int u = 0; u < 8; ++u) {
                const bool vis = __any_sync(0xffffffffu, cls[u] & 1u), odd = __any_sync(0xffffffffu, cls[u] & 2u);
                if (lane == 0 && tile0 + u * GA_WARPS < t_end) sflag[idx + u] = vis ? (odd || p.max_bias > 0.0f ? 1 : 2) : (idx + u < nearly ? 1 : 0);
            }
        }
        __syncwarp();
    }
    unsigned livebits = 0u;                                  // bit s: the tile in stage s is live
    int issue_idx = 0;                                       // index of the next tile to issue in this warp's own sequence
    auto issue = [&](int tile, int stage) {
        const bool live = sflag[issue_idx++] != 0;
        livebits = (livebits & ~(1u << stage)) | ((live ? 1u : 0u) << stage);
        if (live) fetch(tile, stage);
    };
    livebits = (1u << nearly) - 1u; issue_idx = nearly;

    // Q rows -> f16 A fragments (the reference converts the query row to f16 too, ops.cpp:8117-8118)
    unsigned qa[KS][4];
    {
        const float * qra = (const float *) (p.q + ta * p.q_nb1 + (int64_t) ha * p.q_nb2 + iq3 * p.q_nb3);
        const float * qrb = (const float *) (p.q + tb * p.q_nb1 + (int64_t) hb * p.q_nb2 + iq3 * p.q_nb3);
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
            const float2 z = make_float2(0.f, 0.f);
            const float2 a0 = va ? *(const float2 *) (qra + 16 * ks + 2 * t) : z, a2 = va ? *(const float2 *) (qra + 16 * ks + 8 + 2 * t) : z;
            const float2 a1 = vb ? *(const float2 *) (qrb + 16 * ks + 2 * t) : z, a3 = vb ? *(const float2 *) (qrb + 16 * ks + 8 + 2 * t) : z;
            qa[ks][0] = pack_h2(a0.x, a0.y); qa[ks][1] = pack_h2(a1.x, a1.y); qa[ks][2] = pack_h2(a2.x, a2.y); qa[ks][3] = pack_h2(a3.x, a3.y);
        }
    }

    float Ma = -INFINITY, Mb = -INFINITY, la = 0.0f, lb = 0.0f;
    float o[NT][4];
#pragma unroll
    for (int i = 0; i < NT; ++i) o[i][0] = o[i][1] = o[i][2] = o[i][3] = 0.0f;

    // ldmatrix lane addressing inside a box: matrix mi = lane >> 3, row-in-matrix = lane & 7
    const int mi = lane >> 3, lr = lane & 7;
    // K (non-transposed): one x4 covers ONE n-tile (8 keys) and TWO k-steps: 16-byte chunk 4 j + mi of the key's row
    // V (transposed): one x4 covers both key halves of TWO dv n-tiles: keys 8 (mi & 1) + lr, chunk = 2 j + (mi >> 1)
    int stage = 0, cons_idx = 0; unsigned phase = 0u;
    for (int tile = t_begin + warp; tile < t_end; tile += GA_WARPS, ++cons_idx) {
        if ((livebits >> stage) & 1u) {
            const int k0 = tile * GA_TILE;
            // mask terms of a mixed tile first: their latency overlaps the wait for the tile
            float mk[2][4];
            const bool mixed = sflag[cons_idx] == 1;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                mk[nt][0] = mk[nt][1] = mk[nt][2] = mk[nt][3] = 0.0f;
                if (mixed) {
                    const int key = k0 + 8 * nt + 2 * t;
                    float2 fa = make_float2(0.f, 0.f), fb = make_float2(0.f, 0.f);
                    if (p.mask) {
                        if (key + 1 < p.KV) { fa = __half22float2(*(const __half2 *) (ma + key)); fb = __half22float2(*(const __half2 *) (mb + key)); }
                        else if (key < p.KV) { fa.x = __half2float(ma[key]); fb.x = __half2float(mb[key]); }
                    }
                    mk[nt][0] = key     < p.KV ? sla * fa.x : -INFINITY; mk[nt][1] = key + 1 < p.KV ? sla * fa.y : -INFINITY;
                    mk[nt][2] = key     < p.KV ? slb * fb.x : -INFINITY; mk[nt][3] = key + 1 < p.KV ? slb * fb.y : -INFINITY;
                }
            }
            mbar_wait(bar0 + 8 * stage, (phase >> stage) & 1u); phase ^= 1u << stage;
            const uint32_t st = ring + stage * STAGE;

            // S = Q K^T for 16 keys
            float s[2][4], s2[2][4];                                             // two accumulators per n-tile: half the mma dependency chain
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) { s[nt][0] = s[nt][1] = s[nt][2] = s[nt][3] = 0.0f; s2[nt][0] = s2[nt][1] = s2[nt][2] = s2[nt][3] = 0.0f; }
#pragma unroll
            for (int j = 0; j < KS / 2; ++j) {
                const int chunk = 4 * j + mi;                                   // 16-byte chunk of the D*2-byte row
                const uint32_t box = st + (chunk >> 3) * BOX;
#pragma unroll
                for (int nt = 0; nt < 2; ++nt) {
                    const int row = 8 * nt + lr;
                    unsigned b0, b1, b2, b3;
                    ldsm_x4(box + row * 128 + (((chunk & 7) ^ (row & 7)) << 4), b0, b1, b2, b3);
                    mma_f16(s[nt], qa[2 * j][0], qa[2 * j][1], qa[2 * j][2], qa[2 * j][3], b0, b1);
                    mma_f16(s2[nt], qa[2 * j + 1][0], qa[2 * j + 1][1], qa[2 * j + 1][2], qa[2 * j + 1][3], b2, b3);
                }
            }
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) { s[nt][0] += s2[nt][0]; s[nt][1] += s2[nt][1]; s[nt][2] += s2[nt][2]; s[nt][3] += s2[nt][3]; }
            // scale, softcap, mask (ops.cpp:8176-8184); rows g (values 0, 1) and g + 8 (values 2, 3)
            float mxa = -INFINITY, mxb = -INFINITY;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                    float v = s[nt][e] * p.scale;
                    if (p.softcap != 0.0f) v = p.softcap * tanhf(v);
                    v += mk[nt][e];
                    s[nt][e] = v;
                }
                mxa = fmaxf(mxa, fmaxf(s[nt][0], s[nt][1])); mxb = fmaxf(mxb, fmaxf(s[nt][2], s[nt][3]));
            }
            mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, 1)); mxa = fmaxf(mxa, __shfl_xor_sync(0xffffffffu, mxa, 2));
            mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, 1)); mxb = fmaxf(mxb, __shfl_xor_sync(0xffffffffu, mxb, 2));
            const float Mna = fmaxf(Ma, mxa), Mnb = fmaxf(Mb, mxb);
            const float Msa = Mna == -INFINITY ? 0.0f : Mna, Msb = Mnb == -INFINITY ? 0.0f : Mnb;      // all masked so far: keep exp() finite
            const float ca = __expf(Ma - Msa), cb = __expf(Mb - Msb);                                   // Ma == -inf -> 0
            Ma = Mna; Mb = Mnb;
            float pa = 0.0f, pb = 0.0f;
#pragma unroll
            for (int nt = 0; nt < 2; ++nt) {
                s[nt][0] = __expf(s[nt][0] - Msa); s[nt][1] = __expf(s[nt][1] - Msa);
                s[nt][2] = __expf(s[nt][2] - Msb); s[nt][3] = __expf(s[nt][3] - Msb);
                pa += s[nt][0] + s[nt][1]; pb += s[nt][2] + s[nt][3];
            }
            la = la * ca + pa; lb = lb * cb + pb;
            if (__any_sync(0xffffffffu, ca != 1.0f || cb != 1.0f)) {
#pragma unroll
                for (int i = 0; i < NT; ++i) { o[i][0] *= ca; o[i][1] *= ca; o[i][2] *= cb; o[i][3] *= cb; }
            }
            // O += P V: P (16 x 16 keys) as the A operand straight from the S accumulators
            const unsigned p0 = pack_h2(s[0][0], s[0][1]), p1 = pack_h2(s[0][2], s[0][3]), p2 = pack_h2(s[1][0], s[1][1]), p3 = pack_h2(s[1][2], s[1][3]);
#pragma unroll
            for (int j = 0; j < NT / 2; ++j) {
                const int chunk = 2 * j + (mi >> 1);
                const int row = 8 * (mi & 1) + lr;
                unsigned b0, b1, b2, b3;
                ldsm_x4_t(st + (NB + (chunk >> 3)) * BOX + row * 128 + (((chunk & 7) ^ (row & 7)) << 4), b0, b1, b2, b3);
                mma_f16(o[2 * j], p0, p1, p2, p3, b0, b1);
                mma_f16(o[2 * j + 1], p0, p1, p2, p3, b2, b3);
            }
        }
        __syncwarp();
        if (tile_i < t_end) { issue(tile_i, stage); tile_i += GA_WARPS; } else livebits &= ~(1u << stage);
        if (++stage == GA_STAGES) stage = 0;
    }

    // row sums across the four lanes of a row
    la += __shfl_xor_sync(0xffffffffu, la, 1); la += __shfl_xor_sync(0xffffffffu, la, 2);
    lb += __shfl_xor_sync(0xffffffffu, lb, 1); lb += __shfl_xor_sync(0xffffffffu, lb, 2);

    // merge the warps through shared memory (the rings are idle now: every issued tile has been consumed)
    __syncthreads();
    float * sacc = (float *) smem;                                    // [GA_WARPS][16][D]
    float * sms  = sacc + GA_WARPS * 16 * D;                          // [GA_WARPS][16][2]
#pragma unroll
    for (int i = 0; i < NT; ++i) {
        *(float2 *) (sacc + (warp * 16 + ra) * D + 8 * i + 2 * t) = make_float2(o[i][0], o[i][1]);
        *(float2 *) (sacc + (warp * 16 + rb) * D + 8 * i + 2 * t) = make_float2(o[i][2], o[i][3]);
    }
    if (t == 0) { sms[(warp * 16 + ra) * 2] = Ma; sms[(warp * 16 + ra) * 2 + 1] = la; sms[(warp * 16 + rb) * 2] = Mb; sms[(warp * 16 + rb) * 2 + 1] = lb; }
    __syncthreads();
    for (int idx = threadIdx.x; idx < x.nrows * (D / 4); idx += blockDim.x) {
        const int r = idx / (D / 4), d4 = (idx - r * (D / 4)) * 4;
        float Mall = -INFINITY;
#pragma unroll
        for (int w = 0; w < GA_WARPS; ++w) Mall = fmaxf(Mall, sms[(w * 16 + r) * 2]);
        float Sall = 0.0f; float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll
        for (int w = 0; w < GA_WARPS; ++w) {
            const float m = sms[(w * 16 + r) * 2];
            const float wgt = m == -INFINITY ? 0.0f : __expf(m - Mall);
            Sall += sms[(w * 16 + r) * 2 + 1] * wgt;
            const float4 v = *(const float4 *) (sacc + (w * 16 + r) * D + d4);
            acc.x += v.x * wgt; acc.y += v.y * wgt; acc.z += v.z * wgt; acc.w += v.w * wgt;
        }
        const int tok = r / x.G; const int64_t h = (int64_t) hkv * x.G + r % x.G;
        if (x.nsplit > 1) {
            float * part = x.partial + ((((int64_t) iq3 * p.H + h) * p.N + tok) * x.nsplit + split) * (D + 2);
            *(float2 *) (part + d4) = make_float2(acc.x, acc.y); *(float2 *) (part + d4 + 2) = make_float2(acc.z, acc.w);     // (D + 2) floats: 8-byte aligned
            if (d4 == 0) { part[D] = Mall; part[D + 1] = Sall; }
        } else {
            float sink_scale = 1.0f;
            if (p.sinks) {                                               // ops.cpp:8241-8255
                const float sk_ = p.sinks[h];
                if (sk_ > Mall) { sink_scale = __expf(Mall - sk_); Sall = Sall * sink_scale + 1.0f; }
                else            { Sall += __expf(sk_ - Mall); }
            }
            const float f = Sall == 0.0f ? 0.0f : sink_scale / Sall;
            float * drow = (float *) ((char *) p.dst + h * p.d_nb1 + tok * p.d_nb2 + iq3 * p.d_nb3);
            *(float4 *) (drow + d4) = make_float4(acc.x * f, acc.y * f, acc.z * f, acc.w * f);
        }
    }
    if (x.nsplit > 1 && x.counters) {
        // the last CTA of this (kv head, s) to arrive merges the splits: no second launch
        __shared__ unsigned s_last;
        __threadfence();
        __syncthreads();
        unsigned * cnt = x.counters + (size_t) iq3 * gridDim.y + hkv;
        if (threadIdx.x == 0) {
            const unsigned old = atomicAdd(cnt, 1u);
            s_last = old == (unsigned) x.nsplit - 1u;
            if (s_last) *cnt = 0u;                                       // self-resetting: the next launch finds zeros
        }
        __syncthreads();
        if (!s_last) return;
        __threadfence();
        for (int idx = threadIdx.x; idx < x.nrows * (D / 4); idx += blockDim.x) {
            const int r = idx / (D / 4), d4 = (idx - r * (D / 4)) * 4;
            const int tok = r / x.G; const int64_t h = (int64_t) hkv * x.G + r % x.G;
            const float * part = x.partial + (((int64_t) iq3 * p.H + h) * p.N + tok) * x.nsplit * (D + 2);
            float Mall = -INFINITY;
            for (int sp = 0; sp < x.nsplit; ++sp) Mall = fmaxf(Mall, __ldcg(part + sp * (D + 2) + D));
            float Sall = 0.0f; float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
            for (int sp = 0; sp < x.nsplit; ++sp) {
                const float m = __ldcg(part + sp * (D + 2) + D);
                if (m == -INFINITY) continue;
                const float wgt = __expf(m - Mall);
                Sall += __ldcg(part + sp * (D + 2) + D + 1) * wgt;
                const float2 v0 = __ldcg((const float2 *) (part + sp * (D + 2) + d4)), v1 = __ldcg((const float2 *) (part + sp * (D + 2) + d4 + 2));
                acc.x += v0.x * wgt; acc.y += v0.y * wgt; acc.z += v1.x * wgt; acc.w += v1.y * wgt;
            }
            float sink_scale = 1.0f;
            if (p.sinks) {
                const float sk_ = p.sinks[h];
                if (sk_ > Mall) { sink_scale = __expf(Mall - sk_); Sall = Sall * sink_scale + 1.0f; }
                else            { Sall += __expf(sk_ - Mall); }
            }
            const float f = Sall == 0.0f ? 0.0f : sink_scale / Sall;
            float * drow = (float *) ((char *) p.dst + h * p.d_nb1 + tok * p.d_nb2 + iq3 * p.d_nb3);
            *(float4 *) (drow + d4) = make_float4(acc.x * f, acc.y * f, acc.z * f, acc.w * f);
        }
    }
}

// ---- host side -------------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn b200_get_encode_tiled();
int b200_launch_fattn_combine(b200_ctx * ctx, const fattn_args & p, int64_t S, int nsplit, float * partial);      // b200_fattn_dec.cu

bool b200_fattn_gqa_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst) {
    static const int off = getenv("GGML_B200_FATTN_NO_GQA") != nullptr;
    if (off) return false;
    const int64_t D = k->ne[0];
    if ((D != 64 && D != 128) || v->ne[0] != D || k->ne[2] != v->ne[2] || k->ne[3] != v->ne[3]) return false;
    const int64_t G = q->ne[2] / k->ne[2];
    if (q->ne[1] < 1 || q->ne[1] > 8 || q->ne[1] * G > 16) return false;
    // one query token over a short context is pure launch-to-result latency, and there the gather kernel (b200_fattn_dec.cu, more and
    // smaller CTAs, no tensor-map fetch) is ~4 us per layer quicker inside the token's kernel chain (tg128 at n_kv = 256: 662 vs 610 tok/s)
    static const int force = getenv("GGML_B200_FATTN_GQA") != nullptr;
    if (!force && q->ne[1] == 1 && k->ne[1] <= 512) return false;
    if (((uintptr_t) q->data & 7) || (q->nb[1] & 7) || (q->nb[2] & 7) || (q->nb[3] & 7)) return false;
    if (((uintptr_t) k->data & 15) || (k->nb[1] & 15) || (k->nb[2] & 15) || (k->nb[3] & 15)) return false;
    if (((uintptr_t) v->data & 15) || (v->nb[1] & 15) || (v->nb[2] & 15) || (v->nb[3] & 15)) return false;
    if (((uintptr_t) dst->data & 15) || (dst->nb[1] & 15) || (dst->nb[2] & 15) || (dst->nb[3] & 15)) return false;
    if (mask && (((uintptr_t) mask->data & 3) || (mask->nb[1] & 3) || (mask->nb[2] & 3) || (mask->nb[3] & 3))) return false;
    // one mask row per token for the whole group: heads of a group must share it (ne[2] == 1, the llama.cpp case)
    if (mask && mask->ne[2] != 1) return false;
    if (k->ne[1] > ((int64_t) 1 << 30) || k->ne[2] > 65535 || q->ne[3] > 65535) return false;
    return b200_get_encode_tiled() != nullptr;
}

// 4-D fp16 map over a K or V view [D, n_kv, n_head_kv, s]; the two middle dimensions are ordered by stride (the KV cache interleaves
// heads inside a token row, plain tensors do not); box = 64 values x GA_TILE tokens of one head
static bool ga_make_map(encode_tiled_fn encode, CUtensorMap * map, const b200_tensor * t, int * swap, int64_t tokens) {
    *swap = (t->ne[2] > 1 && t->nb[2] < t->nb[1]) ? 1 : 0;
    const int a = *swap ? 2 : 1, b = *swap ? 1 : 2;
    cuuint64_t dim[4] = { (cuuint64_t) t->ne[0], (cuuint64_t) t->ne[a], (cuuint64_t) t->ne[b], (cuuint64_t) t->ne[3] };
    dim[*swap ? 2 : 1] = (cuuint64_t) tokens;                              // (paged cache: the rows of the pool, not the keys attended)
    cuuint64_t str[3] = { (cuuint64_t) t->nb[a], (cuuint64_t) t->nb[b], (cuuint64_t) t->nb[3] };
    if (dim[2] == 1) str[1] = str[0] * dim[1];                             // strides of extent-1 dimensions only have to be well formed
    if (dim[3] == 1) str[2] = str[1] * dim[2];
    cuuint32_t box[4] = { 64, 1, 1, 1 }; box[*swap ? 2 : 1] = GA_TILE;
    const cuuint32_t es[4] = { 1, 1, 1, 1 };
    return encode(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT16, 4, t->data, dim, str, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS;
}

// returns B200_ERR_UNSUPPORTED (without setting the error text) when the tensor maps cannot be built: the caller falls back
int b200_launch_fattn_gqa(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                          const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                          float scale, float max_bias, float softcap, const b200_kv_pages * pages) {
    encode_tiled_fn encode = b200_get_encode_tiled();
    if (!encode) return B200_ERR_UNSUPPORTED;
    // with a page table the rows past the last key of a partly filled tile are real pool memory, not zero fill: whole tiles only
    if (pages && (k->ne[1] % GA_TILE != 0 || pages->page_tokens % GA_TILE != 0)) return B200_ERR_UNSUPPORTED;
    fattn_args p;
    fattn_fill_args(p, q, k, v, mask, sinks, dst, scale, max_bias, softcap);
    fattn_set_pages(p, pages);
    const int64_t S = q->ne[3], Hkv = k->ne[2];
    if (p.N == 0 || p.H == 0 || S == 0) return B200_OK;
    ga_extra x;
    x.G = (int) (p.H / Hkv); x.nrows = (int) p.N * x.G;
    CUtensorMap tk, tv;
    const int64_t rows = pages ? pages->pool_tokens : k->ne[1];
    if (!ga_make_map(encode, &tk, k, &x.k_swap, rows) || !ga_make_map(encode, &tv, v, &x.v_swap, rows)) return B200_ERR_UNSUPPORTED;
    x.ntiles = (int) ((p.KV + GA_TILE - 1) / GA_TILE);
    // A 16-key tile is ~0.9 us of dependent ldmatrix / mma / softmax work for its warp and the split merge costs ~4 us once, so contexts
    // are cut finely: as many splits as one wave of CTAs allows (one CTA per SM, >= 1 tile per warp), then the FEWEST splits that
    // still give the same number of tile rounds per warp (8k keys: 16 even splits, not 18 ragged ones).
    // Measured alternatives, all slower: 4 warps x 2 CTAs per SM; >= 2 / 4 / 8 / 16 tiles per warp; a second wave of CTAs.
    const int64_t groups = Hkv * S;
    int cap = (int) (ctx->sm_count / groups);
    if (cap > x.ntiles / GA_WARPS) cap = x.ntiles / GA_WARPS;
    if (cap < 1) cap = 1;
    auto rounds = [&](int n) { return ((x.ntiles + n - 1) / n + GA_WARPS - 1) / GA_WARPS; };
    int nsplit = cap;
    while (nsplit > 1 && rounds(nsplit - 1) == rounds(cap)) --nsplit;
    x.tiles_per_split = (x.ntiles + nsplit - 1) / nsplit;
    x.nsplit = (x.ntiles + x.tiles_per_split - 1) / x.tiles_per_split;
    x.partial = nullptr; x.counters = nullptr;
    if (x.nsplit > 1) {
        x.partial = (float *) b200_scratch(ctx, 1, (size_t) p.N * p.H * S * x.nsplit * (p.DV + 2) * sizeof(float));
        if (!x.partial) return B200_ERR_ALLOC;
        if (groups <= 16384) {                                             // fixed-size, zeroed once, self-resetting
            const bool fresh = ctx->scratch[5] == nullptr;
            x.counters = (unsigned *) b200_scratch(ctx, 5, 16384 * sizeof(unsigned));
            if (!x.counters) return B200_ERR_ALLOC;
            if (fresh) B200_CUDA_TRY(cudaMemsetAsync(x.counters, 0, 16384 * sizeof(unsigned), ctx->stream));
        }
    }
    const int D = p.DK;
    const size_t stage = (size_t) 2 * (D / 64) * GA_TILE * 128;
    size_t smem = (size_t) GA_WARPS * GA_STAGES * stage;
    const size_t merge = (size_t) GA_WARPS * 16 * (D + 2) * 4;
    if (merge > smem) smem = merge;
    x.flags_per_warp = ((x.tiles_per_split + GA_WARPS - 1) / GA_WARPS + 8 + 15) & ~15;
    smem += GA_WARPS * GA_STAGES * 8 + (size_t) GA_WARPS * x.flags_per_warp + 1024;
    if (smem > 220 * 1024) return B200_ERR_UNSUPPORTED;
    if (!ctx->attr_done[7]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_gqa<64>,  cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_gqa<128>, cudaFuncAttributeMaxDynamicSharedMemorySize, 220 * 1024));
        ctx->attr_done[7] = 1;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) x.nsplit, (unsigned) Hkv, (unsigned) S);
    cfg.blockDim = dim3(GA_WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    // opt-in: with 192 KB of rings per CTA nothing can co-reside, and in the token chain PDL measured 2 % slower
    static const int use_pdl = getenv("GGML_B200_FATTN_PDL") != nullptr;
    cfg.attrs = attr; cfg.numAttrs = (ctx->pdl && use_pdl) ? 1 : 0;
    if (D == 64) B200_LAUNCH(ctx, B200_K_FATTN, cudaLaunchKernelEx(&cfg, k_fattn_gqa<64>, p, x, tk, tv));
    else         B200_LAUNCH(ctx, B200_K_FATTN, cudaLaunchKernelEx(&cfg, k_fattn_gqa<128>, p, x, tk, tv));
    if (x.nsplit > 1 && !x.counters) return b200_launch_fattn_combine(ctx, p, S, x.nsplit, x.partial);
    return B200_OK;
}
