// b200_gemv_stream.cu -- the decode dequant-GEMV (MUL_MAT with one activation column), streamed through the TMA engine.
//
// Replaces the CPU inner loop ggml_vec_dot_q4_K_q8_K / q6_K_q8_K / q8_0_q8_0 / q4_0_q8_0 driven by
// ggml_compute_forward_mul_mat (reference ggml/src/ggml-cpu/ggml-cpu.c:1136-1224, quants.c:115,305,550,708).
//
// Bound: HBM.  Algorithmic bytes per launch = M * K * bytes_per_block / block_size (each weight byte read once).
//
// Why this shape (measured, profiles/r1_run1_*): a plain "one warp per row, 128-bit loads" kernel reached 20 % of HBM on
// B200 -- with ~30 resident warps per SM and two 16-byte loads in flight per lane it cannot cover the ~1 us loaded HBM
// latency at 44 B/ns/SM.  Here every warp is its own streaming engine: lane 0 issues cp.async.bulk copies (the TMA engine,
// no registers, no LSU slots) of RW row-chunks into a private STAGES-deep shared-memory ring, completion is signalled on
// an mbarrier per stage, and the 32 lanes consume a stage with conflict-free 128-bit shared loads while the next
// STAGES-1 stages are in flight (~14 KB per warp, ~160 KB per SM).  The q8_1 activation image (L2 resident, produced by
// b200_quantize.cu) is copied once per CTA into shared memory in the order the lanes read it.
// Rows are split evenly over all warps of the grid (<= 1 row of imbalance), so odd row counts cost nothing.
#include "b200_common.cuh"
#include "b200_ptx.cuh"

#define GS_WARPS  8
#define GS_STAGES 4
#define GS_RW     2       // rows per warp per stage
#define GS_CB     8       // q4_K blocks per row-chunk (1152 B)

struct gs_args {
    const uint8_t * w; size_t w_nb1;        // weight rows (raw GGUF q4_K blocks)
    const int8_t *  aq;                     // q8_1 image: quants [K]
    const __half2 * ads;                    //             (d, s) [K/32]
    float * dst;                            // [M]
    int M; int K;
};

// Lane -> work mapping inside one stage (a chunk of up to 8 blocks of RW rows):
//   block b = lane & 7; iteration it in {0, 1}; c = (lane >> 3) + 4 it is the 16-byte unit of the block's qs,
//   i.e. sub-block pair p = c >> 1 = (lane >> 4) + 2 it, half = (lane >> 3) & 1.
// Every lane of the warp is in the same branch of get_scale_min_k4 (ggml-quants.c:703-710) for a given `it` (p < 2 for
// it = 0, p >= 2 for it = 1), so the 6-bit scale decode is straight-line code with lane-constant shifts.
// Shared-memory reads are conflict-free: a quarter-warp reads 8 blocks at stride 144 B = 36 words (banks 4i..4i+3).
// The activation image is laid out in shared memory as the lanes read it:
//   quants: plane0[c][gb] / plane1[c][gb], 16 B each -- the 16 int8 that meet the low / high nibbles of unit (gb, c);
//   scales: ds4[p][gb] = (dA(2p), s(2p), dA(2p+1)/16, s(2p+1)) as four floats -- the /16 undoes the unshifted high
//   nibbles (dp4a on q & 0xF0F0F0F0 yields 16 x the sum).
template <int NR, int IT>
static __device__ __forceinline__ void gs_q4k_unit(uint32_t blk, int cq, int sh, int shm, int half,
                                                   const int4 & a0, const int4 & a1, const float4 & ds, float ssel,
                                                   float (&accd)[GS_RW], float (&accm)[GS_RW]) {
#pragma unroll
    for (int r = 0; r < NR; ++r) {
        const uint32_t base = blk + r * (GS_CB * 144);
        const int4 hdr = lds128(base);
        const int4 q   = lds128(base + 16 + (cq + 4 * IT) * 16);
        const unsigned s0 = (unsigned) hdr.y, s1 = (unsigned) hdr.z, s2 = (unsigned) hdr.w;
        int sc0, sc1, msel;
        if (IT == 0) {
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
        const float t = (float) (sc0 * i0) * ds.x + (float) (sc1 * i1) * ds.z;
        accd[r] = fmaf(dm.x, t, accd[r]);
        accm[r] = fmaf(dm.y, (float) msel * ssel, accm[r]);
    }
    (void) half;
}

__global__ void __launch_bounds__(GS_WARPS * 32, 2) k_gemv_stream_q4k(const gs_args p) {
    constexpr int CHUNK_BYTES = 144 * GS_CB;
    constexpr int STAGE_BYTES = CHUNK_BYTES * GS_RW;
    extern __shared__ __align__(128) uint8_t smem[];
    // layout: [GS_WARPS][GS_STAGES][STAGE_BYTES] rings | quant planes [K] | ds4 [K/64] float4 | mbarriers
    const int nblocks = p.K >> 8;
    uint8_t *  s_aq  = smem + GS_WARPS * GS_STAGES * STAGE_BYTES;
    float4 *   s_ds  = (float4 *) (s_aq + p.K);
    uint64_t * s_bar = (uint64_t *) (s_ds + nblocks * 4);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ring = smem_u32(smem + warp * GS_STAGES * STAGE_BYTES);
    const uint32_t bar0 = smem_u32(s_bar + warp * GS_STAGES);

    // this warp's rows: an even split of [0, M) over every warp of the grid
    const int gw = blockIdx.x * GS_WARPS + warp, nw = gridDim.x * GS_WARPS;
    const int r_begin = (int) ((int64_t) p.M * gw / nw), r_end = (int) ((int64_t) p.M * (gw + 1) / nw);
    const int rows = r_end - r_begin;
    const int nchunks = (nblocks + GS_CB - 1) / GS_CB;
    const int npairs = (rows + GS_RW - 1) / GS_RW;
    const int nitems = npairs * nchunks;

    if (lane == 0) {
        for (int s = 0; s < GS_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncwarp();

    const uint64_t policy = l2_policy_evict_first();
    const uint8_t * wbase = p.w + (size_t) r_begin * p.w_nb1;
    // producer cursor (uniform across the warp; only lane 0 issues)
    int prp = 0, pkc = 0;
    auto issue = [&](int stage) {
        const int cb = min(GS_CB, nblocks - pkc * GS_CB);
        const int nr = min(GS_RW, rows - prp * GS_RW);
        if (lane == 0) {
            const uint32_t bytes = (uint32_t) cb * 144, bar = bar0 + 8 * stage;
            mbar_arrive_expect_tx(bar, bytes * nr);
            const uint8_t * src = wbase + (size_t) (prp * GS_RW) * p.w_nb1 + (size_t) pkc * CHUNK_BYTES;
            bulk_g2s_hint(ring + stage * STAGE_BYTES, src, bytes, bar, policy);
            if (nr > 1) bulk_g2s_hint(ring + stage * STAGE_BYTES + CHUNK_BYTES, src + p.w_nb1, bytes, bar, policy);
        }
        if (++pkc == nchunks) { pkc = 0; ++prp; }
    };
    const int npro = min(nitems, GS_STAGES);
    for (int i = 0; i < npro; ++i) issue(i);

    // activation image -> shared memory in lane order (overlaps the first bulk copies)
    {
        const int4 * src = (const int4 *) p.aq;
        for (int j = threadIdx.x; j < (p.K >> 4); j += blockDim.x) {       // j: 16-byte word of the natural image
            const int blk32 = j >> 1, hf = j & 1, gb = blk32 >> 3, sub = blk32 & 7;
            const int c = (sub >> 1) * 2 + hf;
            *(int4 *) (s_aq + (sub & 1) * (p.K >> 1) + (c * nblocks + gb) * 16) = src[j];
        }
        for (int j = threadIdx.x; j < (p.K >> 6); j += blockDim.x) {       // j: pair of q8_1 blocks
            const int gb = j >> 2, pp = j & 3;
            const float2 e = __half22float2(p.ads[2 * j]), o = __half22float2(p.ads[2 * j + 1]);
            s_ds[pp * nblocks + gb] = make_float4(e.x, e.y, o.x * 0.0625f, o.y);
        }
    }
    __syncthreads();

    const int bl = lane & 7, cq = lane >> 3, half = cq & 1;
    const int sh = 16 * (lane >> 4), shm = sh + 8 * half;
    const uint32_t aq0 = smem_u32(s_aq), aq1 = aq0 + (p.K >> 1), ads = smem_u32(s_ds);
    float accd[GS_RW], accm[GS_RW];
#pragma unroll
    for (int r = 0; r < GS_RW; ++r) accd[r] = accm[r] = 0.0f;

    int stage = 0, phase = 0, rp = 0, kc = 0;
    for (int item = 0; item < nitems; ++item) {
        const int cb = min(GS_CB, nblocks - kc * GS_CB);
        const int nr = min(GS_RW, rows - rp * GS_RW);
        mbar_wait(bar0 + 8 * stage, (uint32_t) phase);

        if (bl < cb) {
            const int gb = kc * GS_CB + bl;
            const uint32_t blk = ring + stage * STAGE_BYTES + bl * 144;
#pragma unroll
            for (int it = 0; it < 2; ++it) {
                const int c = cq + 4 * it, pp = (lane >> 4) + 2 * it;
                const int4 a0 = lds128(aq0 + (c * nblocks + gb) * 16), a1 = lds128(aq1 + (c * nblocks + gb) * 16);
                const int4 dsr = lds128(ads + (pp * nblocks + gb) * 16);
                const float4 ds = make_float4(__int_as_float(dsr.x), __int_as_float(dsr.y), __int_as_float(dsr.z), __int_as_float(dsr.w));
                const float ssel = half ? ds.w : ds.y;
                if (nr == GS_RW) { if (it == 0) gs_q4k_unit<GS_RW, 0>(blk, cq, sh, shm, half, a0, a1, ds, ssel, accd, accm);
                                   else         gs_q4k_unit<GS_RW, 1>(blk, cq, sh, shm, half, a0, a1, ds, ssel, accd, accm); }
                else             { if (it == 0) gs_q4k_unit<1, 0>(blk, cq, sh, shm, half, a0, a1, ds, ssel, accd, accm);
                                   else         gs_q4k_unit<1, 1>(blk, cq, sh, shm, half, a0, a1, ds, ssel, accd, accm); }
            }
        }
        __syncwarp();                                                     // every lane is done reading this stage
        if (item + GS_STAGES < nitems) issue(stage);

        if (kc == nchunks - 1) {
#pragma unroll
            for (int r = 0; r < GS_RW; ++r) {
                const float v = warp_sum(accd[r] - accm[r]);
                if (lane == 0 && r < nr) p.dst[r_begin + rp * GS_RW + r] = v;
                accd[r] = accm[r] = 0.0f;
            }
            kc = 0; ++rp;
        } else ++kc;
        if (++stage == GS_STAGES) { stage = 0; phase ^= 1; }
    }
}

static size_t gs_smem_bytes(int K) {
    return (size_t) GS_WARPS * GS_STAGES * 144 * GS_CB * GS_RW + K + (size_t) (K >> 8) * 4 * 16 + GS_WARPS * GS_STAGES * 8;
}

// true when the streaming kernel can take this MUL_MAT (one activation column, 2-D weight, aligned rows)
bool b200_gemv_stream_supported(const b200_tensor * w, int64_t ne11, int64_t ne12, int64_t ne13) {
    if (ne11 != 1 || ne12 != 1 || ne13 != 1 || w->ne[2] != 1 || w->ne[3] != 1) return false;
    if (w->type != B200_TYPE_Q4_K) return false;
    if (((uintptr_t) w->data & 15) || (w->nb[1] & 15) || w->ne[0] % 256 != 0 || w->ne[0] > 32768 || w->ne[1] > (1 << 30)) return false;
    return gs_smem_bytes((int) w->ne[0]) <= 110 * 1024;
}

int b200_launch_gemv_stream(b200_ctx * ctx, const b200_tensor * w, const void * image, float * dst) {
    gs_args p;
    const int K = (int) w->ne[0];
    p.w = (const uint8_t *) w->data; p.w_nb1 = w->nb[1];
    p.aq = (const int8_t *) image; p.ads = (const __half2 *) ((const int8_t *) image + K);
    p.dst = dst; p.M = (int) w->ne[1]; p.K = K;
    const size_t smem = gs_smem_bytes(K);
    if (!ctx->attr_done[0]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemv_stream_q4k, cudaFuncAttributeMaxDynamicSharedMemorySize, 110 * 1024));
        ctx->attr_done[0] = 1;
    }
    // persistent-style grid: two CTAs per SM, rows split evenly over all warps; fewer CTAs when there are few rows
    const int64_t want = (p.M + GS_WARPS - 1) / GS_WARPS;
    const unsigned grid = (unsigned) max((int64_t) 1, min(want, (int64_t) ctx->sm_count * 2));
    B200_LAUNCH(ctx, B200_K_GEMV, k_gemv_stream_q4k<<<grid, GS_WARPS * 32, smem, ctx->stream>>>(p));
    return B200_OK;
}
