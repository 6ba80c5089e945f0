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

#define GS_WARPS  4
#define GS_STAGES 6
#define GS_RW     2       // rows per warp per stage

struct gs_args {
    const uint8_t * w; size_t w_nb1;        // weight rows (GGUF blocks, or the group-SoA repack for q6_K/q8_0/q4_0)
    const int8_t *  aq;                     // q8_1 image: quants [K]
    const __half2 * ads;                    //             (d, s) [K/32]
    float * dst;                            // [M]
    int64_t M; int K;
};

template <int TYPE> struct gs_traits;

// ---- q4_K, raw GGUF blocks (144 B, every field 16-byte aligned) ---------------------------------------------------------
// chunk = 8 blocks = 1152 B per row.  unit = 16 B of qs = 16 weights of sub-block 2p (low nibbles) + 16 of 2p+1 (high).
// Shared-memory reads: lanes of a quarter-warp read 16 B at stride 144 B = 36 words -> banks 4i..4i+3: conflict-free.
// Activation image in shared memory: plane 0 = the even q8_1 blocks, plane 1 = the odd ones, so unit U reads
// plane0[16 U .. 16 U + 15] and plane1[16 U ..]: consecutive lanes, consecutive 16-byte words.
template <> struct gs_traits<B200_TYPE_Q4_K> {
    static constexpr int BLOCK_BYTES = 144, CHUNK_BLOCKS = 8, UNITS_PER_BLOCK = 8;
    static __device__ __forceinline__ int act_dst_offset(int j, int K) {     // j = index of a 16-byte word of the natural image
        const int blk = j >> 1;
        return (blk & 1) * (K >> 1) + (blk >> 1) * 32 + (j & 1) * 16;
    }
    // one unit of one row: returns the partial dot product
    static __device__ __forceinline__ float unit(uint32_t row_smem, int b, int c, const int4 & a0, const int4 & a1, const float2 & ds0, const float2 & ds1) {
        const uint32_t blk = row_smem + b * 144;
        const int4 hdr = lds128(blk);
        const int4 q   = lds128(blk + 16 + c * 16);
        const float2 dm = __half22float2(*(const __half2 *) &hdr.x);
        const int p = c >> 1, j0 = 2 * p;
        const unsigned s0 = (unsigned) hdr.y, s1 = (unsigned) hdr.z, s2 = (unsigned) hdr.w;
        int sc0, m0, sc1, m1;                                   // get_scale_min_k4 (ggml-quants.c:703-710) on packed words
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
        int i0 = __dp4a(q.x & 0x0F0F0F0F, a0.x, 0); i0 = __dp4a(q.y & 0x0F0F0F0F, a0.y, i0);
        i0 = __dp4a(q.z & 0x0F0F0F0F, a0.z, i0);    i0 = __dp4a(q.w & 0x0F0F0F0F, a0.w, i0);
        int i1 = __dp4a((q.x >> 4) & 0x0F0F0F0F, a1.x, 0); i1 = __dp4a((q.y >> 4) & 0x0F0F0F0F, a1.y, i1);
        i1 = __dp4a((q.z >> 4) & 0x0F0F0F0F, a1.z, i1);    i1 = __dp4a((q.w >> 4) & 0x0F0F0F0F, a1.w, i1);
        float r = dm.x * ((float) sc0 * ds0.x * (float) i0 + (float) sc1 * ds1.x * (float) i1);
        // the min term uses the whole q8_1 block sum s = d * sum(q): the even half-unit of each sub-block pair adds it
        if ((c & 1) == 0) r -= dm.y * ((float) m0 * ds0.y + (float) m1 * ds1.y);
        return r;
    }
};

template <int TYPE>
__global__ void __launch_bounds__(GS_WARPS * 32) k_gemv_stream(const gs_args p) {
    using T = gs_traits<TYPE>;
    constexpr int CHUNK_BYTES = T::BLOCK_BYTES * T::CHUNK_BLOCKS;
    constexpr int STAGE_BYTES = CHUNK_BYTES * GS_RW;
    extern __shared__ __align__(128) uint8_t smem[];
    // layout: [GS_WARPS][GS_STAGES][STAGE_BYTES] rings | act planes [K] | ds [K/32] half2 | mbarriers
    uint8_t *  ring_all = smem;
    uint8_t *  s_aq     = ring_all + GS_WARPS * GS_STAGES * STAGE_BYTES;
    __half2 *  s_ds     = (__half2 *) (s_aq + p.K);
    uint64_t * s_bar    = (uint64_t *) ((uint8_t *) s_ds + (p.K >> 5) * 4 + ((16 - (((p.K >> 5) * 4) & 15)) & 15));

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ring = smem_u32(ring_all + warp * GS_STAGES * STAGE_BYTES);
    const uint32_t bar0 = smem_u32(s_bar + warp * GS_STAGES);

    // this warp's rows: an even split of [0, M) over every warp of the grid
    const int64_t gw = (int64_t) blockIdx.x * GS_WARPS + warp, nw = (int64_t) gridDim.x * GS_WARPS;
    const int64_t r_begin = p.M * gw / nw, r_end = p.M * (gw + 1) / nw;
    const int nblocks = (TYPE == B200_TYPE_Q4_K || TYPE == B200_TYPE_Q5_K || TYPE == B200_TYPE_Q6_K) ? p.K / 256 : p.K / 32;
    const int nchunks = (nblocks + T::CHUNK_BLOCKS - 1) / T::CHUNK_BLOCKS;
    const int64_t npairs = (r_end - r_begin + GS_RW - 1) / GS_RW;
    const int64_t nitems = npairs * nchunks;

    if (lane == 0) {
        for (int s = 0; s < GS_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncwarp();

    const uint64_t policy = l2_policy_evict_first();
    auto issue = [&](int64_t item) {                                     // lane 0 only
        const int     s  = (int) (item % GS_STAGES);
        const int64_t rp = item / nchunks;
        const int     kc = (int) (item - rp * nchunks);
        const int     cb = min(T::CHUNK_BLOCKS, nblocks - kc * T::CHUNK_BLOCKS);
        const uint32_t bytes = (uint32_t) cb * T::BLOCK_BYTES;
        const int64_t row0 = r_begin + rp * GS_RW;
        const int     nr = (int) min((int64_t) GS_RW, r_end - row0);
        mbar_arrive_expect_tx(bar0 + 8 * s, bytes * nr);
        for (int r = 0; r < nr; ++r)
            bulk_g2s_hint(ring + s * STAGE_BYTES + r * CHUNK_BYTES, p.w + (row0 + r) * p.w_nb1 + (size_t) kc * CHUNK_BYTES, bytes, bar0 + 8 * s, policy);
    };
    if (lane == 0) for (int64_t i = 0; i < nitems && i < GS_STAGES; ++i) issue(i);

    // activation image -> shared memory, permuted into the order the lanes read it (overlaps the first bulk copies)
    {
        const int4 * src = (const int4 *) p.aq;
        for (int j = threadIdx.x; j < p.K / 16; j += blockDim.x) {
            const int4 v = src[j];
            *(int4 *) (s_aq + T::act_dst_offset(j, p.K)) = v;
        }
        for (int j = threadIdx.x; j < p.K / 32; j += blockDim.x) s_ds[j] = p.ads[j];
    }
    __syncthreads();

    const uint32_t aq0 = smem_u32(s_aq), aq1 = aq0 + (p.K >> 1), ads = smem_u32(s_ds);
    float acc[GS_RW];
#pragma unroll
    for (int r = 0; r < GS_RW; ++r) acc[r] = 0.0f;

    for (int64_t item = 0; item < nitems; ++item) {
        const int     s  = (int) (item % GS_STAGES);
        const int64_t rp = item / nchunks;
        const int     kc = (int) (item - rp * nchunks);
        const int     cb = min(T::CHUNK_BLOCKS, nblocks - kc * T::CHUNK_BLOCKS);
        const int64_t row0 = r_begin + rp * GS_RW;
        const int     nr = (int) min((int64_t) GS_RW, r_end - row0);
        mbar_wait(bar0 + 8 * s, (uint32_t) ((item / GS_STAGES) & 1));

        const uint32_t st = ring + s * STAGE_BYTES;
        for (int u = lane; u < cb * T::UNITS_PER_BLOCK; u += 32) {
            const int b = u / T::UNITS_PER_BLOCK, c = u % T::UNITS_PER_BLOCK;
            const int gb = kc * T::CHUNK_BLOCKS + b;                      // block index in the row
            const int U  = gb * T::UNITS_PER_BLOCK + c;
            const int4 a0 = lds128(aq0 + 16 * U), a1 = lds128(aq1 + 16 * U);
            const int2 dsr = lds64(ads + (gb * 8 + 2 * (c >> 1)) * 4);
            const float2 ds0 = __half22float2(*(const __half2 *) &dsr.x), ds1 = __half22float2(*(const __half2 *) &dsr.y);
#pragma unroll
            for (int r = 0; r < GS_RW; ++r)
                if (r < nr) acc[r] += T::unit(st + r * CHUNK_BYTES, b, c, a0, a1, ds0, ds1);
        }
        __syncwarp();                                                     // every lane is done reading this stage
        if (lane == 0 && item + GS_STAGES < nitems) issue(item + GS_STAGES);

        if (kc == nchunks - 1) {
#pragma unroll
            for (int r = 0; r < GS_RW; ++r) {
                const float v = warp_sum(acc[r]);
                if (lane == 0 && r < nr) p.dst[row0 + r] = v;
                acc[r] = 0.0f;
            }
        }
    }
}

static size_t gs_smem_bytes(int chunk_bytes, int K) {
    size_t n = (size_t) GS_WARPS * GS_STAGES * chunk_bytes * GS_RW + K;
    size_t ds = (size_t) (K >> 5) * 4; ds = (ds + 15) & ~(size_t) 15;
    return n + ds + GS_WARPS * GS_STAGES * 8;
}

// true when the streaming kernel can take this MUL_MAT (one activation column, 2-D weight, aligned rows)
bool b200_gemv_stream_supported(const b200_tensor * w, int64_t ne11, int64_t ne12, int64_t ne13) {
    if (ne11 != 1 || ne12 != 1 || ne13 != 1 || w->ne[2] != 1 || w->ne[3] != 1) return false;
    if (w->type != B200_TYPE_Q4_K) return false;
    if (((uintptr_t) w->data & 15) || (w->nb[1] & 15) || w->ne[0] % 256 != 0 || w->ne[0] > 65536) return false;
    return gs_smem_bytes(144 * 8, (int) w->ne[0]) <= 200 * 1024;
}

int b200_launch_gemv_stream(b200_ctx * ctx, const b200_tensor * w, const void * image, float * dst) {
    gs_args p;
    const int K = (int) w->ne[0];
    p.w = (const uint8_t *) w->data; p.w_nb1 = w->nb[1];
    p.aq = (const int8_t *) image; p.ads = (const __half2 *) ((const int8_t *) image + K);
    p.dst = dst; p.M = w->ne[1]; p.K = K;
    const size_t smem = gs_smem_bytes(144 * 8, K);
    if (!ctx->attr_done[0]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemv_stream<B200_TYPE_Q4_K>, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024));
        ctx->attr_done[0] = 1;
    }
    // enough CTAs that every warp gets >= GS_RW rows, at most 3 resident CTAs per SM
    const int64_t want = (p.M + GS_WARPS * GS_RW - 1) / (GS_WARPS * GS_RW);
    const int per_sm = (int) max((size_t) 1, min((size_t) 3, (size_t) (220 * 1024) / (smem + 1024)));
    const int64_t cap = (int64_t) ctx->sm_count * per_sm;
    const unsigned grid = (unsigned) max((int64_t) 1, min(want, cap));
    B200_LAUNCH(ctx, B200_K_GEMV, k_gemv_stream<B200_TYPE_Q4_K><<<grid, GS_WARPS * 32, smem, ctx->stream>>>(p));
    return B200_OK;
}
