// Decode MUL_MAT (one activation column) for raw GGUF block_q4_K weights: TMA-tiled weight stream + int8 tensor-core dot products.
// Replaces the per-row loop ggml_vec_dot_q4_K_q8_K (ggml/src/ggml-cpu/quants.c:550, x86 arch/x86/quants.c:1741) driven by
// ggml_compute_forward_mul_mat (ggml/src/ggml-cpu/ggml-cpu.c:1226-1418), with q8_1 activations (quantize_row_q8_1,
// arch/x86/quants.c:388-490, bit-exact) as ggml-cuda's mul_mat_vec_q does.
//
// Why tensor cores for a GEMV: the kernel is HBM-bound, but the dp4a version (b200_gemv_stream.cu) spends ~94 issue slots per 512
// weight bytes on nibble masks, dp4a and the 6-bit scale decode, and ncu shows the issue slots, not HBM, as the limiter.  One
// mma.sync.m16n8k32 (u8 x s8 -> s32) does a whole 8-row x 64-weight group, which cuts the instruction count ~2.8x:
//   A (16 x 32): rows 0-7  = low  nibbles of qs[32p .. 32p+31] of 8 weight rows  -> sub-block 2p
//                rows 8-15 = high nibbles of the same bytes                      -> sub-block 2p+1
//   B (32 x 8):  column c = the 32 int8 activations of sub-block c of this 256-block (the SAME registers for p = 0..3)
//   C[g][2p] and C[g+8][2p+1] are the two wanted sums; both land in lane (g, t = p), so after the four MMAs of a block lane (g, t)
//   owns sub-blocks 2t, 2t+1 of weight row g and applies their scales: all 32 lanes do scale work, none is idle.
//   Fragment k index 4t+i <-> byte 8t+i and 16+4t+i <-> byte 8t+4+i of the 32-byte group, for A and B alike, so both are LDS.64.
// Data movement: one 2-D TMA box {288 B = 2 blocks, 8 rows} per stage straight from the row-major GGUF tensor (sector-aligned:
// 288 = 9 x 32 B), per-warp rings of 3 stages with one mbarrier each; rows beyond M and blocks beyond K arrive as zeros.
// Work split: a CTA owns a contiguous range of 8-row groups; its (group, 2-block chunk) items go round-robin over the 12 warps, each
// item leaves 8 partial sums in shared memory and the CTA adds them in chunk order at the end (deterministic, no atomics).
#include "b200_common.cuh"
#include "b200_ptx.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <mutex>

#define GQ_WARPS   12
#define GQ_STAGES  3
#define GQ_ROWS    8
#define GQ_CHUNK   288                         // bytes of one row in a stage: two 144-byte blocks
#define GQ_STAGE_BYTES (GQ_ROWS * GQ_CHUNK)    // 2304
#define GQ_MAXSEG  3

struct gq_args {
    float * dst[GQ_MAXSEG]; int M[GQ_MAXSEG]; int nseg; int K;
    const float * x; const int8_t * aq; const __half2 * ads;            // as gs_args (b200_gemv_stream.cu)
    int gmax;                                                           // most 8-row groups any CTA owns
    // row-split exchange, see b200_gemv_stream.cu
    int nll; char * ll_base[B200_MAX_PEERS]; const char * ll_self; const char * arena; unsigned * ctrl;
    struct { float * dst; int M; } full[GQ_MAXSEG];
    int dbg;
    int npf;                                                            // items per warp to prefetch into L2 behind the ring fill
    int npre;                                                           // ring stages to fill before the dependency wait
    int pair; float * glu;                                              // pair: two equal-sized segments (ffn_gate, ffn_up) interleaved group by group so that one CTA owns
                                                                        // the same rows of both, and the epilogue also stores glu[row] = swiglu(gate[row], up[row]) (the GLU node behind them)
    int n_first, skew;                                                  // CTAs 0 .. n_first-1 get (1024 + skew) / 1024 of an even share, the others (1024 - skew) / 1024
    unsigned long long * trace; int trace_cap;                          // b200_debug_trace (measurement only)
};
struct gq_maps { CUtensorMap m[GQ_MAXSEG]; };

static __device__ __forceinline__ void tma_load_2d_hint(uint32_t dst, const void * tmap, int c0, int c1, uint32_t bar, uint64_t policy) {
    asm volatile("cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1, {%3, %4}], [%2], %5;"
                 :: "r"(dst), "l"(tmap), "r"(bar), "r"(c0), "r"(c1), "l"(policy) : "memory");
}
static __device__ __forceinline__ void tma_prefetch_l2_2d(const void * tmap, int c0, int c1) {
    asm volatile("cp.async.bulk.prefetch.tensor.2d.L2.global.tile [%0, {%1, %2}];" :: "l"(tmap), "r"(c0), "r"(c1) : "memory");
}
static __device__ __forceinline__ void mma_u8s8(int (&c)[4], unsigned a0, unsigned a1, unsigned a2, unsigned a3, unsigned b0, unsigned b1) {
    asm volatile("mma.sync.aligned.m16n8k32.row.col.s32.u8.s8.s32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+r"(c[0]), "+r"(c[1]), "+r"(c[2]), "+r"(c[3]) : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

// NW = 12: two CTAs per SM (this kernel's, or this and the PDL-launched next one); NW = 24: ONE CTA per SM for long streams -- the same
// warps per SM, but the activation column is fetched and quantized once per SM (L2 broadcast bandwidth, see b200_gemv_stream.cu BIG)
// PAIR (compile time: the mere presence of the branches in the plain kernel cost 2.4 % of the decode graph): see gq_args::pair
template <int NW, bool PAIR = false>
__global__ void __launch_bounds__(NW * 32, NW == 24 ? 1 : 2) k_gemv_q4k_mma(const gq_args p, const __grid_constant__ gq_maps maps) {
    extern __shared__ __align__(128) uint8_t smem[];
    // layout: [NW][GQ_STAGES][2304] rings | q8 quants [K] | (d, s) float2 [K/32] | partial sums [gmax][nchunks][8] | mbarriers
    const int nblocks = p.K >> 8, nchunks = (nblocks + 1) >> 1;
    uint8_t *  s_q   = smem + NW * GQ_STAGES * GQ_STAGE_BYTES;
    float2 *   s_ds  = (float2 *) (s_q + p.K);
    float *    s_par = (float *) (s_ds + (p.K >> 5));
    uint64_t * s_bar = (uint64_t *) (s_par + (size_t) p.gmax * nchunks * GQ_ROWS);

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ring = smem_u32(smem + warp * GQ_STAGES * GQ_STAGE_BYTES);
    const uint32_t bar0 = smem_u32(s_bar + warp * GQ_STAGES);

    // this CTA's 8-row groups, numbered through the segments
    const int G0 = (p.M[0] + GQ_ROWS - 1) / GQ_ROWS, G1 = p.nseg > 1 ? (p.M[1] + GQ_ROWS - 1) / GQ_ROWS : 0, G2 = p.nseg > 2 ? (p.M[2] + GQ_ROWS - 1) / GQ_ROWS : 0;
    const int G = G0 + G1 + G2;
    // With two CTAs per SM the one that arrived first gets the issue slots first (measured: body 8.6 vs 10.8 us for equal shares), then the
    // other runs on alone with half the warps: the first wave of CTAs (blockIdx < n_first, one per SM) takes a larger share.
    const int64_t wa = 1024 + p.skew, wb = 1024 - p.skew, n1 = p.n_first;
    const int64_t W = n1 * wa + ((int64_t) gridDim.x - n1) * wb;
    auto cut = [&](int64_t b) { const int64_t pre = b <= n1 ? b * wa : n1 * wa + (b - n1) * wb; return (int) ((int64_t) G * pre / W); };
    constexpr bool pair = PAIR;
    const int g_lo = pair ? cut(blockIdx.x) & ~1 : cut(blockIdx.x), g_hi = pair ? cut((int64_t) blockIdx.x + 1) & ~1 : cut((int64_t) blockIdx.x + 1);   // (G is even in pair mode)
    const int nitems_cta = (p.dbg & 4) ? 0 : (g_hi - g_lo) * nchunks;

    unsigned long long * rec = nullptr;
    if (p.trace && threadIdx.x == 0) rec = b200_trace_open(p.trace, p.trace_cap, p.M[0] + p.M[1] + p.M[2], p.K | (1 << 30));
    b200_trace_stamp(rec, 0);
    if (lane == 0) {
        for (int s = 0; s < GQ_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncwarp();
    if (p.dbg & 8) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");      // (old placement of the PDL trigger)

    const uint64_t policy = l2_policy_evict_first();
    int pit = warp, pgl = warp / nchunks, pch = warp - pgl * nchunks;     // producer cursor: next item of this warp to fetch = (group, chunk)
    auto issue = [&](int stage) {
        if (lane == 0) {
            int gg = g_lo + pgl, s = 0;
            if (pair) { s = gg & 1; gg >>= 1; } else if (gg >= G0) { gg -= G0; s = 1; if (gg >= G1) { gg -= G1; s = 2; } }
            const uint32_t bar = bar0 + 8 * stage;
            mbar_arrive_expect_tx(bar, GQ_STAGE_BYTES);
            tma_load_2d_hint(ring + stage * GQ_STAGE_BYTES, &maps.m[s], pch * (GQ_CHUNK / 4), gg * GQ_ROWS, bar, policy);
        }
        pit += NW; pch += NW;
        while (pch >= nchunks) { pch -= nchunks; ++pgl; }
    };
    // weights do not depend on the previous kernel: fill the ring before the dependency wait.  Small matrices (the whole matrix fits the
    // rings: p.npre = 1) fetch only one stage now and the rest BEHIND the activation loads, which otherwise queue up behind ~80 KB of
    // ring fill per CTA (measured: wait -> quantized 3800 -> 2900 cycles, qkv 6.9 -> 5.3 us); long streams need the early start more
    const int npre = p.npre;
    int nfilled = 0;
    for (; nfilled < npre && pit < nitems_cta; ++nfilled) issue(nfilled);
    auto fill_rest = [&]() { for (; nfilled < GQ_STAGES && pit < nitems_cta; ++nfilled) issue(nfilled); };
    // The rings hold ~83 KB per CTA; the activations only exist once the previous kernel has retired and are then quantized
    // (~3 us in all) with nothing consuming, so HBM would idle.  Keep it busy: pull this warp's next items into L2 meanwhile
    // (126 MB); the ring refills then hit L2.
    if (lane == 0) {
        int it = pit;
        for (int i = 0; i < p.npf && it < nitems_cta; ++i, it += NW) {
            const int gl = it / nchunks, ch = it - gl * nchunks;
            int gg = g_lo + gl, s = 0;
            if (pair) { s = gg & 1; gg >>= 1; } else if (gg >= G0) { gg -= G0; s = 1; if (gg >= G1) { gg -= G1; s = 2; } }
            tma_prefetch_l2_2d(&maps.m[s], ch * (GQ_CHUNK / 4), gg * GQ_ROWS);
        }
    }

    b200_trace_stamp(rec, 1);
    asm volatile("griddepcontrol.wait;" ::: "memory");                    // src1 comes from the previous kernel of the stream
    b200_trace_stamp(rec, 2);
    unsigned ll_tag = 0;
    if (p.nll) { asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(ll_tag) : "l"(p.ctrl + 8)); ll_tag += 1u; }

    // activation column -> q8_1 in shared memory, natural order (same arithmetic as q8_1_block<B200_Q8_1_X86>, bit-exact)
    if (p.dbg & 1) {
    } else if (p.aq != nullptr) {
        const int4 * src = (const int4 *) p.aq;
        for (int j = threadIdx.x; j < (p.K >> 4); j += blockDim.x) ((int4 *) s_q)[j] = src[j];
        for (int j = threadIdx.x; j < (p.K >> 5); j += blockDim.x) { float2 v = __half22float2(p.ads[j]); if (j & 1) v.x *= 0.0625f; s_ds[j] = v; }
    } else {
        b200_quantize_column_x86(p.x, p.K, [&](int blk32, int hf, int4 w, float d, float sm) {
            ((int4 *) s_q)[blk32 * 2 + hf] = w;
            if (hf == 0) s_ds[blk32] = make_float2((blk32 & 1) ? d * 0.0625f : d, sm);   // odd sub-blocks meet unshifted high nibbles (16 x)
        }, rec, fill_rest);
    }
    fill_rest();
    // the next kernel may start its prologue (PDL) only now that our activation loads have landed (see b200_gemv_stream.cu)
    if (!(p.dbg & 8)) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads();
    b200_trace_stamp(rec, 3);

    const int g = lane >> 2, t = lane & 3;
    const uint32_t q_sm = smem_u32(s_q) + lane * 8, ds_sm = smem_u32(s_ds) + t * 16;
    const int sh = 16 * (t & 1);
    int stage = 0, phase = 0;
    int gl = warp / nchunks, ch = warp - gl * nchunks;
    // B operand masks: MMA pp only keeps activation columns 2pp, 2pp+1 (this lane feeds column g), so the four MMAs of a block can
    // accumulate into ONE C fragment and lane (g, t) finds its two sums in c[0] (row g, column 2t) and c[3] (row g+8, column 2t+1)
    const unsigned bm0 = (g >> 1) == 0 ? 0xFFFFFFFFu : 0u, bm1 = (g >> 1) == 1 ? 0xFFFFFFFFu : 0u,
                   bm2 = (g >> 1) == 2 ? 0xFFFFFFFFu : 0u, bm3 = (g >> 1) == 3 ? 0xFFFFFFFFu : 0u;
    // get_scale_min_k4 (ggml-quants.c:703-710) for sub-blocks j = 2t, 2t+1, two bytes at a time, with lane-constant selectors so that the
    // decode is branch-free: j < 4 (t < 2): bytes j of scales[0..3] / [4..7]; j >= 4: low / high nibbles of scales[8..11] plus the top
    // two bits of the bytes above
    const bool     hi4 = t >= 2;
    const unsigned mA = hi4 ? 0x0F0Fu : 0x3F3Fu, mB = hi4 ? 0x3030u : 0u;
    const int      shm = hi4 ? sh + 4 : sh;
    for (int it = warp; it < nitems_cta; it += NW) {
        mbar_wait(bar0 + 8 * stage, (uint32_t) phase);
        float acc = 0.0f;
        {
            const uint32_t row = ring + stage * GQ_STAGE_BYTES + g * GQ_CHUNK;
            // (two interleaved MMA chains per item, all loads first, measured slower: 14.7 vs 14.1 us on gate/up)
#pragma unroll
            for (int b = 0; b < 2; ++b) {
                const int blk = ch * 2 + b;
                if (b && blk >= nblocks) break;                           // warp-uniform (odd block count: the box tail is zero fill)
                const uint32_t rb = row + b * 144;
                const int4 hdr = lds128(rb);                              // d | dmin, scales[12]
                const int2 bq  = lds64(q_sm + blk * 256);
                int c[4] = { 0, 0, 0, 0 };
                // A rows 0-7 take the bytes as they are (lo + 16 hi), rows 8-15 the high nibbles in place (16 hi): with the masked B
                // operand c[0] = sum (lo + 16 hi) a(2t), c[2] = sum 16 hi a(2t), c[3] = sum 16 hi a(2t+1); the 16 x is undone by the
                // activation scale of odd sub-blocks
                { const int2 w = lds64(rb + 16 + 8 * t);
                  mma_u8s8(c, (unsigned) w.x, (unsigned) w.x & 0xF0F0F0F0u, (unsigned) w.y, (unsigned) w.y & 0xF0F0F0F0u, (unsigned) bq.x & bm0, (unsigned) bq.y & bm0); }
                { const int2 w = lds64(rb + 48 + 8 * t);
                  mma_u8s8(c, (unsigned) w.x, (unsigned) w.x & 0xF0F0F0F0u, (unsigned) w.y, (unsigned) w.y & 0xF0F0F0F0u, (unsigned) bq.x & bm1, (unsigned) bq.y & bm1); }
                { const int2 w = lds64(rb + 80 + 8 * t);
                  mma_u8s8(c, (unsigned) w.x, (unsigned) w.x & 0xF0F0F0F0u, (unsigned) w.y, (unsigned) w.y & 0xF0F0F0F0u, (unsigned) bq.x & bm2, (unsigned) bq.y & bm2); }
                { const int2 w = lds64(rb + 112 + 8 * t);
                  mma_u8s8(c, (unsigned) w.x, (unsigned) w.x & 0xF0F0F0F0u, (unsigned) w.y, (unsigned) w.y & 0xF0F0F0F0u, (unsigned) bq.x & bm3, (unsigned) bq.y & bm3); }
                const int ilo = c[0] - c[2], ihi = c[3];
                const unsigned xs = hi4 ? (unsigned) hdr.w : (unsigned) hdr.y, xm = hi4 ? (unsigned) hdr.w : (unsigned) hdr.z;
                const unsigned sc2 = ((xs >> sh) & mA)  | (((unsigned) hdr.y >> (sh + 2)) & mB);
                const unsigned m2  = ((xm >> shm) & mA) | (((unsigned) hdr.z >> (sh + 2)) & mB);
                const int4 dsi = lds128(ds_sm + blk * 64);                // dA(2t), s(2t), dA(2t+1) / 16, s(2t+1)
                const float4 ds = make_float4(__int_as_float(dsi.x), __int_as_float(dsi.y), __int_as_float(dsi.z), __int_as_float(dsi.w));
                const float2 dm = make_float2(__half2float(__ushort_as_half((unsigned short) ((unsigned) hdr.x & 0xFFFFu))),
                                              __half2float(__ushort_as_half((unsigned short) ((unsigned) hdr.x >> 16))));
                const float dot = ds.x * (float) (ilo * (int) (sc2 & 0xFFu)) + ds.z * (float) (ihi * (int) (sc2 >> 8));
                const float mn  = ds.y * (float) (m2 & 0xFFu) + ds.w * (float) (m2 >> 8);
                acc += dm.x * dot - dm.y * mn;
            }
        }
        __syncwarp();                                                     // every lane is done reading this stage
        if (pit < nitems_cta) issue(stage);
        acc += __shfl_xor_sync(0xffffffffu, acc, 1);
        acc += __shfl_xor_sync(0xffffffffu, acc, 2);
        if (t == 0) s_par[(gl * nchunks + ch) * GQ_ROWS + g] = acc;
        if (++stage == GQ_STAGES) { stage = 0; phase ^= 1; }
        ch += NW;
        while (ch >= nchunks) { ch -= nchunks; ++gl; }
    }
    b200_trace_stamp(rec, 6);                                             // warp 0: all its items done
    __syncthreads();
    b200_trace_stamp(rec, 7);                                             // every warp done

    // add the chunk partials of every row in chunk order and store (or send) the row
    if (pair) {                                                           // local groups alternate gate, up: one thread finishes both values of a row
        for (int i = threadIdx.x; i < ((g_hi - g_lo) >> 1) * GQ_ROWS; i += blockDim.x) {
            const int pl = i >> 3, r = i & 7;
            float vg = 0.0f, vu = 0.0f;
            if (!(p.dbg & 4)) for (int ch = 0; ch < nchunks; ++ch) { vg += s_par[((2 * pl) * nchunks + ch) * GQ_ROWS + r]; vu += s_par[((2 * pl + 1) * nchunks + ch) * GQ_ROWS + r]; }
            const int rowi = ((g_lo >> 1) + pl) * GQ_ROWS + r;
            if (rowi < p.M[0]) { p.dst[0][rowi] = vg; p.dst[1][rowi] = vu; p.glu[rowi] = b200_swiglu1(vg, vu); }
        }
    } else
    for (int i = threadIdx.x; i < (g_hi - g_lo) * GQ_ROWS; i += blockDim.x) {
        const int gl = i >> 3, r = i & 7;
        float v = 0.0f;
        if (!(p.dbg & 4)) for (int ch = 0; ch < nchunks; ++ch) v += s_par[(gl * nchunks + ch) * GQ_ROWS + r];
        int gg = g_lo + gl, s = 0;
        if (gg >= G0) { gg -= G0; s = 1; if (gg >= G1) { gg -= G1; s = 2; } }
        const int rowi = gg * GQ_ROWS + r;
        if (rowi < p.M[s]) {
            float * d = p.dst[s] + rowi;
            if (p.nll == 0) *d = v;
            else {
                const size_t off = 2 * (size_t) ((const char *) d - p.arena);
                for (int pr = 0; pr < p.nll; ++pr)
                    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(p.ll_base[pr] + off), "r"(__float_as_uint(v)), "r"(ll_tag) : "memory");
            }
        }
    }

    if (p.trace) { __syncthreads(); b200_trace_stamp(rec, 5); }
    if (p.nll) {
        // receive side of the exchange (see gs_args in b200_gemv_stream.cu)
        const int total = p.full[0].M + p.full[1].M + p.full[2].M;
        unsigned long long t0;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
        for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < total; i += gridDim.x * blockDim.x) {
            int sgm = 0, r = i;
            if (r >= p.full[0].M) { r -= p.full[0].M; sgm = 1; if (r >= p.full[1].M) { r -= p.full[1].M; sgm = 2; } }
            float * d = p.full[sgm].dst + r;
            const char * src = p.ll_self + 2 * (size_t) ((const char *) d - p.arena);
            uint2 v;
            while (true) {
                asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(src) : "memory");
                if (v.y == ll_tag) break;
                unsigned long long t1;
                asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
                if (t1 - t0 > 5000000000ull) { p.ctrl[10] = 1u; break; }      // a dead peer must not hang the GPU
            }
            *d = __uint_as_float(v.x);
        }
        __threadfence();
        __syncthreads();
        if (threadIdx.x == 0) {
            const unsigned done = atomicAdd(p.ctrl + 9, 1u);
            if (done == gridDim.x - 1) { p.ctrl[9] = 0u; __threadfence(); *(volatile unsigned *) (p.ctrl + 8) = ll_tag; }
        }
    }
}

// ---- host side -------------------------------------------------------------------------------------------------------------
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn b200_get_encode_tiled();

static size_t gq_smem_bytes(int K, int gmax, int nw = GQ_WARPS) {
    const int nchunks = ((K >> 8) + 1) >> 1;
    return (size_t) nw * GQ_STAGES * GQ_STAGE_BYTES + (size_t) K + (size_t) (K >> 5) * 8 + (size_t) gmax * nchunks * GQ_ROWS * 4 + nw * GQ_STAGES * 8;
}
#define GQ_SMEM_MAX_BIG (200 * 1024)           // the 24-warp shape owns the SM
#define GQ_SMEM_MAX (113 * 1024)               // two CTAs (this kernel's, or this and the PDL-launched next one) per SM

// long streams (> 24 MB) take every warp slot we can register-wise fill: two 12-warp CTAs per SM, or (default) one 24-warp CTA
static bool gq_big(size_t bytes, int K, int64_t G, const b200_ctx * ctx) {
    static const int want = getenv("GGML_B200_GQ_BIG") ? atoi(getenv("GGML_B200_GQ_BIG")) : 0;         // measured slower than two 12-warp CTAs: opt-in
    if (!want || bytes <= ((size_t) 24 << 20)) return false;
    const int64_t grid = G < ctx->sm_count ? G : ctx->sm_count;
    return gq_smem_bytes(K, (int) ((G + grid - 1) / grid), 24) <= GQ_SMEM_MAX_BIG;
}
static int gq_grid(const b200_ctx * ctx, int64_t G, size_t bytes, bool big) {
    // two CTAs per SM for every size (measured: 24 warps per SM beat one CTA + the next kernel's co-resident prologue, wo 5.0 -> 4.1 us)
    static const int per_sm = getenv("GGML_B200_GQ_CTAS") ? atoi(getenv("GGML_B200_GQ_CTAS")) : 2;
    (void) bytes;
    const int64_t cap = (int64_t) ctx->sm_count * (big ? 1 : per_sm);
    return (int) (G < cap ? G : cap);
}

bool b200_gemv_mma_supported(const b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x) {
    // measured (profiles/r1_gemv_mma_microbench.txt): 5.2 vs 4.5 TB/s on a 295 MB matrix, where the instruction count is what limits
    // the dp4a kernel; no gain at <= 33 MB, where launch, quantization and ramp-up latencies dominate.  GGML_B200_GEMV_MMA=1 / 0
    // forces it on for every supported shape / off.
    static const int mode = getenv("GGML_B200_GEMV_MMA") ? atoi(getenv("GGML_B200_GEMV_MMA")) : -1;
    if (mode == 0 || n < 1 || n > GQ_MAXSEG) return false;
    int64_t G = 0; size_t bytes = 0;
    for (int i = 0; i < n; ++i) {
        if (w[i]->type != B200_TYPE_Q4_K || (w[i]->flags & B200_TENSOR_REPACKED) || w[i]->ne[0] != w[0]->ne[0]) return false;
        if (w[i]->ne[1] > (1 << 28) || (w[i]->ne[1] > 0 && (w[i]->nb[1] & 15)) || ((uintptr_t) w[i]->data & 15) || w[i]->ne[2] != 1 || w[i]->ne[3] != 1) return false;
        G += (w[i]->ne[1] + GQ_ROWS - 1) / GQ_ROWS; bytes += (size_t) w[i]->ne[1] * (w[i]->ne[0] >> 8) * 144;
    }
    const int64_t K = w[0]->ne[0];
    if (K % 256 != 0 || K < 512 || K > 32768 || G < 1) return false;
    if (mode < 0 && bytes < ((size_t) 48 << 20) && K > 8192) return false;   // (measured: the dp4a kernel is as quick on mid-size matrices with long rows)
    if (x->type != B200_TYPE_F32 || x->nb[0] != 4 || ((uintptr_t) x->data & 15) || x->ne[1] != 1 || x->ne[2] != 1 || x->ne[3] != 1) return false;
    if (b200_get_encode_tiled() == nullptr) return false;
    if (gq_big(bytes, (int) K, G, ctx)) return true;
    const int grid = gq_grid(ctx, G, bytes, false);
    return gq_smem_bytes((int) K, (int) ((G + grid - 1) / grid * 5 / 4 + 1)) <= GQ_SMEM_MAX;   // (room for a skewed share, see n_first)
}

// tensor maps of the weights are static: a small direct-mapped cache keyed by (data, row stride, K, M)
struct gq_map_entry { const void * data; size_t nb1; int64_t K, M; CUtensorMap map; };
static gq_map_entry g_map_cache[512];
static std::mutex g_map_mutex;

static int gq_weight_map(const b200_tensor * w, CUtensorMap * out) {
    const size_t h = (((uintptr_t) w->data >> 8) ^ ((uintptr_t) w->data >> 20) ^ (size_t) w->ne[1]) & 511;
    std::lock_guard<std::mutex> lock(g_map_mutex);
    gq_map_entry & e = g_map_cache[h];
    if (e.data == w->data && e.nb1 == w->nb[1] && e.K == w->ne[0] && e.M == w->ne[1]) { *out = e.map; return B200_OK; }
    const cuuint64_t dims[2] = { (cuuint64_t) (w->ne[0] >> 8) * 36, (cuuint64_t) (w->ne[1] > 0 ? w->ne[1] : 1) };   // u32 words per row, rows
    const cuuint64_t strides[1] = { (cuuint64_t) (w->ne[1] > 1 ? w->nb[1] : (size_t) (w->ne[0] >> 8) * 144) };
    const cuuint32_t box[2] = { GQ_CHUNK / 4, GQ_ROWS }, es[2] = { 1, 1 };
    const CUresult r = b200_get_encode_tiled()(out, CU_TENSOR_MAP_DATA_TYPE_UINT32, 2, (void *) w->data, dims, strides, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                                               CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) { b200_set_error("gemv_mma: cuTensorMapEncodeTiled failed (%d)", (int) r); return B200_ERR_CUDA; }
    e.data = w->data; e.nb1 = w->nb[1]; e.K = w->ne[0]; e.M = w->ne[1]; e.map = *out;
    return B200_OK;
}

// n <= 3 raw q4_K weights sharing the activation column x; dst[i] is where weight i's rows go (see b200_launch_gemv_stream)
int b200_launch_gemv_mma(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, const void * image, float * const * dst,
                         const b200_rowsplit_peers * peers, float * glu) {
    gq_args p; gq_maps maps;
    if (glu && (n != 2 || peers || w[0]->ne[1] != w[1]->ne[1])) { b200_set_error("gemv_mma: the SwiGLU epilogue takes two weights of equal shape"); return B200_ERR_UNSUPPORTED; }
    p.pair = glu ? 1 : 0; p.glu = glu;
    memset(&maps, 0, sizeof(maps));
    p.nseg = n; p.K = (int) w[0]->ne[0]; p.x = (const float *) x->data;
    p.aq = (const int8_t *) image; p.ads = image ? (const __half2 *) ((const int8_t *) image + p.K) : nullptr;
    int64_t G = 0; size_t bytes = 0;
    for (int i = 0; i < GQ_MAXSEG; ++i) { p.dst[i] = nullptr; p.M[i] = 0; p.full[i].dst = nullptr; p.full[i].M = 0; }
    for (int i = 0; i < n; ++i) {
        p.dst[i] = dst[i]; p.M[i] = (int) w[i]->ne[1];
        G += (w[i]->ne[1] + GQ_ROWS - 1) / GQ_ROWS; bytes += (size_t) w[i]->ne[1] * (w[i]->ne[0] >> 8) * 144;
        if (w[i]->ne[1] > 0) { const int rc = gq_weight_map(w[i], &maps.m[i]); if (rc) return rc; }
    }
    p.nll = 0; p.ll_self = nullptr; p.arena = nullptr; p.ctrl = nullptr;
    for (int i = 0; i < B200_MAX_PEERS; ++i) p.ll_base[i] = nullptr;
    if (peers) {
        p.nll = peers->nll; p.ll_self = (const char *) peers->ll_self; p.arena = (const char *) peers->arena; p.ctrl = peers->ctrl;
        for (int i = 0; i < peers->nll; ++i) p.ll_base[i] = (char *) peers->ll_base[i];
        for (int i = 0; i < n; ++i) { p.full[i].dst = peers->full[i]; p.full[i].M = peers->full_m[i]; }
    }
    { static int dbg = -1; if (dbg < 0) { const char * e = getenv("GGML_B200_GS_DBG"); dbg = e ? atoi(e) : 0; } p.dbg = dbg; }
    { static int npf = -1; if (npf < 0) { const char * e = getenv("GGML_B200_GQ_NPF"); npf = e ? atoi(e) : 0; } p.npf = npf; }
    p.trace = ctx->trace; p.trace_cap = ctx->trace_cap;
    { static int f = -1; if (f < 0) { const char * e = getenv("GGML_B200_GQ_NPRE"); f = e ? atoi(e) : 0; }
      p.npre = f > 0 ? (f > GQ_STAGES ? GQ_STAGES : f) : (bytes > ((size_t) 24 << 20) ? GQ_STAGES : 1); }
    const bool big = !p.pair && gq_big(bytes, p.K, G, ctx);                // (the 24-warp shape is opt-in and has no PAIR variant)
    const int grid = gq_grid(ctx, G, bytes, big);
    { static int sk = -1; if (sk < 0) { const char * e = getenv("GGML_B200_GQ_SKEW"); sk = e ? atoi(e) : 128; }   // measured: 1282 -> 1257 us per token between 0 and 150
      p.n_first = grid > ctx->sm_count ? ctx->sm_count : 0; p.skew = p.n_first ? sk : 0; }
    p.gmax = (int) (((G + grid - 1) / grid) * (1024 + p.skew) / 1024 + 1) + (p.pair ? 2 : 0);      // (pair mode rounds both ends of a share to even)
    const size_t smem = gq_smem_bytes(p.K, p.gmax, big ? 24 : GQ_WARPS);
    static bool attr_done[64] = {};
    if (!attr_done[ctx->device & 63]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemv_q4k_mma<GQ_WARPS>, cudaFuncAttributeMaxDynamicSharedMemorySize, GQ_SMEM_MAX));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemv_q4k_mma<24>, cudaFuncAttributeMaxDynamicSharedMemorySize, GQ_SMEM_MAX_BIG));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemv_q4k_mma<GQ_WARPS, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, GQ_SMEM_MAX));
        attr_done[ctx->device & 63] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) grid);
    cfg.blockDim = dim3((big ? 24 : GQ_WARPS) * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    if (p.pair) B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_gemv_q4k_mma<GQ_WARPS, true>, p, maps));
    else if (big) B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_gemv_q4k_mma<24>, p, maps));
    else     B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_gemv_q4k_mma<GQ_WARPS>, p, maps));
    return B200_OK;
}
