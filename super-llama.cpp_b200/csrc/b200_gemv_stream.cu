// b200_gemv_stream.cu -- the decode dequant-GEMV (MUL_MAT with one activation column), streamed through the TMA engine,
// with the q8_1 activation quantization fused in and up to three weight tensors that share src1 (wq/wk/wv, gate/up)
// served by ONE launch.
//
// Replaces the CPU inner loop ggml_vec_dot_q4_K_q8_K / q6_K_q8_K driven by ggml_compute_forward_mul_mat
// (reference ggml/src/ggml-cpu/ggml-cpu.c:1136-1224, quants.c:550,708) and quantize_row_q8_1
// (ggml-cpu/arch/x86/quants.c:388-449; same arithmetic as b200_quantize.cu, shared through q8_1_block()).
//
// Bound: HBM.  Algorithmic bytes per launch = sum over segments of M * K * bytes_per_block / 256 (each weight byte once).
//
// Shape of the kernel (why: profiles/r1_gemv_stream_v2_*):
//  * every warp is its own streaming engine: lane 0 issues cp.async.bulk copies (TMA engine: no registers, no LSU slots)
//    of GS_RW row-chunks into a private GS_STAGES-deep shared-memory ring, an mbarrier per stage signals arrival, the 32
//    lanes consume a stage with conflict-free 128-bit shared loads while the other stages are in flight;
//  * rows of all segments are split evenly over all warps of a persistent-size grid (two CTAs per SM);
//  * each CTA quantizes the activation column itself (bit-exact q8_1) straight into shared memory, in the order its
//    lanes will read it, while the first ring stages are landing -- no separate quantize launch, no image in HBM;
//  * launched with programmatic dependent launch: barrier init and the weight prefetch (which do not depend on the
//    previous kernel) run under the previous kernel's tail; griddepcontrol.wait guards the first read of src1.
#include "b200_common.cuh"
#include "b200_ptx.cuh"
#include "b200_gemv_traits.cuh"
#include <stdlib.h>

// Two launch shapes (measured, profiles/r1_gemv_*): WIDE = 2 CTAs x 8 warps per SM with deep rings for matrices that
// outlast the rings (best steady-state bandwidth); LEAN = 1 CTA x 12 warps per SM (<= 113 KB smem, <= 85 regs) so that this
// kernel and the PDL-launched next one co-reside on an SM and the next matmul's weights stream in while this one computes.
#define GS_MAXSEG 3

struct gs_seg { const uint8_t * w; size_t nb1; float * dst; int M; };
struct gs_args {
    gs_seg seg[GS_MAXSEG]; int nseg; int Mtot; int K;
    const float * x;                         // f32 activation column: quantized in-kernel when aq == NULL
    const int8_t * aq; const __half2 * ads;  // or a q8_1 image made by k_quantize_q8_1_soa (natural order): [K] quants, [K/32] (d, s)
    // row-split exchange: seg[i].dst points at this rank's rows inside the FULL f32 column full[i], which lives in the exchange
    // arena.  Instead of a plain store, every result row goes as an 8-byte (value, tag) pair into the low-latency mirror of that
    // float in EVERY rank's arena (NVLink peer stores; an 8-byte store is a single transaction, so the tag validates the value
    // and no fence or flag is needed -- a system-scope fence measured ~7 us here).  tag = *expected + 1.  The same kernel then
    // receives: its threads poll the local mirror of all rows of full[i] (from every rank, ours included), write the plain f32
    // column, and the last CTA advances `expected`.  All CTAs of the grid are co-resident (grid <= SMs x CTAs/SM), so the spin
    // cannot starve a producer; peers never wait on us inside the same exchange.
    int nll; char * ll_base[B200_MAX_PEERS]; const char * ll_self; const char * arena; unsigned * ctrl;
    struct { float * dst; int M; } full[GS_MAXSEG];
    // decode MUL_MAT_ID: segment s reads expert ids[id_base + s] of a stacked weight tensor (seg[s].w = expert 0); the ids live on the
    // device and come from an earlier kernel, so in this mode the first weight fetch is issued after griddepcontrol.wait
    const int * ids; int id_base; size_t estride;
    int dbg;                                 // GGML_B200_GS_DBG (measurement only): 1 skip quantization, 2 skip the dot products, 4 no weight traffic
    int npre;                                // ring stages to fill before the dependency wait
    int npf;                                 // GGML_B200_GS_NPF: bytes of this warp's rows (behind the ring fill) to pull into L2 from the prologue
    unsigned long long * trace; int trace_cap;   // b200_debug_trace (measurement only)
};

// ====================================================== the kernel ======================================================
static __device__ __forceinline__ void gs_locate(const gs_args & p, int gr, const uint8_t * & w, float * & dst, int * seg = nullptr) {
    // global row -> (segment, row): at most two compares
    int s = 0, r = gr;
    if (p.nseg > 1 && r >= p.seg[0].M) { r -= p.seg[0].M; s = 1; if (p.nseg > 2 && r >= p.seg[1].M) { r -= p.seg[1].M; s = 2; } }
    w = p.seg[s].w + (size_t) r * p.seg[s].nb1;
    dst = p.seg[s].dst + r;
    if (seg) *seg = s;
}

// SHAPE: 0 = LEAN (12 warps, 1 CTA/SM), 1 = WIDE (8 warps, deep rings, 2 CTAs/SM), 2 = DENSE (12 warps, 2 CTAs/SM: most warps per SM),
// 3 = BIG (16 warps, deep rings, ONE CTA per SM: the warps of WIDE, but the activation column is fetched and quantized once per SM --
// with two CTAs per SM the 57 KB column of a K = 14336 matmul took ~6100 cycles to arrive, L2 broadcast bandwidth, measured)
template <int TYPE, int SHAPE>
__global__ void __launch_bounds__(SHAPE == 3 ? 512 : SHAPE == 1 ? 256 : 384, SHAPE == 3 ? 1 : 2) k_gemv_stream(const gs_args p) {
    using T = gs_traits<TYPE>;
    constexpr int GS_WARPS = SHAPE == 3 ? 16 : SHAPE == 1 ? 8 : 12;
    constexpr int GS_STAGES = T::stages(SHAPE == 3 ? 1 : SHAPE);
    constexpr int CHUNK_BYTES = T::BLOCK_BYTES * GS_CB;
    constexpr int STAGE_BYTES = CHUNK_BYTES * GS_RW;
    extern __shared__ __align__(128) uint8_t smem[];
    // layout: [GS_WARPS][GS_STAGES][STAGE_BYTES] rings | activation image (T::act_bytes) | mbarriers
    const int nblocks = p.K >> 8;
    uint8_t *  s_act = smem + GS_WARPS * GS_STAGES * STAGE_BYTES;
    uint64_t * s_bar = (uint64_t *) (s_act + T::act_bytes(p.K));

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const uint32_t ring = smem_u32(smem + warp * GS_STAGES * STAGE_BYTES);
    const uint32_t bar0 = smem_u32(s_bar + warp * GS_STAGES);

    // this warp's rows: an even split of [0, Mtot) over every warp of the grid
    const int gw = blockIdx.x * GS_WARPS + warp, nw = gridDim.x * GS_WARPS;
    const int r_begin = (int) ((int64_t) p.Mtot * gw / nw), r_end = (int) ((int64_t) p.Mtot * (gw + 1) / nw);
    const int rows = r_end - r_begin;
    const int nchunks = (nblocks + GS_CB - 1) / GS_CB;
    const int npairs = (rows + GS_RW - 1) / GS_RW;
    const int nitems = (p.dbg & 4) ? 0 : npairs * nchunks;

    unsigned long long * rec = nullptr;
    if (p.trace && threadIdx.x == 0) rec = b200_trace_open(p.trace, p.trace_cap, p.Mtot, p.K);
    b200_trace_stamp(rec, 0);
    if (lane == 0) {
        for (int s = 0; s < GS_STAGES; ++s) mbar_init(bar0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncwarp();
    // (dbg & 8: the old placement of the PDL trigger, right at the start)
    if (p.dbg & 8) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

    const uint64_t policy = l2_policy_evict_first();
    int prp = 0, pkc = 0;                                                 // producer cursor (warp-uniform; lane 0 issues)
    auto issue = [&](int stage) {
        const int cb = min(GS_CB, nblocks - pkc * GS_CB);
        const int nr = min(GS_RW, rows - prp * GS_RW);
        if (lane == 0) {
            const uint32_t bytes = (uint32_t) cb * T::BLOCK_BYTES, bar = bar0 + 8 * stage;
            mbar_arrive_expect_tx(bar, bytes * nr);
            for (int r = 0; r < nr; ++r) {
                const uint8_t * w; float * dst; int sg;
                gs_locate(p, r_begin + prp * GS_RW + r, w, dst, &sg);
                if (p.ids) w += (size_t) __ldg(p.ids + p.id_base + sg) * p.estride;
                bulk_g2s_hint(ring + stage * STAGE_BYTES + r * CHUNK_BYTES, w + (size_t) pkc * CHUNK_BYTES, bytes, bar, policy);
            }
        }
        if (++pkc == nchunks) { pkc = 0; ++prp; }
    };
    const int npro = min(nitems, GS_STAGES);
    // weights do not depend on the previous kernel: fill the ring before the dependency wait -- for small matrices (p.npre = 1) only one
    // stage, the rest behind the activation loads (see b200_gemv_mma.cu)
    int nfilled = 0;
    if (!p.ids) for (; nfilled < min(npro, p.npre); ++nfilled) issue(nfilled);
    auto fill_rest = [&]() { for (; nfilled < npro; ++nfilled) issue(nfilled); };
    // The rings hold ~83 KB per CTA.  Until the previous kernel has retired and the activations are quantized nothing is consumed,
    // so nothing more is fetched and HBM idles: pull the rows behind the ring fill into L2 meanwhile (the ring refills then hit L2).
    if (p.npf > 0 && !p.ids && lane == 0 && nitems > npro) {
        const uint32_t rowb = (uint32_t) nblocks * T::BLOCK_BYTES;
        int left = p.npf;
        for (int r = prp * GS_RW; r < rows && left > 0; ++r, left -= (int) rowb) {
            const uint8_t * w; float * dst;
            gs_locate(p, r_begin + r, w, dst);
            asm volatile("cp.async.bulk.prefetch.L2.global [%0], %1;" :: "l"(w), "r"(rowb) : "memory");
        }
    }
    b200_trace_stamp(rec, 1);

    // src1 is produced by the previous kernel in the stream: wait for it (and its memory) before the first read
    asm volatile("griddepcontrol.wait;" ::: "memory");
    b200_trace_stamp(rec, 2);
    unsigned ll_tag = 0;
    if (p.nll) { asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(ll_tag) : "l"(p.ctrl + 8)); ll_tag += 1u; }
    if (p.ids || p.aq != nullptr || (p.dbg & 1)) fill_rest();             // (which experts: only known now)

    // quantize the activation column to q8_1 straight into shared memory, in lane order (b200_quantize_column_x86: same arithmetic
    // as q8_1_block<B200_Q8_1_X86>, bit-exact with the reference)
    if (p.dbg & 1) {
    } else if (p.aq != nullptr) {
        // a ready q8_1 image (one quantize launch instead of one quantization per CTA): permuting copy
        const int4 * src = (const int4 *) p.aq;
        const int nunits = p.K >> 4;
        for (int j0 = threadIdx.x; j0 < nunits; j0 += 4 * blockDim.x) {    // four loads in flight per thread: one L2 round trip for K <= 16384
            int4 v[4]; float2 ds[4];
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int j = j0 + i * blockDim.x;
                if (j < nunits) { v[i] = src[j]; if ((j & 1) == 0) ds[i] = __half22float2(p.ads[j >> 1]); }
            }
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int j = j0 + i * blockDim.x;
                if (j < nunits) T::store_unit(s_act, p.K, nblocks, j >> 1, j & 1, v[i], ds[i].x, ds[i].y);
            }
        }
    } else {
        b200_quantize_column_x86(p.x, p.K, [&](int blk32, int hf, int4 w, float d, float sm) { T::store_unit(s_act, p.K, nblocks, blk32, hf, w, d, sm); }, rec, fill_rest);
    }
    // Let the next kernel of the stream start its prologue (PDL) only now: its CTA lands on this SM and pulls ~80 KB of weights,
    // and loads issued behind that burst wait for it (the activation column took ~1300 cycles to arrive with the trigger at the top
    // of the kernel, measured with b200_debug_trace).  Our own activation loads are done; the next kernel's ring fill overlaps our dot
    // products instead.
    if (!(p.dbg & 8)) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
    __syncthreads();
    b200_trace_stamp(rec, 3);

    const int bl = lane & 7;
    const uint32_t act = smem_u32(s_act);
    float acc[GS_RW];
#pragma unroll
    for (int r = 0; r < GS_RW; ++r) acc[r] = 0.0f;

    int stage = 0, phase = 0, rp = 0, kc = 0;
    for (int item = 0; item < nitems; ++item) {
        const int cb = min(GS_CB, nblocks - kc * GS_CB);
        const int nr = min(GS_RW, rows - rp * GS_RW);
        mbar_wait(bar0 + 8 * stage, (uint32_t) phase);
        if (item == 0) b200_trace_stamp(rec, 4);
        if (bl < cb && !(p.dbg & 2)) {
            const uint32_t st = ring + stage * STAGE_BYTES;
            const int gb = kc * GS_CB + bl;
            if (nr == GS_RW) T::template consume<GS_RW, GS_RW>(st, CHUNK_BYTES, bl, gb, lane, act, p.K, nblocks, acc);
            else             T::template consume<1, GS_RW>(st, CHUNK_BYTES, bl, gb, lane, act, p.K, nblocks, acc);
        }
        __syncwarp();                                                     // every lane is done reading this stage
        if (item + GS_STAGES < nitems) issue(stage);

        if (kc == nchunks - 1) {
#pragma unroll
            for (int r = 0; r < GS_RW; ++r) {
                const float v = warp_sum(acc[r]);
                if (lane == 0 && r < nr) {
                    const uint8_t * w; float * dst;
                    gs_locate(p, r_begin + rp * GS_RW + r, w, dst);
                    if (p.nll == 0) *dst = v;
                    else {
                        const uint2 pk = make_uint2(__float_as_uint(v), ll_tag);
                        const size_t off = 2 * (size_t) ((const char *) dst - p.arena);
                        for (int pr = 0; pr < p.nll; ++pr)
                            asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(p.ll_base[pr] + off), "r"(pk.x), "r"(pk.y) : "memory");
                    }
                }
                acc[r] = 0.0f;
            }
            kc = 0; ++rp;
        } else ++kc;
        if (++stage == GS_STAGES) { stage = 0; phase ^= 1; }
        if (item == 0) b200_trace_stamp(rec, 9);                            // warp 0: first item done
    }
    b200_trace_stamp(rec, 6);                                             // warp 0: all its items done

    if (p.trace) { __syncthreads(); b200_trace_stamp(rec, 5); }
    if (p.nll) {
        // receive side of the exchange (see gs_args)
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

// the receiving side of the row-split exchange: polls the local low-latency mirror until every row carries this exchange's tag
// (rows arrive from all ranks, our own included), writes the plain f32 column, and the last CTA advances `expected`.  Gives up after ~5 s (a dead peer must not hang the GPU).
__global__ void __launch_bounds__(256) k_exchange_unpack(const uint2 * ll, float * dst, int M, unsigned * expected, unsigned * counter, unsigned * error) {
    asm volatile("griddepcontrol.launch_dependents;" ::: "memory");      // the next matmul may prefetch its weights meanwhile
    // the local GEMV (and through it the previous receive, which advanced `expected`) has retired: the PDL chain stays transitive
    asm volatile("griddepcontrol.wait;" ::: "memory");
    unsigned tag;
    asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(tag) : "l"(expected));
    tag += 1u;
    unsigned long long t0;
    asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
    for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < M; i += gridDim.x * blockDim.x) {
        uint2 v;
        while (true) {
            asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v.x), "=r"(v.y) : "l"(ll + i) : "memory");
            if (v.y == tag) break;
            unsigned long long t1;
            asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
            if (t1 - t0 > 5000000000ull) { *error = 1u; break; }
        }
        dst[i] = __uint_as_float(v.x);
    }
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) {
        const unsigned done = atomicAdd(counter, 1u);
        if (done == gridDim.x - 1) { *counter = 0u; __threadfence(); *(volatile unsigned *) expected = tag; }
    }
}
int b200_launch_exchange_unpack(b200_ctx * ctx, const void * ll, float * dst, int64_t M, unsigned * ctrl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) ((M + 255) / 256 < 32 ? (M + 255) / 256 : 32));
    cfg.blockDim = dim3(256);
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    B200_LAUNCH(ctx, B200_K_GLUE, cudaLaunchKernelEx(&cfg, k_exchange_unpack, (const uint2 *) ll, dst, (int) M, ctrl + 8, ctrl + 9, ctrl + 10));
    return B200_OK;
}

template <int TYPE> static size_t gs_smem_bytes(int K, int shape) {
    const int GS_WARPS = shape == 3 ? 16 : shape == 1 ? 8 : 12, GS_STAGES = gs_traits<TYPE>::stages(shape == 3 ? 1 : shape);
    return (size_t) GS_WARPS * GS_STAGES * gs_traits<TYPE>::BLOCK_BYTES * GS_CB * GS_RW + gs_traits<TYPE>::act_bytes(K) + GS_WARPS * GS_STAGES * 8;
}
#define GS_SMEM_MAX (200 * 1024)   // typical shapes need <= 113 KB, which lets two kernels co-reside on an SM

// true when the streaming kernel can take a MUL_MAT of this weight with ONE contiguous f32 activation column
bool b200_gemv_stream_supported(const b200_tensor * w, const b200_tensor * x) {
    if (x->ne[1] != 1 || x->ne[2] != 1 || x->ne[3] != 1 || w->ne[2] != 1 || w->ne[3] != 1) return false;
    if (x->type != B200_TYPE_F32 || x->nb[0] != 4 || ((uintptr_t) x->data & 15)) return false;
    if (((uintptr_t) w->data & 15) || (w->nb[1] & 15) || w->ne[0] % 256 != 0 || w->ne[0] > 32768 || w->ne[1] > (1 << 28)) return false;
    const int K = (int) w->ne[0];
    if (w->type == B200_TYPE_Q4_K) return !(w->flags & B200_TENSOR_REPACKED) && gs_smem_bytes<B200_TYPE_Q4_K>(K, 0) <= GS_SMEM_MAX;
    if (w->type == B200_TYPE_Q6_K) return (w->flags & B200_TENSOR_REPACKED) && K % 2048 == 0 && gs_smem_bytes<B200_TYPE_Q6_K>(K, 0) <= GS_SMEM_MAX;
    if (w->type == B200_TYPE_Q4_0) return (w->flags & B200_TENSOR_REPACKED) && gs_smem_bytes<B200_TYPE_Q4_0>(K, 0) <= GS_SMEM_MAX;
    if (w->type == B200_TYPE_Q8_0) return (w->flags & B200_TENSOR_REPACKED) && gs_smem_bytes<B200_TYPE_Q8_0>(K, 0) <= GS_SMEM_MAX;
    return false;
}

template <int TYPE, int SHAPE>
static int gs_launch_cfg(b200_ctx * ctx, const gs_args & p, int attr_slot) {
    constexpr int WARPS = SHAPE == 3 ? 16 : SHAPE == 1 ? 8 : 12;
    constexpr bool WIDE = SHAPE == 1 || SHAPE == 2;                       // two CTAs per SM
    const size_t smem = gs_smem_bytes<TYPE>(p.K, SHAPE);
    static bool attr_done[64] = {};                                       // per kernel instantiation and device
    (void) attr_slot;
    if (!attr_done[ctx->device & 63]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_gemv_stream<TYPE, SHAPE>, cudaFuncAttributeMaxDynamicSharedMemorySize, WIDE ? 113 * 1024 : GS_SMEM_MAX));
        attr_done[ctx->device & 63] = true;
    }
    // persistent-size grid, rows split evenly over all warps; fewer CTAs when there are few rows
    const int64_t want = (p.Mtot + WARPS - 1) / WARPS;
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) max((int64_t) 1, min(want, (int64_t) ctx->sm_count * (WIDE ? 2 : 1))));
    cfg.blockDim = dim3(WARPS * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_gemv_stream<TYPE, SHAPE>, p));
    return B200_OK;
}
template <int TYPE>
static int gs_launch(b200_ctx * ctx, const gs_args & p, int attr_slot) {
    // WIDE once the weights outlast what the rings of one wave can hold (~22 MB), if its smem fits two CTAs per SM
    const size_t bytes = (size_t) p.Mtot * (p.K >> 8) * gs_traits<TYPE>::BLOCK_BYTES;
    static int force = -1;
    if (force < 0) { const char * e = getenv("GGML_B200_GEMV_SHAPE"); force = e ? (e[0] == 'w' ? 1 : e[0] == 'l' ? 2 : e[0] == 'd' ? 3 : e[0] == '2' ? 4 : 0) : 0; }
    bool wide = bytes > ((size_t) 24 << 20);
    if (force == 1) wide = true; else if (force == 2) wide = false;
    // two-CTA shapes: DENSE (24 warps per SM) measured 8-10 % faster than WIDE in isolation for K <= 8192, but in the token's
    // kernel chain it leaves no room for the next kernel's PDL prologue and the whole step got 3 % slower: opt-in only
    const bool dense = force == 3;
    if ((wide || force == 3) && dense && gs_smem_bytes<TYPE>(p.K, 2) <= 113 * 1024) return gs_launch_cfg<TYPE, 2>(ctx, p, attr_slot);
    // long streams: BIG (one 16-warp CTA per SM); GGML_B200_GEMV_SHAPE=2 keeps the two 8-warp CTAs of WIDE
    if (wide && force != 4 && gs_smem_bytes<TYPE>(p.K, 3) <= GS_SMEM_MAX) return gs_launch_cfg<TYPE, 3>(ctx, p, attr_slot);
    if (wide && gs_smem_bytes<TYPE>(p.K, 1) <= 113 * 1024) return gs_launch_cfg<TYPE, 1>(ctx, p, attr_slot);
    return gs_launch_cfg<TYPE, 0>(ctx, p, attr_slot + 4);
}

// n <= 3 weights of the same type sharing the activation column x; dst[i] is the f32 column of weight i
int b200_launch_gemv_stream(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, const void * image, float * const * dst,
                            const b200_rowsplit_peers * peers, const b200_gemv_ids * idm) {
    if (!idm && b200_gemv_mma_supported(ctx, w, n, x)) return b200_launch_gemv_mma(ctx, w, n, x, image, dst, peers);   // raw q4_K: tensor-core dots
    gs_args p;
    p.ids = idm ? idm->ids : nullptr; p.id_base = idm ? idm->base : 0; p.estride = idm ? idm->stride : 0;
    p.nll = 0; p.ll_self = nullptr; p.arena = nullptr; p.ctrl = nullptr;
    for (int i = 0; i < B200_MAX_PEERS; ++i) p.ll_base[i] = nullptr;
    for (int i = 0; i < GS_MAXSEG; ++i) { p.full[i].dst = nullptr; p.full[i].M = 0; }
    if (peers) {
        p.nll = peers->nll; p.ll_self = (const char *) peers->ll_self; p.arena = (const char *) peers->arena; p.ctrl = peers->ctrl;
        for (int i = 0; i < peers->nll; ++i) p.ll_base[i] = (char *) peers->ll_base[i];
        for (int i = 0; i < n; ++i) { p.full[i].dst = peers->full[i]; p.full[i].M = peers->full_m[i]; }
    }
    { static int dbg = -1; if (dbg < 0) { const char * e = getenv("GGML_B200_GS_DBG"); dbg = e ? atoi(e) : 0; } p.dbg = dbg; }
    { static int npf = -1; if (npf < 0) { const char * e = getenv("GGML_B200_GS_NPF"); npf = e ? atoi(e) : 0; } p.npf = npf; }
    p.trace = ctx->trace; p.trace_cap = ctx->trace_cap;
    p.nseg = n; p.Mtot = 0; p.x = (const float *) x->data; p.K = (int) w[0]->ne[0];
    p.aq = (const int8_t *) image; p.ads = image ? (const __half2 *) ((const int8_t *) image + p.K) : nullptr;
    if (image && idm && idm->ads) p.ads = (const __half2 *) idm->ads;
    for (int i = 0; i < n; ++i) {
        p.seg[i].w = (const uint8_t *) w[i]->data; p.seg[i].nb1 = w[i]->nb[1]; p.seg[i].dst = dst[i]; p.seg[i].M = (int) w[i]->ne[1];
        p.Mtot += p.seg[i].M;
    }
    for (int i = n; i < GS_MAXSEG; ++i) { p.seg[i].w = nullptr; p.seg[i].nb1 = 0; p.seg[i].dst = nullptr; p.seg[i].M = 0; }
    { static int f = -1; if (f < 0) { const char * e = getenv("GGML_B200_GS_NPRE"); f = e ? atoi(e) : 0; }
      p.npre = f > 0 ? f : ((size_t) p.Mtot * b200_row_size(w[0]->type, p.K) > ((size_t) 24 << 20) ? 8 : 1); }
    if (w[0]->type == B200_TYPE_Q4_K) return gs_launch<B200_TYPE_Q4_K>(ctx, p, 8);     // attr_done slots 8..11 wide, 12..15 lean
    if (w[0]->type == B200_TYPE_Q6_K) return gs_launch<B200_TYPE_Q6_K>(ctx, p, 9);
    if (w[0]->type == B200_TYPE_Q4_0) return gs_launch<B200_TYPE_Q4_0>(ctx, p, 10);
    if (w[0]->type == B200_TYPE_Q8_0) return gs_launch<B200_TYPE_Q8_0>(ctx, p, 11);
    b200_set_error("gemv_stream: unsupported type %d", w[0]->type);
    return B200_ERR_UNSUPPORTED;
}
