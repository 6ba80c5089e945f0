// b200_chain.cu -- the persistent decode chain: several dependent MUL_MATs of ONE token, with the glue between them, as one kernel.
//
// Replaces, for one activation column, the node sequences of a llama layer that the reference runs node by node
// (src/models/llama.cpp:40-130: wo -> ADD -> RMS_NORM -> MUL -> ffn_gate / ffn_up -> GLU(SWIGLU) -> ffn_down -> ADD -> RMS_NORM -> MUL ->
// wq / wk / wv; ggml-cpu: ggml_compute_forward_mul_mat ggml-cpu.c:1226-1418, rms_norm ops.cpp:3560-3640, glu), i.e. everything between
// two attention operators.
//
// Why (profiles/r1_gemv_mma_microbench.txt): as separate launches every decode matmul pays ~1.4 us of launch hand-over,
// ~0.7 us until its activation column has arrived and 0.5-1.5 us of q8_1 quantization before the first weight byte is used, and its
// shared-memory ring only holds what ONE launch prefetched -- the matrices under ~35 MB (4 of 5 per layer) ran at 35-45 % of HBM.
// Here the weight stream never stops at an operator boundary:
//   * one CTA per SM, persistent for the whole chain.  A PRODUCER warp walks the chain's weight matrices in their fixed order and
//     pulls this CTA's row tiles (CH_R rows x 2048 weights per stage) with cp.async.bulk into ONE shared-memory ring (~190 KB per SM,
//     ~28 MB over the GPU = ~4 us of HBM time of lookahead); it depends on nothing but ring space, so it runs ahead through the
//     boundaries (and, launched with PDL, under the tail of the attention kernel in front);
//   * CH_NW CONSUMER warps take the stages round-robin: conflict-free 128-bit shared loads, dp4a on nibbles masked in place
//     (b200_gemv_traits.cuh), a 6-shuffle reduction, partial sums into a per-CTA table that is summed in a fixed order (deterministic);
//   * between two phases the activation goes through L2: results are stored, a grid-wide arrival counter (red.release / ld.acquire)
//     tells every CTA when the column is complete, and each CTA then builds its own q8_1 image in shared memory -- residual ADD,
//     RMS_NORM and the norm weight folded into that prologue (one pass, 16 values per thread), SwiGLU into the epilogue of gate / up.
// Bound: HBM.  Algorithmic bytes = sum over the chain's weights of M * K * bytes_per_block / 256.
#include "b200_common.cuh"
#include "b200_ptx.cuh"
#include "b200_gemv_traits.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#define CH_R       4       // weight rows per tile
// consumer warps: template parameter NW of the kernel (CH_NW / CH_NT inside it); the dots are latency-bound per warp (~20 clk per instruction
// at 4 warps per scheduler, measured), so the warp count is what buys consumer throughput
#define CH_MAXPH   B200_CHAIN_MAX_PHASES
#define CH_MAXSEG  3
#define CH_PSUM    2560    // floats in the partial-sum table: stages per CTA and phase <= CH_PSUM / CH_R
#define CH_Q4K_CHUNK 1152  // 8 blocks of 144 B
#define CH_Q6K_CHUNK 1680  // one repacked group of 8 blocks

struct ch_seg { const uint8_t * w; size_t nb1; float * dst; int M; int q6; int map; };      // map: index of the weight's 2-D tensor map
struct ch_phase {
    int prologue, epilogue, K, nseg;
    ch_seg seg[CH_MAXSEG];
    const float * in0, * in1, * normw;
    float * out_add, * out_norm, * out_mul, * glu;
    float eps;
    int nunits;            // units of CH_R rows (SWIGLU: a unit is CH_R rows of gate AND of up)
    int off;               // unit u belongs to CTA (u + off) % gridDim.x
    int nch;               // K / 2048
    unsigned nch_inv;      // ceil(65536 / nch): x / nch == (x * nch_inv) >> 16 for x < 4096
    int types;             // bit 0: a q4_K activation image is needed, bit 1: a q6_K one
    int in1_early;         // in1 is not written by the phase right before: it may be fetched before the hand-over
};
#define CH_MAXMAPS 8
struct ch_args {
    CUtensorMap map[CH_MAXMAPS];   // per weight matrix: rows of nb1 bytes as 8-byte elements, box = one K chunk x CH_R rows: ONE request per stage
    ch_phase ph[CH_MAXPH];
    int nph, slot_bytes, nslot, act_bytes;
    unsigned * sync;       // [0] phase arrivals, [1] CTAs done, [2] error code
    int dbg;               // GGML_B200_CHAIN_DBG (measurement only): 1 skip the dot products, 2 no weight traffic
    unsigned long long * trace; int trace_cap;   // b200_debug_trace (measurement only): per CTA and phase, globaltimer at 5 points
};

static __device__ __forceinline__ unsigned long long ch_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
// every wait in this kernel is bounded: a stuck chain must fail the launch, not hang the GPU
#define CH_TIMEOUT_NS 4000000000ull
static __device__ __noinline__ void ch_die(unsigned * sync, unsigned code) { sync[2] = code; __threadfence(); __trap(); }
static __device__ __forceinline__ void ch_mbar_wait(uint32_t bar, uint32_t parity, unsigned * sync, unsigned code) {
    unsigned n = 0;
    unsigned long long t0 = 0;                       // (%globaltimer is a slow read: only touched after ~1000 failed polls)
    while (!mbar_try_wait_hint(bar, parity, 1000000u)) {
        if ((++n & 0x3ffu) == 0) { const unsigned long long t = ch_now(); if (!t0) t0 = t; else if (t - t0 > CH_TIMEOUT_NS) ch_die(sync, code); }
    }
}
static __device__ __forceinline__ float4 ch_ldcg4(const float * p) {      // produced by other SMs during this launch: never from L1
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
template <int NT> static __device__ __forceinline__ void ch_bar_n() { asm volatile("bar.sync 1, %0;" :: "n"(NT) : "memory"); }
#define ch_bar_consumers() ch_bar_n<CH_NT>()

// unit -> (first tile's segment, first row inside it); PLAIN phases concatenate the segments' rows, SWIGLU phases pair them
static __device__ __forceinline__ void ch_locate(const ch_phase & ph, int u, int tl, int & sg, int & row) {
    row = u * CH_R;
    if (ph.epilogue == B200_CHAIN_Y_SWIGLU) { sg = tl; return; }
    sg = 0;
    if (ph.nseg > 1 && row >= ph.seg[0].M) { row -= ph.seg[0].M; sg = 1; if (ph.nseg > 2 && row >= ph.seg[1].M) { row -= ph.seg[1].M; sg = 2; } }
}

// the q8_1 image(s) of 16 values held by this thread (unit u16 = 16 consecutive activations; lanes 2i, 2i+1 form one 32-value block)
static __device__ __forceinline__ void ch_quant_store(const float (&f)[16], int lane, int u16, bool live, int K, int nblocks, int types, uint8_t * s_act) {
    int4 w; float d; unsigned usum;
    b200_q8_1_half16(f, lane, w, d, usum);
    if (!live) return;
    const float d16 = __half2float(__float2half_rn(d)), s16 = __half2float(__float2half_rn(__fmul_rn(d, (float) (int) usum)));
    uint8_t * a = s_act;
    if (types & 1) { gs_traits<B200_TYPE_Q4_K>::store_unit(a, K, nblocks, u16 >> 1, u16 & 1, w, d16, s16); a += gs_traits<B200_TYPE_Q4_K>::act_bytes(K); }
    if (types & 2)   gs_traits<B200_TYPE_Q6_K>::store_unit(a, K, nblocks, u16 >> 1, u16 & 1, w, d16, s16);
}

template <int CH_NW>
__global__ void __launch_bounds__(CH_NW * 32, 1) k_decode_chain(const __grid_constant__ ch_args P) {
    constexpr int CH_NT = CH_NW * 32;
    extern __shared__ __align__(128) uint8_t smem[];
    // layout: ring [nslot][slot_bytes] | activation image(s) [act_bytes] | psum [CH_PSUM] f32 | red [32] f32 | full[nslot] mbarriers
    uint8_t *  s_act  = smem + (size_t) P.nslot * P.slot_bytes;
    float *    s_psum = (float *) (s_act + P.act_bytes);
    float *    s_red  = s_psum + CH_PSUM;
    uint64_t * s_bar  = (uint64_t *) (s_red + 32);
    const uint32_t ring = smem_u32(smem), full0 = smem_u32(s_bar);
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, tid = threadIdx.x;
    const int G = gridDim.x, cta = blockIdx.x;

    if (tid == 0) {
        for (int s = 0; s < P.nslot; ++s) mbar_init(full0 + 8 * s, 1);
        mbar_fence_init();
    }
    __syncthreads();

    // ------------------------------------------------ the weight stream ------------------------------------------------
    // The CTA's stages form ONE sequence over all phases (phase by phase, unit by unit, tile by tile, K chunk by K chunk); stage s lives in
    // ring slot s % nslot, and slot q belongs to warp q % CH_NW for the whole launch.  A warp that has consumed stage s re-arms the slot with
    // stage s + nslot at once -- whatever phase that belongs to -- so at a phase boundary the ring already holds (or is receiving) the next
    // phase's first nslot stages.  Sixteen issuers, each overlapping its copies with the others' dot products: a dedicated producer
    // thread managed one stage per ~650 ns (an in-order chain of ~150 dependent instructions, measured), i.e. 2.3 TB/s over the GPU.
    int cum[CH_MAXPH + 1];                            // first stage of every phase in this CTA's sequence
    cum[0] = 0;
#pragma unroll
    for (int p = 0; p < CH_MAXPH; ++p) {
        int n = 0;
        if (p < P.nph) {
            const ch_phase & ph = P.ph[p];
            int first = cta - ph.off; if (first < 0) first += G;
            const int nu = first < ph.nunits ? (ph.nunits - first + G - 1) / G : 0;
            n = nu * (ph.epilogue == B200_CHAIN_Y_SWIGLU ? 2 : 1) * ph.nch;
        }
        cum[p + 1] = cum[p] + n;
    }
    const int total = (P.dbg & 2) ? 0 : cum[CH_MAXPH];
    // stage sl of phase ph -> (segment, first row, K chunk): no division (nch_inv), SWIGLU tiles alternate gate / up
    auto decode = [&](const ch_phase & ph, int sl, int & sg, int & row, int & c, int & k) {
        const int tile = (int) (((unsigned) sl * ph.nch_inv) >> 16);
        c = sl - tile * ph.nch;
        int first = cta - ph.off; if (first < 0) first += G;
        if (ph.epilogue == B200_CHAIN_Y_SWIGLU) { k = tile >> 1; sg = tile & 1; row = (first + k * G) * CH_R; }
        else { k = tile; ch_locate(ph, first + k * G, 0, sg, row); }
    };
    auto issue = [&](int s, int slot) {               // one lane: fetch stage s into its slot -- ONE tensor request of CH_R row chunks
        int p = 0, base = 0;
#pragma unroll
        for (int q = 1; q < CH_MAXPH; ++q) if (s >= cum[q]) { p = q; base = cum[q]; }
        const ch_phase & ph = P.ph[p];
        int sg, row, c, k; decode(ph, s - base, sg, row, c, k);
        const ch_seg & S = ph.seg[sg];
        const uint32_t cb = S.q6 ? CH_Q6K_CHUNK : CH_Q4K_CHUNK;
        const uint32_t bar = full0 + 8 * slot, dst = ring + slot * P.slot_bytes;
        if (P.trace && cta == 0 && s < 1024) P.trace[32 * (1 + P.trace_cap) + 1024 + s] = ch_now();   // (measurement only)
        mbar_arrive_expect_tx(bar, cb * CH_R);
        tma_load_2d(dst, &P.map[S.map], c * (int) (cb >> 3), row, bar);
    };
    // first lap: weights do not depend on the kernel in front (PDL): the ring fills under its tail
    if (lane == 0) for (int q = warp; q < P.nslot && q < total; q += CH_NW) issue(q, q);
    // this warp's cursor in the stage sequence: it owns the slots q with q % CH_NW == warp and meets them lap after lap
    int my_slot = warp, my_lap = 0;

    // ------------------------------------------------ consumers ------------------------------------------------
    unsigned long long * rec = nullptr;               // (measurement only) words 2 + 5 p + j: phase p entered, input complete, image built, stages done, published
    if (P.trace && tid == 0) rec = b200_trace_open(P.trace, P.trace_cap, P.nph, 0xC4A1);
#define CH_STAMP(j) do { if (rec) rec[2 + 5 * p + (j)] = ch_now(); } while (0)
    int s_base = 0;                                   // stages of this CTA before the current phase
    unsigned target = 0;
    for (int p = 0; p < P.nph; ++p) {
        const ch_phase & ph = P.ph[p];
        const int K = ph.K, nblocks = K >> 8, nu16 = K >> 4;
        CH_STAMP(0);
        // what does not depend on the phase before is fetched ahead of the hand-over: the norm weight, and the residual if it is older
        // (every CTA reads the same lines at the same moment: ~9 B/clk per SM of L2 broadcast bandwidth, measured -- bytes on the critical path count)
        const bool normp = ph.prologue != B200_CHAIN_X_PLAIN, addp = ph.prologue == B200_CHAIN_X_ADD_NORM_MUL, live = normp && tid < nu16;
        float4 gpre[4], bpre[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            gpre[j] = bpre[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live) gpre[j] = __ldg((const float4 *) (ph.normw + (size_t) tid * 16 + 4 * j));
        }
        if (p == 0) asm volatile("griddepcontrol.wait;" ::: "memory");
        if (live && addp && ph.in1_early) {
#pragma unroll
            for (int j = 0; j < 4; ++j) bpre[j] = ch_ldcg4(ph.in1 + (size_t) tid * 16 + 4 * j);
        }
        // (1) the phase's input is complete: the previous kernel (phase 0), or every CTA's results of the phase before
        if (p > 0) {
            if (tid == 0) {
                unsigned long long t0 = 0;
                unsigned v, n = 0;
                while (true) {
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(P.sync) : "memory");
                    if (v >= target) break;
                    if ((++n & 0x3ffu) == 0) { const unsigned long long t = ch_now(); if (!t0) t0 = t; else if (t - t0 > CH_TIMEOUT_NS) ch_die(P.sync, 0x200u + p); }
                }
            }
            ch_bar_consumers();
        }
        CH_STAMP(1);
        // (2) this CTA's q8_1 image of the activation column
        if (ph.prologue == B200_CHAIN_X_PLAIN) {
            float4 v[4], nx[4];
            int base = tid & ~31;
            if (base < nu16) {
#pragma unroll
                for (int j = 0; j < 4; ++j) v[j] = base + lane < nu16 ? ch_ldcg4(ph.in0 + (size_t) (base + lane) * 16 + 4 * j) : make_float4(0.f, 0.f, 0.f, 0.f);
            }
            while (base < nu16) {
                const int u = base + lane, nb = base + CH_NT;
                if (nb < nu16) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) nx[j] = nb + lane < nu16 ? ch_ldcg4(ph.in0 + (size_t) (nb + lane) * 16 + 4 * j) : make_float4(0.f, 0.f, 0.f, 0.f);
                }
                const float f[16] = { v[0].x, v[0].y, v[0].z, v[0].w, v[1].x, v[1].y, v[1].z, v[1].w, v[2].x, v[2].y, v[2].z, v[2].w, v[3].x, v[3].y, v[3].z, v[3].w };
                ch_quant_store(f, lane, u, u < nu16, K, nblocks, ph.types, s_act);
                if (nb < nu16) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) v[j] = nx[j];
                }
                base = nb;
            }
        } else {
            // h = in0 (+ in1); n = h / sqrt(mean(h^2) + eps); x = n * normw.  K <= 16 * CH_NT: thread = one 16-value unit
            const bool add = addp;
            float h[16], g[16];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                float4 a = make_float4(0.f, 0.f, 0.f, 0.f);
                const float4 gg = gpre[j];
                if (live) {
                    a = ch_ldcg4(ph.in0 + (size_t) tid * 16 + 4 * j);
                    if (add) { const float4 b = ph.in1_early ? bpre[j] : ch_ldcg4(ph.in1 + (size_t) tid * 16 + 4 * j); a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w; }
                }
                h[4 * j] = a.x; h[4 * j + 1] = a.y; h[4 * j + 2] = a.z; h[4 * j + 3] = a.w;
                g[4 * j] = gg.x; g[4 * j + 1] = gg.y; g[4 * j + 2] = gg.z; g[4 * j + 3] = gg.w;
            }
            float ss = 0.0f;
#pragma unroll
            for (int i = 0; i < 16; ++i) ss = fmaf(h[i], h[i], ss);
            ss = warp_sum(ss);
            if (lane == 0) s_red[warp] = ss;
            ch_bar_consumers();
            float tot = 0.0f;
#pragma unroll
            for (int i = 0; i < CH_NW; ++i) tot += s_red[i];
            const float scale = rsqrtf(tot / (float) K + ph.eps);
            float x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = h[i] * scale;
            // the three nodes' values are stored once, each by another CTA
            if (live) {
                if (add && ph.out_add && cta == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) *(float4 *) (ph.out_add + (size_t) tid * 16 + 4 * j) = make_float4(h[4 * j], h[4 * j + 1], h[4 * j + 2], h[4 * j + 3]);
                }
                if (ph.out_norm && cta == 1 % G) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) *(float4 *) (ph.out_norm + (size_t) tid * 16 + 4 * j) = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
                }
            }
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] *= g[i];
            if (live && ph.out_mul && cta == 2 % G) {
#pragma unroll
                for (int j = 0; j < 4; ++j) *(float4 *) (ph.out_mul + (size_t) tid * 16 + 4 * j) = make_float4(x[4 * j], x[4 * j + 1], x[4 * j + 2], x[4 * j + 3]);
            }
            ch_quant_store(x, lane, tid, live, K, nblocks, ph.types, s_act);
        }
        ch_bar_consumers();
        CH_STAMP(2);
        // the next kernel of the stream may start its prologue once the last phase is under way (PDL)
        if (p == P.nph - 1) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

        // (3) this warp's stages of the phase
        int first = cta - ph.off; if (first < 0) first += G;
        const int nunits_l = first < ph.nunits ? (ph.nunits - first + G - 1) / G : 0;
        const int nst = total ? cum[p + 1] - cum[p] : 0;
        const uint32_t act4 = smem_u32(s_act), act6 = act4 + ((ph.types & 1) ? (uint32_t) gs_traits<B200_TYPE_Q4_K>::act_bytes(K) : 0u);
        // A ring slot always belongs to the same warp (slot % CH_NW), whatever the lap: a warp then meets the laps of a slot in order, and its
        // parity wait can never take the slot's previous, still incomplete fill for its own (the fills of one lap complete in any order).
        while (true) {
            const int s = my_lap * P.nslot + my_slot, sl = s - s_base;
            if (sl >= nst) break;
            int sg, row, c, k; decode(ph, sl, sg, row, c, k);
            const bool q6 = ph.seg[sg].q6 != 0;
            ch_mbar_wait(full0 + 8 * my_slot, (uint32_t) my_lap & 1u, P.sync, 0x300u + p);
            if (P.trace && cta == 0 && lane == 0 && s < 1024) P.trace[32 * (1 + P.trace_cap) + s] = ch_now();   // (measurement only)
            float acc[CH_R];
#pragma unroll
            for (int r = 0; r < CH_R; ++r) acc[r] = 0.0f;
            const uint32_t st = ring + my_slot * P.slot_bytes;
            const int bl = lane & 7, gb = c * GS_CB + bl;
            if (!(P.dbg & 1)) {
                if (q6) gs_traits<B200_TYPE_Q6_K>::consume<CH_R, CH_R>(st, CH_Q6K_CHUNK, bl, gb, lane, act6, K, nblocks, acc);
                else    gs_traits<B200_TYPE_Q4_K>::consume<CH_R, CH_R>(st, CH_Q4K_CHUNK, bl, gb, lane, act4, K, nblocks, acc);
            }
            __syncwarp();                              // every lane is done reading the slot
            if (lane == 0 && s + P.nslot < total) issue(s + P.nslot, my_slot);
            // 4 sums over 32 lanes in 6 shuffles: halve the number of values a lane carries, then a plain butterfly
            const bool hi16 = lane & 16, hi8 = lane & 8;
            float a0 = hi16 ? acc[2] : acc[0], b0 = hi16 ? acc[0] : acc[2];
            float a1 = hi16 ? acc[3] : acc[1], b1 = hi16 ? acc[1] : acc[3];
            a0 += __shfl_xor_sync(0xffffffffu, b0, 16); a1 += __shfl_xor_sync(0xffffffffu, b1, 16);   // lanes < 16: rows 0, 1; lanes >= 16: rows 2, 3
            float a = hi8 ? a1 : a0, b = hi8 ? a0 : a1;
            a += __shfl_xor_sync(0xffffffffu, b, 8);                                                    // row = 2 * (lane >> 4) + ((lane >> 3) & 1)
            a += __shfl_xor_sync(0xffffffffu, a, 4); a += __shfl_xor_sync(0xffffffffu, a, 2); a += __shfl_xor_sync(0xffffffffu, a, 1);
            if ((lane & 7) == 0) s_psum[sl * CH_R + (lane >> 3)] = a;
            my_slot += CH_NW;
            if (my_slot >= P.nslot) { my_slot = warp; ++my_lap; }
        }
        s_base += nst;
        ch_bar_consumers();
        CH_STAMP(3);

        // (4) sum the K chunks of every row in chunk order, finish the node(s), publish
        if (ph.epilogue == B200_CHAIN_Y_SWIGLU) {
            for (int t = tid; t < nunits_l * CH_R; t += CH_NT) {
                const int k = t / CH_R, r = t - k * CH_R;
                float gsum = 0.0f, usum = 0.0f;
                for (int c = 0; c < ph.nch; ++c) { gsum += s_psum[((2 * k) * ph.nch + c) * CH_R + r]; usum += s_psum[((2 * k + 1) * ph.nch + c) * CH_R + r]; }
                const int row = (first + k * G) * CH_R + r;
                ph.seg[0].dst[row] = gsum; ph.seg[1].dst[row] = usum;
                ph.glu[row] = b200_swiglu1(gsum, usum);
            }
        } else {
            for (int t = tid; t < nunits_l * CH_R; t += CH_NT) {
                const int k = t / CH_R, r = t - k * CH_R;
                float sum = 0.0f;
                for (int c = 0; c < ph.nch; ++c) sum += s_psum[(k * ph.nch + c) * CH_R + r];
                int sg, row; ch_locate(ph, first + k * G, 0, sg, row);
                ph.seg[sg].dst[row + r] = sum;
            }
        }
        if (p + 1 < P.nph) {
            ch_bar_consumers();                       // every thread's stores are ordered before thread 0's release (CTA barrier), whose
            if (tid == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(P.sync) : "memory");   // release (cumulative over the barrier) makes them visible GPU-wide
            target += (unsigned) G;
        }
        CH_STAMP(4);
    }
    // the last CTA to finish re-arms the counters for the next launch (every CTA has passed every wait by then)
    if (P.nph > 1) {
        ch_bar_consumers();
        if (tid == 0) {
            const unsigned done = atomicAdd(P.sync + 1, 1u);
            if (done == (unsigned) G - 1) { P.sync[0] = 0u; P.sync[1] = 0u; __threadfence(); }
        }
    }
}

// ---- host -----------------------------------------------------------------------------------------------------------------------
static bool ch_col_ok(const b200_tensor * t, int64_t n) {
    return t && t->type == B200_TYPE_F32 && t->nb[0] == 4 && t->ne[0] == n && t->ne[1] == 1 && t->ne[2] == 1 && t->ne[3] == 1 && ((uintptr_t) t->data & 15) == 0;
}
static bool ch_weight_ok(const b200_tensor * w) {
    if (!w || w->ne[2] != 1 || w->ne[3] != 1 || ((uintptr_t) w->data & 15) || (w->nb[1] & 15) || w->ne[0] % 2048 != 0 || w->ne[0] > 32768 || w->ne[1] % CH_R != 0 || w->ne[1] <= 0) return false;
    if (w->type == B200_TYPE_Q4_K) return !(w->flags & B200_TENSOR_REPACKED) && w->nb[0] == 144;
    if (w->type == B200_TYPE_Q6_K) return (w->flags & B200_TENSOR_REPACKED) && w->nb[0] == 210;
    return false;
}

// fills the kernel arguments; B200_ERR_UNSUPPORTED (with a reason) when the chain is not one this kernel takes
typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn b200_get_encode_tiled();            // b200_gemm.cu

static int ch_nw() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_CHAIN_WARPS"); v = e ? atoi(e) : 16; if (v != 16 && v != 24 && v != 28) v = 16; } return v; }
static int ch_build(b200_ctx * ctx, const b200_chain_phase * phases, int n, ch_args & A, size_t & smem) {
    const int nw = ch_nw();
    if (!phases || n < 1 || n > CH_MAXPH) { b200_set_error("decode_chain: 1..%d phases", CH_MAXPH); return B200_ERR_UNSUPPORTED; }
    const int G = ctx->sm_count;
    if (G < 3) { b200_set_error("decode_chain: needs at least 3 SMs"); return B200_ERR_UNSUPPORTED; }
    memset(&A, 0, sizeof(A));
    A.nph = n;
    int off = 0, slot = 0, act = 0, nmaps = 0;
    encode_tiled_fn encode = b200_get_encode_tiled();
    if (!encode) { b200_set_error("decode_chain: cuTensorMapEncodeTiled not available from the driver"); return B200_ERR_UNSUPPORTED; }
    for (int p = 0; p < n; ++p) {
        const b200_chain_phase & s = phases[p];
        ch_phase & d = A.ph[p];
        if (s.n_w < 1 || s.n_w > CH_MAXSEG) { b200_set_error("decode_chain: phase %d has %d weights", p, s.n_w); return B200_ERR_UNSUPPORTED; }
        const int64_t K = s.w[0] ? s.w[0]->ne[0] : 0;
        d.prologue = s.prologue; d.epilogue = s.epilogue; d.K = (int) K; d.nseg = s.n_w; d.eps = s.eps; d.nch = (int) (K / 2048);
        d.nch_inv = d.nch > 0 ? (unsigned) ((65536 + d.nch - 1) / d.nch) : 0;
        int64_t Mtot = 0;
        for (int i = 0; i < s.n_w; ++i) {
            if (!ch_weight_ok(s.w[i]) || s.w[i]->ne[0] != K) { b200_set_error("decode_chain: phase %d weight %d is not raw q4_K / repacked q6_K with K %% 2048 == 0, M %% 4 == 0", p, i); return B200_ERR_UNSUPPORTED; }
            if (!ch_col_ok(s.dst[i], s.w[i]->ne[1])) { b200_set_error("decode_chain: phase %d result %d is not an f32 column of %lld", p, i, (long long) s.w[i]->ne[1]); return B200_ERR_UNSUPPORTED; }
            d.seg[i].w = (const uint8_t *) s.w[i]->data; d.seg[i].nb1 = s.w[i]->nb[1]; d.seg[i].dst = (float *) s.dst[i]->data; d.seg[i].M = (int) s.w[i]->ne[1];
            d.seg[i].q6 = s.w[i]->type == B200_TYPE_Q6_K;
            d.types |= d.seg[i].q6 ? 2 : 1;
            const int cb = d.seg[i].q6 ? CH_Q6K_CHUNK : CH_Q4K_CHUNK;
            slot = max(slot, (cb * CH_R + 127) & ~127);             // tensor-copy destinations are 128-byte aligned
            if (nmaps == CH_MAXMAPS) { b200_set_error("decode_chain: more than %d weight matrices", CH_MAXMAPS); return B200_ERR_UNSUPPORTED; }
            {   // the matrix as [M rows][nb1 / 8 elements of 8 bytes]; box = one K chunk (cb bytes) of CH_R rows
                const cuuint64_t gdim[2] = { (cuuint64_t) (s.w[i]->nb[1] / 8), (cuuint64_t) s.w[i]->ne[1] };
                const cuuint64_t gstr[1] = { (cuuint64_t) s.w[i]->nb[1] };
                const cuuint32_t box[2] = { (cuuint32_t) (cb / 8), CH_R }, es[2] = { 1, 1 };
                if (s.w[i]->nb[1] % 16 || (int64_t) s.w[i]->nb[1] < K / 256 * (d.seg[i].q6 ? 210 : 144) ||
                    encode(&A.map[nmaps], CU_TENSOR_MAP_DATA_TYPE_UINT64, 2, s.w[i]->data, gdim, gstr, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                           CU_TENSOR_MAP_SWIZZLE_NONE, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) != CUDA_SUCCESS) {
                    b200_set_error("decode_chain: phase %d weight %d: tensor map encode failed", p, i); return B200_ERR_UNSUPPORTED;
                }
                d.seg[i].map = nmaps++;
            }
            Mtot += s.w[i]->ne[1];
        }
        if (s.epilogue == B200_CHAIN_Y_SWIGLU) {
            if (s.n_w != 2 || s.w[0]->ne[1] != s.w[1]->ne[1] || !ch_col_ok(s.glu, s.w[0]->ne[1])) { b200_set_error("decode_chain: phase %d: SWIGLU wants gate, up of equal M and a glu column", p); return B200_ERR_UNSUPPORTED; }
            d.glu = (float *) s.glu->data; d.nunits = (int) (s.w[0]->ne[1] / CH_R);
        } else if (s.epilogue == B200_CHAIN_Y_PLAIN) d.nunits = (int) (Mtot / CH_R);
        else { b200_set_error("decode_chain: phase %d: unknown epilogue %d", p, s.epilogue); return B200_ERR_UNSUPPORTED; }
        if (!ch_col_ok(s.in0, K)) { b200_set_error("decode_chain: phase %d: in0 is not an aligned f32 column of %lld", p, (long long) K); return B200_ERR_UNSUPPORTED; }
        d.in0 = (const float *) s.in0->data;
        if (s.prologue == B200_CHAIN_X_ADD_NORM_MUL || s.prologue == B200_CHAIN_X_NORM_MUL) {
            const bool add = s.prologue == B200_CHAIN_X_ADD_NORM_MUL;
            if (K > 16 * 32 * nw || (add && !ch_col_ok(s.in1, K)) || !s.normw || s.normw->type != B200_TYPE_F32 || s.normw->ne[0] != K || s.normw->nb[0] != 4 || ((uintptr_t) s.normw->data & 15) ||
                (s.out_add && !ch_col_ok(s.out_add, K)) || (s.out_norm && !ch_col_ok(s.out_norm, K)) || (s.out_mul && !ch_col_ok(s.out_mul, K))) {
                b200_set_error("decode_chain: phase %d: norm prologue wants aligned f32 columns of K <= %d", p, 16 * 32 * nw); return B200_ERR_UNSUPPORTED;
            }
            d.in1 = add ? (const float *) s.in1->data : nullptr; d.normw = (const float *) s.normw->data;
            d.in1_early = 1;
            if (add && p > 0) {
                const b200_chain_phase & q = phases[p - 1];
                const b200_tensor * outs[7] = { q.dst[0], q.n_w > 1 ? q.dst[1] : nullptr, q.n_w > 2 ? q.dst[2] : nullptr, q.out_add, q.out_norm, q.out_mul, q.glu };
                const char * a0 = (const char *) s.in1->data, * a1 = a0 + K * 4;
                for (const b200_tensor * o : outs) if (o) { const char * b0 = (const char *) o->data, * b1 = b0 + b200_tensor_span(o); if (a0 < b1 && b0 < a1) d.in1_early = 0; }
            }
            d.out_add = add && s.out_add ? (float *) s.out_add->data : nullptr;
            d.out_norm = s.out_norm ? (float *) s.out_norm->data : nullptr; d.out_mul = s.out_mul ? (float *) s.out_mul->data : nullptr;
        } else if (s.prologue != B200_CHAIN_X_PLAIN) { b200_set_error("decode_chain: phase %d: unknown prologue %d", p, s.prologue); return B200_ERR_UNSUPPORTED; }
        d.off = off;
        off = (off + d.nunits) % G;
        const int tpu = s.epilogue == B200_CHAIN_Y_SWIGLU ? 2 : 1;
        const int stages_max = ((d.nunits + G - 1) / G) * tpu * d.nch;
        if (stages_max * CH_R > CH_PSUM || stages_max >= 4096) { b200_set_error("decode_chain: phase %d: %d stages per CTA exceed the partial-sum table", p, stages_max); return B200_ERR_UNSUPPORTED; }
        int a = 0;
        if (d.types & 1) a += (int) gs_traits<B200_TYPE_Q4_K>::act_bytes((int) K);
        if (d.types & 2) a += (int) gs_traits<B200_TYPE_Q6_K>::act_bytes((int) K);
        act = max(act, a);
    }
    act = (act + 127) & ~127;
    const size_t fixed = (size_t) act + CH_PSUM * 4 + 32 * 4, smem_max = 227 * 1024;
    int nslot = (int) ((smem_max - fixed) / ((size_t) slot + 8));
    if (nslot > 255) nslot = 255;
    { static int cap = -1; if (cap < 0) { const char * e = getenv("GGML_B200_CHAIN_SLOTS"); cap = e ? atoi(e) : 0; } if (cap > 0 && nslot > cap) nslot = cap; }
    if (nslot < 16) { b200_set_error("decode_chain: shared memory holds only %d stages", nslot); return B200_ERR_UNSUPPORTED; }
    A.slot_bytes = slot; A.nslot = nslot; A.act_bytes = act;
    smem = (size_t) nslot * slot + fixed + (size_t) nslot * 8;
    return B200_OK;
}

extern "C" int b200_decode_chain_supported(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases) {
    if (!ctx) return 0;
    ch_args A; size_t smem;
    return ch_build(ctx, phases, n_phases, A, smem) == B200_OK;
}

extern "C" int b200_decode_chain(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases) {
    if (!ctx) { b200_set_error("decode_chain: null context"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    ch_args A;
    size_t smem;
    int rc = ch_build(ctx, phases, n_phases, A, smem);
    if (rc != B200_OK) return rc;
    // the arrival counters: scratch slot 7, zeroed when it is (re)allocated, re-armed by the kernel itself
    const bool fresh = ctx->scratch[7] == nullptr;
    unsigned * sync = (unsigned *) b200_scratch(ctx, 7, 256);
    if (!sync) return B200_ERR_ALLOC;
    if (fresh) B200_CUDA_TRY(cudaMemsetAsync(sync, 0, 256, ctx->stream));
    A.sync = sync;
    A.trace = ctx->trace; A.trace_cap = ctx->trace_cap;
    { static int dbg = -1; if (dbg < 0) { const char * e = getenv("GGML_B200_CHAIN_DBG"); dbg = e ? atoi(e) : 0; } A.dbg = dbg; }
    const int nw = ch_nw();
    static bool attr_done[64] = {};
    if (!attr_done[ctx->device & 63]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<16>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<24>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<28>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_done[ctx->device & 63] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) ctx->sm_count);
    cfg.blockDim = dim3(nw * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    if (nw == 28)      B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<28>, A));
    else if (nw == 24) B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<24>, A));
    else               B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<16>, A));
    for (int p = 0; p < n_phases; ++p) {
        const b200_chain_phase & s = phases[p];
        for (int i = 0; i < s.n_w; ++i) b200_note_write(ctx, s.dst[i]->data, b200_tensor_span(s.dst[i]));
        b200_tensor * outs[4] = { s.out_add, s.out_norm, s.out_mul, s.glu };
        for (b200_tensor * t : outs) if (t) b200_note_write(ctx, t->data, b200_tensor_span(t));
    }
    return B200_OK;
}
