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
//   * one CTA per SM, persistent for the whole chain.  This CTA's row tiles of ALL the chain's matrices (CH_R rows x 2048 weights per
//     stage) form one sequence that is pulled through ONE shared-memory ring (up to ~200 KB per SM) by the TMA engine, one 2-D tensor
//     request per stage: a slot is re-armed with the stage nslot ahead the moment it has been consumed, whatever phase that stage
//     belongs to, so the stream runs ahead through the boundaries (and, launched with PDL, under the tail of the kernel in front).
//     Measured ceilings (profiles/r2_membw_ceilings.txt): 7.2 TB/s read-only with <= 64 requests of >= 4.6 KB in flight per SM; with 128
//     requests of 1152 B in flight the same engine delivers 2.8 TB/s -- few, large requests;
//   * every warp owns the ring slots q % CH_NW == warp: conflict-free 128-bit shared loads, dp4a on nibbles masked in place
//     (b200_gemv_traits.cuh), a 6-shuffle reduction, partial sums into a per-CTA table that is summed in a fixed order (deterministic);
//   * between two phases the activation goes through L2: results are stored, a grid-wide arrival counter (red.release / ld.acquire)
//     tells every CTA when the column is complete, and each CTA then builds its own q8_1 image in shared memory -- residual ADD,
//     RMS_NORM and the norm weight folded into that prologue (one pass, 16 values per thread), SwiGLU into the epilogue of gate / up.
//     A second hand-over form is kept behind GGML_B200_CHAIN_HANDOVER=ll: the validity flag INSIDE the data (8-byte (value, tag)
//     stores are single transactions, the reader polls the data itself, the residual stays in registers, gate / up's epilogue publishes
//     q8_1 blocks so that ffn_down fetches 1.25 bytes per activation).  It measured no faster (see ch_force_barrier) and is not the default.
//   * owner-side hand-over (default where the chain has the llama shape; CH_IN_BQ8 / CH_IN_BQ8N): a boundary cost 4.5-6 us with every CTA
//     fetching the whole f32 column (57 KB per CTA in front of ffn_down: 8.5 MB of L2 traffic), normalising and quantizing it 148 times over
//     (profiles/r2_run2_chain_bench_trace.txt: publish 1.5-2.8, wait 0.9, build 2.0-2.5 us).  Now the CTA that OWNS 32 consecutive result rows
//     -- a whole q8_1 block of the next phase's activation -- finishes them itself: residual ADD (the residual of its rows lives in a
//     register for the whole launch), multiplication with the norm weight, sum of squares of its rows, q8_1 quantization, and publishes
//     32 quants + (d, sum) + the partial sum of squares.  RMS_NORM's scale is one scalar of the whole column: q8_1(scale * y) has the quants
//     of q8_1(y) and the block scale scale * d, so the readers add up K / 32 partial sums, fetch 1.25 bytes per activation and only rescale d.
// Bound: HBM.  Algorithmic bytes = sum over the chain's weights of M * K * bytes_per_block / 256.
#include "b200_common.cuh"
#include "b200_ptx.cuh"
#include "b200_gemv_traits.cuh"
#include <cuda.h>
#include <stdlib.h>
#include <string.h>

#define CH_R       4       // weight rows per tile
#define CH_MAXPH   B200_CHAIN_MAX_PHASES
#define CH_MAXSEG  3
#define CH_MAXMAPS 8
#define CH_MAXST   1024    // stages per CTA and launch (the stage table)
#define CH_PSUM    2560    // floats in the partial-sum table: stages per CTA and phase <= CH_PSUM / CH_R
#define CH_Q4K_CHUNK 1152  // 8 blocks of 144 B
#define CH_Q6K_CHUNK 1680  // one repacked group of 8 blocks
#define CH_LL_BYTES  (256 * 1024)   // hand-over area per phase: 8 bytes per f32 (K <= 32768), or 80 bytes per q8_1 block

// how a phase gets its activation column (in0)
enum { CH_IN_EXTERNAL = 0,   // written before this launch: plain loads
       CH_IN_BARRIER  = 1,   // written by the phase before, plain f32: grid-wide arrival counter, then plain loads (general fallback)
       CH_IN_LL_F32   = 2,   // (value, tag) pairs of the phase before
       CH_IN_LL_Q8    = 3,   // q8_1 blocks of the SWIGLU phase before (10 (word, tag) pairs per block) + (value, tag) pairs of its tail rows
       CH_IN_BQ8      = 4,   // arrival counter, then plain loads of the q8_1 blocks the SWIGLU phase before quantized itself (+ its f32 tail rows)
       CH_IN_BQ8N     = 5,   // arrival counter, then q8_1 blocks of (result + residual) * norm weight and per-block sums of squares from the row owners
       CH_IN_LQ8N     = 6 }; // row split only: the same records from the row owners of EVERY rank as 11 (word, tag) pairs per block (8 quant words, d, sum, sum of squares), polled
#define CH_LQ_REC 96         // bytes per CH_IN_LQ8N record (12 pairs, the last one unused)
// (The two owner-side forms were also tried with the validity flag in the data -- 12 (word, tag) pairs per block, no fence in the producer, no
// counter, the readers poll the records: 47.0 against 44.0 us per 8B layer in the same build.  148 x 512 polling threads compete with the
// producers' stores, and the poll loops cost the kernel 14 registers.  Removed.)
// hand-over area of a phase in the owner-side forms: 32 quants per block | (d f32, sum i32) per block | sum of squares per block | f32 tail rows
#define CH_HQ    0
#define CH_HDS   (64 * 1024)
#define CH_HSS   (80 * 1024)
#define CH_HTAIL (96 * 1024)

struct ch_seg { const uint8_t * w; size_t nb1; float * dst; int M; int q6; int map; int row0; };      // map: index of the weight's 2-D tensor map; row0: row-split launches: the row of the FULL result this rank's row 0 is (dst already points there)
struct ch_phase {
    int prologue, epilogue, K, nseg;
    ch_seg seg[CH_MAXSEG];
    const float * in0, * in1, * normw;
    float * out_add, * out_norm, * out_mul, * glu;
    float eps;
    // units of CH_R rows (SWIGLU: CH_R rows of gate AND of up).  Units [0, ng8) go to the CTAs in groups of 8 consecutive units (32 rows
    // = whole q8_1 blocks: only there a CTA can quantize its own results), ngrp = ng8 / 8 groups in all; units [ng8, nunits) one by one, round-robin
    int nunits, ng8, ngrp;
    int off, toff;         // group g belongs to CTA (g + off) % G, tail unit u to CTA (u - ng8 + toff) % G
    // CH_IN_BQ8N as out_mode: what the NEXT phase's ADD + RMS_NORM + MUL need from the row owner
    const float * nx_normw, * nx_in1; float * nx_out_add;
    int nch;               // K / 2048
    unsigned nch_inv;      // ceil(65536 / nch): x / nch == (x * nch_inv) >> 16 for x < 4096
    int types;             // bit 0: a q4_K activation image is needed, bit 1: a q6_K one
    int in_mode, out_mode; // CH_IN_* of this phase's in0 / of the next phase's (what the epilogue has to publish)
    int in1_kept;          // in1 is the out_add column of an earlier phase of this launch: still in registers
    int nq8;               // CH_IN_LL_Q8: the first nq8 activations arrive as q8_1 blocks, the rest as f32 pairs
};
// the tail of the last phase (b200_chain_rope): ROPE of the rows of result i and / or their f16 copy into a KV-cache row
struct ch_tail { int kind, nodst, head_dim, n_dims; float fs, ts, af; const int32_t * pos; const float * ff; float * out; const long long * cidx; char * cache; size_t cnb1; };
struct ch_args {
    CUtensorMap map[CH_MAXMAPS];   // per weight matrix: rows of nb1 bytes as 8-byte elements, box = one K chunk x CH_R rows: ONE request per stage
    ch_phase ph[CH_MAXPH];
    int nph, slot_bytes, nslot, act_bytes;
    int pre_in1;           // the phase whose addend (in1) comes from outside the launch and is fetched at the very start, or -1
    int tail_phase;        // the phase whose results get tail[] applied by their row owners (the last one), or -1
    ch_tail tail[CH_MAXSEG];
    int own_in1;           // CH_IN_BQ8N: the producing phase whose row owners fetch their rows of an addend from outside at the very start, or -1
    unsigned * sync;       // [0] phase arrivals (CH_IN_BARRIER), [1] CTAs done, [2] error code, [3] hand-over epoch, [4] abort
    uint8_t * ll;          // hand-over areas, CH_LL_BYTES per phase
    // Row split over several GPUs (b200_decode_chain_split; one process per GPU, every rank launches the same chain on its own rows of every
    // weight): a result row goes as a (value, tag) pair into the hand-over area of EVERY rank (NVLink peer stores: an aligned 8-byte store is
    // one transaction, the tag validates the value, no fence and no counter -- a system-scope fence measured ~7 us), and every rank's CTAs poll
    // their own copy exactly as they do on one GPU: the all-gather IS the hand-over.  llr[r] = rank r's areas as mapped here (llr[rank] == ll).
    // The last phase's results are needed in full by whatever follows the launch: published the same way into the area behind the phases'
    // (fin_off), and unpacked into the plain f32 columns fin[i].dst by all CTAs before the kernel ends.
    unsigned * epoch;      // row split: the launch counter the tags come from lives in the last 64 bytes of this rank's OWN hand-over area, so it counts
                           // the launches through THESE areas only -- the ranks' tags stay equal whatever else a rank launches on its context
    int nrank; uint8_t * llr[B200_MAX_PEERS]; int poll_ns;      // poll_ns: GGML_B200_CHAIN_POLL_NS, pause between two failed polls of a record (row split)
    int fin_total; unsigned fin_off; struct { float * dst; int M; int base; } fin[CH_MAXSEG];
    int dbg;               // GGML_B200_CHAIN_DBG (measurement only): 1 skip the dot products, 2 no weight traffic
    unsigned long long * trace; int trace_cap;   // b200_debug_trace (measurement only): per CTA and phase, globaltimer at 5 points
};

static __device__ __forceinline__ unsigned long long ch_now() { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); return t; }
// every wait in this kernel is bounded: a stuck chain must fail the launch, not hang the GPU
#define CH_TIMEOUT_NS 2000000000ull
// A wait that runs out records where (sync[2] = code, first one wins) and raises the launch-wide abort word sync[4]; every other wait sees
// it within its next 1024 polls and gives up too, so the kernel ENDS (with garbage results) and b200_decode_chain_error() reports it.
struct ch_watch {                                     // (%globaltimer is a slow read: only touched every 1024 failed polls)
    unsigned n = 0; unsigned long long t0 = 0;
    __device__ __forceinline__ bool spin(unsigned * sync, unsigned code) {         // true: give up
        if ((++n & 0x3ffu) != 0) return false;
        unsigned ab;
        asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(ab) : "l"(sync + 4));
        if (ab) return true;
        const unsigned long long t = ch_now();
        if (!t0) { t0 = t; return false; }
        if (t - t0 <= CH_TIMEOUT_NS) return false;
        atomicCAS(sync + 2, 0u, code); atomicExch(sync + 4, 1u);
        return true;
    }
};
static __device__ __forceinline__ void ch_mbar_wait(uint32_t bar, uint32_t parity, unsigned * sync, unsigned code) {
    ch_watch w;
    while (!mbar_try_wait_hint(bar, parity, 1000000u)) if (w.spin(sync, code)) break;
}
static __device__ __forceinline__ float4 ch_ldcg4(const float * p) {      // produced by other SMs during this launch: never from L1
    float4 r;
    asm volatile("ld.global.cg.v4.f32 {%0, %1, %2, %3}, [%4];" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(p));
    return r;
}
// hand-over pairs: an aligned 8-byte store is one transaction, so the tag validates the value next to it
static __device__ __forceinline__ void ch_ll_store(void * p, unsigned v, unsigned tag) {
    asm volatile("st.volatile.global.v2.u32 [%0], {%1, %2};" :: "l"(p), "r"(v), "r"(tag) : "memory");
}
static __device__ __forceinline__ uint4 ch_ll_load2(const void * p) {      // two pairs: (x = value, y = tag), (z = value, w = tag)
    uint4 r;
    asm volatile("ld.volatile.global.v4.u32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p) : "memory");
    return r;
}
// the same pair into every rank's copy of a hand-over area (one GPU: one plain store)
template <bool SP> static __device__ __forceinline__ void ch_ll_pub(const ch_args & P, size_t off, unsigned v, unsigned tag) {
    if (!SP || P.nrank <= 1) { ch_ll_store(P.ll + off, v, tag); return; }
    for (int r = 0; r < P.nrank; ++r) ch_ll_store(P.llr[r] + off, v, tag);
}
template <int NT> static __device__ __forceinline__ void ch_bar_n() { asm volatile("bar.sync 1, %0;" :: "n"(NT) : "memory"); }
#define ch_bar_consumers() ch_bar_n<CH_NT>()

// what a CTA owns of a phase: fg / ft = its first group / first tail unit, gl = its whole groups, nu = its units
struct ch_own { int fg, ft, gl, nu; };
static __device__ __forceinline__ ch_own ch_owned(const ch_phase & ph, int cta, int G) {
    ch_own o;
    o.fg = cta - ph.off;  if (o.fg < 0) o.fg += G;
    o.ft = cta - ph.toff; if (o.ft < 0) o.ft += G;
    o.gl = o.fg < ph.ngrp ? (ph.ngrp - o.fg + G - 1) / G : 0;
    const int nt = ph.nunits - ph.ng8;
    o.nu = o.gl * 8 + (o.ft < nt ? (nt - o.ft + G - 1) / G : 0);
    return o;
}
// local unit k of a CTA -> global unit (see ch_phase)
static __device__ __forceinline__ int ch_unit(const ch_phase & ph, int k, const ch_own & o, int G) {
    const int kg = o.gl * 8;
    return k < kg ? ((k >> 3) * G + o.fg) * 8 + (k & 7) : ph.ng8 + o.ft + (k - kg) * G;
}
static __device__ __forceinline__ int4 ch_ldcg_i4(const void * p) {
    int4 r;
    asm volatile("ld.global.cg.v4.s32 {%0, %1, %2, %3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
static __device__ __forceinline__ int2 ch_ldcg_i2(const void * p) {
    int2 r;
    asm volatile("ld.global.cg.v2.s32 {%0, %1}, [%2];" : "=r"(r.x), "=r"(r.y) : "l"(p));
    return r;
}
static __device__ __forceinline__ float ch_ldcg_f(const void * p) {
    float r;
    asm volatile("ld.global.cg.f32 %0, [%1];" : "=f"(r) : "l"(p));
    return r;
}
// unit -> (segment, first row inside it); PLAIN phases concatenate the segments' rows, SWIGLU phases pair them (sg = tile parity)
static __device__ __forceinline__ void ch_locate(const ch_phase & ph, int u, int & sg, int & row) {
    row = u * CH_R; sg = 0;
    if (ph.nseg > 1 && row >= ph.seg[0].M) { row -= ph.seg[0].M; sg = 1; if (ph.nseg > 2 && row >= ph.seg[1].M) { row -= ph.seg[1].M; sg = 2; } }
}

// 16 quants + (d, s) of one half block into this CTA's activation image(s)
static __device__ __forceinline__ void ch_store_unit(int u16, int4 w, float d16, float s16, int K, int nblocks, int types, uint8_t * s_act) {
    uint8_t * a = s_act;
    if (types & 1) { gs_traits<B200_TYPE_Q4_K>::store_unit(a, K, nblocks, u16 >> 1, u16 & 1, w, d16, s16); a += gs_traits<B200_TYPE_Q4_K>::act_bytes(K); }
    if (types & 2)   gs_traits<B200_TYPE_Q6_K>::store_unit(a, K, nblocks, u16 >> 1, u16 & 1, w, d16, s16);
}
// the q8_1 image(s) of 16 values held by this thread (unit u16 = 16 consecutive activations; lanes 2i, 2i+1 form one 32-value block)
static __device__ __forceinline__ void ch_quant_store(const float (&f)[16], int lane, int u16, bool live, int K, int nblocks, int types, uint8_t * s_act) {
    int4 w; float d; unsigned usum;
    b200_q8_1_half16(f, lane, w, d, usum);
    if (live) ch_store_unit(u16, w, __half2float(__float2half_rn(d)), __half2float(__float2half_rn(__fmul_rn(d, (float) (int) usum))), K, nblocks, types, s_act);
}


// One warp, one q4_K stage (CH_R = 4 rows x 8 blocks of 256 weights, rows CH_Q4K_CHUNK bytes apart at st): lane = (block bl = lane & 7,
// pair bit = (lane >> 3) & 1, row half = lane >> 4).  In pass `it` the lane takes the sub-block pair pp = 2 it + pair bit -- 64 weights, the
// 32 bytes qs[32 pp ..] whose low nibbles are sub-block 2 pp and whose high nibbles are sub-block 2 pp + 1 (ggml-quants.c:1352-1374) -- of its
// block in the TWO rows 2 (lane >> 4) + {0, 1}: all lanes sit in the same branch of get_scale_min_k4 (ggml-quants.c:703-710), one lane owns
// whole sub-blocks (both 6-bit scale / min pairs are decoded once, not once per 16-byte half as gs_traits does), and the two row sums of a
// lane drop the first step of the cross-lane reduction.  acc[0], acc[1]: rows 2 (lane >> 4), 2 (lane >> 4) + 1.
// Activation image: gs_traits<B200_TYPE_Q4_K>::store_unit (planes of 16-byte units [c][gb], then (dA(2p), s(2p), dA(2p+1) / 16, s(2p+1))[p][gb]).
static __device__ __forceinline__ void ch_consume_q4k(uint32_t st, int lane, uint32_t act, int K, int nblocks, int c, float (&acc)[2]) {
    const int bl = lane & 7, gb = c * GS_CB + bl, pb = (lane >> 3) & 1;
    const uint32_t rows = st + (uint32_t) (lane >> 4) * (2 * CH_Q4K_CHUNK) + bl * 144;
#pragma unroll
    for (int it = 0; it < 2; ++it) {
        const int pp = 2 * it + pb, sh = 16 * pb;
        const int4 a0l = lds128(act + ((2 * pp) * nblocks + gb) * 16), a0h = lds128(act + ((2 * pp + 1) * nblocks + gb) * 16);
        const int4 a1l = lds128(act + (K >> 1) + ((2 * pp) * nblocks + gb) * 16), a1h = lds128(act + (K >> 1) + ((2 * pp + 1) * nblocks + gb) * 16);
        const int4 dsr = lds128(act + K + (pp * nblocks + gb) * 16);
        const float dA0 = __int_as_float(dsr.x), sA0 = __int_as_float(dsr.y), dA1 = __int_as_float(dsr.z), sA1 = __int_as_float(dsr.w);
#pragma unroll
        for (int r = 0; r < 2; ++r) {
            const uint32_t base = rows + r * CH_Q4K_CHUNK;
            const int4 hdr = lds128(base);
            const int4 qa  = lds128(base + 16 + pp * 32), qb = lds128(base + 32 + pp * 32);
            const unsigned s0 = (unsigned) hdr.y, s1 = (unsigned) hdr.z, s2 = (unsigned) hdr.w;
            int sc0, sc1, m0, m1;
            if (it == 0) {
                sc0 = (s0 >> sh) & 63;  sc1 = (s0 >> (sh + 8)) & 63;
                m0  = (s1 >> sh) & 63;  m1  = (s1 >> (sh + 8)) & 63;
            } else {
                sc0 = ((s2 >> sh) & 0xF)        | (((s0 >> (sh + 6)) & 3) << 4);
                sc1 = ((s2 >> (sh + 8)) & 0xF)  | (((s0 >> (sh + 14)) & 3) << 4);
                m0  = ((s2 >> (sh + 4)) & 0xF)  | (((s1 >> (sh + 6)) & 3) << 4);
                m1  = ((s2 >> (sh + 12)) & 0xF) | (((s1 >> (sh + 14)) & 3) << 4);
            }
            int i0 = __dp4a(qa.x & 0x0F0F0F0F, a0l.x, 0), j0 = __dp4a(qb.x & 0x0F0F0F0F, a0h.x, 0);
            i0 = __dp4a(qa.y & 0x0F0F0F0F, a0l.y, i0); j0 = __dp4a(qb.y & 0x0F0F0F0F, a0h.y, j0);
            i0 = __dp4a(qa.z & 0x0F0F0F0F, a0l.z, i0); j0 = __dp4a(qb.z & 0x0F0F0F0F, a0h.z, j0);
            i0 = __dp4a(qa.w & 0x0F0F0F0F, a0l.w, i0); j0 = __dp4a(qb.w & 0x0F0F0F0F, a0h.w, j0);
            int i1 = dp4a_us((unsigned) qa.x & 0xF0F0F0F0u, a1l.x, 0), j1 = dp4a_us((unsigned) qb.x & 0xF0F0F0F0u, a1h.x, 0);
            i1 = dp4a_us((unsigned) qa.y & 0xF0F0F0F0u, a1l.y, i1); j1 = dp4a_us((unsigned) qb.y & 0xF0F0F0F0u, a1h.y, j1);
            i1 = dp4a_us((unsigned) qa.z & 0xF0F0F0F0u, a1l.z, i1); j1 = dp4a_us((unsigned) qb.z & 0xF0F0F0F0u, a1h.z, j1);
            i1 = dp4a_us((unsigned) qa.w & 0xF0F0F0F0u, a1l.w, i1); j1 = dp4a_us((unsigned) qb.w & 0xF0F0F0F0u, a1h.w, j1);
            const float2 dm = __half22float2(*(const __half2 *) &hdr.x);
            const float t = (float) (sc0 * (i0 + j0)) * dA0 + (float) (sc1 * (i1 + j1)) * dA1;      // (dA1 carries the 1/16 of the unshifted high nibbles)
            const float u = (float) m0 * sA0 + (float) m1 * sA1;
            acc[r] = fmaf(dm.x, t, acc[r]);
            acc[r] = fmaf(-dm.y, u, acc[r]);
        }
    }
}

template <int CH_NW, bool TR, bool SP = false>      // TR: the measurement hooks (b200_debug_trace, GGML_B200_CHAIN_DBG) are compiled in; SP: row split over several GPUs
__global__ void __launch_bounds__(CH_NW * 32, 1) k_decode_chain(const __grid_constant__ ch_args P) {
    constexpr int CH_NT = CH_NW * 32;
    extern __shared__ __align__(128) uint8_t smem[];
    // layout: ring [nslot][slot_bytes] | activation image(s) [act_bytes] | psum [CH_PSUM] f32 | red [32] f32 | full[nslot] mbarriers | stage table [CH_MAXST] u32
    uint8_t *  s_act  = smem + (size_t) P.nslot * P.slot_bytes;
    float *    s_psum = (float *) (s_act + P.act_bytes);
    float *    s_red  = s_psum + CH_PSUM;
    uint64_t * s_bar  = (uint64_t *) (s_red + 32);
    uint32_t * s_tab  = (uint32_t *) (s_bar + P.nslot);
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
    // phase's first nslot stages.  Sixteen issuers, each overlapping its request with the others' dot products: a dedicated producer
    // thread managed one stage per ~650 ns (an in-order chain of ~150 dependent instructions, measured), i.e. 2.3 TB/s over the GPU.
    int cum[CH_MAXPH + 1];                            // first stage of every phase in this CTA's sequence
    cum[0] = 0;
#pragma unroll
    for (int p = 0; p < CH_MAXPH; ++p) {
        int n = 0;
        if (p < P.nph) {
            const ch_phase & ph = P.ph[p];
            n = ch_owned(ph, cta, G).nu * (ph.epilogue == B200_CHAIN_Y_SWIGLU ? 2 : 1) * ph.nch;
        }
        cum[p + 1] = cum[p] + n;
    }
    const int total = (TR && (P.dbg & 2)) ? 0 : cum[CH_MAXPH];
    // stage sl of phase ph -> (segment, first row, K chunk, local unit): no division (nch_inv), SWIGLU tiles alternate gate / up
    auto decode = [&](const ch_phase & ph, int sl, int & sg, int & row, int & c, int & k) {
        const int tile = (int) (((unsigned) sl * ph.nch_inv) >> 16);
        c = sl - tile * ph.nch;
        const ch_own o = ch_owned(ph, cta, G);
        if (ph.epilogue == B200_CHAIN_Y_SWIGLU) { k = tile >> 1; sg = tile & 1; row = ch_unit(ph, k, o, G) * CH_R; }
        else { k = tile; ch_locate(ph, ch_unit(ph, k, o, G), sg, row); }
    };
    // The stage table: what stage s of this CTA fetches and for whom, worked out ONCE by as many threads as there are stages (the walk from
    // a stage number to its tensor coordinates is ~120 dependent instructions; done at every refill by one lane it was a fifth of the loop).
    // Entry: first row (18 bits) | tensor map (3) << 18 | q6_K << 21 | K chunk (5) << 22 | segment (2) << 27.
    for (int s = tid; s < total; s += CH_NT) {
        int p = 0, base = 0;
#pragma unroll
        for (int q = 1; q < CH_MAXPH; ++q) if (s >= cum[q]) { p = q; base = cum[q]; }
        const ch_phase & ph = P.ph[p];
        int sg, row, c, k; decode(ph, s - base, sg, row, c, k);
        const ch_seg & S = ph.seg[sg];
        s_tab[s] = (uint32_t) row | ((uint32_t) S.map << 18) | ((uint32_t) (S.q6 != 0) << 21) | ((uint32_t) c << 22) | ((uint32_t) sg << 27);
    }
    __syncthreads();
    auto issue = [&](int s, int slot) {               // one lane: fetch stage s into its slot -- ONE tensor request of CH_R row chunks
        const uint32_t e = s_tab[s];
        const uint32_t cb = (e >> 21) & 1u ? CH_Q6K_CHUNK : CH_Q4K_CHUNK;
        const uint32_t bar = full0 + 8 * slot, dst = ring + slot * P.slot_bytes;
        if (TR && P.trace && cta == 0 && s < 1024) P.trace[32 * (1 + P.trace_cap) + 1024 + s] = ch_now();   // (measurement only)
        mbar_arrive_expect_tx(bar, cb * CH_R);
        tma_load_2d(dst, &P.map[(e >> 18) & 7u], (int) (((e >> 22) & 31u) * (cb >> 3)), (int) (e & 0x3ffffu), bar);
    };
    // first lap: weights do not depend on the kernel in front (PDL): the ring fills under its tail
    if (lane == 0) for (int q = warp; q < P.nslot && q < total; q += CH_NW) issue(q, q);
    // this warp's cursor in the stage sequence: it owns the slots q with q % CH_NW == warp and meets them lap after lap.  (Stages dealt
    // round-robin over the warps instead -- slot s % nslot handed from warp to warp behind a per-slot "lap issued" word -- measured 2-3 us per
    // layer SLOWER although it balances short phases: a slow warp then stalls the warp that inherits its slot.)
    int my_slot = warp, my_lap = 0;

    // ------------------------------------------------ the phases ------------------------------------------------
    unsigned long long * rec = nullptr;               // (measurement only) words 2 + 5 p + j: phase p entered, input complete, image built, stages done, published
    if (TR && P.trace && tid == 0) rec = b200_trace_open(P.trace, P.trace_cap, P.nph, 0xC4A1);
#define CH_STAMP(j) do { if (TR && rec) rec[2 + 5 * p + (j)] = ch_now(); } while (0)
    int s_base = 0;                                   // stages of this CTA before the current phase
    unsigned target = 0, tag = 0;
    // this thread's 16 values of the NEXT residual ADD's addend: fetched right at the start when it comes from outside, afterwards the sum
    // itself (out_add) for the ADD after that.  It never has to be read back from memory in mid-launch: the graph allocator likes to hand
    // an addend's block, dead after its ADD, to one of the results of the very phase that adds it
    float hkeep[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) hkeep[i] = 0.0f;
    // owner-side hand-over (CH_IN_BQ8N): thread t < 32 * (groups of this CTA) is row t of those groups in every producing phase -- same rows
    // each time -- and carries that row's residual from one ADD to the next in a register
    float keep = 0.0f;
    for (int p = 0; p < P.nph; ++p) {
        const ch_phase & ph = P.ph[p];
        const int K = ph.K, nblocks = K >> 8, nu16 = K >> 4;
        const ch_own own = ch_owned(ph, cta, G);
        CH_STAMP(0);
        // what does not depend on the phase before is fetched ahead of the hand-over: the norm weight, and a residual from outside
        // (every CTA reads the same lines at the same moment: ~9 B/clk per SM of L2 broadcast bandwidth, measured -- bytes on the critical path count)
        const bool own_in = ph.in_mode == CH_IN_BQ8N || (SP && ph.in_mode == CH_IN_LQ8N);   // the row owners of the phase before did ADD, norm weight and quantization
        const bool normp = ph.prologue != B200_CHAIN_X_PLAIN && !own_in, addp = ph.prologue == B200_CHAIN_X_ADD_NORM_MUL && !own_in, live = normp && tid < nu16;
        const bool own_out = (ph.out_mode == CH_IN_BQ8N || (SP && ph.out_mode == CH_IN_LQ8N)) && tid < own.gl * 32;      // (warp-uniform: whole groups)
        float nwpre = 0.0f;                                      // the next phase's norm weight of this thread's row
        if (own_out) nwpre = __ldg(ph.nx_normw + ch_unit(ph, tid >> 2, own, G) * CH_R + (tid & 3));
        float4 gpre[4], bpre[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            gpre[j] = bpre[j] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (live) gpre[j] = __ldg((const float4 *) (ph.normw + (size_t) tid * 16 + 4 * j));
        }
        if (p == 0) {
            asm volatile("griddepcontrol.wait;" ::: "memory");
            asm volatile("ld.volatile.global.u32 %0, [%1];" : "=r"(tag) : "l"(SP ? P.epoch : P.sync + 3));
            tag += 1u;                                // this launch's hand-over tag: differs from whatever an earlier launch left in the areas
            if (P.pre_in1 >= 0 && tid < (P.ph[P.pre_in1].K >> 4)) {
#pragma unroll
                for (int j = 0; j < 4; ++j) { const float4 v = ch_ldcg4(P.ph[P.pre_in1].in1 + (size_t) tid * 16 + 4 * j); hkeep[4 * j] = v.x; hkeep[4 * j + 1] = v.y; hkeep[4 * j + 2] = v.z; hkeep[4 * j + 3] = v.w; }
            }
            if (P.own_in1 >= 0) {                             // the first owner-side ADD's addend comes from outside: this thread's row of it
                const ch_phase & pq = P.ph[P.own_in1];
                const ch_own oq = ch_owned(pq, cta, G);
                if (tid < oq.gl * 32) keep = ch_ldcg_f(pq.nx_in1 + ch_unit(pq, tid >> 2, oq, G) * CH_R + (tid & 3));
            }
        }
        if (live && addp) {
            if (ph.in1_kept || p == P.pre_in1) {
#pragma unroll
                for (int j = 0; j < 4; ++j) bpre[j] = make_float4(hkeep[4 * j], hkeep[4 * j + 1], hkeep[4 * j + 2], hkeep[4 * j + 3]);
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) bpre[j] = ch_ldcg4(ph.in1 + (size_t) tid * 16 + 4 * j);
            }
        }
        // (1) CH_IN_BARRIER only: every CTA has stored and released the results of the phase before
        if (ph.in_mode == CH_IN_BARRIER || ph.in_mode == CH_IN_BQ8 || ph.in_mode == CH_IN_BQ8N) {
            if (tid == 0) {
                ch_watch w; unsigned v;
                while (true) {
                    asm volatile("ld.acquire.gpu.global.u32 %0, [%1];" : "=r"(v) : "l"(P.sync) : "memory");
                    if (v >= target) break;
                    if (w.spin(P.sync, 0x200u + p)) break;
                }
            }
            ch_bar_consumers();
        }
        CH_STAMP(1);
        // (2) this CTA's q8_1 image of the activation column
        const uint8_t * llin = P.ll + (size_t) (p > 0 ? p - 1 : 0) * CH_LL_BYTES;
        if (SP && ph.in_mode == CH_IN_LQ8N) {
            // the owner-side records of all ranks, validated word by word by their tags: this thread's units (K <= 64 CH_NT), then the norm's scale as below
            int4 w[4]; int2 ds[4];
            float ssp = 0.0f;
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int u = tid + i * CH_NT;
                w[i] = make_int4(0, 0, 0, 0); ds[i] = make_int2(0, 0);
                if (u < nu16) {
                    const uint8_t * rec = llin + (size_t) (u >> 1) * CH_LQ_REC;
                    ch_watch wt;
                    while (true) {
                        const uint4 q0 = ch_ll_load2(rec + (u & 1) * 32), q1 = ch_ll_load2(rec + (u & 1) * 32 + 16), dd = ch_ll_load2(rec + 64);
                        unsigned sv = 0, st = tag;
                        if (!(u & 1)) asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(sv), "=r"(st) : "l"(rec + 80) : "memory");
                        if (q0.y == tag && q0.w == tag && q1.y == tag && q1.w == tag && dd.y == tag && dd.w == tag && st == tag) {
                            w[i] = make_int4((int) q0.x, (int) q0.z, (int) q1.x, (int) q1.z); ds[i] = make_int2((int) dd.x, (int) dd.z);
                            if (!(u & 1)) ssp += __uint_as_float(sv);
                            break;
                        }
                        if (wt.spin(P.sync, 0x900u + p)) break;
                        if (P.poll_ns) __nanosleep(P.poll_ns);
                    }
                }
            }
            ssp = warp_sum(ssp);
            if (lane == 0) s_red[warp] = ssp;
            ch_bar_consumers();
            float tot = 0.0f;
#pragma unroll
            for (int i = 0; i < CH_NW; ++i) tot += s_red[i];
            const float scale = rsqrtf(tot / (float) K + ph.eps);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int u = tid + i * CH_NT;
                if (u < nu16) {
                    const float d = __int_as_float(ds[i].x) * scale;
                    ch_store_unit(u, w[i], __half2float(__float2half_rn(d)), __half2float(__float2half_rn(__fmul_rn(d, (float) ds[i].y))), K, nblocks, ph.types, s_act);
                }
            }
        } else if (own_in) {
            // (result + residual) * norm weight arrives as q8_1 blocks whose scale still lacks RMS_NORM's factor; the block sums of squares give it.
            // This thread's units (K <= 64 CH_NT) are in flight while the sum is reduced
            const uint8_t * hq = llin + CH_HQ, * hds = llin + CH_HDS, * hss = llin + CH_HSS;
            int4 w[4]; int2 ds[4];
            float ssp = 0.0f;
            {
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int u = tid + i * CH_NT;
                    w[i] = make_int4(0, 0, 0, 0); ds[i] = make_int2(0, 0);
                    if (u < nu16) { w[i] = ch_ldcg_i4(hq + (size_t) u * 16); ds[i] = ch_ldcg_i2(hds + (size_t) (u >> 1) * 8); }
                }
                for (int b = tid; b < (K >> 5); b += CH_NT) ssp += ch_ldcg_f(hss + (size_t) b * 4);
            }
            ssp = warp_sum(ssp);
            if (lane == 0) s_red[warp] = ssp;
            ch_bar_consumers();
            float tot = 0.0f;
#pragma unroll
            for (int i = 0; i < CH_NW; ++i) tot += s_red[i];
            const float scale = rsqrtf(tot / (float) K + ph.eps);
#pragma unroll
            for (int i = 0; i < 4; ++i) {
                const int u = tid + i * CH_NT;
                if (u < nu16) {
                    const float d = __int_as_float(ds[i].x) * scale;
                    ch_store_unit(u, w[i], __half2float(__float2half_rn(d)), __half2float(__float2half_rn(__fmul_rn(d, (float) ds[i].y))), K, nblocks, ph.types, s_act);
                }
            }
        } else if (ph.in_mode == CH_IN_BQ8) {
            // the SWIGLU phase before quantized its own 32-row blocks: 32 quants + (d, sum) per block, plain loads behind the arrival counter
            const int nq16 = ph.nq8 >> 4;
            const uint8_t * hq = llin + CH_HQ, * hds = llin + CH_HDS;
            for (int u0 = tid; u0 < nq16; u0 += 4 * CH_NT) {
                int4 w[4]; int2 ds[4];
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int u = u0 + i * CH_NT;
                    w[i] = make_int4(0, 0, 0, 0); ds[i] = make_int2(0, 0);
                    if (u < nq16) { w[i] = ch_ldcg_i4(hq + (size_t) u * 16); ds[i] = ch_ldcg_i2(hds + (size_t) (u >> 1) * 8); }
                }
#pragma unroll
                for (int i = 0; i < 4; ++i) {
                    const int u = u0 + i * CH_NT;
                    if (u < nq16) {
                        const float d = __int_as_float(ds[i].x);
                        ch_store_unit(u, w[i], __half2float(__float2half_rn(d)), __half2float(__float2half_rn(__fmul_rn(d, (float) ds[i].y))), K, nblocks, ph.types, s_act);
                    }
                }
            }
            // its tail rows (not whole blocks per CTA) come as f32: at most a few units, one warp
            const int ntail16 = nu16 - nq16;
            if (ntail16 > 0 && warp == CH_NW - 1) {
                const float * tl = (const float *) (llin + CH_HTAIL);
                for (int u0 = 0; u0 < ntail16; u0 += 32) {
                    const int u = u0 + lane; const bool on = u < ntail16;
                    float f[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) f[i] = 0.0f;
                    if (on) {
#pragma unroll
                        for (int j = 0; j < 4; ++j) { const float4 v = ch_ldcg4(tl + (size_t) u * 16 + 4 * j); f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w; }
                    }
                    ch_quant_store(f, lane, nq16 + u, on, K, nblocks, ph.types, s_act);
                }
            }
        } else if (!normp && ph.in_mode == CH_IN_LL_Q8) {
            // the SWIGLU phase before quantized its own 32-row blocks: 10 (word, tag) pairs per block (8 quant words, (d, s), pad)
            const int nq16 = ph.nq8 >> 4;
            for (int u = tid; u < nq16; u += CH_NT) {
                const uint8_t * b = llin + (size_t) (u >> 1) * 80;
                uint4 q0, q1, ds;
                ch_watch w;
                while (true) {
                    q0 = ch_ll_load2(b + (u & 1) * 32); q1 = ch_ll_load2(b + (u & 1) * 32 + 16); ds = ch_ll_load2(b + 64);
                    if (q0.y == tag && q0.w == tag && q1.y == tag && q1.w == tag && ds.y == tag) break;
                    if (w.spin(P.sync, 0x400u + p)) break;
                    if (SP && P.poll_ns) __nanosleep(P.poll_ns);
                }
                const float2 f = __half22float2(*(const __half2 *) &ds.x);
                ch_store_unit(u, make_int4((int) q0.x, (int) q0.z, (int) q1.x, (int) q1.z), f.x, f.y, K, nblocks, ph.types, s_act);
            }
            // its tail rows (not whole blocks per CTA) come as f32 pairs behind the block area: at most a few units, one warp
            const int ntail16 = nu16 - nq16;
            if (ntail16 > 0 && warp == CH_NW - 1) {
                const uint8_t * tl = llin + (size_t) (ph.nq8 >> 5) * 80;
                for (int u0 = 0; u0 < ntail16; u0 += 32) {
                    const int u = u0 + lane; const bool on = u < ntail16;
                    float f[16];
#pragma unroll
                    for (int i = 0; i < 16; ++i) f[i] = 0.0f;
                    if (on) {
                        ch_watch w;
                        while (true) {
                            bool ok = true;
#pragma unroll
                            for (int j = 0; j < 8; ++j) { const uint4 v = ch_ll_load2(tl + ((size_t) u * 16 + 2 * j) * 8); f[2 * j] = __uint_as_float(v.x); f[2 * j + 1] = __uint_as_float(v.z); ok = ok && v.y == tag && v.w == tag; }
                            if (ok) break;
                            if (w.spin(P.sync, 0x500u + p)) break;
                        }
                    }
                    ch_quant_store(f, lane, nq16 + u, on, K, nblocks, ph.types, s_act);
                }
            }
        } else if (!normp) {
            // plain column (from outside, after the barrier, or as (value, tag) pairs): double-buffered rounds of 16 values per thread
            const bool pairs = ph.in_mode == CH_IN_LL_F32;
            auto load16 = [&](int u, float (&f)[16]) {
                if (!pairs) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) { const float4 v = ch_ldcg4(ph.in0 + (size_t) u * 16 + 4 * j); f[4 * j] = v.x; f[4 * j + 1] = v.y; f[4 * j + 2] = v.z; f[4 * j + 3] = v.w; }
                } else {
                    ch_watch w;
                    while (true) {
                        bool ok = true;
#pragma unroll
                        for (int j = 0; j < 8; ++j) { const uint4 v = ch_ll_load2(llin + ((size_t) u * 16 + 2 * j) * 8); f[2 * j] = __uint_as_float(v.x); f[2 * j + 1] = __uint_as_float(v.z); ok = ok && v.y == tag && v.w == tag; }
                        if (ok) break;
                        if (w.spin(P.sync, 0x600u + p)) break;
                    }
                    if (SP && P.nrank > 1 && cta == 0) {            // row split: the gathered column also becomes the plain f32 tensor it is in the graph
#pragma unroll
                        for (int j = 0; j < 4; ++j) *(float4 *) ((float *) ph.in0 + (size_t) u * 16 + 4 * j) = make_float4(f[4 * j], f[4 * j + 1], f[4 * j + 2], f[4 * j + 3]);
                    }
                }
            };
            for (int base = tid & ~31; base < nu16; base += CH_NT) {
                const int u = base + lane;
                float f[16];
#pragma unroll
                for (int i = 0; i < 16; ++i) f[i] = 0.0f;
                if (u < nu16) load16(u, f);
                ch_quant_store(f, lane, u, u < nu16, K, nblocks, ph.types, s_act);
            }
        } else {
            // h = in0 (+ in1); n = h / sqrt(mean(h^2) + eps); x = n * normw.  K <= 16 * CH_NT: thread = one 16-value unit
            float h[16], g[16];
            if (live && ph.in_mode == CH_IN_LL_F32) {
                ch_watch w;
                while (true) {
                    bool ok = true;
#pragma unroll
                    for (int j = 0; j < 8; ++j) { const uint4 v = ch_ll_load2(llin + ((size_t) tid * 16 + 2 * j) * 8); h[2 * j] = __uint_as_float(v.x); h[2 * j + 1] = __uint_as_float(v.z); ok = ok && v.y == tag && v.w == tag; }
                    if (ok) break;
                    if (w.spin(P.sync, 0x700u + p)) break;
                }
                if (SP && P.nrank > 1 && cta == 0) {
#pragma unroll
                    for (int j = 0; j < 4; ++j) *(float4 *) ((float *) ph.in0 + (size_t) tid * 16 + 4 * j) = make_float4(h[4 * j], h[4 * j + 1], h[4 * j + 2], h[4 * j + 3]);
                }
            } else {
#pragma unroll
                for (int j = 0; j < 4; ++j) {
                    const float4 a = live ? ch_ldcg4(ph.in0 + (size_t) tid * 16 + 4 * j) : make_float4(0.f, 0.f, 0.f, 0.f);
                    h[4 * j] = a.x; h[4 * j + 1] = a.y; h[4 * j + 2] = a.z; h[4 * j + 3] = a.w;
                }
            }
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                if (!live) { h[4 * j] = h[4 * j + 1] = h[4 * j + 2] = h[4 * j + 3] = 0.0f; }
                h[4 * j] += bpre[j].x; h[4 * j + 1] += bpre[j].y; h[4 * j + 2] += bpre[j].z; h[4 * j + 3] += bpre[j].w;   // (zeros without ADD)
                g[4 * j] = gpre[j].x; g[4 * j + 1] = gpre[j].y; g[4 * j + 2] = gpre[j].z; g[4 * j + 3] = gpre[j].w;
            }
            if (addp) {
#pragma unroll
                for (int i = 0; i < 16; ++i) hkeep[i] = h[i];
            }
            if (rec && p == 1 && h[0] == 12345.678f) rec[25] = 1;       // (measurement only: the loads have landed when the stamp below is taken)
            if (rec && p == 1) rec[26] = ch_now();
            float ss = 0.0f;
#pragma unroll
            for (int i = 0; i < 16; ++i) ss = fmaf(h[i], h[i], ss);
            ss = warp_sum(ss);
            if (lane == 0) s_red[warp] = ss;
            ch_bar_consumers();
            float tot = 0.0f;
#pragma unroll
            for (int i = 0; i < CH_NW; ++i) tot += s_red[i];
            if (rec && p == 1) rec[27] = ch_now();
            const float scale = rsqrtf(tot / (float) K + ph.eps);
            float x[16];
#pragma unroll
            for (int i = 0; i < 16; ++i) x[i] = h[i] * scale;
            // the three nodes' values are stored once, each by another CTA
            if (live) {
                if (addp && ph.out_add && cta == 0) {
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
            if (rec && p == 1) rec[28] = ch_now();
        }
        ch_bar_consumers();
        CH_STAMP(2);
        // the next kernel of the stream may start its prologue once the last phase is under way (PDL)
        if (p == P.nph - 1) asm volatile("griddepcontrol.launch_dependents;" ::: "memory");

        // (3) this warp's stages of the phase
        const int nunits_l = own.nu;
        const int nst = total ? cum[p + 1] - cum[p] : 0;
        const uint32_t act4 = smem_u32(s_act), act6 = act4 + ((ph.types & 1) ? (uint32_t) gs_traits<B200_TYPE_Q4_K>::act_bytes(K) : 0u);
        // A ring slot always belongs to the same warp (slot % CH_NW), whatever the lap: a warp then meets the laps of a slot in order, and its
        // parity wait can never take the slot's previous, still incomplete fill for its own (the fills of one lap complete in any order).
        while (true) {
            const int s = my_lap * P.nslot + my_slot, sl = s - s_base;
            if (sl >= nst) break;
            const uint32_t e = s_tab[s];
            const bool q6 = (e >> 21) & 1u;
            const int c = (int) ((e >> 22) & 31u);
            ch_mbar_wait(full0 + 8 * my_slot, (uint32_t) my_lap & 1u, P.sync, 0x300u + p);
            if (TR && P.trace && cta == 0 && lane == 0 && s < 1024) P.trace[32 * (1 + P.trace_cap) + s] = ch_now();   // (measurement only)
            const uint32_t st = ring + my_slot * P.slot_bytes;
            float a0, a1;                              // this lane's sums of rows 2 (lane >> 4) + {0, 1}
            if (q6) {
                float acc[CH_R];
#pragma unroll
                for (int r = 0; r < CH_R; ++r) acc[r] = 0.0f;
                const int bl = lane & 7, gb = c * GS_CB + bl;
                if (!(TR && (P.dbg & 1))) gs_traits<B200_TYPE_Q6_K>::consume<CH_R, CH_R>(st, CH_Q6K_CHUNK, bl, gb, lane, act6, K, nblocks, acc);
                // every lane has all four rows: halve the number of values a lane carries
                const bool hi16 = lane & 16;
                a0 = hi16 ? acc[2] : acc[0]; a1 = hi16 ? acc[3] : acc[1];
                const float b0 = hi16 ? acc[0] : acc[2], b1 = hi16 ? acc[1] : acc[3];
                a0 += __shfl_xor_sync(0xffffffffu, b0, 16); a1 += __shfl_xor_sync(0xffffffffu, b1, 16);   // lanes < 16: rows 0, 1; lanes >= 16: rows 2, 3
            } else {
                float acc[2] = { 0.0f, 0.0f };
                if (!(TR && (P.dbg & 1))) ch_consume_q4k(st, lane, act4, K, nblocks, c, acc);
                a0 = acc[0]; a1 = acc[1];
            }
            __syncwarp();                              // every lane is done reading the slot
            if (lane == 0 && s + P.nslot < total) issue(s + P.nslot, my_slot);
            const bool hi8 = lane & 8;
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

        // (4) sum the K chunks of every row in chunk order, finish the node(s), publish for the next phase
        uint8_t * llout = P.ll + (size_t) p * CH_LL_BYTES;
        if (ph.epilogue == B200_CHAIN_Y_SWIGLU) {
            auto row_of = [&](int t, float & gsum, float & usum) {       // local row t = 4 k + r
                const int k = t / CH_R, r = t - k * CH_R;
                gsum = 0.0f; usum = 0.0f;
                for (int c = 0; c < ph.nch; ++c) { gsum += s_psum[((2 * k) * ph.nch + c) * CH_R + r]; usum += s_psum[((2 * k + 1) * ph.nch + c) * CH_R + r]; }
                return ch_unit(ph, k, own, G) * CH_R + r;
            };
            const int nrow_g = (ph.out_mode == CH_IN_LL_Q8 || ph.out_mode == CH_IN_BQ8) ? own.gl * 32 : 0;
            // the CTA's whole 32-row blocks: one warp per block, lane = row: gate, up, glu, and the block's q8_1 form for ffn_down
            for (int b0 = warp * 32; b0 < nrow_g; b0 += CH_NT) {
                float gsum, usum;
                const int row = row_of(b0 + lane, gsum, usum);
                const float v = b200_swiglu1(gsum, usum);
                ph.seg[0].dst[row] = gsum; ph.seg[1].dst[row] = usum; ph.glu[row] = v;
                int q, isum; float d;
                q8_1_block<B200_Q8_1_X86>(v, q, d, isum);
                unsigned wq = (unsigned) q & 0xffu;
                wq |= __shfl_down_sync(0xffffffffu, wq, 1) << 8;
                wq |= __shfl_down_sync(0xffffffffu, wq, 2) << 16;
                if (ph.out_mode == CH_IN_BQ8) {                            // (the warp's 32 rows are one aligned block)
                    if ((lane & 3) == 0) *(unsigned *) (llout + CH_HQ + (size_t) (row >> 5) * 32 + lane) = wq;
                    if (lane == 1) *(int2 *) (llout + CH_HDS + (size_t) (row >> 5) * 8) = make_int2(__float_as_int(d), isum);
                } else {
                    const size_t blk = (size_t) p * CH_LL_BYTES + (size_t) (((SP ? ph.seg[0].row0 : 0) + row) >> 5) * 80;
                    if ((lane & 3) == 0) ch_ll_pub<SP>(P, blk + (lane >> 2) * 8, wq, tag);
                    if (lane == 1) { const __half2 ds = q8_1_ds(d, isum); ch_ll_pub<SP>(P, blk + 64, *(const unsigned *) &ds, tag); }
                }
            }
            for (int t = nrow_g + tid; t < nunits_l * CH_R; t += CH_NT) {
                float gsum, usum;
                const int row = row_of(t, gsum, usum);
                const float v = b200_swiglu1(gsum, usum);
                ph.seg[0].dst[row] = gsum; ph.seg[1].dst[row] = usum; ph.glu[row] = v;
                if (ph.out_mode == CH_IN_LL_F32) ch_ll_pub<SP>(P, (size_t) p * CH_LL_BYTES + (size_t) ((SP ? ph.seg[0].row0 : 0) + row) * 8, __float_as_uint(v), tag);
                else if (ph.out_mode == CH_IN_LL_Q8) ch_ll_store(llout + (size_t) (ph.ng8 * CH_R >> 5) * 80 + (size_t) (row - ph.ng8 * CH_R) * 8, __float_as_uint(v), tag);
                else if (ph.out_mode == CH_IN_BQ8) *(float *) (llout + CH_HTAIL + (size_t) (row - ph.ng8 * CH_R) * 4) = v;
            }
        } else if (ph.out_mode == CH_IN_BQ8N || (SP && ph.out_mode == CH_IN_LQ8N)) {
            // thread = one row of this CTA's groups (whole q8_1 blocks of the next phase's activation; warp = group): the matmul result, the
            // residual sum h, and h * norm weight as a q8_1 block with the block's sum of squares of h next to it
            if (own_out) {
                const int k = tid >> 2, r = tid & 3;
                float sum = 0.0f;
                for (int c = 0; c < ph.nch; ++c) sum += s_psum[(k * ph.nch + c) * CH_R + r];
                const int row = ch_unit(ph, k, own, G) * CH_R + r;
                ph.seg[0].dst[row] = sum;
                const float h = sum + keep;
                if (ph.nx_out_add) ph.nx_out_add[row] = h;
                keep = h;
                const float ss = warp_sum(h * h);
                int q, isum; float d;
                q8_1_block<B200_Q8_1_X86>(h * nwpre, q, d, isum);
                unsigned wq = (unsigned) q & 0xffu;
                wq |= __shfl_down_sync(0xffffffffu, wq, 1) << 8;
                wq |= __shfl_down_sync(0xffffffffu, wq, 2) << 16;
                if (SP && ph.out_mode == CH_IN_LQ8N) {
                    const size_t rec = (size_t) p * CH_LL_BYTES + (size_t) ((ph.seg[0].row0 + row) >> 5) * CH_LQ_REC;
                    if ((lane & 3) == 0) ch_ll_pub<SP>(P, rec + (lane >> 2) * 8, wq, tag);
                    if (lane == 1) ch_ll_pub<SP>(P, rec + 64, (unsigned) __float_as_int(d), tag);
                    if (lane == 2) ch_ll_pub<SP>(P, rec + 72, (unsigned) isum, tag);
                    if (lane == 3) ch_ll_pub<SP>(P, rec + 80, __float_as_uint(ss), tag);
                } else {
                    if ((lane & 3) == 0) *(unsigned *) (llout + CH_HQ + (size_t) (row >> 5) * 32 + lane) = wq;
                    if (lane == 1) *(int2 *) (llout + CH_HDS + (size_t) (row >> 5) * 8) = make_int2(__float_as_int(d), isum);
                    if (lane == 2) *(float *) (llout + CH_HSS + (size_t) (row >> 5) * 4) = ss;
                }
            }
        } else if (p == P.tail_phase) {
            // the layer's last matmuls with what the graph does to their rows next: ROPE (adjacent pairs: both rows of a pair sit in one unit,
            // i.e. in neighbouring lanes) and the f16 copy into the KV-cache row the SET_ROWS index names.  Same expressions as gl_rope_rows (b200_glue.cu)
            for (int t0 = warp * 32; t0 < nunits_l * CH_R; t0 += CH_NT) {
                const int t = t0 + lane; const bool on = t < nunits_l * CH_R;
                const int k = t / CH_R, r = t - k * CH_R;
                float sum = 0.0f;
                int sg = 0, row = 0;
                if (on) {
                    for (int c = 0; c < ph.nch; ++c) sum += s_psum[(k * ph.nch + c) * CH_R + r];
                    ch_locate(ph, ch_unit(ph, k, own, G), sg, row);
                }
                const float other = __shfl_xor_sync(0xffffffffu, sum, 1);
                if (!on) continue;
                const int rw = row + r;
                const ch_tail & T = P.tail[sg];
                const bool rope = T.kind == B200_CHAIN_T_ROPE || T.kind == B200_CHAIN_T_ROPE_CACHE;
                if (!(rope && T.out == ph.seg[sg].dst) && !T.nodst) ph.seg[sg].dst[rw] = sum;        // (a ROPE in place leaves only its own value there)
                float y = sum;
                if (rope) {
                    const int d = rw % T.head_dim, ip = d >> 1;
                    if (d < T.n_dims) {
                        const float pos = (float) __ldg(T.pos);
                        const float theta = T.fs * (pos * powf(T.ts, (float) ip) / (T.ff ? __ldg(T.ff + ip) : 1.0f));
                        float sn, cs; sincosf(theta, &sn, &cs); sn *= T.af; cs *= T.af;
                        const float x0 = (d & 1) ? other : sum, x1 = (d & 1) ? sum : other;
                        y = (d & 1) ? x0 * sn + x1 * cs : x0 * cs - x1 * sn;
                    }
                    T.out[rw] = y;
                }
                if (T.kind == B200_CHAIN_T_ROPE_CACHE || T.kind == B200_CHAIN_T_CACHE) {
                    const long long crow = *T.cidx;
                    ((__half *) (T.cache + (size_t) crow * T.cnb1))[rw] = __float2half_rn(y);
                }
            }
        } else {
            for (int t = tid; t < nunits_l * CH_R; t += CH_NT) {
                const int k = t / CH_R, r = t - k * CH_R;
                float sum = 0.0f;
                for (int c = 0; c < ph.nch; ++c) sum += s_psum[(k * ph.nch + c) * CH_R + r];
                int sg, row; ch_locate(ph, ch_unit(ph, k, own, G), sg, row);
                ph.seg[sg].dst[row + r] = sum;
                if (ph.out_mode == CH_IN_LL_F32) ch_ll_pub<SP>(P, (size_t) p * CH_LL_BYTES + (size_t) ((SP ? ph.seg[0].row0 : 0) + row + r) * 8, __float_as_uint(sum), tag);   // (single-segment phases only)
                else if (SP && P.fin_total && p == P.nph - 1) ch_ll_pub<SP>(P, (size_t) P.fin_off + (size_t) (P.fin[sg].base + ph.seg[sg].row0 + row + r) * 8, __float_as_uint(sum), tag);
            }
        }
        if (ph.out_mode == CH_IN_BARRIER || ph.out_mode == CH_IN_BQ8 || ph.out_mode == CH_IN_BQ8N) {
            ch_bar_consumers();                       // every thread's stores are ordered before thread 0's release (CTA barrier), whose
            if (tid == 0) asm volatile("red.release.gpu.global.add.u32 [%0], 1;" :: "l"(P.sync) : "memory");   // cumulativity makes them visible GPU-wide
            target += (unsigned) G;
        }
        CH_STAMP(4);
    }
    // row split: the last phase's results, gathered from every rank, as plain f32 columns (this rank's own rows were stored above: skipped)
    if (SP && P.fin_total) {
        const uint8_t * fin = P.ll + P.fin_off;
        for (int i = cta * CH_NT + tid; i < P.fin_total; i += G * CH_NT) {
            int sg = 0;
            if (i >= P.fin[1].base && P.fin[1].M) sg = 1;
            if (i >= P.fin[2].base && P.fin[2].M) sg = 2;
            const int row = i - P.fin[sg].base;
            const ch_seg & S = P.ph[P.nph - 1].seg[sg];
            if (row >= S.row0 && row < S.row0 + S.M) continue;
            ch_watch w; unsigned v, t;
            while (true) {
                asm volatile("ld.volatile.global.v2.u32 {%0, %1}, [%2];" : "=r"(v), "=r"(t) : "l"(fin + (size_t) i * 8) : "memory");
                if (t == tag) break;
                if (w.spin(P.sync, 0x800u)) break;
            }
            P.fin[sg].dst[row] = __uint_as_float(v);
        }
    }
    // the last CTA to finish re-arms the counters and advances the epoch for the next launch (every CTA has passed every wait by then)
    if (P.nph > 1) {
        ch_bar_consumers();
        if (tid == 0) {
            const unsigned done = atomicAdd(P.sync + 1, 1u);
            if (done == (unsigned) G - 1) { P.sync[0] = 0u; P.sync[1] = 0u; if (SP) *P.epoch = tag; else P.sync[3] = tag; __threadfence(); }
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
static int64_t s_rows_max(const b200_chain_phase & s) { int64_t m = 0; for (int i = 0; i < s.n_w; ++i) m = s.w[i]->ne[1] > m ? s.w[i]->ne[1] : m; return m; }
static bool ch_overlap(const void * a, size_t na, const b200_tensor * t) {
    if (!a || !t) return false;
    const char * a0 = (const char *) a, * b0 = (const char *) t->data;
    return a0 < b0 + b200_tensor_span(t) && b0 < a0 + na;
}

typedef CUresult (*encode_tiled_fn)(CUtensorMap *, CUtensorMapDataType, cuuint32_t, void *, const cuuint64_t *, const cuuint64_t *,
                                    const cuuint32_t *, const cuuint32_t *, CUtensorMapInterleave, CUtensorMapSwizzle,
                                    CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
encode_tiled_fn b200_get_encode_tiled();            // b200_gemm.cu

static int ch_nw() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_CHAIN_WARPS"); v = e ? atoi(e) : 16; if (v != 16 && v != 24) v = 16; } return v; }
// GGML_B200_CHAIN_HANDOVER=ll switches to the flag-in-data hand-over.  Default: the arrival counter.  Measured on Llama-3-8B layers (scripts/chain_bench.py
// --trace): a poll of the data costs a full L2 round trip under the weight stream (~1.2 us) and the first one usually comes too early, so the
// flag-in-data form (0.5 + 2.6 + 2.3 us per boundary) is no faster than release / acquire on one counter (1.5 + 1.3 + 2.0 us).
static bool ch_force_barrier() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_CHAIN_HANDOVER"); v = !(e && e[0] == 'l'); } return v != 0; }
// GGML_B200_CHAIN_OWNER=0 (measurement only) keeps every boundary in the plain form: f32 column through memory, every CTA normalises and quantizes it
static bool ch_owner_on() { static int v = -1; if (v < 0) { const char * e = getenv("GGML_B200_CHAIN_OWNER"); v = !(e && e[0] == '0'); } return v != 0; }
// whole 32-row groups dealt round-robin over G CTAs: the busiest CTA may carry at most 16 % more than the average
static bool ch_groups_balanced(int ngroups, int G) { const int per = (ngroups + G - 1) / G; return ngroups > 0 && (long long) per * G * 100 <= (long long) ngroups * 116 && per <= 16; }

// fills the kernel arguments; B200_ERR_UNSUPPORTED (with a reason) when the chain is not one this kernel takes
static size_t kb_of(int K) { return (size_t) K * 4; }
static bool ch_is_own_norm(int mode) { return mode == CH_IN_BQ8N || mode == CH_IN_LQ8N; }
// row split (sp): result i of phase p is the FULL column dst[i]; this rank's weight holds its rows [row_lo, row_lo + w->ne[1])
static bool ch_col_part_ok(const b200_tensor * t, int64_t lo, int64_t n) {
    return t && t->type == B200_TYPE_F32 && t->nb[0] == 4 && lo >= 0 && lo % 32 == 0 && t->ne[0] >= lo + n && t->ne[1] == 1 && t->ne[2] == 1 && t->ne[3] == 1 && ((uintptr_t) t->data & 15) == 0;
}
static int ch_build(b200_ctx * ctx, const b200_chain_phase * phases, int n, ch_args & A, size_t & smem, const b200_chain_split * sp = nullptr) {
    if (sp && (sp->world < 1 || sp->world > B200_MAX_PEERS || sp->rank < 0 || sp->rank >= sp->world)) { b200_set_error("decode_chain_split: bad world / rank"); return B200_ERR_INVALID; }
    if (!phases || n < 1 || n > CH_MAXPH) { b200_set_error("decode_chain: 1..%d phases", CH_MAXPH); return B200_ERR_UNSUPPORTED; }
    const int G = ctx->sm_count, nw = sp ? 16 : ch_nw();
    if (G < 3) { b200_set_error("decode_chain: needs at least 3 SMs"); return B200_ERR_UNSUPPORTED; }
    memset(&A, 0, sizeof(A));
    A.nph = n;
    int off = 0, slot = 0, act = 0, nmaps = 0;
    encode_tiled_fn encode = b200_get_encode_tiled();
    if (!encode) { b200_set_error("decode_chain: cuTensorMapEncodeTiled not available from the driver"); return B200_ERR_UNSUPPORTED; }
    // ---- pass 1: shapes, weights, tensor maps
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
            const int64_t lo = sp ? sp->row_lo[p][i] : 0;
            if (sp ? !ch_col_part_ok(s.dst[i], lo, s.w[i]->ne[1]) : !ch_col_ok(s.dst[i], s.w[i]->ne[1])) { b200_set_error("decode_chain: phase %d result %d is not an f32 column of %lld", p, i, (long long) s.w[i]->ne[1]); return B200_ERR_UNSUPPORTED; }
            d.seg[i].w = (const uint8_t *) s.w[i]->data; d.seg[i].nb1 = s.w[i]->nb[1]; d.seg[i].dst = (float *) s.dst[i]->data + lo; d.seg[i].M = (int) s.w[i]->ne[1]; d.seg[i].row0 = (int) lo;
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
            const int64_t lo = sp ? sp->row_lo[p][0] : 0;
            if (s.n_w != 2 || s.w[0]->ne[1] != s.w[1]->ne[1] || (sp ? (!ch_col_part_ok(s.glu, lo, s.w[0]->ne[1]) || sp->row_lo[p][1] != lo) : !ch_col_ok(s.glu, s.w[0]->ne[1]))) { b200_set_error("decode_chain: phase %d: SWIGLU wants gate, up of equal M and a glu column", p); return B200_ERR_UNSUPPORTED; }
            d.glu = (float *) s.glu->data + lo; d.nunits = (int) (s.w[0]->ne[1] / CH_R);
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
            d.out_add = add && s.out_add ? (float *) s.out_add->data : nullptr;
            d.out_norm = s.out_norm ? (float *) s.out_norm->data : nullptr; d.out_mul = s.out_mul ? (float *) s.out_mul->data : nullptr;
        } else if (s.prologue != B200_CHAIN_X_PLAIN) { b200_set_error("decode_chain: phase %d: unknown prologue %d", p, s.prologue); return B200_ERR_UNSUPPORTED; }
        int a = 0;
        if (d.types & 1) a += (int) gs_traits<B200_TYPE_Q4_K>::act_bytes((int) K);
        if (d.types & 2) a += (int) gs_traits<B200_TYPE_Q6_K>::act_bytes((int) K);
        act = max(act, a);
    }
    // ---- the tail of the last phase
    A.tail_phase = -1;
    for (int p = 0; p < n; ++p) {
        const b200_chain_phase & s = phases[p];
        bool any = false;
        for (int i = 0; i < s.n_w; ++i) any = any || (s.tail[i] && s.tail[i]->kind != B200_CHAIN_T_NONE);
        if (!any) continue;
        if (sp) { b200_set_error("decode_chain_split: no ROPE / KV-cache tail on a row-split chain"); return B200_ERR_UNSUPPORTED; }
        if (p != n - 1 || s.epilogue != B200_CHAIN_Y_PLAIN) { b200_set_error("decode_chain: phase %d: only the last phase, without GLU, takes a tail", p); return B200_ERR_UNSUPPORTED; }
        A.tail_phase = p;
        for (int i = 0; i < s.n_w; ++i) {
            ch_tail & T = A.tail[i];
            const b200_chain_rope * r = s.tail[i];
            if (!r || r->kind == B200_CHAIN_T_NONE) { T.kind = B200_CHAIN_T_NONE; continue; }
            const int64_t M = s.w[i]->ne[1];
            const bool rope = r->kind == B200_CHAIN_T_ROPE || r->kind == B200_CHAIN_T_ROPE_CACHE, cache = r->kind == B200_CHAIN_T_ROPE_CACHE || r->kind == B200_CHAIN_T_CACHE;
            if ((!rope && !cache) || (rope && (r->head_dim <= 0 || r->head_dim % 2 || M % r->head_dim || r->n_dims < 0 || r->n_dims % 2 || r->n_dims > r->head_dim || !r->pos || !r->out)) ||
                (cache && (!r->cache || !r->cache_idx || r->cache_nb1 < (size_t) M * 2 || ((uintptr_t) r->cache & 1) || (r->cache_nb1 & 1)))) {
                b200_set_error("decode_chain: phase %d: malformed tail of result %d", p, i); return B200_ERR_INVALID;
            }
            T.kind = r->kind; T.nodst = r->skip_dst; T.head_dim = r->head_dim; T.n_dims = r->n_dims; T.fs = r->freq_scale; T.ts = r->theta_scale; T.af = r->attn_factor;
            T.pos = r->pos; T.ff = r->freq_factors; T.out = rope ? r->out : nullptr; T.cidx = (const long long *) r->cache_idx; T.cache = (char *) r->cache; T.cnb1 = r->cache_nb1;
        }
    }
    // ---- pass 2: how every phase gets its inputs.  The flag-in-data hand-over needs in0 to be exactly what the phase before publishes
    // (its single result column, or its glu) and in1 to come from outside or from the last residual sum (kept in registers); anything
    // else that was written earlier in this launch forces the fence-and-counter form for the whole chain.
    bool ll = sp ? true : !ch_force_barrier();       // across GPUs only the flag-in-data form exists
    int last_add = -1;
    A.pre_in1 = -1; A.own_in1 = -1;
    int dkind[CH_MAXPH] = {}, in1_class[CH_MAXPH] = {}, add_before[CH_MAXPH] = {};   // in0 = the phase before's single result (1) / glu (2); addend: outside (0), the last ADD's sum (1), other (2)
    for (int p = 0; p < n; ++p) {
        const b200_chain_phase & s = phases[p];
        ch_phase & d = A.ph[p];
        const size_t kb = (size_t) d.K * 4;
        d.in_mode = CH_IN_EXTERNAL; d.in1_kept = 0;
        bool in1_inside = false;
        add_before[p] = last_add;
        for (int q = 0; q < p; ++q) {
            const b200_chain_phase & e = phases[q];
            const b200_tensor * outs[7] = { e.dst[0], e.n_w > 1 ? e.dst[1] : nullptr, e.n_w > 2 ? e.dst[2] : nullptr, e.out_add, e.out_norm, e.out_mul, e.glu };
            for (int o = 0; o < 7; ++o) {
                if (!outs[o]) continue;
                if (ch_overlap(s.in0->data, kb, outs[o])) {
                    const bool direct = q == p - 1 && s.in0->data == outs[o]->data &&
                                        ((o == 0 && e.n_w == 1 && e.epilogue == B200_CHAIN_Y_PLAIN) || (o == 6 && e.epilogue == B200_CHAIN_Y_SWIGLU));
                    d.in_mode = CH_IN_BARRIER;
                    if (direct) dkind[p] = o == 0 ? 1 : 2;
                    if (direct && ll) d.in_mode = (!sp && o == 6 && s.prologue == B200_CHAIN_X_PLAIN && d.K % 32 == 0) ? CH_IN_LL_Q8 : CH_IN_LL_F32;
                    else ll = false;
                }
                if (d.in1 && ch_overlap(s.in1->data, kb, outs[o])) {
                    if (o == 3 && q == last_add && s.in1->data == outs[o]->data && phases[q].w[0]->ne[0] == d.K) d.in1_kept = 1;
                    else { ll = false; in1_inside = true; }
                }
                if (q < p - 1 && s.normw && ch_overlap(s.normw->data, kb, outs[o])) ll = false;
            }
        }
        in1_class[p] = d.in1_kept ? 1 : in1_inside ? 2 : 0;
        if (s.prologue == B200_CHAIN_X_ADD_NORM_MUL) {
            if (last_add < 0 && !in1_inside && !d.in1_kept) A.pre_in1 = p;        // the chain's first ADD, addend from outside: fetched at the start
            last_add = p;
        }
    }
    for (int p = 0; p < n; ++p) {
        ch_phase & d = A.ph[p];
        if (!ll && d.in_mode != CH_IN_EXTERNAL) d.in_mode = CH_IN_BARRIER;
        if (p > 0) A.ph[p - 1].out_mode = d.in_mode == CH_IN_EXTERNAL ? (ll ? CH_IN_EXTERNAL : CH_IN_BARRIER) : d.in_mode;
    }
    if (sp) {
        // every boundary must be a direct hand-over (the phase before's single result or its glu), and a rank's rows of a handed-over column
        // one contiguous range: then the (value, tag) pairs of all ranks tile the next phase's activation
        for (int p = 1; p < n; ++p) if (!ll || A.ph[p].in_mode != CH_IN_LL_F32) { b200_set_error("decode_chain_split: phase %d does not take its activation straight from the phase before", p); return B200_ERR_UNSUPPORTED; }
        if (phases[n - 1].epilogue != B200_CHAIN_Y_PLAIN) { b200_set_error("decode_chain_split: the last phase must be a plain one"); return B200_ERR_UNSUPPORTED; }
        A.nrank = sp->world;
        for (int r = 0; r < sp->world; ++r) { if (!sp->ll[r]) { b200_set_error("decode_chain_split: hand-over area of rank %d missing", r); return B200_ERR_INVALID; } A.llr[r] = (uint8_t *) sp->ll[r]; }
        int base = 0;
        for (int i = 0; i < phases[n - 1].n_w; ++i) { A.fin[i].dst = (float *) phases[n - 1].dst[i]->data; A.fin[i].M = (int) phases[n - 1].dst[i]->ne[0]; A.fin[i].base = base; base += A.fin[i].M; }
        // a last phase whose weights every rank holds in full (wq / wk / wv replicated: 14 MB per 8B layer against an exchange of ~8 us) needs no gather
        bool whole = true;
        for (int i = 0; i < phases[n - 1].n_w; ++i) whole = whole && A.ph[n - 1].seg[i].row0 == 0 && A.ph[n - 1].seg[i].M == A.fin[i].M;
        A.fin_total = sp->world > 1 && !whole ? base : 0;
        A.fin_off = (unsigned) ((size_t) CH_MAXPH * CH_LL_BYTES);
        A.epoch = (unsigned *) ((uint8_t *) sp->ll[sp->rank] + sp->ll_bytes - 64);
        if (sp->ll_bytes % 64 != 0 || (size_t) A.fin_off + (size_t) base * 8 + 64 > sp->ll_bytes) { b200_set_error("decode_chain_split: hand-over area of %llu bytes is too small", (unsigned long long) sp->ll_bytes); return B200_ERR_INVALID; }
        // owner-side hand-over where a boundary has the llama shape, as pass 2b below does on one GPU, with the records validated by tags
        // (CH_IN_LL_Q8 behind SWIGLU, CH_IN_LQ8N in front of an ADD + RMS_NORM + MUL): every CTA polls 1.25-3 bytes per activation instead
        // of 8 and does not normalise and quantize the column itself.  GGML_B200_CHAIN_SPLIT_OWNER=0 keeps every boundary as f32 pairs.
        static const bool owner_off = [] { const char * e = getenv("GGML_B200_CHAIN_SPLIT_OWNER"); return e && e[0] == '0'; }();
        int prev_own = -1;
        for (int p = 1; p < n && !owner_off; ++p) {
            const b200_chain_phase & s = phases[p], & e = phases[p - 1];
            ch_phase & d = A.ph[p], & pd = A.ph[p - 1];
            if (d.K % 32 != 0 || pd.nunits % 8 != 0) continue;
            bool lo_ok = true;
            for (int i = 0; i < pd.nseg; ++i) lo_ok = lo_ok && pd.seg[i].row0 % 32 == 0;
            if (!lo_ok) continue;
            if (dkind[p] == 2 && s.prologue == B200_CHAIN_X_PLAIN) { d.in_mode = pd.out_mode = CH_IN_LL_Q8; continue; }
            if (dkind[p] != 1 || s.prologue != B200_CHAIN_X_ADD_NORM_MUL || s.out_norm || s.out_mul || pd.nseg != 1) continue;
            if (s.out_add && (ch_overlap(s.out_add->data, kb_of(d.K), e.in0) || ch_overlap(s.out_add->data, kb_of(d.K), e.in1) || ch_overlap(s.out_add->data, kb_of(d.K), e.normw))) continue;
            const bool first = in1_class[p] == 0 && add_before[p] < 0 && prev_own < 0;
            const bool next  = in1_class[p] == 1 && prev_own >= 0 && add_before[p] == prev_own + 1 && A.ph[prev_own].nunits == pd.nunits && A.ph[prev_own].seg[0].row0 == pd.seg[0].row0;
            if (!first && !next) continue;
            d.in_mode = pd.out_mode = CH_IN_LQ8N;
            const int r0 = pd.seg[0].row0;                 // the owners index these by their LOCAL row
            pd.nx_normw = (const float *) s.normw->data + r0; pd.nx_in1 = (const float *) s.in1->data + r0; pd.nx_out_add = s.out_add ? (float *) s.out_add->data + r0 : nullptr;
            if (first) { A.own_in1 = p - 1; if (A.pre_in1 == p) A.pre_in1 = -1; }
            d.in1_kept = 0;
            prev_own = p - 1;
        }
        // an ADD that expects the last sum in every CTA's registers cannot follow an ADD that only the row owners computed
        for (int p = 1; p < n; ++p)
            if (A.ph[p].in1_kept && A.ph[p].in_mode != CH_IN_LQ8N && add_before[p] >= 1 && A.ph[add_before[p]].in_mode == CH_IN_LQ8N) {
                b200_set_error("decode_chain_split: phase %d adds a sum that only the row owners hold", p); return B200_ERR_UNSUPPORTED;
            }
    } else A.nrank = 1;
    if (!ll) {
        for (int p = 0; p + 1 < n; ++p) A.ph[p].out_mode = CH_IN_BARRIER;             // one counter for the whole launch: every boundary arrives
        for (int p = 1; p < n; ++p) {
            A.ph[p].in_mode = CH_IN_BARRIER;
        }
    }
    // ---- pass 2b: owner-side hand-over where a boundary has the llama shape (see the head of this file)
    if (!ll && ch_owner_on()) {
        int plan[CH_MAXPH] = {}, prev_own = -1;       // prev_own: the phase that produced the last owner-side ADD
        for (int p = 1; p < n; ++p) {
            const b200_chain_phase & s = phases[p], & e = phases[p - 1];
            const ch_phase & d = A.ph[p], & pd = A.ph[p - 1];
            if (dkind[p] == 2 && s.prologue == B200_CHAIN_X_PLAIN && d.K % 32 == 0) { plan[p] = CH_IN_BQ8; continue; }
            if (dkind[p] != 1 || s.prologue != B200_CHAIN_X_ADD_NORM_MUL || s.out_norm || s.out_mul || d.K % 32 != 0 || pd.nseg != 1 || !ch_groups_balanced(d.K / 32, G)) continue;
            // the sum is stored while other CTAs may still read the producing phase's inputs
            if (s.out_add && (ch_overlap(s.out_add->data, kb_of(d.K), e.in0) || ch_overlap(s.out_add->data, kb_of(d.K), e.in1) || ch_overlap(s.out_add->data, kb_of(d.K), e.normw))) continue;
            const int ownn = CH_IN_BQ8N;
            if (in1_class[p] == 0 && add_before[p] < 0 && prev_own < 0) { plan[p] = ownn; prev_own = p - 1; }
            else if (in1_class[p] == 1 && prev_own >= 0 && add_before[p] == prev_own + 1 && A.ph[prev_own].nunits == pd.nunits) { plan[p] = ownn; prev_own = p - 1; }
        }
        // an ADD that expects the last sum in every CTA's registers cannot follow an ADD that only the row owners computed
        bool ok = true;
        for (int p = 1; p < n; ++p)
            if (A.ph[p].in1_kept && !ch_is_own_norm(plan[p]) && add_before[p] >= 1 && ch_is_own_norm(plan[add_before[p]])) ok = false;
        for (int p = 1; p < n; ++p) {
            if (!plan[p] || (ch_is_own_norm(plan[p]) && !ok)) continue;
            const b200_chain_phase & s = phases[p];
            ch_phase & d = A.ph[p], & pd = A.ph[p - 1];
            d.in_mode = plan[p]; pd.out_mode = plan[p];
            if (ch_is_own_norm(plan[p])) {
                pd.nx_normw = (const float *) s.normw->data; pd.nx_in1 = (const float *) s.in1->data; pd.nx_out_add = s.out_add ? (float *) s.out_add->data : nullptr;
                if (in1_class[p] == 0) { A.own_in1 = p - 1; if (A.pre_in1 == p) A.pre_in1 = -1; }
                d.in1_kept = 0;
            }
        }
    }
    // ---- pass 3: who owns which rows
    int stages_total = 0;
    for (int p = 0; p < n; ++p) {
        ch_phase & d = A.ph[p];
        d.ng8 = 0; d.ngrp = 0;
        if (sp && d.out_mode == CH_IN_LL_Q8) {
            d.ngrp = d.nunits / 8;                    // row split: every rank's rows as whole blocks (no f32 tail rows in the middle of the column)
            A.ph[p + 1].nq8 = A.ph[p + 1].K;
        } else if (d.out_mode == CH_IN_LL_Q8 || d.out_mode == CH_IN_BQ8) {
            d.ngrp = (d.nunits / 8) / G * G;          // whole 32-row blocks, the same number for every CTA; the remainder goes round-robin
            A.ph[p + 1].nq8 = d.ngrp * 8 * CH_R;
        } else if (ch_is_own_norm(d.out_mode)) d.ngrp = d.nunits / 8;      // every row: the norm needs all sums of squares from the owners
        d.ng8 = d.ngrp * 8;
        d.off = 0; d.toff = off;
        if (sp && d.out_mode == CH_IN_LL_Q8) { d.off = off; off = (off + d.ngrp) % G; }      // few groups per rank: start where the phase before stopped
        off = (off + d.nunits - d.ng8) % G;
        if (ch_is_own_norm(d.out_mode)) off = d.ngrp % G;                   // round-robin units of the next phase start with the CTAs this one left idle
        const int tpu = d.epilogue == B200_CHAIN_Y_SWIGLU ? 2 : 1;
        const int stages_max = ((d.ngrp + G - 1) / G * 8 + (d.nunits - d.ng8 + G - 1) / G) * tpu * d.nch;
        if (stages_max * CH_R > CH_PSUM || stages_max >= 4096) { b200_set_error("decode_chain: phase %d: %d stages per CTA exceed the partial-sum table", p, stages_max); return B200_ERR_UNSUPPORTED; }
        stages_total += stages_max;
        if (d.nch > 31 || s_rows_max(phases[p]) >= (1 << 18)) { b200_set_error("decode_chain: phase %d: matrix too large for the stage table", p); return B200_ERR_UNSUPPORTED; }
        if ((size_t) d.K * 8 > CH_LL_BYTES) { b200_set_error("decode_chain: phase %d: K = %d exceeds the hand-over area", p, d.K); return B200_ERR_UNSUPPORTED; }
    }
    if (stages_total > CH_MAXST) { b200_set_error("decode_chain: %d stages per CTA exceed the stage table", stages_total); return B200_ERR_UNSUPPORTED; }
    act = (act + 127) & ~127;
    const size_t fixed = (size_t) act + CH_PSUM * 4 + 32 * 4 + CH_MAXST * 4, smem_max = 227 * 1024;
    int nslot = (int) ((smem_max - fixed) / ((size_t) slot + 8));
    if (nslot > 255) nslot = 255;
    { static int cap = -1; if (cap < 0) { const char * e = getenv("GGML_B200_CHAIN_SLOTS"); cap = e ? atoi(e) : 0; } if (cap > 0 && nslot > cap) nslot = cap; }
    if (nslot < nw) { b200_set_error("decode_chain: shared memory holds only %d stages", nslot); return B200_ERR_UNSUPPORTED; }
    A.slot_bytes = slot; A.nslot = nslot; A.act_bytes = act;
    smem = (size_t) nslot * slot + fixed + (size_t) nslot * 8;
    return B200_OK;
}

// 0, or the code of the first wait that ran out in a chain launched on this context (0x100 ring slot, 0x200 arrival counter, 0x300 stage,
// 0x400-0x700 hand-over polls; + phase); synchronizes.  A nonzero code means the results of that launch are garbage.
extern "C" int b200_decode_chain_error(b200_ctx * ctx) {
    if (!ctx || !ctx->scratch[7]) return 0;
    unsigned w[8] = {};
    if (cudaSetDevice(ctx->device) != cudaSuccess || cudaStreamSynchronize(ctx->stream) != cudaSuccess ||
        cudaMemcpy(w, ctx->scratch[7], sizeof(w), cudaMemcpyDeviceToHost) != cudaSuccess) { cudaGetLastError(); return -1; }
    if (w[2] || w[4]) {                                  // re-arm for the next launch
        cudaMemset(ctx->scratch[7], 0, 256);
        b200_set_error("decode_chain: a wait timed out (code 0x%x)", w[2]);
    }
    return (int) w[2];
}

extern "C" int b200_decode_chain_supported(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases) {
    if (!ctx) return 0;
    ch_args A; size_t smem;
    return ch_build(ctx, phases, n_phases, A, smem) == B200_OK;
}

static int ch_launch(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases, const b200_chain_split * sp);
extern "C" int b200_decode_chain(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases) { return ch_launch(ctx, phases, n_phases, nullptr); }
extern "C" int b200_decode_chain_split(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases, const b200_chain_split * split) {
    if (!split) { b200_set_error("decode_chain_split: null split"); return B200_ERR_INVALID; }
    return ch_launch(ctx, phases, n_phases, split);
}
extern "C" size_t b200_decode_chain_split_bytes(int64_t last_rows) { return (((size_t) CH_MAXPH * CH_LL_BYTES + (size_t) (last_rows > 0 ? last_rows : 0) * 8 + 63) & ~(size_t) 63) + 64; }
extern "C" int b200_decode_chain_split_supported(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases, const b200_chain_split * split) {
    if (!ctx || !split) return 0;
    ch_args A; size_t smem;
    return ch_build(ctx, phases, n_phases, A, smem, split) == B200_OK;
}
static int ch_launch(b200_ctx * ctx, const b200_chain_phase * phases, int n_phases, const b200_chain_split * sp) {
    if (!ctx) { b200_set_error("decode_chain: null context"); return B200_ERR_INVALID; }
    B200_CUDA_TRY(cudaSetDevice(ctx->device));
    ch_args A;
    size_t smem;
    int rc = ch_build(ctx, phases, n_phases, A, smem, sp);
    if (rc != B200_OK) return rc;
    // counters + hand-over areas: scratch slot 7, zeroed when it is (re)allocated; the kernel re-arms the counters and advances the epoch itself
    const size_t sbytes = 256 + (size_t) CH_MAXPH * CH_LL_BYTES;
    const bool fresh = ctx->scratch[7] == nullptr || ctx->scratch_size[7] < sbytes;
    uint8_t * sc = (uint8_t *) b200_scratch(ctx, 7, sbytes);
    if (!sc) return B200_ERR_ALLOC;
    if (fresh) B200_CUDA_TRY(cudaMemsetAsync(sc, 0, sbytes, ctx->stream));
    A.sync = (unsigned *) sc; A.ll = sp ? (uint8_t *) sp->ll[sp->rank] : sc + 256;      // (row split: the areas live where the peers can reach them)
    if (!sp) A.llr[0] = A.ll;
    // a failed poll of a record is followed by a pause: 148 CTAs x 256 threads re-reading the same lines compete with the peers' incoming stores
    // (two GPUs, 8B: 666 tok/s without, 665 / 674 / 683 / 679 / 668 / 630 with 50 / 200 / 500 / 1000 / 2000 / 4000 ns)
    { static int pn = -1; if (pn < 0) { const char * e = getenv("GGML_B200_CHAIN_POLL_NS"); pn = e ? atoi(e) : 500; } A.poll_ns = pn; }
    A.trace = ctx->trace; A.trace_cap = ctx->trace_cap;
    { static int dbg = -1; if (dbg < 0) { const char * e = getenv("GGML_B200_CHAIN_DBG"); dbg = e ? atoi(e) : 0; } A.dbg = dbg; }
    const int nw = ch_nw();
    { static int said = 0; if (said < 6 && getenv("GGML_B200_DEBUG")) { ++said; char m[128]; int o = 0; for (int p = 0; p < n_phases; ++p) o += snprintf(m + o, sizeof(m) - o, " %d>%d", A.ph[p].in_mode, A.ph[p].out_mode);
        fprintf(stderr, "ggml-b200: decode chain of %d phases, hand-over in>out per phase:%s, tail phase %d, %d ring slots\n", n_phases, m, A.tail_phase, A.nslot); } }
    static bool attr_done[64] = {};
    if (!attr_done[ctx->device & 63]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<16, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<24, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<16, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<24, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
        attr_done[ctx->device & 63] = true;
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned) ctx->sm_count);
    cfg.blockDim = dim3((sp ? 16 : nw) * 32);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    const bool tr = A.trace != nullptr || A.dbg != 0;
    if (sp) {
        static bool sp_attr[64] = {};
        if (!sp_attr[ctx->device & 63]) { B200_CUDA_TRY(cudaFuncSetAttribute(k_decode_chain<16, false, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024)); sp_attr[ctx->device & 63] = true; }
        B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<16, false, true>, A));
    } else
    if (nw == 24) { if (tr) B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<24, true>, A)); else B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<24, false>, A)); }
    else          { if (tr) B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<16, true>, A)); else B200_LAUNCH(ctx, B200_K_GEMV, cudaLaunchKernelEx(&cfg, k_decode_chain<16, false>, A)); }
    for (int p = 0; p < n_phases; ++p) {
        const b200_chain_phase & s = phases[p];
        for (int i = 0; i < s.n_w; ++i) b200_note_write(ctx, s.dst[i]->data, b200_tensor_span(s.dst[i]));
        b200_tensor * outs[4] = { s.out_add, s.out_norm, s.out_mul, s.glu };
        for (b200_tensor * t : outs) if (t) b200_note_write(ctx, t->data, b200_tensor_span(t));
    }
    return B200_OK;
}
