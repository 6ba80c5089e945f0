// b200_fattn_q.cu -- GGML_OP_FLASH_ATTN_EXT over a QUANTIZED KV cache: K and V as q8_0 / q4_0 blocks (llama.cpp -ctk / -ctv), SURVEY 8(f) N4.
//
// Semantics: ggml_compute_forward_flash_attn_ext_f16_one_chunk, reference ggml/src/ggml-cpu/ops.cpp:8047-8272.  With a quantized K the
// reference converts the query row to K's vec_dot_type -- q8_0 for both q8_0 and q4_0 (ops.cpp:8090-8094, 8150-8156) -- and scores with the
// integer block dot product (ggml_vec_dot_q8_0_q8_0 / q4_0_q8_0); a quantized V row is expanded to f32 (v_to_float, ops.cpp:8210-8216) and
// accumulated in f32.  The decode kernel below does exactly that: the query rows are quantized to q8_0 once per CTA, K blocks are dotted with
// dp4a, V blocks are expanded in registers.  ggml-cuda instead converts the WHOLE cache to f16 first (fattn-common.cuh:825-878): 1.06 bytes
// read + 2 written + 2 read per element, against 1.06 here.
//
//  * k_fattn_kvq<TK, TV, D, R> -- query batches of <= 8 tokens.  One CTA per (kv head, group of R query rows of that kv head's GQA group,
//    KV split); every warp stages its tiles of 32 keys -- the K and V rows as they lie in the cache -- into a two-stage shared-memory ring with
//    cp.async, one tile ahead; lane = key for the scores (34 words of its K row for q8_0 / head 128), online softmax per row across the
//    warp, probabilities through shared memory, lane = element group for P.V, four keys per round; 4 warps take tiles round-robin and merge
//    through shared memory, KV splits merge in k_fattn_combine (b200_fattn_dec.cu).  Bound: HBM; algorithmic bytes per kv head and layer:
//    n_kv * (row_size(TK, D) + row_size(TV, D)) -- 2 * 136 B per key at head 128 / q8_0, each read once.
//  * larger batches (prefill): k_dequant_kv expands this call's K / V views to f16 rows in scratch memory and the f16 kernels run on those
//    (the tensor-core kernel is compute bound there, the conversion is ~1 % of its time).
#include "b200_common.cuh"
#include "b200_fattn.cuh"

#define FQ_WARPS 4

struct fq_extra { int nsplit, tiles_per_split, ntiles, gqa, rows; float * partial; };

template <int T> struct fq_type;
template <> struct fq_type<B200_TYPE_Q8_0> { static constexpr int BS = 34; };
template <> struct fq_type<B200_TYPE_Q4_0> { static constexpr int BS = 18; };

static __device__ __forceinline__ float fq_half_at(const uint32_t * W, int byte) {     // fp16 at a 2-byte aligned offset of a word array
    const uint32_t w = W[byte >> 2];
    return __half2float(__ushort_as_half((unsigned short) ((byte & 2) ? (w >> 16) : (w & 0xffffu))));
}
static __device__ __forceinline__ uint32_t fq_word_at(const uint32_t * W, int byte) {  // 4 bytes at a 2-byte aligned offset
    return (byte & 2) ? __funnelshift_r(W[byte >> 2], W[(byte >> 2) + 1], 16) : W[byte >> 2];
}

static __device__ __forceinline__ uint32_t fq_lds32(uint32_t a) { uint32_t v; asm volatile("ld.shared.u32 %0, [%1];" : "=r"(v) : "r"(a)); return v; }
static __device__ __forceinline__ unsigned short fq_lds16(uint32_t a) { unsigned short v; asm volatile("ld.shared.u16 %0, [%1];" : "=h"(v) : "r"(a)); return v; }
// 32 key rows of NWORDS 32-bit words each (row r of the tile at g + r * nb1) -> shared memory, rows NWORDS * 4 bytes apart, asynchronously:
// consecutive lanes copy consecutive words of a row.  Rows beyond the cache's end are not touched
template <int NWORDS>
static __device__ __forceinline__ void fq_stage(uint32_t dst, const char * g, size_t nb1, int rows, int lane) {
    for (int r = 0; r < rows; ++r) {
        const char * gr = g + (size_t) r * nb1;
#pragma unroll
        for (int j = 0; j < (NWORDS + 31) / 32; ++j) {
            const int w = lane + 32 * j;
            if (w < NWORDS) asm volatile("cp.async.ca.shared.global [%0], [%1], 4;" :: "r"(dst + (uint32_t) (r * NWORDS + w) * 4u), "l"(gr + 4 * w) : "memory");
        }
    }
}

// element e (0 .. D/32-1 .. ) of lane: which element of the V row a lane accumulates.  q8_0: lane owns the 16-bit units 2 lane (+ 64 for
// D = 128); q4_0: one 16-bit unit of nibble bytes = elements (j, j + 1, j + 16, j + 17) of a block
template <int TV, int D> struct fq_vmap;
template <int D> struct fq_vmap<B200_TYPE_Q8_0, D> {
    static constexpr int NE = D / 32;   // accumulators per lane: 2 per 64 elements
    static __device__ __forceinline__ bool active(int) { return true; }
    static __device__ __forceinline__ int elem(int lane, int e) { return (e >> 1) * 64 + 2 * lane + (e & 1); }
    static __device__ __forceinline__ void load(uint32_t vrow, int lane, float (&v)[NE]) {      // vrow: the row in shared memory
#pragma unroll
        for (int h = 0; h < D / 64; ++h) {
            const int b = 2 * h + (lane >> 4);
            const unsigned short qs = fq_lds16(vrow + b * 34 + 2 + 2 * (lane & 15));
            const float d = __half2float(__ushort_as_half(fq_lds16(vrow + b * 34)));
            v[2 * h]     = d * (float) (int) (signed char) (qs & 0xff);
            v[2 * h + 1] = d * (float) (int) (signed char) (qs >> 8);
        }
    }
};
template <int D> struct fq_vmap<B200_TYPE_Q4_0, D> {
    static constexpr int NE = 4;
    static __device__ __forceinline__ bool active(int lane) { return lane < D / 4; }
    static __device__ __forceinline__ int elem(int lane, int e) { return (lane >> 3) * 32 + 2 * (lane & 7) + (e & 1) + (e >> 1) * 16; }
    static __device__ __forceinline__ void load(uint32_t vrow, int lane, float (&v)[NE]) {
        if (!active(lane)) { v[0] = v[1] = v[2] = v[3] = 0.0f; return; }
        const int b = lane >> 3;
        const unsigned short qs = fq_lds16(vrow + b * 18 + 2 + 2 * (lane & 7));
        const float d = __half2float(__ushort_as_half(fq_lds16(vrow + b * 18)));
        v[0] = d * (float) ((int) (qs & 0xf) - 8);        v[1] = d * (float) ((int) ((qs >> 8) & 0xf) - 8);
        v[2] = d * (float) ((int) ((qs >> 4) & 0xf) - 8); v[3] = d * (float) ((int) (qs >> 12) - 8);
    }
};

template <int TK, int TV, int D, int R>
__global__ void __launch_bounds__(FQ_WARPS * 32) k_fattn_kvq(const fattn_args p, const fq_extra x) {
    constexpr int NB = D / 32, KBS = fq_type<TK>::BS, NW = NB * KBS / 4;       // blocks per row, K block bytes, words per K row
    constexpr int VBS = fq_type<TV>::BS, NWV = NB * VBS / 4;                   // V block bytes, words per V row
    constexpr int NE = fq_vmap<TV, D>::NE;
    // per warp: two stages of (32 K rows | 32 V rows) as they lie in the cache, filled by cp.async one tile ahead of the arithmetic.
    // (One stage and four CTAs per SM instead of two: 249 against 195 us at 131 k keys -- the kernel is bound by its ~1800 instructions per
    // tile, two thirds of them the P.V loop, not by latency.)
    extern __shared__ __align__(16) uint8_t fq_ring[];
    constexpr uint32_t STAGE = 32u * (NW + NWV) * 4u;
    __shared__ __align__(16) int   sqw[R][D / 4];            // query rows as q8_0 quants
    __shared__ float sdq[R][NB];                             // their block scales (fp16-rounded, as the reference stores them)
    __shared__ int   ssq[R][NB];                             // sum of a block's quants (q4_0: the -8 offset)
    __shared__ __align__(16) float sp[FQ_WARPS][32][R];      // probabilities of a warp's current tile
    __shared__ float sacc[FQ_WARPS][R][D];
    __shared__ float sms[FQ_WARPS][R][2];

    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
    const int split = blockIdx.x % x.nsplit, rg = blockIdx.x / x.nsplit;
    const int64_t hk = blockIdx.y, iq3 = blockIdx.z;
    const int nr = min(R, x.rows - rg * R);                  // live rows of this CTA

    // ---- the query rows -> q8_0 (quantize_row_q8_0 as the x86 CPU backend runs it, arch/x86/quants.c:290-385): warp w takes rows w, w + FQ_WARPS, ...
    constexpr int EPL = D / 32, LPB = 32 / EPL;              // elements per lane, lanes per 32-element block
    for (int r = warp; r < R; r += FQ_WARPS) {
        float v[EPL];
        if (r < nr) {
            const int rho = rg * R + r; const int64_t n = rho / x.gqa, h = hk * x.gqa + rho % x.gqa;
            const float * qrow = (const float *) (p.q + n * p.q_nb1 + h * p.q_nb2 + iq3 * p.q_nb3);
#pragma unroll
            for (int e = 0; e < EPL; ++e) v[e] = qrow[lane * EPL + e];
        } else {
#pragma unroll
            for (int e = 0; e < EPL; ++e) v[e] = 0.0f;
        }
        float amax = 0.0f;
#pragma unroll
        for (int e = 0; e < EPL; ++e) amax = fmaxf(amax, fabsf(v[e]));
#pragma unroll
        for (int o = 1; o < LPB; o <<= 1) amax = fmaxf(amax, __shfl_xor_sync(0xffffffffu, amax, o));
        const float d = __fdiv_rn(amax, 127.0f), id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;      // the x86 body: 127 / amax, nearest even
        int q[EPL], sum = 0; unsigned packed = 0;
#pragma unroll
        for (int e = 0; e < EPL; ++e) { q[e] = __float2int_rn(__fmul_rn(v[e], id)); sum += q[e]; packed |= ((unsigned) q[e] & 0xffu) << (8 * (e & 3)); }
#pragma unroll
        for (int o = 1; o < LPB; o <<= 1) sum += __shfl_xor_sync(0xffffffffu, sum, o);
        if (EPL == 4) sqw[r][lane] = (int) packed;
        else {                                               // EPL == 2: two lanes share a word
            const unsigned other = __shfl_xor_sync(0xffffffffu, packed, 1);
            if (!(lane & 1)) sqw[r][lane >> 1] = (int) (packed | (other << 16));
        }
        if ((lane % LPB) == 0) { sdq[r][lane / LPB] = __half2float(__float2half_rn(d)); ssq[r][lane / LPB] = sum; }
    }
    __syncthreads();

    const char * kbase = p.k + hk * p.k_nb2 + (iq3 / p.rk3) * p.k_nb3;
    const char * vbase = p.v + hk * p.v_nb2 + (iq3 / p.rv3) * p.v_nb3;
    float slope[R]; int64_t moff[R];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const int rho = rg * R + min(r, nr - 1); const int64_t n = rho / x.gqa, h = hk * x.gqa + rho % x.gqa;
        slope[r] = fattn_slope(p, (uint32_t) h);
        moff[r] = n * p.m_nb1 + (h % p.m_ne2) * p.m_nb2 + (iq3 % p.m_ne3) * p.m_nb3;
    }

    float M[R], S[R], acc[R][NE];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        M[r] = -INFINITY; S[r] = 0.0f;
#pragma unroll
        for (int e = 0; e < NE; ++e) acc[r][e] = 0.0f;
    }

    const int t0 = split * x.tiles_per_split, t1 = min(x.ntiles, t0 + x.tiles_per_split);
    const uint32_t ring = (uint32_t) __cvta_generic_to_shared(fq_ring) + (uint32_t) warp * 2u * STAGE;
    auto stage_tile = [&](int t, int st) {                   // whole warp; one commit group per tile (possibly empty)
        if (t < t1) {
            const int rows = (int) min((int64_t) 32, p.KV - (int64_t) t * 32);
            fq_stage<NW>(ring + st * STAGE, kbase + (size_t) t * 32 * p.k_nb1, p.k_nb1, rows, lane);
            fq_stage<NWV>(ring + st * STAGE + 32u * NW * 4u, vbase + (size_t) t * 32 * p.v_nb1, p.v_nb1, rows, lane);
        }
        asm volatile("cp.async.commit_group;" ::: "memory");
    };
    stage_tile(t0 + warp, 0);
    int st = 0;
    for (int t = t0 + warp; t < t1; t += FQ_WARPS, st ^= 1) {
        stage_tile(t + FQ_WARPS, st ^ 1);                    // the next tile's bytes travel while this one is multiplied
        asm volatile("cp.async.wait_group 1;" ::: "memory");
        __syncwarp();
        const uint32_t kst = ring + st * STAGE, vst = kst + 32u * NW * 4u;
        const int64_t ic = (int64_t) t * 32 + lane;
        const bool valid = ic < p.KV;
        float mv[R]; bool any = false;
#pragma unroll
        for (int r = 0; r < R; ++r) {
            mv[r] = -INFINITY;
            if (valid && r < nr) mv[r] = p.mask ? slope[r] * __half2float(*(const __half *) (p.mask + moff[r] + ic * 2)) : 0.0f;
            any = any || mv[r] != -INFINITY;
        }
        float s[R];
#pragma unroll
        for (int r = 0; r < R; ++r) s[r] = 0.0f;
        if (any) {
            uint32_t W[NW + 1];
#pragma unroll
            for (int i = 0; i < NW; ++i) W[i] = fq_lds32(kst + (uint32_t) (lane * NW + i) * 4u);
            W[NW] = 0;
#pragma unroll
            for (int b = 0; b < NB; ++b) {
                const float dk = fq_half_at(W, b * KBS);
                int isum[R];
#pragma unroll
                for (int r = 0; r < R; ++r) isum[r] = 0;
                if (TK == B200_TYPE_Q8_0) {
#pragma unroll
                    for (int i4 = 0; i4 < 2; ++i4) {
                        int kw[4];
#pragma unroll
                        for (int i = 0; i < 4; ++i) kw[i] = (int) fq_word_at(W, b * KBS + 2 + 4 * (4 * i4 + i));
#pragma unroll
                        for (int r = 0; r < R; ++r) {
                            const int4 qq = *(const int4 *) &sqw[r][b * 8 + 4 * i4];
                            isum[r] = __dp4a(kw[0], qq.x, isum[r]); isum[r] = __dp4a(kw[1], qq.y, isum[r]);
                            isum[r] = __dp4a(kw[2], qq.z, isum[r]); isum[r] = __dp4a(kw[3], qq.w, isum[r]);
                        }
                    }
                } else {                                     // q4_0: byte j = element j (low nibble), element j + 16 (high), value - 8 (ggml-quants.c:262-280)
                    int lo[4], hi[4];
#pragma unroll
                    for (int i = 0; i < 4; ++i) { const uint32_t w = fq_word_at(W, b * KBS + 2 + 4 * i); lo[i] = (int) (w & 0x0F0F0F0Fu); hi[i] = (int) ((w >> 4) & 0x0F0F0F0Fu); }
#pragma unroll
                    for (int r = 0; r < R; ++r) {
                        const int4 q0 = *(const int4 *) &sqw[r][b * 8], q1 = *(const int4 *) &sqw[r][b * 8 + 4];
                        int a = __dp4a(lo[0], q0.x, 0); a = __dp4a(lo[1], q0.y, a); a = __dp4a(lo[2], q0.z, a); a = __dp4a(lo[3], q0.w, a);
                        a = __dp4a(hi[0], q1.x, a); a = __dp4a(hi[1], q1.y, a); a = __dp4a(hi[2], q1.z, a); a = __dp4a(hi[3], q1.w, a);
                        isum[r] = a - 8 * ssq[r][b];
                    }
                }
#pragma unroll
                for (int r = 0; r < R; ++r) s[r] = fmaf(dk * sdq[r][b], (float) isum[r], s[r]);
            }
        }
        // scale / softcap / mask, online softmax per row across the 32 keys of the tile (ops.cpp:8158-8208)
        bool tile_live = false;
        float pr[R];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            float sv = -INFINITY;
            if (mv[r] != -INFINITY) {
                sv = s[r] * p.scale;
                if (p.softcap != 0.0f) sv = p.softcap * tanhf(sv);
                sv += mv[r];
            }
            const float mt = warp_max(sv);
            pr[r] = 0.0f;
            if (mt != -INFINITY) {                           // (warp-uniform)
                const float Mn = fmaxf(M[r], mt), ms = expf(M[r] - Mn);      // M == -inf -> 0
                pr[r] = sv == -INFINITY ? 0.0f : expf(sv - Mn);
                S[r] = S[r] * ms + warp_sum(pr[r]);
                M[r] = Mn;
#pragma unroll
                for (int e = 0; e < NE; ++e) acc[r][e] *= ms;
                tile_live = true;
            }
        }
        bool nz = false;
#pragma unroll
        for (int r = 0; r < R; ++r) { sp[warp][lane][r] = pr[r]; nz = nz || pr[r] != 0.0f; }
        const unsigned live = tile_live ? __ballot_sync(0xffffffffu, nz) : 0u;       // (a tile masked for every row: nothing to add; warp-uniform)
        __syncwarp();
        if (live) {
            // every key of the tile that exists (its row was staged: finite values), four at a time so that their shared-memory loads are in
            // flight together; a masked key carries p = 0.  (One key per round, picked from the ballot, left the loop waiting on one load
            // after the other: 483 GB/s at 131 k keys.)
            const int rows = (int) min((int64_t) 32, p.KV - (int64_t) t * 32);
            for (int j0 = 0; j0 < rows; j0 += 4) {
                float v[4][NE], pj[4][R];
#pragma unroll
                for (int u = 0; u < 4; ++u) {
                    const int j = min(j0 + u, rows - 1);             // (a repeated last key gets p = 0 below)
                    fq_vmap<TV, D>::load(vst + (uint32_t) j * NWV * 4u, lane, v[u]);
#pragma unroll
                    for (int r4 = 0; r4 < R; r4 += 4) {
                        const float4 f = *(const float4 *) &sp[warp][j][r4];
                        const bool on = j0 + u < rows;
                        pj[u][r4] = on ? f.x : 0.0f; pj[u][r4 + 1] = on ? f.y : 0.0f; pj[u][r4 + 2] = on ? f.z : 0.0f; pj[u][r4 + 3] = on ? f.w : 0.0f;
                    }
                }
#pragma unroll
                for (int u = 0; u < 4; ++u)
#pragma unroll
                    for (int r = 0; r < R; ++r)
#pragma unroll
                        for (int e = 0; e < NE; ++e) acc[r][e] = fmaf(pj[u][r], v[u][e], acc[r][e]);
            }
        }
        __syncwarp();                                        // every lane is done with this stage before the next tile lands in it
    }
    asm volatile("cp.async.wait_group 0;" ::: "memory");

    // ---- merge the warps
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (fq_vmap<TV, D>::active(lane)) {
#pragma unroll
            for (int e = 0; e < NE; ++e) sacc[warp][r][fq_vmap<TV, D>::elem(lane, e)] = acc[r][e];
        }
        if (lane == 0) { sms[warp][r][0] = M[r]; sms[warp][r][1] = S[r]; }
    }
    __syncthreads();
    for (int r = 0; r < nr; ++r) {
        float Mall = -INFINITY;
#pragma unroll
        for (int w = 0; w < FQ_WARPS; ++w) Mall = fmaxf(Mall, sms[w][r][0]);
        float Sall = 0.0f, wgt[FQ_WARPS];
#pragma unroll
        for (int w = 0; w < FQ_WARPS; ++w) { wgt[w] = sms[w][r][0] == -INFINITY ? 0.0f : expf(sms[w][r][0] - Mall); Sall += sms[w][r][1] * wgt[w]; }
        const int rho = rg * R + r; const int64_t n = rho / x.gqa, h = hk * x.gqa + rho % x.gqa;
        if (x.nsplit > 1) {
            float * part = x.partial + ((((int64_t) iq3 * p.H + h) * p.N + n) * x.nsplit + split) * (D + 2);
            for (int d = threadIdx.x; d < D; d += blockDim.x) {
                float o = 0.0f;
#pragma unroll
                for (int w = 0; w < FQ_WARPS; ++w) o += sacc[w][r][d] * wgt[w];
                part[d] = o;
            }
            if (threadIdx.x == 0) { part[D] = Mall; part[D + 1] = Sall; }
            continue;
        }
        float sink_scale = 1.0f;
        if (p.sinks) {                                       // ops.cpp:8241-8255
            const float sk = p.sinks[h];
            if (sk > Mall) { sink_scale = expf(Mall - sk); Sall = Sall * sink_scale + 1.0f; }
            else           { Sall += expf(sk - Mall); }
        }
        const float Sinv = Sall == 0.0f ? 0.0f : 1.0f / Sall;
        float * drow = (float *) ((char *) p.dst + h * p.d_nb1 + n * p.d_nb2 + iq3 * p.d_nb3);
        for (int d = threadIdx.x; d < D; d += blockDim.x) {
            float o = 0.0f;
#pragma unroll
            for (int w = 0; w < FQ_WARPS; ++w) o += sacc[w][r][d] * wgt[w];
            drow[d] = o * sink_scale * Sinv;
        }
    }
}

// ---- prefill: this call's K or V view -> f16 rows [ne3][ne2][ne1][D] (dequantize_row_q8_0 / q4_0, ggml-quants.c:262-280, 342-354); thread = block
template <int T>
__global__ void __launch_bounds__(256) k_dequant_kv(const char * src, size_t nb1, size_t nb2, size_t nb3, int64_t ne1, int64_t ne2, int nblk, int64_t total, __half * dst) {
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int b = (int) (i % nblk); int64_t r = i / nblk;
    const int64_t i1 = r % ne1; r /= ne1;
    const int64_t i2 = r % ne2, i3 = r / ne2;
    const unsigned short * s = (const unsigned short *) (src + i1 * nb1 + i2 * nb2 + i3 * nb3 + (size_t) b * fq_type<T>::BS);
    const __half d = __ushort_as_half(__ldg(s));
    const float df = __half2float(d);
    __align__(16) __half out[32];
    if (T == B200_TYPE_Q8_0) {
#pragma unroll
        for (int j = 0; j < 16; ++j) { const unsigned short w = __ldg(s + 1 + j); out[2 * j] = __float2half_rn(df * (float) (int) (signed char) (w & 0xff)); out[2 * j + 1] = __float2half_rn(df * (float) (int) (signed char) (w >> 8)); }
    } else {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            const unsigned short w = __ldg(s + 1 + j);
            out[2 * j]      = __float2half_rn(df * (float) ((int) (w & 0xf) - 8));        out[2 * j + 1]      = __float2half_rn(df * (float) ((int) ((w >> 8) & 0xf) - 8));
            out[2 * j + 16] = __float2half_rn(df * (float) ((int) ((w >> 4) & 0xf) - 8)); out[2 * j + 17] = __float2half_rn(df * (float) ((int) (w >> 12) - 8));
        }
    }
    uint4 * o = (uint4 *) (dst + i * 32);
#pragma unroll
    for (int j = 0; j < 4; ++j) o[j] = ((const uint4 *) out)[j];
}

static bool fq_quant(int t) { return t == B200_TYPE_Q8_0 || t == B200_TYPE_Q4_0; }

// K / V types this file serves (b200_flash_attn_ext and the plugin's supports_op ask)
bool b200_fattn_kvq_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v) {
    if (!fq_quant(k->type) || !fq_quant(v->type)) return false;
    const int64_t D = k->ne[0];
    if ((D != 64 && D != 128) || v->ne[0] != D || q->ne[0] != D) return false;
    if (k->nb[0] != b200_type_size(k->type) || v->nb[0] != b200_type_size(v->type)) return false;
    if ((k->nb[1] & 3) || (k->nb[2] & 3) || (k->nb[3] & 3) || ((uintptr_t) k->data & 3)) return false;      // the lane reads its K row as 32-bit words
    if ((v->nb[1] & 3) || (v->nb[2] & 3) || (v->nb[3] & 3) || ((uintptr_t) v->data & 3)) return false;      // ... and both are staged in 4-byte copies
    if (k->nb[1] < b200_row_size(k->type, D) || v->nb[1] < b200_row_size(v->type, D)) return false;
    if (q->ne[2] % k->ne[2] != 0 || k->ne[2] != v->ne[2] || k->ne[3] != v->ne[3]) return false;
    return true;
}

template <int TK, int TV, int D>
static int fq_launch(b200_ctx * ctx, const fattn_args & p, const fq_extra & x, dim3 grid) {
    constexpr size_t ring = (size_t) FQ_WARPS * 2 * 32 * ((D / 32) * (fq_type<TK>::BS + fq_type<TV>::BS));      // two stages of K and V rows per warp
    static bool attr_done[64] = {};
    if (!attr_done[ctx->device & 63]) {
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_kvq<TK, TV, D, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ring));
        B200_CUDA_TRY(cudaFuncSetAttribute(k_fattn_kvq<TK, TV, D, 8>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) ring));
        attr_done[ctx->device & 63] = true;
    }
    if (x.rows <= 4) B200_LAUNCH(ctx, B200_K_FATTN, k_fattn_kvq<TK, TV, D, 4><<<grid, FQ_WARPS * 32, ring, ctx->stream>>>(p, x));
    else             B200_LAUNCH(ctx, B200_K_FATTN, k_fattn_kvq<TK, TV, D, 8><<<grid, FQ_WARPS * 32, ring, ctx->stream>>>(p, x));
    return B200_OK;
}
template <int TK, int TV>
static int fq_launch_d(b200_ctx * ctx, const fattn_args & p, const fq_extra & x, dim3 grid) {
    return p.DK == 64 ? fq_launch<TK, TV, 64>(ctx, p, x, grid) : fq_launch<TK, TV, 128>(ctx, p, x, grid);
}

int b200_launch_fattn_combine(b200_ctx * ctx, const fattn_args & p, int64_t S, int nsplit, float * partial);      // b200_fattn_dec.cu

static int fq_dequant(b200_ctx * ctx, const b200_tensor * t, __half * dst, b200_tensor * out) {
    const int nblk = (int) (t->ne[0] / 32);
    const int64_t total = (int64_t) nblk * t->ne[1] * t->ne[2] * t->ne[3];
    const unsigned grid = (unsigned) ((total + 255) / 256);
    if (t->type == B200_TYPE_Q8_0) B200_LAUNCH(ctx, B200_K_CONVERT, k_dequant_kv<B200_TYPE_Q8_0><<<grid, 256, 0, ctx->stream>>>((const char *) t->data, t->nb[1], t->nb[2], t->nb[3], t->ne[1], t->ne[2], nblk, total, dst));
    else                           B200_LAUNCH(ctx, B200_K_CONVERT, k_dequant_kv<B200_TYPE_Q4_0><<<grid, 256, 0, ctx->stream>>>((const char *) t->data, t->nb[1], t->nb[2], t->nb[3], t->ne[1], t->ne[2], nblk, total, dst));
    *out = *t;
    out->data = dst; out->type = B200_TYPE_F16; out->flags = 0;
    out->nb[0] = 2; out->nb[1] = (size_t) t->ne[0] * 2; out->nb[2] = out->nb[1] * t->ne[1]; out->nb[3] = out->nb[2] * t->ne[2];
    return B200_OK;
}

int b200_launch_fattn_kvq(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                          const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst, float scale, float max_bias, float softcap) {
    if (!b200_fattn_kvq_supported(q, k, v)) { b200_set_error("flash_attn_ext: quantized K/V outside q8_0 / q4_0, head 64 / 128"); return B200_ERR_UNSUPPORTED; }
    fattn_args p;
    fattn_fill_args(p, q, k, v, mask, sinks, dst, scale, max_bias, softcap);
    const int64_t S = q->ne[3];
    if (p.N == 0 || p.H == 0 || S == 0) return B200_OK;
    static const int force_conv = getenv("GGML_B200_FATTN_KVQ_CONVERT") ? atoi(getenv("GGML_B200_FATTN_KVQ_CONVERT")) : 0;   // (measurement only: 1 = always convert)
    if (p.N > 8 || force_conv) {
        // prefill: expand K and V to f16 rows in scratch, then the f16 kernels
        const size_t kel = (size_t) k->ne[0] * k->ne[1] * k->ne[2] * k->ne[3], vel = (size_t) v->ne[0] * v->ne[1] * v->ne[2] * v->ne[3];
        const size_t koff = 0, voff = (kel * 2 + 255) & ~(size_t) 255;
        char * sc = (char *) b200_scratch(ctx, 4, voff + vel * 2);
        if (!sc) return B200_ERR_ALLOC;
        b200_tensor kf, vf;
        int rc = fq_dequant(ctx, k, (__half *) (sc + koff), &kf); if (rc != B200_OK) return rc;
        rc = fq_dequant(ctx, v, (__half *) (sc + voff), &vf);     if (rc != B200_OK) return rc;
        return b200_flash_attn_ext(ctx, q, &kf, &vf, mask, sinks, dst, scale, max_bias, softcap);
    }
    if (k->ne[2] > 65535 || S > 65535) { b200_set_error("fattn: kv heads or S exceed grid limits"); return B200_ERR_UNSUPPORTED; }
    fq_extra x;
    x.gqa = (int) (q->ne[2] / k->ne[2]);
    x.rows = (int) (p.N * x.gqa);
    const int R = x.rows <= 4 ? 4 : 8, ngroups = (x.rows + R - 1) / R;
    x.ntiles = (int) ((p.KV + 31) / 32);
    // KV splits until the grid covers the SMs twice (two CTAs per SM are resident), at least 8 tiles (256 keys, two per warp) per split
    const int64_t base = (int64_t) ngroups * k->ne[2] * S;
    int nsplit = (int) ((2 * ctx->sm_count + base - 1) / base);
    if (nsplit > x.ntiles / 8) nsplit = x.ntiles / 8;
    if (nsplit < 1) nsplit = 1;
    x.tiles_per_split = (x.ntiles + nsplit - 1) / nsplit;
    x.nsplit = (x.ntiles + x.tiles_per_split - 1) / x.tiles_per_split;
    x.partial = nullptr;
    if (x.nsplit > 1) {
        x.partial = (float *) b200_scratch(ctx, 1, (size_t) p.N * p.H * S * x.nsplit * (p.DV + 2) * sizeof(float));
        if (!x.partial) return B200_ERR_ALLOC;
    }
    const dim3 grid((unsigned) (ngroups * x.nsplit), (unsigned) k->ne[2], (unsigned) S);
    int rc;
    if (k->type == B200_TYPE_Q8_0) rc = v->type == B200_TYPE_Q8_0 ? fq_launch_d<B200_TYPE_Q8_0, B200_TYPE_Q8_0>(ctx, p, x, grid) : fq_launch_d<B200_TYPE_Q8_0, B200_TYPE_Q4_0>(ctx, p, x, grid);
    else                           rc = v->type == B200_TYPE_Q8_0 ? fq_launch_d<B200_TYPE_Q4_0, B200_TYPE_Q8_0>(ctx, p, x, grid) : fq_launch_d<B200_TYPE_Q4_0, B200_TYPE_Q4_0>(ctx, p, x, grid);
    if (rc != B200_OK) return rc;
    if (x.nsplit > 1) return b200_launch_fattn_combine(ctx, p, S, x.nsplit, x.partial);
    return B200_OK;
}
