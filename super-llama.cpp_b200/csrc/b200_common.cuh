// b200_common.cuh -- shared device/host definitions for the ggml-b200 kernels.
// GGUF block layouts restated from the reference (ggml/src/ggml-common.h:170-344); byte-exact
// input format, the only thing we share with ggml-cuda is the file format.
#pragma once
#include <cuda_runtime.h>
#include <cuda_fp16.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>
#include "../../include/ggml-b200.h"

#define B200_WARP 32
#define QK 32          // elements per q4_0 / q8_0 / q8_1 block
#define QKK 256        // elements per K-quant super-block

struct __align__(2) blk_q4_0 { __half d; uint8_t qs[16]; };                                 // 18 B
struct __align__(2) blk_q8_0 { __half d; int8_t  qs[32]; };                                 // 34 B
struct __align__(4) blk_q8_1 { __half d; __half s; int8_t qs[32]; };                        // 36 B
struct __align__(4) blk_q4_K { __half d; __half dmin; uint8_t scales[12]; uint8_t qs[128]; };              // 144 B
struct __align__(4) blk_q5_K { __half d; __half dmin; uint8_t scales[12]; uint8_t qh[32]; uint8_t qs[128]; }; // 176 B
struct __align__(2) blk_q6_K { uint8_t ql[128]; uint8_t qh[64]; int8_t scales[16]; __half d; };            // 210 B
static_assert(sizeof(blk_q4_0) == 18 && sizeof(blk_q8_0) == 34 && sizeof(blk_q8_1) == 36, "block size");
static_assert(sizeof(blk_q4_K) == 144 && sizeof(blk_q5_K) == 176 && sizeof(blk_q6_K) == 210, "block size");

// bytes per block / elements per block for a b200_type (0 if not a block type we know)
static inline __host__ __device__ int b200_blck_size(int type) {
    switch (type) {
        case B200_TYPE_F32: case B200_TYPE_F16: case B200_TYPE_BF16: case B200_TYPE_I32: return 1;
        case B200_TYPE_Q4_0: case B200_TYPE_Q8_0: case B200_TYPE_Q8_1: return 32;
        case B200_TYPE_Q4_K: case B200_TYPE_Q5_K: case B200_TYPE_Q6_K: return 256;
    }
    return 0;
}
static inline __host__ __device__ size_t b200_type_size(int type) {
    switch (type) {
        case B200_TYPE_F32: case B200_TYPE_I32: return 4;
        case B200_TYPE_F16: case B200_TYPE_BF16: return 2;
        case B200_TYPE_Q4_0: return 18; case B200_TYPE_Q8_0: return 34; case B200_TYPE_Q8_1: return 36;
        case B200_TYPE_Q4_K: return 144; case B200_TYPE_Q5_K: return 176; case B200_TYPE_Q6_K: return 210;
    }
    return 0;
}
static inline size_t b200_row_size(int type, int64_t ne0) {
    return (size_t)(ne0 / b200_blck_size(type)) * b200_type_size(type);
}

// ---- host-side context ------------------------------------------------------------------
// kernel kinds for the per-kernel event profile (b200_ctx_profile_*), also the index of launch counters
enum b200_kernel_kind {
    B200_K_QUANTIZE = 0, B200_K_GEMV = 1, B200_K_GEMV_ID = 2, B200_K_GEMM = 3, B200_K_FATTN = 4, B200_K_FATTN_COMBINE = 5,
    B200_K_MATMUL_FLOAT = 6, B200_K_GLUE = 7, B200_K_MMID_SORT = 8, B200_K_CONVERT = 9, B200_K_COUNT = 10,
};
struct b200_prof_rec { int kind; cudaEvent_t e0, e1; };
struct b200_ctx {
    int          device;
    cudaStream_t stream;
    int          sm_count;
    // scratch arena for activation quantization / split-KV partials / sorted ids (grows on demand)
    void *       scratch[8];
    size_t       scratch_size[8];
    uint64_t     scratch_gen;       // bumped whenever a scratch block moves (recorded CUDA graphs hold the old address)
    uint64_t     launches;
    uint64_t     launches_kind[B200_K_COUNT];
    // optional per-kernel timing with CUDA events on the launching stream (bench.py's roofline leg)
    int            profile;
    b200_prof_rec * prof; int prof_n, prof_cap;
    int            prof_open;        // index of the record whose e1 is pending, or -1
    // CUDA graph capture state (b200_graph_*)
    int          capturing;
    // the q8_1 image of the most recently quantized activation tensor (scratch slot 0): wq/wk/wv and gate/up share theirs
    struct { const void * data; int64_t ne[4]; size_t nb[4]; size_t span; int variant; int valid; } qcache;
    // per-tile liveness flags of the most recently scanned attention mask (scratch slot 2)
    struct { const void * data; int64_t ne[4]; size_t nb[4]; size_t span; int64_t n; int valid; } mcache;
    int          attr_done[16];
    int          pdl;                // launch the streaming GEMV with programmatic dependent launch (GGML_B200_NO_PDL=1 disables)
    int          pdl_hold;           // != 0: pdl is off for ONE launch (the first behind a cross-stream event wait: its edge must be a full one) and comes back as this
    // device-side timeline of the decode GEMV kernels (b200_debug_trace; measurement only, NULL in production)
    unsigned long long * trace; int trace_cap;
};
// one record per CTA, 32 x u64 (stamps 6.. are kernel-specific fine-grain points): [0] Mtot << 32 | K, [1] smid << 32 | blockIdx, then (globaltimer ns, clock64) pairs at:
// CTA start, prologue prefetch issued, griddepcontrol.wait returned, activations quantized, first stage landed (warp 0), CTA end
#define B200_TRACE_WORDS 32
static __device__ __forceinline__ void b200_trace_stamp(unsigned long long * rec, int i) {
    if (rec) { unsigned long long t; asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t)); rec[2 + 2 * i] = t; rec[3 + 2 * i] = (unsigned long long) clock64(); }
}
static __device__ __forceinline__ unsigned long long * b200_trace_open(unsigned long long * trace, int cap, int Mtot, int K) {
    if (!trace) return nullptr;
    const unsigned long long slot = atomicAdd(trace, 1ull);
    if (slot >= (unsigned long long) cap) return nullptr;
    unsigned smid; asm volatile("mov.u32 %0, %%smid;" : "=r"(smid));
    unsigned long long * rec = trace + B200_TRACE_WORDS * (1 + slot);
    rec[0] = ((unsigned long long) Mtot << 32) | (unsigned) K; rec[1] = ((unsigned long long) smid << 32) | blockIdx.x;
    return rec;
}

void   b200_set_error(const char * fmt, ...);
void * b200_scratch(b200_ctx * ctx, int slot, size_t nbytes);   // returns device ptr, NULL on failure
void   b200_note_write(b200_ctx * ctx, const void * ptr, size_t nbytes);   // invalidates the activation cache on overlap
size_t b200_tensor_span(const b200_tensor * t);
void   b200_prof_begin(b200_ctx * ctx, int kind);               // no-ops unless ctx->profile
void   b200_prof_end(b200_ctx * ctx);

#define B200_CUDA_TRY(expr) do { cudaError_t e_ = (expr); if (e_ != cudaSuccess) { \
        b200_set_error("%s:%d: %s -> %s", __FILE__, __LINE__, #expr, cudaGetErrorString(e_)); return B200_ERR_CUDA; } } while (0)
// every kernel launch is written   B200_LAUNCH(ctx, KIND, kernel<<<grid, block, smem, ctx->stream>>>(args));
#define B200_LAUNCH(ctx, kind, ...) do { if ((ctx)->profile) b200_prof_begin((ctx), (kind)); __VA_ARGS__; \
        if ((ctx)->profile) b200_prof_end(ctx); (ctx)->launches++; (ctx)->launches_kind[kind]++; \
        if ((ctx)->pdl_hold) { (ctx)->pdl = (ctx)->pdl_hold; (ctx)->pdl_hold = 0; } \
        cudaError_t e_ = cudaGetLastError(); if (e_ != cudaSuccess) { \
        b200_set_error("%s:%d: kernel launch -> %s", __FILE__, __LINE__, cudaGetErrorString(e_)); return B200_ERR_CUDA; } } while (0)

// ---- small device helpers ------------------------------------------------------------------
static __device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
static __device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}
// streaming 128-bit load that does not pollute L1 (weights are read exactly once)
static __device__ __forceinline__ int4 ld_stream_16(const void * p) {
    int4 r;
    asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(r.x), "=r"(r.y), "=r"(r.z), "=r"(r.w) : "l"(p));
    return r;
}
static __device__ __forceinline__ int ld_stream_4(const void * p) {
    int r;
    asm volatile("ld.global.nc.L1::no_allocate.s32 %0, [%1];" : "=r"(r) : "l"(p));
    return r;
}
static __device__ __forceinline__ int ld_u16x2(const void * p) {   // 4 bytes from a 2-byte aligned address
    const uint16_t * q = (const uint16_t *) p;
    return (int) ((uint32_t) q[0] | ((uint32_t) q[1] << 16));
}

// ---- q8_1 block quantization, bit-exact with the reference CPU code (see b200_quantize.cu for the citations) ----
// float -> int32 the way x86 cvtps2dq / cvttss2si do it: out-of-range (a denormal amax makes id = inf) and NaN
// give the "integer indefinite" 0x80000000 instead of CUDA's saturation.
static __device__ __forceinline__ int cvt_x86(float r) { return fabsf(r) < 2147483648.0f ? __float2int_rn(r) : (int) 0x80000000u; }

// MAXPS / MAXSS and the scalar MAX() macro alike: a > b ? a : b, i.e. the SECOND operand whenever one of them is NaN
static __device__ __forceinline__ float mx_x86(float a, float b) { return a > b ? a : b; }
// amax of one 32-value block held one value per lane.  NaN-free blocks: the plain maximum.  With NaNs the reference's result depends on
// where they sit: the AVX2 body folds four 8-lane vectors with _mm256_max_ps, then high half against low half, movehl, movehdup
// (arch/x86/quants.c:402-412); the scalar code runs MAX(amax, fabsf(v)) left to right (ggml-quants.c:233-236).  Same order here.
template <int VARIANT>
static __device__ __forceinline__ float q8_1_amax(float v) {
    const float a = fabsf(v);
    if (!__any_sync(0xffffffffu, v != v)) return warp_max(a);
    if (VARIANT == B200_Q8_1_X86) {
        float t = mx_x86(a, __shfl_down_sync(0xffffffffu, a, 8));          // lanes 0-7: lane j folds x[j], x[8+j], x[16+j], x[24+j]
        t = mx_x86(t, __shfl_down_sync(0xffffffffu, a, 16));
        t = mx_x86(t, __shfl_down_sync(0xffffffffu, a, 24));
        t = mx_x86(__shfl_down_sync(0xffffffffu, t, 4), t);                // lanes 0-3: max(high half, low half)
        t = mx_x86(t, __shfl_down_sync(0xffffffffu, t, 2));                // lanes 0-1: max(max4, movehl)
        t = mx_x86(t, __shfl_down_sync(0xffffffffu, t, 1));                // lane 0:    max_ss(max4, movehdup)
        return __shfl_sync(0xffffffffu, t, 0);
    }
    float amax = 0.0f;
    for (int j = 0; j < 32; ++j) amax = mx_x86(amax, __shfl_sync(0xffffffffu, a, j));
    return amax;
}

// fp32 -> fp16 as vcvtps2ph gives it for the NaNs a q8_1 block can hold: a NaN PROPAGATED from a (default, positive) NaN input -> 0x7e00,
// a NaN GENERATED by an invalid operation (inf * 0) is x86's negative "real indefinite" -> 0xfe00.  CUDA arithmetic returns the canonical
// 0x7fffffff for both, so the caller says which one it is.  (Payloads of non-default input NaNs are not reproduced.)
static __device__ __forceinline__ __half f2h_q8_1(float v, bool generated) {
    return v != v ? __ushort_as_half((unsigned short) (generated ? 0xfe00 : 0x7e00)) : __float2half_rn(v);
}
// the (d, s) pair of a block: s = d * sum is a propagated NaN when d is one, a generated one for inf * 0
static __device__ __forceinline__ __half2 q8_1_ds(float d, int isum) {
    return __halves2half2(f2h_q8_1(d, false), f2h_q8_1(__fmul_rn(d, (float) isum), !(d != d)));
}

// one warp, one block: returns this lane's quant, the block scale d (fp32) and the integer sum of quants
template <int VARIANT>
static __device__ __forceinline__ void q8_1_block(float v, int & q, float & d, int & isum) {
    const float amax = q8_1_amax<VARIANT>(v);
    d = __fdiv_rn(amax, 127.0f);
    unsigned usum;
    if (VARIANT == B200_Q8_1_X86) {
        const float id  = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
        const int   q32 = cvt_x86(rintf(__fmul_rn(v, id)));      // round-half-even, as _mm256_round_ps(NEAREST) + cvtps_epi32
        q    = max(-128, min(127, q32));                         // packs_epi32/epi16 saturate
        usum = (unsigned) q32;                                   // ... but the AVX2 body sums the int32 lanes before packing
    } else {
        const float id  = d != 0.0f ? __fdiv_rn(1.0f, d) : 0.0f;
        const int   q32 = cvt_x86(roundf(__fmul_rn(v, id)));     // half away from zero, as roundf(); then C's (int8_t) cast
        q    = (int) (int8_t) (q32 & 0xff);
        usum = (unsigned) q;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) usum += __shfl_xor_sync(0xffffffffu, usum, o);
    isum = (int) usum;
}

// The arithmetic core of the CTA-wide quantizers below: this lane holds 16 consecutive values f (one half of a 32-value block), lane ^ 1 the other half.  Returns the
// 16 quants (w), the block scale d = amax / 127 (fp32, before its fp16 round trip) and the block's int32 quant sum as the x86 code forms it.
static __device__ __forceinline__ void b200_q8_1_half16(const float (&f)[16], int lane, int4 & w, float & d, unsigned & usum) {
    // amax over the |bit patterns| as integers: same order as the floats, and a NaN input shows up as amax > inf
    unsigned ab[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) ab[i] = max(__float_as_uint(f[2 * i]) & 0x7fffffffu, __float_as_uint(f[2 * i + 1]) & 0x7fffffffu);
    unsigned abits = max(max(max(ab[0], ab[1]), max(ab[2], ab[3])), max(max(ab[4], ab[5]), max(ab[6], ab[7])));
    abits = max(abits, __shfl_xor_sync(0xffffffffu, abits, 1));
    const bool  nan_in = abits > 0x7f800000u;
    // with a NaN in the block the reference's amax depends on where it sits (q8_1_amax): same fold order, rare path
    float amax = __uint_as_float(abits);
    if (__any_sync(0xffffffffu, nan_in)) {
        float lo[16], hi[16];                                         // x[0..15] (even lane of the pair), x[16..31] (odd lane)
#pragma unroll
        for (int i = 0; i < 16; ++i) {
            const float o = __shfl_xor_sync(0xffffffffu, f[i], 1);
            lo[i] = fabsf((lane & 1) ? o : f[i]); hi[i] = fabsf((lane & 1) ? f[i] : o);
        }
        float m[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) m[j] = mx_x86(mx_x86(mx_x86(lo[j], lo[8 + j]), hi[j]), hi[8 + j]);
        amax = mx_x86(mx_x86(mx_x86(m[4], m[0]), mx_x86(m[6], m[2])), mx_x86(mx_x86(m[5], m[1]), mx_x86(m[7], m[3])));
    }
    d = __fdiv_rn(amax, 127.0f);
    const float id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
    int * wp = (int *) &w;
    // finite inputs and a finite 127/amax: |x * id| <= 127 (1 + 2^-23), so every quant is in [-127, 127]: nothing saturates, the
    // x86 "integer indefinite" cannot occur and the int32 sum equals the sum of the packed bytes
    const bool fast = !nan_in && abits < 0x7f800000u && (__float_as_uint(id) & 0x7fffffffu) < 0x7f800000u;
    if (__all_sync(0xffffffffu, fast)) {
        int acc = 0;
#pragma unroll
        for (int i = 0; i < 4; ++i) {
            const int q0 = __float2int_rn(__fmul_rn(f[4 * i], id)),     q1 = __float2int_rn(__fmul_rn(f[4 * i + 1], id));
            const int q2 = __float2int_rn(__fmul_rn(f[4 * i + 2], id)), q3 = __float2int_rn(__fmul_rn(f[4 * i + 3], id));
            int hi, wd;
            asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(hi) : "r"(q3), "r"(q2), "r"(0));
            asm("cvt.pack.sat.s8.s32.b32 %0, %1, %2, %3;" : "=r"(wd) : "r"(q1), "r"(q0), "r"(hi));
            wp[i] = wd;
            acc = __dp4a(wd, 0x01010101, acc);
        }
        usum = (unsigned) acc;
    } else {
        int q[16];
#pragma unroll
        for (int i = 0; i < 16; ++i) q[i] = cvt_x86(__fmul_rn(f[i], id));   // cvt.rni rounds half to even itself: same as rintf() first
        usum = 0u;                                                    // int32 lanes summed before saturation (AVX2 body)
#pragma unroll
        for (int i = 0; i < 16; ++i) usum += (unsigned) q[i];
#pragma unroll
        for (int i = 0; i < 4; ++i)
            wp[i] = (int) (((unsigned) max(-128, min(127, q[4 * i])) & 0xffu)            | (((unsigned) max(-128, min(127, q[4 * i + 1])) & 0xffu) << 8) |
                           (((unsigned) max(-128, min(127, q[4 * i + 2])) & 0xffu) << 16) | (((unsigned) max(-128, min(127, q[4 * i + 3])) & 0xffu) << 24));
    }
    usum += __shfl_xor_sync(0xffffffffu, usum, 1);
}
struct b200_no_hook { __device__ __forceinline__ void operator()() const {} };
// after_loads() runs once, right after the first round of loads has been issued (a place to start bulk traffic BEHIND them)
// SwiGLU as the GLU node computes it (ggml-cpu/ops.cpp glu, vec.h silu: x / (1 + exp(-x))), shared by k_swiglu and the GEMV epilogue that finishes gate/up
static __device__ __forceinline__ float b200_swiglu1(float x, float y) { return x / (1.0f + expf(-x)) * y; }
// CTA-wide q8_1 quantization (x86 variant, bit-exact with q8_1_block<B200_Q8_1_X86>) of one f32 column x[0, K) for the decode GEMV
// kernels: TWO lanes per 32-value block, 16 values (one 16-byte unit of quants) per lane, so a block costs one shuffle for amax and
// one for the sum and the 16 conversions of a lane are independent instructions (the 8-lanes-per-block form spent ~2000 cycles per
// CTA in shuffle / division latency chains, measured with b200_debug_trace).  Rounds are double-buffered: the next round's loads are
// in flight while this one converts.  store(blk32, hf, quants, d, s): hf = which half of block blk32, d and s after their fp16 round trip.
template <class F, class H = b200_no_hook>
static __device__ __forceinline__ void b200_quantize_column_x86(const float * x, int K, F store, unsigned long long * rec = nullptr, H after_loads = H()) {
    const int nunits = K >> 4, lane = threadIdx.x & 31;
    const float4 * x4 = (const float4 *) x;
    // every CTA of the grid reads the same small column while gigabytes of weights stream through L2 (evict_first): ask L2 to keep it
    uint64_t keep;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(keep));
    auto ldx = [&](int i) {
        float4 r;
        asm volatile("ld.global.L2::cache_hint.v4.f32 {%0, %1, %2, %3}, [%4], %5;" : "=f"(r.x), "=f"(r.y), "=f"(r.z), "=f"(r.w) : "l"(x4 + i), "l"(keep));
        return r;
    };
    int base = (int) (threadIdx.x & ~31u);                               // warp-uniform; K % 32 == 0 keeps lane pairs together
    float4 v[4];
    if (base < nunits) {
#pragma unroll
        for (int j = 0; j < 4; ++j) v[j] = base + lane < nunits ? ldx((base + lane) * 4 + j) : make_float4(0.f, 0.f, 0.f, 0.f);
    }
    after_loads();
    while (base < nunits) {
        const int u = base + lane, nb = base + (int) blockDim.x;
        float4 nx[4];
        if (nb < nunits) {
#pragma unroll
            for (int j = 0; j < 4; ++j) nx[j] = nb + lane < nunits ? ldx((nb + lane) * 4 + j) : make_float4(0.f, 0.f, 0.f, 0.f);
        }
        const float f[16] = { v[0].x, v[0].y, v[0].z, v[0].w, v[1].x, v[1].y, v[1].z, v[1].w, v[2].x, v[2].y, v[2].z, v[2].w, v[3].x, v[3].y, v[3].z, v[3].w };
        int4 w; float d; unsigned usum;
        b200_q8_1_half16(f, lane, w, d, usum);
        if (rec) b200_trace_stamp(rec, 8);                               // (measurement only) the column's first loads have landed
        if (u < nunits) store(u >> 1, u & 1, w, __half2float(__float2half_rn(d)), __half2float(__float2half_rn(__fmul_rn(d, (float) (int) usum))));
        if (nb < nunits) {
#pragma unroll
            for (int j = 0; j < 4; ++j) v[j] = nx[j];
        }
        base = nb;
    }
}

// ---- internal launchers (defined in the .cu files) ----------------------------------------
int b200_launch_quantize_q8_1(b200_ctx * ctx, const float * x, size_t x_row_stride, void * y, int64_t k, int64_t nrows, int variant);
int b200_launch_quantize_q8_1_soa(b200_ctx * ctx, const float * x, size_t nb1, size_t nb2, size_t nb3,
                                  int64_t k, int64_t ne1, int64_t ne2, int64_t ne3, void * image, int variant);
static inline size_t b200_q8_1_soa_size(int64_t k, int64_t ncol) { return (size_t) ncol * k + (size_t) ncol * (k / 32) * 4; }
int b200_launch_gemv_q(b200_ctx * ctx, const b200_tensor * w, const void * image, int64_t ne11, int64_t ne12, int64_t ne13, b200_tensor * dst);
int b200_launch_gemv_id_q(b200_ctx * ctx, const b200_tensor * as, const void * image, int64_t ne11, int64_t n_tokens,
                          const b200_tensor * ids, b200_tensor * dst);
int b200_launch_fattn_vec(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                          const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                          float scale, float max_bias, float softcap);
bool b200_fattn_dec_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v);
int  b200_launch_fattn_dec(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                           const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                           float scale, float max_bias, float softcap, const b200_kv_pages * pages = nullptr);
bool b200_fattn_gqa_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst);
int  b200_launch_fattn_gqa(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                           const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst, float scale, float max_bias, float softcap,
                           const b200_kv_pages * pages = nullptr);
bool b200_fattn_tc_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v, const b200_tensor * mask, const b200_tensor * dst);
int  b200_launch_fattn_tc(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                          const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                          float scale, float max_bias, float softcap);     // B200_ERR_UNSUPPORTED = use another kernel
bool b200_fattn_kvq_supported(const b200_tensor * q, const b200_tensor * k, const b200_tensor * v);      // q8_0 / q4_0 K and V (b200_fattn_q.cu)
int  b200_launch_fattn_kvq(b200_ctx * ctx, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                           const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst, float scale, float max_bias, float softcap);
bool b200_gemv_stream_supported(const b200_tensor * w, const b200_tensor * x);
// peers of a row-split MUL_MAT: the low-latency mirror (8 bytes per arena float) of every rank's arena, our own arena and its
// control words, and the full f32 columns (inside our arena) that the kernel receives after it has sent its own rows
struct b200_rowsplit_peers { int nll; void * ll_base[B200_MAX_PEERS]; const void * ll_self; const void * arena; unsigned * ctrl; float * full[3]; int full_m[3]; };
// decode MUL_MAT_ID through the streaming GEMV: segment s reads expert ids[base + s] (device memory) at byte stride `stride`
struct b200_gemv_ids { const int * ids; int base; size_t stride; const void * ads; };   // ads: (d, s) pairs of this column when `image` is one column of a wider q8_1 image
int  b200_launch_gemv_stream(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, const void * image, float * const * dst,
                             const b200_rowsplit_peers * peers = nullptr, const b200_gemv_ids * idm = nullptr);
bool b200_gemv_mma_supported(const b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x);
int  b200_launch_gemv_mma(b200_ctx * ctx, const b200_tensor * const * w, int n, const b200_tensor * x, const void * image, float * const * dst,
                          const b200_rowsplit_peers * peers, float * glu = nullptr);   // glu: n == 2 (gate, up), also stores swiglu(gate, up)
int  b200_launch_exchange_unpack(b200_ctx * ctx, const void * ll, float * dst, int64_t M, unsigned * ctrl);
#define B200_GEMV_FUSED_QUANT_MAX_K 16384   // above this, one quantize launch + image beats one quantization per CTA
#define B200_GEMV_MAX_COLS 8
bool b200_gemm_supported(const b200_tensor * w, const b200_tensor * x);
int  b200_launch_gemm_q(b200_ctx * ctx, const b200_tensor * w, const void * image, int64_t N, b200_tensor * dst);   // image: bf16 [N][K]
int  b200_launch_gemm_multi(b200_ctx * ctx, const b200_tensor * const * w, int n, const void * image, int64_t N, b200_tensor * const * dst);
int  b200_launch_act_bf16(b200_ctx * ctx, const float * x, size_t nb1, int64_t K, int64_t N, void * image);
#define B200_ACT_BF16 2                     // activation-image kind next to the two q8_1 variants
bool b200_gemm_id_supported(const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids);
int  b200_launch_gemm_id_q(b200_ctx * ctx, const b200_tensor * as, const b200_tensor * b, const b200_tensor * ids, b200_tensor * dst);
int  b200_launch_matmul_float(b200_ctx * ctx, const b200_tensor * w, const b200_tensor * x, b200_tensor * dst);
