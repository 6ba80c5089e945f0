// b200_glue.cu -- the cheap ops between the hot-path nodes of a llama graph: RMS_NORM, ADD / MUL / SUB / DIV with
// broadcast, ROPE (normal and neox), SET_ROWS (the KV-cache write), CPY / CONT / DUP, GLU (SwiGLU), GET_ROWS (f32 rows), SCALE, SOFT_MAX, SILU.
// They exist so that ggml_backend_sched keeps a whole decode graph on the device instead of bouncing every second node through the
// CPU backend.  Plain bandwidth-trivial kernels over ggml's strided views; semantics follow the reference's CPU operators
// (ggml/src/ggml-cpu/ops.cpp: rms_norm 3560-3640, rope 5524-5790, set_rows, glu; binary-ops.cpp; ggml.c:4255-4269 for the YaRN
// correction range).  What is not listed in b200_glue_supports() stays on the CPU backend.
#include "b200_glue.cuh"
#include "../../include/ggml-b200-abi.h"
#include <math.h>
#include <string.h>

// every glue kernel is launched with programmatic dependent launch: the next kernel (usually a GEMV that prefetches its weights) may
// start at once, and we wait for the kernel in front of us before touching memory
#define GL_PDL() do { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); asm volatile("griddepcontrol.wait;" ::: "memory"); } while (0)

struct gl_view { char * p; int64_t ne[4]; size_t nb[4]; };
static gl_view gl_of(const ggml_tensor * t) {
    gl_view v; v.p = (char *) t->data;
    for (int i = 0; i < 4; ++i) { v.ne[i] = t->ne[i]; v.nb[i] = t->nb[i]; }
    return v;
}
static bool gl_contig_rows(const ggml_tensor * t, size_t es) { return t->nb[0] == es; }
static bool gl_contiguous(const ggml_tensor * t, size_t es) {
    return t->nb[0] == es && t->nb[1] == es * t->ne[0] && t->nb[2] == t->nb[1] * t->ne[1] && t->nb[3] == t->nb[2] * t->ne[2];
}
static int64_t gl_nrows(const ggml_tensor * t) { return t->ne[1] * t->ne[2] * t->ne[3]; }
static int64_t gl_nelem(const ggml_tensor * t) { return t->ne[0] * gl_nrows(t); }
static size_t gl_esize(int type) { return type == GGML_TYPE_F32 || type == GGML_TYPE_I32 ? 4 : type == GGML_TYPE_F16 || type == GGML_TYPE_BF16 ? 2 : type == GGML_TYPE_I64 ? 8 : 0; }

// row index -> (i1, i2, i3) of a view
static __device__ __forceinline__ void gl_unrow(int64_t r, const gl_view & v, int64_t & i1, int64_t & i2, int64_t & i3) {
    i1 = r % v.ne[1]; r /= v.ne[1]; i2 = r % v.ne[2]; i3 = r / v.ne[2];
}

// Launch shape of the row-wise kernels.  A decode graph runs every one of them on a single row of 4096 .. 14336 floats: a loop of
// dependent load -> store rounds in one CTA costs a memory latency per round (k_swiglu measured 60 us that way), so a row is cut into
// chunks of GL_CHUNK elements, one CTA each, and every thread issues all its loads (one float4 when the row is 16-byte aligned)
// before its first store.  blockIdx.x = row * chunks + chunk.
#define GL_CHUNK 1024
static bool gl_vec_ok(const ggml_tensor * t) {                   // rows of float4: 16-byte aligned base and strides, ne0 % 4 == 0
    return t->ne[0] % 4 == 0 && ((uintptr_t) t->data & 15) == 0 && t->nb[1] % 16 == 0 && t->nb[2] % 16 == 0 && t->nb[3] % 16 == 0;
}
static __device__ __forceinline__ float block_sum(float s, float * red) {          // red: 32 floats of shared memory
    s = warp_sum(s);
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    s = (int) threadIdx.x < nw ? red[threadIdx.x] : 0.0f;
    if (threadIdx.x < 32) { s = warp_sum(s); if (threadIdx.x == 0) red[0] = s; }
    __syncthreads();
    return red[0];
}

// ---- RMS_NORM: y = x / sqrt(mean(x^2) + eps), one CTA per row; optionally followed by MUL with the norm weight (every llama norm):
// one pass, both nodes' outputs written.  CACHED: the row (<= 8 * blockDim elements, float4-aligned) stays in registers between the
// sum and the scaling ----------------------------------------------------------------------------------------------------------------
// k_add_rms_norm_mul: the residual ADD in front of the norm taken along (x = a + b stored to the ADD node's dst u, then as below): the
// three nodes between two matmuls of a llama layer in one launch
__global__ void __launch_bounds__(1024) k_add_rms_norm_mul(const gl_view a, const gl_view b, const gl_view u, const gl_view y, const gl_view w, const gl_view z, float eps, int write_y) {
    GL_PDL();
    int64_t i1, i2, i3; gl_unrow(blockIdx.x, a, i1, i2, i3);
    const float * ar = (const float *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    const float * br = (const float *) (b.p + i1 * b.nb[1] + i2 * b.nb[2] + i3 * b.nb[3]);
    float *       ur = (float *) (u.p + i1 * u.nb[1] + i2 * u.nb[2] + i3 * u.nb[3]);
    float *       yr = (float *) (y.p + i1 * y.nb[1] + i2 * y.nb[2] + i3 * y.nb[3]);
    const float * wr = (const float *) (w.p + (i1 % w.ne[1]) * w.nb[1] + (i2 % w.ne[2]) * w.nb[2] + (i3 % w.ne[3]) * w.nb[3]);
    float *       zr = (float *) (z.p + i1 * z.nb[1] + i2 * z.nb[2] + i3 * z.nb[3]);
    __shared__ float red[32];
    const int64_t n0 = a.ne[0];
    float4 v[2], g[2];
    float s = 0.0f;
    #pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int64_t i = ((int64_t) k * blockDim.x + threadIdx.x) * 4;
        const float4 p = i < n0 ? *(const float4 *) (ar + i) : make_float4(0, 0, 0, 0), q = i < n0 ? *(const float4 *) (br + i) : make_float4(0, 0, 0, 0);
        g[k] = i < n0 ? *(const float4 *) (wr + i) : make_float4(0, 0, 0, 0);
        v[k] = make_float4(p.x + q.x, p.y + q.y, p.z + q.z, p.w + q.w);
    }
    #pragma unroll
    for (int k = 0; k < 2; ++k) { s = fmaf(v[k].x, v[k].x, s); s = fmaf(v[k].y, v[k].y, s); s = fmaf(v[k].z, v[k].z, s); s = fmaf(v[k].w, v[k].w, s); }
    const float scale = rsqrtf(block_sum(s, red) / (float) n0 + eps);
    #pragma unroll
    for (int k = 0; k < 2; ++k) {
        const int64_t i = ((int64_t) k * blockDim.x + threadIdx.x) * 4;
        if (i >= n0) continue;
        *(float4 *) (ur + i) = v[k];
        const float4 o = make_float4(v[k].x * scale, v[k].y * scale, v[k].z * scale, v[k].w * scale);
        if (write_y) *(float4 *) (yr + i) = o;
        *(float4 *) (zr + i) = make_float4(o.x * g[k].x, o.y * g[k].y, o.z * g[k].z, o.w * g[k].w);
    }
}

template <bool MUL, bool CACHED>
__global__ void __launch_bounds__(1024) k_rms_norm(const gl_view x, const gl_view y, const gl_view w, const gl_view z, float eps, int write_y) {
    GL_PDL();
    int64_t i1, i2, i3; gl_unrow(blockIdx.x, x, i1, i2, i3);
    const float * xr = (const float *) (x.p + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float *       yr = (float *) (y.p + i1 * y.nb[1] + i2 * y.nb[2] + i3 * y.nb[3]);
    const float * wr = MUL ? (const float *) (w.p + (i1 % w.ne[1]) * w.nb[1] + (i2 % w.ne[2]) * w.nb[2] + (i3 % w.ne[3]) * w.nb[3]) : nullptr;
    float *       zr = MUL ? (float *) (z.p + i1 * z.nb[1] + i2 * z.nb[2] + i3 * z.nb[3]) : nullptr;
    __shared__ float red[32];
    const int64_t n0 = x.ne[0];
    if (CACHED) {
        float4 v[2], g[2];
        float s = 0.0f;
        #pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int64_t i = ((int64_t) k * blockDim.x + threadIdx.x) * 4;
            v[k] = i < n0 ? *(const float4 *) (xr + i) : make_float4(0, 0, 0, 0);
            if (MUL) g[k] = i < n0 ? *(const float4 *) (wr + i) : make_float4(0, 0, 0, 0);
        }
        #pragma unroll
        for (int k = 0; k < 2; ++k) { s = fmaf(v[k].x, v[k].x, s); s = fmaf(v[k].y, v[k].y, s); s = fmaf(v[k].z, v[k].z, s); s = fmaf(v[k].w, v[k].w, s); }
        const float scale = rsqrtf(block_sum(s, red) / (float) n0 + eps);
        #pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int64_t i = ((int64_t) k * blockDim.x + threadIdx.x) * 4;
            if (i >= n0) continue;
            const float4 o = make_float4(v[k].x * scale, v[k].y * scale, v[k].z * scale, v[k].w * scale);   // the RMS_NORM node's value, rounded as the unfused kernel rounds it
            if (!MUL || write_y) *(float4 *) (yr + i) = o;
            if (MUL) *(float4 *) (zr + i) = make_float4(o.x * g[k].x, o.y * g[k].y, o.z * g[k].z, o.w * g[k].w);
        }
    } else {
        float s = 0.0f;
        for (int64_t i = threadIdx.x; i < n0; i += blockDim.x) { const float v = xr[i]; s = fmaf(v, v, s); }
        const float scale = rsqrtf(block_sum(s, red) / (float) n0 + eps);
        for (int64_t i = threadIdx.x; i < n0; i += blockDim.x) {
            const float v = xr[i] * scale;
            if (!MUL || write_y) yr[i] = v;
            if (MUL) zr[i] = v * wr[i];
        }
    }
}

// ---- ADD / SUB / MUL / DIV with src1 broadcast over src0 (ggml_can_repeat) ------------------------------------------------------------
template <int OP> static __device__ __forceinline__ float gl_bin(float x, float y) { return OP == 0 ? x + y : OP == 1 ? x - y : OP == 2 ? x * y : x / y; }
template <int OP, bool VEC>
__global__ void __launch_bounds__(256) k_binary(const gl_view a, const gl_view b, const gl_view d, int chunks) {
    GL_PDL();
    const int64_t row = blockIdx.x / chunks; const int64_t c0 = (int64_t) (blockIdx.x - row * chunks) * GL_CHUNK;
    int64_t i1, i2, i3; gl_unrow(row, d, i1, i2, i3);
    const float * ar = (const float *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]);
    const float * br = (const float *) (b.p + (i1 % b.ne[1]) * b.nb[1] + (i2 % b.ne[2]) * b.nb[2] + (i3 % b.ne[3]) * b.nb[3]);
    float *       dr = (float *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const int64_t n0 = d.ne[0], nb0 = b.ne[0];
    if (VEC) {                                                                    // (nb0 == n0)
        const int64_t i = c0 + threadIdx.x * 4;
        if (i >= n0) return;
        const float4 x = *(const float4 *) (ar + i), y = *(const float4 *) (br + i);
        *(float4 *) (dr + i) = make_float4(gl_bin<OP>(x.x, y.x), gl_bin<OP>(x.y, y.y), gl_bin<OP>(x.z, y.z), gl_bin<OP>(x.w, y.w));
    } else {
        float x[4], y[4];
        #pragma unroll
        for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < n0) { x[k] = ar[i]; y[k] = br[nb0 == n0 ? i : i % nb0]; } }
        #pragma unroll
        for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < n0) dr[i] = gl_bin<OP>(x[k], y[k]); }
    }
}

// ---- SCALE: y = x * s + b (contiguous) ---------------------------------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_scale(const float * x, float * y, int64_t n, float s, float b) {
    GL_PDL();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) y[i] = fmaf(x[i], s, b);
}

// ---- the MoE router's small ops (src/llama-graph.cpp:1097-1300: argsort_top_k of the expert probabilities, sum / clamp of the chosen weights) ----
// ARGSORT (ggml-cpu/ops.cpp ggml_compute_forward_argsort_f32): dst[rank] = index.  One CTA per row, thread i ranks element i by counting
// the elements that go before it (ties: lower index first, i.e. a stable order); rows of at most 1024 elements.
__global__ void __launch_bounds__(1024) k_argsort(const gl_view s, const gl_view d, int desc) {
    GL_PDL();
    __shared__ float row[1024];
    int64_t i1, i2, i3; gl_unrow(blockIdx.x, s, i1, i2, i3);
    const float * sr = (const float *) (s.p + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3]);
    int32_t * dr = (int32_t *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const int n = (int) s.ne[0], i = threadIdx.x;
    if (i < n) row[i] = sr[i];
    __syncthreads();
    if (i >= n) return;
    const float v = row[i];
    int rank = 0;
    for (int j = 0; j < n; ++j) { const float w = row[j]; rank += desc ? (w > v || (w == v && j < i)) : (w < v || (w == v && j < i)); }
    dr[rank] = i;
}
// SUM_ROWS: dst[0, i1, i2, i3] = sum over i0; one warp per row
__global__ void __launch_bounds__(256) k_sum_rows(const gl_view s, const gl_view d, int64_t nrows) {
    GL_PDL();
    const int64_t r = (int64_t) blockIdx.x * 8 + (threadIdx.x >> 5);
    if (r >= nrows) return;
    int64_t i1, i2, i3; gl_unrow(r, s, i1, i2, i3);
    const float * sr = (const float *) (s.p + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3]);
    float acc = 0.0f;
    for (int64_t i = threadIdx.x & 31; i < s.ne[0]; i += 32) acc += sr[i];
    acc = warp_sum(acc);
    if ((threadIdx.x & 31) == 0) *(float *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]) = acc;
}
// CLAMP (contiguous)
__global__ void __launch_bounds__(256) k_clamp(const float * x, float * y, int64_t n, float lo, float hi) {
    GL_PDL();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const float v = x[i]; y[i] = v < lo ? lo : v > hi ? hi : v; }
}

// ---- ROPE (GGML_ROPE_TYPE_NORMAL = 0: pairs (2i, 2i+1); NEOX = 2: pairs (i, i + n_dims/2)), YaRN as ops.cpp:5524-5547 -----------------
struct gl_rope { int n_dims, neox; float freq_scale, ext_factor, attn_factor, theta_scale, corr0, corr1; const int32_t * pos; const float * ff; };
// head rows [h0, h1) of token i2, batch i3: the angle of a pair depends on the position and the pair only, so it is evaluated once and
// applied to every head of the group (prefill: 8 heads per CTA; decode: one, for the parallelism).  cache(h): NULL or the f16 KV-cache row
// segment that takes a copy of head h's result
template <class C>
static __device__ __forceinline__ void gl_rope_rows(const gl_view & x, const gl_view & y, const gl_rope & r, int64_t h0, int64_t h1, int64_t i2, int64_t i3, C cache) {
    const char * xb = x.p + i2 * x.nb[2] + i3 * x.nb[3];
    char *       yb = y.p + i2 * y.nb[2] + i3 * y.nb[3];
    const float p = (float) r.pos[i2];
    for (int64_t ip = threadIdx.x; ip < x.ne[0] / 2; ip += blockDim.x) {       // pair index
        const int64_t i0 = 2 * ip;
        if (i0 >= r.n_dims) {                                                   // channels beyond n_dims pass through
            for (int64_t h = h0; h < h1; ++h) {
                const float * xr = (const float *) (xb + h * x.nb[1]); float * yr = (float *) (yb + h * y.nb[1]);
                const float x0 = xr[i0], x1 = xr[i0 + 1];
                yr[i0] = x0; yr[i0 + 1] = x1;
                __half * ch = cache(h);
                if (ch) { ch[i0] = __float2half_rn(x0); ch[i0 + 1] = __float2half_rn(x1); }
            }
            continue;
        }
        const float theta_extrap = p * powf(r.theta_scale, (float) ip) / (r.ff ? r.ff[ip] : 1.0f);
        const float theta_interp = r.freq_scale * theta_extrap;
        float theta = theta_interp, mscale = r.attn_factor;
        if (r.ext_factor != 0.0f) {
            const float yy = ((float) ip - r.corr0) / fmaxf(0.001f, r.corr1 - r.corr0);
            const float ramp_mix = (1.0f - fminf(1.0f, fmaxf(0.0f, yy))) * r.ext_factor;
            theta = theta_interp * (1.0f - ramp_mix) + theta_extrap * ramp_mix;
            mscale *= 1.0f + 0.1f * logf(1.0f / r.freq_scale);
        }
        float sn, cs; sincosf(theta, &sn, &cs); sn *= mscale; cs *= mscale;
        const int64_t ia = r.neox ? ip : i0, ib = r.neox ? ip + r.n_dims / 2 : i0 + 1;
        // heads in groups of 8: all 16 loads of a group are issued before the first store (one head after the other cost a memory latency
        // per head: ROPE was 0.66 ms of a 13.5 ms pp512 ubatch)
        for (int64_t hg = h0; hg < h1; hg += 8) {
            float x0[8], x1[8];
#pragma unroll
            for (int j = 0; j < 8; ++j) if (hg + j < h1) { const float * xr = (const float *) (xb + (hg + j) * x.nb[1]); x0[j] = xr[ia]; x1[j] = xr[ib]; }
#pragma unroll
            for (int j = 0; j < 8; ++j) if (hg + j < h1) {
                const int64_t h = hg + j;
                float * yr = (float *) (yb + h * y.nb[1]);
                const float y0 = x0[j] * cs - x1[j] * sn, y1 = x0[j] * sn + x1[j] * cs;
                yr[ia] = y0; yr[ib] = y1;
                __half * ch = cache(h);
                if (ch) { ch[ia] = __float2half_rn(y0); ch[ib] = __float2half_rn(y1); }
            }
        }
    }
}
struct gl_no_cache { __device__ __forceinline__ __half * operator()(int64_t) const { return nullptr; } };
static int gl_rope_heads_per_cta(const ggml_tensor * t) { return t->ne[2] * t->ne[3] >= 64 ? (int) (t->ne[1] < 8 ? t->ne[1] : 8) : 1; }
__global__ void __launch_bounds__(128) k_rope(const gl_view x, const gl_view y, const gl_rope r, int hb) {
    GL_PDL();
    const int64_t h0 = (int64_t) blockIdx.x * hb, h1 = h0 + hb < x.ne[1] ? h0 + hb : x.ne[1];
    gl_rope_rows(x, y, r, h0, h1, blockIdx.y, blockIdx.z, gl_no_cache());       // head group, token, batch
}

// ---- SET_ROWS: dst row idx[i1, i2 % ne11, i3 % ne12] <- src row (i1, i2, i3), f32 -> f32 / f16 ---------------------------------------------
template <typename T> static __device__ __forceinline__ T gl_cvt(float v);
template <> __device__ __forceinline__ float  gl_cvt<float>(float v)  { return v; }
template <> __device__ __forceinline__ __half gl_cvt<__half>(float v) { return __float2half_rn(v); }
template <typename T, bool VEC>
__global__ void __launch_bounds__(256) k_set_rows(const gl_view s, const gl_view idx, const gl_view d, int chunks) {
    GL_PDL();
    const int64_t r = blockIdx.x / chunks; const int64_t c0 = (int64_t) (blockIdx.x - r * chunks) * GL_CHUNK;
    int64_t i1, i2, i3; gl_unrow(r, s, i1, i2, i3);
    const int64_t row = *(const int64_t *) (idx.p + i1 * idx.nb[0] + (i2 % idx.ne[1]) * idx.nb[1] + (i3 % idx.ne[2]) * idx.nb[2]);
    const float * sr = (const float *) (s.p + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3]);
    T * dr = (T *) (d.p + row * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const int64_t n0 = s.ne[0];
    if (VEC) {                                                                    // src rows float4-aligned, dst rows 16-byte aligned
        const int64_t i = c0 + threadIdx.x * 4;
        if (i >= n0) return;
        const float4 v = *(const float4 *) (sr + i);
        if (sizeof(T) == 4) *(float4 *) (dr + i) = v;
        else { __half2 lo = __floats2half2_rn(v.x, v.y), hi = __floats2half2_rn(v.z, v.w); uint2 o; o.x = *(unsigned *) &lo; o.y = *(unsigned *) &hi; *(uint2 *) (dr + i) = o; }
    } else {
        float v[4];
        #pragma unroll
        for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < n0) v[k] = sr[i]; }
        #pragma unroll
        for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < n0) dr[i] = gl_cvt<T>(v[k]); }
    }
}

// SET_ROWS into a quantized cache (llama.cpp -ctk / -ctv q8_0, q4_0): thread = one 32-element block, quantized as ggml-cpu's set_rows does
// through from_float (ops.cpp, ggml_compute_forward_set_rows_f32): q8_0 with the x86 body's arithmetic, q4_0 by quantize_row_q4_0_ref
// (ggml/src/ggml-quants.c:38-74; it has no SIMD form)
template <int Q4>
__global__ void __launch_bounds__(256) k_set_rows_q(const gl_view s, const gl_view idx, const gl_view d, int nblk, int64_t total) {
    GL_PDL();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= total) return;
    const int b = (int) (i % nblk); const int64_t r = i / nblk;
    int64_t i1, i2, i3; gl_unrow(r, s, i1, i2, i3);
    const int64_t row = *(const int64_t *) (idx.p + i1 * idx.nb[0] + (i2 % idx.ne[1]) * idx.nb[1] + (i3 % idx.ne[2]) * idx.nb[2]);
    const float * sr = (const float *) (s.p + i1 * s.nb[1] + i2 * s.nb[2] + i3 * s.nb[3]) + b * 32;
    unsigned short * o = (unsigned short *) (d.p + row * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3] + (size_t) b * (Q4 ? 18 : 34));
    float v[32];
    #pragma unroll
    for (int j = 0; j < 32; ++j) v[j] = sr[j];
    if (!Q4) {
        float amax = 0.0f;
        #pragma unroll
        for (int j = 0; j < 32; ++j) amax = fmaxf(amax, fabsf(v[j]));
        // as the x86 quantize_row_q8_0 the CPU backend runs (arch/x86/quants.c:290-385): id = 127 / amax, round to nearest even
        const float dd = __fdiv_rn(amax, 127.0f), id = amax != 0.0f ? __fdiv_rn(127.0f, amax) : 0.0f;
        o[0] = __half_as_ushort(__float2half_rn(dd));
        #pragma unroll
        for (int j = 0; j < 16; ++j) {
            const int q0 = __float2int_rn(__fmul_rn(v[2 * j], id)), q1 = __float2int_rn(__fmul_rn(v[2 * j + 1], id));
            o[1 + j] = (unsigned short) ((q0 & 0xff) | ((q1 & 0xff) << 8));
        }
    } else {
        float amax = 0.0f, mx = 0.0f;
        #pragma unroll
        for (int j = 0; j < 32; ++j) if (amax < fabsf(v[j])) { amax = fabsf(v[j]); mx = v[j]; }
        const float dd = mx / -8.0f, id = dd != 0.0f ? 1.0f / dd : 0.0f;
        o[0] = __half_as_ushort(__float2half_rn(dd));
        #pragma unroll
        for (int j = 0; j < 8; ++j) {
            unsigned w = 0;
            #pragma unroll
            for (int k = 0; k < 2; ++k) {
                const int lo = min(15, (int) (signed char) (v[2 * j + k] * id + 8.5f)), hi = min(15, (int) (signed char) (v[2 * j + k + 16] * id + 8.5f));
                w |= (unsigned) ((lo & 0xf) | ((hi & 0xf) << 4)) << (8 * k);
            }
            o[1 + j] = (unsigned short) w;
        }
    }
}

// ---- GET_ROWS (f32 rows): dst[:, i10, i11, i12] = src0[:, ids[i10, i11, i12], i11, i12] -------------------------------------------------
__global__ void __launch_bounds__(256) k_get_rows_f32(const gl_view s, const gl_view idx, const gl_view d, int chunks) {
    GL_PDL();
    const int64_t r = blockIdx.x / chunks; const int64_t c0 = (int64_t) (blockIdx.x - r * chunks) * GL_CHUNK;
    int64_t i10, i11, i12; gl_unrow(r, d, i10, i11, i12);
    const int32_t row = *(const int32_t *) (idx.p + i10 * idx.nb[0] + i11 * idx.nb[1] + i12 * idx.nb[2]);
    const float * sr = (const float *) (s.p + (int64_t) row * s.nb[1] + i11 * s.nb[2] + i12 * s.nb[3]);
    float * dr = (float *) (d.p + i10 * d.nb[1] + i11 * d.nb[2] + i12 * d.nb[3]);
    float v[4];
    #pragma unroll
    for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < d.ne[0]) v[k] = sr[i]; }
    #pragma unroll
    for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < d.ne[0]) dr[i] = v[k]; }
}

// ---- CPY / CONT / DUP: same logical element order, either side strided; f32 / f16 in, f32 / f16 out -----------------------------------------
template <typename TS, typename TD>
__global__ void __launch_bounds__(256) k_cpy(const gl_view s, const gl_view d, int64_t n) {
    GL_PDL();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    // element i in ggml order (dim 0 fastest) of EACH tensor: the shapes may differ as long as the counts agree (ggml_cpy)
    int64_t r = i; const int64_t s0 = r % s.ne[0]; r /= s.ne[0]; const int64_t s1 = r % s.ne[1]; r /= s.ne[1]; const int64_t s2 = r % s.ne[2]; const int64_t s3 = r / s.ne[2];
    r = i;         const int64_t d0 = r % d.ne[0]; r /= d.ne[0]; const int64_t d1 = r % d.ne[1]; r /= d.ne[1]; const int64_t d2 = r % d.ne[2]; const int64_t d3 = r / d.ne[2];
    const TS v = *(const TS *) (s.p + s0 * s.nb[0] + s1 * s.nb[1] + s2 * s.nb[2] + s3 * s.nb[3]);
    *(TD *) (d.p + d0 * d.nb[0] + d1 * d.nb[1] + d2 * d.nb[2] + d3 * d.nb[3]) = (TD) v;
}

// ---- GLU, SwiGLU: dst = silu(a) * b, a / b either two tensors (split) or the halves of one row; silu(x) = x / (1 + exp(-x)) ------------------
static __device__ __forceinline__ float gl_swiglu1(float x, float y) { return b200_swiglu1(x, y); }      // (b200_common.cuh: the GEMV's fused quantizer computes the same)
template <bool VEC>
__global__ void __launch_bounds__(256) k_swiglu(const gl_view a, const gl_view b, const gl_view d, int64_t aoff, int64_t boff, int chunks) {
    GL_PDL();
    const int64_t row = blockIdx.x / chunks; const int64_t c0 = (int64_t) (blockIdx.x - row * chunks) * GL_CHUNK;
    int64_t i1, i2, i3; gl_unrow(row, d, i1, i2, i3);
    const float * ar = (const float *) (a.p + i1 * a.nb[1] + i2 * a.nb[2] + i3 * a.nb[3]) + aoff;
    const float * br = (const float *) (b.p + i1 * b.nb[1] + i2 * b.nb[2] + i3 * b.nb[3]) + boff;
    float *       dr = (float *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const int64_t n0 = d.ne[0];
    if (VEC) {
        const int64_t i = c0 + threadIdx.x * 4;
        if (i >= n0) return;
        const float4 x = *(const float4 *) (ar + i), y = *(const float4 *) (br + i);
        *(float4 *) (dr + i) = make_float4(gl_swiglu1(x.x, y.x), gl_swiglu1(x.y, y.y), gl_swiglu1(x.z, y.z), gl_swiglu1(x.w, y.w));
    } else {
        float x[4], y[4];
        #pragma unroll
        for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < n0) { x[k] = ar[i]; y[k] = br[i]; } }
        #pragma unroll
        for (int k = 0; k < 4; ++k) { const int64_t i = c0 + k * 256 + threadIdx.x; if (i < n0) dr[i] = gl_swiglu1(x[k], y[k]); }
    }
}

// ---- ROPE(Q) + ROPE(K) + SET_ROWS(K -> cache) + SET_ROWS(V -> cache): the four nodes between the QKV matmuls and the attention of a
// llama layer as one launch.  CTAs [0, nq): one Q head row each; [nq, nq + nk): one K head row each, whose result also goes (f16) to
// the cache row idx[token] that the SET_ROWS node behind it names; the rest: chunks of the V rows (as k_set_rows) -----------------------------
struct gl_rope_kv {
    gl_view qx, qy; gl_rope qr; int nq, qhb, qng;   // nq CTAs = tokens * qng head groups of qhb heads
    gl_view kx, ky; gl_rope kr; int nk, khb, kng;
    gl_view kidx, kd; int64_t k_ne0;              // SET_ROWS(K): its src is ky seen as rows of k_ne0 floats (k_ne0 % ky.ne[0] == 0)
    gl_view vs, vidx, vd; int vchunks;
};
__global__ void __launch_bounds__(128) k_rope_kv(const gl_rope_kv a) {
    GL_PDL();
    int b = blockIdx.x;
    if (b < a.nq) {
        const int64_t g = b % a.qng, t = b / a.qng, h0 = g * a.qhb, h1 = h0 + a.qhb < a.qx.ne[1] ? h0 + a.qhb : a.qx.ne[1];
        gl_rope_rows(a.qx, a.qy, a.qr, h0, h1, t % a.qx.ne[2], t / a.qx.ne[2], gl_no_cache());
        return;
    }
    b -= a.nq;
    if (b < a.nk) {
        const int64_t g = b % a.kng, t = b / a.kng, h0 = g * a.khb, h1 = h0 + a.khb < a.kx.ne[1] ? h0 + a.khb : a.kx.ne[1];
        auto cache = [&](int64_t h) {
            const int64_t e = (t * a.kx.ne[1] + h) * a.ky.ne[0];                 // first element of this head row in the contiguous result
            const int64_t r = e / a.k_ne0, c = e - r * a.k_ne0;
            const int64_t row = *(const int64_t *) (a.kidx.p + r * a.kidx.nb[0]);
            return (__half *) (a.kd.p + row * a.kd.nb[1]) + c;
        };
        gl_rope_rows(a.kx, a.ky, a.kr, h0, h1, t % a.kx.ne[2], t / a.kx.ne[2], cache);
        return;
    }
    b -= a.nk;
    {                                                                            // V: source row r, chunk of GL_CHUNK floats, two float4 per thread
        const int64_t r = b / a.vchunks; const int64_t c0 = (int64_t) (b - r * a.vchunks) * GL_CHUNK;
        const int64_t row = *(const int64_t *) (a.vidx.p + r * a.vidx.nb[0]);
        const float * sr = (const float *) (a.vs.p + r * a.vs.nb[1]);
        __half * dr = (__half *) (a.vd.p + row * a.vd.nb[1]);
        const int64_t n0 = a.vs.ne[0];
        float4 v[2];
        #pragma unroll
        for (int k = 0; k < 2; ++k) { const int64_t i = c0 + (k * 128 + threadIdx.x) * 4; if (i < n0) v[k] = *(const float4 *) (sr + i); }
        #pragma unroll
        for (int k = 0; k < 2; ++k) {
            const int64_t i = c0 + (k * 128 + threadIdx.x) * 4;
            if (i >= n0) continue;
            __half2 lo = __floats2half2_rn(v[k].x, v[k].y), hi = __floats2half2_rn(v[k].z, v[k].w); uint2 o; o.x = *(unsigned *) &lo; o.y = *(unsigned *) &hi;
            *(uint2 *) (dr + i) = o;
        }
    }
}

// ---- SOFT_MAX (ggml_soft_max_ext; ggml-cpu/ops.cpp:5160-5270): dst = softmax(x * scale + slope(head) * mask) over ne0, mask f32 / f16
// broadcast over heads and batch, ALiBi slopes when max_bias > 0, optional per-head sinks.  One CTA per row (what -fa 0 graphs need) -----
static __device__ __forceinline__ float block_max(float v, float * red) {
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    v = (int) threadIdx.x < nw ? red[threadIdx.x] : -INFINITY;
    if (threadIdx.x < 32) { for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o)); if (threadIdx.x == 0) red[0] = v; }
    __syncthreads();
    const float r = red[0];
    __syncthreads();
    return r;
}
template <bool MASK_F16>
__global__ void __launch_bounds__(256) k_soft_max(const gl_view x, const gl_view m, const gl_view d, const float * sinks, float scale, float max_bias, float m0, float m1, unsigned n_head_log2, int has_mask) {
    GL_PDL();
    int64_t i1, i2, i3; gl_unrow(blockIdx.x, x, i1, i2, i3);
    const float * xr = (const float *) (x.p + i1 * x.nb[1] + i2 * x.nb[2] + i3 * x.nb[3]);
    float *       dr = (float *) (d.p + i1 * d.nb[1] + i2 * d.nb[2] + i3 * d.nb[3]);
    const char *  mr = has_mask ? m.p + i1 * m.nb[1] + (i2 % m.ne[2]) * m.nb[2] + (i3 % m.ne[3]) * m.nb[3] : nullptr;
    const unsigned h = (unsigned) i2;
    const float slope = max_bias > 0.0f ? (h < n_head_log2 ? powf(m0, (float) (h + 1)) : powf(m1, (float) (2 * (h - n_head_log2) + 1))) : 1.0f;
    auto val = [&](int64_t i) {
        float v = xr[i] * scale;
        if (mr) v += slope * (MASK_F16 ? __half2float(((const __half *) mr)[i]) : ((const float *) mr)[i]);
        return v;
    };
    __shared__ float red[32];
    const int64_t n0 = x.ne[0];
    float mx = -INFINITY;
    for (int64_t i = threadIdx.x; i < n0; i += blockDim.x) mx = fmaxf(mx, val(i));
    mx = block_max(mx, red);
    if (sinks) mx = fmaxf(mx, sinks[i2]);
    float sum = 0.0f;
    for (int64_t i = threadIdx.x; i < n0; i += blockDim.x) { const float v = val(i); sum += v == -INFINITY ? 0.0f : expf(v - mx); }
    sum = block_sum(sum, red);
    if (sinks) sum += expf(sinks[i2] - mx);
    const float inv = 1.0f / sum;
    for (int64_t i = threadIdx.x; i < n0; i += blockDim.x) { const float v = val(i); dr[i] = v == -INFINITY ? 0.0f : expf(v - mx) * inv; }
}

// ---- UNARY: SILU (x / (1 + exp(-x)), ggml-cpu/vec.h), contiguous f32 ----------------------------------------------------------------------
__global__ void __launch_bounds__(256) k_silu(const float * x, float * y, int64_t n) {
    GL_PDL();
    const int64_t i = (int64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) { const float v = x[i]; y[i] = v / (1.0f + expf(-v)); }
}

template <typename... KArgs, typename... Args>
static cudaError_t gl_launch(b200_ctx * ctx, void (*kernel)(KArgs...), dim3 grid, unsigned block, Args... args) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = grid; cfg.blockDim = dim3(block); cfg.dynamicSmemBytes = 0; cfg.stream = ctx->stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr; cfg.numAttrs = ctx->pdl ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

// ---- which nodes we take ---------------------------------------------------------------------------------------------------------------
static bool gl_same_shape(const ggml_tensor * a, const ggml_tensor * b) { return a->ne[0] == b->ne[0] && a->ne[1] == b->ne[1] && a->ne[2] == b->ne[2] && a->ne[3] == b->ne[3]; }
static bool gl_can_repeat(const ggml_tensor * b, const ggml_tensor * a) {           // b broadcasts over a (ggml_can_repeat)
    for (int i = 0; i < 4; ++i) if (b->ne[i] <= 0 || a->ne[i] % b->ne[i] != 0) return false;
    return true;
}
static bool gl_grid_ok(const ggml_tensor * t) { return gl_nrows(t) * ((t->ne[0] + GL_CHUNK - 1) / GL_CHUNK) <= 0x7fffffff && t->ne[2] <= 65535 && t->ne[3] <= 65535; }

bool b200_glue_supports(const struct ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * b = op->src[1];
    switch ((int) op->op) {
        case GGML_OP_RMS_NORM:
            return a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && gl_contig_rows(a, 4) && gl_contig_rows(op, 4) && gl_grid_ok(op);
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV:
            return a->type == GGML_TYPE_F32 && b->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && gl_contig_rows(a, 4) && gl_contig_rows(b, 4) &&
                   gl_contig_rows(op, 4) && gl_same_shape(a, op) && gl_can_repeat(b, a) && gl_grid_ok(op);
        case GGML_OP_SCALE:
            return a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && gl_contiguous(a, 4) && gl_contiguous(op, 4);
        case GGML_OP_ARGSORT:
            return a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_I32 && gl_contig_rows(a, 4) && gl_contig_rows(op, 4) && a->ne[0] <= 1024 && gl_same_shape(a, op) && gl_nrows(op) <= 0x7fffffff;
        case GGML_OP_SUM_ROWS:
            return a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && gl_contig_rows(a, 4) && op->ne[0] == 1 && gl_nrows(a) == gl_nrows(op) && gl_nrows(op) <= 0x7fffffff;
        case GGML_OP_CLAMP:
            return a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && gl_contiguous(a, 4) && gl_contiguous(op, 4);
        case GGML_OP_UNARY:
            return op->op_params[0] == GGML_UNARY_OP_SILU && a->type == GGML_TYPE_F32 && op->type == GGML_TYPE_F32 && gl_contiguous(a, 4) && gl_contiguous(op, 4);
        case GGML_OP_SOFT_MAX: {
            const ggml_tensor * sk = op->src[2];
            if (a->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || !gl_contig_rows(a, 4) || !gl_contiguous(op, 4) || !gl_same_shape(a, op)) return false;
            if (b && ((b->type != GGML_TYPE_F32 && b->type != GGML_TYPE_F16) || b->nb[0] != gl_esize(b->type) || b->ne[0] != a->ne[0] || b->ne[1] < a->ne[1] ||
                      b->ne[2] <= 0 || b->ne[3] <= 0 || a->ne[2] % b->ne[2] != 0 || a->ne[3] % b->ne[3] != 0)) return false;
            if (sk && (sk->type != GGML_TYPE_F32 || sk->ne[0] != a->ne[2] || sk->nb[0] != 4)) return false;
            return gl_nrows(op) <= 0x7fffffff;
        }
        case GGML_OP_ROPE: {
            const int mode = op->op_params[2];
            if (mode != 0 && mode != 2) return false;                       // normal and neox; mrope / vision stay on the CPU backend
            if (a->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || b->type != GGML_TYPE_I32 || !gl_contig_rows(a, 4) || !gl_contig_rows(op, 4)) return false;
            if (op->src[2] && (op->src[2]->type != GGML_TYPE_F32 || op->src[2]->ne[0] < op->op_params[1] / 2)) return false;
            return a->ne[0] % 2 == 0 && op->op_params[1] % 2 == 0 && op->op_params[1] <= a->ne[0] && a->ne[1] <= 0x7fffffff && gl_grid_ok(op);
        }
        case GGML_OP_SET_ROWS:
            if (op->type == GGML_TYPE_Q8_0 || op->type == GGML_TYPE_Q4_0)      // the quantized KV cache
                return a->type == GGML_TYPE_F32 && b->type == GGML_TYPE_I64 && gl_contig_rows(a, 4) && a->ne[0] % 32 == 0 && a->ne[0] == op->ne[0] &&
                       op->nb[0] == (op->type == GGML_TYPE_Q8_0 ? 34u : 18u) && !(op->nb[1] & 1) && !(op->nb[2] & 1) && !(op->nb[3] & 1) && !((uintptr_t) op->data & 1) &&
                       gl_nelem(a) / 32 < ((int64_t) 1 << 38);
            return a->type == GGML_TYPE_F32 && b->type == GGML_TYPE_I64 && (op->type == GGML_TYPE_F16 || op->type == GGML_TYPE_F32) &&
                   gl_contig_rows(a, 4) && op->nb[0] == gl_esize(op->type) && gl_grid_ok(a);
        case GGML_OP_GET_ROWS:
            return a->type == GGML_TYPE_F32 && b->type == GGML_TYPE_I32 && op->type == GGML_TYPE_F32 && gl_contig_rows(a, 4) && gl_contig_rows(op, 4) && gl_grid_ok(op);
        case GGML_OP_CPY: case GGML_OP_CONT: case GGML_OP_DUP: {
            const ggml_tensor * d = op;                                          // (ggml_cpy(a, b): the node is a view of b)
            const bool ts = a->type == GGML_TYPE_F32 || a->type == GGML_TYPE_F16, td = d->type == GGML_TYPE_F32 || d->type == GGML_TYPE_F16;
            return ts && td && gl_nelem(a) == gl_nelem(d) && gl_nelem(a) < ((int64_t) 1 << 40);
        }
        case GGML_OP_GLU:
            if (op->op_params[0] != GGML_GLU_OP_SWIGLU || a->type != GGML_TYPE_F32 || op->type != GGML_TYPE_F32 || !gl_contig_rows(a, 4) || !gl_contig_rows(op, 4)) return false;
            if (b && (b->type != GGML_TYPE_F32 || !gl_contig_rows(b, 4) || !gl_same_shape(a, b))) return false;
            return gl_grid_ok(op);
        default: return false;
    }
}

// RMS_NORM node + the MUL node right behind it that scales it by a weight row: both outputs from one kernel.  Returns
// B200_ERR_UNSUPPORTED when the pair does not have that shape (the caller then runs the nodes one by one).
int b200_glue_rms_norm_mul(b200_ctx * ctx, struct ggml_tensor * norm, struct ggml_tensor * mul) {
    if (norm->op != GGML_OP_RMS_NORM || mul->op != GGML_OP_MUL || !b200_glue_supports(norm) || !b200_glue_supports(mul)) return B200_ERR_UNSUPPORTED;
    if (mul->src[0] != norm) return B200_ERR_UNSUPPORTED;
    const ggml_tensor * w = mul->src[1];
    if (w->ne[0] != norm->ne[0] || w == norm) return B200_ERR_UNSUPPORTED;
    // the graph allocator may give the MUL the norm's own memory (in place): then only the product is stored; any other overlap: no fusion
    const char * pa = (const char *) norm->data, * pb = (const char *) mul->data;
    const size_t na = (size_t) gl_nrows(norm) * norm->nb[1], nb = (size_t) gl_nrows(mul) * mul->nb[1];
    const bool same = pa == pb && norm->nb[1] == mul->nb[1] && norm->nb[2] == mul->nb[2] && norm->nb[3] == mul->nb[3];
    if (!same && pa < pb + nb && pb < pa + na) return B200_ERR_UNSUPPORTED;
    if (!gl_contiguous(norm, 4) && !same) return B200_ERR_UNSUPPORTED;
    float eps; memcpy(&eps, norm->op_params, 4);
    const ggml_tensor * x = norm->src[0];
    if (x->ne[0] <= 8192 && gl_vec_ok(x) && gl_vec_ok(norm) && gl_vec_ok(w) && gl_vec_ok(mul)) {
        const unsigned block = x->ne[0] <= 1024 ? 128 : x->ne[0] <= 2048 ? 256 : x->ne[0] <= 4096 ? 512 : 1024;
        B200_LAUNCH(ctx, B200_K_GLUE, gl_launch(ctx, k_rms_norm<true, true>, dim3((unsigned) gl_nrows(norm)), block, gl_of(x), gl_of(norm), gl_of(w), gl_of(mul), eps, same ? 0 : 1));
    } else B200_LAUNCH(ctx, B200_K_GLUE, gl_launch(ctx, k_rms_norm<true, false>, dim3((unsigned) gl_nrows(norm)), 256, gl_of(x), gl_of(norm), gl_of(w), gl_of(mul), eps, same ? 0 : 1));
    return B200_OK;
}

static gl_rope gl_rope_params(const ggml_tensor * op) {
    gl_rope r;
    float freq_base, beta_fast, beta_slow;
    r.n_dims = op->op_params[1]; r.neox = op->op_params[2] == 2;
    const int n_ctx_orig = op->op_params[4];
    memcpy(&freq_base, op->op_params + 5, 4);     memcpy(&r.freq_scale, op->op_params + 6, 4);
    memcpy(&r.ext_factor, op->op_params + 7, 4);  memcpy(&r.attn_factor, op->op_params + 8, 4);
    memcpy(&beta_fast, op->op_params + 9, 4);     memcpy(&beta_slow, op->op_params + 10, 4);
    r.theta_scale = powf(freq_base, -2.0f / r.n_dims);
    // ggml_rope_yarn_corr_dims (ggml.c:4255-4269)
    auto corr = [&](float n_rot) { return r.n_dims * logf(n_ctx_orig / (n_rot * 2.0f * (float) M_PI)) / (2.0f * logf(freq_base)); };
    r.corr0 = fmaxf(0.0f, floorf(corr(beta_fast))); r.corr1 = fminf((float) r.n_dims - 1.0f, ceilf(corr(beta_slow)));
    r.pos = (const int32_t *) op->src[1]->data; r.ff = op->src[2] ? (const float *) op->src[2]->data : nullptr;
    return r;
}

// byte spans of two strided tensors do not meet
static bool gl_span_disjoint(const ggml_tensor * a, const ggml_tensor * b) {
    const char * pa = (const char *) a->data, * pb = (const char *) b->data;
    size_t na = gl_esize(a->type), nb = gl_esize(b->type);
    for (int i = 0; i < 4; ++i) { na += (size_t) (a->ne[i] - 1) * a->nb[i]; nb += (size_t) (b->ne[i] - 1) * b->nb[i]; }
    return pa + na <= pb || pb + nb <= pa;
}
static bool gl_same_view(const ggml_tensor * a, const ggml_tensor * b) { return a->data == b->data && a->nb[1] == b->nb[1] && a->nb[2] == b->nb[2] && a->nb[3] == b->nb[3]; }

// [ROPE(Q)] + ROPE(K) + SET_ROWS(K) + SET_ROWS(V) as one launch (rq may be NULL); B200_ERR_UNSUPPORTED when the nodes do not have the
// shape of a llama layer's KV-cache update
// in_regs (the decode chain's tail): the three inputs never come from memory there, so an output may lie over another section's input -- the
// allocator hands ROPE(Q)'s result the block Kcur has just left -- and dead_in[i] reports which inputs (0 Q, 1 K, 2 V) are overwritten that way
static int gl_rope_kv_check(struct ggml_tensor * rq, struct ggml_tensor * rk, struct ggml_tensor * sk, struct ggml_tensor * sv, bool in_regs = false, bool * dead_in = nullptr) {
    if (!rk || !sk || !sv || rk->op != GGML_OP_ROPE || sk->op != GGML_OP_SET_ROWS || sv->op != GGML_OP_SET_ROWS || (rq && rq->op != GGML_OP_ROPE)) { b200_set_error("rope_kv: check 1"); return B200_ERR_UNSUPPORTED; }
    if (!b200_glue_supports(rk) || !b200_glue_supports(sk) || !b200_glue_supports(sv) || (rq && !b200_glue_supports(rq))) { b200_set_error("rope_kv: check 2"); return B200_ERR_UNSUPPORTED; }
    const ggml_tensor * ks = sk->src[0], * vs = sv->src[0], * kidx = sk->src[1], * vidx = sv->src[1];
    // SET_ROWS(K) stores exactly the ROPE(K) result, seen as 2-D rows; both caches f16, one index per source row
    if (ks->data != rk->data || !gl_contiguous(rk, 4) || !gl_contiguous(ks, 4) || gl_nelem(ks) != gl_nelem(rk) || ks->ne[2] != 1 || ks->ne[3] != 1) { b200_set_error("rope_kv: check 3"); return B200_ERR_UNSUPPORTED; }
    if (ks->ne[0] % rk->ne[0] != 0 || sk->type != GGML_TYPE_F16 || sv->type != GGML_TYPE_F16 || vs->ne[2] != 1 || vs->ne[3] != 1) { b200_set_error("rope_kv: check 4"); return B200_ERR_UNSUPPORTED; }
    if (kidx->ne[0] != ks->ne[1] || kidx->ne[1] != 1 || kidx->ne[2] != 1 || vidx->ne[0] != vs->ne[1] || vidx->ne[1] != 1 || vidx->ne[2] != 1) { b200_set_error("rope_kv: check 5"); return B200_ERR_UNSUPPORTED; }
    if (((uintptr_t) sk->data & 1) || sk->nb[1] % 2 || !gl_vec_ok(vs) || ((uintptr_t) sv->data & 15) || sv->nb[1] % 16) { b200_set_error("rope_kv: check 6"); return B200_ERR_UNSUPPORTED; }
    // independent sections: nothing one of them writes may be something another reads or writes (rope in place is fine)
    const ggml_tensor * outs[4] = { rq, rk, sk, sv };
    const ggml_tensor * ins[4]  = { rq ? rq->src[0] : nullptr, rk->src[0], nullptr, vs };
    for (int o = 0; o < 4; ++o) for (int i = 0; i < 4; ++i) {
        if (!outs[o] || o == i) continue;
        if (outs[i] && o < i && !gl_span_disjoint(outs[o], outs[i])) { b200_set_error("rope_kv: output %d (%s) overlaps output %d (%s)", o, outs[o]->name, i, outs[i]->name); return B200_ERR_UNSUPPORTED; }
        if (ins[i] && !gl_span_disjoint(outs[o], ins[i])) {
            if (in_regs && o < 2) { if (dead_in) dead_in[i == 3 ? 2 : i] = true; continue; }       // (a rope result over a dead matmul result; the cache tensors never)
            b200_set_error("rope_kv: output %d (%s) overlaps input %d (%s)", o, outs[o]->name, i, ins[i]->name); return B200_ERR_UNSUPPORTED;
        }
    }
    for (int o = 0; o < 4; ++o) if (outs[o] && (!gl_span_disjoint(outs[o], kidx) || !gl_span_disjoint(outs[o], vidx) || !gl_span_disjoint(outs[o], rk->src[1]))) { b200_set_error("rope_kv: check 9"); return B200_ERR_UNSUPPORTED; }
    return B200_OK;
}
// The same four nodes as the tail of a decode chain (b200_chain_rope x 3: Q, K, V): one token, adjacent-pair ROPE without YaRN mixing.  The
// chain kernel's row owners then rotate and store the rows themselves (b200_chain.cu); B200_ERR_UNSUPPORTED = keep them as a launch of their own
int b200_glue_rope_kv_plan(struct ggml_tensor * rq, struct ggml_tensor * rk, struct ggml_tensor * sk, struct ggml_tensor * sv, b200_chain_rope * out, bool * dead_in) {
    if (!rq) { b200_set_error("rope_kv plan: no ROPE(Q)"); return B200_ERR_UNSUPPORTED; }
    dead_in[0] = dead_in[1] = dead_in[2] = false;
    const int rc = gl_rope_kv_check(rq, rk, sk, sv, true, dead_in);
    if (rc != B200_OK) return rc;
    const ggml_tensor * ks = sk->src[0], * vs = sv->src[0], * kidx = sk->src[1], * vidx = sv->src[1];
    const ggml_tensor * ropes[2] = { rq, rk };
    for (int i = 0; i < 2; ++i) {
        const ggml_tensor * r = ropes[i];
        const gl_rope g = gl_rope_params(r);
        if (r->ne[2] != 1 || r->ne[3] != 1 || g.neox || r->op_params[2] != 0 || g.ext_factor != 0.0f || g.n_dims % 2 || g.n_dims > r->ne[0] || r->ne[0] % 2 ||
            !gl_contiguous(r, 4) || !gl_contiguous(r->src[0], 4) || r->src[1]->type != GGML_TYPE_I32) { b200_set_error("rope_kv plan: not a one-token adjacent-pair rope"); return B200_ERR_UNSUPPORTED; }
        b200_chain_rope & o = out[i];
        memset(&o, 0, sizeof o);
        o.kind = i == 0 ? B200_CHAIN_T_ROPE : B200_CHAIN_T_ROPE_CACHE;
        o.head_dim = (int32_t) r->ne[0]; o.n_dims = g.n_dims; o.freq_scale = g.freq_scale; o.theta_scale = g.theta_scale; o.attn_factor = g.attn_factor;
        o.pos = g.pos; o.freq_factors = g.ff; o.out = (float *) r->data;
    }
    if (ks->ne[1] != 1 || vs->ne[1] != 1 || kidx->type != GGML_TYPE_I64 || vidx->type != GGML_TYPE_I64) { b200_set_error("rope_kv plan: more than one cache row"); return B200_ERR_UNSUPPORTED; }
    out[1].cache_idx = (const int64_t *) kidx->data; out[1].cache = sk->data; out[1].cache_nb1 = sk->nb[1];
    memset(&out[2], 0, sizeof out[2]);
    out[2].kind = B200_CHAIN_T_CACHE; out[2].cache_idx = (const int64_t *) vidx->data; out[2].cache = sv->data; out[2].cache_nb1 = sv->nb[1];
    return B200_OK;
}

int b200_glue_rope_kv(b200_ctx * ctx, struct ggml_tensor * rq, struct ggml_tensor * rk, struct ggml_tensor * sk, struct ggml_tensor * sv) {
    const int rcc = gl_rope_kv_check(rq, rk, sk, sv);
    if (rcc != B200_OK) return rcc;
    const ggml_tensor * ks = sk->src[0], * vs = sv->src[0], * kidx = sk->src[1], * vidx = sv->src[1];
    gl_rope_kv a;
    memset(&a, 0, sizeof a);
    if (rq) {
        a.qx = gl_of(rq->src[0]); a.qy = gl_of(rq); a.qr = gl_rope_params(rq);
        a.qhb = gl_rope_heads_per_cta(rq); a.qng = (int) ((rq->ne[1] + a.qhb - 1) / a.qhb); a.nq = (int) (rq->ne[2] * rq->ne[3]) * a.qng;
    }
    a.kx = gl_of(rk->src[0]); a.ky = gl_of(rk); a.kr = gl_rope_params(rk);
    a.khb = gl_rope_heads_per_cta(rk); a.kng = (int) ((rk->ne[1] + a.khb - 1) / a.khb); a.nk = (int) (rk->ne[2] * rk->ne[3]) * a.kng;
    a.kidx = gl_of(kidx); a.kd = gl_of(sk); a.k_ne0 = ks->ne[0];
    a.vs = gl_of(vs); a.vidx = gl_of(vidx); a.vd = gl_of(sv); a.vchunks = (int) ((vs->ne[0] + GL_CHUNK - 1) / GL_CHUNK);
    if (gl_nrows(rk) > 0x3fffffff || (rq && gl_nrows(rq) > 0x3fffffff)) return B200_ERR_UNSUPPORTED;
    const int64_t grid = (int64_t) a.nq + a.nk + vs->ne[1] * a.vchunks;
    if (grid > 0x7fffffff) return B200_ERR_UNSUPPORTED;
    B200_LAUNCH(ctx, B200_K_GLUE, gl_launch(ctx, k_rope_kv, dim3((unsigned) grid), 128, a));
    return B200_OK;
}

// ADD + RMS_NORM + MUL (residual add, norm, norm weight) in one launch; B200_ERR_UNSUPPORTED when the triple does not fit
int b200_glue_add_rms_norm_mul(b200_ctx * ctx, struct ggml_tensor * add, struct ggml_tensor * norm, struct ggml_tensor * mul) {
    if (add->op != GGML_OP_ADD || norm->op != GGML_OP_RMS_NORM || mul->op != GGML_OP_MUL || norm->src[0] != add || mul->src[0] != norm) return B200_ERR_UNSUPPORTED;
    if (!b200_glue_supports(add) || !b200_glue_supports(norm) || !b200_glue_supports(mul)) return B200_ERR_UNSUPPORTED;
    const ggml_tensor * a = add->src[0], * b = add->src[1], * w = mul->src[1];
    if (!gl_same_shape(a, b) || w->ne[0] != norm->ne[0] || w == norm || w == add || add->ne[0] > 8192) return B200_ERR_UNSUPPORTED;
    const ggml_tensor * all[6] = { a, b, add, norm, mul, w };
    for (int i = 0; i < 6; ++i) if (!gl_vec_ok(all[i])) return B200_ERR_UNSUPPORTED;
    // every thread reads its elements of a, b, w before it writes the same elements of add / norm / mul: an output may coincide exactly
    // with a or b (in-place nodes) or be disjoint from the inputs, nothing in between; of the outputs only norm and mul may coincide
    const ggml_tensor * ins[3] = { a, b, w }, * outs[3] = { add, norm, mul };
    for (int o = 0; o < 3; ++o) for (int i = 0; i < 3; ++i) if (!gl_span_disjoint(outs[o], ins[i]) && !(i < 2 && gl_same_view(outs[o], ins[i]))) return B200_ERR_UNSUPPORTED;
    if (!gl_span_disjoint(add, norm) || !gl_span_disjoint(add, mul)) return B200_ERR_UNSUPPORTED;
    const bool same = gl_same_view(norm, mul);
    if (!same && !gl_span_disjoint(norm, mul)) return B200_ERR_UNSUPPORTED;
    float eps; memcpy(&eps, norm->op_params, 4);
    const unsigned block = add->ne[0] <= 1024 ? 128 : add->ne[0] <= 2048 ? 256 : add->ne[0] <= 4096 ? 512 : 1024;
    B200_LAUNCH(ctx, B200_K_GLUE, gl_launch(ctx, k_add_rms_norm_mul, dim3((unsigned) gl_nrows(add)), block, gl_of(a), gl_of(b), gl_of(add), gl_of(norm), gl_of(w), gl_of(mul), eps, same ? 0 : 1));
    return B200_OK;
}

// ---- dispatch -------------------------------------------------------------------------------------------------------------------------
#define GL_LAUNCH(...) B200_LAUNCH(ctx, B200_K_GLUE, __VA_ARGS__)
int b200_glue_compute(b200_ctx * ctx, struct ggml_tensor * op) {
    const ggml_tensor * a = op->src[0], * b = op->src[1];
    cudaStream_t st = ctx->stream;
    const unsigned rows = (unsigned) gl_nrows(op);
    const int chunks = (int) ((op->ne[0] + GL_CHUNK - 1) / GL_CHUNK);       // CTAs per row of the row-wise kernels
    switch ((int) op->op) {
        case GGML_OP_RMS_NORM: {
            float eps; memcpy(&eps, op->op_params, 4);
            if (a->ne[0] <= 8192 && gl_vec_ok(a) && gl_vec_ok(op)) {
                const unsigned block = a->ne[0] <= 1024 ? 128 : a->ne[0] <= 2048 ? 256 : a->ne[0] <= 4096 ? 512 : 1024;
                GL_LAUNCH(gl_launch(ctx, k_rms_norm<false, true>, dim3(rows), block, gl_of(a), gl_of(op), gl_of(op), gl_of(op), eps, 1));
            } else GL_LAUNCH(gl_launch(ctx, k_rms_norm<false, false>, dim3(rows), 256, gl_of(a), gl_of(op), gl_of(op), gl_of(op), eps, 1));
            return B200_OK;
        }
        case GGML_OP_ADD: case GGML_OP_SUB: case GGML_OP_MUL: case GGML_OP_DIV: {
            const bool vec = b->ne[0] == op->ne[0] && gl_vec_ok(a) && gl_vec_ok(b) && gl_vec_ok(op);
            const dim3 grid(rows * chunks);
            #define GL_BIN(OP) do { if (vec) GL_LAUNCH(gl_launch(ctx, k_binary<OP, true>, grid, 256, gl_of(a), gl_of(b), gl_of(op), chunks)); \
                                    else     GL_LAUNCH(gl_launch(ctx, k_binary<OP, false>, grid, 256, gl_of(a), gl_of(b), gl_of(op), chunks)); } while (0)
            if (op->op == GGML_OP_ADD) GL_BIN(0); else if (op->op == GGML_OP_SUB) GL_BIN(1); else if (op->op == GGML_OP_MUL) GL_BIN(2); else GL_BIN(3);
            #undef GL_BIN
            return B200_OK;
        }
        case GGML_OP_SCALE: {
            float s, bias; memcpy(&s, op->op_params, 4); memcpy(&bias, op->op_params + 1, 4);
            const int64_t n = gl_nelem(op);
            GL_LAUNCH(gl_launch(ctx, k_scale, dim3((unsigned) ((n + 255) / 256)), 256, (const float *) a->data, (float *) op->data, n, s, bias));
            return B200_OK;
        }
        case GGML_OP_ARGSORT:
            GL_LAUNCH(gl_launch(ctx, k_argsort, dim3(rows), (unsigned) ((a->ne[0] + 31) / 32 * 32), gl_of(a), gl_of(op), (int) (op->op_params[0] == 1)));     // GGML_SORT_ORDER_DESC = 1
            return B200_OK;
        case GGML_OP_SUM_ROWS:
            GL_LAUNCH(gl_launch(ctx, k_sum_rows, dim3((unsigned) ((gl_nrows(a) + 7) / 8)), 256, gl_of(a), gl_of(op), gl_nrows(a)));
            return B200_OK;
        case GGML_OP_CLAMP: {
            float lo, hi; memcpy(&lo, op->op_params, 4); memcpy(&hi, op->op_params + 1, 4);
            const int64_t n = gl_nelem(op);
            GL_LAUNCH(gl_launch(ctx, k_clamp, dim3((unsigned) ((n + 255) / 256)), 256, (const float *) a->data, (float *) op->data, n, lo, hi));
            return B200_OK;
        }
        case GGML_OP_UNARY: {
            const int64_t n = gl_nelem(op);
            GL_LAUNCH(gl_launch(ctx, k_silu, dim3((unsigned) ((n + 255) / 256)), 256, (const float *) a->data, (float *) op->data, n));
            return B200_OK;
        }
        case GGML_OP_SOFT_MAX: {
            float scale, max_bias; memcpy(&scale, op->op_params, 4); memcpy(&max_bias, op->op_params + 1, 4);
            const unsigned n_head = (unsigned) a->ne[2], n_head_log2 = 1u << (unsigned) floorf(log2f((float) n_head));
            const float m0 = powf(2.0f, -max_bias / n_head_log2), m1 = powf(2.0f, -(max_bias / 2.0f) / n_head_log2);
            const float * sinks = op->src[2] ? (const float *) op->src[2]->data : nullptr;
            const gl_view mv = b ? gl_of(b) : gl_of(a);
            if (b && b->type == GGML_TYPE_F16) GL_LAUNCH(gl_launch(ctx, k_soft_max<true>, dim3(rows), 256, gl_of(a), mv, gl_of(op), sinks, scale, max_bias, m0, m1, n_head_log2, 1));
            else                               GL_LAUNCH(gl_launch(ctx, k_soft_max<false>, dim3(rows), 256, gl_of(a), mv, gl_of(op), sinks, scale, max_bias, m0, m1, n_head_log2, b ? 1 : 0));
            return B200_OK;
        }
        case GGML_OP_ROPE: {
            const gl_rope r = gl_rope_params(op);
            const int hb = gl_rope_heads_per_cta(op);
            GL_LAUNCH(gl_launch(ctx, k_rope, dim3((unsigned) ((op->ne[1] + hb - 1) / hb), (unsigned) op->ne[2], (unsigned) op->ne[3]), 128, gl_of(a), gl_of(op), r, hb));
            return B200_OK;
        }
        case GGML_OP_SET_ROWS: {
            if (op->type == GGML_TYPE_Q8_0 || op->type == GGML_TYPE_Q4_0) {
                const int nblk = (int) (a->ne[0] / 32); const int64_t total = gl_nrows(a) * nblk;
                if (total == 0) return B200_OK;
                const dim3 qgrid((unsigned) ((total + 255) / 256));
                if (op->type == GGML_TYPE_Q8_0) GL_LAUNCH(gl_launch(ctx, k_set_rows_q<0>, qgrid, 256, gl_of(a), gl_of(b), gl_of(op), nblk, total));
                else                            GL_LAUNCH(gl_launch(ctx, k_set_rows_q<1>, qgrid, 256, gl_of(a), gl_of(b), gl_of(op), nblk, total));
                return B200_OK;
            }
            const unsigned srows = (unsigned) gl_nrows(a);
            const int sch = (int) ((a->ne[0] + GL_CHUNK - 1) / GL_CHUNK);
            const bool vec = gl_vec_ok(a) && ((uintptr_t) op->data & 15) == 0 && op->nb[1] % 16 == 0 && op->nb[2] % 16 == 0 && op->nb[3] % 16 == 0;
            const dim3 grid(srows * sch);
            if (op->type == GGML_TYPE_F16) { if (vec) GL_LAUNCH(gl_launch(ctx, k_set_rows<__half, true>, grid, 256, gl_of(a), gl_of(b), gl_of(op), sch));
                                             else     GL_LAUNCH(gl_launch(ctx, k_set_rows<__half, false>, grid, 256, gl_of(a), gl_of(b), gl_of(op), sch)); }
            else                           { if (vec) GL_LAUNCH(gl_launch(ctx, k_set_rows<float, true>, grid, 256, gl_of(a), gl_of(b), gl_of(op), sch));
                                             else     GL_LAUNCH(gl_launch(ctx, k_set_rows<float, false>, grid, 256, gl_of(a), gl_of(b), gl_of(op), sch)); }
            return B200_OK;
        }
        case GGML_OP_GET_ROWS:
            GL_LAUNCH(gl_launch(ctx, k_get_rows_f32, dim3(rows * chunks), 256, gl_of(a), gl_of(b), gl_of(op), chunks));
            return B200_OK;
        case GGML_OP_CPY: case GGML_OP_CONT: case GGML_OP_DUP: {
            const int64_t n = gl_nelem(a);
            const unsigned grid = (unsigned) ((n + 255) / 256);
            const bool sf = a->type == GGML_TYPE_F32, df = op->type == GGML_TYPE_F32;
            if (sf && df)       GL_LAUNCH(gl_launch(ctx, k_cpy<float, float>, dim3(grid), 256, gl_of(a), gl_of(op), n));
            else if (sf && !df) GL_LAUNCH(gl_launch(ctx, k_cpy<float, __half>, dim3(grid), 256, gl_of(a), gl_of(op), n));
            else if (!sf && df) GL_LAUNCH(gl_launch(ctx, k_cpy<__half, float>, dim3(grid), 256, gl_of(a), gl_of(op), n));
            else                GL_LAUNCH(gl_launch(ctx, k_cpy<__half, __half>, dim3(grid), 256, gl_of(a), gl_of(op), n));
            return B200_OK;
        }
        case GGML_OP_GLU: {
            const bool swapped = op->op_params[1] != 0;
            const int64_t nc = op->ne[0];
            const ggml_tensor * bb = b ? b : a;
            const int64_t aoff = b ? 0 : swapped ? nc : 0, boff = b ? 0 : swapped ? 0 : nc;
            const bool vec = nc % 4 == 0 && gl_vec_ok(a) && gl_vec_ok(bb) && gl_vec_ok(op);
            const dim3 grid(rows * chunks);
            if (vec) GL_LAUNCH(gl_launch(ctx, k_swiglu<true>, grid, 256, gl_of(a), gl_of(bb), gl_of(op), aoff, boff, chunks));
            else     GL_LAUNCH(gl_launch(ctx, k_swiglu<false>, grid, 256, gl_of(a), gl_of(bb), gl_of(op), aoff, boff, chunks));
            return B200_OK;
        }
        default:
            b200_set_error("glue: op %d not implemented", (int) op->op);
            return B200_ERR_UNSUPPORTED;
    }
}
