// b200_fattn.cuh -- argument block shared by the FLASH_ATTN_EXT kernels (semantics: ggml/src/ggml-cpu/ops.cpp:8047-8272).
#pragma once
#include "b200_common.cuh"
#include <math.h>

struct fattn_args {
    const char * q; size_t q_nb1, q_nb2, q_nb3;
    const char * k; size_t k_nb1, k_nb2, k_nb3;
    const char * v; size_t v_nb1, v_nb2, v_nb3;
    const char * mask; size_t m_nb1, m_nb2, m_nb3; int64_t m_ne2, m_ne3;
    const float * sinks;
    float * dst; size_t d_nb1, d_nb2, d_nb3;       // dst[DV, H, N, S]
    int DK, DV;
    int64_t N, H, KV;
    int64_t rk2, rk3, rv2, rv3;                    // broadcast ratios
    float scale, max_bias, softcap, m0, m1; uint32_t n_head_log2;
    const int32_t * pages; int page_shift;         // paged KV cache (b200_flash_attn_ext_paged): logical page -> physical page; NULL = contiguous
};


static inline void fattn_fill_args(fattn_args & p, const b200_tensor * q, const b200_tensor * k, const b200_tensor * v,
                                   const b200_tensor * mask, const b200_tensor * sinks, b200_tensor * dst,
                                   float scale, float max_bias, float softcap) {
    p.q = (const char *) q->data; p.q_nb1 = q->nb[1]; p.q_nb2 = q->nb[2]; p.q_nb3 = q->nb[3];
    p.k = (const char *) k->data; p.k_nb1 = k->nb[1]; p.k_nb2 = k->nb[2]; p.k_nb3 = k->nb[3];
    p.v = (const char *) v->data; p.v_nb1 = v->nb[1]; p.v_nb2 = v->nb[2]; p.v_nb3 = v->nb[3];
    if (mask) { p.mask = (const char *) mask->data; p.m_nb1 = mask->nb[1]; p.m_nb2 = mask->nb[2]; p.m_nb3 = mask->nb[3]; p.m_ne2 = mask->ne[2]; p.m_ne3 = mask->ne[3]; }
    else      { p.mask = nullptr; p.m_nb1 = p.m_nb2 = p.m_nb3 = 0; p.m_ne2 = p.m_ne3 = 1; }
    p.sinks = sinks ? (const float *) sinks->data : nullptr;
    p.dst = (float *) dst->data; p.d_nb1 = dst->nb[1]; p.d_nb2 = dst->nb[2]; p.d_nb3 = dst->nb[3];
    p.DK = (int) k->ne[0]; p.DV = (int) v->ne[0];
    p.N = q->ne[1]; p.H = q->ne[2]; p.KV = k->ne[1];
    p.rk2 = q->ne[2] / k->ne[2]; p.rk3 = q->ne[3] / k->ne[3];
    p.rv2 = q->ne[2] / v->ne[2]; p.rv3 = q->ne[3] / v->ne[3];
    if (softcap != 0.0f) scale /= softcap;                  // ops.cpp:8107-8109
    p.scale = scale; p.max_bias = max_bias; p.softcap = softcap;
    p.n_head_log2 = 1u << (uint32_t) floorf(log2f((float) p.H));
    p.m0 = powf(2.0f, -(max_bias) / p.n_head_log2);
    p.m1 = powf(2.0f, -(max_bias / 2.0f) / p.n_head_log2);
    p.pages = nullptr; p.page_shift = 0;
}
static inline void fattn_set_pages(fattn_args & p, const b200_kv_pages * pg) {
    if (!pg) return;
    p.pages = pg->table; p.page_shift = 0;
    while ((1 << p.page_shift) < pg->page_tokens) ++p.page_shift;
}

// pool row of logical key ic (identity without a page table)
static __device__ __forceinline__ int64_t fattn_row(const fattn_args & p, int64_t ic) {
    if (!p.pages) return ic;
    return ((int64_t) p.pages[ic >> p.page_shift] << p.page_shift) | (ic & (((int64_t) 1 << p.page_shift) - 1));
}

// ALiBi slope of head h (ops.cpp:8138-8139)
static __device__ __forceinline__ float fattn_slope(const fattn_args & p, uint32_t h) {
    return p.max_bias > 0.0f ? (h < p.n_head_log2 ? powf(p.m0, (float) (h + 1)) : powf(p.m1, (float) (2 * (h - p.n_head_log2) + 1))) : 1.0f;
}
