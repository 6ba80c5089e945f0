// ref_driver.c -- TEST INFRASTRUCTURE ONLY.  A thin C driver over the UNMODIFIED reference library
// (oracle/_ref/<isa>/libggml-ref.so, compiled in place from /root/reference/ggml by oracle/Makefile):
// it builds one-node ggml graphs and runs them on the reference's own ggml-cpu backend, so tests and
// bench.py's cpu_baseline / --impl reference leg can call the reference's hot path through flat pointers
// (ctypes).  All tensors are contiguous.  Nothing of the reference is copied here; only its public API is used
// (ggml/include/ggml.h, ggml-cpu.h).
#include "ggml.h"
#include "ggml-cpu.h"
#include <string.h>
#include <stdlib.h>
#include <stdint.h>

void quantize_row_q8_1_ref(const float * x, void * y, int64_t k);    // ggml/src/ggml-quants.h (exported, GGML_API)

static int g_init = 0;
static void ensure_init(void) { if (!g_init) { ggml_cpu_init(); ggml_time_init(); g_init = 1; } }

// ggml-cpu's own from_float for `type` (q8_0, q8_1, q8_K ...): ggml-cpu.c:204-391
int refd_from_float(int type, const float * x, void * y, int64_t k) {
    ensure_init();
    const struct ggml_type_traits_cpu * t = ggml_get_type_traits_cpu((enum ggml_type) type);
    if (!t || !t->from_float) return -1;
    t->from_float(x, y, k);
    return 0;
}
void refd_quantize_row_q8_1_ref(const float * x, void * y, int64_t k) { quantize_row_q8_1_ref(x, y, k); }

// deterministic weight quantizer used by llama-quantize and test-backend-ops (ggml.c ggml_quantize_chunk)
size_t refd_quantize_chunk(int type, const float * src, void * dst, int64_t nrows, int64_t n_per_row) {
    ensure_init();
    return ggml_quantize_chunk((enum ggml_type) type, src, dst, 0, nrows, n_per_row, NULL);
}
int refd_dequantize_row(int type, const void * x, float * y, int64_t k) {
    ensure_init();
    const struct ggml_type_traits * t = ggml_get_type_traits((enum ggml_type) type);
    if (!t || !t->to_float) return -1;
    t->to_float(x, y, k);
    return 0;
}
size_t refd_row_size(int type, int64_t k) { return ggml_row_size((enum ggml_type) type, k); }

static struct ggml_context * make_ctx(size_t data_bytes) {
    struct ggml_init_params ip = { data_bytes + (64u << 20), NULL, false };
    return ggml_init(ip);
}

// runs the graph `reps` times, returns the best wall time of one run in microseconds (or -1)
static double run_graph(struct ggml_context * ctx, struct ggml_tensor * out, int n_threads, int reps) {
    struct ggml_cgraph * gf = ggml_new_graph(ctx);
    ggml_build_forward_expand(gf, out);
    struct ggml_cplan plan = ggml_graph_plan(gf, n_threads, NULL);
    uint8_t * work = plan.work_size ? (uint8_t *) malloc(plan.work_size) : NULL;
    plan.work_data = work;
    double best = -1.0;
    for (int r = 0; r < (reps > 0 ? reps : 1); ++r) {
        const int64_t t0 = ggml_time_us();
        if (ggml_graph_compute(gf, &plan) != GGML_STATUS_SUCCESS) { best = -1.0; break; }
        const double dt = (double) (ggml_time_us() - t0);
        if (best < 0 || dt < best) best = dt;
    }
    free(work);
    return best;
}

// dst[M, N, ne12, ne13] = w[K, M, ne02, ne03]^T x[K, N, ne12, ne13]; returns best-of-reps microseconds, <0 on error
double refd_mul_mat(int type, const void * w, int64_t K, int64_t M, int64_t ne02, int64_t ne03,
                    const float * x, int64_t N, int64_t ne12, int64_t ne13, float * dst, int n_threads, int reps) {
    ensure_init();
    const size_t wbytes = ggml_row_size((enum ggml_type) type, K) * M * ne02 * ne03;
    const size_t xbytes = (size_t) K * N * ne12 * ne13 * 4, dbytes = (size_t) M * N * ne12 * ne13 * 4;
    struct ggml_context * ctx = make_ctx(wbytes + xbytes + dbytes + (size_t) K * N * ne12 * ne13 * 2 + (size_t) n_threads * 4096);
    if (!ctx) return -1.0;
    struct ggml_tensor * tw = ggml_new_tensor_4d(ctx, (enum ggml_type) type, K, M, ne02, ne03);
    struct ggml_tensor * tx = ggml_new_tensor_4d(ctx, GGML_TYPE_F32, K, N, ne12, ne13);
    memcpy(tw->data, w, wbytes); memcpy(tx->data, x, xbytes);
    struct ggml_tensor * out = ggml_mul_mat(ctx, tw, tx);
    const double us = run_graph(ctx, out, n_threads, reps);
    if (us >= 0 && dst) memcpy(dst, out->data, dbytes);
    ggml_free(ctx);
    return us;
}

// as[K, M, n_expert], b[K, ne11, n_tokens], ids[n_used, n_tokens] -> dst[M, n_used, n_tokens]
double refd_mul_mat_id(int type, const void * as, int64_t K, int64_t M, int64_t n_expert,
                       const float * b, int64_t ne11, int64_t n_tokens, const int32_t * ids, int64_t n_used,
                       float * dst, int n_threads, int reps) {
    ensure_init();
    const size_t wbytes = ggml_row_size((enum ggml_type) type, K) * M * n_expert;
    const size_t bbytes = (size_t) K * ne11 * n_tokens * 4, dbytes = (size_t) M * n_used * n_tokens * 4;
    struct ggml_context * ctx = make_ctx(wbytes + bbytes + dbytes + (size_t) K * n_used * n_tokens * 4 + (size_t) n_expert * n_used * n_tokens * 16 + (1u << 20));
    if (!ctx) return -1.0;
    struct ggml_tensor * tas = ggml_new_tensor_3d(ctx, (enum ggml_type) type, K, M, n_expert);
    struct ggml_tensor * tb  = ggml_new_tensor_3d(ctx, GGML_TYPE_F32, K, ne11, n_tokens);
    struct ggml_tensor * tid = ggml_new_tensor_2d(ctx, GGML_TYPE_I32, n_used, n_tokens);
    memcpy(tas->data, as, wbytes); memcpy(tb->data, b, bbytes); memcpy(tid->data, ids, (size_t) n_used * n_tokens * 4);
    struct ggml_tensor * out = ggml_mul_mat_id(ctx, tas, tb, tid);
    const double us = run_graph(ctx, out, n_threads, reps);
    if (us >= 0 && dst) memcpy(dst, out->data, dbytes);
    ggml_free(ctx);
    return us;
}

// q[DK, N, H, S] f32; k[DK, KV, Hk, Sk], v[DV, KV, Hk, Sk] of kv_type; mask[KV, Nm, Hm, Sm] f16 or NULL;
// sinks[H] or NULL; dst[DV, H, N, S].  prec: 0 default, 10 = GGML_PREC_F32.
double refd_flash_attn_ext(const float * q, const void * k, const void * v, int kv_type, const void * mask, const float * sinks,
                           float * dst, int64_t DK, int64_t DV, int64_t N, int64_t H, int64_t S,
                           int64_t KV, int64_t Hk, int64_t Sk, int64_t Nm, int64_t Hm, int64_t Sm,
                           float scale, float max_bias, float logit_softcap, int prec, int n_threads, int reps) {
    ensure_init();
    const size_t qb = (size_t) DK * N * H * S * 4;
    const size_t kb = ggml_row_size((enum ggml_type) kv_type, DK) * KV * Hk * Sk, vb = ggml_row_size((enum ggml_type) kv_type, DV) * KV * Hk * Sk;
    const size_t mb = mask ? (size_t) KV * Nm * Hm * Sm * 2 : 0, db = (size_t) DV * H * N * S * 4;
    struct ggml_context * ctx = make_ctx(qb + kb + vb + mb + db + (size_t) n_threads * (DK + 2 * DV + 64) * 4 + (1u << 20));
    if (!ctx) return -1.0;
    struct ggml_tensor * tq = ggml_new_tensor_4d(ctx, GGML_TYPE_F32, DK, N, H, S);
    struct ggml_tensor * tk = ggml_new_tensor_4d(ctx, (enum ggml_type) kv_type, DK, KV, Hk, Sk);
    struct ggml_tensor * tv = ggml_new_tensor_4d(ctx, (enum ggml_type) kv_type, DV, KV, Hk, Sk);
    struct ggml_tensor * tm = mask ? ggml_new_tensor_4d(ctx, GGML_TYPE_F16, KV, Nm, Hm, Sm) : NULL;
    memcpy(tq->data, q, qb); memcpy(tk->data, k, kb); memcpy(tv->data, v, vb);
    if (tm) memcpy(tm->data, mask, mb);
    struct ggml_tensor * out = ggml_flash_attn_ext(ctx, tq, tk, tv, tm, scale, max_bias, logit_softcap);
    if (prec) ggml_flash_attn_ext_set_prec(out, (enum ggml_prec) prec);
    if (sinks) {
        struct ggml_tensor * ts = ggml_new_tensor_1d(ctx, GGML_TYPE_F32, H);
        memcpy(ts->data, sinks, (size_t) H * 4);
        ggml_flash_attn_ext_add_sinks(out, ts);
    }
    const double us = run_graph(ctx, out, n_threads, reps);
    if (us >= 0 && dst) memcpy(dst, out->data, db);
    ggml_free(ctx);
    return us;
}


// ---- persistent single-op graphs, for the timed CPU baseline (bench.py): weights are set once, then the op is
// ---- re-run on new activations without re-copying anything ------------------------------------------------
typedef struct {
    struct ggml_context * ctx; struct ggml_cgraph * gf; struct ggml_cplan plan; uint8_t * work;
    struct ggml_tensor * in[5]; struct ggml_tensor * out;
} refd_op;

static refd_op * op_finish(refd_op * op, int n_threads) {
    op->gf = ggml_new_graph(op->ctx);
    ggml_build_forward_expand(op->gf, op->out);
    op->plan = ggml_graph_plan(op->gf, n_threads, NULL);
    op->work = op->plan.work_size ? (uint8_t *) malloc(op->plan.work_size) : NULL;
    op->plan.work_data = op->work;
    return op;
}
void * refd_mm_create(int type, int64_t K, int64_t M, int64_t N, int n_threads) {
    ensure_init();
    refd_op * op = (refd_op *) calloc(1, sizeof(refd_op));
    op->ctx = make_ctx(ggml_row_size((enum ggml_type) type, K) * M + (size_t) (K + M) * N * 4);
    if (!op->ctx) { free(op); return NULL; }
    op->in[0] = ggml_new_tensor_2d(op->ctx, (enum ggml_type) type, K, M);
    op->in[1] = ggml_new_tensor_2d(op->ctx, GGML_TYPE_F32, K, N);
    memset(op->in[1]->data, 0, ggml_nbytes(op->in[1]));
    op->out = ggml_mul_mat(op->ctx, op->in[0], op->in[1]);
    return op_finish(op, n_threads);
}
// q[DK,N,H,1] f32, k/v[DK,KV,Hk,1] f16, mask[KV,Nm,1,1] f16 (Nm = 0: no mask)
void * refd_fa_create(int64_t DK, int64_t DV, int64_t N, int64_t H, int64_t KV, int64_t Hk, int64_t Nm, float scale, int n_threads) {
    ensure_init();
    refd_op * op = (refd_op *) calloc(1, sizeof(refd_op));
    op->ctx = make_ctx((size_t) (DK + DV) * N * H * 4 + (size_t) (DK + DV) * KV * Hk * 2 + (size_t) KV * Nm * 2);
    if (!op->ctx) { free(op); return NULL; }
    op->in[0] = ggml_new_tensor_4d(op->ctx, GGML_TYPE_F32, DK, N, H, 1);
    op->in[1] = ggml_new_tensor_4d(op->ctx, GGML_TYPE_F16, DK, KV, Hk, 1);
    op->in[2] = ggml_new_tensor_4d(op->ctx, GGML_TYPE_F16, DV, KV, Hk, 1);
    op->in[3] = Nm > 0 ? ggml_new_tensor_4d(op->ctx, GGML_TYPE_F16, KV, Nm, 1, 1) : NULL;
    for (int i = 0; i < 4; ++i) if (op->in[i]) memset(op->in[i]->data, 0, ggml_nbytes(op->in[i]));
    op->out = ggml_flash_attn_ext(op->ctx, op->in[0], op->in[1], op->in[2], op->in[3], scale, 0.0f, 0.0f);
    return op_finish(op, n_threads);
}
void * refd_op_input(void * h, int i, size_t * nbytes) {
    refd_op * op = (refd_op *) h;
    if (i < 0 || i > 4 || !op->in[i]) return NULL;
    if (nbytes) *nbytes = ggml_nbytes(op->in[i]);
    return op->in[i]->data;
}
void * refd_op_output(void * h, size_t * nbytes) {
    refd_op * op = (refd_op *) h;
    if (nbytes) *nbytes = ggml_nbytes(op->out);
    return op->out->data;
}
// one execution; returns microseconds or -1
double refd_op_run(void * h) {
    refd_op * op = (refd_op *) h;
    const int64_t t0 = ggml_time_us();
    if (ggml_graph_compute(op->gf, &op->plan) != GGML_STATUS_SUCCESS) return -1.0;
    return (double) (ggml_time_us() - t0);
}
void refd_op_free(void * h) {
    refd_op * op = (refd_op *) h;
    if (!op) return;
    free(op->work); ggml_free(op->ctx); free(op);
}
