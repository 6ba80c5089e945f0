// fusion_probe.cpp -- the start of a llama layer as llama.cpp hands it to a backend (src/models/llama.cpp, build_attn with a KV cache):
//   ADD(residual) RMS_NORM MUL(norm weight)  Q = wq x  RESHAPE ROPE(Q)  V = wv x  RESHAPE  K = wk x  RESHAPE ROPE(K)  SET_ROWS(K) SET_ROWS(V)
//   gate = wg x  up = wu x  GLU(SWIGLU)  down = wd glu  ADD(residual)           (the FFN, fed from the same norm to keep the probe small)
// driven through the plugin ABI only (include/ggml-b200-abi.h, dlopen): graph_optimize, then graph_compute five times with new inputs,
// positions and cache rows each time (first run launch by launch, second recorded into a CUDA graph, then replays).  Every output of
// every round is appended to <out file>; tests/test_gpu_backend_ops.py runs the probe with the plugin's fusions on and with
// GGML_B200_NO_FUSE=1 GGML_B200_GRAPHS=0 and requires the two files to be identical bit for bit.
// usage: fusion_probe <libggml-b200.so> <out file> <wv type: 12 (q4_K) | 14 (q6_K)>
#include "ggml-b200-abi.h"
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <vector>

static size_t tsize(int t) { return t == GGML_TYPE_Q6_K ? 210 : t == GGML_TYPE_Q4_K ? 144 : t == GGML_TYPE_F16 ? 2 : t == GGML_TYPE_I64 ? 8 : 4; }
static int    tblck(int t) { return t == GGML_TYPE_Q6_K || t == GGML_TYPE_Q4_K ? 256 : 1; }
#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "fusion_probe: FAILED %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

struct arena { ggml_backend_buffer_t buf; char * base; size_t used, size; };
static ggml_tensor * mk(std::vector<ggml_tensor *> & all, arena & a, int type, int64_t n0, int64_t n1, int64_t n2, const char * name) {
    ggml_tensor * t = new ggml_tensor; memset(t, 0, sizeof *t);
    t->type = (ggml_type) type; t->buffer = a.buf; t->op = GGML_OP_NONE;
    t->ne[0] = n0; t->ne[1] = n1; t->ne[2] = n2; t->ne[3] = 1;
    t->nb[0] = tsize(type); t->nb[1] = (size_t) (n0 / tblck(type)) * tsize(type); t->nb[2] = t->nb[1] * n1; t->nb[3] = t->nb[2] * n2;
    t->data = a.base + a.used; a.used += (t->nb[3] + 4096 + 255) / 256 * 256;
    if (a.used > a.size) { fprintf(stderr, "fusion_probe: arena too small\n"); exit(2); }
    snprintf(t->name, sizeof t->name, "%s", name);
    a.buf->iface.init_tensor(a.buf, t);
    all.push_back(t);
    return t;
}
static ggml_tensor * view(std::vector<ggml_tensor *> & all, ggml_tensor * src, int64_t n0, int64_t n1, int64_t n2, const char * name) {   // contiguous reshape
    ggml_tensor * t = new ggml_tensor; memset(t, 0, sizeof *t);
    t->type = src->type; t->buffer = src->buffer; t->op = GGML_OP_RESHAPE; t->src[0] = src; t->view_src = src->view_src ? src->view_src : src; t->data = src->data;
    t->ne[0] = n0; t->ne[1] = n1; t->ne[2] = n2; t->ne[3] = 1;
    t->nb[0] = tsize(src->type); t->nb[1] = t->nb[0] * n0; t->nb[2] = t->nb[1] * n1; t->nb[3] = t->nb[2] * n2;
    snprintf(t->name, sizeof t->name, "%s", name);
    all.push_back(t);
    return t;
}
static void rope_params(ggml_tensor * t, int n_dims, int mode) {
    const float freq_base = 500000.0f, freq_scale = 1.0f, ext = 0.0f, attn = 1.0f, bf = 32.0f, bs = 1.0f;
    t->op_params[1] = n_dims; t->op_params[2] = mode; t->op_params[4] = 8192;
    memcpy(t->op_params + 5, &freq_base, 4); memcpy(t->op_params + 6, &freq_scale, 4); memcpy(t->op_params + 7, &ext, 4);
    memcpy(t->op_params + 8, &attn, 4); memcpy(t->op_params + 9, &bf, 4); memcpy(t->op_params + 10, &bs, 4);
}

int main(int argc, char ** argv) {
    if (argc < 4) { fprintf(stderr, "usage: %s plugin out wv_type\n", argv[0]); return 2; }
    const int vtype = atoi(argv[3]);
    void * h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    auto init = (ggml_backend_reg_t (*)()) dlsym(h, "ggml_backend_init");
    REQUIRE(init);
    ggml_backend_reg_t reg = init();
    REQUIRE(reg && reg->iface.get_device_count(reg) >= 1);
    ggml_backend_dev_t dev = reg->iface.get_device(reg, 0);
    ggml_backend_t be = dev->iface.init_backend(dev, nullptr);
    REQUIRE(be);
    ggml_backend_buffer_type_t buft = dev->iface.get_buffer_type(dev);
    const int64_t E = 4096, HD = 128, NH = 32, NKV = 8, KV = NKV * HD, NCTX = 512, F = 14336;
    arena wa, ca;
    wa.size = 192u << 20; wa.buf = buft->iface.alloc_buffer(buft, wa.size); ca.size = 16u << 20; ca.buf = buft->iface.alloc_buffer(buft, ca.size);
    REQUIRE(wa.buf && ca.buf);
    wa.base = (char *) wa.buf->iface.get_base(wa.buf); wa.used = 0; ca.base = (char *) ca.buf->iface.get_base(ca.buf); ca.used = 0;
    wa.buf->usage = GGML_BACKEND_BUFFER_USAGE_WEIGHTS;
    std::vector<ggml_tensor *> all;
    // weights
    ggml_tensor * wq = mk(all, wa, GGML_TYPE_Q4_K, E, E, 1, "wq"), * wk = mk(all, wa, GGML_TYPE_Q4_K, E, KV, 1, "wk"), * wv = mk(all, wa, vtype, E, KV, 1, "wv");
    ggml_tensor * wg = mk(all, wa, GGML_TYPE_Q4_K, E, F, 1, "wg"), * wu = mk(all, wa, GGML_TYPE_Q4_K, E, F, 1, "wu"), * wd = mk(all, wa, vtype, F, E, 1, "wd");
    ggml_tensor * nw = mk(all, wa, GGML_TYPE_F32, E, 1, 1, "norm_w");
    uint32_t s = 4242;
    for (ggml_tensor * w : { wq, wk, wv, wg, wu, wd }) {
        const size_t ts = tsize(w->type), n = w->nb[2];
        std::vector<uint8_t> host(n);
        for (auto & b : host) { s = s * 1664525u + 1013904223u; b = (uint8_t) (s >> 24); }
        const size_t doff = w->type == GGML_TYPE_Q6_K ? 208 : 0;
        for (size_t i = 0; i < n / ts; ++i) {
            host[i * ts + doff] = (uint8_t) (i * 37); host[i * ts + doff + 1] = (uint8_t) (0x18 + i % 4);
            if (w->type == GGML_TYPE_Q4_K) { host[i * ts + 2] = (uint8_t) (i * 11); host[i * ts + 3] = (uint8_t) (0x14 + i % 4); }
        }
        wa.buf->iface.set_tensor(wa.buf, w, host.data(), 0, n);
    }
    std::vector<float> f(E);
    for (auto & v : f) { s = s * 1664525u + 1013904223u; v = 0.5f + (s >> 8) % 1000 / 1000.0f; }
    wa.buf->iface.set_tensor(wa.buf, nw, f.data(), 0, E * 4);
    // inputs, caches
    ggml_tensor * ra = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "resid_a"), * rb = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "resid_b");
    ggml_tensor * pos = mk(all, ca, GGML_TYPE_I32, 1, 1, 1, "pos"), * kidx = mk(all, ca, GGML_TYPE_I64, 1, 1, 1, "k_idxs"), * vidx = mk(all, ca, GGML_TYPE_I64, 1, 1, 1, "v_idxs");
    ggml_tensor * kc = mk(all, ca, GGML_TYPE_F16, KV, NCTX, 1, "cache_k"), * vc = mk(all, ca, GGML_TYPE_F16, KV, NCTX, 1, "cache_v");
    // nodes, in llama.cpp's order
    ggml_tensor * add = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "ffn_inp");   add->op = GGML_OP_ADD; add->src[0] = ra; add->src[1] = rb;
    ggml_tensor * nrm = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "norm");      nrm->op = GGML_OP_RMS_NORM; nrm->src[0] = add; { float eps = 1e-5f; memcpy(nrm->op_params, &eps, 4); }
    ggml_tensor * x   = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "attn_norm"); x->op = GGML_OP_MUL; x->src[0] = nrm; x->src[1] = nw;
    ggml_tensor * q   = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "Qcur");      q->op = GGML_OP_MUL_MAT; q->src[0] = wq; q->src[1] = x;
    ggml_tensor * q3  = view(all, q, HD, NH, 1, "Qcur3");
    ggml_tensor * qr  = mk(all, ca, GGML_TYPE_F32, HD, NH, 1, "Qrope");   qr->op = GGML_OP_ROPE; qr->src[0] = q3; qr->src[1] = pos; rope_params(qr, (int) HD, 0);
    ggml_tensor * v   = mk(all, ca, GGML_TYPE_F32, KV, 1, 1, "Vcur");     v->op = GGML_OP_MUL_MAT; v->src[0] = wv; v->src[1] = x;
    ggml_tensor * v3  = view(all, v, HD, NKV, 1, "Vcur3");
    ggml_tensor * k   = mk(all, ca, GGML_TYPE_F32, KV, 1, 1, "Kcur");     k->op = GGML_OP_MUL_MAT; k->src[0] = wk; k->src[1] = x;
    ggml_tensor * k3  = view(all, k, HD, NKV, 1, "Kcur3");
    ggml_tensor * kr  = mk(all, ca, GGML_TYPE_F32, HD, NKV, 1, "Krope");  kr->op = GGML_OP_ROPE; kr->src[0] = k3; kr->src[1] = pos; rope_params(kr, (int) HD, 0);
    ggml_tensor * k2  = view(all, kr, KV, 1, 1, "Krope2d");
    ggml_tensor * v2  = view(all, v3, KV, 1, 1, "Vcur2d");
    ggml_tensor * sk  = view(all, kc, KV, NCTX, 1, "set_k"); sk->op = GGML_OP_SET_ROWS; sk->src[0] = k2; sk->src[1] = kidx; sk->src[2] = nullptr;
    ggml_tensor * sv  = view(all, vc, KV, NCTX, 1, "set_v"); sv->op = GGML_OP_SET_ROWS; sv->src[0] = v2; sv->src[1] = vidx;
    sk->view_src = kc; sv->view_src = vc;
    ggml_tensor * ga  = mk(all, ca, GGML_TYPE_F32, F, 1, 1, "ffn_gate"); ga->op = GGML_OP_MUL_MAT; ga->src[0] = wg; ga->src[1] = x;
    ggml_tensor * up  = mk(all, ca, GGML_TYPE_F32, F, 1, 1, "ffn_up");   up->op = GGML_OP_MUL_MAT; up->src[0] = wu; up->src[1] = x;
    ggml_tensor * glu = mk(all, ca, GGML_TYPE_F32, F, 1, 1, "ffn_swiglu"); glu->op = GGML_OP_GLU; glu->src[0] = ga; glu->src[1] = up; glu->op_params[0] = GGML_GLU_OP_SWIGLU;
    ggml_tensor * dn  = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "ffn_out");  dn->op = GGML_OP_MUL_MAT; dn->src[0] = wd; dn->src[1] = glu;
    ggml_tensor * res = mk(all, ca, GGML_TYPE_F32, E, 1, 1, "l_out");    res->op = GGML_OP_ADD; res->src[0] = dn; res->src[1] = add;
    ggml_tensor * nodes[] = { add, nrm, x, q, q3, qr, v, v3, k, k3, kr, k2, v2, sk, sv, ga, up, glu, dn, res };
    const int nn = (int) (sizeof nodes / sizeof nodes[0]);
    for (int i = 0; i < nn; ++i) if (nodes[i]->op != GGML_OP_RESHAPE) REQUIRE(dev->iface.supports_op(dev, nodes[i]));
    ggml_cgraph g; memset(&g, 0, sizeof g); g.size = nn; g.n_nodes = nn; g.nodes = nodes;
    if (be->iface.graph_optimize) be->iface.graph_optimize(be, &g);
    printf("fusion_probe: node order after graph_optimize:");
    for (int i = 0; i < nn; ++i) printf(" %s", nodes[i]->name);
    printf("\n");
    // dependencies still hold
    for (int i = 0; i < nn; ++i) for (int sidx = 0; sidx < GGML_MAX_SRC; ++sidx) {
        const ggml_tensor * sr = nodes[i]->src[sidx];
        if (!sr || sr->op == GGML_OP_NONE) continue;
        bool before = false; for (int j = 0; j < i; ++j) before |= nodes[j] == sr;
        REQUIRE(before);
    }
    FILE * out = fopen(argv[2], "wb");
    REQUIRE(out);
    std::vector<uint8_t> zero(kc->nb[2], 0);
    ca.buf->iface.set_tensor(ca.buf, kc, zero.data(), 0, kc->nb[2]); ca.buf->iface.set_tensor(ca.buf, vc, zero.data(), 0, vc->nb[2]);
    for (int it = 0; it < 5; ++it) {
        std::vector<float> a(E), b(E);
        for (int64_t i = 0; i < E; ++i) { s = s * 1664525u + 1013904223u; a[i] = ((int) (s >> 8) % 2001 - 1000) / 400.0f; s = s * 1664525u + 1013904223u; b[i] = ((int) (s >> 8) % 2001 - 1000) / 700.0f; }
        const int32_t p = 17 + 61 * it; const int64_t row_k = (5 + 97 * it) % NCTX, row_v = (11 + 31 * it) % NCTX;
        be->iface.set_tensor_async(be, ra, a.data(), 0, E * 4); be->iface.set_tensor_async(be, rb, b.data(), 0, E * 4);
        be->iface.set_tensor_async(be, pos, &p, 0, 4); be->iface.set_tensor_async(be, kidx, &row_k, 0, 8); be->iface.set_tensor_async(be, vidx, &row_v, 0, 8);
        REQUIRE(be->iface.graph_compute(be, &g) == GGML_STATUS_SUCCESS);
        be->iface.synchronize(be);
        for (ggml_tensor * t : { add, x, q, k, v, qr, kr, kc, vc, ga, up, glu, dn, res }) {
            std::vector<uint8_t> hb(t->nb[3]);
            ca.buf->iface.get_tensor(ca.buf, t, hb.data(), 0, hb.size());
            fwrite(hb.data(), 1, hb.size(), out);
            if (it == 4 && (t == kc || t == vc)) {                              // the five rows written, nothing else
                int nz_rows = 0;
                for (int64_t r = 0; r < NCTX; ++r) { bool nz = false; for (size_t i = 0; i < t->nb[1]; ++i) nz |= hb[r * t->nb[1] + i] != 0; nz_rows += nz; }
                REQUIRE(nz_rows == 5);
            }
        }
    }
    fclose(out);
    printf("fusion_probe: OK\n");
    be->iface.free(be);
    return 0;
}
