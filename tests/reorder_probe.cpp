// reorder_probe.cpp -- host-only check of the plugin's graph_optimize (no GPU needed: the function only looks at the node list).
// Builds the node list of llama layers as llama.cpp emits them (Q = wq x, RESHAPE, ROPE(Q), V = wv x, RESHAPE, K = wk x, RESHAPE, ROPE(K),
// SET_ROWS x 2, ..., up = wu h, gate = wg h, GLU, down) with fabricated tensors, calls ggml_backend_b200_graph_optimize through
// get_proc_address and prints the new order; tests/test_abi.py checks adjacency, type grouping and that every node still comes after
// its sources.  Also: a matmul whose weight is NOT a plain weights-buffer leaf (a KV-cache view) must stay where it is.
// usage: reorder_probe <libggml-b200.so> <wv type: 12 | 14>
#include "ggml-b200-abi.h"
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>
#include <utility>

static std::vector<ggml_tensor *> all;
static ggml_tensor * T(const char * name, int type, int op, ggml_tensor * s0 = nullptr, ggml_tensor * s1 = nullptr, ggml_backend_buffer * buf = nullptr) {
    ggml_tensor * t = new ggml_tensor; memset(t, 0, sizeof *t);
    t->type = (ggml_type) type; t->op = (ggml_op) op; t->src[0] = s0; t->src[1] = s1; t->buffer = buf;
    t->ne[0] = 64; t->ne[1] = t->ne[2] = t->ne[3] = 1; t->nb[0] = 4; t->nb[1] = t->nb[2] = t->nb[3] = 256;
    if (op == GGML_OP_RESHAPE || op == GGML_OP_VIEW) t->view_src = s0 && s0->view_src ? s0->view_src : s0;
    snprintf(t->name, sizeof t->name, "%s", name);
    all.push_back(t);
    return t;
}

int main(int argc, char ** argv) {
    if (argc < 3) return 2;
    const int vtype = atoi(argv[2]);
    void * h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    auto init = (ggml_backend_reg_t (*)()) dlsym(h, "ggml_backend_init");
    if (!init) return 2;
    ggml_backend_reg_t reg = init();
    auto optimize = (void (*)(ggml_cgraph *)) reg->iface.get_proc_address(reg, "ggml_backend_b200_graph_optimize");
    if (!optimize) { fprintf(stderr, "no ggml_backend_b200_graph_optimize\n"); return 1; }
    ggml_backend_buffer wbuf; memset(&wbuf, 0, sizeof wbuf); wbuf.usage = GGML_BACKEND_BUFFER_USAGE_WEIGHTS;
    ggml_backend_buffer abuf; memset(&abuf, 0, sizeof abuf); abuf.usage = GGML_BACKEND_BUFFER_USAGE_COMPUTE;
    std::vector<ggml_tensor *> nodes;
    ggml_tensor * pos = T("pos", GGML_TYPE_I32, GGML_OP_NONE, nullptr, nullptr, &abuf), * idx = T("idx", GGML_TYPE_I64, GGML_OP_NONE, nullptr, nullptr, &abuf);
    ggml_tensor * kcache = T("cache_k", GGML_TYPE_F16, GGML_OP_NONE, nullptr, nullptr, &abuf), * vcache = T("cache_v", GGML_TYPE_F16, GGML_OP_NONE, nullptr, nullptr, &abuf);
    ggml_tensor * inp = T("inp", GGML_TYPE_F32, GGML_OP_NONE, nullptr, nullptr, &abuf);
    ggml_tensor * cur = inp;
    for (int il = 0; il < 2; ++il) {
        char n[64];
        auto W = [&](const char * base, int type) { snprintf(n, sizeof n, "%s%d", base, il); return T(n, type, GGML_OP_NONE, nullptr, nullptr, &wbuf); };
        auto N = [&](const char * base, int op, ggml_tensor * s0, ggml_tensor * s1 = nullptr) { snprintf(n, sizeof n, "%s%d", base, il); ggml_tensor * t = T(n, GGML_TYPE_F32, op, s0, s1, &abuf); nodes.push_back(t); return t; };
        ggml_tensor * x = N("norm", GGML_OP_RMS_NORM, cur);
        ggml_tensor * q = N("Q", GGML_OP_MUL_MAT, W("wq", GGML_TYPE_Q4_K), x), * q3 = N("Q3_", GGML_OP_RESHAPE, q), * qr = N("Qrope", GGML_OP_ROPE, q3, pos);
        ggml_tensor * v = N("V", GGML_OP_MUL_MAT, W("wv", il == 1 ? vtype : GGML_TYPE_Q4_K), x), * v3 = N("V3_", GGML_OP_RESHAPE, v);
        ggml_tensor * k = N("K", GGML_OP_MUL_MAT, W("wk", GGML_TYPE_Q4_K), x), * k3 = N("K3_", GGML_OP_RESHAPE, k), * kr = N("Krope", GGML_OP_ROPE, k3, pos);
        ggml_tensor * sk = N("setK", GGML_OP_SET_ROWS, kr, idx), * sv = N("setV", GGML_OP_SET_ROWS, v3, idx);
        sk->view_src = kcache; sv->view_src = vcache;
        // attention without FLASH_ATTN_EXT: kq = K-cache view x Q -- same src[1] position as a weight matmul would have, but src[0] is not a weight
        ggml_tensor * kview = N("kview", GGML_OP_VIEW, kcache); kview->src[1] = sk;
        ggml_tensor * kq = N("KQ", GGML_OP_MUL_MAT, kview, qr);
        ggml_tensor * o = N("O", GGML_OP_MUL_MAT, W("wo", GGML_TYPE_Q4_K), kq);
        ggml_tensor * hh = N("ffn_inp", GGML_OP_ADD, o, cur), * hn = N("ffn_norm", GGML_OP_RMS_NORM, hh);
        ggml_tensor * up = N("up", GGML_OP_MUL_MAT, W("wu", GGML_TYPE_Q4_K), hn), * ga = N("gate", GGML_OP_MUL_MAT, W("wg", GGML_TYPE_Q4_K), hn);
        ggml_tensor * gl = N("glu", GGML_OP_GLU, ga, up);
        ggml_tensor * dn = N("down", GGML_OP_MUL_MAT, W("wd", il == 1 ? vtype : GGML_TYPE_Q4_K), gl);
        cur = N("l_out", GGML_OP_ADD, dn, hh);
    }
    ggml_cgraph g; memset(&g, 0, sizeof g); g.size = g.n_nodes = (int) nodes.size(); g.nodes = nodes.data();
    optimize(&g);
    printf("order:");
    for (int i = 0; i < g.n_nodes; ++i) printf(" %s", g.nodes[i]->name);
    printf("\n");
    for (int i = 0; i < g.n_nodes; ++i) for (int s = 0; s < GGML_MAX_SRC; ++s) {
        const ggml_tensor * sr = g.nodes[i]->src[s];
        if (!sr || sr->op == GGML_OP_NONE) continue;
        bool before = false; for (int j = 0; j < i; ++j) before |= g.nodes[j] == sr;
        if (!before) { printf("BROKEN: %s runs before its source %s\n", g.nodes[i]->name, sr->name); return 1; }
    }
    printf("dependencies: OK\n");

    // ---- the key of the plugin's CUDA-graph cache: same graph -> same key; any change of a pointer, shape, stride, op or op_param -> new key;
    // prefill-sized graphs are not eligible for recording
    auto identity = (int (*)(const ggml_cgraph *, uint64_t *, uint64_t *)) reg->iface.get_proc_address(reg, "ggml_backend_b200_graph_identity");
    if (!identity) { fprintf(stderr, "no ggml_backend_b200_graph_identity\n"); return 1; }
    for (ggml_tensor * t : all) t->buffer = nullptr;                 // (fabricated buffers have no buffer type to inspect)
    static char mem[1 << 16]; size_t off = 0;
    for (ggml_tensor * t : all) if (!t->view_src) { t->data = mem + off; off += 256; }
    for (ggml_tensor * t : all) if (t->view_src) t->data = t->view_src->data;
    uint64_t a1, a2, b1, b2;
    const int e0 = identity(&g, &a1, &a2), e1 = identity(&g, &b1, &b2);
    printf("identity: eligible=%d stable=%d\n", e0, e1 == e0 && a1 == b1 && a2 == b2);
    int changed = 0, tried = 0;
    auto differs = [&]() { uint64_t c1, c2; identity(&g, &c1, &c2); ++tried; if (c1 != a1 || c2 != a2) ++changed; };
    ggml_tensor * n5 = nullptr;                                        // a node that launches work (views are keyed by pointer and shape only)
    for (int i = 0; i < g.n_nodes && !n5; ++i) if (g.nodes[i]->op == GGML_OP_ROPE) n5 = g.nodes[i];
    { void * d = n5->data; n5->data = (char *) d + 64; differs(); n5->data = d; }                        // a node moved
    { int64_t v = n5->ne[1]; n5->ne[1] = 2; differs(); n5->ne[1] = v; }                                    // shape
    { size_t v = n5->nb[1]; n5->nb[1] = 512; differs(); n5->nb[1] = v; }                                   // stride
    { int32_t v = n5->op_params[3]; n5->op_params[3] = 77; differs(); n5->op_params[3] = v; }              // op parameter
    { ggml_op v = g.nodes[1]->op; g.nodes[1]->op = GGML_OP_ADD; differs(); g.nodes[1]->op = v; }           // operator
    { ggml_tensor * src = g.nodes[1]->src[0]; void * d = src->data; src->data = (char *) d + 128; differs(); src->data = d; }   // a source (weight) moved
    { ggml_tensor * t = g.nodes[1]->src[1]; g.nodes[1]->src[1] = nullptr; differs(); g.nodes[1]->src[1] = t; }               // a source dropped
    { std::swap(g.nodes[1], g.nodes[2]); differs(); std::swap(g.nodes[1], g.nodes[2]); }                   // node order
    { --g.n_nodes; differs(); ++g.n_nodes; }                                                               // node count
    uint64_t z1, z2; identity(&g, &z1, &z2);
    printf("identity: %d of %d mutations changed the key, restored=%d\n", changed, tried, z1 == a1 && z2 == a2);
    for (int i = 0; i < g.n_nodes; ++i) if (g.nodes[i]->op == GGML_OP_MUL_MAT) { g.nodes[i]->ne[1] = 512; break; }          // a prefill-sized matmul
    printf("identity: prefill eligible=%d\n", identity(&g, &z1, &z2));
    return 0;
}
