// upload_probe.cpp -- drives the plugin the way llama_model_loader does without mmap (src/llama-model-loader.cpp:1136-1186): a weight
// tensor of a USAGE_WEIGHTS buffer uploaded with ggml_backend_tensor_set_async in ascending chunks.  Checks that the chunked upload
// ends in the device layout (load-time repack), that get_tensor hands back the GGUF bytes, and that MUL_MAT over the chunk-uploaded
// weight is bit-identical to MUL_MAT over the same weight uploaded in one write.  Built only from include/ggml-b200-abi.h (dlopen of
// the plugin, no ggml library): usage  upload_probe <libggml-b200.so> <type: 8|2|14> <K> <M> <chunk bytes>
#include "ggml-b200-abi.h"
#include <dlfcn.h>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <cstdint>
#include <vector>

static size_t tsize(int t) { return t == GGML_TYPE_Q6_K ? 210 : t == GGML_TYPE_Q8_0 ? 34 : t == GGML_TYPE_Q4_0 ? 18 : t == GGML_TYPE_Q4_K ? 144 : 4; }
static int    tblck(int t) { return t == GGML_TYPE_Q6_K || t == GGML_TYPE_Q4_K ? 256 : t == GGML_TYPE_F32 ? 1 : 32; }

static ggml_tensor make(int type, int64_t ne0, int64_t ne1, ggml_backend_buffer_t buf, void * data, const char * name) {
    ggml_tensor t; memset(&t, 0, sizeof t);
    t.type = (ggml_type) type; t.buffer = buf; t.data = data; t.op = GGML_OP_NONE;
    t.ne[0] = ne0; t.ne[1] = ne1; t.ne[2] = t.ne[3] = 1;
    t.nb[0] = tsize(type); t.nb[1] = (size_t) (ne0 / tblck(type)) * tsize(type); t.nb[2] = t.nb[1] * ne1; t.nb[3] = t.nb[2];
    snprintf(t.name, sizeof t.name, "%s", name);
    return t;
}
#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "upload_probe: FAILED %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

int main(int argc, char ** argv) {
    if (argc < 6) { fprintf(stderr, "usage: %s plugin type K M chunk\n", argv[0]); return 2; }
    const int type = atoi(argv[2]); const int64_t K = atoll(argv[3]), M = atoll(argv[4]); const size_t chunk = (size_t) atoll(argv[5]);
    void * h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    auto init = (ggml_backend_reg_t (*)()) dlsym(h, "ggml_backend_init");
    REQUIRE(init);
    ggml_backend_reg_t reg = init();
    REQUIRE(reg && reg->iface.get_device_count(reg) >= 1);
    ggml_backend_dev_t dev = reg->iface.get_device(reg, 0);
    auto is_repacked = (int (*)(const ggml_tensor *)) reg->iface.get_proc_address(reg, "ggml_backend_b200_tensor_is_repacked");
    REQUIRE(is_repacked);
    ggml_backend_dev_props props; dev->iface.get_props(dev, &props);
    ggml_backend_t be = dev->iface.init_backend(dev, nullptr);
    REQUIRE(be);
    ggml_backend_buffer_type_t buft = dev->iface.get_buffer_type(dev);

    const size_t wbytes = (size_t) (K / tblck(type)) * tsize(type) * M;
    const size_t walloc = wbytes + 4096;
    ggml_backend_buffer_t wbuf = buft->iface.alloc_buffer(buft, 2 * walloc + 256);
    ggml_backend_buffer_t abuf = buft->iface.alloc_buffer(buft, (size_t) (K + 3 * M) * 4 + 2048);
    REQUIRE(wbuf && abuf);
    wbuf->usage = GGML_BACKEND_BUFFER_USAGE_WEIGHTS;                 // as ggml_backend_buffer_set_usage does before the upload
    char * wb = (char *) wbuf->iface.get_base(wbuf), * ab = (char *) abuf->iface.get_base(abuf);
    ggml_tensor w1 = make(type, K, M, wbuf, wb, "w_chunked"), w2 = make(type, K, M, wbuf, wb + (walloc + 255) / 256 * 256, "w_whole");
    wbuf->iface.init_tensor(wbuf, &w1); wbuf->iface.init_tensor(wbuf, &w2);

    // random blocks with small finite fp16 scales
    std::vector<uint8_t> host(wbytes); uint32_t s = 12345;
    for (auto & b : host) { s = s * 1664525u + 1013904223u; b = (uint8_t) (s >> 24); }
    const size_t ts = tsize(type), doff = type == GGML_TYPE_Q6_K ? 208 : 0;
    for (size_t i = 0; i < wbytes / ts; ++i) {
        host[i * ts + doff] = (uint8_t) (i * 37); host[i * ts + doff + 1] = (uint8_t) (0x20 + i % 8);                 // fp16 ~ 2^-7 .. 2^-5
        if (type == GGML_TYPE_Q4_K) { host[i * ts + 2] = (uint8_t) (i * 11); host[i * ts + 3] = (uint8_t) (0x1c + i % 4); }
    }
    std::vector<float> x(K); for (int64_t i = 0; i < K; ++i) { s = s * 1664525u + 1013904223u; x[i] = ((int) (s >> 8) % 2001 - 1000) / 500.0f; }

    // (1) chunked, ascending, through the backend's async path, one event per chunk as the loader does
    ggml_backend_event_t ev = props.caps.events ? dev->iface.event_new(dev) : nullptr;
    size_t nchunks = 0;
    for (size_t off = 0; off < wbytes; off += chunk, ++nchunks) {
        const size_t n = wbytes - off < chunk ? wbytes - off : chunk;
        be->iface.set_tensor_async(be, &w1, host.data() + off, off, n);
        if (ev) be->iface.event_record(be, ev);
    }
    if (ev) dev->iface.event_synchronize(dev, ev);
    be->iface.synchronize(be);
    // (2) one whole-tensor write
    wbuf->iface.set_tensor(wbuf, &w2, host.data(), 0, wbytes);
    const int r1 = is_repacked(&w1), r2 = is_repacked(&w2);
    printf("upload_probe: type %d K %lld M %lld: %zu chunks of %zu bytes; repacked chunked=%d whole=%d\n", type, (long long) K, (long long) M, nchunks, chunk, r1, r2);
    REQUIRE(r1 == r2);                                               // the chunked upload ends in the same layout as the whole write

    // (3) the host sees GGUF bytes either way
    std::vector<uint8_t> back(wbytes);
    wbuf->iface.get_tensor(wbuf, &w1, back.data(), 0, wbytes);
    REQUIRE(memcmp(back.data(), host.data(), wbytes) == 0);
    be->iface.get_tensor_async(be, &w2, back.data(), 0, wbytes); be->iface.synchronize(be);
    REQUIRE(memcmp(back.data(), host.data(), wbytes) == 0);
    REQUIRE(is_repacked(&w1) == r1);

    // (4) MUL_MAT over both: bit-identical
    ggml_tensor xa = make(GGML_TYPE_F32, K, 1, abuf, ab, "x");
    ggml_tensor y1 = make(GGML_TYPE_F32, M, 1, abuf, ab + (K * 4 + 255) / 256 * 256, "y1");
    ggml_tensor y2 = make(GGML_TYPE_F32, M, 1, abuf, (char *) y1.data + (M * 4 + 255) / 256 * 256, "y2");
    abuf->iface.set_tensor(abuf, &xa, x.data(), 0, K * 4);
    y1.op = y2.op = GGML_OP_MUL_MAT; y1.src[0] = &w1; y1.src[1] = &xa; y2.src[0] = &w2; y2.src[1] = &xa;
    REQUIRE(dev->iface.supports_op(dev, &y1));
    ggml_tensor * nodes[2] = { &y1, &y2 };
    ggml_cgraph g; memset(&g, 0, sizeof g); g.size = 2; g.n_nodes = 2; g.nodes = nodes;
    REQUIRE(be->iface.graph_compute(be, &g) == GGML_STATUS_SUCCESS);
    be->iface.synchronize(be);
    std::vector<float> o1(M), o2(M);
    abuf->iface.get_tensor(abuf, &y1, o1.data(), 0, M * 4); abuf->iface.get_tensor(abuf, &y2, o2.data(), 0, M * 4);
    REQUIRE(memcmp(o1.data(), o2.data(), M * 4) == 0);
    // (5) the same two-node graph again with new activations each time: the plugin records it into a CUDA graph on its second
    // sight and replays it afterwards (be_graph_compute); a one-node graph is never recorded and serves as the check
    ggml_tensor y3 = make(GGML_TYPE_F32, M, 1, abuf, (char *) y2.data + (M * 4 + 255) / 256 * 256, "y3");
    y3.op = GGML_OP_MUL_MAT; y3.src[0] = &w2; y3.src[1] = &xa;
    ggml_tensor * one[1] = { &y3 };
    ggml_cgraph g1; memset(&g1, 0, sizeof g1); g1.size = 1; g1.n_nodes = 1; g1.nodes = one;
    for (int it = 0; it < 5; ++it) {
        for (int64_t i = 0; i < K; ++i) { s = s * 1664525u + 1013904223u; x[i] = ((int) (s >> 8) % 2001 - 1000) / 250.0f; }
        be->iface.set_tensor_async(be, &xa, x.data(), 0, K * 4);
        REQUIRE(be->iface.graph_compute(be, &g) == GGML_STATUS_SUCCESS);
        REQUIRE(be->iface.graph_compute(be, &g1) == GGML_STATUS_SUCCESS);
        be->iface.synchronize(be);
        std::vector<float> o3(M);
        abuf->iface.get_tensor(abuf, &y1, o1.data(), 0, M * 4); abuf->iface.get_tensor(abuf, &y2, o2.data(), 0, M * 4); abuf->iface.get_tensor(abuf, &y3, o3.data(), 0, M * 4);
        REQUIRE(memcmp(o1.data(), o3.data(), M * 4) == 0);
        REQUIRE(memcmp(o2.data(), o3.data(), M * 4) == 0);
    }
    double sum = 0; for (float v : o1) sum += v;
    REQUIRE(sum == sum);
    printf("upload_probe: OK (sum of outputs %.6g)\n", sum);
    if (ev) dev->iface.event_free(dev, ev);
    abuf->iface.free_buffer(abuf); wbuf->iface.free_buffer(wbuf);
    be->iface.free(be);
    return 0;
}
