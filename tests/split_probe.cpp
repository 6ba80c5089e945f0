// split_probe.cpp -- drives the plugin's split-mode=row buffer type the way llama_model_loader and ggml_backend_sched do
// (src/llama-model.cpp:393-411 asks the registry for "ggml_backend_split_buffer_type"; the CUDA backend's answer is ggml-cuda.cu:777-1000):
// a weight of that buffer type is written whole, read back (GGUF bytes, whatever layout the slices have on the devices), and used as src0
// of MUL_MAT nodes computed by EVERY device's backend in turn as the main device, for a decode column and a prefill batch; the results
// must equal (NMSE < 1e-10; the count of differing words is printed) the same MUL_MAT over the weight in a plain device buffer.  Built only from include/ggml-b200-abi.h (dlopen of
// the plugin): usage  split_probe <libggml-b200.so> <type: 12|14|8|2> <K> <M> [tensor_split weights ...]
// Run with GGML_B200_FAKE_DEVICES=N on a one-GPU box.
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
#define REQUIRE(c) do { if (!(c)) { fprintf(stderr, "split_probe: FAILED %s (line %d)\n", #c, __LINE__); return 1; } } while (0)

int main(int argc, char ** argv) {
    if (argc < 5) { fprintf(stderr, "usage: %s plugin type K M [split ...]\n", argv[0]); return 2; }
    const int type = atoi(argv[2]); const int64_t K = atoll(argv[3]), M = atoll(argv[4]);
    void * h = dlopen(argv[1], RTLD_NOW | RTLD_LOCAL);
    if (!h) { fprintf(stderr, "dlopen: %s\n", dlerror()); return 2; }
    auto init = (ggml_backend_reg_t (*)()) dlsym(h, "ggml_backend_init");
    REQUIRE(init);
    ggml_backend_reg_t reg = init();
    const int ndev = (int) reg->iface.get_device_count(reg);
    REQUIRE(ndev >= 2);
    auto split_fn = (ggml_backend_split_buffer_type_t) reg->iface.get_proc_address(reg, "ggml_backend_split_buffer_type");
    REQUIRE(split_fn);
    std::vector<float> ts(16, 0.0f);
    for (int i = 5; i < argc && i - 5 < 16; ++i) ts[i - 5] = (float) atof(argv[i]);

    const size_t wbytes = (size_t) (K / tblck(type)) * tsize(type) * M;
    std::vector<uint8_t> host(wbytes); uint32_t s = 4242;
    for (auto & b : host) { s = s * 1664525u + 1013904223u; b = (uint8_t) (s >> 24); }
    const size_t tsz = tsize(type), doff = type == GGML_TYPE_Q6_K ? 208 : 0;
    for (size_t i = 0; i < wbytes / tsz; ++i) {
        host[i * tsz + doff] = (uint8_t) (i * 37); host[i * tsz + doff + 1] = (uint8_t) (0x20 + i % 8);
        if (type == GGML_TYPE_Q4_K) { host[i * tsz + 2] = (uint8_t) (i * 11); host[i * tsz + 3] = (uint8_t) (0x1c + i % 4); }
    }
    const int64_t NB = 96;                                    // the prefill batch
    std::vector<float> x(K * NB); for (auto & v : x) { s = s * 1664525u + 1013904223u; v = ((int) (s >> 8) % 2001 - 1000) / 500.0f; }

    for (int main_dev = 0; main_dev < ndev; ++main_dev) {
        ggml_backend_dev_t dev = reg->iface.get_device(reg, main_dev);
        ggml_backend_buffer_type_t sbuft = split_fn(main_dev, ts.data());
        REQUIRE(sbuft && sbuft == split_fn(main_dev, ts.data()));                 // one object per (main device, split)
        REQUIRE(sbuft->device == dev && dev->iface.supports_buft(dev, sbuft) && !sbuft->iface.is_host(sbuft));
        for (int other = 0; other < ndev; ++other) if (other != main_dev) REQUIRE(!reg->iface.get_device(reg, other)->iface.supports_buft(reg->iface.get_device(reg, other), sbuft));
        ggml_backend_t be = dev->iface.init_backend(dev, nullptr);
        REQUIRE(be);
        ggml_backend_buffer_type_t buft = dev->iface.get_buffer_type(dev);
        ggml_tensor ws = make(type, K, M, nullptr, nullptr, "w_split");
        const size_t salloc = sbuft->iface.get_alloc_size(sbuft, &ws);
        REQUIRE(salloc >= wbytes);
        ggml_backend_buffer_t sbuf = sbuft->iface.alloc_buffer(sbuft, salloc);
        ggml_backend_buffer_t pbuf = buft->iface.alloc_buffer(buft, wbytes + 8192);
        ggml_backend_buffer_t abuf = buft->iface.alloc_buffer(buft, (size_t) (K + 2 * M) * 4 * (NB + 1) + 8192);
        REQUIRE(sbuf && pbuf && abuf);
        sbuf->usage = pbuf->usage = GGML_BACKEND_BUFFER_USAGE_WEIGHTS;
        ws.buffer = sbuf; ws.data = sbuf->iface.get_base(sbuf);
        ggml_tensor wp = make(type, K, M, pbuf, pbuf->iface.get_base(pbuf), "w_plain");
        REQUIRE(sbuf->iface.init_tensor(sbuf, &ws) == GGML_STATUS_SUCCESS && ws.extra != nullptr);
        pbuf->iface.init_tensor(pbuf, &wp);
        sbuf->iface.set_tensor(sbuf, &ws, host.data(), 0, wbytes);
        pbuf->iface.set_tensor(pbuf, &wp, host.data(), 0, wbytes);
        std::vector<uint8_t> back(wbytes);
        sbuf->iface.get_tensor(sbuf, &ws, back.data(), 0, wbytes);
        REQUIRE(memcmp(back.data(), host.data(), wbytes) == 0);                    // the host sees GGUF bytes, rows in order

        char * ab = (char *) abuf->iface.get_base(abuf);
        for (int64_t N : { (int64_t) 1, NB }) {
            ggml_tensor xa = make(GGML_TYPE_F32, K, N, abuf, ab, "x");
            ggml_tensor y1 = make(GGML_TYPE_F32, M, N, abuf, ab + ((size_t) K * N * 4 + 255) / 256 * 256, "y_split");
            ggml_tensor y2 = make(GGML_TYPE_F32, M, N, abuf, (char *) y1.data + ((size_t) M * N * 4 + 255) / 256 * 256, "y_plain");
            abuf->iface.set_tensor(abuf, &xa, x.data(), 0, (size_t) K * N * 4);
            y1.op = y2.op = GGML_OP_MUL_MAT; y1.src[0] = &ws; y1.src[1] = &xa; y2.src[0] = &wp; y2.src[1] = &xa;
            REQUIRE(dev->iface.supports_op(dev, &y1) && dev->iface.supports_op(dev, &y2));
            ggml_tensor bad = y1; bad.op = GGML_OP_ADD;
            REQUIRE(!dev->iface.supports_op(dev, &bad));                           // split tensors feed MUL_MAT only
            ggml_tensor * nodes[2] = { &y1, &y2 };
            ggml_cgraph g; memset(&g, 0, sizeof g); g.size = 2; g.n_nodes = 2; g.nodes = nodes;
            for (int rep = 0; rep < 3; ++rep) {                                    // (the second and third time around the plugin would replay a recording if it made one)
                REQUIRE(be->iface.graph_compute(be, &g) == GGML_STATUS_SUCCESS);
                be->iface.synchronize(be);
                std::vector<float> o1((size_t) M * N), o2((size_t) M * N);
                abuf->iface.get_tensor(abuf, &y1, o1.data(), 0, o1.size() * 4); abuf->iface.get_tensor(abuf, &y2, o2.data(), 0, o2.size() * 4);
                size_t diff = 0; double sum = 0, e2 = 0, r2 = 0;
                for (size_t i = 0; i < o1.size(); ++i) { diff += memcmp(&o1[i], &o2[i], 4) != 0; sum += o2[i]; e2 += ((double) o1[i] - o2[i]) * ((double) o1[i] - o2[i]); r2 += (double) o2[i] * o2[i]; }
                if (rep == 0) printf("split_probe: main device %d of %d, type %d, %lld x %lld, n = %lld: %zu of %zu words differ from the unsplit result (sum %.6g)\n", main_dev, ndev, type,
                                     (long long) K, (long long) M, (long long) N, diff, o1.size(), sum);
                // a row slice is a smaller matrix: the kernels may tile it (and order a row's partial sums) differently, nothing more
                REQUIRE(sum == sum && r2 > 0 && e2 / r2 < 1e-10);
            }
        }
        abuf->iface.free_buffer(abuf); pbuf->iface.free_buffer(pbuf); sbuf->iface.free_buffer(sbuf);
        be->iface.free(be);
    }
    printf("split_probe: OK\n");
    return 0;
}
