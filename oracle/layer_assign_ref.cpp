// layer_assign_ref.cpp -- TEST INFRASTRUCTURE (oracle/): drives the reference's OWN llama_multi_gpu_manager (src/llama-multi-gpu.{h,cpp},
// compiled in place into oracle/_ref/avx2/libllama-ref.so) with fabricated ggml devices whose get_memory reports the sizes given on the
// command line, and prints the layer assignments of distribute_layers (src/llama-multi-gpu.cpp:156-192).  tests/golden/make_layer_golden.py
// turns its output into the committed fixture that pins both oracle/layer_assign.py and the product's b200_distribute_layers.
// usage: layer_assign_ref <strategy 0..5> <tp_size> <n_layers> <gpu free MB, comma separated> <layer size MB, comma separated (may be shorter than n_layers)>
#include "llama-multi-gpu.h"
#include "ggml-backend-impl.h"
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

struct fake { std::string name; size_t free_b, total_b; };
static const char * f_name(ggml_backend_dev_t d) { return ((fake *) d->context)->name.c_str(); }
static void f_memory(ggml_backend_dev_t d, size_t * fr, size_t * to) { *fr = ((fake *) d->context)->free_b; *to = ((fake *) d->context)->total_b; }
static enum ggml_backend_dev_type f_type(ggml_backend_dev_t) { return GGML_BACKEND_DEVICE_TYPE_GPU; }

static std::vector<size_t> parse(const char * s) {
    std::vector<size_t> v;
    if (!*s || !strcmp(s, "-")) return v;
    char * dup = strdup(s);
    for (char * t = strtok(dup, ","); t; t = strtok(nullptr, ",")) v.push_back((size_t) (atof(t) * 1024.0 * 1024.0));
    free(dup);
    return v;
}

int main(int argc, char ** argv) {
    if (argc < 6) return 2;
    llama_multi_gpu_manager::config cfg;
    cfg.strategy = (llama_multi_gpu_strategy) atoi(argv[1]);
    cfg.tensor_parallel_size = atoi(argv[2]);
    const int n_layers = atoi(argv[3]);
    const std::vector<size_t> gpu_free = parse(argv[4]), sizes = parse(argv[5]);
    std::vector<fake> fakes(gpu_free.size());
    std::vector<ggml_backend_device> devs(gpu_free.size());
    std::vector<ggml_backend_dev_t> ptrs;
    for (size_t i = 0; i < gpu_free.size(); ++i) {
        fakes[i] = { "FAKE" + std::to_string(i), gpu_free[i], gpu_free[i] + (64u << 20) };
        memset(&devs[i], 0, sizeof devs[i]);
        devs[i].iface.get_name = f_name; devs[i].iface.get_memory = f_memory; devs[i].iface.get_type = f_type; devs[i].context = &fakes[i];
        ptrs.push_back(&devs[i]);
    }
    llama_multi_gpu_manager mgr(cfg);
    mgr.initialize(ptrs);
    alignas(64) static char dummy_model[64];                      // distribute_layers never touches the model argument
    const auto as = mgr.distribute_layers(*reinterpret_cast<const llama_model *>(dummy_model), n_layers, sizes);
    printf("n=%zu\n", as.size());
    for (const auto & a : as) {
        printf("%d %d %zu %d", a.layer_id, a.gpu_id, a.memory_required, a.is_split ? 1 : 0);
        for (int g : a.split_gpu_ids) printf(" %d", g);
        printf("\n");
    }
    return 0;
}
