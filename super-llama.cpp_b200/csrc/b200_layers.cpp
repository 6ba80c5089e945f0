// b200_layers.cpp -- SURVEY 8(a) A9: which GPU(s) each transformer layer goes to.  Host-only planning, the same three policies as the
// fork's llama_multi_gpu_manager::distribute_layers (src/llama-multi-gpu.cpp:156-332): round-robin (:194-214), greedy "most free memory
// left" (:216-275; first GPU wins ties, GPU 0 when nothing is left anywhere), tensor-parallel bookkeeping (:277-332: every layer split
// across the first tp GPUs, GPU 0 coordinating, memory charged as size / tp per GPU; tp < 2 falls back to the greedy policy).  Any other
// strategy value plans like MEMORY_BALANCED, as the reference's switch does (:170-185).  It moves no data; b200_split_row_range and the
// row-split matmuls (b200_api.cu) are what act on a plan.
#include "../../include/ggml-b200.h"
#include <vector>

void b200_set_error(const char * fmt, ...);

static void plan_memory_balanced(int n_layers, const uint64_t * sizes, int n_sizes, int n_gpus, const uint64_t * gpu_free, b200_layer_assignment * out, uint64_t * allocated) {
    for (int i = 0; i < n_layers; ++i) {
        const uint64_t sz = i < n_sizes ? sizes[i] : 0;
        int best = 0; uint64_t best_left = 0;
        for (int g = 0; g < n_gpus; ++g) {
            const uint64_t left = gpu_free[g] > allocated[g] ? gpu_free[g] - allocated[g] : 0;
            if (left > best_left) { best_left = left; best = g; }
        }
        out[i].layer_id = i; out[i].gpu_id = best; out[i].memory_required = sz; out[i].is_split = 0; out[i].n_split = 0;
        allocated[best] += sz;
    }
}

extern "C" B200_API int b200_distribute_layers(int strategy, int tensor_parallel_size, int n_layers, const uint64_t * layer_sizes, int n_sizes,
                                               int n_gpus, const uint64_t * gpu_free_memory, b200_layer_assignment * out, uint64_t * gpu_allocated) {
    if (n_layers < 0 || n_sizes < 0 || n_gpus < 0 || (n_layers > 0 && !out) || (n_sizes > 0 && !layer_sizes) || (n_gpus > 0 && !gpu_free_memory)) {
        b200_set_error("distribute_layers: bad arguments"); return B200_ERR_INVALID;
    }
    if (n_gpus == 0) return 0;                                   // "no GPUs available for layer distribution": an empty plan (:163-166)
    std::vector<uint64_t> local(n_gpus, 0);
    uint64_t * allocated = gpu_allocated ? gpu_allocated : local.data();
    for (int g = 0; g < n_gpus; ++g) allocated[g] = 0;
    for (int i = 0; i < n_layers; ++i) for (int k = 0; k < B200_MAX_PEERS; ++k) out[i].split_gpu_ids[k] = -1;
    if (strategy == B200_LAYERS_ROUND_ROBIN) {
        for (int i = 0; i < n_layers; ++i) {
            out[i].layer_id = i; out[i].gpu_id = i % n_gpus; out[i].memory_required = i < n_sizes ? layer_sizes[i] : 0; out[i].is_split = 0; out[i].n_split = 0;
            allocated[out[i].gpu_id] += out[i].memory_required;
        }
        return n_layers;
    }
    const int tp = tensor_parallel_size < n_gpus ? tensor_parallel_size : n_gpus;
    if (strategy == B200_LAYERS_TENSOR_PARALLEL && tp >= 2) {
        if (tp > B200_MAX_PEERS) { b200_set_error("distribute_layers: tensor_parallel_size %d exceeds %d", tp, B200_MAX_PEERS); return B200_ERR_UNSUPPORTED; }
        for (int i = 0; i < n_layers; ++i) {
            out[i].layer_id = i; out[i].gpu_id = 0; out[i].memory_required = i < n_sizes ? layer_sizes[i] : 0; out[i].is_split = 1; out[i].n_split = tp;
            for (int g = 0; g < tp; ++g) { out[i].split_gpu_ids[g] = g; allocated[g] += out[i].memory_required / (uint64_t) tp; }
        }
        return n_layers;
    }
    plan_memory_balanced(n_layers, layer_sizes, n_sizes, n_gpus, gpu_free_memory, out, allocated);
    return n_layers;
}

// The plan as llama.cpp takes it: llama_model_params::tensor_split, "proportion of the model (layers or rows) to offload to each GPU"
// (include/llama.h) -- for -sm layer the share of layers per GPU, for a tensor-parallel plan the share of every split matrix's rows
// (what b200_split_row_range / ggml_backend_split_buffer_type consume).  tensor_split[n_gpus], sums to 1 (all 0 for an empty plan).
extern "C" B200_API int b200_plan_tensor_split(const b200_layer_assignment * plan, int n, int n_gpus, float * tensor_split) {
    if (n < 0 || n_gpus <= 0 || !tensor_split || (n > 0 && !plan)) { b200_set_error("plan_tensor_split: bad arguments"); return B200_ERR_INVALID; }
    std::vector<double> share(n_gpus, 0.0);
    for (int i = 0; i < n; ++i) {
        if (plan[i].is_split && plan[i].n_split > 0) {
            for (int k = 0; k < plan[i].n_split && k < B200_MAX_PEERS; ++k) {
                const int g = plan[i].split_gpu_ids[k];
                if (g < 0 || g >= n_gpus) { b200_set_error("plan_tensor_split: GPU %d outside 0..%d", g, n_gpus - 1); return B200_ERR_INVALID; }
                share[g] += 1.0 / plan[i].n_split;
            }
        } else {
            if (plan[i].gpu_id < 0 || plan[i].gpu_id >= n_gpus) { b200_set_error("plan_tensor_split: GPU %d outside 0..%d", plan[i].gpu_id, n_gpus - 1); return B200_ERR_INVALID; }
            share[plan[i].gpu_id] += 1.0;
        }
    }
    for (int g = 0; g < n_gpus; ++g) tensor_split[g] = n > 0 ? (float) (share[g] / n) : 0.0f;
    return B200_OK;
}

// what a binding checks its mirrors of the public structs against (include/ggml-b200.h: b200_abi_sizeof)
extern "C" B200_API size_t b200_abi_sizeof(int which) {
    switch (which) {
        case 0: return sizeof(b200_tensor);
        case 1: return sizeof(b200_chain_rope);
        case 2: return sizeof(b200_chain_phase);
        case 3: return sizeof(b200_chain_split);
        case 4: return sizeof(b200_kv_pages);
        case 5: return sizeof(b200_load_item);
        case 6: return sizeof(b200_load_stats);
        case 7: return sizeof(b200_exchange);
        case 8: return sizeof(b200_layer_assignment);
    }
    return 0;
}
