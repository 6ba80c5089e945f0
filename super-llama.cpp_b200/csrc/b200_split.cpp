// b200_split.cpp -- split-mode=row buffer type ("ggml_backend_split_buffer_type" proc address).
// Placeholder until the NCCL row-split path lands: returns NULL, which llama-model.cpp treats as
// "row split not available" (src/llama-model.cpp:393-411).
#include "b200_common.cuh"
#include "../../include/ggml-b200-abi.h"

ggml_backend_buffer_type_t b200_split_buffer_type(int, const float *) { return nullptr; }
