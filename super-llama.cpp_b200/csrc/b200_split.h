// b200_split.h -- split-mode=row inside one process (the "ggml_backend_split_buffer_type" the reference's loader asks the registry for,
// src/llama-model.cpp:393-411): internal interface between the buffer type (b200_split.cpp) and the plugin (b200_backend.cpp).
#pragma once
#include "b200_common.cuh"
#include "../../include/ggml-b200-abi.h"

#define B200_SPLIT_MAX_DEVICES 16

// rows [lo, hi) of a weight live on logical device i (hi == lo: none); `repacked`: the slice is in the device layout of b200_repack
struct b200_split_slice { void * data; int64_t lo, hi; int repacked; };
struct b200_split_extra { int n; b200_split_slice s[B200_SPLIT_MAX_DEVICES]; };          // hangs off ggml_tensor.extra (set by init_tensor)

ggml_backend_buffer_type_t b200_split_buffer_type(int main_device, const float * tensor_split);
bool b200_buft_is_split(ggml_backend_buffer_type_t buft);
static inline const b200_split_extra * b200_split_of(const ggml_tensor * w) {
    return (w && w->buffer && b200_buft_is_split(w->buffer->buft)) ? (const b200_split_extra *) w->extra : nullptr;
}

// provided by b200_backend.cpp: the registry's devices (logical index -> CUDA device; several logical devices may share one GPU with
// GGML_B200_FAKE_DEVICES, which is how a one-GPU box exercises this path)
int                b200_logical_devices();
int                b200_logical_to_cuda(int i);
ggml_backend_dev_t b200_logical_dev(int i);
b200_ctx *         b200_util_ctx(int cuda_device);
int                b200_repack_policy(const ggml_backend_buffer * b);      // 1: weights of this buffer are re-laid out at load time
