// b200_glue.cuh -- cheap non-hot-path ggml ops so a whole llama graph stays on the device.
// Declarations only; kernels live in b200_glue.cu.  All take ggml-style strided views.
#pragma once
#include "b200_common.cuh"

// returns true if (op, types, layout) is implemented by b200_glue_compute
struct ggml_tensor;
bool b200_glue_supports(const struct ggml_tensor * op);
int  b200_glue_compute(b200_ctx * ctx, struct ggml_tensor * op);
// RMS_NORM + the MUL by the norm weight right behind it as one kernel; B200_ERR_UNSUPPORTED when the pair does not fit
int  b200_glue_rms_norm_mul(b200_ctx * ctx, struct ggml_tensor * norm, struct ggml_tensor * mul);
// ADD + RMS_NORM + MUL (the residual add in front of a llama norm) as one launch, same contract
int  b200_glue_add_rms_norm_mul(b200_ctx * ctx, struct ggml_tensor * add, struct ggml_tensor * norm, struct ggml_tensor * mul);
// [ROPE(Q)] + ROPE(K) + SET_ROWS(K result -> f16 cache) + SET_ROWS(V -> f16 cache) as one launch (rq may be NULL), same contract
int  b200_glue_rope_kv(b200_ctx * ctx, struct ggml_tensor * rq, struct ggml_tensor * rk, struct ggml_tensor * sk, struct ggml_tensor * sv);
// the same four nodes described as the tail of a decode chain (out[0..2]: Q, K, V; b200_chain_rope, include/ggml-b200.h); nothing is launched.
// dead_in[0..2]: the Q / K / V matmul result is overwritten by another section's ROPE result (the allocator reused a dead block): do not store it
int  b200_glue_rope_kv_plan(struct ggml_tensor * rq, struct ggml_tensor * rk, struct ggml_tensor * sk, struct ggml_tensor * sv, b200_chain_rope * out, bool * dead_in);
