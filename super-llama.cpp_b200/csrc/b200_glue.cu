// b200_glue.cu -- non-hot-path ggml ops (see b200_glue.cuh).  Filled in after the hot path is green.
#include "b200_glue.cuh"
#include "../../include/ggml-b200-abi.h"

bool b200_glue_supports(const struct ggml_tensor *) { return false; }
int  b200_glue_compute(b200_ctx *, struct ggml_tensor * op) { b200_set_error("op %d not implemented", (int) op->op); return B200_ERR_UNSUPPORTED; }
