// abi_probe.cpp -- prints the binary layout of the ggml plugin ABI as seen through either the reference's own
// headers (-DPROBE_REFERENCE, -I/root/reference/ggml/include -I/root/reference/ggml/src) or the restatement in
// include/ggml-b200-abi.h.  tests/test_abi.py diffs the two outputs (and a committed copy of the reference's).
#include <cstdio>
#include <cstddef>
#ifdef PROBE_REFERENCE
#include "ggml.h"
#include "ggml-backend.h"
#include "ggml-backend-impl.h"
#include "ggml-impl.h"
#else
#include "ggml-b200-abi.h"
#endif
#define SZ(T)      printf("sizeof(" #T ")=%zu\n", sizeof(T))
#define OFF(T, f)  printf("offsetof(" #T "," #f ")=%zu\n", offsetof(T, f))
#define VAL(v)     printf(#v "=%d\n", (int) (v))
int main() {
    SZ(ggml_tensor); OFF(ggml_tensor, type); OFF(ggml_tensor, buffer); OFF(ggml_tensor, ne); OFF(ggml_tensor, nb); OFF(ggml_tensor, op);
    OFF(ggml_tensor, op_params); OFF(ggml_tensor, flags); OFF(ggml_tensor, src); OFF(ggml_tensor, view_src); OFF(ggml_tensor, view_offs);
    OFF(ggml_tensor, data); OFF(ggml_tensor, name); OFF(ggml_tensor, extra);
    OFF(ggml_cgraph, n_nodes); OFF(ggml_cgraph, nodes); OFF(ggml_cgraph, leafs);
    SZ(ggml_backend_buffer_type_i); SZ(ggml_backend_buffer_type); OFF(ggml_backend_buffer_type, device); OFF(ggml_backend_buffer_type, context);
    SZ(ggml_backend_buffer_i); SZ(ggml_backend_buffer); OFF(ggml_backend_buffer, buft); OFF(ggml_backend_buffer, context);
    OFF(ggml_backend_buffer, size); OFF(ggml_backend_buffer, usage);
    OFF(ggml_backend_buffer_i, init_tensor); OFF(ggml_backend_buffer_i, set_tensor); OFF(ggml_backend_buffer_i, cpy_tensor); OFF(ggml_backend_buffer_i, reset);
    SZ(ggml_backend_i); OFF(ggml_backend_i, graph_compute); OFF(ggml_backend_i, event_record); OFF(ggml_backend_i, graph_optimize);
    SZ(ggml_backend); OFF(ggml_backend, iface); OFF(ggml_backend, device); OFF(ggml_backend, context);
    SZ(ggml_backend_event); SZ(ggml_backend_device_i); OFF(ggml_backend_device_i, supports_op); OFF(ggml_backend_device_i, event_new);
    SZ(ggml_backend_device); OFF(ggml_backend_device, reg); OFF(ggml_backend_device, context);
    SZ(ggml_backend_reg_i); SZ(ggml_backend_reg); OFF(ggml_backend_reg, iface); OFF(ggml_backend_reg, context);
    SZ(ggml_backend_dev_props); OFF(ggml_backend_dev_props, caps); SZ(ggml_backend_dev_caps); SZ(ggml_backend_feature);
    VAL(GGML_BACKEND_API_VERSION); VAL(GGML_MAX_DIMS); VAL(GGML_MAX_SRC); VAL(GGML_MAX_OP_PARAMS); VAL(GGML_MAX_NAME);
    VAL(GGML_TYPE_F32); VAL(GGML_TYPE_F16); VAL(GGML_TYPE_Q4_0); VAL(GGML_TYPE_Q8_0); VAL(GGML_TYPE_Q8_1); VAL(GGML_TYPE_Q4_K);
    VAL(GGML_TYPE_Q5_K); VAL(GGML_TYPE_Q6_K); VAL(GGML_TYPE_Q8_K); VAL(GGML_TYPE_I32); VAL(GGML_TYPE_BF16); VAL(GGML_TYPE_COUNT);
    VAL(GGML_OP_NONE); VAL(GGML_OP_DUP); VAL(GGML_OP_ADD); VAL(GGML_OP_ADD_ID); VAL(GGML_OP_SUB); VAL(GGML_OP_MUL); VAL(GGML_OP_DIV);
    VAL(GGML_OP_SUM_ROWS); VAL(GGML_OP_RMS_NORM); VAL(GGML_OP_MUL_MAT); VAL(GGML_OP_MUL_MAT_ID); VAL(GGML_OP_SCALE); VAL(GGML_OP_CPY);
    VAL(GGML_OP_CONT); VAL(GGML_OP_RESHAPE); VAL(GGML_OP_VIEW); VAL(GGML_OP_PERMUTE); VAL(GGML_OP_TRANSPOSE); VAL(GGML_OP_GET_ROWS);
    VAL(GGML_OP_SET_ROWS); VAL(GGML_OP_SOFT_MAX); VAL(GGML_OP_ROPE); VAL(GGML_OP_CLAMP); VAL(GGML_OP_ARGSORT); VAL(GGML_OP_TOP_K);
    VAL(GGML_OP_FLASH_ATTN_EXT); VAL(GGML_OP_UNARY); VAL(GGML_OP_GLU); VAL(GGML_OP_COUNT);
    VAL(GGML_UNARY_OP_SILU); VAL(GGML_GLU_OP_REGLU); VAL(GGML_GLU_OP_GEGLU); VAL(GGML_GLU_OP_SWIGLU);
    VAL(GGML_STATUS_ALLOC_FAILED); VAL(GGML_STATUS_FAILED); VAL(GGML_STATUS_SUCCESS); VAL(GGML_STATUS_ABORTED);
    VAL(GGML_PREC_DEFAULT); VAL(GGML_PREC_F32);
    VAL(GGML_BACKEND_BUFFER_USAGE_ANY); VAL(GGML_BACKEND_BUFFER_USAGE_WEIGHTS); VAL(GGML_BACKEND_BUFFER_USAGE_COMPUTE);
    VAL(GGML_BACKEND_DEVICE_TYPE_CPU); VAL(GGML_BACKEND_DEVICE_TYPE_GPU); VAL(GGML_BACKEND_DEVICE_TYPE_IGPU); VAL(GGML_BACKEND_DEVICE_TYPE_ACCEL);
    return 0;
}
