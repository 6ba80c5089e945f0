// b200_gemm.cu -- prefill GEMM (MUL_MAT / MUL_MAT_ID with more than 8 activation columns).
// Placeholder translation unit until the tcgen05/TMEM kernel lands: reports "not supported" so the
// dispatcher keeps using the column-blocked GEMV (correct, re-reads the weights once per 8 columns).
#include "b200_common.cuh"

bool b200_gemm_supported(const b200_tensor *, const b200_tensor *) { return false; }
int  b200_launch_gemm_q(b200_ctx *, const b200_tensor *, const b200_tensor *, b200_tensor *) { return B200_ERR_UNSUPPORTED; }
bool b200_gemm_id_supported(const b200_tensor *, const b200_tensor *, const b200_tensor *) { return false; }
int  b200_launch_gemm_id_q(b200_ctx *, const b200_tensor *, const b200_tensor *, const b200_tensor *, b200_tensor *) { return B200_ERR_UNSUPPORTED; }
