// umma_peak.cu -- bare tcgen05.mma issue-rate ceilings of one B200 (measurement only, not product code): kind::f16 (bf16 in, f32 accumulate)
// and kind::i8 (s8 in, s32 accumulate), M = 128, N = 256, operands resident in shared memory (no loads in the loop), two alternating
// TMEM accumulators, one CTA per SM (grid = SM count) and two.  The numbers bound what any kernel built on these instructions can reach:
// the int8 figure is the denominator for an int8 variant of the prefill GEMM (ggml-cuda's mmq.cuh works on int8 tensor cores).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a -Isuper-llama.cpp_b200/csrc scripts/umma_peak.cu -o gpurun_out/umma_peak && gpurun_out/umma_peak
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include "b200_ptx.cuh"

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

static __device__ __forceinline__ void umma_i8(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc, uint32_t accumulate) {
    asm volatile("{\n .reg .pred p;\n setp.ne.b32 p, %4, 0;\n tcgen05.mma.cta_group::1.kind::i8 [%0], %1, %2, %3, p;\n}"
                 :: "r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate) : "memory");
}
// instruction descriptor, kind::i8: D s32 (c_format 2), A / B signed 8-bit (format 1), both K-major (cute/arch/mma_sm100_desc.hpp InstrDescriptor)
static __host__ __device__ uint32_t umma_idesc_i8(int M, int N) {
    return (2u << 4) | (1u << 7) | (1u << 10) | ((uint32_t) (N >> 3) << 17) | ((uint32_t) (M >> 4) << 24);
}

constexpr int BM = 128, BN = 256;
// A: 128 rows x 128 B, B: 256 rows x 128 B, K-major SWIZZLE_128B atoms (8 rows x 128 B); one 128-byte row = 64 bf16 or 128 s8 of K
template <bool I8>
__global__ void __launch_bounds__(128, 1) k_umma(int iters, int fill, int * out) {
    extern __shared__ uint8_t raw[];
    uint8_t * smem = (uint8_t *) (((uintptr_t) raw + 1023) & ~(uintptr_t) 1023);
    uint8_t * sa = smem, * sb = smem + BM * 128;
    __shared__ uint64_t bar; __shared__ uint32_t tmem_base;
    for (int i = threadIdx.x; i < (BM + BN) * 128; i += blockDim.x) {
        if (I8) smem[i] = (uint8_t) fill;
        else if ((i & 1) == 0) *(uint16_t *) (smem + i) = fill == 1 ? 0x3f80 : (uint16_t) (0x3c00 + (i * 2654435761u >> 22));      // bf16 1.0 or noise
    }
    if (threadIdx.x == 0) { mbar_init(smem_u32(&bar), 1); mbar_fence_init(); }
    fence_proxy_async();
    __syncthreads();
    if (threadIdx.x < 32) tmem_alloc(smem_u32(&tmem_base), 512);
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = tmem_base;
    if (threadIdx.x == 0) {
        const uint64_t da = umma_desc_sw128(smem_u32(sa)), db = umma_desc_sw128(smem_u32(sb));
        const uint32_t idesc = I8 ? umma_idesc_i8(BM, BN) : umma_idesc_bf16(BM, BN);
        for (int it = 0; it < iters; ++it) {
            const uint32_t d = tmem + (uint32_t) ((it & 1) * BN);
#pragma unroll
            for (int k = 0; k < 4; ++k) {                                     // 4 K-steps of 32 B inside the 128-byte swizzle atom
                if (I8) umma_i8(d, da + 2 * k, db + 2 * k, idesc, (uint32_t) ((it >> 1) | k));
                else    umma_f16(d, da + 2 * k, db + 2 * k, idesc, (uint32_t) ((it >> 1) | k));
            }
        }
        umma_commit(smem_u32(&bar));
    }
    mbar_wait(smem_u32(&bar), 0);
    tc_fence_after();
    // read back one accumulator word per thread so the result is observable (row = thread, column 0 of accumulator 0)
    uint32_t r[16];
    tmem_ld16(tmem + ((uint32_t) (threadIdx.x & ~31) << 16), r);
    tmem_ld_wait();
    if (blockIdx.x == 0 && threadIdx.x < 4) out[threadIdx.x] = (int) r[0];
    tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) tmem_dealloc(tmem, 512);
}

int main() {
    cudaDeviceProp prop; CK(cudaGetDeviceProperties(&prop, 0));
    const int sms = prop.multiProcessorCount;
    int * out; CK(cudaMalloc(&out, 64)); CK(cudaMemset(out, 0, 64));
    const size_t smem = (BM + BN) * 128 + 1024;
    CK(cudaFuncSetAttribute(k_umma<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    CK(cudaFuncSetAttribute(k_umma<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int) smem));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    // sanity: all-ones operands, 2 accumulation rounds into accumulator 0 (iterations 0 and 2) x 4 K-steps: s8 -> 2 * 4 * 32 = 256, bf16 -> 2 * 4 * 16 = 128
    int h[4];
    k_umma<true><<<1, 128, smem>>>(4, 1, out); CK(cudaDeviceSynchronize()); CK(cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost));
    printf("check kind::i8  (all ones, expect 256): %d %d %d %d\n", h[0], h[1], h[2], h[3]);
    k_umma<false><<<1, 128, smem>>>(4, 1, out); CK(cudaDeviceSynchronize()); CK(cudaMemcpy(h, out, 16, cudaMemcpyDeviceToHost));
    printf("check kind::f16 (all ones, expect 128.0): %g %g %g %g\n", *(float *) &h[0], *(float *) &h[1], *(float *) &h[2], *(float *) &h[3]);
    for (int cps : { 1 }) for (int i8 : { 0, 1 }) {
        const int iters = 20000, grid = sms * cps;
        auto launch = [&] { if (i8) k_umma<true><<<grid, 128, smem>>>(iters, 0x55, out); else k_umma<false><<<grid, 128, smem>>>(iters, 0, out); };
        launch(); CK(cudaDeviceSynchronize());
        float best = 1e30f;
        for (int rep = 0; rep < 5; ++rep) {
            CK(cudaEventRecord(e0)); launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
            float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); if (ms < best) best = ms;
        }
        CK(cudaGetLastError());
        const double kper = i8 ? 32.0 : 16.0;
        const double ops = 2.0 * BM * BN * kper * 4.0 * iters * grid;
        printf("tcgen05.mma cta_group::1 kind::%s  M=128 N=256 K=%d  %d CTA/SM (grid %d): %8.1f %s  (%.3f ms for %d x 4 instructions per CTA; %.1f ns per instruction)\n",
               i8 ? "i8 " : "f16", (int) kper, cps, grid, ops / (best * 1e-3) / 1e12, i8 ? "TOP/s" : "TFLOP/s", best, iters, best * 1e6 / (iters * 4.0));
    }
    return 0;
}
