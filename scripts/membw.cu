// membw.cu -- read-bandwidth ceilings of one B200 for the access patterns the decode kernels use (measurement only, not product code).
//   nvcc -O3 -std=c++17 -gencode arch=compute_100a,code=sm_100a scripts/membw.cu -o gpurun_out/membw && gpurun_out/membw
// (1) plain 128-bit loads, grid-stride; (2) cp.async.bulk into a shared-memory ring, nothing consumed: W issuing warps per CTA, each with
// D stages of B bytes in flight, either walking its own contiguous range (layout 0) or interleaved with all other warps of the GPU in
// tiles of B bytes (layout 1, what a persistent kernel with round-robin tiles does).  Buffer 4 GiB >> 126 MB L2.
#include <cuda_runtime.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>

#define CK(x) do { cudaError_t e_ = (x); if (e_ != cudaSuccess) { printf("%s -> %s\n", #x, cudaGetErrorString(e_)); exit(1); } } while (0)

__global__ void k_ldg(const int4 * __restrict__ p, size_t n16, int * out) {
    int4 acc = make_int4(0, 0, 0, 0);
    const size_t stride = (size_t) gridDim.x * blockDim.x;
    size_t i = (size_t) blockIdx.x * blockDim.x + threadIdx.x;
    for (; i + 7 * stride < n16; i += 8 * stride) {
        int4 v[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) asm volatile("ld.global.nc.L1::no_allocate.v4.s32 {%0,%1,%2,%3}, [%4];" : "=r"(v[j].x), "=r"(v[j].y), "=r"(v[j].z), "=r"(v[j].w) : "l"(p + i + j * stride));
#pragma unroll
        for (int j = 0; j < 8; ++j) { acc.x ^= v[j].x; acc.y ^= v[j].y; acc.z ^= v[j].z; acc.w ^= v[j].w; }
    }
    if ((acc.x ^ acc.y ^ acc.z ^ acc.w) == 0x12345678) *out = 1;
}

static __device__ __forceinline__ uint32_t smem_u32(const void * p) { return (uint32_t) __cvta_generic_to_shared(p); }

template <bool HINT>
__global__ void __launch_bounds__(1024, 1) k_bulk(const uint8_t * __restrict__ p, size_t nbytes, int B, int D, int layout) {
    extern __shared__ __align__(128) uint8_t smem[];
    const int W = blockDim.x >> 5, warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint64_t * bars = (uint64_t *) (smem + (size_t) W * D * B);
    const uint32_t ring = smem_u32(smem + (size_t) warp * D * B), bar0 = smem_u32(bars + warp * D);
    if (lane == 0) {
        for (int s = 0; s < D; ++s) asm volatile("mbarrier.init.shared::cta.b64 [%0], 1;" :: "r"(bar0 + 8 * s) : "memory");
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncwarp();
    if (lane != 0) return;
    uint64_t pol;
    asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(pol));
    const size_t gw = (size_t) blockIdx.x * W + warp, nw = (size_t) gridDim.x * W;
    const size_t ntiles = nbytes / B, per = ntiles / nw;            // tiles per warp
    auto addr = [&](size_t i) { return p + (layout ? (i * nw + gw) : (gw * per + i)) * (size_t) B; };
    auto issue = [&](size_t i, int s) {
        const uint32_t bar = bar0 + 8 * s, dst = ring + s * B;
        asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" :: "r"(bar), "r"(B) : "memory");
        if (HINT) asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" :: "r"(dst), "l"(addr(i)), "r"(B), "r"(bar), "l"(pol) : "memory");
        else      asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" :: "r"(dst), "l"(addr(i)), "r"(B), "r"(bar) : "memory");
    };
    for (int s = 0; s < D && (size_t) s < per; ++s) issue(s, s);
    int s = 0; uint32_t ph = 0;
    for (size_t i = 0; i < per; ++i) {
        uint32_t ok = 0;
        while (!ok) asm volatile("{\n .reg .pred q;\n mbarrier.try_wait.parity.shared::cta.b64 q, [%1], %2;\n selp.u32 %0, 1, 0, q;\n}" : "=r"(ok) : "r"(bar0 + 8 * s), "r"(ph) : "memory");
        if (i + D < per) issue(i + D, s);
        if (++s == D) { s = 0; ph ^= 1; }
    }
}

int main() {
    const size_t nbytes = (size_t) 4 << 30;
    uint8_t * buf; int * out;
    CK(cudaMalloc(&buf, nbytes)); CK(cudaMalloc(&out, 4));
    CK(cudaMemset(buf, 1, nbytes));
    cudaEvent_t e0, e1; CK(cudaEventCreate(&e0)); CK(cudaEventCreate(&e1));
    auto timeit = [&](auto launch) {
        launch(); CK(cudaDeviceSynchronize());
        CK(cudaEventRecord(e0)); for (int r = 0; r < 3; ++r) launch(); CK(cudaEventRecord(e1)); CK(cudaEventSynchronize(e1));
        float ms; CK(cudaEventElapsedTime(&ms, e0, e1)); CK(cudaGetLastError());
        return nbytes / (ms / 3 * 1e-3) / 1e9;
    };
    for (int cps : { 4, 8, 16 }) printf("ldg.128  %2d CTAs x 256 thr per SM: %7.0f GB/s\n", cps, timeit([&] { k_ldg<<<148 * cps, 256>>>((const int4 *) buf, nbytes / 16, out); }));
    CK(cudaFuncSetAttribute(k_bulk<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    CK(cudaFuncSetAttribute(k_bulk<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, 227 * 1024));
    struct cfg { int W, B, D, layout, hint; };
    const cfg cfgs[] = { {16, 1152, 8, 0, 1}, {16, 1152, 8, 1, 1}, {16, 4608, 2, 0, 1}, {16, 4608, 2, 1, 1}, {16, 4608, 2, 1, 0}, {16, 9216, 1, 1, 1}, {8, 9216, 2, 1, 1},
                         {4, 9216, 5, 1, 1}, {4, 16384, 3, 1, 1}, {2, 32768, 3, 1, 1}, {1, 32768, 6, 1, 1}, {32, 4608, 1, 1, 1}, {32, 2304, 2, 1, 1}, {16, 2304, 5, 0, 1}, {16, 2304, 5, 1, 0},
                         {16, 4608, 1, 1, 1}, {8, 4608, 2, 1, 1}, {16, 1152, 4, 1, 1}, {16, 1152, 2, 1, 1} };
    for (const cfg & c : cfgs) {
        const size_t smem = (size_t) c.W * c.D * c.B + (size_t) c.W * c.D * 8;
        if (smem > 227 * 1024) { printf("skip W=%d B=%d D=%d\n", c.W, c.B, c.D); continue; }
        const double gbs = c.hint ? timeit([&] { k_bulk<true><<<148, c.W * 32, smem>>>(buf, nbytes, c.B, c.D, c.layout); })
                                  : timeit([&] { k_bulk<false><<<148, c.W * 32, smem>>>(buf, nbytes, c.B, c.D, c.layout); });
        printf("bulk W=%2d warps  B=%5d bytes  D=%d stages (%3d KB in flight / SM)  %s  %s: %7.0f GB/s\n", c.W, c.B, c.D, c.W * c.D * c.B / 1024,
               c.layout ? "interleaved tiles" : "contiguous/warp  ", c.hint ? "evict_first" : "no hint    ", gbs);
    }
    return 0;
}
