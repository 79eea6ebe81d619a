// Standalone micro-benchmark (not part of libpt_stage1.so): cycles per tcgen05.mma (cta_group::1, M = 128, K = 16,
// bf16 -> f32) for combinations of operand major-ness and N, issued back to back by one thread from shared memory.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -o tools/bin/mma_probe tools/mma_probe.cu
#include <cstdio>
#include <cstdint>
#include <cuda_runtime.h>
#include "../projectiontrainer_b200/csrc/sm100_ptx.cuh"
using namespace pt;

struct Cfg { int N, a_mn, b_mn, a_lbo, b_lbo, a_kstep, b_kstep, nmma; };

__global__ void __launch_bounds__(128, 1) probe(Cfg c, long long* out) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~uintptr_t(1023));
    __shared__ uint64_t bar;
    __shared__ uint32_t slot;
    for (int i = threadIdx.x; i < 160 * 1024 / 4; i += blockDim.x) reinterpret_cast<uint32_t*>(smem)[i] = 0x3c003c00u;
    if (threadIdx.x == 0) { ptx::mbar_init(&bar, 1); ptx::fence_mbar_init(); }
    if (threadIdx.x < 32) { ptx::tmem_alloc(&slot, 512); ptx::tmem_relinquish(); }
    ptx::fence_proxy_async_smem();
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tm = slot;
    if (threadIdx.x == 0) {
        const uint32_t idesc = ptx::make_idesc_bf16(128, c.N, c.a_mn, c.b_mn);
        const uint32_t a0 = ptx::smem_u32(smem), b0 = ptx::smem_u32(smem + 64 * 1024);
        for (int rep = 0; rep < 3; ++rep) {
            const long long t0 = clock64();
            for (int i = 0; i < c.nmma; ++i) {
                const int k = i & 3;
                const uint64_t ad = ptx::make_smem_desc_sw128(a0 + k * c.a_kstep, c.a_lbo, 1024);
                const uint64_t bd = ptx::make_smem_desc_sw128(b0 + k * c.b_kstep, c.b_lbo, 1024);
                ptx::mma_bf16_ss(tm, ad, bd, idesc, i != 0);
            }
            const long long t1 = clock64();
            ptx::mma_commit(&bar);
            ptx::mbar_wait(&bar, rep & 1);
            const long long t2 = clock64();
            out[rep * 2] = t1 - t0; out[rep * 2 + 1] = t2 - t0;
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (threadIdx.x < 32) { ptx::tc_fence_after(); ptx::tmem_dealloc(tm, 512); }
}

int main() {
    long long* d; cudaMalloc(&d, 64);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, 200 * 1024);
    const int nm = 64;
    // K-major: k-step 32 B inside the 128-byte row, LBO unused; MN-major: k-step 16 rows x 128 B, LBO = slab stride
    struct { const char* name; Cfg c; } cases[] = {
        {"A K-major , B K-major , N= 64", {64, 0, 0, 0, 0, 32, 32, nm}},
        {"A K-major , B K-major , N=128", {128, 0, 0, 0, 0, 32, 32, nm}},
        {"A K-major , B K-major , N=256", {256, 0, 0, 0, 0, 32, 32, nm}},
        {"A MN-major, B MN-major, N= 64", {64, 1, 1, 16384, 0, 2048, 2048, nm}},
        {"A MN-major, B K-major , N= 64", {64, 1, 0, 16384, 0, 2048, 32, nm}},
        {"A K-major , B MN-major, N= 64", {64, 0, 1, 0, 0, 32, 2048, nm}},
        {"A MN-major, B MN-major, N=128", {128, 1, 1, 16384, 8192, 2048, 2048, nm}},
        {"A MN-major, B MN-major, N=256", {256, 1, 1, 16384, 8192, 2048, 2048, nm}},
        {"A K-major , B MN-major, N=256", {256, 0, 1, 0, 8192, 32, 2048, nm}},
        {"A K-major , B MN-major, N=128", {128, 0, 1, 0, 8192, 32, 2048, nm}},
    };
    for (auto& cs : cases) {
        probe<<<1, 128, 200 * 1024>>>(cs.c, d);
        cudaError_t e = cudaDeviceSynchronize();
        if (e != cudaSuccess) { printf("%s: %s\n", cs.name, cudaGetErrorString(e)); return 1; }
        long long h[6]; cudaMemcpy(h, d, 48, cudaMemcpyDeviceToHost);
        printf("%s: %d MMAs  issue %lld cyc  retire %lld cyc  = %.1f cyc/MMA (128 x N x 16), floor max(32, N/2)\n", cs.name, nm,
               h[4], h[5], double(h[5]) / nm);
    }
    return 0;
}
