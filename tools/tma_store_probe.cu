// Standalone bring-up probe: TMA store of a [128 keys][G heads][pos positions] bf16 smem tile (128-byte rows) into a
// transposed-score tensor dS^T[b, head, key, pos] through a 5-D tensor map, variants by argv[1]:
//   0: dims (pos, G, key, hkv, B) swizzle 128B (what attn_gemma_bwd2 wants)   1: same, no swizzle
//   2: 3-D map (pos, G, key) only                                              3: dims (pos, key, G, ..) (monotonic strides; layout check only)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O2 -std=c++17 -o tools/bin/tma_store_probe tools/tma_store_probe.cu -lcuda
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <vector>
#include <cuda.h>
#include <cuda_runtime.h>
#include <cuda_bf16.h>

__global__ void k5(const __grid_constant__ CUtensorMap tm, int c0, int c1, int c2, int c3, int c4, int rank) {
    extern __shared__ uint8_t raw[];
    uint8_t* sm = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(raw) + 1023) & ~uintptr_t(1023));
    __nv_bfloat16* t = reinterpret_cast<__nv_bfloat16*>(sm);
    for (int i = threadIdx.x; i < 128 * 64; i += blockDim.x) t[i] = __float2bfloat16(float(i % 251));   // dense image
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    __syncthreads();
    if (threadIdx.x == 0) {
        uint32_t sa = static_cast<uint32_t>(__cvta_generic_to_shared(sm));
        if (rank == 5)
            asm volatile("cp.async.bulk.tensor.5d.global.shared::cta.bulk_group [%0, {%2, %3, %4, %5, %6}], [%1];" ::
                         "l"(reinterpret_cast<uint64_t>(&tm)), "r"(sa), "r"(c0), "r"(c1), "r"(c2), "r"(c3), "r"(c4) : "memory");
        else
            asm volatile("cp.async.bulk.tensor.3d.global.shared::cta.bulk_group [%0, {%2, %3, %4}], [%1];" ::
                         "l"(reinterpret_cast<uint64_t>(&tm)), "r"(sa), "r"(c0), "r"(c1), "r"(c2) : "memory");
        asm volatile("cp.async.bulk.commit_group;" ::: "memory");
        asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
    }
}

typedef CUresult (*Enc)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*, const cuuint64_t*,
                        const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                        CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

int main(int argc, char** argv) {
    const int variant = argc > 1 ? atoi(argv[1]) : 0;
    const int B = 2, hkv = 1, G = 4, S = 200, ldT = 200, pos = 16;
    const size_t n = size_t(B) * hkv * G * S * ldT;
    __nv_bfloat16* d; cudaMalloc(&d, n * 2); cudaMemset(d, 0, n * 2);
    void* sym = nullptr; cudaDriverEntryPointQueryResult q;
    cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q);
    Enc enc = reinterpret_cast<Enc>(sym);
    CUtensorMap tm;
    cuuint32_t es[5] = {1, 1, 1, 1, 1};
    CUresult r;
    int rank = 5;
    if (variant == 2) {
        rank = 3;
        cuuint64_t dims[3] = {(cuuint64_t)S, (cuuint64_t)G, (cuuint64_t)S};
        cuuint64_t st[2] = {(cuuint64_t)S * ldT * 2, (cuuint64_t)ldT * 2};
        cuuint32_t box[3] = {(cuuint32_t)pos, (cuuint32_t)G, 128};
        r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else if (variant == 3) {
        cuuint64_t dims[5] = {(cuuint64_t)S, (cuuint64_t)S, (cuuint64_t)G, (cuuint64_t)hkv, (cuuint64_t)B};
        cuuint64_t st[4] = {(cuuint64_t)ldT * 2, (cuuint64_t)S * ldT * 2, (cuuint64_t)G * S * ldT * 2, (cuuint64_t)hkv * G * S * ldT * 2};
        cuuint32_t box[5] = {(cuuint32_t)pos, 128, (cuuint32_t)G, 1, 1};
        r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    } else {
        cuuint64_t dims[5] = {(cuuint64_t)S, (cuuint64_t)G, (cuuint64_t)S, (cuuint64_t)hkv, (cuuint64_t)B};
        cuuint64_t st[4] = {(cuuint64_t)S * ldT * 2, (cuuint64_t)ldT * 2, (cuuint64_t)G * S * ldT * 2, (cuuint64_t)hkv * G * S * ldT * 2};
        cuuint32_t box[5] = {(cuuint32_t)pos, (cuuint32_t)G, 128, 1, 1};
        r = enc(&tm, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, d, dims, st, box, es, CU_TENSOR_MAP_INTERLEAVE_NONE,
                variant == 1 ? CU_TENSOR_MAP_SWIZZLE_NONE : CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    }
    printf("variant %d: encode rc %d\n", variant, (int)r);
    if (r != CUDA_SUCCESS) return 1;
    cudaFuncSetAttribute(k5, cudaFuncAttributeMaxDynamicSharedMemorySize, 20 * 1024);
    const int c_pos = 32, c_key = variant == 3 ? 0 : 128;     // second key tile: rows 128..199 exist, 200..255 clipped
    if (variant == 3) k5<<<1, 128, 20 * 1024>>>(tm, c_pos, 128, 0, 0, 1, 5);
    else k5<<<1, 128, 20 * 1024>>>(tm, c_pos, 0, c_key, 0, 1, rank);
    cudaError_t e = cudaDeviceSynchronize();
    printf("variant %d: kernel %s\n", variant, cudaGetErrorString(e));
    if (e != cudaSuccess) return 2;
    std::vector<__nv_bfloat16> h(n);
    cudaMemcpy(h.data(), d, n * 2, cudaMemcpyDeviceToHost);
    // expected (variants 0/1/2): element (b=1 [or 0 for rank 3], head g, key 128+kr, pos 32+p) = image[kr*64 + g*16 + p] (unswizzled view)
    int bad = 0, nz = 0;
    const int bb = variant == 2 ? 0 : 1;
    for (size_t i = 0; i < n; ++i) nz += __bfloat162float(h[i]) != 0.f;
    if (variant != 3)
        for (int g = 0; g < G; ++g) for (int kr = 0; kr < 72; ++kr) for (int p = 0; p < pos; ++p) {
            const float got = __bfloat162float(h[((size_t(bb) * G + g) * S + 128 + kr) * ldT + c_pos + p]);
            const int lin = kr * 64 + g * 16 + p;
            // swizzled source: 16-byte chunk index (lin % 64) / 8 is XORed with kr & 7 when the map swizzles
            int col = lin % 64;
            if (variant != 1) col = (((col / 8) ^ (kr & 7)) * 8) + col % 8;
            const float want = float((kr * 64 + col) % 251);
            if (got != want && bad++ < 5) printf("  mismatch g %d key %d pos %d: got %g want %g\n", g, kr, p, got, want);
        }
    printf("variant %d: nonzero %d (expect %d), mismatches %d\n", variant, nz, variant == 3 ? -1 : 72 * 64 - 0, bad);
    return 0;
}
