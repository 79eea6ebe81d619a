// Shared device/host helpers for the Stage-1 hot-path kernels (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <cuda.h>
#include <stdint.h>
#include <stdio.h>
#include <string.h>
#include <math.h>

#include "../../include/pt_stage1.h"

namespace pt {

// ---- error plumbing (thread-local string behind pt_last_error) -------------
void set_error(const char* fmt, ...);
int check_launch(const char* what);

#define PT_REQUIRE(cond, ...)                                   \
    do {                                                        \
        if (!(cond)) {                                          \
            pt::set_error(__VA_ARGS__);                         \
            return PT_ERR_INVALID;                              \
        }                                                       \
    } while (0)

constexpr int kNumSMsFallback = 148;
int num_sms();

// ---- programmatic dependent launch (pt_set_dependent_launch) -----------------
// A kernel launched through launch_pdl may become resident while the kernel before it in the stream is still running
// (as soon as every CTA of that kernel has executed pdl_trigger() or exited): its prologue — barrier init, TMEM
// allocation, tensor-map prefetch, index arithmetic — overlaps the predecessor's tail.  Such a kernel must execute
// pdl_wait() in EVERY CTA before its first global-memory access that is not to its own parameters (the wait returns
// once the predecessor grid has completed and its writes are visible), and before it exits.  Both instructions are
// no-ops in a kernel that was launched the ordinary way.
enum { kPdlGemm = 1, kPdlNorm = 2, kPdlDecodeAttn = 4 };   // bits of pt_set_dependent_launch
bool pdl_enabled(int kind);
#ifdef __CUDACC__
__device__ __forceinline__ void pdl_trigger() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(int kind, void (*kernel)(KArgs...), dim3 grid, dim3 block, size_t smem, cudaStream_t st,
                              Args&&... args) {
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = grid;
    cfg.blockDim = block;
    cfg.dynamicSmemBytes = smem;
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled(kind) ? 1 : 0;
    return cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
}
#endif

// ---- small device helpers ---------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ float warp_max(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    return v;
}

// Block-wide sum for blockDim.x <= 1024 (multiple of 32). `sh` needs 32 floats.
__device__ __forceinline__ float block_sum(float v, float* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    float r = (lane < nw) ? sh[lane] : 0.f;
    r = warp_sum(r);
    return r;
}
__device__ __forceinline__ float block_max(float v, float* sh) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    v = warp_max(v);
    __syncthreads();
    if (lane == 0) sh[w] = v;
    __syncthreads();
    const int nw = (blockDim.x + 31) >> 5;
    float r = (lane < nw) ? sh[lane] : -INFINITY;
    r = warp_max(r);
    return r;
}

__device__ __forceinline__ float bf16_lo(uint32_t u) { return __uint_as_float(u << 16); }
__device__ __forceinline__ float bf16_hi(uint32_t u) { return __uint_as_float(u & 0xffff0000u); }
__device__ __forceinline__ uint32_t pack_bf16(float a, float b) {
    __nv_bfloat162 h = __floats2bfloat162_rn(a, b);
    return *reinterpret_cast<uint32_t*>(&h);
}

__device__ __forceinline__ float tanh_fast(float x) {
    float y;
    asm("tanh.approx.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}
// tanh-GELU (SigLIP MLP, Gemma3 GeGLU): 0.5 z (1 + tanh(sqrt(2/pi)(z + 0.044715 z^3)))
__device__ __forceinline__ float gelu_tanh_f(float z) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float u = k0 * (z + k1 * z * z * z);
    return 0.5f * z * (1.f + tanhf(u));
}
// same with the hardware tanh.approx (GEMM epilogues; |err| ~ 1e-3 relative, below bf16 rounding)
__device__ __forceinline__ float gelu_tanh_fast_f(float z) {
    const float u = 0.7978845608028654f * fmaf(0.044715f * z * z, z, z);
    const float h = 0.5f * z;
    return fmaf(h, tanh_fast(u), h);
}
__device__ __forceinline__ float gelu_tanh_grad_f(float z) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    float u = k0 * (z + k1 * z * z * z);
    float t = tanhf(u);
    float du = k0 * (1.f + 3.f * k1 * z * z);
    return 0.5f * (1.f + t) + 0.5f * z * (1.f - t * t) * du;
}
// tanh-GELU value and derivative sharing one hardware tanh (fused GeGLU-backward GEMM epilogue).
//   a = (1 + tanh u) / 2,  u = k0 z (1 + k1 z^2):   gelu = z a,   gelu' = a (1 + z (1 - a) (2 k0 + 6 k0 k1 z^2))
// (1 - tanh^2 = 4 a (1 - a)); 11 FP32 instructions + 1 MUFU per element.
__device__ __forceinline__ void gelu_tanh_fast_both(float z, float& val, float& grad) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    const float z2 = z * z;
    const float t = tanh_fast(z * fmaf(k0 * k1, z2, k0));
    const float a = fmaf(0.5f, t, 0.5f);
    val = z * a;
    const float r = z * fmaf(6.f * k0 * k1, z2, 2.f * k0);
    grad = a * fmaf(r, 1.f - a, 1.f);
}
// Gated-unit activation with the constants as data, so one GEMM epilogue serves GeGLU and SwiGLU:
//   a = (1 + tanh u) / 2,  u = z (c0 + c1 z^2):  act = z a,  act' = a (1 + z (1 - a) (c2 + c3 z^2))
// tanh-GELU: c = (k0, k0 k1, 2 k0, 6 k0 k1);  SiLU (sigmoid(z) = (1 + tanh(z / 2)) / 2): c = (0.5, 0, 1, 0).
struct GluConsts { float c0, c1, c2, c3; };
__host__ __device__ inline GluConsts glu_consts(bool silu) {
    const float k0 = 0.7978845608028654f, k1 = 0.044715f;
    return silu ? GluConsts{0.5f, 0.f, 1.f, 0.f} : GluConsts{k0, k0 * k1, 2.f * k0, 6.f * k0 * k1};
}
__device__ __forceinline__ float glu_act_fast(float z, const GluConsts& c) {
    const float u = z * fmaf(c.c1, z * z, c.c0);
    const float h = 0.5f * z;
    return fmaf(h, tanh_fast(u), h);
}
__device__ __forceinline__ void glu_act_fast_both(float z, const GluConsts& c, float& val, float& grad) {
    const float z2 = z * z;
    const float t = tanh_fast(z * fmaf(c.c1, z2, c.c0));
    const float a = fmaf(0.5f, t, 0.5f);
    val = z * a;
    const float r = z * fmaf(c.c3, z2, c.c2);
    grad = a * fmaf(r, 1.f - a, 1.f);
}
// The same on two values per instruction: Blackwell's packed-f32 FMA pipe (FFMA2 / FMUL2 via fma/mul .f32x2 on a
// 64-bit register pair) halves the issue slots of the epilogue, which is what bounds the GeGLU-backward GEMM.
// A bf16x2 word unpacks straight into such a pair {w << 16, w & 0xffff0000}.  Per-element operations and their
// rounding are those of glu_act_fast_both.
__device__ __forceinline__ uint64_t f2_pack(float lo, float hi) {
    uint64_t r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(lo), "f"(hi));
    return r;
}
__device__ __forceinline__ void f2_unpack(uint64_t v, float& lo, float& hi) {
    asm("mov.b64 {%0, %1}, %2;" : "=f"(lo), "=f"(hi) : "l"(v));
}
__device__ __forceinline__ uint64_t f2_mul(uint64_t a, uint64_t b) {
    uint64_t r;
    asm("mul.rn.ftz.f32x2 %0, %1, %2;" : "=l"(r) : "l"(a), "l"(b));
    return r;
}
__device__ __forceinline__ uint64_t f2_fma(uint64_t a, uint64_t b, uint64_t c) {
    uint64_t r;
    asm("fma.rn.ftz.f32x2 %0, %1, %2, %3;" : "=l"(r) : "l"(a), "l"(b), "l"(c));
    return r;
}
struct GluConsts2 { uint64_t c0, c1, c2, c3, half, one, neg_one; };
__device__ __forceinline__ GluConsts2 glu_consts2(const GluConsts& c) {
    return GluConsts2{f2_pack(c.c0, c.c0), f2_pack(c.c1, c.c1), f2_pack(c.c2, c.c2), f2_pack(c.c3, c.c3),
                      f2_pack(0.5f, 0.5f), f2_pack(1.f, 1.f), f2_pack(-1.f, -1.f)};
}
__device__ __forceinline__ void glu_act_fast_both2(uint64_t z, const GluConsts2& c, uint64_t& val, uint64_t& grad) {
    const uint64_t z2 = f2_mul(z, z);
    float w0, w1;
    f2_unpack(f2_mul(z, f2_fma(c.c1, z2, c.c0)), w0, w1);
    const uint64_t t = f2_pack(tanh_fast(w0), tanh_fast(w1));
    const uint64_t a = f2_fma(c.half, t, c.half);
    val = f2_mul(z, a);
    const uint64_t r = f2_mul(z, f2_fma(c.c3, z2, c.c2));
    grad = f2_mul(a, f2_fma(r, f2_fma(a, c.neg_one, c.one), c.one));
}
// value only (GeGLU-forward epilogue): act = z a on a packed pair, 5 packed FP instructions + 2 MUFU per pair
__device__ __forceinline__ uint64_t glu_act_fast2(uint64_t z, const GluConsts2& c) {
    float w0, w1;
    f2_unpack(f2_mul(z, f2_fma(c.c1, f2_mul(z, z), c.c0)), w0, w1);
    const uint64_t h = f2_mul(c.half, z);
    return f2_fma(h, f2_pack(tanh_fast(w0), tanh_fast(w1)), h);
}
// {round_bf16(a), round_bf16(b)} as a packed pair: one cvt.rn.bf16x2 and the two-instruction unpack
__device__ __forceinline__ uint64_t round_bf16_pair(float a, float b) {
    const uint32_t w = pack_bf16(a, b);
    return f2_pack(bf16_lo(w), bf16_hi(w));
}
// exact SiLU for the unfused kernels
__device__ __forceinline__ float silu_f(float z) { return z / (1.f + __expf(-z)); }
__device__ __forceinline__ float silu_grad_f(float z) {
    const float s = 1.f / (1.f + __expf(-z));
    return s * (1.f + z * (1.f - s));
}
__device__ __forceinline__ float round_bf16(float x) { return __bfloat162float(__float2bfloat16_rn(x)); }
// exact erf-GELU (projector): 0.5 z (1 + erf(z / sqrt 2))
__device__ __forceinline__ float gelu_erf_f(float z) {
    return 0.5f * z * (1.f + erff(z * 0.7071067811865476f));
}
__device__ __forceinline__ float gelu_erf_grad_f(float z) {
    float cdf = 0.5f * (1.f + erff(z * 0.7071067811865476f));
    float pdf = 0.3989422804014327f * __expf(-0.5f * z * z);
    return cdf + z * pdf;
}

}  // namespace pt
