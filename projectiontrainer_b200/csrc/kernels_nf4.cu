// NF4 (4-bit NormalFloat, double-quantised) frozen LLM weights (second half).
//
//   ref:Stage1/train_projection_stage1.py:192-210 — `--enable_qlora` loads the frozen LLM through bitsandbytes:
//     BitsAndBytesConfig(load_in_4bit=True, bnb_4bit_quant_type="nf4", bnb_4bit_compute_dtype=bf16,
//                        bnb_4bit_use_double_quant=True)
//   bitsandbytes 0.45 (not in /root/reference, not in this image; restated from its published source):
//     functional.py quantize_4bit / dequantize_4bit / quantize_blockwise / create_dynamic_map,
//     csrc/kernels.cu kQuantizeBlockwise (blocksize 64, reciprocal-multiply), dQuantizeNF4 (midpoint tree),
//     dQuantize<0> (bisection over the 8-bit dynamic map), kDequantizeBlockwise.
//
// B200-first use of the format: 180 GB of HBM holds the bf16 copy of every LLM the reference's scripts name, so
// the 4-bit codes are expanded ONCE at load time (these kernels) and the step runs the ordinary bf16 tcgen05 GEMMs
// on exactly the values bitsandbytes' Linear4bit would feed its GEMM:
//     w_bf16 -> fp16 -> codes (64-blocks) + 8-bit absmax (256-blocks) + offset -> fp16 -> bf16.
// All three kernels are byte work bounded by one pass over the weight in HBM.
#include <cuda_fp16.h>

#include "pt_common.cuh"

#define ST(s) static_cast<cudaStream_t>(s)

namespace pt {

__constant__ float kNf4[16] = {-1.0f, -0.6961928009986877f, -0.5250730514526367f, -0.39491748809814453f,
                               -0.28444138169288635f, -0.18477343022823334f, -0.09105003625154495f, 0.0f,
                               0.07958029955625534f, 0.16093020141124725f, 0.24611230194568634f,
                               0.33791524171829224f, 0.44070982933044434f, 0.5626170039176941f,
                               0.7229568362236023f, 1.0f};

// dQuantizeNF4: the code index is the number of midpoints strictly below x (every comparison is `>`); a NaN
// (all-zero block: 0 * inf) fails them all and lands on code 0.  Evaluated with one table look-up and ONE exact
// comparison: k = round(128 x + 128) (taken from the mantissa of an FMA against 2^23: a single, exact rounding)
// names a cell of width 1/128, narrower than any gap between midpoints, so a cell holds at most one of them;
// kNf4Lut[k] = number of midpoints below the cell's lower edge, and `x > mid[that count]` settles the one midpoint
// the cell may contain.  An x that sits exactly on a cell edge may go to either cell: both give the same count,
// because no midpoint lies on an edge (nearest: 4.4e-4 away).  x is clamped to [-1, 1] first (+-inf from a
// subnormal absmax; NaN -> -1 -> code 0), which does not change the count.
constexpr float kNf4MidHost[15] = {-0.8480964004993439f, -0.6106329262256622f, -0.4599952697753906f,
                                   -0.33967943489551544f, -0.23460740596055984f, -0.13791173323988914f,
                                   -0.045525018125772476f, 0.03979014977812767f, 0.1202552504837513f,
                                   0.2035212516784668f, 0.2920137718319893f, 0.3893125355243683f,
                                   0.5016634166240692f, 0.6427869200706482f, 0.8614784181118011f};
struct Nf4Tables {
    float mid[16];
    unsigned char lut[512];
};
constexpr Nf4Tables make_nf4_tables() {
    Nf4Tables t{};
    for (int i = 0; i < 15; ++i) t.mid[i] = kNf4MidHost[i];
    t.mid[15] = 2.0f;
    for (int k = 0; k <= 256; ++k) {
        const double edge = (k - 128.5) / 128.0;
        unsigned char c = 0;
        for (int i = 0; i < 15; ++i) c += static_cast<double>(kNf4MidHost[i]) < edge ? 1 : 0;
        t.lut[k] = c;
    }
    return t;
}
__constant__ Nf4Tables kNf4Tables = make_nf4_tables();

__device__ __forceinline__ unsigned nf4_code(const Nf4Tables& t, float x) {
    const float xc = fminf(fmaxf(x, -1.0f), 1.0f);
    const unsigned k = __float_as_uint(__fmaf_rn(xc, 128.0f, 8388608.0f + 128.0f)) & 0x1ffu;
    const unsigned c = t.lut[k];
    return c + (xc > t.mid[c] ? 1u : 0u);
}

// Eight lanes per 64-value block (a warp covers four blocks, a CTA 32): each lane loads 8 consecutive bf16 as one
// 16-byte word and stores its 4 packed bytes as one 32-bit word, so that enough bytes are in flight per SM to
// stream at HBM speed.
__global__ void __launch_bounds__(256)
nf4_quantize_kernel(const __nv_bfloat16* __restrict__ w, long long nblocks, int via_fp16,
                    uint8_t* __restrict__ packed, float* __restrict__ absmax) {
    __shared__ Nf4Tables tab;
    if (threadIdx.x < sizeof(Nf4Tables) / 4)
        reinterpret_cast<unsigned*>(&tab)[threadIdx.x] = reinterpret_cast<const unsigned*>(&kNf4Tables)[threadIdx.x];
    __syncthreads();
    const long long t = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;      // one thread per 8 values
    const long long blk = t >> 3;
    const bool on = blk < nblocks;
    uint4 raw = make_uint4(0u, 0u, 0u, 0u);
    if (on) raw = reinterpret_cast<const uint4*>(w)[t];
    const unsigned words[4] = {raw.x, raw.y, raw.z, raw.w};
    float v[8];
    float m = 0.0f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        v[2 * j] = __uint_as_float(words[j] << 16);                // bf16 -> f32 is a shift
        v[2 * j + 1] = __uint_as_float(words[j] & 0xffff0000u);
    }
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        if (via_fp16) {                                            // packed conversions: full-rate instructions
            const float2 r = __half22float2(__floats2half2_rn(v[2 * j], v[2 * j + 1]));
            v[2 * j] = r.x; v[2 * j + 1] = r.y;
        }
        m = fmaxf(m, fmaxf(fabsf(v[2 * j]), fabsf(v[2 * j + 1])));
    }
#pragma unroll
    for (int o = 4; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    const float inv = __fdiv_rn(1.0f, m);
    unsigned out = 0;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        const unsigned hi = nf4_code(tab, __fmul_rn(v[2 * j], inv)), lo = nf4_code(tab, __fmul_rn(v[2 * j + 1], inv));
        out |= ((hi << 4) | lo) << (8 * j);                        // byte j of the little-endian word
    }
    if (on) {
        reinterpret_cast<unsigned*>(packed)[t] = out;
        if ((threadIdx.x & 7) == 0) absmax[blk] = m;
    }
}

// dQuantize<0>: seven-step bisection over the sorted 256-entry map, then the nearer of the landing code and its
// bracketing neighbour.
__device__ __forceinline__ unsigned dquantize8(const float* code, float x) {
    int pivot = 127, upper_pivot = 255, lower_pivot = 0;
    float lower = -1.0f, upper = 1.0f, val = code[pivot];
#pragma unroll
    for (int i = 64; i > 0; i >>= 1) {
        if (x > val) { lower_pivot = pivot; lower = val; pivot += i; }
        else { upper_pivot = pivot; upper = val; pivot -= i; }
        val = code[pivot];
    }
    if (upper_pivot == 255) upper = code[upper_pivot];
    if (lower_pivot == 0) lower = code[lower_pivot];
    if (x > val) return x > __fmul_rn(__fadd_rn(upper, val), 0.5f) ? upper_pivot : pivot;
    return x < __fmul_rn(__fadd_rn(lower, val), 0.5f) ? lower_pivot : pivot;
}

// One CTA of 256 threads per 256 first-level absmax values: (absmax - offset) -> 8-bit dynamic codes + absmax2.
__global__ void __launch_bounds__(256)
nf4_absmax_quantize_kernel(const float* __restrict__ absmax, long long nb, float offset,
                           const float* __restrict__ code256, uint8_t* __restrict__ qabsmax,
                           float* __restrict__ absmax2) {
    __shared__ float code[256];
    __shared__ float red[8];
    code[threadIdx.x] = code256[threadIdx.x];
    const long long i = static_cast<long long>(blockIdx.x) * 256 + threadIdx.x;
    const bool on = i < nb;
    const float v = on ? __fadd_rn(absmax[i], -offset) : 0.0f;
    float m = fabsf(v);
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) m = fmaxf(m, __shfl_xor_sync(0xffffffffu, m, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = m;
    __syncthreads();
    m = red[0];
#pragma unroll
    for (int k = 1; k < 8; ++k) m = fmaxf(m, red[k]);
    if (on) qabsmax[i] = static_cast<uint8_t>(dquantize8(code, __fmul_rn(v, __fdiv_rn(1.0f, m))));
    if (threadIdx.x == 0) absmax2[blockIdx.x] = m;
}

// A CTA expands 1024 packed words (8192 values); a thread takes words tid, tid + 256, ... so that loads and the
// 16-byte stores stay coalesced, and issues its four loads (and their statistics) before the arithmetic so that
// enough bytes are in flight.  absmax = code[q] * absmax2 (fp32 product), + offset (fp32 sum), value = NF4[c] *
// absmax rounded to fp16 (the dtype bitsandbytes quantised from) and then to bf16 (the GEMM's dtype).
__global__ void __launch_bounds__(256)
nf4_dequantize_kernel(const uint8_t* __restrict__ packed, const uint8_t* __restrict__ qabsmax,
                      const float* __restrict__ absmax2, float offset, const float* __restrict__ code256,
                      long long nwords, int via_fp16, __nv_bfloat16* __restrict__ out) {
    __shared__ float nf4[16];
    if (threadIdx.x < 16) nf4[threadIdx.x] = kNf4[threadIdx.x];
    __syncthreads();
    const long long base = static_cast<long long>(blockIdx.x) * 1024 + threadIdx.x;
    unsigned word[4];
    float a[4];
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const long long t = base + u * 256;
        word[u] = 0u; a[u] = 0.0f;
        if (t < nwords) {
            const long long blk = t >> 3;
            word[u] = reinterpret_cast<const unsigned*>(packed)[t];
            a[u] = __fadd_rn(__fmul_rn(__ldg(code256 + qabsmax[blk]), absmax2[blk >> 8]), offset);
        }
    }
#pragma unroll
    for (int u = 0; u < 4; ++u) {
        const long long t = base + u * 256;
        unsigned res[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const unsigned byte = (word[u] >> (8 * j)) & 0xffu;
            float hi = __fmul_rn(nf4[byte >> 4], a[u]), lo = __fmul_rn(nf4[byte & 15u], a[u]);
            if (via_fp16) {
                const float2 r = __half22float2(__floats2half2_rn(hi, lo));
                hi = r.x; lo = r.y;
            }
            const __nv_bfloat162 pr = __floats2bfloat162_rn(hi, lo);
            res[j] = *reinterpret_cast<const unsigned*>(&pr);
        }
        if (t < nwords) reinterpret_cast<uint4*>(out)[t] = make_uint4(res[0], res[1], res[2], res[3]);
    }
}

}  // namespace pt

using namespace pt;

extern "C" int pt_nf4_quantize(const void* w_bf16, long long n, int via_fp16, uint8_t* packed, float* absmax,
                               void* stream) {
    PT_REQUIRE(w_bf16 && packed && absmax, "pt_nf4_quantize: null pointer");
    PT_REQUIRE(n > 0 && n % 64 == 0, "pt_nf4_quantize: the element count must be a positive multiple of 64");
    PT_REQUIRE(reinterpret_cast<uintptr_t>(w_bf16) % 16 == 0 && reinterpret_cast<uintptr_t>(packed) % 4 == 0,
               "pt_nf4_quantize: the weight must be 16-byte aligned and the code buffer 4-byte aligned");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long nblocks = n / 64;
    nf4_quantize_kernel<<<static_cast<unsigned>((nblocks + 31) / 32), 256, 0, ST(stream)>>>(
        static_cast<const __nv_bfloat16*>(w_bf16), nblocks, via_fp16, packed, absmax);
    return check_launch("nf4_quantize");
}

extern "C" int pt_nf4_absmax_quantize(const float* absmax, long long nb, float offset, const float* code256,
                                      uint8_t* qabsmax, float* absmax2, void* stream) {
    PT_REQUIRE(absmax && code256 && qabsmax && absmax2, "pt_nf4_absmax_quantize: null pointer");
    PT_REQUIRE(nb > 0, "pt_nf4_absmax_quantize: no blocks");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    nf4_absmax_quantize_kernel<<<static_cast<unsigned>((nb + 255) / 256), 256, 0, ST(stream)>>>(
        absmax, nb, offset, code256, qabsmax, absmax2);
    return check_launch("nf4_absmax_quantize");
}

extern "C" int pt_nf4_dequantize(const uint8_t* packed, const uint8_t* qabsmax, const float* absmax2, float offset,
                                 const float* code256, long long n, int via_fp16, void* out_bf16, void* stream) {
    PT_REQUIRE(packed && qabsmax && absmax2 && code256 && out_bf16, "pt_nf4_dequantize: null pointer");
    PT_REQUIRE(n > 0 && n % 64 == 0, "pt_nf4_dequantize: the element count must be a positive multiple of 64");
    PT_REQUIRE(reinterpret_cast<uintptr_t>(out_bf16) % 16 == 0 && reinterpret_cast<uintptr_t>(packed) % 4 == 0,
               "pt_nf4_dequantize: the output must be 16-byte aligned and the code buffer 4-byte aligned");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long nwords = n / 8;
    nf4_dequantize_kernel<<<static_cast<unsigned>((nwords + 1023) / 1024), 256, 0, ST(stream)>>>(
        packed, qabsmax, absmax2, offset, code256, nwords, via_fp16, static_cast<__nv_bfloat16*>(out_bf16));
    return check_launch("nf4_dequantize");
}
