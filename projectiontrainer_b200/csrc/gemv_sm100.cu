// Linear layers of the cached decode step at SMALL batch (M <= 16 rows: the reference's inference scripts generate for
// one image at a time, ref:Stage1/inference_stage1.py:150-193): y[m][n] = sum_k x[m][k] W[n][k].
//
// The 256-row tcgen05 tile wastes 15/16 of its A operand here and costs ~7 us per launch whatever the size; these
// products are pure weight streaming (HBM-bound: 2 bytes per weight, M <= 16 FMAs each).  So the WEIGHT rows become the
// M dimension of mma.sync.m16n8k16 (16 output features per warp tile), the batch rows its N dimension (one or two
// 8-column tiles), and every weight byte is read exactly once, straight from global memory into the A fragments:
// the contraction index is permuted inside each 32-wide chunk so that a lane's fragments of two k-steps are 16 contiguous
// bytes of its weight row (the same permutation on x leaves the sum unchanged).  A CTA owns 16 features; its 8 warps
// take the 32-wide chunks of K round-robin and add up in shared memory.  Epilogues: bf16 / f32 store, f32 atomic add
// (split-K slices in blockIdx.y, for the o / down projections whose N is small: the accumulator of
// pt_rmsnorm_residual_fwd_acc), and GeGLU / SwiGLU (the CTA streams gate rows n and up rows ff + n and stores
// act(gate) * up; pre-activations rounded to bf16 first, as the GEMM's fused epilogue does).
#include "pt_common.cuh"

namespace pt {

constexpr int kGemvWarps = 8;
enum { GEMV_BF16 = 0, GEMV_F32 = 1, GEMV_F32_ATOMIC = 2, GEMV_GEGLU = 3, GEMV_SWIGLU = 4 };

__device__ __forceinline__ uint4 ldg_stream128(const void* p) {     // weights: read once, keep them out of L1
    uint4 v;
    asm volatile("ld.global.nc.L1::no_allocate.v4.u32 {%0,%1,%2,%3}, [%4];"
                 : "=r"(v.x), "=r"(v.y), "=r"(v.z), "=r"(v.w) : "l"(p));
    return v;
}
__device__ __forceinline__ void mma16816_gemv(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                              uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
                 "{%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}

struct GemvArgs {
    const __nv_bfloat16* x; long long lda; int M;
    const __nv_bfloat16* W; long long ldb; int N, K;     // N = output features (GLU modes: ff; W holds 2 ff rows)
    void* out; long long ldc;
    int chunks_per_slice;                                // 32-wide K chunks per blockIdx.y slice
};

template <int MT, int MODE>
__global__ void __launch_bounds__(kGemvWarps * 32) gemv_small_m_kernel(const GemvArgs p) {
    constexpr bool GLU = MODE == GEMV_GEGLU || MODE == GEMV_SWIGLU;
    constexpr int NS = GLU ? 2 : 1;                      // weight row sets: gate, up
    __shared__ float red[kGemvWarps][NS][16][MT * 8 + 1];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const int r = lane >> 2, j = lane & 3;
    const int f0 = blockIdx.x * 16;
    const int nchunks = p.K / 32;
    const int c0 = blockIdx.y * p.chunks_per_slice;
    const int c1 = min(nchunks, c0 + p.chunks_per_slice);
    // rows f0 + r and f0 + r + 8 of each set (clamped: the tail tile's extra rows are computed and dropped)
    const __nv_bfloat16* wr[NS][2];
#pragma unroll
    for (int s = 0; s < NS; ++s) {
        const long long base = static_cast<long long>(s) * p.N;
        wr[s][0] = p.W + (base + min(f0 + r, p.N - 1)) * p.ldb + j * 8;
        wr[s][1] = p.W + (base + min(f0 + r + 8, p.N - 1)) * p.ldb + j * 8;
    }
    const __nv_bfloat16* xr[MT];
    bool xon[MT];
#pragma unroll
    for (int t = 0; t < MT; ++t) {
        xon[t] = t * 8 + r < p.M;
        xr[t] = p.x + static_cast<long long>(xon[t] ? t * 8 + r : 0) * p.lda + j * 8;
    }
    float acc[NS][MT][4];
#pragma unroll
    for (int s = 0; s < NS; ++s)
#pragma unroll
        for (int t = 0; t < MT; ++t) { acc[s][t][0] = 0.f; acc[s][t][1] = 0.f; acc[s][t][2] = 0.f; acc[s][t][3] = 0.f; }

    pdl_trigger();
    constexpr int U = GLU ? 2 : 4;                       // chunks in flight per warp
    for (int c = c0 + warp; c < c1; c += kGemvWarps * U) {
        uint4 a[U][NS][2];
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int cu = c + u * kGemvWarps;
#pragma unroll
            for (int s = 0; s < NS; ++s) {
                a[u][s][0] = cu < c1 ? ldg_stream128(wr[s][0] + cu * 32) : make_uint4(0u, 0u, 0u, 0u);
                a[u][s][1] = cu < c1 ? ldg_stream128(wr[s][1] + cu * 32) : make_uint4(0u, 0u, 0u, 0u);
            }
        }
        if (c == c0 + warp) pdl_wait();                  // the weights are constants; x is the predecessor's output
#pragma unroll
        for (int u = 0; u < U; ++u) {
            const int cu = c + u * kGemvWarps;
            if (cu < c1) {
#pragma unroll
                for (int t = 0; t < MT; ++t) {
                    const uint4 xb = xon[t] ? *reinterpret_cast<const uint4*>(xr[t] + cu * 32) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
                    for (int s = 0; s < NS; ++s) {
                        mma16816_gemv(acc[s][t], a[u][s][0].x, a[u][s][1].x, a[u][s][0].y, a[u][s][1].y, xb.x, xb.y);
                        mma16816_gemv(acc[s][t], a[u][s][0].z, a[u][s][1].z, a[u][s][0].w, a[u][s][1].w, xb.z, xb.w);
                    }
                }
            }
        }
    }
    pdl_wait();                                          // (a warp without chunks has not waited yet)
    // C fragment: c0, c1 = (feature r, batch 2j, 2j+1), c2, c3 = (feature r + 8, same batch columns)
#pragma unroll
    for (int s = 0; s < NS; ++s)
#pragma unroll
        for (int t = 0; t < MT; ++t) {
            red[warp][s][r][t * 8 + 2 * j] = acc[s][t][0];
            red[warp][s][r][t * 8 + 2 * j + 1] = acc[s][t][1];
            red[warp][s][r + 8][t * 8 + 2 * j] = acc[s][t][2];
            red[warp][s][r + 8][t * 8 + 2 * j + 1] = acc[s][t][3];
        }
    __syncthreads();
    for (int e = threadIdx.x; e < 16 * MT * 8; e += kGemvWarps * 32) {
        const int f = e & 15, m = e >> 4;                // 16 consecutive features per batch row
        if (m >= p.M || f0 + f >= p.N) continue;
        float v[NS];
#pragma unroll
        for (int s = 0; s < NS; ++s) {
            float a = 0.f;
#pragma unroll
            for (int w = 0; w < kGemvWarps; ++w) a += red[w][s][f][m];
            v[s] = a;
        }
        const long long o = static_cast<long long>(m) * p.ldc + f0 + f;
        if constexpr (MODE == GEMV_BF16) {
            static_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(v[0]);
        } else if constexpr (MODE == GEMV_F32) {
            static_cast<float*>(p.out)[o] = v[0];
        } else if constexpr (MODE == GEMV_F32_ATOMIC) {
            atomicAdd(static_cast<float*>(p.out) + o, v[0]);
        } else {
            const float g = round_bf16(v[0]), u = round_bf16(v[1]);
            const float a = MODE == GEMV_SWIGLU ? silu_f(g) : gelu_tanh_fast_f(g);
            static_cast<__nv_bfloat16*>(p.out)[o] = __float2bfloat16_rn(a * u);
        }
    }
}

template <int MT, int MODE>
static int launch_gemv(const GemvArgs& a, int slices, cudaStream_t st) {
    const dim3 grid((a.N + 15) / 16, slices);
    const cudaError_t e = launch_pdl(kPdlGemm, gemv_small_m_kernel<MT, MODE>, grid, dim3(kGemvWarps * 32), 0, st, a);
    if (e != cudaSuccess) { set_error("gemv_small_m launch: %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
    return check_launch("gemv_small_m");
}

}  // namespace pt

using namespace pt;

extern "C" int pt_gemv_small_m(const void* x, long long lda, int M, const void* W, long long ldb, int N, int K, void* out,
                               long long ldc, int mode, int ksplit, void* stream) {
    PT_REQUIRE(x && W && out, "pt_gemv_small_m: null pointer");
    PT_REQUIRE(M >= 1 && M <= 32, "pt_gemv_small_m: M must be 1 .. 32 (larger batches take pt_gemm_bf16)");
    PT_REQUIRE(N >= 1 && K >= 32 && K % 32 == 0, "pt_gemv_small_m: K must be a positive multiple of 32");
    PT_REQUIRE(lda % 8 == 0 && ldb % 8 == 0 && lda >= K && ldb >= K, "pt_gemv_small_m: row strides must be multiples of 8, >= K");
    PT_REQUIRE((reinterpret_cast<uintptr_t>(x) | reinterpret_cast<uintptr_t>(W)) % 16 == 0,
               "pt_gemv_small_m: x and W must be 16-byte aligned");
    PT_REQUIRE(mode >= GEMV_BF16 && mode <= GEMV_SWIGLU, "pt_gemv_small_m: mode must be 0 .. 4");
    PT_REQUIRE(ldc >= N, "pt_gemv_small_m: ldc < N");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const int nchunks = K / 32;
    int slices = 1;
    if (mode == GEMV_F32_ATOMIC) {
        slices = ksplit > 0 ? ksplit : 1;
        if (ksplit <= 0) {                               // about two CTAs per SM
            const int tiles = (N + 15) / 16;
            slices = (2 * num_sms() + tiles - 1) / tiles;
        }
        if (slices > nchunks / kGemvWarps) slices = nchunks / kGemvWarps > 0 ? nchunks / kGemvWarps : 1;
    } else {
        PT_REQUIRE(ksplit <= 1, "pt_gemv_small_m: K slices need mode 2 (f32 atomic add)");
    }
    GemvArgs a;
    a.x = static_cast<const __nv_bfloat16*>(x); a.lda = lda; a.M = M;
    a.W = static_cast<const __nv_bfloat16*>(W); a.ldb = ldb; a.N = N; a.K = K;
    a.out = out; a.ldc = ldc;
    a.chunks_per_slice = (nchunks + slices - 1) / slices;
    slices = (nchunks + a.chunks_per_slice - 1) / a.chunks_per_slice;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
#define PT_GEMV(MT)                                                                              \
    switch (mode) {                                                                              \
        case GEMV_BF16: return launch_gemv<MT, GEMV_BF16>(a, slices, st);                        \
        case GEMV_F32: return launch_gemv<MT, GEMV_F32>(a, slices, st);                          \
        case GEMV_F32_ATOMIC: return launch_gemv<MT, GEMV_F32_ATOMIC>(a, slices, st);            \
        case GEMV_GEGLU: return launch_gemv<MT, GEMV_GEGLU>(a, slices, st);                      \
        default: return launch_gemv<MT, GEMV_SWIGLU>(a, slices, st);                             \
    }
    if (M <= 8) { PT_GEMV(1) } else if (M <= 16) { PT_GEMV(2) } else { PT_GEMV(4) }
#undef PT_GEMV
}
