// Bandwidth-bound normalisation kernels: LayerNorm (SigLIP) and the Gemma3
// (1+w) RMSNorm family, forward and backward.  One warp per row, 128-bit loads,
// fp32 statistics, rows re-read from L1/L2 rather than cached in registers so
// any row length that is a multiple of 8 works.
#include "pt_common.cuh"

namespace pt {

constexpr int kWarpsPerBlock = 8;

struct RowMap {     // output row r  ->  input row (r / per) * stride + off + r % per
    int per, stride, off;
    __device__ __forceinline__ long long map(int r) const {
        return per > 0 ? static_cast<long long>(r / per) * stride + off + r % per : r;
    }
};

__device__ __forceinline__ void load8(const float* p, float (&v)[8]) {
    const float4 a = *reinterpret_cast<const float4*>(p);
    const float4 b = *reinterpret_cast<const float4*>(p + 4);
    v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
__device__ __forceinline__ void load8(const __nv_bfloat16* p, float (&v)[8]) {
    const uint4 u = *reinterpret_cast<const uint4*>(p);
    v[0] = bf16_lo(u.x); v[1] = bf16_hi(u.x); v[2] = bf16_lo(u.y); v[3] = bf16_hi(u.y);
    v[4] = bf16_lo(u.z); v[5] = bf16_hi(u.z); v[6] = bf16_lo(u.w); v[7] = bf16_hi(u.w);
}
__device__ __forceinline__ void store8(__nv_bfloat16* p, const float (&v)[8]) {
    uint4 u;
    u.x = pack_bf16(v[0], v[1]); u.y = pack_bf16(v[2], v[3]);
    u.z = pack_bf16(v[4], v[5]); u.w = pack_bf16(v[6], v[7]);
    *reinterpret_cast<uint4*>(p) = u;
}
__device__ __forceinline__ void store8(float* p, const float (&v)[8]) {
    *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
    *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}

// ---- LayerNorm forward: f32 in, bf16 out; optionally drops token 0 of every image -------------
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
layernorm_fwd_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                     const float* __restrict__ beta, __nv_bfloat16* __restrict__ out, int rows_out,
                     int D, float eps, RowMap in_map) {
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (r >= rows_out) return;
    const int lane = threadIdx.x & 31;
    const float* xr = x + in_map.map(r) * D;
    float s = 0.f;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8]; load8(xr + i, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) s += v[j];
    }
    const float mean = warp_sum(s) / D;
    float q = 0.f;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8]; load8(xr + i, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float d = v[j] - mean; q += d * d; }
    }
    const float rstd = rsqrtf(warp_sum(q) / D + eps);
    __nv_bfloat16* orow = out + static_cast<long long>(r) * D;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8], g[8], b[8], o[8];
        load8(xr + i, v); load8(gamma + i, g); load8(beta + i, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[j] - mean) * rstd * g[j] + b[j];
        store8(orow + i, o);
    }
}

// Block-per-row variant (D <= 4096): 8 columns per thread, the row is read once and stays in registers,
// mean and variance are reduced through shared memory.
__global__ void __launch_bounds__(512)
layernorm_fwd_row_kernel(const float* __restrict__ x, const float* __restrict__ gamma,
                         const float* __restrict__ beta, __nv_bfloat16* __restrict__ out, int D, float eps,
                         RowMap in_map) {
    __shared__ float red[16];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const int r = blockIdx.x;
    const int i = threadIdx.x * 8;
    const bool on = i < D;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
    if (on) load8(x + in_map.map(r) * D + i, v);
    float s = ((v[0] + v[1]) + (v[2] + v[3])) + ((v[4] + v[5]) + (v[6] + v[7]));
    s = warp_sum(s);
    if (lane == 0) red[warp] = s;
    __syncthreads();
    float tot = 0.f;
    for (int w = 0; w < nwarps; ++w) tot += red[w];
    __syncthreads();
    const float mean = tot / D;
    float q = 0.f;
    if (on) {
#pragma unroll
        for (int j = 0; j < 8; ++j) { const float d = v[j] - mean; q += d * d; }
    }
    q = warp_sum(q);
    if (lane == 0) red[warp] = q;
    __syncthreads();
    tot = 0.f;
    for (int w = 0; w < nwarps; ++w) tot += red[w];
    const float rstd = rsqrtf(tot / D + eps);
    if (on) {
        float g[8], b[8], o[8];
        load8(gamma + i, g); load8(beta + i, b);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = (v[j] - mean) * rstd * g[j] + b[j];
        store8(out + static_cast<long long>(r) * D + i, o);
    }
}

// ---- RMSNorm forward: out = x * rsqrt(mean x^2 + eps) * (1 + w) ------------------------------
template <typename TIn>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
rmsnorm_fwd_kernel(const TIn* __restrict__ x, const float* __restrict__ w,
                   __nv_bfloat16* __restrict__ out, int rows_out, int D, float eps, RowMap in_map) {
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (r >= rows_out) return;
    const int lane = threadIdx.x & 31;
    const TIn* xr = x + in_map.map(r) * D;
    float q = 0.f;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8]; load8(xr + i, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) q += v[j] * v[j];
    }
    const float rstd = rsqrtf(warp_sum(q) / D + eps);
    __nv_bfloat16* orow = out + static_cast<long long>(r) * D;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8], g[8], o[8];
        load8(xr + i, v); load8(w + i, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = v[j] * rstd * (1.f + g[j]);
        store8(orow + i, o);
    }
}

// ---- fused sandwich step: x_out = x_in + RMS(y; w_post); nxt = RMS(x_out; w_next) ------------
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
rmsnorm_residual_fwd_kernel(const __nv_bfloat16* __restrict__ y, const float* __restrict__ w_post,
                            const float* __restrict__ x_in, float* __restrict__ x_out,
                            const float* __restrict__ w_next, __nv_bfloat16* __restrict__ nxt,
                            int rows, int D, float eps) {
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    const long long base = static_cast<long long>(r) * D;
    float q = 0.f;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8]; load8(y + base + i, v);
#pragma unroll
        for (int j = 0; j < 8; ++j) q += v[j] * v[j];
    }
    const float rstd = rsqrtf(warp_sum(q) / D + eps);
    float q2 = 0.f;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8], g[8], xi[8], o[8];
        load8(y + base + i, v); load8(w_post + i, g); load8(x_in + base + i, xi);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            // HF rounds the norm output to the activation dtype before the residual add
            const float n = __bfloat162float(__float2bfloat16(v[j] * rstd * (1.f + g[j])));
            o[j] = xi[j] + n;
            q2 += o[j] * o[j];
        }
        store8(x_out + base + i, o);
    }
    if (!nxt) return;
    const float rstd2 = rsqrtf(warp_sum(q2) / D + eps);
    __syncwarp();
    for (int i = lane * 8; i < D; i += 256) {
        float v[8], g[8], o[8];
        load8(x_out + base + i, v); load8(w_next + i, g);   // same lane wrote these addresses
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = v[j] * rstd2 * (1.f + g[j]);
        store8(nxt + base + i, o);
    }
}

// Block-per-row variant of the sandwich step (D <= 4096): every operand is read once, the row stays in registers.
// TY = float: y is a split-K accumulator (f32 partial sums added up by the GEMM slices): it is rounded to bf16 as the
// GEMM's own bf16 output would have been, and CLEARED after the read, ready for the next GEMM's slices.
template <typename TY>
__global__ void __launch_bounds__(512)
rmsnorm_residual_fwd_row_kernel(TY* __restrict__ y, const float* __restrict__ w_post,
                                const float* x_in, float* x_out, const float* __restrict__ w_next,
                                __nv_bfloat16* __restrict__ nxt, int D, float eps) {
    __shared__ float red[16];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const long long base = static_cast<long long>(blockIdx.x) * D;
    const int i = threadIdx.x * 8;
    const bool on = i < D;
    float v[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f}, xi[8], g[8];
    float q = 0.f;
    pdl_trigger();
    pdl_wait();
    if (on) {
        load8(y + base + i, v); load8(x_in + base + i, xi); load8(w_post + i, g);
        if constexpr (sizeof(TY) == 4) {
            const float z[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
            store8(y + base + i, z);
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = round_bf16(v[j]);
        }
#pragma unroll
        for (int j = 0; j < 8; ++j) q += v[j] * v[j];
    }
    q = warp_sum(q);
    if (lane == 0) red[warp] = q;
    __syncthreads();
    float tot = 0.f;
    for (int w = 0; w < nwarps; ++w) tot += red[w];
    __syncthreads();
    const float rstd = rsqrtf(tot / D + eps);
    float q2 = 0.f;
    if (on) {
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            // HF rounds the norm output to the activation dtype before the residual add
            const float n = __bfloat162float(__float2bfloat16(v[j] * rstd * (1.f + g[j])));
            v[j] = xi[j] + n;
            q2 += v[j] * v[j];
        }
        store8(x_out + base + i, v);
    }
    if (!nxt) return;
    q2 = warp_sum(q2);
    if (lane == 0) red[warp] = q2;
    __syncthreads();
    tot = 0.f;
    for (int w = 0; w < nwarps; ++w) tot += red[w];
    const float rstd2 = rsqrtf(tot / D + eps);
    if (on) {
        float o[8];
        load8(w_next + i, g);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = v[j] * rstd2 * (1.f + g[j]);
        store8(nxt + base + i, o);
    }
}

// ---- RMSNorm backward ------------------------------------------------------------------------
//   n = u r,  r = rsqrt(mean u^2 + eps),  out = n (1 + w)
//   dn = g (1 + w);  du = r (dn - n mean(dn n))
// mode 0: dst_bf16[row] = du     mode 1: dst_f32[map(row)] = add_f32[row] + du  (add may be null / alias)
template <typename TG, typename TU>
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
rmsnorm_bwd_kernel(const TG* __restrict__ g, const TU* __restrict__ u, const float* __restrict__ w,
                   __nv_bfloat16* dst_bf16, float* dst_f32, const float* add_f32, int rows, int D,
                   float eps, RowMap u_map, RowMap dst_map) {
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (r >= rows) return;
    const int lane = threadIdx.x & 31;
    const TG* gr = g + static_cast<long long>(r) * D;
    const TU* ur = u + u_map.map(r) * D;
    float q = 0.f, dot = 0.f;
    for (int i = lane * 8; i < D; i += 256) {
        float gv[8], uv[8], wv[8];
        load8(gr + i, gv); load8(ur + i, uv); load8(w + i, wv);
#pragma unroll
        for (int j = 0; j < 8; ++j) { q += uv[j] * uv[j]; dot += gv[j] * (1.f + wv[j]) * uv[j]; }
    }
    q = warp_sum(q); dot = warp_sum(dot);
    const float rstd = rsqrtf(q / D + eps);
    const float c = dot * rstd * rstd * rstd / D;     // mean(dn n) r / ... folded: du = r dn - u c
    const long long drow = dst_map.map(r) * D;
    for (int i = lane * 8; i < D; i += 256) {
        float gv[8], uv[8], wv[8], o[8];
        load8(gr + i, gv); load8(ur + i, uv); load8(w + i, wv);
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * gv[j] * (1.f + wv[j]) - uv[j] * c;
        if (dst_bf16) {
            store8(dst_bf16 + drow + i, o);
        } else {
            if (add_f32) {
                float a[8]; load8(add_f32 + drow + i, a);
#pragma unroll
                for (int j = 0; j < 8; ++j) o[j] += a[j];
            }
            store8(dst_f32 + drow + i, o);
        }
    }
}

// Two consecutive norm backwards of the sandwich-norm decoder layer in one pass over the row
// (HF:models/gemma3/modeling_gemma3.py:417-432 reversed): the gradient g1 of a PRE-norm output flows into the
// f32 residual-stream gradient, which is immediately the incoming gradient of the preceding POST-norm:
//     dx   += rmsnorm_bwd(g1; u1, w1)            (u1 = f32 residual stream the pre-norm read)
//     out2  = rmsnorm_bwd(dx; u2, w2)  as bf16   (u2 = bf16 branch output the post-norm read)
// The updated dx row stays in registers between the two (NCH chunks of 256 columns per warp).
// One block per row, 8 columns per thread (blockDim = D / 8 rounded up to a warp): few registers, many rows
// in flight per SM; the two (sum u^2, sum dn u) reductions go through shared memory.
__device__ __forceinline__ void block_sum2(float& a, float& b, float (*red)[2], int lane, int warp, int nwarps) {
    a = warp_sum(a); b = warp_sum(b);
    if (lane == 0) { red[warp][0] = a; red[warp][1] = b; }
    __syncthreads();
    float x = 0.f, y = 0.f;
    for (int w = 0; w < nwarps; ++w) { x += red[w][0]; y += red[w][1]; }
    __syncthreads();
    a = x; b = y;
}

__global__ void __launch_bounds__(320)
rmsnorm_bwd_pair_kernel(const __nv_bfloat16* g1, const float* __restrict__ u1,
                        const float* __restrict__ w1, float* dx, const __nv_bfloat16* __restrict__ u2,
                        const float* __restrict__ w2, __nv_bfloat16* out2 /* may alias g1 */, int D, float eps) {
    __shared__ float red[10][2];
    const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5, nwarps = blockDim.x >> 5;
    const long long base = static_cast<long long>(blockIdx.x) * D;
    const int i = threadIdx.x * 8;
    const bool on = i < D;
    float dn[8], uu[8], a[8], wv[8];
    float q = 0.f, dot = 0.f;
    if (on) {
        float gv[8];
        load8(g1 + base + i, gv); load8(u1 + base + i, uu); load8(w1 + i, wv); load8(dx + base + i, a);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            dn[j] = gv[j] * (1.f + wv[j]);
            q += uu[j] * uu[j];
            dot += dn[j] * uu[j];
        }
    }
    block_sum2(q, dot, red, lane, warp, nwarps);
    float rstd = rsqrtf(q / D + eps);
    float cc = dot * rstd * rstd * rstd / D;
    q = 0.f; dot = 0.f;
    if (on) {
#pragma unroll
        for (int j = 0; j < 8; ++j) a[j] += rstd * dn[j] - uu[j] * cc;
        store8(dx + base + i, a);
        load8(u2 + base + i, uu); load8(w2 + i, wv);
#pragma unroll
        for (int j = 0; j < 8; ++j) {
            dn[j] = a[j] * (1.f + wv[j]);
            q += uu[j] * uu[j];
            dot += dn[j] * uu[j];
        }
    }
    block_sum2(q, dot, red, lane, warp, nwarps);
    rstd = rsqrtf(q / D + eps);
    cc = dot * rstd * rstd * rstd / D;
    if (on) {
        float o[8];
#pragma unroll
        for (int j = 0; j < 8; ++j) o[j] = rstd * dn[j] - uu[j] * cc;
        store8(out2 + base + i, o);
    }
}

// ---- row gather + cast: f32 strided rows -> compact bf16 rows ----------------------------------
__global__ void __launch_bounds__(kWarpsPerBlock * 32)
gather_cast_rows_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ out, int rows_out, int D,
                        RowMap in_map) {
    const int r = blockIdx.x * kWarpsPerBlock + (threadIdx.x >> 5);
    if (r >= rows_out) return;
    const int lane = threadIdx.x & 31;
    const float* xr = x + in_map.map(r) * D;
    __nv_bfloat16* orow = out + static_cast<long long>(r) * D;
    for (int i = lane * 8; i < D; i += 256) {
        float v[8]; load8(xr + i, v);
        store8(orow + i, v);
    }
}

static inline int row_blocks(int rows) { return (rows + kWarpsPerBlock - 1) / kWarpsPerBlock; }

}  // namespace pt

using namespace pt;

extern "C" int pt_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* out_bf16,
                                int rows_out, int D, float eps, int group_rows_out, int group_rows_in,
                                int group_row0, void* stream) {
    PT_REQUIRE(x && gamma && beta && out_bf16, "pt_layernorm_fwd: null pointer");
    PT_REQUIRE(rows_out > 0 && D > 0 && D % 8 == 0, "pt_layernorm_fwd: D must be a positive multiple of 8");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    RowMap m{group_rows_out, group_rows_in, group_row0};
    if (D <= 4096)
        layernorm_fwd_row_kernel<<<rows_out, ((D / 8 + 31) / 32) * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            x, gamma, beta, static_cast<__nv_bfloat16*>(out_bf16), D, eps, m);
    else
        layernorm_fwd_kernel<<<row_blocks(rows_out), kWarpsPerBlock * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            x, gamma, beta, static_cast<__nv_bfloat16*>(out_bf16), rows_out, D, eps, m);
    return check_launch("layernorm_fwd");
}

extern "C" int pt_rmsnorm_fwd(const void* x, int x_dtype, const float* w, void* out_bf16, int rows_out,
                              int D, float eps, int group_rows_out, int group_rows_in, int group_row0,
                              void* stream) {
    PT_REQUIRE(x && w && out_bf16, "pt_rmsnorm_fwd: null pointer");
    PT_REQUIRE(rows_out > 0 && D > 0 && D % 8 == 0, "pt_rmsnorm_fwd: D must be a positive multiple of 8");
    PT_REQUIRE(x_dtype == PT_BF16 || x_dtype == PT_F32, "pt_rmsnorm_fwd: x_dtype must be bf16 or f32");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    RowMap m{group_rows_out, group_rows_in, group_row0};
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    auto* o = static_cast<__nv_bfloat16*>(out_bf16);
    if (x_dtype == PT_F32)
        rmsnorm_fwd_kernel<float><<<row_blocks(rows_out), kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const float*>(x), w, o, rows_out, D, eps, m);
    else
        rmsnorm_fwd_kernel<__nv_bfloat16><<<row_blocks(rows_out), kWarpsPerBlock * 32, 0, st>>>(
            static_cast<const __nv_bfloat16*>(x), w, o, rows_out, D, eps, m);
    return check_launch("rmsnorm_fwd");
}

extern "C" int pt_rmsnorm_residual_fwd(const void* y_bf16, const float* w_post, const float* x_in,
                                       float* x_out, const float* w_next, void* next_bf16, int rows,
                                       int D, float eps, void* stream) {
    PT_REQUIRE(y_bf16 && w_post && x_in && x_out, "pt_rmsnorm_residual_fwd: null pointer");
    PT_REQUIRE((w_next == nullptr) == (next_bf16 == nullptr), "pt_rmsnorm_residual_fwd: w_next/next mismatch");
    PT_REQUIRE(rows > 0 && D > 0 && D % 8 == 0, "pt_rmsnorm_residual_fwd: D must be a positive multiple of 8");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    if (D <= 4096)
        launch_pdl(kPdlNorm, rmsnorm_residual_fwd_row_kernel<const __nv_bfloat16>, dim3(rows), dim3(((D / 8 + 31) / 32) * 32), 0,
                   static_cast<cudaStream_t>(stream), static_cast<const __nv_bfloat16*>(y_bf16), w_post, x_in, x_out,
                   w_next, static_cast<__nv_bfloat16*>(next_bf16), D, eps);
    else
        rmsnorm_residual_fwd_kernel<<<row_blocks(rows), kWarpsPerBlock * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(y_bf16), w_post, x_in, x_out, w_next,
            static_cast<__nv_bfloat16*>(next_bf16), rows, D, eps);
    return check_launch("rmsnorm_residual_fwd");
}

extern "C" int pt_rmsnorm_residual_fwd_acc(float* y_acc, const float* w_post, const float* x_in, float* x_out,
                                           const float* w_next, void* next_bf16, int rows, int D, float eps,
                                           void* stream) {
    PT_REQUIRE(y_acc && w_post && x_in && x_out, "pt_rmsnorm_residual_fwd_acc: null pointer");
    PT_REQUIRE((w_next == nullptr) == (next_bf16 == nullptr), "pt_rmsnorm_residual_fwd_acc: w_next/next mismatch");
    PT_REQUIRE(rows > 0 && D > 0 && D % 8 == 0 && D <= 4096,
               "pt_rmsnorm_residual_fwd_acc: D must be a positive multiple of 8, at most 4096");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    launch_pdl(kPdlNorm, rmsnorm_residual_fwd_row_kernel<float>, dim3(rows), dim3(((D / 8 + 31) / 32) * 32), 0,
               static_cast<cudaStream_t>(stream), y_acc, w_post, x_in, x_out, w_next,
               static_cast<__nv_bfloat16*>(next_bf16), D, eps);
    return check_launch("rmsnorm_residual_fwd_acc");
}

extern "C" int pt_rmsnorm_bwd(const void* g, int g_dtype, const void* u, int u_dtype, const float* w,
                              void* dst, int dst_dtype, const float* add_f32, int rows, int D, float eps,
                              int u_group_rows, int u_group_stride, int u_group_row0,
                              int dst_group_rows, int dst_group_stride, int dst_group_row0,
                              void* stream) {
    PT_REQUIRE(g && u && w && dst, "pt_rmsnorm_bwd: null pointer");
    PT_REQUIRE(rows > 0 && D > 0 && D % 8 == 0, "pt_rmsnorm_bwd: D must be a positive multiple of 8");
    PT_REQUIRE(dst_dtype == PT_BF16 || dst_dtype == PT_F32, "pt_rmsnorm_bwd: bad dst dtype");
    PT_REQUIRE(dst_dtype == PT_F32 || add_f32 == nullptr, "pt_rmsnorm_bwd: add only with f32 dst");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    RowMap um{u_group_rows, u_group_stride, u_group_row0};
    RowMap dm{dst_group_rows, dst_group_stride, dst_group_row0};
    __nv_bfloat16* db = dst_dtype == PT_BF16 ? static_cast<__nv_bfloat16*>(dst) : nullptr;
    float* df = dst_dtype == PT_F32 ? static_cast<float*>(dst) : nullptr;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int nb = row_blocks(rows), nt = kWarpsPerBlock * 32;
#define PT_LAUNCH_RB(TG, TU)                                                                      \
    rmsnorm_bwd_kernel<TG, TU><<<nb, nt, 0, st>>>(static_cast<const TG*>(g), static_cast<const TU*>(u), \
                                                  w, db, df, add_f32, rows, D, eps, um, dm)
    if (g_dtype == PT_F32 && u_dtype == PT_BF16) PT_LAUNCH_RB(float, __nv_bfloat16);
    else if (g_dtype == PT_BF16 && u_dtype == PT_F32) PT_LAUNCH_RB(__nv_bfloat16, float);
    else if (g_dtype == PT_BF16 && u_dtype == PT_BF16) PT_LAUNCH_RB(__nv_bfloat16, __nv_bfloat16);
    else if (g_dtype == PT_F32 && u_dtype == PT_F32) PT_LAUNCH_RB(float, float);
    else { set_error("pt_rmsnorm_bwd: unsupported dtype combination"); return PT_ERR_INVALID; }
#undef PT_LAUNCH_RB
    return check_launch("rmsnorm_bwd");
}

extern "C" int pt_gather_cast_rows(const float* x, void* out_bf16, int rows_out, int D, int group_rows_out,
                                   int group_rows_in, int group_row0, void* stream) {
    PT_REQUIRE(x && out_bf16 && rows_out > 0 && D > 0 && D % 8 == 0, "pt_gather_cast_rows: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    RowMap m{group_rows_out, group_rows_in, group_row0};
    gather_cast_rows_kernel<<<row_blocks(rows_out), kWarpsPerBlock * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        x, static_cast<__nv_bfloat16*>(out_bf16), rows_out, D, m);
    return check_launch("gather_cast_rows");
}

extern "C" int pt_rmsnorm_bwd_pair(const void* g1_bf16, const float* u1, const float* w1, float* dx,
                                   const void* u2_bf16, const float* w2, void* out2_bf16, int rows, int D,
                                   float eps, void* stream) {
    PT_REQUIRE(g1_bf16 && u1 && w1 && dx && u2_bf16 && w2 && out2_bf16, "pt_rmsnorm_bwd_pair: null pointer");
    PT_REQUIRE(rows > 0 && D > 0 && D % 8 == 0 && D <= 2560, "pt_rmsnorm_bwd_pair: D must be a multiple of 8, <= 2560");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const __nv_bfloat16* g1 = static_cast<const __nv_bfloat16*>(g1_bf16);
    const __nv_bfloat16* u2 = static_cast<const __nv_bfloat16*>(u2_bf16);
    __nv_bfloat16* o2 = static_cast<__nv_bfloat16*>(out2_bf16);
    rmsnorm_bwd_pair_kernel<<<rows, ((D / 8 + 31) / 32) * 32, 0, st>>>(g1, u1, w1, dx, u2, w2, o2, D, eps);
    return check_launch("rmsnorm_bwd_pair");
}
