// Remaining bandwidth-bound kernels of the Stage-1 step: patchify, GeGLU fwd/bwd, erf-GELU
// backward, embedding gather + splice + masks/labels, vocabulary cross-entropy (stats + gradient),
// column sums (bias grads), squared-norm reduction and the fused clip + AdamW update.
#include "pt_common.cuh"

namespace pt {

// ---- patchify: f32 [B,3,H,W] -> bf16 [B*gy*gx, 3*p*p], column order (c, py, px) ---------------
__global__ void patchify_kernel(const float* __restrict__ px, __nv_bfloat16* __restrict__ out, int B,
                                int H, int W, int p) {
    const int gx = W / p, gy = H / p, cols = 3 * p * p;
    const long long n8 = static_cast<long long>(B) * gy * gx * cols / 8;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long e = i * 8;
        const int col = static_cast<int>(e % cols);
        const long long row = e / cols;
        const int c = col / (p * p), py = (col / p) % p, x0 = col % p;
        const int tx = static_cast<int>(row % gx), ty = static_cast<int>((row / gx) % gy);
        const int b = static_cast<int>(row / (static_cast<long long>(gx) * gy));
        const float* src = px + ((static_cast<long long>(b) * 3 + c) * H + ty * p + py) * W + tx * p + x0;
        const float4 a = *reinterpret_cast<const float4*>(src);
        const float4 d = *reinterpret_cast<const float4*>(src + 4);
        uint4 u;
        u.x = pack_bf16(a.x, a.y); u.y = pack_bf16(a.z, a.w);
        u.z = pack_bf16(d.x, d.y); u.w = pack_bf16(d.z, d.w);
        *reinterpret_cast<uint4*>(out + e) = u;
    }
}

// ---- GeGLU: act = gelu_tanh(g) * u with gu = [g | u] per row -----------------------------------
template <bool SILU>
__global__ void geglu_fwd_kernel(const __nv_bfloat16* __restrict__ gu, __nv_bfloat16* __restrict__ act,
                                 long long rows, int ff) {
    auto f = [](float z) { return SILU ? silu_f(z) : gelu_tanh_f(z); };
    const int f8 = ff / 8;
    const long long n = rows * f8;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / f8;
        const int c = static_cast<int>(i - r * f8) * 8;
        const uint4 g = *reinterpret_cast<const uint4*>(gu + r * 2 * ff + c);
        const uint4 u = *reinterpret_cast<const uint4*>(gu + r * 2 * ff + ff + c);
        const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            o[j] = pack_bf16(f(bf16_lo(gw[j])) * bf16_lo(uw[j]), f(bf16_hi(gw[j])) * bf16_hi(uw[j]));
        *reinterpret_cast<uint4*>(act + r * ff + c) = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// d_gu = [ d_act * u * gelu'(g) | d_act * gelu(g) ]
template <bool SILU>
__global__ void geglu_bwd_kernel(const __nv_bfloat16* __restrict__ dact, const __nv_bfloat16* __restrict__ gu,
                                 __nv_bfloat16* __restrict__ dgu, long long rows, int ff) {
    auto f = [](float z) { return SILU ? silu_f(z) : gelu_tanh_f(z); };
    auto df = [](float z) { return SILU ? silu_grad_f(z) : gelu_tanh_grad_f(z); };
    const int f8 = ff / 8;
    const long long n = rows * f8;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const long long r = i / f8;
        const int c = static_cast<int>(i - r * f8) * 8;
        const uint4 g = *reinterpret_cast<const uint4*>(gu + r * 2 * ff + c);
        const uint4 u = *reinterpret_cast<const uint4*>(gu + r * 2 * ff + ff + c);
        const uint4 d = *reinterpret_cast<const uint4*>(dact + r * ff + c);
        const uint32_t gw[4] = {g.x, g.y, g.z, g.w}, uw[4] = {u.x, u.y, u.z, u.w}, dw[4] = {d.x, d.y, d.z, d.w};
        uint32_t og[4], ou[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float g0 = bf16_lo(gw[j]), g1 = bf16_hi(gw[j]);
            const float u0 = bf16_lo(uw[j]), u1 = bf16_hi(uw[j]);
            const float d0 = bf16_lo(dw[j]), d1 = bf16_hi(dw[j]);
            og[j] = pack_bf16(d0 * u0 * df(g0), d1 * u1 * df(g1));
            ou[j] = pack_bf16(d0 * f(g0), d1 * f(g1));
        }
        *reinterpret_cast<uint4*>(dgu + r * 2 * ff + c) = make_uint4(og[0], og[1], og[2], og[3]);
        *reinterpret_cast<uint4*>(dgu + r * 2 * ff + ff + c) = make_uint4(ou[0], ou[1], ou[2], ou[3]);
    }
}

// dZ = dH * gelu_erf'(Z)   (projector hidden layer)
__global__ void gelu_erf_bwd_kernel(const __nv_bfloat16* dH, const __nv_bfloat16* __restrict__ Z,
                                    __nv_bfloat16* dZ, long long n8) {   // dZ may alias dH
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n8;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const uint4 h = *reinterpret_cast<const uint4*>(dH + i * 8);
        const uint4 z = *reinterpret_cast<const uint4*>(Z + i * 8);
        const uint32_t hw[4] = {h.x, h.y, h.z, h.w}, zw[4] = {z.x, z.y, z.z, z.w};
        uint32_t o[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            o[j] = pack_bf16(bf16_lo(hw[j]) * gelu_erf_grad_f(bf16_lo(zw[j])),
                             bf16_hi(hw[j]) * gelu_erf_grad_f(bf16_hi(zw[j])));
        *reinterpret_cast<uint4*>(dZ + i * 8) = make_uint4(o[0], o[1], o[2], o[3]);
    }
}

// ---- embedding gather * scale, spliced behind the image tokens; masks, labels, targets ---------
// grid = B*T blocks (one text token each) for the embedding rows, plus B blocks for the integers.
__global__ void embed_splice_kernel(const long long* __restrict__ token_ids, const long long* __restrict__ labels,
                                    const __nv_bfloat16* __restrict__ embed, float scale, long long pad_id,
                                    int has_pad, int B, int T, int n_img, int h, float* __restrict__ x0,
                                    long long* __restrict__ attn_mask, long long* __restrict__ lm_labels,
                                    uint8_t* __restrict__ keyvalid, int* __restrict__ targets, long long vocab,
                                    int* __restrict__ err) {
    const int S = n_img + T;
    if (blockIdx.x < B * T) {
        const int b = blockIdx.x / T, t = blockIdx.x - b * T;
        long long id = token_ids[blockIdx.x];
        if (id < 0 || id >= vocab) {        // nn.Embedding raises here; flag it (PT_ERRFLAG_TOKEN_ID) and read row 0
            if (threadIdx.x == 0 && err) atomicOr(err, PT_ERRFLAG_TOKEN_ID);
            id = 0;
        }
        const __nv_bfloat16* e = embed + id * h;
        float* dst = x0 + (static_cast<long long>(b) * S + n_img + t) * h;
        for (int i = threadIdx.x * 8; i < h; i += blockDim.x * 8) {
            const uint4 u = *reinterpret_cast<const uint4*>(e + i);
            // HF multiplies in the embedding dtype: bf16(e) * bf16(scale) rounded to bf16
            float v[8] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y),
                          bf16_lo(u.z), bf16_hi(u.z), bf16_lo(u.w), bf16_hi(u.w)};
#pragma unroll
            for (int j = 0; j < 8; ++j) v[j] = __bfloat162float(__float2bfloat16(v[j] * scale));
            *reinterpret_cast<float4*>(dst + i) = make_float4(v[0], v[1], v[2], v[3]);
            *reinterpret_cast<float4*>(dst + i + 4) = make_float4(v[4], v[5], v[6], v[7]);
        }
    } else {
        const int b = blockIdx.x - B * T;
        for (int j = threadIdx.x; j < S; j += blockDim.x) {
            long long m = 1, lab = -100;
            if (j >= n_img) {
                const long long id = token_ids[static_cast<long long>(b) * T + j - n_img];
                m = (!has_pad || id != pad_id) ? 1 : 0;
                lab = labels[static_cast<long long>(b) * T + j - n_img];
            }
            attn_mask[static_cast<long long>(b) * S + j] = m;
            lm_labels[static_cast<long long>(b) * S + j] = lab;
            keyvalid[static_cast<long long>(b) * S + j] = static_cast<uint8_t>(m);
            // shifted target for position j is the label of j+1
            long long tg = -100;
            if (j + 1 < S && j + 1 >= n_img) tg = labels[static_cast<long long>(b) * T + j + 1 - n_img];
            targets[static_cast<long long>(b) * S + j] = static_cast<int>(tg);
        }
    }
}

// ---- cross-entropy over the vocabulary -----------------------------------------------------------
// stats: one block per row; lse, per-row loss, and (loss_sum, n_valid) accumulated with atomics.
__global__ void __launch_bounds__(1024)
ce_stats_kernel(const __nv_bfloat16* __restrict__ logits, long long ld, const int* __restrict__ targets,
                int V, float* __restrict__ lse_out, float* __restrict__ acc /* [loss_sum, n_valid] */,
                int* __restrict__ err) {
    __shared__ float sh[32];
    const int r = blockIdx.x;
    const int tgt = targets[r];
    if (tgt >= V && threadIdx.x == 0 && err) atomicOr(err, PT_ERRFLAG_TARGET);   // CrossEntropyLoss raises here
    if (tgt < 0 || tgt >= V) { if (threadIdx.x == 0) lse_out[r] = 0.f; return; }
    const __nv_bfloat16* row = logits + r * ld;
    // one pass over the row (it is 512 KB at V = 262144: a second pass re-reads it through L2): per thread a
    // running maximum and a sum rescaled whenever the maximum moves, one rescale per 8 logits
    float m = -INFINITY, s = 0.f;
#pragma unroll 4
    for (int i = threadIdx.x * 8; i < V; i += blockDim.x * 8) {
        const uint4 u = *reinterpret_cast<const uint4*>(row + i);
        const float v[8] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y),
                            bf16_lo(u.z), bf16_hi(u.z), bf16_lo(u.w), bf16_hi(u.w)};
        const float mx = fmaxf(fmaxf(fmaxf(v[0], v[1]), fmaxf(v[2], v[3])), fmaxf(fmaxf(v[4], v[5]), fmaxf(v[6], v[7])));
        const float mn = fmaxf(m, mx);
        s *= __expf(m - mn);
        m = mn;
        s += (__expf(v[0] - mn) + __expf(v[1] - mn)) + (__expf(v[2] - mn) + __expf(v[3] - mn)) +
             (__expf(v[4] - mn) + __expf(v[5] - mn)) + (__expf(v[6] - mn) + __expf(v[7] - mn));
    }
    const float mt = m;
    m = block_max(m, sh);
    s *= __expf(mt - m);              // a thread that saw no element has mt = -inf, s = 0: contributes 0
    s = block_sum(s, sh);
    if (threadIdx.x == 0) {
        const float lse = m + logf(s);
        lse_out[r] = lse;
        atomicAdd(acc, lse - __bfloat162float(row[tgt]));
        atomicAdd(acc + 1, 1.f);
    }
}

// The same statistics from the per-tile partials the logits GEMM left behind (PT_ACT_CE_STATS): stats
// [parts][rows][2] = (max, sum exp(x - max)) per row and 128-column half tile.  Block = 32 rows x 8 part lanes; every
// warp-iteration reads 32 consecutive rows of one part (256 contiguous bytes).
__global__ void __launch_bounds__(256)
ce_combine_kernel(const float2* __restrict__ stats, int parts, int rows, const __nv_bfloat16* __restrict__ logits,
                  long long ld, const int* __restrict__ targets, int V, float* __restrict__ lse_out,
                  float* __restrict__ acc, int* __restrict__ err) {
    __shared__ float sm[8][32], ss[8][32];
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    const int r = blockIdx.x * 32 + lane;
    float m = -INFINITY, s = 0.f;
    if (r < rows) {
#pragma unroll 4
        for (int pidx = w; pidx < parts; pidx += 8) {
            const float2 v = stats[static_cast<long long>(pidx) * rows + r];
            const float mn = fmaxf(m, v.x);
            if (mn > -INFINITY) {
                s = s * __expf(m - mn) + v.y * __expf(v.x - mn);
                m = mn;
            }
        }
    }
    sm[w][lane] = m; ss[w][lane] = s;
    __syncthreads();
    if (w != 0 || r >= rows) return;
    for (int k = 1; k < 8; ++k) {
        const float m2 = sm[k][lane], s2 = ss[k][lane];
        const float mn = fmaxf(m, m2);
        if (mn > -INFINITY) { s = s * __expf(m - mn) + s2 * __expf(m2 - mn); m = mn; }
    }
    const int tgt = targets[r];
    if (tgt >= V && err) atomicOr(err, PT_ERRFLAG_TARGET);
    if (tgt < 0 || tgt >= V) { lse_out[r] = 0.f; return; }
    const float lse = m + logf(s);
    lse_out[r] = lse;
    atomicAdd(acc, lse - __bfloat162float(logits[r * ld + tgt]));
    atomicAdd(acc + 1, 1.f);
}

// grad: dlogits = (softmax - onehot) * gscale / n_valid, in place over the bf16 logits.
__global__ void __launch_bounds__(1024)
ce_grad_kernel(__nv_bfloat16* __restrict__ logits, long long ld, const int* __restrict__ targets, int V,
               const float* __restrict__ lse, const float* __restrict__ acc, float gscale) {
    const int r = blockIdx.y;
    const int tgt = targets[r];
    __nv_bfloat16* row = logits + r * ld;
    const int i = (blockIdx.x * blockDim.x + threadIdx.x) * 8;
    if (i >= V) return;
    if (tgt < 0 || tgt >= V) { *reinterpret_cast<uint4*>(row + i) = make_uint4(0, 0, 0, 0); return; }
    const float k = gscale / fmaxf(acc[1], 1.f);
    const float l = lse[r];
    const uint4 u = *reinterpret_cast<const uint4*>(row + i);
    float v[8] = {bf16_lo(u.x), bf16_hi(u.x), bf16_lo(u.y), bf16_hi(u.y),
                  bf16_lo(u.z), bf16_hi(u.z), bf16_lo(u.w), bf16_hi(u.w)};
#pragma unroll
    for (int j = 0; j < 8; ++j) v[j] = (__expf(v[j] - l) - ((i + j) == tgt ? 1.f : 0.f)) * k;
    uint4 o;
    o.x = pack_bf16(v[0], v[1]); o.y = pack_bf16(v[2], v[3]);
    o.z = pack_bf16(v[4], v[5]); o.w = pack_bf16(v[6], v[7]);
    *reinterpret_cast<uint4*>(row + i) = o;
}

__global__ void ce_finalize_kernel(const float* __restrict__ acc, float* __restrict__ loss) {
    // no valid label at all: HF's mean over an empty set is NaN (HF:loss/loss_utils.py:28-45) — keep it visible
    loss[0] = acc[1] > 0.f ? acc[0] / acc[1] : __int_as_float(0x7fc00000);
}

// ---- column sums of a bf16 [M,N] matrix into f32 [N] (bias gradients) ---------------------------
__global__ void __launch_bounds__(256)
colsum_kernel(const __nv_bfloat16* __restrict__ x, long long ld, float* __restrict__ out, int M, int N,
              int rows_per_block) {
    __shared__ float sh[8][64];
    const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
    const int col = blockIdx.x * 64 + tx * 2;
    const int r0 = blockIdx.y * rows_per_block;
    const int r1 = min(M, r0 + rows_per_block);
    float a = 0.f, b = 0.f;
    if (col < N) {
        for (int r = r0 + ty; r < r1; r += 8) {
            const uint32_t u = *reinterpret_cast<const uint32_t*>(x + r * ld + col);
            a += bf16_lo(u); b += bf16_hi(u);
        }
    }
    sh[ty][tx * 2] = a; sh[ty][tx * 2 + 1] = b;
    __syncthreads();
    if (threadIdx.x < 64) {
        float s = 0.f;
#pragma unroll
        for (int j = 0; j < 8; ++j) s += sh[j][threadIdx.x];
        const int c = blockIdx.x * 64 + threadIdx.x;
        if (c < N) atomicAdd(out + c, s);
    }
}

// ---- sum of squares of a flat f32 buffer --------------------------------------------------------
__global__ void __launch_bounds__(256)
sumsq_kernel(const float* __restrict__ g, long long n, float* __restrict__ out) {
    __shared__ float sh[32];
    float s = 0.f;
    const long long n4 = n / 4;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const float4 v = *reinterpret_cast<const float4*>(g + i * 4);
        s += v.x * v.x + v.y * v.y + v.z * v.z + v.w * v.w;
    }
    if (blockIdx.x == 0 && threadIdx.x == 0)
        for (long long i = n4 * 4; i < n; ++i) s += g[i] * g[i];
    s = block_sum(s, sh);
    if (threadIdx.x == 0) atomicAdd(out, s);
}

// ---- fused gradient clip + AdamW + bf16 shadow refresh -----------------------------------------
// g_eff = g * grad_scale * min(1, max_norm / (sqrt(sumsq) * grad_scale + 1e-6))
__global__ void __launch_bounds__(256)
clip_adamw_kernel(float* __restrict__ p, const float* __restrict__ g, float* __restrict__ m,
                  float* __restrict__ v, __nv_bfloat16* __restrict__ shadow, long long n,
                  const float* __restrict__ sumsq, float grad_scale, float max_norm, float lr, float b1,
                  float b2, float eps, float wd, float bc1, float bc2) {
    const float total = sqrtf(sumsq[0]) * grad_scale;
    const float coef = grad_scale * fminf(1.f, max_norm / (total + 1e-6f));
    const float inv_sqrt_bc2 = rsqrtf(bc2);
    const float step = lr / bc1;
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        const float gi = g[i] * coef;
        float pi = p[i] * (1.f - lr * wd);
        const float mi = b1 * m[i] + (1.f - b1) * gi;
        const float vi = b2 * v[i] + (1.f - b2) * gi * gi;
        pi -= step * mi / (sqrtf(vi) * inv_sqrt_bc2 + eps);
        p[i] = pi; m[i] = mi; v[i] = vi;
        if (shadow) shadow[i] = __float2bfloat16(pi);
    }
}

__global__ void cast_f32_bf16_kernel(const float* __restrict__ x, __nv_bfloat16* __restrict__ y, long long n) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n;
         i += static_cast<long long>(gridDim.x) * blockDim.x)
        y[i] = __float2bfloat16(x[i]);
}

// out[i] = bf16(sum_s x[s * slice + i]): the f32 partials of a split-K GEMM (slice % 4 == 0), 16-byte loads,
// the S loads of one output quad independent of each other.
__global__ void sum_slices_kernel(const float4* __restrict__ x, long long n4, int S, uint2* __restrict__ out) {
    for (long long i = static_cast<long long>(blockIdx.x) * blockDim.x + threadIdx.x; i < n4;
         i += static_cast<long long>(gridDim.x) * blockDim.x) {
        float4 a = x[i];
#pragma unroll 4
        for (int s = 1; s < S; ++s) {
            const float4 b = x[s * n4 + i];
            a.x += b.x; a.y += b.y; a.z += b.z; a.w += b.w;
        }
        out[i] = make_uint2(pack_bf16(a.x, a.y), pack_bf16(a.z, a.w));
    }
}

static inline int grid_for(long long n, int threads) {
    long long b = (n + threads - 1) / threads;
    const long long cap = static_cast<long long>(num_sms()) * 16;
    return static_cast<int>(b < 1 ? 1 : (b > cap ? cap : b));
}

}  // namespace pt

using namespace pt;
#define ST(s) static_cast<cudaStream_t>(s)
#define BF(p) static_cast<__nv_bfloat16*>(p)
#define CBF(p) static_cast<const __nv_bfloat16*>(p)

extern "C" int pt_patchify(const float* pixels, void* cols_bf16, int B, int H, int W, int patch, void* stream) {
    PT_REQUIRE(pixels && cols_bf16, "pt_patchify: null pointer");
    PT_REQUIRE(B > 0 && patch % 8 == 0 && H % patch == 0 && W % patch == 0 && W % 4 == 0,
               "pt_patchify: patch must be a multiple of 8 dividing H and W");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long n8 = static_cast<long long>(B) * 3 * H * W / 8;
    patchify_kernel<<<grid_for(n8, 256), 256, 0, ST(stream)>>>(pixels, BF(cols_bf16), B, H, W, patch);
    return check_launch("patchify");
}

extern "C" int pt_geglu_fwd(const void* gu, void* act, long long rows, int ff, void* stream) {
    PT_REQUIRE(gu && act && rows > 0 && ff > 0 && ff % 8 == 0, "pt_geglu_fwd: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    geglu_fwd_kernel<false><<<grid_for(rows * (ff / 8), 256), 256, 0, ST(stream)>>>(CBF(gu), BF(act), rows, ff);
    return check_launch("geglu_fwd");
}

extern "C" int pt_swiglu_fwd(const void* gu, void* act, long long rows, int ff, void* stream) {
    PT_REQUIRE(gu && act && rows > 0 && ff > 0 && ff % 8 == 0, "pt_swiglu_fwd: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    geglu_fwd_kernel<true><<<grid_for(rows * (ff / 8), 256), 256, 0, ST(stream)>>>(CBF(gu), BF(act), rows, ff);
    return check_launch("swiglu_fwd");
}

extern "C" int pt_geglu_bwd(const void* dact, const void* gu, void* dgu, long long rows, int ff, void* stream) {
    PT_REQUIRE(dact && gu && dgu && rows > 0 && ff > 0 && ff % 8 == 0, "pt_geglu_bwd: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    geglu_bwd_kernel<false><<<grid_for(rows * (ff / 8), 256), 256, 0, ST(stream)>>>(CBF(dact), CBF(gu), BF(dgu), rows, ff);
    return check_launch("geglu_bwd");
}

extern "C" int pt_swiglu_bwd(const void* dact, const void* gu, void* dgu, long long rows, int ff, void* stream) {
    PT_REQUIRE(dact && gu && dgu && rows > 0 && ff > 0 && ff % 8 == 0, "pt_swiglu_bwd: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    geglu_bwd_kernel<true><<<grid_for(rows * (ff / 8), 256), 256, 0, ST(stream)>>>(CBF(dact), CBF(gu), BF(dgu), rows, ff);
    return check_launch("swiglu_bwd");
}

extern "C" int pt_gelu_erf_bwd(const void* dH, const void* Z, void* dZ, long long n, void* stream) {
    PT_REQUIRE(dH && Z && dZ && n > 0 && n % 8 == 0, "pt_gelu_erf_bwd: n must be a positive multiple of 8");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    gelu_erf_bwd_kernel<<<grid_for(n / 8, 256), 256, 0, ST(stream)>>>(CBF(dH), CBF(Z), BF(dZ), n / 8);
    return check_launch("gelu_erf_bwd");
}

extern "C" int pt_embed_splice(const int64_t* token_ids, const int64_t* labels, const void* embed_bf16,
                               float scale, int64_t pad_id, int has_pad, int B, int T, int n_img, int h,
                               float* x0, int64_t* attn_mask, int64_t* lm_labels, uint8_t* keyvalid,
                               int32_t* targets, int64_t vocab, int32_t* err_flags, void* stream) {
    PT_REQUIRE(token_ids && labels && embed_bf16 && x0 && attn_mask && lm_labels && keyvalid && targets,
               "pt_embed_splice: null pointer");
    PT_REQUIRE(B > 0 && T > 0 && n_img >= 0 && h > 0 && h % 8 == 0 && vocab > 0, "pt_embed_splice: bad shape");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    embed_splice_kernel<<<B * T + B, 128, 0, ST(stream)>>>(
        reinterpret_cast<const long long*>(token_ids), reinterpret_cast<const long long*>(labels),
        CBF(embed_bf16), scale, pad_id, has_pad, B, T, n_img, h, x0,
        reinterpret_cast<long long*>(attn_mask), reinterpret_cast<long long*>(lm_labels), keyvalid, targets,
        static_cast<long long>(vocab), err_flags);
    return check_launch("embed_splice");
}

extern "C" int pt_ce_from_partials(void* logits_bf16, long long ld, const int32_t* targets, int rows, int V,
                                   const float* stats, int parts, float* lse, float* acc2, float* loss, float gscale,
                                   int32_t* err_flags, void* stream) {
    PT_REQUIRE(logits_bf16 && targets && stats && lse && acc2 && loss, "pt_ce_from_partials: null pointer");
    PT_REQUIRE(rows > 0 && V > 0 && V % 8 == 0 && ld % 8 == 0, "pt_ce_from_partials: V and ld must be multiples of 8");
    PT_REQUIRE(parts == 2 * ((V + 255) / 256), "pt_ce_from_partials: parts must be 2 * ceil(V / 256) (got %d for V = %d)", parts, V);
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = ST(stream);
    if (cudaMemsetAsync(acc2, 0, 2 * sizeof(float), st) != cudaSuccess) return check_launch("ce acc memset");
    ce_combine_kernel<<<(rows + 31) / 32, 256, 0, st>>>(reinterpret_cast<const float2*>(stats), parts, rows,
                                                        CBF(logits_bf16), ld, targets, V, lse, acc2, err_flags);
    ce_finalize_kernel<<<1, 1, 0, st>>>(acc2, loss);
    dim3 grid((V / 8 + 255) / 256, rows);
    ce_grad_kernel<<<grid, 256, 0, st>>>(BF(logits_bf16), ld, targets, V, lse, acc2, gscale);
    return check_launch("ce_from_partials");
}

extern "C" int pt_ce_fwd_bwd(void* logits_bf16, long long ld, const int32_t* targets, int rows, int V,
                             float* lse, float* acc2, float* loss, float gscale, int32_t* err_flags, void* stream) {
    PT_REQUIRE(logits_bf16 && targets && lse && acc2 && loss, "pt_ce_fwd_bwd: null pointer");
    PT_REQUIRE(rows > 0 && V > 0 && V % 8 == 0 && ld % 8 == 0, "pt_ce_fwd_bwd: V and ld must be multiples of 8");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = ST(stream);
    if (cudaMemsetAsync(acc2, 0, 2 * sizeof(float), st) != cudaSuccess) return check_launch("ce memset");
    ce_stats_kernel<<<rows, 1024, 0, st>>>(CBF(logits_bf16), ld, targets, V, lse, acc2, err_flags);
    ce_finalize_kernel<<<1, 1, 0, st>>>(acc2, loss);
    dim3 grid((V / 8 + 255) / 256, rows);
    ce_grad_kernel<<<grid, 256, 0, st>>>(BF(logits_bf16), ld, targets, V, lse, acc2, gscale);
    return check_launch("ce_fwd_bwd");
}

extern "C" int pt_colsum(const void* x_bf16, long long ld, float* out, int M, int N, void* stream) {
    PT_REQUIRE(x_bf16 && out && M > 0 && N > 0 && N % 2 == 0 && ld % 2 == 0, "pt_colsum: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = ST(stream);
    if (cudaMemsetAsync(out, 0, N * sizeof(float), st) != cudaSuccess) return check_launch("colsum memset");
    const int rpb = 512;
    dim3 grid((N + 63) / 64, (M + rpb - 1) / rpb);
    colsum_kernel<<<grid, 256, 0, st>>>(CBF(x_bf16), ld, out, M, N, rpb);
    return check_launch("colsum");
}

extern "C" int pt_sumsq(const float* g, long long n, float* out, void* stream) {
    PT_REQUIRE(g && out && n > 0, "pt_sumsq: bad arguments");
    PT_REQUIRE((reinterpret_cast<uintptr_t>(g) & 15) == 0, "pt_sumsq: buffer must be 16B aligned");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = ST(stream);
    if (cudaMemsetAsync(out, 0, sizeof(float), st) != cudaSuccess) return check_launch("sumsq memset");
    sumsq_kernel<<<grid_for(n / 4 + 1, 256), 256, 0, st>>>(g, n, out);
    return check_launch("sumsq");
}

extern "C" int pt_clip_adamw(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n,
                             const float* sumsq, float grad_scale, float max_norm, float lr, float beta1,
                             float beta2, float eps, float weight_decay, int step, void* stream) {
    PT_REQUIRE(p && g && m && v && sumsq && n > 0 && step >= 1, "pt_clip_adamw: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const float bc1 = 1.f - powf(beta1, static_cast<float>(step));
    const float bc2 = 1.f - powf(beta2, static_cast<float>(step));
    clip_adamw_kernel<<<grid_for(n, 256), 256, 0, ST(stream)>>>(p, g, m, v, BF(shadow_bf16), n, sumsq, grad_scale,
                                                                 max_norm, lr, beta1, beta2, eps, weight_decay, bc1, bc2);
    return check_launch("clip_adamw");
}

extern "C" int pt_cast_f32_bf16(const float* x, void* y_bf16, long long n, void* stream) {
    PT_REQUIRE(x && y_bf16 && n > 0, "pt_cast_f32_bf16: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cast_f32_bf16_kernel<<<grid_for(n, 256), 256, 0, ST(stream)>>>(x, BF(y_bf16), n);
    return check_launch("cast_f32_bf16");
}

extern "C" int pt_sum_slices(const float* x, long long slice_elems, int n_slices, void* out_bf16, void* stream) {
    PT_REQUIRE(x && out_bf16 && slice_elems > 0 && n_slices >= 1, "pt_sum_slices: bad arguments");
    PT_REQUIRE(slice_elems % 4 == 0, "pt_sum_slices: slice_elems must be a multiple of 4 (%lld)", slice_elems);
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long n4 = slice_elems / 4;
    sum_slices_kernel<<<grid_for(n4, 256), 256, 0, ST(stream)>>>(reinterpret_cast<const float4*>(x), n4, n_slices,
                                                                 reinterpret_cast<uint2*>(out_bf16));
    return check_launch("sum_slices");
}
