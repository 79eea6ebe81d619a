// Bandwidth-bound pieces of Gemma3 / SigLIP attention around the tensor-core GEMMs:
//   - per-head q/k RMSNorm fused with RoPE (forward and backward)
//   - masked row softmax (causal AND sliding-window AND key-valid) and its backward
#include "pt_common.cuh"

namespace pt {

constexpr int kAttnWarps = 8;
// below this many token rows a warp per (row, head) fills the SMs better than a warp per row (decode steps)
constexpr int kRowKernelMinRows = 148 * 8 * 4;

__device__ __forceinline__ void ld4(const __nv_bfloat16* p, float (&v)[4]) {
    const uint2 u = *reinterpret_cast<const uint2*>(p);
    v[0] = bf16_lo(u.x); v[1] = bf16_hi(u.x); v[2] = bf16_lo(u.y); v[3] = bf16_hi(u.y);
}
__device__ __forceinline__ void st4(__nv_bfloat16* p, const float (&v)[4]) {
    uint2 u; u.x = pack_bf16(v[0], v[1]); u.y = pack_bf16(v[2], v[3]);
    *reinterpret_cast<uint2*>(p) = u;
}

// One warp per (token row, head).  Heads 0..hq-1 are q heads, hq..hq+hkv-1 are k heads of the
// fused qkv row [ (hq + 2 hkv) * hd ].  Lane l owns pairs (i, i + hd/2) for i in 4l..4l+3 (+128 j).
//   n = u rsqrt(mean u^2 + eps) (1 + w);  y[i] = n[i] c - n[i+h] s;  y[i+h] = n[i+h] c + n[i] s
__global__ void __launch_bounds__(kAttnWarps * 32)
qknorm_rope_fwd_kernel(const __nv_bfloat16* __restrict__ qkv, const float* __restrict__ wq,
                       const float* __restrict__ wk, const float* __restrict__ cs,
                       const float* __restrict__ sn, __nv_bfloat16* __restrict__ q_out,
                       __nv_bfloat16* __restrict__ k_out, int rows, int S, int hq, int hkv, int hd,
                       float eps) {
    const int item = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
    const int nh = hq + hkv;
    if (item >= rows * nh) return;
    const int row = item / nh, head = item - row * nh;
    const int lane = threadIdx.x & 31;
    const int half = hd >> 1;
    const __nv_bfloat16* u = qkv + static_cast<long long>(row) * (hq + 2 * hkv) * hd + head * hd;
    const float* w = head < hq ? wq : wk;
    __nv_bfloat16* out = head < hq ? q_out + (static_cast<long long>(row) * hq + head) * hd
                                   : k_out + (static_cast<long long>(row) * hkv + (head - hq)) * hd;
    const int pos = row % S;
    float q = 0.f;
    for (int i = lane * 4; i < half; i += 128) {
        float a[4], b[4]; ld4(u + i, a); ld4(u + half + i, b);
#pragma unroll
        for (int j = 0; j < 4; ++j) q += a[j] * a[j] + b[j] * b[j];
    }
    const float rstd = rsqrtf(warp_sum(q) / hd + eps);
    for (int i = lane * 4; i < half; i += 128) {
        float a[4], b[4], ya[4], yb[4];
        ld4(u + i, a); ld4(u + half + i, b);
        const float4 c = *reinterpret_cast<const float4*>(cs + static_cast<long long>(pos) * half + i);
        const float4 s = *reinterpret_cast<const float4*>(sn + static_cast<long long>(pos) * half + i);
        const float4 w1 = *reinterpret_cast<const float4*>(w + i);
        const float4 w2 = *reinterpret_cast<const float4*>(w + half + i);
        const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {s.x, s.y, s.z, s.w};
        const float g1[4] = {w1.x, w1.y, w1.z, w1.w}, g2[4] = {w2.x, w2.y, w2.z, w2.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float n1 = a[j] * rstd * (1.f + g1[j]), n2 = b[j] * rstd * (1.f + g2[j]);
            ya[j] = n1 * cc[j] - n2 * ss[j];
            yb[j] = n2 * cc[j] + n1 * ss[j];
        }
        st4(out + i, ya); st4(out + half + i, yb);
    }
}

// Backward of the above: dy (q_out / k_out layout) -> d_qkv (q and k column ranges of the fused row).
__global__ void __launch_bounds__(kAttnWarps * 32)
qknorm_rope_bwd_kernel(const __nv_bfloat16* __restrict__ dq, const __nv_bfloat16* __restrict__ dk,
                       const __nv_bfloat16* __restrict__ qkv, const float* __restrict__ wq,
                       const float* __restrict__ wk, const float* __restrict__ cs,
                       const float* __restrict__ sn, __nv_bfloat16* __restrict__ d_qkv, int rows,
                       int S, int hq, int hkv, int hd, float eps) {
    const int item = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
    const int nh = hq + hkv;
    if (item >= rows * nh) return;
    const int row = item / nh, head = item - row * nh;
    const int lane = threadIdx.x & 31;
    const int half = hd >> 1;
    const long long off = static_cast<long long>(row) * (hq + 2 * hkv) * hd + head * hd;
    const __nv_bfloat16* u = qkv + off;
    __nv_bfloat16* du = d_qkv + off;
    const float* w = head < hq ? wq : wk;
    const __nv_bfloat16* dy = head < hq ? dq + (static_cast<long long>(row) * hq + head) * hd
                                        : dk + (static_cast<long long>(row) * hkv + (head - hq)) * hd;
    const int pos = row % S;
    float q = 0.f, dot = 0.f;
    for (int i = lane * 4; i < half; i += 128) {
        float a[4], b[4], da[4], db[4];
        ld4(u + i, a); ld4(u + half + i, b); ld4(dy + i, da); ld4(dy + half + i, db);
        const float4 c = *reinterpret_cast<const float4*>(cs + static_cast<long long>(pos) * half + i);
        const float4 s = *reinterpret_cast<const float4*>(sn + static_cast<long long>(pos) * half + i);
        const float4 w1 = *reinterpret_cast<const float4*>(w + i);
        const float4 w2 = *reinterpret_cast<const float4*>(w + half + i);
        const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {s.x, s.y, s.z, s.w};
        const float g1[4] = {w1.x, w1.y, w1.z, w1.w}, g2[4] = {w2.x, w2.y, w2.z, w2.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float dn1 = (da[j] * cc[j] + db[j] * ss[j]) * (1.f + g1[j]);
            const float dn2 = (db[j] * cc[j] - da[j] * ss[j]) * (1.f + g2[j]);
            q += a[j] * a[j] + b[j] * b[j];
            dot += dn1 * a[j] + dn2 * b[j];
        }
    }
    q = warp_sum(q); dot = warp_sum(dot);
    const float rstd = rsqrtf(q / hd + eps);
    const float c3 = dot * rstd * rstd * rstd / hd;
    for (int i = lane * 4; i < half; i += 128) {
        float a[4], b[4], da[4], db[4], oa[4], ob[4];
        ld4(u + i, a); ld4(u + half + i, b); ld4(dy + i, da); ld4(dy + half + i, db);
        const float4 c = *reinterpret_cast<const float4*>(cs + static_cast<long long>(pos) * half + i);
        const float4 s = *reinterpret_cast<const float4*>(sn + static_cast<long long>(pos) * half + i);
        const float4 w1 = *reinterpret_cast<const float4*>(w + i);
        const float4 w2 = *reinterpret_cast<const float4*>(w + half + i);
        const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {s.x, s.y, s.z, s.w};
        const float g1[4] = {w1.x, w1.y, w1.z, w1.w}, g2[4] = {w2.x, w2.y, w2.z, w2.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float dn1 = (da[j] * cc[j] + db[j] * ss[j]) * (1.f + g1[j]);
            const float dn2 = (db[j] * cc[j] - da[j] * ss[j]) * (1.f + g2[j]);
            oa[j] = rstd * dn1 - a[j] * c3;
            ob[j] = rstd * dn2 - b[j] * c3;
        }
        st4(du + i, oa); st4(du + half + i, ob);
    }
}

// Single-pass variants for head_dim <= 256 (one 4-wide slice of each half per lane): every operand is loaded
// once and stays in registers across the reduction.
__global__ void __launch_bounds__(kAttnWarps * 32)
qknorm_rope_fwd_reg_kernel(const __nv_bfloat16* __restrict__ qkv, const float* __restrict__ wq,
                           const float* __restrict__ wk, const float* __restrict__ cs,
                           const float* __restrict__ sn, __nv_bfloat16* __restrict__ q_out,
                           __nv_bfloat16* __restrict__ k_out, int rows, int S, int hq, int hkv, int hd, float eps) {
    const int item = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
    const int nh = hq + hkv;
    if (item >= rows * nh) return;
    const int row = item / nh, head = item - row * nh;
    const int lane = threadIdx.x & 31;
    const int half = hd >> 1;
    const int i = lane * 4;
    const bool on = i < half;
    const __nv_bfloat16* u = qkv + static_cast<long long>(row) * (hq + 2 * hkv) * hd + head * hd;
    const float* w = head < hq ? wq : wk;
    __nv_bfloat16* out = head < hq ? q_out + (static_cast<long long>(row) * hq + head) * hd
                                   : k_out + (static_cast<long long>(row) * hkv + (head - hq)) * hd;
    const long long tb = static_cast<long long>(row % S) * half + i;
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
    float4 c = make_float4(0.f, 0.f, 0.f, 0.f), s = c, w1 = c, w2 = c;
    if (on) {
        ld4(u + i, a); ld4(u + half + i, b);
        c = *reinterpret_cast<const float4*>(cs + tb);
        s = *reinterpret_cast<const float4*>(sn + tb);
        w1 = *reinterpret_cast<const float4*>(w + i);
        w2 = *reinterpret_cast<const float4*>(w + half + i);
    }
    float q = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) q += a[j] * a[j] + b[j] * b[j];
    const float rstd = rsqrtf(warp_sum(q) / hd + eps);
    if (on) {
        const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {s.x, s.y, s.z, s.w};
        const float g1[4] = {w1.x, w1.y, w1.z, w1.w}, g2[4] = {w2.x, w2.y, w2.z, w2.w};
        float ya[4], yb[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const float n1 = a[j] * rstd * (1.f + g1[j]), n2 = b[j] * rstd * (1.f + g2[j]);
            ya[j] = n1 * cc[j] - n2 * ss[j];
            yb[j] = n2 * cc[j] + n1 * ss[j];
        }
        st4(out + i, ya); st4(out + half + i, yb);
    }
}

__global__ void __launch_bounds__(kAttnWarps * 32)
qknorm_rope_bwd_reg_kernel(const __nv_bfloat16* __restrict__ dq, const __nv_bfloat16* __restrict__ dk,
                           const __nv_bfloat16* __restrict__ qkv, const float* __restrict__ wq,
                           const float* __restrict__ wk, const float* __restrict__ cs,
                           const float* __restrict__ sn, __nv_bfloat16* __restrict__ d_qkv, int rows,
                           int S, int hq, int hkv, int hd, float eps) {
    const int item = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
    const int nh = hq + hkv;
    if (item >= rows * nh) return;
    const int row = item / nh, head = item - row * nh;
    const int lane = threadIdx.x & 31;
    const int half = hd >> 1;
    const int i = lane * 4;
    const bool on = i < half;
    const long long off = static_cast<long long>(row) * (hq + 2 * hkv) * hd + head * hd;
    const __nv_bfloat16* u = qkv + off;
    __nv_bfloat16* du = d_qkv + off;
    const float* w = head < hq ? wq : wk;
    const __nv_bfloat16* dy = head < hq ? dq + (static_cast<long long>(row) * hq + head) * hd
                                        : dk + (static_cast<long long>(row) * hkv + (head - hq)) * hd;
    const long long tb = static_cast<long long>(row % S) * half + i;
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f}, dn1[4] = {0.f, 0.f, 0.f, 0.f}, dn2[4] = {0.f, 0.f, 0.f, 0.f};
    if (on) {
        float da[4], db[4];
        ld4(u + i, a); ld4(u + half + i, b); ld4(dy + i, da); ld4(dy + half + i, db);
        const float4 c = *reinterpret_cast<const float4*>(cs + tb);
        const float4 s = *reinterpret_cast<const float4*>(sn + tb);
        const float4 w1 = *reinterpret_cast<const float4*>(w + i);
        const float4 w2 = *reinterpret_cast<const float4*>(w + half + i);
        const float cc[4] = {c.x, c.y, c.z, c.w}, ss[4] = {s.x, s.y, s.z, s.w};
        const float g1[4] = {w1.x, w1.y, w1.z, w1.w}, g2[4] = {w2.x, w2.y, w2.z, w2.w};
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            dn1[j] = (da[j] * cc[j] + db[j] * ss[j]) * (1.f + g1[j]);
            dn2[j] = (db[j] * cc[j] - da[j] * ss[j]) * (1.f + g2[j]);
        }
    }
    float q = 0.f, dot = 0.f;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
        q += a[j] * a[j] + b[j] * b[j];
        dot += dn1[j] * a[j] + dn2[j] * b[j];
    }
    q = warp_sum(q); dot = warp_sum(dot);
    const float rstd = rsqrtf(q / hd + eps);
    const float c3 = dot * rstd * rstd * rstd / hd;
    if (on) {
        float oa[4], ob[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            oa[j] = rstd * dn1[j] - a[j] * c3;
            ob[j] = rstd * dn2[j] - b[j] * c3;
        }
        st4(du + i, oa); st4(du + half + i, ob);
    }
}

// Row variant of the forward (head_dim <= 256): one warp walks every q and k head of a token row, so the row's
// cos/sin slice and the two norm gains are loaded once instead of once per head, and the next head's operands are
// in flight while the current head's reduction completes (33.7 -> 29.0 us per launch at 22496 rows x 5 heads).
__global__ void __launch_bounds__(kAttnWarps * 32)
qknorm_rope_fwd_row_kernel(const __nv_bfloat16* __restrict__ qkv, const float* __restrict__ wq,
                           const float* __restrict__ wk, const float* __restrict__ cs,
                           const float* __restrict__ sn, __nv_bfloat16* __restrict__ q_out,
                           __nv_bfloat16* __restrict__ k_out, int rows, int S, int hq, int hkv, int hd, float eps) {
    const int row = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
    if (row >= rows) return;
    const int lane = threadIdx.x & 31;
    const int half = hd >> 1, nh = hq + hkv;
    const int i = lane * 4;
    const bool on = i < half;
    const __nv_bfloat16* u = qkv + static_cast<long long>(row) * (hq + 2 * hkv) * hd;
    __nv_bfloat16* oq = q_out + static_cast<long long>(row) * hq * hd;
    __nv_bfloat16* ok = k_out + static_cast<long long>(row) * hkv * hd;
    const long long tb = static_cast<long long>(row % S) * half + i;
    float cc[4] = {0.f, 0.f, 0.f, 0.f}, ss[4] = {0.f, 0.f, 0.f, 0.f};
    float gq1[4] = {0.f, 0.f, 0.f, 0.f}, gq2[4] = {0.f, 0.f, 0.f, 0.f}, gk1[4] = {0.f, 0.f, 0.f, 0.f}, gk2[4] = {0.f, 0.f, 0.f, 0.f};
    float a[4] = {0.f, 0.f, 0.f, 0.f}, b[4] = {0.f, 0.f, 0.f, 0.f};
    if (on) {
        ld4(u + i, a); ld4(u + half + i, b);
        const float4 c = *reinterpret_cast<const float4*>(cs + tb);
        const float4 s = *reinterpret_cast<const float4*>(sn + tb);
        const float4 q1 = *reinterpret_cast<const float4*>(wq + i), q2 = *reinterpret_cast<const float4*>(wq + half + i);
        const float4 k1 = *reinterpret_cast<const float4*>(wk + i), k2 = *reinterpret_cast<const float4*>(wk + half + i);
        cc[0] = c.x; cc[1] = c.y; cc[2] = c.z; cc[3] = c.w; ss[0] = s.x; ss[1] = s.y; ss[2] = s.z; ss[3] = s.w;
        gq1[0] = 1.f + q1.x; gq1[1] = 1.f + q1.y; gq1[2] = 1.f + q1.z; gq1[3] = 1.f + q1.w;
        gq2[0] = 1.f + q2.x; gq2[1] = 1.f + q2.y; gq2[2] = 1.f + q2.z; gq2[3] = 1.f + q2.w;
        gk1[0] = 1.f + k1.x; gk1[1] = 1.f + k1.y; gk1[2] = 1.f + k1.z; gk1[3] = 1.f + k1.w;
        gk2[0] = 1.f + k2.x; gk2[1] = 1.f + k2.y; gk2[2] = 1.f + k2.z; gk2[3] = 1.f + k2.w;
    }
    for (int head = 0; head < nh; ++head) {
        float an[4] = {0.f, 0.f, 0.f, 0.f}, bn[4] = {0.f, 0.f, 0.f, 0.f};
        if (on && head + 1 < nh) { ld4(u + (head + 1) * hd + i, an); ld4(u + (head + 1) * hd + half + i, bn); }
        float q = 0.f;
#pragma unroll
        for (int j = 0; j < 4; ++j) q += a[j] * a[j] + b[j] * b[j];
        const float rstd = rsqrtf(warp_sum(q) / hd + eps);
        if (on) {
            const bool isq = head < hq;
            __nv_bfloat16* out = isq ? oq + head * hd : ok + (head - hq) * hd;
            float ya[4], yb[4];
#pragma unroll
            for (int j = 0; j < 4; ++j) {
                const float n1 = a[j] * rstd * (isq ? gq1[j] : gk1[j]), n2 = b[j] * rstd * (isq ? gq2[j] : gk2[j]);
                ya[j] = n1 * cc[j] - n2 * ss[j];
                yb[j] = n2 * cc[j] + n1 * ss[j];
            }
            st4(out + i, ya); st4(out + half + i, yb);
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) { a[j] = an[j]; b[j] = bn[j]; }
    }
}

// ---- masked softmax over score rows -------------------------------------------------------------
// scores f32 [Z][Sq][ld] (already scaled) -> P bf16 [Z][Sq][ld]; columns >= Sk and masked keys get 0.
// mask: key k allowed for query q iff (!causal || k <= q) && (window <= 0 || k > q - window)
//       && (keyvalid == null || keyvalid[(z / heads_per_batch) * Sk + k]).
__global__ void __launch_bounds__(kAttnWarps * 32)
softmax_fwd_kernel(const float* __restrict__ sc, __nv_bfloat16* __restrict__ P,
                   const uint8_t* __restrict__ keyvalid, int Z, int Sq, int Sk, int ld,
                   int heads_per_batch, int causal, int window) {
    const long long item = static_cast<long long>(blockIdx.x) * kAttnWarps + (threadIdx.x >> 5);
    if (item >= static_cast<long long>(Z) * Sq) return;
    const int z = static_cast<int>(item / Sq), qi = static_cast<int>(item - static_cast<long long>(z) * Sq);
    const int lane = threadIdx.x & 31;
    const float* srow = sc + item * ld;
    __nv_bfloat16* prow = P + item * ld;
    const uint8_t* kv = keyvalid ? keyvalid + static_cast<long long>(z / heads_per_batch) * Sk : nullptr;
    const int k_hi = causal ? min(qi, Sk - 1) : Sk - 1;                 // inclusive
    const int k_lo = (window > 0) ? max(0, qi - window + 1) : 0;
    float m = -INFINITY, l = 0.f;
    for (int k = k_lo + lane; k <= k_hi; k += 32) {
        if (kv && !kv[k]) continue;
        const float v = srow[k];
        if (v > m) { l = l * __expf(m - v) + 1.f; m = v; }
        else l += __expf(v - m);
    }
    const float mm = warp_max(m);
    l = (m == -INFINITY) ? 0.f : l * __expf(m - mm);
    l = warp_sum(l);
    const float inv = l > 0.f ? 1.f / l : 0.f;
    for (int k = lane * 2; k < ld; k += 64) {
        float o[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const int kk = k + j;
            const bool ok = kk >= k_lo && kk <= k_hi && (!kv || kv[kk]);
            o[j] = ok ? __expf(srow[kk] - mm) * inv : 0.f;
        }
        *reinterpret_cast<uint32_t*>(prow + k) = pack_bf16(o[0], o[1]);
    }
}

// Same contract, rows of up to 1024 columns held in registers: one global read of the in-band scores, one write.
// Lane l owns the column pairs 2 l + 64 i, i = 0..15.
__global__ void __launch_bounds__(kAttnWarps * 32)
softmax_fwd_reg_kernel(const float* __restrict__ sc, __nv_bfloat16* __restrict__ P,
                       const uint8_t* __restrict__ keyvalid, int Z, int Sq, int Sk, int ld,
                       int heads_per_batch, int causal, int window) {
    const long long item = static_cast<long long>(blockIdx.x) * kAttnWarps + (threadIdx.x >> 5);
    if (item >= static_cast<long long>(Z) * Sq) return;
    const int z = static_cast<int>(item / Sq), qi = static_cast<int>(item - static_cast<long long>(z) * Sq);
    const int lane = threadIdx.x & 31;
    const float* srow = sc + item * ld;
    __nv_bfloat16* prow = P + item * ld;
    const uint8_t* kv = keyvalid ? keyvalid + static_cast<long long>(z / heads_per_batch) * Sk : nullptr;
    const int k_hi = causal ? min(qi, Sk - 1) : Sk - 1;                 // inclusive
    const int k_lo = (window > 0) ? max(0, qi - window + 1) : 0;
    float v[16][2];
    float m = -INFINITY;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int k = 2 * lane + 64 * i;
        v[i][0] = -INFINITY; v[i][1] = -INFINITY;
        if (k + 1 >= k_lo && k <= k_hi) {
            const float2 f = *reinterpret_cast<const float2*>(srow + k);
            if (k >= k_lo && (!kv || kv[k])) v[i][0] = f.x;
            if (k + 1 <= k_hi && (!kv || kv[k + 1])) v[i][1] = f.y;
        }
        m = fmaxf(m, fmaxf(v[i][0], v[i][1]));
    }
    m = warp_max(m);
    float l = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        v[i][0] = (v[i][0] == -INFINITY) ? 0.f : __expf(v[i][0] - m);
        v[i][1] = (v[i][1] == -INFINITY) ? 0.f : __expf(v[i][1] - m);
        l += v[i][0] + v[i][1];
    }
    l = warp_sum(l);
    const float inv = l > 0.f ? 1.f / l : 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int k = 2 * lane + 64 * i;
        if (k < ld) *reinterpret_cast<uint32_t*>(prow + k) = pack_bf16(v[i][0] * inv, v[i][1] * inv);
    }
}

__global__ void __launch_bounds__(kAttnWarps * 32)
softmax_bwd_reg_kernel(const __nv_bfloat16* __restrict__ P, const float* __restrict__ dP,
                       __nv_bfloat16* __restrict__ dS, long long rows, int Sk, int ld, float scale,
                       int Sq, int causal, int window) {
    const long long item = static_cast<long long>(blockIdx.x) * kAttnWarps + (threadIdx.x >> 5);
    if (item >= rows) return;
    const int lane = threadIdx.x & 31;
    const __nv_bfloat16* prow = P + item * ld;
    const float* drow = dP + item * ld;
    __nv_bfloat16* orow = dS + item * ld;
    // columns outside [k_lo, k_hi] hold P = 0 and dP is undefined there (band-skipped tiles): not read
    const int qi = Sq > 0 ? static_cast<int>(item % Sq) : 0;
    const int k_hi = (Sq > 0 && causal) ? min(qi, Sk - 1) : Sk - 1;
    const int k_lo = (Sq > 0 && window > 0) ? max(0, qi - window + 1) : 0;
    uint32_t pu[16];
    float2 d[16];
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int k = 2 * lane + 64 * i;
        const bool in = k + 1 >= k_lo && k <= k_hi;
        pu[i] = in ? *reinterpret_cast<const uint32_t*>(prow + k) : 0u;
        d[i] = in ? *reinterpret_cast<const float2*>(drow + k) : make_float2(0.f, 0.f);
    }
    float dot = 0.f;
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const float p0 = bf16_lo(pu[i]), p1 = bf16_hi(pu[i]);
        d[i].x = p0 != 0.f ? d[i].x : 0.f;
        d[i].y = p1 != 0.f ? d[i].y : 0.f;
        dot = fmaf(p0, d[i].x, dot);
        dot = fmaf(p1, d[i].y, dot);
    }
    dot = warp_sum(dot);
#pragma unroll
    for (int i = 0; i < 16; ++i) {
        const int k = 2 * lane + 64 * i;
        const float p0 = bf16_lo(pu[i]), p1 = bf16_hi(pu[i]);
        if (k < ld)
            *reinterpret_cast<uint32_t*>(orow + k) =
                pack_bf16(p0 != 0.f ? scale * p0 * (d[i].x - dot) : 0.f, p1 != 0.f ? scale * p1 * (d[i].y - dot) : 0.f);
    }
}

// dS[z][q][k] = scale * P (dP - sum_k P dP); written bf16 over all ld columns (0 where P == 0).
__global__ void __launch_bounds__(kAttnWarps * 32)
softmax_bwd_kernel(const __nv_bfloat16* __restrict__ P, const float* __restrict__ dP,
                   __nv_bfloat16* __restrict__ dS, long long rows, int Sk, int ld, float scale) {
    const long long item = static_cast<long long>(blockIdx.x) * kAttnWarps + (threadIdx.x >> 5);
    if (item >= rows) return;
    const int lane = threadIdx.x & 31;
    const __nv_bfloat16* prow = P + item * ld;
    const float* drow = dP + item * ld;
    __nv_bfloat16* orow = dS + item * ld;
    float dot = 0.f;
    for (int k = lane * 2; k < ld; k += 64) {
        const uint32_t u = *reinterpret_cast<const uint32_t*>(prow + k);
        const float p0 = bf16_lo(u), p1 = bf16_hi(u);
        if (p0 != 0.f) dot += p0 * drow[k];
        if (p1 != 0.f) dot += p1 * drow[k + 1];
    }
    dot = warp_sum(dot);
    for (int k = lane * 2; k < ld; k += 64) {
        const uint32_t u = *reinterpret_cast<const uint32_t*>(prow + k);
        const float p0 = bf16_lo(u), p1 = bf16_hi(u);
        const float o0 = p0 != 0.f ? scale * p0 * (drow[k] - dot) : 0.f;
        const float o1 = p1 != 0.f ? scale * p1 * (drow[k + 1] - dot) : 0.f;
        *reinterpret_cast<uint32_t*>(orow + k) = pack_bf16(o0, o1);
    }
}

// delta[(b*hq + head)*S + s] = sum_d dO[b,s,head,d] * O[b,s,head,d]  (= sum_k P dP of that score row).
// One warp per (token row, head).
__global__ void __launch_bounds__(kAttnWarps * 32)
attn_delta_kernel(const __nv_bfloat16* __restrict__ dO, const __nv_bfloat16* __restrict__ O,
                  float* __restrict__ delta, int rows, int S, int hq, int hd) {
    const int item = blockIdx.x * kAttnWarps + (threadIdx.x >> 5);
    if (item >= rows * hq) return;
    const int row = item / hq, head = item - row * hq;
    const int lane = threadIdx.x & 31;
    const long long off = (static_cast<long long>(row) * hq + head) * hd;
    float acc = 0.f;
    for (int i = lane * 4; i < hd; i += 128) {
        float a[4], b[4]; ld4(dO + off + i, a); ld4(O + off + i, b);
#pragma unroll
        for (int j = 0; j < 4; ++j) acc = fmaf(a[j], b[j], acc);
    }
    acc = warp_sum(acc);
    const int b = row / S, s = row - b * S;
    if (lane == 0) delta[(static_cast<long long>(b) * hq + head) * S + s] = acc;
}

}  // namespace pt

using namespace pt;

extern "C" int pt_qknorm_rope_fwd(const void* qkv, const float* wq, const float* wk, const float* cos_t,
                                  const float* sin_t, void* q_out, void* k_out, int rows, int S, int hq,
                                  int hkv, int hd, float eps, void* stream) {
    PT_REQUIRE(qkv && wq && wk && cos_t && sin_t && q_out && k_out, "pt_qknorm_rope_fwd: null pointer");
    PT_REQUIRE(rows > 0 && S > 0 && hq > 0 && hkv > 0 && hd % 8 == 0, "pt_qknorm_rope_fwd: bad shape");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long items = static_cast<long long>(rows) * (hq + hkv);
    const int nb = static_cast<int>((items + kAttnWarps - 1) / kAttnWarps);
    if (hd <= 256 && rows >= kRowKernelMinRows)
        qknorm_rope_fwd_row_kernel<<<(rows + kAttnWarps - 1) / kAttnWarps, kAttnWarps * 32, 0,
                                     static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(qkv), wq, wk, cos_t, sin_t, static_cast<__nv_bfloat16*>(q_out),
            static_cast<__nv_bfloat16*>(k_out), rows, S, hq, hkv, hd, eps);
    else if (hd <= 256)
        qknorm_rope_fwd_reg_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(qkv), wq, wk, cos_t, sin_t, static_cast<__nv_bfloat16*>(q_out),
            static_cast<__nv_bfloat16*>(k_out), rows, S, hq, hkv, hd, eps);
    else
        qknorm_rope_fwd_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(qkv), wq, wk, cos_t, sin_t, static_cast<__nv_bfloat16*>(q_out),
            static_cast<__nv_bfloat16*>(k_out), rows, S, hq, hkv, hd, eps);
    return check_launch("qknorm_rope_fwd");
}

extern "C" int pt_qknorm_rope_bwd(const void* dq, const void* dk, const void* qkv, const float* wq,
                                  const float* wk, const float* cos_t, const float* sin_t, void* d_qkv,
                                  int rows, int S, int hq, int hkv, int hd, float eps, void* stream) {
    PT_REQUIRE(dq && dk && qkv && wq && wk && cos_t && sin_t && d_qkv, "pt_qknorm_rope_bwd: null pointer");
    PT_REQUIRE(rows > 0 && S > 0 && hq > 0 && hkv > 0 && hd % 8 == 0, "pt_qknorm_rope_bwd: bad shape");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long items = static_cast<long long>(rows) * (hq + hkv);
    const int nb = static_cast<int>((items + kAttnWarps - 1) / kAttnWarps);
    if (hd <= 256)      // (a warp-per-row backward measured slower than this: 44.8 vs 40.6 us, 80 registers)
        qknorm_rope_bwd_reg_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(dq), static_cast<const __nv_bfloat16*>(dk),
            static_cast<const __nv_bfloat16*>(qkv), wq, wk, cos_t, sin_t, static_cast<__nv_bfloat16*>(d_qkv),
            rows, S, hq, hkv, hd, eps);
    else
        qknorm_rope_bwd_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(dq), static_cast<const __nv_bfloat16*>(dk),
            static_cast<const __nv_bfloat16*>(qkv), wq, wk, cos_t, sin_t, static_cast<__nv_bfloat16*>(d_qkv),
            rows, S, hq, hkv, hd, eps);
    return check_launch("qknorm_rope_bwd");
}

extern "C" int pt_softmax_fwd(const float* scores, void* P_bf16, const uint8_t* keyvalid, int Z, int Sq,
                              int Sk, int ld, int heads_per_batch, int causal, int window, void* stream) {
    PT_REQUIRE(scores && P_bf16, "pt_softmax_fwd: null pointer");
    PT_REQUIRE(Z > 0 && Sq > 0 && Sk > 0 && ld >= Sk && ld % 2 == 0 && heads_per_batch > 0,
               "pt_softmax_fwd: bad shape (ld must be even and >= Sk)");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long items = static_cast<long long>(Z) * Sq;
    const int nb = static_cast<int>((items + kAttnWarps - 1) / kAttnWarps);
    if (ld <= 1024)
        softmax_fwd_reg_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            scores, static_cast<__nv_bfloat16*>(P_bf16), keyvalid, Z, Sq, Sk, ld, heads_per_batch, causal, window);
    else
        softmax_fwd_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            scores, static_cast<__nv_bfloat16*>(P_bf16), keyvalid, Z, Sq, Sk, ld, heads_per_batch, causal, window);
    return check_launch("softmax_fwd");
}

extern "C" int pt_softmax_bwd(const void* P_bf16, const float* dP, void* dS_bf16, long long rows, int Sk,
                              int ld, float scale, int Sq, int causal, int window, void* stream) {
    PT_REQUIRE(P_bf16 && dP && dS_bf16, "pt_softmax_bwd: null pointer");
    PT_REQUIRE(rows > 0 && Sk > 0 && ld >= Sk && ld % 2 == 0, "pt_softmax_bwd: bad shape");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const int nb = static_cast<int>((rows + kAttnWarps - 1) / kAttnWarps);
    if (ld <= 1024)
        softmax_bwd_reg_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(P_bf16), dP, static_cast<__nv_bfloat16*>(dS_bf16), rows, Sk, ld, scale,
            Sq, causal, window);
    else
        softmax_bwd_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
            static_cast<const __nv_bfloat16*>(P_bf16), dP, static_cast<__nv_bfloat16*>(dS_bf16), rows, Sk, ld, scale);
    return check_launch("softmax_bwd");
}

extern "C" int pt_attn_delta(const void* dO_bf16, const void* O_bf16, float* delta, int rows, int S, int hq, int hd,
                             void* stream) {
    PT_REQUIRE(dO_bf16 && O_bf16 && delta, "pt_attn_delta: null pointer");
    PT_REQUIRE(rows > 0 && S > 0 && rows % S == 0 && hq > 0 && hd % 4 == 0, "pt_attn_delta: bad shape");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const long long items = static_cast<long long>(rows) * hq;
    const int nb = static_cast<int>((items + kAttnWarps - 1) / kAttnWarps);
    attn_delta_kernel<<<nb, kAttnWarps * 32, 0, static_cast<cudaStream_t>(stream)>>>(
        static_cast<const __nv_bfloat16*>(dO_bf16), static_cast<const __nv_bfloat16*>(O_bf16), delta, rows, S, hq, hd);
    return check_launch("attn_delta");
}
