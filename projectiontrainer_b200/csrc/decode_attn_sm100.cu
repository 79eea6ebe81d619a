// One cached decode step of Gemma3 / Qwen3 attention in ONE launch per layer (replaces, in the graphed decode step,
// pt_qknorm_rope_fwd + two cache row copies + the score GEMM + pt_softmax_fwd + the P V GEMM: 6 launches of a few
// microseconds each, whose launch gaps — not their work — were the cost of a 26-layer step).
//
// What one (sample, kv head) does (HF:models/gemma3/modeling_gemma3.py:278-300,362-384 with a DynamicCache, the
// per-token forward behind ref:Stage1/inference_stage1.py:170 `generate`):
//   q_g  = rope(rmsnorm(q head g))  for the G = hq / hkv query heads of the group, k_new likewise, v_new as is
//   K[b, slot] = k_new, V[b, slot] = v_new                                     (the cache append)
//   o_g  = softmax_k(scale * q_g . K[b, k]) V[b, k]   over  k in [lo, slot],  valid[b, k] != 0,
//          lo = window > 0 ? max(0, slot - window + 1) : 0
// Layout: the keys of one (sample, kv head) are split, in tiles of 64, over a CLUSTER of nsplit CTAs
// (flash-decoding).  A CTA stages a K/V tile in shared memory with 16-byte cp.async (K rows padded by 16 bytes: the
// per-key reads below are bank-conflict free), then
//   scores: warp w takes keys 8w..8w+7 of the tile: S[G x 8] = Q[G x HD] K^T on mma.sync m16n8k16 (the G <= 8 query
//           heads of the group are rows 0..G-1 of the 16-row A operand, kept in registers for the whole launch);
//   softmax: warp g updates head g's running (max, sum) over the tile and turns the scores into bf16 probabilities;
//   P V:    warp w takes channel blocks of 8: O[G x 8] += P[G x 64] V[64 x 8], V fragments by ldmatrix.trans.
// (tcgen05 does not fit here: its M is 64+ rows and a decode step has G = 1..8 per K/V row — the kernel is bound by
// the K/V bytes, the tensor cores only have to keep the arithmetic out of the way, which mma.sync does.)
// The CTAs of the cluster merge their (max, sum, o) through distributed shared memory into rank 0, which writes
// the bf16 output row.  The CTA that owns the tile containing `slot` appends this token's k / v to the cache.
// HBM-bound by the K/V rows it reads: 2 * (slot - lo + 1) * HD * 2 bytes per (sample, kv head).
// (Versions without tensor cores — a warp per key with shuffle reductions, then a thread per (key, head) — were
// bound by latency / the FMA pipe at 0.65 and 1.6 TB/s; see profiles/r02_decode_attn.txt.)
#include <cooperative_groups.h>

#include "pt_common.cuh"

namespace cg = cooperative_groups;

namespace pt {

constexpr int kDecThreads = 256;
constexpr int kDecWarps = kDecThreads / 32;
constexpr int kDecTile = 32;                            // keys per shared-memory tile
constexpr int kDecStages = 3;                           // tiles in flight per CTA (cp.async ring)

template <int EPL>
__device__ __forceinline__ void ld_row(const __nv_bfloat16* p, float (&v)[EPL]) {
    if constexpr (EPL == 8) {
        const uint4 u = __ldg(reinterpret_cast<const uint4*>(p));
        v[0] = bf16_lo(u.x); v[1] = bf16_hi(u.x); v[2] = bf16_lo(u.y); v[3] = bf16_hi(u.y);
        v[4] = bf16_lo(u.z); v[5] = bf16_hi(u.z); v[6] = bf16_lo(u.w); v[7] = bf16_hi(u.w);
    } else if constexpr (EPL == 4) {
        const uint2 u = __ldg(reinterpret_cast<const uint2*>(p));
        v[0] = bf16_lo(u.x); v[1] = bf16_hi(u.x); v[2] = bf16_lo(u.y); v[3] = bf16_hi(u.y);
    } else {
        const uint32_t u = __ldg(reinterpret_cast<const uint32_t*>(p));
        v[0] = bf16_lo(u); v[1] = bf16_hi(u);
    }
}

__device__ __forceinline__ void cp_async16(void* smem_dst, const void* gmem_src) {
    const uint32_t d = static_cast<uint32_t>(__cvta_generic_to_shared(smem_dst));
    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d), "l"(gmem_src) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() { asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory"); }

struct DecodeAttnArgs {
    const __nv_bfloat16* qkv;       // [B, (hq + 2 hkv) * HD]: this token's fused projection
    const float *wq, *wk;           // [HD] norm gains (applied as 1 + w)
    const float *cs, *sn;           // [rows, HD/2] rotary tables
    const long long* pos;           // [B] table row of this token, per sample
    const long long* slot;          // [1] cache slot of this token
    const uint8_t* valid;           // [B, Smax] key validity (prompt pads = 0), or null
    __nv_bfloat16 *K, *V;           // [B, Smax, hkv * HD]
    __nv_bfloat16* out;             // [B, hq * HD]
    int B, Smax, hq, hkv, window, table_rows;
    float eps, scale;
};

// shared-memory carve-up (bytes), the same on the host (size) and the device (offsets)
template <int HD, int G>
struct DecSmem {
    static constexpr int KROW = HD + 8;                                     // bf16 per staged K / V row (16-byte pad:
                                                                            // fragment loads hit 32 distinct banks)
    static constexpr int PROW = kDecTile + 8;                               // bf16 per probability row
    static constexpr int oK = 0;
    static constexpr int stage = kDecTile * KROW * 2;                       // one K (or V) tile
    static constexpr int oV = oK + kDecStages * stage;
    static constexpr int oQ = oV + kDecStages * stage;                      // bf16 [G][HD]
    static constexpr int oKnew = oQ + G * HD * 2;                           // bf16 [HD]
    static constexpr int oVnew = oKnew + HD * 2;                            // bf16 [HD]
    static constexpr int oSc = oVnew + HD * 2;                              // f32 [G][tile]: scores
    static constexpr int oP = oSc + G * kDecTile * 4;                       // bf16 [G][PROW]: probabilities
    static constexpr int oCorr = oP + G * PROW * 2;                         // f32 [G]
    static constexpr int oRml = oCorr + 16 * 4;                             // f32 [G][2]  merged (max, sum)
    static constexpr int oRo = oRml + 16 * 4;                               // f32 [G][HD] merged o
    static constexpr int bytes = oRo + G * HD * 4;
};

// D += A B for one m16n8k16 bf16 tile (A row-major 16x16, B "col": 16x8 given as [n][k] rows), f32 accumulate
__device__ __forceinline__ void mma16816(float (&d)[4], uint32_t a0, uint32_t a1, uint32_t a2, uint32_t a3,
                                         uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k16.row.col.f32.bf16.bf16.f32 {%0,%1,%2,%3}, {%4,%5,%6,%7}, {%8,%9}, "
                 "{%0,%1,%2,%3};"
                 : "+f"(d[0]), "+f"(d[1]), "+f"(d[2]), "+f"(d[3])
                 : "r"(a0), "r"(a1), "r"(a2), "r"(a3), "r"(b0), "r"(b1));
}
// two 8x8 bf16 blocks of a row-major [k][n] tile, transposed into the B fragment of m16n8k16 (rows = k)
__device__ __forceinline__ void ldmatrix_x2_trans(uint32_t& b0, uint32_t& b1, const void* smem_row) {
    const uint32_t a = static_cast<uint32_t>(__cvta_generic_to_shared(smem_row));
    asm volatile("ldmatrix.sync.aligned.m8n8.x2.trans.shared.b16 {%0,%1}, [%2];" : "=r"(b0), "=r"(b1) : "r"(a));
}

template <int HD, int G>
__global__ void __launch_bounds__(kDecThreads) decode_attn_kernel(const DecodeAttnArgs p) {
    using L = DecSmem<HD, G>;
    constexpr int EPL = HD >= 64 ? HD / 32 : 2;         // norm + RoPE: channels per lane
    constexpr int LA = HD / EPL;                        // lanes that hold channels (32 from HD = 64 up)
    constexpr int HALF = HD / 2;
    constexpr int T = kDecTile, KROW = L::KROW;
    constexpr int PROW = L::PROW;
    constexpr int CH = HD / 8;                          // 16-byte chunks per row
    constexpr int KS = HD / 16;                         // k-steps of the score MMAs
    constexpr int NTW = HD >= 64 ? HD / 64 : 1;         // 8-channel blocks of O per warp (block = warp + 8 t)
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oK);
    __nv_bfloat16* Vs = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oV);
    __nv_bfloat16* qs = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oQ);
    __nv_bfloat16* knew = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oKnew);
    __nv_bfloat16* vnew = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oVnew);
    float* sc = reinterpret_cast<float*>(smem_raw + L::oSc);
    __nv_bfloat16* Pb = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oP);
    float* corr_s = reinterpret_cast<float*>(smem_raw + L::oCorr);
    float* r_ml = reinterpret_cast<float*>(smem_raw + L::oRml);
    float* r_o = reinterpret_cast<float*>(smem_raw + L::oRo);

    cg::cluster_group cluster = cg::this_cluster();
    const int nsplit = static_cast<int>(cluster.num_blocks()), rank = static_cast<int>(cluster.block_rank());
    const int item = blockIdx.x / nsplit;
    const int b = item / p.hkv, kvh = item - b * p.hkv;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    const int slot = static_cast<int>(*p.slot);
    long long trow = p.pos[b];
    if (slot < 0 || slot >= p.Smax) return;             // the whole grid reads the same slot: nobody waits below
    trow = trow < 0 ? 0 : (trow >= p.table_rows ? p.table_rows - 1 : trow);
    const int kvd = p.hkv * HD;
    const __nv_bfloat16* row = p.qkv + static_cast<long long>(b) * (p.hq + 2 * p.hkv) * HD;

    // ---- this token: q heads of the group (bf16 like pt_qknorm_rope_fwd's q_out), k, v
    {
        const bool on = lane < LA;
        float cc[EPL], ss[EPL];
#pragma unroll
        for (int j = 0; j < EPL; ++j) { cc[j] = 0.f; ss[j] = 0.f; }
        if (on) {   // channel i < HD/2 pairs with i + HD/2: lane ^ (LA/2), same j
            const int ti = (lane & (LA / 2 - 1)) * EPL;
            const float sgn = lane < LA / 2 ? -1.f : 1.f;
#pragma unroll
            for (int j = 0; j < EPL; ++j) {
                cc[j] = p.cs[trow * HALF + ti + j];
                ss[j] = sgn * p.sn[trow * HALF + ti + j];
            }
        }
        for (int h = warp; h < G + 1; h += kDecWarps) {
            const bool isq = h < G;
            const __nv_bfloat16* u = row + (isq ? kvh * G + h : p.hq + kvh) * HD;
            const float* w = isq ? p.wq : p.wk;
            float a[EPL];
#pragma unroll
            for (int j = 0; j < EPL; ++j) a[j] = 0.f;
            if (on) ld_row<EPL>(u + lane * EPL, a);
            float q = 0.f;
#pragma unroll
            for (int j = 0; j < EPL; ++j) q += a[j] * a[j];
            const float rstd = rsqrtf(warp_sum(q) / HD + p.eps);
#pragma unroll
            for (int j = 0; j < EPL; ++j) {
                const float n = on ? a[j] * rstd * (1.f + w[lane * EPL + j]) : 0.f;
                const float o = __shfl_xor_sync(0xffffffffu, n, LA / 2);
                const float y = n * cc[j] + o * ss[j];
                if (on) {
                    (isq ? qs + h * HD : knew)[lane * EPL + j] = __float2bfloat16_rn(y);
                }
            }
        }
        if (tid < CH)
            reinterpret_cast<uint4*>(vnew)[tid] =
                __ldg(reinterpret_cast<const uint4*>(row + (p.hq + p.hkv + kvh) * HD) + tid);
    }
    __syncthreads();

    // keys lo .. slot: cache rows, except `slot` itself = this token (staged from knew / vnew, appended to the cache)
    const int lo = p.window > 0 ? max(0, slot - p.window + 1) : 0;
    const int nkeys = slot - lo + 1;
    const int ntiles = (nkeys + T - 1) / T;
    const long long cbase = (static_cast<long long>(b) * p.Smax) * kvd + kvh * HD;
    const uint8_t* vb = p.valid ? p.valid + static_cast<long long>(b) * p.Smax : nullptr;
    const int fr = lane >> 2, fk = (lane & 3) * 2;      // fragment row (query head / key) and column pair
    uint32_t qa[KS][2];                                  // A fragments of Q: rows >= G (and 8..15) are zero
#pragma unroll
    for (int ks = 0; ks < KS; ++ks) {
        qa[ks][0] = fr < G ? *reinterpret_cast<const uint32_t*>(qs + fr * HD + ks * 16 + fk) : 0u;
        qa[ks][1] = fr < G ? *reinterpret_cast<const uint32_t*>(qs + fr * HD + ks * 16 + 8 + fk) : 0u;
    }
    float acc[NTW][4];
#pragma unroll
    for (int t = 0; t < NTW; ++t) { acc[t][0] = 0.f; acc[t][1] = 0.f; acc[t][2] = 0.f; acc[t][3] = 0.f; }
    float m_run = -INFINITY, l_run = 0.f;               // head `warp`'s running state (warps < G)

    // stage one tile (cache rows by cp.async; `slot` = this token, also appended to the cache; rows past it zeroed:
    // P is 0 there and V must not hold NaN bits)
    auto stage_tile = [&](int tile, int st) {
        if (tile < ntiles) {
            __nv_bfloat16* Kd = Ks + st * (T * KROW);
            __nv_bfloat16* Vd = Vs + st * (T * KROW);
            const int k0 = lo + tile * T;
            for (int c = tid; c < T * CH; c += kDecThreads) {
                const int r = c / CH, ci = c - r * CH;
                const int key = k0 + r;
                const long long off = cbase + static_cast<long long>(key) * kvd + ci * 8;
                if (key < slot) {
                    cp_async16(Kd + r * KROW + ci * 8, p.K + off);
                    cp_async16(Vd + r * KROW + ci * 8, p.V + off);
                } else if (key == slot) {
                    const uint4 kk = reinterpret_cast<const uint4*>(knew)[ci], vv = reinterpret_cast<const uint4*>(vnew)[ci];
                    *reinterpret_cast<uint4*>(Kd + r * KROW + ci * 8) = kk;
                    *reinterpret_cast<uint4*>(Vd + r * KROW + ci * 8) = vv;
                    *reinterpret_cast<uint4*>(p.K + off) = kk;
                    *reinterpret_cast<uint4*>(p.V + off) = vv;
                } else {
                    *reinterpret_cast<uint4*>(Kd + r * KROW + ci * 8) = make_uint4(0u, 0u, 0u, 0u);
                    *reinterpret_cast<uint4*>(Vd + r * KROW + ci * 8) = make_uint4(0u, 0u, 0u, 0u);
                }
            }
        }
        cp_async_commit();                               // one group per call, empty or not: the wait counts groups
    };
#pragma unroll
    for (int s = 0; s < kDecStages - 1; ++s) stage_tile(rank + s * nsplit, s);

    int it = 0;
    for (int tile = rank; tile < ntiles; tile += nsplit, ++it) {
        const int k0 = lo + tile * T;
        const int tn = min(T, nkeys - tile * T);
        const __nv_bfloat16* Kt = Ks + (it % kDecStages) * (T * KROW);
        const __nv_bfloat16* Vt = Vs + (it % kDecStages) * (T * KROW);
        cp_async_wait<kDecStages - 2>();                 // this tile has landed (the younger groups may be in flight)
        __syncthreads();                                 // ... for every thread; and the previous tile is consumed
        stage_tile(tile + (kDecStages - 1) * nsplit, (it + kDecStages - 1) % kDecStages);
        if (warp < T / 8) {   // scores of keys 8 warp .. 8 warp + 7
            const int r0 = warp * 8 + fk;
            bool ok0 = false, ok1 = false;
            if (fr < G) {
                ok0 = r0 < tn && (vb == nullptr || vb[k0 + r0] != 0);
                ok1 = r0 + 1 < tn && (vb == nullptr || vb[k0 + r0 + 1] != 0);
            }
            const __nv_bfloat16* kr = Kt + (warp * 8 + fr) * KROW + fk;
            float c[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
            for (int ks = 0; ks < KS; ++ks) {
                const uint32_t b0 = *reinterpret_cast<const uint32_t*>(kr + ks * 16);
                const uint32_t b1 = *reinterpret_cast<const uint32_t*>(kr + ks * 16 + 8);
                mma16816(c, qa[ks][0], 0u, qa[ks][1], 0u, b0, b1);
            }
            if (fr < G) {
                sc[fr * T + r0] = ok0 ? c[0] * p.scale : -INFINITY;
                sc[fr * T + r0 + 1] = ok1 ? c[1] * p.scale : -INFINITY;
            }
        }
        __syncthreads();
        if (warp < G) {                                  // head `warp`: running max / sum, scores -> probabilities
            static_assert(T == 32, "one score per lane");
            const float s0 = sc[warp * T + lane];
            const float mn = fmaxf(m_run, warp_max(s0));
            const float corr = m_run == -INFINITY ? 0.f : __expf(m_run - mn);
            const float p0 = s0 == -INFINITY ? 0.f : __expf(s0 - mn);
            Pb[warp * PROW + lane] = __float2bfloat16_rn(p0);
            l_run = l_run * corr + warp_sum(p0);
            m_run = mn;
            if (lane == 0) corr_s[warp] = corr;
        }
        __syncthreads();
        {   // O[head, 8-channel blocks of this warp] = corr * O + P V
            const float cg_ = fr < G ? corr_s[fr] : 0.f;
#pragma unroll
            for (int t = 0; t < NTW; ++t) { acc[t][0] *= cg_; acc[t][1] *= cg_; }
#pragma unroll
            for (int kk = 0; kk < T / 16; ++kk) {
                if (kk * 16 < tn) {
                    const uint32_t a0 = fr < G ? *reinterpret_cast<const uint32_t*>(Pb + fr * PROW + kk * 16 + fk) : 0u;
                    const uint32_t a2 = fr < G ? *reinterpret_cast<const uint32_t*>(Pb + fr * PROW + kk * 16 + 8 + fk) : 0u;
#pragma unroll
                    for (int t = 0; t < NTW; ++t) {
                        const int nt = warp + kDecWarps * t;
                        if (nt < HD / 8) {
                            uint32_t b0, b1;
                            ldmatrix_x2_trans(b0, b1, Vt + (kk * 16 + (lane & 15)) * KROW + nt * 8);
                            mma16816(acc[t], a0, 0u, a2, 0u, b0, b1);
                        }
                    }
                }
            }
        }
    }
    cp_async_wait<0>();

    // ---- this CTA's (max, sum, o) for the cluster merge
    if (fr < G) {
#pragma unroll
        for (int t = 0; t < NTW; ++t) {
            const int nt = warp + kDecWarps * t;
            if (nt < HD / 8) {
                r_o[fr * HD + nt * 8 + fk] = acc[t][0];
                r_o[fr * HD + nt * 8 + fk + 1] = acc[t][1];
            }
        }
    }
    if (warp < G && lane == 0) { r_ml[warp * 2] = m_run; r_ml[warp * 2 + 1] = l_run; }
    // ---- merge the CTAs of the cluster into rank 0 (reads of the partners' shared memory)
    cluster.sync();
    if (rank == 0) {
        __nv_bfloat16* out = p.out + (static_cast<long long>(b) * p.hq + kvh * G) * HD;
        for (int idx = tid; idx < G * HD / 2; idx += kDecThreads) {
            const int g = (idx * 2) / HD;
            float M = -INFINITY;
            for (int r = 0; r < nsplit; ++r) M = fmaxf(M, cluster.map_shared_rank(r_ml, r)[g * 2]);
            float a0 = 0.f, a1 = 0.f, Lsum = 0.f;
            for (int r = 0; r < nsplit; ++r) {
                const float* ml = cluster.map_shared_rank(r_ml, r);
                const float mr = ml[g * 2];
                const float f = mr == -INFINITY ? 0.f : __expf(mr - M);
                const float2 v = *reinterpret_cast<const float2*>(cluster.map_shared_rank(r_o, r) + idx * 2);
                a0 += f * v.x; a1 += f * v.y; Lsum += f * ml[g * 2 + 1];
            }
            const float inv = 1.f / Lsum;               // key `slot` is always visible: Lsum > 0
            *reinterpret_cast<uint32_t*>(out + idx * 2) = pack_bf16(a0 * inv, a1 * inv);
        }
    }
    cluster.sync();                                      // partners keep their shared memory until rank 0 has read it
}

template <int HD, int G>
static int launch_decode_attn(const DecodeAttnArgs& a, int nsplit, cudaStream_t stream) {
    const size_t smem = DecSmem<HD, G>::bytes;
    if (smem > 48 * 1024) {
        static bool done[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 64 && !done[dev]) {
            const cudaError_t e = cudaFuncSetAttribute(decode_attn_kernel<HD, G>,
                                                       cudaFuncAttributeMaxDynamicSharedMemorySize, static_cast<int>(smem));
            if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(decode_attn): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
            done[dev] = true;
        }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(a.B * a.hkv * nsplit));
    cfg.blockDim = dim3(kDecThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = static_cast<unsigned>(nsplit);
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    const cudaError_t e = cudaLaunchKernelEx(&cfg, decode_attn_kernel<HD, G>, a);
    if (e != cudaSuccess) { set_error("decode_attn launch: %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
    return check_launch("decode_attn");
}

template <int HD>
static int dispatch_group(const DecodeAttnArgs& a, int G, int nsplit, cudaStream_t stream) {
    switch (G) {
        case 1: return launch_decode_attn<HD, 1>(a, nsplit, stream);
        case 2: return launch_decode_attn<HD, 2>(a, nsplit, stream);
        case 4: return launch_decode_attn<HD, 4>(a, nsplit, stream);
        case 8: return launch_decode_attn<HD, 8>(a, nsplit, stream);
    }
    return PT_ERR_INVALID;
}

}  // namespace pt

using namespace pt;

extern "C" int pt_decode_attn_supported(int hq, int hkv, int hd) {
    if (hq <= 0 || hkv <= 0 || hq % hkv) return 0;
    const int G = hq / hkv;
    return (hd == 16 || hd == 32 || hd == 64 || hd == 128 || hd == 256) && (G == 1 || G == 2 || G == 4 || G == 8);
}

extern "C" int pt_decode_attn(const void* qkv, const float* wq, const float* wk, const float* cos_t,
                              const float* sin_t, int table_rows, const long long* pos, const long long* slot,
                              const uint8_t* valid, void* K, void* V, void* out, int B, int Smax, int hq, int hkv,
                              int hd, int window, float eps, float scale, int nsplit, void* stream) {
    PT_REQUIRE(qkv && wq && wk && cos_t && sin_t && pos && slot && K && V && out, "pt_decode_attn: null pointer");
    PT_REQUIRE(B > 0 && Smax > 0 && table_rows > 0, "pt_decode_attn: bad shape");
    PT_REQUIRE(pt_decode_attn_supported(hq, hkv, hd),
               "pt_decode_attn: head_dim must be 16..256 (power of two) and hq / hkv one of 1, 2, 4, 8");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    if (nsplit <= 0) {                                   // a CTA per SM or more (measured: profiles/r02_decode_attn.txt)
        nsplit = 1;
        while (nsplit < 8 && B * hkv * nsplit < 148) nsplit *= 2;
    }
    PT_REQUIRE(nsplit == 1 || nsplit == 2 || nsplit == 4 || nsplit == 8, "pt_decode_attn: nsplit must be 1, 2, 4 or 8");
    DecodeAttnArgs a;
    a.qkv = static_cast<const __nv_bfloat16*>(qkv); a.wq = wq; a.wk = wk; a.cs = cos_t; a.sn = sin_t;
    a.pos = pos; a.slot = slot; a.valid = valid;
    a.K = static_cast<__nv_bfloat16*>(K); a.V = static_cast<__nv_bfloat16*>(V);
    a.out = static_cast<__nv_bfloat16*>(out);
    a.B = B; a.Smax = Smax; a.hq = hq; a.hkv = hkv; a.window = window; a.table_rows = table_rows;
    a.eps = eps; a.scale = scale;
    const int G = hq / hkv;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    switch (hd) {
        case 16: return dispatch_group<16>(a, G, nsplit, st);
        case 32: return dispatch_group<32>(a, G, nsplit, st);
        case 64: return dispatch_group<64>(a, G, nsplit, st);
        case 128: return dispatch_group<128>(a, G, nsplit, st);
        case 256: return dispatch_group<256>(a, G, nsplit, st);
    }
    return PT_ERR_INVALID;
}
