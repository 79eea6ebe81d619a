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
// Layout: the keys of one (sample, kv head) are split, in tiles of 32, over a CLUSTER of nsplit CTAs
// (flash-decoding).  A CTA stages K/V tiles in a two-stage shared-memory ring with 16-byte cp.async (unpadded rows,
// chunks XOR-swizzled so that the fragment reads below are bank-conflict free), then, per tile,
//   scores: warp w takes keys 8 (w & 3) .. + 7 and half of the contraction (w >> 2):
//           S[G x 8] = Q[G x HD] K^T on mma.sync m16n8k16 (the G <= 8 query heads of the group are rows 0..G-1 of the
//           16-row A operand, kept in registers for the whole launch; the contraction index is permuted inside each
//           32-wide chunk so that a lane's K fragments of two k-steps are one 16-byte read);
//   softmax + P V: every warp turns the scores of its lanes' query head into the A fragments of the P V product in
//           registers (a quad of lanes holds the 32 scores of a head: two shuffle steps give the tile maximum and
//           sum; every warp keeps the same running (max, sum)), then takes channel blocks of 8:
//           O[G x 8] += P[G x 32] V[32 x 8], V fragments by ldmatrix.trans.  Two block barriers per tile.
// (tcgen05 does not fit here: its M is 64+ rows and a decode step has G = 1..8 per K/V row; the tensor cores only
// have to keep the arithmetic out of the way, which mma.sync does.)
// The CTAs of the cluster merge their (max, sum, o) through distributed shared memory into rank 0, which writes
// the bf16 output row.  The CTA that owns the tile containing `slot` appends this token's k / v to the cache.
// Algorithmic bytes: the K/V rows it reads, 2 * (slot - lo + 1) * HD * 2 per (sample, kv head) — and that is what it
// reads (ncu: 42.22 MB against 42.0 MB at B = 64).  It runs at 2.8 of 7.7 TB/s there: bound by the instruction stream
// of a tile at two warps per scheduler and by ~5.5 us that do not depend on the key count, not by HBM.
// (Versions without tensor cores — a warp per key with shuffle reductions, then a thread per (key, head) — were
// bound by latency / the FMA pipe at 0.65 and 1.6 TB/s; see profiles/r02_decode_attn.txt.)
#include <cooperative_groups.h>

#include "pt_common.cuh"

namespace cg = cooperative_groups;

namespace pt {

constexpr int kDecThreads = 256;
constexpr int kDecWarps = kDecThreads / 32;
constexpr int kDecTile = 32;                            // keys per shared-memory tile
constexpr int kDecStages = 2;                           // tiles in flight per CTA (cp.async ring)

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
    int vm_bytes;                   // shared-memory bytes for the key-visibility table (0: read `valid` per tile)
    float eps, scale;
};

// shared-memory carve-up (bytes), the same on the host (size) and the device (offsets)
template <int HD, int G>
struct DecSmem {
    // bf16 per staged row.  HD >= 64: rows are unpadded and their 16-byte chunks XOR-swizzled — K: chunk ^ 4 (row & 1)
    // (a quarter-warp reads 4 chunks of an even and an odd row: disjoint banks), V: chunk ^ (row & 7) (ldmatrix reads
    // one chunk column of 8 rows).  Below 64 (tests only): 16-byte pad, 4-byte fragment reads.
    static constexpr int KROW = HD >= 64 ? HD : HD + 8;
    static constexpr int VROW = HD >= 64 ? HD : HD + 8;
    static constexpr int SROW = kDecTile + 8;                               // f32 per score row (bank spread)
    static constexpr int kstage = kDecTile * KROW * 2;
    static constexpr int vstage = kDecTile * VROW * 2;
    static constexpr int oK = 0;
    static constexpr int oV = oK + kDecStages * kstage;
    static constexpr int oQ = oV + kDecStages * vstage;                     // bf16 [G][HD]
    static constexpr int oKnew = oQ + G * HD * 2;                           // bf16 [HD]
    static constexpr int oVnew = oKnew + HD * 2;                            // bf16 [HD]
    static constexpr int oSc = oVnew + HD * 2;                              // f32 [2][G][SROW]: score partials
    static constexpr int oVm = oSc + 2 * G * SROW * 4;                      // u8 [vm_bytes]: key visibility, this CTA's tiles
    // after the key-visibility bytes (vm_bytes, a multiple of 16, chosen by the host):
    static constexpr int rml = 16 * 4;                                      // f32 [G][2]  merged (max, sum)
    static constexpr int ro = G * HD * 4;                                   // f32 [G][HD] merged o
    static constexpr int fixed = oVm + rml + ro;
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
__device__ __forceinline__ float ex2(float x) {
    float y;
    asm("ex2.approx.ftz.f32 %0, %1;" : "=f"(y) : "f"(x));
    return y;
}

template <int HD, int G>
__global__ void __launch_bounds__(kDecThreads) decode_attn_kernel(const DecodeAttnArgs p) {
    using L = DecSmem<HD, G>;
    constexpr int EPL = HD >= 64 ? HD / 32 : 2;         // norm + RoPE: channels per lane
    constexpr int LA = HD / EPL;                        // lanes that hold channels (32 from HD = 64 up)
    constexpr int HALF = HD / 2;
    constexpr int T = kDecTile, KROW = L::KROW, VROW = L::VROW, SROW = L::SROW;
    constexpr int CH = HD / 8;                          // 16-byte chunks per row
    constexpr bool kWide = HD >= 64;                    // scores: 16-byte fragments, k-steps split over warp pairs
    constexpr int RSTEP = kDecThreads / CH;             // rows between a thread's chunks of a tile
    constexpr int NJ = (T + RSTEP - 1) / RSTEP;         // chunks per thread per tile (per operand)
    constexpr int KS = HD / 16;                         // k-steps of the score MMAs
    constexpr int NPW = kWide ? HD / 64 : 0;            // pairs of k-steps per warp (wide path)
    constexpr int NTW = HD >= 64 ? HD / 64 : 1;         // 8-channel blocks of O per warp (block = warp + 8 t)
    static_assert(T == 32, "a quad of lanes holds the 32 scores of a head");
    extern __shared__ __align__(16) uint8_t smem_raw[];
    __nv_bfloat16* Ks = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oK);
    __nv_bfloat16* Vs = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oV);
    __nv_bfloat16* qs = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oQ);
    __nv_bfloat16* knew = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oKnew);
    __nv_bfloat16* vnew = reinterpret_cast<__nv_bfloat16*>(smem_raw + L::oVnew);
    float* sc = reinterpret_cast<float*>(smem_raw + L::oSc);
    uint8_t* vm = smem_raw + L::oVm;
    float* r_ml = reinterpret_cast<float*>(smem_raw + L::oVm + p.vm_bytes);
    float* r_o = r_ml + 16;

    cg::cluster_group cluster = cg::this_cluster();
    const int nsplit = static_cast<int>(cluster.num_blocks()), rank = static_cast<int>(cluster.block_rank());
    const int item = blockIdx.x / nsplit;
    const int b = item / p.hkv, kvh = item - b * p.hkv;
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    pdl_trigger();
    pdl_wait();                                          // everything read below was written by earlier kernels
    const int slot = static_cast<int>(*p.slot);
    long long trow = p.pos[b];
    if (slot < 0 || slot >= p.Smax) return;             // the whole grid reads the same slot: nobody waits below
    trow = trow < 0 ? 0 : (trow >= p.table_rows ? p.table_rows - 1 : trow);
    const int kvd = p.hkv * HD;
    const __nv_bfloat16* row = p.qkv + static_cast<long long>(b) * (p.hq + 2 * p.hkv) * HD;

    // keys lo .. slot: cache rows, except `slot` itself = this token (staged from knew / vnew, appended to the cache)
    const int lo = p.window > 0 ? max(0, slot - p.window + 1) : 0;
    const int nkeys = slot - lo + 1;
    const int ntiles = (nkeys + T - 1) / T;
    const uint8_t* vb = p.valid ? p.valid + static_cast<long long>(b) * p.Smax : nullptr;
    const bool vm_on = p.vm_bytes > 0;

    auto kchunk = [](int r, int ci) { return kWide ? ci ^ ((r & 1) << 2) : ci; };
    auto vchunk = [](int r, int ci) { return kWide ? ci ^ (r & 7) : ci; };
    static_assert(!kWide || RSTEP % 8 == 0, "a thread's rows of a tile share their swizzle");
    // ---- staging.  A thread copies the 16-byte chunk `sci` of rows sr0, sr0 + RSTEP, ... of a tile.
    const int sr0 = tid / CH, sci = tid - sr0 * CH;
    const long long cbase = (static_cast<long long>(b) * p.Smax) * kvd + kvh * HD;
    const long long gofs0 = cbase + static_cast<long long>(lo + sr0) * kvd + sci * 8;
    // the cache rows of a tile (every row of every tile but the last one)
    auto stage_rows = [&](int tile, int st) {
        if (tile < ntiles) {
            __nv_bfloat16* Kd = Ks + st * (T * KROW) + sr0 * KROW + kchunk(sr0, sci) * 8;   // rows sr0 + j RSTEP: the
            __nv_bfloat16* Vd = Vs + st * (T * VROW) + sr0 * VROW + vchunk(sr0, sci) * 8;   // same swizzle (RSTEP % 8 == 0)
            const long long g0 = gofs0 + static_cast<long long>(tile) * T * kvd;
            const int nrow = min(T, slot - (lo + tile * T));            // rows before `slot`
#pragma unroll
            for (int j = 0; j < NJ; ++j) {
                if (sr0 + j * RSTEP < nrow) {
                    cp_async16(Kd + j * RSTEP * KROW, p.K + g0 + static_cast<long long>(j) * RSTEP * kvd);
                    cp_async16(Vd + j * RSTEP * VROW, p.V + g0 + static_cast<long long>(j) * RSTEP * kvd);
                }
            }
        }
        cp_async_commit();                               // one group per call, empty or not: the wait counts groups
    };
    // the last tile: row `slot` = this token (also appended to the cache), the rows past it zeroed (P is 0 there
    // and V must not hold NaN bits)
    auto patch_rows = [&](int tile, int st) {
        if (tile == ntiles - 1) {
            const int k0 = lo + tile * T;
            for (int c = tid; c < T * CH; c += kDecThreads) {
                const int r = c / CH, ci = c - r * CH;
                const int key = k0 + r;
                if (key < slot) continue;
                uint4 kk = make_uint4(0u, 0u, 0u, 0u), vv = kk;
                if (key == slot) {
                    kk = reinterpret_cast<const uint4*>(knew)[ci];
                    vv = reinterpret_cast<const uint4*>(vnew)[ci];
                    const long long off = cbase + static_cast<long long>(key) * kvd + ci * 8;
                    *reinterpret_cast<uint4*>(p.K + off) = kk;
                    *reinterpret_cast<uint4*>(p.V + off) = vv;
                }
                *reinterpret_cast<uint4*>(Ks + st * (T * KROW) + r * KROW + kchunk(r, ci) * 8) = kk;
                *reinterpret_cast<uint4*>(Vs + st * (T * VROW) + r * VROW + vchunk(r, ci) * 8) = vv;
            }
        }
    };
    // the first tiles are on their way while this token's heads are normalised
#pragma unroll
    for (int s = 0; s < kDecStages - 1; ++s) stage_rows(rank + s * nsplit, s);

    if (vm_on) {    // visibility of this CTA's keys, local tile by local tile
        const int nloc = ntiles > rank ? (ntiles - rank + nsplit - 1) / nsplit : 0;
        for (int i = tid; i < nloc * T && i < p.vm_bytes; i += kDecThreads) {
            const int key = lo + (rank + (i / T) * nsplit) * T + (i % T);
            vm[i] = key <= slot && (vb == nullptr || vb[key] != 0);
        }
    }

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
                if (on) (isq ? qs + h * HD : knew)[lane * EPL + j] = __float2bfloat16_rn(y);
            }
        }
        if (tid < CH)
            reinterpret_cast<uint4*>(vnew)[tid] =
                __ldg(reinterpret_cast<const uint4*>(row + (p.hq + p.hkv + kvh) * HD) + tid);
    }
    __syncthreads();
#pragma unroll
    for (int s = 0; s < kDecStages - 1; ++s) patch_rows(rank + s * nsplit, s);

    const int fr = lane >> 2, fk = (lane & 3) * 2;      // fragment row (query head / key) and column pair
    // A fragments of Q (rows >= G and 8..15 are zero).  Wide path: warp w scores keys 8 (w & 3) .. + 7 over the
    // k-step pairs NPW (w >> 2) .. + NPW - 1; the contraction index is permuted inside each pair so that a lane's two
    // fragments of two k-steps are 16 contiguous bytes (the same permutation on Q and K leaves the sum unchanged).
    const int nt_qk = warp & 3, kh = warp >> 2;
    uint4 qw[kWide ? NPW : 1];
    uint32_t qa[kWide ? 1 : KS][2];
    if constexpr (kWide) {
#pragma unroll
        for (int g = 0; g < NPW; ++g)
            qw[g] = fr < G ? *reinterpret_cast<const uint4*>(qs + fr * HD + (kh * NPW + g) * 32 + (lane & 3) * 8)
                           : make_uint4(0u, 0u, 0u, 0u);
    } else {
#pragma unroll
        for (int ks = 0; ks < KS; ++ks) {
            qa[ks][0] = fr < G ? *reinterpret_cast<const uint32_t*>(qs + fr * HD + ks * 16 + fk) : 0u;
            qa[ks][1] = fr < G ? *reinterpret_cast<const uint32_t*>(qs + fr * HD + ks * 16 + 8 + fk) : 0u;
        }
    }
    float acc[NTW][4];
#pragma unroll
    for (int t = 0; t < NTW; ++t) { acc[t][0] = 0.f; acc[t][1] = 0.f; acc[t][2] = 0.f; acc[t][3] = 0.f; }
    float m_run = -INFINITY, l_run = 0.f;               // running state of head `fr`, log2 domain (every warp the same)
    const float sl2 = p.scale * 1.4426950408889634f;

    int it = 0;
    for (int tile = rank; tile < ntiles; tile += nsplit, ++it) {
        const int st = it % kDecStages;
        const __nv_bfloat16* Kt = Ks + st * (T * KROW);
        const __nv_bfloat16* Vt = Vs + st * (T * VROW);
        cp_async_wait<kDecStages - 2>();                 // this tile has landed (the younger groups may be in flight)
        __syncthreads();                                 // ... for every thread; and the previous tile is consumed
        {
            const int nxt = tile + (kDecStages - 1) * nsplit, nst = (it + kDecStages - 1) % kDecStages;
            stage_rows(nxt, nst);
            patch_rows(nxt, nst);
        }
        if (kWide || warp < T / 8) {   // score partials of keys 8 nt_qk .. + 7
            const int r0 = nt_qk * 8 + fk;
            float c0, c1;
            if constexpr (kWide) {
                const __nv_bfloat16* kr = Kt + (nt_qk * 8 + fr) * KROW;
                float ca[4] = {0.f, 0.f, 0.f, 0.f}, cb[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int g = 0; g < NPW; ++g) {
                    const uint4 kk =
                        *reinterpret_cast<const uint4*>(kr + kchunk(fr, (kh * NPW + g) * 4 + (lane & 3)) * 8);
                    mma16816(ca, qw[g].x, 0u, qw[g].y, 0u, kk.x, kk.y);
                    mma16816(cb, qw[g].z, 0u, qw[g].w, 0u, kk.z, kk.w);
                }
                c0 = ca[0] + cb[0]; c1 = ca[1] + cb[1];
            } else {
                const __nv_bfloat16* kr = Kt + (nt_qk * 8 + fr) * KROW + fk;
                float ca[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
                for (int ks = 0; ks < KS; ++ks)
                    mma16816(ca, qa[ks][0], 0u, qa[ks][1], 0u, *reinterpret_cast<const uint32_t*>(kr + ks * 16),
                             *reinterpret_cast<const uint32_t*>(kr + ks * 16 + 8));
                c0 = ca[0]; c1 = ca[1];
            }
            if (fr < G) {
                float2 o2 = make_float2(c0 * sl2, c1 * sl2);
                if (!kWide || kh == 0) {                 // the mask rides on the first partial
                    bool ok0, ok1;
                    if (vm_on) {
                        const uchar2 v2 = *reinterpret_cast<const uchar2*>(vm + it * T + r0);
                        ok0 = v2.x != 0; ok1 = v2.y != 0;
                    } else {
                        const int k0 = lo + tile * T;
                        ok0 = k0 + r0 <= slot && (vb == nullptr || vb[k0 + r0] != 0);
                        ok1 = k0 + r0 + 1 <= slot && (vb == nullptr || vb[k0 + r0 + 1] != 0);
                    }
                    if (!ok0) o2.x = -INFINITY;
                    if (!ok1) o2.y = -INFINITY;
                }
                *reinterpret_cast<float2*>(sc + ((kWide ? kh : 0) * G + fr) * SROW + r0) = o2;
            }
        }
        __syncthreads();
        {   // head fr: the tile's scores -> running (max, sum) and the A fragments of P; O = corr * O + P V
            float sv[T / 4];
#pragma unroll
            for (int kk = 0; kk < T / 16; ++kk) {
                float2 lo2 = make_float2(-INFINITY, -INFINITY), hi2 = lo2;
                if (fr < G) {
                    lo2 = *reinterpret_cast<const float2*>(sc + fr * SROW + kk * 16 + fk);
                    hi2 = *reinterpret_cast<const float2*>(sc + fr * SROW + kk * 16 + 8 + fk);
                    if constexpr (kWide) {
                        const float2 l1 = *reinterpret_cast<const float2*>(sc + (G + fr) * SROW + kk * 16 + fk);
                        const float2 h1 = *reinterpret_cast<const float2*>(sc + (G + fr) * SROW + kk * 16 + 8 + fk);
                        lo2.x += l1.x; lo2.y += l1.y; hi2.x += h1.x; hi2.y += h1.y;
                    }
                }
                sv[4 * kk] = lo2.x; sv[4 * kk + 1] = lo2.y; sv[4 * kk + 2] = hi2.x; sv[4 * kk + 3] = hi2.y;
            }
            float mx = sv[0];
#pragma unroll
            for (int j = 1; j < T / 4; ++j) mx = fmaxf(mx, sv[j]);
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 1));
            mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, 2));
            const float mn = fmaxf(m_run, mx);
            const float mref = mn == -INFINITY ? 0.f : mn;   // nothing visible yet: every 2^(-inf - 0) below is 0
            const float corr = ex2(m_run - mref);
            float ls = 0.f;
#pragma unroll
            for (int j = 0; j < T / 4; ++j) {
                sv[j] = ex2(sv[j] - mref);
                ls += sv[j];
            }
            ls += __shfl_xor_sync(0xffffffffu, ls, 1);
            ls += __shfl_xor_sync(0xffffffffu, ls, 2);
            l_run = l_run * corr + ls;
            m_run = mn;
#pragma unroll
            for (int t = 0; t < NTW; ++t) { acc[t][0] *= corr; acc[t][1] *= corr; }
            const int tn = min(T, nkeys - tile * T);
            const __nv_bfloat16* vr = Vt + (lane & 15) * VROW;
#pragma unroll
            for (int kk = 0; kk < T / 16; ++kk) {
                if (kk * 16 < tn) {
                    const uint32_t a0 = pack_bf16(sv[4 * kk], sv[4 * kk + 1]);
                    const uint32_t a2 = pack_bf16(sv[4 * kk + 2], sv[4 * kk + 3]);
#pragma unroll
                    for (int t = 0; t < NTW; ++t) {
                        if (warp + kDecWarps * t < HD / 8) {
                            uint32_t b0, b1;
                            ldmatrix_x2_trans(b0, b1, vr + kk * 16 * VROW + vchunk(lane, warp + kDecWarps * t) * 8);
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
    if (warp == 0 && fr < G && fk == 0) { r_ml[fr * 2] = m_run; r_ml[fr * 2 + 1] = l_run; }
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
                const float f = ex2(ml[g * 2] - M);      // a partner without a visible key: 2^(-inf) = 0
                const float2 v = *reinterpret_cast<const float2*>(cluster.map_shared_rank(r_o, r) + idx * 2);
                a0 += f * v.x; a1 += f * v.y; Lsum += f * ml[g * 2 + 1];
            }
            const float inv = 1.f / Lsum;               // key `slot` is always visible: M is finite, Lsum > 0
            *reinterpret_cast<uint32_t*>(out + idx * 2) = pack_bf16(a0 * inv, a1 * inv);
        }
    }
    cluster.sync();                                      // partners keep their shared memory until rank 0 has read it
}

template <int HD, int G>
static int launch_decode_attn(const DecodeAttnArgs& a_in, int nsplit, cudaStream_t stream) {
    DecodeAttnArgs a = a_in;
    const int nloc = ((a.Smax + kDecTile - 1) / kDecTile + nsplit - 1) / nsplit;   // tiles of one CTA, at most
    a.vm_bytes = nloc * kDecTile <= 32 * 1024 ? nloc * kDecTile : 0;
    const size_t smem = static_cast<size_t>(DecSmem<HD, G>::fixed) + a.vm_bytes;
    {   // opt in once per device to the largest size a launch can ask for
        static bool done[64] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        if (dev < 0 || dev >= 64 || !done[dev]) {
            const cudaError_t e = cudaFuncSetAttribute(decode_attn_kernel<HD, G>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                                       DecSmem<HD, G>::fixed + 32 * 1024);
            if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(decode_attn): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
            if (dev >= 0 && dev < 64) done[dev] = true;
        }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3(static_cast<unsigned>(a.B * a.hkv * nsplit));
    cfg.blockDim = dim3(kDecThreads);
    cfg.dynamicSmemBytes = smem;
    cfg.stream = stream;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = static_cast<unsigned>(nsplit);
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl_enabled(kPdlDecodeAttn) ? 2 : 1;
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
