// Fused causal / sliding-window GQA attention forward for the Gemma3 decoder (head_dim 256; also instantiated
// for head_dim 128, the Qwen3 geometry), flash style:
//     O = softmax(mask(Q K^T * scale)) V,   LSE = log2-domain log-sum-exp of every score row
//     HF:models/gemma3/modeling_gemma3.py:278-300,370-384
// One pass over the key tiles of a query block with an online softmax: S_j = Q K_j^T lands in TMEM, the softmax
// warps turn it into un-normalised probabilities p = exp2((s - m) c) against the running row max m (the O
// accumulator in TMEM is rescaled only when some row's max grew by more than 2^8: a warp-uniform, rare event),
// O += P_j V_j, and the epilogue divides by the row sum.  Neither scores nor probabilities reach HBM: the
// backward (attn_gemma_bwd_sm100.cu) recomputes P from Q, K and the saved LSE.
//
// Work item = (batch, kv head, block of 128 stacked query rows): the G = hq/hkv query heads that share a kv
// head are stacked with their positions, row r <-> (position s0 + r / G, head kvh*G + r % G), by a 5-D TMA box
// over the rotated queries [B*S, hq*hd], so one K/V tile serves all G heads.  Only key tiles inside the causal
// (and sliding-window) band of the block are visited.  Roles: warps 0-3 softmax (thread = row), warp 4 MMA
// issuer, warp 5 TMA producer; every hand-off is an mbarrier (TMA tx, tcgen05.commit or 128 thread arrivals).
// TMEM: S0 [0,64) S1 [64,128) O [128,128+hd).
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"
#include "attn_gemma_maps.cuh"

namespace pt {

constexpr int kGthreads = 192;
constexpr int kGpBytes = kGqRows * kGkTile * 2;     // 16 KB
// head_dim 256: Q 64 KB + ring of 4 x 32 KB + P 16 KB;  head_dim 128 (Qwen3): 32 KB + 4 x 16 KB + 16 KB
constexpr int gemma_smem(int hd) { return kGqRows * hd * 2 + 4 * kGkTile * hd * 2 + kGpBytes + 256 + 1024; }

struct AttnGemmaDev {
    int B, S, hq, hkv, G, window;     // window 0 = causal only
    int pos_per_blk, nqb;             // 128 / G query positions per block; blocks per (batch, kv head)
    float scale_log2e;
    const unsigned long long* kvbits; // [B][ceil(S/64)] key-valid bit masks (bit i = key 64 w + i), or null
    int kvwords;
    float* lse;                       // [B*hq, S] log2-domain log-sum-exp (m c + log2 l; +inf: no visible key), or null
    __nv_bfloat16* O;                 // [B*S, hq*hd]
};

// keyvalid bytes [B][S] -> bit masks [B][ceil(S/64)]
__global__ void pack_keyvalid_kernel(const uint8_t* __restrict__ kv, unsigned long long* __restrict__ bits, int B,
                                     int S, int words) {
    const int idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx >= B * words) return;
    const int b = idx / words, w = idx - b * words;
    unsigned long long m = 0ull;
    for (int i = 0; i < 64; ++i) {
        const int k = w * 64 + i;
        if (k < S && kv[static_cast<long long>(b) * S + k]) m |= 1ull << i;
    }
    bits[idx] = m;
}

__device__ __forceinline__ void gemma_item(const AttnGemmaDev& p, int item, int& b, int& kvh, int& qb, int& jlo, int& nt) {
    const int per = p.B * p.hkv;
    const int rev = item / per;                   // heaviest (latest) query blocks first
    const int rem = item - rev * per;
    qb = p.nqb - 1 - rev;
    b = rem / p.hkv; kvh = rem - b * p.hkv;
    const int s0 = qb * p.pos_per_blk;
    const int s1 = min(p.S - 1, s0 + p.pos_per_blk - 1);
    jlo = p.window > 0 ? max(0, s0 - p.window + 1) / kGkTile : 0;
    nt = s1 / kGkTile - jlo + 1;
}

template <int HD>
__global__ void __launch_bounds__(kGthreads, 1)
attn_gemma_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmK,
                      const __grid_constant__ CUtensorMap tmV, const AttnGemmaDev p) {
    constexpr int kQB = kGqRows * HD * 2;          // Q block: HD/64 slabs (64 columns) of 128 rows x 128 B
    constexpr int kKVB = kGkTile * HD * 2;         // K or V tile: HD/64 slabs of 64 rows x 128 B
    constexpr int kSlabs = HD / 64;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sQ = smem;
    uint8_t* sKV = sQ + kQB;                // ring of 4 tiles: K_j, V_j, K_j+1, V_j+1
    uint8_t* sP = sKV + 4 * kKVB;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kGpBytes);
    uint64_t* q_full = bars;            // TMA -> MMA
    uint64_t* q_empty = bars + 1;       // MMA (commit after the item's last S) -> TMA
    uint64_t* ld_full = bars + 2;       // [4] TMA -> MMA (K or V tile in ring stage s)
    uint64_t* ld_empty = bars + 6;      // [4] commit after the MMAs that read stage s
    uint64_t* s_full = bars + 10;       // [2] commit after S_j -> softmax
    uint64_t* s_free = bars + 12;       // [2] softmax (128 arrivals) -> MMA
    uint64_t* p_full = bars + 14;       // softmax (128 arrivals) -> MMA
    uint64_t* pv_done = bars + 15;      // commit after PV_j -> softmax
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 16);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 128) {
        ptx::prefetch_tmap(&tmQ); ptx::prefetch_tmap(&tmK); ptx::prefetch_tmap(&tmV);
        ptx::mbar_init(q_full, 1); ptx::mbar_init(q_empty, 1);
        for (int s = 0; s < 4; ++s) { ptx::mbar_init(&ld_full[s], 1); ptx::mbar_init(&ld_empty[s], 1); }
        for (int s = 0; s < 2; ++s) { ptx::mbar_init(&s_full[s], 1); ptx::mbar_init(&s_free[s], 128); }
        ptx::mbar_init(p_full, 128);
        ptx::mbar_init(pv_done, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 4) {
        ptx::tmem_alloc(tmem_slot, 512);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tS[2] = {tmem_base, tmem_base + kGkTile};
    const uint32_t tO = tmem_base + 2 * kGkTile;
    const int items = p.B * p.hkv * p.nqb;

    if (warp == 5) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t n_item = 0, lc = 0;                  // lc = tiles loaded so far -> ring stage / parity
            for (int jj = 0; jj * static_cast<int>(gridDim.x) < items; ++jj, ++n_item) {
                const int item = snake_item(jj, blockIdx.x, gridDim.x);
                if (item >= items) break;          // (only in the last run)
                int b, kvh, qb, jlo, nt;
                gemma_item(p, item, b, kvh, qb, jlo, nt);
                ptx::mbar_wait(q_empty, (n_item & 1) ^ 1);           // previous item's S MMAs retired
                ptx::mbar_arrive_expect_tx(q_full, kQB);
#pragma unroll
                for (int c = 0; c < kSlabs; ++c)
                    ptx::tma_load_5d(&tmQ, q_full, sQ + c * (kGqRows * 128), c * 64, 0, qb * p.pos_per_blk, kvh, b);
                auto load_tile = [&](const CUtensorMap* tm, int j) {
                    const int st = lc & 3;
                    ptx::mbar_wait(&ld_empty[st], ((lc >> 2) & 1) ^ 1);
                    ptx::mbar_arrive_expect_tx(&ld_full[st], kKVB);
#pragma unroll
                    for (int c = 0; c < kSlabs; ++c)
                        ptx::tma_load_5d(tm, &ld_full[st], sKV + st * kKVB + c * (kGkTile * 128), c * 64,
                                         (jlo + j) * kGkTile, kvh, b, 0);
                    ++lc;
                };
                for (int j = 0; j < nt; ++j) { load_tile(&tmK, j); load_tile(&tmV, j); }
            }
        }
    } else if (warp == 4) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idescS = ptx::make_idesc_bf16(kGqRows, kGkTile, 0, 0);
            const uint32_t idescO = ptx::make_idesc_bf16(kGqRows, HD, 0, 1);      // B = V, MN-major
            uint32_t n_item = 0, base = 0 /* ring index of the item's first tile */, sc = 0, pc = 0;
            auto issue_S = [&](uint32_t li) {              // li = ring index of the K tile
                const int st = li & 3, sb = sc & 1;
                ptx::mbar_wait(&ld_full[st], (li >> 2) & 1);
                ptx::mbar_wait(&s_free[sb], ((sc >> 1) & 1) ^ 1);
                ptx::tc_fence_after();
                const uint32_t qa = ptx::smem_u32(sQ), ka = ptx::smem_u32(sKV + st * kKVB);
#pragma unroll
                for (int k = 0; k < HD / 16; ++k) {
                    const uint64_t qdesc = ptx::make_smem_desc_sw128(qa + (k >> 2) * (kGqRows * 128) + (k & 3) * 32, 0, 1024);
                    const uint64_t kdesc = ptx::make_smem_desc_sw128(ka + (k >> 2) * (kGkTile * 128) + (k & 3) * 32, 0, 1024);
                    ptx::mma_bf16_ss(tS[sb], qdesc, kdesc, idescS, k != 0);
                }
                ptx::mma_commit(&s_full[sb]);
                ptx::mma_commit(&ld_empty[st]);
                ++sc;
            };
            for (int jj = 0; jj * static_cast<int>(gridDim.x) < items; ++jj, ++n_item) {
                const int item = snake_item(jj, blockIdx.x, gridDim.x);
                if (item >= items) break;          // (only in the last run)
                int b, kvh, qb, jlo, nt;
                gemma_item(p, item, b, kvh, qb, jlo, nt);
                ptx::mbar_wait(q_full, n_item & 1);
                issue_S(base);                                            // K_j at base + 2j, V_j at base + 2j + 1
                if (nt == 1) ptx::mma_commit(q_empty);
                for (int j = 0; j < nt; ++j) {
                    if (j + 1 < nt) {
                        issue_S(base + 2 * (j + 1));
                        if (j + 2 == nt) ptx::mma_commit(q_empty);        // last S of the item issued: Q is free
                    }
                    const uint32_t lv = base + 2 * j + 1;
                    const int sv = lv & 3;
                    ptx::mbar_wait(p_full, pc & 1);
                    ptx::mbar_wait(&ld_full[sv], (lv >> 2) & 1);
                    ptx::tc_fence_after();
                    const uint32_t pa = ptx::smem_u32(sP), va = ptx::smem_u32(sKV + sv * kKVB);
#pragma unroll
                    for (int k = 0; k < kGkTile / 16; ++k) {
                        const uint64_t pdesc = ptx::make_smem_desc_sw128(pa + k * 32, 0, 1024);
                        const uint64_t vdesc = ptx::make_smem_desc_sw128(va + k * 16 * 128, kGkTile * 128, 1024);
                        ptx::mma_bf16_ss(tO, pdesc, vdesc, idescO, (j | k) != 0);
                    }
                    ptx::mma_commit(pv_done);
                    ptx::mma_commit(&ld_empty[sv]);
                    ++pc;
                }
                base += 2 * nt;
            }
        }
    } else {
        // ===================== softmax warps: thread = stacked query row =====================
        const int row = threadIdx.x;                               // TMEM lane
        const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
        const uint32_t p_row = ptx::smem_u32(sP) + row * 128;
        const float c2 = p.scale_log2e;
        uint32_t sc = 0, pc = 0;
        for (int jj = 0; jj * static_cast<int>(gridDim.x) < items; ++jj) {
            const int item = snake_item(jj, blockIdx.x, gridDim.x);
            if (item >= items) break;              // (only in the last run)
            int b, kvh, qb, jlo, nt;
            gemma_item(p, item, b, kvh, qb, jlo, nt);
            const int s = qb * p.pos_per_blk + row / p.G;           // query position of this row
            const int head = kvh * p.G + row % p.G;
            const bool live = s < p.S;
            const int k_hi = live ? s : -1;                          // visible keys: k_lo <= k <= k_hi
            const int k_lo = p.window > 0 ? max(0, s - p.window + 1) : 0;
            const unsigned long long* kvw = p.kvbits ? p.kvbits + static_cast<long long>(b) * p.kvwords : nullptr;

            float m_run = -INFINITY, l_run = 0.f;
            for (int j = 0; j < nt; ++j) {
                // the key-padding word is fetched BEFORE the wait for the S tile (an L2-latency load right in
                // front of its first use was the top stall of the first version of this kernel)
                const unsigned long long mask =
                    gemma_tile_mask(kvw ? __ldg(kvw + jlo + j) : ~0ull, (jlo + j) * kGkTile, k_lo, k_hi);
                uint32_t v[2][32];
                {
                    const int sb = sc & 1;
                    ptx::mbar_wait(&s_full[sb], (sc >> 1) & 1);
                    ptx::tc_fence_after();
                    ptx::tmem_ld_32x32(tS[sb] + lane_off, v[0]);
                    ptx::tmem_ld_32x32(tS[sb] + lane_off + 32, v[1]);
                    ptx::tmem_ld_wait();
                    ptx::tc_fence_before();
                    ptx::mbar_arrive(&s_free[sb]);
                    ++sc;
                }
                // tiles strictly inside the band are fully visible for every row of the warp: no per-element select
                if (!__all_sync(0xffffffffu, mask == ~0ull)) {
#pragma unroll
                    for (int c = 0; c < 2; ++c)
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (!((mask >> (c * 32 + i)) & 1ull)) v[c][i] = __float_as_uint(-INFINITY);
                }
                float mxs[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) mxs[i] = -INFINITY;
#pragma unroll
                for (int c = 0; c < 2; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i) mxs[i & 7] = fmaxf(mxs[i & 7], __uint_as_float(v[c][i]));
                const float mx = fmaxf(fmaxf(fmaxf(mxs[0], mxs[1]), fmaxf(mxs[2], mxs[3])),
                                       fmaxf(fmaxf(mxs[4], mxs[5]), fmaxf(mxs[6], mxs[7])));
                bool pv_waited = false;
                if (j == 0) {
                    m_run = mx;                                       // PV_0 overwrites O: nothing to rescale
                } else {
                    const float m_new = fmaxf(m_run, mx);
                    // lazy rescale: only when some row's max grew by more than 2^8 (m_run = -inf: first visible key)
                    const bool grow = m_new > m_run && (m_new - m_run) * c2 > 8.f;
                    if (__any_sync(0xffffffffu, grow)) {
                        const float alpha = m_new == -INFINITY ? 1.f : ptx::exp2_fast((m_run - m_new) * c2);
                        ptx::mbar_wait(pv_done, pc & 1); ++pc;        // O += P_{j-1} V_{j-1} has retired
                        pv_waited = true;
                        ptx::tc_fence_after();
#pragma unroll 1
                        for (int c = 0; c < HD / 32; ++c) {
                            uint32_t o[32];
                            ptx::tmem_ld_32x32(tO + lane_off + c * 32, o);
                            ptx::tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                            ptx::tmem_st_32x32(tO + lane_off + c * 32, o);
                        }
                        ptx::tmem_st_wait();
                        l_run *= alpha;
                        m_run = m_new;
                    }
                }
                // p = exp2(s c - m c): at most 2^8 under the lazy max; masked scores are -inf -> exactly 0
                const float mb = m_run == -INFINITY ? 0.f : m_run * c2;
                float ls[4] = {0.f, 0.f, 0.f, 0.f};
                uint32_t pk[32];
#pragma unroll
                for (int c = 0; c < 2; ++c)
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float e0 = ptx::exp2_fast(fmaf(__uint_as_float(v[c][2 * i]), c2, -mb));
                        const float e1 = ptx::exp2_fast(fmaf(__uint_as_float(v[c][2 * i + 1]), c2, -mb));
                        ls[i & 3] += e0 + e1;
                        pk[c * 16 + i] = pack_bf16(e0, e1);
                    }
                l_run += (ls[0] + ls[1]) + (ls[2] + ls[3]);
                // the single P buffer is free once the previous tile's PV has retired
                if (j > 0 && !pv_waited) { ptx::mbar_wait(pv_done, pc & 1); ++pc; }
#pragma unroll
                for (int q = 0; q < 8; ++q)
                    ptx::sts128(p_row + ((q ^ (row & 7)) << 4), pk[4 * q], pk[4 * q + 1], pk[4 * q + 2], pk[4 * q + 3]);
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before();
                ptx::mbar_arrive(p_full);
            }
            // ---- epilogue: O / l -> bf16 -> global, LSE
            ptx::mbar_wait(pv_done, pc & 1); ++pc;
            ptx::tc_fence_after();
            const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
            if (live && p.lse)
                p.lse[(static_cast<long long>(b) * p.hq + head) * p.S + s] =
                    l_run > 0.f ? fmaf(m_run, c2, log2f(l_run)) : INFINITY;
            __nv_bfloat16* op = p.O + (static_cast<long long>(b) * p.S + (live ? s : 0)) * (p.hq * HD) + head * HD;
#pragma unroll 1
            for (int c = 0; c < HD / 32; ++c) {
                uint32_t o[32];
                ptx::tmem_ld_32x32(tO + lane_off + c * 32, o);
                ptx::tmem_ld_wait();
                if (live) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        uint4 u;
                        u.x = pack_bf16(__uint_as_float(o[8 * q]) * inv, __uint_as_float(o[8 * q + 1]) * inv);
                        u.y = pack_bf16(__uint_as_float(o[8 * q + 2]) * inv, __uint_as_float(o[8 * q + 3]) * inv);
                        u.z = pack_bf16(__uint_as_float(o[8 * q + 4]) * inv, __uint_as_float(o[8 * q + 5]) * inv);
                        u.w = pack_bf16(__uint_as_float(o[8 * q + 6]) * inv, __uint_as_float(o[8 * q + 7]) * inv);
                        *reinterpret_cast<uint4*>(op + c * 32 + q * 8) = u;
                    }
                }
            }
            ptx::tc_fence_before();
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 4) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, 512);
    }
}

}  // namespace pt

using namespace pt;

extern "C" int pt_pack_keyvalid(const uint8_t* keyvalid, void* kvbits, int B, int S, void* stream) {
    PT_REQUIRE(keyvalid && kvbits && B > 0 && S > 0, "pt_pack_keyvalid: bad arguments");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const int words = (S + 63) / 64, n = B * words;
    pack_keyvalid_kernel<<<(n + 127) / 128, 128, 0, static_cast<cudaStream_t>(stream)>>>(
        keyvalid, static_cast<unsigned long long*>(kvbits), B, S, words);
    return check_launch("pack_keyvalid");
}

extern "C" int pt_attn_gemma_fwd(const void* q_rot, const void* k_rot, const void* qkv, const void* kvbits,
                                 float* lse, void* O_bf16, int B, int S, int hq, int hkv, int hd,
                                 int window, float scale, void* stream) {
    PT_REQUIRE(q_rot && k_rot && qkv && O_bf16, "pt_attn_gemma_fwd: null pointer");
    PT_REQUIRE(hd == 256 || hd == 128, "pt_attn_gemma_fwd: head_dim must be 256 or 128 (got %d)", hd);
    PT_REQUIRE(B > 0 && S > 0 && hq > 0 && hkv > 0 && hq % hkv == 0 && window >= 0, "pt_attn_gemma_fwd: bad shape");
    const int G = hq / hkv;
    PT_REQUIRE(G <= 128 && (G & (G - 1)) == 0, "pt_attn_gemma_fwd: hq / hkv must be a power of two <= 128 (got %d)", G);
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int QKV = (hq + 2 * hkv) * hd;
    const int pos = kGqRows / G;
    CUtensorMap tq, tk, tv;
    int rc = gemma_stacked_map(&tq, q_rot, B, S, hq, hkv, hd, "attn_gemma Q");
    if (rc != PT_OK) return rc;
    rc = gemma_kv_map(&tk, k_rot, static_cast<long long>(hkv) * hd, B, S, hkv, hd, "attn_gemma K");
    if (rc != PT_OK) return rc;
    // values in place inside the fused qkv activation [B*S, QKV] at column (hq + hkv) * hd
    rc = gemma_kv_map(&tv, static_cast<const __nv_bfloat16*>(qkv) + static_cast<long long>(hq + hkv) * hd, QKV, B, S,
                      hkv, hd, "attn_gemma V");
    if (rc != PT_OK) return rc;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(attn_gemma_fwd_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             gemma_smem(256));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(attn_gemma_fwd_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     gemma_smem(128));
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(attn_gemma): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
        attr_done = true;
    }
    AttnGemmaDev p;
    p.B = B; p.S = S; p.hq = hq; p.hkv = hkv; p.G = G; p.window = window;
    p.pos_per_blk = pos;
    p.nqb = (S + pos - 1) / pos;
    p.scale_log2e = scale * 1.4426950408889634f;
    p.kvwords = (S + 63) / 64;
    p.kvbits = static_cast<const unsigned long long*>(kvbits);
    p.lse = lse;
    p.O = static_cast<__nv_bfloat16*>(O_bf16);
    const long long items = static_cast<long long>(B) * hkv * p.nqb;
    PT_REQUIRE(items < (1ll << 31), "pt_attn_gemma_fwd: too many work items");
    const int grid = items < num_sms() ? static_cast<int>(items) : num_sms();
    if (hd == 256)
        attn_gemma_fwd_kernel<256><<<grid, kGthreads, gemma_smem(256), st>>>(tq, tk, tv, p);
    else
        attn_gemma_fwd_kernel<128><<<grid, kGthreads, gemma_smem(128), st>>>(tq, tk, tv, p);
    return check_launch("attn_gemma_fwd_kernel");
}
