// Fused causal / sliding-window GQA attention BACKWARD for the Gemma3 decoder (head_dim 256; also instantiated
// for head_dim 128, the Qwen3 geometry), flash style — autograd through HF:models/gemma3/modeling_gemma3.py:278-300,
// 370-384 as invoked by ref:Stage1/projector_trainer.py:236-237:
//     P  = exp2(Q K^T c - LSE)            (recomputed; the forward saves only the log-sum-exp of every row)
//     dP = dO V^T,   dS = scale * P * (dP - delta),   delta = rowsum(dO * O)
//     dV = P^T dO,   dK = dS^T Q          (summed over the G = hq/hkv query heads that share a kv head)
// and dS (bf16, band only) goes to HBM for the one contraction that is stationary in the OTHER operand,
// dQ = dS K (a band-limited GEMM, engine.py).  Neither P nor dP ever reaches HBM.
//
// Key/value stationary: work item = (batch, kv head, tile of 64 keys); its K and V tiles stay in shared memory
// while the item walks the blocks of 128 stacked query rows (row r <-> position s0 + r / G, head kvh*G + r % G:
// the same 5-D TMA box as the forward, so the sum over the G heads is part of the contraction) that can see
// those keys under the causal / window band.  Per block:
//     S  = Q K^T, dP = dO V^T            tcgen05.mma 128 x 64 x hd, accumulators in TMEM
//     softmax warps (thread = query row): P, dS -> bf16 -> swizzled smem (and dS -> global)
//     dV^T += dO^T P,  dK^T += Q^T dS    tcgen05.mma (hd/128) x [128 x 64 x 128]: the accumulators are held
//                                         TRANSPOSED, [head_dim, key], so that M is 128 with 64-key tiles and
//                                         S + dP + dV^T + dK^T fit the 512 TMEM columns at head_dim 256
// Q and dO blocks are read through the same smem image twice: K-major as the A operand of S / dP, and MN-major
// (64-column slabs) as the A operand of dK^T / dV^T; P and dS are MN-major B operands.
// With head_dim 256 the resident K / V tiles (64 KB), one Q and one dO block (128 KB) and P / dS (32 KB) fill
// the SM's shared memory, so there is one block in flight; within it the loads, the two early and the two late
// MMA groups and the two softmax phases overlap pairwise (dO lands first -> dP; Q -> S; P -> dV^T runs while dS
// is computed; dO of the next block loads while dK^T runs).
// Roles: warps 0-7 softmax (two threads per TMEM lane = stacked query row, 32 key columns each; at item end
// lane = head_dim row of the transposed accumulators, which go back to [key][head_dim] rows through smem),
// warp 8 MMA issuer, warp 9 TMA producer.
// TMEM columns: S [0,64) dP [64,128) dV^T [128,128+64 MB) dK^T [..+64 MB), MB = hd / 128.
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"
#include "attn_gemma_maps.cuh"

namespace pt {

constexpr int kBwThreads = 320;         // 8 softmax warps + MMA issuer + TMA producer
constexpr int kBwPBytes = kGqRows * kGkTile * 2;            // 16 KB: P or dS block, one 128-byte row per query row
// K + V tiles, Q + dO blocks, P + dS, barriers, alignment slack
constexpr int gemma_bwd_smem(int hd) { return 2 * kGkTile * hd * 2 + 2 * kGqRows * hd * 2 + 2 * kBwPBytes + 256 + 1024; }

struct AttnGemmaBwdDev {
    int B, S, hq, hkv, G, window;
    int pos_per_blk, nqb, nkt;        // 128 / G positions per query block; query blocks; key tiles
    int lds;                          // row stride of dS (elements)
    float scale, scale_log2e;
    const unsigned long long* kvbits; // [B][ceil(S/64)] key-valid bit masks or null
    int kvwords;
    const float* lse;                 // [B*hq, S] from the forward
    const float* delta;               // [B*hq, S] rowsum(dO * O)
    __nv_bfloat16* dS;                // [B*hq, S, lds]
    __nv_bfloat16* dK;                // element (b*S + k, kvh*hd + d) at dK[(b*S + k) * ld_dk + kvh*hd + d]
    __nv_bfloat16* dV;
    long long ld_dk, ld_dv;
    long long* trace;                 // tools/attn_bwd_trace.py: SM-clock timestamps of CTA 0's hand-offs, or null
    int trace_iters;
};

// timeline probe (debug hook, null in the product path): event e of block `it` of CTA 0
__device__ __forceinline__ void bwd_trace(const AttnGemmaBwdDev& p, uint32_t it, int e) {
    if (p.trace != nullptr && blockIdx.x == 0 && it < static_cast<uint32_t>(p.trace_iters))
        p.trace[it * 16 + e] = clock64();
}

// item -> (batch, kv head, key tile, first query block, number of query blocks).  Items are ordered by key tile:
// early keys are seen by the most queries, so the heaviest items start first.
__device__ __forceinline__ void gemma_bwd_item(const AttnGemmaBwdDev& p, int item, int& b, int& kvh, int& kt, int& qb0,
                                               int& nqb) {
    const int per = p.B * p.hkv;
    kt = item / per;
    const int rem = item - kt * per;
    b = rem / p.hkv; kvh = rem - b * p.hkv;
    const int k0 = kt * kGkTile;
    const int k1 = min(p.S - 1, k0 + kGkTile - 1);
    qb0 = k0 / p.pos_per_blk;                                  // first block holding a position >= k0
    int qb1 = p.nqb - 1;
    if (p.window > 0) qb1 = min(qb1, (k1 + p.window - 1) / p.pos_per_blk);   // last position that sees key k1
    nqb = qb1 - qb0 + 1;
}

template <int HD>
__global__ void __launch_bounds__(kBwThreads, 1)
attn_gemma_bwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                      const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                      const AttnGemmaBwdDev p) {
    constexpr int kQB = kGqRows * HD * 2;          // Q / dO block: HD/64 slabs (64 columns) of 128 rows x 128 B
    constexpr int kKVB = kGkTile * HD * 2;         // K / V tile: HD/64 slabs of 64 rows x 128 B
    constexpr int kSlabs = HD / 64;
    constexpr int kSlabQ = kGqRows * 128;          // bytes of one 64-column slab of a Q / dO block
    constexpr int MB = HD / 128;                   // 128-row M blocks of the transposed accumulators
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sK = smem;
    uint8_t* sV = sK + kKVB;
    uint8_t* sQ = sV + kKVB;
    uint8_t* sdO = sQ + kQB;
    uint8_t* sP = sdO + kQB;
    uint8_t* sdS = sP + kBwPBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sdS + kBwPBytes);
    uint64_t* kv_full = bars;           // TMA -> MMA, once per item
    uint64_t* q_full = bars + 1;        // TMA -> MMA, once per block
    uint64_t* do_full = bars + 2;
    uint64_t* q_empty = bars + 3;       // commit after dK^T (last reader of Q; K / V / P / dS are free then too)
    uint64_t* do_empty = bars + 4;      // commit after dV^T (last reader of dO)
    uint64_t* s_full = bars + 5;        // commit after S -> softmax
    uint64_t* dp_full = bars + 6;       // commit after dP -> softmax
    uint64_t* p_full = bars + 7;        // softmax (256 arrivals) -> MMA: P in smem, S read
    uint64_t* ds_full = bars + 8;       // softmax (256 arrivals) -> MMA: dS in smem, dP read
    uint64_t* acc_full = bars + 9;      // commit after the item's last dK^T -> epilogue
    uint64_t* acc_free = bars + 10;     // epilogue (256 arrivals) -> MMA: accumulators read
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 11);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 256) {
        ptx::prefetch_tmap(&tmQ); ptx::prefetch_tmap(&tmdO); ptx::prefetch_tmap(&tmK); ptx::prefetch_tmap(&tmV);
        ptx::mbar_init(kv_full, 1); ptx::mbar_init(q_full, 1); ptx::mbar_init(do_full, 1);
        ptx::mbar_init(q_empty, 1); ptx::mbar_init(do_empty, 1);
        ptx::mbar_init(s_full, 1); ptx::mbar_init(dp_full, 1);
        ptx::mbar_init(p_full, 256); ptx::mbar_init(ds_full, 256);
        ptx::mbar_init(acc_full, 1); ptx::mbar_init(acc_free, 256);
        ptx::fence_mbar_init();
    }
    constexpr uint32_t kTmemCols = HD == 256 ? 512 : 256;
    if (warp == 8) {
        ptx::tmem_alloc(tmem_slot, kTmemCols);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tS = tmem_base, tdP = tmem_base + kGkTile;
    const uint32_t tdV = tmem_base + 2 * kGkTile, tdK = tdV + MB * kGkTile;
    const int items = p.B * p.hkv * p.nkt;

    if (warp == 9) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            uint32_t n_item = 0, it = 0;                    // it = blocks loaded so far (all items)
            for (int jj = 0; jj * static_cast<int>(gridDim.x) < items; ++jj, ++n_item) {
                const int item = snake_item(jj, blockIdx.x, gridDim.x);
                if (item >= items) break;          // (only in the last run)
                int b, kvh, kt, qb0, nqb;
                gemma_bwd_item(p, item, b, kvh, kt, qb0, nqb);
                for (int i = 0; i < nqb; ++i, ++it) {
                    const int s0 = (qb0 + i) * p.pos_per_blk;
                    // dO first: its buffer is released one MMA group earlier (after dV^T) than Q's (after dK^T)
                    ptx::mbar_wait(do_empty, (it & 1) ^ 1);
                    bwd_trace(p, it, 0);
                    if (i == 0) {
                        // every MMA of the previous item has retired (q_empty below was its last commit, and the
                        // wait on it precedes this point for it > 0 ... see the q_empty wait of the last block)
                        ptx::mbar_wait(q_empty, (it & 1) ^ 1);
                        ptx::mbar_arrive_expect_tx(kv_full, 2 * kKVB);
#pragma unroll
                        for (int c = 0; c < kSlabs; ++c) {
                            ptx::tma_load_5d(&tmV, kv_full, sV + c * (kGkTile * 128), c * 64, kt * kGkTile, kvh, b, 0);
                            ptx::tma_load_5d(&tmK, kv_full, sK + c * (kGkTile * 128), c * 64, kt * kGkTile, kvh, b, 0);
                        }
                    }
                    ptx::mbar_arrive_expect_tx(do_full, kQB);
#pragma unroll
                    for (int c = 0; c < kSlabs; ++c)
                        ptx::tma_load_5d(&tmdO, do_full, sdO + c * kSlabQ, c * 64, 0, s0, kvh, b);
                    if (i != 0) ptx::mbar_wait(q_empty, (it & 1) ^ 1);
                    bwd_trace(p, it, 1);
                    ptx::mbar_arrive_expect_tx(q_full, kQB);
#pragma unroll
                    for (int c = 0; c < kSlabs; ++c)
                        ptx::tma_load_5d(&tmQ, q_full, sQ + c * kSlabQ, c * 64, 0, s0, kvh, b);
                }
            }
        }
    } else if (warp == 8) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idescS = ptx::make_idesc_bf16(kGqRows, kGkTile, 0, 0);     // A = Q / dO, B = K / V: K-major
            const uint32_t idescT = ptx::make_idesc_bf16(128, kGkTile, 1, 1);         // A = Q^T / dO^T, B = dS / P: MN-major
            const uint32_t qa = ptx::smem_u32(sQ), doa = ptx::smem_u32(sdO);
            const uint32_t ka = ptx::smem_u32(sK), va = ptx::smem_u32(sV);
            const uint32_t pa = ptx::smem_u32(sP), dsa = ptx::smem_u32(sdS);
            uint32_t n_item = 0, it = 0;
            // D[128 x 64] (+)= A[128 x hd] B[64 x hd]^T, both K-major in 64-column slabs
            auto issue_scores = [&](uint32_t d, uint32_t a, uint32_t bb) {
#pragma unroll
                for (int k = 0; k < HD / 16; ++k) {
                    const uint64_t adesc = ptx::make_smem_desc_sw128(a + (k >> 2) * kSlabQ + (k & 3) * 32, 0, 1024);
                    const uint64_t bdesc = ptx::make_smem_desc_sw128(bb + (k >> 2) * (kGkTile * 128) + (k & 3) * 32, 0, 1024);
                    ptx::mma_bf16_ss(d, adesc, bdesc, idescS, k != 0);
                }
            };
            // D^T[hd x 64] (+)= A[128 rows x hd]^T B[128 rows x 64]: contraction over the 128 query rows;
            // A read MN-major (slab of 64 head-dim columns = 64 M elements, slab stride kSlabQ), B MN-major (one slab)
            auto issue_grad = [&](uint32_t d, uint32_t a, uint32_t bb, bool first) {
#pragma unroll
                for (int h = 0; h < MB; ++h)
#pragma unroll
                    for (int k = 0; k < kGqRows / 16; ++k) {
                        const uint64_t adesc = ptx::make_smem_desc_sw128(a + 2 * h * kSlabQ + k * 16 * 128, kSlabQ, 1024);
                        const uint64_t bdesc = ptx::make_smem_desc_sw128(bb + k * 16 * 128, 0, 1024);
                        ptx::mma_bf16_ss(d + h * kGkTile, adesc, bdesc, idescT, !(first && k == 0));
                    }
            };
            for (int jj = 0; jj * static_cast<int>(gridDim.x) < items; ++jj, ++n_item) {
                const int item = snake_item(jj, blockIdx.x, gridDim.x);
                if (item >= items) break;          // (only in the last run)
                int b, kvh, kt, qb0, nqb;
                gemma_bwd_item(p, item, b, kvh, kt, qb0, nqb);
                ptx::mbar_wait(kv_full, n_item & 1);
                for (int i = 0; i < nqb; ++i, ++it) {
                    const uint32_t ph = it & 1;
                    ptx::mbar_wait(do_full, ph);
                    bwd_trace(p, it, 2);
                    ptx::tc_fence_after();
                    issue_scores(tdP, doa, va);
                    ptx::mma_commit(dp_full);
                    ptx::mbar_wait(q_full, ph);
                    bwd_trace(p, it, 3);
                    ptx::tc_fence_after();
                    issue_scores(tS, qa, ka);
                    ptx::mma_commit(s_full);
                    if (i == 0) {                       // previous item's accumulators have been read
                        ptx::mbar_wait(acc_free, (n_item & 1) ^ 1);
                        ptx::tc_fence_after();
                    }
                    ptx::mbar_wait(p_full, ph);
                    bwd_trace(p, it, 4);
                    ptx::tc_fence_after();
                    issue_grad(tdV, doa, pa, i == 0);
                    ptx::mma_commit(do_empty);
                    ptx::mbar_wait(ds_full, ph);
                    bwd_trace(p, it, 5);
                    ptx::tc_fence_after();
                    issue_grad(tdK, qa, dsa, i == 0);
                    ptx::mma_commit(q_empty);
                }
                ptx::mma_commit(acc_full);
            }
        }
    } else {
        // ===================== softmax warps: two threads per stacked query row =====================
        // warps 0-3 own key columns [0,32) of rows 32 (warp & 3) .. +31, warps 4-7 columns [32,64) (a warp may touch
        // TMEM lanes 32 (warp % 4) .. +31 only).  The backward needs no cross-thread row reduction: LSE and delta
        // are per-row inputs.
        const int row = (warp & 3) * 32 + lane;                    // TMEM lane
        const int half = warp >> 2;
        const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const uint32_t p_row = ptx::smem_u32(sP) + row * 128;
        const uint32_t ds_row = ptx::smem_u32(sdS) + row * 128;
        const float c2 = p.scale_log2e;
        const bool tr0 = threadIdx.x == 0;
        uint32_t n_item = 0, it = 0;
        for (int jj = 0; jj * static_cast<int>(gridDim.x) < items; ++jj, ++n_item) {
                const int item = snake_item(jj, blockIdx.x, gridDim.x);
                if (item >= items) break;          // (only in the last run)
            int b, kvh, kt, qb0, nqb;
            gemma_bwd_item(p, item, b, kvh, kt, qb0, nqb);
            const int k0 = kt * kGkTile;
            const unsigned long long kw = p.kvbits ? __ldg(p.kvbits + static_cast<long long>(b) * p.kvwords + kt) : ~0ull;
            const int head = kvh * p.G + row % p.G;
            for (int i = 0; i < nqb; ++i, ++it) {
                const uint32_t ph = it & 1;
                const int s = (qb0 + i) * p.pos_per_blk + row / p.G;         // query position of this row
                const bool live = s < p.S;
                const long long ridx = (static_cast<long long>(b) * p.hq + head) * p.S + (live ? s : 0);
                // row statistics: in flight while the block's MMAs run
                const float lse = live ? __ldg(p.lse + ridx) : INFINITY;
                const float dl = live ? __ldg(p.delta + ridx) : 0.f;
                const uint32_t mask = static_cast<uint32_t>(
                    gemma_tile_mask(kw, k0, p.window > 0 ? max(0, s - p.window + 1) : 0, live ? s : -1) >> (half * 32));
                const bool all_vis = __all_sync(0xffffffffu, mask == 0xffffffffu);

                // ---- P = exp2(S c - LSE) -> bf16 -> smem (B operand of dV^T += dO^T P)
                float pf[32];
                {
                    uint32_t v[32];
                    ptx::mbar_wait(s_full, ph);
                    if (tr0) bwd_trace(p, it, 6);
                    ptx::tc_fence_after();
                    ptx::tmem_ld_32x32(tS + lane_off + half * 32, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 32; ++j) pf[j] = ptx::exp2_fast(fmaf(__uint_as_float(v[j]), c2, -lse));
                }
                if (!all_vis) {
#pragma unroll
                    for (int j = 0; j < 32; ++j)
                        if (!((mask >> j) & 1u)) pf[j] = 0.f;
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    ptx::sts128(p_row + (((half * 4 + q) ^ (row & 7)) << 4), pack_bf16(pf[8 * q], pf[8 * q + 1]),
                                pack_bf16(pf[8 * q + 2], pf[8 * q + 3]), pack_bf16(pf[8 * q + 4], pf[8 * q + 5]),
                                pack_bf16(pf[8 * q + 6], pf[8 * q + 7]));
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before();
                ptx::mbar_arrive(p_full);
                if (tr0) bwd_trace(p, it, 7);

                // ---- dS = scale P (dP - delta) -> bf16 -> smem (B operand of dK^T += Q^T dS) and global
                uint32_t dk[16];
                ptx::mbar_wait(dp_full, ph);
                if (tr0) bwd_trace(p, it, 8);
                ptx::tc_fence_after();
                {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32(tdP + lane_off + half * 32, v);
                    ptx::tmem_ld_wait();
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float d0 = p.scale * pf[2 * j] * (__uint_as_float(v[2 * j]) - dl);
                        const float d1 = p.scale * pf[2 * j + 1] * (__uint_as_float(v[2 * j + 1]) - dl);
                        dk[j] = pack_bf16(d0, d1);
                    }
                }
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    ptx::sts128(ds_row + (((half * 4 + q) ^ (row & 7)) << 4), dk[4 * q], dk[4 * q + 1], dk[4 * q + 2],
                                dk[4 * q + 3]);
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before();
                ptx::mbar_arrive(ds_full);
                if (tr0) bwd_trace(p, it, 9);
                if (live) {
                    __nv_bfloat16* g = p.dS + ridx * p.lds + k0 + half * 32;
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        if (k0 + half * 32 + q * 8 < p.lds)
                            *reinterpret_cast<uint4*>(g + q * 8) = make_uint4(dk[4 * q], dk[4 * q + 1], dk[4 * q + 2], dk[4 * q + 3]);
                }
            }
            if (tr0) bwd_trace(p, it - 1, 10);
            // ---- epilogue.  The accumulators are transposed: TMEM lane = head-dim row (h * 128 + row), column =
            // key; this thread owns keys half * 32 .. +31.  They are turned back into [key][head_dim] rows through
            // shared memory (the P and dS blocks: 32 KB, idle until this warp group writes the next item's first P)
            // and leave as 16-byte coalesced stores.
            ptx::mbar_wait(acc_full, n_item & 1);
            ptx::tc_fence_after();
            const int nk = min(kGkTile, p.S - k0);
            uint8_t* stage = sP;                                    // [64 keys][HD] bf16
#pragma unroll 1
            for (int t = 0; t < 2; ++t) {
                const uint32_t tacc = t == 0 ? tdV : tdK;
#pragma unroll 1
                for (int h = 0; h < MB; ++h) {
                    uint32_t v[32];
                    ptx::tmem_ld_32x32(tacc + h * kGkTile + lane_off + half * 32, v);
                    ptx::tmem_ld_wait();
                    __nv_bfloat16* st = reinterpret_cast<__nv_bfloat16*>(stage) + (half * 32) * HD + h * 128 + row;
#pragma unroll
                    for (int j = 0; j < 32; ++j) st[j * HD] = __float2bfloat16(__uint_as_float(v[j]));
                }
                if (t == 1) {                                       // both accumulators are in registers / smem now
                    ptx::tc_fence_before();
                    ptx::mbar_arrive(acc_free);
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");
                __nv_bfloat16* outp = t == 0 ? p.dV : p.dK;
                const long long ldo = t == 0 ? p.ld_dv : p.ld_dk;
                __nv_bfloat16* o = outp + (static_cast<long long>(b) * p.S + k0) * ldo + kvh * HD;
                constexpr int kChunks = HD / 8;                     // 16-byte chunks per key row
                for (int idx = threadIdx.x; idx < nk * kChunks; idx += 256) {
                    const int key = idx / kChunks, c = idx - key * kChunks;
                    *reinterpret_cast<uint4*>(o + key * ldo + c * 8) =
                        *reinterpret_cast<const uint4*>(stage + (key * HD + c * 8) * 2);
                }
                asm volatile("bar.sync 1, 256;" ::: "memory");
            }
            if (tr0) bwd_trace(p, it - 1, 11);
        }
    }
    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 8) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, kTmemCols);
    }
}

}  // namespace pt

using namespace pt;

static long long* g_bwd_trace = nullptr;
static int g_bwd_trace_iters = 0;
// debug hook for tools/attn_bwd_trace.py (not part of the product path): device buffer of 16 * iters int64
extern "C" int pt_attn_gemma_bwd_set_trace(void* buf, int iters) {
    g_bwd_trace = static_cast<long long*>(buf);
    g_bwd_trace_iters = buf ? iters : 0;
    return PT_OK;
}

extern "C" int pt_attn_gemma_bwd(const void* q_rot, const void* k_rot, const void* qkv, const void* dO_bf16,
                                 const void* kvbits, const float* lse, const float* delta, void* dS_bf16,
                                 void* dK_bf16, long long ld_dk, void* dV_bf16, long long ld_dv, int B, int S, int hq,
                                 int hkv, int hd, int window, float scale, int lds, void* stream) {
    PT_REQUIRE(q_rot && k_rot && qkv && dO_bf16 && lse && delta && dS_bf16 && dK_bf16 && dV_bf16,
               "pt_attn_gemma_bwd: null pointer");
    PT_REQUIRE(hd == 256 || hd == 128, "pt_attn_gemma_bwd: head_dim must be 256 or 128 (got %d)", hd);
    PT_REQUIRE(B > 0 && S > 0 && hq > 0 && hkv > 0 && hq % hkv == 0 && window >= 0, "pt_attn_gemma_bwd: bad shape");
    const int G = hq / hkv;
    PT_REQUIRE(G <= 128 && (G & (G - 1)) == 0, "pt_attn_gemma_bwd: hq / hkv must be a power of two <= 128 (got %d)", G);
    PT_REQUIRE(lds >= S && lds % 8 == 0, "pt_attn_gemma_bwd: lds must be >= S and a multiple of 8");
    PT_REQUIRE(ld_dk >= static_cast<long long>(hkv) * hd && ld_dv >= static_cast<long long>(hkv) * hd,
               "pt_attn_gemma_bwd: dK / dV row strides must cover hkv * hd columns");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int QKV = (hq + 2 * hkv) * hd;
    CUtensorMap tq, tdo, tk, tv;
    int rc = gemma_stacked_map(&tq, q_rot, B, S, hq, hkv, hd, "attn_gemma_bwd Q");
    if (rc != PT_OK) return rc;
    rc = gemma_stacked_map(&tdo, dO_bf16, B, S, hq, hkv, hd, "attn_gemma_bwd dO");
    if (rc != PT_OK) return rc;
    rc = gemma_kv_map(&tk, k_rot, static_cast<long long>(hkv) * hd, B, S, hkv, hd, "attn_gemma_bwd K");
    if (rc != PT_OK) return rc;
    rc = gemma_kv_map(&tv, static_cast<const __nv_bfloat16*>(qkv) + static_cast<long long>(hq + hkv) * hd, QKV, B, S,
                      hkv, hd, "attn_gemma_bwd V");
    if (rc != PT_OK) return rc;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(attn_gemma_bwd_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             gemma_bwd_smem(256));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(attn_gemma_bwd_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     gemma_bwd_smem(128));
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(attn_gemma_bwd): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
        attr_done = true;
    }
    AttnGemmaBwdDev p;
    p.B = B; p.S = S; p.hq = hq; p.hkv = hkv; p.G = G; p.window = window;
    p.pos_per_blk = kGqRows / G;
    p.nqb = (S + p.pos_per_blk - 1) / p.pos_per_blk;
    p.nkt = (S + kGkTile - 1) / kGkTile;
    p.lds = lds;
    p.scale = scale;
    p.scale_log2e = scale * 1.4426950408889634f;
    p.kvwords = (S + 63) / 64;
    p.kvbits = static_cast<const unsigned long long*>(kvbits);
    p.lse = lse; p.delta = delta;
    p.dS = static_cast<__nv_bfloat16*>(dS_bf16);
    p.dK = static_cast<__nv_bfloat16*>(dK_bf16);
    p.dV = static_cast<__nv_bfloat16*>(dV_bf16);
    p.ld_dk = ld_dk; p.ld_dv = ld_dv;
    p.trace = g_bwd_trace; p.trace_iters = g_bwd_trace_iters;
    const long long items = static_cast<long long>(B) * hkv * p.nkt;
    PT_REQUIRE(items < (1ll << 31), "pt_attn_gemma_bwd: too many work items");
    const int grid = items < num_sms() ? static_cast<int>(items) : num_sms();
    if (hd == 256)
        attn_gemma_bwd_kernel<256><<<grid, kBwThreads, gemma_bwd_smem(256), st>>>(tq, tdo, tk, tv, p);
    else
        attn_gemma_bwd_kernel<128><<<grid, kBwThreads, gemma_bwd_smem(128), st>>>(tq, tdo, tk, tv, p);
    return check_launch("attn_gemma_bwd_kernel");
}
