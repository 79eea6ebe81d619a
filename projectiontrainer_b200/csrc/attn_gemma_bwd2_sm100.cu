// Fused causal / sliding-window GQA attention BACKWARD, second generation: a CTA PAIR per tile of 128 keys
// (autograd through HF:models/gemma3/modeling_gemma3.py:278-300, 370-384 as invoked by
// ref:Stage1/projector_trainer.py:236-237; head_dim 256, also instantiated for 128):
//     P  = exp2(Q K^T c - LSE),  dP = dO V^T,  dS = scale * P * (dP - delta)
//     dV = P^T dO,  dK = dS^T Q   (summed over the G = hq/hkv query heads of a kv head);  dQ = dS K is a band GEMM
//
// Why a pair.  At head_dim 256 one SM cannot hold what a key/value-stationary backward wants: dK and dV of 128 keys
// fill all 512 TMEM columns, and K + V tiles of 128 keys fill more than half of shared memory.  The first kernel
// (attn_gemma_bwd_sm100.cu) therefore works on 64-key tiles with transposed accumulators; its 128 x 64 x 16 MMAs
// read 6 KB of operands from smem per 32 tensor cycles and it has room for ONE query block in flight (measured:
// 6600 cycles per block, tensor pipe 70 % busy at 60 % of its rate — profiles/r02_attn_bwd_timeline_v2.txt).
// Here the two SMs of a cluster split the work BY OPERAND instead:
//     CTA 0 keeps the K tile  [128 keys x hd]:  S^T  = K Q_j^T   ->  P^T  ->  dV += P^T  dO_j
//     CTA 1 keeps the V tile  [128 keys x hd]:  dP^T = V dO_j^T  ->  dS^T ->  dK += dS^T Q_j
// per block j of 64 stacked query rows.  Each CTA owns one 128-key accumulator (hd TMEM columns, natural
// orientation: lane = key) plus a double-buffered 128 x 64 score tile, keeps one resident tile, and streams the
// query-side blocks through two 2-stage TMA rings.  The only exchange is the 16 KB P^T tile, which CTA 0 pushes
// into CTA 1's shared memory with one bulk DSMEM copy per block (cp.async.bulk.shared::cluster, completion on
// CTA 1's mbarrier); CTA 1 hands the buffer back with a remote mbarrier arrive.  Every MMA is one of the two the
// forward kernel issues (R E^T with 128-row / 64-row K-major slabs; X L with X a 128 x 64 K-major tile and L a
// 64-row tile read MN-major, N = hd), so the grad MMAs run 128 x hd x 16 — tensor-bound, not smem-bound.
// dS leaves TRANSPOSED, dS^T [B*hq, key, pos] (bf16): the dS^T operand tile of the dK product — 128 keys x 64
// positions of one head, 128-byte rows — goes out with one TMA store per block; the dQ GEMM reads it as an MN-major
// A operand.
//
// Roles per CTA: warps 0-7 elementwise (two threads per key row, 32 query columns each), warp 8 MMA issuer,
// warps 9 / 11 TMA producers (score-side / gradient-side ring), warp 10 copy engine driver (CTA 0: DSMEM push of P^T; CTA 1: dS^T TMA store + buffer hand-back).
// A block is 64 consecutive positions of ONE query head (the G heads of the kv head are separate blocks over the same
// resident tile: the sum over heads is a sum over blocks), so that a dS^T tile is contiguous per key in global memory.
#include <cstdlib>
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"
#include "attn_gemma_maps.cuh"

namespace pt {

constexpr int kB2Threads = 384;
constexpr int kB2Keys = 128;            // keys per item (resident tile rows, accumulator lanes)
constexpr int kB2Rows = 64;             // stacked query rows per block
constexpr int kB2XBytes = kB2Keys * kB2Rows * 2;      // 16 KB: P^T / dS^T tile, one 128-byte row per key
// resident tile + 2 x early block + 2 x late block + 2 x X tile + barriers + alignment slack
constexpr int gemma_bwd2_smem(int hd) { return kB2Keys * hd * 2 + 4 * kB2Rows * hd * 2 + 2 * kB2XBytes + 512 + 1024; }

struct AttnGemmaBwd2Dev {
    int B, S, hq, hkv, G, window;
    int nqb, nkt;                     // 64-position query blocks per head; 128-key tiles
    float scale, scale_log2e;
    const unsigned long long* kvbits; // [B][ceil(S/64)] key-valid bit masks or null
    int kvwords;
    const float* lse;                 // [B*hq, S] from the forward
    const float* delta;               // [B*hq, S] rowsum(dO * O)
    __nv_bfloat16* dK;                // element (b*S + k, kvh*hd + d) at dK[(b*S + k) * ld_dk + kvh*hd + d]
    __nv_bfloat16* dV;
    long long ld_dk, ld_dv;
    long long* trace;                 // tools/attn_bwd_trace.py --pair: SM-clock timestamps of cluster 0, or null
    int trace_iters;
    int dbg;                          // PT_BWD2_DBG bit mask (bring-up aid): 1 no DSMEM copy, 2 no dS^T store, 4 no epilogue
                                      // stores, 8 no statistics loads
};

// timeline probe (debug hook, null in the product path): event e of block `it`, per CTA rank of cluster 0
__device__ __forceinline__ void bwd2_trace(const AttnGemmaBwd2Dev& p, uint32_t rank, uint32_t it, int e) {
    if (p.trace != nullptr && blockIdx.x < 2 && it < static_cast<uint32_t>(p.trace_iters))
    {
        unsigned long long t;       // %globaltimer (ns): comparable between the two SMs of the pair, unlike clock64
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
        p.trace[(it * 2 + rank) * 16 + e] = static_cast<long long>(t);
    }
}

// item -> (batch, kv head, key tile, first 64-position block, blocks per head); an item walks G * nqb blocks
__device__ __forceinline__ void gemma_bwd2_item(const AttnGemmaBwd2Dev& p, int item, int& b, int& kvh, int& kt,
                                                int& qb0, int& nqb) {
    const int per = p.B * p.hkv;
    kt = item / per;                                          // early key tiles (seen by most queries) first
    const int rem = item - kt * per;
    b = rem / p.hkv; kvh = rem - b * p.hkv;
    const int k0 = kt * kB2Keys;
    const int k1 = min(p.S - 1, k0 + kB2Keys - 1);
    qb0 = k0 / kB2Rows;
    int qb1 = p.nqb - 1;
    if (p.window > 0) qb1 = min(qb1, (k1 + p.window - 1) / kB2Rows);
    nqb = qb1 - qb0 + 1;
}

template <int HD>
__global__ void __launch_bounds__(kB2Threads, 1)
attn_gemma_bwd2_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmdO,
                       const __grid_constant__ CUtensorMap tmK, const __grid_constant__ CUtensorMap tmV,
                       const __grid_constant__ CUtensorMap tmST, const AttnGemmaBwd2Dev p) {
    constexpr int kRB = kB2Keys * HD * 2;          // resident tile: HD/64 slabs of 128 rows x 128 B
    constexpr int kBB = kB2Rows * HD * 2;          // query-side block: HD/64 slabs of 64 rows x 128 B
    constexpr int kSlabs = HD / 64;
    constexpr int kSlabR = kB2Keys * 128, kSlabB = kB2Rows * 128;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sR = smem;
    uint8_t* sE = sR + kRB;                 // [2] operand of the score product (CTA 0: Q block, CTA 1: dO block)
    uint8_t* sL = sE + 2 * kBB;             // [2] operand of the gradient product (CTA 0: dO block, CTA 1: Q block)
    uint8_t* sX = sL + 2 * kBB;             // [2] CTA 0: P^T (MMA operand + source of the push); CTA 1: received P^T,
                                            //     turned into dS^T in place (MMA operand)
    uint64_t* bars = reinterpret_cast<uint64_t*>(sX + 2 * kB2XBytes);
    uint64_t* r_full = bars;                // TMA -> MMA, once per item
    uint64_t* r_empty = bars + 1;           // commit after the item's last score MMA -> TMA
    uint64_t* e_full = bars + 2;            // [2]
    uint64_t* e_empty = bars + 4;           // [2] commit after the score MMAs of the block
    uint64_t* l_full = bars + 6;            // [2]
    uint64_t* l_empty = bars + 8;           // [2] commit after the gradient MMAs of the block
    uint64_t* t_full = bars + 10;           // [2] commit: score tile in TMEM
    uint64_t* t_free = bars + 12;           // [2] 256 arrivals: score tile read
    uint64_t* x_full = bars + 14;           // [2] 256 arrivals: X tile written (MMA and copy warp wait)
    uint64_t* x_empty = bars + 16;          // [2] commit after the gradient MMAs that read the X tile
    uint64_t* copy_done = bars + 18;        // [2] CTA 0's: 8 remote arrivals from CTA 1: the pushed P^T has landed (source free)
    uint64_t* acc_full = bars + 20;         // commit after the item's last gradient MMA
    uint64_t* acc_free = bars + 21;         // 256 arrivals: accumulator read
    uint64_t* recv_free = bars + 22;        // [2] CTA 0's: remote arrival from CTA 1's copy warp: its X[b] has been consumed
    uint64_t* recv_full = bars + 24;        // [2] CTA 1's: expect_tx + the DSMEM copy's complete_tx
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 26);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    if (threadIdx.x == 256) {
        ptx::prefetch_tmap(&tmQ); ptx::prefetch_tmap(&tmdO); ptx::prefetch_tmap(&tmK); ptx::prefetch_tmap(&tmV);
        ptx::prefetch_tmap(&tmST);
        ptx::mbar_init(r_full, 1); ptx::mbar_init(r_empty, 1);
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&e_full[s], 1); ptx::mbar_init(&e_empty[s], 1);
            ptx::mbar_init(&l_full[s], 1); ptx::mbar_init(&l_empty[s], 1);
            ptx::mbar_init(&t_full[s], 1); ptx::mbar_init(&t_free[s], 256);
            ptx::mbar_init(&x_full[s], 256); ptx::mbar_init(&x_empty[s], 1); ptx::mbar_init(&copy_done[s], 8);
            ptx::mbar_init(&recv_free[s], 1); ptx::mbar_init(&recv_full[s], 1);
        }
        ptx::mbar_init(acc_full, 1); ptx::mbar_init(acc_free, 256);
        ptx::fence_mbar_init();
    }
    constexpr uint32_t kTmemCols = HD == 256 ? 512 : 256;
    if (warp == 8) {
        ptx::tmem_alloc(tmem_slot, kTmemCols);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    ptx::cluster_sync_all();                // both CTAs' barriers exist before any remote signal / copy
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tT[2] = {tmem_base, tmem_base + kB2Rows};
    const uint32_t tAcc = tmem_base + 2 * kB2Rows;
    const int items = p.B * p.hkv * p.nkt;
    const int cid = static_cast<int>(ptx::cluster_id_x()), ncl = static_cast<int>(ptx::num_clusters_x());

    if (warp == 9) {
        // ===================== TMA producer: resident tile + score-product operand (E ring) =====================
        if (lane == 0) {
            const CUtensorMap* tmR = rank == 0 ? &tmK : &tmV;
            const CUtensorMap* tmE = rank == 0 ? &tmQ : &tmdO;
            uint32_t n_item = 0, it = 0;
            for (int jj = 0; jj * ncl < items; ++jj) {
                const int item = snake_item(jj, cid, ncl);
                if (item >= items) continue;
                int b, kvh, kt, qb0, nqb;
                gemma_bwd2_item(p, item, b, kvh, kt, qb0, nqb);
                ptx::mbar_wait(r_empty, (n_item & 1) ^ 1);
                ptx::mbar_arrive_expect_tx(r_full, kRB);
#pragma unroll
                for (int c = 0; c < kSlabs; ++c)
                    ptx::tma_load_5d(tmR, r_full, sR + c * kSlabR, c * 64, kt * kB2Keys, kvh, b, 0);
                const int nblk = p.G * nqb;
                for (int i = 0; i < nblk; ++i, ++it) {
                    const int g = i / nqb, s0 = (qb0 + i - g * nqb) * kB2Rows, head = kvh * p.G + g, st = it & 1;
                    ptx::mbar_wait(&e_empty[st], ((it >> 1) & 1) ^ 1);
                    bwd2_trace(p, rank, it, 0);
                    ptx::mbar_arrive_expect_tx(&e_full[st], kBB);
#pragma unroll
                    for (int c = 0; c < kSlabs; ++c)
                        ptx::tma_load_5d(tmE, &e_full[st], sE + st * kBB + c * kSlabB, c * 64, s0, head, b, 0);
                }
                ++n_item;
            }
        }
    } else if (warp == 11) {
        // ===================== TMA producer: gradient-product operand (L ring) =====================
        // (its buffers free up later than the E ring's: one thread for both would hold the next E load back)
        if (lane == 0) {
            const CUtensorMap* tmL = rank == 0 ? &tmdO : &tmQ;
            uint32_t it = 0;
            for (int jj = 0; jj * ncl < items; ++jj) {
                const int item = snake_item(jj, cid, ncl);
                if (item >= items) continue;
                int b, kvh, kt, qb0, nqb;
                gemma_bwd2_item(p, item, b, kvh, kt, qb0, nqb);
                const int nblk = p.G * nqb;
                for (int i = 0; i < nblk; ++i, ++it) {
                    const int g = i / nqb, s0 = (qb0 + i - g * nqb) * kB2Rows, head = kvh * p.G + g, st = it & 1;
                    ptx::mbar_wait(&l_empty[st], ((it >> 1) & 1) ^ 1);
                    bwd2_trace(p, rank, it, 1);
                    ptx::mbar_arrive_expect_tx(&l_full[st], kBB);
#pragma unroll
                    for (int c = 0; c < kSlabs; ++c)
                        ptx::tma_load_5d(tmL, &l_full[st], sL + st * kBB + c * kSlabB, c * 64, s0, head, b, 0);
                }
            }
        }
    } else if (warp == 8) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idescT = ptx::make_idesc_bf16(kB2Keys, kB2Rows, 0, 0);   // R E^T: both K-major
            const uint32_t idescA = ptx::make_idesc_bf16(kB2Keys, HD, 0, 1);        // X L: L read MN-major (N = hd)
            const uint32_t ra = ptx::smem_u32(sR);
            uint32_t n_item = 0, it = 0;
            auto scores = [&](uint32_t j) {                 // global block index j
                const int st = j & 1;
                ptx::mbar_wait(&e_full[st], (j >> 1) & 1);
                ptx::mbar_wait(&t_free[st], ((j >> 1) & 1) ^ 1);
                bwd2_trace(p, rank, j, 2);
                ptx::tc_fence_after();
                const uint32_t ea = ptx::smem_u32(sE + st * kBB);
#pragma unroll
                for (int k = 0; k < HD / 16; ++k) {
                    const uint64_t adesc = ptx::make_smem_desc_sw128(ra + (k >> 2) * kSlabR + (k & 3) * 32, 0, 1024);
                    const uint64_t bdesc = ptx::make_smem_desc_sw128(ea + (k >> 2) * kSlabB + (k & 3) * 32, 0, 1024);
                    ptx::mma_bf16_ss(tT[st], adesc, bdesc, idescT, k != 0);
                }
                ptx::mma_commit(&t_full[st]);
                ptx::mma_commit(&e_empty[st]);
            };
            for (int jj = 0; jj * ncl < items; ++jj) {
                const int item = snake_item(jj, cid, ncl);
                if (item >= items) continue;
                int b, kvh, kt, qb0, nqb;
                gemma_bwd2_item(p, item, b, kvh, kt, qb0, nqb);
                ptx::mbar_wait(r_full, n_item & 1);
                const int nblk = p.G * nqb;
                scores(it);
                if (nblk == 1) ptx::mma_commit(r_empty);
                for (int i = 0; i < nblk; ++i, ++it) {
                    if (i + 1 < nblk) {
                        scores(it + 1);                      // overlaps the elementwise work on block it
                        if (i + 2 == nblk) ptx::mma_commit(r_empty);
                    }
                    const int st = it & 1;
                    const int xb = it & 1;
                    if (i == 0) {                           // previous item's accumulator has been read
                        ptx::mbar_wait(acc_free, (n_item & 1) ^ 1);
                    }
                    ptx::mbar_wait(&x_full[xb], (it >> 1) & 1);
                    bwd2_trace(p, rank, it, 3);
                    ptx::mbar_wait(&l_full[st], (it >> 1) & 1);
                    bwd2_trace(p, rank, it, 4);
                    ptx::tc_fence_after();
                    const uint32_t xa = ptx::smem_u32(sX + xb * kB2XBytes), la = ptx::smem_u32(sL + st * kBB);
#pragma unroll
                    for (int k = 0; k < kB2Rows / 16; ++k) {
                        const uint64_t adesc = ptx::make_smem_desc_sw128(xa + k * 32, 0, 1024);
                        const uint64_t bdesc = ptx::make_smem_desc_sw128(la + k * 16 * 128, kSlabB, 1024);
                        ptx::mma_bf16_ss(tAcc, adesc, bdesc, idescA, (i | k) != 0);
                    }
                    ptx::mma_commit(&x_empty[xb]);
                    ptx::mma_commit(&l_empty[st]);
                    bwd2_trace(p, rank, it, 5);
                }
                ptx::mma_commit(acc_full);
                ++n_item;
            }
        }
    } else if (warp == 10) {
        // ===================== copy engine driver =====================
        // CTA 0: pushes every finished P^T tile into the same X buffer of CTA 1 (bulk DSMEM copy, completion on CTA 1's
        // recv_full) as soon as CTA 1 has released that buffer.  CTA 1: arms recv_full for the tile to come and hands
        // each buffer back to CTA 0 once its gradient MMAs have read the dS^T that replaced P^T in it.
        if (lane == 0) {
            uint32_t it = 0;
            const uint32_t peer_x = ptx::mapa(ptx::smem_u32(sX), 1);
            const uint32_t peer_full = ptx::mapa(ptx::smem_u32(recv_full), 1);
            const uint32_t peer_free = ptx::mapa(ptx::smem_u32(recv_free), 0);
            for (int jj = 0; jj * ncl < items; ++jj) {
                const int item = snake_item(jj, cid, ncl);
                if (item >= items) continue;
                int b, kvh, kt, qb0, nqb;
                gemma_bwd2_item(p, item, b, kvh, kt, qb0, nqb);
                const int nblk = p.G * nqb;
                for (int i = 0; i < nblk; ++i, ++it) {
                    const int xb = it & 1;
                    const uint32_t par = (it >> 1) & 1;
                    if (rank == 0) {
                        ptx::mbar_wait(&x_full[xb], par);
                        bwd2_trace(p, rank, it, 6);
                        ptx::mbar_wait(&recv_free[xb], par ^ 1);
                        bwd2_trace(p, rank, it, 7);
                        if (!(p.dbg & 1))
                            ptx::dsmem_bulk_copy(peer_x + xb * kB2XBytes, sX + xb * kB2XBytes, kB2XBytes, peer_full + xb * 8);
                    } else {
                        if (p.dbg & 1) ptx::mbar_arrive(&recv_full[xb]); else ptx::mbar_arrive_expect_tx(&recv_full[xb], kB2XBytes);
                        // dS^T tile [128 keys][64 positions of one head] -> dS^T[b*hq + head, key, pos]: the MMA operand
                        // tile itself is the TMA box (128-byte rows, 128-byte swizzle)
                        ptx::mbar_wait(&x_full[xb], par);
                        if (!(p.dbg & 2)) {
                            const int g = i / nqb;
                            ptx::tma_store_3d(&tmST, sX + xb * kB2XBytes, (qb0 + i - g * nqb) * kB2Rows, kt * kB2Keys,
                                              b * p.hq + kvh * p.G + g);
                            ptx::bulk_commit();
                        }
                        ptx::mbar_wait(&x_empty[xb], par);       // the gradient MMAs have read the tile ...
                        ptx::bulk_wait_read<0>();                // ... and so has the store: hand the buffer back
                        ptx::mbar_arrive_cluster(peer_free + xb * 8);
                    }
                }
            }
            ptx::bulk_wait_all<0>();
        }
    } else {
        // ===================== elementwise warps: two threads per key row =====================
        // warps 0-3 own query columns [0,32) of key rows 32 (warp & 3) .. +31, warps 4-7 columns [32,64); column c of a
        // block is position s0 + c of the block's head.  LSE / delta are per COLUMN here: lane l of a warp fetches the
        // value of column half*32 + l (one block ahead: its L2 latency overlaps the block at hand) and the warp shares
        // them by shuffle.
        const int row = (warp & 3) * 32 + lane;                    // TMEM lane = key inside the tile
        const int half = warp >> 2;
        const uint32_t lane_off = static_cast<uint32_t>((warp & 3) * 32) << 16;
        const float c2 = p.scale_log2e;
        const int my_c = half * 32 + lane;                         // the column whose statistics this lane fetches
        const float* stat = rank == 0 ? p.lse : p.delta;
        const float dead = rank == 0 ? INFINITY : 0.f;             // dead columns: P = exp2(x - inf) = 0 / delta 0
        const uint32_t peer_copy_done = ptx::mapa(ptx::smem_u32(copy_done), 0);
        auto fetch_stat = [&](int b, int head, int s0) -> float {
            return (s0 + my_c < p.S && !(p.dbg & 8)) ? __ldg(stat + (static_cast<long long>(b) * p.hq + head) * p.S + s0 + my_c) : dead;
        };
        uint32_t n_item = 0, it = 0;
        for (int jj = 0; jj * ncl < items; ++jj) {
                const int item = snake_item(jj, cid, ncl);
                if (item >= items) continue;
            int b, kvh, kt, qb0, nqb;
            gemma_bwd2_item(p, item, b, kvh, kt, qb0, nqb);
            const int nblk = p.G * nqb;
            const int key = kt * kB2Keys + row;
            bool key_ok = key < p.S;
            if (key_ok && p.kvbits) {
                const unsigned long long w = __ldg(p.kvbits + static_cast<long long>(b) * p.kvwords + (key >> 6));
                key_ok = (w >> (key & 63)) & 1ull;
            }
            // positions that see this key: key <= s < key + window (and s < S)
            const int s_lo = key, s_hi = min(p.S - 1, p.window > 0 ? key + p.window - 1 : p.S - 1);
            float stat_next = fetch_stat(b, kvh * p.G, qb0 * kB2Rows);
            for (int i = 0; i < nblk; ++i, ++it) {
                const int g = i / nqb, s0 = (qb0 + i - g * nqb) * kB2Rows;
                const int tb = it & 1;
                const int xb = it & 1;
                const uint32_t xpar = (it >> 1) & 1;
                // this lane's column statistic (LSE on CTA 0, delta on CTA 1), and the next block's request
                const float my_stat = stat_next;
                if (i + 1 < nblk) {
                    const int g1 = (i + 1) / nqb;
                    stat_next = fetch_stat(b, kvh * p.G + g1, (qb0 + i + 1 - g1 * nqb) * kB2Rows);
                }
                // visibility of this thread's 32 columns: positions [s_lo, s_hi]
                uint32_t mask = 0;
                if (key_ok) {
                    const int lo = max(0, s_lo - s0), hi = min(kB2Rows - 1, s_hi - s0);
                    if (hi >= lo) {
                        const unsigned long long m = ((hi - lo + 1 >= 64) ? ~0ull : ((1ull << (hi - lo + 1)) - 1ull)) << lo;
                        mask = static_cast<uint32_t>(m >> (half * 32));
                    }
                }
                uint32_t v[32];
                ptx::mbar_wait(&t_full[tb], (it >> 1) & 1);
                if (threadIdx.x == 0) bwd2_trace(p, rank, it, 8);
                ptx::tc_fence_after();
                ptx::tmem_ld_32x32(tT[tb] + lane_off + half * 32, v);
                ptx::tmem_ld_wait();
                ptx::tc_fence_before();
                ptx::mbar_arrive(&t_free[tb]);
                uint32_t out[16];
                if (rank == 0) {
                    // P^T = exp2(S^T c - LSE[column]); masked -> exactly 0
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float l0 = __shfl_sync(0xffffffffu, my_stat, 2 * j), l1 = __shfl_sync(0xffffffffu, my_stat, 2 * j + 1);
                        float p0 = ptx::exp2_fast(fmaf(__uint_as_float(v[2 * j]), c2, -l0));
                        float p1 = ptx::exp2_fast(fmaf(__uint_as_float(v[2 * j + 1]), c2, -l1));
                        if (!((mask >> (2 * j)) & 1u)) p0 = 0.f;
                        if (!((mask >> (2 * j + 1)) & 1u)) p1 = 0.f;
                        out[j] = pack_bf16(p0, p1);
                    }
                } else {
                    // dS^T = scale * P^T * (dP^T - delta[column]) with the P^T tile CTA 0 pushed into X[0]
                    ptx::mbar_wait(&recv_full[xb], xpar);
                    if (threadIdx.x == 0) bwd2_trace(p, rank, it, 9);
                    const uint32_t pr = ptx::smem_u32(sX + xb * kB2XBytes) + row * 128;
                    uint32_t pw[16];
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        const float4 f = ptx::lds128(pr + (((half * 4 + q) ^ (row & 7)) << 4));
                        pw[4 * q] = __float_as_uint(f.x); pw[4 * q + 1] = __float_as_uint(f.y);
                        pw[4 * q + 2] = __float_as_uint(f.z); pw[4 * q + 3] = __float_as_uint(f.w);
                    }
#pragma unroll
                    for (int j = 0; j < 16; ++j) {
                        const float d0 = __shfl_sync(0xffffffffu, my_stat, 2 * j), d1 = __shfl_sync(0xffffffffu, my_stat, 2 * j + 1);
                        out[j] = pack_bf16(p.scale * bf16_lo(pw[j]) * (__uint_as_float(v[2 * j]) - d0),
                                           p.scale * bf16_hi(pw[j]) * (__uint_as_float(v[2 * j + 1]) - d1));
                    }
                    // the tile has landed: CTA 0 may refill the buffer it was pushed from (one arrival per warp)
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive_cluster(peer_copy_done + xb * 8);
                }
                if (threadIdx.x == 0) bwd2_trace(p, rank, it, 10);
                if (rank == 0) {
                    // the buffer is free once the gradient MMAs of block it - 2 have read it and its push has landed
                    ptx::mbar_wait(&x_empty[xb], xpar ^ 1);
                    ptx::mbar_wait(&copy_done[xb], xpar ^ 1);
                }   // (CTA 1 writes dS^T over the P^T it has just read: the buffer is its own until the MMAs are done)
                if (threadIdx.x == 0) bwd2_trace(p, rank, it, 11);
                const uint32_t xr = ptx::smem_u32(sX + xb * kB2XBytes) + row * 128;
#pragma unroll
                for (int q = 0; q < 4; ++q)
                    ptx::sts128(xr + (((half * 4 + q) ^ (row & 7)) << 4), out[4 * q], out[4 * q + 1], out[4 * q + 2], out[4 * q + 3]);
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before();
                ptx::mbar_arrive(&x_full[xb]);
                if (threadIdx.x == 0) bwd2_trace(p, rank, it, 12);
            }
            if (threadIdx.x == 0) bwd2_trace(p, rank, it - 1, 13);
            // ---- epilogue: accumulator [128 keys x hd] (lane = key) -> bf16 -> this key's row of dV (CTA 0) / dK (CTA 1)
            ptx::mbar_wait(acc_full, n_item & 1);
            ptx::tc_fence_after();
            __nv_bfloat16* outp = rank == 0 ? p.dV : p.dK;
            const long long ldo = rank == 0 ? p.ld_dv : p.ld_dk;
            __nv_bfloat16* o = outp + (static_cast<long long>(b) * p.S + (key < p.S ? key : 0)) * ldo + kvh * HD + half * (HD / 2);
#pragma unroll 1
            for (int c = 0; c < HD / 64; ++c) {
                uint32_t a[32];
                ptx::tmem_ld_32x32(tAcc + lane_off + half * (HD / 2) + c * 32, a);
                ptx::tmem_ld_wait();
                if (key < p.S && !(p.dbg & 4)) {
#pragma unroll
                    for (int q = 0; q < 4; ++q) {
                        uint4 u;
                        u.x = pack_bf16(__uint_as_float(a[8 * q]), __uint_as_float(a[8 * q + 1]));
                        u.y = pack_bf16(__uint_as_float(a[8 * q + 2]), __uint_as_float(a[8 * q + 3]));
                        u.z = pack_bf16(__uint_as_float(a[8 * q + 4]), __uint_as_float(a[8 * q + 5]));
                        u.w = pack_bf16(__uint_as_float(a[8 * q + 6]), __uint_as_float(a[8 * q + 7]));
                        *reinterpret_cast<uint4*>(o + c * 32 + q * 8) = u;
                    }
                }
            }
            ptx::tc_fence_before();
            ptx::mbar_arrive(acc_free);
            if (threadIdx.x == 0) bwd2_trace(p, rank, it - 1, 14);
            ++n_item;
        }
    }
    ptx::tc_fence_before();
    ptx::cluster_sync_all();                // no CTA leaves while its peer may still signal / copy into it
    if (warp == 8) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, kTmemCols);
    }
}

}  // namespace pt

using namespace pt;

static long long* g_bwd2_trace = nullptr;
static int g_bwd2_trace_iters = 0;
// debug hook for tools/attn_bwd_trace.py --pair (not part of the product path): device buffer of 32 * iters int64
extern "C" int pt_attn_gemma_bwd_pair_set_trace(void* buf, int iters) {
    g_bwd2_trace = static_cast<long long*>(buf);
    g_bwd2_trace_iters = buf ? iters : 0;
    return PT_OK;
}

extern "C" int pt_attn_gemma_bwd_pair(const void* q_rot, const void* k_rot, const void* qkv, const void* dO_bf16,
                                      const void* kvbits, const float* lse, const float* delta, void* dST_bf16,
                                      void* dK_bf16, long long ld_dk, void* dV_bf16, long long ld_dv, int B, int S,
                                      int hq, int hkv, int hd, int window, float scale, int ldt, void* stream) {
    PT_REQUIRE(q_rot && k_rot && qkv && dO_bf16 && lse && delta && dST_bf16 && dK_bf16 && dV_bf16,
               "pt_attn_gemma_bwd_pair: null pointer");
    PT_REQUIRE(hd == 256 || hd == 128, "pt_attn_gemma_bwd_pair: head_dim must be 256 or 128 (got %d)", hd);
    PT_REQUIRE(B > 0 && S > 0 && hq > 0 && hkv > 0 && hq % hkv == 0 && window >= 0, "pt_attn_gemma_bwd_pair: bad shape");
    const int G = hq / hkv;
    PT_REQUIRE(ldt >= S && ldt % 8 == 0, "pt_attn_gemma_bwd_pair: ldt must be >= S and a multiple of 8");
    PT_REQUIRE(ld_dk >= static_cast<long long>(hkv) * hd && ld_dv >= static_cast<long long>(hkv) * hd &&
                   ld_dk % 8 == 0 && ld_dv % 8 == 0,
               "pt_attn_gemma_bwd_pair: dK / dV row strides must cover hkv * hd columns and be multiples of 8");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    const int QKV = (hq + 2 * hkv) * hd;
    CUtensorMap tq, tdo, tk, tv, tst;
    int rc = gemma_head_rows_map(&tq, q_rot, B, S, hq, hd, kB2Rows, "attn_gemma_bwd_pair Q");
    if (rc != PT_OK) return rc;
    rc = gemma_head_rows_map(&tdo, dO_bf16, B, S, hq, hd, kB2Rows, "attn_gemma_bwd_pair dO");
    if (rc != PT_OK) return rc;
    rc = gemma_dst_map(&tst, dST_bf16, ldt, B, S, hq, "attn_gemma_bwd_pair dS^T");
    if (rc != PT_OK) return rc;
    rc = gemma_kv_map(&tk, k_rot, static_cast<long long>(hkv) * hd, B, S, hkv, hd, "attn_gemma_bwd_pair K", kB2Keys);
    if (rc != PT_OK) return rc;
    rc = gemma_kv_map(&tv, static_cast<const __nv_bfloat16*>(qkv) + static_cast<long long>(hq + hkv) * hd, QKV, B, S,
                      hkv, hd, "attn_gemma_bwd_pair V", kB2Keys);
    if (rc != PT_OK) return rc;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(attn_gemma_bwd2_kernel<256>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                             gemma_bwd2_smem(256));
        if (e == cudaSuccess)
            e = cudaFuncSetAttribute(attn_gemma_bwd2_kernel<128>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                     gemma_bwd2_smem(128));
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(attn_gemma_bwd2): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
        attr_done = true;
    }
    AttnGemmaBwd2Dev p;
    p.B = B; p.S = S; p.hq = hq; p.hkv = hkv; p.G = G; p.window = window;
    p.nqb = (S + kB2Rows - 1) / kB2Rows;
    p.nkt = (S + kB2Keys - 1) / kB2Keys;
    p.scale = scale;
    p.scale_log2e = scale * 1.4426950408889634f;
    p.kvwords = (S + 63) / 64;
    p.kvbits = static_cast<const unsigned long long*>(kvbits);
    p.lse = lse; p.delta = delta;
    p.dK = static_cast<__nv_bfloat16*>(dK_bf16);
    p.dV = static_cast<__nv_bfloat16*>(dV_bf16);
    p.ld_dk = ld_dk; p.ld_dv = ld_dv;
    p.trace = g_bwd2_trace; p.trace_iters = g_bwd2_trace_iters;
    { const char* e = getenv("PT_BWD2_DBG"); p.dbg = e ? atoi(e) : 0; }
    const long long items = static_cast<long long>(B) * hkv * p.nkt;
    PT_REQUIRE(items < (1ll << 31), "pt_attn_gemma_bwd_pair: too many work items");
    const int max_clusters = num_sms() / 2;
    const int clusters = items < max_clusters ? static_cast<int>(items) : max_clusters;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(2 * clusters, 1, 1);
    cfg.blockDim = dim3(kB2Threads, 1, 1);
    cfg.dynamicSmemBytes = gemma_bwd2_smem(hd);
    cfg.stream = st;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2; attr[0].val.clusterDim.y = 1; attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    cudaError_t e = hd == 256 ? cudaLaunchKernelEx(&cfg, attn_gemma_bwd2_kernel<256>, tq, tdo, tk, tv, tst, p)
                              : cudaLaunchKernelEx(&cfg, attn_gemma_bwd2_kernel<128>, tq, tdo, tk, tv, tst, p);
    if (e != cudaSuccess) { set_error("attn_gemma_bwd2_kernel: launch failed: %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
    return PT_OK;
}
