// CTA-pair bf16 GEMM for sm_100a: tcgen05.mma.cta_group::2 (UMMA M = 256 over two SMs of a TPC).
//
// One cluster of two CTAs owns a 256 x BN output tile.  CTA r of the pair stages rows
// [128 r, 128 r + 128) of the A tile and rows [BN/2 r, BN/2 (r+1)) of the B tile, so every operand byte
// is fetched from L2 once per PAIR (half the L2->SM traffic per flop of the single-CTA kernel, which is
// what bounded it), and the pair's accumulator lives in both CTAs' TMEM (128 lanes x BN columns each).
//
//   warp 0      TMA producer (one lane, both CTAs): cp.async.bulk.tensor ... cta_group::2, bytes credited
//               to the LEADER CTA's full barrier
//   warp 1      MMA issuer (one lane, leader CTA only); tcgen05.commit multicast frees the smem stage /
//               publishes the accumulator in BOTH CTAs
//   warp 2      TMEM allocator (cta_group::2, both CTAs)
//   warps 4..11 epilogue: tcgen05.ld -> registers -> (alpha, bias, act) -> per-warp swizzled smem
//               staging -> coalesced 16-byte global stores (+ residual), rows of this CTA's half
//
// Two accumulator stages (2 x BN <= 512 TMEM columns) overlap the epilogue of tile i with the main loop
// of tile i+1.  Persistent: cluster c walks tiles c, c + #clusters, ...; inside one batch slice the
// tile order runs fastest along the dimension with fewer tiles, so the operand panel that is re-read
// across concurrently running tiles is the small one and stays in L2.
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"

namespace pt {

constexpr int kGemm2Threads = 384;     // 4 control warps + 8 epilogue warps
constexpr int kEpiWarps = 8;
constexpr int kEpiStageBytes = 4096;   // per epilogue warp: 32 rows x 128 B
constexpr int kSmemTotal = 232448;
// internal activation code: the GeGLU-backward epilogue on the load/store unit (first version, kept as the A/B
// reference: PT_GEMM_GBWD_TMA=0); PT_ACT_GEGLU_BWD itself moves its tiles with TMA
constexpr int kActGbwdLsu = 100;
// Epilogues that move their tiles with TMA (tma_epilogue(ACT)): per epilogue warp two buffers of up to two 32 x 32
// bf16 tiles (PT_ACT_GEGLU_BWD: gate | up), 64-byte rows, SWIZZLE_64B
constexpr int kGbwdBufBytes = 4096;
constexpr int kGbwdStageBytes = 2 * kGbwdBufBytes;
__host__ __device__ constexpr bool tma_epilogue(int act) {
    return act == PT_ACT_GEGLU_BWD || act == PT_ACT_GELU_ERF_BWD || act == PT_ACT_ATTN_DELTA;
}

template <int BN, int ACT = PT_ACT_NONE>
struct Gemm2Cfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = (BN / 2) * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int EPI_WARP_BYTES = tma_epilogue(ACT) ? kGbwdStageBytes : kEpiStageBytes;
    static constexpr int EPI_BYTES = kEpiWarps * EPI_WARP_BYTES;
    static constexpr int BAR_BYTES = 512;
    static constexpr int STAGES_RAW = (kSmemTotal - 1024 - EPI_BYTES - BAR_BYTES) / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + EPI_BYTES + BAR_BYTES + 1024;
    static constexpr int ACC_STRIDE = 256;   // TMEM columns per accumulator stage
};

// ---- epilogue helper --------------------------------------------------------------------------
// Thread `lane` holds the raw f32 accumulators of 32 consecutive columns of row `lane` of a 32-row slab.
// They are transposed through a per-warp XOR-swizzled smem tile (32 rows x 128 B, conflict-free both
// ways) so that everything touching global memory is row-contiguous: in the second phase lane l owns
// the column quad 4 (l & 7) .. +3 of rows (l >> 3) + 4 it, it = 0..7 — its bias quad is loaded once, the
// eight residual quads are loaded up front (independent loads; C may alias R element-wise), and each
// row segment is written as 128 contiguous bytes (f32) or 64 (bf16) by 8 neighbouring lanes.
// Column quads are all-or-nothing (N is rounded up to a multiple of 4: see pt_gemm_bf16's contract), which
// keeps this straight-line code small — the epilogue is instruction-issue and I-cache sensitive.
struct EpiArgs {
    uint8_t* C;           // byte address of element (row 0 of the slab, column col0)
    uint8_t* Cpre;        // same position (bf16) or nullptr
    const uint8_t* R;     // same position or nullptr
    const float* bias;    // bias + col0 or nullptr
    long long c_step, p_step, r_step;   // bytes per 4 rows
    long long c_row, p_row, r_row;      // bytes per row
    int c_f32, r_f32;
    int c_atomic;         // split-K launch: f32 C is accumulated into with vector atomics instead of stored
    float alpha;
    int rows_valid;       // rows of the slab that exist (1..32)
    int row_lo;           // rows of the slab below this one belong to the previous (overlapped) tile
    int cols_q;           // columns of the chunk to produce, multiple of 4 (4..32)
};

template <int ACT>
__device__ __forceinline__ void epi_chunk(const uint32_t (&v)[32], uint32_t stg, int lane, const EpiArgs& o) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
        ptx::sts128(stg + lane * 128 + ((j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncwarp();
    const int ch = lane & 7, c = ch * 4, rsub = lane >> 3;
    const bool col_ok = c < o.cols_q;
    float4 b4 = make_float4(0.f, 0.f, 0.f, 0.f);
    if (o.bias && col_ok) b4 = __ldg(reinterpret_cast<const float4*>(o.bias + c));
    float4 res[8];
    if (o.R) {
        const uint8_t* rp = o.R + rsub * o.r_row + c * (o.r_f32 ? 4 : 2);
#pragma unroll
        for (int it = 0; it < 8; ++it) {
            const int r = it * 4 + rsub;
            res[it] = make_float4(0.f, 0.f, 0.f, 0.f);
            if (col_ok && r >= o.row_lo && r < o.rows_valid) {
                if (o.r_f32) {
                    res[it] = *reinterpret_cast<const float4*>(rp);
                } else {
                    const uint2 q = *reinterpret_cast<const uint2*>(rp);
                    res[it] = make_float4(bf16_lo(q.x), bf16_hi(q.x), bf16_lo(q.y), bf16_hi(q.y));
                }
            }
            rp += o.r_step;
        }
    }
    uint8_t* cp = o.C + rsub * o.c_row + c * (o.c_f32 ? 4 : 2);
    uint8_t* pp = o.Cpre + rsub * o.p_row + c * 2;
    const uint32_t sp = stg + rsub * 128;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int r = it * 4 + rsub;
        const float4 a = ptx::lds128(sp + it * 512 + ((ch ^ (r & 7)) << 4));
        const bool ok = col_ok && r >= o.row_lo && r < o.rows_valid;
        float x0 = fmaf(a.x, o.alpha, b4.x), x1 = fmaf(a.y, o.alpha, b4.y);
        float x2 = fmaf(a.z, o.alpha, b4.z), x3 = fmaf(a.w, o.alpha, b4.w);
        if (o.Cpre != nullptr) {
            uint2 u; u.x = pack_bf16(x0, x1); u.y = pack_bf16(x2, x3);
            if (ok) *reinterpret_cast<uint2*>(pp) = u;
            pp += o.p_step;
        }
        if (ACT == PT_ACT_GELU_TANH) {
            x0 = gelu_tanh_fast_f(x0); x1 = gelu_tanh_fast_f(x1);
            x2 = gelu_tanh_fast_f(x2); x3 = gelu_tanh_fast_f(x3);
        } else if (ACT == PT_ACT_GELU_ERF) {
            x0 = gelu_erf_f(x0); x1 = gelu_erf_f(x1); x2 = gelu_erf_f(x2); x3 = gelu_erf_f(x3);
        }
        if (o.R) { x0 += res[it].x; x1 += res[it].y; x2 += res[it].z; x3 += res[it].w; }
        if (o.c_f32) {
            if (ok) {
                if (o.c_atomic) atomicAdd(reinterpret_cast<float4*>(cp), make_float4(x0, x1, x2, x3));
                else *reinterpret_cast<float4*>(cp) = make_float4(x0, x1, x2, x3);
            }
        } else {
            uint2 u; u.x = pack_bf16(x0, x1); u.y = pack_bf16(x2, x3);
            if (ok) *reinterpret_cast<uint2*>(cp) = u;
        }
        cp += o.c_step;
    }
    __syncwarp();
}

// GeGLU backward fused into the down-projection input gradient: the accumulator chunk is d_act; the saved
// pre-activations (gate at GU, up at GU + ff) are read in the row-contiguous second-phase layout and the two
// halves of d_gu are written at C and C + ff:  d_gate = d_act * up * gelu'(gate),  d_up = d_act * gelu(gate).
// (d_act stays in f32: the unfused path rounds it to bf16 between the GEMM and geglu_bwd.)
struct EpiGbwd {
    uint8_t* C;           // d_gu, byte address of (row 0 of the slab, gate column col0)
    const uint8_t* GU;    // gu, same position
    long long c_row, g_row, ff_bytes;
    int rows_valid, row_lo, cols_q;
    GluConsts glu;
};

__device__ __forceinline__ void epi_chunk_gbwd(const uint32_t (&v)[32], uint32_t stg, int lane, const EpiGbwd& o) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
        ptx::sts128(stg + lane * 128 + ((j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncwarp();
    const int ch = lane & 7, c = ch * 4, rsub = lane >> 3;
    const bool col_ok = c < o.cols_q;
    uint2 gq[8], uq[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int r = it * 4 + rsub;
        gq[it] = make_uint2(0u, 0u); uq[it] = make_uint2(0u, 0u);
        if (col_ok && r >= o.row_lo && r < o.rows_valid) {
            const uint8_t* gp = o.GU + r * o.g_row + c * 2;
            gq[it] = *reinterpret_cast<const uint2*>(gp);
            uq[it] = *reinterpret_cast<const uint2*>(gp + o.ff_bytes);
        }
    }
    const uint32_t sp = stg + rsub * 128;
    const GluConsts2 k2 = glu_consts2(o.glu);
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int r = it * 4 + rsub;
        const float4 a = ptx::lds128(sp + it * 512 + ((ch ^ (r & 7)) << 4));
        const bool ok = col_ok && r >= o.row_lo && r < o.rows_valid;
        // two column pairs per lane and row, each as one packed-f32 register pair (FFMA2 / FMUL2)
        const uint32_t gw[2] = {gq[it].x, gq[it].y}, uw[2] = {uq[it].x, uq[it].y};
        const uint64_t d2[2] = {f2_pack(a.x, a.y), f2_pack(a.z, a.w)};
        uint32_t wg[2], wu[2];
#pragma unroll
        for (int j = 0; j < 2; ++j) {
            const uint64_t g2 = f2_pack(bf16_lo(gw[j]), bf16_hi(gw[j]));
            const uint64_t u2 = f2_pack(bf16_lo(uw[j]), bf16_hi(uw[j]));
            uint64_t val, grad;
            glu_act_fast_both2(g2, k2, val, grad);
            float x0, x1;
            f2_unpack(f2_mul(f2_mul(d2[j], u2), grad), x0, x1);
            wg[j] = pack_bf16(x0, x1);
            f2_unpack(f2_mul(d2[j], val), x0, x1);
            wu[j] = pack_bf16(x0, x1);
        }
        if (ok) {
            uint8_t* cp = o.C + r * o.c_row + c * 2;
            *reinterpret_cast<uint2*>(cp) = make_uint2(wg[0], wg[1]);
            *reinterpret_cast<uint2*>(cp + o.ff_bytes) = make_uint2(wu[0], wu[1]);
        }
    }
    __syncwarp();
}

// Softmax backward fused into the dP = dO V^T product (HF eager attention under autograd): the accumulator
// chunk is dP; with the saved probabilities P (same [row, column] position) and delta[row] = sum_k P dP
// (= rowsum(dO * O), computed beforehand) the epilogue writes dS = alpha * P * (dP - delta) as bf16.
struct EpiSbwd {
    uint8_t* C;           // dS, byte address of (row 0 of the slab, column col0)
    const uint8_t* P;     // saved probabilities, same position
    const float* delta;   // delta of row 0 of the slab
    long long c_row, p_row;
    float alpha;
    int rows_valid, row_lo, cols_q;
};

__device__ __forceinline__ void epi_chunk_sbwd(const uint32_t (&v)[32], uint32_t stg, int lane, const EpiSbwd& o) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
        ptx::sts128(stg + lane * 128 + ((j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncwarp();
    const int ch = lane & 7, c = ch * 4, rsub = lane >> 3;
    const bool col_ok = c < o.cols_q;
    uint2 pq[8];
    float dl[8];
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int r = it * 4 + rsub;
        pq[it] = make_uint2(0u, 0u); dl[it] = 0.f;
        if (col_ok && r >= o.row_lo && r < o.rows_valid) {
            pq[it] = *reinterpret_cast<const uint2*>(o.P + r * o.p_row + c * 2);
            dl[it] = __ldg(o.delta + r);
        }
    }
    const uint32_t sp = stg + rsub * 128;
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int r = it * 4 + rsub;
        const float4 a = ptx::lds128(sp + it * 512 + ((ch ^ (r & 7)) << 4));
        if (col_ok && r >= o.row_lo && r < o.rows_valid) {
            const float p0 = bf16_lo(pq[it].x), p1 = bf16_hi(pq[it].x), p2 = bf16_lo(pq[it].y), p3 = bf16_hi(pq[it].y);
            uint2 w;
            // p == 0 marks masked positions, where the accumulator may be anything: force an exact zero
            w.x = pack_bf16(p0 != 0.f ? o.alpha * p0 * (a.x - dl[it]) : 0.f, p1 != 0.f ? o.alpha * p1 * (a.y - dl[it]) : 0.f);
            w.y = pack_bf16(p2 != 0.f ? o.alpha * p2 * (a.z - dl[it]) : 0.f, p3 != 0.f ? o.alpha * p3 * (a.w - dl[it]) : 0.f);
            *reinterpret_cast<uint2*>(o.C + r * o.c_row + c * 2) = w;
        }
    }
    __syncwarp();
}

// Same with P and delta already in registers: the caller loads the NEXT chunk's probabilities before it waits
// for the current chunk's accumulator, so their L2 latency is off the critical path (with K = head_dim the
// tile's MMAs are short and this epilogue is what the kernel waits for).
__device__ __forceinline__ void epi_chunk_sbwd_pre(const uint32_t (&v)[32], uint32_t stg, int lane, const EpiSbwd& o,
                                                   const uint2 (&pq)[8], const float (&dl)[8]) {
#pragma unroll
    for (int j = 0; j < 8; ++j)
        ptx::sts128(stg + lane * 128 + ((j ^ (lane & 7)) << 4), v[4 * j], v[4 * j + 1], v[4 * j + 2], v[4 * j + 3]);
    __syncwarp();
    const int ch = lane & 7, c = ch * 4, rsub = lane >> 3;
    const bool col_ok = c < o.cols_q;
    const uint32_t sp = stg + rsub * 128;
    const uint64_t al2 = f2_pack(o.alpha, o.alpha), one2 = f2_pack(1.f, 1.f);
#pragma unroll
    for (int it = 0; it < 8; ++it) {
        const int r = it * 4 + rsub;
        const float4 a = ptx::lds128(sp + it * 512 + ((ch ^ (r & 7)) << 4));
        if (col_ok && r >= o.row_lo && r < o.rows_valid) {
            // (alpha p) (dP - delta) on the packed-f32 pipe, two columns per instruction; p == 0 still selects 0
            const float p0 = bf16_lo(pq[it].x), p1 = bf16_hi(pq[it].x), p2 = bf16_lo(pq[it].y), p3 = bf16_hi(pq[it].y);
            const uint64_t nd = f2_pack(-dl[it], -dl[it]);
            float r0, r1, r2, r3;
            f2_unpack(f2_mul(f2_mul(al2, f2_pack(p0, p1)), f2_fma(f2_pack(a.x, a.y), one2, nd)), r0, r1);
            f2_unpack(f2_mul(f2_mul(al2, f2_pack(p2, p3)), f2_fma(f2_pack(a.z, a.w), one2, nd)), r2, r3);
            uint2 w;
            w.x = pack_bf16(p0 != 0.f ? r0 : 0.f, p1 != 0.f ? r1 : 0.f);
            w.y = pack_bf16(p2 != 0.f ? r2 : 0.f, p3 != 0.f ? r3 : 0.f);
            *reinterpret_cast<uint2*>(o.C + r * o.c_row + c * 2) = w;
        }
    }
    __syncwarp();
}

// ---- kernel ----------------------------------------------------------------------------------------
template <int BN, int ACT>
__global__ void __launch_bounds__(kGemm2Threads, 1)
gemm2_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                  const __grid_constant__ CUtensorMap tmAux, const __grid_constant__ CUtensorMap tmC,
                  const __grid_constant__ CUtensorMap tmA2, const __grid_constant__ CUtensorMap tmB2, const GemmDev p) {
    using Cfg = Gemm2Cfg<BN, ACT>;
    constexpr int STAGES = Cfg::STAGES;
    constexpr int BNH = BN / 2;                 // B rows staged by each CTA
    constexpr int TM = 2 * BM;                  // tile rows per cluster

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;
    uint8_t* smem_epi = smem + STAGES * Cfg::STAGE_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_epi + Cfg::EPI_BYTES);
    uint64_t* full_bar = bars;                       // leader's are the live ones
    uint64_t* empty_bar = bars + STAGES;             // per CTA (multicast commit)
    uint64_t* tfull_bar = bars + 2 * STAGES;         // per CTA (multicast commit)
    uint64_t* tempty_bar = bars + 2 * STAGES + 2;    // leader's are the live ones
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);
    [[maybe_unused]] uint64_t* gbwd_bar = bars + 2 * STAGES + 5;     // [kEpiWarps][2]: TMA gate / up tiles landed

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;
    const uint32_t rank = ptx::cluster_ctarank();
    const bool leader = rank == 0;

    const int mt = (p.M + TM - 1) / TM;
    const int nt = (p.N + BN - 1) / BN;
    const int tiles_per_z = mt * nt;
    // split-K (GemmDev::ksplit > 1): the tile list repeats once per K slice; slice s contracts over K blocks
    // [s * kb_per, min(nkb, (s + 1) * kb_per)) and its epilogue adds into C
    const int tiles_per_split = tiles_per_z * p.nz[0] * p.nz[1] * p.nz[2];
    const int total_tiles = tiles_per_split * (p.ksplit > 1 ? p.ksplit : 1);
    const int nkb = (p.K + BK - 1) / BK;
    // second contraction segment (GemmDev::K2): nkb2 more K blocks per tile, read through tmA2 / tmB2
    const int nkb2 = (p.K2 + BK - 1) / BK;
    const int k_iters = nkb * p.kg + nkb2;
    const int cid = static_cast<int>(ptx::cluster_id_x());
    const int ncl = static_cast<int>(ptx::num_clusters_x());

    pdl_trigger();                                // the next kernel's prologue may overlap this one (pt_common.cuh)
    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
        if (p.K2 > 0) { ptx::prefetch_tmap(&tmA2); ptx::prefetch_tmap(&tmB2); }
        if constexpr (tma_epilogue(ACT)) { ptx::prefetch_tmap(&tmAux); ptx::prefetch_tmap(&tmC); }
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full_bar[s], 2);      // leader's expect_tx arrive + peer's plain arrive
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&tfull_bar[s], 1);
            ptx::mbar_init(&tempty_bar[s], 2 * kEpiWarps);
        }
        if constexpr (tma_epilogue(ACT))
            for (int s = 0; s < 2 * kEpiWarps; ++s) ptx::mbar_init(&gbwd_bar[s], 1);
        ptx::fence_mbar_init();
    }
    ptx::cluster_sync_all();                      // peer barriers initialised before any remote signal
    if (warp == 2) {
        ptx::tmem_alloc_2sm(tmem_slot, 512);
        ptx::tmem_relinquish_2sm();
    }
    ptx::tc_fence_before();
    ptx::cluster_sync_all();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    pdl_wait();                                   // operands and C belong to the kernels before this one until here

    auto decode = [&](int t, int& m_blk, int& n_blk, int (&z)[3]) {
        if (p.ksplit > 1) t %= tiles_per_split;
        int zz = t / tiles_per_z;
        const int r = t - zz * tiles_per_z;
        if (p.raster_n) { m_blk = r / nt; n_blk = r - m_blk * nt; }
        else            { n_blk = r / mt; m_blk = r - n_blk * mt; }
        z[0] = zz % p.nz[0]; zz /= p.nz[0];
        z[1] = zz % p.nz[1]; zz /= p.nz[1];
        z[2] = zz;
    };

    if (warp == 0) {
        // ===================== TMA producer (both CTAs) =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = cid; t < total_tiles; t += ncl) {
                int m_blk, n_blk, z[3];
                decode(t, m_blk, n_blk, z);
                // tensor-map coordinates of the three extra dims, kept in registers (no indexed local
                // arrays in this loop: the producer is the pipeline's critical single thread)
                const int z0 = z[0], z1 = z[1], z2 = z[2];
                int ea0, ea1, ea2, eb0, eb1, eb2;
                if (p.kg_slot >= 0) {          // dims: (k-group, z0, z1)
                    ea0 = 0; ea1 = (p.a_emask & 2) ? z0 : 0; ea2 = (p.a_emask & 4) ? z1 : 0;
                    eb0 = 0; eb1 = (p.b_emask & 2) ? z0 : 0; eb2 = (p.b_emask & 4) ? z1 : 0;
                } else {                        // dims: (z0, z1, z2)
                    ea0 = (p.a_emask & 1) ? z0 : 0; ea1 = (p.a_emask & 2) ? z1 : 0; ea2 = (p.a_emask & 4) ? z2 : 0;
                    eb0 = (p.b_emask & 1) ? z0 : 0; eb1 = (p.b_emask & 2) ? z1 : 0; eb2 = (p.b_emask & 4) ? z2 : 0;
                }
                const bool kga = p.kg_slot >= 0 && (p.a_emask & 1), kgb = p.kg_slot >= 0 && (p.b_emask & 1);
                const int tile_m = m_blk == mt - 1 ? p.m_last : m_blk * TM;
                const int tile_n = n_blk == nt - 1 ? p.n_last : n_blk * BN;
                int kb_lo = 0, kb_hi = nkb;
                if (p.band && band_span(p, tile_m, tile_n, TM, BN, nkb, kb_lo, kb_hi)) continue;
                if (p.ksplit > 1) { kb_lo = (t / tiles_per_split) * p.kb_per; kb_hi = min(nkb, kb_lo + p.kb_per); }
                const int main_iters = (kb_hi - kb_lo) * p.kg;
                const int m0 = tile_m + static_cast<int>(rank) * BM;
                // fused GeGLU: the pair's tile is [128 gate rows | the 128 up rows of the same ff range]
                // half-width last n tile (GemmDev::n_half): a full in-bounds box whose second half holds this CTA's
                // BNH / 2 rows of B
                const bool half = p.n_half && n_blk == nt - 1;
                const int n0 = ACT == PT_ACT_GEGLU
                                   ? n_blk * BNH + static_cast<int>(rank) * p.ff
                                   : half ? tile_n - BNH / 2 + static_cast<int>(rank) * (BNH / 2)
                                          : tile_n + static_cast<int>(rank) * BNH;
                // one pipeline stage: wait for the slot, post the byte count, issue this CTA's A and B boxes
                auto fill = [&](const CUtensorMap* mA, const CUtensorMap* mB, int kc, int a0, int a1, int a2, int b0,
                                int b1, int b2) {
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    const uint32_t fb = ptx::mapa(ptx::smem_u32(&full_bar[stage]), 0);
                    if (leader) ptx::mbar_arrive_expect_tx(&full_bar[stage], 2 * Cfg::STAGE_BYTES);
                    uint8_t* sa = smem_a + stage * Cfg::A_BYTES;
                    uint8_t* sb = smem_b + stage * Cfg::B_BYTES;
                    if (!p.a_mn) {
                        ptx::tma_load_5d_2sm(mA, fb, sa, kc, m0, a0, a1, a2);
                    } else {
#pragma unroll
                        for (int c = 0; c < BM / 64; ++c)
                            ptx::tma_load_5d_2sm(mA, fb, sa + c * (BK * 128), m0 + c * 64, kc, a0, a1, a2);
                    }
                    if (!p.b_mn) {
                        ptx::tma_load_5d_2sm(mB, fb, sb, kc, n0, b0, b1, b2);
                    } else {
#pragma unroll
                        for (int c = 0; c < BNH / 64; ++c)
                            ptx::tma_load_5d_2sm(mB, fb, sb + c * (BK * 128), n0 + c * 64, kc, b0, b1, b2);
                    }
                    if (!leader) ptx::mbar_arrive_cluster(fb);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                };
                int g = 0, kb = kb_lo;
                for (int it = 0; it < main_iters; ++it) {
                    if (kga) ea0 = g;
                    if (kgb) eb0 = g;
                    fill(&tmA, &tmB, kb * BK, ea0, ea1, ea2, eb0, eb1, eb2);
                    if (++kb == kb_hi) { kb = kb_lo; ++g; }
                }
                // K blocks past the main contraction come from the second operand pair (no batch dims)
                for (int it = 0; it < nkb2; ++it) fill(&tmA2, &tmB2, it * BK, 0, 0, 0, 0, 0, 0);
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer (leader CTA only) =====================
        if (leader && lane == 0) {
            const uint32_t idesc = ptx::make_idesc_bf16(TM, BN, p.a_mn, p.b_mn);
            const uint32_t a_lbo = p.a_mn ? BK * 128 : 0, b_lbo = p.b_mn ? BK * 128 : 0;
            const uint32_t a_kstep = p.a_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
            const uint32_t b_kstep = p.b_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            int last_acc = -1;
            uint32_t last_phase = 0;
            const uint32_t idesc_half = ptx::make_idesc_bf16(TM, BN / 2, p.a_mn, p.b_mn);
            for (int t = cid; t < total_tiles; t += ncl) {
                int tile_iters = k_iters;
                if (p.ksplit > 1) {
                    const int lo = (t / tiles_per_split) * p.kb_per;
                    tile_iters = min(nkb, lo + p.kb_per) - lo;
                }
                bool half = false;
                if (p.n_half) {
                    int m_blk, n_blk, z[3];
                    decode(t, m_blk, n_blk, z);
                    half = n_blk == nt - 1;
                }
                // the B rows of a half tile are the second half of the staged box: 64 rows x 128 B (K-major) or the
                // second 64-column chunk, BK x 128 B (MN-major) — 8 KB further on either way
                const uint32_t b_skip = half ? (BNH / 2) * 128 : 0;
                const uint32_t idesc_t = half ? idesc_half : idesc;
                if (p.band) {
                    int m_blk, n_blk, z[3];
                    decode(t, m_blk, n_blk, z);
                    int kb_lo, kb_hi;
                    if (band_span(p, m_blk == mt - 1 ? p.m_last : m_blk * TM, n_blk == nt - 1 ? p.n_last : n_blk * BN,
                                  TM, BN, nkb, kb_lo, kb_hi)) continue;
                    tile_iters = (kb_hi - kb_lo) * p.kg + nkb2;
                }
                ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * Cfg::ACC_STRIDE;
                for (int it = 0; it < tile_iters; ++it) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t adesc = ptx::make_smem_desc_sw128(
                        ptx::smem_u32(smem_a + stage * Cfg::A_BYTES), a_lbo, 1024);
                    const uint64_t bdesc = ptx::make_smem_desc_sw128(
                        ptx::smem_u32(smem_b + stage * Cfg::B_BYTES) + b_skip, b_lbo, 1024);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        ptx::mma_bf16_ss_2sm(d_tmem, adesc + k * a_kstep, bdesc + k * b_kstep, idesc_t,
                                             (it | k) != 0);
                    }
                    ptx::mma_commit_2sm_mc(&empty_bar[stage], 3);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit_2sm_mc(&tfull_bar[acc], 3);
                last_acc = acc; last_phase = acc_phase;
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
            // the peer's epilogue arrives remotely on our tempty barriers: drain before teardown
            if (last_acc >= 0) ptx::mbar_wait(&tempty_bar[last_acc], last_phase);
        }
    } else if (warp >= 4) {
        // ===================== epilogue (both CTAs, rows of this CTA's half) =====================
        const int ew = warp - 4;
        const int q = warp & 3;                 // TMEM lane quarter this warp may read
        const int hf = ew >> 2;                 // which half of the BN columns
        const uint32_t stg = ptx::smem_u32(smem_epi + ew * Cfg::EPI_WARP_BYTES);
        int acc = 0;
        uint32_t acc_phase = 0;
        [[maybe_unused]] uint32_t gb_phase = 0;      // bit b: parity of this warp's TMA buffer b
        [[maybe_unused]] bool gb_pref = false;       // chunk 0 of the tile at hand has been requested already
        for (int t = cid; t < total_tiles; t += ncl) {
            int m_blk, n_blk, z[3];
            decode(t, m_blk, n_blk, z);
            const long long coff = z[0] * p.c_zs[0] + z[1] * p.c_zs[1] + z[2] * p.c_zs[2];
            const long long roff = z[0] * p.r_zs[0] + z[1] * p.r_zs[1] + z[2] * p.r_zs[2];
            const int tile_m = m_blk == mt - 1 ? p.m_last : m_blk * TM;
            const int tile_n = n_blk == nt - 1 ? p.n_last : n_blk * BN;
            if (p.band == PT_BAND_SCORES) {
                int kb_lo, kb_hi;
                if (band_span(p, tile_m, tile_n, TM, BN, nkb, kb_lo, kb_hi)) continue;
            }
            const int row0 = tile_m + static_cast<int>(rank) * BM + q * 32;        // first row of the slab
            const int rows_valid = min(32, p.M - row0);                            // may be <= 0
            const int row_lo = max(0, m_blk * TM - row0);                          // rows below: previous tile's

            if constexpr (ACT == kActGbwdLsu) {
                // pull this warp's slab of the saved gate / up pre-activations into L2 while the tile's MMAs run
                const int pr = row0 + lane;
                if (pr < p.M && pr >= m_blk * TM) {
                    const uint8_t* g = static_cast<const uint8_t*>(p.aux) +
                                       (static_cast<long long>(pr) * p.ld_aux + tile_n + hf * BNH) * 2;
#pragma unroll
                    for (int b = 0; b < BNH * 2; b += 128) {
                        if (tile_n + hf * BNH + b / 2 < p.N) {
                            ptx::prefetch_l2(g + b);
                            ptx::prefetch_l2(g + static_cast<long long>(p.ff) * 2 + b);
                        }
                    }
                }
            }
            if constexpr (ACT == PT_ACT_SOFTMAX_BWD) {
                // same for the saved probabilities this warp's slab will be multiplied with
                const int pr = row0 + lane;
                if (pr < p.M && pr >= m_blk * TM) {
                    const uint8_t* g = static_cast<const uint8_t*>(p.aux) +
                                       (coff / p.ldc * p.ld_aux + static_cast<long long>(pr) * p.ld_aux + tile_n + hf * BNH) * 2;
#pragma unroll
                    for (int b = 0; b < BNH * 2; b += 128)
                        if (tile_n + hf * BNH + b / 2 < p.N) ptx::prefetch_l2(g + b);
                }
            }
            constexpr int NCH = BNH / 32;
            // softmax backward: this lane's delta values and the first chunk's probabilities, fetched while the
            // tile's MMAs are still running
            [[maybe_unused]] uint2 sb_pq[8];
            [[maybe_unused]] float sb_dl[8];
            [[maybe_unused]] const uint8_t* sb_P = nullptr;
            [[maybe_unused]] auto sb_load = [&](int c, uint2 (&dst)[8]) {
                const int col0 = tile_n + hf * BNH + c * 32;
                const int cols_q = min(32, ((p.N + 3) & ~3) - col0);
                const int cq = (lane & 7) * 4, rsub = lane >> 3;
                const bool col_live = col0 < p.N && col0 >= n_blk * BN && cq < cols_q;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int r = it * 4 + rsub;
                    dst[it] = make_uint2(0u, 0u);
                    if (col_live && r >= row_lo && r < rows_valid)
                        dst[it] = *reinterpret_cast<const uint2*>(sb_P + static_cast<long long>(r) * p.ld_aux * 2 +
                                                                  (c * 32 + cq) * 2);
                }
            };
            if constexpr (ACT == PT_ACT_SOFTMAX_BWD) {
                sb_P = static_cast<const uint8_t*>(p.aux) +
                       (coff / p.ldc * p.ld_aux + static_cast<long long>(row0) * p.ld_aux + tile_n + hf * BNH) * 2;
                const float* dbase = p.bias + coff / p.ldc + row0;
                const int rsub = lane >> 3;
#pragma unroll
                for (int it = 0; it < 8; ++it) {
                    const int r = it * 4 + rsub;
                    sb_dl[it] = (r >= row_lo && r < rows_valid) ? __ldg(dbase + r) : 0.f;
                }
                sb_load(0, sb_pq);
            }
            if constexpr (tma_epilogue(ACT)) {
                // ---- Epilogues with TMA-moved tiles (GeGLU backward; erf-GELU backward; attention delta).  The
                // accumulator chunk (lane = row, 32 columns) is used in the layout tcgen05.ld delivers; the saved
                // tile(s) of the same 32 x 32 block (gate and up / Z / O) arrive by TMA in a 64-byte-swizzled buffer
                // (conflict-free 16-byte row reads), the results are written over them and leave by TMA store.  No
                // transpose through smem, no global LD/ST instructions, no per-row address arithmetic: about half
                // the instructions of the load/store-unit version of the GeGLU backward (165 -> 114 registers).
                constexpr int NCHg = BNH / 32;
                constexpr int NT = ACT == PT_ACT_GEGLU_BWD ? 2 : 1;         // tiles per chunk (in and out)
                const bool live = row0 < p.M;                               // warp-uniform
                uint8_t* wst = smem_epi + ew * Cfg::EPI_WARP_BYTES;
                uint64_t* gb = gbwd_bar + 2 * ew;
                auto request = [&](int buf, int r0, int c0) {               // lane 0: saved tile(s) -> buffer buf
                    ptx::mbar_arrive_expect_tx(&gb[buf], NT * (kGbwdBufBytes / 2));
                    ptx::tma_load_2d(&tmAux, &gb[buf], wst + buf * kGbwdBufBytes, c0, r0);
                    if constexpr (NT == 2)
                        ptx::tma_load_2d(&tmAux, &gb[buf], wst + buf * kGbwdBufBytes + kGbwdBufBytes / 2, c0 + p.ff, r0);
                };
                const int colb = tile_n + hf * BNH;
                if (live && !gb_pref && lane == 0) {
                    ptx::bulk_wait_read<0>();
                    request(0, row0, colb);
                }
                ptx::mbar_wait(&tfull_bar[acc], acc_phase);
                ptx::tc_fence_after();
                const uint32_t taddr = tmem_base + acc * Cfg::ACC_STRIDE + (static_cast<uint32_t>(q * 32) << 16);
                [[maybe_unused]] const GluConsts2 k2 = glu_consts2(p.glu);
                [[maybe_unused]] float dsum = 0.f;
                const uint32_t sw = (lane >> 1) & 3;                        // SWIZZLE_64B: chunk ^= (row >> 1) & 3
#pragma unroll 1
                for (int c = 0; c < NCHg; ++c) {
                    const int buf = c & 1;
                    uint32_t v[32];
                    if (live) {
                        ptx::tmem_ld_32x32(taddr + hf * BNH + c * 32, v);
                        if (lane == 0 && c + 1 < NCHg) {
                            ptx::bulk_wait_read<0>();                       // chunk c-1's store has read buffer buf^1
                            request(buf ^ 1, row0, colb + (c + 1) * 32);
                        }
                        ptx::mbar_wait(&gb[buf], (gb_phase >> buf) & 1u);
                        gb_phase ^= 1u << buf;
                        ptx::tmem_ld_wait();
                    }
                    if (c == NCHg - 1) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&tempty_bar[acc]), 0));
                    }
                    if (!live) continue;
                    const uint32_t ga = ptx::smem_u32(wst + buf * kGbwdBufBytes) + lane * 64, ua = ga + kGbwdBufBytes / 2;
#pragma unroll
                    for (int j = 0; j < 4; ++j) {                           // 16-byte chunk j = columns 8 j .. 8 j + 7
                        const uint32_t off = (static_cast<uint32_t>(j) ^ sw) << 4;
                        const float4 gq = ptx::lds128(ga + off);
                        const uint32_t gw[4] = {__float_as_uint(gq.x), __float_as_uint(gq.y), __float_as_uint(gq.z), __float_as_uint(gq.w)};
                        uint32_t og[4];
                        if constexpr (ACT == PT_ACT_GEGLU_BWD) {
                            const float4 uq = ptx::lds128(ua + off);
                            const uint32_t uw[4] = {__float_as_uint(uq.x), __float_as_uint(uq.y), __float_as_uint(uq.z), __float_as_uint(uq.w)};
                            uint32_t ou[4];
#pragma unroll
                            for (int w = 0; w < 4; ++w) {
                                const uint64_t g2 = f2_pack(bf16_lo(gw[w]), bf16_hi(gw[w]));
                                const uint64_t u2 = f2_pack(bf16_lo(uw[w]), bf16_hi(uw[w]));
                                const uint64_t d2 = f2_pack(__uint_as_float(v[8 * j + 2 * w]), __uint_as_float(v[8 * j + 2 * w + 1]));
                                uint64_t val, grad;
#ifdef PT_GBWD_CHEAP            // probe build (tools/): the epilogue's floor without the activation arithmetic
                                val = g2; grad = g2;
#else
                                glu_act_fast_both2(g2, k2, val, grad);
#endif
                                float x0, x1;
                                f2_unpack(f2_mul(f2_mul(d2, u2), grad), x0, x1);
                                og[w] = pack_bf16(x0, x1);
                                f2_unpack(f2_mul(d2, val), x0, x1);
                                ou[w] = pack_bf16(x0, x1);
                            }
                            ptx::sts128(ua + off, ou[0], ou[1], ou[2], ou[3]);
                        } else if constexpr (ACT == PT_ACT_GELU_ERF_BWD) {
                            // dZ = dH * gelu_erf'(Z): backward of nn.GELU() (ref:Stage1/projectors.py:18), exact erf
#pragma unroll
                            for (int w = 0; w < 4; ++w)
                                og[w] = pack_bf16(__uint_as_float(v[8 * j + 2 * w]) * gelu_erf_grad_f(bf16_lo(gw[w])),
                                                  __uint_as_float(v[8 * j + 2 * w + 1]) * gelu_erf_grad_f(bf16_hi(gw[w])));
                        } else {
                            // dO tile as bf16 + this row's share of delta = sum_d dO[d] * O[d] (on the rounded dO, as
                            // pt_attn_delta computes it)
#pragma unroll
                            for (int w = 0; w < 4; ++w) {
                                og[w] = pack_bf16(__uint_as_float(v[8 * j + 2 * w]), __uint_as_float(v[8 * j + 2 * w + 1]));
                                dsum = fmaf(bf16_lo(og[w]), bf16_lo(gw[w]), dsum);
                                dsum = fmaf(bf16_hi(og[w]), bf16_hi(gw[w]), dsum);
                            }
                        }
                        ptx::sts128(ga + off, og[0], og[1], og[2], og[3]);
                    }
                    ptx::fence_proxy_async_smem();
                    __syncwarp();
                    if (lane == 0) {
                        ptx::tma_store_2d(&tmC, wst + buf * kGbwdBufBytes, colb + c * 32, row0);
                        if constexpr (NT == 2)
                            ptx::tma_store_2d(&tmC, wst + buf * kGbwdBufBytes + kGbwdBufBytes / 2, colb + c * 32 + p.ff, row0);
                        ptx::bulk_commit();
                    }
                }
                if constexpr (ACT == PT_ACT_ATTN_DELTA) {
                    // this warp covered BNH columns of one head (hd % BNH == 0): add the row's partial to delta[b, head, s]
                    const int r = row0 + lane;
                    if (live && r < p.M && r >= m_blk * TM) {               // (rows a shifted last tile repeats: counted once)
                        const int bb = r / p.dl_S, ss = r - bb * p.dl_S, head = colb / p.dl_hd;
                        atomicAdd(p.delta + (static_cast<long long>(bb) * p.dl_hq + head) * p.dl_S + ss, dsum);
                    }
                }
                // request the next tile's first chunk now: its L2 latency then overlaps that tile's hand-over
                gb_pref = false;
                if (t + ncl < total_tiles) {
                    int m2, n2, z2[3];
                    decode(t + ncl, m2, n2, z2);
                    const int r2 = (m2 == mt - 1 ? p.m_last : m2 * TM) + static_cast<int>(rank) * BM + q * 32;
                    const int c2 = (n2 == nt - 1 ? p.n_last : n2 * BN) + hf * BNH;
                    if (r2 < p.M) {
                        gb_pref = true;
                        if (lane == 0) {
                            // buffer 0 was last read by the store of the last even chunk: the most recent group
                            // (odd last chunk -> buffer 1) may still be pending, nothing older
                            if (((NCHg - 1) & 1) && live) ptx::bulk_wait_read<1>(); else ptx::bulk_wait_read<0>();
                            request(0, r2, c2);
                        }
                    }
                }
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
                continue;
            }
            ptx::mbar_wait(&tfull_bar[acc], acc_phase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + acc * Cfg::ACC_STRIDE + (static_cast<uint32_t>(q * 32) << 16);
            if constexpr (ACT == PT_ACT_GEGLU) {
                // accumulator columns [0,128) = gate, [128,256) = up of ff columns n_blk*128 .. +127; this warp
                // produces two 32-column chunks: gate and up pre-activations -> C, gelu_tanh(gate)*up -> aux
                const bool live = rows_valid > row_lo;                                   // warp-uniform
                const GluConsts2 k2 = glu_consts2(p.glu);
#pragma unroll 1
                for (int c2 = 0; c2 < 2; ++c2) {
                    const int cg = (hf * 2 + c2) * 32;
                    uint32_t vg[32], vu[32];
                    if (live) {
                        ptx::tmem_ld_32x32(taddr + cg, vg);
                        ptx::tmem_ld_32x32(taddr + BNH + cg, vu);
                        ptx::tmem_ld_wait();
                    }
                    if (c2 == 1) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&tempty_bar[acc]), 0));
                    }
                    if (!live) continue;
                    const int colg = n_blk * BNH + cg;
                    EpiArgs o;
                    o.Cpre = nullptr; o.R = nullptr; o.bias = nullptr;
                    o.p_row = 0; o.p_step = 0; o.r_row = 0; o.r_step = 0;
                    o.c_f32 = 0; o.r_f32 = 0; o.alpha = 1.f; o.c_atomic = 0;
                    o.rows_valid = rows_valid; o.row_lo = row_lo; o.cols_q = 32;
                    o.c_row = p.ldc * 2; o.c_step = 4 * o.c_row;
                    uint8_t* crow = static_cast<uint8_t*>(p.C) + (coff + static_cast<long long>(row0) * p.ldc) * 2;
                    o.C = crow + static_cast<long long>(colg) * 2;
                    epi_chunk<PT_ACT_NONE>(vg, stg, lane, o);
                    // on the packed-f32 pipe (FFMA2 / FMUL2): two columns per instruction
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        const uint64_t g2 = round_bf16_pair(__uint_as_float(vg[j]), __uint_as_float(vg[j + 1]));
                        const uint64_t u2 = round_bf16_pair(__uint_as_float(vu[j]), __uint_as_float(vu[j + 1]));
                        float a0, a1;
                        f2_unpack(f2_mul(glu_act_fast2(g2, k2), u2), a0, a1);
                        vg[j] = __float_as_uint(a0); vg[j + 1] = __float_as_uint(a1);
                    }
                    o.C = crow + static_cast<long long>(p.ff + colg) * 2;
                    epi_chunk<PT_ACT_NONE>(vu, stg, lane, o);
                    o.c_row = p.ld_aux * 2; o.c_step = 4 * o.c_row;
                    o.C = static_cast<uint8_t*>(p.aux) + (static_cast<long long>(row0) * p.ld_aux + colg) * 2;
                    epi_chunk<PT_ACT_NONE>(vg, stg, lane, o);
                }
            } else if constexpr (ACT == PT_ACT_QKNORM_ROPE) {
                // The 256-wide tile n_blk is head n_blk of the fused q|k|v row (q heads, then k heads, then v heads).
                // A warp owns 32 rows (TMEM lanes) and may read every column of them: it sums the squares of the whole
                // head itself (both warps of a row quadrant do: no exchange), then, like the GeGLU epilogue, takes two
                // 32-column chunks of the first half together with their rotary partners 128 columns further on.
                const bool live = rows_valid > row_lo;                                   // warp-uniform
                const int hq_ = p.dl_hq, hkv_ = (p.N / 256 - hq_) / 2;
                const bool rot = n_blk < hq_ + hkv_;
                float rstd = 0.f;
                if (live && rot) {
                    float ssq = 0.f;
#pragma unroll 1
                    for (int c8 = 0; c8 < 8; ++c8) {
                        uint32_t v[32];
                        ptx::tmem_ld_32x32(taddr + c8 * 32, v);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int j = 0; j < 32; ++j) {
                            const float x = round_bf16(__uint_as_float(v[j]));   // what the stored projection holds
                            ssq = fmaf(x, x, ssq);
                        }
                    }
                    rstd = rsqrtf(ssq * (1.f / 256.f) + p.qn_eps);
                }
                const float* gw = p.bias + (n_blk < hq_ ? 0 : 256);
                const float* cs = static_cast<const float*>(p.R) + static_cast<long long>((row0 + lane) % p.dl_S) * 256;
#pragma unroll 1
                for (int c2 = 0; c2 < 2; ++c2) {
                    const int cg = (hf * 2 + c2) * 32;
                    uint32_t v1[32], v2[32];
                    if (live) {
                        ptx::tmem_ld_32x32(taddr + cg, v1);
                        ptx::tmem_ld_32x32(taddr + BNH + cg, v2);
                        ptx::tmem_ld_wait();
                    }
                    if (c2 == 1) {
                        ptx::tc_fence_before();
                        __syncwarp();
                        if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&tempty_bar[acc]), 0));
                    }
                    if (!live) continue;
                    EpiArgs o;
                    o.Cpre = nullptr; o.R = nullptr; o.bias = nullptr;
                    o.p_row = 0; o.p_step = 0; o.r_row = 0; o.r_step = 0;
                    o.c_f32 = 0; o.r_f32 = 0; o.alpha = 1.f; o.c_atomic = 0;
                    o.rows_valid = rows_valid; o.row_lo = row_lo; o.cols_q = 32;
                    o.c_row = p.ldc * 2; o.c_step = 4 * o.c_row;
                    uint8_t* crow = static_cast<uint8_t*>(p.C) + (coff + static_cast<long long>(row0) * p.ldc) * 2;
                    o.C = crow + static_cast<long long>(n_blk * 256 + cg) * 2;
                    epi_chunk<PT_ACT_NONE>(v1, stg, lane, o);
                    o.C = crow + static_cast<long long>(n_blk * 256 + BNH + cg) * 2;
                    epi_chunk<PT_ACT_NONE>(v2, stg, lane, o);
                    if (!rot) continue;
#pragma unroll
                    for (int j = 0; j < 32; j += 4) {
                        const float4 c4 = *reinterpret_cast<const float4*>(cs + cg + j);
                        const float4 s4 = *reinterpret_cast<const float4*>(cs + 128 + cg + j);
                        const float4 g1 = *reinterpret_cast<const float4*>(gw + cg + j);
                        const float4 g2 = *reinterpret_cast<const float4*>(gw + 128 + cg + j);
                        const float cc[4] = {c4.x, c4.y, c4.z, c4.w}, ss[4] = {s4.x, s4.y, s4.z, s4.w};
                        const float w1[4] = {g1.x, g1.y, g1.z, g1.w}, w2[4] = {g2.x, g2.y, g2.z, g2.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            const float n1 = round_bf16(__uint_as_float(v1[j + e])) * rstd * (1.f + w1[e]);
                            const float n2 = round_bf16(__uint_as_float(v2[j + e])) * rstd * (1.f + w2[e]);
                            v1[j + e] = __float_as_uint(n1 * cc[e] - n2 * ss[e]);
                            v2[j + e] = __float_as_uint(n2 * cc[e] + n1 * ss[e]);
                        }
                    }
                    const bool isq = n_blk < hq_;
                    const long long ldo = isq ? p.ld_aux : static_cast<long long>(hkv_) * 256;
                    uint8_t* orow = (isq ? static_cast<uint8_t*>(p.aux) : reinterpret_cast<uint8_t*>(p.delta)) +
                                    (static_cast<long long>(row0) * ldo + (isq ? n_blk : n_blk - hq_) * 256) * 2;
                    o.c_row = ldo * 2; o.c_step = 4 * o.c_row;
                    o.C = orow + static_cast<long long>(cg) * 2;
                    epi_chunk<PT_ACT_NONE>(v1, stg, lane, o);
                    o.C = orow + static_cast<long long>(BNH + cg) * 2;
                    epi_chunk<PT_ACT_NONE>(v2, stg, lane, o);
                }
            } else {
            [[maybe_unused]] float ce_m = -INFINITY, ce_s = 0.f;       // PT_ACT_CE_STATS: this lane's row, this half tile
#pragma unroll 1
            for (int c = 0; c < NCH; ++c) {
                const int col_local = hf * BNH + c * 32;
                const int col0 = tile_n + col_local;
                const bool live = rows_valid > row_lo && col0 < p.N && col0 >= n_blk * BN;   // warp-uniform
                uint32_t v[32];
                if (live) ptx::tmem_ld_32x32(taddr + col_local, v);
                [[maybe_unused]] uint2 sb_pn[8];
                if constexpr (ACT == PT_ACT_SOFTMAX_BWD) {
                    if (c + 1 < NCH) sb_load(c + 1, sb_pn);          // next chunk's P: in flight during this chunk
                }
                if (live) ptx::tmem_ld_wait();
                if (c == NCH - 1) {
                    // accumulator stage fully read by this warp: hand it back to the leader's MMA warp
                    ptx::tc_fence_before();
                    __syncwarp();
                    if (lane == 0) ptx::mbar_arrive_cluster(ptx::mapa(ptx::smem_u32(&tempty_bar[acc]), 0));
                }
                if (!live) {
                    if constexpr (ACT == PT_ACT_SOFTMAX_BWD) {
#pragma unroll
                        for (int it = 0; it < 8; ++it) sb_pq[it] = sb_pn[it];
                    }
                    continue;
                }
                if constexpr (ACT == PT_ACT_CE_STATS) {
                    // running maximum and sum of exponentials over the chunk's columns, from the logits as they
                    // will be stored (rounded to bf16); lane = row, so no exchange between lanes is needed
                    const int ncol = min(32, p.N - col0);
                    float x[32], cm = -INFINITY;
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        x[j] = __bfloat162float(__float2bfloat16_rn(__uint_as_float(v[j])));
                        if (j < ncol) cm = fmaxf(cm, x[j]);
                    }
                    const float mn = fmaxf(ce_m, cm);
                    const float k = 1.4426950408889634f, mk = -mn * k;
                    float s0 = 0.f, s1 = 0.f;
#pragma unroll
                    for (int j = 0; j < 32; j += 2) {
                        if (j < ncol) s0 += exp2f(fmaf(x[j], k, mk));
                        if (j + 1 < ncol) s1 += exp2f(fmaf(x[j + 1], k, mk));
                    }
                    ce_s = ce_s * exp2f((ce_m - mn) * k) + (s0 + s1);
                    ce_m = mn;
                }
                if constexpr (ACT == PT_ACT_SOFTMAX_BWD) {
                    EpiSbwd o;
                    const long long co = coff + static_cast<long long>(row0) * p.ldc + col0;
                    o.c_row = p.ldc * 2; o.p_row = p.ld_aux * 2;
                    o.C = static_cast<uint8_t*>(p.C) + co * 2;
                    o.P = static_cast<const uint8_t*>(p.aux) +
                          (coff / p.ldc * p.ld_aux + static_cast<long long>(row0) * p.ld_aux + col0) * 2;
                    o.delta = p.bias + coff / p.ldc + row0;
                    o.alpha = p.alpha;
                    o.rows_valid = rows_valid; o.row_lo = row_lo;
                    o.cols_q = min(32, ((p.N + 3) & ~3) - col0);
                    epi_chunk_sbwd_pre(v, stg, lane, o, sb_pq, sb_dl);
#pragma unroll
                    for (int it = 0; it < 8; ++it) sb_pq[it] = sb_pn[it];
                    continue;
                }
                if constexpr (ACT == kActGbwdLsu) {
                    EpiGbwd o;
                    o.c_row = p.ldc * 2; o.g_row = p.ld_aux * 2; o.ff_bytes = static_cast<long long>(p.ff) * 2;
                    o.C = static_cast<uint8_t*>(p.C) + (static_cast<long long>(row0) * p.ldc + col0) * 2;
                    o.GU = static_cast<const uint8_t*>(p.aux) + (static_cast<long long>(row0) * p.ld_aux + col0) * 2;
                    o.rows_valid = rows_valid; o.row_lo = row_lo;
                    o.cols_q = min(32, ((p.N + 3) & ~3) - col0);
                    o.glu = p.glu;
                    epi_chunk_gbwd(v, stg, lane, o);
                    continue;
                }
                EpiArgs o;
                const long long co = coff + static_cast<long long>(row0) * p.ldc + col0;
                const int ce = p.c_f32 ? 4 : 2, re = p.r_f32 ? 4 : 2;
                o.C = static_cast<uint8_t*>(p.C) + co * ce;
                o.c_row = p.ldc * ce; o.c_step = 4 * o.c_row;
                o.Cpre = p.C_pre ? static_cast<uint8_t*>(p.C_pre) + co * 2 : nullptr;
                o.p_row = p.ldc * 2; o.p_step = 4 * o.p_row;
                o.R = p.R ? static_cast<const uint8_t*>(p.R) +
                                (roff + static_cast<long long>(row0) * p.ldr + col0) * re : nullptr;
                o.r_row = p.ldr * re; o.r_step = 4 * o.r_row;
                o.bias = p.bias ? p.bias + col0 : nullptr;
                o.c_f32 = p.c_f32; o.r_f32 = p.r_f32; o.alpha = p.alpha;
                o.c_atomic = p.ksplit > 1;
                o.rows_valid = rows_valid; o.row_lo = row_lo;
                o.cols_q = min(32, ((p.N + 3) & ~3) - col0);
                epi_chunk<ACT>(v, stg, lane, o);
            }
            if constexpr (ACT == PT_ACT_CE_STATS) {
                // one (max, sum) pair per row and half tile, part-major so that a warp's 32 rows are contiguous
                if (lane >= row_lo && lane < rows_valid) {
                    const long long part = static_cast<long long>(n_blk) * 2 + hf;
                    reinterpret_cast<float2*>(p.delta)[part * p.M + row0 + lane] = make_float2(ce_m, ce_s);
                }
            }
            }   // !GEGLU
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    if constexpr (tma_epilogue(ACT)) {
        if (warp >= 4 && lane == 0) ptx::bulk_wait_all<0>();     // this warp's TMA stores have been performed
    }
    ptx::tc_fence_before();
    ptx::cluster_sync_all();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc_2sm(tmem_base, 512);
    }
}

// ---- host ----------------------------------------------------------------------------------------
template <int BN, int ACT>
static int launch_gemm2_t(const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& t_aux,
                          const CUtensorMap& t_c, const CUtensorMap& ta2, const CUtensorMap& tb2, const GemmDev& p,
                          int tiles, cudaStream_t st) {
    using Cfg = Gemm2Cfg<BN, ACT>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(gemm2_bf16_kernel<BN, ACT>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) {
            set_error("cudaFuncSetAttribute(gemm2<%d>): %s", BN, cudaGetErrorString(e));
            return PT_ERR_CUDA;
        }
        attr_done = true;
    }
    const int max_clusters = num_sms() / 2;
    const int clusters = tiles < max_clusters ? tiles : max_clusters;
    cudaLaunchConfig_t cfg;
    memset(&cfg, 0, sizeof(cfg));
    cfg.gridDim = dim3(2 * clusters, 1, 1);
    cfg.blockDim = dim3(kGemm2Threads, 1, 1);
    cfg.dynamicSmemBytes = Cfg::SMEM_BYTES;
    cfg.stream = st;
    cudaLaunchAttribute attr[2];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 2;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    attr[1].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[1].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    // (the delta epilogue's launch follows a memset of its output, not a kernel: launched the ordinary way)
    cfg.numAttrs = pdl_enabled(kPdlGemm) && ACT != PT_ACT_ATTN_DELTA ? 2 : 1;
    cudaError_t e = cudaLaunchKernelEx(&cfg, gemm2_bf16_kernel<BN, ACT>, ta, tb, t_aux, t_c, ta2, tb2, p);
    if (e != cudaSuccess) {
        set_error("gemm2_bf16_kernel<%d>: launch failed: %s", BN, cudaGetErrorString(e));
        return PT_ERR_CUDA;
    }
    return PT_OK;
}

int gemm2_tiles(const GemmDev& p, int bn) {
    const long long mt = (p.M + 2 * BM - 1) / (2 * BM), nt = (p.N + bn - 1) / bn;
    const long long t = mt * nt * p.nz[0] * p.nz[1] * p.nz[2] * (p.ksplit > 1 ? p.ksplit : 1);
    return t < (1ll << 31) ? static_cast<int>(t) : -1;
}

// PT_GEMM_GBWD_TMA=0: the first-generation (load/store unit) GeGLU-backward epilogue, for A/B probes
static int gbwd_tma_enabled() {
    static const int v = [] { const char* e = getenv("PT_GEMM_GBWD_TMA"); return !(e && e[0] == '0'); }();
    return v;
}

int launch_gemm2(int bn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& t_aux, const CUtensorMap& t_c,
                 const CUtensorMap& ta2, const CUtensorMap& tb2, const GemmDev& p, cudaStream_t st) {
    const int tiles = gemm2_tiles(p, bn);
    if (tiles <= 0) { set_error("gemm2: bad tile count"); return PT_ERR_INVALID; }
#define PT_G2(BN_)                                                                      \
    case BN_:                                                                           \
        if (p.act == PT_ACT_GELU_TANH) return launch_gemm2_t<BN_, PT_ACT_GELU_TANH>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st); \
        if (p.act == PT_ACT_GELU_ERF) return launch_gemm2_t<BN_, PT_ACT_GELU_ERF>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st);   \
        if (p.act == PT_ACT_GEGLU_BWD)                                                                      \
            return gbwd_tma_enabled() ? launch_gemm2_t<BN_, PT_ACT_GEGLU_BWD>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st) \
                                      : launch_gemm2_t<BN_, kActGbwdLsu>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st);  \
        if (p.act == PT_ACT_SOFTMAX_BWD) return launch_gemm2_t<BN_, PT_ACT_SOFTMAX_BWD>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st); \
        if (p.act == PT_ACT_GELU_ERF_BWD) return launch_gemm2_t<BN_, PT_ACT_GELU_ERF_BWD>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st); \
        if (p.act == PT_ACT_ATTN_DELTA) return launch_gemm2_t<BN_, PT_ACT_ATTN_DELTA>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st); \
        return launch_gemm2_t<BN_, PT_ACT_NONE>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st);
    if (p.act == PT_ACT_CE_STATS) {
        if (bn != 256) { set_error("gemm2: the CE-statistics epilogue needs BN = 256"); return PT_ERR_INVALID; }
        return launch_gemm2_t<256, PT_ACT_CE_STATS>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st);
    }
    if (p.act == PT_ACT_QKNORM_ROPE) {
        if (bn != 256) { set_error("gemm2: the q/k-norm + RoPE epilogue needs BN = 256"); return PT_ERR_INVALID; }
        return launch_gemm2_t<256, PT_ACT_QKNORM_ROPE>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st);
    }
    if (p.act == PT_ACT_GEGLU) {
        if (bn != 256) { set_error("gemm2: the fused GeGLU epilogue needs BN = 256"); return PT_ERR_INVALID; }
        return launch_gemm2_t<256, PT_ACT_GEGLU>(ta, tb, t_aux, t_c, ta2, tb2, p, tiles, st);
    }
    switch (bn) {
        PT_G2(256) PT_G2(192) PT_G2(128) PT_G2(64)
        default: set_error("gemm2: unsupported BN %d", bn); return PT_ERR_INVALID;
    }
#undef PT_G2
}

}  // namespace pt
