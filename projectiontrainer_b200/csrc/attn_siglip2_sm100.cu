// Fused non-causal attention forward for the SigLIP tower, second generation (head_dim 64):
//     O[b, q, h, :] = softmax(Q K^T * scale) V        HF:models/siglip/modeling_siglip.py:293-306
//
// (A first version kept K/V of a whole (image, head) resident in smem: 224 KB, hence one CTA — one softmax
// warpgroup — per SM; tensor pipe and softmax warps took turns at ~190 TFLOP/s.)  Here K/V tiles of 96 keys stream through a TMA ring (3 K stages, 2 V stages) (they are
// L2 hits: 144 KB per (image, head), re-read once per 128-row query block), smem drops to ~109 KB and TMEM to
// 256 columns, so TWO CTAs are co-resident per SM and one CTA's softmax overlaps the other's MMAs.
//
// Work item = (image, head, 128-row query block); persistent CTAs walk items.  Per item and 96-key tile j:
//     control thread : TMA K_j / V_j (ring), tcgen05.mma S_j = Q K_j^T (128 x 96 x 64) -> TMEM (2 S buffers)
//     4 softmax warps: one tcgen05.ld of the 96 scores of "their" row, row max, LAZY rescale of O (only when
//                      some row's max grew by more than 2^8 — warp-uniform decision), exp2, P_j -> bf16 ->
//                      swizzled smem (UMMA K-major layout)
//     control thread : tcgen05.mma O += P_j V_j (128 x 64 x 96), V as MN-major operand
// and finally O / rowsum -> bf16 -> global.  TMEM columns: S0 [0,96) S1 [96,192) O [192,256).
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"

namespace pt {

constexpr int kBqRows = 128;
constexpr int kBkTile = 96;
constexpr int kBhd = 64;
constexpr int kBkStages = 3;                          // K tile j+2 is consumed (S_{j+2}) right after PV_j is issued
constexpr int kBvStages = 2;                          // V tile j is consumed two tiles after its load is issued
constexpr int kBthreads = 160;                        // 4 softmax warps + 1 control warp
constexpr int kBqBytes = kBqRows * kBhd * 2;          // 16 KB
constexpr int kBkvBytes = kBkTile * kBhd * 2;         // 12 KB per K or V tile
constexpr int kBpBytes = 2 * kBqRows * 128;           // two 64-column swizzle atoms (the second half used)
constexpr int kBsmem = kBqBytes + (kBkStages + kBvStages) * kBkvBytes + kBpBytes + 256 + 1024;   // 109.25 KB
static_assert(2 * (kBsmem + 1024) <= 232448, "two CTAs must fit one SM");
constexpr float kLazyRescaleLog2 = 8.f;

struct AttnSiglip2Dev {
    int B, N, H;
    int nqb, nkt;
    float scale_log2e;
    __nv_bfloat16* out;   // [B*N, H*64]
};

__global__ void __launch_bounds__(kBthreads, 2)
attn_siglip2_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                        const AttnSiglip2Dev p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sQ = smem;
    uint8_t* sK = sQ + kBqBytes;
    uint8_t* sV = sK + kBkStages * kBkvBytes;
    uint8_t* sP = sV + kBvStages * kBkvBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kBpBytes);
    uint64_t* q_full = bars;                  // TMA -> MMA
    uint64_t* k_full = bars + 1;              // [3] TMA -> MMA
    uint64_t* v_full = bars + 4;              // [2] TMA -> MMA
    uint64_t* s_full = bars + 6;              // [2] MMA -> softmax
    uint64_t* p_full = bars + 8;              // softmax (128 arrivals) -> MMA
    uint64_t* pv_done = bars + 9;             // MMA -> softmax / control
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 10);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 128) {
        ptx::prefetch_tmap(&tmQ);
        ptx::prefetch_tmap(&tmKV);
        ptx::mbar_init(q_full, 1);
        for (int s = 0; s < kBkStages; ++s) ptx::mbar_init(&k_full[s], 1);
        for (int s = 0; s < kBvStages; ++s) ptx::mbar_init(&v_full[s], 1);
        ptx::mbar_init(&s_full[0], 1); ptx::mbar_init(&s_full[1], 1);
        ptx::mbar_init(p_full, 128);
        ptx::mbar_init(pv_done, 1);
        ptx::fence_mbar_init();
    }
    if (warp == 4) {
        ptx::tmem_alloc(tmem_slot, 256);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;
    const uint32_t tS[2] = {tmem_base, tmem_base + kBkTile};
    const uint32_t tO = tmem_base + 2 * kBkTile;

    const int items = p.B * p.H * p.nqb;
    const int d = p.H * kBhd;

    if (warp == 4) {
        // ===================== control thread: TMA + MMA =====================
        if (lane == 0) {
            const uint32_t idescS = ptx::make_idesc_bf16(kBqRows, kBkTile, 0, 0);
            const uint32_t idescO = ptx::make_idesc_bf16(kBqRows, kBhd, 0, 1);      // B = V, MN-major
            uint32_t ph_q = 0, ph_k = 0, ph_v = 0 /* bit s = parity of stage s */, ph_p = 0, ph_pv = 0;
            for (int item = blockIdx.x; item < items; item += gridDim.x) {
                const int b = item / (p.H * p.nqb);
                const int r = item - b * (p.H * p.nqb);
                const int h = r / p.nqb, qb = r - h * p.nqb;
                auto load_k = [&](int j) {
                    const int s = j % kBkStages;
                    ptx::mbar_arrive_expect_tx(&k_full[s], kBkvBytes);
                    ptx::tma_load_5d(&tmKV, &k_full[s], sK + s * kBkvBytes, d + h * kBhd, j * kBkTile, b, 0, 0);
                };
                auto load_v = [&](int j) {
                    const int s = j % kBvStages;
                    ptx::mbar_arrive_expect_tx(&v_full[s], kBkvBytes);
                    ptx::tma_load_5d(&tmKV, &v_full[s], sV + s * kBkvBytes, 2 * d + h * kBhd, j * kBkTile, b, 0, 0);
                };
                auto wait_k = [&](int j) {
                    const int s = j % kBkStages;
                    ptx::mbar_wait(&k_full[s], (ph_k >> s) & 1u);
                    ph_k ^= 1u << s;
                };
                auto wait_v = [&](int j) {
                    const int s = j % kBvStages;
                    ptx::mbar_wait(&v_full[s], (ph_v >> s) & 1u);
                    ph_v ^= 1u << s;
                };
                const uint64_t qdesc = ptx::make_smem_desc_sw128(ptx::smem_u32(sQ), 0, 1024);
                auto issue_S = [&](int j) {
                    const uint64_t kdesc =
                        ptx::make_smem_desc_sw128(ptx::smem_u32(sK + (j % kBkStages) * kBkvBytes), 0, 1024);
#pragma unroll
                    for (int k = 0; k < kBhd / 16; ++k)
                        ptx::mma_bf16_ss(tS[j & 1], qdesc + k * 2, kdesc + k * 2, idescS, k != 0);
                    ptx::mma_commit(&s_full[j & 1]);
                };
                // every MMA of the previous item has retired (its last pv_done was waited for below and
                // tcgen05 operations of one thread complete in order): Q and all K/V stages are free
                ptx::mbar_arrive_expect_tx(q_full, kBqBytes);
                ptx::tma_load_5d(&tmQ, q_full, sQ, h * kBhd, qb * kBqRows, b, 0, 0);
                for (int j = 0; j < kBkStages && j < p.nkt; ++j) load_k(j);
                for (int j = 0; j < kBvStages && j < p.nkt; ++j) load_v(j);
                ptx::mbar_wait(q_full, ph_q); ph_q ^= 1;
                wait_k(0);
                ptx::tc_fence_after();
                issue_S(0);
                if (p.nkt > 1) { wait_k(1); ptx::tc_fence_after(); issue_S(1); }
                for (int j = 0; j < p.nkt; ++j) {
                    ptx::mbar_wait(p_full, ph_p); ph_p ^= 1;
                    wait_v(j);
                    ptx::tc_fence_after();
                    const uint32_t vbase = ptx::smem_u32(sV + (j % kBvStages) * kBkvBytes);
#pragma unroll
                    for (int k = 0; k < kBkTile / 16; ++k) {
                        const uint64_t pdesc = ptx::make_smem_desc_sw128(
                            ptx::smem_u32(sP + (k >> 2) * (kBqRows * 128) + (k & 3) * 32), 0, 1024);
                        const uint64_t vdesc = ptx::make_smem_desc_sw128(vbase + k * 16 * 128, 0, 1024);
                        ptx::mma_bf16_ss(tO, pdesc, vdesc, idescO, (j | k) != 0);
                    }
                    ptx::mma_commit(pv_done);
                    if (j + 2 < p.nkt) { wait_k(j + 2); ptx::tc_fence_after(); issue_S(j + 2); }
                    ptx::mbar_wait(pv_done, ph_pv); ph_pv ^= 1;
                    // S_j and PV_j have retired: K stage j % 3 and V stage j % 2 are free
                    if (j + kBkStages < p.nkt) load_k(j + kBkStages);
                    if (j + kBvStages < p.nkt) load_v(j + kBvStages);
                }
            }
        }
    } else {
        // ===================== softmax warps: thread = query row =====================
        const int row = threadIdx.x;                               // TMEM lane
        const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
        const uint32_t p_row = ptx::smem_u32(sP) + row * 128;
        uint32_t ph_s = 0 /* bit i = parity of s_full[i] */, ph_pv = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const int b = item / (p.H * p.nqb);
            const int r = item - b * (p.H * p.nqb);
            const int h = r / p.nqb, qb = r - h * p.nqb;
            float m_run = -INFINITY, l_run = 0.f;
            for (int j = 0; j < p.nkt; ++j) {
                const int sb = j & 1;
                ptx::mbar_wait(&s_full[sb], (ph_s >> sb) & 1u); ph_s ^= 1u << sb;
                ptx::tc_fence_after();
                const int kvalid = min(kBkTile, p.N - j * kBkTile);      // keys of this tile that exist
                uint32_t v[3][32];
#pragma unroll
                for (int c = 0; c < 3; ++c) ptx::tmem_ld_32x32(tS[sb] + lane_off + c * 32, v[c]);
                ptx::tmem_ld_wait();
                if (kvalid < kBkTile) {
#pragma unroll
                    for (int c = 0; c < 3; ++c)
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (c * 32 + i >= kvalid) v[c][i] = __float_as_uint(-INFINITY);
                }
                // eight independent max chains (a single chain of 96 dependent FMNMX is pure latency)
                float mxs[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) mxs[i] = -INFINITY;
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int i = 0; i < 32; ++i)
                        mxs[i & 7] = fmaxf(mxs[i & 7], __uint_as_float(v[c][i]));
                const float mx = fmaxf(fmaxf(fmaxf(mxs[0], mxs[1]), fmaxf(mxs[2], mxs[3])),
                                       fmaxf(fmaxf(mxs[4], mxs[5]), fmaxf(mxs[6], mxs[7])));
                bool pv_waited = false;
                if (j == 0) {
                    m_run = mx;
                } else {
                    const float m_new = fmaxf(m_run, mx);
                    const bool grow = (m_new - m_run) * p.scale_log2e > kLazyRescaleLog2;
                    if (__any_sync(0xffffffffu, grow)) {
                        // rescale O (warp-collective TMEM access): needs the previous tile's PV retired
                        const float alpha = ptx::exp2_fast((m_run - m_new) * p.scale_log2e);
                        ptx::mbar_wait(pv_done, ph_pv); ph_pv ^= 1;
                        pv_waited = true;
                        ptx::tc_fence_after();
#pragma unroll
                        for (int c = 0; c < kBhd / 32; ++c) {
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
                // p = exp2(s * scale*log2e - m * scale*log2e) (<= 2^8 under the lazy max), bf16 pairs
                const float mb = m_run * p.scale_log2e;
                float ls[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};      // independent partial row sums
                uint32_t pk[48];
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int i = 0; i < 16; ++i) {
                        const float e0 = ptx::exp2_fast(fmaf(__uint_as_float(v[c][2 * i]), p.scale_log2e, -mb));
                        const float e1 = ptx::exp2_fast(fmaf(__uint_as_float(v[c][2 * i + 1]), p.scale_log2e, -mb));
                        ls[i & 7] += e0 + e1;
                        pk[c * 16 + i] = pack_bf16(e0, e1);
                    }
                l_run += ((ls[0] + ls[1]) + (ls[2] + ls[3])) + ((ls[4] + ls[5]) + (ls[6] + ls[7]));
                // the single P buffer is free once the previous tile's PV has retired
                if (j > 0 && !pv_waited) { ptx::mbar_wait(pv_done, ph_pv); ph_pv ^= 1; }
                // columns c*32 .. +31 of this row: atom (c*32)/64, 16-byte chunks ((c&1)*4 + q) ^ (row&7)
#pragma unroll
                for (int c = 0; c < 3; ++c)
#pragma unroll
                    for (int q = 0; q < 4; ++q)
                        ptx::sts128(p_row + (c >> 1) * (kBqRows * 128) + ((((c & 1) * 4 + q) ^ (row & 7)) << 4),
                                    pk[c * 16 + 4 * q], pk[c * 16 + 4 * q + 1], pk[c * 16 + 4 * q + 2],
                                    pk[c * 16 + 4 * q + 3]);
                ptx::fence_proxy_async_smem();
                ptx::tc_fence_before();
                ptx::mbar_arrive(p_full);
            }
            // epilogue: O / l -> bf16 -> global
            ptx::mbar_wait(pv_done, ph_pv); ph_pv ^= 1;
            ptx::tc_fence_after();
            const int qrow = qb * kBqRows + row;
            const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
            __nv_bfloat16* op = p.out + (static_cast<long long>(b) * p.N + qrow) * d + h * kBhd;
#pragma unroll
            for (int c = 0; c < kBhd / 32; ++c) {
                uint32_t o[32];
                ptx::tmem_ld_32x32(tO + lane_off + c * 32, o);
                ptx::tmem_ld_wait();
                if (qrow < p.N) {
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
        ptx::tmem_dealloc(tmem_base, 256);
    }
}

}  // namespace pt

using namespace pt;

extern "C" int pt_attn_siglip_fwd(const void* qkv_bf16, void* out_bf16, int B, int N, int H, int hd, float scale,
                                  void* stream) {
    PT_REQUIRE(qkv_bf16 && out_bf16, "pt_attn_siglip_fwd: null pointer");
    PT_REQUIRE(hd == kBhd, "pt_attn_siglip_fwd: head_dim must be 64 (got %d)", hd);
    PT_REQUIRE(B > 0 && H > 0 && N > 0, "pt_attn_siglip_fwd: bad shape");
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const int d = H * kBhd;
    const int ext[3] = {B, 1, 1};
    const long long est[3] = {static_cast<long long>(N) * 3 * d, 0, 0};
    CUtensorMap tq, tkv;
    int rc = encode_operand(&tq, qkv_bf16, 0, N, 3 * d, 3 * d, ext, est, kBqRows, "qkv(Q)");
    if (rc != PT_OK) return rc;
    rc = encode_operand(&tkv, qkv_bf16, 0, N, 3 * d, 3 * d, ext, est, kBkTile, "qkv(KV)");
    if (rc != PT_OK) return rc;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(attn_siglip2_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kBsmem);
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(attn_siglip2): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
        attr_done = true;
    }
    AttnSiglip2Dev p;
    p.B = B; p.N = N; p.H = H;
    p.nqb = (N + kBqRows - 1) / kBqRows;
    p.nkt = (N + kBkTile - 1) / kBkTile;
    p.scale_log2e = scale * 1.4426950408889634f;
    p.out = static_cast<__nv_bfloat16*>(out_bf16);
    const long long items = static_cast<long long>(B) * H * p.nqb;
    PT_REQUIRE(items < (1ll << 31), "pt_attn_siglip_fwd: too many work items");
    const int slots = 2 * num_sms();                       // two co-resident CTAs per SM
    const int grid = items < slots ? static_cast<int>(items) : slots;
    attn_siglip2_fwd_kernel<<<grid, kBthreads, kBsmem, static_cast<cudaStream_t>(stream)>>>(tq, tkv, p);
    return check_launch("attn_siglip2_fwd_kernel");
}
