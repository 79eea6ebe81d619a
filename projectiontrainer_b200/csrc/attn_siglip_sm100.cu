// Fused non-causal multi-head attention forward for the SigLIP tower (head_dim 64, up to 576 keys):
//     O[b, q, h, :] = softmax(Q K^T * scale) V        HF:models/siglip/modeling_siglip.py:293-306
// reading q/k/v in place from the fused qkv activation [B*N, 3*d] and writing the head-concatenated
// context [B*N, d].  Replaces (GEMM -> f32 scores in HBM -> softmax kernel -> bf16 P in HBM -> GEMM).
//
// One persistent CTA per SM walks (image, head) pairs.  For a pair, K and V of the whole sequence
// (<= 576 x 64 bf16 = 72 KB each) are TMA-loaded ONCE into 128B-swizzled smem and reused by every 128-row
// query block.  Per query block and per 192-key tile:
//     control thread : tcgen05.mma  S = Q K_j^T   (128 x 192 x 64)  -> TMEM, two S buffers
//     4 softmax warps: tcgen05.ld S, running row max / sum (online softmax, exp2), rescale O in TMEM
//                      (tcgen05.ld/st), P_j -> bf16 -> smem in the UMMA K-major swizzled layout
//     control thread : tcgen05.mma  O += P_j V_j  (128 x 64 x 192), V read as an MN-major operand
// and finally O / rowsum -> bf16 -> global.  TMEM: S0 [0,192) S1 [192,384) O [384,448).
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"

namespace pt {

constexpr int kAqRows = 128;          // query rows per block (UMMA M)
constexpr int kAkTile = 192;          // keys per tile (UMMA N of S, K of PV)
constexpr int kAhd = 64;              // head dim
constexpr int kAmaxKeys = 576;
constexpr int kAthreads = 160;        // 4 softmax warps + 1 control warp
constexpr int kAqBytes = kAqRows * kAhd * 2;            // 16 KB
constexpr int kAkvBytes = kAmaxKeys * kAhd * 2;         // 72 KB
constexpr int kApBytes = kAqRows * kAkTile * 2;         // 48 KB
constexpr int kAsmem = 2 * kAqBytes + 2 * kAkvBytes + kApBytes + 256 + 1024;

struct AttnSiglipDev {
    int B, N, H;          // images, tokens per image, heads
    int nqb, nkt;         // query blocks, key tiles
    float scale_log2e;    // scale * log2(e)
    __nv_bfloat16* out;   // [B*N, H*64]
};

__global__ void __launch_bounds__(kAthreads, 1)
attn_siglip_fwd_kernel(const __grid_constant__ CUtensorMap tmQ, const __grid_constant__ CUtensorMap tmKV,
                       const AttnSiglipDev p) {
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* sQ = smem;                               // 2 buffers
    uint8_t* sK = smem + 2 * kAqBytes;
    uint8_t* sV = sK + kAkvBytes;
    uint8_t* sP = sV + kAkvBytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(sP + kApBytes);
    uint64_t* q_full = bars;          // [2] TMA -> MMA
    uint64_t* kv_full = bars + 2;     // TMA -> MMA
    uint64_t* s_full = bars + 3;      // [2] MMA -> softmax
    uint64_t* p_full = bars + 5;      // softmax (128 arrivals) -> MMA
    uint64_t* pv_done = bars + 6;     // MMA -> softmax / control
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 8);

    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    if (threadIdx.x == 128) {
        ptx::prefetch_tmap(&tmQ);
        ptx::prefetch_tmap(&tmKV);
        ptx::mbar_init(&q_full[0], 1); ptx::mbar_init(&q_full[1], 1);
        ptx::mbar_init(kv_full, 1);
        ptx::mbar_init(&s_full[0], 1); ptx::mbar_init(&s_full[1], 1);
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
    const uint32_t tS[2] = {tmem_base, tmem_base + kAkTile};
    const uint32_t tO = tmem_base + 2 * kAkTile;

    const int items = p.B * p.H;
    const int d = p.H * kAhd;

    if (warp == 4) {
        // ===================== control thread: TMA + MMA =====================
        if (lane == 0) {
            const uint32_t idescS = ptx::make_idesc_bf16(kAqRows, kAkTile, 0, 0);
            const uint32_t idescO = ptx::make_idesc_bf16(kAqRows, kAhd, 0, 1);      // B = V, MN-major
            uint32_t ph_q[2] = {0, 0}, ph_kv = 0, ph_p = 0, ph_pv = 0;
            for (int item = blockIdx.x; item < items; item += gridDim.x) {
                const int b = item / p.H, h = item - b * p.H;
                // K / V of this (image, head).  The previous pair's MMAs have all retired: this thread waits
                // for every pv_done below and tcgen05 operations of one thread complete in order.
                ptx::mbar_arrive_expect_tx(kv_full, 2 * p.nkt * kAkTile * kAhd * 2);
                for (int j = 0; j < p.nkt; ++j) {
                    ptx::tma_load_5d(&tmKV, kv_full, sK + j * kAkTile * 128, d + h * kAhd, j * kAkTile, b, 0, 0);
                    ptx::tma_load_5d(&tmKV, kv_full, sV + j * kAkTile * 128, 2 * d + h * kAhd, j * kAkTile, b, 0, 0);
                }
                // first Q block
                ptx::mbar_arrive_expect_tx(&q_full[0], kAqBytes);
                ptx::tma_load_5d(&tmQ, &q_full[0], sQ, h * kAhd, 0, b, 0, 0);
                ptx::mbar_wait(kv_full, ph_kv); ph_kv ^= 1;
                for (int qb = 0; qb < p.nqb; ++qb) {
                    const int qbuf = qb & 1;
                    if (qb + 1 < p.nqb) {
                        // the other Q buffer was last read by the S MMAs of block qb-1, all of which retired
                        // before that block's final pv_done (commits are ordered), already waited below
                        ptx::mbar_arrive_expect_tx(&q_full[qbuf ^ 1], kAqBytes);
                        ptx::tma_load_5d(&tmQ, &q_full[qbuf ^ 1], sQ + (qbuf ^ 1) * kAqBytes, h * kAhd,
                                         (qb + 1) * kAqRows, b, 0, 0);
                    }
                    ptx::mbar_wait(&q_full[qbuf], ph_q[qbuf]); ph_q[qbuf] ^= 1;
                    ptx::tc_fence_after();
                    const uint64_t qdesc = ptx::make_smem_desc_sw128(ptx::smem_u32(sQ + qbuf * kAqBytes), 0, 1024);
                    auto issue_S = [&](int j) {
                        const uint64_t kdesc =
                            ptx::make_smem_desc_sw128(ptx::smem_u32(sK + j * kAkTile * 128), 0, 1024);
#pragma unroll
                        for (int k = 0; k < kAhd / 16; ++k)
                            ptx::mma_bf16_ss(tS[j & 1], qdesc + k * 2, kdesc + k * 2, idescS, k != 0);
                        ptx::mma_commit(&s_full[j & 1]);
                    };
                    issue_S(0);
                    if (p.nkt > 1) issue_S(1);
                    for (int j = 0; j < p.nkt; ++j) {
                        ptx::mbar_wait(p_full, ph_p); ph_p ^= 1;
                        ptx::tc_fence_after();
#pragma unroll
                        for (int k = 0; k < kAkTile / 16; ++k) {
                            const uint64_t pdesc = ptx::make_smem_desc_sw128(
                                ptx::smem_u32(sP + (k >> 2) * (kAqRows * 128) + (k & 3) * 32), 0, 1024);
                            const uint64_t vdesc = ptx::make_smem_desc_sw128(
                                ptx::smem_u32(sV + (j * kAkTile + k * 16) * 128), 0, 1024);
                            ptx::mma_bf16_ss(tO, pdesc, vdesc, idescO, (j | k) != 0);
                        }
                        ptx::mma_commit(pv_done);
                        if (j + 2 < p.nkt) issue_S(j + 2);
                        // (the softmax warps cannot produce P_{j+1} before this PV retires, so waiting here
                        //  costs nothing and keeps this thread's view of the barrier phase exact)
                        ptx::mbar_wait(pv_done, ph_pv); ph_pv ^= 1;
                    }
                }
            }
        }
    } else {
        // ===================== softmax warps: thread = query row =====================
        const int row = threadIdx.x;                               // TMEM lane
        const uint32_t lane_off = static_cast<uint32_t>(warp * 32) << 16;
        uint32_t ph_s[2] = {0, 0}, ph_pv = 0;
        for (int item = blockIdx.x; item < items; item += gridDim.x) {
            const int b = item / p.H, h = item - b * p.H;
            for (int qb = 0; qb < p.nqb; ++qb) {
                float m_run = -INFINITY, l_run = 0.f;
                for (int j = 0; j < p.nkt; ++j) {
                    const int sb = j & 1;
                    ptx::mbar_wait(&s_full[sb], ph_s[sb]); ph_s[sb] ^= 1;
                    ptx::tc_fence_after();
                    const int kvalid = min(kAkTile, p.N - j * kAkTile);      // keys of this tile that exist
                    // pass 1: tile row max
                    float mx = -INFINITY;
#pragma unroll 1
                    for (int c = 0; c < kAkTile / 32; ++c) {
                        if (c * 32 >= kvalid) break;
                        uint32_t v[32];
                        ptx::tmem_ld_32x32(tS[sb] + lane_off + c * 32, v);
                        ptx::tmem_ld_wait();
#pragma unroll
                        for (int i = 0; i < 32; ++i)
                            if (c * 32 + i < kvalid) mx = fmaxf(mx, __uint_as_float(v[i]));
                    }
                    const float m_new = fmaxf(m_run, mx);
                    const float alpha = (m_run == -INFINITY) ? 0.f : ptx::exp2_fast((m_run - m_new) * p.scale_log2e);
                    // O and the single P buffer are free once the previous tile's PV has retired
                    if (j > 0) {
                        ptx::mbar_wait(pv_done, ph_pv); ph_pv ^= 1;
                        ptx::tc_fence_after();
#pragma unroll
                        for (int c = 0; c < kAhd / 32; ++c) {
                            uint32_t o[32];
                            ptx::tmem_ld_32x32(tO + lane_off + c * 32, o);
                            ptx::tmem_ld_wait();
#pragma unroll
                            for (int i = 0; i < 32; ++i) o[i] = __float_as_uint(__uint_as_float(o[i]) * alpha);
                            ptx::tmem_st_32x32(tO + lane_off + c * 32, o);
                        }
                        ptx::tmem_st_wait();
                    }
                    // pass 2: p = exp2((s - m) * scale*log2e) -> bf16 -> swizzled smem; row sum
                    const float mb = m_new * p.scale_log2e;
                    float lsum = 0.f;
#pragma unroll 1
                    for (int c = 0; c < kAkTile / 32; ++c) {
                        uint32_t v[32];
                        if (c * 32 < kvalid) {
                            ptx::tmem_ld_32x32(tS[sb] + lane_off + c * 32, v);
                            ptx::tmem_ld_wait();
                        }
                        uint32_t pk[16];
#pragma unroll
                        for (int i = 0; i < 16; ++i) {
                            float e0 = 0.f, e1 = 0.f;
                            if (c * 32 + 2 * i < kvalid)
                                e0 = ptx::exp2_fast(fmaf(__uint_as_float(v[2 * i]), p.scale_log2e, -mb));
                            if (c * 32 + 2 * i + 1 < kvalid)
                                e1 = ptx::exp2_fast(fmaf(__uint_as_float(v[2 * i + 1]), p.scale_log2e, -mb));
                            pk[i] = pack_bf16(e0, e1);
                            // sum what the tensor core will see (bf16-rounded), like the unfused path
                            lsum += bf16_lo(pk[i]) + bf16_hi(pk[i]);
                        }
                        // columns c*32 .. +31 of row `row`: atom (c*32)/64, 16-byte chunks ((c&1)*4 + q) ^ (row&7)
                        const uint32_t base = ptx::smem_u32(sP) + (c >> 1) * (kAqRows * 128) + row * 128;
#pragma unroll
                        for (int q = 0; q < 4; ++q)
                            ptx::sts128(base + ((((c & 1) * 4 + q) ^ (row & 7)) << 4), pk[4 * q], pk[4 * q + 1],
                                        pk[4 * q + 2], pk[4 * q + 3]);
                    }
                    l_run = l_run * alpha + lsum;
                    m_run = m_new;
                    ptx::fence_proxy_async_smem();
                    ptx::tc_fence_before();
                    ptx::mbar_arrive(p_full);
                }
                // epilogue: O / l -> bf16 -> global
                ptx::mbar_wait(pv_done, ph_pv); ph_pv ^= 1;
                ptx::tc_fence_after();
                const int qrow = qb * kAqRows + row;
                const float inv = l_run > 0.f ? 1.f / l_run : 0.f;
                __nv_bfloat16* op = p.out + (static_cast<long long>(b) * p.N + qrow) * d + h * kAhd;
#pragma unroll
                for (int c = 0; c < kAhd / 32; ++c) {
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

extern "C" int pt_attn_siglip_fwd_v1(const void* qkv_bf16, void* out_bf16, int B, int N, int H, int hd, float scale,
                                  void* stream) {
    PT_REQUIRE(qkv_bf16 && out_bf16, "pt_attn_siglip_fwd_v1: null pointer");
    PT_REQUIRE(hd == kAhd, "pt_attn_siglip_fwd_v1: head_dim must be 64 (got %d)", hd);
    PT_REQUIRE(B > 0 && H > 0 && N > 0 && N <= kAmaxKeys, "pt_attn_siglip_fwd_v1: need 0 < N <= %d (N=%d)", kAmaxKeys, N);
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;
    const int d = H * kAhd;
    const int ext[3] = {B, 1, 1};
    const long long est[3] = {static_cast<long long>(N) * 3 * d, 0, 0};
    CUtensorMap tq, tkv;
    int rc = encode_operand(&tq, qkv_bf16, 0, N, 3 * d, 3 * d, ext, est, kAqRows, "qkv(Q)");
    if (rc != PT_OK) return rc;
    rc = encode_operand(&tkv, qkv_bf16, 0, N, 3 * d, 3 * d, ext, est, kAkTile, "qkv(KV)");
    if (rc != PT_OK) return rc;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(attn_siglip_fwd_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, kAsmem);
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(attn_siglip): %s", cudaGetErrorString(e)); return PT_ERR_CUDA; }
        attr_done = true;
    }
    AttnSiglipDev p;
    p.B = B; p.N = N; p.H = H;
    p.nqb = (N + kAqRows - 1) / kAqRows;
    p.nkt = (N + kAkTile - 1) / kAkTile;
    p.scale_log2e = scale * 1.4426950408889634f;
    p.out = static_cast<__nv_bfloat16*>(out_bf16);
    const int items = B * H;
    const int grid = items < num_sms() ? items : num_sms();
    attn_siglip_fwd_kernel<<<grid, kAthreads, kAsmem, static_cast<cudaStream_t>(stream)>>>(tq, tkv, p);
    return check_launch("attn_siglip_fwd_kernel");
}
