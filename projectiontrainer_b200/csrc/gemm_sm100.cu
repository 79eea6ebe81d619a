// Persistent, warp-specialised bf16 GEMM for sm_100a.
//
//   warp 0      TMA producer   (one lane): global -> 128B-swizzled smem ring
//   warp 1      MMA issuer     (one lane): tcgen05.mma 128 x BN x 16, f32 accumulators in TMEM
//   warp 2      TMEM allocator
//   warps 4..7  epilogue: tcgen05.ld -> registers -> (alpha, bias, act, residual) -> global
//
// Two TMEM accumulator stages let the epilogue of tile i overlap the main loop
// of tile i+1.  Grid = min(#tiles, #SMs); each CTA walks tiles t = bid, bid+grid, ...
// (m fastest so CTAs running together share the same B panel in L2).
//
// Operands may be K-major ([rows][K], box 64k x rows) or MN-major ([K][rows],
// 64k x 64mn boxes); both land in smem in the canonical SWIZZLE_128B layouts
// that the UMMA shared-memory descriptors name.
#include <cstdlib>
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"

namespace pt {

constexpr int kGemmThreads = 256;
constexpr int kAccStride = 256;   // TMEM columns per accumulator stage
constexpr int kTmemCols = 512;
constexpr int kSmemBudget = 232448 - 1024 - 256;   // dyn smem minus align slack and barriers

template <int BN>
struct GemmCfg {
    static constexpr int A_BYTES = BM * BK * 2;
    static constexpr int B_BYTES = BN * BK * 2;
    static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
    static constexpr int STAGES_RAW = kSmemBudget / STAGE_BYTES;
    static constexpr int STAGES = STAGES_RAW > 8 ? 8 : STAGES_RAW;
    static constexpr int SMEM_BYTES = STAGES * STAGE_BYTES + 1024 + 256;
};

template <int BN>
__global__ void __launch_bounds__(kGemmThreads, 1)
gemm_bf16_kernel(const __grid_constant__ CUtensorMap tmA, const __grid_constant__ CUtensorMap tmB,
                 const GemmDev p) {
    using Cfg = GemmCfg<BN>;
    constexpr int STAGES = Cfg::STAGES;

    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>(
        (reinterpret_cast<uintptr_t>(smem_raw) + 1023) & ~static_cast<uintptr_t>(1023));
    uint8_t* smem_a = smem;
    uint8_t* smem_b = smem + STAGES * Cfg::A_BYTES;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem + STAGES * Cfg::STAGE_BYTES);
    uint64_t* full_bar = bars;
    uint64_t* empty_bar = bars + STAGES;
    uint64_t* tfull_bar = bars + 2 * STAGES;
    uint64_t* tempty_bar = bars + 2 * STAGES + 2;
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(bars + 2 * STAGES + 4);

    const int warp = threadIdx.x >> 5;
    const int lane = threadIdx.x & 31;

    const int mt = (p.M + BM - 1) / BM;
    const int nt = (p.N + BN - 1) / BN;
    const int tiles_per_z = mt * nt;
    const int total_tiles = tiles_per_z * p.nz[0] * p.nz[1] * p.nz[2];
    const int nkb = (p.K + BK - 1) / BK;
    const int k_iters = nkb * p.kg;

    if (warp == 0 && lane == 0) {
        ptx::prefetch_tmap(&tmA);
        ptx::prefetch_tmap(&tmB);
    }
    if (warp == 1 && lane == 0) {
        for (int s = 0; s < STAGES; ++s) {
            ptx::mbar_init(&full_bar[s], 1);
            ptx::mbar_init(&empty_bar[s], 1);
        }
        for (int s = 0; s < 2; ++s) {
            ptx::mbar_init(&tfull_bar[s], 1);
            ptx::mbar_init(&tempty_bar[s], 4);
        }
        ptx::fence_mbar_init();
    }
    if (warp == 2) {
        ptx::tmem_alloc(tmem_slot, kTmemCols);
        ptx::tmem_relinquish();
    }
    ptx::tc_fence_before();
    __syncthreads();
    ptx::tc_fence_after();
    const uint32_t tmem_base = *tmem_slot;

    if (warp == 0) {
        // ===================== TMA producer =====================
        if (lane == 0) {
            int stage = 0;
            uint32_t phase = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                int zz = t / tiles_per_z, r = t - zz * tiles_per_z;
                const int n_blk = p.raster_n ? r % nt : r / mt;
                const int m_blk = p.raster_n ? r / nt : r - n_blk * mt;
                int z[3];
                z[0] = zz % p.nz[0]; zz /= p.nz[0];
                z[1] = zz % p.nz[1]; zz /= p.nz[1];
                z[2] = zz;
                int ea[3] = {0, 0, 0}, eb[3] = {0, 0, 0};
#pragma unroll
                for (int i = 0; i < 3; ++i) {
                    const int s = p.z_slot[i];
                    if (s >= 0) {
                        if ((p.a_emask >> s) & 1) ea[s] = z[i];
                        if ((p.b_emask >> s) & 1) eb[s] = z[i];
                    }
                }
                for (int it = 0; it < k_iters; ++it) {
                    const int g = it / nkb, kb = it - g * nkb;
                    if (p.kg_slot >= 0) {
                        if ((p.a_emask >> p.kg_slot) & 1) ea[p.kg_slot] = g;
                        if ((p.b_emask >> p.kg_slot) & 1) eb[p.kg_slot] = g;
                    }
                    ptx::mbar_wait(&empty_bar[stage], phase ^ 1);
                    ptx::mbar_arrive_expect_tx(&full_bar[stage], Cfg::STAGE_BYTES);
                    uint8_t* sa = smem_a + stage * Cfg::A_BYTES;
                    uint8_t* sb = smem_b + stage * Cfg::B_BYTES;
                    if (!p.a_mn) {
                        ptx::tma_load_5d(&tmA, &full_bar[stage], sa, kb * BK, m_blk * BM, ea[0],
                                         ea[1], ea[2]);
                    } else {
#pragma unroll
                        for (int c = 0; c < BM / 64; ++c)
                            ptx::tma_load_5d(&tmA, &full_bar[stage], sa + c * (BK * 128),
                                             m_blk * BM + c * 64, kb * BK, ea[0], ea[1], ea[2]);
                    }
                    if (!p.b_mn) {
                        ptx::tma_load_5d(&tmB, &full_bar[stage], sb, kb * BK, n_blk * BN, eb[0],
                                         eb[1], eb[2]);
                    } else {
#pragma unroll
                        for (int c = 0; c < BN / 64; ++c)
                            ptx::tma_load_5d(&tmB, &full_bar[stage], sb + c * (BK * 128),
                                             n_blk * BN + c * 64, kb * BK, eb[0], eb[1], eb[2]);
                    }
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
            }
        }
    } else if (warp == 1) {
        // ===================== MMA issuer =====================
        if (lane == 0) {
            const uint32_t idesc = ptx::make_idesc_bf16(BM, BN, p.a_mn, p.b_mn);
            // K-major: 8-row groups 1024 B apart (SBO); LBO unused with swizzle.
            // MN-major: 64-wide MN chunks BK*128 B apart (LBO); 8-k-row groups 1024 B apart (SBO).
            const uint32_t a_lbo = p.a_mn ? BK * 128 : 0, b_lbo = p.b_mn ? BK * 128 : 0;
            const uint32_t a_kstep = p.a_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
            const uint32_t b_kstep = p.b_mn ? (UMMA_K * 128) >> 4 : (UMMA_K * 2) >> 4;
            int stage = 0;
            uint32_t phase = 0;
            int acc = 0;
            uint32_t acc_phase = 0;
            for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
                ptx::mbar_wait(&tempty_bar[acc], acc_phase ^ 1);
                ptx::tc_fence_after();
                const uint32_t d_tmem = tmem_base + acc * kAccStride;
                for (int it = 0; it < k_iters; ++it) {
                    ptx::mbar_wait(&full_bar[stage], phase);
                    ptx::tc_fence_after();
                    const uint64_t adesc = ptx::make_smem_desc_sw128(
                        ptx::smem_u32(smem_a + stage * Cfg::A_BYTES), a_lbo, 1024);
                    const uint64_t bdesc = ptx::make_smem_desc_sw128(
                        ptx::smem_u32(smem_b + stage * Cfg::B_BYTES), b_lbo, 1024);
#pragma unroll
                    for (int k = 0; k < BK / UMMA_K; ++k) {
                        ptx::mma_bf16_ss(d_tmem, adesc + k * a_kstep, bdesc + k * b_kstep, idesc,
                                         (it | k) != 0);
                    }
                    ptx::mma_commit(&empty_bar[stage]);
                    if (++stage == STAGES) { stage = 0; phase ^= 1; }
                }
                ptx::mma_commit(&tfull_bar[acc]);
                if (++acc == 2) { acc = 0; acc_phase ^= 1; }
            }
        }
    } else if (warp >= 4) {
        // ===================== epilogue =====================
        const int ew = warp - 4;   // == warp % 4: the TMEM lane quarter this warp may read
        int acc = 0;
        uint32_t acc_phase = 0;
        for (int t = blockIdx.x; t < total_tiles; t += gridDim.x) {
            int zz = t / tiles_per_z, r = t - zz * tiles_per_z;
            const int n_blk = p.raster_n ? r % nt : r / mt;
            const int m_blk = p.raster_n ? r / nt : r - n_blk * mt;
            const int z0 = zz % p.nz[0]; zz /= p.nz[0];
            const int z1 = zz % p.nz[1]; zz /= p.nz[1];
            const int z2 = zz;
            const long long coff = z0 * p.c_zs[0] + z1 * p.c_zs[1] + z2 * p.c_zs[2];
            const long long roff = z0 * p.r_zs[0] + z1 * p.r_zs[1] + z2 * p.r_zs[2];
            const int row = m_blk * BM + ew * 32 + lane;
            const bool row_ok = row < p.M;

            ptx::mbar_wait(&tfull_bar[acc], acc_phase);
            ptx::tc_fence_after();
            const uint32_t taddr = tmem_base + acc * kAccStride + (static_cast<uint32_t>(ew * 32) << 16);
#pragma unroll 1
            for (int c = 0; c < BN / 32; ++c) {
                uint32_t v[32];
                ptx::tmem_ld_32x32(taddr + c * 32, v);
                ptx::tmem_ld_wait();
                const int col0 = n_blk * BN + c * 32;
                if (row_ok && col0 < p.N) {
#pragma unroll
                    for (int g8 = 0; g8 < 4; ++g8) {
                        const int col = col0 + g8 * 8;
                        if (col >= p.N) break;
                        float f[8], pre[8];
#pragma unroll
                        for (int j = 0; j < 8; ++j) f[j] = __uint_as_float(v[g8 * 8 + j]) * p.alpha;
                        const bool full8 = col + 8 <= p.N;
                        if (p.bias) {
#pragma unroll
                            for (int j = 0; j < 8; ++j)
                                if (full8 || col + j < p.N) f[j] += __ldg(p.bias + col + j);
                        }
#pragma unroll
                        for (int j = 0; j < 8; ++j) pre[j] = f[j];
                        if (p.act != PT_ACT_NONE) {
#pragma unroll
                            for (int j = 0; j < 8; ++j) f[j] = apply_act(f[j], p.act);
                        }
                        if (p.R) {
                            const long long ro = roff + static_cast<long long>(row) * p.ldr + col;
                            if (p.r_f32) {
                                const float* rp = static_cast<const float*>(p.R) + ro;
                                if (full8) {
                                    const float4 r0 = *reinterpret_cast<const float4*>(rp);
                                    const float4 r1 = *reinterpret_cast<const float4*>(rp + 4);
                                    f[0] += r0.x; f[1] += r0.y; f[2] += r0.z; f[3] += r0.w;
                                    f[4] += r1.x; f[5] += r1.y; f[6] += r1.z; f[7] += r1.w;
                                } else {
                                    for (int j = 0; j < 8 && col + j < p.N; ++j) f[j] += rp[j];
                                }
                            } else {
                                const __nv_bfloat16* rp = static_cast<const __nv_bfloat16*>(p.R) + ro;
                                if (full8) {
                                    const uint4 u = *reinterpret_cast<const uint4*>(rp);
                                    f[0] += bf16_lo(u.x); f[1] += bf16_hi(u.x);
                                    f[2] += bf16_lo(u.y); f[3] += bf16_hi(u.y);
                                    f[4] += bf16_lo(u.z); f[5] += bf16_hi(u.z);
                                    f[6] += bf16_lo(u.w); f[7] += bf16_hi(u.w);
                                } else {
                                    for (int j = 0; j < 8 && col + j < p.N; ++j)
                                        f[j] += __bfloat162float(rp[j]);
                                }
                            }
                        }
                        const long long co = coff + static_cast<long long>(row) * p.ldc + col;
                        if (p.c_f32) {
                            float* cp = static_cast<float*>(p.C) + co;
                            if (full8) {
                                *reinterpret_cast<float4*>(cp) = make_float4(f[0], f[1], f[2], f[3]);
                                *reinterpret_cast<float4*>(cp + 4) = make_float4(f[4], f[5], f[6], f[7]);
                            } else {
                                for (int j = 0; j < 8 && col + j < p.N; ++j) cp[j] = f[j];
                            }
                        } else {
                            __nv_bfloat16* cp = static_cast<__nv_bfloat16*>(p.C) + co;
                            if (full8) {
                                uint4 u;
                                u.x = pack_bf16(f[0], f[1]); u.y = pack_bf16(f[2], f[3]);
                                u.z = pack_bf16(f[4], f[5]); u.w = pack_bf16(f[6], f[7]);
                                *reinterpret_cast<uint4*>(cp) = u;
                            } else {
                                for (int j = 0; j < 8 && col + j < p.N; ++j)
                                    cp[j] = __float2bfloat16(f[j]);
                            }
                        }
                        if (p.C_pre) {
                            __nv_bfloat16* pp = static_cast<__nv_bfloat16*>(p.C_pre) + co;
                            if (full8) {
                                uint4 u;
                                u.x = pack_bf16(pre[0], pre[1]); u.y = pack_bf16(pre[2], pre[3]);
                                u.z = pack_bf16(pre[4], pre[5]); u.w = pack_bf16(pre[6], pre[7]);
                                *reinterpret_cast<uint4*>(pp) = u;
                            } else {
                                for (int j = 0; j < 8 && col + j < p.N; ++j)
                                    pp[j] = __float2bfloat16(pre[j]);
                            }
                        }
                    }
                }
            }
            ptx::tc_fence_before();
            __syncwarp();
            if (lane == 0) ptx::mbar_arrive(&tempty_bar[acc]);
            if (++acc == 2) { acc = 0; acc_phase ^= 1; }
        }
    }

    ptx::tc_fence_before();
    __syncthreads();
    if (warp == 2) {
        ptx::tc_fence_after();
        ptx::tmem_dealloc(tmem_base, kTmemCols);
    }
}


// ---------------------------------------------------------------------------
// Host side: tensor-map encoding + launch
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*,
                                  const cuuint64_t*, const cuuint64_t*, const cuuint32_t*,
                                  const cuuint32_t*, CUtensorMapInterleave, CUtensorMapSwizzle,
                                  CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn get_encode_fn() {
    static EncodeTiledFn fn = nullptr;
    if (fn) return fn;
    void* sym = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &sym, cudaEnableDefault, &q) !=
            cudaSuccess || q != cudaDriverEntryPointSuccess || !sym) {
        return nullptr;
    }
    fn = reinterpret_cast<EncodeTiledFn>(sym);
    return fn;
}

// One operand = a 5-D bf16 tensor map {inner, outer, e0, e1, e2}.
//   K-major : inner = K (box 64), outer = rows (box `rows_box`)
//   MN-major: inner = rows (box 64), outer = K (box 64)
int encode_operand(CUtensorMap* out, const void* base, int mn_major, int rows, int K,
                          long long ld, const int ext[3], const long long est[3], int rows_box,
                          const char* name) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return PT_ERR_CUDA; }
    if ((reinterpret_cast<uintptr_t>(base) & 15) || (ld & 7)) {
        set_error("gemm operand %s: base must be 16B aligned and ld a multiple of 8 (ld=%lld)", name, ld);
        return PT_ERR_INVALID;
    }
    cuuint64_t dims[5];
    cuuint64_t strides[4];
    cuuint32_t box[5] = {64, 64, 1, 1, 1};
    cuuint32_t estr[5] = {1, 1, 1, 1, 1};
    if (!mn_major) {
        dims[0] = static_cast<cuuint64_t>(K); dims[1] = static_cast<cuuint64_t>(rows);
        box[1] = static_cast<cuuint32_t>(rows_box);
    } else {
        dims[0] = static_cast<cuuint64_t>(rows); dims[1] = static_cast<cuuint64_t>(K);
    }
    strides[0] = static_cast<cuuint64_t>(ld) * 2;
    for (int i = 0; i < 3; ++i) {
        if (est[i] != 0) {
            if (est[i] & 7) {
                set_error("gemm operand %s: batch stride %d (%lld) must be a multiple of 8", name, i, est[i]);
                return PT_ERR_INVALID;
            }
            dims[2 + i] = static_cast<cuuint64_t>(ext[i]);
            strides[1 + i] = static_cast<cuuint64_t>(est[i]) * 2;
        } else {
            dims[2 + i] = 1;
            strides[1 + i] = strides[0];
        }
    }
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), dims, strides,
                     box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B,
                     CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(%s) failed: CUresult %d (dims %llu %llu %llu %llu %llu, "
                  "strides %llu %llu %llu %llu)", name, static_cast<int>(r),
                  (unsigned long long)dims[0], (unsigned long long)dims[1], (unsigned long long)dims[2],
                  (unsigned long long)dims[3], (unsigned long long)dims[4],
                  (unsigned long long)strides[0], (unsigned long long)strides[1],
                  (unsigned long long)strides[2], (unsigned long long)strides[3]);
        return PT_ERR_CUDA;
    }
    return PT_OK;
}

// Generic 5-D bf16 tiled map with 128-byte swizzle (attention kernels build their own geometries).
int encode_tiled_bf16_2d(CUtensorMap* out, const void* base, long long cols, long long rows, long long ld,
                         int box_cols, int box_rows, int swizzle_bytes, const char* name) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return PT_ERR_CUDA; }
    cuuint64_t d[2] = {static_cast<cuuint64_t>(cols), static_cast<cuuint64_t>(rows)};
    cuuint64_t st[1] = {static_cast<cuuint64_t>(ld) * 2};
    cuuint32_t bx[2] = {static_cast<cuuint32_t>(box_cols), static_cast<cuuint32_t>(box_rows)}, es[2] = {1, 1};
    const CUtensorMapSwizzle sw = swizzle_bytes == 128 ? CU_TENSOR_MAP_SWIZZLE_128B
                                  : swizzle_bytes == 64 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_NONE;
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), d, st, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, sw, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(%s) failed: CUresult %d (dims %lld %lld, ld %lld, box %d %d, swizzle %d)", name,
                  static_cast<int>(r), cols, rows, ld, box_cols, box_rows, swizzle_bytes);
        return PT_ERR_CUDA;
    }
    return PT_OK;
}

int encode_tiled_bf16_5d(CUtensorMap* out, const void* base, const unsigned long long dims[5],
                         const unsigned long long strides_bytes[4], const unsigned int box[5], const char* name) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return PT_ERR_CUDA; }
    cuuint64_t d[5], st[4];
    cuuint32_t bx[5], es[5] = {1, 1, 1, 1, 1};
    for (int i = 0; i < 5; ++i) { d[i] = dims[i]; bx[i] = box[i]; }
    for (int i = 0; i < 4; ++i) st[i] = strides_bytes[i];
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 5, const_cast<void*>(base), d, st, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(%s) failed: CUresult %d (dims %llu %llu %llu %llu %llu, strides %llu %llu "
                  "%llu %llu, box %u %u %u %u %u)", name, static_cast<int>(r), dims[0], dims[1], dims[2], dims[3],
                  dims[4], strides_bytes[0], strides_bytes[1], strides_bytes[2], strides_bytes[3], box[0], box[1],
                  box[2], box[3], box[4]);
        return PT_ERR_CUDA;
    }
    return PT_OK;
}

template <int BN>
static int launch_gemm(const CUtensorMap& ta, const CUtensorMap& tb, const GemmDev& p, int tiles,
                       cudaStream_t st) {
    using Cfg = GemmCfg<BN>;
    static bool attr_done = false;
    if (!attr_done) {
        cudaError_t e = cudaFuncSetAttribute(gemm_bf16_kernel<BN>,
                                             cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM_BYTES);
        if (e != cudaSuccess) { set_error("cudaFuncSetAttribute(gemm<%d>): %s", BN, cudaGetErrorString(e)); return PT_ERR_CUDA; }
        attr_done = true;
    }
    int grid = tiles < num_sms() ? tiles : num_sms();
    gemm_bf16_kernel<BN><<<grid, kGemmThreads, Cfg::SMEM_BYTES, st>>>(ta, tb, p);
    return check_launch("gemm_bf16_kernel");
}

// Pick the N tile that wastes the least tensor-core work (ties -> larger tile).
static int pick_bn(int M, int N, long long nbatch) {
    const int cands[4] = {256, 192, 128, 64};
    const int sms = num_sms();
    double best = 1e30; int best_bn = 256;
    const long long mt = (M + BM - 1) / BM;
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        const long long nt = (N + bn - 1) / bn;
        const long long tiles = mt * nt * nbatch;
        const long long waves = (tiles + sms - 1) / sms;
        // cost ~ waves * (mainloop ∝ bn, with a fixed per-tile overhead)
        double cost = static_cast<double>(waves) * (bn + 24.0);
        if (cost < best * 0.999) { best = cost; best_bn = bn; }
    }
    return best_bn;
}

// N tile for the CTA-pair kernel: 74 clusters, each tile 256 x bn; MN-major B needs bn/2 % 64 == 0.
static int pick_bn2(int M, int N, long long nbatch, int b_mn) {
    const int cands[4] = {256, 192, 128, 64};
    const int slots = num_sms() / 2;
    double best = 1e30; int best_bn = 256;
    const long long mt = (M + 2 * BM - 1) / (2 * BM);
    for (int i = 0; i < 4; ++i) {
        const int bn = cands[i];
        if (b_mn && (bn / 2) % 64 != 0) continue;
        const long long nt = (N + bn - 1) / bn;
        const long long tiles = mt * nt * nbatch;
        const long long waves = (tiles + slots - 1) / slots;
        // narrower tiles re-read A from smem more often per flop: measured ~1.4 PF/s at bn 256 but ~1.0 at 128
        // (operand reads + TMA writes exceed the 128 B/clk smem port), hence the per-width factor
        const double f = bn == 256 ? 1.0 : bn == 192 ? 1.15 : bn == 128 ? 1.5 : 2.5;
        double cost = static_cast<double>(waves) * (bn * f + 16.0);
        if (cost < best * 0.999) { best = cost; best_bn = bn; }
    }
    return best_bn;
}

// PT_GEMM_HALF_TILE=0 in the environment: shifted full-width last n tile everywhere (A/B probes)
static int half_tile_enabled() {
    static const int v = [] { const char* e = getenv("PT_GEMM_HALF_TILE"); return !(e && e[0] == '0'); }();
    return v;
}
int g_gemm_impl = 0;   // 0 auto (CTA-pair kernel), 1 single-CTA kernel, 2 CTA-pair kernel

}  // namespace pt

extern "C" int pt_gemm_bf16(const pt_gemm_args* a, void* stream) {
    using namespace pt;
    PT_REQUIRE(a != nullptr, "pt_gemm_bf16: null args");
    PT_REQUIRE(a->M > 0 && a->N > 0 && a->K > 0, "pt_gemm_bf16: M,N,K must be positive (%d,%d,%d)", a->M, a->N, a->K);
    PT_REQUIRE(a->A && a->B && a->C, "pt_gemm_bf16: null operand");
    PT_REQUIRE(a->nz[0] >= 1 && a->nz[1] >= 1 && a->nz[2] >= 1, "pt_gemm_bf16: batch extents must be >= 1");
    PT_REQUIRE(a->c_dtype == PT_BF16 || a->c_dtype == PT_F32, "pt_gemm_bf16: c_dtype must be bf16 or f32");
    const int kg = a->kg > 1 ? a->kg : 1;
    PT_REQUIRE(kg == 1 || a->nz[2] == 1, "pt_gemm_bf16: kg > 1 requires nz[2] == 1");
    const int cvec = a->c_dtype == PT_F32 ? 4 : 8;
    PT_REQUIRE((a->ldc % cvec) == 0 && (reinterpret_cast<uintptr_t>(a->C) & 15) == 0,
               "pt_gemm_bf16: C must be 16B aligned with ldc %% %d == 0 (ldc=%lld)", cvec, (long long)a->ldc);
    for (int i = 0; i < 3; ++i)
        PT_REQUIRE((a->c_zs[i] % cvec) == 0, "pt_gemm_bf16: c batch stride %d must be a multiple of %d", i, cvec);
    if (a->C_pre) PT_REQUIRE(a->c_dtype == PT_BF16 || (a->ldc % 8) == 0, "pt_gemm_bf16: C_pre needs ldc %% 8 == 0");
    if (a->residual) {
        const int rvec = a->r_dtype == PT_F32 ? 4 : 8;
        PT_REQUIRE(a->r_dtype == PT_BF16 || a->r_dtype == PT_F32, "pt_gemm_bf16: r_dtype must be bf16 or f32");
        PT_REQUIRE((a->ldr % rvec) == 0 && (reinterpret_cast<uintptr_t>(a->residual) & 15) == 0,
                   "pt_gemm_bf16: residual must be 16B aligned with ldr %% %d == 0", rvec);
        for (int i = 0; i < 3; ++i)
            PT_REQUIRE((a->r_zs[i] % rvec) == 0, "pt_gemm_bf16: residual batch stride %d misaligned", i);
    }
    PT_REQUIRE((a->N % 4) == 0 || ((!a->bias || a->act == PT_ACT_SOFTMAX_BWD) && !a->residual && !a->C_pre),
               "pt_gemm_bf16: bias / residual / C_pre need N %% 4 == 0 (N=%d)", a->N);
    PT_REQUIRE(a->act >= PT_ACT_NONE && a->act <= PT_ACT_SWIGLU_BWD, "pt_gemm_bf16: bad act %d", a->act);
    const bool silu = a->act == PT_ACT_SWIGLU || a->act == PT_ACT_SWIGLU_BWD;
    const bool geglu = a->act == PT_ACT_GEGLU || a->act == PT_ACT_SWIGLU;
    const bool geglu_bwd = a->act == PT_ACT_GEGLU_BWD || a->act == PT_ACT_SWIGLU_BWD;
    if (geglu || geglu_bwd) {
        PT_REQUIRE(a->aux != nullptr && (reinterpret_cast<uintptr_t>(a->aux) & 15) == 0 && a->ld_aux % 8 == 0,
                   "pt_gemm_bf16: fused GeGLU needs a 16B-aligned aux matrix with ld_aux %% 8 == 0");
        PT_REQUIRE(a->c_dtype == PT_BF16 && !a->bias && !a->residual && !a->C_pre && a->alpha == 1.0f,
                   "pt_gemm_bf16: fused GeGLU takes bf16 C and no bias / residual / C_pre / alpha");
        PT_REQUIRE(a->nz[0] * a->nz[1] * a->nz[2] == 1 && kg == 1, "pt_gemm_bf16: fused GeGLU is not batched");
        PT_REQUIRE(g_gemm_impl != 1, "pt_gemm_bf16: fused GeGLU epilogues exist in the CTA-pair kernel only");
        const int ff = geglu ? a->N / 2 : a->N;
        PT_REQUIRE(ff % 128 == 0 && (!geglu || a->N == 2 * ff), "pt_gemm_bf16: fused GeGLU needs ff %% 128 == 0 (ff=%d)", ff);
        PT_REQUIRE(a->ldc >= 2 * ff && a->ld_aux >= (geglu ? ff : 2 * ff), "pt_gemm_bf16: fused GeGLU: ldc / ld_aux too small");
    }
    if (pt_device_check() != PT_OK) return PT_ERR_CUDA;

    GemmDev p;
    memset(&p, 0, sizeof(p));
    p.M = a->M; p.N = a->N; p.K = a->K;
    p.kg = kg;
    p.a_mn = a->a_mn_major ? 1 : 0;
    p.b_mn = a->b_mn_major ? 1 : 0;
    int ext[3]; long long ast[3], bst[3];
    if (kg > 1) {
        p.kg_slot = 0; p.z_slot[0] = 1; p.z_slot[1] = 2; p.z_slot[2] = -1;
        ext[0] = kg; ext[1] = a->nz[0]; ext[2] = a->nz[1];
        ast[0] = a->a_kgs; ast[1] = a->a_zs[0]; ast[2] = a->a_zs[1];
        bst[0] = a->b_kgs; bst[1] = a->b_zs[0]; bst[2] = a->b_zs[1];
    } else {
        p.kg_slot = -1;
        for (int i = 0; i < 3; ++i) { p.z_slot[i] = i; ext[i] = a->nz[i]; ast[i] = a->a_zs[i]; bst[i] = a->b_zs[i]; }
    }
    for (int i = 0; i < 3; ++i) {
        p.nz[i] = a->nz[i];
        if (ast[i] != 0) p.a_emask |= 1 << i;
        if (bst[i] != 0) p.b_emask |= 1 << i;
        p.c_zs[i] = a->c_zs[i]; p.r_zs[i] = a->r_zs[i];
    }
    p.C = a->C; p.C_pre = a->C_pre; p.bias = a->bias; p.R = a->residual;
    p.ldc = a->ldc; p.ldr = a->ldr;
    p.c_f32 = a->c_dtype == PT_F32; p.r_f32 = a->r_dtype == PT_F32;
    p.act = geglu ? PT_ACT_GEGLU : (geglu_bwd ? PT_ACT_GEGLU_BWD : a->act);     // SwiGLU = same epilogues, other constants
    p.alpha = a->alpha;
    p.glu = glu_consts(silu);
    if (a->act == PT_ACT_SOFTMAX_BWD) {
        PT_REQUIRE(a->aux && a->bias && a->c_dtype == PT_BF16 && !a->residual && !a->C_pre && kg == 1,
                   "pt_gemm_bf16: fused softmax backward needs aux = P, bias = delta, bf16 C, no residual / C_pre");
        PT_REQUIRE(g_gemm_impl != 1, "pt_gemm_bf16: the fused softmax-backward epilogue exists in the CTA-pair kernel only");
        PT_REQUIRE(a->ld_aux % 4 == 0 && (reinterpret_cast<uintptr_t>(a->aux) & 7) == 0, "pt_gemm_bf16: P misaligned");
        for (int i = 0; i < 3; ++i)
            PT_REQUIRE(a->c_zs[i] % a->ldc == 0, "pt_gemm_bf16: fused softmax backward needs c_zs %% ldc == 0");
    }
    p.aux = a->aux; p.ld_aux = a->ld_aux; p.ff = geglu ? a->N / 2 : a->N;
    PT_REQUIRE(a->band >= PT_BAND_NONE && a->band <= PT_BAND_K_BY_KEY && a->band_window >= 0,
               "pt_gemm_bf16: bad band %d / window %d", a->band, a->band_window);
    p.band = a->band; p.band_w = a->band_window;     // (the single-CTA kernel ignores it: a pure work-skipping hint)

    const long long nbatch = static_cast<long long>(a->nz[0]) * a->nz[1] * a->nz[2];
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUtensorMap ta, tb;
    if (g_gemm_impl != 1) {
        // CTA-pair kernel (gemm2_sm100.cu): 256 x bn tiles, one per cluster of two SMs
        int bn = geglu ? 256 : pick_bn2(a->M, a->N, nbatch, p.b_mn);
        if (!geglu && a->bn_hint != 0) {
            PT_REQUIRE(a->bn_hint == 64 || a->bn_hint == 128 || a->bn_hint == 192 || a->bn_hint == 256,
                       "pt_gemm_bf16: bn_hint must be 0, 64, 128, 192 or 256");
            PT_REQUIRE(!p.b_mn || (a->bn_hint / 2) % 64 == 0, "pt_gemm_bf16: an MN-major B operand needs bn_hint 128 or 256");
            bn = a->bn_hint;
        }
        const long long mt = (a->M + 2 * BM - 1) / (2 * BM), nt = (a->N + bn - 1) / bn;
        PT_REQUIRE(mt * nt * nbatch < (1ll << 31), "pt_gemm_bf16: too many tiles");
        p.raster_n = nt <= mt;
        // (a shifted start must keep TMA's 16-byte global alignment when it is the contiguous dimension)
        p.m_last = (a->M >= 2 * BM && (!p.a_mn || a->M % 8 == 0)) ? a->M - 2 * BM : static_cast<int>(mt - 1) * 2 * BM;
        p.n_last = (a->N >= bn && a->N % 32 == 0) ? a->N - bn : static_cast<int>(nt - 1) * bn;
        // (pays from K ~ 1536 on: 255 vs 266 us at K = 6912, 507 vs 527 at 13 824, but 54 vs 51 us at K = 1024, where
        //  the tile is bound by its operand loads and epilogue, which the half tile does not shorten)
        p.n_half = half_tile_enabled() && bn == 256 && nt >= 2 && a->N % bn == bn / 2 && p.band == 0 &&
                   static_cast<long long>(a->K) * (a->kg > 0 ? a->kg : 1) >= 1536 &&
                   (p.act == PT_ACT_NONE || p.act == PT_ACT_GELU_TANH || p.act == PT_ACT_GELU_ERF);
        if (p.n_half) p.n_last = static_cast<int>(nt - 1) * bn;
        int rc = encode_operand(&ta, a->A, p.a_mn, a->M, a->K, a->lda, ext, ast, BM, "A");
        if (rc != PT_OK) return rc;
        rc = encode_operand(&tb, a->B, p.b_mn, a->N, a->K, a->ldb, ext, bst, bn / 2, "B");
        if (rc != PT_OK) return rc;
        CUtensorMap t_aux = ta, t_c = ta;
        if (geglu_bwd) {
            // the GeGLU-backward epilogue moves its gate / up tiles in and its d_gate / d_up tiles out with TMA
            rc = encode_tiled_bf16_2d(&t_aux, a->aux, 2ll * p.ff, a->M, a->ld_aux, 32, 32, 64, "gu (epilogue)");
            if (rc != PT_OK) return rc;
            rc = encode_tiled_bf16_2d(&t_c, a->C, 2ll * p.ff, a->M, a->ldc, 32, 32, 64, "d_gu (epilogue)");
            if (rc != PT_OK) return rc;
        }
        return launch_gemm2(bn, ta, tb, t_aux, t_c, p, st);
    }
    const int bn = pick_bn(a->M, a->N, nbatch);
    const long long tiles_ll = static_cast<long long>((a->M + BM - 1) / BM) * ((a->N + bn - 1) / bn) * nbatch;
    PT_REQUIRE(tiles_ll < (1ll << 31), "pt_gemm_bf16: too many tiles");
    const int tiles = static_cast<int>(tiles_ll);
    p.raster_n = (a->N + bn - 1) / bn <= (a->M + BM - 1) / BM;

    int rc = encode_operand(&ta, a->A, p.a_mn, a->M, a->K, a->lda, ext, ast, BM, "A");
    if (rc != PT_OK) return rc;
    rc = encode_operand(&tb, a->B, p.b_mn, a->N, a->K, a->ldb, ext, bst, bn, "B");
    if (rc != PT_OK) return rc;
    switch (bn) {
        case 256: return launch_gemm<256>(ta, tb, p, tiles, st);
        case 192: return launch_gemm<192>(ta, tb, p, tiles, st);
        case 128: return launch_gemm<128>(ta, tb, p, tiles, st);
        default:  return launch_gemm<64>(ta, tb, p, tiles, st);
    }
}

extern "C" int pt_set_gemm_impl(int impl) {
    if (impl < 0 || impl > 2) { pt::set_error("pt_set_gemm_impl: impl must be 0 (auto), 1 or 2"); return PT_ERR_INVALID; }
    pt::g_gemm_impl = impl;
    return PT_OK;
}
