// Host side of pt_gemm_bf16: argument checks, tensor-map encoding, tile-width choice, launch of the CTA-pair
// tcgen05 kernel (gemm2_sm100.cu).
//
// Operands may be K-major ([rows][K], box 64k x rows) or MN-major ([K][rows], 64k x 64mn boxes); both land in
// smem in the canonical SWIZZLE_128B layouts that the UMMA shared-memory descriptors name.
#include <cstdlib>
#include "pt_common.cuh"
#include "sm100_ptx.cuh"
#include "gemm_common.cuh"

namespace pt {

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

int encode_tiled_bf16_3d(CUtensorMap* out, const void* base, const unsigned long long dims[3],
                         const unsigned long long strides_bytes[2], const unsigned int box[3], const char* name) {
    EncodeTiledFn enc = get_encode_fn();
    if (!enc) { set_error("cuTensorMapEncodeTiled entry point not available"); return PT_ERR_CUDA; }
    cuuint64_t d[3] = {dims[0], dims[1], dims[2]}, st[2] = {strides_bytes[0], strides_bytes[1]};
    cuuint32_t bx[3] = {box[0], box[1], box[2]}, es[3] = {1, 1, 1};
    CUresult r = enc(out, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 3, const_cast<void*>(base), d, st, bx, es,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_256B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS) {
        set_error("cuTensorMapEncodeTiled(%s) failed: CUresult %d (dims %llu %llu %llu, strides %llu %llu, box %u %u %u)",
                  name, static_cast<int>(r), dims[0], dims[1], dims[2], strides_bytes[0], strides_bytes[1], box[0],
                  box[1], box[2]);
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
    PT_REQUIRE(a->act >= PT_ACT_NONE && a->act <= PT_ACT_QKNORM_ROPE, "pt_gemm_bf16: bad act %d", a->act);
    if (a->act == PT_ACT_CE_STATS) {
        PT_REQUIRE(a->aux2 != nullptr && (reinterpret_cast<uintptr_t>(a->aux2) & 7) == 0,
                   "pt_gemm_bf16: PT_ACT_CE_STATS needs aux2 = the f32 partials buffer [2 * ceil(N / 256)][M][2]");
        PT_REQUIRE(a->c_dtype == PT_BF16 && !a->bias && !a->residual && !a->C_pre && a->alpha == 1.0f && a->K2 == 0 &&
                       a->nz[0] * a->nz[1] * a->nz[2] == 1 && kg == 1 && a->band == PT_BAND_NONE && a->N >= 256,
                   "pt_gemm_bf16: PT_ACT_CE_STATS takes bf16 C, N >= 256, no bias / residual / C_pre / alpha / K2 and is not batched");
    }
    const bool tile_epi = a->act == PT_ACT_GELU_ERF_BWD || a->act == PT_ACT_ATTN_DELTA;
    if (tile_epi) {
        PT_REQUIRE(a->aux != nullptr && (reinterpret_cast<uintptr_t>(a->aux) & 15) == 0 && a->ld_aux % 8 == 0 &&
                       a->ld_aux >= a->N, "pt_gemm_bf16: act %d needs a 16B-aligned aux matrix [M, N] with ld_aux %% 8 == 0", a->act);
        PT_REQUIRE(a->c_dtype == PT_BF16 && !a->bias && !a->residual && !a->C_pre && a->alpha == 1.0f && a->N % 32 == 0,
                   "pt_gemm_bf16: act %d takes bf16 C, N %% 32 == 0 and no bias / residual / C_pre / alpha", a->act);
        PT_REQUIRE(a->nz[0] * a->nz[1] * a->nz[2] == 1 && kg == 1 && a->band == PT_BAND_NONE, "pt_gemm_bf16: act %d is not batched", a->act);
    }
    if (a->act == PT_ACT_ATTN_DELTA) {
        PT_REQUIRE(a->aux2 != nullptr && a->aux_i0 > 0 && a->aux_i1 > 0 && a->aux_i1 % 128 == 0 && a->N % 256 == 0 &&
                       a->N % a->aux_i1 == 0 && a->M % a->aux_i0 == 0,
                   "pt_gemm_bf16: PT_ACT_ATTN_DELTA needs aux2 = delta, aux_i0 = S (M %% S == 0), aux_i1 = head_dim "
                   "(%% 128 == 0) and N %% 256 == 0");
    }
    if (a->act == PT_ACT_QKNORM_ROPE) {
        PT_REQUIRE(a->aux && a->aux2 && a->bias && a->residual && a->r_dtype == PT_F32 && a->aux_i0 > 0 && a->aux_i1 > 0,
                   "pt_gemm_bf16: PT_ACT_QKNORM_ROPE needs aux = q out, aux2 = k out, bias = gains, residual = f32 rotary "
                   "table, aux_i0 = table rows, aux_i1 = query heads");
        PT_REQUIRE(a->N % 256 == 0 && a->N / 256 > a->aux_i1 && (a->N / 256 - a->aux_i1) % 2 == 0,
                   "pt_gemm_bf16: PT_ACT_QKNORM_ROPE needs N = (hq + 2 hkv) * 256");
        PT_REQUIRE(a->c_dtype == PT_BF16 && !a->C_pre && a->alpha == 1.0f && a->nz[0] * a->nz[1] * a->nz[2] == 1 &&
                       kg == 1 && a->band == PT_BAND_NONE && a->ksplit <= 1,
                   "pt_gemm_bf16: PT_ACT_QKNORM_ROPE takes bf16 C, no C_pre / alpha / batch / K groups / band / split-K");
        PT_REQUIRE(a->ld_aux >= a->aux_i1 * 256 && a->ld_aux % 8 == 0 && (reinterpret_cast<uintptr_t>(a->aux) & 15) == 0 &&
                       (reinterpret_cast<uintptr_t>(a->aux2) & 15) == 0 && (reinterpret_cast<uintptr_t>(a->residual) & 15) == 0,
                   "pt_gemm_bf16: PT_ACT_QKNORM_ROPE: aux / aux2 / table misaligned or ld_aux too small");
    }
    const bool silu = a->act == PT_ACT_SWIGLU || a->act == PT_ACT_SWIGLU_BWD;
    const bool geglu = a->act == PT_ACT_GEGLU || a->act == PT_ACT_SWIGLU;
    const bool geglu_bwd = a->act == PT_ACT_GEGLU_BWD || a->act == PT_ACT_SWIGLU_BWD;
    if (geglu || geglu_bwd) {
        PT_REQUIRE(a->aux != nullptr && (reinterpret_cast<uintptr_t>(a->aux) & 15) == 0 && a->ld_aux % 8 == 0,
                   "pt_gemm_bf16: fused GeGLU needs a 16B-aligned aux matrix with ld_aux %% 8 == 0");
        PT_REQUIRE(a->c_dtype == PT_BF16 && !a->bias && !a->residual && !a->C_pre && a->alpha == 1.0f,
                   "pt_gemm_bf16: fused GeGLU takes bf16 C and no bias / residual / C_pre / alpha");
        PT_REQUIRE(a->nz[0] * a->nz[1] * a->nz[2] == 1 && kg == 1, "pt_gemm_bf16: fused GeGLU is not batched");
        const int ff = geglu ? a->N / 2 : a->N;
        PT_REQUIRE(ff % 128 == 0 && (!geglu || a->N == 2 * ff), "pt_gemm_bf16: fused GeGLU needs ff %% 128 == 0 (ff=%d)", ff);
        PT_REQUIRE(a->ldc >= 2 * ff && a->ld_aux >= (geglu ? ff : 2 * ff), "pt_gemm_bf16: fused GeGLU: ldc / ld_aux too small");
    }
    PT_REQUIRE(a->K2 >= 0, "pt_gemm_bf16: K2 must be >= 0");
    if (a->K2 > 0) {
        PT_REQUIRE(a->A2 && a->B2, "pt_gemm_bf16: K2 > 0 needs A2 and B2");
        PT_REQUIRE(a->nz[0] * a->nz[1] * a->nz[2] == 1 && kg == 1 && a->band == PT_BAND_NONE,
                   "pt_gemm_bf16: the second contraction segment is not batched, grouped or banded");
    }
    if (a->ksplit > 1 || a->ksplit == -1) {
        PT_REQUIRE(a->c_dtype == PT_F32 && !a->bias && !a->residual && !a->C_pre && a->act == PT_ACT_NONE &&
                       a->band == PT_BAND_NONE && kg == 1 && a->K2 == 0,
                   "pt_gemm_bf16: split-K needs f32 C and no bias / act / residual / C_pre / band / K groups / K2");
    } else {
        PT_REQUIRE(a->ksplit == 0 || a->ksplit == 1, "pt_gemm_bf16: bad ksplit %d", a->ksplit);
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
        PT_REQUIRE(a->ld_aux % 4 == 0 && (reinterpret_cast<uintptr_t>(a->aux) & 7) == 0, "pt_gemm_bf16: P misaligned");
        for (int i = 0; i < 3; ++i)
            PT_REQUIRE(a->c_zs[i] % a->ldc == 0, "pt_gemm_bf16: fused softmax backward needs c_zs %% ldc == 0");
    }
    p.aux = a->aux; p.ld_aux = a->ld_aux; p.ff = geglu ? a->N / 2 : a->N;
    if (a->act == PT_ACT_ATTN_DELTA) {
        p.delta = static_cast<float*>(a->aux2);
        p.dl_S = a->aux_i0; p.dl_hd = a->aux_i1; p.dl_hq = a->N / a->aux_i1;
    }
    if (a->act == PT_ACT_CE_STATS) p.delta = static_cast<float*>(a->aux2);
    if (a->act == PT_ACT_QKNORM_ROPE) {
        p.delta = static_cast<float*>(a->aux2);              // (bf16 rows: the kernel casts)
        p.dl_S = a->aux_i0; p.dl_hq = a->aux_i1; p.dl_hd = 256; p.qn_eps = a->eps;
    }
    PT_REQUIRE(a->band >= PT_BAND_NONE && a->band <= PT_BAND_K_BY_KEY && a->band_window >= 0,
               "pt_gemm_bf16: bad band %d / window %d", a->band, a->band_window);
    p.band = a->band; p.band_w = a->band_window;

    const long long nbatch = static_cast<long long>(a->nz[0]) * a->nz[1] * a->nz[2];
    cudaStream_t st = static_cast<cudaStream_t>(stream);
    CUtensorMap ta, tb;
    // CTA-pair kernel (gemm2_sm100.cu): 256 x bn tiles, one per cluster of two SMs
    const bool wide_only = geglu || a->act == PT_ACT_ATTN_DELTA || a->act == PT_ACT_CE_STATS || a->act == PT_ACT_QKNORM_ROPE;
    int bn = wide_only ? 256 : pick_bn2(a->M, a->N, nbatch, p.b_mn);
    if (!wide_only && a->bn_hint != 0) {
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
    if (a->ksplit > 1 || a->ksplit == -1) {
        const int nkb = (a->K + BK - 1) / BK;
        const long long tiles = mt * nt * nbatch, slots = num_sms() / 2;
        long long want = a->ksplit;
        if (a->ksplit == -1) {
            // the slice count that fills the waves best: a launch of tiles * n clusters' worth of work runs
            // ceil(tiles * n / slots) waves; n between "one wave" and four times that, smaller n on ties
            const long long lo = tiles >= slots ? 1 : (slots + tiles - 1) / tiles;
            double best = -1.0;
            want = lo;
            for (long long n = lo; n <= 4 * lo && n <= nkb; ++n) {
                const long long waves = (tiles * n + slots - 1) / slots;
                const double eff = static_cast<double>(tiles * n) / static_cast<double>(waves * slots);
                if (eff > best + 0.02) { best = eff; want = n; }
            }
        }
        if (want > nkb) want = nkb;
        p.kb_per = static_cast<int>((nkb + want - 1) / want);
        p.ksplit = (nkb + p.kb_per - 1) / p.kb_per;          // every slice non-empty
        PT_REQUIRE(tiles * p.ksplit < (1ll << 31), "pt_gemm_bf16: too many tiles");
    }
    CUtensorMap t_aux = ta, t_c = ta, ta2 = ta, tb2 = tb;
    p.K2 = a->K2;
    if (a->K2 > 0) {
        const int one[3] = {1, 1, 1};
        const long long none[3] = {0, 0, 0};
        rc = encode_operand(&ta2, a->A2, p.a_mn, a->M, a->K2, a->lda2, one, none, BM, "A2");
        if (rc != PT_OK) return rc;
        rc = encode_operand(&tb2, a->B2, p.b_mn, a->N, a->K2, a->ldb2, one, none, bn / 2, "B2");
        if (rc != PT_OK) return rc;
    }
    if (geglu_bwd) {
        // the GeGLU-backward epilogue moves its gate / up tiles in and its d_gate / d_up tiles out with TMA
        rc = encode_tiled_bf16_2d(&t_aux, a->aux, 2ll * p.ff, a->M, a->ld_aux, 32, 32, 64, "gu (epilogue)");
        if (rc != PT_OK) return rc;
        rc = encode_tiled_bf16_2d(&t_c, a->C, 2ll * p.ff, a->M, a->ldc, 32, 32, 64, "d_gu (epilogue)");
        if (rc != PT_OK) return rc;
    } else if (tile_epi) {
        rc = encode_tiled_bf16_2d(&t_aux, a->aux, a->N, a->M, a->ld_aux, 32, 32, 64, "aux (epilogue)");
        if (rc != PT_OK) return rc;
        rc = encode_tiled_bf16_2d(&t_c, a->C, a->N, a->M, a->ldc, 32, 32, 64, "C (epilogue)");
        if (rc != PT_OK) return rc;
        if (a->act == PT_ACT_ATTN_DELTA &&
            cudaMemsetAsync(p.delta, 0, static_cast<size_t>(a->M) * p.dl_hq * sizeof(float), st) != cudaSuccess)
            return check_launch("attn delta memset");
    }
    return launch_gemm2(bn, ta, tb, t_aux, t_c, ta2, tb2, p, st);
}
