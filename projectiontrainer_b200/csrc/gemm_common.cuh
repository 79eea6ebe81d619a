// Shared between the single-CTA (gemm_sm100.cu) and CTA-pair (gemm2_sm100.cu) GEMM kernels.
#pragma once
#include "pt_common.cuh"

namespace pt {

constexpr int BM = 128;       // rows of A per CTA (UMMA M = 128 for cta_group::1, 256 over a CTA pair)
constexpr int BK = 64;        // 64 bf16 = one 128-byte swizzle row
constexpr int UMMA_K = 16;

struct GemmDev {
    int M, N, K;
    int nz[3];        // batch extents used for tile enumeration
    int kg;           // K groups (>=1)
    int a_mn, b_mn;
    int a_emask, b_emask;   // bit i set: tensor-map extra dim i takes a coordinate
    int kg_slot;      // which extra dim is the K-group (-1: none)
    int z_slot[3];    // which extra dim batch dim i maps to
    void* C;
    void* C_pre;
    const float* bias;
    const void* R;
    long long ldc, ldr;
    long long c_zs[3], r_zs[3];
    int c_f32, r_f32, act;
    float alpha;
    int raster_n;     // tile order inside one batch slice: 1 = n fastest, 0 = m fastest
    // CTA-pair kernel: first row / column of the LAST m / n tile.  When the extent is not a multiple of the
    // tile, the last tile is shifted back to end exactly at the edge (its loads stay in bounds: TMA boxes
    // that are partly or wholly out of bounds were measured to stall the pipeline) and its epilogue skips
    // the rows / columns the previous tile already produced.
    int m_last, n_last;
    // CTA-pair kernel, 256-wide tiles, N % 256 == 128 (the decoder's h = 1152): the last n tile runs as a
    // 128-wide MMA on the 128 columns that are left instead of a shifted full tile that recomputes 128 columns
    // of its neighbour (10 % of the MMA work of an N = 1152 GEMM).  Each CTA still loads a full, in-bounds
    // 128-row B box placed so that its 64 rows are the box's second half.
    int n_half;
    // fused GeGLU epilogues (PT_ACT_GEGLU / PT_ACT_GEGLU_BWD): second matrix, its row stride, and ff
    void* aux;
    long long ld_aux;
    int ff;
    GluConsts glu;    // activation of the fused gated-unit epilogues (tanh-GELU or SiLU)
    // (PT_ACT_CE_STATS re-uses `delta` for its f32 partials [2 * nt][M][2])
    // PT_ACT_ATTN_DELTA: delta [B*hq, S] f32 (zeroed by the host wrapper), row r = b * dl_S + s, head = column / dl_hd
    float* delta;
    int dl_S, dl_hd, dl_hq;
    // causal / sliding-window band of self-attention products (pt_gemm_args.band): work that only meets
    // masked (query, key) pairs is not issued.  band_w = window (0: unbounded below).
    int band, band_w;
    // second contraction segment (pt_gemm_args.K2): the accumulator also takes sum_k A2[m][k] * B2[n][k] over K2,
    // read through a second pair of tensor maps after the main K blocks (LoRA: [x | t] . [W | B]^T in one pass)
    int K2;
    // split-K: > 1 cuts the K blocks into ksplit slices of kb_per blocks, each slice a tile of its own whose epilogue
    // adds (vector atomics) into the f32 C — thin weight-gradient products whose output has fewer tiles than the
    // GPU has SM pairs (LoRA: dA [64, in], dB_i [out_i, r] contracted over all B*S rows)
    int ksplit, kb_per;
    // PT_ACT_QKNORM_ROPE re-uses aux / ld_aux (rotated q), delta (rotated k, bf16), bias (gains [q | k]), R (rotary
    // table [S][cos | sin]), dl_S (S) and dl_hq (query heads); its epsilon:
    float qn_eps;
};

// K-block range [lo, hi) a tile must contract over, or skip = true when the whole tile is masked.
// Tile rows [tm, tm + TM), columns [tn, tn + BN); nkb = K blocks of BK.
__device__ __forceinline__ bool band_span(const GemmDev& p, int tm, int tn, int TM, int BN, int nkb, int& lo, int& hi) {
    lo = 0; hi = nkb;
    if (p.band == PT_BAND_SCORES) {          // rows = queries, columns = keys: key k allowed iff q - W < k <= q
        if (tn > tm + TM - 1) return true;
        if (p.band_w > 0 && tn + BN - 1 <= tm - p.band_w) return true;
    } else if (p.band == PT_BAND_K_BY_QUERY) {   // rows = queries, contraction over keys
        hi = min(nkb, (tm + TM - 1) / BK + 1);
        if (p.band_w > 0) lo = max(0, tm - p.band_w + 1) / BK;
    } else if (p.band == PT_BAND_K_BY_KEY) {     // rows = keys, contraction over queries: k <= q < k + W
        lo = tm / BK;
        if (p.band_w > 0) hi = min(nkb, (tm + TM - 1 + p.band_w - 1) / BK + 1);
    }
    return false;
}

__device__ __forceinline__ float apply_act(float v, int act) {
    if (act == PT_ACT_GELU_TANH) {
        const float k0 = 0.7978845608028654f, k1 = 0.044715f;
        float u = k0 * (v + k1 * v * v * v);
        return 0.5f * v * (1.f + tanh_fast(u));
    } else if (act == PT_ACT_GELU_ERF) {
        return gelu_erf_f(v);
    }
    return v;
}


// host helpers (gemm_sm100.cu)
int encode_operand(CUtensorMap* out, const void* base, int mn_major, int rows, int K, long long ld,
                   const int ext[3], const long long est[3], int rows_box, const char* name);
int encode_tiled_bf16_5d(CUtensorMap* out, const void* base, const unsigned long long dims[5],
                         const unsigned long long strides_bytes[4], const unsigned int box[5], const char* name);
// row-major bf16 matrix [rows, cols] with row stride ld (elements) as a 2-D map; box = box_cols x box_rows;
// swizzle_bytes in {0, 64, 128} (box_cols * 2 must equal it when non-zero)
int encode_tiled_bf16_2d(CUtensorMap* out, const void* base, long long cols, long long rows, long long ld,
                         int box_cols, int box_rows, int swizzle_bytes, const char* name);
// generic bf16 map of rank 3 (dims / box innermost first; strides_bytes for dims 1, 2), 128-byte swizzle
int encode_tiled_bf16_3d(CUtensorMap* out, const void* base, const unsigned long long dims[3],
                         const unsigned long long strides_bytes[2], const unsigned int box[3], const char* name);
// CTA-pair kernel (gemm2_sm100.cu): bn in {64,128,192,256}; tensor map of B must use box rows bn/2.
// t_aux / t_c: 2-D maps (32 x 32 boxes, 64-byte swizzle) of the aux matrix and of C for the epilogues that move their
// tiles with TMA (PT_ACT_GEGLU_BWD); any valid map otherwise (unused).
int launch_gemm2(int bn, const CUtensorMap& ta, const CUtensorMap& tb, const CUtensorMap& t_aux, const CUtensorMap& t_c,
                 const CUtensorMap& ta2, const CUtensorMap& tb2,
                 const GemmDev& p, cudaStream_t st);
int gemm2_tiles(const GemmDev& p, int bn);

}  // namespace pt
