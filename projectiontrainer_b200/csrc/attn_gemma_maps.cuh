// Tensor maps shared by the fused Gemma3 attention forward and backward (attn_gemma_sm100.cu,
// attn_gemma_bwd_sm100.cu): stacked-query boxes and 64-key K / V tiles, 128-byte swizzle.
#pragma once
#include "pt_common.cuh"
#include "gemm_common.cuh"

namespace pt {

constexpr int kGqRows = 128;      // stacked query rows per block: 128 / G positions x G heads
constexpr int kGkTile = 64;       // keys per tile

// [B*S, hq*hd] (rotated queries, dO, ...) viewed as (hd, G, S, hkv, B); box = 64 columns x G heads x 128/G
// positions, so that row r of the box is (position s0 + r / G, head kvh*G + r % G).  Positions >= S read as 0.
inline int gemma_stacked_map(CUtensorMap* out, const void* base, int B, int S, int hq, int hkv, int hd,
                             const char* name) {
    const int G = hq / hkv, pos = kGqRows / G;
    const unsigned long long dims[5] = {(unsigned long long)hd, (unsigned long long)G, (unsigned long long)S,
                                        (unsigned long long)hkv, (unsigned long long)B};
    const unsigned long long str[4] = {(unsigned long long)hd * 2, (unsigned long long)hq * hd * 2,
                                       (unsigned long long)G * hd * 2, (unsigned long long)S * hq * hd * 2};
    const unsigned int box[5] = {64, (unsigned)G, (unsigned)pos, 1, 1};
    return encode_tiled_bf16_5d(out, base, dims, str, box, name);
}

// per-kv-head rows inside [B*S, ld] at `base` (column offset folded in by the caller) viewed as
// (hd, S, hkv, B, 1); box = 64 columns x 64 keys.  Keys >= S read as 0.
inline int gemma_kv_map(CUtensorMap* out, const void* base, long long ld, int B, int S, int hkv, int hd,
                        const char* name, int box_rows = kGkTile) {
    const unsigned long long dims[5] = {(unsigned long long)hd, (unsigned long long)S, (unsigned long long)hkv,
                                        (unsigned long long)B, 1};
    const unsigned long long str[4] = {(unsigned long long)ld * 2, (unsigned long long)hd * 2,
                                       (unsigned long long)S * ld * 2, (unsigned long long)S * ld * 2 * B};
    const unsigned int box[5] = {64, static_cast<unsigned>(box_rows), 1, 1, 1};
    return encode_tiled_bf16_5d(out, base, dims, str, box, name);
}

// [B*S, hq*hd] viewed as (hd, S, hq, B, 1): box = 64 columns x `rows` positions of ONE head.  Positions >= S read 0.
inline int gemma_head_rows_map(CUtensorMap* out, const void* base, int B, int S, int hq, int hd, int rows,
                               const char* name) {
    const unsigned long long dims[5] = {(unsigned long long)hd, (unsigned long long)S, (unsigned long long)hq,
                                        (unsigned long long)B, 1};
    const unsigned long long str[4] = {(unsigned long long)hq * hd * 2, (unsigned long long)hd * 2,
                                       (unsigned long long)S * hq * hd * 2, (unsigned long long)S * hq * hd * 2 * B};
    const unsigned int box[5] = {64, (unsigned)rows, 1, 1, 1};
    return encode_tiled_bf16_5d(out, base, dims, str, box, name);
}

// dS^T [B*hq, S keys, ldT] (element (head_b, key, pos)) as (pos, key, head_b); box = 64 positions x 128 keys: the smem
// image is a [128 keys][64 positions] bf16 tile of 128-byte rows, i.e. the K-major MMA operand tile itself.
inline int gemma_dst_map(CUtensorMap* out, const void* base, long long ldT, int B, int S, int hq, const char* name) {
    const unsigned long long dims[3] = {(unsigned long long)S, (unsigned long long)S, (unsigned long long)B * hq};
    const unsigned long long str[2] = {(unsigned long long)ldT * 2, (unsigned long long)S * ldT * 2};
    const unsigned int box[3] = {64, 128, 1};
    return encode_tiled_bf16_3d(out, base, dims, str, box, name);
}

// Work items are sorted by falling length.  CTA (or cluster) c of n takes one item of every run of n items, walking
// the runs back and forth (c, 2n - 1 - c, 2n + c, ...), so that the long item of one run meets a short one of the next:
// the busiest CTA then carries ~10 % more than the mean instead of ~35 % with plain round-robin (attention backward,
// config 3: 68 vs 84 blocks, mean 62).
__device__ __forceinline__ int snake_item(int j, int c, int n) { return j * n + ((j & 1) ? n - 1 - c : c); }

// visibility mask of the 64 keys k0 .. k0+63 for a query that sees keys k_lo <= k <= k_hi, on top of the
// key-padding word kw
__device__ __forceinline__ unsigned long long gemma_tile_mask(unsigned long long kw, int k0, int k_lo, int k_hi) {
    const int hi = k_hi - k0, lo = k_lo - k0;
    if (hi < 0 || lo > 63) return 0ull;
    if (hi < 63) kw &= (2ull << hi) - 1ull;
    if (lo > 0) kw &= ~0ull << lo;
    return kw;
}

}  // namespace pt
