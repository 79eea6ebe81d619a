/*
 * pt_stage1.h — C ABI of libpt_stage1.so: the B200 (sm_100a) kernels behind the
 * Stage-1 projector-alignment step of SabaPivot/ProjectionTrainer.
 *
 * The reference has no FFI of its own (pure Python on HF transformers); each
 * entry point below names the reference / HF call site whose arithmetic it
 * replaces.  "ref:" paths are relative to the reference repo, "HF:" paths are
 * relative to the transformers package the reference calls into.
 *
 * Conventions
 *   - plain C, no torch types: raw device pointers, explicit sizes/strides
 *     (strides in ELEMENTS unless a name ends in _bytes), explicit stream
 *     (a cudaStream_t passed as void*; NULL = legacy default stream);
 *   - every call returns 0 (PT_OK) or a negative pt_status; the message of the
 *     last failure on the calling thread is pt_last_error();
 *   - the caller owns all memory; no call allocates or synchronises;
 *   - no CPU fallback exists: on a machine without an sm_100 device every
 *     compute entry fails with PT_ERR_CUDA.
 */
#ifndef PT_STAGE1_H_
#define PT_STAGE1_H_

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
    PT_OK = 0,
    PT_ERR_INVALID = -1, /* bad argument / unsupported shape */
    PT_ERR_CUDA = -2,    /* CUDA runtime or driver failure */
} pt_status;

typedef enum { PT_BF16 = 0, PT_F32 = 1, PT_I64 = 2, PT_I32 = 3 } pt_dtype;
typedef enum {
    PT_ACT_NONE = 0, PT_ACT_GELU_TANH = 1, PT_ACT_GELU_ERF = 2,
    /* GeGLU fused into the gate||up projection (HF:models/gemma3/modeling_gemma3.py:127-129): N = 2*ff,
     * B = [gate rows ; up rows]; C [M,2*ff] bf16 receives the pre-activations (kept for the backward) and
     * aux [M,ff] bf16 (row stride ld_aux) receives gelu_tanh(gate) * up.  Needs ff % 128 == 0. */
    PT_ACT_GEGLU = 3,
    /* Backward of the above fused into the down-projection input gradient: the accumulator is
     * d_act [M, N = ff]; aux = the saved pre-activations [M,2*ff] bf16 (row stride ld_aux, read only);
     * C [M,2*ff] bf16 receives [d_act * up * gelu_tanh'(gate) | d_act * gelu_tanh(gate)]. */
    PT_ACT_GEGLU_BWD = 4,
    /* Softmax backward fused into dP = dO V^T (HF eager attention under autograd): aux = the saved
     * probabilities P, bf16, laid out exactly like C (rows contiguous over all batches, row stride ld_aux,
     * c_zs multiples of ldc); bias = delta, f32, one value per C row (row index = element offset / ldc) with
     * delta[r] = sum_k P[r,k] dP[r,k] = rowsum(dO * O) (pt_attn_delta); C (bf16) receives
     * alpha * P * (dP - delta), exactly 0 where P == 0. */
    PT_ACT_SOFTMAX_BWD = 5,
    /* PT_ACT_GEGLU / PT_ACT_GEGLU_BWD with SiLU in place of tanh-GELU: the SwiGLU MLP of the Qwen3 decoder
     * (HF:models/qwen3/modeling_qwen3.py Qwen3MLP: down(silu(gate) * up)); same operands and layout. */
    PT_ACT_SWIGLU = 6,
    PT_ACT_SWIGLU_BWD = 7,
    /* Backward of nn.GELU() (ref:Stage1/projectors.py:18) fused into the projector's second-layer input gradient:
     * the accumulator is dH [M, N]; aux = the saved pre-activations Z [M, N] bf16 (row stride ld_aux, read only);
     * C [M, N] bf16 receives dH * gelu_erf'(Z) (exact erf).  Not batched; N % 32 == 0. */
    PT_ACT_GELU_ERF_BWD = 8,
    /* delta = rowsum(dO * O) of the attention backward fused into the o-projection input gradient: the accumulator
     * is dO [M = B*S, N = hq*hd]; aux = the attention output O, bf16, same shape (row stride ld_aux); C (bf16)
     * receives dO and aux2 (f32 [B*hq, S], zero-filled by pt_gemm_bf16 itself) receives
     * delta[(b*hq + head)*S + s] = sum_d bf16(dO[b*S+s, head*hd + d]) * O[...], with S = aux_i0, hd = aux_i1
     * (hd % 128 == 0, N % 256 == 0).  Replaces pt_attn_delta + one read of dO and O. */
    PT_ACT_ATTN_DELTA = 9,
    /* Cross-entropy statistics of the LM head (HF:models/gemma3/modeling_gemma3.py:649-656 + HF:loss/loss_utils.py:45-67)
     * in the epilogue of the logits GEMM: C [M, N] bf16 receives the logits as usual and aux2 (f32
     * [2 * ceil(N / 256)][M][2]) receives, per row and per 128-column half tile, the running maximum and
     * sum(exp(x - max)) of the bf16-ROUNDED logits (what a pass over the stored logits would see).
     * pt_ce_from_partials turns them into the row log-sum-exps: one 2.1 GB read of the logits less per step.
     * No bias / residual / C_pre / alpha, not batched; 256-wide tiles. */
    PT_ACT_CE_STATS = 10,
    /* Per-head q/k RMSNorm + rotary embedding (pt_qknorm_rope_fwd, HF:models/gemma3/modeling_gemma3.py:362-368,233-263)
     * in the epilogue of the fused q|k|v projection, head_dim 256 (a 256-wide tile IS one head): C [M, N] bf16
     * receives the projection as usual (the backward reads it), aux [M, aux_i1 * 256] bf16 (row stride ld_aux)
     * receives the rotated q heads and aux2 [M, hkv * 256] bf16 (dense) the rotated k heads, hkv = (N / 256 - aux_i1) / 2.
     * bias = the norm gains f32 [512] = [q gains | k gains] (applied as 1 + w); residual = the rotary table f32
     * [aux_i0, 256] = [cos (128) | sin (128)] per position, position of row r = r %% aux_i0; eps = the norm's epsilon.
     * The sum of squares runs over the bf16-rounded projection, as the separate kernel's does.  Not batched. */
    PT_ACT_QKNORM_ROPE = 11
} pt_act;

typedef enum {
    PT_BAND_NONE = 0,
    PT_BAND_SCORES = 1,      /* rows = queries, columns = keys (Q K^T, dO V^T): masked tiles skipped */
    PT_BAND_K_BY_QUERY = 2,  /* rows = queries, contraction over keys (P V, dS K) */
    PT_BAND_K_BY_KEY = 3     /* rows = keys, contraction over queries (P^T dO, dS^T Q) */
} pt_band;

const char* pt_last_error(void);
int pt_version(void);
/* 0 when a compute-capability 10.x device is current, PT_ERR_CUDA otherwise. */
int pt_device_check(void);
/* Programmatic dependent launch for the kernels that support it: `kinds` is a bit set — 1: pt_gemm_bf16, 2:
 * pt_rmsnorm_residual_fwd[_acc] on rows of <= 4096, 4: pt_decode_attn.  A kernel of an enabled kind may become
 * resident while its predecessor in the stream drains, its prologue (barrier init, TMEM allocation, tensor-map
 * prefetch) overlapping the predecessor's tail; every such kernel waits for the predecessor's completion before its
 * first dependent memory access, so results do not change.  Process-wide, 0 by default; returns the previous setting.
 * No counterpart in the reference (CUDA launch plumbing). */
int pt_set_dependent_launch(int kinds);

/* ------------------------------------------------------------------------- *
 * Tensor-core GEMM (tcgen05.mma, TMEM accumulators, TMA-fed smem pipeline).
 *
 *   C[z][m][n] = epilogue( alpha * sum_k A[z][m][k] * B[z][n][k] )
 *   epilogue(v) = act(v + bias[n]) (+ residual[z][m][n])
 *
 * A is [M,K] and B is [N,K], both bf16 with K contiguous ("Y = X W^T" with
 * nn.Linear weight layout).  Replaces every torch.nn.functional.linear /
 * torch.matmul on the path: ref:Stage1/projectors.py:16-20 (projector),
 * HF:models/siglip/modeling_siglip.py:124-130,285-309,324-326,
 * HF:models/gemma3/modeling_gemma3.py:127-129,357-387,652 and their autograd
 * input-gradient products.  With b_mn_major / a_mn_major set the operand is
 * instead stored [K,N] / [K,M] with N / M contiguous (used for dX = dY W and
 * for weight gradients without a transpose copy).
 *
 * The batch index z = (z0,z1,z2) (z0 fastest).  An operand whose stride for a
 * batch dim is 0 is broadcast along it.  Alignment: A/B base pointers 16 bytes,
 * lda/ldb and all A/B batch strides multiples of 8 elements; ldc (and ldr)
 * multiples of 8 (bf16) or 4 (f32) elements with 16-byte aligned bases.  Output columns are produced in
 * quads: when N % 4 != 0 the columns N .. roundup4(N)-1 of C (which exist, since ldc >= roundup4(N)) are
 * overwritten too (with zeros when there is no bias), and bias / residual require N % 4 == 0.
 * ------------------------------------------------------------------------- */
typedef struct {
    int32_t M, N, K;
    int32_t nz[3];            /* batch extents, each >= 1 */

    const void* A;            /* bf16 */
    int64_t lda;              /* row stride (elements) */
    int64_t a_zs[3];          /* batch strides (elements); 0 = broadcast */
    int32_t a_mn_major;       /* 0: A[m][k] (K contiguous); 1: A[k][m] */

    const void* B;            /* bf16 */
    int64_t ldb;
    int64_t b_zs[3];
    int32_t b_mn_major;       /* 0: B[n][k]; 1: B[k][n] */

    void* C;                  /* bf16 or f32, row-major [M,N] */
    int32_t c_dtype;          /* PT_BF16 | PT_F32 */
    int64_t ldc;
    int64_t c_zs[3];

    void* C_pre;              /* optional bf16 [M,N] (same ldc/c_zs): v + bias before act */

    const float* bias;        /* optional f32 [N] */
    const void* residual;     /* optional [M,N], added after act; may alias C */
    int32_t r_dtype;          /* PT_BF16 | PT_F32 */
    int64_t ldr;
    int64_t r_zs[3];

    int32_t act;              /* pt_act */
    float alpha;              /* scale on the accumulator */
    /* Optional grouped contraction: the K loop runs over kg groups of K each,
     * group g reading A + g*a_kgs / B + g*b_kgs (GQA: sum over the q-heads that
     * share one kv-head).  kg <= 1 disables it; with kg > 1, nz[2] must be 1. */
    int32_t kg;
    int64_t a_kgs, b_kgs;
    /* Causal / sliding-window band of a self-attention product (M = K or M = N = sequence length): tiles and
     * K blocks that only meet masked (query, key) pairs are skipped.  Their outputs are NOT written
     * (PT_BAND_SCORES) — consumers must ignore masked positions, as pt_softmax_fwd / pt_softmax_bwd do — and
     * the skipped K blocks must hold zeros in the probability operand (they do: the softmax kernels write 0).
     * band_window = sliding window (key k visible to query q iff q - window < k <= q), 0 = causal only. */
    int32_t band;             /* pt_band */
    int32_t band_window;
    /* Second matrix of the PT_ACT_GEGLU / PT_ACT_GEGLU_BWD epilogues (see pt_act); NULL otherwise. */
    void* aux;
    int64_t ld_aux;
    /* Output-tile width override for the CTA-pair kernel: 0 = cost model, else 64 / 128 / 192 / 256 (a tuning
     * hint for few-tile problems whose last wave would leave most clusters idle; results do not depend on it). */
    int32_t bn_hint;
    /* PT_ACT_ATTN_DELTA only (see pt_act): f32 output [B*hq, S], sequence length, head dim; NULL / 0 otherwise. */
    void* aux2;
    int32_t aux_i0, aux_i1;
    /* Optional second contraction segment: the accumulator also takes sum_k A2[m][k] * B2[n][k] over K2 before the
     * epilogue, i.e. C = epilogue(alpha * ([A | A2] . [B | B2]^T)).  A2 / B2 are bf16 with the SAME major-ness as
     * A / B (a_mn_major / b_mn_major), row strides lda2 / ldb2, no batch dims (nz = 1, kg <= 1, band none); K2 = 0
     * disables it.  This is how a LoRA adapter rides the base projection: A2 = t = x A_lora^T, B2 = B_lora
     * (peft's `result + lora_B(lora_A(x)) * scaling`, ref:Stage2/train_vqa_stage2.py:225-250), and for the input
     * gradient A2 = dt, B2 = A_lora read MN-major — one pass over the output, every fused epilogue still applies. */
    const void* A2;
    int64_t lda2;
    const void* B2;
    int64_t ldb2;
    int32_t K2;
    /* Split-K: 0 / 1 = off.  n > 1 cuts the contraction into (at most) n slices that run as separate tiles and ADD
     * their partial products into C with f32 atomics: C += alpha * A . B^T, so the caller clears (or pre-loads) C.
     * -1 = pick n so that the launch has about one tile per SM pair.  Needs f32 C and no bias / act / residual /
     * C_pre / band / K groups / K2.  For products with a long contraction and a small output (the adapters' weight
     * gradients); the summation order, hence the last bits of C, varies from run to run. */
    int32_t ksplit;
    float eps;                /* PT_ACT_QKNORM_ROPE: RMSNorm epsilon */
} pt_gemm_args;

int pt_gemm_bf16(const pt_gemm_args* args, void* stream);


/* ------------------------------------------------------------------------- *
 * Bandwidth-bound kernels.  Row-group remapping: when group_rows > 0, logical
 * row r lives at physical row (r / group_rows) * group_stride + group_row0 +
 * r % group_rows of the strided side (used to drop token 0 of every image,
 * ref:Stage1/projector_trainer.py:173, and to address only label rows).
 * ------------------------------------------------------------------------- */

/* f32 [B,3,H,W] -> bf16 [B*(H/p)*(W/p), 3*p*p], column order (c,py,px): the im2col of
 * Conv2d(3,d,p,stride=p) — HF:models/siglip/modeling_siglip.py:124-130,178-180. */
int pt_patchify(const float* pixels, void* cols_bf16, int B, int H, int W, int patch, void* stream);

/* LayerNorm (biased variance), f32 in, bf16 out — HF:models/siglip/modeling_siglip.py:348,357,618. */
int pt_layernorm_fwd(const float* x, const float* gamma, const float* beta, void* out_bf16,
                     int rows_out, int D, float eps, int group_rows_out, int group_rows_in,
                     int group_row0, void* stream);

/* Gemma3 RMSNorm: x * rsqrt(mean x^2 + eps) * (1 + w), fp32 inside — HF:models/gemma3/modeling_gemma3.py:138-146. */
int pt_rmsnorm_fwd(const void* x, int x_dtype, const float* w, void* out_bf16, int rows_out, int D,
                   float eps, int group_rows_out, int group_rows_in, int group_row0, void* stream);

/* Sandwich-norm residual step of the decoder layer (HF:models/gemma3/modeling_gemma3.py:417-432):
 *   x_out = x_in + RMS(y; w_post);   next = RMS(x_out; w_next)   (w_next/next may both be NULL) */
int pt_rmsnorm_residual_fwd(const void* y_bf16, const float* w_post, const float* x_in, float* x_out,
                            const float* w_next, void* next_bf16, int rows, int D, float eps,
                            void* stream);
/* The same step behind a split-K GEMM (the decode step's o-projection and down-projection: 64 rows, where one
 * CTA pair per 64 output columns would stream the whole contraction alone): y_acc f32 [rows, D] holds the sum of
 * the GEMM slices' atomic adds; it is rounded to bf16 (what the GEMM's bf16 output would have been), used as y
 * above, and CLEARED in the same pass — the caller zeroes it once, every later GEMM finds it empty.  D <= 4096. */
int pt_rmsnorm_residual_fwd_acc(float* y_acc, const float* w_post, const float* x_in, float* x_out,
                                const float* w_next, void* next_bf16, int rows, int D, float eps, void* stream);

/* RMSNorm input gradient.  g: grad of the norm output; u: saved norm input (u_group_* remap);
 * dst bf16: dst = du;  dst f32: dst[remap(r)] = (add ? add[remap(r)] : 0) + du. */
int pt_rmsnorm_bwd(const void* g, int g_dtype, const void* u, int u_dtype, const float* w, void* dst,
                   int dst_dtype, const float* add_f32, int rows, int D, float eps, int u_group_rows,
                   int u_group_stride, int u_group_row0, int dst_group_rows, int dst_group_stride,
                   int dst_group_row0, void* stream);
/* Two consecutive norm backwards of the sandwich-norm layer in one pass (D <= 2560):
 *   dx += rmsnorm_bwd(g1; u1, w1);  out2 = bf16(rmsnorm_bwd(dx; u2, w2))
 * g1 bf16 [rows,D] (gradient of a pre-norm output), u1 f32 (the residual stream it normalised), dx f32 in/out,
 * u2 bf16 (the branch output the preceding post-norm normalised). */
int pt_rmsnorm_bwd_pair(const void* g1_bf16, const float* u1, const float* w1, float* dx, const void* u2_bf16,
                        const float* w2, void* out2_bf16, int rows, int D, float eps, void* stream);

/* Per-head q/k RMSNorm + rotary embedding on the fused qkv row [(hq+2hkv)*hd]; position = row % S;
 * cos/sin are f32 [S, hd/2] — HF:models/gemma3/modeling_gemma3.py:362-368,233-263. */
int pt_qknorm_rope_fwd(const void* qkv, const float* wq, const float* wk, const float* cos_t,
                       const float* sin_t, void* q_out, void* k_out, int rows, int S, int hq, int hkv,
                       int hd, float eps, void* stream);
int pt_qknorm_rope_bwd(const void* dq, const void* dk, const void* qkv, const float* wq, const float* wk,
                       const float* cos_t, const float* sin_t, void* d_qkv, int rows, int S, int hq,
                       int hkv, int hd, float eps, void* stream);

/* Row softmax of pre-scaled f32 scores [Z,Sq,ld] -> bf16 probabilities [Z,Sq,ld]; key k is allowed
 * for query q iff (!causal || k <= q) && (window <= 0 || k > q - window) && keyvalid[z/heads_per_batch][k]
 * — HF:masking_utils.py:74-99,148-159; HF:models/gemma3/modeling_gemma3.py:370-384 (SigLIP: no mask). */
int pt_softmax_fwd(const float* scores, void* P_bf16, const uint8_t* keyvalid, int Z, int Sq, int Sk,
                   int ld, int heads_per_batch, int causal, int window, void* stream);
/* dS = scale * P * (dP - rowsum(P dP)).  Row r belongs to query r % Sq; with Sq > 0 the same causal / window
 * band as the forward bounds the columns that are read (P is 0 and dP may be unwritten outside it, see
 * pt_gemm_args.band); Sq = 0 reads whole rows. */
int pt_softmax_bwd(const void* P_bf16, const float* dP, void* dS_bf16, long long rows, int Sk, int ld,
                   float scale, int Sq, int causal, int window, void* stream);

/* Fused causal / sliding-window GQA attention forward of the Gemma3 decoder, head_dim 256 or 128, flash style
 * (HF:models/gemma3/modeling_gemma3.py:278-300,370-384; mask HF:masking_utils.py:74-99,148-159):
 *   O[b*S + q, head*hd ..] = sum_k softmax_k(mask(q_rot . k_rot * scale)) V     (bf16, [B*S, hq*hd])
 *   lse[(b*hq + head)*S + q] = log2(sum_k 2^(score * scale * log2 e))           (f32; +inf for a row without a
 *                              visible key; NULL = not wanted: inference)
 * q_rot [B*S, hq*hd], k_rot [B*S, hkv*hd] are the outputs of pt_qknorm_rope_fwd; V is read in place from the
 * fused activation qkv [B*S, (hq+2hkv)*hd].  kvbits: optional key-padding bit masks from pt_pack_keyvalid
 * (NULL = every key valid).  window 0 = causal only.  hq/hkv a power of two.
 * One pass over the key tiles with an online softmax; scores and probabilities never reach HBM. */
int pt_attn_gemma_fwd(const void* q_rot, const void* k_rot, const void* qkv, const void* kvbits, float* lse,
                      void* O_bf16, int B, int S, int hq, int hkv, int hd, int window, float scale, void* stream);
/* One cached decode step of the same attention in one launch (the per-token forward of `generate`,
 * ref:Stage1/inference_stage1.py:170 -> HF:models/gemma3/modeling_gemma3.py:278-300,362-384 with a DynamicCache):
 * per (sample, kv head): q/k RMSNorm + RoPE of this token's heads from the fused row qkv [B, (hq+2hkv)*hd]
 * (as pt_qknorm_rope_fwd; table row pos[b] of cos_t / sin_t [table_rows, hd/2]), the cache append
 * K[b, slot] = k_rot, V[b, slot] = v (K, V bf16 [B, Smax, hkv*hd]; slot = *slot, a device scalar so a captured graph
 * serves every token), and out[b, head*hd ..] = softmax_k(scale * q_rot . K[b, k]) V[b, k] over the keys
 * k in [window > 0 ? max(0, slot - window + 1) : 0, slot] with valid[b*Smax + k] != 0 (NULL = all valid).
 * The keys are split over a cluster of nsplit CTAs (1, 2, 4, 8; <= 0: chosen from B * hkv) that merge through
 * distributed shared memory.  hd in {16, 32, 64, 128, 256}, hq / hkv in {1, 2, 4, 8}
 * (pt_decode_attn_supported); other geometries are rejected — the caller keeps the GEMM + softmax launches. */
int pt_decode_attn_supported(int hq, int hkv, int hd);
int pt_decode_attn(const void* qkv, const float* wq, const float* wk, const float* cos_t, const float* sin_t,
                   int table_rows, const long long* pos, const long long* slot, const uint8_t* valid, void* K,
                   void* V, void* out_bf16, int B, int Smax, int hq, int hkv, int hd, int window, float eps,
                   float scale, int nsplit, void* stream);
/* Linear layer at small batch (M <= 32 rows; the decode step of the reference's one-image-at-a-time inference,
 * ref:Stage1/inference_stage1.py:150-193 -> torch.nn.functional.linear in HF:models/gemma3/modeling_gemma3.py:127-129,
 * 357-387, 652): out[m][n] = sum_k x[m][k] W[n][k], x bf16 [M, K] (row stride lda), W bf16 [N, K] (row stride ldb),
 * K % 32 == 0, pure weight streaming on mma.sync with the weight rows as the MMA's M dimension.
 * mode 0: out bf16 [M, N]; 1: out f32; 2: out f32 += (atomic; ksplit K slices, <= 0: chosen; the accumulator of
 * pt_rmsnorm_residual_fwd_acc); 3 / 4: GeGLU / SwiGLU — W holds 2 N rows [gate rows ; up rows], out bf16 [M, N] =
 * act(bf16(gate)) * bf16(up) like pt_gemm_bf16's PT_ACT_GEGLU / PT_ACT_SWIGLU epilogue (the pre-activations are not
 * stored: inference). */
int pt_gemv_small_m(const void* x, long long lda, int M, const void* W, long long ldb, int N, int K, void* out,
                    long long ldc, int mode, int ksplit, void* stream);
/* The next token of every row of `generate` in one launch (HF:generation/utils.py `_sample` with the logits
 * processors / warpers of `_get_logits_processor`, as called at ref:Stage1/inference_stage1.py:176-186,
 * ref:Stage1/projector_trainer.py:386-394): repetition penalty over prev[b, 0..nprev) (x < 0 ? x * rp : x / rp; NULL or
 * rp == 1: none) -> / temperature -> top-k (values below the k-th largest go, ties with it stay; 1 <= top_k <= 1024)
 * -> top-p (ascending cumulative probability <= 1 - top_p goes, the largest stays; >= 1: none) -> softmax -> one draw
 * by inverse CDF over the surviving tokens in descending order at u[b] in [0, 1) (torch.multinomial's distribution,
 * not its random stream); top_k == 0: argmax of the penalised logits (smallest index among ties).  Then HF's
 * bookkeeping: a row with done[b] != 0 emits pad_id (if >= 0); out[b * ld_out] = token; next[b] = token (optional);
 * done[b] |= token == eos_id (if >= 0).  logits f32 [B, ld >= V]; with a repetition penalty the penalised entries of
 * logits are overwritten with their penalised values (the selection then reads the row as it is). */
int pt_sample_token(float* logits, long long ld, int B, int V, const long long* prev, long long ld_prev, int nprev,
                    float repetition_penalty, float temperature, int top_k, float top_p, const float* u, uint8_t* done,
                    long long pad_id, long long eos_id, long long* out, long long ld_out, long long* next, void* stream);
/* Fused backward of the same attention (autograd through HF:models/gemma3/modeling_gemma3.py:370-384, invoked at
 * ref:Stage1/projector_trainer.py:236-237), key/value stationary, P recomputed from q_rot, k_rot and lse:
 *   dV[b*S + k, kvh*hd ..] = sum_{heads of kvh} P^T dO          (bf16, row stride ld_dv elements)
 *   dK[b*S + k, kvh*hd ..] = sum_{heads of kvh} dS^T q_rot      (bf16, row stride ld_dk; gradient w.r.t. k_rot)
 *   dS[(b*hq + head), q, k] = scale * P * (dO V^T - delta)      (bf16, [B*hq, S, lds]; written only inside the
 *                             (query block, key tile) pairs of the causal / window band — the caller zero-fills
 *                             the buffer once per window value; masked positions inside get exact zeros)
 * delta [B*hq, S] = rowsum(dO * O) (pt_attn_delta).  dQ = dS k_rot is left to a band-limited pt_gemm_bf16
 * (PT_BAND_K_BY_QUERY): the one contraction that is stationary in the other operand. */
int pt_attn_gemma_bwd(const void* q_rot, const void* k_rot, const void* qkv, const void* dO_bf16, const void* kvbits,
                      const float* lse, const float* delta, void* dS_bf16, void* dK_bf16, long long ld_dk,
                      void* dV_bf16, long long ld_dv, int B, int S, int hq, int hkv, int hd, int window, float scale,
                      int lds, void* stream);
/* The same backward on a CTA PAIR per tile of 128 keys (attn_gemma_bwd2_sm100.cu: one SM keeps the K tile and forms
 * P^T and dV, its partner keeps the V tile and forms dS^T and dK; P^T crosses by a bulk DSMEM copy) — about twice as
 * fast at head_dim 256.  Difference to pt_attn_gemma_bwd: dS leaves TRANSPOSED:
 *   dST[(b*hq + head), k, q] = scale * P[q, k] * (dO V^T - delta)[q, k]     (bf16, [B*hq, S keys, ldt]; written only
 *                              inside the (128-key tile, 64-position block) pairs of the band; zero-fill once per
 *                              window value)
 * so dQ = dS k_rot is a band-limited pt_gemm_bf16 with dST as an MN-major A operand (a_mn_major = 1). */
int pt_attn_gemma_bwd_pair(const void* q_rot, const void* k_rot, const void* qkv, const void* dO_bf16,
                           const void* kvbits, const float* lse, const float* delta, void* dST_bf16, void* dK_bf16,
                           long long ld_dk, void* dV_bf16, long long ld_dv, int B, int S, int hq, int hkv, int hd,
                           int window, float scale, int ldt, void* stream);
/* keyvalid u8 [B,S] (0 = padded key, ref:Stage1/projector_trainer.py:205-208) -> u64 [B, ceil(S/64)] bit masks
 * (bit i of word w = key 64 w + i; keys >= S are 0). */
int pt_pack_keyvalid(const uint8_t* keyvalid, void* kvbits, int B, int S, void* stream);

/* delta[(b*hq + head)*S + s] = sum_d dO[b,s,head,d] * O[b,s,head,d] over dO, O bf16 [rows = B*S, hq*hd]: the
 * row term sum_k P dP of the softmax backward, for pt_gemm_bf16's PT_ACT_SOFTMAX_BWD epilogue. */
int pt_attn_delta(const void* dO_bf16, const void* O_bf16, float* delta, int rows, int S, int hq, int hd,
                  void* stream);

/* Fused non-causal attention forward of the SigLIP tower (HF:models/siglip/modeling_siglip.py:293-306):
 * out[b, q, h*hd .. ] = softmax(q k^T * scale) v, with q/k/v read in place from the fused activation
 * qkv [B*N, 3*H*hd] (column blocks q | k | v, head h at columns h*hd) and out [B*N, H*hd], both bf16.
 * tcgen05 flash-style kernel (96-key K/V tiles streamed by TMA, online softmax with lazy rescale, two CTAs
 * per SM); requires hd == 64, any N. */
int pt_attn_siglip_fwd(const void* qkv_bf16, void* out_bf16, int B, int N, int H, int hd, float scale, void* stream);

/* GeGLU: act = gelu_tanh(g) * u with gu = [g | u] per row — HF:models/gemma3/modeling_gemma3.py:127-129. */
int pt_geglu_fwd(const void* gu, void* act, long long rows, int ff, void* stream);
int pt_geglu_bwd(const void* dact, const void* gu, void* dgu, long long rows, int ff, void* stream);
/* SwiGLU (Qwen3MLP): the same two entries with silu(g) = g * sigmoid(g) in place of gelu_tanh(g). */
int pt_swiglu_fwd(const void* gu, void* act, long long rows, int ff, void* stream);
int pt_swiglu_bwd(const void* dact, const void* gu, void* dgu, long long rows, int ff, void* stream);

/* dZ = dH * d/dz[0.5 z (1 + erf(z/sqrt2))] — backward of nn.GELU() in ref:Stage1/projectors.py:18. */
int pt_gelu_erf_bwd(const void* dH, const void* Z, void* dZ, long long n, void* stream);

/* Device-side input validation.  nn.Embedding and CrossEntropyLoss raise on an out-of-range index; a kernel
 * cannot, so it ORs a bit into the caller's int32 flag word (nullable) and keeps its reads in bounds; the host
 * checks the word where it already synchronises on the loss (Stage1Engine.check_errors). */
#define PT_ERRFLAG_TOKEN_ID 1   /* pt_embed_splice: a token id outside [0, vocab): row 0 was read instead */
#define PT_ERRFLAG_TARGET   2   /* pt_ce_fwd_bwd: a target >= V: the row was left out of the loss */

/* Text half of the splice (ref:Stage1/projector_trainer.py:186-220): x0[b, n_img+t, :] =
 * bf16(embed[token_ids[b,t]] * scale), embed bf16 [vocab, h]; attention mask / lm labels (int64, bit-exact),
 * key-valid bytes and shifted int32 targets (HF:loss/loss_utils.py:57-60) for all S = n_img+T positions. */
int pt_embed_splice(const int64_t* token_ids, const int64_t* labels, const void* embed_bf16, float scale,
                    int64_t pad_id, int has_pad, int B, int T, int n_img, int h, float* x0,
                    int64_t* attn_mask, int64_t* lm_labels, uint8_t* keyvalid, int32_t* targets,
                    int64_t vocab, int32_t* err_flags, void* stream);

/* Masked causal-LM cross-entropy over bf16 logits [rows, V] (HF:loss/loss_utils.py:28-67): writes
 * lse[rows], acc2 = {sum of row losses, #valid rows}, loss = mean (NaN when no row has a label, as HF's mean
 * over an empty set); then overwrites the logits with dlogits = (softmax - onehot) * gscale / n_valid (zero rows
 * where target < 0). */
int pt_ce_fwd_bwd(void* logits_bf16, long long ld, const int32_t* targets, int rows, int V, float* lse,
                  float* acc2, float* loss, float gscale, int32_t* err_flags, void* stream);

/* out[n] = sum_m x[m][n] (bias gradients). */
int pt_colsum(const void* x_bf16, long long ld, float* out, int M, int N, void* stream);
/* The rest of the shifted masked cross-entropy when the logits GEMM ran with PT_ACT_CE_STATS: combines the per-tile
 * partials (stats: f32 [parts][rows][2], parts = 2 * ceil(V / 256)) into lse [rows], forms the loss
 * (mean over targets >= 0 of lse - logit[target]; NaN when there is none) and overwrites the logits with
 * dlogits = (softmax - onehot) * gscale / n_valid, exactly as pt_ce_fwd_bwd does after its own statistics pass. */
int pt_ce_from_partials(void* logits_bf16, long long ld, const int32_t* targets, int rows, int V, const float* stats,
                        int parts, float* lse, float* acc2, float* loss, float gscale, int32_t* err_flags, void* stream);
/* out[0] = sum g^2 */
int pt_sumsq(const float* g, long long n, float* out, void* stream);
/* clip_grad_norm_(max_norm) + AdamW step on a flat f32 buffer, refreshing a bf16 shadow copy
 * (ref:Stage1/projector_trainer.py:241-243).  sumsq: device scalar from pt_sumsq of the raw grads;
 * grad_scale multiplies grads first (1/world after an allreduce-sum).  step is 1-based. */
int pt_clip_adamw(float* p, const float* g, float* m, float* v, void* shadow_bf16, long long n,
                  const float* sumsq, float grad_scale, float max_norm, float lr, float beta1, float beta2,
                  float eps, float weight_decay, int step, void* stream);
int pt_cast_f32_bf16(const float* x, void* y_bf16, long long n, void* stream);
/* out[i] = bf16(sum_s x[s * slice_elems + i]), i < slice_elems (% 4 == 0): reduces the f32 partials of a split-K
 * pt_gemm_bf16 launch (K slices as a batch dim) — the LM-head input gradient dlogits @ E of
 * ref:Stage1/projector_trainer.py:236 (accelerator.backward through HF:models/gemma3/modeling_gemma3.py lm_head). */
int pt_sum_slices(const float* x, long long slice_elems, int n_slices, void* out_bf16, void* stream);
/* out[r] = bf16(x[remap(r)]) — e.g. the image-token rows of dL/dX0 handed to the projector backward. */
int pt_gather_cast_rows(const float* x, void* out_bf16, int rows_out, int D, int group_rows_out,
                        int group_rows_in, int group_row0, void* stream);

/* ---- image preprocessing in front of the vision tower -------------------------------
 * ref:Stage1/train_projection_stage1.py:99-104 and ref:Stage2/dataset.py:90-95:
 *     image = image.resize((S, S)); pixel_values = processor(images=image, return_tensors="pt").pixel_values
 * i.e. Pillow's antialiased BICUBIC resize (src/libImaging/Resample.c: 8-bit fixed point, horizontal then
 * vertical pass, each rounded to uint8) followed by HF SiglipImageProcessor rescale (x * rescale in float64 ->
 * float32) and normalize ((x - mean) / std in float32).  src_rgb: device uint8 [H, W, 3] with row stride
 * src_row_stride bytes; dst_chw: device f32 [3, out_h, out_w]; dst_u8_hwc (nullable): the resized uint8 image
 * [out_h, out_w, 3] (bit-exact with Pillow).  workspace: device scratch of at least
 * pt_image_preprocess_workspace(H, W, out_h, out_w) bytes, 16-byte aligned. */
long long pt_image_preprocess_workspace(int H, int W, int out_h, int out_w);
int pt_image_preprocess(const uint8_t* src_rgb, int H, int W, long long src_row_stride, float* dst_chw,
                        uint8_t* dst_u8_hwc, int out_h, int out_w, float mean0, float mean1, float mean2,
                        float std0, float std1, float std2, double rescale, void* workspace,
                        long long workspace_bytes, void* stream);
/* pt_image_preprocess without a caller workspace and without any allocation.  No counterpart in the reference: added
 * for callers that cannot call cudaMalloc (code under CUDA graph capture, tight
 * serving loops, harnesses that forbid it).  The scratch (tap tables + the horizontally resampled intermediate) lives
 * in a file-scope device array of pt_image_preprocess_static_max_bytes() = 8 MiB that every GPU loading the library
 * holds until exit; a resize whose pt_image_preprocess_workspace() exceeds it is rejected with PT_ERR_INVALID.  Same
 * kernels, same results as pt_image_preprocess.  All calls on a device share the array: calls that overlap on the GPU
 * (different streams without an ordering between them) race; calls on one stream are safe.  Nothing needs clearing
 * between calls (every byte read was written earlier in the same call). */
long long pt_image_preprocess_static_max_bytes(void);
int pt_image_preprocess_static(const uint8_t* src_rgb, int H, int W, long long src_row_stride, float* dst_chw,
                               uint8_t* dst_u8_hwc, int out_h, int out_w, float mean0, float mean1, float mean2,
                               float std0, float std1, float std2, double rescale, void* stream);

/* ---- baseline JPEG decode in front of pt_image_preprocess (the decode half) ---------------
 * ref:Stage1/train_projection_stage1.py:99 and ref:Stage2/dataset.py:86-90: `Image.open(path).convert("RGB")`, i.e.
 * Pillow's libjpeg-turbo decode of a baseline file: Huffman decode, de-quantise, islow integer IDCT, fancy chroma
 * up-sampling (h2v1 / h2v2), JFIF YCbCr -> RGB; bit-exact with Pillow.  Host part (no GPU needed, meant for the
 * data-loader workers):
 *   pt_jpeg_plan   parses n files (pointers + sizes); fills widths[n], heights[n], rgb_offsets[n + 1] (byte offset of
 *                  image i in the packed destination, [n] = its total size) and the sizes of the blob and the
 *                  device workspace
 *   pt_jpeg_pack   writes the blob (16-byte aligned host buffer, ideally pinned): per-image geometry, quantisation
 *                  and Huffman look-up tables, and the entropy-coded data with the FF00 stuffing removed, cut at the
 *                  restart markers
 * Device part:
 *   pt_jpeg_decode_batch   host_blob = the packed blob (its header is read on the host), dev_blob = the same bytes
 *                  in device memory; dst_rgb = device uint8, image i at rgb_offsets[i] as [height][width][3];
 *                  workspace = device scratch (256-byte aligned).  One Huffman launch for the whole batch (one warp
 *                  per restart segment; a file without restart markers is one segment), then IDCT and colour
 *                  launches per image.
 * Rejected with PT_ERR_INVALID (nothing falls back): progressive / lossless / arithmetic / 12-bit files, CMYK,
 * multi-scan files, chroma sub-sampling other than 4:4:4 / 4:2:2 / 4:2:0. */
int pt_jpeg_plan(const uint8_t* const* files, const int64_t* sizes, int n, int32_t* widths, int32_t* heights,
                 int64_t* rgb_offsets, int64_t* blob_bytes, int64_t* workspace_bytes);
int pt_jpeg_pack(const uint8_t* const* files, const int64_t* sizes, int n, uint8_t* host_blob, int64_t blob_bytes);
int pt_jpeg_decode_batch(const uint8_t* host_blob, const uint8_t* dev_blob, uint8_t* dst_rgb, void* workspace,
                         int64_t workspace_bytes, void* stream);

/* ---- NF4 frozen LLM weights -----------------------------------------------------------
 * ref:Stage1/train_projection_stage1.py:192-210: `--enable_qlora` loads the frozen LLM through bitsandbytes with
 * BitsAndBytesConfig(load_in_4bit, "nf4", compute dtype bf16, double quant).  These entry points restate
 * bitsandbytes 0.45 quantize_4bit / dequantize_4bit for that configuration (blocksize 64, nested 8-bit absmax
 * with blocksize 256 and a mean offset); the codes are expanded to bf16 once at load time.
 * via_fp16 != 0 reproduces transformers' float16 override of bitsandbytes loads (bf16 -> fp16 before quantising,
 * fp16 -> bf16 after de-quantising).  n: elements of the weight, a multiple of 64.
 *   pt_nf4_quantize:        w bf16 [n] -> packed uint8 [n/2] (first of a pair in the high nibble), absmax f32 [n/64]
 *   pt_nf4_absmax_quantize: absmax f32 [nb] - offset -> qabsmax uint8 [nb], absmax2 f32 [ceil(nb/256)];
 *                           code256: device f32 [256], the sorted dynamic map (create_dynamic_map)
 *   pt_nf4_dequantize:      -> out bf16 [n] = bf16(fp16(NF4[c] * (code256[q] * absmax2 + offset))) */
int pt_nf4_quantize(const void* w_bf16, long long n, int via_fp16, uint8_t* packed, float* absmax, void* stream);
int pt_nf4_absmax_quantize(const float* absmax, long long nb, float offset, const float* code256,
                           uint8_t* qabsmax, float* absmax2, void* stream);
int pt_nf4_dequantize(const uint8_t* packed, const uint8_t* qabsmax, const float* absmax2, float offset,
                      const float* code256, long long n, int via_fp16, void* out_bf16, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PT_STAGE1_H_ */
