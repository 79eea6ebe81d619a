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
typedef enum { PT_ACT_NONE = 0, PT_ACT_GELU_TANH = 1, PT_ACT_GELU_ERF = 2 } pt_act;

const char* pt_last_error(void);
int pt_version(void);
/* 0 when a compute-capability 10.x device is current, PT_ERR_CUDA otherwise. */
int pt_device_check(void);

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
 * multiples of 8 (bf16) or 4 (f32) elements with 16-byte aligned bases.
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
} pt_gemm_args;

int pt_gemm_bf16(const pt_gemm_args* args, void* stream);

#ifdef __cplusplus
}
#endif
#endif /* PT_STAGE1_H_ */
