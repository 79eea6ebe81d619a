"""CPU oracle for the NF4 (4-bit NormalFloat, double-quantised) frozen-weight format.  TEST INFRASTRUCTURE ONLY.

Only tests/ may import this module; the product path (projectiontrainer_b200/) never does.

PARITY UNPINNED.  `--enable_qlora` makes the reference load the frozen LLM through bitsandbytes
(ref:Stage1/train_projection_stage1.py:192-210: `BitsAndBytesConfig(load_in_4bit=True, bnb_4bit_quant_type="nf4",
bnb_4bit_compute_dtype=bf16, bnb_4bit_use_double_quant=True)`).  bitsandbytes (requirements.txt pins 0.45.5) is not
under /root/reference and is not installed in this image, so nothing here could be checked against its outputs; the
functions restate its published algorithm from knowledge of `bitsandbytes/functional.py` (`quantize_4bit`,
`dequantize_4bit`, `quantize_blockwise`, `create_dynamic_map`) and `csrc/kernels.cu` (`kQuantizeBlockwise`,
`dQuantizeNF4`, `dQuantize<0>`, `kDequantizeBlockwise`).  The one published known-answer that IS pinned is the 16-entry
NF4 code book (QLoRA, Dettmers et al. 2023, appendix E), checked in tests/test_nf4_oracle.py together with the
structural properties the algorithm guarantees.

What the reference computes with such a layer (bitsandbytes `Linear4bit.forward` -> `matmul_4bit`): the input is cast
to the compute dtype (bf16), the weight is de-quantised to the dtype it was quantised from and cast to the input's
dtype, then a plain GEMM.  transformers overrides `torch_dtype=None` with float16 for bitsandbytes loads, so the chain
for a bf16 checkpoint is   w_bf16 -> fp16 -> NF4 blocks of 64 (+ 8-bit absmax, blocks of 256) -> fp16 -> bf16 -> GEMM.
"""
from __future__ import annotations

import numpy as np
import torch

# QLoRA appendix E / bitsandbytes `get_4bit_type("nf4")`
NF4_TABLE = np.array([-1.0, -0.6961928009986877, -0.5250730514526367, -0.39491748809814453, -0.28444138169288635,
                      -0.18477343022823334, -0.09105003625154495, 0.0, 0.07958029955625534, 0.16093020141124725,
                      0.24611230194568634, 0.33791524171829224, 0.44070982933044434, 0.5626170039176941,
                      0.7229568362236023, 1.0], dtype=np.float32)
# dQuantizeNF4's decision tree compares against the midpoints of neighbouring codes with `>` (a tie goes down)
NF4_MIDPOINTS = ((NF4_TABLE[:-1].astype(np.float64) + NF4_TABLE[1:].astype(np.float64)) / 2).astype(np.float32)
BLOCK, BLOCK2 = 64, 256


def dynamic_map() -> np.ndarray:
    """`create_dynamic_map(signed=True, max_exponent_bits=7, total_bits=8)`: for each of 7 decades 10^-6 .. 10^0,
    2^i bucket means of linspace(0.1, 1, 2^i + 1), both signs; plus 0 and 1.0; sorted.  256 float32 values,
    asymmetric (1.0 is a code, -1.0 is not)."""
    data = []
    for i in range(7):
        boundaries = torch.linspace(0.1, 1, 2 ** i + 1)
        means = (boundaries[:-1] + boundaries[1:]) / 2.0
        data += ((10 ** (-6 + i)) * means).tolist()
        data += (-(10 ** (-6 + i)) * means).tolist()
    data += [0, 1.0]
    assert len(data) == 256
    data.sort()
    return torch.tensor(data, dtype=torch.float32).numpy()


def _ftz(x):
    """bitsandbytes' CUDA sources are built with --use_fast_math, i.e. -ftz=true: single-precision operations flush
    subnormal inputs and results to zero (sign kept).  Applied after every fp32 operation restated below."""
    x = np.asarray(x, dtype=np.float32)
    return np.where(np.abs(x) < np.float32(2.0 ** -126), np.copysign(np.float32(0), x), x).astype(np.float32)


def _round_dtype(x: np.ndarray, dtype: str) -> np.ndarray:
    t = torch.from_numpy(np.ascontiguousarray(x, dtype=np.float32))
    return t.to({"fp16": torch.float16, "bf16": torch.bfloat16}[dtype]).float().numpy()


def quantize_nf4(w: np.ndarray):
    """kQuantizeBlockwise<NF4>, blocksize 64: w float32 [n] (n % 64 == 0) -> packed uint8 [n/2] (first value of a pair
    in the high nibble), absmax float32 [n/64].  The kernel multiplies by the reciprocal of absmax."""
    w = _ftz(w).reshape(-1, BLOCK)
    absmax = np.abs(w).max(axis=1)
    with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
        inv = _ftz(np.float32(1.0) / absmax)
        x = _ftz(w * inv[:, None])
    code = np.searchsorted(NF4_MIDPOINTS, x, side="left").astype(np.uint8)     # number of midpoints below x
    code[np.isnan(x)] = 0                                                       # all-zero block: every `>` is false
    code = code.reshape(-1, 2)
    return ((code[:, 0] << 4) | code[:, 1]).astype(np.uint8), absmax


def _dquantize8(code: np.ndarray, x: np.ndarray) -> np.ndarray:
    """kernels.cu `dQuantize<0>`: seven-step bisection over the sorted 256-entry map, then the nearer of the landing
    code and its bracketing neighbour."""
    n = x.shape[0]
    pivot = np.full(n, 127, dtype=np.int64)
    upper_pivot, lower_pivot = np.full(n, 255, dtype=np.int64), np.zeros(n, dtype=np.int64)
    lower, upper = np.full(n, -1.0, dtype=np.float32), np.full(n, 1.0, dtype=np.float32)
    val = code[pivot]
    step = 64
    while step > 0:
        up = x > val
        lower_pivot = np.where(up, pivot, lower_pivot)
        lower = np.where(up, val, lower)
        upper_pivot = np.where(up, upper_pivot, pivot)
        upper = np.where(up, upper, val)
        pivot = np.where(up, pivot + step, pivot - step)
        val = code[pivot]
        step >>= 1
    upper = np.where(upper_pivot == 255, code[255], upper)
    lower = np.where(lower_pivot == 0, code[0], lower)
    above = x > val
    mid_up = ((upper + val) * np.float32(0.5)).astype(np.float32)
    mid_lo = ((lower + val) * np.float32(0.5)).astype(np.float32)
    return np.where(above, np.where(x > mid_up, upper_pivot, pivot),
                    np.where(x < mid_lo, lower_pivot, pivot)).astype(np.uint8)


def double_quant(absmax: np.ndarray, offset=None):
    """`quantize_4bit(compress_statistics=True)`: offset = mean(absmax); (absmax - offset) through the blockwise 8-bit
    dynamic quantiser, blocksize 256.  -> qabsmax uint8 [nb], absmax2 float32 [ceil(nb/256)], offset float32."""
    code = dynamic_map()
    if offset is None:          # tests pass the device's fp32 mean: its summation order is not the CPU's
        offset = torch.from_numpy(absmax).mean().item()
    offset = np.float32(offset)
    a = _ftz(absmax - offset)
    nb = a.shape[0]
    q = np.empty(nb, dtype=np.uint8)
    absmax2 = np.empty((nb + BLOCK2 - 1) // BLOCK2, dtype=np.float32)
    for b in range(absmax2.shape[0]):
        blk = a[b * BLOCK2:(b + 1) * BLOCK2]
        absmax2[b] = np.abs(blk).max()
        with np.errstate(divide="ignore", invalid="ignore", over="ignore"):
            x = _ftz(blk * _ftz(np.float32(1.0) / absmax2[b]))
        q[b * BLOCK2:(b + 1) * BLOCK2] = _dquantize8(code, x)
    return q, absmax2, offset


def dequantize_nf4(packed, qabsmax, absmax2, offset, out_dtype: str = "fp16") -> np.ndarray:
    """`dequantize_4bit` with a nested state: absmax = code[q] * absmax2[block / 256] (fp32), + offset (fp32), then
    NF4[c] * absmax rounded to the dtype the weight was quantised from.  Returns those values as float32 [n]."""
    code = dynamic_map()
    nb = qabsmax.shape[0]
    a = _ftz(code[qabsmax] * np.repeat(absmax2, BLOCK2)[:nb])
    a = _ftz(a + np.float32(offset))
    c = np.stack([packed >> 4, packed & 0x0F], axis=1).reshape(-1, BLOCK)
    v = _ftz(NF4_TABLE[c] * a[:, None]).reshape(-1)
    return _round_dtype(v, out_dtype)


def roundtrip(w_bf16: np.ndarray, via_fp16: bool = True, offset=None) -> np.ndarray:
    """The values the reference's GEMM sees for a bf16 checkpoint weight loaded in 4 bit (see the module docstring):
    float32 array holding bf16-representable numbers."""
    w = np.asarray(w_bf16, dtype=np.float32).reshape(-1)
    src = _round_dtype(w, "fp16") if via_fp16 else w
    packed, absmax = quantize_nf4(src)
    q, absmax2, offset = double_quant(absmax, offset)
    deq = dequantize_nf4(packed, q, absmax2, offset, "fp16" if via_fp16 else "bf16")
    return _round_dtype(deq, "bf16").reshape(np.shape(w_bf16))
