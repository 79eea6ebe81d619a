"""Python-side launchers: torch tensors in, C-ABI calls out.

torch is used only for device memory and the current stream; all arithmetic
happens in libpt_stage1.so.  Every function launches on torch's current CUDA
stream and never synchronises.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import ACT_GELU_ERF, ACT_GELU_TANH, ACT_NONE, PT_BF16, PT_F32, GemmArgs

_DT = {torch.bfloat16: PT_BF16, torch.float32: PT_F32}


def _stream() -> C.c_void_p:
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t: torch.Tensor | None) -> C.c_void_p:
    return C.c_void_p(0 if t is None else t.data_ptr())


def _req_cuda(*ts):
    for t in ts:
        if t is not None and not t.is_cuda:
            raise RuntimeError("projectiontrainer_b200 ops need CUDA tensors (no CPU path)")


def gemm_raw(*, M, N, K, A, lda, B, ldb, Cout, ldc, nz=(1, 1, 1), a_zs=(0, 0, 0), b_zs=(0, 0, 0),
             c_zs=(0, 0, 0), a_mn=False, b_mn=False, bias=None, residual=None, ldr=0,
             r_zs=(0, 0, 0), act=ACT_NONE, alpha=1.0, c_pre=None, kg=1, a_kgs=0, b_kgs=0,
             a_off=0, b_off=0, c_off=0, r_off=0):
    """Fully general launcher; A/B/Cout are tensors used only as base pointers
    (+ element offsets), all geometry comes from the explicit arguments."""
    _req_cuda(A, B, Cout, bias, residual, c_pre)
    if A.dtype != torch.bfloat16 or B.dtype != torch.bfloat16:
        raise RuntimeError("gemm operands must be bf16")
    g = GemmArgs()
    g.M, g.N, g.K = int(M), int(N), int(K)
    g.nz[:] = [int(v) for v in nz]
    g.A = A.data_ptr() + 2 * int(a_off)
    g.lda = int(lda)
    g.a_zs[:] = [int(v) for v in a_zs]
    g.a_mn_major = int(bool(a_mn))
    g.B = B.data_ptr() + 2 * int(b_off)
    g.ldb = int(ldb)
    g.b_zs[:] = [int(v) for v in b_zs]
    g.b_mn_major = int(bool(b_mn))
    g.C = Cout.data_ptr() + Cout.element_size() * int(c_off)
    g.c_dtype = _DT[Cout.dtype]
    g.ldc = int(ldc)
    g.c_zs[:] = [int(v) for v in c_zs]
    g.C_pre = None if c_pre is None else c_pre.data_ptr() + 2 * int(c_off)
    if bias is not None:
        if bias.dtype != torch.float32:
            raise RuntimeError("gemm bias must be f32")
        g.bias = bias.data_ptr()
    if residual is not None:
        g.residual = residual.data_ptr() + residual.element_size() * int(r_off)
        g.r_dtype = _DT[residual.dtype]
        g.ldr = int(ldr)
        g.r_zs[:] = [int(v) for v in r_zs]
    g.act = int(act)
    g.alpha = float(alpha)
    g.kg = int(kg)
    g.a_kgs = int(a_kgs)
    g.b_kgs = int(b_kgs)
    _lib.call("pt_gemm_bf16", C.byref(g), _stream())


def linear(x: torch.Tensor, w: torch.Tensor, *, bias=None, act=ACT_NONE, residual=None,
           out=None, out_dtype=torch.bfloat16, c_pre=None, alpha=1.0) -> torch.Tensor:
    """y[M,N] = act(x[M,K] @ w[N,K]^T + bias) (+ residual).  x, w bf16 row-major 2-D."""
    M, K = x.shape
    N, K2 = w.shape
    assert K == K2 and x.stride(1) == 1 and w.stride(1) == 1
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    gemm_raw(M=M, N=N, K=K, A=x, lda=x.stride(0), B=w, ldb=w.stride(0), Cout=out,
             ldc=out.stride(0), bias=bias, residual=residual,
             ldr=0 if residual is None else residual.stride(0), act=act, alpha=alpha, c_pre=c_pre)
    return out


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, *, out=None, out_dtype=torch.bfloat16,
                 residual=None) -> torch.Tensor:
    """dx[M,K] = dy[M,N] @ w[N,K] without transposing w (B operand MN-major)."""
    M, N = dy.shape
    N2, K = w.shape
    assert N == N2 and dy.stride(1) == 1 and w.stride(1) == 1
    if out is None:
        out = torch.empty((M, K), dtype=out_dtype, device=dy.device)
    gemm_raw(M=M, N=K, K=N, A=dy, lda=dy.stride(0), B=w, ldb=w.stride(0), b_mn=True, Cout=out,
             ldc=out.stride(0), residual=residual,
             ldr=0 if residual is None else residual.stride(0))
    return out


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, *, out=None) -> torch.Tensor:
    """dw[N,K] = dy[M,N]^T @ x[M,K] (f32 out), both operands MN-major, no transposes."""
    M, N = dy.shape
    M2, K = x.shape
    assert M == M2 and dy.stride(1) == 1 and x.stride(1) == 1
    if out is None:
        out = torch.empty((N, K), dtype=torch.float32, device=dy.device)
    gemm_raw(M=N, N=K, K=M, A=dy, lda=dy.stride(0), a_mn=True, B=x, ldb=x.stride(0), b_mn=True,
             Cout=out, ldc=out.stride(0))
    return out
