"""Python-side launchers: torch tensors in, C-ABI calls out.

torch is used only for device memory and the current stream; all arithmetic
happens in libpt_stage1.so.  Every function launches on torch's current CUDA
stream and never synchronises.
"""
from __future__ import annotations

import ctypes as C

import torch

from . import _lib
from ._lib import BAND_K_BY_KEY, BAND_K_BY_QUERY, BAND_NONE, BAND_SCORES  # noqa: F401
from ._lib import ACT_QKNORM_ROPE, ACT_CE_STATS, ACT_ATTN_DELTA, ACT_GELU_ERF_BWD, ACT_SWIGLU, ACT_SWIGLU_BWD, ACT_GEGLU, ACT_GEGLU_BWD, ACT_SOFTMAX_BWD, ACT_GELU_ERF, ACT_GELU_TANH, ACT_NONE, PT_BF16, PT_F32, GemmArgs

_DT = {torch.bfloat16: PT_BF16, torch.float32: PT_F32}

# When set to a list, every GEMM launch is bracketed by CUDA events on the launching stream and
# (flops, start, end, shape) is appended (bench.py's roofline pass; never set on the timed path).
GEMM_PROFILE = None


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
             a_off=0, b_off=0, c_off=0, r_off=0, aux=None, ld_aux=0, band=0, band_window=0, bn_hint=0,
             aux2=None, aux_i0=0, aux_i1=0, A2=None, lda2=0, B2=None, ldb2=0, K2=0, ksplit=0, eps=0.0):
    """Fully general launcher; A/B/Cout are tensors used only as base pointers
    (+ element offsets), all geometry comes from the explicit arguments."""
    _req_cuda(A, B, Cout, bias, residual, c_pre, aux)
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
    g.bn_hint = int(bn_hint)
    g.band = int(band)
    g.band_window = int(band_window)
    if aux is not None:
        g.aux = aux.data_ptr()
        g.ld_aux = int(ld_aux)
    if aux2 is not None:
        g.aux2 = aux2.data_ptr()
    g.aux_i0, g.aux_i1 = int(aux_i0), int(aux_i1)
    if K2:
        _req_cuda(A2, B2)
        if A2.dtype != torch.bfloat16 or B2.dtype != torch.bfloat16:
            raise RuntimeError("gemm operands must be bf16")
        g.A2, g.lda2, g.B2, g.ldb2, g.K2 = A2.data_ptr(), int(lda2), B2.data_ptr(), int(ldb2), int(K2)
    g.ksplit = int(ksplit)
    g.eps = float(eps)
    if GEMM_PROFILE is not None:
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        _lib.call("pt_gemm_bf16", C.byref(g), _stream())
        e1.record()
        nb = int(nz[0]) * int(nz[1]) * int(nz[2])
        GEMM_PROFILE.append((2.0 * M * N * (K * nb * max(1, int(kg)) + int(K2)), e0, e1, (M, N, K, nb, int(kg))))
        return
    _lib.call("pt_gemm_bf16", C.byref(g), _stream())


def _seg2(M: int, N: int, a2, b2, b_mn: bool) -> dict:
    """Second contraction segment (pt_gemm_args.K2) of a linear / dgrad launch: a2 [M, K2] K-major and b2 either
    [N, K2] (forward: the LoRA B matrices) or [K2, N] read MN-major (input gradient: the LoRA A matrices)."""
    if a2 is None:
        return {}
    K2 = a2.shape[1]
    assert a2.shape[0] == M and a2.stride(1) == 1 and b2.stride(1) == 1
    assert tuple(b2.shape) == ((K2, N) if b_mn else (N, K2)), (tuple(a2.shape), tuple(b2.shape), N)
    return dict(A2=a2, lda2=a2.stride(0), B2=b2, ldb2=b2.stride(0), K2=K2)


def linear(x: torch.Tensor, w: torch.Tensor, *, bias=None, act=ACT_NONE, residual=None,
           out=None, out_dtype=torch.bfloat16, c_pre=None, alpha=1.0, x2=None, w2=None, ksplit=0,
           bn_hint=0) -> torch.Tensor:
    """y[M,N] = act(alpha * (x[M,K] @ w[N,K]^T (+ x2[M,K2] @ w2[N,K2]^T)) + bias) (+ residual).  bf16 row-major 2-D
    operands."""
    M, K = x.shape
    N, K2 = w.shape
    assert K == K2 and x.stride(1) == 1 and w.stride(1) == 1
    if out is None:
        out = torch.empty((M, N), dtype=out_dtype, device=x.device)
    gemm_raw(M=M, N=N, K=K, A=x, lda=x.stride(0), B=w, ldb=w.stride(0), Cout=out,
             ldc=out.stride(0), bias=bias, residual=residual,
             ldr=0 if residual is None else residual.stride(0), act=act, alpha=alpha, c_pre=c_pre, ksplit=ksplit,
             bn_hint=bn_hint, **_seg2(M, N, x2, w2, False))
    return out


def linear_dgrad(dy: torch.Tensor, w: torch.Tensor, *, out=None, out_dtype=torch.bfloat16,
                 residual=None, alpha=1.0, dy2=None, w2=None, ksplit=0) -> torch.Tensor:
    """dx[M,K] = alpha * (dy[M,N] @ w[N,K] (+ dy2[M,N2] @ w2[N2,K])) (+ residual) without transposing w (B operands
    MN-major)."""
    M, N = dy.shape
    N2, K = w.shape
    assert N == N2 and dy.stride(1) == 1 and w.stride(1) == 1
    if out is None:
        out = torch.empty((M, K), dtype=out_dtype, device=dy.device)
    gemm_raw(M=M, N=K, K=N, A=dy, lda=dy.stride(0), B=w, ldb=w.stride(0), b_mn=True, Cout=out,
             ldc=out.stride(0), residual=residual, alpha=alpha, ksplit=ksplit,
             ldr=0 if residual is None else residual.stride(0), **_seg2(M, K, dy2, w2, True))
    return out


def linear_dgrad_splitk(dy: torch.Tensor, w: torch.Tensor, *, out: torch.Tensor, partials: torch.Tensor,
                        splits: int) -> torch.Tensor:
    """linear_dgrad for a very long contraction and a small output (LM head: dy [R,V] @ w [V,h]): the K range is cut
    into `splits` slices run as a batch dim of ONE launch on 256-wide tiles (f32 partials [splits, M, K_out]), then
    summed into the bf16 result.  80 tiles x 8 slices fill 74 clusters to 96 %, where the single launch has to take
    128-wide tiles (144 tiles) and is bound by operand ingest per flop: 2395 -> 1875 us at R=4096, V=262144, h=1152."""
    M, N = dy.shape
    N2, K = w.shape
    assert N == N2 and dy.stride(1) == 1 and w.stride(1) == 1 and N % (splits * 64) == 0 and (M * K) % 4 == 0
    assert partials.dtype == torch.float32 and partials.numel() >= splits * M * K and out.is_contiguous()
    ks = N // splits
    gemm_raw(M=M, N=K, K=ks, A=dy, lda=dy.stride(0), B=w, ldb=w.stride(0), b_mn=True, Cout=partials, ldc=K,
             nz=(splits, 1, 1), a_zs=(ks, 0, 0), b_zs=(ks * w.stride(0), 0, 0), c_zs=(M * K, 0, 0), bn_hint=256)
    _lib.call("pt_sum_slices", _ptr(partials), C.c_longlong(M * K), int(splits), _ptr(out), _stream())
    return out


def geglu_fusable(ff: int) -> bool:
    """The fused GeGLU GEMM epilogues tile ff in 128-column blocks (CTA-pair kernel only)."""
    return ff % 128 == 0


def linear_geglu(x: torch.Tensor, w_gu: torch.Tensor, gu: torch.Tensor, act: torch.Tensor, silu: bool = False,
                 x2=None, w2=None):
    """gu[M,2ff] = x @ w_gu^T (gate | up pre-activations, kept for the backward) and, in the same
    epilogue, act[M,ff] = gelu_tanh(gate) * up   (HF:models/gemma3/modeling_gemma3.py:127-129)."""
    M, K = x.shape
    N = w_gu.shape[0]
    gemm_raw(M=M, N=N, K=K, A=x, lda=x.stride(0), B=w_gu, ldb=w_gu.stride(0), Cout=gu, ldc=gu.stride(0),
             act=ACT_SWIGLU if silu else ACT_GEGLU, aux=act, ld_aux=act.stride(0), **_seg2(M, N, x2, w2, False))


def linear_qknorm_rope(x: torch.Tensor, w_qkv: torch.Tensor, qkv: torch.Tensor, q_out: torch.Tensor, k_out: torch.Tensor,
                       gains: torch.Tensor, table: torch.Tensor, hq: int, eps: float, x2=None, w2=None):
    """qkv[M, (hq+2hkv)*256] = x @ w_qkv^T (kept for the backward) and, in the same epilogue, the per-head q/k RMSNorm +
    rotary embedding of pt_qknorm_rope_fwd: q_out [M, hq*256], k_out [M, hkv*256].  gains f32 [512] = [q | k] norm
    gains; table f32 [S, 256] = [cos | sin] per position (position of row r = r % S).  head_dim 256 only."""
    M, K = x.shape
    N = w_qkv.shape[0]
    hkv = (N // 256 - hq) // 2
    if gains.dtype != torch.float32 or gains.numel() != 512 or table.dtype != torch.float32 or table.shape[1] != 256 \
            or not table.is_contiguous() or tuple(q_out.shape) != (M, hq * 256) or tuple(k_out.shape) != (M, hkv * 256) \
            or not k_out.is_contiguous() or q_out.dtype != torch.bfloat16 or k_out.dtype != torch.bfloat16:
        raise RuntimeError("linear_qknorm_rope: gains f32 [512], table f32 [S, 256], q_out / k_out bf16 [M, heads * 256]")
    gemm_raw(M=M, N=N, K=K, A=x, lda=x.stride(0), B=w_qkv, ldb=w_qkv.stride(0), Cout=qkv, ldc=qkv.stride(0),
             act=ACT_QKNORM_ROPE, aux=q_out, ld_aux=q_out.stride(0), aux2=k_out, aux_i0=table.shape[0], aux_i1=hq,
             bias=gains, residual=table, ldr=256, eps=eps, **_seg2(M, N, x2, w2, False))


def linear_dgrad_geglu_bwd(dy: torch.Tensor, w_down: torch.Tensor, gu: torch.Tensor, d_gu: torch.Tensor,
                           silu: bool = False, dy2=None, w2=None):
    """d_act = dy[M,h] @ w_down[h,ff] formed in TMEM only; the epilogue turns it into
    d_gu[M,2ff] = [d_act * up * gelu_tanh'(gate) | d_act * gelu_tanh(gate)] reading the saved gu."""
    M, N = dy.shape
    ff = w_down.shape[1]
    gemm_raw(M=M, N=ff, K=N, A=dy, lda=dy.stride(0), B=w_down, ldb=w_down.stride(0), b_mn=True, Cout=d_gu,
             ldc=d_gu.stride(0), act=ACT_SWIGLU_BWD if silu else ACT_GEGLU_BWD, aux=gu, ld_aux=gu.stride(0),
             **_seg2(M, ff, dy2, w2, True))


def linear_dgrad_gelu_erf_bwd(dy: torch.Tensor, w: torch.Tensor, Z: torch.Tensor, out: torch.Tensor) -> torch.Tensor:
    """out[M,K] = (dy[M,N] @ w[N,K]) * gelu_erf'(Z[M,K]): the projector's second-layer input gradient with the backward
    of nn.GELU() (ref:Stage1/projectors.py:18) in the GEMM epilogue — dH never reaches HBM.  `out` must not alias Z (edge tiles overlap their neighbours)."""
    M, N = dy.shape
    N2, K = w.shape
    assert N == N2 and Z.shape == (M, K) and out.shape == (M, K) and Z.stride(1) == 1 and out.stride(1) == 1
    assert out.data_ptr() != Z.data_ptr()
    gemm_raw(M=M, N=K, K=N, A=dy, lda=dy.stride(0), B=w, ldb=w.stride(0), b_mn=True, Cout=out, ldc=out.stride(0),
             act=ACT_GELU_ERF_BWD, aux=Z, ld_aux=Z.stride(0))
    return out


def linear_dgrad_attn_delta(dy: torch.Tensor, w: torch.Tensor, O: torch.Tensor, out: torch.Tensor,
                            delta: torch.Tensor, S: int, hd: int, dy2=None, w2=None) -> torch.Tensor:
    """out[M, hq*hd] = dy[M,N] @ w[N, hq*hd] (the o-projection input gradient dO) and, in the same epilogue,
    delta[(b*hq + head)*S + s] = sum_d dO * O over each head's hd columns (f32 [B*hq*S]; replaces attn_delta)."""
    M, N = dy.shape
    N2, K = w.shape
    assert N == N2 and O.shape == (M, K) and out.shape == (M, K) and delta.dtype == torch.float32
    assert delta.numel() >= M * (K // hd)
    _req_cuda(delta)
    gemm_raw(M=M, N=K, K=N, A=dy, lda=dy.stride(0), B=w, ldb=w.stride(0), b_mn=True, Cout=out, ldc=out.stride(0),
             act=ACT_ATTN_DELTA, aux=O, ld_aux=O.stride(0), aux2=delta, aux_i0=S, aux_i1=hd,
             **_seg2(M, K, dy2, w2, True))
    return out


def attn_delta_fusable(hq: int, hd: int) -> bool:
    """The delta epilogue works on 256-wide output tiles whose 128-column halves must lie inside one head."""
    return hd % 128 == 0 and (hq * hd) % 256 == 0


def linear_wgrad(dy: torch.Tensor, x: torch.Tensor, *, out=None, alpha=1.0, accumulate: bool = False) -> torch.Tensor:
    """dw[N,K] = alpha * dy[M,N]^T @ x[M,K] (f32 out), both operands MN-major, no transposes.  dy / x / out may be
    column slices of wider matrices (LoRA: one adapter's columns of a packed projection).  accumulate=True:
    dw += ... as a split-K launch (pt_gemm_args.ksplit = -1) — for small outputs contracted over many rows, where one
    tile per output block would leave most SMs idle; the caller clears `out` first."""
    M, N = dy.shape
    M2, K = x.shape
    assert M == M2 and dy.stride(1) == 1 and x.stride(1) == 1
    if out is None:
        out = torch.empty((N, K), dtype=torch.float32, device=dy.device)
    assert out.stride(1) == 1 and tuple(out.shape) == (N, K)
    gemm_raw(M=N, N=K, K=M, A=dy, lda=dy.stride(0), a_mn=True, B=x, ldb=x.stride(0), b_mn=True,
             Cout=out, ldc=out.stride(0), alpha=alpha, ksplit=-1 if accumulate else 0)
    return out


# --------------------------------------------------------------------------- bandwidth kernels
def _i64(v):
    return C.c_int64(int(v))


def _f(v):
    return C.c_float(float(v))


def patchify(pixels: torch.Tensor, patch: int, out: torch.Tensor):
    B, _, H, W = pixels.shape
    _req_cuda(pixels, out)
    _lib.call("pt_patchify", _ptr(pixels), _ptr(out), B, H, W, int(patch), _stream())


def layernorm_fwd(x, gamma, beta, out, eps, *, group=(0, 0, 0)):
    _req_cuda(x, out)
    _lib.call("pt_layernorm_fwd", _ptr(x), _ptr(gamma), _ptr(beta), _ptr(out), out.shape[0],
              x.shape[-1], _f(eps), *map(int, group), _stream())


def rmsnorm_fwd(x, w, out, eps, *, group=(0, 0, 0)):
    _req_cuda(x, out)
    _lib.call("pt_rmsnorm_fwd", _ptr(x), _DT[x.dtype], _ptr(w), _ptr(out), out.shape[0],
              x.shape[-1], _f(eps), *map(int, group), _stream())


def rmsnorm_residual_fwd(y, w_post, x_in, x_out, w_next, nxt, eps):
    _req_cuda(y, x_in, x_out)
    _lib.call("pt_rmsnorm_residual_fwd", _ptr(y), _ptr(w_post), _ptr(x_in), _ptr(x_out),
              _ptr(w_next), _ptr(nxt), y.shape[0], y.shape[1], _f(eps), _stream())


def rmsnorm_residual_fwd_acc(y_acc, w_post, x_in, x_out, w_next, nxt, eps):
    """rmsnorm_residual_fwd behind a split-K GEMM: y_acc f32 [rows, D] (the slices' atomic sums) is rounded to bf16,
    used, and cleared for the next GEMM in the same pass."""
    _req_cuda(y_acc, x_in, x_out)
    if y_acc.dtype != torch.float32 or not y_acc.is_contiguous():
        raise RuntimeError("rmsnorm_residual_fwd_acc: y_acc must be contiguous float32")
    _lib.call("pt_rmsnorm_residual_fwd_acc", _ptr(y_acc), _ptr(w_post), _ptr(x_in), _ptr(x_out),
              _ptr(w_next), _ptr(nxt), y_acc.shape[0], y_acc.shape[1], _f(eps), _stream())


def rmsnorm_bwd(g, u, w, dst, eps, *, add=None, rows=None, u_group=(0, 0, 0), dst_group=(0, 0, 0)):
    _req_cuda(g, u, dst)
    rows = g.shape[0] if rows is None else rows
    _lib.call("pt_rmsnorm_bwd", _ptr(g), _DT[g.dtype], _ptr(u), _DT[u.dtype], _ptr(w), _ptr(dst),
              _DT[dst.dtype], _ptr(add), int(rows), g.shape[-1], _f(eps), *map(int, u_group),
              *map(int, dst_group), _stream())


def rmsnorm_bwd_pair(g1, u1, w1, dx, u2, w2, out2, eps):
    _req_cuda(g1, u1, dx, u2, out2)
    _lib.call("pt_rmsnorm_bwd_pair", _ptr(g1), _ptr(u1), _ptr(w1), _ptr(dx), _ptr(u2), _ptr(w2), _ptr(out2),
              g1.shape[0], g1.shape[-1], _f(eps), _stream())


def qknorm_rope_fwd(qkv, wq, wk, cos_t, sin_t, q_out, k_out, S, hq, hkv, hd, eps):
    _req_cuda(qkv, q_out, k_out)
    _lib.call("pt_qknorm_rope_fwd", _ptr(qkv), _ptr(wq), _ptr(wk), _ptr(cos_t), _ptr(sin_t),
              _ptr(q_out), _ptr(k_out), qkv.shape[0], int(S), hq, hkv, hd, _f(eps), _stream())


def qknorm_rope_bwd(dq, dk, qkv, wq, wk, cos_t, sin_t, d_qkv, S, hq, hkv, hd, eps):
    _req_cuda(dq, dk, qkv, d_qkv)
    _lib.call("pt_qknorm_rope_bwd", _ptr(dq), _ptr(dk), _ptr(qkv), _ptr(wq), _ptr(wk), _ptr(cos_t),
              _ptr(sin_t), _ptr(d_qkv), qkv.shape[0], int(S), hq, hkv, hd, _f(eps), _stream())


def attn_siglip_fwd(qkv, out, B, N, H, hd, scale):
    """Fused non-causal attention over the fused qkv activation [B*N, 3*H*hd] -> out [B*N, H*hd]."""
    _req_cuda(qkv, out)
    _lib.call("pt_attn_siglip_fwd", _ptr(qkv), _ptr(out), int(B), int(N), int(H), int(hd), _f(scale), _stream())


def softmax_fwd(scores, P, keyvalid, Z, Sq, Sk, ld, heads_per_batch, causal, window):
    _req_cuda(scores, P, keyvalid)
    _lib.call("pt_softmax_fwd", _ptr(scores), _ptr(P), _ptr(keyvalid), int(Z), int(Sq), int(Sk),
              int(ld), int(heads_per_batch), int(causal), int(window), _stream())


GEMV_BF16, GEMV_F32, GEMV_F32_ATOMIC, GEMV_GEGLU, GEMV_SWIGLU = range(5)


GEMV_MAX_ROWS = 32


def gemv_ok(M: int, K: int, max_rows: int = GEMV_MAX_ROWS) -> bool:
    return 1 <= M <= min(max_rows, GEMV_MAX_ROWS) and K >= 32 and K % 32 == 0


def gemv(x: torch.Tensor, w: torch.Tensor, out: torch.Tensor, mode: int = GEMV_BF16, ksplit: int = 0):
    """out[M, N] = x[M, K] @ w[N, K]^T for M <= 16 (weight streaming; pt_gemv_small_m).  GLU modes: w holds
    [gate rows ; up rows] (2 N rows) and out = act(gate) * up."""
    _req_cuda(x, w, out)
    M, K = x.shape
    N = w.shape[0] // 2 if mode in (GEMV_GEGLU, GEMV_SWIGLU) else w.shape[0]
    want = torch.float32 if mode in (GEMV_F32, GEMV_F32_ATOMIC) else torch.bfloat16
    if x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16 or out.dtype != want or w.shape[1] != K or \
            x.stride(1) != 1 or w.stride(1) != 1 or out.stride(1) != 1 or tuple(out.shape) != (M, N):
        raise RuntimeError("gemv: bf16 row-major x [M, K], w [N, K]; out [M, N] bf16 (modes 0, 3, 4) or f32 (1, 2)")
    _lib.call("pt_gemv_small_m", _ptr(x), C.c_longlong(x.stride(0)), int(M), _ptr(w), C.c_longlong(w.stride(0)), int(N),
              int(K), _ptr(out), C.c_longlong(out.stride(0)), int(mode), int(ksplit), _stream())
    return out


def sample_token(logits, *, out, step, next_tokens=None, prev=None, repetition_penalty=1.0, temperature=1.0, top_k=0,
                 top_p=1.0, u=None, done=None, pad_id=None, eos_id=None):
    """The next token of every row (HF's processor order: repetition penalty, temperature, top-k, top-p, one draw at
    u[b]; top_k = 0: argmax) and HF's finished-row bookkeeping, written to out[:, step] (i64 [B, n]) and next_tokens.
    logits f32 [B, V] (row stride >= V); prev = out: the tokens out[:, :step] are the penalised ones — their logits
    are overwritten with the penalised values."""
    _req_cuda(logits, out, next_tokens, prev, u, done)
    B, V = logits.shape
    if logits.dtype != torch.float32 or logits.stride(1) != 1 or out.dtype != torch.int64 or out.stride(1) != 1:
        raise RuntimeError("sample_token: logits must be f32 rows, out int64 rows")
    if (done is not None and done.dtype != torch.uint8) or (u is not None and (u.dtype != torch.float32 or u.numel() < B)):
        raise RuntimeError("sample_token: done must be uint8 [B], u float32 [B]")
    if prev is not None and (prev.dtype != torch.int64 or prev.stride(1) != 1):
        raise RuntimeError("sample_token: prev must be int64 rows")
    _lib.call("pt_sample_token", _ptr(logits), C.c_longlong(logits.stride(0)), int(B), int(V), _ptr(prev),
              C.c_longlong(0 if prev is None else prev.stride(0)), int(step if prev is not None else 0),
              _f(repetition_penalty), _f(temperature), int(top_k), _f(top_p), _ptr(u), _ptr(done),
              C.c_longlong(-1 if pad_id is None else int(pad_id)), C.c_longlong(-1 if eos_id is None else int(eos_id)),
              C.c_void_p(out.data_ptr() + 8 * int(step)), C.c_longlong(out.stride(0)), _ptr(next_tokens), _stream())


def decode_attn_supported(hq: int, hkv: int, hd: int) -> bool:
    return bool(_lib.lib().pt_decode_attn_supported(int(hq), int(hkv), int(hd)))


def decode_attn(qkv, wq, wk, cos_t, sin_t, pos, slot, valid, K, V, out, hq, hkv, hd, window, eps, scale, nsplit=0):
    """One cached decode step of attention in one launch: q/k norm + RoPE of this token (table row pos[b]), the
    cache append at *slot, attention over the visible cache slots.  qkv bf16 [B, (hq+2hkv)*hd]; K, V bf16
    [B, Smax, hkv*hd]; pos i64 [B]; slot i64 [1]; valid u8 [B, Smax] or None; out bf16 [B, hq*hd]."""
    _req_cuda(qkv, K, V, out, pos, slot, valid, cos_t, sin_t)
    B, Smax = K.shape[0], K.shape[1]
    if pos.dtype != torch.int64 or slot.dtype != torch.int64 or (valid is not None and valid.dtype != torch.uint8):
        raise RuntimeError("decode_attn: pos / slot must be int64, valid uint8")
    _lib.call("pt_decode_attn", _ptr(qkv), _ptr(wq), _ptr(wk), _ptr(cos_t), _ptr(sin_t), int(cos_t.shape[0]),
              _ptr(pos), _ptr(slot), _ptr(valid), _ptr(K), _ptr(V), _ptr(out), int(B), int(Smax), int(hq), int(hkv),
              int(hd), int(window), _f(eps), _f(scale), int(nsplit), _stream())


def softmax_bwd(P, dP, dS, rows, Sk, ld, scale, Sq=0, causal=0, window=0):
    _req_cuda(P, dP, dS)
    _lib.call("pt_softmax_bwd", _ptr(P), _ptr(dP), _ptr(dS), C.c_longlong(int(rows)), int(Sk),
              int(ld), _f(scale), int(Sq), int(causal), int(window), _stream())


def attn_gemma_fusable(hq: int, hkv: int, hd: int) -> bool:
    G = hq // max(1, hkv)
    return hd in (128, 256) and hq % hkv == 0 and G <= 128 and (G & (G - 1)) == 0


def pack_keyvalid(keyvalid, kvbits):
    """u8 [B,S] key-valid flags -> u64 [B, ceil(S/64)] bit masks for attn_gemma_fwd."""
    _req_cuda(keyvalid, kvbits)
    _lib.call("pt_pack_keyvalid", _ptr(keyvalid), _ptr(kvbits), keyvalid.shape[0], keyvalid.shape[1], _stream())


def attn_gemma_fwd(q_rot, k_rot, qkv, kvbits, lse, O, B, S, hq, hkv, hd, window, scale):
    """Fused causal / sliding-window GQA attention forward (hd 256 / 128), one pass with an online softmax:
    writes O and (unless None) the log2-domain log-sum-exp of every score row, f32 [B*hq, S].
    kvbits: packed key-padding masks (pack_keyvalid) or None."""
    _req_cuda(q_rot, k_rot, qkv, kvbits, lse, O)
    _lib.call("pt_attn_gemma_fwd", _ptr(q_rot), _ptr(k_rot), _ptr(qkv), _ptr(kvbits), _ptr(lse),
              _ptr(O), int(B), int(S), int(hq), int(hkv), int(hd), int(window), _f(scale), _stream())


def attn_gemma_bwd(q_rot, k_rot, qkv, dO, kvbits, lse, delta, dS, dK, dV, B, S, hq, hkv, hd, window, scale, lds):
    """Fused backward of attn_gemma_fwd: dK (gradient w.r.t. k_rot), dV and dS (bf16 [B*hq, S, lds], band only:
    zero-fill once).  dK / dV may be column slices of wider row-major buffers (their row stride is passed on)."""
    _req_cuda(q_rot, k_rot, qkv, dO, kvbits, lse, delta, dS, dK, dV)
    assert dK.stride(1) == 1 and dV.stride(1) == 1
    _lib.call("pt_attn_gemma_bwd", _ptr(q_rot), _ptr(k_rot), _ptr(qkv), _ptr(dO), _ptr(kvbits), _ptr(lse),
              _ptr(delta), _ptr(dS), _ptr(dK), C.c_longlong(dK.stride(0)), _ptr(dV), C.c_longlong(dV.stride(0)),
              int(B), int(S), int(hq), int(hkv), int(hd), int(window), _f(scale), int(lds), _stream())


def attn_gemma_bwd_pair_ok(hq: int, hkv: int, hd: int) -> bool:
    G = hq // max(1, hkv)
    return hd in (128, 256) and hq % hkv == 0 and G >= 1


def attn_gemma_bwd_pair(q_rot, k_rot, qkv, dO, kvbits, lse, delta, dST, dK, dV, B, S, hq, hkv, hd, window, scale, ldt):
    """CTA-pair backward of attn_gemma_fwd: dK, dV and dS^T (bf16 [B*hq, S keys, ldt], band only: zero-fill once)."""
    _req_cuda(q_rot, k_rot, qkv, dO, kvbits, lse, delta, dST, dK, dV)
    assert dK.stride(1) == 1 and dV.stride(1) == 1
    _lib.call("pt_attn_gemma_bwd_pair", _ptr(q_rot), _ptr(k_rot), _ptr(qkv), _ptr(dO), _ptr(kvbits), _ptr(lse),
              _ptr(delta), _ptr(dST), _ptr(dK), C.c_longlong(dK.stride(0)), _ptr(dV), C.c_longlong(dV.stride(0)),
              int(B), int(S), int(hq), int(hkv), int(hd), int(window), _f(scale), int(ldt), _stream())


def attn_dq_from_dst(dST, k_rot, dQ, B, S, hq, hkv, hd, window, ldt):
    """dQ[b, q, head, :] = sum_k dS^T[b, head, k, q] K[b, k, kvh, :]: band-limited GEMM, dS^T read MN-major."""
    G = hq // hkv
    gemm_raw(M=S, N=hd, K=S, A=dST, lda=ldt, a_mn=True, B=k_rot, ldb=hkv * hd, b_mn=True, Cout=dQ, ldc=hq * hd,
             nz=(G, hkv, B), a_zs=(S * ldt, G * S * ldt, hq * S * ldt), b_zs=(0, hd, S * hkv * hd),
             c_zs=(hd, G * hd, S * hq * hd), band=BAND_K_BY_QUERY, band_window=window)


def attn_delta(dO, O, delta, S, hq, hd):
    _req_cuda(dO, O, delta)
    _lib.call("pt_attn_delta", _ptr(dO), _ptr(O), _ptr(delta), dO.shape[0], int(S), int(hq), int(hd), _stream())


def geglu_fwd(gu, act, silu=False):
    _req_cuda(gu, act)
    _lib.call("pt_swiglu_fwd" if silu else "pt_geglu_fwd", _ptr(gu), _ptr(act), C.c_longlong(gu.shape[0]), act.shape[1], _stream())


def geglu_bwd(dact, gu, dgu, silu=False):
    _req_cuda(dact, gu, dgu)
    _lib.call("pt_swiglu_bwd" if silu else "pt_geglu_bwd", _ptr(dact), _ptr(gu), _ptr(dgu), C.c_longlong(gu.shape[0]),
              dact.shape[1], _stream())


def gelu_erf_bwd(dH, Z, dZ):
    _req_cuda(dH, Z, dZ)
    _lib.call("pt_gelu_erf_bwd", _ptr(dH), _ptr(Z), _ptr(dZ), C.c_longlong(dH.numel()), _stream())


ERRFLAG_TOKEN_ID, ERRFLAG_TARGET = 1, 2


def embed_splice(token_ids, labels, embed, scale, pad_id, n_img, x0, attn_mask, lm_labels,
                 keyvalid, targets, err=None):
    """`err` (int32 [1], optional): ERRFLAG_TOKEN_ID is OR-ed in when a token id lies outside the table."""
    _req_cuda(token_ids, labels, embed, x0, attn_mask, lm_labels, keyvalid, targets, err)
    B, T = token_ids.shape
    has_pad = pad_id is not None
    _lib.call("pt_embed_splice", _ptr(token_ids), _ptr(labels), _ptr(embed), _f(scale),
              _i64(pad_id if has_pad else 0), int(has_pad), B, T, int(n_img), embed.shape[1],
              _ptr(x0), _ptr(attn_mask), _ptr(lm_labels), _ptr(keyvalid), _ptr(targets),
              _i64(embed.shape[0]), _ptr(err), _stream())


def ce_fwd_bwd(logits, targets, lse, acc2, loss, gscale, err=None):
    """`err` (int32 [1], optional): ERRFLAG_TARGET is OR-ed in when a target is >= V."""
    _req_cuda(logits, targets, lse, acc2, loss, err)
    rows, V = logits.shape
    _lib.call("pt_ce_fwd_bwd", _ptr(logits), C.c_longlong(logits.stride(0)), _ptr(targets), rows, V,
              _ptr(lse), _ptr(acc2), _ptr(loss), _f(gscale), _ptr(err), _stream())


def ce_stats_parts(V: int) -> int:
    """Partials per row that the CE-statistics epilogue writes: one per 128-column half of every 256-wide tile."""
    return 2 * ((V + 255) // 256)


def linear_ce_stats(x: torch.Tensor, w: torch.Tensor, out: torch.Tensor, stats: torch.Tensor) -> torch.Tensor:
    """logits[M, V] = x[M, K] @ w[V, K]^T (bf16, the LM head) with, in the same epilogue, the per-row (max, sum exp)
    partials of the rounded logits into stats (f32 [ce_stats_parts(V), M, 2]) — ce_from_partials finishes the loss."""
    M, K = x.shape
    V = w.shape[0]
    assert stats.dtype == torch.float32 and stats.numel() >= ce_stats_parts(V) * M * 2 and V >= 256
    _req_cuda(stats)
    gemm_raw(M=M, N=V, K=K, A=x, lda=x.stride(0), B=w, ldb=w.stride(0), Cout=out, ldc=out.stride(0),
             act=ACT_CE_STATS, aux2=stats)
    return out


def ce_from_partials(logits, targets, stats, lse, acc2, loss, gscale, err=None):
    """ce_fwd_bwd without its pass over the logits for the statistics (they come from linear_ce_stats)."""
    _req_cuda(logits, targets, stats, lse, acc2, loss, err)
    rows, V = logits.shape
    _lib.call("pt_ce_from_partials", _ptr(logits), C.c_longlong(logits.stride(0)), _ptr(targets), int(rows), int(V),
              _ptr(stats), int(ce_stats_parts(V)), _ptr(lse), _ptr(acc2), _ptr(loss), _f(gscale), _ptr(err), _stream())


def colsum(x, out):
    _req_cuda(x, out)
    _lib.call("pt_colsum", _ptr(x), C.c_longlong(x.stride(0)), _ptr(out), x.shape[0], x.shape[1],
              _stream())


def sumsq(g, out):
    _req_cuda(g, out)
    _lib.call("pt_sumsq", _ptr(g), C.c_longlong(g.numel()), _ptr(out), _stream())


def clip_adamw(p, g, m, v, shadow, sumsq_t, *, grad_scale, max_norm, lr, beta1, beta2, eps, wd, step):
    _req_cuda(p, g, m, v, shadow, sumsq_t)
    _lib.call("pt_clip_adamw", _ptr(p), _ptr(g), _ptr(m), _ptr(v), _ptr(shadow),
              C.c_longlong(p.numel()), _ptr(sumsq_t), _f(grad_scale), _f(max_norm), _f(lr),
              _f(beta1), _f(beta2), _f(eps), _f(wd), int(step), _stream())


def cast_f32_bf16(x, y):
    _req_cuda(x, y)
    _lib.call("pt_cast_f32_bf16", _ptr(x), _ptr(y), C.c_longlong(x.numel()), _stream())


def gather_cast_rows(x, out, *, group=(0, 0, 0)):
    _req_cuda(x, out)
    _lib.call("pt_gather_cast_rows", _ptr(x), _ptr(out), out.shape[0], out.shape[1],
              *map(int, group), _stream())


def image_preprocess_workspace(H: int, W: int, out_h: int, out_w: int) -> int:
    n = int(_lib.lib().pt_image_preprocess_workspace(int(H), int(W), int(out_h), int(out_w)))
    if n < 0:
        raise RuntimeError(f"pt_image_preprocess_workspace: bad sizes {(H, W, out_h, out_w)}")
    return n


def image_preprocess(src_u8_hwc, dst_chw, workspace, *, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5),
                     rescale: float = 1 / 255, dst_u8=None):
    """Decoded RGB uint8 [H, W, 3] (CUDA) -> pixel_values f32 [3, S, S]: Pillow's antialiased bicubic
    `Image.resize((S, S))` + SiglipImageProcessor rescale / normalize (ref:Stage1/train_projection_stage1.py:99-104,
    ref:Stage2/dataset.py:90-95).  `dst_u8` (optional, uint8 [S, S, 3]) receives the resized image itself."""
    _req_cuda(src_u8_hwc, dst_chw, workspace, dst_u8)
    H, W, ch = src_u8_hwc.shape
    if ch != 3 or src_u8_hwc.dtype != torch.uint8 or src_u8_hwc.stride(2) != 1 or src_u8_hwc.stride(1) != 3:
        raise RuntimeError("image_preprocess: source must be uint8 [H, W, 3] with packed pixels")
    _, oh, ow = dst_chw.shape
    _lib.call("pt_image_preprocess", _ptr(src_u8_hwc), int(H), int(W), C.c_longlong(int(src_u8_hwc.stride(0))),
              _ptr(dst_chw), _ptr(dst_u8), int(oh), int(ow), _f(mean[0]), _f(mean[1]), _f(mean[2]),
              _f(std[0]), _f(std[1]), _f(std[2]), C.c_double(float(rescale)), _ptr(workspace),
              C.c_longlong(int(workspace.numel() * workspace.element_size())), _stream())


# ----------------------------------------------------------------------------- JPEG decode
def _jpeg_args(files):
    n = len(files)
    keep, addr, length = [], [], []
    for f in files:
        if isinstance(f, torch.Tensor):
            if f.dtype != torch.uint8 or f.is_cuda or not f.is_contiguous():
                raise RuntimeError("jpeg_pack: file tensors must be contiguous host uint8")
            keep.append(f); addr.append(f.data_ptr()); length.append(f.numel())
        else:
            b = bytes(f) if not isinstance(f, bytes) else f
            keep.append(b); addr.append(C.cast(C.c_char_p(b), C.c_void_p).value); length.append(len(b))
    return keep, (C.c_void_p * n)(*addr), (C.c_int64 * n)(*length)


def jpeg_plan(files) -> list:
    """Host-side header parse only: [(H, W)] per file; raises RuntimeError naming the reason for a file the GPU
    decoder does not take (not a JPEG, progressive, CMYK, unusual sub-sampling, ...)."""
    n = len(files)
    keep, ptrs, sizes = _jpeg_args(files)
    w, h = (C.c_int32 * n)(), (C.c_int32 * n)()
    offs = (C.c_int64 * (n + 1))()
    blob_b, ws_b = C.c_int64(0), C.c_int64(0)
    _lib.check(_lib.lib().pt_jpeg_plan(ptrs, sizes, n, w, h, offs, C.byref(blob_b), C.byref(ws_b)), "pt_jpeg_plan")
    return [(int(h[i]), int(w[i])) for i in range(n)]


def jpeg_pack(files, pin: bool = True) -> dict:
    """Host half of the GPU JPEG decode (runs without a GPU; meant for DataLoader workers): `files` = encoded baseline
    JPEG files as bytes / bytearray / uint8 tensors.  Returns {"blob": uint8 host tensor (pinned when possible),
    "sizes": [(H, W)], "rgb_offsets": [n + 1], "workspace_bytes"} for jpeg_decode."""
    n = len(files)
    keep, ptrs, sizes = _jpeg_args(files)
    w, h = (C.c_int32 * n)(), (C.c_int32 * n)()
    offs = (C.c_int64 * (n + 1))()
    blob_b, ws_b = C.c_int64(0), C.c_int64(0)
    L = _lib.lib()
    _lib.check(L.pt_jpeg_plan(ptrs, sizes, n, w, h, offs, C.byref(blob_b), C.byref(ws_b)), "pt_jpeg_plan")
    blob = torch.empty(blob_b.value, dtype=torch.uint8)
    if pin and torch.cuda.is_available():
        blob = blob.pin_memory()
    _lib.check(L.pt_jpeg_pack(ptrs, sizes, n, C.c_void_p(blob.data_ptr()), C.c_int64(blob_b.value)), "pt_jpeg_pack")
    return {"blob": blob, "sizes": [(int(h[i]), int(w[i])) for i in range(n)], "rgb_offsets": [int(o) for o in offs],
            "workspace_bytes": int(ws_b.value)}


def jpeg_decode(packed: dict, device, workspace: torch.Tensor | None = None) -> list:
    """Device half: uploads the blob and decodes the batch; returns one uint8 [H, W, 3] CUDA tensor per file (views of
    one buffer), bit-exact with `PIL.Image.open(f).convert("RGB")`."""
    blob = packed["blob"]
    dev_blob = blob.to(device, non_blocking=True)
    need = packed["workspace_bytes"]
    if workspace is None or workspace.numel() < need:
        workspace = torch.empty(need, dtype=torch.uint8, device=device)
    offs = packed["rgb_offsets"]
    dst = torch.empty(offs[-1], dtype=torch.uint8, device=device)
    _lib.call("pt_jpeg_decode_batch", C.c_void_p(blob.data_ptr()), _ptr(dev_blob), _ptr(dst), _ptr(workspace),
              C.c_int64(workspace.numel()), _stream())
    return [dst[offs[i]:offs[i] + 3 * hh * ww].view(hh, ww, 3) for i, (hh, ww) in enumerate(packed["sizes"])]


# ----------------------------------------------------------------------------- NF4 frozen weights
_NF4_DYNAMIC_MAP: dict = {}


def nf4_dynamic_map(device) -> torch.Tensor:
    """bitsandbytes `create_dynamic_map(signed=True, max_exponent_bits=7, total_bits=8)`: the sorted 256-entry code
    book of the nested (second-level) absmax quantiser — 2^i bucket means of linspace(0.1, 1, 2^i + 1) for each of
    the decades 10^-6 .. 10^0, both signs, plus 0 and 1.0."""
    key = str(device)
    if key not in _NF4_DYNAMIC_MAP:
        data: list = []
        for i in range(7):
            bounds = torch.linspace(0.1, 1, 2 ** i + 1)
            means = (bounds[:-1] + bounds[1:]) / 2.0
            data += ((10 ** (-6 + i)) * means).tolist()
            data += (-(10 ** (-6 + i)) * means).tolist()
        data += [0, 1.0]
        data.sort()
        _NF4_DYNAMIC_MAP[key] = torch.tensor(data, dtype=torch.float32, device=device)
    return _NF4_DYNAMIC_MAP[key]


def nf4_quantize(w: torch.Tensor, via_fp16: bool = True) -> dict:
    """bf16 weight (any shape, numel % 64 == 0) -> bitsandbytes-style nested NF4 state: `packed` uint8 [n/2],
    `qabsmax` uint8 [n/64], `absmax2` f32 [ceil(n/64/256)], `offset` (python float, the fp32 mean of the first-level
    absmax), `code256`, `shape`, `via_fp16` (ref:Stage1/train_projection_stage1.py:195-200)."""
    _req_cuda(w)
    if w.dtype != torch.bfloat16 or not w.is_contiguous():
        raise RuntimeError("nf4_quantize: contiguous bf16 weight expected")
    n = w.numel()
    if n == 0 or n % 64:
        raise RuntimeError(f"nf4_quantize: numel {n} is not a positive multiple of 64")
    nb = n // 64
    packed = torch.empty(n // 2, dtype=torch.uint8, device=w.device)
    absmax = torch.empty(nb, dtype=torch.float32, device=w.device)
    _lib.call("pt_nf4_quantize", _ptr(w), C.c_longlong(n), int(bool(via_fp16)), _ptr(packed), _ptr(absmax), _stream())
    offset = float(absmax.mean().item())
    code = nf4_dynamic_map(w.device)
    qabsmax = torch.empty(nb, dtype=torch.uint8, device=w.device)
    absmax2 = torch.empty((nb + 255) // 256, dtype=torch.float32, device=w.device)
    _lib.call("pt_nf4_absmax_quantize", _ptr(absmax), C.c_longlong(nb), _f(offset), _ptr(code), _ptr(qabsmax),
              _ptr(absmax2), _stream())
    return {"packed": packed, "qabsmax": qabsmax, "absmax2": absmax2, "offset": offset, "code256": code,
            "shape": tuple(w.shape), "via_fp16": bool(via_fp16), "absmax": absmax}


def nf4_dequantize(state: dict, out: torch.Tensor | None = None) -> torch.Tensor:
    """Nested NF4 state -> the bf16 weight bitsandbytes' Linear4bit feeds its GEMM."""
    packed = state["packed"]
    _req_cuda(packed, state["qabsmax"], state["absmax2"], state["code256"], out)
    n = packed.numel() * 2
    if out is None:
        out = torch.empty(state["shape"], dtype=torch.bfloat16, device=packed.device)
    if out.numel() != n or out.dtype != torch.bfloat16 or not out.is_contiguous():
        raise RuntimeError("nf4_dequantize: output must be contiguous bf16 of the quantised weight's size")
    _lib.call("pt_nf4_dequantize", _ptr(packed), _ptr(state["qabsmax"]), _ptr(state["absmax2"]), _f(state["offset"]),
              _ptr(state["code256"]), C.c_longlong(n), int(bool(state["via_fp16"])), _ptr(out), _stream())
    return out
