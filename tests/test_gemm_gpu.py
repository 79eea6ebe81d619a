"""tcgen05 GEMM vs torch fp32 matmul on bf16-rounded inputs (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _rand(shape, dev, scale=1.0):
    return (torch.randn(shape, device=dev, dtype=torch.float32) * scale).to(torch.bfloat16)


def _check(out, ref, tol=2e-2, name=""):
    out = out.float()
    err = (out - ref).abs().max().item()
    den = ref.abs().max().item() + 1e-6
    assert err / den < tol, f"{name}: max err {err} vs scale {den}"


@pytest.mark.parametrize("M,N,K", [(128, 256, 64), (256, 256, 256), (575, 1152, 1024), (600, 384, 320), (4000, 1152, 6912),
                                   (1000, 1536, 1152), (130, 64, 768), (384, 192, 2048),
                                   (4096, 1024, 4096), (18400, 1152, 1280), (22496, 1536, 1152),
                                   (100, 40, 72), (257, 328, 8)])
def test_gemm_tn(cuda_dev, M, N, K):
    from projectiontrainer_b200 import ops
    torch.manual_seed(M + N + K)
    x, w = _rand((M, K), cuda_dev), _rand((N, K), cuda_dev, 0.05)
    ref = x.float() @ w.float().t()
    _check(ops.linear(x, w), ref, name="bf16 out")
    _check(ops.linear(x, w, out_dtype=torch.float32), ref, tol=1e-3, name="f32 out")


def test_gemm_epilogues(cuda_dev):
    from projectiontrainer_b200 import ops
    torch.manual_seed(0)
    M, N, K = 700, 1152, 512
    x, w = _rand((M, K), cuda_dev), _rand((N, K), cuda_dev, 0.05)
    b = torch.randn(N, device=cuda_dev)
    res = torch.randn(M, N, device=cuda_dev)
    z = x.float() @ w.float().t() + b
    pre = torch.empty(M, N, device=cuda_dev, dtype=torch.bfloat16)
    y = ops.linear(x, w, bias=b, act=ops.ACT_GELU_ERF, c_pre=pre)
    _check(y, torch.nn.functional.gelu(z), name="gelu_erf")
    _check(pre, z, name="pre")
    y = ops.linear(x, w, bias=b, act=ops.ACT_GELU_TANH, out_dtype=torch.float32)
    _check(y, torch.nn.functional.gelu(z, approximate="tanh"), tol=2e-3, name="gelu_tanh")
    y = ops.linear(x, w, bias=b, residual=res, out_dtype=torch.float32)
    _check(y, z + res, tol=1e-3, name="residual f32")
    acc = res.clone()
    ops.linear(x, w, residual=acc, out=acc, alpha=0.5)
    _check(acc, 0.5 * (z - b) + res, tol=1e-3, name="in-place accumulate")


@pytest.mark.parametrize("M,N,K", [(256, 256, 256), (575, 1024, 1152), (1000, 10240, 1152),
                                   (703, 256, 703), (1000, 1152, 1536), (300, 384, 200), (22496, 1152, 1024)])
def test_gemm_dgrad_mn_major_b(cuda_dev, M, N, K):
    """dx[M,N] = dy[M,K] @ W[K,N] with W row-major (B operand MN-major)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(1)
    Kp = (K + 7) // 8 * 8
    dy = _rand((M, Kp), cuda_dev)[:, :K]
    w = _rand((K, N), cuda_dev, 0.05)
    ref = dy.float() @ w.float()
    out = torch.empty(M, N, device=cuda_dev, dtype=torch.float32)
    ops.gemm_raw(M=M, N=N, K=K, A=dy, lda=dy.stride(0), B=w, ldb=w.stride(0), b_mn=True,
                 Cout=out, ldc=N)
    _check(out, ref, tol=1e-3, name="dgrad")


@pytest.mark.parametrize("M,V,h,S", [(64, 512, 64, 8), (300, 4096, 328, 4), (1000, 16384, 1152, 8),
                                     (24, 2048, 256, 2)])
def test_gemm_dgrad_split_k(cuda_dev, M, V, h, S):
    """LM-head input gradient as a split-K launch (K slices as a batch dim, f32 partials) + pt_sum_slices:
    same result as the single launch, and both match the f32 matmul."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(V + h)
    Vp = V + 8                                  # padded row stride, as the engine's logits buffer has
    dy = _rand((M, Vp), cuda_dev, 0.1)[:, :V]
    w = _rand((V, h), cuda_dev, 0.05)
    ref = dy.float() @ w.float()
    one = ops.linear_dgrad(dy, w)
    part = torch.full((S, M, h), float("nan"), device=cuda_dev, dtype=torch.float32)
    out = torch.empty(M, h, device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_dgrad_splitk(dy, w, out=out, partials=part, splits=S)
    assert torch.isfinite(part).all()
    _check(part.sum(0), ref, tol=1e-3, name="split-K partials")
    _check(out, ref, name="split-K bf16")
    # one bf16 rounding of (nearly) the same f32 sum on both sides
    assert (out.float() - one.float()).abs().max().item() <= 2 ** -7 * ref.abs().max().item()


@pytest.mark.parametrize("T,N,K", [(256, 128, 256), (1000, 1152, 1024), (18400, 1152, 256),
                                   (703, 703, 256)])
def test_gemm_wgrad_mn_major_ab(cuda_dev, T, N, K):
    """dw[N,K] = dy[T,N]^T @ x[T,K] (both operands MN-major, contraction over rows)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(2)
    Np = (N + 7) // 8 * 8
    dy = _rand((T, Np), cuda_dev)[:, :N]
    x = _rand((T, K), cuda_dev)
    ref = dy.float().t() @ x.float()
    out = torch.empty(N, K, device=cuda_dev, dtype=torch.float32)
    ops.gemm_raw(M=N, N=K, K=T, A=dy, lda=dy.stride(0), a_mn=True, B=x, ldb=x.stride(0),
                 b_mn=True, Cout=out, ldc=K)
    _check(out, ref, tol=1e-3, name="wgrad")


def test_gemm_batched_gqa_scores(cuda_dev):
    """S[b,h,q,k] = scale * Q[b,q,h,:] . K[b,k,h//G,:]  (strided 3-D batch, broadcast kv)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(3)
    Bn, S, Hq, Hkv, D = 3, 703, 4, 2, 256
    G = Hq // Hkv
    q = _rand((Bn, S, Hq, D), cuda_dev)
    k = _rand((Bn, S, Hkv, D), cuda_dev)
    Sp = 704
    out = torch.zeros(Bn, Hq, S, Sp, device=cuda_dev, dtype=torch.float32)
    # z0 = head within group, z1 = kv head, z2 = batch
    ops.gemm_raw(M=S, N=S, K=D, A=q, lda=Hq * D, B=k, ldb=Hkv * D, Cout=out, ldc=Sp,
                 nz=(G, Hkv, Bn), a_zs=(D, G * D, S * Hq * D), b_zs=(0, D, S * Hkv * D),
                 c_zs=(S * Sp, G * S * Sp, Hq * S * Sp), alpha=1.0 / 16)
    kk = k.float().repeat_interleave(G, dim=2)
    ref = torch.einsum("bqhd,bkhd->bhqk", q.float(), kk) / 16
    _check(out[..., :S], ref, tol=1e-3, name="scores")
    assert out[..., S:].abs().max().item() == 0.0


def test_gemm_grouped_contraction(cuda_dev):
    """dV[b,k,kvh,:] = sum_{g,q} P[b,kvh*G+g,q,k] dO[b,q,kvh*G+g,:]  (kg groups, MN-major A and B)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(4)
    Bn, S, Hq, Hkv, D = 2, 703, 4, 2, 256
    G = Hq // Hkv
    Sp = 704
    P = torch.zeros(Bn, Hq, S, Sp, device=cuda_dev, dtype=torch.bfloat16)
    P[..., :S] = _rand((Bn, Hq, S, S), cuda_dev, 0.1)
    dO = _rand((Bn, S, Hq, D), cuda_dev)
    out = torch.empty(Bn, S, Hkv, D, device=cuda_dev, dtype=torch.bfloat16)
    ops.gemm_raw(M=S, N=D, K=S, A=P, lda=Sp, a_mn=True, B=dO, ldb=Hq * D, b_mn=True, Cout=out,
                 ldc=Hkv * D, nz=(Hkv, Bn, 1), a_zs=(G * S * Sp, Hq * S * Sp, 0),
                 b_zs=(G * D, S * Hq * D, 0), c_zs=(D, S * Hkv * D, 0),
                 kg=G, a_kgs=S * Sp, b_kgs=D)
    Pf = P[..., :S].float().view(Bn, Hkv, G, S, S)
    dOf = dO.float().view(Bn, S, Hkv, G, D)
    ref = torch.einsum("bhgqk,bqhgd->bkhd", Pf, dOf)
    _check(out, ref, name="grouped")


@pytest.mark.parametrize("M,ff,h", [(700, 384, 192), (22496, 6912, 1152), (64, 128, 64), (3, 256, 128)])
def test_gemm_fused_geglu_fwd_bwd(cuda_dev, M, ff, h):
    """GeGLU fused into the gate||up epilogue and its backward fused into the down-dgrad epilogue, against
    the unfused kernels (same bf16 rounding points) and a torch fp32 reference."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(M + ff)
    x, w_gu = _rand((M, h), cuda_dev), _rand((2 * ff, h), cuda_dev, 0.08)
    gu = torch.empty(M, 2 * ff, device=cuda_dev, dtype=torch.bfloat16)
    act = torch.empty(M, ff, device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_geglu(x, w_gu, gu, act)
    gu_ref = x.float() @ w_gu.float().t()
    _check(gu, gu_ref, name="gu")
    g, u = gu.float()[:, :ff], gu.float()[:, ff:]
    _check(act, torch.nn.functional.gelu(g, approximate="tanh") * u, tol=1e-2, name="act vs torch on the stored gu")
    gu2, act2 = ops.linear(x, w_gu), torch.empty_like(act)
    ops.geglu_fwd(gu2, act2)
    assert torch.equal(gu2, gu)
    _check(act, act2.float(), tol=1e-2, name="act vs unfused kernel")
    # backward
    dy, w_down = _rand((M, h), cuda_dev), _rand((h, ff), cuda_dev, 0.08)
    d_gu = torch.empty(M, 2 * ff, device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_dgrad_geglu_bwd(dy, w_down, gu, d_gu)
    d_act = ops.linear_dgrad(dy, w_down)
    d_gu2 = torch.empty_like(d_gu)
    ops.geglu_bwd(d_act, gu, d_gu2)
    _check(d_gu, d_gu2.float(), tol=1e-2, name="d_gu vs unfused kernel")
    gt = g.clone().requires_grad_(True)
    ut = u.clone().requires_grad_(True)
    (torch.nn.functional.gelu(gt, approximate="tanh") * ut).backward(d_act.float())
    _check(d_gu, torch.cat([gt.grad, ut.grad], 1), tol=1e-2, name="d_gu vs torch autograd")


@pytest.mark.parametrize("M,N,K", [(18400, 10240, 1152), (575, 640, 256), (70, 32, 64), (300, 2560, 320)])
def test_gemm_fused_gelu_erf_backward_epilogue(cuda_dev, M, N, K):
    """dZ = (dP W2) * gelu_erf'(Z) in the GEMM epilogue (TMA-moved Z / dZ tiles) vs the unfused pair and torch autograd."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(M + N)
    dy, w = _rand((M, K), cuda_dev), _rand((K, N), cuda_dev, 0.05)
    Z = _rand((M, N), cuda_dev, 1.5)
    out = torch.full((M, N), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_dgrad_gelu_erf_bwd(dy, w, Z, out)
    dH = dy.float() @ w.float()
    zr = Z.float().requires_grad_(True)
    torch.nn.functional.gelu(zr).backward(dH)
    _check(out, zr.grad, tol=1e-2, name="vs autograd")
    un = ops.linear_dgrad(dy, w)
    ops.gelu_erf_bwd(un, Z, un)
    _check(out, un.float(), tol=1e-2, name="vs unfused kernels")


@pytest.mark.parametrize("B,S,hq,hd,K", [(32, 703, 4, 256, 1152), (2, 100, 4, 256, 128), (3, 77, 8, 128, 256), (1, 300, 2, 128, 64)])
def test_gemm_fused_attention_delta_epilogue(cuda_dev, B, S, hq, hd, K):
    """dO = d_y W_o with delta[b, head, s] = sum_d dO * O in the epilogue vs linear_dgrad + pt_attn_delta."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(S + hq)
    M, N = B * S, hq * hd
    dy, w = _rand((M, K), cuda_dev), _rand((K, N), cuda_dev, 0.05)
    O = _rand((M, N), cuda_dev)
    dO = torch.full((M, N), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    delta = torch.full((B * hq * S,), float("nan"), device=cuda_dev)
    ops.linear_dgrad_attn_delta(dy, w, O, dO, delta, S, hd)
    ref = ops.linear_dgrad(dy, w)
    assert torch.equal(dO, ref)
    d_ref = torch.empty_like(delta)
    ops.attn_delta(ref, O, d_ref, S, hq, hd)
    d64 = (ref.double() * O.double()).view(B, S, hq, hd).sum(-1).permute(0, 2, 1).reshape(-1)
    assert torch.isfinite(delta).all()
    assert (delta.double() - d64).abs().max().item() < 1e-4 * d64.abs().max().item() + 1e-5
    assert (delta - d_ref).abs().max().item() < 1e-4 * d_ref.abs().max().item() + 1e-5


@pytest.mark.parametrize("S,W", [(703, 0), (703, 512), (300, 64), (1000, 100)])
def test_gemm_attention_bands_skip_only_masked_work(cuda_dev, S, W):
    """pt_gemm_args.band: causal / sliding-window tile and K-block skipping must not change any value a masked
    consumer can see (scores inside the band; P V, P^T dO with P zero outside the band)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(S + W)
    hd, G = 64, 2
    ld = (S + 7) // 8 * 8
    q, k = _rand((G, S, hd), cuda_dev), _rand((S, hd), cuda_dev)
    qi, ki = torch.arange(S, device=cuda_dev)[:, None], torch.arange(S, device=cuda_dev)[None, :]
    allowed = (ki <= qi) & ((ki > qi - W) if W else torch.ones_like(ki <= qi))
    # scores: batched over G, B shared
    sc = torch.full((G, S, ld), float("nan"), device=cuda_dev)
    ops.gemm_raw(M=S, N=S, K=hd, A=q, lda=hd, B=k, ldb=hd, Cout=sc, ldc=ld, nz=(G, 1, 1), a_zs=(S * hd, 0, 0),
                 c_zs=(S * ld, 0, 0), alpha=0.5, band=ops.BAND_SCORES, band_window=W)
    ref = 0.5 * torch.einsum("gqd,kd->gqk", q.float(), k.float())
    got = sc[:, :, :S]
    assert bool(torch.isfinite(got[:, allowed]).all())
    _check(got[:, allowed], ref[:, allowed], tol=1e-3, name="scores in band")
    # P (zero outside the band) times V, contraction over keys
    P = torch.zeros(G, S, ld, device=cuda_dev, dtype=torch.bfloat16)
    P[:, :, :S] = (torch.rand(G, S, S, device=cuda_dev) * allowed).to(torch.bfloat16)
    v = _rand((S, hd), cuda_dev)
    o = torch.empty(G, S, hd, device=cuda_dev, dtype=torch.bfloat16)
    ops.gemm_raw(M=S, N=hd, K=S, A=P, lda=ld, B=v, ldb=hd, b_mn=True, Cout=o, ldc=hd, nz=(G, 1, 1),
                 a_zs=(S * ld, 0, 0), c_zs=(S * hd, 0, 0), band=ops.BAND_K_BY_QUERY, band_window=W)
    _check(o, torch.einsum("gqk,kd->gqd", P[:, :, :S].float(), v.float()), name="P V")
    # P^T dO summed over the G heads (grouped contraction), contraction over queries
    do = _rand((G, S, hd), cuda_dev)
    dv = torch.empty(S, hd, device=cuda_dev, dtype=torch.bfloat16)
    ops.gemm_raw(M=S, N=hd, K=S, A=P, lda=ld, a_mn=True, B=do, ldb=hd, b_mn=True, Cout=dv, ldc=hd, kg=G,
                 a_kgs=S * ld, b_kgs=S * hd, band=ops.BAND_K_BY_KEY, band_window=W)
    _check(dv, torch.einsum("gqk,gqd->kd", P[:, :, :S].float(), do.float()), name="P^T dO")


@pytest.mark.parametrize("S,W", [(703, 0), (703, 512), (90, 16)])
def test_gemm_fused_softmax_backward_epilogue(cuda_dev, S, W):
    """dS = scale * P * (dO V^T - delta) in the GEMM epilogue (delta = rowsum(dO * O)) vs torch autograd
    through softmax(QK^T * scale) V on the same bf16-rounded tensors."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(S + W)
    Bn, hq, hd, scale = 2, 2, 64, 0.125
    ld = (S + 7) // 8 * 8
    q, k, v = (_rand((Bn, S, hq, hd), cuda_dev), _rand((Bn, S, 1, hd), cuda_dev), _rand((Bn, S, 1, hd), cuda_dev))
    qi, ki = torch.arange(S, device=cuda_dev)[:, None], torch.arange(S, device=cuda_dev)[None, :]
    allowed = (ki <= qi) & ((ki > qi - W) if W else (ki <= qi))
    qf = q.float().permute(0, 2, 1, 3)                               # [B, hq, S, hd]
    kf, vf = k.float()[:, :, 0], v.float()[:, :, 0]                  # [B, S, hd]
    sc = torch.einsum("bhqd,bkd->bhqk", qf, kf) * scale
    Pf = torch.softmax(sc.masked_fill(~allowed, float("-inf")), dim=-1)
    P = torch.zeros(Bn * hq, S, ld, device=cuda_dev, dtype=torch.bfloat16)
    P[:, :, :S] = Pf.reshape(Bn * hq, S, S).to(torch.bfloat16)
    Pb = P[:, :, :S].float().reshape(Bn, hq, S, S)
    O = torch.einsum("bhqk,bkd->bqhd", Pb, vf).to(torch.bfloat16)    # [B, S, hq, hd]
    dO = _rand((Bn, S, hq, hd), cuda_dev)
    delta = torch.empty(Bn * hq * S, device=cuda_dev)
    ops.attn_delta(dO.view(Bn * S, hq * hd), O.view(Bn * S, hq * hd), delta, S, hq, hd)
    ref_delta = (dO.float() * O.float()).sum(-1).permute(0, 2, 1).reshape(-1)
    _check(delta, ref_delta, tol=1e-3, name="delta")
    dS = torch.full((Bn * hq, S, ld), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.gemm_raw(M=S, N=S, K=hd, A=dO, lda=hq * hd, B=v, ldb=hd, Cout=dS, ldc=ld, nz=(hq, 1, Bn),
                 a_zs=(hd, 0, S * hq * hd), b_zs=(0, 0, S * hd), c_zs=(S * ld, 0, hq * S * ld), alpha=scale,
                 band=ops.BAND_SCORES, band_window=W, act=ops.ACT_SOFTMAX_BWD, aux=P, ld_aux=ld, bias=delta)
    dP = torch.einsum("bqhd,bkd->bhqk", dO.float(), vf)
    ref = scale * Pb * (dP - (Pb * dP).sum(-1, keepdim=True))
    got = dS[:, :, :S].float().reshape(Bn, hq, S, S)
    inband = allowed[None, None].expand_as(got)
    assert bool(torch.isfinite(got[inband]).all())
    _check(got[inband], ref[inband], tol=2e-2, name="dS in band")


@pytest.mark.parametrize("M,N,K,K2", [(703, 1536, 1152, 64), (22496, 1152, 1024, 64), (300, 256, 200, 128),
                                      (1000, 13824, 1152, 64), (128, 64, 64, 64)])
def test_gemm_second_contraction_segment(cuda_dev, M, N, K, K2):
    """C = [A | A2] . [B | B2]^T (pt_gemm_args.K2): the LoRA ride-along, forward (all operands K-major) and input
    gradient (B operands MN-major), against the two products summed in fp32."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(M + K2)
    x, w = _rand((M, K), cuda_dev), _rand((N, K), cuda_dev, 0.05)
    t, b = _rand((M, K2), cuda_dev), _rand((N, K2), cuda_dev, 0.05)
    ref = x.float() @ w.float().t() + t.float() @ b.float().t()
    _check(ops.linear(x, w, x2=t, w2=b), ref, name="fwd bf16")
    _check(ops.linear(x, w, x2=t, w2=b, out_dtype=torch.float32), ref, tol=1e-3, name="fwd f32")
    res = torch.randn(M, N, device=cuda_dev)
    _check(ops.linear(x, w, x2=t, w2=b, residual=res, out_dtype=torch.float32), ref + res, tol=1e-3, name="fwd + residual")
    # dx[M, K] = dy[M, N] @ w[N, K] + dt[M, K2] @ a[K2, K]
    dy, dt, a = _rand((M, N), cuda_dev), _rand((M, K2), cuda_dev), _rand((K2, K), cuda_dev, 0.05)
    ref = dy.float() @ w.float() + dt.float() @ a.float()
    _check(ops.linear_dgrad(dy, w, dy2=dt, w2=a, out_dtype=torch.float32), ref, tol=1e-3, name="dgrad f32")
    _check(ops.linear_dgrad(dy, w, dy2=dt, w2=a), ref, name="dgrad bf16")
    # without a segment the launcher is what it was
    _check(ops.linear_dgrad(dy, w, out_dtype=torch.float32), dy.float() @ w.float(), tol=1e-3, name="no segment")


def test_gemm_second_segment_under_fused_epilogues(cuda_dev):
    """The segment feeds the same accumulator the fused epilogues read: GeGLU forward / backward and the attention
    delta epilogue with adapters riding along."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(3)
    M, h, ff, R = 600, 256, 512, 64
    x, w = _rand((M, h), cuda_dev), _rand((2 * ff, h), cuda_dev, 0.06)
    t, b = _rand((M, R), cuda_dev), _rand((2 * ff, R), cuda_dev, 0.06)
    gu, act = torch.empty(M, 2 * ff, device=cuda_dev, dtype=torch.bfloat16), torch.empty(M, ff, device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_geglu(x, w, gu, act, x2=t, w2=b)
    gu_ref = x.float() @ w.float().t() + t.float() @ b.float().t()
    _check(gu, gu_ref, name="gu")
    g_, u_ = gu.float()[:, :ff], gu.float()[:, ff:]
    _check(act, torch.nn.functional.gelu(g_, approximate="tanh") * u_, name="act")
    # down-projection input gradient with the GeGLU backward in the epilogue
    dy, wd = _rand((M, h), cuda_dev), _rand((h, ff), cuda_dev, 0.06)
    dt, a = _rand((M, R), cuda_dev), _rand((R, ff), cuda_dev, 0.06)
    d_gu = torch.empty(M, 2 * ff, device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_dgrad_geglu_bwd(dy, wd, gu, d_gu, dy2=dt, w2=a)
    d_act = dy.float() @ wd.float() + dt.float() @ a.float()
    gr = g_.clone().requires_grad_(True)
    ur = u_.clone().requires_grad_(True)
    (torch.nn.functional.gelu(gr, approximate="tanh") * ur).backward(d_act)
    _check(d_gu[:, :ff], gr.grad, name="d_gate")
    _check(d_gu[:, ff:], ur.grad, name="d_up")
    # o-projection input gradient + delta
    B, S, hq, hd = 2, 300, 2, 128
    dy, wo = _rand((B * S, h), cuda_dev), _rand((h, hq * hd), cuda_dev, 0.06)
    dt, a = _rand((B * S, R), cuda_dev), _rand((R, hq * hd), cuda_dev, 0.06)
    O = _rand((B * S, hq * hd), cuda_dev)
    d_o = torch.empty(B * S, hq * hd, device=cuda_dev, dtype=torch.bfloat16)
    delta = torch.empty(B * hq * S, device=cuda_dev)
    ops.linear_dgrad_attn_delta(dy, wo, O, d_o, delta, S, hd, dy2=dt, w2=a)
    ref = dy.float() @ wo.float() + dt.float() @ a.float()
    _check(d_o, ref, name="d_o")
    dref = (ref * O.float()).view(B, S, hq, hd).sum(-1).permute(0, 2, 1).reshape(-1)
    _check(delta, dref, tol=2e-2, name="delta")


def test_gemm_wgrad_on_column_slices(cuda_dev):
    """One adapter's weight gradient: dB_i = dy[:, rows_i]^T t[:, i*r:(i+1)*r] written into a diagonal block of a
    wider f32 matrix — operand and output views with their parents' row strides, 16-column output."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(4)
    T, out, R, r = 1406, 1536, 64, 16
    dy, t = _rand((T, out), cuda_dev), _rand((T, R), cuda_dev)
    gB = torch.full((out, R), 7.0, device=cuda_dev)
    blocks = [(0, 1024), (1024, 256), (1280, 256)]
    for i, (r0, n) in enumerate(blocks):
        ops.linear_wgrad(dy[:, r0:r0 + n], t[:, i * r:(i + 1) * r], out=gB[r0:r0 + n, i * r:(i + 1) * r])
        ref = dy[:, r0:r0 + n].float().t() @ t[:, i * r:(i + 1) * r].float()
        _check(gB[r0:r0 + n, i * r:(i + 1) * r], ref, tol=2e-3, name=f"block {i}")
    mask = torch.ones(out, R, dtype=torch.bool, device=cuda_dev)
    for i, (r0, n) in enumerate(blocks):
        mask[r0:r0 + n, i * r:(i + 1) * r] = False
    assert (gB[mask] == 7.0).all()          # nothing outside the diagonal blocks was written
    dt, x = _rand((T, R), cuda_dev), _rand((T, 1152), cuda_dev)
    _check(ops.linear_wgrad(dt, x, alpha=2.0), 2.0 * dt.float().t() @ x.float(), tol=2e-3, name="dA")


@pytest.mark.parametrize("T,N,K", [(22496, 64, 1152), (1406, 1024, 16), (22496, 16, 64), (703, 256, 256), (100, 64, 72)])
def test_gemm_split_k_accumulates_into_f32_output(cuda_dev, T, N, K):
    """ksplit: the contraction is cut into slices that run as separate tiles and add into C with f32 atomics
    (C += alpha * A^T B): thin weight gradients contracted over all rows of the batch."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(T + N)
    dy, x = _rand((T, N), cuda_dev), _rand((T, K), cuda_dev)
    ref = dy.float().t() @ x.float()
    out = torch.zeros(N, K, device=cuda_dev)
    ops.linear_wgrad(dy, x, out=out, accumulate=True)
    _check(out, ref, tol=2e-3, name="auto split")
    ops.linear_wgrad(dy, x, out=out, accumulate=True, alpha=0.5)             # adds to what is there
    _check(out, 1.5 * ref, tol=2e-3, name="accumulate")
    # explicit slice counts, K-major operands, a contraction that is not a multiple of the K block
    Tp = (T + 7) // 8 * 8                       # row stride of a K-major operand: a multiple of 8 elements
    a, b = _rand((N, Tp), cuda_dev)[:, :T], _rand((K, Tp), cuda_dev)[:, :T]
    ref = a.float() @ b.float().t()
    for ks in (2, 7, 1000):
        out = torch.zeros(N, K, device=cuda_dev)
        ops.gemm_raw(M=N, N=K, K=T, A=a, lda=Tp, B=b, ldb=Tp, Cout=out, ldc=K, ksplit=ks)
        _check(out, ref, tol=2e-3, name=f"ksplit {ks}")
    with pytest.raises(RuntimeError, match="split-K"):
        ops.gemm_raw(M=N, N=K, K=T, A=a, lda=Tp, B=b, ldb=Tp, Cout=out.to(torch.bfloat16), ldc=K, ksplit=2)


@pytest.mark.parametrize("M,S,K,hq,hkv,seg", [(1406, 703, 1152, 4, 1, False), (300, 100, 512, 8, 4, False),
                                               (22496, 703, 1152, 4, 1, False), (831, 831, 1152, 4, 1, True)])
def test_qknorm_rope_epilogue_equals_the_separate_kernel(cuda_dev, M, S, K, hq, hkv, seg):
    """PT_ACT_QKNORM_ROPE: per-head q/k RMSNorm + RoPE in the epilogue of the q|k|v projection (a 256-wide tile is one
    head) against the projection followed by pt_qknorm_rope_fwd.  The stored projection must be bit-identical; the
    rotated heads differ only through the order of the 256-term sum of squares (one bf16 ulp at most)."""
    from projectiontrainer_b200 import ops
    hd = 256
    N = (hq + 2 * hkv) * hd
    g = torch.Generator(device=cuda_dev).manual_seed(M + K)
    x = torch.randn(M, K, generator=g, device=cuda_dev).bfloat16()
    w = (torch.randn(N, K, generator=g, device=cuda_dev) * K ** -0.5).bfloat16()
    wq, wk = torch.randn(hd, generator=g, device=cuda_dev) * 0.3, torch.randn(hd, generator=g, device=cuda_dev) * 0.3
    ang = torch.rand(S, hd // 2, generator=g, device=cuda_dev) * 6.28
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    kw = {}
    if seg:
        kw = dict(x2=torch.randn(M, 64, generator=g, device=cuda_dev).bfloat16(),
                  w2=(torch.randn(N, 64, generator=g, device=cuda_dev) * 0.05).bfloat16())
    qkv_ref = ops.linear(x, w, **kw)
    q_ref = torch.empty(M, hq * hd, device=cuda_dev, dtype=torch.bfloat16)
    k_ref = torch.empty(M, hkv * hd, device=cuda_dev, dtype=torch.bfloat16)
    ops.qknorm_rope_fwd(qkv_ref, wq, wk, cos, sin, q_ref, k_ref, S, hq, hkv, hd, 1e-6)
    qkv = torch.full((M, N), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    q = torch.full((M, hq * hd), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    k = torch.full((M, hkv * hd), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_qknorm_rope(x, w, qkv, q, k, torch.cat([wq, wk]), torch.cat([cos, sin], 1).contiguous(), hq, 1e-6, **kw)
    torch.cuda.synchronize()
    assert torch.equal(qkv, qkv_ref)
    for got, ref, name in ((q, q_ref, "q"), (k, k_ref, "k")):
        assert bool(torch.isfinite(got.float()).all()), name
        d = (got.float() - ref.float()).abs()
        assert float(d.max()) <= 2 ** -7 * float(ref.float().abs().max()), name
        assert float((d > 0).float().mean()) < 0.02, name          # almost every element is bit-identical
