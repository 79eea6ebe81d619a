"""Fused SigLIP attention kernel vs fp32 torch attention on the same bf16-rounded q/k/v (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,N,H,amp", [(2, 576, 16, 1.5), (3, 100, 2, 1.5), (1, 193, 3, 1.5), (2, 384, 4, 1.5),
                                       (1, 9, 1, 1.5), (40, 576, 16, 1.5), (2, 576, 4, 6.0), (2, 1000, 2, 3.0)])
def test_attn_siglip_fwd(cuda_dev, B, N, H, amp):
    """amp = 6 drives score ranges far past the lazy-rescale threshold (row max grows by > 2^8 between tiles)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(N + H)
    hd, d = 64, H * 64
    qkv = (torch.randn(B * N, 3 * d, device=cuda_dev) * amp).to(torch.bfloat16)
    out = torch.full((B * N, d), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.attn_siglip_fwd(qkv, out, B, N, H, hd, hd ** -0.5)
    torch.cuda.synchronize()
    q, k, v = (t.reshape(B, N, H, hd).transpose(1, 2) for t in qkv.float().split(d, dim=1))
    ref = torch.softmax(q @ k.transpose(-1, -2) * hd ** -0.5, dim=-1) @ v
    ref = ref.transpose(1, 2).reshape(B * N, d)
    assert torch.isfinite(out.float()).all()
    err = (out.float() - ref).abs().max().item()
    assert err < 2e-2 * ref.abs().max().item() + 1e-3, err
    cos = torch.nn.functional.cosine_similarity(out.float().flatten(), ref.flatten(), dim=0).item()
    assert cos > 0.9995, cos


GEMMA_GEOMS = [(2, 200, 4, 1, 0, False), (2, 200, 4, 1, 64, True), (1, 703, 4, 1, 512, True),
               (2, 300, 8, 4, 0, True), (3, 831, 8, 4, 100, False), (1, 40, 2, 2, 0, True),
               (2, 1247, 4, 1, 512, True), (1, 1247, 8, 4, 1024, True)]


def _gemma_case(dev, B, S, hq, hkv, W, pad, hd, amp=1.0):
    """bf16 q_rot / k_rot / fused qkv (only its v columns matter), key-valid flags and the fp32 torch view of them."""
    torch.manual_seed(S + hq + W)
    G = hq // hkv
    QKV = (hq + 2 * hkv) * hd
    q = (torch.randn(B * S, hq * hd, device=dev) * amp).to(torch.bfloat16)
    k = (torch.randn(B * S, hkv * hd, device=dev) * amp).to(torch.bfloat16)
    qkv = (torch.randn(B * S, QKV, device=dev)).to(torch.bfloat16)
    keyvalid = torch.ones(B, S, dtype=torch.uint8, device=dev)
    if pad:
        for b in range(B):
            keyvalid[b, :7 * b + 3] = 0
        keyvalid[0, S // 2] = 0                      # an isolated masked key inside the band
    qf = q.float().view(B, S, hq, hd).transpose(1, 2)                                  # [B,hq,S,hd]
    kf = k.float().view(B, S, hkv, hd).transpose(1, 2)                                 # [B,hkv,S,hd]
    vf = qkv.float().view(B, S, QKV)[:, :, (hq + hkv) * hd:].reshape(B, S, hkv, hd).transpose(1, 2)
    qi, ki = torch.arange(S, device=dev)[:, None], torch.arange(S, device=dev)[None, :]
    allowed = (ki <= qi) & ((ki > qi - W) if W else (ki <= qi))
    allowed = allowed[None, None] & keyvalid.bool()[:, None, None, :]
    return q, k, qkv, keyvalid, qf, kf, vf, allowed, G, QKV


@pytest.mark.parametrize("B,S,hq,hkv,W,pad", GEMMA_GEOMS)
@pytest.mark.parametrize("hd", [256, 128])
@pytest.mark.parametrize("amp", [1.0, 4.0], ids=["amp1", "amp4"])
def test_attn_gemma_fwd(cuda_dev, B, S, hq, hkv, W, pad, hd, amp):
    """Fused causal / sliding-window GQA attention (one pass, online softmax) vs fp32 torch on the same
    bf16-rounded q/k/v: O and the log2-domain log-sum-exp of every row.  amp = 4 drives the scores far past the
    lazy-rescale threshold (a row max growing by more than 2^8 between key tiles)."""
    from projectiontrainer_b200 import ops
    q, k, qkv, keyvalid, qf, kf, vf, allowed, G, QKV = _gemma_case(cuda_dev, B, S, hq, hkv, W, pad, hd, amp)
    kvbits = torch.empty(B * ((S + 63) // 64), dtype=torch.int64, device=cuda_dev)
    lse = torch.full((B * hq, S), float("nan"), dtype=torch.float32, device=cuda_dev)
    O = torch.full((B * S, hq * hd), float("nan"), dtype=torch.bfloat16, device=cuda_dev)
    scale = 1.0 / 16
    ops.pack_keyvalid(keyvalid, kvbits)
    ops.attn_gemma_fwd(q, k, qkv, kvbits, lse, O, B, S, hq, hkv, hd, W, scale)
    torch.cuda.synchronize()
    sc = (qf @ kf.repeat_interleave(G, dim=1).transpose(-1, -2)) * scale
    scm = sc.masked_fill(~allowed, float("-inf"))
    Pref = torch.nan_to_num(torch.softmax(scm, dim=-1), nan=0.0)
    Oref = (Pref @ vf.repeat_interleave(G, dim=1)).transpose(1, 2).reshape(B * S, hq * hd)
    rows_with_keys = allowed.any(-1).expand(B, hq, S)
    live = rows_with_keys.transpose(1, 2).reshape(B * S, hq)[:, :, None].expand(B * S, hq, hd).reshape(B * S, hq * hd)
    got = O.float()
    assert bool(torch.isfinite(got[live]).all())
    err = (got[live] - Oref[live]).abs().max().item()
    assert err < 2e-2 * Oref.abs().max().item() + 1e-3, err
    lse_ref = torch.logsumexp(scm, dim=-1) * 1.4426950408889634                    # log2 domain
    lg = lse.view(B, hq, S)
    assert bool((lg[~rows_with_keys] == float("inf")).all())                       # no visible key: +inf
    assert float((lg[rows_with_keys] - lse_ref[rows_with_keys]).abs().max()) < 2e-2
    # inference call: no LSE wanted
    O2 = torch.empty_like(O)
    ops.attn_gemma_fwd(q, k, qkv, kvbits, None, O2, B, S, hq, hkv, hd, W, scale)
    torch.cuda.synchronize()
    assert torch.equal(O2[live], O[live])


@pytest.mark.parametrize("B,S,hq,hkv,W,pad", GEMMA_GEOMS)
@pytest.mark.parametrize("hd", [256, 128])
@pytest.mark.parametrize("pair", [True, False], ids=["cta_pair", "single_cta"])
def test_attn_gemma_bwd(cuda_dev, B, S, hq, hkv, W, pad, hd, pair):
    """Fused key/value-stationary backward (P recomputed from the forward's log-sum-exp) + the band-limited dQ GEMM
    vs fp32 torch autograd through softmax attention on the same bf16-rounded q/k/v/dO: dQ, dK, dV (max error
    < 2e-2 of range, cosine > 0.9995), exact zeros on masked dS entries and on rows without a visible key."""
    from projectiontrainer_b200 import ops
    q, k, qkv, keyvalid, qf, kf, vf, allowed, G, QKV = _gemma_case(cuda_dev, B, S, hq, hkv, W, pad, hd)
    ld = (S + 7) // 8 * 8
    kvbits = torch.empty(B * ((S + 63) // 64), dtype=torch.int64, device=cuda_dev)
    lse = torch.empty((B * hq, S), dtype=torch.float32, device=cuda_dev)
    O = torch.zeros((B * S, hq * hd), dtype=torch.bfloat16, device=cuda_dev)
    scale = 1.0 / 16
    ops.pack_keyvalid(keyvalid, kvbits)
    ops.attn_gemma_fwd(q, k, qkv, kvbits, lse, O, B, S, hq, hkv, hd, W, scale)
    dO = torch.randn(B * S, hq * hd, device=cuda_dev).to(torch.bfloat16)
    rows_with_keys = allowed.any(-1).expand(B, hq, S)                                      # [B,hq,S]
    live_rows = rows_with_keys.transpose(1, 2).reshape(B * S, hq)
    # rows without a visible key carry no gradient in the engine (their O is masked out downstream): zero dO there
    dO = (dO.view(B * S, hq, hd) * live_rows[:, :, None]).reshape(B * S, hq * hd).contiguous()
    O = (O.view(B * S, hq, hd) * live_rows[:, :, None]).reshape(B * S, hq * hd).contiguous()
    delta = torch.empty(B * hq * S, dtype=torch.float32, device=cuda_dev)
    ops.attn_delta(dO, O, delta, S, hq, hd)
    dS = torch.zeros(B * hq, S, ld, dtype=torch.bfloat16, device=cuda_dev)
    dK = torch.full((B * S, hkv * hd), float("nan"), dtype=torch.bfloat16, device=cuda_dev)
    d_qkv = torch.full((B * S, QKV), 7.0, dtype=torch.bfloat16, device=cuda_dev)
    dQ = torch.full((B * S, hq * hd), float("nan"), dtype=torch.bfloat16, device=cuda_dev)
    if pair:        # CTA pair per 128 keys; dS leaves transposed ([head, key, query]) for an MN-major dQ GEMM
        ops.attn_gemma_bwd_pair(q, k, qkv, dO, kvbits, lse, delta, dS, dK, d_qkv[:, (hq + hkv) * hd:], B, S, hq, hkv,
                                hd, W, scale, ld)
        ops.attn_dq_from_dst(dS, k, dQ, B, S, hq, hkv, hd, W, ld)
    else:
        ops.attn_gemma_bwd(q, k, qkv, dO, kvbits, lse, delta, dS, dK, d_qkv[:, (hq + hkv) * hd:], B, S, hq, hkv, hd, W,
                           scale, ld)
        ops.gemm_raw(M=S, N=hd, K=S, A=dS, lda=ld, B=k, ldb=hkv * hd, b_mn=True, Cout=dQ, ldc=hq * hd, nz=(G, hkv, B),
                     a_zs=(S * ld, G * S * ld, hq * S * ld), b_zs=(0, hd, S * hkv * hd), c_zs=(hd, G * hd, S * hq * hd),
                     band=ops.BAND_K_BY_QUERY, band_window=W)
    torch.cuda.synchronize()
    # fp32 autograd reference
    qr, kr, vr = qf.clone().requires_grad_(True), kf.clone().requires_grad_(True), vf.clone().requires_grad_(True)
    sc = (qr @ kr.repeat_interleave(G, dim=1).transpose(-1, -2)) * scale
    P = torch.nan_to_num(torch.softmax(sc.masked_fill(~allowed, float("-inf")), dim=-1), nan=0.0)
    out = (P @ vr.repeat_interleave(G, dim=1)).transpose(1, 2).reshape(B * S, hq * hd)
    out.backward(dO.float())
    dq_ref = qr.grad.transpose(1, 2).reshape(B * S, hq * hd)
    dk_ref = kr.grad.transpose(1, 2).reshape(B * S, hkv * hd)
    dv_ref = vr.grad.transpose(1, 2).reshape(B * S, hkv * hd)
    assert (d_qkv[:, :(hq + hkv) * hd] == 7.0).all()                 # q / k columns belong to the caller
    dV = d_qkv[:, (hq + hkv) * hd:].float()
    for name, got, ref in (("dV", dV, dv_ref), ("dK", dK.float(), dk_ref), ("dQ", dQ.float(), dq_ref)):
        assert bool(torch.isfinite(got).all()), name
        err = (got - ref).abs().max().item()
        assert err < 2e-2 * ref.abs().max().item() + 1e-4, (name, err, ref.abs().max().item())
        cos = torch.nn.functional.cosine_similarity(got.flatten(), ref.flatten(), dim=0).item()
        assert cos > 0.9995, (name, cos)
    dSg = dS[:, :, :S].float().view(B, hq, S, S)
    if pair:
        dSg = dSg.transpose(-1, -2)
    assert float(dSg[~allowed.expand_as(dSg)].abs().max()) == 0.0    # masked / out of band: exactly 0
    assert float(dQ.float().view(B * S, hq, hd)[~live_rows].abs().max() if (~live_rows).any() else 0.0) == 0.0


@pytest.mark.parametrize("rows,S,hq,hkv,hd", [(6000, 750, 4, 1, 256), (5000, 500, 8, 4, 128), (300, 100, 4, 1, 256),
                                              (257, 257, 2, 2, 64), (96, 32, 2, 1, 512), (4800, 160, 3, 2, 72)])
def test_qknorm_rope_fwd_bwd(cuda_dev, rows, S, hq, hkv, hd):
    """q/k RMS-norm ((1 + w) gain) + rotate-half RoPE over the fused qkv activation, forward and backward, vs fp32
    torch autograd on the same bf16-rounded inputs.  Covers the warp-per-row kernel (rows >= 4736, head_dim <=
    256), the warp-per-(row, head) register kernel and the generic head_dim > 256 loop."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + hd)
    nh, half, eps = hq + hkv, hd // 2, 1e-6
    qkv = torch.randn(rows, (hq + 2 * hkv) * hd, device=cuda_dev).to(torch.bfloat16)
    wq, wk = torch.randn(hd, device=cuda_dev) * 0.3, torch.randn(hd, device=cuda_dev) * 0.3
    ang = torch.rand(S, half, device=cuda_dev) * 6.28
    cs, sn = ang.cos().contiguous(), ang.sin().contiguous()
    q_out = torch.full((rows, hq * hd), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    k_out = torch.full((rows, hkv * hd), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.qknorm_rope_fwd(qkv, wq, wk, cs, sn, q_out, k_out, S, hq, hkv, hd, eps)

    x = qkv.float()[:, :nh * hd].reshape(rows, nh, hd).requires_grad_(True)
    w = torch.cat([wq.expand(hq, hd), wk.expand(hkv, hd)], dim=0)
    n = x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * (1 + w)
    pos = torch.arange(rows, device=cuda_dev) % S
    c, s = cs[pos][:, None, :], sn[pos][:, None, :]
    y = torch.cat([n[..., :half] * c - n[..., half:] * s, n[..., half:] * c + n[..., :half] * s], dim=-1)
    got = torch.cat([q_out.float().reshape(rows, hq, hd), k_out.float().reshape(rows, hkv, hd)], dim=1)
    assert torch.isfinite(got).all()
    assert (got - y).abs().max().item() < 2e-2 * y.abs().max().item()

    dy = torch.randn(rows, nh, hd, device=cuda_dev).to(torch.bfloat16)
    y.backward(dy.float())
    dq, dk = dy[:, :hq].reshape(rows, hq * hd).contiguous(), dy[:, hq:].reshape(rows, hkv * hd).contiguous()
    d_qkv = torch.full_like(qkv, 7.0)
    ops.qknorm_rope_bwd(dq, dk, qkv, wq, wk, cs, sn, d_qkv, S, hq, hkv, hd, eps)
    torch.cuda.synchronize()
    gd = d_qkv.float()[:, :nh * hd].reshape(rows, nh, hd)
    assert (gd - x.grad).abs().max().item() < 2e-2 * x.grad.abs().max().item()
    assert torch.nn.functional.cosine_similarity(gd.flatten(), x.grad.flatten(), dim=0).item() > 0.9999
    assert (d_qkv[:, nh * hd:] == 7.0).all()           # the v columns belong to the caller
