"""pt_decode_attn — one cached decode step of GQA attention in one launch (q/k RMSNorm + RoPE, cache append,
flash-decoding over a CTA cluster) — against a plain PyTorch fp32 restatement of HF's per-token attention
(HF:models/gemma3/modeling_gemma3.py:278-300,362-384), through the C-ABI.

Tolerance: q_rot / k_rot are rounded to bf16 in both (the kernel keeps pt_qknorm_rope_fwd's rounding), the cache and
the output are bf16; scores, softmax and P V accumulate in f32 in both, so the outputs differ by summation order and
one bf16 rounding: |diff| <= 2^-7 * max|ref| per call."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _ref(qkv, wq, wk, cos, sin, pos, slot, valid, K, V, hq, hkv, hd, window, eps, scale):
    B = qkv.shape[0]
    G = hq // hkv
    x = qkv.float().view(B, hq + 2 * hkv, hd)

    def norm_rope(u, w):                                   # u [B, n, hd]
        n = u * torch.rsqrt(u.pow(2).mean(-1, keepdim=True) + eps) * (1.0 + w)
        c, s = cos[pos][:, None, :], sin[pos][:, None, :]
        a, b = n[..., :hd // 2], n[..., hd // 2:]
        return torch.cat([a * c - b * s, b * c + a * s], -1).bfloat16().float()

    q, k = norm_rope(x[:, :hq], wq), norm_rope(x[:, hq:hq + hkv], wk)
    v = x[:, hq + hkv:]
    Kn, Vn = K.clone(), V.clone()
    Kn[:, slot] = k.reshape(B, hkv * hd).bfloat16()
    Vn[:, slot] = v.reshape(B, hkv * hd).bfloat16()
    Smax = K.shape[1]
    ar = torch.arange(Smax, device=qkv.device)
    vis = (ar <= slot)[None, :].expand(B, Smax).clone()
    if window > 0:
        vis &= (ar > slot - window)[None, :]
    if valid is not None:
        vis &= valid.bool()
    Kf = Kn.float().view(B, Smax, hkv, hd).repeat_interleave(G, 2)      # [B, Smax, hq, hd]
    Vf = Vn.float().view(B, Smax, hkv, hd).repeat_interleave(G, 2)
    sc = torch.einsum("bhd,bshd->bhs", q, Kf) * scale
    sc = sc.masked_fill(~vis[:, None, :], float("-inf"))
    o = torch.einsum("bhs,bshd->bhd", sc.softmax(-1), Vf)
    return o.reshape(B, hq * hd), Kn, Vn


def _case(dev, B, hq, hkv, hd, Smax, slot, window, pads, seed, nsplit=0):
    from projectiontrainer_b200 import ops
    g = torch.Generator(device=dev).manual_seed(seed)
    rnd = lambda *s: torch.randn(*s, generator=g, device=dev)
    qkv = rnd(B, (hq + 2 * hkv) * hd).bfloat16()
    wq, wk = rnd(hd) * 0.3, rnd(hd) * 0.3
    ang = torch.rand(Smax, hd // 2, generator=g, device=dev) * 6.28
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    K = rnd(B, Smax, hkv * hd).bfloat16()
    V = rnd(B, Smax, hkv * hd).bfloat16()
    valid = torch.ones(B, Smax, dtype=torch.uint8, device=dev)
    if pads:
        for b in range(B):
            valid[b, :min((b * 3) % 7, slot)] = 0
    pos = (torch.full((B,), slot, device=dev) - (1 - valid[:, :max(slot, 1)].long()).sum(1)).clamp_(min=0)
    slot_t = torch.tensor([slot], dtype=torch.int64, device=dev)
    eps, scale = 1e-6, hd ** -0.5
    ref, Kr, Vr = _ref(qkv, wq, wk, cos, sin, pos, slot, valid if pads else None, K, V, hq, hkv, hd, window, eps, scale)
    out = torch.full((B, hq * hd), float("nan"), dtype=torch.bfloat16, device=dev)
    ops.decode_attn(qkv, wq, wk, cos, sin, pos, slot_t, valid if pads else None, K, V, out, hq, hkv, hd, window, eps,
                    scale, nsplit=nsplit)
    torch.cuda.synchronize()
    tag = (B, hq, hkv, hd, Smax, slot, window, pads, nsplit)
    assert bool(torch.isfinite(out.float()).all()), tag
    err = float((out.float() - ref).abs().max()) / max(float(ref.abs().max()), 1e-6)
    assert err <= 2 ** -7, (tag, err)
    # the cache append: V row bit-exact, K row = the bf16 rotated key (one bf16 ulp where the f32 order differs)
    assert torch.equal(V, Vr), tag
    kd = (K.float() - Kr.float()).abs()
    assert float(kd.max()) <= 2 ** -7 * float(Kr.float().abs().max()), tag
    rows = torch.ones(Smax, dtype=torch.bool, device=dev)
    rows[slot] = False
    assert torch.equal(K[:, rows], Kr[:, rows]), tag     # nothing but row `slot` was written


@pytest.mark.parametrize("hd", [16, 32, 64, 128, 256])
@pytest.mark.parametrize("G", [1, 2, 4, 8])
def test_every_geometry_matches_the_fp32_restatement(cuda_dev, hd, G):
    hkv = 2
    _case(cuda_dev, B=3, hq=hkv * G, hkv=hkv, hd=hd, Smax=72, slot=37, window=0, pads=False, seed=hd + G)
    _case(cuda_dev, B=3, hq=hkv * G, hkv=hkv, hd=hd, Smax=72, slot=53, window=20, pads=True, seed=hd + G + 1)


@pytest.mark.parametrize("nsplit", [1, 2, 4, 8])
@pytest.mark.parametrize("slot", [0, 1, 7, 40, 299])
def test_cluster_sizes_and_cache_fill_levels(cuda_dev, nsplit, slot):
    """slot 0: only this token's key (every CTA but rank 0 has nothing: the merges take empty partners); slot 1..7:
    fewer keys than warps; 299 = the last slot of the cache."""
    for window in (0, 5, 64):
        _case(cuda_dev, B=2, hq=4, hkv=1, hd=256, Smax=300, slot=slot, window=window, pads=slot >= 7, seed=slot,
              nsplit=nsplit)
        _case(cuda_dev, B=2, hq=8, hkv=2, hd=128, Smax=300, slot=slot, window=window, pads=False, seed=slot + 1,
              nsplit=nsplit)


def test_configs4_geometry_and_a_full_window(cuda_dev):
    """Gemma3-1B decode at the bench's shape (B = 64, hq 4 / hkv 1 / hd 256, 640-slot cache, window 512) and
    Qwen3-8B's (hq 32 / hkv 8 / hd 128)."""
    _case(cuda_dev, B=64, hq=4, hkv=1, hd=256, Smax=640, slot=633, window=512, pads=True, seed=5)
    _case(cuda_dev, B=64, hq=4, hkv=1, hd=256, Smax=640, slot=633, window=0, pads=True, seed=6)
    _case(cuda_dev, B=8, hq=32, hkv=8, hd=128, Smax=640, slot=600, window=0, pads=False, seed=7)


def test_cache_longer_than_the_visibility_table(cuda_dev):
    """Past 32 K keys per CTA the key-visibility bytes are read from `valid` tile by tile instead of from the table in
    shared memory (33 000 slots on one CTA); the same cache through 8 CTAs uses the table."""
    for nsplit in (1, 8):
        _case(cuda_dev, B=1, hq=2, hkv=1, hd=16, Smax=33000, slot=32990, window=0, pads=True, seed=11, nsplit=nsplit)
    _case(cuda_dev, B=2, hq=4, hkv=2, hd=64, Smax=33000, slot=32999, window=4096, pads=False, seed=12, nsplit=1)


def test_rejects_what_it_does_not_take(cuda_dev):
    from projectiontrainer_b200 import ops
    assert not ops.decode_attn_supported(6, 2, 256) and not ops.decode_attn_supported(4, 2, 96)
    assert not ops.decode_attn_supported(32, 2, 128) and ops.decode_attn_supported(32, 8, 128)
    z = torch.zeros(1, 8 * 96, dtype=torch.bfloat16, device=cuda_dev)
    kv = torch.zeros(1, 8, 2 * 96, dtype=torch.bfloat16, device=cuda_dev)
    f = torch.zeros(96, device=cuda_dev)
    t = torch.zeros(8, 48, device=cuda_dev)
    i = torch.zeros(1, dtype=torch.int64, device=cuda_dev)
    with pytest.raises(RuntimeError, match="head_dim"):
        ops.decode_attn(z, f, f, t, t, i, i, None, kv, kv, z, 4, 2, 96, 0, 1e-6, 1.0)
    with pytest.raises(RuntimeError, match="nsplit"):
        z2 = torch.zeros(1, 8 * 64, dtype=torch.bfloat16, device=cuda_dev)
        kv2 = torch.zeros(1, 8, 2 * 64, dtype=torch.bfloat16, device=cuda_dev)
        ops.decode_attn(z2, f, f, t, t, i, i, None, kv2, kv2, z2, 4, 2, 64, 0, 1e-6, 1.0, nsplit=3)
