"""Fused SigLIP attention kernel vs fp32 torch attention on the same bf16-rounded q/k/v (GPU only)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("B,N,H", [(2, 576, 16), (3, 100, 2), (1, 193, 3), (2, 384, 4), (1, 9, 1), (40, 576, 16)])
def test_attn_siglip_fwd(cuda_dev, B, N, H):
    from projectiontrainer_b200 import ops
    torch.manual_seed(N + H)
    hd, d = 64, H * 64
    qkv = (torch.randn(B * N, 3 * d, device=cuda_dev) * 1.5).to(torch.bfloat16)
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
