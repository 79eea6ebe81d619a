"""pt_set_dependent_launch: kernels launched with the programmatic-dependent-launch attribute (their prologues overlap
the predecessor's tail; each waits for the predecessor before its first dependent access) give bit-identical results
to ordinary launches — a chain in which every kernel reads what the one before it wrote, eager and as a CUDA graph."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _chain(dev, kinds, graph):
    from projectiontrainer_b200 import _lib, ops
    g = torch.Generator(device=dev).manual_seed(3)
    rnd = lambda *s: torch.randn(*s, generator=g, device=dev)
    B, h, hq, hkv, hd, Smax, slot = 16, 512, 4, 2, 64, 96, 70
    x0 = rnd(B, h)
    a0 = rnd(B, h).bfloat16()
    Ws = [(rnd(h, h) * h ** -0.5).bfloat16() for _ in range(6)]
    w_qkv = (rnd((hq + 2 * hkv) * hd, h) * h ** -0.5).bfloat16()
    wn = [rnd(h) * 0.2 for _ in range(12)]
    ang = torch.rand(Smax, hd // 2, generator=g, device=dev)
    cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
    K0, V0 = rnd(B, Smax, hkv * hd).bfloat16(), rnd(B, Smax, hkv * hd).bfloat16()
    pos = torch.full((B,), slot, dtype=torch.int64, device=dev)
    slot_t = torch.tensor([slot], dtype=torch.int64, device=dev)
    wq = rnd(hd) * 0.1
    y = torch.empty(B, h, device=dev, dtype=torch.bfloat16)
    x = [torch.empty(B, h, device=dev) for _ in range(2)]
    a = torch.empty(B, h, device=dev, dtype=torch.bfloat16)
    qkv = torch.empty(B, (hq + 2 * hkv) * hd, device=dev, dtype=torch.bfloat16)
    o = torch.empty(B, hq * hd, device=dev, dtype=torch.bfloat16)
    K, V = K0.clone(), V0.clone()

    def run():
        a.copy_(a0); x[0].copy_(x0); K.copy_(K0); V.copy_(V0)
        for i, W in enumerate(Ws):                  # GEMM -> norm -> GEMM -> ...: every launch depends on the last
            ops.linear(a, W, out=y)
            ops.rmsnorm_residual_fwd(y, wn[2 * i], x[i % 2], x[(i + 1) % 2], wn[2 * i + 1], a, 1e-6)
        ops.linear(a, w_qkv, out=qkv)
        ops.decode_attn(qkv, wq, wq, cos, sin, pos, slot_t, None, K, V, o, hq, hkv, hd, 0, 1e-6, hd ** -0.5)
        ops.linear(o[:, :h // 2].contiguous() if False else o, Ws[0][:, :hq * hd].contiguous(), out=y)

    was = _lib.lib().pt_set_dependent_launch(kinds)
    try:
        run()
        torch.cuda.synchronize()
        if graph:
            gr = torch.cuda.CUDAGraph()
            with torch.cuda.graph(gr):
                run()
            for _ in range(3):
                gr.replay()
        torch.cuda.synchronize()
    finally:
        assert _lib.lib().pt_set_dependent_launch(was) == kinds
    return y.clone(), x[0].clone(), o.clone(), K.clone()


@pytest.mark.parametrize("graph", [False, True], ids=["eager", "graph"])
def test_dependent_launch_changes_no_result(cuda_dev, graph):
    ref = _chain(cuda_dev, 0, graph)
    for kinds in (1, 2, 4, 7):
        got = _chain(cuda_dev, kinds, graph)
        for r, t in zip(ref, got):
            assert torch.equal(r, t), kinds
