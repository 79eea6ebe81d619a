"""pt_gemv_small_m — the decode step's linear layers at batch <= 16 as weight streaming on mma.sync — against
pt_gemm_bf16 (the tcgen05 GEMM, itself pinned to fp32 references in test_gemm_gpu.py) and a plain fp32 product.
Tolerance: both accumulate bf16 products in f32, in different orders: 2^-8 relative to the row scale for bf16
outputs (one rounding), 1e-5 for f32 outputs."""
import pytest
import torch

pytestmark = pytest.mark.gpu

SHAPES = [(1536, 1152), (1152, 1024), (1152, 6912), (64, 32), (1000, 96), (262144 // 16, 1152)]


@pytest.mark.parametrize("M", [1, 3, 8, 9, 16, 17, 32])
@pytest.mark.parametrize("N,K", SHAPES)
def test_store_modes_match_the_fp32_product(cuda_dev, M, N, K):
    from projectiontrainer_b200 import ops
    g = torch.Generator(device=cuda_dev).manual_seed(M + N + K)
    x = torch.randn(M, K + 8, generator=g, device=cuda_dev).bfloat16()[:, :K]        # row stride != K
    w = (torch.randn(N, K, generator=g, device=cuda_dev) * K ** -0.5).bfloat16()
    ref = x.float() @ w.float().t()
    out = torch.full((M, N), float("nan"), device=cuda_dev)
    ops.gemv(x, w, out, ops.GEMV_F32)
    assert float((out - ref).abs().max()) <= 1e-5 * max(1.0, float(ref.abs().max())) + 1e-5
    ob = torch.full((M, N + 8), float("nan"), device=cuda_dev, dtype=torch.bfloat16)[:, :N]
    ops.gemv(x, w, ob)
    assert float((ob.float() - ref).abs().max()) <= 2 ** -8 * float(ref.abs().max()) + 1e-6
    if N % 16 == 0 and K % 64 == 0 and N >= 128:           # the tensor-core GEMM's own alignment rules
        assert float((ob.float() - ops.linear(x.contiguous(), w).float()).abs().max()) <= 2 ** -7 * float(ref.abs().max())


@pytest.mark.parametrize("M", [1, 5, 16, 32])
@pytest.mark.parametrize("ksplit", [0, 1, 3, 27])
def test_atomic_slices_accumulate(cuda_dev, M, ksplit):
    from projectiontrainer_b200 import ops
    g = torch.Generator(device=cuda_dev).manual_seed(M + ksplit)
    N, K = 1152, 6912
    x = torch.randn(M, K, generator=g, device=cuda_dev).bfloat16()
    w = (torch.randn(N, K, generator=g, device=cuda_dev) * K ** -0.5).bfloat16()
    ref = x.float() @ w.float().t()
    acc = torch.ones(M, N, device=cuda_dev)               # adds onto what is there
    ops.gemv(x, w, acc, ops.GEMV_F32_ATOMIC, ksplit=ksplit)
    assert float((acc - 1.0 - ref).abs().max()) <= 2e-5 * max(1.0, float(ref.abs().max()))
    with pytest.raises(RuntimeError, match="slices"):
        ops.gemv(x, w, torch.empty(M, N, device=cuda_dev), ops.GEMV_F32, ksplit=4)


@pytest.mark.parametrize("silu", [False, True], ids=["geglu", "swiglu"])
@pytest.mark.parametrize("M", [1, 7, 12, 29])
def test_glu_modes_match_the_fused_gemm_epilogue(cuda_dev, M, silu):
    from projectiontrainer_b200 import ops
    g = torch.Generator(device=cuda_dev).manual_seed(M)
    K, ff = 1152, 6912
    x = torch.randn(M, K, generator=g, device=cuda_dev).bfloat16()
    w = (torch.randn(2 * ff, K, generator=g, device=cuda_dev) * K ** -0.5).bfloat16()
    act = torch.full((M, ff), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.gemv(x, w, act, ops.GEMV_SWIGLU if silu else ops.GEMV_GEGLU)
    gu = torch.empty(M, 2 * ff, device=cuda_dev, dtype=torch.bfloat16)
    act_ref = torch.empty(M, ff, device=cuda_dev, dtype=torch.bfloat16)
    ops.linear_geglu(x, w, gu, act_ref, silu=silu)
    pre = (x.float() @ w.float().t()).bfloat16().float()
    gate, up = pre[:, :ff], pre[:, ff:]
    want = (torch.nn.functional.silu(gate) if silu else torch.nn.functional.gelu(gate, approximate="tanh")) * up
    scale = float(want.abs().max())
    assert float((act.float() - want).abs().max()) <= 2 ** -6 * scale
    assert float((act.float() - act_ref.float()).abs().max()) <= 2 ** -6 * scale


def test_rejects(cuda_dev):
    from projectiontrainer_b200 import ops
    x = torch.zeros(33, 64, device=cuda_dev, dtype=torch.bfloat16)
    w = torch.zeros(32, 64, device=cuda_dev, dtype=torch.bfloat16)
    with pytest.raises(RuntimeError, match="1 .. 32"):
        ops.gemv(x, w, torch.empty(33, 32, device=cuda_dev, dtype=torch.bfloat16))
    with pytest.raises(RuntimeError, match="multiple of 32"):
        ops.gemv(x[:4, :48].contiguous(), w[:, :48].contiguous(), torch.empty(4, 32, device=cuda_dev, dtype=torch.bfloat16))
    assert ops.gemv_ok(32, 1152) and not ops.gemv_ok(33, 1152) and not ops.gemv_ok(4, 48) and not ops.gemv_ok(17, 1152, 16)
