"""Normalisation kernels through the C ABI vs fp32 torch (values) and torch autograd (gradients), at the odd row
counts of the real step (575 / 703 / 831 tokens per sample; 18 400 / 22 496 rows per batch) and the three widths the
towers use (SigLIP 1024, Gemma3-1B 1152, Gemma3-4B 2560) — SURVEY.md §4, VERDICT r1 "bandwidth kernels have no unit
tests".  bf16 outputs are held to bf16 rounding of the fp32 reference; f32 outputs to 1e-5 relative."""
import pytest
import torch

pytestmark = pytest.mark.gpu

ROWS_D = [(575, 1024), (703, 1152), (831, 2560), (18400, 1024), (22496, 1152), (5, 64), (1, 2560)]


def _bf16_close(got, ref, what=""):
    """got: bf16 tensor from a kernel; ref: fp32 reference.  One bf16 ulp of slack on top of exact rounding."""
    g, r = got.float(), ref.float()
    assert torch.isfinite(g).all(), what
    tol = 2.0 ** -7 * r.abs() + 1e-6
    bad = (g - r).abs() > tol
    assert not bad.any(), (what, int(bad.sum()), float((g - r).abs().max()))


def _rms(x, w, eps):
    x = x.float()
    return x * torch.rsqrt(x.pow(2).mean(-1, keepdim=True) + eps) * (1 + w.float())


@pytest.mark.parametrize("rows,D", ROWS_D)
def test_layernorm_fwd(cuda_dev, rows, D):
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + D)
    x = torch.randn(rows, D, device=cuda_dev) * 3 + 0.5
    g, b = torch.randn(D, device=cuda_dev), torch.randn(D, device=cuda_dev)
    out = torch.full((rows, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.layernorm_fwd(x, g, b, out, 1e-6)
    ref = torch.nn.functional.layer_norm(x, (D,), g, b, 1e-6)
    _bf16_close(out, ref, "layernorm")


@pytest.mark.parametrize("B,N,D", [(3, 576, 1024), (2, 10, 64)])
def test_layernorm_fwd_drops_token0(cuda_dev, B, N, D):
    """group remap = the reference's `last_hidden_state[:, 1:, :]` (ref:Stage1/projector_trainer.py:173)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(N)
    x = torch.randn(B * N, D, device=cuda_dev)
    g, b = torch.randn(D, device=cuda_dev), torch.randn(D, device=cuda_dev)
    out = torch.full((B * (N - 1), D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.layernorm_fwd(x, g, b, out, 1e-6, group=(N - 1, N, 1))
    ref = torch.nn.functional.layer_norm(x, (D,), g, b, 1e-6).view(B, N, D)[:, 1:].reshape(-1, D)
    _bf16_close(out, ref, "layernorm drop-first")


@pytest.mark.parametrize("rows,D", ROWS_D)
@pytest.mark.parametrize("xdt", [torch.float32, torch.bfloat16])
def test_rmsnorm_fwd(cuda_dev, rows, D, xdt):
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows * 3 + D)
    x = (torch.randn(rows, D, device=cuda_dev) * 2).to(xdt)
    w = torch.randn(D, device=cuda_dev) * 0.3
    out = torch.full((rows, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.rmsnorm_fwd(x, w, out, 1e-6)
    _bf16_close(out, _rms(x, w, 1e-6), "rmsnorm")


def test_rmsnorm_fwd_label_rows(cuda_dev):
    """group remap (T, S, lab0): only the rows whose shifted target can be a label go through the final norm."""
    from projectiontrainer_b200 import ops
    B, S, T, lab0, D = 3, 703, 128, 574, 1152
    x = torch.randn(B * S, D, device=cuda_dev)
    w = torch.randn(D, device=cuda_dev) * 0.3
    out = torch.full((B * T, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.rmsnorm_fwd(x, w, out, 1e-6, group=(T, S, lab0))
    ref = _rms(x, w, 1e-6).view(B, S, D)[:, lab0:lab0 + T].reshape(-1, D)
    _bf16_close(out, ref, "rmsnorm label rows")


@pytest.mark.parametrize("rows,D", ROWS_D)
@pytest.mark.parametrize("with_next", [True, False])
def test_rmsnorm_residual_fwd(cuda_dev, rows, D, with_next):
    """x_out = x_in + RMS(y; w_post); next = RMS(x_out; w_next) (HF:models/gemma3/modeling_gemma3.py:417-432)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + 7 * D)
    y = torch.randn(rows, D, device=cuda_dev).to(torch.bfloat16)
    x_in = torch.randn(rows, D, device=cuda_dev) * 4
    wp, wn = torch.randn(D, device=cuda_dev) * 0.3, torch.randn(D, device=cuda_dev) * 0.3
    x_out = torch.full((rows, D), float("nan"), device=cuda_dev)
    nxt = torch.full((rows, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.rmsnorm_residual_fwd(y, wp, x_in, x_out, wn if with_next else None, nxt if with_next else None, 1e-6)
    # HF rounds the post-norm output to the activation dtype before the residual add; the kernel keeps it in f32:
    # the two differ by at most one bf16 rounding of the branch
    branch = _rms(y, wp, 1e-6)
    ref_out = x_in + branch
    assert (x_out - ref_out).abs().max().item() <= 2.0 ** -8 * branch.abs().max().item() + 1e-6
    if with_next:
        _bf16_close(nxt, _rms(x_out, wn, 1e-6), "next pre-norm")
    else:
        assert torch.isnan(nxt.float()).all()          # untouched


@pytest.mark.parametrize("rows,D", [(64, 1152), (3, 128), (64, 2560), (8, 4096)])
@pytest.mark.parametrize("with_next", [True, False])
def test_rmsnorm_residual_fwd_acc_equals_the_bf16_input_kernel_and_clears(cuda_dev, rows, D, with_next):
    """The split-K accumulator variant: y_acc (f32 sums of the GEMM slices) is rounded to bf16, used exactly as
    pt_rmsnorm_residual_fwd uses a bf16 y (bit-equal outputs), and left zeroed; and through a split-K GEMM it
    reproduces GEMM -> bf16 y -> norm up to the summation order of the slices."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + D)
    y_acc = torch.randn(rows, D, device=cuda_dev)
    x_in = torch.randn(rows, D, device=cuda_dev) * 4
    wp, wn = torch.randn(D, device=cuda_dev) * 0.3, torch.randn(D, device=cuda_dev) * 0.3
    outs = []
    for acc in (False, True):
        x_out = torch.full((rows, D), float("nan"), device=cuda_dev)
        nxt = torch.full((rows, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
        if acc:
            y = y_acc.clone()
            ops.rmsnorm_residual_fwd_acc(y, wp, x_in, x_out, wn if with_next else None, nxt if with_next else None, 1e-6)
            assert int(torch.count_nonzero(y)) == 0
        else:
            ops.rmsnorm_residual_fwd(y_acc.bfloat16(), wp, x_in, x_out, wn if with_next else None,
                                     nxt if with_next else None, 1e-6)
        outs.append((x_out, nxt))
    assert torch.equal(outs[0][0], outs[1][0])
    if with_next:
        assert torch.equal(outs[0][1], outs[1][1])
    # behind a split-K GEMM, twice in a row on the same accumulator (the second finds it cleared)
    K = 1024
    a = torch.randn(rows, K, device=cuda_dev).bfloat16()
    w = (torch.randn(D, K, device=cuda_dev) * K ** -0.5).bfloat16()
    y_ref = ops.linear(a, w)
    x_ref = torch.empty(rows, D, device=cuda_dev)
    ops.rmsnorm_residual_fwd(y_ref, wp, x_in, x_ref, None, None, 1e-6)
    acc = torch.zeros(rows, D, device=cuda_dev)
    for _ in range(2):
        ops.linear(a, w, out=acc, ksplit=-1)
        x_out = torch.empty(rows, D, device=cuda_dev)
        ops.rmsnorm_residual_fwd_acc(acc, wp, x_in, x_out, None, None, 1e-6)
        branch = (x_ref - x_in).abs().max().item()
        assert (x_out - x_ref).abs().max().item() <= 2.0 ** -7 * branch
    with pytest.raises(RuntimeError, match="4096"):
        big = torch.zeros(2, 4104, device=cuda_dev)
        ops.rmsnorm_residual_fwd_acc(big, torch.zeros(4104, device=cuda_dev), big.clone(), big.clone(), None, None, 1e-6)


def _rms_bwd_ref(g, u, w, eps):
    u = u.float().clone().requires_grad_(True)
    _rms(u, w, eps).backward(g.float())
    return u.grad


@pytest.mark.parametrize("rows,D", ROWS_D)
@pytest.mark.parametrize("gdt,udt,ddt", [(torch.bfloat16, torch.float32, torch.float32),
                                         (torch.float32, torch.bfloat16, torch.bfloat16),
                                         (torch.bfloat16, torch.bfloat16, torch.bfloat16)])
def test_rmsnorm_bwd(cuda_dev, rows, D, gdt, udt, ddt):
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + D + 1)
    g = torch.randn(rows, D, device=cuda_dev).to(gdt)
    u = (torch.randn(rows, D, device=cuda_dev) * 2).to(udt)
    w = torch.randn(D, device=cuda_dev) * 0.3
    ref = _rms_bwd_ref(g, u, w, 1e-6)
    if ddt == torch.float32:
        add = torch.randn(rows, D, device=cuda_dev)
        dst = torch.full((rows, D), float("nan"), device=cuda_dev)
        ops.rmsnorm_bwd(g, u, w, dst, 1e-6, add=add)
        assert (dst - (add + ref)).abs().max().item() <= 2e-5 * (add + ref).abs().max().item() + 1e-6
        dst2 = torch.full((rows, D), float("nan"), device=cuda_dev)
        ops.rmsnorm_bwd(g, u, w, dst2, 1e-6)
        assert (dst2 - ref).abs().max().item() <= 2e-5 * ref.abs().max().item() + 1e-6
    else:
        dst = torch.full((rows, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
        ops.rmsnorm_bwd(g, u, w, dst, 1e-6)
        _bf16_close(dst, ref, "rmsnorm_bwd")


def test_rmsnorm_bwd_label_rows_scatter(cuda_dev):
    """Final-norm backward: R = B*T gradient rows, the norm input and the f32 destination addressed through the
    (T, S, lab0) remap; rows outside the label span keep the caller's values."""
    from projectiontrainer_b200 import ops
    B, S, T, lab0, D = 2, 703, 128, 574, 1152
    g = torch.randn(B * T, D, device=cuda_dev).to(torch.bfloat16)
    u = torch.randn(B * S, D, device=cuda_dev)
    w = torch.randn(D, device=cuda_dev) * 0.3
    dst = torch.zeros(B * S, D, device=cuda_dev)
    ops.rmsnorm_bwd(g, u, w, dst, 1e-6, rows=B * T, u_group=(T, S, lab0), dst_group=(T, S, lab0))
    u_sel = u.view(B, S, D)[:, lab0:lab0 + T].reshape(-1, D)
    ref = torch.zeros(B, S, D, device=cuda_dev)
    ref[:, lab0:lab0 + T] = _rms_bwd_ref(g, u_sel, w, 1e-6).view(B, T, D)
    assert (dst - ref.view(-1, D)).abs().max().item() <= 2e-5 * ref.abs().max().item() + 1e-6
    assert (dst.view(B, S, D)[:, :lab0] == 0).all() and (dst.view(B, S, D)[:, lab0 + T:] == 0).all()


@pytest.mark.parametrize("rows,D", [r for r in ROWS_D if r[1] <= 2560])
def test_rmsnorm_bwd_pair(cuda_dev, rows, D):
    """dx += rmsnorm_bwd(g1; u1, w1); out2 = bf16(rmsnorm_bwd(dx; u2, w2)): pre-norm backward into the f32 gradient
    stream + the preceding post-norm backward, vs two autograd calls."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + 2 * D)
    g1 = torch.randn(rows, D, device=cuda_dev).to(torch.bfloat16)
    u1 = torch.randn(rows, D, device=cuda_dev) * 3
    u2 = torch.randn(rows, D, device=cuda_dev).to(torch.bfloat16)
    w1, w2 = torch.randn(D, device=cuda_dev) * 0.3, torch.randn(D, device=cuda_dev) * 0.3
    dx0 = torch.randn(rows, D, device=cuda_dev)
    dx = dx0.clone()
    out2 = torch.full((rows, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.rmsnorm_bwd_pair(g1, u1, w1, dx, u2, w2, out2, 1e-6)
    dx_ref = dx0 + _rms_bwd_ref(g1, u1, w1, 1e-6)
    assert (dx - dx_ref).abs().max().item() <= 2e-5 * dx_ref.abs().max().item() + 1e-6
    _bf16_close(out2, _rms_bwd_ref(dx_ref, u2, w2, 1e-6), "rmsnorm_bwd_pair out2")
