"""The remaining bandwidth kernels through the C ABI: the splice (INT parts bit-exact), the masked
vocabulary cross-entropy, clip + AdamW, and the small reductions / casts — each against plain fp32 torch on the same
inputs ("bandwidth kernels have no unit tests")."""
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


# ------------------------------------------------------------------------------------------------ splice
def _splice_ref(ids, labels, embed, scale, pad_id, n_img):
    """ref:Stage1/projector_trainer.py:186-220 + HF:models/gemma3/modeling_gemma3.py:102-113, HF:loss/loss_utils.py:57-60."""
    B, T = ids.shape
    S = n_img + T
    e = (embed[ids] * torch.tensor(scale, dtype=torch.bfloat16, device=ids.device)).float()    # bf16 product, as HF
    text_mask = (ids != pad_id).long() if pad_id is not None else torch.ones_like(ids)
    attn = torch.cat([torch.ones(B, n_img, dtype=torch.int64, device=ids.device), text_mask], 1)
    lm = torch.cat([torch.full((B, n_img), -100, dtype=torch.int64, device=ids.device), labels], 1)
    tgt = torch.cat([lm[:, 1:], torch.full((B, 1), -100, dtype=torch.int64, device=ids.device)], 1).int()
    return e, attn, lm, tgt


@pytest.mark.parametrize("case", ["no_pad", "left_pad", "all_but_one_pad", "pad_id_none", "right_pad"])
@pytest.mark.parametrize("B,T,n_img,h,V", [(3, 12, 7, 64, 512), (2, 128, 575, 1152, 4096), (1, 1, 0, 8, 16)])
def test_embed_splice_bit_exact(cuda_dev, case, B, T, n_img, h, V):
    from projectiontrainer_b200 import ops
    torch.manual_seed(T + h)
    pad = 0
    embed = torch.randn(V, h, device=cuda_dev).to(torch.bfloat16)
    ids = torch.randint(3, V, (B, T), device=cuda_dev)
    if case == "left_pad":
        for b in range(B):
            ids[b, :min(T - 1, 2 * b + 1)] = pad
    elif case == "right_pad":
        for b in range(B):
            ids[b, T - min(T - 1, b + 1):] = pad
    elif case == "all_but_one_pad":
        ids[:, :T - 1] = pad
    pad_id = None if case == "pad_id_none" else pad
    labels = ids.clone()
    if pad_id is not None:
        labels[ids == pad_id] = -100                    # ref:Stage1/train_projection_stage1.py:108-112
    scale = float(torch.tensor(h ** 0.5, dtype=torch.bfloat16).float())
    S = n_img + T
    x0 = torch.full((B * S, h), 7.0, device=cuda_dev)
    attn = torch.full((B, S), -1, dtype=torch.int64, device=cuda_dev)
    lm = torch.full((B, S), -1, dtype=torch.int64, device=cuda_dev)
    kv = torch.full((B, S), 9, dtype=torch.uint8, device=cuda_dev)
    tgt = torch.full((B, S), -1, dtype=torch.int32, device=cuda_dev)
    err = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
    ops.embed_splice(ids, labels, embed, scale, pad_id, n_img, x0, attn, lm, kv, tgt, err=err)
    e_ref, attn_ref, lm_ref, tgt_ref = _splice_ref(ids, labels, embed, scale, pad_id, n_img)
    assert torch.equal(attn, attn_ref) and torch.equal(lm, lm_ref) and torch.equal(tgt, tgt_ref)      # INT: bit-exact
    assert torch.equal(kv.long(), attn_ref)
    x = x0.view(B, S, h)
    assert torch.equal(x[:, n_img:], e_ref)                                   # exact copy of the producer's values
    assert (x[:, :n_img] == 7.0).all()                                        # image rows belong to the projector
    assert int(err.item()) == 0


def test_embed_splice_flags_out_of_range_ids(cuda_dev):
    """nn.Embedding raises on an id outside the table; the kernel reads row 0 instead and raises the flag."""
    from projectiontrainer_b200 import ops
    V, h, B, T = 100, 64, 2, 5
    embed = torch.randn(V, h, device=cuda_dev).to(torch.bfloat16)
    ids = torch.randint(3, V, (B, T), device=cuda_dev)
    ids[1, 2] = V + 5
    ids[0, 0] = -3
    S = T
    bufs = (torch.zeros(B * S, h, device=cuda_dev), torch.zeros(B, S, dtype=torch.int64, device=cuda_dev),
            torch.zeros(B, S, dtype=torch.int64, device=cuda_dev), torch.zeros(B, S, dtype=torch.uint8, device=cuda_dev),
            torch.zeros(B, S, dtype=torch.int32, device=cuda_dev))
    err = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
    ops.embed_splice(ids, ids, embed, 1.0, None, 0, *bufs, err=err)
    assert int(err.item()) & ops.ERRFLAG_TOKEN_ID
    assert torch.isfinite(bufs[0]).all()


# ------------------------------------------------------------------------------------------------ cross-entropy
@pytest.mark.parametrize("rows,V,frac_ignored", [(64, 1000, 0.3), (257, 4096, 0.0), (16, 262208, 0.5), (128, 262144, 0.1),
                                                 (5, 8, 0.4)])
def test_ce_fwd_bwd(cuda_dev, rows, V, frac_ignored):
    """loss / lse / dlogits of the shifted masked CE (HF:loss/loss_utils.py:28-67) vs torch on the same bf16 logits;
    V = 262 208 (Gemma3-4B) is a multiple of 8 but of nothing larger than 64."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + V)
    ld = (V + 7) // 8 * 8 + 8                       # padded row stride
    buf = torch.zeros(rows, ld, device=cuda_dev, dtype=torch.bfloat16)
    logits = buf[:, :V]
    logits.copy_((torch.randn(rows, V, device=cuda_dev) * 3).to(torch.bfloat16))
    tgt = torch.randint(0, V, (rows,), device=cuda_dev, dtype=torch.int32)
    tgt[torch.rand(rows, device=cuda_dev) < frac_ignored] = -100
    tgt[0] = V - 1
    ref_in = logits.float().clone().requires_grad_(True)
    gscale = 0.5
    ref_loss = torch.nn.functional.cross_entropy(ref_in, tgt.long(), ignore_index=-100)
    (ref_loss * gscale).backward()
    lse = torch.empty(rows, device=cuda_dev)
    acc2, loss = torch.empty(2, device=cuda_dev), torch.empty(1, device=cuda_dev)
    err = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
    ops.ce_fwd_bwd(logits, tgt, lse, acc2, loss, gscale, err=err)
    valid = tgt >= 0
    assert abs(loss.item() - ref_loss.item()) < 2e-5 * abs(ref_loss.item())
    assert int(acc2[1].item()) == int(valid.sum())
    lse_ref = torch.logsumexp(ref_in.detach(), dim=-1)
    assert (lse[valid] - lse_ref[valid]).abs().max().item() < 1e-4
    got = logits.float()                             # overwritten in place by dlogits
    assert (got[~valid] == 0).all()
    d = (got - ref_in.grad).abs().max().item()
    assert d <= 2.0 ** -8 * ref_in.grad.abs().max().item() + 1e-9, d
    assert (buf[:, V:] == 0).all()                   # the pad columns of the row stride are not touched
    assert int(err.item()) == 0


def test_ce_all_rows_ignored_is_nan_and_bad_target_is_flagged(cuda_dev):
    from projectiontrainer_b200 import ops
    rows, V = 8, 64
    logits = torch.randn(rows, V, device=cuda_dev).to(torch.bfloat16)
    tgt = torch.full((rows,), -100, dtype=torch.int32, device=cuda_dev)
    lse, acc2, loss = torch.empty(rows, device=cuda_dev), torch.empty(2, device=cuda_dev), torch.empty(1, device=cuda_dev)
    err = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
    ops.ce_fwd_bwd(logits, tgt, lse, acc2, loss, 1.0, err=err)
    assert math.isnan(loss.item())                   # HF: mean over an empty set
    assert (logits.float() == 0).all() and int(err.item()) == 0
    logits = torch.randn(rows, V, device=cuda_dev).to(torch.bfloat16)
    tgt[3], tgt[4] = 5, V + 1
    ops.ce_fwd_bwd(logits, tgt, lse, acc2, loss, 1.0, err=err)
    assert int(err.item()) & ops.ERRFLAG_TARGET and int(acc2[1].item()) == 1 and math.isfinite(loss.item())


# ------------------------------------------------------------------------------------------------ optimizer
@pytest.mark.parametrize("n,clip_active", [(22293632 // 64, True), (4099, False), (8, True)])
def test_clip_adamw_matches_torch(cuda_dev, n, clip_active):
    """clip_grad_norm_(5.0) + AdamW over three steps on a flat buffer (ref:Stage1/projector_trainer.py:75-79, 241-243)."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(n)
    p0 = torch.randn(n, device=cuda_dev)
    ref_p = torch.nn.Parameter(p0.clone())
    opt = torch.optim.AdamW([ref_p], lr=1e-3, weight_decay=0.01)
    p, m, v = p0.clone(), torch.zeros(n, device=cuda_dev), torch.zeros(n, device=cuda_dev)
    shadow = torch.empty(n, device=cuda_dev, dtype=torch.bfloat16)
    ss = torch.zeros(1, device=cuda_dev)
    world = 2
    for step in range(1, 4):
        g = torch.randn(n, device=cuda_dev) * (20.0 if clip_active else 1e-3)
        lr = 1e-3 * (0.5 + 0.1 * step)
        for grp in opt.param_groups:
            grp["lr"] = lr
        ref_p.grad = g.clone()
        norm = torch.nn.utils.clip_grad_norm_([ref_p], 5.0)
        assert (norm > 5.0) == clip_active
        opt.step()
        # the engine hands the kernel the all-reduced SUM over ranks and folds 1/world in (grad_scale)
        gs = g * world
        ops.sumsq(gs, ss)
        assert abs(ss.item() - float((gs.double() ** 2).sum())) < 1e-4 * float((gs.double() ** 2).sum())
        ops.clip_adamw(p, gs, m, v, shadow, ss, grad_scale=1.0 / world, max_norm=5.0, lr=lr, beta1=0.9, beta2=0.999,
                       eps=1e-8, wd=0.01, step=step)
        assert (p - ref_p.detach()).abs().max().item() < 2e-6 * max(1.0, p0.abs().max().item())
        assert torch.equal(shadow, p.to(torch.bfloat16))


# ------------------------------------------------------------------------------------------------ small kernels
@pytest.mark.parametrize("M,N,ld_pad", [(18400, 1152, 0), (575, 10240, 0), (703, 64, 8), (1, 2, 0)])
def test_colsum(cuda_dev, M, N, ld_pad):
    from projectiontrainer_b200 import ops
    torch.manual_seed(M + N)
    buf = torch.randn(M, N + ld_pad, device=cuda_dev).to(torch.bfloat16)
    x = buf[:, :N]
    out = torch.full((N,), float("nan"), device=cuda_dev)
    ops.colsum(x, out)
    ref = x.double().sum(0)
    assert (out.double() - ref).abs().max().item() < 1e-4 * math.sqrt(M) + 1e-5


@pytest.mark.parametrize("slices,elems", [(8, 4096 * 1152), (3, 12), (1, 4)])
def test_sum_slices(cuda_dev, slices, elems):
    from projectiontrainer_b200 import _lib, ops
    import ctypes as C
    torch.manual_seed(slices)
    x = torch.randn(slices, elems, device=cuda_dev)
    out = torch.full((elems,), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    _lib.call("pt_sum_slices", ops._ptr(x), C.c_longlong(elems), slices, ops._ptr(out), ops._stream())
    ref = x.sum(0)
    assert (out.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item() + 1e-6


def test_gather_cast_rows(cuda_dev):
    """The image-token rows of dL/dX0 (rows 0..574 of every sample) as bf16: the projector backward's input."""
    from projectiontrainer_b200 import ops
    B, S, n_img, D = 3, 703, 575, 1152
    x = torch.randn(B * S, D, device=cuda_dev)
    out = torch.full((B * n_img, D), float("nan"), device=cuda_dev, dtype=torch.bfloat16)
    ops.gather_cast_rows(x, out, group=(n_img, S, 0))
    assert torch.equal(out, x.view(B, S, D)[:, :n_img].reshape(-1, D).to(torch.bfloat16))


@pytest.mark.parametrize("n", [18400 * 10240 // 16, 4096, 8])
def test_gelu_erf_bwd(cuda_dev, n):
    """dZ = dH * gelu_erf'(Z): backward of nn.GELU() (ref:Stage1/projectors.py:18) vs autograd."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(n)
    z = (torch.randn(n, device=cuda_dev) * 2).to(torch.bfloat16)
    dh = torch.randn(n, device=cuda_dev).to(torch.bfloat16)
    zr = z.float().requires_grad_(True)
    torch.nn.functional.gelu(zr).backward(dh.float())
    dz = torch.empty_like(dh)
    ops.gelu_erf_bwd(dh, z, dz)
    assert (dz.float() - zr.grad).abs().max().item() <= 2.0 ** -7 * zr.grad.abs().max().item() + 1e-6
    ops.gelu_erf_bwd(dh, z, dh)                       # in place, as the engine calls it
    assert torch.equal(dh, dz)


def test_cast_and_patchify(cuda_dev):
    from projectiontrainer_b200 import ops
    x = torch.randn(1000003, device=cuda_dev)
    y = torch.empty(x.numel(), device=cuda_dev, dtype=torch.bfloat16)
    ops.cast_f32_bf16(x, y)
    assert torch.equal(y, x.to(torch.bfloat16))
    # im2col of Conv2d(3, d, 16, stride 16): column order (c, py, px) = conv.weight.flatten(1)
    B, H, W, p = 2, 64, 96, 16
    px = torch.randn(B, 3, H, W, device=cuda_dev)
    cols = torch.empty(B * (H // p) * (W // p), 3 * p * p, device=cuda_dev, dtype=torch.bfloat16)
    ops.patchify(px, p, cols)
    ref = torch.nn.functional.unfold(px, kernel_size=p, stride=p).transpose(1, 2).reshape(-1, 3 * p * p)
    assert torch.equal(cols, ref.to(torch.bfloat16))


# ------------------------------------------------------------------------------------------------ CE statistics in the LM-head epilogue
@pytest.mark.parametrize("rows,V,K", [(64, 1024, 128), (300, 304, 64), (257, 1312, 192), (128, 262208, 64),
                                      (4096, 4096, 1152), (96, 151936, 64)])
def test_lm_head_gemm_with_ce_statistics_epilogue(cuda_dev, rows, V, K):
    """PT_ACT_CE_STATS + pt_ce_from_partials == the logits GEMM followed by pt_ce_fwd_bwd: same logits, same lse,
    loss and dlogits; against torch's cross-entropy on the stored bf16 logits.  Vocabularies that are not multiples of
    the 256-wide tile (shifted last tile: 1312, 262208 = Gemma3-4B, 151936 = Qwen3; columns past the end: 304) and
    row counts that are not multiples of the 256-row tile."""
    from projectiontrainer_b200 import ops
    torch.manual_seed(rows + V)
    x = (torch.randn(rows, K, device=cuda_dev) * 1.5).to(torch.bfloat16)
    w = (torch.randn(V, K, device=cuda_dev) * (3.0 / K ** 0.5)).to(torch.bfloat16)
    tgt = torch.randint(0, V, (rows,), device=cuda_dev, dtype=torch.int32)
    tgt[torch.rand(rows, device=cuda_dev) < 0.25] = -100
    tgt[0] = V - 1
    ld = (V + 7) // 8 * 8 + 8
    mk = lambda: torch.zeros(rows, ld, device=cuda_dev, dtype=torch.bfloat16)[:, :V]
    # reference path: plain GEMM, then the statistics pass over the stored logits
    lg_a = ops.linear(x, w, out=mk())
    stored = lg_a.clone()
    lse_a, acc_a, loss_a = torch.empty(rows, device=cuda_dev), torch.empty(2, device=cuda_dev), torch.empty(1, device=cuda_dev)
    ops.ce_fwd_bwd(lg_a, tgt, lse_a, acc_a, loss_a, 0.5)
    # fused path
    stats = torch.full((ops.ce_stats_parts(V), rows, 2), float("nan"), device=cuda_dev)
    lg_b = ops.linear_ce_stats(x, w, mk(), stats)
    assert torch.equal(lg_b, stored)                              # the epilogue stores the same logits
    lse_b, acc_b, loss_b = torch.empty(rows, device=cuda_dev), torch.empty(2, device=cuda_dev), torch.empty(1, device=cuda_dev)
    err = torch.zeros(1, dtype=torch.int32, device=cuda_dev)
    ops.ce_from_partials(lg_b, tgt, stats, lse_b, acc_b, loss_b, 0.5, err=err)
    valid = tgt >= 0
    ref = stored.float()
    lse_ref = torch.logsumexp(ref, dim=-1)
    assert (lse_b[valid] - lse_ref[valid]).abs().max().item() < 1e-4
    assert (lse_b[valid] - lse_a[valid]).abs().max().item() < 1e-4
    ref_loss = torch.nn.functional.cross_entropy(ref, tgt.long(), ignore_index=-100)
    assert abs(loss_b.item() - ref_loss.item()) < 2e-5 * abs(ref_loss.item())
    assert int(acc_b[1].item()) == int(valid.sum()) and int(err.item()) == 0
    d = (lg_b.float() - lg_a.float()).abs().max().item()
    assert d <= 2.0 ** -8 * lg_a.float().abs().max().item() + 1e-12, d          # dlogits of both paths
    assert (lg_b.float()[~valid] == 0).all()
    assert not torch.isnan(stats).any()                           # every (part, row) pair was written
