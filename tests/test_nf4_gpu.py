"""NF4 frozen-weight kernels through the C ABI vs oracle/nf4_oracle.py: codes, statistics and expanded bf16 weights
bit for bit; the Stage-1 step on NF4-stored LLM weights vs the fp32 oracle on the same weights (GPU only)."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _weights(kind, shape, seed):
    g = torch.Generator().manual_seed(seed)
    w = torch.randn(shape, generator=g) * 0.03
    if kind == "outliers":
        w[::7, ::13] *= 40.0
    elif kind == "tiny":                     # below fp16's normal range: the fp16 detour loses bits / flushes
        w = w * 1e-6
    elif kind == "subnormal":                # absmax below fp32's normal range: 1 / absmax overflows to +inf
        w = w * 1e-38
    elif kind == "zero_blocks":
        w.view(-1, 64)[::3] = 0.0
    elif kind == "constant":                 # every block has the same absmax: the second level sees all zeros
        w = torch.full(shape, 0.0123)
    return w.to(torch.bfloat16)


@pytest.mark.parametrize("via_fp16", [True, False], ids=["via_fp16", "direct"])
@pytest.mark.parametrize("kind,shape", [("gauss", (1152, 256)), ("gauss", (6912, 1152)), ("outliers", (1024, 1152)),
                                        ("tiny", (256, 128)), ("subnormal", (128, 128)), ("zero_blocks", (512, 192)), ("constant", (64, 64)),
                                        ("gauss", (1, 64)), ("gauss", (257, 64 * 3))])
def test_nf4_quantize_dequantize_bit_exact_vs_oracle(cuda_dev, kind, shape, via_fp16):
    from oracle import nf4_oracle as N
    from projectiontrainer_b200 import ops
    w = _weights(kind, shape, seed=shape[0] + shape[1])
    st = ops.nf4_quantize(w.to(cuda_dev), via_fp16)
    out = ops.nf4_dequantize(st)
    torch.cuda.synchronize()
    assert np.array_equal(st["code256"].cpu().numpy(), N.dynamic_map())
    src = w.float()
    src = src.to(torch.float16).float() if via_fp16 else src
    packed, absmax = N.quantize_nf4(src.numpy().reshape(-1))
    assert np.array_equal(st["absmax"].cpu().numpy(), absmax)
    assert np.array_equal(st["packed"].cpu().numpy(), packed)
    assert abs(st["offset"] - float(absmax.mean())) <= 1e-6 * abs(float(absmax.mean())) + 1e-30
    q, absmax2, _ = N.double_quant(absmax, offset=st["offset"])          # same fp32 offset: summation order differs
    assert np.array_equal(st["absmax2"].cpu().numpy(), absmax2)
    assert np.array_equal(st["qabsmax"].cpu().numpy(), q)
    ref = N.dequantize_nf4(packed, q, absmax2, st["offset"], "fp16" if via_fp16 else "bf16")
    ref = torch.from_numpy(ref).to(torch.bfloat16)
    assert out.shape == w.shape and torch.equal(out.cpu().view(torch.int16), ref.reshape(shape).view(torch.int16))


def test_nf4_rejects_bad_inputs(cuda_dev):
    from projectiontrainer_b200 import ops
    with pytest.raises(RuntimeError, match="multiple of 64"):
        ops.nf4_quantize(torch.zeros(3, 33, dtype=torch.bfloat16, device=cuda_dev))
    with pytest.raises(RuntimeError, match="bf16"):
        ops.nf4_quantize(torch.zeros(64, dtype=torch.float32, device=cuda_dev))
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.nf4_quantize(torch.zeros(64, dtype=torch.bfloat16))


@pytest.mark.parametrize("recipe", ["gemma3", "qwen3"])
def test_stage1_step_on_nf4_frozen_llm_vs_oracle(cuda_dev, recipe):
    """`--enable_qlora` in Stage 1 (ref:Stage1/train_projection_stage1.py:192-210): the LLM stays frozen, its decoder
    Linear weights are NF4.  Engine on `nf4_frozen_llm_` weights vs the fp32 oracle holding the oracle's own NF4
    round trip of the same weights, quantised per HF module (q, k, v, o, gate, up, down separately)."""
    from oracle import nf4_oracle as N
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200 import weights as Wt
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    vkw = dict(hidden=128, ff=256, layers=2, heads=2, image=160, patch=16)
    if recipe == "qwen3":
        lkw = dict(vocab=1024, hidden=256, ff=384, layers=4, hq=8, hkv=2, hd=128, window=0, q_scalar=128.0, pattern=1,
                   sandwich_norm=False, silu=True, scale_embed=False, tie_lm_head=False)
    else:
        lkw = dict(vocab=1024, hidden=256, ff=384, layers=6, hq=4, hkv=1, hd=256, window=16, q_scalar=256.0)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    lc = GemmaCfg(**lkw)
    Wg = {k: v.clone() for k, v in Wl.items()}
    report = Wt.nf4_frozen_llm_(Wg, lc, cuda_dev)
    assert len(report) == 4 * lc.layers and all(0.05 < e < 0.15 for e in report.values()), report
    # the oracle's copy: per-module round trip of the bf16 weights, offsets taken from the device run's statistics
    q_rows, kv_rows = lc.hq * lc.hd, lc.hkv * lc.hd
    bq = lambda t: t.to(torch.bfloat16).float()
    Wv_q = {k: (bq(v) if v.dim() == 2 and k != "pos" else v) for k, v in Wv.items()}
    Wl_q = {k: (bq(v) if v.dim() == 2 else v) for k, v in Wl.items()}      # what the engine holds without NF4
    Wo = dict(Wl_q)
    for l in range(lc.layers):
        for name, splits in ((f"l{l}.qkv_w", (q_rows, kv_rows, kv_rows)), (f"l{l}.o_w", None),
                             (f"l{l}.gu_w", (lc.ff, lc.ff)), (f"l{l}.down_w", None)):
            w = Wl[name].to(torch.bfloat16).float()
            parts, r0 = [], 0
            for rows in (splits or (w.shape[0],)):
                parts.append(torch.from_numpy(N.roundtrip(w[r0:r0 + rows].numpy())))
                r0 += rows
            Wo[name] = torch.cat(parts, 0)
            mism = (Wo[name] != Wg[name].float().cpu()).float().mean().item()
            assert mism < 2e-3, (name, mism)          # identical up to the fp32 mean's summation order
            Wo[name] = Wg[name].float().cpu()         # and then exactly the engine's weights for the step check
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wg, lc, cuda_dev),
                         E.ProjectorState(P, cuda_dev), cuda_dev)
    px, ids, lab = O.synthetic_batch(3, 24, ovc, olc, seed=77, left_pad=True)
    ref = O.stage1_step(px, ids, lab, Wv_q, P, Wo, ovc, olc, scale_dtype=torch.bfloat16)
    loss = eng.forward_backward(px.to(cuda_dev), ids.to(cuda_dev), lab.to(cuda_dev), pad_id=olc.pad_id)
    assert abs(loss.item() - ref["loss"]) / ref["loss"] < 1e-2
    for k, gr in eng.proj.grads().items():
        cos = torch.nn.functional.cosine_similarity(gr.cpu().flatten().float(), ref["grads"][k].flatten(), dim=0)
        assert cos.item() > 0.999, (k, cos.item())
    # and the quantisation really changed the model: the un-quantised oracle gives a different loss
    base = O.stage1_step(px, ids, lab, Wv_q, P, Wl_q, ovc, olc, scale_dtype=torch.bfloat16)
    assert abs(base["loss"] - ref["loss"]) > 1e-4
