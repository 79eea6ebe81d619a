"""Parity of the CUDA Stage-1 step with the reference (GPU only).  Two checkers: the committed golden
vectors recorded from the UNMODIFIED reference trainer (tests/golden, fp32 HF on CPU) and the CPU oracle
on the same seeded inputs.  Tolerances are north_star's for the bf16 path: splice / masks / labels
bit-exact, loss within 1e-2 relative, projector-gradient cosine >= 0.999."""
import math
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float(torch.dot(a, b) / (a.norm() * b.norm()))


def _engine(case, dev):
    from oracle import stage1_oracle as O
    from oracle.make_golden import CASES
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    vkw, lkw, B, T, left_pad, accum, epochs, warm = CASES[case]
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    px, ids, lab = O.synthetic_batch(B, T, ovc, olc, left_pad=bool(left_pad))
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), dev), E.PackedGemma(Wl, GemmaCfg(**lkw), dev),
                         E.ProjectorState(P, dev), dev, grad_accum=accum)
    return eng, (px, ids, lab), olc, np.load(GOLDEN / f"{case}.npz")


@pytest.mark.parametrize("case", ["tiny_pad_gqa", "tiny_nopad_mqa", "tiny_qwen3_gqa"])
def test_step_matches_reference_golden_tiny(cuda_dev, case):
    eng, (px, ids, lab), lc, g = _engine(case, cuda_dev)
    perm = torch.from_numpy(g["perms"][0])
    px, ids, lab = px[perm].to(cuda_dev), ids[perm].to(cuda_dev), lab[perm].to(cuda_dev)
    B, T = ids.shape
    S = eng.n_img + T
    patch = eng.vision_forward(px)
    x0 = eng.buf("x0", (B * S, lc.hidden), torch.float32)
    Z, Hm = eng.projector_forward(patch, x0, rows_per_img=eng.n_img, out_batch_stride=S * lc.hidden)
    am, ll, kv, tg = eng.splice(ids, lab, x0, lc.pad_id)
    torch.cuda.synchronize()
    # integer work: bit-exact against what the reference handed to the language model
    assert am.dtype == torch.int64 and ll.dtype == torch.int64
    assert np.array_equal(am.cpu().numpy(), g["attention_mask"])
    assert np.array_equal(ll.cpu().numpy(), g["lm_labels"])
    assert _cos(patch.float().cpu(), torch.from_numpy(g["patch"])) > 0.999
    assert _cos(x0.cpu(), torch.from_numpy(g["spliced"])) > 0.9995
    loss = eng.llm_forward_loss(x0, kv, tg, B, S, T)
    dx = eng.llm_backward()
    eng.projector_backward(dx, patch, Z, Hm, B, S)
    torch.cuda.synchronize()
    ref_loss = float(g["step0.loss"])
    assert abs(loss.item() - ref_loss) / ref_loss < 1e-2            # north_star bf16 tolerance
    for k, gr in eng.proj.grads().items():
        assert _cos(gr.cpu(), torch.from_numpy(g[f"step0.grad.{k}"])) > 0.999, k


@pytest.mark.parametrize("case", ["config1_full", "config3_b2_full"])
def test_step_matches_reference_golden_full_size(cuda_dev, case):
    """Full model size (SigLIP-L/16-384 + MLP + Gemma3-1B, left pads) against the reference's own fp32 CPU run
    (oracle/make_golden.py --full): BASELINE.json configs[0] (B=2, T=64, S=639) and the BENCHMARKED geometry of
    configs[2] (T=128, S=703 > window 512) at B=2; we run bf16 on the GPU."""
    eng, (px, ids, lab), lc, g = _engine(case, cuda_dev)
    perm = torch.from_numpy(g["perms"][0])
    T = ids.shape[1]
    loss = eng.forward_backward(px[perm].to(cuda_dev), ids[perm].to(cuda_dev), lab[perm].to(cuda_dev), lc.pad_id)
    torch.cuda.synchronize()
    ref_loss = float(g["step0.loss"])
    assert abs(loss.item() - ref_loss) / ref_loss < 1e-2
    am = eng.buf("s_mask", (2, eng.n_img + T), torch.int64)
    ll = eng.buf("s_labels", (2, eng.n_img + T), torch.int64)
    assert np.array_equal(am.cpu().numpy(), g["attention_mask"]) and np.array_equal(ll.cpu().numpy(), g["lm_labels"])
    for k, gr in eng.proj.grads().items():
        key = f"step0.grad.{k}"
        flat = gr.cpu().flatten()
        if key in g:
            ref, ours = torch.from_numpy(g[key]).flatten(), flat
        else:                                   # large tensors are stored as a 262144-entry random sample
            ref, ours = torch.from_numpy(g[key + "__sample"]), flat[torch.from_numpy(g[key + "__idx"])]
            assert abs(float(flat.double().norm()) - float(g[key + "__l2"])) / float(g[key + "__l2"]) < 2e-2, k
        assert _cos(ours, ref) > 0.999, k


def test_trainer_accepts_hf_modules_front_door(cuda_dev, tmp_path):
    """what `train_projection_stage1.py:178-252, 338-356` hands the trainer is an HF `SiglipModel` (or
    `SiglipVisionModel`) and an HF `Gemma3ForCausalLM`.  Passed as weight containers they must (1) pack to exactly
    the buffers the pre-packed path holds and (2) train to the same losses and learning rates."""
    from oracle import stage1_oracle as O
    from oracle.make_golden import CASES, build_hf_gemma, build_hf_siglip
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.accelerator_setup import Stage1Accelerator
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200.projector_trainer import ProjectionTrainerStage1
    from projectiontrainer_b200.projectors import MLPProjector
    case = "tiny_pad_gqa"
    vkw, lkw, B, T, left_pad, accum, epochs, warm = CASES[case]
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    px, ids, lab = O.synthetic_batch(B, T, ovc, olc, left_pad=True)

    class DS(torch.utils.data.Dataset):
        def __len__(self): return B
        def __getitem__(self, i): return {"pixel_values": px[i], "token_ids": ids[i], "labels": lab[i]}

    class Tok:
        pad_token_id, eos_token_id = olc.pad_id, 1

    def run(vision, llm, out):
        proj = MLPProjector(ovc.hidden, olc.hidden)
        proj.load_state_dict(P)
        acc = Stage1Accelerator(gradient_accumulation_steps=1, device=cuda_dev)
        tr = ProjectionTrainerStage1(acc, vision, llm, proj, None, Tok(), DS(), output_dir=str(out), batch_size=B,
                                     learning_rate=1e-3, weight_decay=0.01, num_epochs=3,
                                     gradient_accumulation_steps=1, warmup_ratio=0.34, save_every_n_epochs=0)
        tr.train()
        steps = [v for s, v in acc.logged if "train/batch_loss" in v]
        return tr, [(v["train/batch_loss"], v["train/learning_rate"]) for v in steps], proj

    hf_vision, hf_llm = build_hf_siglip(ovc, Wv), build_hf_gemma(olc, Wl)          # HF SiglipVisionModel / Gemma3ForCausalLM
    tr_hf, traj_hf, proj_hf = run(hf_vision, hf_llm, tmp_path / "hf")
    tr_pk, traj_pk, proj_pk = run(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev),
                                  E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev), tmp_path / "packed")
    # (1) the packed frozen weights are identical, tensor by tensor
    a, b = tr_hf.engine, tr_pk.engine
    assert a.vcfg == b.vcfg and a.lcfg == b.lcfg
    for name in ("patch_w", "patch_b", "pos", "post_g", "post_b"):
        assert torch.equal(getattr(a.vis, name), getattr(b.vis, name)), name
    for la, lb in zip(a.vis.layers, b.vis.layers):
        assert all(torch.equal(la[k], lb[k]) for k in la)
    assert torch.equal(a.llm.embed, b.llm.embed) and torch.equal(a.llm.final_norm, b.llm.final_norm)
    assert a.llm.embed_scale == b.llm.embed_scale and a.llm.lm_head.data_ptr() == a.llm.embed.data_ptr()
    for la, lb in zip(a.llm.layers, b.llm.layers):
        assert all((la[k] is None and lb[k] is None) or torch.equal(la[k], lb[k]) for k in la)
    # (2) same trajectory up to the summation order of the atomics in the loss / bias-gradient reductions
    assert len(traj_hf) == 3
    for (l1, lr1), (l2, lr2) in zip(traj_hf, traj_pk):
        assert lr1 == lr2 and abs(l1 - l2) < 1e-5 * abs(l2), (traj_hf, traj_pk)
    for k, v in proj_hf.state_dict().items():
        d1, d2 = v.cpu().double() - P[k].double(), proj_pk.state_dict()[k].cpu().double() - P[k].double()
        assert float((d1 - d2).norm() / d2.norm()) < 2e-2, k        # an entry with a ~0 gradient may flip its Adam sign
    # and it is the reference's trajectory (recorded from its own trainer on the same HF modules)
    g = np.load(GOLDEN / f"{case}.npz")
    # (the golden run of this case used grad_accum 2 / 2 epochs / lr 1e-3 / no warm-up: only step 0's loss is comparable)
    assert abs(traj_hf[0][0] - float(g["step0.loss"])) / float(g["step0.loss"]) < 1e-2
    # a SiglipModel-like container (`.vision_model` attribute, train_projection_stage1.py:247) is accepted too
    class SiglipModelLike(torch.nn.Module):
        def __init__(self, vm):
            super().__init__()
            self.vision_model = vm
    from projectiontrainer_b200 import weights as Wt
    W2, c2 = Wt.siglip_from_hf(SiglipModelLike(hf_vision.vision_model).vision_model)
    assert c2 == a.vcfg and torch.equal(W2["l0.qkv_w"].to(torch.bfloat16).to(cuda_dev), a.vis.layers[0]["qkv_w"])
    # unsupported containers are rejected loudly, not silently mis-packed
    with pytest.raises(Exception):
        ProjectionTrainerStage1(Stage1Accelerator(device=cuda_dev), torch.nn.Linear(4, 4), hf_llm,
                                MLPProjector(ovc.hidden, olc.hidden), None, Tok(), DS(), output_dir=str(tmp_path / "x"))


def test_trainer_dropin_trajectory_tiny(cuda_dev, tmp_path):
    """The drop-in `ProjectionTrainerStage1` + `MLPProjector` reproduce the reference trainer's recorded
    run (losses per step, schedule, artefacts) on the tiny fixture."""
    import json
    from oracle import stage1_oracle as O
    from oracle.make_golden import CASES
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.accelerator_setup import Stage1Accelerator
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200.projector_trainer import ProjectionTrainerStage1
    from projectiontrainer_b200.projectors import MLPProjector
    case = "tiny_nopad_mqa"
    vkw, lkw, B, T, left_pad, accum, epochs, warm = CASES[case]
    g = np.load(GOLDEN / f"{case}.npz")
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    px, ids, lab = O.synthetic_batch(B, T, ovc, olc, left_pad=bool(left_pad))
    proj = MLPProjector(ovc.hidden, olc.hidden)
    proj.load_state_dict(P)

    class DS(torch.utils.data.Dataset):
        def __len__(self): return B
        def __getitem__(self, i): return {"pixel_values": px[i], "token_ids": ids[i], "labels": lab[i]}

    class Tok:
        pad_token_id, eos_token_id = olc.pad_id, 1

    acc = Stage1Accelerator(gradient_accumulation_steps=accum, device=cuda_dev)
    tr = ProjectionTrainerStage1(acc, E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev),
                                 E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev), proj, None, Tok(), DS(),
                                 output_dir=str(tmp_path), batch_size=B, learning_rate=float(g["base_lr"]),
                                 weight_decay=float(g["weight_decay"]), num_epochs=epochs,
                                 gradient_accumulation_steps=accum, warmup_ratio=warm, save_every_n_epochs=0)
    assert tr.max_train_steps == epochs * math.ceil(1 / accum)
    tr.train()
    steps = [v for s, v in acc.logged if "train/batch_loss" in v]
    assert len(steps) == epochs
    for i, v in enumerate(steps):
        ref = float(g[f"step{i}.loss"])
        assert abs(v["train/batch_loss"] - ref) / ref < 1e-2, i
        assert abs(v["train/learning_rate"] - float(g[f"step{i}.lr_after"])) < 1e-12, i
    sd = torch.load(tmp_path / f"projector_epoch_{epochs}.bin")
    assert list(sd.keys()) == list(E.ProjectorState.KEYS) and sd["model.0.weight"].dtype == torch.float32
    assert json.loads((tmp_path / "projector_config.json").read_text()) == {"vision_dim": ovc.hidden, "llm_dim": olc.hidden}
    # the nn.Module the caller holds sees the trained values (shared storage) and they moved like the reference's
    for k, p in proj.state_dict().items():
        assert torch.equal(p.cpu(), sd[k])
        ours, ref = p.cpu() - P[k], torch.from_numpy(g[f"step{epochs - 1}.param.{k}"]) - P[k]
        # Adam moves every entry by ~lr * sign-like(g): an entry whose gradient is ~0 can flip under bf16
        # noise (see tests/test_oracle_golden.py), so the bound on the small bias vectors is loose
        assert _cos(ours, ref) > (0.98 if ours.numel() > 1000 else 0.9), k


def test_mlp_projector_module_autograd(cuda_dev):
    from projectiontrainer_b200.projectors import MLPProjector
    torch.manual_seed(0)
    m = MLPProjector(64, 96).to(cuda_dev)
    x = torch.randn(3, 37, 64, device=cuda_dev)
    y = m(x)
    w = torch.randn_like(y)
    (y * w).sum().backward()
    ref = MLPProjector(64, 96).to(cuda_dev)
    ref.load_state_dict(m.state_dict())
    yr = ref.model(x.to(torch.bfloat16).float())
    (yr * w).sum().backward()
    assert y.dtype == torch.float32 and _cos(y, yr) > 0.9999
    for (k, p), (_, q) in zip(m.named_parameters(), ref.named_parameters()):
        assert _cos(p.grad, q.grad) > 0.999, k


@pytest.mark.parametrize("fused", [True, False], ids=["fused_epilogues", "unfused_kernels"])
def test_step_hd256_gqa_window_pads_vs_oracle(cuda_dev, fused, monkeypatch):
    """The fused paths that only engage at head_dim 256 / ff % 128 == 0 (flash-style Gemma attention forward,
    softmax backward in the dO V^T epilogue, GeGLU epilogues, attention bands) on a reduced-width decoder with the
    4B geometry class (GQA with 2 kv heads, sliding window shorter than the sequence, linear-scaled global RoPE,
    left pads), checked against the CPU oracle holding the same bf16-rounded matrices; the unfused kernels are
    run on the same inputs as a second witness."""
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    if not fused:
        for k in ("PT_STAGE1_FUSE_ATTN_FWD", "PT_STAGE1_FUSE_SOFTMAX_BWD", "PT_STAGE1_FUSE_GEGLU"):
            monkeypatch.setenv(k, "0")
    vkw = dict(hidden=128, ff=256, layers=2, heads=2, image=160, patch=16)          # 100 patches -> 99 image tokens
    lkw = dict(vocab=1024, hidden=256, ff=256, layers=7, hq=4, hkv=2, hd=256, window=40, q_scalar=256.0,
               rope_factor_global=8.0)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    px, ids, lab = O.synthetic_batch(3, 48, ovc, olc, left_pad=True)
    q = lambda t: t.to(torch.bfloat16).float()
    Wv_q = {k: (q(v) if v.dim() == 2 and k != "pos" else v) for k, v in Wv.items()}
    Wl_q = {k: (q(v) if v.dim() == 2 else v) for k, v in Wl.items()}
    ref = O.stage1_step(px, ids, lab, Wv_q, P, Wl_q, ovc, olc, scale_dtype=torch.bfloat16)
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev),
                         E.ProjectorState(P, cuda_dev), cuda_dev)
    assert eng.fuse_attn_fwd == fused and eng.fuse_geglu == fused and eng.fuse_softmax_bwd == fused
    loss = eng.forward_backward(px.to(cuda_dev), ids.to(cuda_dev), lab.to(cuda_dev), olc.pad_id)
    torch.cuda.synchronize()
    assert abs(loss.item() - ref["loss"]) / ref["loss"] < 1e-2
    for k, gr in eng.proj.grads().items():
        assert _cos(gr.cpu(), ref["grads"][k]) > 0.999, k


def test_train_step_graph_replay_and_staged_host_batches_match_eager(cuda_dev):
    """`train_step` through the CUDA graph with host batches prefetched by `stage_batch` (the e2e path of
    bench.py and of the trainer's look-ahead loop) follows the same trajectory as eager launches on device
    batches: three optimizer steps on two alternating batches."""
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    vkw = dict(hidden=64, ff=128, layers=2, heads=1, image=64, patch=16)
    lkw = dict(vocab=512, hidden=128, ff=128, layers=6, hq=2, hkv=1, hd=256, window=8, q_scalar=256.0)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    batches = []
    for seed in (11, 12):
        px, ids, lab = O.synthetic_batch(2, 10, ovc, olc, seed=seed, left_pad=True)
        batches.append({"pixel_values": px, "token_ids": ids, "labels": lab})

    def run(use_graph, staged):
        eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev),
                             E.ProjectorState(P, cuda_dev), cuda_dev)
        eng.use_graph = use_graph
        losses = []
        for i in range(4):
            b = batches[i & 1]
            if staged:
                b = eng.stage_batch({k: v.pin_memory() for k, v in b.items()})
            else:
                b = {k: v.to(cuda_dev) for k, v in b.items()}
            losses.append(float(eng.train_step(b, 1e-3, pad_id=olc.pad_id).item()))
        torch.cuda.synchronize()
        return losses, eng.proj.flat.clone()

    l0, p0 = run(False, False)
    l1, p1 = run(True, True)
    assert np.allclose(l0, l1, rtol=1e-5, atol=1e-6), (l0, l1)
    assert l0[2] < l0[0]                                   # the same batch, two optimizer steps later
    assert _cos(p0.cpu(), p1.cpu()) > 1 - 1e-6


@pytest.mark.parametrize("fused", [True, False], ids=["fused_kernels", "unfused_kernels"])
def test_step_qwen3_recipe_hd128_vs_oracle(cuda_dev, fused, monkeypatch):
    """The Qwen3 decoder recipe (pre-norm only: residual adds in the o / down projection epilogues, SwiGLU fused
    into the gate||up and down-dgrad GEMMs, per-head q/k norms, unscaled embeddings, untied LM head, full
    attention, GQA 4:1, head_dim 128) on a reduced-width decoder with left pads, against the CPU oracle holding
    the same bf16-rounded matrices (the oracle's Qwen3 recipe is pinned to the reference trainer driving HF
    Qwen3ForCausalLM: tests/golden/tiny_qwen3_gqa.npz)."""
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    if not fused:
        for k in ("PT_STAGE1_FUSE_ATTN_FWD", "PT_STAGE1_FUSE_SOFTMAX_BWD", "PT_STAGE1_FUSE_GEGLU"):
            monkeypatch.setenv(k, "0")
    vkw = dict(hidden=128, ff=256, layers=2, heads=2, image=160, patch=16)
    lkw = dict(vocab=1024, hidden=256, ff=384, layers=4, hq=8, hkv=2, hd=128, window=0, q_scalar=128.0, pattern=1,
               sandwich_norm=False, silu=True, scale_embed=False, tie_lm_head=False)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    assert "lm_head" in Wl and "l0.post_attn_norm" not in Wl
    px, ids, lab = O.synthetic_batch(3, 40, ovc, olc, left_pad=True)
    q = lambda t: t.to(torch.bfloat16).float()
    Wv_q = {k: (q(v) if v.dim() == 2 and k != "pos" else v) for k, v in Wv.items()}
    Wl_q = {k: (q(v) if v.dim() == 2 else v) for k, v in Wl.items()}
    ref = O.stage1_step(px, ids, lab, Wv_q, P, Wl_q, ovc, olc, scale_dtype=torch.bfloat16)
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev),
                         E.ProjectorState(P, cuda_dev), cuda_dev)
    assert eng.fuse_geglu == fused and eng.fuse_attn_fwd == fused          # head_dim 128 instantiation of the fused attention
    for it in range(3):                                       # eager, graph capture, graph replay
        loss = eng._graphed_forward_backward({"pixel_values": px.to(cuda_dev), "token_ids": ids.to(cuda_dev),
                                              "labels": lab.to(cuda_dev)}, olc.pad_id)
        torch.cuda.synchronize()
        assert abs(loss.item() - ref["loss"]) / ref["loss"] < 1e-2, (it, loss.item(), ref["loss"])
        for k, gr in eng.proj.grads().items():
            assert _cos(gr.cpu(), ref["grads"][k]) > 0.999, (it, k)


def test_qwen3_recipe_prefill_and_cached_decode_vs_oracle(cuda_dev):
    """Inference on the Qwen3 recipe: prefill logits and KV-cached decode steps against the oracle's cache-free
    forward (teacher-forced on the oracle's greedy tokens)."""
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    vkw = dict(hidden=64, ff=128, layers=1, heads=1, image=64, patch=16)
    lkw = dict(vocab=512, hidden=128, ff=256, layers=3, hq=4, hkv=2, hd=64, window=0, q_scalar=64.0, pattern=1,
               sandwich_norm=False, silu=True, scale_embed=False, tie_lm_head=False)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    px, ids, _ = O.synthetic_batch(2, 6, ovc, olc, left_pad=False)
    q = lambda t: t.to(torch.bfloat16).float()
    Wv_q = {k: (q(v) if v.dim() == 2 and k != "pos" else v) for k, v in Wv.items()}
    Wl_q = {k: (q(v) if v.dim() == 2 else v) for k, v in Wl.items()}
    new = 6
    toks, ref_logits = O.greedy_generate(px, ids, Wv_q, P, Wl_q, ovc, olc, new, eos_id=-1, pad_id=olc.pad_id,
                                         scale_dtype=torch.bfloat16)
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev),
                         E.ProjectorState(P, cuda_dev), cuda_dev)
    logits = eng.prefill(px, ids, max_new_tokens=new)
    for t in range(new):
        lg, r = logits.float().cpu(), ref_logits[:, t]
        assert _cos(lg, r) > 0.999, t
        assert (lg - r).abs().max() / r.abs().max() < 3e-2, t
        if t + 1 < new:
            logits = eng.decode_step(toks[:, t].to(cuda_dev))


class _CollectingDataset(torch.utils.data.Dataset):
    """Every item forces a full garbage collection in the worker that produces it."""

    def __len__(self):
        return 6

    def __getitem__(self, i):
        import gc
        gc.collect()
        return torch.full((2,), float(i))


def test_forked_loader_workers_do_not_collect_inherited_cuda_garbage(cuda_dev):
    """The trainers fork DataLoader workers at the start of every epoch.  Unreachable CUDA objects waiting for the
    cycle collector at that moment (a dropped geometry's graph) would be destroyed by the CHILD's collector,
    which aborts the worker; `_ShardedLoader` collects and freezes before the fork (regression: a worker died with
    'Fatal Python error: Aborted ... Garbage-collecting' when the trainer test ran late in the suite)."""
    import gc
    from torch.utils.data import DataLoader
    from projectiontrainer_b200.projector_trainer import _ShardedLoader

    class Cycle:
        pass

    x = torch.ones(8, device=cuda_dev)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        y = x * 2
    graph.replay()
    torch.cuda.synchronize()
    c = Cycle()
    c.me, c.graph, c.event, c.pinned = c, graph, torch.cuda.Event(), torch.empty(1 << 20).pin_memory()
    c.event.record()
    gc.collect()
    gc.collect()                                  # age the cycle into the oldest generation, then orphan it
    del c, graph, y
    loader = _ShardedLoader(DataLoader(_CollectingDataset(), batch_size=2, num_workers=2), 0, 1)
    got = torch.cat([b for b in loader])
    assert got.shape == (6, 2) and got[:, 0].tolist() == [0.0, 1.0, 2.0, 3.0, 4.0, 5.0]
