"""Stage-2 LoRA path on the CUDA kernels (second half), GPU only.  Checkers: the vectors
recorded from the reference's `VQATrainerStage2.train()` with `enable_qlora=True` (tests/golden/vqa_lora_*.npz) and the
CPU oracle on the same inputs.  bf16 tolerances: loss 1e-2 relative, gradient cosine per adapter group."""
import math
from pathlib import Path

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"


def _cos(a, b):
    a, b = a.double().flatten(), b.double().flatten()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-300))


def _to_peft(L: dict) -> dict:
    """oracle keys `l<i>.<target>.A` -> peft keys."""
    from projectiontrainer_b200.lora import LoraState
    out = {}
    for k, v in L.items():
        li, t, w = k.split(".")
        out[LoraState.key(int(li[1:]), t, w)] = v
    return out


def _by_oracle_key(lo, buf) -> dict:
    from projectiontrainer_b200.lora import LoraState
    sd = lo.state_dict(buf)
    out = {}
    for li, groups in enumerate(lo.layers):
        for g in groups.values():
            for t, _, _, _ in g.parts:
                for w in "AB":
                    out[f"l{li}.{t}.{w}"] = sd[LoraState.key(li, t, w)]
    return out


def _group_cos(ours: dict, ref: dict, which: str) -> float:
    """Cosine over all adapter matrices of one kind (A or B) taken as one vector."""
    ks = sorted(k for k in ours if k.endswith("." + which))
    a = torch.cat([ours[k].detach().cpu().float().flatten() for k in ks])
    b = torch.cat([torch.as_tensor(ref[k]).float().flatten() for k in ks])
    return _cos(a, b)


class _Tok:
    def __init__(self, pad, side):
        self.pad_token_id, self.eos_token_id, self.padding_side = pad, 1, side


def _collated(samples, idx, tok):
    from projectiontrainer_b200.vqa_trainer import vqa_collate_fn
    px, qs, as_ = samples
    return vqa_collate_fn([{"pixel_values": px[i], "question_input_ids": qs[i], "answer_input_ids": as_[i]}
                           for i in idx], tok)


@pytest.mark.parametrize("case", ["vqa_lora_all_right", "vqa_lora_qvd_left_frozen_proj"])
def test_lora_engine_trajectory_matches_reference_golden(cuda_dev, case):
    """The reference's recorded `enable_qlora=True` run through `train_step`: micro-batch losses, accumulated adapter
    (and projector) gradients at each clip, one clip per module + AdamW, adapter weights at the end, validation loss."""
    from oracle import stage1_oracle as O
    from oracle import stage2_oracle as O2
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200.lora import LoraState
    from projectiontrainer_b200.vqa_trainer import accumulate_sync_flag, vqa_step_inputs
    vkw, lkw = O2.LORA_CASES[case][:2]
    r, alpha, targets, b_std, train_proj = O2.LORA_CASES[case][-1]
    g = dict(np.load(GOLDEN / f"{case}.npz", allow_pickle=False))
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    L0 = O2.init_lora(olc, r, targets, b_std=b_std)
    train = O2.synthetic_vqa_samples(int(g["n_train"]), ovc, olc, seed=4321)
    val = O2.synthetic_vqa_samples(int(g["n_val"]), ovc, olc, seed=9876)
    GA, B, epochs = int(g["grad_accum"]), int(g["B"]), int(g["epochs"])
    tok = _Tok(olc.pad_id, str(g["padding_side"]))
    lcfg = GemmaCfg(**lkw)
    lo = LoraState(lcfg, cuda_dev, r=r, alpha=alpha, targets=targets)
    lo.load_state_dict(_to_peft(L0))
    assert lo.num_parameters() == sum(v.numel() for v in L0.values())
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, lcfg, cuda_dev),
                         E.ProjectorState(P, cuda_dev), cuda_dev, loss_divisor=GA * int(g["backward_divisor"]),
                         lora=lo, train_projector=train_proj)
    n_batches = math.ceil(int(g["n_train"]) / B)
    total = epochs * math.ceil(n_batches / GA)
    warm = math.ceil(float(g["warmup_ratio"]) * total)
    mi, step, micro = 0, 0, 0
    for ep in range(epochs):
        for bi in range(n_batches):
            sync, micro = accumulate_sync_flag(micro, GA, bi == n_batches - 1)
            assert sync == bool(g[f"micro{mi}.sync"])
            cb = _collated(train, g[f"micro{mi}.idx"].tolist(), tok)
            batch = vqa_step_inputs(cb, olc.pad_id, 1)
            lr = O.cosine_lr(step, float(g["base_lr"]), warm, total)
            loss = eng.train_step(batch, lr, pad_id=olc.pad_id, weight_decay=float(g["weight_decay"]),
                                  max_norm=1.0, sync_gradients=sync)
            torch.cuda.synchronize()
            ref_loss = float(g[f"micro{mi}.loss"])
            assert abs(loss.item() - ref_loss) / ref_loss < 1e-2, (mi, loss.item(), ref_loss)
            if sync:
                ours = _by_oracle_key(lo, lo.grad)
                ref = {k: g[f"step{step}.lora_grad.{k}"] for k in ours}
                for w in "AB":
                    c = _group_cos(ours, ref, w)
                    assert c > (0.999 if step == 0 else 0.99), (step, w, c)
                if train_proj:
                    for k, gr in eng.proj.grads().items():
                        # identical parameters at step 0; afterwards the two trajectories differ by Adam's sign-like
                        # first updates on near-zero gradient entries (bf16 noise), in adapters AND projector here
                        c = _cos(gr.cpu(), torch.from_numpy(g[f"step{step}.grad.{k}"]))
                        assert c > (0.999 if step == 0 else 0.98), (step, k, c)
                step += 1
            mi += 1
    assert step == int(g["n_steps"])
    final = _by_oracle_key(lo, lo.flat)
    for w in "AB":
        ks = sorted(k for k in final if k.endswith("." + w))
        ours = torch.cat([(final[k].cpu() - L0[k]).flatten() for k in ks])
        ref = torch.cat([(torch.from_numpy(g[f"step{step - 1}.lora_param.{k}"]) - L0[k]).flatten() for k in ks])
        assert _cos(ours, ref) > 0.97, (w, _cos(ours, ref))
    # block layout: nothing outside the adapters' own entries ever moved
    assert int((lo.flat != 0).sum().item()) <= lo.num_parameters()
    if not train_proj:
        for k, p in eng.proj.params().items():
            assert torch.equal(p.cpu(), P[k]), k
    vt = _Tok(olc.pad_id, "left")
    n_val, losses = int(g["n_val"]), []
    for s in range(0, n_val, B):
        vb = vqa_step_inputs(_collated(val, list(range(s, min(s + B, n_val))), vt), olc.pad_id, 1)
        losses.append(eng.eval_loss(vb, pad_id=olc.pad_id).item())
    assert abs(np.mean(losses) - float(g["val_loss"][-1])) / float(g["val_loss"][-1]) < 1e-2


@pytest.mark.parametrize("targets,train_proj", [(None, True), (("k_proj", "o_proj", "up_proj"), False)])
def test_lora_step_hd256_fused_path_vs_oracle(cuda_dev, targets, train_proj):
    """The flagship kernels with adapters riding along (fused attention fwd / pair bwd, GeGLU epilogues, attention-delta
    epilogue, all with the second contraction segment) under the Stage-2 layout, through the CUDA graph: loss, adapter
    gradients and projector gradients against the CPU oracle holding the same bf16-rounded matrices."""
    from oracle import stage1_oracle as O
    from oracle import stage2_oracle as O2
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200.lora import TARGETS, LoraState
    from projectiontrainer_b200.vqa_trainer import vqa_step_inputs
    targets = targets or TARGETS
    vkw = dict(hidden=128, ff=256, layers=2, heads=2, image=160, patch=16)
    lkw = dict(vocab=1024, hidden=256, ff=256, layers=7, hq=4, hkv=2, hd=256, window=40, q_scalar=256.0,
               rope_factor_global=8.0)
    r, alpha = 16, 32.0
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    qz = lambda t: t.to(torch.bfloat16).float()
    L0 = {k: qz(v) for k, v in O2.init_lora(olc, r, targets, b_std=0.02).items()}
    samples = O2.synthetic_vqa_samples(3, ovc, olc, seed=77, q_len=(3, 14), a_len=(5, 30))
    cb = _collated(samples, [0, 1, 2], _Tok(olc.pad_id, "left"))
    Wv_q = {k: (qz(v) if v.dim() == 2 and k != "pos" else v) for k, v in Wv.items()}
    Wl_q = {k: (qz(v) if v.dim() == 2 else v) for k, v in Wl.items()}
    ref = O2.vqa_micro_step(cb["pixel_values"], cb["question_input_ids"], cb["answer_input_ids"], Wv_q, P, Wl_q,
                            ovc, olc, scale_dtype=torch.bfloat16, lora=L0, lora_scale=alpha / r,
                            train_projector=train_proj)
    lcfg = GemmaCfg(**lkw)
    lo = LoraState(lcfg, cuda_dev, r=r, alpha=alpha, targets=targets)
    lo.load_state_dict(_to_peft(L0))
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, lcfg, cuda_dev),
                         E.ProjectorState(P, cuda_dev), cuda_dev, lora=lo, train_projector=train_proj)
    assert eng.fuse_attn_fwd and eng.fuse_geglu and eng.fuse_attn_delta
    batch = vqa_step_inputs(cb, olc.pad_id, bucket=32)
    for it in range(3):          # eager, capture, replay
        lo.grad.fill_(float("nan"))            # every adapter entry's gradient is rewritten by each step
        loss = eng._graphed_forward_backward({k: (v.to(cuda_dev) if torch.is_tensor(v) else v)
                                              for k, v in batch.items()}, olc.pad_id)
        torch.cuda.synchronize()
        assert abs(loss.item() - ref["loss"]) / ref["loss"] < 1e-2, (it, loss.item(), ref["loss"])
        ours = _by_oracle_key(lo, lo.grad)
        for w in "AB":
            c = _group_cos(ours, ref["lora_grads"], w)
            assert c > 0.999, (it, w, c)
        worst = min(_cos(ours[k].cpu(), ref["lora_grads"][k]) for k in ours)
        assert worst > 0.99, (it, worst)
        if train_proj:
            for k, gr in eng.proj.grads().items():
                assert _cos(gr.cpu(), ref["grads"][k]) > 0.999, (it, k)
    # merged weights (generation): W + (alpha / r) B A per packed projection, against the oracle's merge
    Wm = O2.lora_merge(Wl_q, L0, olc, alpha / r)
    for li, Lm in enumerate(eng.merge_adapters()):
        for key in ("qkv_w", "o_w", "gu_w", "down_w"):
            d = (Lm[key].float().cpu() - Wm[f"l{li}.{key}"]).abs().max().item()
            assert d <= 2.0 ** -8 * Wm[f"l{li}.{key}"].abs().max().item(), (li, key, d)
    # eval / generation run with the adapters: the eval loss equals the training forward's
    ev = eng.eval_loss(batch, pad_id=olc.pad_id).item()
    assert abs(ev - ref["loss"]) / ref["loss"] < 1e-2
    px = cb["pixel_values"].to(cuda_dev)
    gen = eng.generate(px, cb["question_input_ids"], max_new_tokens=4, prompt_pad_id=olc.pad_id)
    with eng.merged(eng.merge_adapters()):
        gen2 = eng.generate(px, cb["question_input_ids"], max_new_tokens=4, prompt_pad_id=olc.pad_id)
    assert torch.equal(gen, gen2) and eng.lora is lo


def test_qlora_trainer_dropin_end_to_end_tiny(cuda_dev, tmp_path):
    """`VQATrainerStage2(enable_qlora=True)` through the real DataLoader: adapters (peft init: B = 0) and the projector
    train, the checkpoint holds peft's adapter files, and a second trainer resumed from them
    (`--resume_qlora_adapter_path`) starts from the same adapter weights and the same validation loss."""
    import json
    from oracle import stage1_oracle as O
    from oracle import stage2_oracle as O2
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.accelerator_setup import Stage1Accelerator
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200.projectors import MLPProjector
    from projectiontrainer_b200.train_vqa_stage2 import load_pretrained_projector, load_qlora_adapter
    from projectiontrainer_b200.vqa_trainer import VQATrainerStage2
    vkw, lkw = O2.LORA_CASES["vqa_lora_all_right"][:2]
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    train = O2.synthetic_vqa_samples(6, ovc, olc, seed=4321)
    val = O2.synthetic_vqa_samples(2, ovc, olc, seed=9876)

    class DS(torch.utils.data.Dataset):
        def __init__(self, s): self.s = s
        def __len__(self): return len(self.s[1])
        def __getitem__(self, i):
            return {"pixel_values": self.s[0][i], "question_input_ids": self.s[1][i], "answer_input_ids": self.s[2][i]}

    class Tok(_Tok):
        saved = None
        def batch_decode(self, ids, skip_special_tokens=True):
            return [" ".join(str(int(t)) for t in row) for row in ids]
        def save_pretrained(self, path):
            Tok.saved = path

    def make(out_dir, proj_sd, lora_cfg, lora_sd, epochs):
        proj = MLPProjector(ovc.hidden, olc.hidden)
        proj.load_state_dict(proj_sd)
        acc = Stage1Accelerator(gradient_accumulation_steps=2, device=cuda_dev)
        tr = VQATrainerStage2(acc, E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev),
                              E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev), proj, Tok(olc.pad_id, "right"), DS(train),
                              DS(val), str(out_dir), 2, 1e-3, 0.01, epochs, 2, 0.25, freeze_vision_encoder=True,
                              freeze_projection_layer=False, freeze_llm=True, enable_qlora=True,
                              train_ve_first_epoch=False, wandb_project="none", lora_config=lora_cfg,
                              lora_state_dict=lora_sd)
        real_eval = tr.evaluate
        tr.evaluate = lambda ep, gs: real_eval(ep, gs, max_new_tokens=3)
        return tr, acc, proj

    cfg = {"r": 8, "lora_alpha": 16, "target_modules": ["q_proj", "v_proj", "gate_proj", "down_proj"], "lora_dropout": 0.05}
    tr, acc, proj = make(tmp_path / "a", P, cfg, None, 2)
    lo = tr.engine.lora
    assert lo.r == 8 and lo.targets == ("q_proj", "v_proj", "gate_proj", "down_proj") and lo.dropout == 0.05
    sd0 = {k: v.cpu().clone() for k, v in lo.state_dict().items()}
    assert all((v == 0).all() for k, v in sd0.items() if "lora_B" in k)          # peft init
    tr.train()
    losses = [v["train/batch_loss"] for s, v in acc.logged if "train/batch_loss" in v]
    vals = [v["val/loss"] for s, v in acc.logged if "val/loss" in v]
    assert len(losses) == 6 and all(math.isfinite(x) for x in losses) and len(vals) == 2
    sd1 = {k: v.cpu().clone() for k, v in lo.state_dict().items()}
    assert all(not torch.equal(sd1[k], sd0[k]) for k in sd0)                     # every adapter matrix moved
    ck = tmp_path / "a" / "checkpoint-epoch_2"
    ad = ck / "language_model"
    assert Tok.saved == str(ad)                                                  # ref :740-741
    meta = json.loads((ad / "adapter_config.json").read_text())
    assert meta["peft_type"] == "LORA" and meta["r"] == 8 and meta["lora_alpha"] == 16 and meta["lora_dropout"] == 0.05
    cfg2, sd2 = load_qlora_adapter(str(ad))
    assert set(sd2) == set(sd1) and all(torch.equal(sd2[k], sd1[k]) for k in sd1)
    opt = torch.load(ck / "optimizer.bin")
    assert opt["lora"]["step"] == 4 and set(opt["lora"]["exp_avg"]) == set(sd1) and opt["step"] == 4
    back = load_pretrained_projector(str(ck / "projection_layer"))
    # resumed trainer: same adapters, same validation loss as the first trainer ended with
    tr2, acc2, _ = make(tmp_path / "b", back.state_dict(), cfg2, sd2, 1)
    assert all(torch.equal(v.cpu(), sd1[k]) for k, v in tr2.engine.lora.state_dict().items())
    v2 = tr2.evaluate(0, 0)
    assert abs(v2 - vals[-1]) / vals[-1] < 1e-3, (v2, vals)


@pytest.mark.parametrize("fused", [True, False], ids=["fused_kernels", "unfused_kernels"])
def test_lora_qwen3_recipe_projector_frozen_vs_oracle(cuda_dev, fused, monkeypatch):
    """What the reference's Stage-2 launch script runs (ref:Stage2/run_vqa_train_stage2.sh: Qwen3 + `--enable_qlora`,
    projector frozen): the Qwen3 decoder recipe (pre-norm, residual adds in the o / down epilogues, SwiGLU, untied LM
    head, head_dim 128) with LoRA on all seven projections and no projector gradient, Stage-2 layout, against the CPU
    oracle on the same bf16-rounded matrices; fused and unfused kernel paths."""
    from oracle import stage1_oracle as O
    from oracle import stage2_oracle as O2
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200.lora import LoraState
    from projectiontrainer_b200.vqa_trainer import vqa_step_inputs
    if not fused:
        for k in ("PT_STAGE1_FUSE_ATTN_FWD", "PT_STAGE1_FUSE_SOFTMAX_BWD", "PT_STAGE1_FUSE_GEGLU", "PT_STAGE1_FUSE_CE_STATS"):
            monkeypatch.setenv(k, "0")
    vkw = dict(hidden=128, ff=256, layers=2, heads=2, image=160, patch=16)
    lkw = dict(vocab=1024, hidden=256, ff=384, layers=4, hq=8, hkv=2, hd=128, window=0, q_scalar=128.0, pattern=1,
               sandwich_norm=False, silu=True, scale_embed=False, tie_lm_head=False)
    r, alpha = 16, 32.0
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    qz = lambda t: t.to(torch.bfloat16).float()
    L0 = {k: qz(v) for k, v in O2.init_lora(olc, r, O2.ALL_TARGETS, b_std=0.02).items()}
    samples = O2.synthetic_vqa_samples(3, ovc, olc, seed=78, q_len=(3, 14), a_len=(5, 30))
    cb = _collated(samples, [0, 1, 2], _Tok(olc.pad_id, "left"))
    Wv_q = {k: (qz(v) if v.dim() == 2 and k != "pos" else v) for k, v in Wv.items()}
    Wl_q = {k: (qz(v) if v.dim() == 2 else v) for k, v in Wl.items()}
    ref = O2.vqa_micro_step(cb["pixel_values"], cb["question_input_ids"], cb["answer_input_ids"], Wv_q, P, Wl_q,
                            ovc, olc, scale_dtype=torch.bfloat16, lora=L0, lora_scale=alpha / r, train_projector=False)
    lcfg = GemmaCfg(**lkw)
    lo = LoraState(lcfg, cuda_dev, r=r, alpha=alpha)
    lo.load_state_dict(_to_peft(L0))
    eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, lcfg, cuda_dev),
                         E.ProjectorState(P, cuda_dev), cuda_dev, lora=lo, train_projector=False)
    assert eng.fuse_geglu == fused and eng.fuse_attn_fwd == fused
    batch = vqa_step_inputs(cb, olc.pad_id, bucket=32)
    eng.proj.grad.fill_(123.0)
    for it in range(3):
        lo.grad.fill_(float("nan"))
        loss = eng._graphed_forward_backward({k: (v.to(cuda_dev) if torch.is_tensor(v) else v)
                                              for k, v in batch.items()}, olc.pad_id)
        torch.cuda.synchronize()
        assert abs(loss.item() - ref["loss"]) / ref["loss"] < 1e-2, (it, loss.item(), ref["loss"])
        ours = _by_oracle_key(lo, lo.grad)
        for w in "AB":
            c = _group_cos(ours, ref["lora_grads"], w)
            assert c > 0.999, (it, w, c)
        assert min(_cos(ours[k].cpu(), ref["lora_grads"][k]) for k in ours) > 0.99
    assert (eng.proj.grad == 123.0).all()                 # frozen projector: its backward never ran
    p0 = {k: v.clone() for k, v in eng.proj.params().items()}
    eng.optimizer_step(1e-3, max_norm=1.0)
    assert all(torch.equal(v, p0[k]) for k, v in eng.proj.params().items()) and lo.step == 1


def test_thin_adapter_products_split_k_at_small_batch(cuda_dev):
    """At small batch sizes (the reference's Stage-2 default is 4 per GPU) the rank-wide products t = x A^T and
    dt = dy B have few row tiles and a long contraction: the engine runs them split-K through an f32 scratch matrix.
    Same numbers as the direct launch, forward and input-gradient form, scale included."""
    from projectiontrainer_b200 import engine as E, ops
    eng = E.Stage1Engine.__new__(E.Stage1Engine)
    eng._ws, eng.device = {}, cuda_dev
    torch.manual_seed(0)
    M, K, R = 3324, 4096, 64
    x = torch.randn(M, K, device=cuda_dev).to(torch.bfloat16)
    a = (torch.randn(R, K, device=cuda_dev) * 0.02).to(torch.bfloat16)
    out = torch.empty(M, R, device=cuda_dev, dtype=torch.bfloat16)
    eng._thin(ops.linear, x, a, out, 2.0)
    assert ("lora_thin_f32", (M, R), torch.float32) in eng._ws                      # the split-K route was taken
    ref = 2.0 * x.float() @ a.float().t()
    assert (out.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item()
    direct = ops.linear(x, a, alpha=2.0)
    assert (out.float() - direct.float()).abs().max().item() <= 2.0 ** -7 * ref.abs().max().item()
    dy = torch.randn(M, 12288, device=cuda_dev).to(torch.bfloat16)
    b = (torch.randn(12288, R, device=cuda_dev) * 0.02).to(torch.bfloat16)
    eng._thin(ops.linear_dgrad, dy, b, out, 0.5)
    ref = 0.5 * dy.float() @ b.float()
    assert (out.float() - ref).abs().max().item() <= 2.0 ** -8 * ref.abs().max().item()
    # many row tiles: the direct launch
    eng._ws.clear()
    xl = torch.randn(22496, 1152, device=cuda_dev).to(torch.bfloat16)
    al = (torch.randn(R, 1152, device=cuda_dev) * 0.02).to(torch.bfloat16)
    outl = torch.empty(22496, R, device=cuda_dev, dtype=torch.bfloat16)
    eng._thin(ops.linear, xl, al, outl, 1.0)
    assert not eng._ws
