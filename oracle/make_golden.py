#!/usr/bin/env python
"""Generates tests/golden/*.npz by running the UNMODIFIED reference in this container.

TEST INFRASTRUCTURE ONLY (see oracle/stage1_oracle.py header).  Needs /root/reference and HF
transformers; it does not travel to the GPU box — only the vectors it writes do.

What runs: the reference's own `ProjectionTrainerStage1.train()` (Stage1/projector_trainer.py:124)
and `MLPProjector` (Stage1/projectors.py:4), imported from /root/reference/Stage1, driving HF
`SiglipVisionModel` and `Gemma3ForCausalLM` built from explicit configs on CPU in fp32.  The only
stand-ins are a single-process accelerator duck type (accelerate is not installed here; the
14 members the trainer touches) and an in-memory dataset with the batch schema of
Stage1/train_projection_stage1.py:114-118.  Weights come from oracle.stage1_oracle.init_* (seeded,
so tests can rebuild them) and are loaded INTO the HF modules; inputs from synthetic_batch.

Recorded per optimizer step: the loss the trainer logs, the four projector gradients as they are
when clip_grad_norm_ is entered, the learning rate, and the parameters after optimizer.step().
From step 0 also: vision features after the `[:, 1:, :]` slice, projector output, spliced
embeddings, attention mask and LM labels exactly as handed to the language model.

    python oracle/make_golden.py            # tiny cases (seconds)
    python oracle/make_golden.py --full     # + config 1 at full model size (minutes, ~6 GB RAM)
"""
from __future__ import annotations

import argparse
import logging
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference/Stage1")

from oracle import stage1_oracle as O  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"

CASES = {
    # name: (siglip kwargs, gemma kwargs, B, T, left_pad, grad_accum, epochs, warmup_ratio)
    "tiny_pad_gqa": (dict(hidden=64, ff=128, layers=2, heads=2, image=64, patch=16),
                     dict(vocab=512, hidden=64, ff=128, layers=6, hq=4, hkv=2, hd=16, window=8,
                          q_scalar=16.0, rope_factor_global=8.0), 3, 12, True, 2, 2, 0.0),
    "tiny_nopad_mqa": (dict(hidden=64, ff=128, layers=1, heads=4, image=48, patch=16),
                       dict(vocab=304, hidden=48, ff=96, layers=7, hq=4, hkv=1, hd=32, window=5,
                            q_scalar=32.0), 2, 9, False, 1, 3, 0.34),
    "config1_full": ({}, {}, 2, 64, True, 2, 1, 0.0),
    # the BENCHMARKED geometry (BASELINE.json configs[2]: T = 128, S = 703 > window 512) at full model size, B = 2
    "config3_b2_full": ({}, {}, 2, 128, True, 1, 1, 0.0),
    # Qwen3 decoder recipe (the reference's launch script trains against Qwen/Qwen3-8B:
    # run_projection_train_stage1.sh:12): pre-norm only, SwiGLU, q/k norms, unscaled embeddings, untied LM head,
    # full attention everywhere, GQA; the trainer takes its non-Gemma embedding branch (projector_trainer.py:190-195)
    "tiny_qwen3_gqa": (dict(hidden=64, ff=128, layers=2, heads=2, image=64, patch=16),
                       dict(vocab=512, hidden=64, ff=128, layers=3, hq=4, hkv=2, hd=32, window=0, q_scalar=32.0,
                            pattern=1, sandwich_norm=False, silu=True, scale_embed=False, tie_lm_head=False),
                       3, 12, True, 1, 3, 0.34),
}


# ------------------------------------------------------------------ oracle weights -> HF modules
def build_hf_siglip(cfg: O.SiglipCfg, W: dict):
    from transformers import SiglipVisionConfig, SiglipVisionModel
    hc = SiglipVisionConfig(hidden_size=cfg.hidden, intermediate_size=cfg.ff, num_hidden_layers=cfg.layers,
                            num_attention_heads=cfg.heads, image_size=cfg.image, patch_size=cfg.patch,
                            layer_norm_eps=cfg.eps)
    m = SiglipVisionModel(hc)
    d = cfg.hidden
    sd = {"embeddings.patch_embedding.weight": W["patch_w"].reshape(d, 3, cfg.patch, cfg.patch),
          "embeddings.patch_embedding.bias": W["patch_b"],
          "embeddings.position_embedding.weight": W["pos"],
          "post_layernorm.weight": W["post_g"], "post_layernorm.bias": W["post_b"]}
    for l in range(cfg.layers):
        s, p = f"encoder.layers.{l}.", f"l{l}."
        sd[s + "layer_norm1.weight"], sd[s + "layer_norm1.bias"] = W[p + "ln1_g"], W[p + "ln1_b"]
        sd[s + "layer_norm2.weight"], sd[s + "layer_norm2.bias"] = W[p + "ln2_g"], W[p + "ln2_b"]
        for i, n in enumerate("qkv"):
            sd[s + f"self_attn.{n}_proj.weight"] = W[p + "qkv_w"][i * d:(i + 1) * d]
            sd[s + f"self_attn.{n}_proj.bias"] = W[p + "qkv_b"][i * d:(i + 1) * d]
        sd[s + "self_attn.out_proj.weight"], sd[s + "self_attn.out_proj.bias"] = W[p + "o_w"], W[p + "o_b"]
        sd[s + "mlp.fc1.weight"], sd[s + "mlp.fc1.bias"] = W[p + "fc1_w"], W[p + "fc1_b"]
        sd[s + "mlp.fc2.weight"], sd[s + "mlp.fc2.bias"] = W[p + "fc2_w"], W[p + "fc2_b"]
    missing, unexpected = m.vision_model.load_state_dict(sd, strict=False)
    assert not unexpected, unexpected
    assert all(k.startswith("head.") for k in missing), missing   # pooling head: unused by Stage 1
    return m.eval()


def build_hf_gemma(cfg: O.GemmaCfg, W: dict):
    from transformers import Gemma3ForCausalLM, Gemma3TextConfig
    kw = dict(vocab_size=cfg.vocab, hidden_size=cfg.hidden, intermediate_size=cfg.ff,
              num_hidden_layers=cfg.layers, num_attention_heads=cfg.hq, num_key_value_heads=cfg.hkv,
              head_dim=cfg.hd, sliding_window=cfg.window, query_pre_attn_scalar=int(cfg.q_scalar),
              max_position_embeddings=32768, rms_norm_eps=cfg.eps, pad_token_id=cfg.pad_id,
              rope_parameters={
                  "full_attention": ({"rope_type": "linear", "factor": cfg.rope_factor_global,
                                      "rope_theta": cfg.theta_global} if cfg.rope_factor_global != 1.0
                                     else {"rope_type": "default", "rope_theta": cfg.theta_global}),
                  "sliding_attention": {"rope_type": "default", "rope_theta": cfg.theta_local}})
    hc = Gemma3TextConfig(**kw)
    m = Gemma3ForCausalLM(hc)
    qd, kd = cfg.hq * cfg.hd, cfg.hkv * cfg.hd
    sd = {"model.embed_tokens.weight": W["embed"], "lm_head.weight": W["embed"],
          "model.norm.weight": W["final_norm"]}
    for l in range(cfg.layers):
        s, p = f"model.layers.{l}.", f"l{l}."
        sd[s + "input_layernorm.weight"] = W[p + "in_norm"]
        sd[s + "post_attention_layernorm.weight"] = W[p + "post_attn_norm"]
        sd[s + "pre_feedforward_layernorm.weight"] = W[p + "pre_ffn_norm"]
        sd[s + "post_feedforward_layernorm.weight"] = W[p + "post_ffn_norm"]
        sd[s + "self_attn.q_norm.weight"], sd[s + "self_attn.k_norm.weight"] = W[p + "q_norm"], W[p + "k_norm"]
        sd[s + "self_attn.q_proj.weight"] = W[p + "qkv_w"][:qd]
        sd[s + "self_attn.k_proj.weight"] = W[p + "qkv_w"][qd:qd + kd]
        sd[s + "self_attn.v_proj.weight"] = W[p + "qkv_w"][qd + kd:]
        sd[s + "self_attn.o_proj.weight"] = W[p + "o_w"]
        sd[s + "mlp.gate_proj.weight"] = W[p + "gu_w"][:cfg.ff]
        sd[s + "mlp.up_proj.weight"] = W[p + "gu_w"][cfg.ff:]
        sd[s + "mlp.down_proj.weight"] = W[p + "down_w"]
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    assert m.lm_head.weight.data_ptr() == m.model.embed_tokens.weight.data_ptr() or \
        torch.equal(m.lm_head.weight, m.model.embed_tokens.weight)
    return m.eval()


def build_hf_qwen3(cfg: O.GemmaCfg, W: dict):
    """HF Qwen3ForCausalLM holding the oracle's weights; norm gains are stored as (gain - 1) in W."""
    from transformers import Qwen3Config, Qwen3ForCausalLM
    hc = Qwen3Config(vocab_size=cfg.vocab, hidden_size=cfg.hidden, intermediate_size=cfg.ff,
                     num_hidden_layers=cfg.layers, num_attention_heads=cfg.hq, num_key_value_heads=cfg.hkv,
                     head_dim=cfg.hd, rms_norm_eps=cfg.eps, max_position_embeddings=32768,
                     tie_word_embeddings=cfg.tie_lm_head, use_sliding_window=False, attention_bias=False,
                     pad_token_id=cfg.pad_id,
                     rope_parameters={"rope_type": "default", "rope_theta": cfg.theta_global})
    m = Qwen3ForCausalLM(hc)
    assert m.config.hidden_act == "silu" and all(t == "full_attention" for t in m.config.layer_types)
    qd, kd = cfg.hq * cfg.hd, cfg.hkv * cfg.hd
    sd = {"model.embed_tokens.weight": W["embed"], "lm_head.weight": W.get("lm_head", W["embed"]),
          "model.norm.weight": 1 + W["final_norm"]}
    for l in range(cfg.layers):
        s, p = f"model.layers.{l}.", f"l{l}."
        sd[s + "input_layernorm.weight"] = 1 + W[p + "in_norm"]
        sd[s + "post_attention_layernorm.weight"] = 1 + W[p + "pre_ffn_norm"]
        sd[s + "self_attn.q_norm.weight"], sd[s + "self_attn.k_norm.weight"] = 1 + W[p + "q_norm"], 1 + W[p + "k_norm"]
        sd[s + "self_attn.q_proj.weight"] = W[p + "qkv_w"][:qd]
        sd[s + "self_attn.k_proj.weight"] = W[p + "qkv_w"][qd:qd + kd]
        sd[s + "self_attn.v_proj.weight"] = W[p + "qkv_w"][qd + kd:]
        sd[s + "self_attn.o_proj.weight"] = W[p + "o_w"]
        sd[s + "mlp.gate_proj.weight"] = W[p + "gu_w"][:cfg.ff]
        sd[s + "mlp.up_proj.weight"] = W[p + "gu_w"][cfg.ff:]
        sd[s + "mlp.down_proj.weight"] = W[p + "down_w"]
    missing, unexpected = m.load_state_dict(sd, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    return m.eval()


# ------------------------------------------------------------------ stand-ins the trainer needs
class Recorder:
    def __init__(self):
        self.steps = []          # dicts: loss, grads, lr, params
        self.first = {}          # intermediates of step 0
        self.perms = []


class SingleProcessAccelerator:
    """The 14 accelerator members Stage1/projector_trainer.py touches, one process."""

    def __init__(self, grad_accum: int, rec: Recorder, projector_ref):
        self.device = torch.device("cpu")
        self.is_main_process = True
        self.num_processes = 1
        self.process_index = 0
        self.gradient_accumulation_steps = grad_accum
        self.sync_gradients = True
        self.trackers = []
        self._rec, self._proj = rec, projector_ref
        self._cur = None

    def prepare(self, *objs):
        return objs if len(objs) > 1 else objs[0]

    def prepare_model(self, m, device_placement=True):
        return m

    def unwrap_model(self, m):
        return m

    def backward(self, loss):
        loss.backward()

    def clip_grad_norm_(self, params, max_norm):
        params = list(params)
        self._cur = {"grads": {n: p.grad.detach().clone() for n, p in self._proj.named_parameters()}}
        return torch.nn.utils.clip_grad_norm_(params, max_norm)

    def gather(self, t):
        t = t.detach().reshape(1)
        if self._cur is not None and "loss" not in self._cur:
            # first gather after the optimizer step (projector_trainer.py:251): parameters are updated
            self._cur["loss"] = float(t[0])
            self._cur["params"] = {n: p.detach().clone() for n, p in self._proj.named_parameters()}
        return t

    def log(self, d, step=None):
        if self._cur is not None and "train/learning_rate" in d:
            self._cur["lr_after"] = float(d["train/learning_rate"])
            self._rec.steps.append(self._cur)
            self._cur = None


class ListDataset(torch.utils.data.Dataset):
    def __init__(self, pixels, ids, labels):
        self.items = [{"pixel_values": pixels[i], "token_ids": ids[i], "labels": labels[i], "_idx": i}
                      for i in range(pixels.shape[0])]

    def __len__(self):
        return len(self.items)

    def __getitem__(self, i):
        return self.items[i]


class Tok:
    def __init__(self, pad):
        self.pad_token_id, self.eos_token_id = pad, 1


def run_case(name: str, out_dir: Path):
    vkw, lkw, B, T, left_pad, accum, epochs, warm = CASES[name]
    import projector_trainer as ref_trainer      # /root/reference/Stage1/projector_trainer.py
    from projectors import MLPProjector          # /root/reference/Stage1/projectors.py
    logging.getLogger("projector_trainer").setLevel(logging.WARNING)

    vc, lc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    t0 = time.time()
    Wv, Wl, P = O.init_siglip(vc), O.init_gemma(lc), O.init_projector(vc.hidden, lc.hidden)
    vision = build_hf_siglip(vc, Wv)
    llm = build_hf_gemma(lc, Wl) if lc.sandwich_norm else build_hf_qwen3(lc, Wl)
    proj = MLPProjector(vc.hidden, lc.hidden)
    proj.load_state_dict(P)
    pixels, ids, labels = O.synthetic_batch(B, T, vc, lc, left_pad=left_pad)
    print(f"[{name}] models built in {time.time() - t0:.1f}s", flush=True)

    rec = Recorder()
    acc = SingleProcessAccelerator(accum, rec, proj)

    # hooks: what the reference hands to / gets from the three modules on the first step
    def vis_hook(mod, args, kwargs, out):
        if "patch" not in rec.first:
            rec.first["patch"] = out.last_hidden_state[:, 1:, :].detach().clone()
    def proj_hook(mod, args, out):
        if "projected" not in rec.first:
            rec.first["projected"] = out.detach().clone()
    def llm_pre(mod, args, kwargs):
        if "spliced" not in rec.first:
            rec.first["spliced"] = kwargs["inputs_embeds"].detach().clone()
            rec.first["attention_mask"] = kwargs["attention_mask"].clone()
            rec.first["lm_labels"] = kwargs["labels"].clone()
    def emb_hook(mod, args, out):
        rec.perms.append(args[0].clone())
    vision.vision_model.register_forward_hook(vis_hook, with_kwargs=True)
    proj.register_forward_hook(proj_hook)
    llm.register_forward_pre_hook(llm_pre, with_kwargs=True)
    llm.get_input_embeddings().register_forward_hook(emb_hook)

    # the DataLoader shuffles (projector_trainer.py:56-61) and spawns 2 workers; keep it as is
    torch.manual_seed(7)
    trainer = ref_trainer.ProjectionTrainerStage1(
        acc, vision, llm, proj, processor=None, tokenizer=Tok(lc.pad_id),
        train_dataset=ListDataset(pixels, ids, labels), val_dataset=None,
        output_dir=str(out_dir / "_tmp_ckpt"), batch_size=B, learning_rate=1e-3, weight_decay=0.01,
        num_epochs=epochs, gradient_accumulation_steps=accum, warmup_ratio=warm, save_every_n_epochs=0)
    t0 = time.time()
    trainer.train()
    print(f"[{name}] reference train(): {len(rec.steps)} step(s) in {time.time() - t0:.1f}s, "
          f"losses {[round(s['loss'], 5) for s in rec.steps]}", flush=True)
    assert len(rec.steps) == epochs

    # sample order the shuffling loader produced at each step (match on token rows)
    perms = []
    for seen in rec.perms:
        perm = [next(j for j in range(B) if torch.equal(ids[j], row)) for row in seen]
        perms.append(perm)
    full = name.endswith("_full")
    save = {"case": np.array(name), "B": B, "T": T, "left_pad": int(left_pad), "grad_accum": accum,
            "epochs": epochs, "warmup_ratio": warm, "base_lr": 1e-3, "weight_decay": 0.01,
            "perms": np.array(perms, dtype=np.int64),
            "attention_mask": rec.first["attention_mask"].numpy(),
            "lm_labels": rec.first["lm_labels"].numpy(),
            "weights_abs_sum": np.array([float(sum(v.double().abs().sum() for v in W.values()))
                                         for W in (Wv, Wl, P)])}
    rng = np.random.default_rng(0)

    def put(key, t: torch.Tensor):
        a = t.detach().float().numpy()
        if full and a.size > 300_000:
            idx = np.sort(rng.choice(a.size, 262_144, replace=False))
            save[key + "__idx"] = idx.astype(np.int64)
            save[key + "__sample"] = a.reshape(-1)[idx]
            save[key + "__l2"] = np.array(float(np.sqrt((a.astype(np.float64) ** 2).sum())))
            save[key + "__shape"] = np.array(a.shape, dtype=np.int64)
        else:
            save[key] = a
    for k in ("patch", "projected", "spliced"):
        put(k, rec.first[k])
    for i, s in enumerate(rec.steps):
        save[f"step{i}.loss"] = np.array(s["loss"])
        save[f"step{i}.lr_after"] = np.array(s["lr_after"])
        for n, g in s["grads"].items():
            put(f"step{i}.grad.{n}", g)
        if not full and i == len(rec.steps) - 1:
            for n, p in s["params"].items():
                put(f"step{i}.param.{n}", p)
    out_dir.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(out_dir / f"{name}.npz", **save)
    import shutil
    shutil.rmtree(out_dir / "_tmp_ckpt", ignore_errors=True)
    print(f"[{name}] wrote {out_dir / (name + '.npz')} "
          f"({(out_dir / (name + '.npz')).stat().st_size / 1e6:.2f} MB)", flush=True)


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--full", action="store_true")
    ap.add_argument("--only", default=None)
    a = ap.parse_args()
    torch.set_num_threads(8)
    names = [a.only] if a.only else [n for n in CASES if a.full or not n.endswith("_full")]
    for n in names:
        run_case(n, GOLDEN)
