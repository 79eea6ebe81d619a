"""CPU oracle for the Stage-2 VQA step: frozen LLM, optionally with LoRA adapters.  TEST INFRASTRUCTURE ONLY.

Only tests/ may import this module (see oracle/stage1_oracle.py for the rule); the product path
(projectiontrainer_b200/) never does.

Restates, on top of the Stage-1 oracle's towers, what ref:Stage2/trainer.py does when the vision
encoder and the LLM are frozen and the projector trains (`--unfreeze_projection_layer`, no
`--enable_qlora`, no `--unfreeze_llm`):

  * ref:Stage2/trainer.py:18-61   `vqa_collate_fn` (manual left/right padding of questions and answers)
  * ref:Stage2/trainer.py:336-395 `[image][question][answer]` splice, attention mask, answer-only labels
  * ref:Stage2/trainer.py:397-416 logits -> fp32, shifted `nn.CrossEntropyLoss()` (mean over labels != -100)
  * ref:Stage2/trainer.py:306, 418-441 `accelerator.accumulate`, loss / GA, clip_grad_norm_(1.0), AdamW,
    cosine schedule with warm-up

Parity pin: oracle/make_golden_stage2.py runs the reference's own `VQATrainerStage2.train()` (imported
from /root/reference/Stage2/trainer.py) here on HF SigLIP / Gemma3 and records tests/golden/vqa_*.npz;
tests/test_stage2_oracle_golden.py checks this restatement against those vectors.
"""
from __future__ import annotations

import numpy as np
import torch

from . import stage1_oracle as O


# fixtures recorded by oracle/make_golden_stage2.py:
# name: (siglip kwargs, gemma kwargs, n_train, n_val, batch, padding_side, grad_accum, epochs, warmup_ratio)
GOLDEN_CASES = {
    "vqa_tiny_right_gqa": (dict(hidden=64, ff=128, layers=2, heads=2, image=64, patch=16),
                           dict(vocab=512, hidden=64, ff=128, layers=6, hq=4, hkv=2, hd=16, window=8,
                                q_scalar=16.0, rope_factor_global=8.0), 5, 3, 2, "right", 2, 2, 0.25),
    "vqa_tiny_left_mqa": (dict(hidden=64, ff=128, layers=1, heads=4, image=48, patch=16),
                          dict(vocab=304, hidden=48, ff=96, layers=7, hq=4, hkv=1, hd=32, window=5,
                               q_scalar=32.0), 6, 2, 3, "left", 1, 2, 0.0),
}


# LoRA fixtures (oracle/make_golden_stage2.py, `enable_qlora=True` with a hand-rolled adapter wrapper — peft is not in
# this image): the tuple above + (rank, lora_alpha, target modules, std of the initial B, projector trains?)
ALL_TARGETS = ("q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj")
LORA_CASES = {
    "vqa_lora_all_right": (dict(hidden=64, ff=128, layers=2, heads=2, image=64, patch=16),
                           dict(vocab=512, hidden=64, ff=128, layers=4, hq=4, hkv=2, hd=16, window=8,
                                q_scalar=16.0, rope_factor_global=8.0), 5, 2, 2, "right", 2, 2, 0.25,
                           (8, 16.0, ALL_TARGETS, 0.02, True)),
    "vqa_lora_qvd_left_frozen_proj": (dict(hidden=64, ff=128, layers=1, heads=4, image=48, patch=16),
                                      dict(vocab=304, hidden=48, ff=96, layers=3, hq=4, hkv=1, hd=32, window=5,
                                           q_scalar=32.0), 4, 2, 2, "left", 1, 2, 0.0,
                                      (16, 32.0, ("q_proj", "v_proj", "down_proj"), 0.05, False)),
}


# --------------------------------------------------------------------------- LoRA adapters
def lora_target_rows(cfg: O.GemmaCfg):
    """target module -> (packed weight of the Stage-1 oracle, first row, rows, input width)."""
    qd, kd = cfg.hq * cfg.hd, cfg.hkv * cfg.hd
    return {"q_proj": ("qkv_w", 0, qd, cfg.hidden), "k_proj": ("qkv_w", qd, kd, cfg.hidden),
            "v_proj": ("qkv_w", qd + kd, kd, cfg.hidden), "o_proj": ("o_w", 0, cfg.hidden, qd),
            "gate_proj": ("gu_w", 0, cfg.ff, cfg.hidden), "up_proj": ("gu_w", cfg.ff, cfg.ff, cfg.hidden),
            "down_proj": ("down_w", 0, cfg.hidden, cfg.ff)}


def init_lora(cfg: O.GemmaCfg, r: int, targets, seed: int = 7, b_std: float = 0.0) -> dict:
    """{"l<i>.<target>.A": [r, in], "l<i>.<target>.B": [out, r]}: A as peft initialises it (kaiming_uniform_ with
    a = sqrt(5), i.e. U(+-1/sqrt(in))), B zero as in peft or N(0, b_std) so that a fixture exercises every gradient
    path from the first step (with B = 0 the first dA is identically zero)."""
    g = torch.Generator().manual_seed(seed)
    rows = lora_target_rows(cfg)
    out = {}
    for i in range(cfg.layers):
        for t in targets:
            _, _, n, in_dim = rows[t]
            out[f"l{i}.{t}.A"] = (torch.rand(r, in_dim, generator=g) * 2 - 1) / in_dim ** 0.5
            out[f"l{i}.{t}.B"] = torch.randn(n, r, generator=g) * b_std
    return out


def lora_merge(Wl: dict, lora: dict, cfg: O.GemmaCfg, scale: float) -> dict:
    """peft `Linear.forward`: result = base(x) + lora_B(lora_A(dropout(x))) * scaling, dropout p = 0, written on the
    weights: W + scale * B A for every adapted target (the same function of A and B, so autograd through this gives
    the adapter gradients).  ref:Stage2/train_vqa_stage2.py:225-250 builds the adapters."""
    rows = lora_target_rows(cfg)
    W = dict(Wl)
    for i in range(cfg.layers):
        by_w: dict = {}
        for t, (wname, r0, n, _) in rows.items():
            if f"l{i}.{t}.A" in lora:
                by_w.setdefault(wname, []).append((r0, n, lora[f"l{i}.{t}.B"] @ lora[f"l{i}.{t}.A"]))
        for wname, parts in by_w.items():
            base = Wl[f"l{i}.{wname}"]
            zero = torch.zeros_like(base)
            pieces, at = [], 0
            for r0, n, d in sorted(parts, key=lambda q: q[0]):
                if r0 > at:
                    pieces.append(zero[at:r0])
                pieces.append(scale * d)
                at = r0 + n
            if at < base.shape[0]:
                pieces.append(zero[at:])
            W[f"l{i}.{wname}"] = base + torch.cat(pieces, dim=0)
    return W


# --------------------------------------------------------------------------- collate (integer, bit-exact)
def vqa_collate(questions: list, answers: list, pad_id: int, padding_side: str = "right"):
    """ref:Stage2/trainer.py:18-61 — pad every question to the batch's longest question and every
    answer to the longest answer, pads on `padding_side`.  Lists of 1-D int64 arrays in, two int64
    matrices out."""
    def pad(seqs):
        L = max(len(s) for s in seqs)
        out = np.full((len(seqs), L), pad_id, dtype=np.int64)
        for i, s in enumerate(seqs):
            s = np.asarray(s, dtype=np.int64)
            if padding_side == "left":
                out[i, L - len(s):] = s
            else:
                out[i, :len(s)] = s
        return out
    return pad(questions), pad(answers)


# --------------------------------------------------------------------------- splice (integer, bit-exact)
def vqa_masks_labels(q_ids: np.ndarray, a_ids: np.ndarray, n_img: int, pad_id: int):
    """ref:Stage2/trainer.py:370-395 — attention mask 1 for image tokens, (id != pad) for question and
    answer; labels -100 for image and question, the answer ids with pads replaced by -100."""
    B = q_ids.shape[0]
    attn = np.concatenate([np.ones((B, n_img), dtype=np.int64), (q_ids != pad_id).astype(np.int64),
                           (a_ids != pad_id).astype(np.int64)], axis=1)
    a_lab = a_ids.copy()
    a_lab[a_lab == pad_id] = -100
    labels = np.concatenate([np.full((B, n_img), -100, dtype=np.int64),
                             np.full(q_ids.shape, -100, dtype=np.int64), a_lab], axis=1)
    return attn, labels


def vqa_forward(pixels, q_ids, a_ids, Wv, P, Wl, vcfg: O.SiglipCfg, lcfg: O.GemmaCfg, scale_dtype=None,
                lora: dict | None = None, lora_scale: float = 1.0):
    """ref:Stage2/trainer.py:311-416 — frozen SigLIP, `[:, 1:, :]`, projector, Gemma3 embedding * scale
    for question and answer, concatenation, frozen Gemma3, fp32 shifted CE.  Returns (loss tensor,
    dict of what the reference hands to the language model)."""
    with torch.no_grad():
        patch = O.siglip_forward(pixels, Wv, vcfg)[:, 1:, :]
        q_emb = O.embed_tokens(q_ids, Wl["embed"], lcfg.hidden, scale_dtype, lcfg.scale_embed)
        a_emb = O.embed_tokens(a_ids, Wl["embed"], lcfg.hidden, scale_dtype, lcfg.scale_embed)
    proj = O.projector_forward(patch, P)
    X0 = torch.cat([proj, q_emb.to(proj.dtype), a_emb.to(proj.dtype)], dim=1)
    attn, labels = vqa_masks_labels(q_ids.numpy(), a_ids.numpy(), proj.shape[1], lcfg.pad_id)
    if lora is not None:
        Wl = lora_merge(Wl, lora, lcfg, lora_scale)
    hidden = O.gemma3_hidden(X0, torch.from_numpy(attn), Wl, lcfg)
    loss, n_valid = O.causal_lm_loss(hidden, Wl, lcfg, labels)
    return loss, dict(spliced=X0, attention_mask=attn, lm_labels=labels, n_valid=n_valid,
                      projected=proj, patch=patch)


def vqa_micro_step(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, loss_divisor: float = 1.0,
                   scale_dtype=None, lora: dict | None = None, lora_scale: float = 1.0, train_projector: bool = True):
    """One micro-batch forward+backward: the loss as logged (undivided) and the projector (and adapter)
    gradients of loss / loss_divisor (ref:Stage2/trainer.py:418-420)."""
    P = {k: v.detach().clone().requires_grad_(train_projector) for k, v in P.items()}
    if lora is not None:
        lora = {k: v.detach().clone().requires_grad_(True) for k, v in lora.items()}
    loss, aux = vqa_forward(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, scale_dtype, lora, lora_scale)
    (loss / loss_divisor).backward()
    out = {"loss": float(loss.detach()),
           "grads": {k: v.grad.detach() for k, v in P.items()} if train_projector else {}}
    if lora is not None:
        out["lora_grads"] = {k: v.grad.detach() for k, v in lora.items()}
    out.update({k: (v.detach() if torch.is_tensor(v) else v) for k, v in aux.items()})
    return out


@torch.no_grad()
def vqa_eval_loss(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, scale_dtype=None, lora=None, lora_scale=1.0) -> float:
    """ref:Stage2/trainer.py:517-590 — the validation loss of one batch (same arithmetic, no backward)."""
    loss, _ = vqa_forward(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, scale_dtype, lora, lora_scale)
    return float(loss)


# --------------------------------------------------------------------------- accumulation schedule
def sync_flags(n_batches: int, grad_accum: int) -> list:
    """accelerate's `Accelerator.accumulate()` as ref:Stage2/trainer.py:306 uses it: gradients
    synchronise (and the optimizer steps) on every `grad_accum`-th micro-batch and on the last batch
    of the loader, where the micro-batch counter restarts."""
    flags, step = [], 0
    for i in range(n_batches):
        if i == n_batches - 1:
            step = 0
            flags.append(True)
        else:
            step += 1
            flags.append(step % grad_accum == 0)
    return flags


# --------------------------------------------------------------------------- synthetic VQA samples
def synthetic_vqa_samples(n: int, vcfg: O.SiglipCfg, lcfg: O.GemmaCfg, seed: int = 4321,
                          q_len=(2, 7), a_len=(3, 10)):
    """Ragged (image, question, answer) triplets with the schema of ref:Stage2/dataset.py:116-120:
    question ids without special tokens, answer ids starting with BOS(2); no id equals pad."""
    g = torch.Generator().manual_seed(seed)
    pixels = torch.rand(n, 3, vcfg.image, vcfg.image, generator=g) * 2 - 1
    qs, as_ = [], []
    for _ in range(n):
        lq = int(torch.randint(q_len[0], q_len[1] + 1, (1,), generator=g))
        la = int(torch.randint(a_len[0], a_len[1] + 1, (1,), generator=g))
        qs.append(torch.randint(3, lcfg.vocab, (lq,), generator=g))
        a = torch.randint(3, lcfg.vocab, (la,), generator=g)
        a[0] = 2
        as_.append(a)
    return pixels, qs, as_
