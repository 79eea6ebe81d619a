"""CPU oracle for the Stage-2 VQA step, frozen-LLM variant.  TEST INFRASTRUCTURE ONLY.

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


def vqa_forward(pixels, q_ids, a_ids, Wv, P, Wl, vcfg: O.SiglipCfg, lcfg: O.GemmaCfg, scale_dtype=None):
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
    hidden = O.gemma3_hidden(X0, torch.from_numpy(attn), Wl, lcfg)
    loss, n_valid = O.causal_lm_loss(hidden, Wl, lcfg, labels)
    return loss, dict(spliced=X0, attention_mask=attn, lm_labels=labels, n_valid=n_valid,
                      projected=proj, patch=patch)


def vqa_micro_step(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, loss_divisor: float = 1.0,
                   scale_dtype=None):
    """One micro-batch forward+backward: the loss as logged (undivided) and the projector gradients
    of loss / loss_divisor (ref:Stage2/trainer.py:418-420)."""
    P = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    loss, aux = vqa_forward(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, scale_dtype)
    (loss / loss_divisor).backward()
    out = {"loss": float(loss.detach()), "grads": {k: v.grad.detach() for k, v in P.items()}}
    out.update({k: (v.detach() if torch.is_tensor(v) else v) for k, v in aux.items()})
    return out


@torch.no_grad()
def vqa_eval_loss(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, scale_dtype=None) -> float:
    """ref:Stage2/trainer.py:517-590 — the validation loss of one batch (same arithmetic, no backward)."""
    loss, _ = vqa_forward(pixels, q_ids, a_ids, Wv, P, Wl, vcfg, lcfg, scale_dtype)
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
