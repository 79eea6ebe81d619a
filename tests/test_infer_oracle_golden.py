"""Pins the oracle's inference restatement (prefill logits + cache-free greedy decode) against vectors
recorded from HF `generate` driven exactly as ref:Stage1/inference_stage1.py:150-193 drives it
(oracle/make_golden_infer.py; KV cache, sliding-window layers and stopping criteria are HF's).  CPU only."""
from pathlib import Path

import numpy as np
import pytest
import torch

from oracle import stage1_oracle as O
from oracle.make_golden_infer import CASES, LEFT_PADS

GOLDEN = Path(__file__).resolve().parent / "golden"


def load(name):
    g = np.load(GOLDEN / f"{name}.npz")
    vkw, lkw, B, Tp, new = CASES[name]
    vc, lc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(vc), O.init_gemma(lc), O.init_projector(vc.hidden, lc.hidden)
    px, ids, _ = O.synthetic_batch(B, max(Tp, 1), vc, lc, left_pad=False)
    for b, k in enumerate(LEFT_PADS.get(name, [])):
        ids[b, :k] = lc.pad_id
    prompt = ids[:, :Tp] if Tp else None
    if Tp:
        assert np.array_equal(prompt.numpy(), g["prompt_ids"])
    return g, vc, lc, Wv, Wl, P, px, prompt, new


@pytest.mark.parametrize("name", list(CASES))
def test_prefill_logits_and_greedy_tokens_match_hf_generate(name):
    g, vc, lc, Wv, Wl, P, px, prompt, new = load(name)
    X0 = O.prefill_inputs(px, prompt, Wv, P, Wl, vc, lc)
    np.testing.assert_allclose(X0[:, :g["projected"].shape[1]].numpy(), g["projected"], rtol=0, atol=2e-5)
    padded = name in LEFT_PADS
    if padded:      # a plain forward keeps arange positions under a padded mask (F3); only generate moves them
        attn = torch.cat([torch.ones(X0.shape[0], X0.shape[1] - prompt.shape[1], dtype=torch.int64),
                          (prompt != lc.pad_id).long()], dim=1)
        hid = O.gemma3_hidden(X0, attn, Wl, lc)
        lg = (O.rms_norm(hid[:, -1, :], Wl["final_norm"], lc.eps) @ Wl["embed"].T).float()
    else:
        lg = O.last_position_logits(X0, Wl, lc)
    ref = g["prefill_logits"]
    assert np.abs(lg.numpy() - ref).max() / np.abs(ref).max() < 1e-4
    toks, _ = O.greedy_generate(px, prompt, Wv, P, Wl, vc, lc, new, eos_id=1, pad_id=lc.pad_id,
                                mask_prompt_pads=padded)
    assert toks.dtype == torch.int64
    assert np.array_equal(toks.numpy(), g["generated"])              # integer work: exact


@pytest.mark.parametrize("name", list(CASES))
def test_eos_stops_and_pads_like_hf(name):
    g, vc, lc, Wv, Wl, P, px, prompt, new = load(name)
    toks, _ = O.greedy_generate(px, prompt, Wv, P, Wl, vc, lc, new, eos_id=int(g["eos_variant_id"]),
                                pad_id=lc.pad_id, mask_prompt_pads=name in LEFT_PADS)
    assert np.array_equal(toks.numpy(), g["generated_eos"])
