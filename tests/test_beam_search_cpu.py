"""Beam-search bookkeeping of `Stage1Engine.generate(num_beams=3)` pinned to HF `generate(num_beams=3)` token for
token (tests/golden/infer_*.npz, recorded by oracle/make_golden_infer.py from HF transformers on the reference's
inference inputs).  CPU only: the method runs on a stand-in engine whose prefill / decode step take their logits from
the fp32 CPU oracle, so only the control logic is under test (candidate ranking, finished pool, length penalty, early
stop heuristic, fill value, cache re-ordering); the decode arithmetic itself is covered by tests/test_infer_gpu.py."""
from pathlib import Path
from types import SimpleNamespace

import numpy as np
import pytest
import torch

from oracle import stage1_oracle as O
from oracle.make_golden_infer import CASES, LEFT_PADS

GOLDEN = Path(__file__).resolve().parent / "golden"


class _OracleEngine(SimpleNamespace):
    """prefill / decode_step of the engine, computed by the oracle; the token history of every row lives in the fake
    KV cache so that the beam re-ordering under test moves it."""

    def __init__(self, name):
        vkw, lkw, B, Tp, new = CASES[name]
        self.vc, self.lc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
        self.Wv, self.Wl, self.P = O.init_siglip(self.vc), O.init_gemma(self.lc), O.init_projector(self.vc.hidden, self.lc.hidden)
        px, ids, _ = O.synthetic_batch(B, max(Tp, 1), self.vc, self.lc, left_pad=False)
        self.pixels, self.prompt, self.new = px, (ids[:, :Tp] if Tp else None), new
        self.attn = None
        if name in LEFT_PADS:
            for b, k in enumerate(LEFT_PADS[name]):
                self.prompt[b, :k] = self.lc.pad_id
        self.device, self.lora = torch.device("cpu"), None

    def _logits(self):
        hist, n = self.cache["k"][0], self.cache["n"]
        X = self.X0
        if n:
            X = torch.cat([X, O.embed_tokens(hist[:, :n].long(), self.Wl["embed"], self.lc.hidden)], dim=1)
        attn = None
        if self.attn0 is not None:
            attn = torch.cat([self.attn0, torch.ones(X.shape[0], n, dtype=torch.long)], dim=1)
        return O.last_position_logits(X, self.Wl, self.lc, attn)

    def prefill(self, pixels, prompt_ids, max_new_tokens=0, prompt_pad_id=None):
        self.X0 = O.prefill_inputs(pixels, prompt_ids, self.Wv, self.P, self.Wl, self.vc, self.lc)
        self.attn0 = None
        if prompt_pad_id is not None and prompt_ids is not None:
            n_img = self.X0.shape[1] - prompt_ids.shape[1]
            self.attn0 = torch.cat([torch.ones(pixels.shape[0], n_img, dtype=torch.long), (prompt_ids != prompt_pad_id).long()], 1)
        R = pixels.shape[0]
        k_all, v_all = torch.zeros(1, R, max_new_tokens + 1), torch.zeros(1, R, 1)      # [layers, rows, ...] like the engine's
        self.cache = {"k": list(k_all.unbind(0)), "v": list(v_all.unbind(0)), "k_all": k_all, "v_all": v_all, "n": 0}
        return self._logits()

    def decode_step(self, tokens):
        self.cache["k"][0][:, self.cache["n"]] = tokens.float()
        self.cache["n"] += 1
        return self._logits()


def _beam(eng, eos, **kw):
    from projectiontrainer_b200.engine import Stage1Engine
    args = dict(max_new_tokens=eng.new, do_sample=False, temperature=1.0, top_p=1.0, top_k=0, pad_token_id=eng.lc.pad_id,
                eos_token_id=eos, generator=None, prompt_pad_id=None, num_beams=3, length_penalty=1.0)
    args.update(kw)
    return Stage1Engine._beam_generate(eng, eng.pixels, eng.prompt, **args)


@pytest.mark.parametrize("name", list(CASES))
def test_beam_search_matches_hf_generate(name):
    g = np.load(GOLDEN / f"{name}.npz")
    eng = _OracleEngine(name)
    pad = eng.lc.pad_id if name in LEFT_PADS else None
    out = _beam(eng, 1, prompt_pad_id=pad)
    assert np.array_equal(out.numpy(), g["beam3"]), (out.tolist(), g["beam3"].tolist())
    for k, e in enumerate(g["beam3_eos_ids"].tolist()):          # hypotheses that finish at different lengths
        out = _beam(eng, int(e), prompt_pad_id=pad)
        ref = g[f"beam3_eos{k}"]
        assert np.array_equal(out.numpy(), ref), (k, out.tolist(), ref.tolist())


def test_beam_sampling_draws_valid_sequences():
    """num_beams=3 with do_sample=True (the Stage-2 validation's call): candidates are drawn without replacement from
    the warped distribution; outputs are well-formed and reproducible under a seeded generator."""
    eng = _OracleEngine("infer_tiny_prompt_mqa")
    kw = dict(do_sample=True, top_p=0.9, top_k=50)
    a = _beam(eng, 1, generator=torch.Generator().manual_seed(3), **kw)
    b = _beam(eng, 1, generator=torch.Generator().manual_seed(3), **kw)
    c = _beam(eng, 1, generator=torch.Generator().manual_seed(4), **kw)
    assert torch.equal(a, b) and a.shape[0] == eng.pixels.shape[0] and 1 <= a.shape[1] <= eng.new
    assert a.min() >= 0 and a.max() < eng.lc.vocab and (a.shape != c.shape or not torch.equal(a, c))


def test_top_p_on_the_top_k_candidates_equals_hf_warpers_on_the_whole_row():
    """Beam sampling applies HF's TopKLogitsWarper then TopPLogitsWarper (min_tokens_to_keep = 2).  After top-k the
    finite entries of a row are the k candidates, so top-p is evaluated on them alone (a threshold value) instead of
    through a sort of the whole vocabulary: the surviving set must be the one HF's formulation leaves."""
    g = torch.Generator().manual_seed(0)
    for V, k, top_p, temp in ((5000, 50, 0.9, 1.0), (5000, 50, 0.5, 0.7), (300, 2, 0.9, 1.0), (40, 50, 0.3, 1.3)):
        lp = torch.log_softmax(torch.randn(7, V, generator=g) * 3, dim=-1) / temp
        # HF's formulation (the previous implementation)
        kth = lp.topk(min(max(k, 2), V), dim=1).values[:, -1:]
        ref = lp.masked_fill(lp < kth, float("-inf"))
        srt, idx = ref.sort(dim=1, descending=False)
        rm = srt.softmax(dim=1).cumsum(dim=1) <= (1.0 - top_p)
        rm[:, -2:] = False
        ref = ref.masked_fill(rm.scatter(1, idx, rm), float("-inf"))
        # the engine's: warp the top-k candidates only, nothing of size [rows, V] after the top-k
        vals, tix = lp.topk(min(max(k, 2), V), dim=1)
        asc = vals.flip(1)
        rm2 = asc.softmax(dim=1).cumsum(dim=1) <= (1.0 - top_p)
        rm2[:, -2:] = False
        vals = vals.masked_fill(rm2.flip(1), float("-inf"))
        new = torch.full_like(lp, float("-inf")).scatter(1, tix, vals)
        assert torch.equal(torch.isfinite(ref), torch.isfinite(new)), (V, k, top_p)
        assert torch.equal(ref[torch.isfinite(ref)], new[torch.isfinite(new)])
        assert int(torch.isfinite(new).sum(1).min()) >= 2


@pytest.mark.parametrize("name", list(CASES))
def test_repetition_penalty_matches_hf_generate(name):
    """The Stage-2 inference scripts' call: 3 beams with repetition_penalty 1.8 and length_penalty 1.2
    (ref:Stage2/inference_vqa_stage2.py:180-195), and greedy decoding with the penalty — HF applies the penalty to the
    log-probabilities of every beam over the tokens that beam has generated."""
    from projectiontrainer_b200.engine import Stage1Engine
    g = np.load(GOLDEN / f"{name}.npz")
    eng = _OracleEngine(name)
    pad = eng.lc.pad_id if name in LEFT_PADS else None
    out = _beam(eng, int(g["beam3_eos_ids"][1]), prompt_pad_id=pad, repetition_penalty=1.8, length_penalty=1.2)
    assert np.array_equal(out.numpy(), g["beam3_rep"]), (out.tolist(), g["beam3_rep"].tolist())
    eng.fuse_sampler = False                   # the torch restatement of the token choice (the kernel needs a GPU)
    out = Stage1Engine.generate(eng, eng.pixels, eng.prompt, max_new_tokens=eng.new, do_sample=False,
                                repetition_penalty=1.8, pad_token_id=eng.lc.pad_id, eos_token_id=1, prompt_pad_id=pad)
    assert np.array_equal(out.numpy(), g["greedy_rep"]), (out.tolist(), g["greedy_rep"].tolist())
