#!/usr/bin/env python
"""Generates tests/golden/infer_*.npz: the reference's Stage-1 INFERENCE path run in this container.

TEST INFRASTRUCTURE ONLY (see oracle/stage1_oracle.py header).  Needs /root/reference and HF transformers.

The reference's `inference_stage1.py::main()` loads checkpoints from the hub (no network here), so the
generator executes its inference block (ref:Stage1/inference_stage1.py:150-193) call for call on modules
built from explicit configs: `vision_encoder.vision_model(pixel_values=...)` -> `[:, 1:, :]` -> the
reference's own `MLPProjector` (imported from /root/reference/Stage1/projectors.py) -> all-ones mask ->
`llm_model.generate(inputs_embeds=..., attention_mask=..., max_new_tokens=..., do_sample=False, ...)`
(HF's real generation loop with its KV cache, sliding-window cache layers and stopping criteria), CPU fp32.
Recorded: the generated ids as `generate` returns them (new tokens only, pad after eos), the last-position
logits of the prefill (what the first token is chosen from) and, for the prompt variant of BASELINE.json
configs[4] (image tokens + a text prompt), the same with prompt embeddings appended.

    python oracle/make_golden_infer.py
"""
from __future__ import annotations

import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference/Stage1")

from oracle import stage1_oracle as O  # noqa: E402
from oracle.make_golden import build_hf_gemma, build_hf_siglip  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"

CASES = {
    # name: (siglip kwargs, gemma kwargs, B, prompt tokens, max_new_tokens)
    "infer_tiny_gqa": (dict(hidden=64, ff=128, layers=2, heads=2, image=64, patch=16),
                       dict(vocab=512, hidden=64, ff=128, layers=6, hq=4, hkv=2, hd=16, window=8,
                            q_scalar=16.0, rope_factor_global=8.0), 3, 0, 12),
    "infer_tiny_prompt_mqa": (dict(hidden=64, ff=128, layers=1, heads=4, image=48, patch=16),
                              dict(vocab=304, hidden=48, ff=96, layers=7, hq=4, hkv=1, hd=32, window=5,
                                   q_scalar=32.0), 2, 6, 10),
    # Stage-2 evaluation prompt (ref:Stage2/trainer.py:594-611): [image][left-padded question], mask 0 on pads
    "infer_tiny_leftpad_gqa": (dict(hidden=64, ff=128, layers=2, heads=2, image=64, patch=16),
                               dict(vocab=512, hidden=64, ff=128, layers=6, hq=4, hkv=2, hd=16, window=8,
                                    q_scalar=16.0, rope_factor_global=8.0), 3, 7, 10),
}
LEFT_PADS = {"infer_tiny_leftpad_gqa": [0, 2, 4]}


@torch.no_grad()
def run_case(name: str):
    from projectors import MLPProjector          # /root/reference/Stage1/projectors.py
    vkw, lkw, B, Tp, new = CASES[name]
    vc, lc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(vc), O.init_gemma(lc), O.init_projector(vc.hidden, lc.hidden)
    vision, llm = build_hf_siglip(vc, Wv), build_hf_gemma(lc, Wl)
    proj = MLPProjector(vc.hidden, lc.hidden)
    proj.load_state_dict(P)
    proj.eval()
    pixels, ids, _ = O.synthetic_batch(B, max(Tp, 1), vc, lc, left_pad=False)

    # ---- ref:Stage1/inference_stage1.py:152-187
    vision_outputs = vision.vision_model(pixel_values=pixels, return_dict=True)
    patch_embeddings = vision_outputs.last_hidden_state[:, 1:, :]
    projected = proj(patch_embeddings)
    inputs_embeds = projected
    if Tp:
        inputs_embeds = torch.cat([projected, llm.get_input_embeddings()(ids[:, :Tp])], dim=1)
    attention_mask = torch.ones(inputs_embeds.shape[:2], dtype=torch.long)
    if name in LEFT_PADS:
        for b, k in enumerate(LEFT_PADS[name]):
            ids[b, :k] = lc.pad_id
        q = ids[:, :Tp]
        # the Gemma3 branch of the reference indexes the embedding weight and applies embed_scale by hand
        # (ref:Stage2/trainer.py:543-549); same values as the embedding layer
        emb = llm.get_input_embeddings()
        inputs_embeds = torch.cat([projected, emb.weight[q] * emb.embed_scale.clone().detach()], dim=1)
        attention_mask = torch.cat([torch.ones(projected.shape[:2], dtype=torch.long), (q != lc.pad_id).long()], dim=1)
    prefill_logits = llm(inputs_embeds=inputs_embeds, attention_mask=attention_mask).logits[:, -1, :].float()
    outputs = llm.generate(inputs_embeds=inputs_embeds, attention_mask=attention_mask, max_new_tokens=new,
                           pad_token_id=lc.pad_id, eos_token_id=1, do_sample=False, temperature=None,
                           top_p=None, top_k=None, repetition_penalty=1.0)
    # forced-eos variant: make token `first` the eos id so stopping / padding behaviour is exercised
    first = int(outputs[0, 1]) if outputs.shape[1] > 1 else int(outputs[0, 0])
    outputs_eos = llm.generate(inputs_embeds=inputs_embeds, attention_mask=attention_mask, max_new_tokens=new,
                               pad_token_id=lc.pad_id, eos_token_id=first, do_sample=False, temperature=None,
                               top_p=None, top_k=None, repetition_penalty=1.0)
    print(f"[{name}] generated {tuple(outputs.shape)}: {outputs.tolist()}  | eos={first}: {outputs_eos.tolist()}")
    # beam search (the deterministic half of what Stage-2 validation asks for: ref:Stage2/trainer.py:603-611 calls
    # generate(num_beams=3, do_sample=True, ...); sampled outputs cannot be recorded, the beam bookkeeping can)
    beam = llm.generate(inputs_embeds=inputs_embeds, attention_mask=attention_mask, max_new_tokens=new,
                        pad_token_id=lc.pad_id, eos_token_id=1, do_sample=False, num_beams=3, temperature=None,
                        top_p=None, top_k=None, repetition_penalty=1.0)
    # eos variants: ids that occur early / late in the beam outputs, so that hypotheses finish at different lengths
    beam_eos_ids = [int(beam[0, 2]), int(beam[-1, min(5, beam.shape[1] - 1)])]
    beam_eos = [llm.generate(inputs_embeds=inputs_embeds, attention_mask=attention_mask, max_new_tokens=new,
                             pad_token_id=lc.pad_id, eos_token_id=e, do_sample=False, num_beams=3, temperature=None,
                             top_p=None, top_k=None, repetition_penalty=1.0) for e in beam_eos_ids]
    # the Stage-2 inference scripts' call (ref:Stage2/inference_vqa_stage2.py:180-195, defaults :316-319): beams with a
    # repetition penalty and a length penalty; also greedy with the penalty (ref:Stage1/inference_stage1.py:185)
    beam_rep = llm.generate(inputs_embeds=inputs_embeds, attention_mask=attention_mask, max_new_tokens=new,
                            pad_token_id=lc.pad_id, eos_token_id=beam_eos_ids[1], do_sample=False, num_beams=3,
                            temperature=None, top_p=None, top_k=None, repetition_penalty=1.8, length_penalty=1.2)
    greedy_rep = llm.generate(inputs_embeds=inputs_embeds, attention_mask=attention_mask, max_new_tokens=new,
                              pad_token_id=lc.pad_id, eos_token_id=1, do_sample=False, temperature=None,
                              top_p=None, top_k=None, repetition_penalty=1.8)
    print(f"[{name}] beam3 {beam.tolist()} | eos={beam_eos_ids}: {[b.tolist() for b in beam_eos]} | "
          f"rep 1.8 beams {beam_rep.tolist()} greedy {greedy_rep.tolist()}")
    np.savez_compressed(GOLDEN / f"{name}.npz", case=np.array(name), B=B, prompt_tokens=Tp, max_new_tokens=new,
                        prompt_ids=ids[:, :Tp].numpy(), prefill_logits=prefill_logits.numpy(),
                        generated=outputs.numpy(), eos_variant_id=first, generated_eos=outputs_eos.numpy(),
                        projected=projected.numpy(), beam3=beam.numpy(), beam3_eos_ids=np.array(beam_eos_ids),
                        beam3_eos0=beam_eos[0].numpy(), beam3_eos1=beam_eos[1].numpy(),
                        beam3_rep=beam_rep.numpy(), greedy_rep=greedy_rep.numpy())


if __name__ == "__main__":
    torch.set_num_threads(8)
    torch.manual_seed(0)
    for n in ([sys.argv[1]] if len(sys.argv) > 1 else CASES):
        run_case(n)
