"""GPU probe: the "existing Blackwell library kernels" bar of SURVEY.md §8(d) — the reference's Stage-1 step executed
the way the reference executes it (HF transformers modules, PyTorch autograd, cuBLASLt GEMMs, SDPA attention, bf16
autocast with the frozen towers held in bf16, fp32 projector + torch.optim.AdamW) on the same B200, same config as
bench.py (BASELINE.json configs[2]: SigLIP-L/16-384 + MLP(1024->10240->1152) + Gemma3-1B, B=32, T=128, S=703), random
init, synthetic data.  The loop body restates ref:Stage1/projector_trainer.py:157-245 (the reference itself is not on
the GPU box); none of this repository's kernels run here.  Not a bench.py arm: a number to put beside it."""
import argparse
import json
import sys
from pathlib import Path

import torch
import torch.nn as nn

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--tokens", type=int, default=128)
ap.add_argument("--steps", type=int, default=5)
ap.add_argument("--warmup", type=int, default=3)
a = ap.parse_args()

from transformers import Gemma3ForCausalLM, Gemma3TextConfig, SiglipVisionConfig, SiglipVisionModel

dev = torch.device("cuda:0")
torch.manual_seed(0)
vc = SiglipVisionConfig(hidden_size=1024, intermediate_size=4096, num_hidden_layers=24, num_attention_heads=16,
                        image_size=384, patch_size=16)
lc = Gemma3TextConfig(vocab_size=262144, hidden_size=1152, intermediate_size=6912, num_hidden_layers=26,
                      num_attention_heads=4, num_key_value_heads=1, head_dim=256, sliding_window=512,
                      query_pre_attn_scalar=256, max_position_embeddings=32768, pad_token_id=0)
with torch.device(dev):
    vision = SiglipVisionModel(vc).to(torch.bfloat16).eval()
    llm = Gemma3ForCausalLM(lc).to(torch.bfloat16).eval()
    proj = nn.Sequential(nn.Linear(1024, 10240), nn.GELU(), nn.Linear(10240, 1152)).float()
for m in (vision, llm):
    m.requires_grad_(False)
opt = torch.optim.AdamW(proj.parameters(), lr=1e-4, weight_decay=0.01)
B, T = a.batch, a.tokens
g = torch.Generator(device=dev).manual_seed(1234)
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, lc.vocab_size, (B, T), generator=g, device=dev)
ids[:, 0] = 2
labels = ids.clone()
embed = llm.get_input_embeddings()


def step():
    with torch.autocast("cuda", dtype=torch.bfloat16):
        with torch.no_grad():
            patch = vision.vision_model(pixel_values=px.to(torch.bfloat16)).last_hidden_state[:, 1:, :]
        projected = proj(patch).float()                        # accelerate returns fp32 from the prepared projector
        with torch.no_grad():
            text = embed(ids)
        inputs = torch.cat([projected, text.to(projected.dtype)], dim=1)
        n_img = projected.shape[1]
        mask = torch.cat([torch.ones(B, n_img, dtype=torch.long, device=dev), (ids != 0).long()], dim=1)
        lm_labels = torch.cat([torch.full((B, n_img), -100, dtype=torch.long, device=dev), labels], dim=1)
        out = llm(inputs_embeds=inputs, attention_mask=mask, labels=lm_labels, return_dict=True)
        loss = out.loss
    (loss / 2).backward()
    torch.nn.utils.clip_grad_norm_(proj.parameters(), 5.0)
    opt.step()
    opt.zero_grad()
    return loss


for _ in range(a.warmup):
    loss = step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.steps):
    loss = step()
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.steps
import transformers
print(json.dumps({"config": f"HF transformers {transformers.__version__} + torch {torch.__version__} (cuBLASLt, "
                  f"{getattr(llm.config, '_attn_implementation', '?')} attention, bf16 autocast), Stage-1 step, SigLIP-L + MLP "
                  f"+ Gemma3-1B, B={B}, T={T}, random init", "ms_per_step": ms, "images_per_s": B / ms * 1e3,
                  "loss": float(loss.item()), "mem_GB": torch.cuda.max_memory_allocated() / 2 ** 30}), flush=True)
