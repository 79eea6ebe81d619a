"""GPU probe: BASELINE.json configs[1] (SigLIP-L forward only, B=128) and configs[4] (prefill, 575 image
tokens + 64-token prompt, Gemma3-1B, B=64) through the engine; CUDA-event timing, JSON lines."""
import sys, json, argparse
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine
from projectiontrainer_b200.flops import stage1_flops_per_image

ap = argparse.ArgumentParser()
ap.add_argument("--iters", type=int, default=5)
ap.add_argument("--decode", type=int, default=24)
a = ap.parse_args()
dev = torch.device("cuda:0")
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev, seed=0)
g = torch.Generator(device=dev).manual_seed(1)
fl = stage1_flops_per_image(SIGLIP_L, GEMMA3_1B, 128)

def timed(fn, iters):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters

B = 128
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ms = timed(lambda: eng.vision_forward(px, drop_first=False), a.iters)
print(json.dumps({"config": "configs[1] SigLIP-L/16-384 forward, bf16, B=128", "ms": ms, "images_per_s": B / ms * 1e3,
                  "tflops": B * fl["vision"] / ms / 1e9}), flush=True)
B, Tp = 64, 64
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, GEMMA3_1B.vocab, (B, Tp), generator=g, device=dev)
ms = timed(lambda: eng.prefill(px, ids, max_new_tokens=a.decode), a.iters)
print(json.dumps({"config": "configs[4] prefill 575 image + 64 prompt tokens, Gemma3-1B, B=64 (vision + projector + LLM, "
                  "LM head on last position)", "ms": ms, "images_per_s": B / ms * 1e3}), flush=True)
lg = eng.prefill(px, ids, max_new_tokens=a.decode)
nxt = lg.argmax(-1)
for _ in range(3):                      # eager warm-up, graph capture, first replay
    nxt = eng.decode_step(nxt).argmax(-1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.decode - 4):
    nxt = eng.decode_step(nxt).argmax(-1)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / (a.decode - 4)
print(json.dumps({"config": "cached decode step after that prefill, B=64 (one CUDA-graph replay per token)", "ms_per_token_step": ms,
                  "tokens_per_s": B / ms * 1e3}), flush=True)

# ---- BASELINE.json configs[3]: Gemma3-4B text decoder (h 2560, ff 10240, L 34, 8/4 heads x 256, window 1024,
# linear-scaled global RoPE), MLP(1024 -> 10240 -> 2560), T=256, B=16 per GPU: one full training step
del eng
torch.cuda.empty_cache()
from projectiontrainer_b200.config import GEMMA3_4B
eng = build_engine(SIGLIP_L, GEMMA3_4B, dev, seed=0)
B, T = 16, 256
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, GEMMA3_4B.vocab, (B, T), generator=g, device=dev); ids[:, 0] = 2
batch = {"pixel_values": px, "token_ids": ids, "labels": ids.clone()}
losses = []
def step4():
    losses.append(eng.train_step(batch, 1e-4, pad_id=GEMMA3_4B.pad_id))
ms = timed(step4, 3)
fl4 = stage1_flops_per_image(SIGLIP_L, GEMMA3_4B, T)["train"]
print(json.dumps({"config": "configs[3] Stage-1 step SigLIP-L + MLP(1024->2560) + Gemma3-4B text, T=256, B=16, 1 GPU",
                  "ms_per_step": ms, "images_per_s": B / ms * 1e3, "step_tflops": B * fl4 / ms / 1e9,
                  "first_loss": float(losses[0]), "last_loss": float(losses[-1]),
                  "mem_gb": torch.cuda.max_memory_allocated() / 1e9}), flush=True)
