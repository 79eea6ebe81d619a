"""GPU probe: the Stage-1 step against a frozen Qwen3-8B-shaped decoder (the LLM of the reference's launch script,
run_projection_train_stage1.sh:12; bf16 weights, random init), SigLIP-L/16-384, B=8, T=128 -> S=703."""
import sys, json, argparse
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200.config import QWEN3_8B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=8)
ap.add_argument("--iters", type=int, default=4)
a = ap.parse_args()
dev = torch.device("cuda:0")
lc = QWEN3_8B
eng = build_engine(SIGLIP_L, lc, dev, seed=0)
g = torch.Generator(device=dev).manual_seed(1)
B, T = a.batch, 128
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, lc.vocab, (B, T), generator=g, device=dev)
ids[:, 0] = 2
batch = {"pixel_values": px, "token_ids": ids, "labels": ids.clone()}
losses = [eng.train_step(batch, 1e-4, pad_id=lc.pad_id).item() for _ in range(3)]      # eager, capture, replay
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.iters):
    loss = eng.train_step(batch, 1e-4, pad_id=lc.pad_id)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.iters
S = eng.n_img + T
M = B * S
# decoder GEMM flops: forward + input-gradient backward = 2 passes over every frozen matrix, 2 flops per MAC
per_tok = 2 * (lc.hidden * (lc.hq + 2 * lc.hkv) * lc.hd + lc.hq * lc.hd * lc.hidden + 3 * lc.hidden * lc.ff) * lc.layers
attn = 4 * S * lc.hq * lc.hd * lc.layers / 2            # QK^T + PV, causal half, per token
flops = 2 * M * (per_tok + 2.5 * attn)
print(json.dumps({"config": f"Stage-1 step, frozen Qwen3-8B-shaped decoder (36 layers, h 4096, ff 12288, GQA 32/8, hd 128), "
                  f"SigLIP-L, B={B}, T={T}, S={S}, bf16, CUDA-graph replay", "ms_per_step": ms,
                  "images_per_s": B / ms * 1e3, "decoder_tflops_est": flops / ms / 1e9, "losses": losses,
                  "last_loss": float(loss.item()), "mem_GB": torch.cuda.max_memory_allocated() / 2**30}), flush=True)
