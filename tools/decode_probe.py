"""GPU probe: where the cached decode step's time goes (BASELINE.json configs[4]: B = 64, 575 image + 64 prompt tokens,
Gemma3-1B).  Graph replay time per token, then one eager step with per-GEMM events."""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
from projectiontrainer_b200.config import GEMMA3_1B, GEMMA3_4B, QWEN3_8B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
LLM = {"gemma3_1b": GEMMA3_1B, "gemma3_4b": GEMMA3_4B, "qwen3_8b": QWEN3_8B}[sys.argv[2] if len(sys.argv) > 2 else "gemma3_1b"]
eng = build_engine(SIGLIP_L, LLM, dev)
g = torch.Generator(device=dev).manual_seed(1)
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, LLM.vocab, (B, 64), generator=g, device=dev)
nxt = eng.prefill(px, ids, max_new_tokens=40).argmax(-1)
for _ in range(3):
    nxt = eng.decode_step(nxt).argmax(-1)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(20):
    nxt = eng.decode_step(nxt).argmax(-1)
e1.record(); torch.cuda.synchronize()
print(f"{sys.argv[2] if len(sys.argv) > 2 else 'gemma3_1b'} B={B}: graph replay {e0.elapsed_time(e1) / 20:.3f} ms per token step")
ops.GEMM_PROFILE = []
e0.record()
eng._decode_step_fixed()
e1.record(); torch.cuda.synchronize()
prof, ops.GEMM_PROFILE = ops.GEMM_PROFILE, None
agg = {}
for fl, a, b, sh in prof:
    k = "x".join(map(str, sh[:4]))
    n, t = agg.get(k, (0, 0.0))
    agg[k] = (n + 1, t + a.elapsed_time(b))
tot = sum(t for _, t in agg.values())
print(f"eager step {e0.elapsed_time(e1):.3f} ms, GEMM launches {len(prof)}, GEMM time {tot:.3f} ms")
for k, (n, t) in sorted(agg.items(), key=lambda q: -q[1][1]):
    print(f"  M x N x K x batch {k:28s} n={n:3d}  {t / n * 1e3:8.1f} us each  {t:7.3f} ms")
