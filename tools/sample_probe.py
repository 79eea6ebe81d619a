"""GPU probe: pt_sample_token alone (V = 262144), us per launch, greedy / sampling / sampling + repetition penalty."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
V = 262144
for B in (1, 8, 64):
    g = torch.Generator(device=dev).manual_seed(B)
    logits = torch.randn(B, V, generator=g, device=dev) * 2
    out = torch.zeros(B, 64, dtype=torch.int64, device=dev)
    nxt = torch.zeros(B, dtype=torch.int64, device=dev)
    u = torch.rand(B, generator=g, device=dev)
    done = torch.zeros(B, dtype=torch.uint8, device=dev)
    for name, kw in (("argmax", dict(top_k=0)), ("sample k=50 p=0.9", dict(top_k=50, top_p=0.9, temperature=0.7, u=u)),
                     ("sample + penalty", dict(top_k=50, top_p=0.9, temperature=0.7, u=u, prev=out, repetition_penalty=1.1))):
        f = lambda: ops.sample_token(logits, out=out, step=40, next_tokens=nxt, done=done, pad_id=0, eos_id=1, **kw)
        f(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(50):
            f()
        e1.record(); torch.cuda.synchronize()
        print(f"B={B:2d} {name:20s}: {e0.elapsed_time(e1) / 50 * 1e3:7.1f} us per launch")

# the same on real decode-step logits (random-init Gemma3-1B through the engine)
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
px = torch.rand(1, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, GEMMA3_1B.vocab, (1, 64), generator=g, device=dev)
lg = eng.prefill(px, ids, max_new_tokens=8)
print("logits: std", float(lg.std()), "max", float(lg.max()), "distinct values", int(lg.unique().numel()),
      "values >= 50th largest", int((lg >= lg.topk(50).values[0, -1]).sum()), "stride", lg.stride())
out = torch.zeros(1, 64, dtype=torch.int64, device=dev)
nxt = torch.zeros(1, dtype=torch.int64, device=dev)
u = torch.rand(1, generator=g, device=dev)
for name, kw in (("argmax", dict(top_k=0)), ("sample k=50 p=0.9", dict(top_k=50, top_p=0.9, temperature=0.7, u=u))):
    f = lambda: ops.sample_token(lg, out=out, step=3, next_tokens=nxt, **kw)
    f(); torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(50):
        f()
    e1.record(); torch.cuda.synchronize()
    print(f"engine logits {name:20s}: {e0.elapsed_time(e1) / 50 * 1e3:7.1f} us per launch")
