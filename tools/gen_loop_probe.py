import sys, time
sys.path.insert(0, "/root/repo")
import torch
from projectiontrainer_b200 import ops
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine
dev = torch.device("cuda:0")
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
B = 1
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, GEMMA3_1B.vocab, (B, 64), generator=g, device=dev)
N = 60
for name, kw in (("greedy", dict(top_k=0)), ("sample", dict(top_k=50, top_p=0.9, temperature=0.7)), ("greedy", dict(top_k=0)), ("sample", dict(top_k=50, top_p=0.9, temperature=0.7))):
    logits = eng.prefill(px, ids, max_new_tokens=N + 4)
    out = torch.zeros(B, N + 4, dtype=torch.int64, device=dev)
    nxt = torch.zeros(B, dtype=torch.int64, device=dev)
    u = torch.rand(N + 4, B, generator=g, device=dev)
    done = torch.zeros(B, dtype=torch.uint8, device=dev)
    for step in range(3):
        ops.sample_token(logits, out=out, step=step, next_tokens=nxt, u=u[step] if kw["top_k"] else None, done=done, pad_id=0, eos_id=None, **kw)
        logits = eng.decode_step(nxt)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for step in range(3, N):
        ops.sample_token(logits, out=out, step=step, next_tokens=nxt, u=u[step] if kw["top_k"] else None, done=done, pad_id=0, eos_id=None, **kw)
        logits = eng.decode_step(nxt)
    t1 = time.perf_counter(); torch.cuda.synchronize(); t2 = time.perf_counter()
    print(f"{name}: cpu enqueue {(t1 - t0) / (N - 3) * 1e3:.3f} ms/step, total {(t2 - t0) / (N - 3) * 1e3:.3f} ms/step")
