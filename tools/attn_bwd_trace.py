"""GPU probe: SM-clock timeline of one CTA of the fused Gemma3 attention backward (pt_attn_gemma_bwd_set_trace).
Prints, per query block of CTA 0, the cycle offsets of every hand-off:
  0 producer: dO buffer free     1 producer: Q buffer free
  2 MMA: dO landed -> issue dP   3 MMA: Q landed -> issue S     4 MMA: P ready -> dV^T    5 MMA: dS ready -> dK^T
  6 softmax: S ready             7 softmax: P handed over       8 softmax: dP ready       9 softmax: dS handed over
 10 softmax: item's blocks done  11 softmax: accumulators stored"""
import ctypes as C
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import _lib, ops
dev = torch.device("cuda:0")
B, S, hq, hkv, hd = 32, 703, 4, 1, 256
W = int(sys.argv[1]) if len(sys.argv) > 1 else 0
G, QKV, ld = hq // hkv, (hq + 2 * hkv) * hd, (S + 7) // 8 * 8
q = torch.randn(B * S, hq * hd, device=dev).to(torch.bfloat16)
k = torch.randn(B * S, hkv * hd, device=dev).to(torch.bfloat16)
qkv = torch.randn(B * S, QKV, device=dev).to(torch.bfloat16)
dO = torch.randn(B * S, hq * hd, device=dev).to(torch.bfloat16)
lse = torch.empty(B * hq, S, dtype=torch.float32, device=dev)
delta = torch.zeros(B * hq * S, dtype=torch.float32, device=dev)
O = torch.empty(B * S, hq * hd, dtype=torch.bfloat16, device=dev)
dS = torch.zeros(B * hq, S, ld, dtype=torch.bfloat16, device=dev)
dK = torch.empty(B * S, hkv * hd, dtype=torch.bfloat16, device=dev)
dV = torch.empty(B * S, hkv * hd, dtype=torch.bfloat16, device=dev)
ops.attn_gemma_fwd(q, k, qkv, None, lse, O, B, S, hq, hkv, hd, W, 1 / 16)
run = lambda: ops.attn_gemma_bwd(q, k, qkv, dO, None, lse, delta, dS, dK, dV, B, S, hq, hkv, hd, W, 1 / 16, ld)
for _ in range(3):
    run()
iters = 64
tr = torch.zeros(iters * 16, dtype=torch.int64, device=dev)
_lib.lib().pt_attn_gemma_bwd_set_trace(C.c_void_p(tr.data_ptr()), iters)
run()
torch.cuda.synchronize()
_lib.lib().pt_attn_gemma_bwd_set_trace(C.c_void_p(0), 0)
t = tr.view(iters, 16).cpu()
t0 = int(t[0][t[0] > 0].min())
prev = t0
for i in range(iters):
    row = t[i]
    if int(row.max()) == 0:
        break
    ev = {e: int(row[e]) - t0 for e in range(12) if int(row[e]) > 0}
    start = min(ev.values())
    print(f"blk {i:2d} +{start - (prev - t0):6d} | " + " ".join(f"{e}:{v - start:5d}" for e, v in sorted(ev.items())))
    prev = t0 + start
