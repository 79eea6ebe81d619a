"""GPU probe: SM-clock timeline of cluster 0 of the CTA-pair attention backward (pt_attn_gemma_bwd_pair_set_trace).
Per block and CTA rank: 0 producer: E buffer free  1 producer: L buffer free  2 MMA: issue scores  3 MMA: X tile ready
4 MMA: L landed -> grad  5 MMA: grad issued  6 copy: X ready  7 copy: peer buffer free -> push  8 elementwise: scores ready
9 (CTA 1) P^T received  10 result computed  11 X buffer free  12 X handed over  13 item blocks done  14 accumulator stored"""
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
run = lambda: ops.attn_gemma_bwd_pair(q, k, qkv, dO, None, lse, delta, dS, dK, dV, B, S, hq, hkv, hd, W, 1 / 16, ld)
for _ in range(3):
    run()
iters = 48
tr = torch.zeros(iters * 32, dtype=torch.int64, device=dev)
_lib.lib().pt_attn_gemma_bwd_pair_set_trace(C.c_void_p(tr.data_ptr()), iters)
run()
torch.cuda.synchronize()
_lib.lib().pt_attn_gemma_bwd_pair_set_trace(C.c_void_p(0), 0)
t = tr.view(iters, 2, 16).cpu()
t0 = int(t[t > 0].min())
names = {0: "E-free", 1: "L-free", 2: "scores>", 3: "X-ready", 4: "L-landed", 5: "grad>", 6: "cpy:X-ready", 7: "cpy:push", 8: "ew:T-ready",
         9: "ew:P-recv", 10: "ew:done", 11: "ew:X-free", 12: "ew:handed", 13: "item-end", 14: "acc-stored"}
# absolute timeline (ns since the first event) of blocks 20..25, both CTAs, sorted by time
ev = []
for i in range(20, min(26, iters)):
    for r in range(2):
        for e in range(15):
            v = int(t[i, r, e])
            if v > 0:
                ev.append((v - t0, f"cta{r} blk{i} {names[e]}"))
ev.sort()
base = ev[0][0] if ev else 0
for ts, what in ev:
    print(f"{ts - base:7d} ns  {what}")
