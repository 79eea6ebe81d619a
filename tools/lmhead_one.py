"""GPU probe / ncu target: the LM-head logits GEMM with the cross-entropy statistics in its epilogue
(PT_ACT_CE_STATS; M = 32*128 label rows, N = 262144, K = 1152) next to the plain logits GEMM, CUDA events."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
M, V, h = 4096, 262144, 1152
x = torch.randn(M, h, device=dev).to(torch.bfloat16)
w = (torch.randn(V, h, device=dev) * 0.03).to(torch.bfloat16)
logits = torch.empty(M, V, device=dev, dtype=torch.bfloat16)
stats = torch.empty(ops.ce_stats_parts(V), M, 2, device=dev)
for name, fn in (("plain logits GEMM", lambda: ops.linear(x, w, out=logits)),
                 ("logits GEMM + CE statistics epilogue", lambda: ops.linear_ce_stats(x, w, logits, stats))):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(5): fn()
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 5 * 1e3
    print(f"{name} M={M} N={V} K={h}: {us:.1f} us  {2*M*V*h/us/1e6:.0f} TFLOP/s")
# ncu --set full --clock-control none: -k regex:gemm2 -s 12 -c 1 python tools/lmhead_one.py   (the 13th gemm2 launch = CE-stats variant)
