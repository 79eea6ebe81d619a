"""GPU probe / ncu target: the dominant GEMM of the step — Gemma3 gate||up projection with the fused GeGLU
epilogue (M = 32*703, N = 2*6912, K = 1152) — timed with CUDA events."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
M, ff, h = 32 * 703, 6912, 1152
x = torch.randn(M, h, device=dev).to(torch.bfloat16)
w = (torch.randn(2 * ff, h, device=dev) * 0.03).to(torch.bfloat16)
gu = torch.empty(M, 2 * ff, device=dev, dtype=torch.bfloat16)
act = torch.empty(M, ff, device=dev, dtype=torch.bfloat16)
for _ in range(3): ops.linear_geglu(x, w, gu, act)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): ops.linear_geglu(x, w, gu, act)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 10 * 1e3
print(f"linear_geglu M={M} N={2*ff} K={h}: {us:.1f} us  {2*M*2*ff*h/us/1e6:.0f} TFLOP/s")
