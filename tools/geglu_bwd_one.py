"""GPU probe / ncu target: the Gemma3 down-projection input gradient with the fused GeGLU-backward epilogue
(M = 32*703, N = ff = 6912, K = 1152; the epilogue reads the saved gate||up slab and writes d_gate||d_up),
timed with CUDA events and checked against the unfused pair (linear_dgrad + geglu_bwd)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
M, ff, h = 32 * 703, 6912, 1152
dy = torch.randn(M, h, device=dev).to(torch.bfloat16)
w = (torch.randn(h, ff, device=dev) * 0.03).to(torch.bfloat16)
gu = torch.randn(M, 2 * ff, device=dev).to(torch.bfloat16)
d_gu = torch.empty(M, 2 * ff, device=dev, dtype=torch.bfloat16)
for _ in range(3): ops.linear_dgrad_geglu_bwd(dy, w, gu, d_gu)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(10): ops.linear_dgrad_geglu_bwd(dy, w, gu, d_gu)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 10 * 1e3
print(f"linear_dgrad_geglu_bwd M={M} N={ff} K={h}: {us:.1f} us  {2*M*ff*h/us/1e6:.0f} TFLOP/s  "
      f"(epilogue traffic {M*ff*8/1e6:.0f} MB = {M*ff*8/us/1e3:.0f} GB/s)")
d_act = torch.empty(M, ff, device=dev, dtype=torch.bfloat16)
ref = torch.empty_like(d_gu)
ops.linear_dgrad(dy, w, out=d_act)
ops.geglu_bwd(d_act, gu, ref)
torch.cuda.synchronize()
err = (d_gu.float() - ref.float()).abs().max().item()
print(f"max |fused - unfused| = {err:.4g} (unfused rounds d_act to bf16 first; scale {ref.float().abs().max().item():.3g})")
