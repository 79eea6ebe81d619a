"""GPU probe: time the fused SigLIP attention kernels (B=32, N=576, H=16)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
B, N, H, hd = 32, 576, 16, 64
d = H * hd
qkv = torch.randn(B * N, 3 * d, device=dev).to(torch.bfloat16)
out = torch.empty(B * N, d, device=dev, dtype=torch.bfloat16)
for _v in (0,):
    for _ in range(3): ops.attn_siglip_fwd(qkv, out, B, N, H, hd, hd ** -0.5)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): ops.attn_siglip_fwd(qkv, out, B, N, H, hd, hd ** -0.5)
    e1.record(); torch.cuda.synchronize()
    us = e0.elapsed_time(e1) / 20 * 1e3
    print(f"attn_siglip2: {us:.1f} us  {4 * B * H * N * N * hd / us / 1e6:.0f} TFLOP/s", flush=True)
