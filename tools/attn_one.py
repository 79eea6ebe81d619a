"""GPU probe / ncu target: the Gemma3 attention kernels at config 3 (B=32, S=703, 4:1 heads x 256, window 512):
single-pass forward, CTA-pair backward, dQ GEMM — a few launches each."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
B, S, hq, hkv, hd, W = 32, 703, 4, 1, 256, 512
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
dQ = torch.empty(B * S, hq * hd, dtype=torch.bfloat16, device=dev)
for _ in range(4):
    ops.attn_gemma_fwd(q, k, qkv, None, lse, O, B, S, hq, hkv, hd, W, 1 / 16)
    ops.attn_gemma_bwd_pair(q, k, qkv, dO, None, lse, delta, dS, dK, dV, B, S, hq, hkv, hd, W, 1 / 16, ld)
    ops.attn_dq_from_dst(dS, k, dQ, B, S, hq, hkv, hd, W, ld)
torch.cuda.synchronize()
print("ok")
