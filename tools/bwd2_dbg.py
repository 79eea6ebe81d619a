"""bring-up probe for the CTA-pair attention backward: one small launch, reports errors (PT_BWD2_DBG selects what is skipped)"""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
B, S, hq, hkv, hd, W = 2, 200, 4, 1, 256, 0
G, QKV, ld = hq // hkv, (hq + 2 * hkv) * hd, (S + 7) // 8 * 8
q = torch.randn(B * S, hq * hd, device=dev).to(torch.bfloat16)
k = torch.randn(B * S, hkv * hd, device=dev).to(torch.bfloat16)
qkv = torch.randn(B * S, QKV, device=dev).to(torch.bfloat16)
dO = torch.randn(B * S, hq * hd, device=dev).to(torch.bfloat16)
lse = torch.zeros(B * hq, S, dtype=torch.float32, device=dev) + 20
delta = torch.zeros(B * hq * S, dtype=torch.float32, device=dev)
dS = torch.zeros(B * hq, S, ld, dtype=torch.bfloat16, device=dev)
dK = torch.zeros(B * S, hkv * hd, dtype=torch.bfloat16, device=dev)
dV = torch.zeros(B * S, hkv * hd, dtype=torch.bfloat16, device=dev)
try:
    ops.attn_gemma_bwd_pair(q, k, qkv, dO, None, lse, delta, dS, dK, dV, B, S, hq, hkv, hd, W, 1 / 16, ld)
    torch.cuda.synchronize()
    print("PT_BWD2_DBG", os.environ.get("PT_BWD2_DBG", "0"), "ok", float(dV.float().abs().sum()), float(dK.float().abs().sum()),
          float(dS.float().abs().sum()), flush=True)
except Exception as e:
    print("PT_BWD2_DBG", os.environ.get("PT_BWD2_DBG", "0"), "FAILED", str(e).splitlines()[0], flush=True)
