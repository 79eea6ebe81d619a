"""GPU probe: time the fused Gemma3 attention forward at config 3 (B=32, S=703, hq=4, hkv=1, hd=256)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
B, S, hq, hkv, hd = 32, 703, 4, 1, 256
QKV = (hq + 2 * hkv) * hd
ld = 704
q = torch.randn(B * S, hq * hd, device=dev).to(torch.bfloat16)
k = torch.randn(B * S, hkv * hd, device=dev).to(torch.bfloat16)
qkv = torch.randn(B * S, QKV, device=dev).to(torch.bfloat16)
kv = torch.ones(B, S, dtype=torch.uint8, device=dev)
bits = torch.empty(B * 11, dtype=torch.int64, device=dev)
ops.pack_keyvalid(kv, bits)
P = torch.zeros(B * hq, S, ld, dtype=torch.bfloat16, device=dev)
O = torch.empty(B * S, hq * hd, dtype=torch.bfloat16, device=dev)
for W in (0, 512):
    for _ in range(3): ops.attn_gemma_fwd(q, k, qkv, bits, P, O, B, S, hq, hkv, hd, W, 1 / 16, ld)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): ops.attn_gemma_fwd(q, k, qkv, bits, P, O, B, S, hq, hkv, hd, W, 1 / 16, ld)
    e1.record(); torch.cuda.synchronize()
    print(f"window={W}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us", flush=True)
