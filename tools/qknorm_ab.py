"""GPU probe: q/k RMSNorm + RoPE in the q|k|v projection's epilogue (PT_ACT_QKNORM_ROPE) against the separate
pt_qknorm_rope_fwd launch — the Stage-1 training step (B = 32, T = 128) in ONE process, alternating, graphs recaptured
per setting; kernels per step; and the two GEMM variants alone."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import _lib, ops
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
B, T = 32, 128
bs = []
for s in range(2):
    ids = torch.randint(3, GEMMA3_1B.vocab, (B, T), generator=g, device=dev)
    bs.append({"pixel_values": torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1, "token_ids": ids,
               "labels": ids.clone()})


def train(on, n=10):
    eng.fuse_qknorm = bool(on)
    eng._graphs.clear()
    for i in range(4):
        loss = eng.train_step(bs[i % 2], 1e-4, max_norm=1.0)
    torch.cuda.synchronize()
    n0 = _lib.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        loss = eng.train_step(bs[i % 2], 1e-4, max_norm=1.0)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, float(loss)


for rep in range(3):
    for on in (0, 1):
        ms, loss = train(on)
        print(f"train B=32 fuse_qknorm={on}: {ms:.2f} ms per step (loss {loss:.4f})", flush=True)
# the two variants alone, M = 22496
M, K, hq, hkv, hd, S = 22496, 1152, 4, 1, 256, 703
x = torch.randn(M, K, generator=g, device=dev).bfloat16()
w = (torch.randn((hq + 2 * hkv) * hd, K, generator=g, device=dev) * K ** -0.5).bfloat16()
wq = torch.randn(hd, generator=g, device=dev) * 0.3
ang = torch.rand(S, hd // 2, generator=g, device=dev)
cos, sin = ang.cos().contiguous(), ang.sin().contiguous()
qkv = torch.empty(M, w.shape[0], device=dev, dtype=torch.bfloat16)
q = torch.empty(M, hq * hd, device=dev, dtype=torch.bfloat16)
k = torch.empty(M, hkv * hd, device=dev, dtype=torch.bfloat16)
gains, table = torch.cat([wq, wq]), torch.cat([cos, sin], 1).contiguous()
for name, f in (("linear + pt_qknorm_rope_fwd", lambda: (ops.linear(x, w, out=qkv), ops.qknorm_rope_fwd(qkv, wq, wq, cos, sin, q, k, S, hq, hkv, hd, 1e-6))),
                ("PT_ACT_QKNORM_ROPE", lambda: ops.linear_qknorm_rope(x, w, qkv, q, k, gains, table, hq, 1e-6))):
    for _ in range(3):
        f()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20):
        f()
    e1.record(); torch.cuda.synchronize()
    print(f"{name:30s}: {e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
