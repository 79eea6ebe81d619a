"""GPU probe: LM-head input gradient d_yf[R,h] = dlogits[R,V] @ E[V,h] (R=4096, V=262144, h=1152) — the one launch
of the step on the 128-wide tile — against a split-K form expressed through the batch dims of pt_gemm_bf16
(f32 partials [S,R,h], 256-wide tiles), timed with CUDA events."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
R, V, h = 4096, 262144, 1152
dl = (torch.randn(R, V, device=dev) * 0.01).to(torch.bfloat16)
E = (torch.randn(V, h, device=dev) * 0.03).to(torch.bfloat16)
out = torch.empty(R, h, device=dev, dtype=torch.bfloat16)


def timed(fn, n=5):
    for _ in range(2): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


us = timed(lambda: ops.linear_dgrad(dl, E, out=out))
print(f"linear_dgrad (one launch):            {us:8.1f} us  {2*R*V*h/us/1e6:6.0f} TFLOP/s")
ref = out.float()
for S in (2, 4, 8, 16):
    for bn in (0, 256):
        part = torch.empty(S, R, h, device=dev, dtype=torch.float32)
        Ks = V // S
        f = lambda: ops.gemm_raw(M=R, N=h, K=Ks, A=dl, lda=V, B=E, ldb=h, b_mn=True, Cout=part, ldc=h, nz=(S, 1, 1),
                                 a_zs=(Ks, 0, 0), b_zs=(Ks * h, 0, 0), c_zs=(R * h, 0, 0), bn_hint=bn)
        us = timed(f)
        t_red = timed(lambda: part.sum(0))
        err = (part.sum(0) - ref).abs().max().item()
        print(f"split-K S={S:2d} bn_hint={bn:3d}: {us:8.1f} us  {2*R*V*h/us/1e6:6.0f} TFLOP/s   (+ torch sum of partials {t_red:6.1f} us)  "
              f"max |diff| vs one launch {err:.3g}")
