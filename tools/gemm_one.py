"""GPU probe: time ONE GEMM shape through pt_gemm_bf16 (and cuBLAS beside it); ncu target.
   python tools/gemm_one.py M N K [--bmn] [--f32] [--res] [--bias] [--act 0|1|2] [--impl 0|1|2] [--iters n]"""
import sys, argparse
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
ap = argparse.ArgumentParser()
ap.add_argument("M", type=int); ap.add_argument("N", type=int); ap.add_argument("K", type=int)
ap.add_argument("--bmn", action="store_true"); ap.add_argument("--f32", action="store_true")
ap.add_argument("--res", action="store_true"); ap.add_argument("--bias", action="store_true")
ap.add_argument("--act", type=int, default=0); ap.add_argument("--impl", type=int, default=0)
ap.add_argument("--bn", type=int, default=0); ap.add_argument("--iters", type=int, default=10); ap.add_argument("--cublas", type=int, default=1)
a = ap.parse_args()
dev = torch.device("cuda:0")
torch.manual_seed(0)
M, N, K = a.M, a.N, a.K
x = (torch.randn(M, K, device=dev)).to(torch.bfloat16)
w = (torch.randn(K, N, device=dev) * 0.05).to(torch.bfloat16) if a.bmn else (torch.randn(N, K, device=dev) * 0.05).to(torch.bfloat16)
out = torch.empty(M, N, device=dev, dtype=torch.float32 if a.f32 else torch.bfloat16)
bias = torch.randn(N, device=dev) if a.bias else None
res = out if a.res else None
def run():
    ops.gemm_raw(M=M, N=N, K=K, A=x, lda=K, B=w, ldb=w.stride(0), b_mn=a.bmn, Cout=out, ldc=N, bias=bias,
                 residual=res, ldr=N, act=a.act, bn_hint=a.bn)
for _ in range(3): run()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(a.iters): run()
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.iters
print(f"ours impl={a.impl} bn_hint={a.bn} {M}x{N}x{K} bmn={a.bmn}: {ms*1e3:.1f} us  {2.0*M*N*K/ms/1e9:.1f} TF/s", flush=True)
if a.cublas:
    wt = w if a.bmn else w.t()
    o2 = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    for _ in range(3): torch.matmul(x, wt, out=o2)
    e0.record()
    for _ in range(a.iters): torch.matmul(x, wt, out=o2)
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / a.iters
    print(f"cublas: {ms*1e3:.1f} us  {2.0*M*N*K/ms/1e9:.1f} TF/s")
