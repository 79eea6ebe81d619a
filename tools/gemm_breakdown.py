"""GPU probe: per-shape time / TFLOP/s of every GEMM launch in one Stage-1 step (CUDA events around each
launch), next to torch.matmul (cuBLAS) on the same logical shape as the library bar."""
import sys, collections, argparse
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

ap = argparse.ArgumentParser(); ap.add_argument("--B", type=int, default=32); ap.add_argument("--T", type=int, default=128)
ap.add_argument("--cublas", type=int, default=1)
a = ap.parse_args()
dev = torch.device("cuda:0")
vc, lc = SIGLIP_L, GEMMA3_1B
eng = build_engine(vc, lc, dev)
eng.use_graph = False
g = torch.Generator(device=dev).manual_seed(1234)
px = torch.rand(a.B, 3, vc.image, vc.image, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, lc.vocab, (a.B, a.T), generator=g, device=dev); ids[:, 0] = 2
batch = {"pixel_values": px, "token_ids": ids, "labels": ids.clone()}
for _ in range(2):
    eng.train_step(batch, 1e-4)
torch.cuda.synchronize()
ops.GEMM_PROFILE = []
eng.train_step(batch, 1e-4)
torch.cuda.synchronize()
prof, ops.GEMM_PROFILE = ops.GEMM_PROFILE, None
agg = collections.OrderedDict()
for fl, e0, e1, shape in prof:
    d = agg.setdefault(shape, [0, 0.0, fl])
    d[0] += 1; d[1] += e0.elapsed_time(e1)
tot = sum(v[1] for v in agg.values())
print(f"GEMM launches {len(prof)} total {tot:.2f} ms")
print(f"{'M':>6} {'N':>6} {'K':>6} {'nb':>4} {'kg':>2} {'n':>4} {'ms_tot':>8} {'us_each':>8} {'TF/s':>7} {'cuBLAS TF/s':>11}")
for (M, N, K, nb, kg), (n, ms, fl) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    cb = float("nan")
    if a.cublas:
        try:
            x = torch.randn(nb, M, K * kg, device=dev, dtype=torch.bfloat16); w = torch.randn(nb, N, K * kg, device=dev, dtype=torch.bfloat16)
            wt = w.transpose(1, 2)
            for _ in range(3): torch.matmul(x, wt)
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5): torch.matmul(x, wt)
            e1.record(); torch.cuda.synchronize()
            cb = fl / (e0.elapsed_time(e1) / 5 * 1e-3) / 1e12
            del x, w
        except Exception as ex:
            cb = float("nan")
    print(f"{M:6d} {N:6d} {K:6d} {nb:4d} {kg:2d} {n:4d} {ms:8.3f} {1e3*ms/n:8.1f} {fl/(ms/n*1e-3)/1e12:7.1f} {cb:11.1f}")
