"""GPU probe: the decode step's GEMMs (M = B rows, Gemma3-1B shapes) per tile width and split-K factor, inside a
CUDA graph of 26 launches over 26 different weight matrices (weights larger than L2 in total for the big ones)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 64
L = 26
shapes = {"qkv": (1536, 1152), "o": (1152, 1024), "down": (1152, 6912), "gate_up": (13824, 1152)}
for name, (N, K) in shapes.items():
    x = torch.randn(M, K, device=dev).bfloat16()
    ws = [torch.randn(N, K, device=dev).bfloat16() for _ in range(L)]
    mb = N * K * 2 / 1e6
    for bn in (0, 64, 128, 256):
        for ks in (0, -1, 4, 8, 16):
            if ks > 0 and K // 64 < ks:
                continue
            out = torch.zeros(M, N, device=dev, dtype=torch.float32 if ks else torch.bfloat16)
            def run():
                for w in ws:
                    ops.gemm_raw(M=M, N=N, K=K, A=x, lda=K, B=w, ldb=K, Cout=out, ldc=N, bn_hint=bn, ksplit=ks)
            try:
                run()
            except RuntimeError as e:
                print(name, bn, ks, "rejected:", str(e)[-60:]); continue
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g):
                run()
            g.replay()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(5):
                g.replay()
            e1.record(); torch.cuda.synchronize()
            us = e0.elapsed_time(e1) / (5 * L) * 1e3
            print(f"{name:8s} M={M} N={N} K={K} bn_hint={bn:3d} ksplit={ks:2d}: {us:6.1f} us  ({mb / us * 1e3:5.0f} GB/s of weights)")
