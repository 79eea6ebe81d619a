"""GPU probe: pt_gemv_small_m per shape (decode-step linear layers of Gemma3-1B and Qwen3-8B), M rows, inside a CUDA
graph of L launches over L different weight matrices (more than L2 in total): us per launch and GB/s of weights."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
M = int(sys.argv[1]) if len(sys.argv) > 1 else 1
shapes = [("g1b qkv", 1536, 1152, 0), ("g1b o", 1152, 1024, 2), ("g1b gate_up", 6912, 1152, 3), ("g1b down", 1152, 6912, 2),
          ("q8b qkv", 6144, 4096, 0), ("q8b o", 4096, 4096, 2), ("q8b gate_up", 12288, 4096, 4), ("q8b down", 4096, 12288, 2),
          ("g1b lm_head", 262144, 1152, 1)]
for name, N, K, mode in shapes:
    rows = 2 * N if mode in (3, 4) else N
    mb = rows * K * 2 / 1e6
    L = max(2, min(26, int(1200 / mb)))
    x = torch.randn(M, K, device=dev).bfloat16()
    ws = [torch.randn(rows, K, device=dev).bfloat16() for _ in range(L)]
    out = torch.zeros(M, N, device=dev, dtype=torch.float32 if mode in (1, 2) else torch.bfloat16)
    def run():
        for w in ws:
            ops.gemv(x, w, out, mode)
    run(); torch.cuda.synchronize()
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
    print(f"M={M} {name:12s} N={N:6d} K={K:5d} mode {mode}: {us:7.1f} us  {mb:7.1f} MB  {mb / us * 1e3:6.0f} GB/s")
    del ws
