"""GPU probe: does stream capture record programmatic edges for launches made with the dependent-launch attribute,
and what a chain of small dependent GEMMs gains from it (alternating, one process)."""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import _lib, ops
dev = torch.device("cuda:0")
os.makedirs("gpurun_out", exist_ok=True)
x = torch.randn(64, 1024, device=dev).bfloat16()
ws = [torch.randn(1024, 1024, device=dev).bfloat16() * 0.03 for _ in range(26)]
ys = [torch.empty(64, 1024, device=dev, dtype=torch.bfloat16) for _ in range(2)]


def chain():
    src = x
    for i, w in enumerate(ws):                      # each GEMM reads the previous one's output
        ops.linear(src, w, out=ys[i % 2])
        src = ys[i % 2]


graphs = {}
for on in (0, 1):
    _lib.lib().pt_set_dependent_launch(on)
    chain(); torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    g.enable_debug_mode()
    with torch.cuda.graph(g):
        chain()
    g.debug_dump(f"gpurun_out/pdl_graph_{on}.dot")
    graphs[on] = g
    g.replay(); torch.cuda.synchronize()
    if on == 0:
        ref = ys[(len(ws) - 1) % 2].clone()
    else:
        print("same result:", torch.equal(ref, ys[(len(ws) - 1) % 2]))
for rep in range(3):
    for on in (0, 1):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        graphs[on].replay()
        e0.record()
        for _ in range(20):
            graphs[on].replay()
        e1.record(); torch.cuda.synchronize()
        print(f"pdl={on}: {e0.elapsed_time(e1) / (20 * 26) * 1e3:.2f} us per dependent GEMM (M=64, N=K=1024)")
