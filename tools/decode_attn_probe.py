"""GPU probe: pt_decode_attn alone at the decode bench's geometry (Gemma3-1B: hq 4 / hkv 1 / hd 256; B samples,
Smax-slot cache, this token at `slot`), per cluster size, against the HBM time of the K/V rows it reads."""
import os
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
hq, hkv, hd = (int(v) for v in sys.argv[2:5]) if len(sys.argv) > 4 else (4, 1, 256)
Smax, slot, L = 680, 640, 26
g = torch.Generator(device=dev).manual_seed(0)
qkv = torch.randn(B, (hq + 2 * hkv) * hd, generator=g, device=dev).bfloat16()
w = torch.zeros(hd, device=dev)
ang = torch.rand(Smax, hd // 2, generator=g, device=dev)
cos, sin = ang.cos(), ang.sin()
Ks = [torch.randn(B, Smax, hkv * hd, generator=g, device=dev).bfloat16() for _ in range(L)]   # > L2 in total
Vs = [torch.randn(B, Smax, hkv * hd, generator=g, device=dev).bfloat16() for _ in range(L)]
pos = torch.full((B,), slot, dtype=torch.int64, device=dev)
slot_t = torch.tensor([slot], dtype=torch.int64, device=dev)
valid = torch.ones(B, Smax, dtype=torch.uint8, device=dev)
out = torch.empty(B, hq * hd, dtype=torch.bfloat16, device=dev)
for window in (0, 512):
    keys = slot + 1 if window == 0 else min(window, slot + 1)
    mb = 2 * B * keys * hkv * hd * 2 / 1e6
    for ns in [int(v) for v in os.environ.get("PROBE_NSPLIT", "1,2,4,8").split(",")]:
        def run():
            for li in range(L):
                ops.decode_attn(qkv, w, w, cos, sin, pos, slot_t, valid, Ks[li], Vs[li], out, hq, hkv, hd, window,
                                1e-6, hd ** -0.5, nsplit=ns)
        run()
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()                   # 26 dependent launches back to back, as in the decode step
        with torch.cuda.graph(graph):
            run()
        graph.replay()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(5):
            graph.replay()
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / (5 * L) * 1e3
        print(f"B={B} window={window} nsplit={ns}: {us:7.1f} us per launch, {mb:.1f} MB of K/V rows -> {mb / us * 1e3:.0f} GB/s")
