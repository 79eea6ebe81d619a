"""GPU probe: the down-projection input gradient with the fused GeGLU-backward epilogue (M=22496, N=6912, K=1152 and
the Gemma3-4B shape M=13296, N=10240, K=2560), timed in this process.  Run twice, with PT_GEMM_GBWD_TMA=1 (TMA-moved
epilogue tiles) and =0 (load/store-unit epilogue), back to back on one box for an A/B."""
import os, sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
torch.manual_seed(0)
for (M, ff, h) in ((22496, 6912, 1152), (13296, 10240, 2560)):
    sets = []
    for _ in range(3):
        sets.append(((torch.randn(M, h, device=dev)).to(torch.bfloat16), (torch.randn(h, ff, device=dev) * 0.05).to(torch.bfloat16),
                     torch.randn(M, 2 * ff, device=dev).to(torch.bfloat16), torch.empty(M, 2 * ff, device=dev, dtype=torch.bfloat16)))
    def run(i):
        dy, w, gu, dgu = sets[i % 3]
        ops.linear_dgrad_geglu_bwd(dy, w, gu, dgu)
    def plain(i):
        dy, w, gu, dgu = sets[i % 3]
        ops.linear_dgrad(dy, w, out=dgu[:, :ff])
    for name, fn in (("fused GeGLU-bwd epilogue", run), ("plain dgrad", plain)):
        for i in range(3): fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(30): fn(i)
        e1.record(); torch.cuda.synchronize()
        us = e0.elapsed_time(e1) / 30 * 1e3
        print(f"PT_GEMM_GBWD_TMA={os.environ.get('PT_GEMM_GBWD_TMA', '1')} M={M} N={ff} K={h} {name}: {us:.1f} us "
              f"{2.0 * M * ff * h / us / 1e6:.0f} TFLOP/s", flush=True)
