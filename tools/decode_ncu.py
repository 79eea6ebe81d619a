"""ncu target: ONE cached decode step (BASELINE.json configs[4] geometry: B = 64, 575 image + 64 prompt tokens,
Gemma3-1B) between cudaProfilerStart/Stop.  Run as
  ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file X python tools/decode_ncu.py
"""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, GEMMA3_1B.vocab, (B, 64), generator=g, device=dev)
nxt = eng.prefill(px, ids, max_new_tokens=40).argmax(-1)
nxt = eng.decode_step(nxt).argmax(-1)          # the first call runs the fixed-shape step eagerly (sizes workspaces)
eng.cache["g_tokens"].copy_(nxt.view(B, 1))
torch.cuda.synchronize()
torch.cuda.profiler.start()
eng._decode_step_fixed()
torch.cuda.synchronize()
torch.cuda.profiler.stop()
