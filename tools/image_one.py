"""One image through pt_image_preprocess (for ncu)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np, torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")
H, W, S = 2500, 2048, 384
src = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (H, W, 3), dtype=np.uint8)).to(dev)
ws = torch.empty(ops.image_preprocess_workspace(H, W, S, S), dtype=torch.uint8, device=dev)
out = torch.empty(3, S, S, dtype=torch.float32, device=dev)
for _ in range(3):
    ops.image_preprocess(src, out, ws)
torch.cuda.synchronize()
