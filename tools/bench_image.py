"""GPU probe: image preprocessing kernels (Pillow-exact bicubic resize + SigLIP normalize) on CXR-sized inputs,
against the HBM roofline (algorithmic bytes = source image read once + f32 output written once), with Pillow's
own time for the same image on one host core beside it.  One JSON line per size."""
import sys, json, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from PIL import Image
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
peaks = {}
try:
    peaks = json.load(open(Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"))
except Exception:
    pass
for H, W in ((2500, 2048), (3000, 3000), (1024, 1024), (512, 512)):
    S = 384
    img = np.random.default_rng(0).integers(0, 256, (H, W, 3), dtype=np.uint8)
    n = 16
    srcs = [torch.from_numpy(img).to(dev) for _ in range(n)]          # distinct buffers: 16 x 15 MB > L2 for CXR sizes
    ws = torch.empty(ops.image_preprocess_workspace(H, W, S, S), dtype=torch.uint8, device=dev)
    out = torch.empty(n, 3, S, S, dtype=torch.float32, device=dev)
    for i in range(n):
        ops.image_preprocess(srcs[i], out[i], ws)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    e0.record()
    for _ in range(reps):
        for i in range(n):
            ops.image_preprocess(srcs[i], out[i], ws)
    e1.record()
    torch.cuda.synchronize()
    us = e0.elapsed_time(e1) * 1e3 / (reps * n)
    alg = H * W * 3 + 3 * S * S * 4
    t = time.perf_counter()
    for _ in range(3):
        ref = np.asarray(Image.fromarray(img).resize((S, S)))
    cpu_ms = (time.perf_counter() - t) / 3 * 1e3
    print(json.dumps({"image": [H, W], "out": S, "us_per_image": us, "images_per_s": 1e6 / us,
                      "algorithmic_bytes": alg, "achieved_GBps": alg / us / 1e3,
                      "hbm_peak_GBps": peaks.get("hbm_gbs"),
                      "pillow_resize_ms_1core": cpu_ms}), flush=True)
