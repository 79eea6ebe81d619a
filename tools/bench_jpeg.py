"""GPU probe: baseline-JPEG decode throughput.  A batch of chest-X-ray-sized grayscale files (2500 x 2048) and of
4:2:0 colour files, with and without restart markers: host pack time (one core), device decode time (CUDA events,
upload included / excluded), per-kernel split, against Pillow's decode on one host core.  One JSON line."""
import io
import json
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from PIL import Image
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
N = int(sys.argv[1]) if len(sys.argv) > 1 else 32
H, W = 2500, 2048
rng = np.random.default_rng(0)
yy, xx = np.mgrid[0:H, 0:W]
base = (128 + 90 * np.sin(yy / 97.0) * np.cos(xx / 61.0)).astype(np.float32)


def enc(arr, **kw):
    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


def make(kind, **kw):
    out = []
    for i in range(4):                                  # 4 distinct images, repeated to N
        g = np.clip(base + rng.normal(0, 6 + 2 * i, (H, W)), 0, 255).astype(np.uint8)
        out.append(enc(g, quality=90, **kw) if kind == "gray" else
                   enc(np.stack([g, np.roll(g, 7, 0), np.roll(g, 5, 1)], -1), quality=90, subsampling=2, **kw))
    return [out[i % 4] for i in range(N)]


def ev():
    return torch.cuda.Event(enable_timing=True)


res = {"workload": f"{N} baseline JPEG files of {H} x {W}, quality 90", "cases": {}}
for name, kind, kw in (("gray", "gray", {}), ("gray_rst_1row", "gray", dict(restart_marker_rows=1)),
                       ("ycc420", "rgb", {}), ("ycc420_rst_1row", "rgb", dict(restart_marker_rows=1))):
    files = make(kind, **kw)
    t0 = time.perf_counter()
    packed = ops.jpeg_pack(files)
    t_pack = time.perf_counter() - t0
    ws = torch.empty(packed["workspace_bytes"], dtype=torch.uint8, device=dev)
    for _ in range(2):
        out = ops.jpeg_decode(packed, dev, ws)
    torch.cuda.synchronize()
    e0, e1 = ev(), ev()
    e0.record()
    for _ in range(3):
        out = ops.jpeg_decode(packed, dev, ws)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / 3
    t0 = time.perf_counter()
    for f in files[:4]:
        np.asarray(Image.open(io.BytesIO(f)).convert("RGB"))
    t_pil = (time.perf_counter() - t0) / 4
    assert np.array_equal(out[1].cpu().numpy(), np.asarray(Image.open(io.BytesIO(files[1])).convert("RGB")))
    mb = sum(len(f) for f in files) / 1e6
    res["cases"][name] = {"file_MB_total": round(mb, 1), "host_pack_ms_per_image": round(t_pack / N * 1e3, 2),
                          "gpu_decode_ms_per_batch": round(ms, 2), "gpu_images_per_s": round(N / ms * 1e3, 1),
                          "gpu_out_GBps": round(N * H * W * 3 / ms / 1e6, 1),
                          "pillow_ms_per_image_one_core": round(t_pil * 1e3, 1),
                          "pillow_images_per_s_one_core": round(1 / t_pil, 1)}
print(json.dumps(res))
