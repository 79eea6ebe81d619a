"""GPU probe: one sample of the Stage-2 VQA inference loop (ref:Stage2/inference_vqa_stage2.py:211-298) end to end —
a 2500 x 2048 chest-X-ray-sized JPEG on disk, a 16-token question, the reference's generation call (3 beams, sampling,
temperature 0.3, top_p 0.9, top_k 50, repetition penalty 1.8, length penalty 1.2), 16 new tokens, random-init SigLIP-L +
Gemma3-1B with LoRA adapters folded in — wall-clock ms per sample with Pillow's decode + resize on the host (the
reference's own image path) and with --gpu_decode."""
import os
import sys
import tempfile
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import argparse
import numpy as np
import torch
from PIL import Image
from projectiontrainer_b200 import inference_vqa_stage2 as V

dev = torch.device("cuda:0")
args = V.build_parser().parse_args(["--synthetic"])
engine, data, image_size, tokenize, decode, pad_id, eos_id = V._synthetic(args, dev)
root = tempfile.mkdtemp(prefix="pt_vqa_probe_")
rng = np.random.default_rng(1)
img = (rng.random((2500, 2048)) * 60 + np.linspace(0, 190, 2048)[None, :]).astype(np.uint8)
Image.fromarray(img).save(os.path.join(root, "cxr.jpg"), quality=90)
print("file", os.path.getsize(os.path.join(root, "cxr.jpg")) // 1000, "kB")
sample = {"image": "cxr.jpg", "problem": "what abnormality is visible in this image"}
gen = dict(max_new_tokens=16, do_sample=True, temperature=0.3, top_p=0.9, top_k=50, repetition_penalty=1.8,
           length_penalty=1.2, num_beams=3)
with engine.merged(engine.merge_adapters()):
    for gd in (False, True):
        f = lambda: V.process_sample(sample, root, engine, image_size, tokenize, decode, gen, pad_id, None, gpu_decode=gd)
        f(); f()
        ts = []
        for _ in range(5):
            torch.cuda.synchronize(); t0 = time.perf_counter()
            out = f()
            torch.cuda.synchronize(); ts.append(time.perf_counter() - t0)
        t0 = time.perf_counter(); V.load_pixel_values(engine, os.path.join(root, "cxr.jpg"), image_size, gd); torch.cuda.synchronize()
        ti = time.perf_counter() - t0
        print(f"gpu_decode={gd}: {min(ts) * 1e3:6.1f} ms per sample (image -> pixel_values {ti * 1e3:.1f} ms), 16 tokens: {len(out.split())}")
