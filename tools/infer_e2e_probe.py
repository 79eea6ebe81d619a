"""GPU probe: the reference's single-image inference call end to end (ref:Stage1/inference_stage1.py:150-193: open the
image, processor, vision tower -> [:, 1:, :] -> projector, `generate` with do_sample, temperature 0.7, top_p 0.9,
max_new_tokens 100) on SigLIP-L + Gemma3-1B, random-init weights, one synthetic 1024 x 768 JPEG:
wall-clock ms from the encoded file bytes in host memory to the generated token ids on the host."""
import io
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from PIL import Image
from projectiontrainer_b200 import ops
from projectiontrainer_b200.config import GEMMA3_1B, GEMMA3_4B, QWEN3_8B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
LLM = {"gemma3_1b": GEMMA3_1B, "gemma3_4b": GEMMA3_4B, "qwen3_8b": QWEN3_8B}[sys.argv[1] if len(sys.argv) > 1 else "gemma3_1b"]
NEW = 100
eng = build_engine(SIGLIP_L, LLM, dev)
rng = np.random.default_rng(0)
img = (rng.random((768, 1024, 3)) * 64 + np.linspace(0, 190, 1024)[None, :, None]).astype(np.uint8)
buf = io.BytesIO()
Image.fromarray(img).save(buf, format="JPEG", quality=90)
data = buf.getvalue()
g = torch.Generator(device=dev).manual_seed(0)
prompt = torch.randint(3, LLM.vocab, (1, 16), generator=g, device=dev)


def once():
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    packed = ops.jpeg_pack([torch.frombuffer(bytearray(data), dtype=torch.uint8)], pin=True)
    rgb = ops.jpeg_decode(packed, dev)
    px = eng.preprocess_images(rgb)
    torch.cuda.synchronize()
    t1 = time.perf_counter()
    out = eng.generate(px, prompt, max_new_tokens=NEW, do_sample=True, temperature=0.7, top_p=0.9, pad_token_id=0,
                       eos_token_id=None, generator=g)
    toks = out.cpu()
    t2 = time.perf_counter()
    return (t1 - t0) * 1e3, (t2 - t1) * 1e3, toks


once(); once()
runs = [once() for _ in range(5)]
pre = min(r[0] for r in runs)
gen = min(r[1] for r in runs)
torch.cuda.synchronize(); t0 = time.perf_counter()
eng.generate(eng.preprocess_images(ops.jpeg_decode(ops.jpeg_pack([torch.frombuffer(bytearray(data), dtype=torch.uint8)], pin=True), dev)),
             prompt, max_new_tokens=1, do_sample=True, temperature=0.7, top_p=0.9, pad_token_id=0, eos_token_id=None, generator=g).cpu()
first = (time.perf_counter() - t0) * 1e3
print(f"{sys.argv[1] if len(sys.argv) > 1 else 'gemma3_1b'}: JPEG {len(data) / 1e3:.0f} kB 1024x768 -> decode + resize + normalize "
      f"{pre:.2f} ms; generate {NEW} sampled tokens {gen:.1f} ms ({(gen) / NEW:.2f} ms per token incl. prefill); "
      f"file bytes -> first token {first:.1f} ms; tokens {tuple(runs[0][2].shape)}")
