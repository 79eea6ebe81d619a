"""GPU probe: `generate(num_beams=3, do_sample=True, top_p=0.9, top_k=50)` — the Stage-2 validation call
(ref:Stage2/trainer.py:603-611) — ms per generated token at B images (prefill excluded by differencing two lengths)."""
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 4
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, GEMMA3_1B.vocab, (B, 32), generator=g, device=dev)


def run(n, **kw):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = eng.generate(px, ids, max_new_tokens=n, pad_token_id=0, eos_token_id=None, generator=g, **kw)
    torch.cuda.synchronize()
    return time.perf_counter() - t0, out


for name, kw in (("beam search n=3", dict(num_beams=3, do_sample=False)),
                 ("beam sampling n=3 top_p=0.9 top_k=50", dict(num_beams=3, do_sample=True, top_p=0.9, top_k=50))):
    run(40, **kw); run(8, **kw)
    ms = min((run(40, **kw)[0] - run(8, **kw)[0]) / 32 * 1e3 for _ in range(3))
    print(f"B={B} {name:40s}: {ms:6.3f} ms per token step")
