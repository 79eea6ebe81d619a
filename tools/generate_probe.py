"""GPU probe: `generate` end to end (BASELINE.json configs[4] geometry: B images, 575 image + 64 prompt tokens, Gemma3-1B):
ms per generated token step for greedy decoding and for the reference's sampling call (temperature 0.7, top_p 0.9, HF's
default top_k 50; ref:Stage1/inference_stage1.py:176-186), next to the bare decode step."""
import sys
import time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
NEW = 48
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
ids = torch.randint(3, GEMMA3_1B.vocab, (B, 64), generator=g, device=dev)


def timed(**kw):
    def run(n):
        torch.cuda.synchronize(); t0 = time.perf_counter()
        out = eng.generate(px, ids, max_new_tokens=n, pad_token_id=0, eos_token_id=None, **kw)
        torch.cuda.synchronize()
        return time.perf_counter() - t0, out
    run(NEW); run(8)                                    # warm: graph capture, function attributes
    best = None
    for _ in range(3):
        (d8, _), (dn, out) = run(8), run(NEW)           # the difference leaves the prefill out
        ms = (dn - d8) / (NEW - 8) * 1e3
        best = ms if best is None else min(best, ms)
    return best, out


for name, kw in (("greedy", dict(do_sample=False)),
                 ("sample T=0.7 top_p=0.9 top_k=50", dict(do_sample=True, temperature=0.7, top_p=0.9, generator=g)),
                 ("sample + repetition_penalty 1.1", dict(do_sample=True, temperature=0.7, top_p=0.9,
                                                          repetition_penalty=1.1, generator=g))):
    ms, out = timed(**kw)
    print(f"B={B} {name:34s}: {ms:6.3f} ms per token step  ({B / ms * 1e3:8.0f} tokens/s)  out {tuple(out.shape)}")
