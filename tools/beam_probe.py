"""GPU probe: `generate(num_beams=3, do_sample=True, top_p=0.9, top_k=50)` — the Stage-2 validation call
(ref:Stage2/trainer.py:603-611) — ms per generated token at B images (64 tokens, the prefill timed separately and subtracted)."""
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


NEW = 64


def run(**kw):
    torch.cuda.synchronize(); t0 = time.perf_counter()
    out = eng.generate(px, ids, max_new_tokens=NEW, pad_token_id=0, eos_token_id=None, generator=g, **kw)
    torch.cuda.synchronize()
    return time.perf_counter() - t0, out


def prefill_time(nb):
    rep = lambda t: t.repeat_interleave(nb, dim=0)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    eng.prefill(rep(px), rep(ids), max_new_tokens=NEW)
    torch.cuda.synchronize()
    return time.perf_counter() - t0


for name, kw in (("greedy (no beams)", dict(num_beams=1, do_sample=False)),
                 ("beam search n=3", dict(num_beams=3, do_sample=False)),
                 ("beam sampling n=3 top_p=0.9 top_k=50", dict(num_beams=3, do_sample=True, top_p=0.9, top_k=50)),
                 ("beam sampling + repetition penalty 1.8", dict(num_beams=3, do_sample=True, top_p=0.9, top_k=50,
                                                                 temperature=0.3, repetition_penalty=1.8, length_penalty=1.2))):
    nb = kw["num_beams"]
    run(**kw); run(**kw)                               # same cache geometry every call: one graph capture, then replays
    tp = min(prefill_time(nb) for _ in range(3))
    tt = min(run(**kw)[0] for _ in range(3))
    print(f"B={B} {name:40s}: {(tt - tp) / NEW * 1e3:6.3f} ms per token step (prefill {tp * 1e3:.1f} ms)")
