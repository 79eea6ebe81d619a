"""GPU probe: programmatic dependent launch on / off, in ONE process on one box: the Stage-1 training step
(B = 32, T = 128) and the cached decode step (B = 64), alternating."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import _lib
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
B, T = 32, 128
bs = []
for s in range(2):
    ids = torch.randint(3, GEMMA3_1B.vocab, (B, T), generator=g, device=dev)
    bs.append({"pixel_values": torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1, "token_ids": ids,
               "labels": ids.clone()})


def train(on, n=8):
    _lib.lib().pt_set_dependent_launch(on)
    eng._graphs.clear()                                 # recapture under this setting
    for i in range(4):
        loss = eng.train_step(bs[i % 2], 1e-4, max_norm=1.0)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(n):
        loss = eng.train_step(bs[i % 2], 1e-4, max_norm=1.0)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n, float(loss)


def decode(on, Bd=64, n=20):
    _lib.lib().pt_set_dependent_launch(on)
    px = torch.rand(Bd, 3, 384, 384, generator=g, device=dev) * 2 - 1
    ids = torch.randint(3, GEMMA3_1B.vocab, (Bd, 64), generator=g, device=dev)
    nxt = eng.prefill(px, ids, max_new_tokens=40).argmax(-1)
    eng.cache.update(g_graph=None, g_calls=0)          # recapture under this setting
    for _ in range(3):
        nxt = eng.decode_step(nxt).argmax(-1)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        nxt = eng.decode_step(nxt).argmax(-1)
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n


SETTINGS = (0, 1, 3, 7, 5)     # bits: 1 GEMMs, 2 norm kernels, 4 decode attention


def decode_pair(Bd=64, n=20, reps=3):
    """two caches, one graph each (captured with the setting off / on), replayed alternately from the same slot"""
    px = torch.rand(Bd, 3, 384, 384, generator=g, device=dev) * 2 - 1
    ids = torch.randint(3, GEMMA3_1B.vocab, (Bd, 64), generator=g, device=dev)
    caches = {}
    for on in SETTINGS:
        eng.decode_pdl = on
        eng._cache_key = None                           # a fresh cache (and graph) per setting
        nxt = eng.prefill(px, ids, max_new_tokens=40 + 3 + n * reps).argmax(-1)
        for _ in range(3):
            nxt = eng.decode_step(nxt).argmax(-1)
        caches[on] = (eng.cache, nxt)
    for rep in range(reps):
        for on in SETTINGS:
            eng.cache, nxt = caches[on]
            torch.cuda.synchronize()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for _ in range(n):
                nxt = eng.decode_step(nxt).argmax(-1)
            e1.record(); torch.cuda.synchronize()
            caches[on] = (eng.cache, nxt)
            print(f"decode B={Bd} pdl={on}: {e0.elapsed_time(e1) / n:.3f} ms per token step", flush=True)


if "--train" not in sys.argv:
    for bd in [int(v) for v in sys.argv[1:] if v.isdigit()] or [64, 8]:
        decode_pair(Bd=bd)
else:
    for rep in range(2):
        for on in (0, 1):
            ms, loss = train(on)
            print(f"train  B=32 pdl={on}: {ms:.2f} ms per step (loss {loss:.4f})", flush=True)
