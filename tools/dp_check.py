"""Data-parallel correctness of the one exchange step of the path (VERDICT r1 "multi-GPU
correctness is asserted by nothing on GPUs"): under torch.distributed (NCCL, one process per GPU)

  1. the all-reduced flat projector gradient equals the mean of the per-rank gradients (all-gathered, summed in
     fp64) to within a few fp32 ulps of the largest addend — i.e. DDP's gradient averaging
     (ref:Stage1/accelerator_setup.py:11-16, projector_trainer.py:100-102, 237), and
  2. after K optimizer steps on different per-rank batches the projector parameters (fp32 master, bf16 shadow, AdamW
     moments) are bit-identical on every rank.

`dp_check(eng, batch, ...)` is called by bench.py when --gpus > 1 (key `dp_check` of the JSON line) and by
tests/test_dp_gpu.py; run stand-alone as
    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 tools/dp_check.py
"""
from __future__ import annotations

import os
import sys
from pathlib import Path

sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

import torch
import torch.distributed as dist


def dp_check(eng, batch: dict, pad_id, world: int, steps: int = 3, lr: float = 1e-3) -> dict:
    Pj = eng.proj
    dev = Pj.flat.device
    graph, eng.use_graph = eng.use_graph, False
    # ---- 1. all-reduce == mean of the per-rank gradients
    eng.select_geometry(("dp_check",) + tuple(batch["token_ids"].shape))
    eng.forward_backward(batch["pixel_values"], batch["token_ids"], batch["labels"], pad_id)
    local = Pj.grad.clone()
    gathered = [torch.empty_like(local) for _ in range(world)]
    dist.all_gather(gathered, local)
    mean64 = torch.zeros_like(local, dtype=torch.float64)
    amax = torch.zeros_like(local)
    for g in gathered:
        mean64 += g.double()
        amax = torch.maximum(amax, g.abs())
    mean64 /= world
    reduced = local.clone()
    dist.all_reduce(reduced)                                   # what Stage1Engine.train_step's `allreduce` does
    got = reduced.double() / world                             # 1/world is folded into pt_clip_adamw (grad_scale)
    # error in units of the fp32 spacing at the largest addend of each entry
    ulp = torch.clamp(amax, min=1e-30).double() * 2.0 ** -23
    max_ulp = float(((got - mean64).abs() / ulp).max())
    differs = bool(any(not torch.equal(gathered[0], g) for g in gathered[1:]))     # ranks really saw different data
    # ---- 2. parameters stay identical over K optimizer steps
    for _ in range(steps):
        eng.train_step(batch, lr, pad_id=pad_id, allreduce=lambda g: dist.all_reduce(g), world_size=world)
    same = True
    for t in (Pj.flat, Pj.shadow, Pj.m, Pj.v):
        ref = t.clone()
        dist.broadcast(ref, src=0)
        same = same and torch.equal(ref, t)
    flag = torch.tensor([1 if same else 0], device=dev)
    dist.all_reduce(flag, op=dist.ReduceOp.MIN)
    eng.use_graph = graph
    ok = max_ulp <= 2.0 * world and bool(flag.item()) and (differs or world == 1)
    return {"ok": ok, "allreduce_vs_fp64_mean_max_ulp": max_ulp, "ulp_bound": 2.0 * world,
            "per_rank_grads_differ": differs, "params_identical_after_steps": bool(flag.item()), "steps": steps,
            "world": world, "grad_elems": int(local.numel())}


def main():
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200 import weights as Wt
    rank, world, local = (int(os.environ.get(k, d)) for k, d in (("RANK", "0"), ("WORLD_SIZE", "1"), ("LOCAL_RANK", "0")))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    dist.init_process_group("nccl", device_id=dev)
    # reduced widths, real geometry classes (hd 256 GQA, window < S): seconds per rank
    vc = SiglipCfg(hidden=128, ff=256, layers=2, heads=2, image=160, patch=16)
    lc = GemmaCfg(vocab=1024, hidden=256, ff=256, layers=7, hq=4, hkv=2, hd=256, window=40, q_scalar=256.0,
                  rope_factor_global=8.0)
    eng = E.Stage1Engine(E.PackedSiglip(Wt.random_siglip(vc, dev, 0), vc, dev),
                         E.PackedGemma(Wt.random_gemma(lc, dev, 0), lc, dev),
                         E.ProjectorState(Wt.random_projector(vc.hidden, lc.hidden, dev, 10, 0), dev), dev)
    g = torch.Generator(device=dev).manual_seed(77 + rank)            # a different batch on every rank
    B, T = 3, 48
    px = torch.rand(B, 3, vc.image, vc.image, generator=g, device=dev) * 2 - 1
    ids = torch.randint(3, lc.vocab, (B, T), generator=g, device=dev)
    ids[:, 0] = 2
    out = dp_check(eng, {"pixel_values": px, "token_ids": ids, "labels": ids.clone()}, lc.pad_id, world)
    if rank == 0:
        import json
        print(json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if out["ok"] else 1)


if __name__ == "__main__":
    main()
