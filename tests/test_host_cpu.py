"""CPU-side checks: the C-ABI library loads and exports every symbol include/pt_stage1.h declares (no
compute calls), host logic of the drop-in trainer, and the world_size-2 gloo path of the data-parallel
runtime (sum all-reduce of the flat gradient, loss gather, loader sharding)."""
import ctypes
import math
import os
import re
import socket
from pathlib import Path

import pytest
import torch
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def test_c_abi_exports_every_declared_symbol():
    from projectiontrainer_b200 import build
    lib_path = build.build()
    hdr = (ROOT / "include" / "pt_stage1.h").read_text()
    hdr = re.sub(r"/\*.*?\*/", "", hdr, flags=re.S)
    names = set(re.findall(r"\b(pt_[a-z0-9_]+)\s*\(", hdr))
    assert len(names) >= 20, names
    lib = ctypes.CDLL(str(lib_path))
    missing = [n for n in sorted(names) if not hasattr(lib, n)]
    assert not missing, f"declared in pt_stage1.h but not exported: {missing}"
    lib.pt_last_error.restype = ctypes.c_char_p
    assert lib.pt_version() >= 100


def test_compute_entry_fails_loudly_without_a_b200():
    """No CPU fallback: on a box without an sm_100 device a compute entry returns PT_ERR_CUDA."""
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    from projectiontrainer_b200 import _lib
    L = _lib.lib()
    assert L.pt_device_check() == -2
    assert b"no CPU path" in L.pt_last_error() or b"CUDA" in L.pt_last_error()
    from projectiontrainer_b200 import ops
    with pytest.raises(RuntimeError):
        ops.linear(torch.zeros(8, 8, dtype=torch.bfloat16), torch.zeros(8, 8, dtype=torch.bfloat16))


def test_projector_module_layout_matches_reference():
    from projectiontrainer_b200.projectors import MLPProjector
    m = MLPProjector(32, 48)
    assert list(m.state_dict().keys()) == ["model.0.weight", "model.0.bias", "model.2.weight", "model.2.bias"]
    assert m.model[0].in_features == 32 and m.model[0].out_features == 320 and m.model[2].out_features == 48
    assert all(p.dtype == torch.float32 for p in m.parameters())
    with pytest.raises(RuntimeError):
        m(torch.zeros(1, 4, 32))          # CPU tensors: no fallback path


def test_cosine_schedule_matches_oracle_and_hf_formula():
    from oracle import stage1_oracle as O
    from projectiontrainer_b200.projector_trainer import cosine_with_warmup
    for warm, total in [(0, 10), (3, 10), (2, 3), (5, 5), (0, 1)]:
        for s in range(total + 3):
            assert abs(cosine_with_warmup(s, warm, total) * 1e-3 - O.cosine_lr(s, 1e-3, warm, total)) < 1e-15


def test_sharded_loader_even_batches():
    from projectiontrainer_b200.projector_trainer import _ShardedLoader
    batches = list(range(7))
    for world in (1, 2, 3, 4):
        seen = [list(_ShardedLoader(batches, r, world)) for r in range(world)]
        assert all(len(s) == math.ceil(7 / world) for s in seen)       # same step count on every rank
        assert set(sum(seen, [])) == set(batches)                       # every batch is consumed


def test_sharded_loader_loads_only_its_own_samples():
    """shard at the batch sampler (accelerate's BatchSamplerShard) — a rank must not decode the batches
    of the other ranks.  Every rank sees the same seeded permutation, takes batch i iff i % world == rank, and
    short ranks wrap around to one of the first batches so that all ranks step equally often."""
    import torch
    from projectiontrainer_b200.projector_trainer import sharded_loader

    class Counting(torch.utils.data.Dataset):
        def __init__(self, n):
            self.n, self.calls = n, []
        def __len__(self):
            return self.n
        def __getitem__(self, i):
            self.calls.append(i)
            return {"x": torch.tensor(i)}

    n, bs = 23, 4                                   # 6 batches, the last one short
    for world in (1, 2, 4):
        per_rank, lens = [], []
        for r in range(world):
            ds = Counting(n)
            ld = sharded_loader(ds, batch_size=bs, shuffle=True, rank=r, world=world,
                                generator=torch.Generator().manual_seed(0), num_workers=0)
            assert ld.unsharded_len == 6 and len(ld) == math.ceil(6 / world)
            batches = [b["x"].tolist() for b in ld]
            assert len(batches) == len(ld)
            assert sorted(ds.calls) == sorted(sum(batches, []))          # nothing loaded and thrown away
            assert len(ds.calls) <= math.ceil(6 / world) * bs
            per_rank.append(batches)
            lens.append(len(batches))
        assert len(set(lens)) == 1
        assert set(sum(sum(per_rank, []), [])) == set(range(n))          # together: the whole data set
    # two epochs draw two different permutations, identically on every rank
    lds = [sharded_loader(Counting(n), batch_size=bs, shuffle=True, rank=r, world=2,
                          generator=torch.Generator().manual_seed(0), num_workers=0) for r in range(2)]
    e1 = [[b["x"].tolist() for b in ld] for ld in lds]
    e2 = [[b["x"].tolist() for b in ld] for ld in lds]
    assert e1 != e2 and set(sum(e2[0] + e2[1], [])) == set(range(n))


def test_config_rejects_unsupported_models():
    from projectiontrainer_b200.config import gemma_cfg_from_hf

    class C:
        model_type = "qwen3"
    with pytest.raises(RuntimeError):
        gemma_cfg_from_hf(C())


def test_flops_model_matches_survey():
    from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
    from projectiontrainer_b200.flops import stage1_flops_per_image
    f = stage1_flops_per_image(SIGLIP_L, GEMMA3_1B, 128)
    assert abs(f["train"] / 1e9 - 2637.1) < 0.5 and abs(f["vision"] / 1e9 - 381.4) < 0.2


# ------------------------------------------------------------------------------ world_size 2 (gloo)
def _free_port():
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        return s.getsockname()[1]


def _rank_main(rank, world, port, out):
    os.environ.update(RANK=str(rank), LOCAL_RANK=str(rank), WORLD_SIZE=str(world), MASTER_ADDR="127.0.0.1",
                      MASTER_PORT=str(port))
    from projectiontrainer_b200.accelerator_setup import Stage1Accelerator
    acc = Stage1Accelerator(gradient_accumulation_steps=2, device="cpu", backend="gloo")
    assert acc.num_processes == world and acc.process_index == rank and acc.is_main_process == (rank == 0)
    g = torch.full((1000,), float(rank + 1))
    acc.all_reduce_sum_(g)                                   # the path's one exchange: SUM, mean folded later
    mean = g / world
    losses = acc.gather(torch.tensor(10.0 + rank))
    # Stage 2 (ref:Stage2/trainer.py:306): gradients accumulate locally inside accumulate() windows and are
    # exchanged ONCE per window, at the sync micro-batch; 3 micro-batches with GA=2 -> windows {0,1} and {2}
    from projectiontrainer_b200.vqa_trainer import accumulate_sync_flag
    window, micro, exchanged = torch.zeros(4), 0, []
    for i in range(3):
        sync, micro = accumulate_sync_flag(micro, acc.gradient_accumulation_steps, i == 2)
        window += float(rank + 1) * (i + 1)
        if sync:
            acc.all_reduce_sum_(window)
            exchanged.append(float(window[0]) / world)
            window.zero_()
    assert exchanged == [4.5, 4.5] and acc.backward_divisor == 2
    assert acc.gather_object([f"r{rank}a", f"r{rank}b"]) == ["r0a", "r0b", "r1a", "r1b"]
    out.put((rank, float(mean[0]), losses.tolist()))
    acc.wait_for_everyone()
    acc.end_training()


def test_two_process_allreduce_and_gather_gloo():
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_rank_main, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, mean, losses in res:
        assert mean == 1.5                      # mean over ranks of per-rank gradients (DDP semantics)
        assert losses == [10.0, 11.0]


def test_qwen3_container_is_repacked_like_the_oracle_names_it():
    """weights.llm_from_hf on an HF Qwen3ForCausalLM: gains packed as weight - 1, post_attention_layernorm is the
    MLP's pre-norm, untied head kept, config recipe flags set; and unsupported variants raise."""
    import torch
    from oracle import stage1_oracle as O
    from oracle.make_golden import CASES, build_hf_qwen3
    from projectiontrainer_b200 import weights as Wt
    lkw = CASES["tiny_qwen3_gqa"][1]
    olc = O.GemmaCfg(**lkw)
    Wl = O.init_gemma(olc)
    hf = build_hf_qwen3(olc, Wl)
    W, cfg = Wt.llm_from_hf(hf)
    assert (cfg.sandwich_norm, cfg.silu, cfg.scale_embed, cfg.tie_lm_head, cfg.pattern) == (False, True, False, False, 1)
    assert (cfg.hq, cfg.hkv, cfg.hd, cfg.ff, cfg.q_scalar) == (olc.hq, olc.hkv, olc.hd, olc.ff, float(olc.hd))
    assert all(cfg.is_full(i) for i in range(cfg.layers))
    assert set(W) == set(Wl)
    for k in Wl:
        assert torch.allclose(W[k].float(), Wl[k], atol=1e-7), k
    hf.config.use_sliding_window = True
    with pytest.raises(RuntimeError, match="sliding"):
        Wt.llm_from_hf(hf)
