#!/usr/bin/env python
"""Stage-1 projector-train throughput benchmark (BASELINE.json metric) — see DESIGN.md §Measurement.

    python bench.py --gpus N --steps K --warmup W            # this repo's sm_100a path
    python bench.py --impl reference --gpus N --steps K ...  # CPU reference arm: the reference's own trainer
                                                             # (baseline/reference_arm.py) on the host cores

Workload = BASELINE.json configs[2] ("config 3" in SURVEY.md): SigLIP-L/16-384 + MLP(1024->10240->1152)
+ Gemma3-1B, 128 caption tokens, batch 32 per GPU, bf16, data-parallel; synthetic CXR-shaped data,
random-init weights.  One JSON line on stdout (rank 0).  Besides the contract's keys the line carries
  parity       loss / projector-gradient agreement of the engine with HF fp32 on the same weights and inputs (B = 2
               slice of the benchmark batch, run untimed after the timed region)
  configs      the other BASELINE.json configurations: [1] SigLIP forward B=128, [3] the Gemma3-4B step B=16 per GPU
               (on all --gpus ranks), [4] prefill B=64 + cached decode
  library_bar  the same step through HF / cuBLASLt / SDPA on this GPU (tools/hf_bar.py)
(--no-extras skips these three).
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time
from pathlib import Path

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))
# stdout carries the one JSON line: NCCL's own banner ("NCCL version ...", printed at communicator creation when
# NCCL_DEBUG is VERSION or higher) goes to stderr instead
os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")


# ... and because NCCL 2.28 still printed the banner on stdout under torchrun on the GPU box, file descriptor 1 is
# pointed at stderr for the whole run (C-level writes included) and the JSON line is written to the saved descriptor.
_JSON_FD = None


def _claim_stdout():
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(obj) -> None:
    sys.stdout.flush()
    data = (json.dumps(obj) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.buffer.write(data); sys.stdout.flush()
    else:
        os.write(_JSON_FD, data)

METRIC = "stage1_projector_train_images_per_sec"
UNIT = "images/s"
B_PER_GPU, T_TOK = 32, 128


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=B_PER_GPU)
    ap.add_argument("--tokens", type=int, default=T_TOK)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true", help="skip the parity / configs / library_bar blocks")
    ap.add_argument("--ref-batch", type=int, default=2, help="images per step of the CPU reference arm (bounded sample)")
    ap.add_argument("--profile", action="store_true",
                    help="ncu helper: 1 warm-up step + --steps steps, nothing else, no JSON value")
    ap.add_argument("--cpu-budget-s", type=float, default=240.0)
    ap.add_argument("--graph", type=int, default=1, help="replay forward+backward as a CUDA graph (0 = eager launches)")
    return ap.parse_args()


def workload_config(args, n):
    return {"workload": "BASELINE.json configs[2]: Stage-1 projector training step, SigLIP-L/16-384 + "
                        "MLP(1024->10240->1152) + Gemma3-1B, 128 caption tokens (S=575+128=703), bf16, "
                        "data-parallel",
            "batch_per_gpu": args.batch, "global_batch": args.batch * n, "caption_tokens": args.tokens,
            "parallelism": f"dp{n}", "cuda_graph": bool(args.graph), "l2": "per-step working set (>30 GB of activations) >> 126 MB L2, no flush"}


# --------------------------------------------------------------------------------------------- CPU arm
def cpu_port_sample(budget_s: float, steps: int, warmup: int, tokens: int):
    """Fallback when the reference's own files are not staged under baseline/_ref: the oracle port of the step on
    the host cores, each step = ONE image of the workload (B=1, T=tokens, full-size models, fp32), all host threads."""
    import torch
    from oracle import stage1_oracle as O
    torch.set_num_threads(os.cpu_count() or 1)
    cores = torch.get_num_threads()
    vc, lc = O.SIGLIP_L, O.GEMMA3_1B
    t0 = time.time()
    Wv, Wl = O.init_siglip(vc), O.init_gemma(lc)
    P = O.init_projector(vc.hidden, lc.hidden)
    init_s = time.time() - t0
    times = []
    done_w = 0
    t_start = time.time()
    for i in range(warmup + steps):
        px, ids, lab = O.synthetic_batch(1, tokens, vc, lc, seed=1234 + i)
        t1 = time.time()
        O.stage1_step(px, ids, lab, Wv, P, Wl, vc, lc)
        dt = time.time() - t1
        if i >= warmup:
            times.append(dt)
        else:
            done_w += 1
        if time.time() - t_start > budget_s and len(times) >= 1:
            break
    mean = sum(times) / len(times)
    return dict(value=1.0 / mean, unit=UNIT, cores=cores, kind="port",
                sample=f"{len(times)} timed step(s) of 1 image each (B=1, T={tokens}, S={575 + tokens}, fp32, "
                       f"full-size models, oracle/stage1_oracle.py; {done_w} warm-up; weight init {init_s:.0f}s "
                       f"not timed)", steps=len(times), warmup=done_w, ms_per_step=mean * 1e3)


def cpu_reference_sample(budget_s: float, steps: int, warmup: int, tokens: int, batch: int):
    """The reference's OWN ProjectionTrainerStage1.train() on HF modules (baseline/reference_arm.py, kind
    "reference"); the oracle port (kind "port") only if the staged copy of the reference's files is missing."""
    from baseline import reference_arm
    if reference_arm.available():
        try:
            return reference_arm.run(steps, warmup, batch, tokens, budget_s)
        except Exception as e:        # say which arm ran and why
            cb = cpu_port_sample(budget_s, steps, min(warmup, 1), tokens)
            cb["sample"] += f"; the reference's own trainer failed here ({type(e).__name__}: {e})"
            return cb
    cb = cpu_port_sample(budget_s, steps, min(warmup, 1), tokens)
    cb["sample"] += "; baseline/_ref is not staged (build() copies it where /root/reference exists)"
    return cb


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    cb = cpu_reference_sample(args.cpu_budget_s, args.steps, args.warmup, args.tokens, args.ref_batch)
    k, w, ms = cb.pop("steps"), cb.pop("warmup"), cb.pop("ms_per_step")
    cb.pop("last_loss", None)
    line = {"impl": "reference", "metric": METRIC, "value": cb["value"], "unit": UNIT, "n_gpus": args.gpus,
            "steps": k, "warmup": w, "ms_per_step": ms, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": workload_config(args, args.gpus),
            "cpu_baseline": cb,
            "e2e": {"value": cb["value"], "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    emit(line)


# --------------------------------------------------------------------------------------------- clocks
class ClockSampler:
    Q = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
         "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.rows, self.proc = [], None
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "200", "-i", str(index)], stdout=subprocess.PIPE, text=True)
            self.th = threading.Thread(target=self._read, daemon=True)
            self.th.start()
        except Exception:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.rows.append((time.time(), ln.strip()))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        sm, mx, reasons = [], None, set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ts, ln in self.rows:
            if ts < t0 or ts > t1:
                continue
            f = [x.strip() for x in ln.split(",")]
            try:
                sm.append(float(f[0])); mx = float(f[1])
            except Exception:
                continue
            for nm, val in zip(names, f[3:7]):
                if val.lower().startswith("active"):
                    reasons.add(nm)
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": mx, "reasons": sorted(reasons),
                "samples": len(sm)}


# --------------------------------------------------------------------------------------------- our arm
def run_ours(args):
    import torch
    import torch.distributed as dist
    from projectiontrainer_b200 import _lib, ops
    from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
    from projectiontrainer_b200.engine import build_engine
    from projectiontrainer_b200.flops import stage1_flops_per_image

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise RuntimeError("bench.py (impl=ours) needs a B200: there is no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    allreduce = (lambda g: dist.all_reduce(g)) if world > 1 else None

    vc, lc = SIGLIP_L, GEMMA3_1B
    B, T = args.batch, args.tokens
    eng = build_engine(vc, lc, dev, seed=0, grad_accum=1)
    eng.use_graph = bool(args.graph) and not args.profile
    g = torch.Generator(device=dev).manual_seed(1234 + rank)
    px = torch.rand(B, 3, vc.image, vc.image, generator=g, device=dev) * 2 - 1
    ids = torch.randint(3, lc.vocab, (B, T), generator=g, device=dev)
    ids[:, 0] = 2
    lab = ids.clone()
    dev_batch = {"pixel_values": px, "token_ids": ids, "labels": lab}
    host_batch = {k: v.cpu().pin_memory() for k, v in dev_batch.items()}
    lr = 1e-4

    def sync():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step(batch):
        return eng.train_step(batch, lr, pad_id=lc.pad_id, allreduce=allreduce, world_size=world)

    if args.profile:
        step(dev_batch)
        sync()
        n0 = _lib.launch_count
        for _ in range(args.steps):
            step(dev_batch)
        sync()
        emit({"profile_run": True, "launches_per_step": (_lib.launch_count - n0) // args.steps})
        return
    for _ in range(max(args.warmup, 3)):
        step(dev_batch)
    sync()
    sampler = ClockSampler(local) if rank == 0 else None
    # ---- device-resident timed region
    n0 = _lib.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sync()
    t_wall0 = time.time()
    e0.record()
    for _ in range(args.steps):
        loss = step(dev_batch)
    e1.record()
    sync()
    t_wall1 = time.time()
    ms = e0.elapsed_time(e1)
    launches = _lib.launch_count - n0
    # ---- end-to-end: pinned host inputs copied in, loss read back, every step
    # (the copy of batch i+1 is started on a side stream before step i is launched — the DataLoader-prefetch
    #  pattern of the trainer; every step's 56.7 MB still crosses PCIe inside the timed region)
    for _ in range(2):
        float(step(eng.stage_batch(host_batch)).item())
    sync()
    e2, e3 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e2.record()
    staged = eng.stage_batch(host_batch)
    for i in range(args.steps):
        nxt = eng.stage_batch(host_batch) if i + 1 < args.steps else None
        lv = float(step(staged).item())
        staged = nxt
    e3.record()
    sync()
    ms_e2e = e2.elapsed_time(e3)
    clocks = sampler.stop(t_wall0, t_wall1) if sampler else None
    t = torch.tensor([ms, ms_e2e], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ms_e2e = t.tolist()

    dp = None
    if world > 1:
        # all-reduced gradient == mean of the per-rank gradients; parameters identical on every rank (tools/dp_check.py)
        from tools.dp_check import dp_check
        try:
            dp = dp_check(eng, dev_batch, lc.pad_id, world)
        except Exception as e:
            dp = {"ok": False, "error": f"{type(e).__name__}: {e}"}

    # ---- roofline pass for the dominant kernel (gemm_bf16_kernel): CUDA events around every launch
    roof = None
    if rank == 0:
        ops.GEMM_PROFILE = []
        eng.use_graph = False
        # rank-local pass (no collective: the other ranks do not take part in it) over the forward + backward only:
        # every GEMM of the step is in it, and no optimizer step runs that the other ranks would not see
        eng.forward_backward(px, ids, lab, lc.pad_id)
        torch.cuda.synchronize()
        prof, ops.GEMM_PROFILE = ops.GEMM_PROFILE, None
        fl = sum(p[0] for p in prof)
        tm = sum(p[1].elapsed_time(p[2]) for p in prof)
        peaks = {}
        pk = ROOT / "MEASURED_PEAKS.json"
        which = "fallback(B200_PROFILING.md)"
        peak = 1400.0
        if pk.exists():
            peaks = json.loads(pk.read_text())
            peak, which = float(peaks.get("bf16_tflops_sustained", 1400.0)), "measured(bf16_tflops_sustained)"
        # dominant kernel = the CTA-pair GEMM on its largest single shape of the step, the Gemma3 gate||up
        # projection [B*S, 2*ff, h] (26 launches per step); algorithmic work per launch = 2*M*N*K
        dom_shape = (B * (575 + T), 2 * lc.ff, lc.hidden, 1, 1)
        dom = [p for p in prof if p[3] == dom_shape]
        d_fl = dom[0][0]
        d_ms = sum(p[1].elapsed_time(p[2]) for p in dom) / len(dom)
        traffic, tsrc = None, None
        tj = ROOT / "profiles" / "ncu_traffic.json"
        if tj.exists():
            tr = json.loads(tj.read_text()).get("gemm2_gateup_fwd")
            if tr and tuple(tr.get("shape", ())) == dom_shape[:3]:
                traffic, tsrc = tr["dram_bytes_per_launch"], tr["source"]
        ach = d_fl / (d_ms * 1e-3) / 1e12
        roof = {"bound": "tensor",
                "kernel": "gemm2_bf16_kernel (tcgen05 cta_group::2), Gemma3 gate||up projection "
                          f"M={dom_shape[0]} N={dom_shape[1]} K={dom_shape[2]}"
                          + (" with the fused GeGLU epilogue" if eng.fuse_geglu else ""),
                "achieved": ach, "peak": peak, "peak_source": which, "unit": "TFLOP/s", "frac": ach / peak,
                "launches_per_step": len(dom), "flops_per_launch": d_fl, "avg_launch_ms": d_ms,
                "traffic": traffic, "traffic_source": tsrc,
                "all_gemm_launches": {"launches": len(prof), "gemm_ms_per_step": tm,
                                      "achieved": fl / (tm * 1e-3) / 1e12, "frac": fl / (tm * 1e-3) / 1e12 / peak}}

    peak_sus = float(json.loads((ROOT / "MEASURED_PEAKS.json").read_text()).get("bf16_tflops_sustained", 1400.0)) \
        if (ROOT / "MEASURED_PEAKS.json").exists() else 1400.0
    parity, configs, library = None, None, None
    if not args.no_extras:
        if rank == 0:
            parity = parity_block(eng, dev_batch, vc, lc, dev)
            configs = aux_configs_1b(eng, vc, lc, dev, peak_sus)
    fpi = stage1_flops_per_image(vc, lc, T)["train"]
    del eng
    torch.cuda.empty_cache()
    if not args.no_extras:
        c3 = config3_4b_step(dev, rank, world, allreduce, sync, peak_sus)       # every rank takes part
        if rank == 0:
            configs["configs[3]"] = c3
            configs["next_rows"] = next_rows_block(vc, lc, dev)
            try:
                sm = ClockSampler(local)
                t0 = time.time()
                from tools.hf_bar import library_bar
                library = library_bar(B, T, 3, 2, dev)
                library["clocks"] = sm.stop(t0, time.time())
            except Exception as e:          # a comparison line, not the product: say why it is missing
                library = {"unavailable": f"{type(e).__name__}: {e}"}
    if rank != 0:
        if world > 1:
            dist.barrier()
            dist.destroy_process_group()
        return
    ips = args.steps * B * world / (ms * 1e-3)
    ips_e2e = args.steps * B * world / (ms_e2e * 1e-3)
    h2d = sum(v.numel() * v.element_size() for v in host_batch.values())
    line = {"metric": METRIC, "value": ips, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": max(args.warmup, 3), "ms_per_step": ms / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": workload_config(args, world),
            "e2e": {"value": ips_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
                    "ms_per_step": ms_e2e / args.steps},
            "gpu_launches": launches, "clocks": clocks, "roofline": roof,
            "step_tflops_per_gpu": ips / world * fpi / 1e12, "flops_per_image": fpi, "last_loss": lv}
    if dp is not None:
        line["dp_check"] = dp
    if parity is not None:
        line["parity"] = parity
    if configs is not None:
        line["configs"] = configs
    if library is not None:
        if "value" in library:
            library["ours_over_library_bar"] = (ips / world) / library["value"]
        line["library_bar"] = library
    if world == 1 and not args.no_cpu_baseline:
        # ~20-30 s of CPU work: the reference's own trainer, one untimed warm-up step, then 2 timed steps of 2 images
        cb = cpu_reference_sample(30.0, 2, 1, args.tokens, args.ref_batch)
        for k in ("steps", "warmup", "ms_per_step", "last_loss"):
            cb.pop(k, None)
        line["cpu_baseline"] = cb
    emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


# --------------------------------------------------------------------------------------------- extra blocks
def _timed(fn, iters, warm=2):
    import torch
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def parity_block(eng, dev_batch, vc, lc, dev):
    """Untimed, after the timed region: the first two samples of the benchmark batch through (a) the engine as it
    stands (bf16 kernels, the projector weights the timed steps left behind) and (b) HF SiglipVisionModel +
    Gemma3ForCausalLM in fp32 on this GPU holding the same frozen weights (regenerated from the engine's seed) and the
    same projector.  north_star tolerances: loss 1e-2 relative, gradient cosine >= 0.999."""
    import torch
    import torch.nn as nn
    from projectiontrainer_b200 import weights as Wt
    from tools.hf_bar import hf_from_weights, hf_step
    try:
        px, ids, lab = (dev_batch[k][:2].contiguous() for k in ("pixel_values", "token_ids", "labels"))
        vision, llm = hf_from_weights(Wt.random_siglip(vc, dev, 0), vc, Wt.random_gemma(lc, dev, 0), lc, dev, torch.float32)
        P = {k: v.detach().clone() for k, v in eng.proj.params().items()}
        mid = P["model.0.weight"].shape[0]
        proj = nn.Sequential(nn.Linear(vc.hidden, mid), nn.GELU(), nn.Linear(mid, lc.hidden)).to(dev).float()
        proj.load_state_dict({k.replace("model.", ""): v for k, v in P.items()})
        ref_loss = float(hf_step(vision, llm, proj, px, ids, lab, lc.pad_id, autocast=False).item())
        ref_g = {"model." + k: p.grad.detach() for k, p in proj.named_parameters()}
        del vision, llm
        torch.cuda.empty_cache()
        graph = eng.use_graph
        eng.use_graph = False
        eng.select_geometry(("parity", 2))
        loss = float(eng.forward_backward(px, ids, lab, lc.pad_id).item())
        eng.use_graph = graph
        cos = {}
        for k, g in eng.proj.grads().items():
            a, b = g.double().flatten(), ref_g[k].double().flatten()
            cos[k] = float(torch.dot(a, b) / (a.norm() * b.norm()))
        eng.select_geometry(("default",))
        return {"checker": "HF transformers fp32 (SiglipVisionModel + Gemma3ForCausalLM) on this GPU, same weights, "
                           "first 2 samples of the benchmark batch", "loss": loss, "loss_ref": ref_loss,
                "loss_rel": abs(loss - ref_loss) / abs(ref_loss), "grad_cos": cos, "grad_cos_min": min(cos.values()),
                "tolerance": {"loss_rel": 1e-2, "grad_cos_min": 0.999},
                "ok": abs(loss - ref_loss) / abs(ref_loss) < 1e-2 and min(cos.values()) >= 0.999}
    except Exception as e:
        return {"unavailable": f"{type(e).__name__}: {e}"}


def next_rows_block(vc, lc, dev):
    """SURVEY.md §8(f) rows on the driver-run line (rank 0, after the main engine is gone): the Stage-2 micro-batch
    with LoRA adapters (rank 2: r=16 on the seven decoder projections, ref:Stage2/train_vqa_stage2.py:229-236) next to
    the same micro-batch with the LLM frozen, and the GPU JPEG decode + resize + normalize of one batch of
    chest-X-ray-sized files (rank 3) next to Pillow on one host core.  Short runs: a few seconds each."""
    import io
    import numpy as np
    import torch
    from projectiontrainer_b200 import ops
    from projectiontrainer_b200.engine import build_engine
    from projectiontrainer_b200.lora import LoraState
    out = {}
    try:
        B, T = 32, 128
        g = torch.Generator(device=dev).manual_seed(7)
        batch = {"pixel_values": torch.rand(B, 3, vc.image, vc.image, generator=g, device=dev) * 2 - 1,
                 "token_ids": torch.randint(3, lc.vocab, (B, T), generator=g, device=dev)}
        batch["labels"] = batch["token_ids"].clone()
        batch["labels"][:, :T // 2] = -100
        batch["label_span"] = (T // 2, T - T // 2)
        res = {}
        for name, lora in (("frozen_llm", False), ("lora_r16_all_targets", True)):
            eng = build_engine(vc, lc, dev)
            if lora:
                eng.lora = LoraState(lc, dev, r=16, alpha=32.0)
                mask = torch.zeros_like(eng.lora.flat)
                for v in eng.lora.state_dict(mask).values():
                    v.fill_(1.0)
                eng.lora.flat.add_(torch.randn_like(mask) * 0.01 * mask * (eng.lora.flat == 0))    # B != 0: no zero shortcuts
                eng.lora.refresh_shadow()
            ms = _timed(lambda: eng.train_step(batch, 1e-4, max_norm=1.0), 6, warm=4)
            res[name] = {"ms_per_step": ms, "images_per_s": B / ms * 1e3}
            if lora:
                res[name]["adapter_params"] = eng.lora.num_parameters()
            del eng
            torch.cuda.empty_cache()
        out["stage2_vqa_microbatch"] = {"what": f"SigLIP-L + Gemma3-1B, B={B}, question {T // 2} + answer {T - T // 2} "
                                                "tokens, forward + backward + clip 1.0 + AdamW through the CUDA graph",
                                        **res}
    except Exception as e:
        out["stage2_vqa_microbatch"] = {"error": f"{type(e).__name__}: {e}"}
    try:
        from PIL import Image
        H, W, N, S = 2500, 2048, 16, vc.image
        rng = np.random.default_rng(0)
        yy, xx = np.mgrid[0:H, 0:W]
        base = (128 + 90 * np.sin(yy / 97.0) * np.cos(xx / 61.0)).astype(np.float32)
        files = []
        for i in range(2):
            b = io.BytesIO()
            Image.fromarray(np.clip(base + rng.normal(0, 6 + 2 * i, (H, W)), 0, 255).astype(np.uint8)).save(b, "JPEG", quality=90)
            files.append(b.getvalue())
        files = [files[i % 2] for i in range(N)]
        t0 = time.perf_counter()
        ref = np.asarray(Image.open(io.BytesIO(files[0])).convert("RGB").resize((S, S)))
        t_pil = time.perf_counter() - t0
        t0 = time.perf_counter()
        packed = ops.jpeg_pack(files)
        t_pack = (time.perf_counter() - t0) / N
        ws = torch.empty(packed["workspace_bytes"], dtype=torch.uint8, device=dev)
        ws2 = torch.empty(ops.image_preprocess_workspace(H, W, S, S), dtype=torch.uint8, device=dev)
        px = torch.empty(N, 3, S, S, device=dev)
        u8 = torch.empty(S, S, 3, dtype=torch.uint8, device=dev)

        def run():
            imgs = ops.jpeg_decode(packed, dev, ws)
            for i, im in enumerate(imgs):
                ops.image_preprocess(im, px[i], ws2, dst_u8=u8 if i == 0 else None)
        ms = _timed(run, 3, warm=2)
        out["jpeg_decode_resize_normalize"] = {
            "what": f"{N} baseline grayscale JPEG files of {H} x {W} (quality 90, no restart markers, "
                    f"{len(files[0]) / 1e6:.2f} MB each): upload + Huffman/IDCT/colour + bicubic resize to {S} x {S} + "
                    "normalize on the GPU, vs Pillow decode + resize on one host core",
            "gpu_ms_per_batch": ms, "gpu_images_per_s": N / ms * 1e3, "host_pack_ms_per_image": t_pack * 1e3,
            "pillow_ms_per_image": t_pil * 1e3, "resized_image_equals_pillow": bool(np.array_equal(u8.cpu().numpy(), ref))}
    except Exception as e:
        out["jpeg_decode_resize_normalize"] = {"error": f"{type(e).__name__}: {e}"}
    return out


def aux_configs_1b(eng, vc, lc, dev, peak):
    """BASELINE.json configs[1] (SigLIP-L forward only, B=128) and configs[4] (prefill of 575 image + 64 prompt tokens
    at B=64, LM head on the last position, plus the cached decode step) on the benchmark's engine."""
    import torch
    from projectiontrainer_b200.flops import stage1_flops_per_image
    out = {}
    g = torch.Generator(device=dev).manual_seed(1)
    fl = stage1_flops_per_image(vc, lc, 64)
    try:
        eng.select_geometry(("configs[1]",))
        B = 128
        px = torch.rand(B, 3, vc.image, vc.image, generator=g, device=dev) * 2 - 1
        ms = _timed(lambda: eng.vision_forward(px, drop_first=False), 5)
        tf = B * fl["vision"] / ms / 1e9
        out["configs[1]"] = {"what": "SigLIP ViT-L/16-384 frozen encoder forward only, bf16, B=128, 1 GPU",
                             "ms": ms, "images_per_s": B / ms * 1e3, "tflops": tf, "frac_of_sustained_peak": tf / peak}
        del px
        eng.select_geometry(("configs[4]",))
        B, Tp, nd = 64, 64, 24
        px = torch.rand(B, 3, vc.image, vc.image, generator=g, device=dev) * 2 - 1
        ids = torch.randint(3, lc.vocab, (B, Tp), generator=g, device=dev)
        ms = _timed(lambda: eng.prefill(px, ids, max_new_tokens=nd), 5)
        tf = B * fl["prefill"] / ms / 1e9
        nxt = eng.prefill(px, ids, max_new_tokens=nd).argmax(-1)
        for _ in range(3):                      # eager warm-up, graph capture, first replay
            nxt = eng.decode_step(nxt).argmax(-1)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(nd - 4):
            nxt = eng.decode_step(nxt).argmax(-1)
        e1.record()
        torch.cuda.synchronize()
        msd = e0.elapsed_time(e1) / (nd - 4)
        out["configs[4]"] = {"what": "Stage-1 inference prefill (inference_stage1.py path): vision + projector + 575 "
                                     "image tokens + 64-token prompt through Gemma3-1B, B=64, LM head on the last "
                                     "position; then the KV-cached decode step (one CUDA-graph replay per token)",
                             "prefill_ms": ms, "images_per_s": B / ms * 1e3, "tflops": tf,
                             "frac_of_sustained_peak": tf / peak, "decode_ms_per_token_step": msd,
                             "decode_tokens_per_s": B / msd * 1e3}
        # `generate` end to end with the reference's sampling call (ref:Stage1/inference_stage1.py:176-186: do_sample,
        # temperature 0.7, top_p 0.9, HF's default top_k 50), per generated token: difference of two lengths so that
        # the prefill drops out; includes the token choice (pt_sample_token) and the host loop
        import time

        def gen(n):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            eng.generate(px, ids, max_new_tokens=n, do_sample=True, temperature=0.7, top_p=0.9, pad_token_id=0,
                         eos_token_id=None, generator=g)
            torch.cuda.synchronize()
            return time.perf_counter() - t0
        gen(nd), gen(8)
        msg = min((gen(nd) - gen(8)) / (nd - 8) * 1e3 for _ in range(3))
        out["configs[4]"].update({"generate_sampled_ms_per_token_step": msg,
                                  "generate_sampled_tokens_per_s": B / msg * 1e3})
        # the reference's inference scripts run one image at a time: the same step at B = 1 (linear layers as weight
        # streaming, pt_gemv_small_m)
        nxt = eng.prefill(px[:1], ids[:1], max_new_tokens=nd).argmax(-1)
        for _ in range(3):
            nxt = eng.decode_step(nxt).argmax(-1)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(nd - 4):
            nxt = eng.decode_step(nxt).argmax(-1)
        e1.record()
        torch.cuda.synchronize()
        ms1 = e0.elapsed_time(e1) / (nd - 4)
        out["configs[4]"].update({"decode_b1_ms_per_token_step": ms1, "decode_b1_tokens_per_s": 1e3 / ms1})
        eng.select_geometry(("default",))
    except Exception as e:
        out["error"] = f"{type(e).__name__}: {e}"
    return out


def config3_4b_step(dev, rank, world, allreduce, sync, peak):
    """BASELINE.json configs[3]: SigLIP-L + MLP(1024->10240->2560) + Gemma3-4B text, 256-token captions, batch 16 per
    GPU, data-parallel over all ranks (the gradient all-reduce is inside the timed steps)."""
    import torch
    import torch.distributed as dist
    from projectiontrainer_b200.config import GEMMA3_4B, SIGLIP_L
    from projectiontrainer_b200.engine import build_engine
    from projectiontrainer_b200.flops import stage1_flops_per_image
    try:
        eng = build_engine(SIGLIP_L, GEMMA3_4B, dev, seed=0)
        B, T = 16, 256
        g = torch.Generator(device=dev).manual_seed(4321 + rank)
        px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
        ids = torch.randint(3, GEMMA3_4B.vocab, (B, T), generator=g, device=dev)
        ids[:, 0] = 2
        batch = {"pixel_values": px, "token_ids": ids, "labels": ids.clone()}
        step = lambda: eng.train_step(batch, 1e-4, pad_id=GEMMA3_4B.pad_id, allreduce=allreduce, world_size=world)
        first = float(step().item())
        for _ in range(2):
            step()
        sync()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        n = 5
        for _ in range(n):
            last = step()
        e1.record()
        sync()
        t = torch.tensor([e0.elapsed_time(e1) / n], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
        fl = stage1_flops_per_image(SIGLIP_L, GEMMA3_4B, T)["train"]
        tf = B * fl / ms / 1e9
        out = {"what": "Stage-1 projector training SigLIP-L + MLP(1024->10240->2560) + Gemma3-4B text, 256 caption "
                       f"tokens (S=831), batch 16 per GPU, bf16, dp{world}",
               "ms_per_step": ms, "images_per_s": B * world / ms * 1e3, "tflops_per_gpu": tf,
               "frac_of_sustained_peak": tf / peak, "n_gpus": world, "first_loss": first,
               "last_loss": float(last.item()), "mem_GB": torch.cuda.max_memory_allocated(dev) / 2 ** 30}
        del eng
        torch.cuda.empty_cache()
        return out
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"}


if __name__ == "__main__":
    a = parse()
    _claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_ours(a)
