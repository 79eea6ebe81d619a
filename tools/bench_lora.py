"""GPU probe: what LoRA adapters add to the step.  Stage-2 micro-batch on SigLIP-L + Gemma3-1B (B images, question
and answer of T/2 tokens each -> S = 575 + T), three configurations in one process, alternating:

  frozen       projector trains, no adapters                     (the Stage-1 / frozen-LLM Stage-2 step)
  lora         + LoRA r=16 on all seven decoder projections       (ref:Stage2/train_vqa_stage2.py:229-236)
  lora_only    adapters train, projector frozen

Per configuration: ms per forward+backward+optimizer step through the CUDA graph (CUDA events, max of nothing:
one GPU), kernels per step, and, from one eager pass with per-launch events, the time in the adapter GEMMs
(t = x A^T, dt = dy B, dA, dB) next to the base GEMMs.  Prints one JSON line."""
import json
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import _lib, ops
from projectiontrainer_b200.config import GEMMA3_1B, GEMMA3_4B, QWEN3_8B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine
from projectiontrainer_b200.lora import LoraState

dev = torch.device("cuda:0")
B = int(sys.argv[1]) if len(sys.argv) > 1 and not sys.argv[1].startswith("-") else 32
T = int(sys.argv[2]) if len(sys.argv) > 2 and not sys.argv[2].startswith("-") else 128
STEPS, WARM = 10, 4
LLM_NAME = sys.argv[sys.argv.index("--llm") + 1] if "--llm" in sys.argv else "gemma3_1b"
LLM = {"gemma3_1b": GEMMA3_1B, "gemma3_4b": GEMMA3_4B, "qwen3_8b": QWEN3_8B}[LLM_NAME]


def batch(seed):
    g = torch.Generator(device=dev).manual_seed(seed)
    px = torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1
    ids = torch.randint(3, LLM.vocab, (B, T), generator=g, device=dev)
    lab = ids.clone()
    lab[:, :T // 2] = -100
    return {"pixel_values": px, "token_ids": ids, "labels": lab, "label_span": (T // 2, T - T // 2)}


def run(name, lora, train_projector):
    eng = build_engine(SIGLIP_L, LLM, dev)
    if lora:
        eng.lora = LoraState(LLM, dev, r=16, alpha=32.0)
        # B = 0 at the start would make every dt (and dA) exactly zero: give the adapters a trained-like state
        eng.lora.flat.add_(torch.randn_like(eng.lora.flat) * 0.01 * (eng.lora.flat == 0) *
                           _adapter_mask(eng.lora))
        eng.lora.refresh_shadow()
    eng.train_projector = train_projector
    bs = [batch(s) for s in range(3)]
    for i in range(WARM):
        loss = eng.train_step(bs[i % 3], 1e-4, max_norm=1.0)
    torch.cuda.synchronize()
    n0 = _lib.launch_count
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(STEPS):
        loss = eng.train_step(bs[i % 3], 1e-4, max_norm=1.0)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / STEPS
    launches = (_lib.launch_count - n0) // STEPS
    # per-GEMM times from one eager pass
    ops.GEMM_PROFILE = []
    eng.forward_backward(bs[0]["pixel_values"], bs[0]["token_ids"], bs[0]["labels"], 0, bs[0]["label_span"])
    torch.cuda.synchronize()
    prof, ops.GEMM_PROFILE = ops.GEMM_PROFILE, None
    small = lambda sh: min(sh[0], sh[1]) <= 64 or sh[2] <= 64          # an adapter GEMM has a rank-sized dimension
    t_ad = sum(a.elapsed_time(b) for _, a, b, sh in prof if small(sh))
    t_base = sum(a.elapsed_time(b) for _, a, b, sh in prof if not small(sh))
    out = {"ms_per_step": round(ms, 2), "images_per_s": round(B / ms * 1e3, 1), "kernels_per_step": launches,
           "loss": round(float(loss.item()), 4), "gemm_ms_base": round(t_base, 2), "gemm_ms_adapters": round(t_ad, 2),
           "gemm_launches_adapters": sum(1 for _, _, _, sh in prof if small(sh)),
           "mem_GB": round(torch.cuda.max_memory_allocated(dev) / 2 ** 30, 1)}
    if lora:
        out["adapter_params"] = eng.lora.num_parameters()
        agg = {}
        for _, a, b, sh in prof:
            if small(sh):
                k = "x".join(map(str, sh[:3]))
                n, t = agg.get(k, (0, 0.0))
                agg[k] = (n + 1, t + a.elapsed_time(b))
        out["adapter_gemms_MxNxK"] = {k: [n, round(t / n * 1e3, 1)] for k, (n, t) in sorted(agg.items(), key=lambda q: -q[1][1])}
    del eng
    torch.cuda.empty_cache()
    torch.cuda.reset_peak_memory_stats(dev)
    return out


def _adapter_mask(lo):
    m = torch.zeros_like(lo.flat)
    for v in lo.state_dict(m).values():
        v.fill_(1.0)
    return m


res = {"workload": f"Stage-2 micro-batch, SigLIP-L/16-384 + {LLM_NAME}, B={B}, question {T // 2} + answer {T - T // 2} tokens, "
                   f"S={575 + T}, LoRA r=16 alpha=32 on q,k,v,o,gate,up,down", "steps": STEPS, "warmup": WARM}
for name, lora, tp in (("frozen", False, True), ("lora", True, True)) + ((("lora_only", True, False),) if "--all" in sys.argv else ()):
    res[name] = run(name, lora, tp)
print(json.dumps(res))
