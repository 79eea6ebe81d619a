"""LoRA adapters on the frozen decoder (second half: `--enable_qlora`).

What the reference builds (ref:Stage2/train_vqa_stage2.py:225-250): peft `LoraConfig(r=16, lora_alpha=32,
target_modules=[q,k,v,o,gate,up,down]_proj, lora_dropout=0.05, bias="none")` on every decoder layer, i.e. per target
`y = W x + (alpha / r) * B (A x)` with A [r, in] ~ kaiming_uniform(a=sqrt(5)) and B [out, r] = 0 at the start; the
adapter matrices and the projector are what `optim.AdamW` updates, each module clipped to norm 1.0 on its own
(ref:Stage2/trainer.py:427-441).

How it is laid out here.  The engine keeps q|k|v and gate|up as ONE packed projection each, so the adapters of a packed
projection are packed the same way:

    A_cat [R, in]    rows i*r .. (i+1)*r-1 = A of the i-th adapter of the group           (R = roundup64(n * r))
    B_blk [out, R]   block diagonal: rows of target i, columns i*r .. (i+1)*r-1 = B_i; zero elsewhere

so that  t = s * x A_cat^T  is one thin GEMM and  y = [x | t] . [W | B_blk]^T  is the base GEMM with one more K block
(pt_gemm_args.K2): one pass over the output, every fused epilogue (GeGLU, attention delta, ...) still applies.  The
backward mirrors it: dt = s * dy B_blk, dx = [dy | dt] . [W ; A_cat], dA_cat = dt^T x, dB_i = dy_i^T t_i (written
straight into the diagonal blocks; the off-diagonal entries of B_blk get no gradient and stay exactly zero under
AdamW: zero gradient, zero weight, zero moments).  Masters, gradients and moments are flat fp32 buffers in that block
layout; the bf16 shadow the GEMMs read is refreshed by the optimizer kernel.

Stated difference from the reference: lora_dropout is not applied (p = 0) — a Bernoulli mask on every adapter input
has no counterpart that could be compared, and it would need its own pass over each projection input.
"""
from __future__ import annotations

import math

import torch

from . import ops
from .config import GemmaCfg

F32, BF16 = torch.float32, torch.bfloat16

TARGETS = ("q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj")
# packed projection -> (HF sub-module of the decoder layer, its targets in row order)
GROUPS = (("qkv", "self_attn", ("q_proj", "k_proj", "v_proj")), ("o", "self_attn", ("o_proj",)),
          ("gu", "mlp", ("gate_proj", "up_proj")), ("down", "mlp", ("down_proj",)))


def group_geometry(cfg: GemmaCfg, group: str):
    """(input width, [(target, first output row, rows)]) of a packed projection."""
    qd, kd = cfg.hq * cfg.hd, cfg.hkv * cfg.hd
    if group == "qkv":
        return cfg.hidden, [("q_proj", 0, qd), ("k_proj", qd, kd), ("v_proj", qd + kd, kd)]
    if group == "o":
        return qd, [("o_proj", 0, cfg.hidden)]
    if group == "gu":
        return cfg.hidden, [("gate_proj", 0, cfg.ff), ("up_proj", cfg.ff, cfg.ff)]
    if group == "down":
        return cfg.ff, [("down_proj", 0, cfg.hidden)]
    raise KeyError(group)


class LoraGroup:
    """Views of one (layer, packed projection) into the flat buffers."""
    __slots__ = ("name", "in_dim", "out_dim", "R", "parts", "offA", "offB")

    def __init__(self, name, in_dim, out_dim, R, parts, offA, offB):
        self.name, self.in_dim, self.out_dim, self.R, self.parts = name, in_dim, out_dim, R, parts
        self.offA, self.offB = offA, offB

    def A(self, buf):
        return buf[self.offA:self.offA + self.R * self.in_dim].view(self.R, self.in_dim)

    def B(self, buf):
        return buf[self.offB:self.offB + self.out_dim * self.R].view(self.out_dim, self.R)


class LoraState:
    """All adapters of the decoder: flat fp32 masters / gradients / AdamW moments + the bf16 shadow."""

    def __init__(self, cfg: GemmaCfg, device, r: int = 16, alpha: float = 32.0, targets=TARGETS, seed: int = 0,
                 dropout: float = 0.0):
        if r <= 0 or r % 8:
            raise ValueError(f"LoRA rank {r}: a positive multiple of 8 is needed (16-byte aligned adapter columns)")
        unknown = set(targets) - set(TARGETS)
        if unknown or not targets:
            raise ValueError(f"LoRA target_modules {sorted(unknown)} are not decoder projections {TARGETS}")
        self.cfg, self.r, self.alpha, self.targets = cfg, int(r), float(alpha), tuple(t for t in TARGETS if t in targets)
        self.scale = self.alpha / self.r
        self.dropout = float(dropout)           # recorded for adapter_config.json; not applied (module docstring)
        self.layers, o = [], 0
        for _ in range(cfg.layers):
            groups = {}
            for gname, _, _ in GROUPS:
                in_dim, full = group_geometry(cfg, gname)
                parts = [(t, r0, n, i) for i, (t, r0, n) in enumerate(p for p in full if p[0] in self.targets)]
                if not parts:
                    continue
                out_dim = sum(n for _, _, n in full)
                R = -(-len(parts) * self.r // 64) * 64
                groups[gname] = LoraGroup(gname, in_dim, out_dim, R, parts, o, o + R * in_dim)
                o += R * in_dim + out_dim * R
            self.layers.append(groups)
        self.n = o
        self.flat = torch.zeros(o, dtype=F32, device=device)
        self.grad = torch.zeros(o, dtype=F32, device=device)
        self.m = torch.zeros(o, dtype=F32, device=device)
        self.v = torch.zeros(o, dtype=F32, device=device)
        self.shadow = torch.zeros(o, dtype=BF16, device=device)
        self.sumsq = torch.zeros(1, dtype=F32, device=device)
        self.step = 0
        # peft `reset_lora_parameters`: A ~ kaiming_uniform_(a=sqrt(5)) = U(-1/sqrt(in), 1/sqrt(in)), B = 0
        gen = torch.Generator(device="cpu").manual_seed(seed)
        for groups in self.layers:
            for g in groups.values():
                bound = 1.0 / math.sqrt(g.in_dim)
                for _, _, _, i in g.parts:
                    a = (torch.rand(self.r, g.in_dim, generator=gen) * 2 - 1) * bound
                    g.A(self.flat)[i * self.r:(i + 1) * self.r].copy_(a)
        self.refresh_shadow()

    # ------------------------------------------------------------------ bookkeeping
    def refresh_shadow(self):
        ops.cast_f32_bf16(self.flat, self.shadow)

    def num_parameters(self) -> int:
        """Adapter entries that exist in the reference's model (the padding / off-diagonal zeros do not count)."""
        return sum(self.r * g.in_dim + n * self.r for groups in self.layers for g in groups.values()
                   for _, _, n, _ in g.parts)

    @staticmethod
    def key(li: int, target: str, which: str) -> str:
        """peft's name of adapter matrix `which` ("A" / "B") of `target` in decoder layer li."""
        sub = next(s for _, s, ts in GROUPS if target in ts)
        return f"base_model.model.model.layers.{li}.{sub}.{target}.lora_{which}.weight"

    def state_dict(self, buf=None) -> dict:
        """peft `get_peft_model_state_dict` naming (what `save_pretrained` writes into adapter_model.safetensors);
        the values are views into `buf` (default: the fp32 masters)."""
        buf = self.flat if buf is None else buf
        out = {}
        for li, groups in enumerate(self.layers):
            for g in groups.values():
                A, B = g.A(buf), g.B(buf)
                for t, r0, n, i in g.parts:
                    out[self.key(li, t, "A")] = A[i * self.r:(i + 1) * self.r]
                    out[self.key(li, t, "B")] = B[r0:r0 + n, i * self.r:(i + 1) * self.r]
        return out

    def load_state_dict(self, sd: dict) -> None:
        """Adapter weights by peft key (also accepted: keys carrying the adapter name, `...lora_A.default.weight`)."""
        sd = {k.replace(".default.", "."): v for k, v in sd.items()}
        mine = self.state_dict()
        missing = [k for k in mine if k not in sd]
        if missing:
            raise KeyError(f"adapter state dict lacks {len(missing)} tensors, e.g. {missing[:3]}")
        for k, dst in mine.items():
            if tuple(sd[k].shape) != tuple(dst.shape):
                raise ValueError(f"{k}: shape {tuple(sd[k].shape)} != {tuple(dst.shape)}")
            dst.copy_(sd[k].to(device=dst.device, dtype=F32))
        self.refresh_shadow()

    def adapter_config(self, base_model_name: str = "") -> dict:
        """The fields of peft's adapter_config.json that `PeftModel.from_pretrained` needs."""
        return {"peft_type": "LORA", "task_type": "CAUSAL_LM", "r": self.r, "lora_alpha": self.alpha,
                "lora_dropout": self.dropout, "bias": "none", "target_modules": list(self.targets),
                "base_model_name_or_path": base_model_name, "fan_in_fan_out": False, "inference_mode": True,
                "init_lora_weights": True, "modules_to_save": None}
