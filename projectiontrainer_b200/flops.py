"""Algorithmic FLOPs per image of the Stage-1 step (multiply-add = 2 FLOPs;
necessary work only: LM head on label rows, no pooling head, unmasked attention pairs,
input-gradient-only LLM backward)."""
from __future__ import annotations

from .config import GemmaCfg, SiglipCfg


def pairs(S: int, W: int) -> int:
    return S * (S + 1) // 2 if W >= S else W * (W + 1) // 2 + (S - W) * W


def stage1_flops_per_image(v: SiglipCfg, l: GemmaCfg, T: int, expansion: int = 10) -> dict:
    N, d, f, Lv = v.tokens, v.hidden, v.ff, v.layers
    Tv = N - 1
    S = Tv + T
    m = expansion * d
    h = l.hidden
    vision = 2 * N * (3 * v.patch ** 2) * d + Lv * (8 * N * d * d + 4 * N * N * d + 4 * N * d * f)
    proj_fwd = 2 * Tv * d * m + 2 * Tv * m * h
    proj_bwd = 2 * Tv * d * m + 2 * (2 * Tv * m * h)
    lin = 2 * S * l.layers * (h * l.hq * l.hd + 2 * h * l.hkv * l.hd + l.hq * l.hd * h + 3 * h * l.ff)
    n_full = l.layers // l.pattern
    n_slide = l.layers - n_full
    att = 4 * l.hq * l.hd * (n_full * pairs(S, 1 << 30) + n_slide * pairs(S, l.window))
    head = 2 * T * h * l.vocab
    train = vision + proj_fwd + proj_bwd + (lin + att + head) + (lin + 2 * att + head)
    prefill = vision + proj_fwd + lin + att + 2 * h * l.vocab
    return dict(vision=vision, proj_fwd=proj_fwd, proj_bwd=proj_bwd, lin=lin, att=att, head=head,
                train=train, prefill=prefill)
