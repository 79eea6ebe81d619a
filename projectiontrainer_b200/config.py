"""Model geometry of the two frozen towers on the Stage-1 path."""
from __future__ import annotations

from dataclasses import dataclass


@dataclass
class SiglipCfg:
    """SigLIP vision tower (HF SiglipVisionConfig fields the path depends on)."""
    hidden: int = 1024
    ff: int = 4096
    layers: int = 24
    heads: int = 16
    image: int = 384
    patch: int = 16
    eps: float = 1e-6

    @property
    def grid(self) -> int:
        return self.image // self.patch

    @property
    def tokens(self) -> int:
        return self.grid * self.grid

    @property
    def hd(self) -> int:
        return self.hidden // self.heads


@dataclass
class GemmaCfg:
    """Gemma3 text decoder (HF Gemma3TextConfig fields the path depends on)."""
    vocab: int = 262144
    hidden: int = 1152
    ff: int = 6912
    layers: int = 26
    hq: int = 4
    hkv: int = 1
    hd: int = 256
    window: int = 512
    theta_local: float = 10000.0
    theta_global: float = 1000000.0
    rope_factor_global: float = 1.0
    eps: float = 1e-6
    q_scalar: float = 256.0
    pattern: int = 6
    pad_id: int = 0
    # decoder recipe; the defaults are Gemma3.  Qwen3 (HF:models/qwen3/modeling_qwen3.py): no norm on the
    # branch outputs (pre-norm only), SwiGLU, plain embeddings, full attention in every layer (pattern=1),
    # possibly an untied LM head.  Norm gains are packed as (gain - 1) for both families (weights.py).
    sandwich_norm: bool = True
    silu: bool = False
    scale_embed: bool = True
    tie_lm_head: bool = True

    def is_full(self, i: int) -> bool:
        return (i + 1) % self.pattern == 0


SIGLIP_L = SiglipCfg()
GEMMA3_1B = GemmaCfg()
GEMMA3_4B = GemmaCfg(vocab=262208, hidden=2560, ff=10240, layers=34, hq=8, hkv=4, hd=256,
                     window=1024, rope_factor_global=8.0)


QWEN3_8B = GemmaCfg(vocab=151936, hidden=4096, ff=12288, layers=36, hq=32, hkv=8, hd=128, window=0,
                    theta_global=1000000.0, q_scalar=128.0, pattern=1, pad_id=151643, sandwich_norm=False,
                    silu=True, scale_embed=False, tie_lm_head=False)


def siglip_cfg_from_hf(c) -> SiglipCfg:
    """c: transformers SiglipVisionConfig (or SiglipConfig.vision_config)."""
    if getattr(c, "hidden_act", "gelu_pytorch_tanh") != "gelu_pytorch_tanh":
        raise RuntimeError(f"unsupported SigLIP activation {c.hidden_act!r} (only gelu_pytorch_tanh)")
    return SiglipCfg(hidden=c.hidden_size, ff=c.intermediate_size, layers=c.num_hidden_layers,
                     heads=c.num_attention_heads, image=c.image_size, patch=c.patch_size,
                     eps=c.layer_norm_eps)


def qwen3_cfg_from_hf(c) -> GemmaCfg:
    """c: transformers Qwen3Config (dense).  Rejects features the kernels do not implement."""
    if getattr(c, "hidden_act", "silu") != "silu":
        raise RuntimeError(f"unsupported Qwen3 activation {c.hidden_act!r} (only silu)")
    if getattr(c, "attention_bias", False):
        raise RuntimeError("Qwen3 with attention_bias is not implemented")
    types = list(getattr(c, "layer_types", []) or [])
    if getattr(c, "use_sliding_window", False) or any(t != "full_attention" for t in types):
        raise RuntimeError("Qwen3 with sliding-window layers is not implemented")
    theta = 1000000.0
    rp = getattr(c, "rope_parameters", None)
    if isinstance(rp, dict) and rp:
        if rp.get("rope_type", "default") != "default":
            raise RuntimeError(f"unsupported rope_type {rp.get('rope_type')!r}")
        theta = float(rp.get("rope_theta", theta))
    else:
        if getattr(c, "rope_scaling", None):
            raise RuntimeError("unsupported rope scaling")
        theta = float(getattr(c, "rope_theta", theta))
    hd = getattr(c, "head_dim", None) or c.hidden_size // c.num_attention_heads
    return GemmaCfg(vocab=c.vocab_size, hidden=c.hidden_size, ff=c.intermediate_size, layers=c.num_hidden_layers,
                    hq=c.num_attention_heads, hkv=c.num_key_value_heads, hd=hd, window=0, theta_global=theta,
                    eps=c.rms_norm_eps, q_scalar=float(hd), pattern=1,
                    pad_id=c.pad_token_id if c.pad_token_id is not None else 0, sandwich_norm=False, silu=True,
                    scale_embed=False, tie_lm_head=bool(getattr(c, "tie_word_embeddings", False)))


def gemma_cfg_from_hf(c) -> GemmaCfg:
    """c: transformers Gemma3TextConfig.  Rejects features the kernels do not implement."""
    if getattr(c, "model_type", "") not in ("gemma3_text",):
        raise RuntimeError(f"unsupported language model type {getattr(c, 'model_type', None)!r}: "
                           "this build implements the Gemma3 text and Qwen3 dense decoders only (no fallback)")
    if getattr(c, "attn_logit_softcapping", None) or getattr(c, "final_logit_softcapping", None):
        raise RuntimeError("logit soft-capping is not implemented")
    if getattr(c, "hidden_activation", "gelu_pytorch_tanh") != "gelu_pytorch_tanh":
        raise RuntimeError("only gelu_pytorch_tanh GeGLU is implemented")
    layer_types = list(getattr(c, "layer_types", []) or [])
    pattern = 6
    if layer_types:
        fulls = [i for i, t in enumerate(layer_types) if t == "full_attention"]
        pattern = fulls[0] + 1 if fulls else len(layer_types) + 1
        for i, t in enumerate(layer_types):
            if (t == "full_attention") != ((i + 1) % pattern == 0):
                raise RuntimeError("unsupported layer_types pattern")
    theta_l, theta_g, factor = 10000.0, 1000000.0, 1.0
    rp = getattr(c, "rope_parameters", None)
    if isinstance(rp, dict) and "full_attention" in rp:
        theta_g = float(rp["full_attention"].get("rope_theta", theta_g))
        theta_l = float(rp["sliding_attention"].get("rope_theta", theta_l))
        fa = rp["full_attention"]
        if fa.get("rope_type", "default") == "linear":
            factor = float(fa.get("factor", 1.0))
        elif fa.get("rope_type", "default") != "default":
            raise RuntimeError(f"unsupported rope_type {fa.get('rope_type')!r}")
    else:
        theta_g = float(getattr(c, "rope_theta", theta_g))
        theta_l = float(getattr(c, "rope_local_base_freq", theta_l))
        rs = getattr(c, "rope_scaling", None)
        if rs:
            if rs.get("rope_type", rs.get("type")) != "linear":
                raise RuntimeError("unsupported rope scaling")
            factor = float(rs["factor"])
    return GemmaCfg(vocab=c.vocab_size, hidden=c.hidden_size, ff=c.intermediate_size,
                    layers=c.num_hidden_layers, hq=c.num_attention_heads,
                    hkv=c.num_key_value_heads, hd=c.head_dim, window=c.sliding_window,
                    theta_local=theta_l, theta_global=theta_g, rope_factor_global=factor,
                    eps=c.rms_norm_eps, q_scalar=float(c.query_pre_attn_scalar), pattern=pattern,
                    pad_id=c.pad_token_id if c.pad_token_id is not None else 0)
