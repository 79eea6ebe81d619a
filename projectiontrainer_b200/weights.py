"""Weight containers -> the flat name scheme the engine packs from.

HF modules are accepted as *weight containers only*: their parameters are read once and
repacked (fused qkv / gate-up matrices, bf16 matrices, fp32 vectors); none of their forward
code runs on the hot path.  Unsupported architectures are rejected loudly.
"""
from __future__ import annotations

import torch

from .config import GemmaCfg, SiglipCfg, gemma_cfg_from_hf, qwen3_cfg_from_hf, siglip_cfg_from_hf


def siglip_from_hf(vision_model) -> tuple[dict, SiglipCfg]:
    """vision_model: HF SiglipVisionTransformer (`SiglipModel.vision_model` /
    `SiglipVisionModel.vision_model`), the object ref:Stage1/projector_trainer.py:165-171 calls."""
    cfg = siglip_cfg_from_hf(vision_model.config)
    sd = {k: v.detach() for k, v in vision_model.state_dict().items()}
    d = cfg.hidden
    W = {"patch_w": sd["embeddings.patch_embedding.weight"].reshape(d, -1),
         "patch_b": sd["embeddings.patch_embedding.bias"],
         "pos": sd["embeddings.position_embedding.weight"],
         "post_g": sd["post_layernorm.weight"], "post_b": sd["post_layernorm.bias"]}
    for l in range(cfg.layers):
        s, p = f"encoder.layers.{l}.", f"l{l}."
        W[p + "ln1_g"], W[p + "ln1_b"] = sd[s + "layer_norm1.weight"], sd[s + "layer_norm1.bias"]
        W[p + "ln2_g"], W[p + "ln2_b"] = sd[s + "layer_norm2.weight"], sd[s + "layer_norm2.bias"]
        W[p + "qkv_w"] = torch.cat([sd[s + f"self_attn.{n}_proj.weight"] for n in "qkv"], 0)
        W[p + "qkv_b"] = torch.cat([sd[s + f"self_attn.{n}_proj.bias"] for n in "qkv"], 0)
        W[p + "o_w"], W[p + "o_b"] = sd[s + "self_attn.out_proj.weight"], sd[s + "self_attn.out_proj.bias"]
        W[p + "fc1_w"], W[p + "fc1_b"] = sd[s + "mlp.fc1.weight"], sd[s + "mlp.fc1.bias"]
        W[p + "fc2_w"], W[p + "fc2_b"] = sd[s + "mlp.fc2.weight"], sd[s + "mlp.fc2.bias"]
    return W, cfg


def gemma_from_hf(causal_lm) -> tuple[dict, GemmaCfg]:
    """causal_lm: HF Gemma3ForCausalLM (what AutoModelForCausalLM returns for gemma3_text,
    ref:Stage1/train_projection_stage1.py:204)."""
    cfg = gemma_cfg_from_hf(causal_lm.config)
    sd = {k: v.detach() for k, v in causal_lm.state_dict().items()}
    W = {"embed": sd["model.embed_tokens.weight"], "final_norm": sd["model.norm.weight"]}
    if "lm_head.weight" in sd and sd["lm_head.weight"].data_ptr() != sd["model.embed_tokens.weight"].data_ptr():
        if not torch.equal(sd["lm_head.weight"], sd["model.embed_tokens.weight"]):
            raise RuntimeError("untied lm_head is not supported (Gemma3 ties it to embed_tokens)")
    for l in range(cfg.layers):
        s, p = f"model.layers.{l}.", f"l{l}."
        W[p + "in_norm"] = sd[s + "input_layernorm.weight"]
        W[p + "post_attn_norm"] = sd[s + "post_attention_layernorm.weight"]
        W[p + "pre_ffn_norm"] = sd[s + "pre_feedforward_layernorm.weight"]
        W[p + "post_ffn_norm"] = sd[s + "post_feedforward_layernorm.weight"]
        W[p + "q_norm"], W[p + "k_norm"] = sd[s + "self_attn.q_norm.weight"], sd[s + "self_attn.k_norm.weight"]
        W[p + "qkv_w"] = torch.cat([sd[s + f"self_attn.{n}_proj.weight"] for n in "qkv"], 0)
        W[p + "o_w"] = sd[s + "self_attn.o_proj.weight"]
        W[p + "gu_w"] = torch.cat([sd[s + "mlp.gate_proj.weight"], sd[s + "mlp.up_proj.weight"]], 0)
        W[p + "down_w"] = sd[s + "mlp.down_proj.weight"]
    return W, cfg


def qwen3_from_hf(causal_lm) -> tuple[dict, GemmaCfg]:
    """causal_lm: HF Qwen3ForCausalLM (what AutoModelForCausalLM returns for Qwen/Qwen3-*, the LLM of
    ref:Stage1/run_projection_train_stage1.sh:12).  Qwen3RMSNorm multiplies by `weight`, the norm kernels by
    (1 + gain): gains are packed as weight - 1.  `input_layernorm` -> in_norm, `post_attention_layernorm` ->
    pre_ffn_norm (it is the MLP's pre-norm); there are no norms on the branch outputs."""
    cfg = qwen3_cfg_from_hf(causal_lm.config)
    sd = {k: v.detach() for k, v in causal_lm.state_dict().items()}
    m1 = lambda t: t.float() - 1.0
    W = {"embed": sd["model.embed_tokens.weight"], "final_norm": m1(sd["model.norm.weight"])}
    head = sd.get("lm_head.weight")
    if head is not None and head.data_ptr() != W["embed"].data_ptr() and not torch.equal(head, W["embed"]):
        W["lm_head"] = head
        cfg.tie_lm_head = False
    else:
        cfg.tie_lm_head = True
    for l in range(cfg.layers):
        s, p = f"model.layers.{l}.", f"l{l}."
        W[p + "in_norm"] = m1(sd[s + "input_layernorm.weight"])
        W[p + "pre_ffn_norm"] = m1(sd[s + "post_attention_layernorm.weight"])
        W[p + "q_norm"], W[p + "k_norm"] = m1(sd[s + "self_attn.q_norm.weight"]), m1(sd[s + "self_attn.k_norm.weight"])
        W[p + "qkv_w"] = torch.cat([sd[s + f"self_attn.{n}_proj.weight"] for n in "qkv"], 0)
        W[p + "o_w"] = sd[s + "self_attn.o_proj.weight"]
        W[p + "gu_w"] = torch.cat([sd[s + "mlp.gate_proj.weight"], sd[s + "mlp.up_proj.weight"]], 0)
        W[p + "down_w"] = sd[s + "mlp.down_proj.weight"]
    return W, cfg


class _BaseView:
    """A peft-wrapped causal LM seen as its base model: `state_dict()` with peft's `base_model.model.` prefix and
    `.base_layer` infixes stripped and the adapter tensors left out (those go to lora.LoraState)."""

    def __init__(self, peft_model):
        self._m, self.config = peft_model, peft_model.config

    def state_dict(self):
        out = {}
        for k, v in self._m.state_dict().items():
            if ".lora_" in k:
                continue
            k = k[len("base_model.model."):] if k.startswith("base_model.model.") else k
            out[k.replace(".base_layer.", ".")] = v
        return out


def llm_from_hf(causal_lm) -> tuple[dict, GemmaCfg]:
    """Dispatch on the HF model type: Gemma3 text and Qwen3 dense decoders; anything else raises."""
    if any(getattr(p, "quant_state", None) is not None or p.dtype == torch.uint8 for p in causal_lm.parameters()):
        raise RuntimeError("the language model holds bitsandbytes 4-bit weights: pass the bf16 model — with "
                           "enable_qlora the NF4 quantise / de-quantise pass runs here (weights.nf4_frozen_llm_)")
    if getattr(causal_lm, "peft_config", None):
        causal_lm = _BaseView(causal_lm)
    mt = getattr(causal_lm.config, "model_type", None)
    if mt == "qwen3":
        return qwen3_from_hf(causal_lm)
    return gemma_from_hf(causal_lm)


# ----------------------------------------------------------------------------- synthetic weights
def _r(gen, device, *shape, std=0.02):
    return torch.randn(*shape, generator=gen, device=device, dtype=torch.float32) * std


def random_siglip(cfg: SiglipCfg, device, seed: int = 0) -> dict:
    """Random-init SigLIP weights generated on `device` (benchmarks: no checkpoints, no network)."""
    g = torch.Generator(device=device).manual_seed(seed)
    d, f = cfg.hidden, cfg.ff
    W = {"patch_w": _r(g, device, d, 3 * cfg.patch ** 2, std=(3 * cfg.patch ** 2) ** -0.5),
         "patch_b": _r(g, device, d), "pos": _r(g, device, cfg.tokens, d, std=d ** -0.5),
         "post_g": 1 + _r(g, device, d, std=0.1), "post_b": _r(g, device, d, std=0.1)}
    for l in range(cfg.layers):
        p = f"l{l}."
        W[p + "ln1_g"], W[p + "ln1_b"] = 1 + _r(g, device, d, std=0.1), _r(g, device, d, std=0.1)
        W[p + "ln2_g"], W[p + "ln2_b"] = 1 + _r(g, device, d, std=0.1), _r(g, device, d, std=0.1)
        W[p + "qkv_w"], W[p + "qkv_b"] = _r(g, device, 3 * d, d, std=d ** -0.5), _r(g, device, 3 * d)
        W[p + "o_w"], W[p + "o_b"] = _r(g, device, d, d, std=d ** -0.5), _r(g, device, d)
        W[p + "fc1_w"], W[p + "fc1_b"] = _r(g, device, f, d, std=d ** -0.5), _r(g, device, f)
        W[p + "fc2_w"], W[p + "fc2_b"] = _r(g, device, d, f, std=f ** -0.5), _r(g, device, d)
    return W


def random_gemma(cfg: GemmaCfg, device, seed: int = 0) -> dict:
    g = torch.Generator(device=device).manual_seed(seed + 1)
    h = cfg.hidden
    W = {"embed": _r(g, device, cfg.vocab, h).to(torch.bfloat16), "final_norm": _r(g, device, h, std=0.3)}
    if not cfg.tie_lm_head:
        W["lm_head"] = _r(g, device, cfg.vocab, h).to(torch.bfloat16)
    for l in range(cfg.layers):
        p = f"l{l}."
        for n in (("in_norm", "post_attn_norm", "pre_ffn_norm", "post_ffn_norm") if cfg.sandwich_norm
                  else ("in_norm", "pre_ffn_norm")):
            W[p + n] = _r(g, device, h, std=0.3)
        W[p + "q_norm"], W[p + "k_norm"] = _r(g, device, cfg.hd, std=0.3), _r(g, device, cfg.hd, std=0.3)
        W[p + "qkv_w"] = _r(g, device, (cfg.hq + 2 * cfg.hkv) * cfg.hd, h).to(torch.bfloat16)
        W[p + "o_w"] = _r(g, device, h, cfg.hq * cfg.hd).to(torch.bfloat16)
        W[p + "gu_w"] = _r(g, device, 2 * cfg.ff, h).to(torch.bfloat16)
        W[p + "down_w"] = _r(g, device, h, cfg.ff).to(torch.bfloat16)
    return W


def random_projector(vision_dim: int, llm_dim: int, device, expansion: int = 10, seed: int = 0) -> dict:
    """nn.Linear default init: U(-1/sqrt(fan_in), 1/sqrt(fan_in)) for weight and bias."""
    g = torch.Generator(device=device).manual_seed(seed + 2)
    mid = vision_dim * expansion
    u = lambda *s, bound: (torch.rand(*s, generator=g, device=device) * 2 - 1) * bound
    return {"model.0.weight": u(mid, vision_dim, bound=vision_dim ** -0.5),
            "model.0.bias": u(mid, bound=vision_dim ** -0.5),
            "model.2.weight": u(llm_dim, mid, bound=mid ** -0.5),
            "model.2.bias": u(llm_dim, bound=mid ** -0.5)}


# ----------------------------------------------------------------------------- NF4 frozen LLM (--enable_qlora)
def nf4_frozen_llm_(W: dict, cfg: GemmaCfg, device, via_fp16: bool = True) -> dict:
    """What `--enable_qlora` does to the frozen LLM in Stage 1 (ref:Stage1/train_projection_stage1.py:192-210):
    every decoder `nn.Linear` weight (q/k/v/o/gate/up/down — bitsandbytes leaves embeddings, norms and the LM head
    alone) is stored as double-quantised NF4 and expanded again for the GEMM.  Here each such weight is replaced, in
    place in the packed dict, by its quantise -> de-quantise image in bf16 on `device` (one pass at load time; 180 GB
    of HBM keeps the bf16 copy resident, so the step itself is unchanged).  bitsandbytes quantises each Linear on its
    own (its own 64-blocks, 256-blocks and mean offset): the fused qkv / gate||up matrices are split back into the
    HF modules' row ranges first.  Returns {weight name: relative Frobenius error} for logging."""
    from . import ops
    q_rows, kv_rows = cfg.hq * cfg.hd, cfg.hkv * cfg.hd
    report = {}
    for l in range(cfg.layers):
        p = f"l{l}."
        for name, splits in ((p + "qkv_w", (q_rows, kv_rows, kv_rows)), (p + "o_w", None),
                             (p + "gu_w", (cfg.ff, cfg.ff)), (p + "down_w", None)):
            w = W[name].to(device=device, dtype=torch.bfloat16).contiguous()
            out = torch.empty_like(w)
            r0 = 0
            for rows in (splits or (w.shape[0],)):
                ops.nf4_dequantize(ops.nf4_quantize(w[r0:r0 + rows], via_fp16), out[r0:r0 + rows])
                r0 += rows
            assert r0 == w.shape[0], (name, r0, tuple(w.shape))
            report[name] = float((out.float() - w.float()).norm() / w.float().norm().clamp_min(1e-30))
            W[name] = out
    return report
