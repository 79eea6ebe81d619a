"""HF / PyTorch library path of the Stage-1 step on the GPU, for bench.py's `library_bar` and `parity` blocks and for
tools/bench_hf_gpu.py.  The step body restates ref:Stage1/projector_trainer.py:157-245 on stock HF transformers
modules (cuBLASLt GEMMs, SDPA attention, torch autograd); none of this repository's kernels run here and nothing
here is on the product path.

  hf_from_weights   HF SiglipVisionModel / Gemma3ForCausalLM holding a given weight dict (the package's flat name
                    scheme, weights.py) — so that both sides of a parity check compute with the same parameters
  hf_step           one forward + backward of the reference loop, fp32 or under bf16 autocast
  library_bar       the timed bf16-autocast training step at BASELINE.json configs[2] (random init on the device)
"""
from __future__ import annotations

import torch
import torch.nn as nn


def hf_from_weights(Wv: dict, vc, Wl: dict, lc, device, dtype):
    """vc / lc: projectiontrainer_b200.config.SiglipCfg / GemmaCfg (Gemma3 recipe only)."""
    from transformers import Gemma3ForCausalLM, Gemma3TextConfig, SiglipVisionConfig, SiglipVisionModel
    if not (lc.sandwich_norm and lc.scale_embed and lc.tie_lm_head and not lc.silu):
        raise RuntimeError("hf_from_weights: Gemma3 decoder recipe only")
    d = vc.hidden
    with torch.device(device):
        vision = SiglipVisionModel(SiglipVisionConfig(
            hidden_size=vc.hidden, intermediate_size=vc.ff, num_hidden_layers=vc.layers,
            num_attention_heads=vc.heads, image_size=vc.image, patch_size=vc.patch, layer_norm_eps=vc.eps))
        rope = {"full_attention": ({"rope_type": "linear", "factor": lc.rope_factor_global, "rope_theta": lc.theta_global}
                                   if lc.rope_factor_global != 1.0 else
                                   {"rope_type": "default", "rope_theta": lc.theta_global}),
                "sliding_attention": {"rope_type": "default", "rope_theta": lc.theta_local}}
        llm = Gemma3ForCausalLM(Gemma3TextConfig(
            vocab_size=lc.vocab, hidden_size=lc.hidden, intermediate_size=lc.ff, num_hidden_layers=lc.layers,
            num_attention_heads=lc.hq, num_key_value_heads=lc.hkv, head_dim=lc.hd, sliding_window=lc.window,
            query_pre_attn_scalar=int(lc.q_scalar), max_position_embeddings=32768, rms_norm_eps=lc.eps,
            pad_token_id=lc.pad_id, rope_parameters=rope))
    sv = {"embeddings.patch_embedding.weight": Wv["patch_w"].reshape(d, 3, vc.patch, vc.patch),
          "embeddings.patch_embedding.bias": Wv["patch_b"], "embeddings.position_embedding.weight": Wv["pos"],
          "post_layernorm.weight": Wv["post_g"], "post_layernorm.bias": Wv["post_b"]}
    for l in range(vc.layers):
        s, p = f"encoder.layers.{l}.", f"l{l}."
        sv[s + "layer_norm1.weight"], sv[s + "layer_norm1.bias"] = Wv[p + "ln1_g"], Wv[p + "ln1_b"]
        sv[s + "layer_norm2.weight"], sv[s + "layer_norm2.bias"] = Wv[p + "ln2_g"], Wv[p + "ln2_b"]
        for i, n in enumerate("qkv"):
            sv[s + f"self_attn.{n}_proj.weight"] = Wv[p + "qkv_w"][i * d:(i + 1) * d]
            sv[s + f"self_attn.{n}_proj.bias"] = Wv[p + "qkv_b"][i * d:(i + 1) * d]
        sv[s + "self_attn.out_proj.weight"], sv[s + "self_attn.out_proj.bias"] = Wv[p + "o_w"], Wv[p + "o_b"]
        sv[s + "mlp.fc1.weight"], sv[s + "mlp.fc1.bias"] = Wv[p + "fc1_w"], Wv[p + "fc1_b"]
        sv[s + "mlp.fc2.weight"], sv[s + "mlp.fc2.bias"] = Wv[p + "fc2_w"], Wv[p + "fc2_b"]
    missing, unexpected = vision.vision_model.load_state_dict(sv, strict=False)
    assert not unexpected and all(k.startswith("head.") for k in missing), (missing, unexpected)
    qd, kd = lc.hq * lc.hd, lc.hkv * lc.hd
    sl = {"model.embed_tokens.weight": Wl["embed"], "lm_head.weight": Wl["embed"], "model.norm.weight": Wl["final_norm"]}
    for l in range(lc.layers):
        s, p = f"model.layers.{l}.", f"l{l}."
        sl[s + "input_layernorm.weight"] = Wl[p + "in_norm"]
        sl[s + "post_attention_layernorm.weight"] = Wl[p + "post_attn_norm"]
        sl[s + "pre_feedforward_layernorm.weight"] = Wl[p + "pre_ffn_norm"]
        sl[s + "post_feedforward_layernorm.weight"] = Wl[p + "post_ffn_norm"]
        sl[s + "self_attn.q_norm.weight"], sl[s + "self_attn.k_norm.weight"] = Wl[p + "q_norm"], Wl[p + "k_norm"]
        sl[s + "self_attn.q_proj.weight"] = Wl[p + "qkv_w"][:qd]
        sl[s + "self_attn.k_proj.weight"] = Wl[p + "qkv_w"][qd:qd + kd]
        sl[s + "self_attn.v_proj.weight"] = Wl[p + "qkv_w"][qd + kd:]
        sl[s + "self_attn.o_proj.weight"] = Wl[p + "o_w"]
        sl[s + "mlp.gate_proj.weight"] = Wl[p + "gu_w"][:lc.ff]
        sl[s + "mlp.up_proj.weight"] = Wl[p + "gu_w"][lc.ff:]
        sl[s + "mlp.down_proj.weight"] = Wl[p + "down_w"]
    missing, unexpected = llm.load_state_dict(sl, strict=False)
    assert not unexpected and not missing, (missing, unexpected)
    vision, llm = vision.to(dtype).eval(), llm.to(dtype).eval()
    for m in (vision, llm):
        m.requires_grad_(False)
    return vision, llm


def hf_step(vision, llm, proj, px, ids, labels, pad_id, autocast: bool, loss_div: float = 1.0):
    """ref:Stage1/projector_trainer.py:157-237: returns (loss tensor, nothing else); gradients land on proj."""
    dev = px.device
    B = px.shape[0]
    wdt = next(vision.parameters()).dtype
    ctx = torch.autocast("cuda", dtype=torch.bfloat16) if autocast else torch.autocast("cuda", enabled=False)
    with ctx:
        with torch.no_grad():
            patch = vision.vision_model(pixel_values=px.to(wdt)).last_hidden_state[:, 1:, :]
        projected = proj(patch.to(next(proj.parameters()).dtype) if not autocast else patch).float()
        with torch.no_grad():
            text = llm.get_input_embeddings()(ids)
        inputs = torch.cat([projected, text.to(projected.dtype)], dim=1)
        n_img = projected.shape[1]
        mask = torch.cat([torch.ones(B, n_img, dtype=torch.long, device=dev),
                          (ids != pad_id).long() if pad_id is not None else torch.ones_like(ids)], dim=1)
        lm_labels = torch.cat([torch.full((B, n_img), -100, dtype=torch.long, device=dev), labels], dim=1)
        if not autocast:
            inputs = inputs.to(next(llm.parameters()).dtype)
        loss = llm(inputs_embeds=inputs, attention_mask=mask, labels=lm_labels, return_dict=True).loss
    (loss / loss_div).backward()
    return loss


def library_bar(batch: int, tokens: int, steps: int, warmup: int, device) -> dict:
    """The step executed the reference's way on this GPU: HF modules in bf16 under autocast, fp32 projector,
    torch.optim.AdamW, clip 5.0 — BASELINE.json configs[2] geometry, random init, synthetic data."""
    import transformers
    from transformers import Gemma3ForCausalLM, Gemma3TextConfig, SiglipVisionConfig, SiglipVisionModel
    torch.manual_seed(0)
    with torch.device(device):
        vision = SiglipVisionModel(SiglipVisionConfig(hidden_size=1024, intermediate_size=4096, num_hidden_layers=24,
                                                      num_attention_heads=16, image_size=384, patch_size=16))
        llm = Gemma3ForCausalLM(Gemma3TextConfig(vocab_size=262144, hidden_size=1152, intermediate_size=6912,
                                                 num_hidden_layers=26, num_attention_heads=4, num_key_value_heads=1,
                                                 head_dim=256, sliding_window=512, query_pre_attn_scalar=256,
                                                 max_position_embeddings=32768, pad_token_id=0))
        proj = nn.Sequential(nn.Linear(1024, 10240), nn.GELU(), nn.Linear(10240, 1152)).float()
    vision, llm = vision.to(torch.bfloat16).eval(), llm.to(torch.bfloat16).eval()
    for m in (vision, llm):
        m.requires_grad_(False)
    opt = torch.optim.AdamW(proj.parameters(), lr=1e-4, weight_decay=0.01)
    g = torch.Generator(device=device).manual_seed(1234)
    px = torch.rand(batch, 3, 384, 384, generator=g, device=device) * 2 - 1
    ids = torch.randint(3, 262144, (batch, tokens), generator=g, device=device)
    ids[:, 0] = 2

    def step():
        loss = hf_step(vision, llm, proj, px, ids, ids, 0, autocast=True)
        torch.nn.utils.clip_grad_norm_(proj.parameters(), 5.0)
        opt.step()
        opt.zero_grad()
        return loss

    torch.cuda.reset_peak_memory_stats(device)
    for _ in range(warmup):
        loss = step()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        loss = step()
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    out = {"what": f"the same step through HF transformers {transformers.__version__} + torch {torch.__version__} "
                   f"(cuBLASLt GEMMs, {getattr(llm.config, '_attn_implementation', '?')} attention, bf16 autocast, "
                   f"torch autograd + AdamW) on this GPU, B={batch}, T={tokens}, random init",
           "value": batch / ms * 1e3, "unit": "images/s", "ms_per_step": ms, "steps": steps, "warmup": warmup,
           "loss": float(loss.item()), "mem_GB": torch.cuda.max_memory_allocated(device) / 2 ** 30}
    del vision, llm, proj, opt
    torch.cuda.empty_cache()
    return out
