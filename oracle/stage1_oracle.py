"""CPU oracle for the Stage-1 projector-alignment step.  TEST INFRASTRUCTURE ONLY.

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference
legs may import this module; the product path (projectiontrainer_b200/) never does.

This is a restatement, in plain torch-on-CPU tensor algebra, of the arithmetic the
reference executes for one Stage-1 step.  The reference itself is pure Python that
delegates the model math to HF transformers (not vendored under /root/reference;
pinned transformers==4.51.0, requirements.txt:6; the copy in this image is 5.5.0),
so each function cites both the reference call site ("ref:") and the HF source it
restates ("HF:", relative to the transformers package).

Parity pin: the reference ships no tests or golden vectors, so the
oracle is pinned against outputs of the reference run in the build container:
oracle/make_golden.py imports the reference's own Stage1/projectors.py and executes
Stage1/projector_trainer.py:157-237 semantics on HF SiglipVisionModel /
Gemma3ForCausalLM, and tests/test_oracle_golden.py checks this restatement against
the committed vectors (tests/golden/*.npz).  Gradients come from torch autograd over
the restated forward; the projector backward is also written out by hand
(projector_backward) as the spec the CUDA kernels follow.
"""
from __future__ import annotations

import math
from dataclasses import dataclass, field

import numpy as np
import torch
import torch.nn.functional as F


# --------------------------------------------------------------------------- configs
@dataclass
class SiglipCfg:
    hidden: int = 1024
    ff: int = 4096
    layers: int = 24
    heads: int = 16
    image: int = 384
    patch: int = 16
    eps: float = 1e-6

    @property
    def grid(self):
        return self.image // self.patch

    @property
    def tokens(self):
        return self.grid * self.grid


@dataclass
class GemmaCfg:
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
    rope_factor_global: float = 1.0   # 8.0 for the 4B/12B/27B text configs
    eps: float = 1e-6
    q_scalar: float = 256.0           # query_pre_attn_scalar
    pattern: int = 6                  # layer i is full-attention iff (i+1) % pattern == 0
    pad_id: int = 0
    # decoder recipe.  Defaults = Gemma3.  Qwen3 (HF:models/qwen3/modeling_qwen3.py): no norms on the branch
    # outputs, SwiGLU, unscaled embeddings, every layer full attention (pattern=1), optionally an untied LM
    # head (W["lm_head"]).  Norm gains are stored as (gain - 1) in both families so rms_norm is shared.
    sandwich_norm: bool = True
    silu: bool = False
    scale_embed: bool = True
    tie_lm_head: bool = True

    def is_full(self, i: int) -> bool:
        return (i + 1) % self.pattern == 0


def qwen3_cfg(**kw) -> "GemmaCfg":
    """A Qwen3-recipe decoder config (full attention everywhere, q_scalar = head_dim)."""
    kw.setdefault("pattern", 1)
    kw.setdefault("q_scalar", float(kw.get("hd", 128)))
    kw.setdefault("theta_global", 1000000.0)
    return GemmaCfg(sandwich_norm=False, silu=True, scale_embed=False, **kw)


SIGLIP_L = SiglipCfg()
GEMMA3_1B = GemmaCfg()
GEMMA3_4B = GemmaCfg(vocab=262208, hidden=2560, ff=10240, layers=34, hq=8, hkv=4, hd=256,
                     window=1024, rope_factor_global=8.0)


# --------------------------------------------------------------------------- SigLIP
def gelu_tanh(z):
    # HF:activations.py (gelu_pytorch_tanh) — 0.5 z (1 + tanh(sqrt(2/pi) (z + 0.044715 z^3)))
    return 0.5 * z * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (z + 0.044715 * z ** 3)))


def patchify(pixels: torch.Tensor, patch: int) -> torch.Tensor:
    """[B,3,H,W] -> [B, (H/p)(W/p), 3*p*p] with column order (c, py, px).

    Equals the im2col of Conv2d(3,d,kernel=p,stride=p) against weight.flatten(1)
    (HF:models/siglip/modeling_siglip.py:124-130,178-180)."""
    B, Cc, H, W = pixels.shape
    gy, gx = H // patch, W // patch
    x = pixels.reshape(B, Cc, gy, patch, gx, patch).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(B, gy * gx, Cc * patch * patch)


def layer_norm(x, g, b, eps):
    mu = x.mean(-1, keepdim=True)
    var = ((x - mu) ** 2).mean(-1, keepdim=True)     # biased variance
    return (x - mu) / torch.sqrt(var + eps) * g + b


def siglip_forward(pixels: torch.Tensor, W: dict, cfg: SiglipCfg) -> torch.Tensor:
    """Frozen SigLIP vision tower -> post-LayerNorm features [B, tokens, hidden]
    (ref:Stage1/projector_trainer.py:167-171; HF:models/siglip/modeling_siglip.py:175-185
    embeddings, :340-361 encoder layer, :275-311 attention, :323-327 MLP, :617-618 post LN).
    The pooling head (:620) is not needed by Stage 1 and is not computed."""
    x = patchify(pixels, cfg.patch) @ W["patch_w"].T + W["patch_b"] + W["pos"]
    B, N, d = x.shape
    hd = d // cfg.heads
    for l in range(cfg.layers):
        p = f"l{l}."
        a = layer_norm(x, W[p + "ln1_g"], W[p + "ln1_b"], cfg.eps)
        qkv = a @ W[p + "qkv_w"].T + W[p + "qkv_b"]              # fused [3d, d] = (q; k; v)
        q, k, v = qkv.split(d, dim=-1)
        q = q.reshape(B, N, cfg.heads, hd).transpose(1, 2)
        k = k.reshape(B, N, cfg.heads, hd).transpose(1, 2)
        v = v.reshape(B, N, cfg.heads, hd).transpose(1, 2)
        s = (q @ k.transpose(-1, -2)) * hd ** -0.5
        o = torch.softmax(s.float(), dim=-1).to(x.dtype) @ v
        o = o.transpose(1, 2).reshape(B, N, d)
        x = x + o @ W[p + "o_w"].T + W[p + "o_b"]
        m = layer_norm(x, W[p + "ln2_g"], W[p + "ln2_b"], cfg.eps)
        x = x + gelu_tanh(m @ W[p + "fc1_w"].T + W[p + "fc1_b"]) @ W[p + "fc2_w"].T + W[p + "fc2_b"]
    return layer_norm(x, W["post_g"], W["post_b"], cfg.eps)


# --------------------------------------------------------------------------- projector
def gelu_erf(z):
    return 0.5 * z * (1.0 + torch.erf(z / math.sqrt(2.0)))


def projector_forward(x, P: dict):
    """ref:Stage1/projectors.py:16-20,29 — Linear(+bias) -> exact (erf) GELU -> Linear(+bias)."""
    z = x @ P["model.0.weight"].T + P["model.0.bias"]
    return gelu_erf(z) @ P["model.2.weight"].T + P["model.2.bias"]


def projector_backward(x, P: dict, dY):
    """Hand-written backward of projector_forward (what autograd does at
    ref:Stage1/projector_trainer.py:237); x: [R, vd], dY: [R, h].  No gradient to x."""
    z = x @ P["model.0.weight"].T + P["model.0.bias"]
    hm = gelu_erf(z)
    dW2 = dY.T @ hm
    db2 = dY.sum(0)
    dH = dY @ P["model.2.weight"]
    cdf = 0.5 * (1.0 + torch.erf(z / math.sqrt(2.0)))
    pdf = torch.exp(-0.5 * z * z) / math.sqrt(2.0 * math.pi)
    dZ = dH * (cdf + z * pdf)
    return {"model.0.weight": dZ.T @ x, "model.0.bias": dZ.sum(0),
            "model.2.weight": dW2, "model.2.bias": db2}


# --------------------------------------------------------------------------- splice (integer, bit-exact)
def splice_masks_labels(token_ids: np.ndarray, labels: np.ndarray, n_img: int, pad_id):
    """ref:Stage1/projector_trainer.py:198-220 — attention mask and LM labels of the spliced
    sequence [image tokens ; caption tokens].  int64 in, int64 out, bit-exact."""
    B, T = token_ids.shape
    vis_mask = np.ones((B, n_img), dtype=np.int64)
    if pad_id is None:
        txt_mask = np.ones_like(token_ids, dtype=np.int64)
    else:
        txt_mask = (token_ids != pad_id).astype(np.int64)
    attn = np.concatenate([vis_mask, txt_mask], axis=1)
    lm_labels = np.concatenate([np.full((B, n_img), -100, dtype=labels.dtype), labels], axis=1)
    return attn, lm_labels


def shifted_targets(lm_labels: np.ndarray) -> np.ndarray:
    """HF:loss/loss_utils.py:57-60 — labels = pad(labels,(0,1),-100)[...,1:]."""
    B, S = lm_labels.shape
    out = np.full((B, S), -100, dtype=lm_labels.dtype)
    out[:, :-1] = lm_labels[:, 1:]
    return out


def embed_tokens(token_ids: torch.Tensor, E: torch.Tensor, hidden: int, scale_dtype=None, scale: bool = True):
    """HF:models/gemma3/modeling_gemma3.py:102-113 — embedding * sqrt(hidden), the scale being
    rounded to the embedding dtype first (scale_dtype=torch.bfloat16 emulates a bf16 checkpoint
    whose values are held in a wider tensor: sqrt(1152) = 33.94 -> 34.0)."""
    if not scale:               # Qwen3: plain nn.Embedding
        return E[token_ids]
    scale = torch.tensor(hidden ** 0.5, dtype=scale_dtype or E.dtype).to(E.dtype)
    return E[token_ids] * scale


# --------------------------------------------------------------------------- Gemma3 decoder
def rms_norm(x, w, eps):
    """HF:models/gemma3/modeling_gemma3.py:138-146 — fp32 inside, (1 + w) gain."""
    xf = x.float() if x.dtype in (torch.bfloat16, torch.float16) else x
    y = xf * torch.rsqrt(xf.pow(2).mean(-1, keepdim=True) + eps)
    return (y * (1.0 + w.to(y.dtype))).to(x.dtype)


def rope_tables(S: int, hd: int, theta: float, factor: float = 1.0, dtype=torch.float32):
    """HF:models/gemma3/modeling_gemma3.py:216-230 — cos/sin [S, hd] (each half duplicated)."""
    inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float32) / hd))
    inv = inv / factor
    ang = torch.arange(S, dtype=torch.float32)[:, None] * inv[None, :]
    emb = torch.cat([ang, ang], dim=-1)
    return emb.cos().to(dtype), emb.sin().to(dtype)


def rotate_half(u):
    h = u.shape[-1] // 2
    return torch.cat([-u[..., h:], u[..., :h]], dim=-1)


def allowed_mask(keyvalid: torch.Tensor, S: int, window: int | None):
    """HF:masking_utils.py:74-99,148-159 — causal AND key-not-padded AND (full OR kv > q - W)."""
    q = torch.arange(S)[:, None]
    k = torch.arange(S)[None, :]
    m = k <= q
    if window is not None:
        m = m & (k > q - window)
    return m[None, :, :] & keyvalid.bool()[:, None, :]         # [B, S, S]


def gemma3_hidden(X0, keyvalid, W: dict, cfg: GemmaCfg, pos_ids=None):
    """Decoder stack without the final norm (HF:models/gemma3/modeling_gemma3.py:521-583,
    decoder layer :404-434, attention :346-388, MLP :127-129).  Positions are arange(S) as in a
    plain forward (F3); `pos_ids` i64 [B, S] gives per-sample positions instead (what `generate`
    derives from a padded attention mask, HF:generation/utils.py prepare_inputs_for_generation)."""
    B, S, h = X0.shape
    G = cfg.hq // cfg.hkv
    dt = X0.dtype
    cos_l, sin_l = rope_tables(S, cfg.hd, cfg.theta_local, 1.0, dt)
    cos_g, sin_g = rope_tables(S, cfg.hd, cfg.theta_global, cfg.rope_factor_global, dt)
    if pos_ids is None:
        cos_l, sin_l, cos_g, sin_g = (t[None] for t in (cos_l, sin_l, cos_g, sin_g))
    else:
        cos_l, sin_l, cos_g, sin_g = (t[pos_ids] for t in (cos_l, sin_l, cos_g, sin_g))   # [B, S, hd]
    m_full = allowed_mask(keyvalid, S, None)
    m_slide = allowed_mask(keyvalid, S, cfg.window)
    x = X0
    for l in range(cfg.layers):
        p = f"l{l}."
        full = cfg.is_full(l)
        cos, sin = (cos_g, sin_g) if full else (cos_l, sin_l)
        mask = m_full if full else m_slide
        a = rms_norm(x, W[p + "in_norm"], cfg.eps)
        qkv = a @ W[p + "qkv_w"].T                                # fused [(hq+2hkv)*hd, h]
        q, k, v = qkv.split([cfg.hq * cfg.hd, cfg.hkv * cfg.hd, cfg.hkv * cfg.hd], dim=-1)
        q = q.reshape(B, S, cfg.hq, cfg.hd)
        k = k.reshape(B, S, cfg.hkv, cfg.hd)
        v = v.reshape(B, S, cfg.hkv, cfg.hd)
        q = rms_norm(q, W[p + "q_norm"], cfg.eps)
        k = rms_norm(k, W[p + "k_norm"], cfg.eps)
        q = q * cos[:, :, None, :] + rotate_half(q) * sin[:, :, None, :]
        k = k * cos[:, :, None, :] + rotate_half(k) * sin[:, :, None, :]
        q = q.transpose(1, 2)                                     # [B, hq, S, hd]
        k = k.transpose(1, 2).repeat_interleave(G, dim=1)
        v = v.transpose(1, 2).repeat_interleave(G, dim=1)
        s = (q @ k.transpose(-1, -2)) * cfg.q_scalar ** -0.5
        s = s.masked_fill(~mask[:, None, :, :], float("-inf"))
        pr = torch.softmax(s.float(), dim=-1).to(dt)
        pr = torch.nan_to_num(pr, nan=0.0)                        # fully masked (pad) rows: unused
        o = (pr @ v).transpose(1, 2).reshape(B, S, cfg.hq * cfg.hd)
        o = o @ W[p + "o_w"].T
        x = x + (rms_norm(o, W[p + "post_attn_norm"], cfg.eps) if cfg.sandwich_norm else o)
        f = rms_norm(x, W[p + "pre_ffn_norm"], cfg.eps)            # Qwen3: post_attention_layernorm
        gu = f @ W[p + "gu_w"].T                                  # fused [2*ff, h] = (gate; up)
        g, u = gu.split(cfg.ff, dim=-1)
        mlp = ((F.silu(g) if cfg.silu else gelu_tanh(g)) * u) @ W[p + "down_w"].T
        x = x + (rms_norm(mlp, W[p + "post_ffn_norm"], cfg.eps) if cfg.sandwich_norm else mlp)
    return x


def causal_lm_loss(hidden, W: dict, cfg: GemmaCfg, lm_labels: np.ndarray):
    """HF:models/gemma3/modeling_gemma3.py:583,649-660 + HF:loss/loss_utils.py:28-67 —
    final RMSNorm, tied LM head, fp32 shifted cross-entropy, mean over labels != -100.
    Only rows with a valid shifted target are pushed through the head (same value)."""
    tgt = torch.from_numpy(shifted_targets(lm_labels))
    sel = tgt != -100
    y = rms_norm(hidden, W["final_norm"], cfg.eps)[sel]          # [n, h]
    logits = (y @ W.get("lm_head", W["embed"]).T).float()
    return F.cross_entropy(logits, tgt[sel], reduction="mean"), int(sel.sum())


# --------------------------------------------------------------------------- whole step
def stage1_step(pixels, token_ids, labels, Wv, P, Wl, vcfg: SiglipCfg, lcfg: GemmaCfg,
                grad_accum: int = 1, want_intermediates: bool = False, scale_dtype=None):
    """One Stage-1 forward+backward (ref:Stage1/projector_trainer.py:152-237).
    Returns loss (unscaled, as logged) and the four projector gradients of loss/grad_accum."""
    P = {k: v.detach().clone().requires_grad_(True) for k, v in P.items()}
    with torch.no_grad():
        feats = siglip_forward(pixels, Wv, vcfg)
        patch = feats[:, 1:, :]                      # ref:projector_trainer.py:173 ("discard CLS")
        txt = embed_tokens(token_ids, Wl["embed"], lcfg.hidden, scale_dtype, lcfg.scale_embed)
    proj = projector_forward(patch, P)
    X0 = torch.cat([proj, txt.to(proj.dtype)], dim=1)
    n_img = proj.shape[1]
    attn, lm_labels = splice_masks_labels(token_ids.numpy(), labels.numpy(), n_img, lcfg.pad_id)
    hidden = gemma3_hidden(X0, torch.from_numpy(attn), Wl, lcfg)
    loss, n_valid = causal_lm_loss(hidden, Wl, lcfg, lm_labels)
    (loss / grad_accum).backward()
    out = {"loss": float(loss.detach()), "n_valid": n_valid,
           "grads": {k: v.grad.detach() for k, v in P.items()},
           "attention_mask": attn, "lm_labels": lm_labels}
    if want_intermediates:
        out.update(patch=patch.detach(), projected=proj.detach(), spliced=X0.detach(),
                   hidden=hidden.detach())
    return out


# --------------------------------------------------------------------------- inference (prefill + greedy decode)
def generate_positions(attn: torch.Tensor) -> torch.Tensor:
    """HF:generation/utils.py prepare_inputs_for_generation — inside `generate`, position ids come
    from the attention mask: cumsum(mask) - 1, masked positions set to 1."""
    pos = attn.long().cumsum(-1) - 1
    return pos.masked_fill(attn == 0, 1)


def last_position_logits(X0, Wl: dict, lcfg: GemmaCfg, attn=None):
    """HF Gemma3ForCausalLM.forward on inputs_embeds, logits of the last position
    (HF:models/gemma3/modeling_gemma3.py:649-656) — what `generate` samples the next token from.
    `attn` None = all-ones mask (ref:Stage1/inference_stage1.py:170); a padded mask (Stage-2
    evaluation prompt, ref:Stage2/trainer.py:594-596) also moves the positions as generate does."""
    B, S, _ = X0.shape
    if attn is None:
        hidden = gemma3_hidden(X0, torch.ones(B, S, dtype=torch.int64), Wl, lcfg)
    else:
        hidden = gemma3_hidden(X0, attn, Wl, lcfg, pos_ids=generate_positions(attn))
    y = rms_norm(hidden[:, -1, :], Wl["final_norm"], lcfg.eps)
    return (y @ Wl.get("lm_head", Wl["embed"]).T).float()


@torch.no_grad()
def prefill_inputs(pixels, prompt_ids, Wv, P, Wl, vcfg: SiglipCfg, lcfg: GemmaCfg, scale_dtype=None):
    """ref:Stage1/inference_stage1.py:152-168: patch features without token 0 -> projector ->
    (optionally) prompt embeddings appended."""
    patch = siglip_forward(pixels, Wv, vcfg)[:, 1:, :]
    X0 = projector_forward(patch, P)
    if prompt_ids is not None and prompt_ids.shape[1] > 0:
        X0 = torch.cat([X0, embed_tokens(prompt_ids, Wl["embed"], lcfg.hidden, scale_dtype,
                                         lcfg.scale_embed).to(X0.dtype)], dim=1)
    return X0


@torch.no_grad()
def greedy_generate(pixels, prompt_ids, Wv, P, Wl, vcfg: SiglipCfg, lcfg: GemmaCfg, max_new_tokens: int,
                    eos_id=1, pad_id=0, scale_dtype=None, mask_prompt_pads: bool = False):
    """`generate(inputs_embeds=..., do_sample=False)` (ref:Stage1/inference_stage1.py:176-187) restated
    WITHOUT a KV cache: every step re-runs the full forward over [inputs ; tokens so far].  Returns the
    new tokens i64 [B, n] (pad after eos, early stop when all rows finished) and the last-position
    logits of every step."""
    X = prefill_inputs(pixels, prompt_ids, Wv, P, Wl, vcfg, lcfg, scale_dtype)
    B = X.shape[0]
    done = torch.zeros(B, dtype=torch.bool)
    toks, step_logits = [], []
    attn = None
    if mask_prompt_pads:       # ref:Stage2/trainer.py:594-596: ones for the image, (id != pad) for the question
        attn = torch.cat([torch.ones(B, X.shape[1] - prompt_ids.shape[1], dtype=torch.int64),
                          (prompt_ids != pad_id).long()], dim=1)
    for _ in range(max_new_tokens):
        lg = last_position_logits(X, Wl, lcfg, attn)
        step_logits.append(lg)
        nxt = lg.argmax(dim=-1)
        nxt = torch.where(done, torch.full_like(nxt, pad_id), nxt)
        toks.append(nxt)
        done |= nxt == eos_id
        if bool(done.all()):
            break
        X = torch.cat([X, embed_tokens(nxt[:, None], Wl["embed"], lcfg.hidden, scale_dtype,
                                       lcfg.scale_embed).to(X.dtype)], dim=1)
        if attn is not None:
            attn = torch.cat([attn, torch.ones(B, 1, dtype=torch.int64)], dim=1)
    return torch.stack(toks, dim=1), torch.stack(step_logits, dim=1)


# --------------------------------------------------------------------------- optimiser
def cosine_lr(step: int, base_lr: float, warmup: int, total: int) -> float:
    """HF:optimization.py get_cosine_schedule_with_warmup (ref:projector_trainer.py:91-95)."""
    if step < warmup:
        return base_lr * step / max(1, warmup)
    prog = (step - warmup) / max(1, total - warmup)
    return base_lr * max(0.0, 0.5 * (1.0 + math.cos(math.pi * prog)))


def clip_adamw(params: dict, grads: dict, m: dict, v: dict, step: int, lr: float,
               wd: float = 0.01, max_norm: float = 5.0, b1=0.9, b2=0.999, eps=1e-8):
    """ref:projector_trainer.py:241-243 — clip_grad_norm_(5.0) then torch.optim.AdamW.step
    (decoupled weight decay, bias-corrected).  `step` is 1-based.  Updates in place."""
    total = math.sqrt(sum(float((g.double() ** 2).sum()) for g in grads.values()))
    coef = min(1.0, max_norm / (total + 1e-6))
    for k in params:
        g = grads[k] * coef
        params[k].mul_(1.0 - lr * wd)
        m[k].mul_(b1).add_(g, alpha=1 - b1)
        v[k].mul_(b2).addcmul_(g, g, value=1 - b2)
        bc1, bc2 = 1 - b1 ** step, 1 - b2 ** step
        denom = (v[k].sqrt() / math.sqrt(bc2)).add_(eps)
        params[k].addcdiv_(m[k], denom, value=-lr / bc1)
    return total


# --------------------------------------------------------------------------- weights
def init_siglip(cfg: SiglipCfg, seed: int = 0, dtype=torch.float32) -> dict:
    g = torch.Generator().manual_seed(seed)
    d, f = cfg.hidden, cfg.ff
    r = lambda *s, std=0.02: (torch.randn(*s, generator=g) * std).to(dtype)
    W = {"patch_w": r(d, 3 * cfg.patch ** 2, std=(3 * cfg.patch ** 2) ** -0.5), "patch_b": r(d),
         "pos": r(cfg.tokens, d, std=d ** -0.5),
         "post_g": 1 + r(d, std=0.1), "post_b": r(d, std=0.1)}
    for l in range(cfg.layers):
        p = f"l{l}."
        W[p + "ln1_g"], W[p + "ln1_b"] = 1 + r(d, std=0.1), r(d, std=0.1)
        W[p + "ln2_g"], W[p + "ln2_b"] = 1 + r(d, std=0.1), r(d, std=0.1)
        W[p + "qkv_w"], W[p + "qkv_b"] = r(3 * d, d, std=d ** -0.5), r(3 * d)
        W[p + "o_w"], W[p + "o_b"] = r(d, d, std=d ** -0.5), r(d)
        W[p + "fc1_w"], W[p + "fc1_b"] = r(f, d, std=d ** -0.5), r(f)
        W[p + "fc2_w"], W[p + "fc2_b"] = r(d, f, std=f ** -0.5), r(d)
    return W


def init_gemma(cfg: GemmaCfg, seed: int = 0, dtype=torch.float32) -> dict:
    g = torch.Generator().manual_seed(seed + 1)
    h = cfg.hidden
    r = lambda *s, std=0.02: (torch.randn(*s, generator=g) * std).to(dtype)
    W = {"embed": r(cfg.vocab, h, std=0.02), "final_norm": r(h, std=0.3)}
    if not cfg.tie_lm_head:
        W["lm_head"] = r(cfg.vocab, h, std=0.02)
    for l in range(cfg.layers):
        p = f"l{l}."
        for n in (("in_norm", "post_attn_norm", "pre_ffn_norm", "post_ffn_norm") if cfg.sandwich_norm
                  else ("in_norm", "pre_ffn_norm")):
            W[p + n] = r(h, std=0.3)
        W[p + "q_norm"], W[p + "k_norm"] = r(cfg.hd, std=0.3), r(cfg.hd, std=0.3)
        W[p + "qkv_w"] = r((cfg.hq + 2 * cfg.hkv) * cfg.hd, h, std=0.02)
        W[p + "o_w"] = r(h, cfg.hq * cfg.hd, std=0.02)
        W[p + "gu_w"] = r(2 * cfg.ff, h, std=0.02)
        W[p + "down_w"] = r(h, cfg.ff, std=0.02)
    return W


def init_projector(vision_dim: int, llm_dim: int, expansion: int = 10, seed: int = 0) -> dict:
    """torch default nn.Linear init (kaiming-uniform a=sqrt(5) == U(-1/sqrt(in), 1/sqrt(in)))."""
    g = torch.Generator().manual_seed(seed + 2)
    mid = vision_dim * expansion
    u = lambda *s, bound: (torch.rand(*s, generator=g) * 2 - 1) * bound
    return {"model.0.weight": u(mid, vision_dim, bound=vision_dim ** -0.5),
            "model.0.bias": u(mid, bound=vision_dim ** -0.5),
            "model.2.weight": u(llm_dim, mid, bound=mid ** -0.5),
            "model.2.bias": u(llm_dim, bound=mid ** -0.5)}


def synthetic_batch(B: int, T: int, vcfg: SiglipCfg, lcfg: GemmaCfg, seed: int = 1234,
                    left_pad: bool = False):
    """SURVEY.md §8(d) synthetic CXR-shaped batch: pixels ~ U(-1,1), ids[:,0]=BOS(2),
    rest ~ U{3..V-1}; left_pad gives sample i 5*i leading pads (labels -100 there)."""
    g = torch.Generator().manual_seed(seed)
    pixels = torch.rand(B, 3, vcfg.image, vcfg.image, generator=g) * 2 - 1
    ids = torch.randint(3, lcfg.vocab, (B, T), generator=g)
    ids[:, 0] = 2
    if left_pad:
        for i in range(B):
            k = min(5 * i, T - 2)
            if k > 0:
                ids[i, k:] = ids[i, : T - k].clone()
                ids[i, :k] = lcfg.pad_id
    labels = ids.clone()
    labels[ids == lcfg.pad_id] = -100
    return pixels, ids, labels
