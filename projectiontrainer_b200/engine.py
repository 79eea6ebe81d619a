"""Stage-1 step engine: orchestrates the sm_100a kernels (via the C ABI) for

    frozen SigLIP forward -> projector fwd -> splice -> frozen Gemma3 fwd -> masked CE
    -> input-gradient-only backward -> projector wgrads -> clip + AdamW

mirroring ref:Stage1/projector_trainer.py:152-245.  torch supplies device memory and
streams; all arithmetic of the step runs in the sm_100a kernels (what torch does on the step path is plumbing: one
buffer zero-fill, an int32 target copy, KV-cache slot copies) and there is no CPU fallback.

Numerics (bf16 mode = what `accelerate launch --mixed_precision bf16` gives the reference):
bf16 GEMM operands with fp32 accumulation, fp32 residual streams, fp32 norm / softmax /
cross-entropy internals, fp32 projector master weights, gradients and AdamW state.
"""
from __future__ import annotations

import contextlib
import math
import os
from collections import OrderedDict

import torch

from . import _lib, ops
from .config import GemmaCfg, SiglipCfg
from .ops import ACT_GELU_ERF, ACT_GELU_TANH

BF16, F32 = torch.bfloat16, torch.float32

# NVTX ranges per phase of the step (off by default: PT_STAGE1_NVTX=1 turns them on for nsys / ncu --nvtx timelines)
_NVTX = os.environ.get("PT_STAGE1_NVTX", "0") != "0"


class _phase:
    """`with _phase("llm_backward"):` — an NVTX range when PT_STAGE1_NVTX=1, nothing otherwise."""
    __slots__ = ("name",)

    def __init__(self, name):
        self.name = name

    def __enter__(self):
        if _NVTX:
            torch.cuda.nvtx.range_push("stage1/" + self.name)

    def __exit__(self, *exc):
        if _NVTX:
            torch.cuda.nvtx.range_pop()
        return False


def _pad8(n: int) -> int:
    return (n + 7) // 8 * 8


def rope_tables(S: int, hd: int, theta: float, factor: float, device):
    """cos/sin [S, hd/2] in fp32 (HF:models/gemma3/modeling_gemma3.py:216-230).  One-time setup
    (inv_freq on the host as HF does), not part of the per-step path."""
    inv = 1.0 / (theta ** (torch.arange(0, hd, 2, dtype=torch.float32) / hd)) / factor
    ang = torch.arange(S, dtype=torch.float32)[:, None] * inv[None, :]
    return ang.cos().to(device).contiguous(), ang.sin().to(device).contiguous()


class PackedSiglip:
    """Frozen SigLIP weights: bf16 matrices, fp32 vectors."""

    def __init__(self, W: dict, cfg: SiglipCfg, device):
        self.cfg = cfg
        bf = lambda t: t.detach().to(device=device, dtype=BF16).contiguous()
        f32 = lambda t: t.detach().to(device=device, dtype=F32).contiguous()
        self.patch_w, self.patch_b, self.pos = bf(W["patch_w"]), f32(W["patch_b"]), f32(W["pos"])
        self.post_g, self.post_b = f32(W["post_g"]), f32(W["post_b"])
        self.layers = []
        for l in range(cfg.layers):
            p = f"l{l}."
            self.layers.append(dict(
                ln1_g=f32(W[p + "ln1_g"]), ln1_b=f32(W[p + "ln1_b"]),
                ln2_g=f32(W[p + "ln2_g"]), ln2_b=f32(W[p + "ln2_b"]),
                qkv_w=bf(W[p + "qkv_w"]), qkv_b=f32(W[p + "qkv_b"]),
                o_w=bf(W[p + "o_w"]), o_b=f32(W[p + "o_b"]),
                fc1_w=bf(W[p + "fc1_w"]), fc1_b=f32(W[p + "fc1_b"]),
                fc2_w=bf(W[p + "fc2_w"]), fc2_b=f32(W[p + "fc2_b"])))


class PackedGemma:
    """Frozen Gemma3 text-decoder weights: bf16 matrices, fp32 norm gains."""

    def __init__(self, W: dict, cfg: GemmaCfg, device):
        self.cfg = cfg
        bf = lambda t: t.detach().to(device=device, dtype=BF16).contiguous()
        f32 = lambda t: t.detach().to(device=device, dtype=F32).contiguous()
        self.embed = bf(W["embed"])
        self.lm_head = bf(W["lm_head"]) if "lm_head" in W else self.embed     # tied unless the container says otherwise
        self.final_norm = f32(W["final_norm"])
        # HF casts sqrt(hidden) to the embedding dtype before multiplying (modeling_gemma3.py:102-113);
        # Qwen3 embeddings are not scaled
        self.embed_scale = float(torch.tensor(cfg.hidden ** 0.5, dtype=BF16).float()) if cfg.scale_embed else 1.0
        self.layers = []
        post = (lambda k: f32(W[k])) if cfg.sandwich_norm else (lambda k: None)
        for l in range(cfg.layers):
            p = f"l{l}."
            self.layers.append(dict(
                in_norm=f32(W[p + "in_norm"]), post_attn_norm=post(p + "post_attn_norm"),
                pre_ffn_norm=f32(W[p + "pre_ffn_norm"]), post_ffn_norm=post(p + "post_ffn_norm"),
                q_norm=f32(W[p + "q_norm"]), k_norm=f32(W[p + "k_norm"]),
                qkv_w=bf(W[p + "qkv_w"]), o_w=bf(W[p + "o_w"]),
                gu_w=bf(W[p + "gu_w"]), down_w=bf(W[p + "down_w"])))


class ProjectorState:
    """fp32 master weights, gradients and AdamW moments in flat buffers; bf16 shadow for GEMMs.

    Layout of every flat buffer: [W1 (mid x vd) | b1 (mid) | W2 (h x mid) | b2 (h)]
    = state_dict order model.0.weight, model.0.bias, model.2.weight, model.2.bias
    (ref:Stage1/projectors.py:16-20)."""

    KEYS = ("model.0.weight", "model.0.bias", "model.2.weight", "model.2.bias")

    def __init__(self, P: dict, device):
        shapes = [tuple(P[k].shape) for k in self.KEYS]
        sizes = [math.prod(s) for s in shapes]
        # keep every segment 16-byte aligned
        offs, o = [], 0
        for n in sizes:
            offs.append(o)
            o += (n + 7) // 8 * 8
        self.n = o
        self.shapes, self.sizes, self.offs = shapes, sizes, offs
        self.flat = torch.zeros(o, dtype=F32, device=device)
        self.grad = torch.zeros(o, dtype=F32, device=device)
        self.m = torch.zeros(o, dtype=F32, device=device)
        self.v = torch.zeros(o, dtype=F32, device=device)
        self.shadow = torch.zeros(o, dtype=BF16, device=device)
        for k, off, n, s in zip(self.KEYS, offs, sizes, shapes):
            self.flat[off:off + n].view(s).copy_(P[k].detach().to(device=device, dtype=F32))
        self.refresh_shadow()
        self.sumsq = torch.zeros(1, dtype=F32, device=device)
        self.step = 0

    def view(self, buf, i):
        return buf[self.offs[i]:self.offs[i] + self.sizes[i]].view(self.shapes[i])

    @classmethod
    def from_module(cls, module, device):
        """Adopt an `MLPProjector` (ref:Stage1/projectors.py): its four fp32 parameters are re-pointed at
        views of the flat master buffer and their `.grad` at views of the flat gradient buffer, so
        `state_dict()`, `save_projection` and user code keep seeing live values while the fused step
        updates them in place."""
        sd = {k: v.detach() for k, v in module.state_dict().items()}
        if tuple(sd.keys()) != cls.KEYS:
            raise RuntimeError(f"unsupported projector layout {list(sd.keys())}: expected {list(cls.KEYS)} "
                               "(Linear -> GELU -> Linear, ref:Stage1/projectors.py:16-20)")
        st = cls(sd, device)
        params = dict(module.named_parameters())
        for i, k in enumerate(cls.KEYS):
            params[k].data = st.view(st.flat, i)
            params[k].grad = st.view(st.grad, i)
        st.module = module
        return st

    def refresh_shadow(self):
        ops.cast_f32_bf16(self.flat, self.shadow)

    @property
    def w1(self): return self.view(self.shadow, 0)
    @property
    def w2(self): return self.view(self.shadow, 2)
    @property
    def b1(self): return self.view(self.flat, 1)
    @property
    def b2(self): return self.view(self.flat, 3)

    def params(self):
        return {k: self.view(self.flat, i) for i, k in enumerate(self.KEYS)}

    def grads(self):
        return {k: self.view(self.grad, i) for i, k in enumerate(self.KEYS)}


class Stage1Engine:
    """One data-parallel rank of the Stage-1 step on one B200."""

    def __init__(self, vis: PackedSiglip, llm: PackedGemma, proj: ProjectorState, device,
                 grad_accum: int = 1, loss_divisor: float | None = None, lora=None, train_projector: bool = True):
        self.vis, self.llm, self.proj = vis, llm, proj
        # Stage 2 with --enable_qlora: LoRA adapters on the decoder projections (lora.LoraState); the decoder then
        # also runs weight-gradient GEMMs for the adapter matrices.  train_projector=False freezes the projector
        # (ref:Stage2/trainer.py:214-217) — only valid together with adapters.
        self.lora = lora
        self.train_projector = bool(train_projector)
        if not self.train_projector and lora is None:
            raise ValueError("No trainable parameters found. Check freezing configuration and QLoRA setup.")
        self.vcfg, self.lcfg = vis.cfg, llm.cfg
        self.device = device
        # gradients are those of loss / loss_divisor (the reference divides by gradient_accumulation_steps
        # before backward: Stage1/projector_trainer.py:236, Stage2/trainer.py:419)
        self.loss_divisor = float(grad_accum if loss_divisor is None else loss_divisor)
        self._ws: dict = {}
        # one workspace + graph set per batch geometry (Stage 2 pads per batch: ref:Stage2/trainer.py:18-61);
        # least-recently-used geometries are dropped beyond max_geometries
        self._arenas: OrderedDict = OrderedDict()
        self._geometry = ("default",)
        self.max_geometries = int(os.environ.get("PT_STAGE1_MAX_GEOMETRIES", "6"))
        total = torch.cuda.get_device_properties(device).total_memory if torch.cuda.is_available() else 0
        self.arena_budget_bytes = int(float(os.environ.get("PT_STAGE1_ARENA_GB", "0")) * 2 ** 30) or int(0.6 * total)
        # device-side input validation (token ids / targets out of range): kernels OR bits into this word, the host
        # reads it where it already synchronises on the loss (check_errors)
        self._err = torch.zeros(1, dtype=torch.int32, device=device) if torch.cuda.is_available() else None
        self._acc_grad = None           # gradient accumulated over micro-batches that did not step
        self._acc_pending = False
        self._rope: dict = {}
        self.n_img = self.vcfg.tokens - 1      # ref:projector_trainer.py:173 drops token 0
        self.saved = None
        # CUDA-graph replay of forward+backward (launch-bound inner loop: ~900 launches per step)
        self.use_graph = os.environ.get("PT_STAGE1_GRAPH", "1") != "0"
        # graphs of the current geometry, least recently used first; at most max_graphs stay captured per geometry
        # (a graph pins ~900 kernel nodes; Stage 2 keys it on the bucketed label span as well)
        self._graphs: OrderedDict = OrderedDict()
        self.max_graphs = int(os.environ.get("PT_STAGE1_MAX_GRAPHS", "8"))
        # GeGLU forward / backward fused into the gate||up and down-dgrad GEMM epilogues (needs ff % 128 == 0)
        self.fuse_attn_fwd = ops.attn_gemma_fusable(self.lcfg.hq, self.lcfg.hkv, self.lcfg.hd) and \
            os.environ.get("PT_STAGE1_FUSE_ATTN_FWD", "1") != "0"
        self.bn_hint_dkv = int(os.environ.get("PT_STAGE1_BN_DKV", "0"))     # tile-width hint of the dV / dK GEMMs
        self.fuse_softmax_bwd = os.environ.get("PT_STAGE1_FUSE_SOFTMAX_BWD", "1") != "0"
        # dV / dK / dQ of the attention backward on three streams (fork / join inside the captured graph).  Off by
        # default: measured neutral on B200 (101.3 / 101.9 ms per step without, 100.9 / 102.3 ms with) — the step
        # runs under sw_power_cap, and filling the last-wave bubbles lowered the SM clock (1459 -> 1421 MHz) by as
        # much as it saved.
        self.attn_bwd_streams = os.environ.get("PT_STAGE1_ATTN_BWD_STREAMS", "0") != "0"
        self._side = None
        # LM-head input gradient as a split-K launch (K = vocab) + partial sum; 1 = the single 128-wide-tile launch
        self.lm_head_splitk = int(os.environ.get("PT_STAGE1_LM_HEAD_SPLITK", "8"))
        self.fuse_geglu = ops.geglu_fusable(self.lcfg.ff) and os.environ.get("PT_STAGE1_FUSE_GEGLU", "1") != "0"
        # delta = rowsum(dO * O) in the o-projection dgrad epilogue; erf-GELU backward in the projector dgrad epilogue
        self.fuse_attn_delta = ops.attn_delta_fusable(self.lcfg.hq, self.lcfg.hd) and \
            os.environ.get("PT_STAGE1_FUSE_DELTA", "1") != "0"
        self.fuse_gelu_bwd = os.environ.get("PT_STAGE1_FUSE_GELU_BWD", "1") != "0"
        # cross-entropy statistics (row max / sum exp partials) in the LM-head GEMM epilogue: the separate statistics
        # pass over the 2.1 GB of logits goes away
        self.fuse_ce_stats = self.lcfg.vocab >= 256 and os.environ.get("PT_STAGE1_FUSE_CE_STATS", "1") != "0"
        # decode step: programmatic dependent launch (bit set of pt_set_dependent_launch: 1 GEMMs, 2 norm kernels,
        # 4 attention).  A chain of dependent M = 64 GEMMs gains 1.5 us per launch (9.37 -> 7.86 us), but inside the
        # step the A/B is box dependent (-6 % on one box, 0 .. +3 % on another, profiles/r02_decode_attn.txt): off.
        self.decode_pdl = int(os.environ.get("PT_STAGE1_DECODE_PDL", "0"))
        # graphed decode step: q/k norm + RoPE, cache append and attention over the cache in one launch per layer
        # (0: pt_qknorm_rope_fwd, two cache row copies, score GEMM, softmax, P V GEMM)
        self.fuse_decode_attn = ops.decode_attn_supported(self.lcfg.hq, self.lcfg.hkv, self.lcfg.hd) and \
            os.environ.get("PT_STAGE1_FUSE_DECODE_ATTN", "1") != "0"
        # decode step at batch <= 16: the linear layers as weight-streaming mma.sync kernels (pt_gemv_small_m) instead
        # of the 256-row tcgen05 tile
        self.decode_gemv = os.environ.get("PT_STAGE1_DECODE_GEMV", "1") != "0"
        self.decode_gemv_rows = int(os.environ.get("PT_STAGE1_DECODE_GEMV_ROWS", "16"))
        # q/k RMSNorm + RoPE in the epilogue of the q|k|v projection (PT_ACT_QKNORM_ROPE, head_dim 256).  Built, pinned
        # (tests/test_gemm_gpu.py) and measured: the epilogue of this short-K GEMM is already its bottleneck, and a second
        # pass over TMEM for the sum of squares, the per-row rotary table reads and two more output chunks per pair make
        # it 164 us against 127 us for the projection + pt_qknorm_rope_fwd (step +0.3 .. +2 ms): off
        self.fuse_qknorm = os.environ.get("PT_STAGE1_FUSE_QKNORM", "0") != "0"
        # generate(): the next-token choice (HF processors + draw + bookkeeping) as one launch per token
        # (0: the same logic in torch ops on the [B, V] logits, also what top_k = 0 / > 1024 takes)
        self.fuse_sampler = os.environ.get("PT_STAGE1_FUSE_SAMPLER", "1") != "0"
        # graphed decode step, sandwich-norm decoders: o- and down-projection as split-K GEMMs into an f32 accumulator
        # that the following norm kernel rounds, consumes and clears (0: one CTA pair per 64 output columns streams
        # the whole contraction)
        self.decode_splitk = self.lcfg.sandwich_norm and self.lcfg.hidden <= 4096 and \
            os.environ.get("PT_STAGE1_DECODE_SPLITK", "1") != "0"
        # attention backward on a CTA pair per 128 keys (0: the single-CTA kernel on 64-key tiles)
        self.attn_bwd_pair = ops.attn_gemma_bwd_pair_ok(self.lcfg.hq, self.lcfg.hkv, self.lcfg.hd) and \
            os.environ.get("PT_STAGE1_ATTN_BWD_PAIR", "1") != "0"

    # ------------------------------------------------------------------ workspace
    def buf(self, name, shape, dtype, zero=False):
        key = (name, tuple(shape), dtype)
        t = self._ws.get(key)
        if t is None:
            t = (torch.zeros if zero else torch.empty)(shape, dtype=dtype, device=self.device)
            self._ws[key] = t
        return t

    def select_geometry(self, key):
        """Switch the workspace (and the CUDA graphs captured over it) to the arena of batch geometry
        `key`.  Buffers are keyed by shape, so without this every new padded length would add a full
        activation set; here at most `max_geometries` sets stay alive, oldest first out."""
        if key == self._geometry:
            return
        self._arenas[self._geometry] = (self._ws, self._graphs)
        if key in self._arenas:
            self._ws, self._graphs = self._arenas.pop(key)
        else:
            # a new arena will be about as large as the one being left: keep the total under the byte budget too
            # (long Stage-2 sequences at large batch sizes: six arenas of 40 GB do not fit)
            nbytes = lambda ws: sum(t.numel() * t.element_size() for t in ws.values())
            sizes = [nbytes(ws) for ws, _ in self._arenas.values()]
            estimate = sizes[-1] if sizes else 0
            self._ws, self._graphs = {}, OrderedDict()
            while self._arenas and (len(self._arenas) >= max(1, self.max_geometries) or
                                    sum(sizes) + estimate > self.arena_budget_bytes):
                _, (ws, graphs) = self._arenas.popitem(last=False)
                sizes.pop(0)
                graphs.clear()
                ws.clear()
        self._geometry = key
        self.saved = None

    def arena_bytes(self) -> int:
        """Bytes held by the workspaces of every live geometry (current one included)."""
        nbytes = lambda ws: sum(t.numel() * t.element_size() for t in ws.values())
        return nbytes(self._ws) + sum(nbytes(ws) for ws, _ in self._arenas.values())

    def _enforce_budget(self) -> None:
        """After the current geometry grew (first call at a new label span sizes new CE-side buffers): drop the
        least recently used OTHER geometries until the total is back under the byte budget."""
        while self._arenas and self.arena_bytes() > self.arena_budget_bytes:
            _, (ws, graphs) = self._arenas.popitem(last=False)
            graphs.clear()
            ws.clear()

    def rope(self, S, full: bool):
        key = (S, full)
        if key not in self._rope:
            c = self.lcfg
            self._rope[key] = rope_tables(S, c.hd, c.theta_global if full else c.theta_local,
                                          c.rope_factor_global if full else 1.0, self.device)
        return self._rope[key]

    # ------------------------------------------------------------------ SigLIP forward
    def vision_forward(self, pixels: torch.Tensor, drop_first: bool = True) -> torch.Tensor:
        """pixels f32 [B,3,H,W] -> bf16 features [B*(N-1), d] (token 0 dropped, the reference's
        `last_hidden_state[:, 1:, :]`) or [B*N, d] with drop_first=False."""
        c, V = self.vcfg, self.vis
        B = pixels.shape[0]
        N, d, H, hd = c.tokens, c.hidden, c.heads, c.hd
        M = B * N
        cols = self.buf("v_cols", (M, 3 * c.patch ** 2), BF16)
        ops.patchify(pixels, c.patch, cols)
        x = self.buf("v_x", (M, d), F32)
        # patch-embed GEMM, epilogue adds bias and the position table (broadcast over images)
        ops.gemm_raw(M=N, N=d, K=cols.shape[1], A=cols, lda=cols.shape[1], B=V.patch_w,
                     ldb=cols.shape[1], Cout=x, ldc=d, nz=(B, 1, 1), a_zs=(N * cols.shape[1], 0, 0),
                     c_zs=(N * d, 0, 0), bias=V.patch_b, residual=V.pos, ldr=d, r_zs=(0, 0, 0))
        a = self.buf("v_a", (M, d), BF16)
        qkv = self.buf("v_qkv", (M, 3 * d), BF16)
        ld = _pad8(N)
        fused = hd == 64                    # tcgen05 fused attention kernel; other head dims: GEMM + softmax + GEMM
        if not fused:
            sc = self.buf("v_scores", (B * H, N, ld), F32)
            P = self.buf("v_P", (B * H, N, ld), BF16)
        o = self.buf("v_o", (M, d), BF16)
        h1 = self.buf("v_h1", (M, c.ff), BF16)
        for L in V.layers:
            ops.layernorm_fwd(x, L["ln1_g"], L["ln1_b"], a, c.eps)
            ops.linear(a, L["qkv_w"], bias=L["qkv_b"], out=qkv)
            if fused:
                ops.attn_siglip_fwd(qkv, o, B, N, H, hd, hd ** -0.5)
            else:
                # scores[b,h] = q k^T / sqrt(hd)
                ops.gemm_raw(M=N, N=N, K=hd, A=qkv, lda=3 * d, B=qkv, b_off=d, ldb=3 * d, Cout=sc,
                             ldc=ld, nz=(H, B, 1), a_zs=(hd, N * 3 * d, 0), b_zs=(hd, N * 3 * d, 0),
                             c_zs=(N * ld, H * N * ld, 0), alpha=hd ** -0.5)
                ops.softmax_fwd(sc, P, None, B * H, N, N, ld, H, 0, 0)
                # o[b,:,h,:] = P[b,h] v[b,:,h,:]   (v read in place, MN-major)
                ops.gemm_raw(M=N, N=hd, K=N, A=P, lda=ld, B=qkv, b_off=2 * d, ldb=3 * d, b_mn=True,
                             Cout=o, ldc=d, nz=(H, B, 1), a_zs=(N * ld, H * N * ld, 0),
                             b_zs=(hd, N * 3 * d, 0), c_zs=(hd, N * d, 0))
            ops.linear(o, L["o_w"], bias=L["o_b"], residual=x, out=x)
            ops.layernorm_fwd(x, L["ln2_g"], L["ln2_b"], a, c.eps)
            ops.linear(a, L["fc1_w"], bias=L["fc1_b"], act=ACT_GELU_TANH, out=h1)
            ops.linear(h1, L["fc2_w"], bias=L["fc2_b"], residual=x, out=x)
        if drop_first:
            feat = self.buf("v_feat", (B * (N - 1), d), BF16)
            ops.layernorm_fwd(x, V.post_g, V.post_b, feat, c.eps, group=(N - 1, N, 1))
        else:
            feat = self.buf("v_feat_all", (M, d), BF16)
            ops.layernorm_fwd(x, V.post_g, V.post_b, feat, c.eps)
        return feat

    # ------------------------------------------------------------------ projector forward
    def projector_forward(self, patch: torch.Tensor, out: torch.Tensor, *, rows_per_img: int,
                          out_batch_stride: int, save: bool = True):
        """out[b, 0:rows_per_img, :] (f32, row stride h, image stride out_batch_stride) =
        W2 gelu_erf(W1 patch + b1) + b2   (ref:Stage1/projectors.py:16-20)."""
        Pj = self.proj
        R, vd = patch.shape
        mid, h = Pj.shapes[0][0], Pj.shapes[2][0]
        Z = self.buf("p_Z", (R, mid), BF16) if save else None
        Hm = self.buf("p_H", (R, mid), BF16)
        ops.linear(patch, Pj.w1, bias=Pj.b1, act=ACT_GELU_ERF, out=Hm, c_pre=Z)
        Bn = R // rows_per_img
        ops.gemm_raw(M=rows_per_img, N=h, K=mid, A=Hm, lda=mid, B=Pj.w2, ldb=mid, Cout=out, ldc=h,
                     nz=(Bn, 1, 1), a_zs=(rows_per_img * mid, 0, 0), c_zs=(out_batch_stride, 0, 0),
                     bias=Pj.b2)
        return Z, Hm

    # ------------------------------------------------------------------ Gemma3 attention helpers
    def _attn_geometry(self, B, S):
        c = self.lcfg
        G = c.hq // c.hkv
        ld = _pad8(S)
        QKV = (c.hq + 2 * c.hkv) * c.hd
        return G, ld, QKV

    def _scores_like(self, A, a_off, lda, a_zs, Bm, b_off, ldb, b_zs, out, B, S, alpha, b_mn=False, window=0,
                     **epilogue):
        """out[b, kvh*G+g, q, k] (f32, ld) = alpha * A_row(q) . B_row(k) over hd, batched (g, kvh, b)."""
        c = self.lcfg
        G, ld, _ = self._attn_geometry(B, S)
        ops.gemm_raw(M=S, N=S, K=c.hd, A=A, a_off=a_off, lda=lda, B=Bm, b_off=b_off, ldb=ldb,
                     Cout=out, ldc=ld, nz=(G, c.hkv, B), a_zs=a_zs, b_zs=b_zs,
                     c_zs=(S * ld, G * S * ld, c.hq * S * ld), alpha=alpha, band=ops.BAND_SCORES,
                     band_window=window, **epilogue)

    def _thin(self, fn, x, w, out, alpha):
        """A rank-wide adapter product out[M, R] = alpha * x . w (fn = ops.linear or ops.linear_dgrad).  At small batch
        sizes it has a handful of 256-row tiles for 74 SM pairs and a long contraction: then it runs split-K into an
        f32 scratch matrix (cleared first) that is cast to bf16 afterwards."""
        M, K = x.shape
        tiles = -(-M // 256) * -(-out.shape[1] // 64)
        waves = -(-tiles // 74)
        # few tiles, or a badly filled last wave in front of a long contraction (88 tiles = 1.2 waves at B*S = 22 496)
        if not (K >= 1024 and (tiles * 2 <= 74 or (K >= 4096 and tiles < 0.7 * waves * 74))):
            fn(x, w, out=out, alpha=alpha)
            return
        tmp = self.buf("lora_thin_f32", tuple(out.shape), F32)
        tmp.zero_()
        fn(x, w, out=tmp, alpha=alpha, ksplit=-1)
        ops.cast_f32_bf16(tmp, out)

    # ------------------------------------------------------------------ LLM forward + loss
    def _llm_stack(self, x0: torch.Tensor, keyvalid, B: int, S: int, save: bool, cache=None, pos_ids=None):
        """Decoder stack over x0 f32 [B*S, h] (HF:models/gemma3/modeling_gemma3.py:404-434 per layer).
        Returns the final residual stream (before the final norm) and, with save=True, the per-layer
        activations the input-gradient backward needs.  With `cache` (prefill) the rotated keys and
        the values of every layer are also stored into the KV cache at positions 0..S-1.  Positions
        are arange(S) (F3) unless `pos_ids` i64 [B, S] gives them per sample (generate under a padded
        prompt mask): the RoPE rows are then gathered once per table and looked up per (sample, position)."""
        c, Lm = self.lcfg, self.llm
        h, hd, hq, hkv, ff = c.hidden, c.hd, c.hq, c.hkv, c.ff
        M = B * S
        G, ld, QKV = self._attn_geometry(B, S)
        scale = c.q_scalar ** -0.5
        sv = []
        lo = self.lora
        keep = save and lo is not None          # adapter weight gradients need every projection's input
        abuf = lambda li, which: self.buf(f"l{li}_a{which}" if keep else "l_a", (M, h), BF16)
        a = abuf(0, "in")
        sc = None if self.fuse_attn_fwd else self.buf("l_scores", (B * hq, S, ld), F32)
        kvbits = None
        if self.fuse_attn_fwd and keyvalid is not None:
            kvbits = self.buf("l_kvbits", (B * ((S + 63) // 64),), torch.int64)
            ops.pack_keyvalid(keyvalid.view(B, S), kvbits)
        x = x0
        ops.rmsnorm_fwd(x, Lm.layers[0]["in_norm"], a, c.eps)
        rope_rows, tabs = S, {f: self.rope(S, f) for f in (True, False)}
        if pos_ids is not None:
            flat = pos_ids.reshape(-1)
            tabs = {f: (t[0].index_select(0, flat), t[1].index_select(0, flat)) for f, t in tabs.items()}
            rope_rows = M
        cs_tabs = {}
        if self.fuse_qknorm and hd == 256:      # [cos | sin] per position, as the q|k|v projection's epilogue reads it
            for f, t in tabs.items():
                key = ("cs", S, f)
                if pos_ids is not None or key not in self._rope:
                    cs = torch.cat([t[0], t[1]], dim=1).contiguous()
                    if pos_ids is not None:
                        cs_tabs[f] = cs
                        continue
                    self._rope[key] = cs
                cs_tabs[f] = self._rope[key]
        for li, L in enumerate(Lm.layers):
            tag = f"l{li}_" if save else "l_"
            full = c.is_full(li)
            cos_t, sin_t = tabs[full]
            qkv = self.buf(tag + "qkv", (M, QKV), BF16)
            q_post = self.buf(tag + "q", (M, hq * hd), BF16)
            k_post = self.buf(tag + "k", (M, hkv * hd), BF16)
            # fused attention keeps only the log-sum-exp of every score row (the backward recomputes P);
            # the GEMM + softmax + GEMM fall-back materialises the probabilities
            P = None if self.fuse_attn_fwd else self.buf(tag + "P", (B * hq, S, ld), BF16)
            lse = self.buf(tag + "lse", (B * hq, S), F32) if self.fuse_attn_fwd and save else None
            attn_o = self.buf(tag + "ao", (M, hq * hd), BF16)      # kept: delta = rowsum(dO * O) in the backward
            y_attn = self.buf(tag + "ya", (M, h), BF16) if c.sandwich_norm else None
            x1 = self.buf(tag + "x1", (M, h), F32)
            gu = self.buf(tag + "gu", (M, 2 * ff), BF16)
            y_mlp = self.buf(tag + "ym", (M, h), BF16) if c.sandwich_norm else None
            x2 = self.buf(tag + ("x2" if save else f"x2_{li & 1}"), (M, h), F32)
            act = self.buf(f"l{li}_act" if keep else "l_act", (M, ff), BF16)
            a_in, ts = a, {}

            def adapter(g, x_in, _li=li, _tag=tag, _ts=ts):
                """t = (alpha / r) * x A_cat^T of packed projection g -> the (x2, w2) segment of its base GEMM."""
                grp = None if lo is None else lo.layers[_li].get(g)
                if grp is None:
                    return {}
                t = self.buf(_tag + "t_" + g, (M, grp.R), BF16)
                self._thin(ops.linear, x_in, grp.A(lo.shadow), t, lo.scale)
                _ts[g] = t
                return dict(x2=t, w2=grp.B(lo.shadow))

            if self.fuse_qknorm and hd == 256:
                if "qk_gains" not in L:
                    L["qk_gains"] = torch.cat([L["q_norm"], L["k_norm"]]).contiguous()
                ops.linear_qknorm_rope(a_in, L["qkv_w"], qkv, q_post, k_post, L["qk_gains"], cs_tabs[full], hq, c.eps,
                                       **adapter("qkv", a_in))
            else:
                ops.linear(a_in, L["qkv_w"], out=qkv, **adapter("qkv", a_in))
                ops.qknorm_rope_fwd(qkv, L["q_norm"], L["k_norm"], cos_t, sin_t, q_post, k_post, rope_rows, hq,
                                    hkv, hd, c.eps)
            if cache is not None:      # KV-cache fill: device-to-device copies (plumbing, no arithmetic)
                cache["k"][li][:, :S].copy_(k_post.view(B, S, hkv * hd))
                cache["v"][li][:, :S].copy_(qkv.view(B, S, QKV)[:, :, (hq + hkv) * hd:])
            win = 0 if full else c.window
            if self.fuse_attn_fwd:
                ops.attn_gemma_fwd(q_post, k_post, qkv, kvbits, lse, attn_o, B, S, hq, hkv, hd, win, scale)
            else:
                self._scores_like(q_post, 0, hq * hd, (hd, G * hd, S * hq * hd), k_post, 0, hkv * hd,
                                  (0, hd, S * hkv * hd), sc, B, S, scale, window=win)
                ops.softmax_fwd(sc, P, keyvalid, B * hq, S, S, ld, hq, 1, win)
                # attn_o[b, q, head, :] = P[b, head] V[b, :, kvh, :]   (V in place inside qkv, MN-major)
                ops.gemm_raw(M=S, N=hd, K=S, A=P, lda=ld, B=qkv, b_off=(hq + hkv) * hd, ldb=QKV, b_mn=True,
                             Cout=attn_o, ldc=hq * hd, nz=(G, hkv, B),
                             a_zs=(S * ld, G * S * ld, hq * S * ld), b_zs=(0, hd, S * QKV),
                             c_zs=(hd, G * hd, S * hq * hd), band=ops.BAND_K_BY_QUERY, band_window=win)
            last = li == len(Lm.layers) - 1
            a_ffn = abuf(li, "ffn")
            if c.sandwich_norm:
                ops.linear(attn_o, L["o_w"], out=y_attn, **adapter("o", attn_o))
                ops.rmsnorm_residual_fwd(y_attn, L["post_attn_norm"], x, x1, L["pre_ffn_norm"], a_ffn, c.eps)
            else:       # pre-norm decoder (Qwen3): the residual add rides the o-projection epilogue, in f32
                ops.linear(attn_o, L["o_w"], out=x1, residual=x, **adapter("o", attn_o))
                ops.rmsnorm_fwd(x1, L["pre_ffn_norm"], a_ffn, c.eps)
            if self.fuse_geglu:
                ops.linear_geglu(a_ffn, L["gu_w"], gu, act, silu=c.silu, **adapter("gu", a_ffn))
            else:
                ops.linear(a_ffn, L["gu_w"], out=gu, **adapter("gu", a_ffn))
                ops.geglu_fwd(gu, act, silu=c.silu)
            a = a_ffn if last else abuf(li + 1, "in")
            if c.sandwich_norm:
                ops.linear(act, L["down_w"], out=y_mlp, **adapter("down", act))
                ops.rmsnorm_residual_fwd(y_mlp, L["post_ffn_norm"], x1, x2,
                                         None if last else Lm.layers[li + 1]["in_norm"],
                                         None if last else a, c.eps)
            else:
                ops.linear(act, L["down_w"], out=x2, residual=x1, **adapter("down", act))
                if not last:
                    ops.rmsnorm_fwd(x2, Lm.layers[li + 1]["in_norm"], a, c.eps)
            if save:
                sv.append(dict(x_in=x, qkv=qkv, q=q_post, k=k_post, P=P, lse=lse, kvbits=kvbits, ao=attn_o, ya=y_attn,
                               x1=x1, gu=gu, ym=y_mlp, full=full, a_in=a_in, a_ffn=a_ffn, act=act, t=ts))
            x = x2
        return x, sv

    def llm_forward_loss(self, x0: torch.Tensor, keyvalid: torch.Tensor, targets: torch.Tensor,
                         B: int, S: int, T: int, save: bool = True, lab0: int | None = None) -> torch.Tensor:
        """x0 f32 [B*S, h] spliced embeddings -> device scalar loss; leaves dlogits in place and the
        activations the input-gradient backward needs in self.saved.  Only the T positions
        lab0 .. lab0+T-1 of every sample go through the final norm and the LM head: the ones whose
        shifted target can be a label (Stage 1: the caption, lab0 = S-T-1; Stage 2: the answer)."""
        lab0 = S - T - 1 if lab0 is None else lab0
        c, Lm = self.lcfg, self.llm
        h = c.hidden
        x, sv = self._llm_stack(x0, keyvalid, B, S, save)
        # final norm only on rows that can carry a label: positions lab0 .. lab0+T-1 of every sample
        R = B * T
        yf = self.buf("l_yf", (R, h), BF16)
        ops.rmsnorm_fwd(x, Lm.final_norm, yf, c.eps, group=(T, S, lab0))
        logits = self.buf("l_logits", (R, _pad8(c.vocab)), BF16)[:, :c.vocab]    # row stride padded to 16 B
        stats = None
        if self.fuse_ce_stats:
            stats = self.buf("l_ce_parts", (ops.ce_stats_parts(c.vocab), R, 2), F32)
            ops.linear_ce_stats(yf, Lm.lm_head, logits, stats)
        else:
            ops.linear(yf, Lm.lm_head, out=logits)
        tgt = self.buf("l_tgt", (R,), torch.int32)
        # targets[b, lab0+t] for t in 0..T-1 -> compact [B*T] (int copy: plumbing, not arithmetic)
        tgt.view(B, T).copy_(targets.view(B, S)[:, lab0:lab0 + T])
        lse = self.buf("l_lse", (R,), F32)
        acc2 = self.buf("l_acc2", (2,), F32)
        loss = self.buf("l_loss", (1,), F32)
        if stats is not None:
            ops.ce_from_partials(logits, tgt, stats, lse, acc2, loss, 1.0 / self.loss_divisor, err=self._err)
        else:
            ops.ce_fwd_bwd(logits, tgt, lse, acc2, loss, 1.0 / self.loss_divisor, err=self._err)
        self.saved = dict(layers=sv, x_final=x, logits=logits, B=B, S=S, T=T, lab0=lab0) if save else None
        return loss

    # ------------------------------------------------------------------ LLM backward (input grads only)
    def _attn_dv(self, a, d_o, d_qkv, B, S, G, ld, QKV, win):
        """dV[b,k,kvh,:] = sum_g P[b,kvh*G+g]^T dO[b,:,kvh*G+g,:]  -> v columns of d_qkv"""
        c = self.lcfg
        hd, hq, hkv = c.hd, c.hq, c.hkv
        ops.gemm_raw(M=S, N=hd, K=S, A=a["P"], lda=ld, a_mn=True, B=d_o, ldb=hq * hd, b_mn=True,
                     Cout=d_qkv, c_off=(hq + hkv) * hd, ldc=QKV, nz=(hkv, B, 1),
                     a_zs=(G * S * ld, hq * S * ld, 0), b_zs=(G * hd, S * hq * hd, 0),
                     c_zs=(hd, S * QKV, 0), kg=G, a_kgs=S * ld, b_kgs=hd, band=ops.BAND_K_BY_KEY,
                     band_window=win, bn_hint=self.bn_hint_dkv)

    def llm_backward(self) -> torch.Tensor:
        """Consumes self.saved; returns dL/dx0 as f32 [B*S, h] (weights are frozen: no wgrad GEMMs)."""
        c, Lm, s = self.lcfg, self.llm, self.saved
        B, S, T = s["B"], s["S"], s["T"]
        h, hd, hq, hkv, ff = c.hidden, c.hd, c.hq, c.hkv, c.ff
        M, R = B * S, B * T
        G, ld, QKV = self._attn_geometry(B, S)
        scale = c.q_scalar ** -0.5
        d_yf = self.buf("b_dyf", (R, h), BF16)
        if self.lm_head_splitk > 1 and c.vocab % (self.lm_head_splitk * 64) == 0 and (R * h) % 4 == 0:
            part = self.buf("b_dyf_part", (self.lm_head_splitk, R, h), F32)
            ops.linear_dgrad_splitk(s["logits"], Lm.lm_head, out=d_yf, partials=part, splits=self.lm_head_splitk)
        else:
            ops.linear_dgrad(s["logits"], Lm.lm_head, out=d_yf)        # dlogits @ E
        dx = self.buf("b_dx", (M, h), F32)
        dx.zero_()
        grp = (T, S, s["lab0"])
        ops.rmsnorm_bwd(d_yf, s["x_final"], Lm.final_norm, dx, c.eps, rows=R, u_group=grp, dst_group=grp)
        d_y = self.buf("b_dy", (M, h), BF16)
        d_act = self.buf("b_dact", (M, ff), BF16)
        d_gu = self.buf("b_dgu", (M, 2 * ff), BF16)
        d_o = self.buf("b_do", (M, hq * hd), BF16)
        fused_attn = self.fuse_attn_fwd           # flash-style backward: P recomputed from the saved log-sum-exp
        dP = None if self.fuse_softmax_bwd or fused_attn else self.buf("l_scores", (B * hq, S, ld), F32)
        if fused_attn:
            # the fused backward writes dS only inside the (query block, key tile) pairs of a layer's band and the
            # dQ GEMM reads whole K blocks: one zero-filled buffer per band shape (full-causal / sliding layers)
            tag = "b_dST_" if self.attn_bwd_pair else "b_dS_"
            dS_by_band = {f: self.buf(tag + ("full" if f else "slide"), (B * hq, S, ld), BF16, zero=True)
                          for f in {a["full"] for a in s["layers"]}}
        else:
            dS = self.buf("b_dS", (B * hq, S, ld), BF16)
        delta = self.buf("b_delta", (B * hq * S,), F32)
        dq = self.buf("b_dq", (M, hq * hd), BF16)
        dk = self.buf("b_dk", (M, hkv * hd), BF16)
        d_qkv = self.buf("b_dqkv", (M, QKV), BF16)
        sandwich = c.sandwich_norm
        pair = sandwich and h <= 2560
        side = self.attn_bwd_streams and ops.GEMM_PROFILE is None     # per-launch timing pass: one stream
        if side:
            if self._side is None:
                self._side = (torch.cuda.Stream(self.device), torch.cuda.Stream(self.device),
                              *(torch.cuda.Event() for _ in range(4)))
            sv, sk, ev_do, ev_dv, ev_ds, ev_dk = self._side
            main = torch.cuda.current_stream()
        lo = self.lora
        if lo is not None:
            lo.grad.zero_()         # the adapter weight gradients are accumulated into by split-K launches

        def lora_pre(li, g, dy):
            """dt = (alpha / r) * dy B_blk of packed projection g -> the (dy2, w2) segment of its input-gradient GEMM."""
            grp = None if lo is None else lo.layers[li].get(g)
            if grp is None:
                return {}, None
            dt = self.buf("b_dt_" + g, (M, grp.R), BF16)
            self._thin(ops.linear_dgrad, dy, grp.B(lo.shadow), dt, lo.scale)
            return dict(dy2=dt, w2=grp.A(lo.shadow)), (grp, dt)

        def lora_wgrad(st, dy, x_in, t):
            """dA_cat = dt^T x (rows past the adapters' ranks come out zero: those dt columns are);
            dB_i = dy_i^T t_i into the i-th diagonal block."""
            if st is None:
                return
            # thin outputs contracted over all B*S rows: split-K launches that add into the (cleared) gradient buffer
            grp, dt = st
            ops.linear_wgrad(dt, x_in, out=grp.A(lo.grad), accumulate=True)
            gB, r = grp.B(lo.grad), lo.r
            for _, r0, n, i in grp.parts:
                ops.linear_wgrad(dy[:, r0:r0 + n], t[:, i * r:(i + 1) * r], out=gB[r0:r0 + n, i * r:(i + 1) * r],
                                 accumulate=True)

        for li in range(len(Lm.layers) - 1, -1, -1):
            L, a = Lm.layers[li], s["layers"][li]
            cos_t, sin_t = self.rope(S, a["full"])
            # ---- MLP branch
            if not sandwich:        # pre-norm decoder: the branch output's gradient is the stream's, as bf16
                ops.cast_f32_bf16(dx, d_y)
            elif li == len(Lm.layers) - 1 or not pair:
                ops.rmsnorm_bwd(dx, a["ym"], L["post_ffn_norm"], d_y, c.eps)
            pre, st = lora_pre(li, "down", d_y)
            if self.fuse_geglu:
                ops.linear_dgrad_geglu_bwd(d_y, L["down_w"], a["gu"], d_gu, silu=c.silu, **pre)
            else:
                ops.linear_dgrad(d_y, L["down_w"], out=d_act, **pre)
                ops.geglu_bwd(d_act, a["gu"], d_gu, silu=c.silu)
            lora_wgrad(st, d_y, a["act"], a["t"].get("down"))
            pre, st = lora_pre(li, "gu", d_gu)
            lora_wgrad(st, d_gu, a["a_ffn"], a["t"].get("gu"))
            ops.linear_dgrad(d_gu, L["gu_w"], out=d_y, **pre)
            # ---- attention branch (pre-ffn norm backward + post-attention norm backward in one pass)
            if pair:
                ops.rmsnorm_bwd_pair(d_y, a["x1"], L["pre_ffn_norm"], dx, a["ya"], L["post_attn_norm"], d_y, c.eps)
            else:
                ops.rmsnorm_bwd(d_y, a["x1"], L["pre_ffn_norm"], dx, c.eps, add=dx)
                if sandwich:
                    ops.rmsnorm_bwd(dx, a["ya"], L["post_attn_norm"], d_y, c.eps)
                else:
                    ops.cast_f32_bf16(dx, d_y)
            win = 0 if a["full"] else c.window
            pre, st = lora_pre(li, "o", d_y)
            if fused_attn and self.fuse_attn_delta:
                # o-projection input gradient with delta = rowsum(dO * O) formed in its epilogue
                ops.linear_dgrad_attn_delta(d_y, L["o_w"], a["ao"], d_o, delta, S, hd, **pre)
            else:
                ops.linear_dgrad(d_y, L["o_w"], out=d_o, **pre)
            lora_wgrad(st, d_y, a["ao"], a["t"].get("o"))
            if fused_attn:
                # dK, dV and dS in one key/value-stationary kernel; dQ = dS K as a band GEMM
                dS = dS_by_band[a["full"]]
                if not self.fuse_attn_delta:
                    ops.attn_delta(d_o, a["ao"], delta, S, hq, hd)
                if self.attn_bwd_pair:
                    # a CTA pair per 128 keys (K tile -> P^T, dV on one SM; V tile -> dS^T, dK on its partner);
                    # dS leaves transposed and feeds the dQ GEMM as an MN-major operand
                    ops.attn_gemma_bwd_pair(a["q"], a["k"], a["qkv"], d_o, a["kvbits"], a["lse"], delta, dS, dk,
                                            d_qkv[:, (hq + hkv) * hd:], B, S, hq, hkv, hd, win, scale, ld)
                    ops.attn_dq_from_dst(dS, a["k"], dq, B, S, hq, hkv, hd, win, ld)
                else:
                    ops.attn_gemma_bwd(a["q"], a["k"], a["qkv"], d_o, a["kvbits"], a["lse"], delta, dS, dk,
                                       d_qkv[:, (hq + hkv) * hd:], B, S, hq, hkv, hd, win, scale, ld)
                    ops.gemm_raw(M=S, N=hd, K=S, A=dS, lda=ld, B=a["k"], ldb=hkv * hd, b_mn=True, Cout=dq,
                                 ldc=hq * hd, nz=(G, hkv, B), a_zs=(S * ld, G * S * ld, hq * S * ld),
                                 b_zs=(0, hd, S * hkv * hd), c_zs=(hd, G * hd, S * hq * hd), band=ops.BAND_K_BY_QUERY,
                                 band_window=win)
                ops.qknorm_rope_bwd(dq, dk, a["qkv"], L["q_norm"], L["k_norm"], cos_t, sin_t, d_qkv, S, hq,
                                    hkv, hd, c.eps)
                pre, st = lora_pre(li, "qkv", d_qkv)
                lora_wgrad(st, d_qkv, a["a_in"], a["t"].get("qkv"))
                ops.linear_dgrad(d_qkv, L["qkv_w"], out=d_y, **pre)
                if pair and li > 0:     # input-norm backward + the previous layer's post-ffn norm backward
                    prev = s["layers"][li - 1]
                    ops.rmsnorm_bwd_pair(d_y, a["x_in"], L["in_norm"], dx, prev["ym"],
                                         Lm.layers[li - 1]["post_ffn_norm"], d_y, c.eps)
                else:
                    ops.rmsnorm_bwd(d_y, a["x_in"], L["in_norm"], dx, c.eps, add=dx)
                continue
            # dP[b, head] = dO[b,:,head,:] V[b,:,kvh,:]^T
            if side:        # dV needs only P and dO: it runs beside delta + the dS GEMM
                ev_do.record(main)
                sv.wait_event(ev_do)
                with torch.cuda.stream(sv):
                    self._attn_dv(a, d_o, d_qkv, B, S, G, ld, QKV, win)
                    ev_dv.record(sv)
            if self.fuse_softmax_bwd:
                # dS = scale * P * (dO V^T - delta) in the GEMM epilogue, delta = rowsum(dO * O)
                ops.attn_delta(d_o, a["ao"], delta, S, hq, hd)
                self._scores_like(d_o, 0, hq * hd, (hd, G * hd, S * hq * hd), a["qkv"], (hq + hkv) * hd,
                                  QKV, (0, hd, S * QKV), dS, B, S, scale, window=win,
                                  act=ops.ACT_SOFTMAX_BWD, aux=a["P"], ld_aux=ld, bias=delta)
            else:
                self._scores_like(d_o, 0, hq * hd, (hd, G * hd, S * hq * hd), a["qkv"], (hq + hkv) * hd,
                                  QKV, (0, hd, S * QKV), dP, B, S, 1.0, window=win)
                ops.softmax_bwd(a["P"], dP, dS, B * hq * S, S, ld, scale, Sq=S, causal=1, window=win)
            # The three remaining contractions are independent persistent launches whose last wave leaves most
            # SMs idle (dV / dK: 96 tiles over 74 clusters); on separate streams the next launch's clusters take
            # those SMs as they free up.
            def dk_gemm():
                # dK[b,k,kvh,:] = sum_g dS[b,kvh*G+g]^T Q[b,:,kvh*G+g,:]
                ops.gemm_raw(M=S, N=hd, K=S, A=dS, lda=ld, a_mn=True, B=a["q"], ldb=hq * hd, b_mn=True,
                             Cout=dk, ldc=hkv * hd, nz=(hkv, B, 1), a_zs=(G * S * ld, hq * S * ld, 0),
                             b_zs=(G * hd, S * hq * hd, 0), c_zs=(hd, S * hkv * hd, 0), kg=G,
                             a_kgs=S * ld, b_kgs=hd, band=ops.BAND_K_BY_KEY, band_window=win,
                             bn_hint=self.bn_hint_dkv)
            if side:
                ev_ds.record(main)
                sk.wait_event(ev_ds)
                with torch.cuda.stream(sk):
                    dk_gemm()
                    ev_dk.record(sk)
            else:
                self._attn_dv(a, d_o, d_qkv, B, S, G, ld, QKV, win)
            # dQ[b,q,head,:] = dS[b,head] K[b,:,kvh,:]
            ops.gemm_raw(M=S, N=hd, K=S, A=dS, lda=ld, B=a["k"], ldb=hkv * hd, b_mn=True, Cout=dq,
                         ldc=hq * hd, nz=(G, hkv, B), a_zs=(S * ld, G * S * ld, hq * S * ld),
                         b_zs=(0, hd, S * hkv * hd), c_zs=(hd, G * hd, S * hq * hd), band=ops.BAND_K_BY_QUERY,
                         band_window=win)
            if side:
                main.wait_event(ev_dv)
                main.wait_event(ev_dk)
            else:
                dk_gemm()
            ops.qknorm_rope_bwd(dq, dk, a["qkv"], L["q_norm"], L["k_norm"], cos_t, sin_t, d_qkv, S, hq,
                                hkv, hd, c.eps)
            pre, st = lora_pre(li, "qkv", d_qkv)
            lora_wgrad(st, d_qkv, a["a_in"], a["t"].get("qkv"))
            ops.linear_dgrad(d_qkv, L["qkv_w"], out=d_y, **pre)
            if pair and li > 0:     # input-norm backward + the previous layer's post-ffn norm backward
                prev = s["layers"][li - 1]
                ops.rmsnorm_bwd_pair(d_y, a["x_in"], L["in_norm"], dx, prev["ym"],
                                     Lm.layers[li - 1]["post_ffn_norm"], d_y, c.eps)
            else:
                ops.rmsnorm_bwd(d_y, a["x_in"], L["in_norm"], dx, c.eps, add=dx)
        return dx

    # ------------------------------------------------------------------ projector backward
    def projector_backward(self, dx: torch.Tensor, patch: torch.Tensor, Z: torch.Tensor,
                           Hm: torch.Tensor, B: int, S: int):
        """dL/dX0[:, :n_img] -> the four projector gradients (flat f32 buffer), no gradient to the
        vision features (ref:Stage1/projectors.py:16-20 under autograd, projector_trainer.py:237)."""
        Pj = self.proj
        n_img = self.n_img
        R = B * n_img
        mid, h = Pj.shapes[0][0], Pj.shapes[2][0]
        dP = self.buf("p_dP", (R, h), BF16)
        ops.gather_cast_rows(dx, dP, group=(n_img, S, 0))
        g = Pj.grad
        ops.linear_wgrad(dP, Hm, out=Pj.view(g, 2))                 # dW2 = dP^T Hm
        ops.colsum(dP, Pj.view(g, 3))                               # db2
        dH = self.buf("p_dH", (R, mid), BF16)
        if self.fuse_gelu_bwd and mid % 32 == 0:
            ops.linear_dgrad_gelu_erf_bwd(dP, Pj.w2, Z, dH)         # dZ = (dP W2) * gelu'(Z) in the GEMM epilogue
        else:
            ops.linear_dgrad(dP, Pj.w2, out=dH)                     # dH = dP W2
            ops.gelu_erf_bwd(dH, Z, dH)                             # dZ in place
        ops.linear_wgrad(dH, patch, out=Pj.view(g, 0))              # dW1 = dZ^T patch
        ops.colsum(dH, Pj.view(g, 1))                               # db1

    # ------------------------------------------------------------------ optimizer
    def optimizer_step(self, lr: float, *, weight_decay: float = 0.01, max_norm: float = 5.0,
                       beta1: float = 0.9, beta2: float = 0.999, eps: float = 1e-8,
                       grad_scale: float = 1.0):
        """clip_grad_norm_(5.0) + AdamW on the flat projector buffer
        (ref:Stage1/projector_trainer.py:241-243); refreshes the bf16 shadow weights."""
        # one flat buffer per trainable module, each clipped on its own norm (Stage 2 clips module by module:
        # ref:Stage2/trainer.py:436-439) and updated with the same AdamW hyper-parameters (:145-150)
        for st in self.trainables():
            st.step += 1
            ops.sumsq(st.grad, st.sumsq)
            ops.clip_adamw(st.flat, st.grad, st.m, st.v, st.shadow, st.sumsq, grad_scale=grad_scale,
                           max_norm=max_norm, lr=lr, beta1=beta1, beta2=beta2, eps=eps, wd=weight_decay,
                           step=st.step)

    def trainables(self):
        """The flat-buffer states the optimizer steps: the projector (unless frozen) and the LoRA adapters."""
        return ([self.proj] if self.train_projector else []) + ([self.lora] if self.lora is not None else [])

    # ------------------------------------------------------------------ adapters folded into the weights
    def merge_adapters(self):
        """Decoder layers with W' = W + (alpha / r) * B A in bf16 (peft `merge_and_unload`), for the decode loop: one
        GEMM per packed projection, the base weight entering as the epilogue's residual.  The merged matrices live in
        buffers the engine keeps and refreshes in place on every call (a captured decode graph holds their addresses).
        Returns a layer list for `merged()`; None without adapters."""
        lo = self.lora
        if lo is None:
            return None
        if getattr(self, "_merged_layers", None) is None:
            self._merged_layers = []
            for li, L in enumerate(self.llm.layers):
                Lm = dict(L)
                for g, key in (("qkv", "qkv_w"), ("o", "o_w"), ("gu", "gu_w"), ("down", "down_w")):
                    if lo.layers[li].get(g) is not None:
                        Lm[key] = torch.empty_like(L[key])
                self._merged_layers.append(Lm)
        for li, (L, Lm) in enumerate(zip(self.llm.layers, self._merged_layers)):
            for g, key in (("qkv", "qkv_w"), ("o", "o_w"), ("gu", "gu_w"), ("down", "down_w")):
                grp = lo.layers[li].get(g)
                if grp is not None:
                    ops.linear_dgrad(grp.B(lo.shadow), grp.A(lo.shadow), out=Lm[key], residual=L[key], alpha=lo.scale)
        return self._merged_layers

    @contextlib.contextmanager
    def merged(self, layers):
        """Run with the merged decoder layers of merge_adapters() and the adapter path off (generation)."""
        if layers is None:
            yield
            return
        keep = (self.llm.layers, self.lora)
        self.llm.layers, self.lora = layers, None
        try:
            yield
        finally:
            self.llm.layers, self.lora = keep

    # ------------------------------------------------------------------ whole step
    def splice(self, token_ids: torch.Tensor, labels: torch.Tensor, x0: torch.Tensor, pad_id):
        B, T = token_ids.shape
        S = self.n_img + T
        attn_mask = self.buf("s_mask", (B, S), torch.int64)
        lm_labels = self.buf("s_labels", (B, S), torch.int64)
        keyvalid = self.buf("s_kv", (B, S), torch.uint8)
        targets = self.buf("s_tgt", (B, S), torch.int32)
        ops.embed_splice(token_ids, labels, self.llm.embed, self.llm.embed_scale, pad_id, self.n_img,
                         x0, attn_mask, lm_labels, keyvalid, targets, err=self._err)
        return attn_mask, lm_labels, keyvalid, targets

    def check_errors(self) -> None:
        """Raise if a kernel flagged an out-of-range token id or CE target since the last call (what
        `nn.Embedding` / `CrossEntropyLoss` raise on in the reference).  Costs a 4-byte device read: call it where
        the loss is read back anyway (the trainers do, once per optimizer step)."""
        flags = int(self._err.item())
        if flags:
            self._err.zero_()
            what = []
            if flags & ops.ERRFLAG_TOKEN_ID:
                what.append(f"a token id outside the embedding table [0, {self.llm.embed.shape[0]})")
            if flags & ops.ERRFLAG_TARGET:
                what.append(f"a label >= the vocabulary size {self.lcfg.vocab}")
            raise IndexError("Stage-1 step: " + " and ".join(what) + " (the tokenizer has ids the language model "
                             "does not: resize the embedding table or fix the pad / eos id)")

    def _span(self, T: int, label_span):
        """(first position whose shifted target can be a label, number of such positions) in sequence
        coordinates, from `label_span` = (offset of the first labelled text token, count) in text
        coordinates; None = the whole text (Stage 1)."""
        off, cnt = (0, T) if label_span is None else (int(label_span[0]), int(label_span[1]))
        if off < 0 or cnt <= 0 or off + cnt > T:
            raise ValueError(f"label_span {label_span} does not fit a text of {T} tokens")
        return self.n_img + off - 1, cnt

    def forward_backward(self, pixels: torch.Tensor, token_ids: torch.Tensor, labels: torch.Tensor,
                         pad_id=0, label_span=None) -> torch.Tensor:
        """Steps 1-8 of the reference loop (ref:Stage1/projector_trainer.py:157-237; with `label_span`
        the Stage-2 frozen-LLM step ref:Stage2/trainer.py:311-420): returns the device-resident scalar
        loss; projector gradients (of loss / loss_divisor) land in proj.grad."""
        B, T = token_ids.shape
        S = self.n_img + T
        lab0, Tl = self._span(T, label_span)
        h = self.lcfg.hidden
        with _phase("vision_forward"):
            patch = self.vision_forward(pixels)
        x0 = self.buf("x0", (B * S, h), F32)
        with _phase("projector_forward"):
            Z, Hm = self.projector_forward(patch, x0, rows_per_img=self.n_img, out_batch_stride=S * h)
        with _phase("splice"):
            _, _, keyvalid, targets = self.splice(token_ids, labels, x0, pad_id)
        with _phase("llm_forward_loss"):
            loss = self.llm_forward_loss(x0, keyvalid, targets, B, S, Tl, lab0=lab0)
        with _phase("llm_backward"):
            dx = self.llm_backward()
        if self.train_projector:
            with _phase("projector_backward"):
                self.projector_backward(dx, patch, Z, Hm, B, S)
        return loss

    def preprocess_images(self, images, out: torch.Tensor | None = None, *, mean=(0.5, 0.5, 0.5),
                          std=(0.5, 0.5, 0.5), rescale: float = 1 / 255) -> torch.Tensor:
        """Decoded RGB images (uint8 [H, W, 3] tensors of any size, host or device) -> `pixel_values`
        f32 [B, 3, S, S] on the device: what the reference's datasets do per item on the CPU with
        `image.resize((S, S))` + `processor(images=image)` (ref:Stage1/train_projection_stage1.py:99-104,
        ref:Stage2/dataset.py:90-95), here Pillow-exact on the GPU (`pt_image_preprocess`).  Runs on the
        current stream; host images are copied in first (non_blocking when pinned)."""
        S, dev = self.vcfg.image, self.device
        B = len(images)
        if out is None:
            out = torch.empty(B, 3, S, S, dtype=F32, device=dev)
        for i, im in enumerate(images):
            if im.dim() != 3 or im.shape[2] != 3 or im.dtype != torch.uint8:
                raise ValueError(f"image {i}: expected uint8 [H, W, 3], got {tuple(im.shape)} {im.dtype}")
            src = im if im.is_cuda else im.to(dev, non_blocking=True)
            H, W, _ = src.shape
            need = ops.image_preprocess_workspace(H, W, S, S)
            if not hasattr(self, "_img_ws"):
                self._img_ws = {}
            sid = torch.cuda.current_stream().cuda_stream      # one scratch buffer per stream that calls this
            ws = self._img_ws.get(sid)
            if ws is None or ws.numel() < need:
                ws = self._img_ws[sid] = torch.empty(need, dtype=torch.uint8, device=dev)
            ops.image_preprocess(src.contiguous(), out[i], ws, mean=mean, std=std, rescale=rescale)
        return out

    def raw_images(self, batch: dict):
        """The batch's undecoded-size uint8 images, or None: `images_u8` (list of [H, W, 3] tensors) with the slots
        of files that travel encoded (`jpeg` = ops.jpeg_pack + `index`, image_io.collate_images) filled by the GPU
        JPEG decode on the current stream."""
        raw, pack = batch.get("images_u8"), batch.get("jpeg")
        if pack is None:
            return raw
        if not hasattr(self, "_jpeg_ws"):
            self._jpeg_ws = {}
        sid = torch.cuda.current_stream().cuda_stream
        ws = self._jpeg_ws.get(sid)
        if ws is None or ws.numel() < pack["workspace_bytes"]:
            ws = self._jpeg_ws[sid] = torch.empty(pack["workspace_bytes"], dtype=torch.uint8, device=self.device)
        decoded = ops.jpeg_decode(pack, self.device, ws)
        raw = list(raw) if raw is not None else [None] * len(decoded)
        for i, im in zip(pack.get("index", range(len(decoded))), decoded):
            raw[i] = im
        return raw

    def stage_batch(self, batch: dict) -> dict:
        """Start the host->device copy of a (pinned) host batch on a side stream so that it overlaps the step
        that is running; returns a batch dict to hand to train_step later.  Two staging slots alternate (the
        DataLoader-prefetch pattern: while step i computes, batch i+1 is copied)."""
        if not hasattr(self, "_stage"):
            self._stage = dict(stream=torch.cuda.Stream(device=self.device), slots=[None, None], n=0)
        sg = self._stage
        i = sg["n"] & 1
        sg["n"] += 1
        has_raw = "images_u8" in batch or "jpeg" in batch   # undecoded-size images instead of processor output
        raw = None
        ti, lb = batch["token_ids"], batch["labels"]
        if has_raw:
            pv = torch.empty(ti.shape[0], 3, self.vcfg.image, self.vcfg.image, dtype=F32, device="meta")
        else:
            pv = batch["pixel_values"]
        slot = sg["slots"][i]
        if slot is None or slot["px"].shape != pv.shape or slot["ids"].shape != ti.shape:
            slot = dict(px=torch.empty(pv.shape, dtype=F32, device=self.device),
                        ids=torch.empty(ti.shape, dtype=torch.int64, device=self.device),
                        lab=torch.empty(lb.shape, dtype=torch.int64, device=self.device),
                        event=torch.cuda.Event(), done=torch.cuda.Event())
            slot["done"].record()
            sg["slots"][i] = slot
        with torch.cuda.stream(sg["stream"]):
            sg["stream"].wait_event(slot["done"])       # the step that consumed this slot has read it
            if has_raw:
                raw = self.raw_images(batch)             # JPEG decode of the encoded files, on the copy stream
                self.preprocess_images(raw, slot["px"])  # upload + resize + normalize on the copy stream
            else:
                slot["px"].copy_(pv, non_blocking=True)
            slot["ids"].copy_(ti, non_blocking=True)
            slot["lab"].copy_(lb, non_blocking=True)
            slot["event"].record()
        return {"pixel_values": slot["px"] if has_raw else pv, "token_ids": ti, "labels": lb,
                "_staged": slot, "label_span": batch.get("label_span")}

    def _graphed_forward_backward(self, batch: dict, pad_id) -> torch.Tensor:
        """forward_backward through a CUDA graph keyed on the batch geometry: inputs are copied into
        static device buffers, the first call per geometry runs eagerly (sizes the workspaces), the
        second captures, later calls replay.  The all-reduce and the optimizer stay outside the graph
        (lr and the step count change every step)."""
        pv, ti, lb = batch["pixel_values"], batch["token_ids"], batch["labels"]
        span = batch.get("label_span")
        span = None if span is None else (int(span[0]), int(span[1]))
        key = (tuple(pv.shape), tuple(ti.shape), pad_id, span, self.loss_divisor)
        g = self._graphs.get(key)
        if g is None:
            while len(self._graphs) >= max(1, self.max_graphs):      # least recently used graph of this geometry
                self._graphs.popitem(last=False)
            g = dict(px=torch.empty(pv.shape, dtype=F32, device=self.device),
                     ids=torch.empty(ti.shape, dtype=torch.int64, device=self.device),
                     lab=torch.empty(lb.shape, dtype=torch.int64, device=self.device),
                     graph=None, calls=0, loss=None)
            self._graphs[key] = g
        else:
            self._graphs.move_to_end(key)
        st = batch.get("_staged")
        if st is not None:
            # inputs were prefetched by stage_batch() on the copy stream: order after that copy, then D2D
            torch.cuda.current_stream().wait_event(st["event"])
            pv, ti, lb = st["px"], st["ids"], st["lab"]
        g["px"].copy_(pv, non_blocking=True)
        g["ids"].copy_(ti, non_blocking=True)
        g["lab"].copy_(lb, non_blocking=True)
        if st is not None:
            st["done"].record()                 # staging slot reusable once these copies have run
        g["calls"] += 1
        if g["calls"] == 1:
            n0 = _lib.launch_count
            loss = self.forward_backward(g["px"], g["ids"], g["lab"], pad_id, span)
            g["launches"] = _lib.launch_count - n0
            self._enforce_budget()
            return loss
        if g["graph"] is None:
            graph = torch.cuda.CUDAGraph()
            # thread_local: the DataLoader's pin-memory thread may call cudaHostAlloc while this thread captures
            with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                g["loss"] = self.forward_backward(g["px"], g["ids"], g["lab"], pad_id, span)
            g["graph"] = graph
        g["graph"].replay()
        _lib.launch_count += g["launches"]          # kernels inside the replayed graph
        return g["loss"]

    @torch.no_grad()
    def eval_loss(self, batch: dict, pad_id=0) -> torch.Tensor:
        """Validation forward (ref:Stage1/projector_trainer.py:304-370): the same loss, no backward."""
        dev = self.device
        if "pixel_values" in batch:
            px = batch["pixel_values"].to(dev, dtype=F32, non_blocking=True)
        else:
            px = self.preprocess_images(self.raw_images(batch))
        ids = batch["token_ids"].to(dev, non_blocking=True)
        lab = batch["labels"].to(dev, non_blocking=True)
        B, T = ids.shape
        S = self.n_img + T
        self.select_geometry((B, T))
        lab0, Tl = self._span(T, batch.get("label_span"))
        patch = self.vision_forward(px)
        x0 = self.buf("x0", (B * S, self.lcfg.hidden), F32)
        self.projector_forward(patch, x0, rows_per_img=self.n_img, out_batch_stride=S * self.lcfg.hidden)
        _, _, keyvalid, targets = self.splice(ids, lab, x0, pad_id)
        loss = self.llm_forward_loss(x0, keyvalid, targets, B, S, Tl, save=False, lab0=lab0)
        return loss.clone()

    # ------------------------------------------------------------------ inference: prefill + decode
    @torch.no_grad()
    def prefill(self, pixels: torch.Tensor, prompt_ids: torch.Tensor | None = None,
                max_new_tokens: int = 0, prompt_pad_id: int | None = None) -> torch.Tensor:
        """ref:Stage1/inference_stage1.py:150-175 (and projector_trainer.py:374-392): vision forward,
        `[:, 1:, :]`, projector, optional prompt embeddings appended, all-ones attention mask, frozen
        Gemma3 forward; the LM head runs on the LAST position only.  Returns f32 logits [B, V] and
        leaves a KV cache sized for `max_new_tokens` more positions in self.cache.  With
        `prompt_pad_id` (Stage-2 evaluation, ref:Stage2/trainer.py:594-611: `[image][left-padded
        question]`) prompt pads are masked as keys for the prefill and every decode step, and
        positions follow HF generate (cumsum of the mask - 1)."""
        c = self.lcfg
        dev = self.device
        px = pixels.to(dev, dtype=F32, non_blocking=True)
        B = px.shape[0]
        T = 0 if prompt_ids is None else prompt_ids.shape[1]
        S = self.n_img + T
        h = c.hidden
        patch = self.vision_forward(px)
        x0 = self.buf("x0", (B * S, h), F32)
        self.projector_forward(patch, x0, rows_per_img=self.n_img, out_batch_stride=S * h, save=False)
        keyvalid = pos_ids = None
        if T:
            ids = prompt_ids.to(dev, non_blocking=True)
            # pad_id None: every key valid (all-ones mask, ref:Stage1/inference_stage1.py:170)
            _, _, kv, _ = self.splice(ids, ids, x0, prompt_pad_id)
            if prompt_pad_id is not None:
                keyvalid = kv
                vm = kv.view(B, S).long()
                pos_ids = (vm.cumsum(1) - 1).masked_fill_(vm == 0, 1)   # index plumbing (HF generate positions)
        Smax = _pad8(S + max_new_tokens)
        kvd = c.hkv * c.hd
        key = (B, Smax)
        if getattr(self, "_cache_key", None) != key:
            # one tensor per operand, [layers, B, Smax, kvd]: a layer's cache is a contiguous slice of it, and beam
            # search re-orders the rows of all layers in one pass
            k_all = torch.zeros(c.layers, B, Smax, kvd, dtype=BF16, device=dev)
            v_all = torch.zeros(c.layers, B, Smax, kvd, dtype=BF16, device=dev)
            self.cache = dict(k=list(k_all.unbind(0)), v=list(v_all.unbind(0)), k_all=k_all, v_all=v_all)
            self._cache_key = key
        self.cache.update(pos=S, B=B, Smax=Smax, valid=None, next_pos=None)
        if keyvalid is not None:
            valid = torch.ones(B, Smax, dtype=torch.uint8, device=dev)
            valid[:, :S].copy_(keyvalid.view(B, S))
            self.cache.update(valid=valid, next_pos=keyvalid.view(B, S).long().sum(1))
        # device-side position state of the graphed decode step (one slot index for the whole batch, one RoPE
        # position per sample, key validity per cache slot); allocated once per cache, refilled per prefill
        kc = self.cache
        if kc.get("g_valid") is None or kc["g_valid"].shape != (B, Smax):
            kc.update(g_valid=torch.ones(B, Smax, dtype=torch.uint8, device=dev),
                      g_slot=torch.zeros(1, dtype=torch.int64, device=dev),
                      g_pos=torch.zeros(B, dtype=torch.int64, device=dev),
                      g_ar=torch.arange(Smax, dtype=torch.int64, device=dev), g_graph=None, g_calls=0)
        kc["g_valid"].fill_(1)
        if keyvalid is not None:
            kc["g_valid"][:, :S].copy_(keyvalid.view(B, S))
            kc["g_pos"].copy_(kc["next_pos"])
        else:
            kc["g_pos"].fill_(S)
        kc["g_slot"].fill_(S)
        if "d_yacc" in kc.get("g_ws", {}):          # the split-K accumulator is empty between uses; make sure of it
            kc["g_ws"]["d_yacc"].zero_()            # after a step that did not finish
        x, _ = self._llm_stack(x0, keyvalid, B, S, save=False, cache=self.cache, pos_ids=pos_ids)
        return self._last_logits(x, B, S)

    def _last_logits(self, x: torch.Tensor, B: int, S: int, buf=None, skinny: bool = False) -> torch.Tensor:
        c, Lm = self.lcfg, self.llm
        buf = buf or self.buf
        yf = buf("i_yf", (B, c.hidden), BF16)
        ops.rmsnorm_fwd(x, Lm.final_norm, yf, c.eps, group=(1, S, S - 1))
        logits = buf("i_logits", (B, _pad8(c.vocab)), F32)[:, :c.vocab]
        if skinny:
            ops.gemv(yf, Lm.lm_head, logits, ops.GEMV_F32)
        else:
            ops.linear(yf, Lm.lm_head, out=logits)
        return logits

    @torch.no_grad()
    def decode_step(self, tokens: torch.Tensor) -> torch.Tensor:
        """One cached decode step: tokens i64 [B] at position cache['pos'] -> f32 logits [B, V]
        (HF generate's per-token forward; sliding layers attend to the last `window` keys only).
        By default the step runs as ONE CUDA graph whose position lives on the device (fixed shapes: scores over
        the whole cache, masked), so every token of every sequence replays the same graph; PT_STAGE1_DECODE_GRAPH=0
        selects the eager step with exact-length contractions."""
        if self.lora is not None:
            raise RuntimeError("decode_step with live LoRA adapters: run it inside `with engine.merged(engine."
                               "merge_adapters())` (generate() does)")
        kc = self.cache
        if kc["pos"] >= kc["Smax"]:
            raise RuntimeError("KV cache full: prefill(max_new_tokens=...) was sized too small")
        if os.environ.get("PT_STAGE1_DECODE_GRAPH", "1") == "0":
            return self._decode_step_eager(tokens)
        B = kc["B"]
        if kc.get("g_tokens") is None or kc["g_tokens"].shape[0] != B:
            kc["g_tokens"] = torch.zeros(B, 1, dtype=torch.int64, device=self.device)
        kc["g_tokens"].copy_(tokens.to(self.device).view(B, 1), non_blocking=True)
        variant = (self.fuse_decode_attn, self.decode_splitk, self.decode_pdl, self.decode_gemv)
        if kc.get("g_variant") != variant:          # the captured graph belongs to one variant of the step
            kc.update(g_variant=variant, g_graph=None, g_calls=0)
        kc["g_calls"] += 1
        if kc["g_calls"] == 1:                       # sizes the workspaces
            logits = self._decode_step_fixed()
        else:
            if kc["g_graph"] is None:
                graph = torch.cuda.CUDAGraph()
                with torch.cuda.graph(graph, capture_error_mode="thread_local"):
                    kc["g_logits"] = self._decode_step_fixed()
                kc["g_graph"] = graph
            kc["g_graph"].replay()
            logits = kc["g_logits"]
        kc["pos"] += 1
        if kc.get("next_pos") is not None:
            kc["next_pos"] = kc["next_pos"] + 1
        return logits

    def _decode_step_fixed(self) -> torch.Tensor:
        """The decode step with position-independent shapes: the new key / value go to cache slot g_slot (device
        scalar) and attention sees the slots up to g_slot (in sliding layers from g_slot - window + 1) that are not
        prompt pads; RoPE uses the per-sample positions g_pos.  Per layer the attention is ONE launch
        (pt_decode_attn: norm + RoPE, cache append, flash-decoding over a CTA cluster), or, for geometries it does not
        take, six (scores and P V over all Smax slots, masked in the softmax).  Ends by advancing g_slot / g_pos, so
        a captured graph serves every step."""
        was = _lib.lib().pt_set_dependent_launch(self.decode_pdl)
        try:
            return self._decode_step_fixed_body()
        finally:
            _lib.lib().pt_set_dependent_launch(was)

    def _decode_step_fixed_body(self) -> torch.Tensor:
        c, Lm, kc = self.lcfg, self.llm, self.cache
        B, Smax = kc["B"], kc["Smax"]
        h, hd, hq, hkv, ff = c.hidden, c.hd, c.hq, c.hkv, c.ff
        G, QKV, kvd = hq // hkv, (hq + 2 * hkv) * hd, hkv * hd
        scale = c.q_scalar ** -0.5
        slot, posv, ar = kc["g_slot"], kc["g_pos"], kc["g_ar"]
        ids = kc["g_tokens"]
        ws = kc.setdefault("g_ws", {})      # owned by the cache, not by a batch-geometry arena: the graph keeps pointers

        def dbuf(name, shape, dtype):
            t = ws.get(name)
            if t is None:
                t = ws[name] = torch.empty(shape, dtype=dtype, device=self.device)
            return t
        x = dbuf("d_x", (B, h), F32)
        ops.embed_splice(ids, ids, Lm.embed, Lm.embed_scale, None, 0, x,
                         dbuf("d_mask", (B, 1), torch.int64), dbuf("d_lab", (B, 1), torch.int64),
                         dbuf("d_kv", (B, 1), torch.uint8), dbuf("d_tgt", (B, 1), torch.int32))
        a = dbuf("d_a", (B, h), BF16)
        qkv = dbuf("d_qkv", (B, QKV), BF16)
        attn_o = dbuf("d_o", (B, hq * hd), BF16)
        y = dbuf("d_y", (B, h), BF16)
        x1 = dbuf("d_x1", (B, h), F32)
        gu = dbuf("d_gu", (B, 2 * ff), BF16)
        act = dbuf("d_act", (B, ff), BF16)
        fused = self.fuse_decode_attn
        splitk = self.decode_splitk
        skinny = self.decode_gemv and all(ops.gemv_ok(B, k, self.decode_gemv_rows) for k in (h, hq * hd, ff))
        if splitk and "d_yacc" not in ws:
            ws["d_yacc"] = torch.zeros((B, h), dtype=F32, device=self.device)   # empty between uses, see the kernel
        yacc = ws.get("d_yacc")
        # slices per 128-column tile so that the launch is about one CTA pair per SM pair (measured at B = 64 on
        # Gemma3-1B: down 34.5 us unsplit, 14.9 with the generic wave-fill choice on 64-wide tiles, 11.7 this way)
        sk = lambda K: max(2, min(K // 64, round(74 / -(-h // 128))))
        sk_o, sk_down = sk(hq * hd), sk(ff)
        tabs = {}
        if fused:       # the kernel reads the table row g_pos[b] and masks from g_slot / g_valid itself
            for f in (True, False):
                tabs[f] = self.rope(Smax, f)
        else:
            q_post = dbuf("d_q", (B, hq * hd), BF16)
            k_post = dbuf("d_k", (B, kvd), BF16)
            ld = _pad8(Smax)
            sc = dbuf("d_sc", (B * hq, ld), F32)
            P = dbuf("d_P", (B * hq, ld), BF16)
            # key masks of this step (index plumbing on [B, Smax] bytes): causal, prompt pads, sliding window
            m_full = ((ar <= slot) & kc["g_valid"].bool()).to(torch.uint8)
            m_slide = (m_full.bool() & (ar > slot - c.window)).to(torch.uint8) if c.window > 0 else m_full
            for f in (True, False):
                cos_t, sin_t = self.rope(Smax, f)
                tabs[f] = (cos_t.index_select(0, posv), sin_t.index_select(0, posv))
        ops.rmsnorm_fwd(x, Lm.layers[0]["in_norm"], a, c.eps)
        for li, L in enumerate(Lm.layers):
            full = c.is_full(li)
            if skinny:
                ops.gemv(a, L["qkv_w"], qkv)
            else:
                ops.linear(a, L["qkv_w"], out=qkv)
            K, V = kc["k"][li], kc["v"][li]
            if fused:
                ops.decode_attn(qkv, L["q_norm"], L["k_norm"], tabs[full][0], tabs[full][1], posv, slot,
                                kc["g_valid"], K, V, attn_o, hq, hkv, hd, 0 if full else c.window, c.eps, scale)
            else:
                self._decode_attn_unfused(qkv, L, tabs[full], q_post, k_post, K, V, slot, sc, P,
                                          m_full if full else m_slide, attn_o, B, Smax, ld, scale)
            last = li == len(Lm.layers) - 1
            if splitk:
                if skinny:
                    ops.gemv(attn_o, L["o_w"], yacc, ops.GEMV_F32_ATOMIC)
                else:
                    ops.linear(attn_o, L["o_w"], out=yacc, ksplit=sk_o, bn_hint=128)
                ops.rmsnorm_residual_fwd_acc(yacc, L["post_attn_norm"], x, x1, L["pre_ffn_norm"], a, c.eps)
            elif c.sandwich_norm:
                ops.linear(attn_o, L["o_w"], out=y)
                ops.rmsnorm_residual_fwd(y, L["post_attn_norm"], x, x1, L["pre_ffn_norm"], a, c.eps)
            else:
                if skinny:      # x1 = x + attn_o W_o^T: the K slices add onto a copy of the residual stream
                    x1.copy_(x)
                    ops.gemv(attn_o, L["o_w"], x1, ops.GEMV_F32_ATOMIC)
                else:
                    ops.linear(attn_o, L["o_w"], out=x1, residual=x)
                ops.rmsnorm_fwd(x1, L["pre_ffn_norm"], a, c.eps)
            if skinny:
                ops.gemv(a, L["gu_w"], act, ops.GEMV_SWIGLU if c.silu else ops.GEMV_GEGLU)
            elif self.fuse_geglu:
                ops.linear_geglu(a, L["gu_w"], gu, act, silu=c.silu)
            else:
                ops.linear(a, L["gu_w"], out=gu)
                ops.geglu_fwd(gu, act, silu=c.silu)
            if splitk:
                if skinny:
                    ops.gemv(act, L["down_w"], yacc, ops.GEMV_F32_ATOMIC)
                else:
                    ops.linear(act, L["down_w"], out=yacc, ksplit=sk_down, bn_hint=128)
                ops.rmsnorm_residual_fwd_acc(yacc, L["post_ffn_norm"], x1, x,
                                             None if last else Lm.layers[li + 1]["in_norm"],
                                             None if last else a, c.eps)
            elif c.sandwich_norm:
                ops.linear(act, L["down_w"], out=y)
                ops.rmsnorm_residual_fwd(y, L["post_ffn_norm"], x1, x,
                                         None if last else Lm.layers[li + 1]["in_norm"],
                                         None if last else a, c.eps)
            else:
                if skinny:
                    x.copy_(x1)
                    ops.gemv(act, L["down_w"], x, ops.GEMV_F32_ATOMIC)
                else:
                    ops.linear(act, L["down_w"], out=x, residual=x1)
                if not last:
                    ops.rmsnorm_fwd(x, Lm.layers[li + 1]["in_norm"], a, c.eps)
        logits = self._last_logits(x, B, 1, buf=dbuf, skinny=skinny)
        slot.add_(1)
        posv.add_(1)
        return logits

    def _decode_attn_unfused(self, qkv, L, tab, q_post, k_post, K, V, slot, sc, P, mask, attn_o, B, Smax, ld, scale):
        """The decode attention as six launches (geometries pt_decode_attn does not take, and the A/B of it):
        scores and P V run over all Smax slots, `mask` zeroes what the step may not see."""
        c = self.lcfg
        hd, hq, hkv = c.hd, c.hq, c.hkv
        G, kvd = hq // hkv, hkv * hd
        ops.qknorm_rope_fwd(qkv, L["q_norm"], L["k_norm"], tab[0], tab[1], q_post, k_post, B, hq, hkv, hd, c.eps)
        K.index_copy_(1, slot, k_post.view(B, 1, kvd))
        V.index_copy_(1, slot, qkv[:, (hq + hkv) * hd:].reshape(B, 1, kvd))
        ops.gemm_raw(M=G, N=Smax, K=hd, A=q_post, lda=hd, B=K, ldb=kvd, Cout=sc, ldc=ld,
                     nz=(hkv, B, 1), a_zs=(G * hd, hq * hd, 0), b_zs=(hd, Smax * kvd, 0),
                     c_zs=(G * ld, hq * ld, 0), alpha=scale)
        ops.softmax_fwd(sc, P, mask, B * hq, 1, Smax, ld, hq, 0, 0)
        ops.gemm_raw(M=G, N=hd, K=Smax, A=P, lda=ld, B=V, ldb=kvd, b_mn=True, Cout=attn_o, ldc=hd,
                     nz=(hkv, B, 1), a_zs=(G * ld, hq * ld, 0), b_zs=(hd, Smax * kvd, 0),
                     c_zs=(G * hd, hq * hd, 0))

    def _decode_step_eager(self, tokens: torch.Tensor) -> torch.Tensor:
        c, Lm, kc = self.lcfg, self.llm, self.cache
        B, pos, Smax = kc["B"], kc["pos"], kc["Smax"]
        h, hd, hq, hkv, ff = c.hidden, c.hd, c.hq, c.hkv, c.ff
        G = hq // hkv
        QKV = (hq + 2 * hkv) * hd
        kvd = hkv * hd
        scale = c.q_scalar ** -0.5
        ids = tokens.to(self.device).view(B, 1)
        x = self.buf("d_x", (B, h), F32)
        ops.embed_splice(ids, ids, Lm.embed, Lm.embed_scale, None, 0, x,
                         self.buf("d_mask", (B, 1), torch.int64), self.buf("d_lab", (B, 1), torch.int64),
                         self.buf("d_kv", (B, 1), torch.uint8), self.buf("d_tgt", (B, 1), torch.int32))
        a = self.buf("d_a", (B, h), BF16)
        qkv = self.buf("d_qkv", (B, QKV), BF16)
        q_post = self.buf("d_q", (B, hq * hd), BF16)
        k_post = self.buf("d_k", (B, kvd), BF16)
        ldm = _pad8(Smax)
        sc = self.buf("d_sc", (B * hq, ldm), F32)
        P = self.buf("d_P", (B * hq, ldm), BF16)
        attn_o = self.buf("d_o", (B, hq * hd), BF16)
        y = self.buf("d_y", (B, h), BF16)
        x1 = self.buf("d_x1", (B, h), F32)
        gu = self.buf("d_gu", (B, 2 * ff), BF16)
        act = self.buf("d_act", (B, ff), BF16)
        ops.rmsnorm_fwd(x, Lm.layers[0]["in_norm"], a, c.eps)
        valid, next_pos = kc.get("valid"), kc.get("next_pos")
        rope_rows, tabs = 1, {}
        for f in (True, False):
            cos_t, sin_t = self.rope(Smax, f)
            if next_pos is None:
                tabs[f] = (cos_t[pos:pos + 1], sin_t[pos:pos + 1])
            else:       # per-sample positions (padded prompt): one table row per sample
                tabs[f] = (cos_t.index_select(0, next_pos), sin_t.index_select(0, next_pos))
                rope_rows = B
        for li, L in enumerate(Lm.layers):
            full = c.is_full(li)
            ops.linear(a, L["qkv_w"], out=qkv)
            ops.qknorm_rope_fwd(qkv, L["q_norm"], L["k_norm"], tabs[full][0], tabs[full][1],
                                q_post, k_post, rope_rows, hq, hkv, hd, c.eps)
            K, V = kc["k"][li], kc["v"][li]
            K[:, pos].copy_(k_post)
            V[:, pos].copy_(qkv[:, (hq + hkv) * hd:])
            lo = 0 if full else max(0, pos - c.window + 1)
            n = pos + 1 - lo
            ld = _pad8(n)
            # scores[b, kvh*G+g, k] = scale * q[b, kvh*G+g, :] . K[b, lo+k, kvh, :]
            ops.gemm_raw(M=G, N=n, K=hd, A=q_post, lda=hd, B=K, b_off=lo * kvd, ldb=kvd, Cout=sc, ldc=ld,
                         nz=(hkv, B, 1), a_zs=(G * hd, hq * hd, 0), b_zs=(hd, Smax * kvd, 0),
                         c_zs=(G * ld, hq * ld, 0), alpha=scale)
            kvm = None if valid is None else valid[:, lo:pos + 1].contiguous()
            ops.softmax_fwd(sc, P, kvm, B * hq, 1, n, ld, hq, 0, 0)
            ops.gemm_raw(M=G, N=hd, K=n, A=P, lda=ld, B=V, b_off=lo * kvd, ldb=kvd, b_mn=True,
                         Cout=attn_o, ldc=hd, nz=(hkv, B, 1), a_zs=(G * ld, hq * ld, 0),
                         b_zs=(hd, Smax * kvd, 0), c_zs=(G * hd, hq * hd, 0))
            last = li == len(Lm.layers) - 1
            if c.sandwich_norm:
                ops.linear(attn_o, L["o_w"], out=y)
                ops.rmsnorm_residual_fwd(y, L["post_attn_norm"], x, x1, L["pre_ffn_norm"], a, c.eps)
            else:
                ops.linear(attn_o, L["o_w"], out=x1, residual=x)
                ops.rmsnorm_fwd(x1, L["pre_ffn_norm"], a, c.eps)
            if self.fuse_geglu:
                ops.linear_geglu(a, L["gu_w"], gu, act, silu=c.silu)
            else:
                ops.linear(a, L["gu_w"], out=gu)
                ops.geglu_fwd(gu, act, silu=c.silu)
            if c.sandwich_norm:
                ops.linear(act, L["down_w"], out=y)
                ops.rmsnorm_residual_fwd(y, L["post_ffn_norm"], x1, x,
                                         None if last else Lm.layers[li + 1]["in_norm"],
                                         None if last else a, c.eps)
            else:
                ops.linear(act, L["down_w"], out=x, residual=x1)
                if not last:
                    ops.rmsnorm_fwd(x, Lm.layers[li + 1]["in_norm"], a, c.eps)
        kc["pos"] = pos + 1
        if next_pos is not None:
            kc["next_pos"] = next_pos + 1
        return self._last_logits(x, B, 1)

    @torch.no_grad()
    def generate(self, pixels: torch.Tensor, prompt_ids: torch.Tensor | None = None, *,
                 max_new_tokens: int = 64, do_sample: bool = False, temperature: float = 1.0,
                 top_p: float = 1.0, top_k: int = 50, repetition_penalty: float = 1.0,
                 pad_token_id: int | None = 0, eos_token_id: int | None = 1,
                 generator: torch.Generator | None = None, prompt_pad_id: int | None = None,
                 num_beams: int = 1, length_penalty: float = 1.0) -> torch.Tensor:
        """`language_model.generate(inputs_embeds=projected, attention_mask=ones, ...)` of
        ref:Stage1/inference_stage1.py:176-187 / projector_trainer.py:386-393: returns the NEW tokens
        only, i64 [B, n<=max_new_tokens]; finished rows are filled with pad_token_id; stops early when
        every row has produced eos.  Token selection (HF logits processors: repetition penalty ->
        temperature -> top-k -> top-p -> multinomial, or argmax) and the finished-row bookkeeping are one launch per
        token (pt_sample_token; the draw has torch.multinomial's distribution, not its random stream); top_k = 0 or
        > 1024 with sampling takes the torch restatement of the same logic below."""
        if self.lora is not None:
            # the decode loop reads the packed weights only: fold the adapters in for the length of this call
            kw = dict(max_new_tokens=max_new_tokens, do_sample=do_sample, temperature=temperature, top_p=top_p,
                      top_k=top_k, repetition_penalty=repetition_penalty, pad_token_id=pad_token_id,
                      eos_token_id=eos_token_id, generator=generator, prompt_pad_id=prompt_pad_id,
                      num_beams=num_beams, length_penalty=length_penalty)
            with self.merged(self.merge_adapters()):
                return self.generate(pixels, prompt_ids, **kw)
        if num_beams > 1:
            return self._beam_generate(pixels, prompt_ids, max_new_tokens=max_new_tokens, do_sample=do_sample,
                                       temperature=temperature, top_p=top_p, top_k=top_k,
                                       repetition_penalty=repetition_penalty,
                                       pad_token_id=pad_token_id, eos_token_id=eos_token_id, generator=generator,
                                       prompt_pad_id=prompt_pad_id, num_beams=num_beams, length_penalty=length_penalty)
        logits = self.prefill(pixels, prompt_ids, max_new_tokens=max_new_tokens, prompt_pad_id=prompt_pad_id)
        B = logits.shape[0]
        dev = self.device
        out = torch.full((B, max_new_tokens), pad_token_id if pad_token_id is not None else 0,
                         dtype=torch.int64, device=dev)
        n = 0
        if self.fuse_sampler and (not do_sample or (top_k and 0 < top_k <= 1024)) and temperature > 0 and top_p > 0:
            # one launch per token: processors, draw and finished-row bookkeeping (pt_sample_token); the uniform
            # numbers of all steps come from `generator` up front
            u = None
            if do_sample:
                on_dev = generator is None or generator.device.type == "cuda"
                u = torch.rand((max_new_tokens, B), generator=generator, dtype=F32,
                               device=dev if on_dev else "cpu").to(dev)
            done8 = torch.zeros(B, dtype=torch.uint8, device=dev)
            nxt = torch.empty(B, dtype=torch.int64, device=dev)
            for step in range(max_new_tokens):
                ops.sample_token(logits, out=out, step=step, next_tokens=nxt,
                                 prev=out if repetition_penalty != 1.0 and step > 0 else None,
                                 repetition_penalty=repetition_penalty, temperature=temperature,
                                 top_k=top_k if do_sample else 0, top_p=top_p, u=None if u is None else u[step],
                                 done=done8, pad_id=pad_token_id, eos_id=eos_token_id)
                n = step + 1
                # "has every row emitted eos?" costs a host sync: asked every 8th token, the exact stopping point is
                # recovered from the tokens below (finished rows only emit pad, the uniform numbers are per step)
                if eos_token_id is not None and (n % 8 == 0 or n == max_new_tokens) and bool(done8.all()):
                    break
                if step + 1 < max_new_tokens:
                    logits = self.decode_step(nxt)
            out = out[:, :n]
            if eos_token_id is not None:
                hit = out == eos_token_id
                if bool(hit.any(dim=1).all()):       # HF stops after the step in which the last row emits eos
                    out = out[:, :int(hit.float().argmax(dim=1).max()) + 1]
            return out
        done = torch.zeros(B, dtype=torch.bool, device=dev)
        for step in range(max_new_tokens):
            sc = logits.float().clone()
            if repetition_penalty != 1.0 and step > 0:
                prev = out[:, :step]
                g = sc.gather(1, prev)
                sc.scatter_(1, prev, torch.where(g < 0, g * repetition_penalty, g / repetition_penalty))
            if do_sample:
                if temperature != 1.0:
                    sc = sc / temperature
                if top_k and top_k > 0:
                    kth = sc.topk(min(top_k, sc.shape[1]), dim=1).values[:, -1:]
                    sc = sc.masked_fill(sc < kth, float("-inf"))
                if top_p < 1.0:
                    srt, idx = sc.sort(dim=1, descending=False)
                    cum = srt.softmax(dim=1).cumsum(dim=1)
                    rm = cum <= (1.0 - top_p)
                    rm[:, -1:] = False
                    sc = sc.masked_fill(rm.scatter(1, idx, rm), float("-inf"))
                nxt = torch.multinomial(sc.softmax(dim=1), 1, generator=generator).squeeze(1)
            else:
                nxt = sc.argmax(dim=1)
            if pad_token_id is not None:
                nxt = torch.where(done, torch.full_like(nxt, pad_token_id), nxt)
            out[:, step] = nxt
            n = step + 1
            if eos_token_id is not None:
                done |= nxt == eos_token_id
                if bool(done.all()):
                    break
            if step + 1 < max_new_tokens:
                logits = self.decode_step(nxt)
        return out[:, :n]

    @torch.no_grad()
    def _beam_generate(self, pixels, prompt_ids, *, max_new_tokens, do_sample, temperature, top_p, top_k,
                       pad_token_id, eos_token_id, generator, prompt_pad_id, num_beams, length_penalty,
                       repetition_penalty=1.0):
        """HF `generate(num_beams=n)` (beam search, or beam sampling with do_sample=True: what the Stage-2 validation
        asks for, ref:Stage2/trainer.py:603-611 `num_beams=3, do_sample=True, top_p=0.9, top_k=50`), restated from
        `GenerationMixin._beam_search` of transformers 5.x with early_stopping=False and no prompt ids (generation
        starts from `inputs_embeds`, so lengths count generated tokens only): every sample runs `num_beams` rows of
        the KV-cached decode; per step the log-probabilities of all beams' continuations are ranked (or sampled
        without replacement), 2n candidates are kept, the best n unfinished ones run on, candidates among the first n
        that hit eos / the length limit compete for the n finished slots with score / length^penalty, and the KV
        cache rows are re-ordered to follow their beams.  Bookkeeping is torch on [B, n, V] tensors; the model
        arithmetic is the same decode step as greedy generation."""
        nb, dev = int(num_beams), self.device
        B = pixels.shape[0]
        rep_i = lambda t: None if t is None else t.repeat_interleave(nb, dim=0)
        logits = self.prefill(rep_i(pixels), rep_i(prompt_ids), max_new_tokens=max_new_tokens, prompt_pad_id=prompt_pad_id)
        V, L = logits.shape[1], int(max_new_tokens)
        fill = pad_token_id if pad_token_id else (eos_token_id if eos_token_id is not None else -1)
        keep = 2 * nb                                    # (number of eos ids + 1) * num_beams
        top_mask = torch.arange(keep, device=dev) < nb
        run_seq = torch.full((B, nb, L), fill, dtype=torch.int64, device=dev)
        fin_seq = run_seq.clone()
        run_score = torch.zeros(B, nb, device=dev)
        run_score[:, 1:] = -1e9                          # identical beams at the start: only the first one counts
        fin_score = torch.full((B, nb), -1e9, device=dev)
        fin_len = torch.zeros(B, nb, dtype=torch.int64, device=dev)
        finished = torch.zeros(B, nb, dtype=torch.bool, device=dev)
        improvable = torch.ones(B, 1, dtype=torch.bool, device=dev)
        gather = lambda t, idx: t.gather(1, idx.view(B, -1, *([1] * (t.dim() - 2))).expand(-1, -1, *t.shape[2:]))
        kc = self.cache
        overlap = dev.type == "cuda"
        if overlap:
            stop_h = torch.zeros(1, dtype=torch.bool, pin_memory=True)
            stop_ev = torch.cuda.Event()
        for cur in range(L):
            lp = torch.log_softmax(logits.float(), dim=-1)
            if repetition_penalty != 1.0 and cur > 0:    # HF's processor runs on the log-probabilities of each beam,
                prev = run_seq[:, :, :cur].reshape(B * nb, cur)      # over the tokens that beam has generated
                gsel = lp.gather(1, prev)
                lp = lp.scatter(1, prev, torch.where(gsel < 0, gsel * repetition_penalty, gsel / repetition_penalty))
            small = do_sample and bool(top_k) and top_k > 0
            if small:
                # HF's warpers (min_tokens_to_keep = 2 under beam search) on the top-k candidates only: after
                # TopKLogitsWarper every other entry of the row is -inf (probability 0), so temperature, top-p, the joint
                # softmax over the beams and the draw without replacement see exactly these values.  Nothing of size
                # [rows, V] is touched after the top-k.  (Entries tied with the k-th value beyond the list are left out:
                # HF keeps them; for float log-probabilities that is a set of measure zero.)
                if temperature != 1.0:
                    lp = lp / temperature
                kk = min(max(top_k, 2), V)
                vals, tix = lp.topk(kk, dim=1)                                   # descending, [rows, kk]
                if top_p < 1.0:
                    # ascending cumulative probability <= 1 - top_p goes, the two largest stay
                    asc = vals.flip(1)
                    rm = asc.softmax(dim=1).cumsum(dim=1) <= (1.0 - top_p)
                    rm[:, -2:] = False
                    vals = vals.masked_fill(rm.flip(1), float("-inf"))
                acc = (vals.view(B, nb, kk) + run_score[:, :, None]).view(B, nb * kk)
                pick = torch.multinomial(acc.softmax(dim=-1), keep, generator=generator)
                cand_score = acc.gather(1, pick)
                cand = (pick // kk) * V + tix.view(B, nb * kk).gather(1, pick)      # beam * V + token
            else:
                if do_sample:
                    if temperature != 1.0:
                        lp = lp / temperature
                    if top_p < 1.0:
                        srt, idx = lp.sort(dim=1, descending=False)
                        rm = srt.softmax(dim=1).cumsum(dim=1) <= (1.0 - top_p)
                        rm[:, -2:] = False
                        lp = lp.masked_fill(rm.scatter(1, idx, rm), float("-inf"))
                if do_sample:
                    acc = (lp.view(B, nb, V) + run_score[:, :, None]).view(B, nb * V)
                    cand = torch.multinomial(acc.softmax(dim=-1), keep, generator=generator)
                    cand_score = acc.gather(1, cand)
                else:       # the best `keep` of all beams' continuations are among each beam's own best `keep`
                    vals, tix = lp.topk(keep, dim=1)
                    acc = (vals.view(B, nb, keep) + run_score[:, :, None]).view(B, nb * keep)
                    cand_score, pick = acc.topk(keep, dim=1)
                    cand = (pick // keep) * V + tix.view(B, nb * keep).gather(1, pick)
            src_beam, tok = cand // V, cand % V
            cand_seq = gather(run_seq, src_beam)
            cand_seq[:, :, cur] = tok
            hit = (cur + 1 >= L) | ((tok == eos_token_id) if eos_token_id is not None else torch.zeros_like(tok, dtype=torch.bool))
            # the running beams of the next step: best unfinished candidates
            live_score = cand_score + hit.float() * -1e9
            nxt = live_score.topk(nb, dim=1).indices
            run_seq, run_score, run_src = gather(cand_seq, nxt), live_score.gather(1, nxt), src_beam.gather(1, nxt)
            # finished pool: candidates among the first num_beams that just ended compete with the stored ones
            just = hit & top_mask[None, :]
            fs = cand_score / float(cur + 1) ** length_penalty
            fs = fs + (~improvable).float() * -1e9 + (~just).float() * -1e9
            m_score = torch.cat([fin_score, fs], dim=1)
            pick = m_score.topk(nb, dim=1).indices
            fin_seq = gather(torch.cat([fin_seq, cand_seq], dim=1), pick)
            fin_len = torch.cat([fin_len, torch.full_like(tok, cur + 1)], dim=1).gather(1, pick)
            finished = torch.cat([finished, just], dim=1).gather(1, pick)
            fin_score = m_score.gather(1, pick)
            # can a running beam still beat the worst finished one?  (early_stopping=False heuristic)
            best_run = run_score[:, :1] / float(cur + 1) ** length_penalty
            worst_fin = torch.where(finished, fin_score.min(dim=1, keepdim=True).values, torch.full_like(fin_score, -1e9))
            improvable = improvable & (best_run > worst_fin).any(dim=-1, keepdim=True)
            stop = (~improvable.any()) | hit.all()
            if overlap:                                  # the answer travels to the host behind an event ...
                stop_h.copy_(stop.view(1), non_blocking=True)
                stop_ev.record()
            elif bool(stop):
                break
            if cur + 1 < L:
                # KV rows follow their beams (in place: the decode graph holds the cache addresses)
                rows = (run_src + torch.arange(B, device=dev)[:, None] * nb).view(-1)
                kc["k_all"].copy_(kc["k_all"].index_select(1, rows))
                kc["v_all"].copy_(kc["v_all"].index_select(1, rows))
                logits = self.decode_step(run_seq[:, :, cur].reshape(-1))
            if overlap:                                  # ... while the next decode step is already queued: the host
                stop_ev.synchronize()                    # waits for the flag only and prepares the next step's
                if bool(stop_h[0]):                      # bookkeeping during the decode (a step queued past the
                    break                                # stop is never read)
        out_len = int(fin_len[:, 0].max().item())
        return fin_seq[:, 0, :max(out_len, 1)]

    def train_step(self, batch: dict, lr: float, *, pad_id=0, weight_decay: float = 0.01,
                   allreduce=None, world_size: int = 1, max_norm: float = 5.0,
                   sync_gradients: bool = True) -> torch.Tensor:
        """One micro-batch (and, when `sync_gradients`, one optimizer step) from a host or device batch
        dict with the reference's keys `pixel_values`, `token_ids`, `labels`
        (ref:Stage1/train_projection_stage1.py:114-118) and optionally `label_span` (Stage 2, see
        forward_backward).  Host tensors are copied in (non_blocking from pinned memory).
        `sync_gradients=False` is a micro-batch inside `accelerator.accumulate()`
        (ref:Stage2/trainer.py:306, 422): its gradient is added to the accumulator and neither the
        all-reduce nor the optimizer run.  `allreduce(flat_grad)` sums the flat gradient over ranks
        (NCCL); the 1/world mean is folded into the optimizer kernel.  Returns the device scalar loss
        (no host sync here)."""
        dev = self.device
        if "pixel_values" not in batch:         # decoded images that were not staged: preprocess them now
            batch = {**batch, "pixel_values": self.preprocess_images(self.raw_images(batch))}
        self.select_geometry((tuple(batch["pixel_values"].shape), tuple(batch["token_ids"].shape)))
        if self.use_graph:
            loss = self._graphed_forward_backward(batch, pad_id)
        else:
            st = batch.get("_staged")
            span = batch.get("label_span")
            if st is not None:
                torch.cuda.current_stream().wait_event(st["event"])
                batch = {"pixel_values": st["px"], "token_ids": st["ids"], "labels": st["lab"]}
            px = batch["pixel_values"].to(dev, dtype=F32, non_blocking=True)
            ids = batch["token_ids"].to(dev, non_blocking=True)
            lab = batch["labels"].to(dev, non_blocking=True)
            loss = self.forward_backward(px, ids, lab, pad_id, span)
            if st is not None:
                st["done"].record()
        states = self.trainables()
        if not sync_gradients:
            # flat f32 axpy over the gradient buffers (22M projector entries, + the adapters): bookkeeping between
            # micro-batches
            if self._acc_grad is None:
                self._acc_grad = [torch.zeros_like(st.grad) for st in states]
            for acc, st in zip(self._acc_grad, states):
                if self._acc_pending:
                    acc.add_(st.grad)
                else:
                    acc.copy_(st.grad)
            self._acc_pending = True
            return loss
        if self._acc_pending:
            for acc, st in zip(self._acc_grad, states):
                st.grad.add_(acc)
            self._acc_pending = False
        if allreduce is not None and world_size > 1:
            with _phase("grad_allreduce"):
                for st in states:
                    allreduce(st.grad)
        with _phase("clip_adamw"):
            self.optimizer_step(lr, weight_decay=weight_decay, grad_scale=1.0 / world_size, max_norm=max_norm)
        return loss


def build_engine(vcfg: SiglipCfg, lcfg: GemmaCfg, device, seed: int = 0, grad_accum: int = 1,
                 expansion: int = 10) -> Stage1Engine:
    """Random-init engine for benchmarks and smoke tests (synthetic weights, generated on device)."""
    from . import weights as Wt
    vis = PackedSiglip(Wt.random_siglip(vcfg, device, seed), vcfg, device)
    llm = PackedGemma(Wt.random_gemma(lcfg, device, seed), lcfg, device)
    proj = ProjectorState(Wt.random_projector(vcfg.hidden, lcfg.hidden, device, expansion, seed), device)
    return Stage1Engine(vis, llm, proj, device, grad_accum=grad_accum)
