"""Drop-in for ref:Stage1/projectors.py — same class name, constructor, `.model` Sequential layout and
state-dict keys (`model.0.weight`, `model.0.bias`, `model.2.weight`, `model.2.bias`), fp32 master
parameters.  On a CUDA device `forward` / `backward` run the sm_100a kernels (tcgen05 GEMMs with the
bias + erf-GELU epilogue, fused GELU' and wgrad GEMMs); there is no other compute path: calling it on
CPU tensors raises.  Inside `ProjectionTrainerStage1` the same parameters are driven by the fused step
engine instead (engine.ProjectorState shares their storage)."""
from __future__ import annotations

import torch
import torch.nn as nn

from . import ops

BF16 = torch.bfloat16


class _ProjectorFn(torch.autograd.Function):
    """y = W2 gelu_erf(W1 x + b1) + b2 with bf16 tensor-core GEMMs and fp32 accumulation / output
    (what `accelerate --mixed_precision bf16` gives the reference module); backward as in
    oracle.projector_backward: dW2 = dY^T H, db2 = sum dY, dZ = (dY W2) * gelu'(Z), dW1 = dZ^T x,
    db1 = sum dZ, plus dx = dZ W1 when the input needs a gradient."""

    @staticmethod
    def forward(ctx, x, w1, b1, w2, b2):
        if not x.is_cuda:
            raise RuntimeError("MLPProjector runs on sm_100a CUDA kernels only (no CPU path)")
        shp = x.shape
        x2 = x.reshape(-1, shp[-1]).to(BF16).contiguous()
        w1b, w2b = w1.detach().to(BF16).contiguous(), w2.detach().to(BF16).contiguous()
        R, mid, h = x2.shape[0], w1.shape[0], w2.shape[0]
        Z = torch.empty(R, mid, dtype=BF16, device=x.device)
        H = torch.empty(R, mid, dtype=BF16, device=x.device)
        ops.linear(x2, w1b, bias=b1.detach().float().contiguous(), act=ops.ACT_GELU_ERF, out=H, c_pre=Z)
        y = torch.empty(R, h, dtype=torch.float32, device=x.device)
        ops.linear(H, w2b, bias=b2.detach().float().contiguous(), out=y)
        ctx.save_for_backward(x2, Z, H, w1b, w2b)
        ctx.x_shape, ctx.x_dtype = shp, x.dtype
        return y.reshape(*shp[:-1], h)

    @staticmethod
    def backward(ctx, dy):
        x2, Z, H, w1b, w2b = ctx.saved_tensors
        dyb = dy.reshape(-1, dy.shape[-1]).to(BF16).contiguous()
        dW2 = ops.linear_wgrad(dyb, H)
        db2 = torch.empty(dyb.shape[1], dtype=torch.float32, device=dy.device)
        ops.colsum(dyb, db2)
        dH = ops.linear_dgrad(dyb, w2b)
        ops.gelu_erf_bwd(dH, Z, dH)
        dW1 = ops.linear_wgrad(dH, x2)
        db1 = torch.empty(dH.shape[1], dtype=torch.float32, device=dy.device)
        ops.colsum(dH, db1)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = ops.linear_dgrad(dH, w1b, out_dtype=torch.float32).reshape(ctx.x_shape).to(ctx.x_dtype)
        return dx, dW1, db1, dW2, db2


class MLPProjector(nn.Module):
    """2-layer MLP projecting every vision patch embedding into the LLM embedding space
    (ref:Stage1/projectors.py:4-29): Linear(vision_dim, vision_dim*expansion) -> GELU(erf) ->
    Linear(vision_dim*expansion, llm_dim)."""

    def __init__(self, vision_dim: int, llm_dim: int, expansion_factor: int = 10):
        super().__init__()
        intermediate_dim = vision_dim * expansion_factor
        self.model = nn.Sequential(
            nn.Linear(vision_dim, intermediate_dim),
            nn.GELU(),
            nn.Linear(intermediate_dim, llm_dim),
        )

    def forward(self, x: torch.Tensor) -> torch.Tensor:
        """x: [batch, num_patches, vision_dim] -> [batch, num_patches, llm_dim] (fp32)."""
        l1, l2 = self.model[0], self.model[2]
        return _ProjectorFn.apply(x, l1.weight, l1.bias, l2.weight, l2.bias)
