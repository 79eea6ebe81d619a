"""GPU probe: the HBM-bound kernels of the Stage-1 step, each launched alone at its config-3 shape
(LLM rows M = 32*703 = 22 496, D = 1152; vision rows 18 432, D = 1024; CE over 4096 x 262 144 bf16 logits).
Prints CUDA-event time and algorithmic GB/s per kernel; run under
`ncu --set full --clock-control none -k regex:<name> -s 2 -c 1` for the DRAM-traffic capture
(inputs are far larger than the 126 MB L2, so every launch streams from HBM)."""
import sys, argparse
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops

ap = argparse.ArgumentParser(); ap.add_argument("--reps", type=int, default=3)
a = ap.parse_args()
dev = torch.device("cuda:0")
BF, F32 = torch.bfloat16, torch.float32
M, D, Mv, Dv, R, V = 32 * 703, 1152, 32 * 576, 1024, 4096, 262144
S, hq, hkv, hd = 703, 4, 1, 256
g = torch.Generator(device=dev).manual_seed(7)
rn = lambda *s, dt=BF: torch.randn(*s, generator=g, device=dev, dtype=F32).to(dt)
eps = 1e-6


def timed(name, nbytes, fn):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(a.reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); fn(); e1.record(); torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1) * 1e3)
    us = min(ts)
    print(f"{name:28s} {nbytes / 1e6:9.1f} MB  {us:8.1f} us  {nbytes / us / 1e3:7.0f} GB/s", flush=True)


# rmsnorm_bwd_pair: g1 bf16, u1 f32 (stream), dx f32 in/out, u2 bf16, out2 bf16 -> 18 D M bytes
g1, u1, dx, u2, out2 = rn(M, D), rn(M, D, dt=F32), rn(M, D, dt=F32), rn(M, D), torch.empty(M, D, device=dev, dtype=BF)
w1, w2 = rn(D, dt=F32), rn(D, dt=F32)
timed("rmsnorm_bwd_pair", 18 * D * M, lambda: ops.rmsnorm_bwd_pair(g1, u1, w1, dx, u2, w2, out2, eps))
# rmsnorm_residual_fwd: y bf16, x_in f32 -> x_out f32, nxt bf16 -> 12 D M bytes
y, x_in, x_out, nxt = rn(M, D), rn(M, D, dt=F32), torch.empty(M, D, device=dev, dtype=F32), torch.empty(M, D, device=dev, dtype=BF)
timed("rmsnorm_residual_fwd", 12 * D * M, lambda: ops.rmsnorm_residual_fwd(y, w1, x_in, x_out, w2, nxt, eps))
# layernorm_fwd (SigLIP): f32 in, bf16 out -> 6 D M bytes
xv, ov = rn(Mv, Dv, dt=F32), torch.empty(Mv, Dv, device=dev, dtype=BF)
gam, bet = rn(Dv, dt=F32), rn(Dv, dt=F32)
timed("layernorm_fwd", 6 * Dv * Mv, lambda: ops.layernorm_fwd(xv, gam, bet, ov, eps))
# qk-norm + RoPE
QKV = (hq + 2 * hkv) * hd
qkv, q_out, k_out = rn(M, QKV), torch.empty(M, hq * hd, device=dev, dtype=BF), torch.empty(M, hkv * hd, device=dev, dtype=BF)
wq, wk = rn(hd, dt=F32), rn(hd, dt=F32)
ang = torch.rand(S, hd // 2, generator=g, device=dev) * 6.28
cos_t, sin_t = ang.cos().contiguous(), ang.sin().contiguous()
timed("qknorm_rope_fwd", 5120 * M, lambda: ops.qknorm_rope_fwd(qkv, wq, wk, cos_t, sin_t, q_out, k_out, S, hq, hkv, hd, eps))
dq, dk, d_qkv = rn(M, hq * hd), rn(M, hkv * hd), torch.empty(M, QKV, device=dev, dtype=BF)
timed("qknorm_rope_bwd", 7680 * M, lambda: ops.qknorm_rope_bwd(dq, dk, qkv, wq, wk, cos_t, sin_t, d_qkv, S, hq, hkv, hd, eps))
# attn_delta: dO, O bf16 [M, hq*hd] -> delta f32 [B*hq*S]
dO, O_, delta = rn(M, hq * hd), rn(M, hq * hd), torch.empty(32 * hq * S, device=dev, dtype=F32)
timed("attn_delta", 2 * 2 * M * hq * hd + 4 * M * hq, lambda: ops.attn_delta(dO, O_, delta, S, hq, hd))
# projector GELU backward: dH, Z bf16 in, dZ bf16 out over 18 400 x 10 240
n = 18400 * 10240
dH, Z, dZ = rn(n), rn(n), torch.empty(n, device=dev, dtype=BF)
timed("gelu_erf_bwd", 6 * n, lambda: ops.gelu_erf_bwd(dH, Z, dZ))
del dH, Z, dZ
# clip + AdamW over the 22.3 M projector parameters: p g m v in, p m v + bf16 shadow out
npar = 1024 * 10240 + 10240 + 10240 * 1152 + 1152
p, gr, m_, v_ = rn(npar, dt=F32), rn(npar, dt=F32), torch.zeros(npar, device=dev), torch.zeros(npar, device=dev)
sh, ss = torch.empty(npar, device=dev, dtype=BF), torch.ones(1, device=dev)
timed("clip_adamw", 30 * npar, lambda: ops.clip_adamw(p, gr, m_, v_, sh, ss, grad_scale=1.0, max_norm=5.0, lr=1e-4,
                                                        beta1=0.9, beta2=0.999, eps=1e-8, wd=0.01, step=1))
# masked CE over the vocabulary: statistics pass (read 2 B/logit) + in-place gradient pass (read + write)
logits = (torch.randn(R, V, generator=g, device=dev, dtype=F32) * 2).to(BF)
tg = torch.randint(0, V, (R,), generator=g, device=dev).to(torch.int32)
lse, acc2, loss = torch.empty(R, device=dev), torch.zeros(2, device=dev), torch.zeros(1, device=dev)
timed("ce_fwd_bwd (stats+grad)", 6 * R * V, lambda: ops.ce_fwd_bwd(logits, tg, lse, acc2, loss, 1.0 / R))
