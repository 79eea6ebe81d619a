"""GPU probe: time the fused Gemma3 attention forward and backward at config 3 (B=32, S=703, hq=4, hkv=1, hd=256)
and config 4 (B=16, S=831, 8:4) geometry.  Prints microseconds per launch (CUDA events, 20 launches)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops
dev = torch.device("cuda:0")


def timed(fn, n=20):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(n):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / n * 1e3


for (B, S, hq, hkv, hd, windows) in ((32, 703, 4, 1, 256, (0, 512)), (16, 831, 8, 4, 256, (0, 1024))):
    G, QKV, ld = hq // hkv, (hq + 2 * hkv) * hd, (S + 7) // 8 * 8
    q = torch.randn(B * S, hq * hd, device=dev).to(torch.bfloat16)
    k = torch.randn(B * S, hkv * hd, device=dev).to(torch.bfloat16)
    qkv = torch.randn(B * S, QKV, device=dev).to(torch.bfloat16)
    dO = torch.randn(B * S, hq * hd, device=dev).to(torch.bfloat16)
    kv = torch.ones(B, S, dtype=torch.uint8, device=dev)
    bits = torch.empty(B * ((S + 63) // 64), dtype=torch.int64, device=dev)
    ops.pack_keyvalid(kv, bits)
    lse = torch.empty(B * hq, S, dtype=torch.float32, device=dev)
    delta = torch.empty(B * hq * S, dtype=torch.float32, device=dev)
    O = torch.empty(B * S, hq * hd, dtype=torch.bfloat16, device=dev)
    dS = torch.zeros(B * hq, S, ld, dtype=torch.bfloat16, device=dev)
    dK = torch.empty(B * S, hkv * hd, dtype=torch.bfloat16, device=dev)
    dV = torch.empty(B * S, hkv * hd, dtype=torch.bfloat16, device=dev)
    dQ = torch.empty(B * S, hq * hd, dtype=torch.bfloat16, device=dev)
    for W in windows:
        pairs = sum(min(s + 1, W) if W else s + 1 for s in range(S))
        gf = 4.0 * B * hq * hd * pairs / 1e9
        t_f = timed(lambda: ops.attn_gemma_fwd(q, k, qkv, bits, lse, O, B, S, hq, hkv, hd, W, 1 / 16))
        t_d = timed(lambda: ops.attn_delta(dO, O, delta, S, hq, hd))
        t_b = timed(lambda: ops.attn_gemma_bwd(q, k, qkv, dO, bits, lse, delta, dS, dK, dV, B, S, hq, hkv, hd, W,
                                               1 / 16, ld))
        t_p = timed(lambda: ops.attn_gemma_bwd_pair(q, k, qkv, dO, bits, lse, delta, dS, dK, dV, B, S, hq, hkv, hd, W,
                                                    1 / 16, ld))
        t_q2 = timed(lambda: ops.attn_dq_from_dst(dS, k, dQ, B, S, hq, hkv, hd, W, ld))
        t_q = timed(lambda: ops.gemm_raw(M=S, N=hd, K=S, A=dS, lda=ld, B=k, ldb=hkv * hd, b_mn=True, Cout=dQ,
                                         ldc=hq * hd, nz=(G, hkv, B), a_zs=(S * ld, G * S * ld, hq * S * ld),
                                         b_zs=(0, hd, S * hkv * hd), c_zs=(hd, G * hd, S * hq * hd),
                                         band=ops.BAND_K_BY_QUERY, band_window=W))
        print(f"B={B} S={S} {hq}:{hkv} hd={hd} window={W}: fwd {t_f:.1f} us ({gf / t_f * 1e3:.0f} TF/s)  "
              f"delta {t_d:.1f}  bwd(dK,dV,dS) {t_b:.1f} us ({2 * gf / t_b * 1e3:.0f} TF/s of 4 products)  "
              f"dQ gemm {t_q:.1f} us | pair bwd {t_p:.1f} us ({2 * gf / t_p * 1e3:.0f} TF/s)  dQ gemm (dS^T, MN-major A) {t_q2:.1f} us",
              flush=True)
