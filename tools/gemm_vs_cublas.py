"""GPU probe: every GEMM shape of the Stage-1 step through this repository's
CTA-pair tcgen05 kernel — once with the epilogue the engine actually uses and once as a plain bf16 GEMM — against
cuBLAS (torch.matmul, plain bf16 output) on the same box, same process, alternating, operands rotating through more
than the 126 MB L2.  TFLOP/s = 2 M N K / mean launch time (CUDA events, 20 launches after 3 warm-ups)."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
torch.manual_seed(0)
BF, F32 = torch.bfloat16, torch.float32
ROT = 4


def rnd(*s, scale=1.0, dtype=BF):
    return (torch.randn(*s, device=dev) * scale).to(dtype)


def timeit(fns, iters=20, warm=3):
    """fns: ROT variants of the same launch on different buffers, cycled."""
    for i in range(warm):
        fns[i % len(fns)]()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(iters):
        fns[i % len(fns)]()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters * 1e3


def case(name, M, N, K, *, dgrad=False, wgrad=False, epi=None, count=1):
    """epi: callable (x, w, out) -> launch with the engine's epilogue, or None."""
    fl = 2.0 * M * N * K
    sets = []
    for _ in range(ROT):
        if wgrad:       # dW[N,K] = dy[M2,N]^T x[M2,K]: here (M, N, K) = (out rows, out cols, contraction)
            dy, x = rnd(K, M), rnd(K, N)
            sets.append((dy, x, torch.empty(M, N, device=dev, dtype=F32)))
        elif dgrad:     # dx[M,N] = dy[M,K] w[K,N]
            sets.append((rnd(M, K), rnd(K, N, scale=0.05), torch.empty(M, N, device=dev, dtype=BF)))
        else:
            sets.append((rnd(M, K), rnd(N, K, scale=0.05), torch.empty(M, N, device=dev, dtype=BF)))
    if wgrad:
        plain = [lambda s=s: ops.linear_wgrad(s[0], s[1], out=s[2]) for s in sets]
        cub = [lambda s=s: torch.matmul(s[0].t(), s[1], out=s[2].view(M, N)) if False else torch.matmul(s[0].t(), s[1]) for s in sets]
    elif dgrad:
        plain = [lambda s=s: ops.linear_dgrad(s[0], s[1], out=s[2]) for s in sets]
        cub = [lambda s=s: torch.matmul(s[0], s[1], out=s[2]) for s in sets]
    else:
        plain = [lambda s=s: ops.linear(s[0], s[1], out=s[2]) for s in sets]
        cub = [lambda s=s: torch.matmul(s[0], s[1].t(), out=s[2]) for s in sets]
    t_c = timeit(cub)
    t_p = timeit(plain)
    t_e = None
    if epi is not None:
        extra = [epi(*s) for s in sets]
        t_e = timeit(extra)
    t_c2 = timeit(cub)
    t_c = min(t_c, t_c2)
    tf = lambda t: fl / t / 1e6
    e = f"{t_e:8.1f} {tf(t_e):7.0f} {tf(t_e) / tf(t_c):5.2f}" if t_e else " " * 22
    print(f"{name:34s} {M:6d} {N:6d} {K:6d} x{count:<3d} | ours+epilogue {e} | ours plain {t_p:8.1f} {tf(t_p):7.0f} "
          f"{tf(t_p) / tf(t_c):5.2f} | cuBLAS {t_c:8.1f} {tf(t_c):7.0f}", flush=True)
    del sets
    torch.cuda.empty_cache()


print("shape (M N K x launches per step)                                | us TF/s ratio-to-cuBLAS", flush=True)
M1, M2 = 18432, 22496
bias = lambda n: torch.randn(n, device=dev)
# ---- SigLIP (bias everywhere; f32 residual stream in and out on out-proj / fc2; tanh-GELU on fc1)
case("siglip patch-embed +bias+pos", M1, 1024, 768, count=1,
     epi=lambda x, w, o: (lambda r=torch.randn(M1, 1024, device=dev), b=bias(1024), of=torch.empty(M1, 1024, device=dev):
                          lambda: ops.linear(x, w, bias=b, residual=r, out=of))())
case("siglip qkv +bias", M1, 3072, 1024, count=24,
     epi=lambda x, w, o: (lambda b=bias(3072): lambda: ops.linear(x, w, bias=b, out=o))())
case("siglip out-proj +bias+res(f32)", M1, 1024, 1024, count=24,
     epi=lambda x, w, o: (lambda r=torch.randn(M1, 1024, device=dev), b=bias(1024):
                          lambda: ops.linear(x, w, bias=b, residual=r, out=r))())
case("siglip fc1 +bias+gelu_tanh", M1, 4096, 1024, count=24,
     epi=lambda x, w, o: (lambda b=bias(4096): lambda: ops.linear(x, w, bias=b, act=ops.ACT_GELU_TANH, out=o))())
case("siglip fc2 +bias+res(f32)", M1, 1024, 4096, count=24,
     epi=lambda x, w, o: (lambda r=torch.randn(M1, 1024, device=dev), b=bias(1024):
                          lambda: ops.linear(x, w, bias=b, residual=r, out=r))())
# ---- projector
case("projector L1 +bias+gelu_erf+pre", 18400, 10240, 1024, count=1,
     epi=lambda x, w, o: (lambda b=bias(10240), z=torch.empty(18400, 10240, device=dev, dtype=BF):
                          lambda: ops.linear(x, w, bias=b, act=ops.ACT_GELU_ERF, out=o, c_pre=z))())
case("projector L2 (f32 out)", 18400, 1152, 10240, count=1)
case("projector L2 dgrad", 18400, 10240, 1152, dgrad=True, count=1)
case("projector L2 wgrad", 1152, 10240, 18400, wgrad=True, count=1)
case("projector L1 wgrad", 10240, 1024, 18400, wgrad=True, count=1)
# ---- Gemma3-1B
case("gemma qkv", M2, 1536, 1152, count=26)
case("gemma qkv dgrad", M2, 1152, 1536, dgrad=True, count=26)
case("gemma o-proj", M2, 1152, 1024, count=26)
case("gemma o-proj dgrad", M2, 1024, 1152, dgrad=True, count=26)
case("gemma gate||up +GeGLU", M2, 13824, 1152, count=26,
     epi=lambda x, w, o: (lambda a=torch.empty(M2, 6912, device=dev, dtype=BF): lambda: ops.linear_geglu(x, w, o, a))())
case("gemma gate||up dgrad", M2, 1152, 13824, dgrad=True, count=26)
case("gemma down", M2, 1152, 6912, count=26)
case("gemma down dgrad +GeGLU bwd", M2, 6912, 1152, dgrad=True, count=26,
     epi=lambda dy, w, o: (lambda gu=rnd(M2, 13824), dgu=torch.empty(M2, 13824, device=dev, dtype=BF):
                           lambda: ops.linear_dgrad_geglu_bwd(dy, w, gu, dgu))())
case("lm head", 4096, 262144, 1152, count=1)
