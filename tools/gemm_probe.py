"""GPU probe: correctness diagnostics + timing of the tcgen05 GEMM vs cuBLAS (torch.matmul)."""
import sys, time, json
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
torch.manual_seed(0)


def rnd(*s, scale=1.0):
    return (torch.randn(*s, device=dev) * scale).to(torch.bfloat16)


def diag(name, out, ref):
    out = out.float()
    d = (out - ref).abs()
    print(f"[{name}] max_err={d.max().item():.4e} ref_max={ref.abs().max().item():.4e} "
          f"bad_frac={(d > 1e-2 * ref.abs().max()).float().mean().item():.4f}", flush=True)
    if d.max() > 2e-2 * ref.abs().max():
        bad = (d > 1e-2 * ref.abs().max())
        rows = bad.any(1).nonzero().flatten()[:8].tolist()
        cols = bad.any(0).nonzero().flatten()[:8].tolist()
        print("   first bad rows", rows, "cols", cols)
        print("   out[0,:8]", out[0, :8].tolist())
        print("   ref[0,:8]", ref[0, :8].tolist())


def correctness():
    M, N, K = 256, 256, 256
    x, w = rnd(M, K), rnd(N, K)
    diag("tn", ops.linear(x, w, out_dtype=torch.float32), x.float() @ w.float().t())
    wt = w.t().contiguous()   # [K, N]
    out = torch.empty(M, N, device=dev)
    ops.gemm_raw(M=M, N=N, K=K, A=x, lda=K, B=wt, ldb=N, b_mn=True, Cout=out, ldc=N)
    diag("b_mn", out, x.float() @ w.float().t())
    xt = x.t().contiguous()   # [K, M]
    ops.gemm_raw(M=M, N=N, K=K, A=xt, lda=M, a_mn=True, B=w, ldb=K, Cout=out, ldc=N)
    diag("a_mn", out, x.float() @ w.float().t())
    ops.gemm_raw(M=M, N=N, K=K, A=xt, lda=M, a_mn=True, B=wt, ldb=N, b_mn=True, Cout=out, ldc=N)
    diag("ab_mn", out, x.float() @ w.float().t())


def timeit(fn, iters=20, warm=3):
    for _ in range(warm):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def timing():
    shapes = [("siglip_qkv", 18432, 3072, 1024), ("siglip_fc1", 18432, 4096, 1024),
              ("siglip_fc2", 18432, 1024, 4096), ("proj_l1", 18400, 10240, 1024),
              ("proj_l2", 18400, 1152, 10240), ("gemma_gateup", 22496, 13824, 1152),
              ("gemma_down", 22496, 1152, 6912), ("gemma_qkv", 22496, 1536, 1152),
              ("gemma_o", 22496, 1152, 1024), ("lm_head", 4096, 262144, 1152),
              ("square8k", 8192, 8192, 8192)]
    res = []
    for name, M, N, K in shapes:
        x, w = rnd(M, K), rnd(N, K, scale=0.05)
        out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
        t_mine = timeit(lambda: ops.linear(x, w, out=out))
        t_cublas = timeit(lambda: torch.matmul(x, w.t(), out=out))
        fl = 2.0 * M * N * K
        r = dict(name=name, M=M, N=N, K=K, ms=t_mine, tflops=fl / t_mine / 1e9,
                 cublas_ms=t_cublas, cublas_tflops=fl / t_cublas / 1e9)
        print(json.dumps(r), flush=True)
        res.append(r)
        del x, w, out
    # dgrad (B MN-major) and wgrad (A,B MN-major) flavours
    M, N, K = 22496, 1152, 13824     # dx = dgu @ Wgu
    dy, w = rnd(M, K), rnd(K, N, scale=0.05)
    out = torch.empty(M, N, device=dev, dtype=torch.bfloat16)
    t = timeit(lambda: ops.gemm_raw(M=M, N=N, K=K, A=dy, lda=K, B=w, ldb=N, b_mn=True, Cout=out, ldc=N))
    print(json.dumps(dict(name="gemma_gateup_dgrad_bmn", ms=t, tflops=2.0 * M * N * K / t / 1e9)), flush=True)
    T, N, K = 18400, 1152, 10240
    dy, x = rnd(T, N), rnd(T, K)
    out = torch.empty(N, K, device=dev)
    t = timeit(lambda: ops.linear_wgrad(dy, x, out=out))
    print(json.dumps(dict(name="proj_wgrad2_abmn", ms=t, tflops=2.0 * T * N * K / t / 1e9)), flush=True)
    Path("gpurun_out").mkdir(exist_ok=True)
    Path("gpurun_out/gemm_timing.json").write_text(json.dumps(res, indent=1))


if __name__ == "__main__":
    correctness()
    if "--time" in sys.argv:
        timing()
