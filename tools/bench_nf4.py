"""GPU probe: the three NF4 kernels on LLM-sized weights against the HBM roofline.  Algorithmic bytes per value:
quantise 2 (bf16 in) + 0.5 (codes) + 1/16 (fp32 absmax) = 2.5625; de-quantise 0.5 + 1/64 + 2 (bf16 out) = 2.5156;
the nested-statistics kernel moves 5 bytes per 64 values.  Buffers rotate through more than the L2 holds.  The CPU
column is the numpy oracle (test infrastructure) on one core for a bounded sample.  One JSON line per shape."""
import sys, json, time
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200 import ops, _lib
import ctypes as C

dev = torch.device("cuda:0")
peaks = {}
try:
    peaks = json.load(open(Path(__file__).resolve().parent.parent / "MEASURED_PEAKS.json"))
except Exception:
    pass


def timed(fn, reps):
    fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) * 1e3 / reps


for name, shape in (("Qwen3-8B gate_proj", (12288, 4096)), ("Gemma3-1B gate_proj", (6912, 1152)),
                    ("Gemma3-4B down_proj", (2560, 10240))):
    n = shape[0] * shape[1]
    nbuf = max(2, int(400e6 // (2 * n)) + 1)                       # > 126 MB of L2 between reuses
    ws = [(torch.randn(shape, device=dev) * 0.02).to(torch.bfloat16) for _ in range(nbuf)]
    sts = [ops.nf4_quantize(w) for w in ws]
    outs = [torch.empty_like(w) for w in ws]
    packed = [torch.empty_like(s["packed"]) for s in sts]
    absmax = [torch.empty_like(s["absmax"]) for s in sts]
    code = ops.nf4_dynamic_map(dev)
    it = [0]

    def q():
        i = it[0] % nbuf; it[0] += 1
        _lib.call("pt_nf4_quantize", ops._ptr(ws[i]), C.c_longlong(n), 1, ops._ptr(packed[i]), ops._ptr(absmax[i]), ops._stream())

    def a():
        i = it[0] % nbuf; it[0] += 1
        s = sts[i]
        _lib.call("pt_nf4_absmax_quantize", ops._ptr(s["absmax"]), C.c_longlong(n // 64), ops._f(s["offset"]), ops._ptr(code),
                  ops._ptr(s["qabsmax"]), ops._ptr(s["absmax2"]), ops._stream())

    def d():
        i = it[0] % nbuf; it[0] += 1
        ops.nf4_dequantize(sts[i], outs[i])

    us_q, us_a, us_d = timed(q, 5 * nbuf), timed(a, 5 * nbuf), timed(d, 5 * nbuf)
    cpu = None
    if n <= 8e6:
        from oracle import nf4_oracle as N
        w = ws[0].float().cpu().numpy()
        t = time.perf_counter(); N.roundtrip(w); cpu = (time.perf_counter() - t) * 1e3
    print(json.dumps({"weight": name, "shape": shape, "quantize_us": us_q, "quantize_GBps": n * 2.5625 / us_q / 1e3,
                      "absmax_quantize_us": us_a, "dequantize_us": us_d, "dequantize_GBps": n * 2.515625 / us_d / 1e3,
                      "hbm_peak_GBps": peaks.get("hbm_gbs"), "oracle_roundtrip_ms_1core": cpu}), flush=True)
