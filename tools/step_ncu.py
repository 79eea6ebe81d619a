"""ncu target: ONE Stage-1 training step (BASELINE.json configs[2]: B = 32, T = 128, SigLIP-L + Gemma3-1B) between
cudaProfilerStart/Stop, eager launches (PT_STAGE1_GRAPH=0) so that every kernel is its own row.  Run as
  PT_STAGE1_GRAPH=0 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv \\
      --log-file X python tools/step_ncu.py
and summarise with `python tools/step_ncu.py --summarise X`."""
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))

if "--summarise" in sys.argv:
    import collections, csv, re
    with open(sys.argv[sys.argv.index("--summarise") + 1]) as f:
        rows = list(csv.DictReader(l for l in f if not l.startswith("==")))
    agg, tot, n = collections.OrderedDict(), 0.0, 0
    for r in rows:
        if r.get("Metric Name") != "gpu__time_duration.sum":
            continue
        v = float(r["Metric Value"].replace(",", "")) * {"ns": 1e-3, "us": 1.0, "ms": 1e3}[r["Metric Unit"]]
        k = re.sub(r"^void ", "", re.sub(r"\(.*", "", r["Kernel Name"]))[:80]
        a = agg.setdefault(k, [0, 0.0]); a[0] += 1; a[1] += v; tot += v; n += 1
    print(f"launches {n}  total {tot / 1e3:.2f} ms")
    for k, (c, t) in sorted(agg.items(), key=lambda q: -q[1][1]):
        print(f"{t / 1e3:9.3f} ms {100 * t / tot:5.1f}%  n={c:5d}  avg {t / c:8.1f} us  {k}")
    sys.exit(0)

import torch
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine

dev = torch.device("cuda:0")
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev)
g = torch.Generator(device=dev).manual_seed(1)
B, T = 32, 128
ids = torch.randint(3, GEMMA3_1B.vocab, (B, T), generator=g, device=dev)
batch = {"pixel_values": torch.rand(B, 3, 384, 384, generator=g, device=dev) * 2 - 1, "token_ids": ids, "labels": ids.clone()}
for _ in range(2):
    eng.train_step(batch, 1e-4, max_norm=1.0)
torch.cuda.synchronize()
torch.cuda.profiler.start()
loss = eng.train_step(batch, 1e-4, max_norm=1.0)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print("loss", float(loss))
