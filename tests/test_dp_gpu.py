"""Data-parallel correctness on real GPUs (needs >= 2 devices; skipped on a one-GPU box): spawns
`torchrun --nproc-per-node 2 tools/dp_check.py` — all-reduced gradient == mean of the per-rank gradients, parameters
bit-identical on every rank after three optimizer steps on different per-rank batches."""
import json
import os
import subprocess
import sys
from pathlib import Path

import pytest
import torch

pytestmark = pytest.mark.gpu
ROOT = Path(__file__).resolve().parent.parent


def test_allreduce_is_the_mean_and_ranks_stay_identical():
    if not torch.cuda.is_available() or torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs (gpurun --gpus 2)")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2",
                        "--master-addr", "127.0.0.1", "--master-port", "29611", str(ROOT / "tools" / "dp_check.py")],
                       capture_output=True, text=True, timeout=600, env=env, cwd=str(ROOT))
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    line = [ln for ln in r.stdout.splitlines() if ln.startswith("{")][-1]
    out = json.loads(line)
    assert out["ok"] and out["per_rank_grads_differ"] and out["params_identical_after_steps"], out
