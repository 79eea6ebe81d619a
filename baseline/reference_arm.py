"""CPU reference arm: the reference's OWN `ProjectionTrainerStage1.train()` timed on the host cores.

Runs the unmodified `/root/reference/Stage1/{projector_trainer,projectors}.py` — staged by
`__graft_entry__.build()` as verbatim copies under the git-ignored `baseline/_ref/Stage1/` so that they travel to
the GPU box — driving stock HF `SiglipVisionModel` / `Gemma3ForCausalLM` (random init at the full SigLIP-L/16-384
and Gemma3-1B sizes, fp32, CPU, HF's default attention path) exactly the way `train_projection_stage1.py:338-377`
does.  Nothing of this repo's engine, kernels or oracle is on that path.  The only stand-ins are the ones the
reference needs from its environment: a single-process accelerator duck type (accelerate is not in the image; the
14 members the trainer touches) which also takes the per-step timestamps, and an in-memory dataset
with the batch schema of `train_projection_stage1.py:114-118`.

Used by `bench.py --impl reference` and by the `cpu_baseline` leg of the default bench run.
"""
from __future__ import annotations

import logging
import os
import sys
import tempfile
import time
from pathlib import Path

REF_DIR = Path(__file__).resolve().parent / "_ref" / "Stage1"


def available() -> bool:
    return (REF_DIR / "projector_trainer.py").exists() and (REF_DIR / "projectors.py").exists()


class _BudgetSpent(Exception):
    pass


class _TimingAccelerator:
    """accelerate.Accelerator as far as Stage1/projector_trainer.py uses it, one process, CPU."""

    def __init__(self, torch, grad_accum: int, max_steps: int, budget_s: float, min_steps: int):
        self._torch = torch
        self.device = torch.device("cpu")
        self.is_main_process = True
        self.num_processes = 1
        self.process_index = 0
        self.gradient_accumulation_steps = grad_accum
        self.sync_gradients = True
        self.trackers = []
        self.step_end = []            # wall-clock time at the end of every optimizer step
        self.losses = []
        self._max, self._budget, self._min = max_steps, budget_s, min_steps
        self.t_start = None

    def prepare(self, *objs):
        return objs if len(objs) > 1 else objs[0]

    def prepare_model(self, m, device_placement=True):
        return m

    def unwrap_model(self, m):
        return m

    def backward(self, loss):
        loss.backward()

    def clip_grad_norm_(self, params, max_norm):
        return self._torch.nn.utils.clip_grad_norm_(list(params), max_norm)

    def gather(self, t):
        return t.detach().reshape(1)

    def log(self, d, step=None):
        if "train/learning_rate" not in d:
            return
        now = time.time()
        self.step_end.append(now)
        self.losses.append(float(d["train/batch_loss"]))
        if len(self.step_end) >= self._max or (now - self.t_start > self._budget and len(self.step_end) >= self._min):
            raise _BudgetSpent()


def run(steps: int, warmup: int, batch: int, tokens: int, budget_s: float = 240.0) -> dict:
    """`warmup + steps` optimizer steps of `batch` images with `tokens` caption tokens each through the reference
    trainer; returns images/s over the timed steps (the loop is cut short once `budget_s` of wall clock is spent and
    at least one timed step exists)."""
    import torch
    from torch.utils.data import Dataset
    from transformers import Gemma3ForCausalLM, Gemma3TextConfig, SiglipVisionConfig, SiglipVisionModel
    if not available():
        raise RuntimeError(f"{REF_DIR} is not staged (run __graft_entry__.build() where /root/reference exists)")
    sys.path.insert(0, str(REF_DIR))
    import projector_trainer as ref_trainer          # baseline/_ref/Stage1/projector_trainer.py (verbatim copy)
    from projectors import MLPProjector              # baseline/_ref/Stage1/projectors.py
    assert Path(ref_trainer.__file__).resolve().parent == REF_DIR.resolve(), ref_trainer.__file__
    logging.getLogger("projector_trainer").setLevel(logging.WARNING)

    torch.set_num_threads(os.cpu_count() or 1)
    cores = torch.get_num_threads()
    torch.manual_seed(0)
    t0 = time.time()
    vision = SiglipVisionModel(SiglipVisionConfig(hidden_size=1024, intermediate_size=4096, num_hidden_layers=24,
                                                  num_attention_heads=16, image_size=384, patch_size=16)).eval()
    llm = Gemma3ForCausalLM(Gemma3TextConfig(vocab_size=262144, hidden_size=1152, intermediate_size=6912,
                                             num_hidden_layers=26, num_attention_heads=4, num_key_value_heads=1,
                                             head_dim=256, sliding_window=512, query_pre_attn_scalar=256,
                                             max_position_embeddings=32768, pad_token_id=0)).eval()
    proj = MLPProjector(1024, 1152)
    init_s = time.time() - t0

    n = (warmup + steps) * batch
    g = torch.Generator().manual_seed(1234)
    px = torch.rand(n, 3, 384, 384, generator=g) * 2 - 1
    ids = torch.randint(3, 262144, (n, tokens), generator=g)
    ids[:, 0] = 2

    class Items(Dataset):
        def __len__(self):
            return n

        def __getitem__(self, i):
            return {"pixel_values": px[i], "token_ids": ids[i], "labels": ids[i].clone()}

    class Tok:
        pad_token_id, eos_token_id = 0, 1

    acc = _TimingAccelerator(torch, 1, warmup + steps, budget_s, warmup + 1)
    with tempfile.TemporaryDirectory() as out_dir:
        trainer = ref_trainer.ProjectionTrainerStage1(
            acc, vision, llm, proj, processor=None, tokenizer=Tok(), train_dataset=Items(), val_dataset=None,
            output_dir=out_dir, batch_size=batch, learning_rate=1e-4, weight_decay=0.01, num_epochs=1,
            gradient_accumulation_steps=1, warmup_ratio=0.0, save_every_n_epochs=0)
        acc.t_start = time.time()
        try:
            trainer.train()
        except _BudgetSpent:
            pass
    ends = [acc.t_start] + acc.step_end
    durs = [b - a for a, b in zip(ends[:-1], ends[1:])]
    done_w = min(warmup, len(durs) - 1)
    timed = durs[done_w:]
    mean = sum(timed) / len(timed)
    return dict(value=batch / mean, unit="images/s", cores=cores, kind="reference",
                sample=f"{len(timed)} timed optimizer step(s) of {batch} image(s) each through the reference's own "
                       f"ProjectionTrainerStage1.train() (Stage1/projector_trainer.py, verbatim copy under "
                       f"baseline/_ref) on HF SiglipVisionModel + Gemma3ForCausalLM, full-size random init, fp32, "
                       f"T={tokens}, S={575 + tokens}, {cores} torch threads; {done_w} warm-up step(s); model init "
                       f"{init_s:.0f}s not timed",
                steps=len(timed), warmup=done_w, ms_per_step=mean * 1e3, last_loss=acc.losses[-1])


if __name__ == "__main__":
    import json
    print(json.dumps(run(int(sys.argv[1]) if len(sys.argv) > 1 else 2, 1, 2, 128)))
