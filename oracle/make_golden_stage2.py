#!/usr/bin/env python
"""Generates tests/golden/vqa_*.npz by running the UNMODIFIED reference Stage-2 trainer here.

TEST INFRASTRUCTURE ONLY (see oracle/stage1_oracle.py header).  Needs /root/reference and HF
transformers; only the vectors it writes travel to the GPU box.

What runs: the reference's own `VQATrainerStage2.train()` and `vqa_collate_fn`
(/root/reference/Stage2/trainer.py:18, 63, 268) with `freeze_llm=True, freeze_projection_layer=False,
freeze_vision_encoder=True, enable_qlora=False` — the frozen-LLM variant SURVEY.md §8(f) rank 2 names — and, for the
vqa_lora_* cases, with `enable_qlora=True` on an LLM whose projection Linears are wrapped by a hand-rolled LoRA module
(`LoRALinear` below = peft's forward with dropout 0; peft itself is not in this image) —
on HF `SiglipVisionModel` / `Gemma3ForCausalLM` built from explicit configs on CPU in fp32, with the
reference's `MLPProjector` (Stage1/projectors.py).  Stand-ins, because `accelerate` is not installed in
this image:

  * module `accelerate.utils` with `gather_object` = identity (one process);
  * a single-process accelerator duck type.  Two behaviours of the real library are emulated FROM
    KNOWLEDGE of accelerate's source (pinned accelerate==1.6.0, requirements.txt; not in the container,
    so flagged "verify" in DESIGN.md): `accumulate()` sets `sync_gradients` on every GA-th micro-batch
    and on the loader's last batch (`GradientState.end_of_dataloader`), and `backward()` divides the
    loss by `gradient_accumulation_steps` once more (the trainer has already divided it).  The divisor
    applied is recorded in the file (`backward_divisor`).

Recorded: per micro-batch the sample indices the shuffling loader drew, the attention mask handed to
the language model, the logged loss and the sync flag; per optimizer step the projector gradients as
clip_grad_norm_(1.0) sees them (accumulated), the learning rate and the parameters after the step;
from the first micro-batch the spliced `inputs_embeds`; per epoch the validation loss
(`val/loss`, evaluation collates with left padding: trainer.py:500-501).

    python oracle/make_golden_stage2.py
"""
from __future__ import annotations

import contextlib
import logging
import sys
import time
import types
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
sys.path.insert(0, "/root/reference")
sys.path.insert(0, "/root/reference/Stage1")

from oracle import stage1_oracle as O  # noqa: E402
from oracle import stage2_oracle as O2  # noqa: E402
from oracle.make_golden import build_hf_gemma, build_hf_siglip  # noqa: E402

GOLDEN = ROOT / "tests" / "golden"

CASES = O2.GOLDEN_CASES


def install_accelerate_stub():
    if "accelerate" in sys.modules:
        return
    acc = types.ModuleType("accelerate")
    utils = types.ModuleType("accelerate.utils")
    utils.gather_object = lambda obj: obj
    acc.utils = utils
    sys.modules["accelerate"], sys.modules["accelerate.utils"] = acc, utils


class Recorder:
    def __init__(self):
        self.micro, self.steps, self.val, self.first = [], [], [], {}


class _PreparedLoader:
    """accelerate's DataLoaderShard, one process: flags the last batch (end_of_dataloader)."""

    def __init__(self, loader, acc):
        self.loader, self.acc, self.dataset = loader, acc, loader.dataset

    def __len__(self):
        return len(self.loader)

    def __iter__(self):
        self.acc._end = False
        it = iter(self.loader)
        cur = next(it, None)
        while cur is not None:
            nxt = next(it, None)
            self.acc._end = nxt is None
            yield cur
            cur = nxt
        self.acc._end = False


class SingleProcessAccelerator:
    def __init__(self, grad_accum: int, rec: Recorder, projector_ref, llm_ref=None):
        self._llm = llm_ref
        self.device = torch.device("cpu")
        self.is_main_process = True
        self.num_processes, self.process_index = 1, 0
        self.gradient_accumulation_steps = grad_accum
        self.sync_gradients = True
        self.trackers = []
        self._rec, self._proj = rec, projector_ref
        self._end, self._step, self._in_train, self._cur = False, 0, False, None

    def prepare(self, *objs):
        out = tuple(_PreparedLoader(o, self) if isinstance(o, torch.utils.data.DataLoader) else o for o in objs)
        return out if len(out) > 1 else out[0]

    def prepare_model(self, m, device_placement=True):
        return m

    def unwrap_model(self, m):
        return m

    @contextlib.contextmanager
    def accumulate(self, *models):
        if self._end:
            self._step = 0
            self.sync_gradients = True
        else:
            self._step += 1
            self.sync_gradients = self._step % self.gradient_accumulation_steps == 0
        self._in_train = True
        self._rec.micro.append({"sync": self.sync_gradients})
        try:
            yield
        finally:
            self._in_train = False

    def backward(self, loss):
        (loss / self.gradient_accumulation_steps).backward()

    def clip_grad_norm_(self, params, max_norm):
        assert max_norm == 1.0
        if self._cur is None:       # first module of this optimizer step: nothing has been clipped yet
            self._cur = {"grads": {n: p.grad.detach().clone() for n, p in self._proj.named_parameters()
                                   if p.grad is not None}}
            if self._llm is not None:
                self._cur["lora_grads"] = {n: p.grad.detach().clone() for n, p in self._llm.named_parameters()
                                           if p.requires_grad}
        return torch.nn.utils.clip_grad_norm_(list(params), max_norm)

    def gather(self, t):
        t = t.detach().reshape(1)
        if self._rec.micro and "loss" not in self._rec.micro[-1]:
            self._rec.micro[-1]["loss"] = float(t[0])
        return t

    def log(self, d, step=None):
        if "train/step_loss" in d and self._cur is not None:
            self._cur["lr_after"] = float(d["train/learning_rate"])
            self._cur["params"] = {n: p.detach().clone() for n, p in self._proj.named_parameters()}
            if self._llm is not None:
                self._cur["lora_params"] = {n: p.detach().clone() for n, p in self._llm.named_parameters()
                                            if p.requires_grad}
            self._rec.steps.append(self._cur)
            self._cur = None
        if "val/loss" in d:
            self._rec.val.append(float(d["val/loss"]))

    def wait_for_everyone(self):
        pass

    def save_state(self, path):
        pass

    def save_model(self, model, path):
        pass


class VQAList(torch.utils.data.Dataset):
    def __init__(self, pixels, qs, as_):
        self.items = [{"pixel_values": pixels[i], "question_input_ids": qs[i], "answer_input_ids": as_[i]}
                      for i in range(len(qs))]

    def __len__(self):
        return len(self.items)

    def __getitem__(self, i):
        return self.items[i]


class Tok:
    def __init__(self, pad, side):
        self.pad_token_id, self.eos_token_id, self.padding_side = pad, 1, side

    def batch_decode(self, ids, skip_special_tokens=True):
        return [" ".join(str(int(t)) for t in row) for row in ids]

    def save_pretrained(self, path):
        pass


class LoRALinear(torch.nn.Module):
    """peft.tuners.lora.layer.Linear, forward only, dropout p = 0 (peft is not in this image):
    result = base_layer(x) + lora_B(lora_A(x)) * scaling."""

    def __init__(self, base: torch.nn.Linear, r: int, alpha: float):
        super().__init__()
        self.base_layer = base
        self.lora_A = torch.nn.Linear(base.in_features, r, bias=False)
        self.lora_B = torch.nn.Linear(r, base.out_features, bias=False)
        self.scaling = alpha / r

    def forward(self, x):
        return self.base_layer(x) + self.lora_B(self.lora_A(x)) * self.scaling


def inject_lora(llm, lora: dict, r: int, alpha: float, targets):
    """What `get_peft_model(llm, LoraConfig(...))` does to the module tree (ref:Stage2/train_vqa_stage2.py:225-246):
    every target Linear of every decoder layer is wrapped, the base model is frozen, the adapters train."""
    llm.requires_grad_(False)
    for i, layer in enumerate(llm.model.layers):
        for t in targets:
            parent = layer.self_attn if t in ("q_proj", "k_proj", "v_proj", "o_proj") else layer.mlp
            w = LoRALinear(getattr(parent, t), r, alpha)
            with torch.no_grad():
                w.lora_A.weight.copy_(lora[f"l{i}.{t}.A"])
                w.lora_B.weight.copy_(lora[f"l{i}.{t}.B"])
            setattr(parent, t, w)
    return llm


def _lora_key(name: str) -> str:
    """HF parameter name `model.layers.3.self_attn.q_proj.lora_A.weight` -> oracle key `l3.q_proj.A`."""
    parts = name.split(".")
    return f"l{parts[2]}.{parts[4]}.{parts[5][-1]}"


def run_case(name: str, out_dir: Path):
    lora_cfg = None
    if name in O2.LORA_CASES:
        *base, lora_cfg = O2.LORA_CASES[name]
        vkw, lkw, n_train, n_val, B, side, accum, epochs, warm = base
    else:
        vkw, lkw, n_train, n_val, B, side, accum, epochs, warm = CASES[name]
    import transformers                                  # resolves its own optional-dependency checks first
    from transformers import get_cosine_schedule_with_warmup  # noqa: F401
    install_accelerate_stub()
    from Stage2 import trainer as ref_trainer            # /root/reference/Stage2/trainer.py
    from projectors import MLPProjector                  # /root/reference/Stage1/projectors.py
    logging.getLogger("Stage2.trainer").setLevel(logging.ERROR)

    vc, lc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(vc), O.init_gemma(lc), O.init_projector(vc.hidden, lc.hidden)
    vision, llm = build_hf_siglip(vc, Wv), build_hf_gemma(lc, Wl)
    proj = MLPProjector(vc.hidden, lc.hidden)
    proj.load_state_dict(P)
    px_t, q_t, a_t = O2.synthetic_vqa_samples(n_train, vc, lc, seed=4321)
    px_v, q_v, a_v = O2.synthetic_vqa_samples(n_val, vc, lc, seed=9876)

    train_proj = True
    if lora_cfg is not None:
        r, alpha, targets, b_std, train_proj = lora_cfg
        L0 = O2.init_lora(lc, r, targets, b_std=b_std)
        inject_lora(llm, L0, r, alpha, targets)

    rec = Recorder()
    acc = SingleProcessAccelerator(accum, rec, proj, llm if lora_cfg is not None else None)

    def vis_pre(mod, args, kwargs):
        if acc._in_train:
            pv = kwargs["pixel_values"]
            rec.micro[-1]["idx"] = [next(j for j in range(n_train) if torch.equal(px_t[j], row)) for row in pv]
    def llm_pre(mod, args, kwargs):
        if acc._in_train:
            rec.micro[-1]["attention_mask"] = kwargs["attention_mask"].clone()
            if "spliced" not in rec.first:
                rec.first["spliced"] = kwargs["inputs_embeds"].detach().clone()
    vision.vision_model.register_forward_pre_hook(vis_pre, with_kwargs=True)
    llm.register_forward_pre_hook(llm_pre, with_kwargs=True)

    torch.manual_seed(11)
    tr = ref_trainer.VQATrainerStage2(
        accelerator=acc, vision_encoder=vision, language_model=llm, projection_layer=proj,
        tokenizer=Tok(lc.pad_id, side), train_dataset=VQAList(px_t, q_t, a_t),
        val_dataset=VQAList(px_v, q_v, a_v), output_dir=str(out_dir / "_tmp_vqa"), batch_size=B,
        learning_rate=1e-3, weight_decay=0.01, num_epochs=epochs, gradient_accumulation_steps=accum,
        warmup_ratio=warm, freeze_vision_encoder=True, freeze_projection_layer=not train_proj, freeze_llm=True,
        enable_qlora=lora_cfg is not None, train_ve_first_epoch=False, wandb_project="none")
    t0 = time.time()
    tr.train()
    print(f"[{name}] reference train(): {len(rec.micro)} micro-batches, {len(rec.steps)} optimizer steps, "
          f"val {rec.val} in {time.time() - t0:.1f}s; losses {[round(m['loss'], 5) for m in rec.micro]}",
          flush=True)

    save = {"case": np.array(name), "n_train": n_train, "n_val": n_val, "B": B,
            "padding_side": np.array(side), "grad_accum": accum, "backward_divisor": accum, "epochs": epochs,
            "warmup_ratio": warm, "base_lr": 1e-3, "weight_decay": 0.01, "max_norm": 1.0,
            "n_micro": len(rec.micro), "n_steps": len(rec.steps),
            "val_loss": np.array(rec.val), "spliced0": rec.first["spliced"].float().numpy(),
            "weights_abs_sum": np.array([float(sum(v.double().abs().sum() for v in W.values()))
                                         for W in (Wv, Wl, P)])}
    for i, m in enumerate(rec.micro):
        save[f"micro{i}.idx"] = np.array(m["idx"], dtype=np.int64)
        save[f"micro{i}.attention_mask"] = m["attention_mask"].numpy()
        save[f"micro{i}.loss"] = np.array(m["loss"])
        save[f"micro{i}.sync"] = np.array(int(m["sync"]))
    for i, s in enumerate(rec.steps):
        save[f"step{i}.lr_after"] = np.array(s["lr_after"])
        for n, g in s["grads"].items():
            save[f"step{i}.grad.{n}"] = g.float().numpy()
        for n, g in s.get("lora_grads", {}).items():
            save[f"step{i}.lora_grad.{_lora_key(n)}"] = g.float().numpy()
        if i == len(rec.steps) - 1:
            for n, p in s["params"].items():
                save[f"step{i}.param.{n}"] = p.float().numpy()
            for n, p in s.get("lora_params", {}).items():
                save[f"step{i}.lora_param.{_lora_key(n)}"] = p.float().numpy()
    if lora_cfg is not None:
        save.update({"lora_r": r, "lora_alpha": alpha, "lora_targets": np.array(list(targets)), "lora_b_std": b_std,
                     "train_projector": int(train_proj)})
    out_dir.mkdir(parents=True, exist_ok=True)
    np.savez_compressed(out_dir / f"{name}.npz", **save)
    import shutil
    shutil.rmtree(out_dir / "_tmp_vqa", ignore_errors=True)
    print(f"[{name}] wrote {(out_dir / (name + '.npz')).stat().st_size / 1e6:.2f} MB", flush=True)


if __name__ == "__main__":
    torch.set_num_threads(8)
    only = sys.argv[1] if len(sys.argv) > 1 else None
    for n in ([only] if only else [*CASES, *O2.LORA_CASES]):
        run_case(n, GOLDEN)
