"""Drop-in for ref:Stage2/trainer.py: same `vqa_collate_fn`, same `VQATrainerStage2`
constructor signature, `.train()`, `.evaluate()`, `.save_model()`, artefact names and logged metric names.  The
vision encoder and the LLM's own weights stay frozen; what trains is the projector
(`--unfreeze_projection_layer`) and / or LoRA adapters on the decoder projections (`--enable_qlora`,
ref:Stage2/train_vqa_stage2.py:225-250; lora.py).  The micro-batch body (ref lines 311-420: vision forward,
`[:, 1:, :]`, projector, question/answer embedding, `[image][question][answer]` concatenation, mask, answer-only
labels, fp32 shifted CE, backward) runs in `engine.Stage1Engine` on the sm_100a kernels with the answer rows as its
label span; this file is the host loop: dynamic per-batch padding, `accelerator.accumulate()` bookkeeping, one
clip to 1.0 per trainable module, AdamW, cosine schedule.

Not built (raises NotImplementedError, never falls back): full fine-tuning of the LLM (`--unfreeze_llm`) and the
first-epoch vision-encoder training.  LoRA dropout is not applied (lora.py).  Validation generation is the
reference's call: 3-beam sampling, top-k 50, top-p 0.9 (`engine.generate(num_beams=3, do_sample=True, ...)`, HF's
beam bookkeeping restated and pinned to HF `generate` token for token in tests/test_beam_search_cpu.py)."""
from __future__ import annotations

import json
import logging
import math
import os
from functools import partial

import torch
import torch.nn as nn
from torch.utils.data import DataLoader

from . import weights as Wt
from .engine import PackedGemma, PackedSiglip, ProjectorState, Stage1Engine
from .lora import TARGETS as LORA_TARGETS, LoraState
from .projector_trainer import cosine_with_warmup, sharded_loader

logger = logging.getLogger(__name__)


def vqa_collate_fn(batch, tokenizer):
    """ref:Stage2/trainer.py:18-61 — stack `pixel_values`; pad `question_input_ids` to the longest question
    of the batch and `answer_input_ids` to the longest answer, pads on `tokenizer.padding_side`."""
    pad_id, left = tokenizer.pad_token_id, tokenizer.padding_side == "left"

    def pad_to_longest(seqs):
        longest = max(int(s.shape[0]) for s in seqs)
        out = torch.full((len(seqs), longest), pad_id, dtype=seqs[0].dtype)
        for i, s in enumerate(seqs):
            n = int(s.shape[0])
            if left:
                out[i, longest - n:] = s
            else:
                out[i, :n] = s
        return out

    out = {"question_input_ids": pad_to_longest([it["question_input_ids"] for it in batch]),
           "answer_input_ids": pad_to_longest([it["answer_input_ids"] for it in batch])}
    if "images_u8" in batch[0] or "jpeg_bytes" in batch[0]:
        # GPU preprocessing (vqa_dataset gpu_preprocess / gpu_decode): decoded images of any size, encoded JPEG files
        from .image_io import collate_images
        out.update(collate_images(batch))
    else:
        out["pixel_values"] = torch.stack([it["pixel_values"] for it in batch])
    return out


def vqa_step_inputs(batch: dict, pad_id: int, bucket: int = 1) -> dict:
    """Collated VQA batch -> the engine's batch dict.  Integer work, bit-exact with
    ref:Stage2/trainer.py:366-395: text = [question ; answer], labels = -100 over the question and the
    answer ids with pads -> -100, `label_span` = the columns that can carry a label (the answer; widened to
    bucket boundaries, see below).  `bucket` > 1 appends pad columns
    AFTER the answer so the text length is a multiple of it: masked keys without labels, at positions
    behind every real token, so the loss and gradients are those of the unpadded batch while the number of
    distinct sequence lengths (each owns a workspace and a CUDA graph in the engine) stays small."""
    q, a = batch["question_input_ids"], batch["answer_input_ids"]
    B, Lq = q.shape
    La = a.shape[1]
    T = Lq + La
    Tb = -(-T // bucket) * bucket
    ids = torch.full((B, Tb), pad_id, dtype=torch.int64)
    ids[:, :Lq] = q
    ids[:, Lq:T] = a
    labels = torch.full((B, Tb), -100, dtype=torch.int64)
    labels[:, Lq:T] = torch.where(a == pad_id, torch.full_like(a, -100), a)
    if q.is_pinned():       # DataLoader(pin_memory=True): keep the step's copies asynchronous
        ids, labels = ids.pin_memory(), labels.pin_memory()
    # the span of positions that go through the final norm / LM head / CE is bucketed as well (start rounded down,
    # end = the bucketed text length): the extra rows carry target -100, which the CE kernels skip, so loss and
    # gradients are unchanged, while the engine keys its CUDA graph and its [rows, vocab] logits buffer on the span —
    # exact (Lq, La) pairs would re-capture a graph and add a logits buffer for almost every micro-batch
    off = (Lq // bucket) * bucket
    out = {"token_ids": ids, "labels": labels, "label_span": (off, Tb - off)}
    out.update({k: batch[k] for k in ("pixel_values", "images_u8", "jpeg") if k in batch})
    return out


def accumulate_sync_flag(micro_step: int, grad_accum: int, end_of_dataloader: bool):
    """`accelerate.Accelerator.accumulate()` bookkeeping (ref:Stage2/trainer.py:306): returns
    (sync_gradients, new micro_step).  Gradients synchronise on every `grad_accum`-th micro-batch and on
    the loader's last batch, where the counter restarts."""
    if end_of_dataloader:
        return True, 0
    micro_step += 1
    return micro_step % grad_accum == 0, micro_step


def _peft_adapters(language_model):
    """(LoraConfig fields, adapter state dict by peft key) of a peft-wrapped model, (None, None) for anything else.
    Only attributes are read — peft itself is never imported or run."""
    pc = getattr(language_model, "peft_config", None)
    if not pc:
        return None, None
    c = pc.get("default") or next(iter(pc.values()))
    cfg = {"r": c.r, "lora_alpha": c.lora_alpha, "target_modules": sorted(c.target_modules),
           "lora_dropout": getattr(c, "lora_dropout", 0.0)}
    sd = {n.replace(".default.", "."): p.detach() for n, p in language_model.named_parameters() if ".lora_" in n}
    return cfg, sd


class VQATrainerStage2:
    def __init__(self, accelerator, vision_encoder, language_model, projection_layer, tokenizer, train_dataset,
                 val_dataset, output_dir: str, batch_size: int, learning_rate: float, weight_decay: float,
                 num_epochs: int, gradient_accumulation_steps: int, warmup_ratio: float,
                 freeze_vision_encoder: bool, freeze_projection_layer: bool, freeze_llm: bool,
                 enable_qlora: bool, train_ve_first_epoch: bool, wandb_project: str, lora_config: dict | None = None,
                 lora_state_dict: dict | None = None):
        """`lora_config` / `lora_state_dict` are additions (the reference wraps the LLM with peft before it builds the
        trainer; peft module trees are not executed here): LoraConfig fields `r`, `lora_alpha`, `target_modules`,
        `lora_dropout` (defaults = ref:Stage2/train_vqa_stage2.py:229-236) and adapter weights by peft key
        (`--resume_qlora_adapter_path`).  A peft-wrapped `language_model` is read for both instead."""
        if not freeze_llm and not enable_qlora:
            raise NotImplementedError(
                "Stage 2 with the LLM's own weights trainable (--unfreeze_llm) is not built: the decoder's "
                "weight-gradient GEMMs exist for LoRA adapters only (SURVEY.md §8f); there is no fallback path")
        if train_ve_first_epoch or not freeze_vision_encoder:
            raise NotImplementedError("first-epoch vision-encoder training (--train_ve_first_epoch) is not built: "
                                      "the SigLIP tower is forward-only here")
        if freeze_projection_layer and not enable_qlora:
            # ref:Stage2/trainer.py:242-243
            raise ValueError("No trainable parameters found. Check freezing configuration and QLoRA setup.")
        self.accelerator, self.tokenizer = accelerator, tokenizer
        self.vision_encoder, self.language_model, self.projection_layer = vision_encoder, language_model, projection_layer
        self.device = accelerator.device
        self.output_dir, self.num_epochs, self.batch_size = output_dir, num_epochs, batch_size
        self.wandb_project, self.train_ve_first_epoch, self.enable_qlora = wandb_project, False, bool(enable_qlora)
        self.learning_rate, self.weight_decay = learning_rate, weight_decay
        self.grad_accum = int(gradient_accumulation_steps)
        if self.device.type != "cuda":
            raise RuntimeError("VQATrainerStage2 (B200 build) needs a CUDA device: the step runs on sm_100a "
                               "kernels only, there is no CPU path")
        self.validation_dir = os.path.join(output_dir, "validation_examples")
        if accelerator.is_main_process:
            os.makedirs(self.validation_dir, exist_ok=True)
            os.makedirs(output_dir, exist_ok=True)
        self.bucket = int(os.environ.get("PT_STAGE2_BUCKET", "32"))

        collate = partial(vqa_collate_fn, tokenizer=tokenizer)
        gen = torch.Generator().manual_seed(0)      # same shuffle on every process (accelerate syncs RNG)
        rank, world = accelerator.process_index, accelerator.num_processes
        self.train_loader = sharded_loader(train_dataset, batch_size=batch_size, shuffle=True, rank=rank, world=world,
                                           generator=gen, num_workers=4, pin_memory=True, collate_fn=collate)
        self.val_loader = sharded_loader(val_dataset, batch_size=batch_size, shuffle=False, rank=rank, world=world,
                                         num_workers=4, pin_memory=True, collate_fn=collate)
        # schedule length from the UNSHARDED loader, as the reference computes it before prepare() (:152-157)
        self.max_train_steps = num_epochs * math.ceil(self.train_loader.unsharded_len / self.grad_accum)
        self.num_warmup_steps = math.ceil(warmup_ratio * self.max_train_steps)
        self._sched_step = 0
        self.val_loader.dataset = val_dataset

        vis = vision_encoder if isinstance(vision_encoder, PackedSiglip) else \
            PackedSiglip(*Wt.siglip_from_hf(getattr(vision_encoder, "vision_model", vision_encoder)), self.device)
        peft_cfg, peft_sd = _peft_adapters(language_model) if enable_qlora else (None, None)
        if isinstance(language_model, PackedGemma):
            llm = language_model        # already packed (and, for QLoRA, already through weights.nf4_frozen_llm_)
        else:
            Wl, lc = Wt.llm_from_hf(language_model)
            if enable_qlora:
                # ref:Stage2/train_vqa_stage2.py:168-186: the base weights the adapters ride on are the NF4 images
                err = Wt.nf4_frozen_llm_(Wl, lc, self.device)
                logger.info(f"LLM base weights through NF4: mean relative error {sum(err.values()) / len(err):.4f}")
            llm = PackedGemma(Wl, lc, self.device)
        if isinstance(projection_layer, nn.Module):
            projection_layer.to(self.device).float()      # fp32 master weights (the reference casts the
            projection_layer.requires_grad_(not freeze_projection_layer)   # projector to bf16 under mixed precision)
            proj = ProjectorState.from_module(projection_layer, self.device)
        else:
            proj = projection_layer
        lora = None
        if enable_qlora:
            cfg = {"r": 16, "lora_alpha": 32, "target_modules": list(LORA_TARGETS), "lora_dropout": 0.05}
            cfg.update(peft_cfg or {})
            cfg.update(lora_config or {})
            lora = LoraState(llm.cfg, self.device, r=int(cfg["r"]), alpha=float(cfg["lora_alpha"]),
                             targets=tuple(cfg["target_modules"]), dropout=float(cfg["lora_dropout"]))
            sd = lora_state_dict or peft_sd
            if sd:
                lora.load_state_dict(sd)
            if lora.dropout > 0:
                logger.warning(f"lora_dropout={lora.dropout} is recorded in adapter_config.json but NOT applied "
                               "during training (stated difference from the reference, see lora.py)")
            logger.info(f"trainable params: {lora.num_parameters():,d} (LoRA r={lora.r}, alpha={lora.alpha:g}, "
                        f"targets {list(lora.targets)})")
        ga = self.grad_accum
        self.engine = Stage1Engine(vis, llm, proj, self.device, lora=lora, train_projector=not freeze_projection_layer,
                                   loss_divisor=ga * getattr(accelerator, "backward_divisor", ga))
        self.pad_id = tokenizer.pad_token_id
        if self.pad_id is None:
            raise ValueError("tokenizer.pad_token_id is None: the dataset sets it to eos (ref:Stage2/dataset.py:33-35)")
        logger.info("QLoRA enabled. Trainable parameters determined by PEFT." if enable_qlora
                    else "Freezing LLM parameters (QLoRA disabled).")
        logger.info("Freezing projection_layer parameters." if freeze_projection_layer
                    else "projection_layer parameters are trainable.")

    # accelerate's prepared scheduler ticks num_processes times per optimizer step
    def current_lr(self) -> float:
        return self.learning_rate * cosine_with_warmup(self._sched_step, self.num_warmup_steps, self.max_train_steps)

    def _step_inputs(self, batch):
        return vqa_step_inputs(batch, self.pad_id, self.bucket)

    def _prefetched(self, loader):
        """One-batch look-ahead, yields (engine batch, is_last): the host->device copy of micro-batch i+1
        runs on a side stream while micro-batch i computes."""
        it = iter(loader)
        cur = next(it, None)
        cur = None if cur is None else self.engine.stage_batch(self._step_inputs(cur))
        while cur is not None:
            nxt = next(it, None)
            nxt = None if nxt is None else self.engine.stage_batch(self._step_inputs(nxt))
            yield cur, nxt is None
            cur = nxt

    def train(self):
        acc = self.accelerator
        logger.info(f"Process {acc.process_index}: Starting Stage 2 training for {self.num_epochs} epochs on device {self.device}")
        global_step, micro = 0, 0
        for epoch in range(self.num_epochs):
            epoch_train_loss = 0.0
            for batch, last in self._prefetched(self.train_loader):
                sync, micro = accumulate_sync_flag(micro, self.grad_accum, last)
                loss = self.engine.train_step(batch, self.current_lr(), pad_id=self.pad_id,
                                              weight_decay=self.weight_decay, allreduce=acc.all_reduce_sum_,
                                              world_size=acc.num_processes, max_norm=1.0, sync_gradients=sync)
                micro_loss = acc.gather(loss).mean().item()             # the one host sync of the micro-batch
                self.engine.check_errors()      # out-of-range token ids / labels flagged by the kernels
                if sync:
                    self._sched_step += acc.num_processes
                    global_step += 1
                    epoch_train_loss += micro_loss
                if acc.is_main_process:
                    d = {"train/batch_loss": micro_loss}
                    if sync:
                        d.update({"train/learning_rate": self.current_lr(), "step": global_step,
                                  "train/step_loss": micro_loss})
                    acc.log(d, step=global_step)
            avg_epoch_loss = epoch_train_loss / max(1, len(self.train_loader))      # ref :467 (sync steps only)
            if acc.is_main_process:
                logger.info(f"Epoch {epoch + 1} Training Loss: {avg_epoch_loss:.4f}")
                acc.log({"train/loss": avg_epoch_loss, "train/learning_rate": self.current_lr(),
                         "epoch": epoch + 1}, step=global_step)
            self.evaluate(epoch, global_step)
            if acc.is_main_process:
                self.save_model(os.path.join(self.output_dir, f"checkpoint-epoch_{epoch + 1}"))
        logger.info(f"Process {acc.process_index}: Stage 2 training complete!")

    @torch.no_grad()
    def evaluate(self, epoch, global_step, max_new_tokens: int = 512):
        """ref:Stage2/trainer.py:490-677: validation loss over the whole set (left-padded collation), beam-sampled
        predictions from `[image][question]` (:603-611), up to 20 examples written to validation_examples/."""
        acc, tok = self.accelerator, self.tokenizer
        logger.info(f"Starting evaluation for Epoch {epoch + 1}...")
        original_side = tok.padding_side
        tok.padding_side = "left"
        total_val_loss, n_batches, examples = 0.0, 0, []
        max_examples = min(20, len(self.val_loader.dataset))
        can_decode = hasattr(tok, "batch_decode")
        try:
            for vb in self.val_loader:
                loss = self.engine.eval_loss(self._step_inputs(vb), pad_id=self.pad_id)
                total_val_loss += acc.gather(loss).mean().item()
                n_batches += 1
                if not can_decode:
                    continue
                try:
                    px = vb["pixel_values"] if "pixel_values" in vb else \
                        self.engine.preprocess_images(self.engine.raw_images(vb))
                    gen = self.engine.generate(px, vb["question_input_ids"],
                                               max_new_tokens=max_new_tokens, do_sample=True, num_beams=3, top_p=0.9, top_k=50,
                                               pad_token_id=self.pad_id, eos_token_id=tok.eos_token_id,
                                               prompt_pad_id=self.pad_id)
                    preds = tok.batch_decode(gen.cpu(), skip_special_tokens=True)
                    questions = tok.batch_decode(vb["question_input_ids"], skip_special_tokens=True)
                    refs = tok.batch_decode(vb["answer_input_ids"], skip_special_tokens=True)
                    examples.extend({"epoch": epoch + 1, "question": q, "ground_truth": r, "prediction": p}
                                    for q, r, p in zip(questions, refs, preds))
                except Exception as e:
                    logger.error(f"Error during validation generation/logging: {e}", exc_info=True)
        finally:
            tok.padding_side = original_side
        acc.wait_for_everyone()
        gathered = acc.gather_object(examples) if hasattr(acc, "gather_object") else examples
        gathered = gathered[:max_examples] if acc.is_main_process else []
        avg_val_loss = total_val_loss / max(1, n_batches)
        if acc.is_main_process:
            logger.info(f"Epoch {epoch + 1} Validation Loss: {avg_val_loss:.4f}")
            acc.log({"val/loss": avg_val_loss, "epoch": epoch + 1}, step=global_step)
            if gathered:
                block = [f"Validation Examples for Epoch {epoch + 1}\n", "=" * 80 + "\n\n"]
                for i, ex in enumerate(gathered):
                    block += [f"Example {i + 1}:\n", f"Question: {ex['question']}\n",
                              f"Ground Truth: {ex['ground_truth']}\n", f"Prediction: {ex['prediction']}\n",
                              "=" * 80 + "\n\n"]
                with open(os.path.join(self.validation_dir, f"epoch_{epoch + 1}_examples.txt"), "w",
                          encoding="utf-8") as f:
                    f.writelines(block)
                with open(os.path.join(self.validation_dir, "all_validation_examples.txt"),
                          "a" if epoch > 0 else "w", encoding="utf-8") as f:
                    f.writelines(["\n"] + block)
        logger.info("Evaluation finished.")
        return avg_val_loss

    def save_model(self, path):
        """ref:Stage2/trainer.py:679-768: `language_model/adapter_model.safetensors` + `adapter_config.json` (what
        peft's `save_pretrained` writes, :724-729) when adapters train; `projection_layer/model.safetensors` (what
        `accelerator.save_model` writes and Stage2/train_vqa_stage2.py:25-80 loads back) + `projector_config.json`
        when the projector trains; the optimizer moments stand in for `accelerator.save_state`.  The frozen base
        model is not saved (ref :735-736)."""
        if not self.accelerator.is_main_process:
            return
        from safetensors.torch import save_file
        torch.cuda.synchronize(self.device)
        os.makedirs(path, exist_ok=True)
        eng, opt = self.engine, {"scheduler_step": self._sched_step}
        if eng.lora is not None:
            lo, llm_dir = eng.lora, os.path.join(path, "language_model")
            os.makedirs(llm_dir, exist_ok=True)
            logger.info("Saving QLoRA adapters using PEFT save_pretrained...")
            save_file({k: v.detach().cpu().clone().contiguous() for k, v in lo.state_dict().items()},
                      os.path.join(llm_dir, "adapter_model.safetensors"))
            base = getattr(getattr(self.language_model, "config", None), "_name_or_path", "") or ""
            with open(os.path.join(llm_dir, "adapter_config.json"), "w") as f:
                json.dump(lo.adapter_config(base), f, indent=2)
            if hasattr(self.tokenizer, "save_pretrained"):
                self.tokenizer.save_pretrained(llm_dir)
            host = lambda buf: {n: t.cpu().clone() for n, t in lo.state_dict(buf).items()}
            opt["lora"] = {"exp_avg": host(lo.m), "exp_avg_sq": host(lo.v), "step": lo.step}
            logger.info(f"QLoRA adapters saved to {llm_dir}")
        else:
            logger.info("Language model was frozen and not saved.")
        if eng.train_projector:
            proj = eng.proj
            proj_dir = os.path.join(path, "projection_layer")
            os.makedirs(proj_dir, exist_ok=True)
            save_file({k: v.detach().cpu().clone().contiguous() for k, v in proj.params().items()},
                      os.path.join(proj_dir, "model.safetensors"))
            with open(os.path.join(proj_dir, "projector_config.json"), "w") as f:
                json.dump({"vision_dim": int(proj.shapes[0][1]), "llm_dim": int(proj.shapes[2][0])}, f, indent=4)
            opt.update({"exp_avg": proj.m.cpu(), "exp_avg_sq": proj.v.cpu(), "step": proj.step})
            logger.info(f"Projection layer saved to {proj_dir}")
        torch.save(opt, os.path.join(path, "optimizer.bin"))
