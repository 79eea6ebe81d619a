"""Drop-in for ref:Stage1/projector_trainer.py: same class, constructor signature, `.train()`,
`.save_projection()`, artefact names and logged metric names.  The step body (ref lines 152-245: vision
forward, `[:, 1:, :]` slice, projector, embed + splice, masks/labels, frozen-LLM loss, backward, clip 5.0,
AdamW, cosine schedule) executes in `engine.Stage1Engine` on the sm_100a kernels; this file is the host
loop around it.  Differences that are deliberate and documented in DESIGN.md: no DDP wrappers (one flat
NCCL all-reduce of the projector gradient) and one host read of the loss per step instead of two.  The
validation pass reports the loss and, when the tokenizer can decode, the sampled-generation last-word
accuracy (prefill + KV-cached decode in the engine)."""
from __future__ import annotations

import gc
import json
import logging
import math
import os
import re
import time

import torch
import torch.nn as nn
from torch.utils.data import DataLoader

from . import weights as Wt
from .engine import PackedGemma, PackedSiglip, ProjectorState, Stage1Engine

logger = logging.getLogger(__name__)


def get_last_word(text) -> str:
    """ref:Stage1/projector_trainer.py:132-137."""
    if not text or not isinstance(text, str):
        return ""
    words = re.findall(r"\b\w+\b", text.lower())
    return words[-1] if words else ""


def cosine_with_warmup(step: int, warmup: int, total: int) -> float:
    """lr multiplier of HF `get_cosine_schedule_with_warmup` (ref:projector_trainer.py:91-95)."""
    if step < warmup:
        return step / max(1, warmup)
    progress = (step - warmup) / max(1, total - warmup)
    return max(0.0, 0.5 * (1.0 + math.cos(math.pi * progress)))


def _fork_safe_iter(loader):
    """DataLoader workers are forked when the iterator is built.  A child that inherits unreachable but not yet
    collected CUDA objects (a dropped geometry's graphs, events, pinned blocks) and garbage-collects them calls
    into a CUDA context it does not own and aborts: collect them in this process first, and keep the survivors
    frozen (out of the child's collector) across the fork."""
    gc.collect()
    gc.freeze()
    try:
        return iter(loader)
    finally:
        gc.unfreeze()


class _ShardedBatchSampler(torch.utils.data.Sampler):
    """Batch sampler of ONE process: what accelerate's `BatchSamplerShard` (split_batches=False, even_batches=True)
    hands it.  Every process draws the same index stream (same seeded sampler), cuts it into batches and keeps
    batch i iff i % world == rank, so a process only ever LOADS its own samples (decode / resize / tokenize run for
    1/world of the data set per process); processes that come up short take one of the first `world` batches again so
    that all of them run the same number of optimizer steps."""

    def __init__(self, sampler, batch_size: int, rank: int, world: int):
        self.sampler, self.batch_size, self.rank, self.world = sampler, int(batch_size), int(rank), int(world)

    @property
    def unsharded_len(self) -> int:
        return math.ceil(len(self.sampler) / self.batch_size)

    def __len__(self):
        return math.ceil(self.unsharded_len / self.world)

    def __iter__(self):
        first, mine, batch, i = [], 0, [], 0
        def done(b):
            nonlocal mine, i
            if i < self.world:
                first.append(b)
            take = i % self.world == self.rank
            i += 1
            if take:
                mine += 1
            return take
        for idx in self.sampler:
            batch.append(idx)
            if len(batch) == self.batch_size:
                if done(batch):
                    yield batch
                batch = []
        if batch and done(batch):
            yield batch
        if mine < len(self) and first:
            yield first[self.rank % len(first)]


def sharded_loader(dataset, *, batch_size, shuffle, rank, world, generator=None, **loader_kw):
    """DataLoader over this process's share of the batches (see _ShardedBatchSampler), wrapped for fork-safe
    iteration.  `.unsharded_len` = number of batches of the whole data set (the reference sizes its schedule from
    the loader BEFORE `accelerator.prepare` shards it: ref:Stage1/projector_trainer.py:82-87)."""
    from torch.utils.data import RandomSampler, SequentialSampler
    sampler = RandomSampler(dataset, generator=generator) if shuffle else SequentialSampler(dataset)
    bs = _ShardedBatchSampler(sampler, batch_size, rank, world)
    # the DataLoader draws its workers' base seed from the same generator, as DataLoader(shuffle=True, generator=g) does
    out = _ShardedLoader(DataLoader(dataset, batch_sampler=bs, generator=generator, **loader_kw), rank, world,
                         presharded=True)
    out.unsharded_len = bs.unsharded_len
    return out


class _ShardedLoader:
    """What `accelerator.prepare(dataloader)` gives each process: batch i of the (identically shuffled)
    loader goes to process i % N; short processes wrap around so that every process takes the same
    number of optimizer steps (accelerate's even_batches).  With `presharded` the wrapped DataLoader already
    yields this process's batches only (sharded_loader: nothing is loaded and thrown away); otherwise any iterable of
    batches is filtered here."""

    def __init__(self, loader, rank, world, presharded: bool = False):
        self.loader, self.rank, self.world, self.presharded = loader, rank, world, presharded

    def __len__(self):
        return len(self.loader) if self.presharded else math.ceil(len(self.loader) / self.world)

    def __iter__(self):
        if self.world == 1 or self.presharded:
            yield from _fork_safe_iter(self.loader)
            return
        first, mine = [], 0
        for i, b in enumerate(_fork_safe_iter(self.loader)):
            if i < self.world:
                first.append(b)
            if i % self.world == self.rank:
                mine += 1
                yield b
        if mine < len(self) and first:
            yield first[self.rank % len(first)]


def stage1_collate(items):
    """torch's default collate, except that decoded images of different sizes (`images_u8`, datasets built
    with gpu_preprocess=True) stay a list and encoded JPEG files (`jpeg_bytes`, gpu_decode=True) are packed for the
    GPU decoder (image_io.collate_images)."""
    from torch.utils.data import default_collate
    from .image_io import collate_images, has_raw_images, strip_images
    if not has_raw_images(items):
        return default_collate(items)
    out = default_collate([strip_images(it) for it in items])
    out.update(collate_images(items))
    return out


class ProjectionTrainerStage1:
    def __init__(self, accelerator, vision_encoder, language_model, projection_layer, processor, tokenizer,
                 train_dataset, val_dataset=None, output_dir="./trained_projection_stage1", batch_size=8,
                 learning_rate=1e-4, weight_decay=0.01, num_epochs=10, gradient_accumulation_steps=1,
                 warmup_ratio=0.0, wandb_project="xray_projection_training", save_every_n_epochs=0):
        self.accelerator = accelerator
        self.vision_encoder, self.language_model = vision_encoder, language_model
        self.projection_layer = projection_layer
        self.processor, self.tokenizer = processor, tokenizer
        self.device = accelerator.device
        self.output_dir, self.num_epochs, self.batch_size = output_dir, num_epochs, batch_size
        self.wandb_project, self.save_every_n_epochs = wandb_project, save_every_n_epochs
        self.val_dataset = val_dataset
        self.learning_rate, self.weight_decay = learning_rate, weight_decay
        if self.device.type != "cuda":
            raise RuntimeError("ProjectionTrainerStage1 (B200 build) needs a CUDA device: the Stage-1 step "
                               "runs on sm_100a kernels only, there is no CPU path")
        if accelerator.is_main_process:
            os.makedirs(output_dir, exist_ok=True)

        gen = torch.Generator().manual_seed(0)      # same shuffle on every process (accelerate syncs RNG)
        rank, world = accelerator.process_index, accelerator.num_processes
        self.train_loader = sharded_loader(train_dataset, batch_size=batch_size, shuffle=True, rank=rank, world=world,
                                           generator=gen, num_workers=2, pin_memory=True, collate_fn=stage1_collate)
        self.val_loader = None
        if val_dataset is not None:
            self.val_loader = sharded_loader(val_dataset, batch_size=batch_size, shuffle=False, rank=rank,
                                             world=world, num_workers=2, pin_memory=True, collate_fn=stage1_collate)
        # schedule lengths are computed from the UNSHARDED loader, as the reference does (:82-87)
        steps_per_epoch = math.ceil(self.train_loader.unsharded_len / gradient_accumulation_steps)
        self.max_train_steps = num_epochs * steps_per_epoch
        self.num_warmup_steps = math.ceil(warmup_ratio * self.max_train_steps)
        self._sched_step = 0

        # frozen towers -> packed bf16 weights (read once; none of the HF forward code runs afterwards)
        vis = vision_encoder if isinstance(vision_encoder, PackedSiglip) else \
            PackedSiglip(*Wt.siglip_from_hf(getattr(vision_encoder, "vision_model", vision_encoder)), self.device)
        llm = language_model if isinstance(language_model, PackedGemma) else \
            PackedGemma(*Wt.llm_from_hf(language_model), self.device)
        if isinstance(projection_layer, nn.Module):
            projection_layer.to(self.device).float()
            projection_layer.requires_grad_(True)
            proj = ProjectorState.from_module(projection_layer, self.device)
        else:
            proj = projection_layer
        ga = accelerator.gradient_accumulation_steps
        self.engine = Stage1Engine(vis, llm, proj, self.device,
                                   loss_divisor=ga * getattr(accelerator, "backward_divisor", ga))
        pad = getattr(tokenizer, "pad_token_id", None)
        self.pad_id = pad
        logger.info("Froze vision encoder and language model parameters (packed, no gradients are formed).")

    # ---- schedule: accelerate's prepared scheduler ticks num_processes times per optimizer step
    def current_lr(self) -> float:
        return self.learning_rate * cosine_with_warmup(self._sched_step, self.num_warmup_steps, self.max_train_steps)

    def _prefetched(self, loader):
        """One-batch look-ahead: the host->device copy of batch i+1 runs on a side stream while step i
        computes (the reference gets the same overlap from DataLoader workers + pinned memory)."""
        def on_host(b):
            return "images_u8" in b or "jpeg" in b or not b["pixel_values"].is_cuda
        it = iter(loader)
        cur = next(it, None)
        if cur is not None and on_host(cur):
            cur = self.engine.stage_batch(cur)
        while cur is not None:
            nxt = next(it, None)
            if nxt is not None and on_host(nxt):
                nxt = self.engine.stage_batch(nxt)
            yield cur
            cur = nxt

    def train(self):
        acc = self.accelerator
        logger.info(f"Process {acc.process_index}: Starting Stage 1 training for {self.num_epochs} epochs on {self.device}")
        global_step = 0
        best_val_loss = float("inf")
        for epoch in range(self.num_epochs):
            epoch_train_loss = 0.0
            n_steps = 0
            for batch in self._prefetched(self.train_loader):
                lr = self.current_lr()
                loss = self.engine.train_step(batch, lr, pad_id=self.pad_id, weight_decay=self.weight_decay,
                                              allreduce=acc.all_reduce_sum_, world_size=acc.num_processes)
                self._sched_step += acc.num_processes
                global_step += 1
                n_steps += 1
                avg_loss_step = acc.gather(loss).mean().item()          # the one host sync of the step
                self.engine.check_errors()      # out-of-range token ids / labels flagged by the kernels
                epoch_train_loss += avg_loss_step
                if acc.is_main_process:
                    acc.log({"train/batch_loss": avg_loss_step, "train/learning_rate": self.current_lr(),
                             "step": global_step}, step=global_step)
            steps_per_epoch = math.ceil(len(self.train_loader) / acc.gradient_accumulation_steps)
            avg_epoch = epoch_train_loss / max(1, steps_per_epoch)
            if acc.is_main_process:
                acc.log({"train/epoch_loss": avg_epoch, "epoch": epoch + 1}, step=global_step)
                logger.info(f"Epoch {epoch + 1}/{self.num_epochs} completed. Avg Train Loss: {avg_epoch:.4f}")
                if self.save_every_n_epochs > 0 and (epoch + 1) % self.save_every_n_epochs == 0:
                    self.save_projection(epoch=epoch + 1)
            if self.val_loader is not None:
                val_loss, val_steps, total_samples, correct_last_word = 0.0, 0, 0, 0
                can_decode = hasattr(self.tokenizer, "batch_decode")
                for vb in self.val_loader:
                    if "images_u8" in vb or "jpeg" in vb:
                        vb["pixel_values"] = self.engine.preprocess_images(self.engine.raw_images(vb))
                    val_loss += self.engine.eval_loss(vb, pad_id=self.pad_id).item()
                    val_steps += 1
                    if not can_decode:
                        continue
                    try:    # sampled generation from the image tokens only (ref :374-409)
                        gen = self.engine.generate(vb["pixel_values"], None, max_new_tokens=64, do_sample=True,
                                                   pad_token_id=self.pad_id,
                                                   eos_token_id=getattr(self.tokenizer, "eos_token_id", None))
                        gen_texts = self.tokenizer.batch_decode(gen.cpu(), skip_special_tokens=True)
                        ref_texts = self.tokenizer.batch_decode(vb["token_ids"].cpu().numpy(),
                                                                skip_special_tokens=True)
                        for r, g_ in zip(ref_texts, gen_texts):
                            rl, gl = get_last_word(r), get_last_word(g_)
                            correct_last_word += int(bool(rl) and bool(gl) and rl == gl)
                            total_samples += 1
                    except Exception as e:
                        logger.error(f"Error during generation for accuracy calculation: {e}", exc_info=True)
                avg_val = val_loss / val_steps if val_steps else float("inf")
                last_word_accuracy = (correct_last_word / total_samples) * 100 if total_samples else 0.0
                if avg_val < best_val_loss:
                    best_val_loss = avg_val
                    if acc.is_main_process:
                        self.save_projection(epoch=epoch + 1, is_best=True)
                if acc.is_main_process:
                    logger.info(f"Epoch {epoch + 1}/{self.num_epochs} validation loss: {avg_val:.4f}, last word "
                                f"accuracy: {last_word_accuracy:.2f}% ({correct_last_word}/{total_samples})")
                    acc.log({"validation/loss": avg_val, "validation/last_word_accuracy": last_word_accuracy,
                             "epoch": epoch + 1}, step=global_step)
        logger.info("Stage 1 Training finished.")
        self.save_projection(epoch=self.num_epochs)

    def save_projection(self, epoch=None, is_best=False):
        """Same files as ref:projector_trainer.py:455-521: projector_{best|epoch_N|final}.bin holding the
        fp32 state dict (keys model.0.*, model.2.*) and projector_config.json {vision_dim, llm_dim}."""
        if not self.accelerator.is_main_process:
            return
        torch.cuda.synchronize(self.device)
        proj = self.engine.proj
        name = "projector_best.bin" if is_best else (f"projector_epoch_{epoch}.bin" if epoch is not None
                                                     else "projector_final.bin")
        sd = {k: v.detach().cpu().clone() for k, v in proj.params().items()}
        torch.save(sd, os.path.join(self.output_dir, name))
        cfg = {"vision_dim": int(proj.shapes[0][1]), "llm_dim": int(proj.shapes[2][0])}
        with open(os.path.join(self.output_dir, "projector_config.json"), "w") as f:
            json.dump(cfg, f, indent=4)
        logger.info(f"Projection layer state dict saved to {os.path.join(self.output_dir, name)}")
