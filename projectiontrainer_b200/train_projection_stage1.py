"""Drop-in for ref:Stage1/train_projection_stage1.py: same flags (ref lines 139-159), same dataset item
schema, same model-loading order; the training step itself runs on the B200 kernels.

    torchrun --nproc-per-node N -m projectiontrainer_b200.train_projection_stage1 --image_root ... --train_json ...
    python -m projectiontrainer_b200.train_projection_stage1 --synthetic 64 --disable_wandb     # no files, no weights

`--synthetic N` (an addition) trains on N random CXR-shaped samples with random-init SigLIP-L/16-384 and
Gemma3 weights — there are no checkpoints or datasets on the build / GPU boxes.  `--enable_qlora` (ref:192-210: the
frozen LLM loaded as double-quantised NF4) stores the decoder's Linear weights as NF4 and expands them once at load time
(`weights.nf4_frozen_llm_`); bitsandbytes is not needed."""
from __future__ import annotations

import argparse
import json
import logging
import os

import torch
from torch.utils.data import Dataset

from .accelerator_setup import setup_accelerator_and_logging
from .config import GEMMA3_1B, GEMMA3_4B, QWEN3_8B, SIGLIP_L
from .projector_trainer import ProjectionTrainerStage1
from .projectors import MLPProjector

logger = logging.getLogger(__name__)


class XrayTextPairDataset(Dataset):
    """X-ray image / caption pairs from a JSON list of {"image", "normal_caption"} records
    (ref:train_projection_stage1.py:25-126).  Items: pixel_values f32 [3,S,S] (processor output),
    token_ids i64 [max_length] (tokenizer padded to max_length), labels = token_ids with pad -> -100."""

    def __init__(self, image_root, json_file, processor, tokenizer, img_size, max_length=512, image_root_2=None,
                 gpu_preprocess=False, cache_tokens=False, gpu_decode=False):
        # gpu_preprocess (an addition): items carry the decoded image as `images_u8` uint8 [H, W, 3]; resize +
        # rescale + normalize run on the GPU in the trainer's look-ahead stage (Pillow-exact kernels)
        # gpu_decode (an addition, implies gpu_preprocess): workers only read the file; baseline JPEGs are decoded on
        # the GPU too (image_io.py)
        self.gpu_preprocess, self.gpu_decode = gpu_preprocess or gpu_decode, gpu_decode
        self.image_root, self.image_root_2, self.img_size = image_root, image_root_2, img_size
        self.processor, self.tokenizer, self.max_length = processor, tokenizer, max_length
        self.samples = []
        if json_file:
            with open(json_file, "r", encoding="utf-8") as f:
                self.samples = json.load(f)
        # cache_tokens (an addition): captions are tokenised once, on first use after the
        # sample list is final, into one int64 matrix that forked DataLoader workers share, instead of once per
        # epoch per sample inside the workers
        self.cache_tokens, self._token_matrix = cache_tokens, None

    def _tokens(self, idx):
        if not self.cache_tokens:
            return self.tokenizer(self.samples[idx]["normal_caption"], max_length=self.max_length,
                                  padding="max_length", truncation=True, return_tensors="pt").input_ids.squeeze(0)
        self._ensure_tokens()
        return self._token_matrix[idx].clone()

    def _ensure_tokens(self):
        if self.cache_tokens and self.samples and (self._token_matrix is None or
                                                   self._token_matrix.shape[0] != len(self.samples)):
            enc = self.tokenizer([s["normal_caption"] for s in self.samples], max_length=self.max_length,
                                 padding="max_length", truncation=True, return_tensors="pt")
            self._token_matrix = enc.input_ids.contiguous().share_memory_()

    def __len__(self):
        # the trainer sizes its schedule from len(loader) in the parent process, before any worker is forked:
        # building the cache here makes every worker inherit it instead of rebuilding it
        self._ensure_tokens()
        return len(self.samples)

    def _resolve(self, name):
        for root in (self.image_root, self.image_root_2):
            if root is None:
                continue
            p = os.path.join(root, name)
            if os.path.isdir(p):      # a study directory: take its first image file
                files = sorted(f for f in os.listdir(p) if f.lower().endswith((".jpg", ".jpeg", ".png")))
                if files:
                    return os.path.join(p, files[0])
            if os.path.isfile(p):
                return p
        raise FileNotFoundError(f"Failed to find image file. Started with: {name}. Looked in: "
                                f"{self.image_root}, {self.image_root_2}")

    def __getitem__(self, idx):
        from PIL import Image
        sample = self.samples[idx]
        image = None if self.gpu_decode else Image.open(self._resolve(sample["image"])).convert("RGB")
        token_ids = self._tokens(idx)
        labels = token_ids.clone()
        if self.tokenizer.pad_token_id is not None:
            labels[labels == self.tokenizer.pad_token_id] = -100
        if self.gpu_decode:
            from .image_io import load_image_item
            return {**load_image_item(self._resolve(sample["image"]), True), "token_ids": token_ids, "labels": labels}
        if self.gpu_preprocess:
            import numpy as np
            return {"images_u8": torch.from_numpy(np.array(image, dtype=np.uint8)), "token_ids": token_ids,
                    "labels": labels}
        image_inputs = self.processor(images=image.resize((self.img_size, self.img_size)), return_tensors="pt")
        return {"pixel_values": image_inputs.pixel_values.squeeze(0), "token_ids": token_ids, "labels": labels}


class SyntheticXrayTextDataset(Dataset):
    """CXR-shaped random samples: pixels ~ U(-1,1), ids[0] = BOS(2), rest ~ U{3..V-1}."""

    def __init__(self, n, img_size, tokens, vocab, pad_id=0, seed=1234):
        g = torch.Generator().manual_seed(seed)
        self.px = torch.rand(n, 3, img_size, img_size, generator=g) * 2 - 1
        self.ids = torch.randint(3, vocab, (n, tokens), generator=g)
        self.ids[:, 0] = 2
        self.pad_id = pad_id

    def __len__(self):
        return self.px.shape[0]

    def __getitem__(self, i):
        ids = self.ids[i]
        labels = ids.clone()
        labels[ids == self.pad_id] = -100
        return {"pixel_values": self.px[i], "token_ids": ids, "labels": labels}


class _Tok:
    def __init__(self, pad, eos=1):
        self.pad_token_id, self.eos_token_id = pad, eos


def build_parser():
    p = argparse.ArgumentParser(description="Train Stage 1: Vision-Language Projector Alignment (B200 build)")
    p.add_argument("--image_root", type=str, default=None)
    p.add_argument("--image_root_2", type=str, default=None)
    p.add_argument("--train_json", type=str, default=None)
    p.add_argument("--val_json", type=str, default=None)
    p.add_argument("--train_val_split", type=float, default=0.0)
    p.add_argument("--output_dir", type=str, default="./trained_projection_stage1")
    p.add_argument("--vision_model_name", type=str, default="google/siglip-large-patch16-384")
    p.add_argument("--llm_name", type=str, default="google/gemma-3-1b-it")
    p.add_argument("--batch_size", type=int, default=4)
    p.add_argument("--learning_rate", type=float, default=1e-4)
    p.add_argument("--weight_decay", type=float, default=0.01)
    p.add_argument("--num_epochs", type=int, default=10)
    p.add_argument("--warmup_ratio", type=float, default=0.0)
    p.add_argument("--wandb_project", type=str, default="xray_projection_training_stage1")
    p.add_argument("--wandb_run_name", type=str, default=None)
    p.add_argument("--wandb_log_freq", type=int, default=100)
    p.add_argument("--disable_wandb", action="store_true")
    p.add_argument("--gradient_accumulation_steps", type=int, default=2)
    p.add_argument("--img_size", type=int, default=384)
    p.add_argument("--save_every_n_epochs", type=int, default=2)
    p.add_argument("--enable_qlora", action="store_true")
    p.add_argument("--gpu_decode", action="store_true",
                   help="workers only read the image files; baseline JPEGs are decoded on the GPU (bit-exact with Pillow), "
                        "then resized / normalized there (implies --gpu_preprocess)")
    p.add_argument("--gpu_preprocess", action="store_true",
                   help="resize / rescale / normalize the decoded images on the GPU (Pillow-exact) instead of in the workers")
    p.add_argument("--cache_tokens", action="store_true", help="tokenise every caption once instead of once per epoch")
    p.add_argument("--synthetic", type=int, default=0, help="train on N synthetic samples with random-init models")
    p.add_argument("--synthetic_llm", choices=["1b", "4b", "qwen3_8b"], default="1b")
    p.add_argument("--max_length", type=int, default=512, help="caption tokens (the reference's dataset default)")
    return p


def _nf4(Wt, Wl, lc, dev):
    """ref:195-210 loads the frozen LLM as double-quantised NF4 (bitsandbytes); here the bf16 weights go through the
    same quantise -> de-quantise once at load time and the step runs on the resulting bf16 matrices."""
    import time
    t0 = time.perf_counter()
    errs = list(Wt.nf4_frozen_llm_(Wl, lc, dev).values())
    torch.cuda.synchronize()
    logger.info(f"Frozen LLM weights stored as NF4 (blocksize 64, double quant): {len(errs)} matrices in "
                f"{time.perf_counter() - t0:.2f} s, relative error mean {sum(errs) / len(errs):.4f} max {max(errs):.4f}")


def main(argv=None):
    args = build_parser().parse_args(argv)
    accelerator = setup_accelerator_and_logging(args)
    if args.synthetic > 0:
        from .engine import PackedGemma, PackedSiglip
        from . import weights as Wt
        lc = {"1b": GEMMA3_1B, "4b": GEMMA3_4B, "qwen3_8b": QWEN3_8B}[args.synthetic_llm]
        vc, dev = SIGLIP_L, accelerator.device
        vision_encoder = PackedSiglip(Wt.random_siglip(vc, dev, 0), vc, dev)
        Wl = Wt.random_gemma(lc, dev, 0)
        if args.enable_qlora:
            _nf4(Wt, Wl, lc, dev)
        language_model = PackedGemma(Wl, lc, dev)
        vision_dim, llm_dim = vc.hidden, lc.hidden
        processor, tokenizer = None, _Tok(lc.pad_id)
        T = min(args.max_length, 128)
        train_dataset = SyntheticXrayTextDataset(args.synthetic, vc.image, T, lc.vocab, lc.pad_id)
        val_dataset = None
    else:
        if not (args.image_root and args.train_json):
            raise SystemExit("--image_root and --train_json are required (or use --synthetic N)")
        from transformers import AutoModel, AutoModelForCausalLM, AutoProcessor, AutoTokenizer
        processor = AutoProcessor.from_pretrained(args.vision_model_name)
        vision_encoder = AutoModel.from_pretrained(args.vision_model_name, torch_dtype=torch.bfloat16)
        tokenizer = AutoTokenizer.from_pretrained(args.llm_name, padding_side="left")
        if tokenizer.pad_token is None:
            tokenizer.pad_token = tokenizer.eos_token
        language_model = AutoModelForCausalLM.from_pretrained(args.llm_name, torch_dtype=torch.bfloat16)
        vision_dim = vision_encoder.config.vision_config.hidden_size
        llm_dim = language_model.config.hidden_size
        if args.enable_qlora:
            from .engine import PackedGemma
            from . import weights as Wt
            Wl, lc = Wt.llm_from_hf(language_model)
            _nf4(Wt, Wl, lc, accelerator.device)
            language_model = PackedGemma(Wl, lc, accelerator.device)
        full = XrayTextPairDataset(args.image_root, args.train_json, processor, tokenizer, args.img_size,
                                   max_length=args.max_length, image_root_2=args.image_root_2, gpu_preprocess=args.gpu_preprocess, gpu_decode=args.gpu_decode,
                                   cache_tokens=args.cache_tokens)
        train_dataset, val_dataset = full, None
        if args.val_json:
            val_dataset = XrayTextPairDataset(args.image_root, args.val_json, processor, tokenizer, args.img_size,
                                              max_length=args.max_length, image_root_2=args.image_root_2, gpu_preprocess=args.gpu_preprocess, gpu_decode=args.gpu_decode,
                                   cache_tokens=args.cache_tokens)
        elif args.train_val_split > 0:
            from sklearn.model_selection import train_test_split
            tr, va = train_test_split(full.samples, test_size=args.train_val_split, random_state=42)
            train_dataset = XrayTextPairDataset(args.image_root, None, processor, tokenizer, args.img_size,
                                                max_length=args.max_length, image_root_2=args.image_root_2, gpu_preprocess=args.gpu_preprocess, gpu_decode=args.gpu_decode,
                                   cache_tokens=args.cache_tokens)
            val_dataset = XrayTextPairDataset(args.image_root, None, processor, tokenizer, args.img_size,
                                              max_length=args.max_length, image_root_2=args.image_root_2, gpu_preprocess=args.gpu_preprocess, gpu_decode=args.gpu_decode,
                                   cache_tokens=args.cache_tokens)
            train_dataset.samples, val_dataset.samples = tr, va
        if len(train_dataset) == 0:
            logger.error("Training dataset is empty.")
            return
    projection_layer = MLPProjector(vision_dim, llm_dim)
    trainer = ProjectionTrainerStage1(
        accelerator=accelerator, vision_encoder=vision_encoder, language_model=language_model,
        projection_layer=projection_layer, processor=processor, tokenizer=tokenizer,
        train_dataset=train_dataset, val_dataset=val_dataset, output_dir=args.output_dir,
        batch_size=args.batch_size, learning_rate=args.learning_rate, weight_decay=args.weight_decay,
        num_epochs=args.num_epochs, gradient_accumulation_steps=args.gradient_accumulation_steps,
        warmup_ratio=args.warmup_ratio, wandb_project=args.wandb_project,
        save_every_n_epochs=args.save_every_n_epochs)
    try:
        trainer.train()
    except Exception as e:      # the reference logs and does not re-raise (:378-389)
        logger.error(f"Training failed: {e}", exc_info=True)
    finally:
        accelerator.end_training()


if __name__ == "__main__":
    main()
