"""Drop-in for ref:Stage2/train_vqa_stage2.py: same flags (ref lines 84-120), same projector loading
(`--stage1_projector_path` directory with projector_config.json + .safetensors/.bin), same dataset schema, the
same LoRA configuration under `--enable_qlora` (ref lines 225-250); the step runs on the B200 kernels.

    torchrun --nproc-per-node N -m projectiontrainer_b200.train_vqa_stage2 --image_root ... --train_json ... \
        --val_json ... --stage1_projector_path ... --unfreeze_projection_layer --llm_name google/gemma-3-1b-it
    python -m projectiontrainer_b200.train_vqa_stage2 --synthetic 64 --unfreeze_projection_layer --disable_wandb

`--enable_qlora` puts the decoder's projection weights through NF4 (weights.nf4_frozen_llm_) and trains LoRA adapters
(r=16, alpha=32, all seven projections; lora.py), `--resume_qlora_adapter_path` loads a peft adapter directory.
`--unfreeze_llm` and `--train_ve_first_epoch` are parsed and rejected: full LLM / vision fine-tuning is not built
(see vqa_trainer.py).  `--synthetic N` (an addition) trains on N random triplets with random-init SigLIP-L/16-384 and
Gemma3 weights."""
from __future__ import annotations

import argparse
import json
import logging
import os

import torch

from .accelerator_setup import setup_accelerator_and_logging
from .config import GEMMA3_1B, GEMMA3_4B, QWEN3_8B, SIGLIP_L
from .projectors import MLPProjector
from .vqa_dataset import SyntheticVQADataset, XrayVQADataset
from .vqa_trainer import VQATrainerStage2

logger = logging.getLogger(__name__)


def load_pretrained_projector(projector_path):
    """ref:Stage2/train_vqa_stage2.py:25-80: projector_config.json + the first .safetensors (preferred) or
    .bin weight file of the directory -> MLPProjector."""
    config_path = os.path.join(projector_path, "projector_config.json")
    if not os.path.exists(config_path):
        raise FileNotFoundError(f"Projector config file not found at {config_path}")
    if not os.path.isdir(projector_path):
        raise FileNotFoundError(f"Projector model directory not found at {projector_path}")
    with open(config_path, "r") as f:
        config = json.load(f)
    projector = MLPProjector(vision_dim=config["vision_dim"], llm_dim=config["llm_dim"])
    files = [f for f in os.listdir(projector_path)
             if f.endswith((".bin", ".safetensors")) and "optimizer" not in f and "scheduler" not in f]
    if not files:
        raise FileNotFoundError(f"No model weight file (.bin or .safetensors) found in {projector_path}")
    name = next((f for f in files if f.endswith(".safetensors")), files[0])
    path = os.path.join(projector_path, name)
    if path.endswith(".safetensors"):
        from safetensors.torch import load_file
        state = load_file(path, device="cpu")
    else:
        state = torch.load(path, map_location="cpu")
    projector.load_state_dict(state)
    logger.info(f"Successfully loaded pre-trained projector weights into {type(projector).__name__}.")
    return projector


def load_qlora_adapter(path):
    """ref:Stage2/train_vqa_stage2.py:206-223 (`PeftModel.from_pretrained(llm, path, is_trainable=True)`): the LoRA
    fields of adapter_config.json + the tensors of adapter_model.safetensors / adapter_model.bin, by peft key."""
    if not os.path.isdir(path):
        raise SystemExit(f"Provided adapter path is not a directory: {path}")
    with open(os.path.join(path, "adapter_config.json")) as f:
        c = json.load(f)
    cfg = {"r": c["r"], "lora_alpha": c["lora_alpha"], "target_modules": list(c["target_modules"]),
           "lora_dropout": c.get("lora_dropout", 0.0)}
    st = os.path.join(path, "adapter_model.safetensors")
    if os.path.exists(st):
        from safetensors.torch import load_file
        sd = load_file(st, device="cpu")
    else:
        sd = torch.load(os.path.join(path, "adapter_model.bin"), map_location="cpu")
    logger.info(f"Successfully loaded QLoRA adapter from {path}.")
    return cfg, sd


class _Tok:
    def __init__(self, pad, eos=1, side="right"):
        self.pad_token_id, self.eos_token_id, self.padding_side = pad, eos, side


def build_parser():
    p = argparse.ArgumentParser(description="Train Stage 2: VQA Fine-tuning (B200 build, frozen-LLM variant)")
    p.add_argument("--image_root", type=str, default=None)
    p.add_argument("--image_root_2", type=str, default=None)
    p.add_argument("--train_json", type=str, default=None)
    p.add_argument("--val_json", type=str, default=None)
    p.add_argument("--img_size", type=int, default=384)
    p.add_argument("--max_q_len", type=int, default=128)
    p.add_argument("--max_a_len", type=int, default=512)
    p.add_argument("--vision_model_name", type=str, default="StanfordAIMI/XraySigLIP__vit-l-16-siglip-384__webli")
    p.add_argument("--llm_name", type=str, default="Qwen/Qwen3-8B")
    p.add_argument("--stage1_projector_path", type=str, default=None)
    p.add_argument("--output_dir", type=str, default="./Stage2/trained_vqa_stage2")
    p.add_argument("--batch_size", type=int, default=1)
    p.add_argument("--learning_rate", type=float, default=1e-4)
    p.add_argument("--weight_decay", type=float, default=0.01)
    p.add_argument("--num_epochs", type=int, default=5)
    p.add_argument("--warmup_ratio", type=float, default=0.05)
    p.add_argument("--gradient_accumulation_steps", type=int, default=8)
    p.add_argument("--enable_qlora", action="store_true")
    p.add_argument("--unfreeze_projection_layer", action="store_true")
    p.add_argument("--unfreeze_llm", action="store_true")
    p.add_argument("--train_ve_first_epoch", action="store_true")
    p.add_argument("--resume_qlora_adapter_path", type=str, default=None)
    p.add_argument("--wandb_project", type=str, default="xray_vqa_training_stage2")
    p.add_argument("--wandb_run_name", type=str, default=None)
    p.add_argument("--disable_wandb", action="store_true")
    p.add_argument("--gpu_decode", action="store_true",
                   help="workers only read the image files; baseline JPEGs are decoded on the GPU (bit-exact with Pillow), "
                        "then resized / normalized there (implies --gpu_preprocess)")
    p.add_argument("--gpu_preprocess", action="store_true",
                   help="resize / rescale / normalize the decoded images on the GPU (Pillow-exact) instead of in the workers")
    p.add_argument("--synthetic", type=int, default=0, help="train on N synthetic triplets with random-init models")
    p.add_argument("--synthetic_llm", choices=["1b", "4b", "qwen3_8b"], default="1b")
    return p


def main(argv=None):
    args = build_parser().parse_args(argv)
    if args.unfreeze_llm or args.train_ve_first_epoch:
        raise SystemExit("--unfreeze_llm / --train_ve_first_epoch are not supported by this build (the LLM trains "
                         "through LoRA adapters only; no fallback to another backend)")
    if args.resume_qlora_adapter_path and not args.enable_qlora:
        raise SystemExit("--resume_qlora_adapter_path needs --enable_qlora")
    freeze_proj = not args.unfreeze_projection_layer
    # ref:Stage2/train_vqa_stage2.py:229-236
    lora_config = {"r": 16, "lora_alpha": 32, "lora_dropout": 0.05,
                   "target_modules": ["q_proj", "k_proj", "v_proj", "o_proj", "gate_proj", "up_proj", "down_proj"]}
    lora_sd = None
    if args.resume_qlora_adapter_path:
        lora_config, lora_sd = load_qlora_adapter(args.resume_qlora_adapter_path)
    accelerator = setup_accelerator_and_logging(args)
    if args.synthetic > 0:
        from . import weights as Wt
        from .engine import PackedGemma, PackedSiglip
        lc = {"1b": GEMMA3_1B, "4b": GEMMA3_4B, "qwen3_8b": QWEN3_8B}[args.synthetic_llm]
        vc, dev = SIGLIP_L, accelerator.device
        vision_encoder = PackedSiglip(Wt.random_siglip(vc, dev, 0), vc, dev)
        Wl = Wt.random_gemma(lc, dev, 0)
        if args.enable_qlora:
            Wt.nf4_frozen_llm_(Wl, lc, dev)
        language_model = PackedGemma(Wl, lc, dev)
        tokenizer = _Tok(lc.pad_id)
        projection_layer = MLPProjector(vc.hidden, lc.hidden)
        n_val = max(1, args.synthetic // 8)
        train_dataset = SyntheticVQADataset(args.synthetic, vc.image, lc.vocab)
        val_dataset = SyntheticVQADataset(n_val, vc.image, lc.vocab, seed=4321)
    else:
        if not (args.image_root and args.train_json and args.val_json and args.stage1_projector_path):
            raise SystemExit("--image_root, --train_json, --val_json and --stage1_projector_path are required "
                             "(or use --synthetic N)")
        from transformers import AutoModel, AutoModelForCausalLM, AutoProcessor, AutoTokenizer
        processor = AutoProcessor.from_pretrained(args.vision_model_name)
        vision_encoder = AutoModel.from_pretrained(args.vision_model_name, torch_dtype=torch.bfloat16)
        tokenizer = AutoTokenizer.from_pretrained(args.llm_name)
        language_model = AutoModelForCausalLM.from_pretrained(args.llm_name, torch_dtype=torch.bfloat16)
        if tokenizer.pad_token is None:
            tokenizer.pad_token = tokenizer.eos_token
        projection_layer = load_pretrained_projector(args.stage1_projector_path)
        kw = dict(processor=processor, tokenizer=tokenizer, img_size=args.img_size, max_q_len=args.max_q_len,
                  max_a_len=args.max_a_len, image_root_2=args.image_root_2, gpu_preprocess=args.gpu_preprocess,
                  gpu_decode=args.gpu_decode)
        train_dataset = XrayVQADataset(args.image_root, args.train_json, **kw)
        val_dataset = XrayVQADataset(args.image_root, args.val_json, **kw)
    trainer = VQATrainerStage2(
        accelerator=accelerator, vision_encoder=vision_encoder, language_model=language_model,
        projection_layer=projection_layer, tokenizer=tokenizer, train_dataset=train_dataset,
        val_dataset=val_dataset, output_dir=args.output_dir, batch_size=args.batch_size,
        learning_rate=args.learning_rate, weight_decay=args.weight_decay, num_epochs=args.num_epochs,
        gradient_accumulation_steps=args.gradient_accumulation_steps, warmup_ratio=args.warmup_ratio,
        freeze_vision_encoder=True, freeze_projection_layer=freeze_proj, freeze_llm=True,
        enable_qlora=args.enable_qlora, train_ve_first_epoch=False, wandb_project=args.wandb_project,
        lora_config=lora_config if args.enable_qlora else None, lora_state_dict=lora_sd)
    try:
        trainer.train()
    finally:
        accelerator.end_training()
    logger.info(f"Process {accelerator.process_index}: Stage 2 training script finished!")


if __name__ == "__main__":
    main()
