"""Drop-in for ref:Stage1/inference_stage1.py: same flags, same checkpoint lookup
(`projector_config.json` + `model.safetensors` / `pytorch_model.bin`, plus this trainer's own
`projector_{final,best}.bin`), same key fix-ups (`module.` prefix removed, `model.` prefix added), and the
same inference block (ref lines 150-193): vision tower -> `[:, 1:, :]` -> projector -> all-ones mask ->
`generate(inputs_embeds=projected, ...)` -> decode.  The model arithmetic runs in `engine.Stage1Engine`
(prefill + KV-cached decode on the sm_100a kernels); HF modules are read once as weight containers.
`--synthetic` runs random-init SigLIP-L + Gemma3-1B on a random image (no network / checkpoints needed)."""
from __future__ import annotations

import argparse
import json
import logging
import os

import torch

from . import weights as Wt
from .config import GEMMA3_1B, SIGLIP_L
from .engine import PackedGemma, PackedSiglip, ProjectorState, Stage1Engine
from .projectors import MLPProjector

logging.basicConfig(level=logging.INFO, format="%(asctime)s - %(levelname)s - %(message)s")
logger = logging.getLogger(__name__)


def parse_args(argv=None):
    p = argparse.ArgumentParser(description="Inference using Stage 1 aligned projector.")
    p.add_argument("--vision_model_name", type=str, default="StanfordAIMI/XraySigLIP__vit-b-16-siglip-512__webli")
    p.add_argument("--llm_name", type=str, default="deepseek-ai/DeepSeek-R1-Distill-Qwen-1.5B")
    p.add_argument("--projection_path", type=str, required=False, default=None)
    p.add_argument("--image_path", type=str, required=False, default=None)
    p.add_argument("--device", type=str, default="cuda:0")
    p.add_argument("--max_new_tokens", type=int, default=256)
    p.add_argument("--temperature", type=float, default=0.7)
    p.add_argument("--top_p", type=float, default=0.9)
    p.add_argument("--repetition_penalty", type=float, default=1.0)
    p.add_argument("--do_sample", type=bool, default=True)
    p.add_argument("--synthetic", action="store_true", help="random-init SigLIP-L + Gemma3-1B, random image")
    p.add_argument("--gpu_decode", action="store_true",
                   help="decode a baseline JPEG and resize on the GPU (same pixel values as Pillow + the processor; not a "
                        "reference flag)")
    return p.parse_args(argv)


def load_projector_state(projection_path: str) -> tuple[dict, dict]:
    """ref:Stage1/inference_stage1.py:82-122 (config + weights + key fix-ups)."""
    cfg_path = os.path.join(projection_path, "projector_config.json")
    if not os.path.exists(cfg_path):
        raise FileNotFoundError(f"Projector config (projector_config.json) not found in {projection_path}")
    with open(cfg_path) as f:
        cfg = json.load(f)
    sd = None
    st = os.path.join(projection_path, "model.safetensors")
    if os.path.exists(st):
        from safetensors.torch import load_file
        sd = load_file(st, device="cpu")
    else:
        for name in ("pytorch_model.bin", "projector_final.bin", "projector_best.bin"):
            pth = os.path.join(projection_path, name)
            if os.path.exists(pth):
                sd = torch.load(pth, map_location="cpu")
                logger.info(f"Loaded projector weights from {pth}")
                break
    if sd is None:
        raise FileNotFoundError(f"Projector weights (model.safetensors or pytorch_model.bin) not found in "
                                f"{projection_path}")
    if all(k.startswith("module.") for k in sd):
        sd = {k.replace("module.", "", 1): v for k, v in sd.items()}
    if any(not k.startswith("model.") for k in sd):
        sd = {"model." + k: v for k, v in sd.items()}
    return cfg, sd


def build_engine_from_hf(vision_encoder, llm_model, projection_layer: MLPProjector, device) -> Stage1Engine:
    vis = PackedSiglip(*Wt.siglip_from_hf(getattr(vision_encoder, "vision_model", vision_encoder)), device)
    llm = PackedGemma(*Wt.llm_from_hf(llm_model), device)
    proj = ProjectorState({k: v.detach().float() for k, v in projection_layer.state_dict().items()}, device)
    return Stage1Engine(vis, llm, proj, device)


def run_inference(engine: Stage1Engine, pixel_values: torch.Tensor, args, pad_token_id, eos_token_id):
    """ref:Stage1/inference_stage1.py:150-193 on the engine: returns the generated ids [B, n]."""
    return engine.generate(pixel_values, None, max_new_tokens=args.max_new_tokens, do_sample=args.do_sample,
                           temperature=args.temperature, top_p=args.top_p,
                           repetition_penalty=args.repetition_penalty, pad_token_id=pad_token_id,
                           eos_token_id=eos_token_id)


def main(argv=None):
    args = parse_args(argv)
    device = torch.device(args.device)
    if device.type != "cuda" or not torch.cuda.is_available():
        logger.error("This build runs on a B200 only (sm_100a kernels, no CPU path).")
        raise SystemExit(1)
    logger.info(f"Using device: {device}")
    tokenizer = None
    if args.synthetic:
        vc, lc = SIGLIP_L, GEMMA3_1B
        vis = PackedSiglip(Wt.random_siglip(vc, device, 0), vc, device)
        llm = PackedGemma(Wt.random_gemma(lc, device, 0), lc, device)
        if args.projection_path:
            _, sd = load_projector_state(args.projection_path)
        else:
            sd = Wt.random_projector(vc.hidden, lc.hidden, device)
        engine = Stage1Engine(vis, llm, ProjectorState(sd, device), device)
        pixel_values = torch.rand(1, 3, vc.image, vc.image, device=device) * 2 - 1
        pad_id, eos_id = lc.pad_id, 1
    else:
        if not (args.projection_path and args.image_path):
            logger.error("--projection_path and --image_path are required (or use --synthetic)")
            raise SystemExit(1)
        from PIL import Image
        from transformers import AutoModel, AutoModelForCausalLM, AutoProcessor, AutoTokenizer
        try:
            processor = AutoProcessor.from_pretrained(args.vision_model_name)
            tokenizer = AutoTokenizer.from_pretrained(args.llm_name)
            vision_encoder = AutoModel.from_pretrained(args.vision_model_name, torch_dtype=torch.bfloat16)
            llm_model = AutoModelForCausalLM.from_pretrained(args.llm_name, torch_dtype=torch.bfloat16)
        except Exception as e:
            logger.error(f"Failed to load base models: {e}", exc_info=True)
            raise SystemExit(1)
        if tokenizer.pad_token is None:
            tokenizer.pad_token = tokenizer.eos_token
        try:
            cfg, sd = load_projector_state(args.projection_path)
            projection_layer = MLPProjector(vision_dim=cfg["vision_dim"], llm_dim=cfg["llm_dim"])
            projection_layer.load_state_dict(sd)
        except FileNotFoundError as e:
            logger.error(f"Error loading projection model: {e}")
            raise SystemExit(1)
        engine = build_engine_from_hf(vision_encoder, llm_model, projection_layer, device)
        try:
            if args.gpu_decode:     # file bytes -> GPU decode + the processor's bicubic resize / rescale / normalise
                from .inference_vqa_stage2 import load_pixel_values
                pixel_values = load_pixel_values(engine, args.image_path, engine.vcfg.image, gpu_decode=True)
            else:
                image = Image.open(args.image_path).convert("RGB")
                pixel_values = processor(images=image, return_tensors="pt").pixel_values
        except FileNotFoundError:
            logger.error(f"Image file not found: {args.image_path}")
            raise SystemExit(1)
        pad_id, eos_id = tokenizer.pad_token_id, tokenizer.eos_token_id
    generated_text = "[Inference Error]"
    try:
        ids = run_inference(engine, pixel_values, args, pad_id, eos_id)
        torch.cuda.synchronize(device)
        generated_text = tokenizer.decode(ids[0], skip_special_tokens=True) if tokenizer is not None \
            else " ".join(str(int(t)) for t in ids[0])
    except Exception as e:      # the reference logs and prints the error marker (:188-189)
        logger.error(f"Error during model inference/generation: {e}", exc_info=True)
    print("\n--- Generated Text (Stage 1 Alignment Test) ---")
    print(f"Image: {args.image_path}")
    print(f"Generated Description: {generated_text}")
    print("--- End ---")
    return generated_text


if __name__ == "__main__":
    main()
