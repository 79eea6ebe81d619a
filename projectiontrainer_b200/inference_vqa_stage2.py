"""Drop-in for ref:Stage2/inference_vqa_stage2.py (the JSON-file loop, :211-468): same flags and defaults, same
loading order (processor, vision encoder, base LLM, tokenizer from the adapter directory with the base model as the
fall-back, QLoRA adapters — a missing / broken adapter directory is logged and the base LLM is used, as the reference
does —, `projector_best.bin`), same per-sample block (`process_sample`, ref :211-298): resolve the image path
(`image_root`, or `image_root_2` for "p.../s..." study directories: first .jpg inside), RGB, resize to the processor's
size, tokenise the question without special tokens, `[image][question]` -> `generate(eos, pad, **generation_args)` with
the reference's defaults (3 beams, sampling, temperature 0.3, top_p 0.9, top_k 50, repetition_penalty 1.8,
length_penalty 1.2) -> decode -> the same printed record.

The model arithmetic runs in `engine.Stage1Engine` (prefill + KV-cached decode on the sm_100a kernels; the adapters are
folded into the decoder weights once, before the loop: peft's `merge_and_unload`, which the reference mentions as the
faster option).  HF modules are read once as weight containers.  `--synthetic` runs random-init SigLIP-L + Gemma3-1B
with random adapters on generated images and questions (no network / checkpoints needed)."""
from __future__ import annotations

import argparse
import json
import logging
import os

import torch

from . import weights as Wt
from .config import GEMMA3_1B, SIGLIP_L
from .engine import PackedGemma, PackedSiglip, ProjectorState, Stage1Engine
from .lora import LoraState
from .projectors import MLPProjector
from .train_vqa_stage2 import load_qlora_adapter

logging.basicConfig(level=logging.INFO, format="%(asctime)s - %(levelname)s - %(message)s")
logger = logging.getLogger(__name__)


def build_parser():
    p = argparse.ArgumentParser(description="Run Stage 2 VQA Inference on a JSON file")
    p.add_argument("--vision_model_name", type=str, default="StanfordAIMI/XraySigLIP__vit-l-16-siglip-384__webli")
    p.add_argument("--base_llm_name", type=str, default=None, help="Name or path of the base LLM (e.g. 'Qwen/Qwen3-8B').")
    p.add_argument("--adapter_path", type=str, default=None, help="Path to the trained QLoRA adapter directory.")
    p.add_argument("--projector_path", type=str, default=None, help="Directory holding projector_best.bin")
    p.add_argument("--input_json", type=str, default=None, help="JSON list of {image, problem, normal_caption}")
    p.add_argument("--image_root", type=str, default=None)
    p.add_argument("--image_root_2", type=str, default=None, help="root for 'p10012261/s50349409'-style image entries")
    p.add_argument("--device", type=str, default="cuda" if torch.cuda.is_available() else "cpu")
    p.add_argument("--max_new_tokens", type=int, default=512)
    p.add_argument("--do_sample", type=bool, default=True)
    p.add_argument("--temperature", type=float, default=0.3)
    p.add_argument("--top_p", type=float, default=0.9)
    p.add_argument("--top_k", type=int, default=50)
    p.add_argument("--repetition_penalty", type=float, default=1.8)
    p.add_argument("--length_penalty", type=float, default=1.2)
    p.add_argument("--num_beams", type=int, default=3)
    p.add_argument("--synthetic", action="store_true", help="random-init models, generated images and questions")
    p.add_argument("--gpu_decode", action="store_true",
                   help="decode baseline JPEG files and resize on the GPU (same pixel values as Pillow; not a reference flag)")
    return p


def resolve_image_path(image_filename: str, image_root: str, image_root_2: str | None):
    """ref:Stage2/inference_vqa_stage2.py:222-241.  None when a study directory holds no .jpg."""
    if image_filename.startswith("p") and "/" in image_filename and image_root_2:
        path = os.path.join(image_root_2, image_filename)
        if os.path.isdir(path):
            jpgs = [f for f in os.listdir(path) if f.lower().endswith(".jpg")]
            if not jpgs:
                logger.warning(f"No .jpg files found in directory: {path}. Skipping sample.")
                return None
            path = os.path.join(path, jpgs[0])
        return path
    return os.path.join(image_root, image_filename)


def load_pixel_values(engine: Stage1Engine, image_path: str, image_size: int, gpu_decode: bool = False) -> torch.Tensor:
    """`Image.open(path).convert('RGB').resize((S, S))` + `processor(images=image)` (ref :243-247) -> pixel_values
    [1, 3, S, S] on the device.  The rescale / normalise always runs on the GPU (Pillow-exact).  With `gpu_decode`
    the host only reads the file: a baseline JPEG is decoded on the GPU (bit-exact with Pillow) and resized there with
    Pillow's bicubic arithmetic — the same pixel_values, without the ~50 ms per 2500 x 2048 chest X-ray that Pillow's
    decode + resize cost on a host core; other formats take Pillow as below."""
    import numpy as np
    from PIL import Image
    if gpu_decode and image_size == engine.vcfg.image:
        from . import image_io, ops
        item = image_io.load_image_item(image_path, True)
        if "jpeg_bytes" in item:
            rgb = ops.jpeg_decode(ops.jpeg_pack([item["jpeg_bytes"]], pin=True), engine.device)
            return engine.preprocess_images(rgb)
        return engine.preprocess_images([item["images_u8"]])
    image = Image.open(image_path).convert("RGB").resize((image_size, image_size))
    return engine.preprocess_images([torch.from_numpy(np.array(image, dtype=np.uint8))])


def process_sample(sample, image_root, engine: Stage1Engine, image_size: int, tokenize, decode, generation_args,
                   pad_token_id, eos_token_id, image_root_2=None, generator=None, gpu_decode: bool = False):
    """ref:Stage2/inference_vqa_stage2.py:211-298 on the engine.  `tokenize(text) -> int64 [1, n]` (no special
    tokens), `decode(ids) -> str`.  Returns the generated text, or None for a skipped / failed sample."""
    image_path = None
    try:
        image_filename, question_text = sample.get("image"), sample.get("problem")
        if not image_filename or not question_text:
            logger.warning(f"Skipping sample due to missing 'image' or 'problem': {sample}")
            return None
        image_path = resolve_image_path(image_filename, image_root, image_root_2)
        if image_path is None:
            return None
        pixel_values = load_pixel_values(engine, image_path, image_size, gpu_decode)
        question_tokens = tokenize(question_text).to(engine.device)
        ids = engine.generate(pixel_values, question_tokens, pad_token_id=pad_token_id, eos_token_id=eos_token_id,
                              generator=generator, **generation_args)
        return decode(ids[0])
    except FileNotFoundError:
        logger.error(f"Image file not found: {image_path}. Skipping sample.")
        return None
    except Exception as e:
        logger.error(f"Error processing sample for image {image_path}: {e}", exc_info=True)
        return None


def attach_adapters(engine: Stage1Engine, adapter_path: str) -> bool:
    """`PeftModel.from_pretrained(llm, adapter_path)` (ref :389-398): LoRA state from adapter_config.json +
    adapter_model.safetensors / .bin.  A failure is logged and the base LLM is kept, as the reference does."""
    try:
        cfg, sd = load_qlora_adapter(adapter_path)
        lora = LoraState(engine.llm.cfg, engine.device, r=int(cfg["r"]), alpha=float(cfg["lora_alpha"]),
                         targets=tuple(cfg["target_modules"]), dropout=float(cfg.get("lora_dropout", 0.0)))
        lora.load_state_dict(sd)
        engine.lora = lora
        logger.info(f"QLoRA adapters loaded from {adapter_path}")
        return True
    except (Exception, SystemExit) as e:
        logger.error(f"Failed to load PEFT adapters from {adapter_path}: {e}", exc_info=True)
        logger.warning("Proceeding with the base LLM without adapters.")
        return False


def load_models(args, device):
    """Processor, vision encoder, base LLM, tokenizer (adapter directory first), adapters, `projector_best.bin`
    (ref:Stage2/inference_vqa_stage2.py:347-414, ref:Stage2/inference_generation.py:249-311) -> engine, image size,
    tokenize(text) -> ids [1, n] without special tokens, decode(ids) -> text, pad id, eos id."""
    logger.info("Loading models and processor...")
    try:
        from transformers import AutoModel, AutoModelForCausalLM, AutoProcessor, AutoTokenizer
        processor = AutoProcessor.from_pretrained(args.vision_model_name)
        vision_encoder = AutoModel.from_pretrained(args.vision_model_name, torch_dtype=torch.bfloat16)
        vision_dim = vision_encoder.config.vision_config.hidden_size
        logger.info(f"Loading base LLM: {args.base_llm_name}")
        llm = AutoModelForCausalLM.from_pretrained(args.base_llm_name, torch_dtype=torch.bfloat16, trust_remote_code=True)
        try:
            tokenizer = AutoTokenizer.from_pretrained(args.adapter_path, trust_remote_code=True)
            logger.info(f"Tokenizer loaded from adapter path: {args.adapter_path}")
        except Exception:
            logger.warning(f"Could not load tokenizer from adapter path '{args.adapter_path}'. Loading from base "
                           f"model '{args.base_llm_name}'.")
            tokenizer = AutoTokenizer.from_pretrained(args.base_llm_name, trust_remote_code=True)
        tokenizer.padding_side = "left"
        if tokenizer.pad_token is None:
            tokenizer.pad_token = tokenizer.eos_token
            logger.warning("pad_token was None. Set to eos_token.")
        projection_layer = MLPProjector(vision_dim=vision_dim, llm_dim=llm.config.hidden_size)
        weights_path = os.path.join(args.projector_path, "projector_best.bin")
        if not os.path.exists(weights_path):
            raise FileNotFoundError(f"Projector weights not found at {weights_path}")
        projection_layer.load_state_dict(torch.load(weights_path, map_location="cpu"))
        engine = Stage1Engine(
            PackedSiglip(*Wt.siglip_from_hf(getattr(vision_encoder, "vision_model", vision_encoder)), device),
            PackedGemma(*Wt.llm_from_hf(llm), device),
            ProjectorState({k: v.detach().float() for k, v in projection_layer.state_dict().items()}, device), device)
        attach_adapters(engine, args.adapter_path)
        size = getattr(getattr(processor, "image_processor", processor), "size", {}) or {}
        image_size = size["height"] if "height" in size else 384
        logger.info("Models loaded successfully.")
    except SystemExit:
        raise
    except Exception as e:
        logger.error(f"Failed to load models: {e}", exc_info=True)
        raise SystemExit(1)
    tokenize = lambda text: tokenizer(text, return_tensors="pt", add_special_tokens=False).input_ids
    decode = lambda ids: tokenizer.decode(ids, skip_special_tokens=True)
    return engine, image_size, tokenize, decode, tokenizer.pad_token_id, tokenizer.eos_token_id


def run(engine: Stage1Engine, data, args, image_size, tokenize, decode, pad_token_id, eos_token_id, generator=None):
    generation_args = dict(max_new_tokens=args.max_new_tokens, do_sample=args.do_sample, temperature=args.temperature,
                           top_p=args.top_p, top_k=args.top_k, repetition_penalty=args.repetition_penalty,
                           length_penalty=args.length_penalty, num_beams=args.num_beams)
    predictions = []
    # the adapters are folded into the decoder once for the whole loop (generate() would do it per call)
    with engine.merged(engine.merge_adapters()):
        logger.info("Starting inference loop...")
        for i, sample in enumerate(data):
            logger.info(f"Processing sample {i + 1}/{len(data)}...")
            prediction = process_sample(sample, args.image_root, engine, image_size, tokenize, decode, generation_args,
                                        pad_token_id, eos_token_id, image_root_2=args.image_root_2, generator=generator,
                                        gpu_decode=getattr(args, "gpu_decode", False))
            predictions.append(prediction)
            if prediction is not None:
                print("\n" + "=" * 30 + f" Sample {i + 1} " + "=" * 30)
                print(f"Image Path: {os.path.join(args.image_root, sample.get('image'))}")
                print(f"Question:   {sample.get('problem')}")
                print(f"Prediction: {prediction}")
                print(f"Answer:     {sample.get('normal_caption', 'N/A')}")
                print("=" * 70)
    return predictions


def _synthetic(args, device):
    """Random-init SigLIP-L + Gemma3-1B + random adapters, three generated JPEGs and questions in a temp directory."""
    import tempfile

    import numpy as np
    from PIL import Image
    vc, lc = SIGLIP_L, GEMMA3_1B
    engine = Stage1Engine(PackedSiglip(Wt.random_siglip(vc, device, 0), vc, device),
                          PackedGemma(Wt.random_gemma(lc, device, 0), lc, device),
                          ProjectorState(Wt.random_projector(vc.hidden, lc.hidden, device), device), device)
    lora = LoraState(lc, device, r=16, alpha=32.0)
    g = torch.Generator(device=device).manual_seed(0)
    lora.flat.add_(torch.randn(lora.flat.shape, generator=g, device=device) * 0.01)
    lora.refresh_shadow()
    engine.lora = lora
    root = tempfile.mkdtemp(prefix="pt_vqa_")
    rng = np.random.default_rng(0)
    data = []
    for i in range(3):
        img = (rng.random((300 + 40 * i, 400, 3)) * 255).astype(np.uint8)
        Image.fromarray(img).save(os.path.join(root, f"img_{i}.jpg"), quality=90)
        data.append({"image": f"img_{i}.jpg", "problem": f"question {i} " + "what is shown " * (i + 1),
                     "normal_caption": "n/a"})
    args.image_root = root
    tokenize = lambda text: torch.tensor([[3 + (ord(ch) * 7919) % (lc.vocab - 3) for ch in text][:64]], dtype=torch.int64)
    decode = lambda ids: " ".join(str(int(t)) for t in ids)
    return engine, data, vc.image, tokenize, decode, lc.pad_id, 1


def main(argv=None):
    args = build_parser().parse_args(argv)
    device = torch.device(args.device)
    if device.type != "cuda" or not torch.cuda.is_available():
        logger.error("This build runs on a B200 only (sm_100a kernels, no CPU path).")
        raise SystemExit(1)
    logger.info(f"Using device: {device}")
    if args.synthetic:
        engine, data, image_size, tokenize, decode, pad_id, eos_id = _synthetic(args, device)
        args.max_new_tokens = min(args.max_new_tokens, 24)
    else:
        for name in ("base_llm_name", "adapter_path", "projector_path", "input_json", "image_root"):
            if getattr(args, name) is None:
                logger.error(f"--{name} is required (or use --synthetic)")
                raise SystemExit(1)
        for path, what in ((args.adapter_path, "Adapter path"), (args.projector_path, "Projector path")):
            if not os.path.isdir(path):
                logger.error(f"{what} not found: {path}")
                raise SystemExit(1)
        if not os.path.exists(args.input_json):
            logger.error(f"Input JSON file not found: {args.input_json}")
            raise SystemExit(1)
        if not os.path.isdir(args.image_root):
            logger.error(f"Image root directory not found: {args.image_root}")
            raise SystemExit(1)
        engine, image_size, tokenize, decode, pad_id, eos_id = load_models(args, device)
        try:
            with open(args.input_json, "r", encoding="utf-8") as f:
                data = json.load(f)
            logger.info(f"Loaded {len(data)} samples from {args.input_json}")
        except Exception as e:
            logger.error(f"Failed to load or parse JSON file {args.input_json}: {e}")
            raise SystemExit(1)
    return run(engine, data, args, image_size, tokenize, decode, pad_id, eos_id)


if __name__ == "__main__":
    main()
