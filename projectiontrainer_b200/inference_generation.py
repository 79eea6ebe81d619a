"""Drop-in for ref:Stage2/inference_generation.py: the disease-identification evaluation loop — same flags, same
model loading (shared with `inference_vqa_stage2.load_models`), the fixed prompt of ref :59, the reference's generation
call (`do_sample=True, temperature=0.7, top_p=0.9, num_beams=1, max_new_tokens=--max_length`, ref :78-88), a sample
counted correct when its `normal_caption` occurs in the generated text (case-insensitive, ref :94), the per-sample log
line and the summary (ref :129-199: overall match rate, per-label accuracy, ground-truth distribution, candidate labels
found in the predictions).

The model arithmetic runs in `engine.Stage1Engine` (adapters folded into the decoder once before the loop).
`--synthetic`: random-init SigLIP-L + Gemma3-1B on generated images (no network / checkpoints needed)."""
from __future__ import annotations

import argparse
import json
import logging
import os
from collections import Counter

import torch

from . import inference_vqa_stage2 as V

logger = logging.getLogger(__name__)

PROMPT = "Identify the diseases in this chest X-ray image. Provide your answer in a single word or phrase."


def build_parser():
    p = argparse.ArgumentParser(description="Run Stage 2 Generation Inference on a JSON file")
    p.add_argument("--vision_model_name", type=str, default="StanfordAIMI/XraySigLIP__vit-l-16-siglip-384__webli")
    p.add_argument("--base_llm_name", type=str, default=None)
    p.add_argument("--adapter_path", type=str, default=None, help="QLoRA adapter directory (language_model folder)")
    p.add_argument("--projector_path", type=str, default=None, help="directory holding projector_best.bin")
    p.add_argument("--input_json", type=str, default=None, help="JSON list of {image, normal_caption}")
    p.add_argument("--image_root", type=str, default=None)
    p.add_argument("--max_length", type=int, default=128, help="Maximum length of generated text")
    p.add_argument("--verbose", action="store_true", help="Show detailed results for each image")
    p.add_argument("--candidate_labels", type=str, default=None, help="comma-separated candidate labels of this run")
    p.add_argument("--device", type=str, default="cuda" if torch.cuda.is_available() else "cpu")
    p.add_argument("--synthetic", action="store_true", help="random-init models, generated images")
    p.add_argument("--gpu_decode", action="store_true",
                   help="decode baseline JPEG files and resize on the GPU (same pixel values as Pillow; not a reference flag)")
    return p


def process_sample(sample, image_root, engine, image_size, tokenize, decode, pad_token_id, eos_token_id, max_length=128,
                   generator=None, gpu_decode=False):
    """ref:Stage2/inference_generation.py:24-107 on the engine: {generated_text, normal_caption, is_correct} or None."""
    image_rel_path = sample.get("image")
    normal_caption = sample.get("normal_caption", "")
    if not image_rel_path:
        logger.warning(f"Skipping sample due to missing 'image': {sample}")
        return None
    text = V.process_sample({"image": image_rel_path, "problem": PROMPT}, image_root, engine, 384 if image_size is None
                            else image_size, tokenize, decode,
                            dict(max_new_tokens=max_length, do_sample=True, temperature=0.7, top_p=0.9, num_beams=1),
                            pad_token_id, eos_token_id, generator=generator, gpu_decode=gpu_decode)
    if text is None:
        return None
    return {"generated_text": text, "normal_caption": normal_caption,
            "is_correct": normal_caption.lower() in text.lower() if normal_caption else False}


def summarise(all_results, candidate_labels) -> dict:
    """The numbers of the reference's summary (ref :129-199)."""
    total = len(all_results)
    truth, found_truth, predicted = Counter(), Counter(), Counter()
    for r in all_results:
        text = r.get("generated_text", "").lower()
        gt = r.get("normal_caption")
        if gt:
            truth[gt] += 1
            found_truth[gt] += gt.lower() in text
        for label in candidate_labels:
            predicted[label] += label.lower() in text
    return {"total": total, "correct": sum(1 for r in all_results if r["is_correct"]), "ground_truth": truth,
            "label_correct": found_truth, "predicted": Counter({k: v for k, v in predicted.items() if v})}


def display_summary(all_results, candidate_labels):
    s = summarise(all_results, candidate_labels)
    bar = "=" * 60
    print("\n" + bar + "\nSUMMARY OF RESULTS\n" + bar)
    total = s["total"]
    if total == 0:
        print("No results to display")
        return s
    print(f"Total images processed: {total}")
    print(f"Images classified as {', '.join(candidate_labels)}: {total} (100.00%)")
    print(f"\nImages with ground truth labels: {total}")
    print(f"Predictions matching ground truth (exact label string): {s['correct']} ({s['correct'] / total * 100:.2f}%)")
    by_count = [k for k, _ in s["ground_truth"].most_common()]
    print("\nClassification Accuracy distribution (Ground Truth label found in prediction):")
    for label in by_count:
        n, ok = s["ground_truth"][label], s["label_correct"][label]
        print(f"  {label:<20}: {ok}/{n} ({ok / n * 100 if n else 0:.2f}%)")
    print("\nGround Truth Distribution:")
    for label in by_count:
        print(f"  {label:<40}: {s['ground_truth'][label]} ({s['ground_truth'][label] / total * 100:.2f}%)")
    print("\nPrediction distribution (Candidate label found in prediction):")
    for label, n in s["predicted"].most_common():
        print(f"  {label:<20}: {n} ({n / total * 100:.2f}%)")
    print(bar)
    return s


def main(argv=None):
    args = build_parser().parse_args(argv)
    device = torch.device(args.device)
    if device.type != "cuda" or not torch.cuda.is_available():
        logger.error("This build runs on a B200 only (sm_100a kernels, no CPU path).")
        raise SystemExit(1)
    if args.synthetic:
        engine, data, image_size, tokenize, decode, pad_id, eos_id = V._synthetic(args, device)
        for i, d in enumerate(data):
            d["normal_caption"] = str(3 + i)
        args.max_length = min(args.max_length, 16)
        args.candidate_labels = args.candidate_labels or "3,4,5"
    else:
        for name in ("base_llm_name", "adapter_path", "projector_path", "input_json", "image_root", "candidate_labels"):
            if getattr(args, name) is None:
                logger.error(f"--{name} is required (or use --synthetic)")
                raise SystemExit(1)
        for path, what in ((args.adapter_path, "Adapter path"), (args.projector_path, "Projector path"),
                           (args.image_root, "Image root directory")):
            if not os.path.isdir(path):
                logger.error(f"{what} not found: {path}")
                raise SystemExit(1)
        if not os.path.exists(args.input_json):
            logger.error(f"Input JSON file not found: {args.input_json}")
            raise SystemExit(1)
        engine, image_size, tokenize, decode, pad_id, eos_id = V.load_models(args, device)
        with open(args.input_json, "r", encoding="utf-8") as f:
            data = json.load(f)
        logger.info(f"Loaded {len(data)} samples from {args.input_json}")
    labels = [label.strip() for label in args.candidate_labels.split(",")]
    logger.info(f"Using candidate labels for analysis: {labels}")
    all_results, skipped = [], 0
    with engine.merged(engine.merge_adapters()):
        for sample in data:
            result = process_sample(sample, args.image_root, engine, image_size, tokenize, decode, pad_id, eos_id,
                                    max_length=args.max_length, gpu_decode=args.gpu_decode)
            if result is None:
                skipped += 1
                continue
            all_results.append(result)
            logger.info(f"Image: {sample.get('image')}, Model Answer: {result['generated_text']}, "
                        f"Ground Truth: {result['normal_caption']}")
            if args.verbose:
                print(f"\nGeneration Results for {os.path.basename(sample.get('image'))}:\n" + "-" * 60)
                print(f"Generated Text: {result['generated_text'][:100]}")
                if result["normal_caption"]:
                    print(f"Ground Truth: {result['normal_caption']}\nCorrect: {'Yes' if result['is_correct'] else 'No'}")
                print("-" * 60)
    logger.info(f"Inference completed. Processed: {len(all_results)}, Skipped: {skipped}")
    if all_results:
        return display_summary(all_results, labels)
    logger.info("No samples were processed successfully.")
    return summarise([], labels)


if __name__ == "__main__":
    main()
