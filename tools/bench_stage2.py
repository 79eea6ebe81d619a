"""GPU probe: the Stage-2 frozen-LLM VQA micro-batch on SigLIP-L/16-384 + Gemma3-1B,
synthetic ragged triplets collated / bucketed by the host code of vqa_trainer.py, through train_step with
accumulation (GA=2) and staged host batches; CUDA-event timing, one JSON line."""
import sys, json, argparse
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from projectiontrainer_b200.config import GEMMA3_1B, SIGLIP_L
from projectiontrainer_b200.engine import build_engine
from projectiontrainer_b200.vqa_dataset import SyntheticVQADataset
from projectiontrainer_b200.vqa_trainer import accumulate_sync_flag, vqa_collate_fn, vqa_step_inputs

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=32)
ap.add_argument("--micro", type=int, default=12)
ap.add_argument("--bucket", type=int, default=32)
a = ap.parse_args()
dev = torch.device("cuda:0")
eng = build_engine(SIGLIP_L, GEMMA3_1B, dev, seed=0, grad_accum=4)


class Tok:
    pad_token_id, eos_token_id, padding_side = GEMMA3_1B.pad_id, 1, "right"


ds = SyntheticVQADataset(a.batch * 4, SIGLIP_L.image, GEMMA3_1B.vocab)
batches = []
for i in range(4):
    cb = vqa_collate_fn([ds[j] for j in range(i * a.batch, (i + 1) * a.batch)], Tok())
    cb["pixel_values"] = cb["pixel_values"].pin_memory()
    batches.append(vqa_step_inputs(cb, Tok.pad_token_id, a.bucket))
shapes = [tuple(b["token_ids"].shape) + tuple(b["label_span"]) for b in batches]


def run(n):
    micro = 0
    nxt = eng.stage_batch(batches[0])
    for i in range(n):
        cur = nxt
        if i + 1 < n:
            nxt = eng.stage_batch(batches[(i + 1) % 4])
        sync, micro = accumulate_sync_flag(micro, 2, i == n - 1)
        loss = eng.train_step(cur, 1e-4, pad_id=Tok.pad_token_id, max_norm=1.0, sync_gradients=sync)
    return loss


run(12)          # every geometry: eager, capture, replay
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
loss = run(a.micro)
e1.record()
torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / a.micro
print(json.dumps({"config": f"Stage-2 frozen-LLM VQA micro-batch, SigLIP-L + Gemma3-1B, B={a.batch}, GA=2, ragged "
                  f"q 8-48 / a 16-96 tokens, bucket {a.bucket}, host batches staged", "text_shapes(T,_,Lq,La)": shapes,
                  "ms_per_micro_batch": ms, "images_per_s": a.batch / ms * 1e3, "loss": float(loss.item()),
                  "geometries": len(eng._arenas) + 1}), flush=True)
