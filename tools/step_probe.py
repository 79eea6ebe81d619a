"""GPU probe: engine vs CPU oracle at a tiny (or given) config, stage by stage."""
import sys, time, argparse
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import torch
from oracle import stage1_oracle as O
from projectiontrainer_b200 import engine as E
from projectiontrainer_b200.config import SiglipCfg, GemmaCfg

ap = argparse.ArgumentParser()
ap.add_argument("--size", default="tiny")
ap.add_argument("--B", type=int, default=3)
ap.add_argument("--T", type=int, default=12)
ap.add_argument("--pad", type=int, default=1)
args = ap.parse_args()
dev = torch.device("cuda:0")

if args.size == "tiny":
    vc = dict(hidden=64, ff=128, layers=2, heads=2, image=64, patch=16)
    lc = dict(vocab=512, hidden=64, ff=128, layers=6, hq=4, hkv=2, hd=16, window=8, q_scalar=16.0)
elif args.size == "mid":
    vc = dict(hidden=256, ff=512, layers=2, heads=4, image=128, patch=16)
    lc = dict(vocab=2048, hidden=256, ff=768, layers=6, hq=4, hkv=1, hd=64, window=32, q_scalar=64.0,
              rope_factor_global=8.0)
else:
    vc, lc = {}, {}
ovc, olc = O.SiglipCfg(**vc), O.GemmaCfg(**lc)
pvc, plc = SiglipCfg(**vc), GemmaCfg(**lc)
Wv, Wl = O.init_siglip(ovc), O.init_gemma(olc)
P = O.init_projector(ovc.hidden, olc.hidden)
pixels, ids, labels = O.synthetic_batch(args.B, args.T, ovc, olc, left_pad=bool(args.pad))

def q(t):  # what the engine actually holds: bf16-rounded matrices
    return t.to(torch.bfloat16).float()
Wv_q = {k: (q(v) if v.dim() == 2 and k != "pos" else v) for k, v in Wv.items()}
Wl_q = {k: (q(v) if v.dim() == 2 else v) for k, v in Wl.items()}

t0 = time.time()
ref = O.stage1_step(pixels, ids, labels, Wv_q, P, Wl_q, ovc, olc, want_intermediates=True, scale_dtype=torch.bfloat16)
print(f"oracle: loss={ref['loss']:.6f} n_valid={ref['n_valid']} ({time.time()-t0:.1f}s)", flush=True)

eng = E.Stage1Engine(E.PackedSiglip(Wv, pvc, dev), E.PackedGemma(Wl, plc, dev),
                     E.ProjectorState(P, dev), dev)

def cmp(name, a, b):
    a, b = a.float().cpu().flatten(), b.float().cpu().flatten()
    cos = torch.nn.functional.cosine_similarity(a, b, dim=0).item()
    print(f"  {name:28s} max|d|={((a-b).abs().max().item()):.4e} ref_max={b.abs().max().item():.3e} cos={cos:.6f}", flush=True)

B, T = ids.shape
S = eng.n_img + T
patch = eng.vision_forward(pixels.to(dev))
torch.cuda.synchronize()
cmp("patch features", patch, ref["patch"].reshape(-1, ovc.hidden))
x0 = eng.buf("x0", (B * S, plc.hidden), torch.float32)
Z, Hm = eng.projector_forward(patch, x0, rows_per_img=eng.n_img, out_batch_stride=S * plc.hidden)
am, ll, kv, tg = eng.splice(ids.to(dev), labels.to(dev), x0, plc.pad_id)
torch.cuda.synchronize()
cmp("spliced x0", x0, ref["spliced"].reshape(-1, olc.hidden))
print("  attention_mask exact:", bool((am.cpu().numpy() == ref["attention_mask"]).all()),
      " lm_labels exact:", bool((ll.cpu().numpy() == ref["lm_labels"]).all()))
loss = eng.llm_forward_loss(x0, kv, tg, B, S, T)
torch.cuda.synchronize()
cmp("final hidden", eng.saved["x_final"], ref["hidden"].reshape(-1, olc.hidden))
print(f"  loss engine={loss.item():.6f} oracle={ref['loss']:.6f} rel={(abs(loss.item()-ref['loss'])/abs(ref['loss'])):.3e}")
dx = eng.llm_backward()
eng.projector_backward(dx, patch, Z, Hm, B, S)
torch.cuda.synchronize()
for k, g in eng.proj.grads().items():
    cmp("grad " + k, g, ref["grads"][k])
