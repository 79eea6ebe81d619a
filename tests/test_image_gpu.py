"""Parity of the CUDA image preprocessing through the C ABI: the resized uint8 image
bit-exact with Pillow (the library the reference calls) and the f32 `pixel_values` bit-exact with the oracle's
rescale/normalize of it.  GPU only."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _run(img: np.ndarray, S: int, dev, src=None):
    from projectiontrainer_b200 import ops
    H, W, _ = img.shape
    src = torch.from_numpy(img).to(dev) if src is None else src
    ws = torch.empty(ops.image_preprocess_workspace(H, W, S, S), dtype=torch.uint8, device=dev)
    out = torch.empty(3, S, S, dtype=torch.float32, device=dev)
    u8 = torch.empty(S, S, 3, dtype=torch.uint8, device=dev)
    ops.image_preprocess(src, out, ws, dst_u8=u8)
    torch.cuda.synchronize()
    return out.cpu().numpy(), u8.cpu().numpy()


@pytest.mark.parametrize("shape", [(2500, 2048, 384), (2048, 2500, 384), (1000, 777, 384), (200, 300, 384),
                                   (384, 384, 384), (384, 500, 384), (501, 384, 384), (7, 5, 16), (1, 1, 8),
                                   (4000, 3000, 384), (33, 900, 64)])
def test_resize_normalize_bit_exact_vs_pillow_and_oracle(cuda_dev, shape):
    from PIL import Image
    from oracle import image_oracle as IO
    H, W, S = shape
    rng = np.random.default_rng(H * 31 + W)
    yy, xx = np.mgrid[0:H, 0:W]
    smooth = np.stack([yy * 255 // max(H - 1, 1), xx * 255 // max(W - 1, 1), (yy + xx) % 256], -1).astype(np.uint8)
    edges = np.zeros((H, W, 3), dtype=np.uint8)
    edges[::7] = 255
    edges[:, ::5] = 255
    for img in (rng.integers(0, 256, (H, W, 3), dtype=np.uint8), smooth, edges):
        out, u8 = _run(img, S, cuda_dev)
        ref_u8 = np.asarray(Image.fromarray(img).resize((S, S)))
        assert np.array_equal(u8, ref_u8)                                 # byte work: bit-exact
        assert np.array_equal(out, IO.rescale_normalize(ref_u8))          # f32, same operation order: exact


def test_strided_misaligned_source_rows(cuda_dev):
    """A crop of a wider image: row stride != 3 W and rows starting at every byte alignment."""
    from PIL import Image
    big = np.random.default_rng(3).integers(0, 256, (900, 1203, 3), dtype=np.uint8)
    full = torch.from_numpy(big).to(cuda_dev)
    for x0 in (0, 1, 2, 3, 5):
        crop = full[10:810, x0:x0 + 1001]
        out, u8 = _run(big[10:810, x0:x0 + 1001], 384, cuda_dev, src=crop)
        assert np.array_equal(u8, np.asarray(Image.fromarray(big[10:810, x0:x0 + 1001]).resize((384, 384)))), x0


def test_bad_arguments_fail_loudly(cuda_dev):
    from projectiontrainer_b200 import ops
    src = torch.zeros(16, 16, 3, dtype=torch.uint8, device=cuda_dev)
    out = torch.empty(3, 8, 8, dtype=torch.float32, device=cuda_dev)
    small = torch.empty(16, dtype=torch.uint8, device=cuda_dev)
    with pytest.raises(RuntimeError, match="workspace too small"):
        ops.image_preprocess(src, out, small)
    with pytest.raises(RuntimeError):
        ops.image_preprocess(src.float(), out, small)
    with pytest.raises(RuntimeError, match="CUDA"):
        ops.image_preprocess(src.cpu(), out, small)


def test_engine_preprocess_images_feeds_the_step(cuda_dev):
    """`Stage1Engine.preprocess_images` on ragged host images == the dataset's PIL + processor path, and the
    batch dict with `images_u8` trains like the one with `pixel_values`."""
    from PIL import Image
    from oracle import image_oracle as IO
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    vkw = dict(hidden=64, ff=128, layers=1, heads=1, image=64, patch=16)
    lkw = dict(vocab=512, hidden=128, ff=128, layers=6, hq=2, hkv=1, hd=256, window=8, q_scalar=256.0)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    rng = np.random.default_rng(9)
    raws = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in ((300, 200), (64, 64), (129, 500))]
    ref = np.stack([IO.rescale_normalize(np.asarray(Image.fromarray(r).resize((64, 64)))) for r in raws])
    _, ids, lab = O.synthetic_batch(3, 10, ovc, olc, left_pad=True)

    def run(batch):
        eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev),
                             E.ProjectorState(P, cuda_dev), cuda_dev)
        losses = [eng.train_step(eng.stage_batch(batch), 1e-3, pad_id=olc.pad_id).item() for _ in range(3)]
        return eng, losses

    eng, l_raw = run({"images_u8": [torch.from_numpy(r) for r in raws], "token_ids": ids, "labels": lab})
    px = eng.preprocess_images([torch.from_numpy(r) for r in raws])
    torch.cuda.synchronize()
    assert np.array_equal(px.cpu().numpy(), ref)
    _, l_px = run({"pixel_values": torch.from_numpy(ref), "token_ids": ids, "labels": lab})
    assert np.allclose(l_raw, l_px, rtol=1e-6)
