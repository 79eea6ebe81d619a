"""GPU baseline-JPEG decode (decode half) through the C ABI against Pillow — the library the
reference's datasets call (`Image.open(path).convert("RGB")`, ref:Stage1/train_projection_stage1.py:99): every stage is
integer work with libjpeg's constants, so the bar is bit-exact.  Files are produced here with Pillow's encoder."""
import io

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _encode(arr, **kw):
    from PIL import Image, ImageFile
    ImageFile.MAXBLOCK = max(ImageFile.MAXBLOCK, 1 << 24)      # optimize=True needs the whole file in one buffer
    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


def _pillow(data):
    from PIL import Image
    return np.asarray(Image.open(io.BytesIO(data)).convert("RGB"))


def _images(H, W, rng):
    yy, xx = np.mgrid[0:H, 0:W]
    smooth = np.stack([yy * 255 // max(H - 1, 1), xx * 255 // max(W - 1, 1), (yy * 3 + xx * 5) % 256], -1).astype(np.uint8)
    noise = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
    edges = np.zeros((H, W, 3), dtype=np.uint8)
    edges[::7, :, 0] = 255
    edges[:, ::5, 1] = 255
    edges[H // 3:, W // 2:, 2] = 200
    return smooth, noise, edges


@pytest.mark.parametrize("H,W", [(64, 64), (40, 56), (33, 17), (1, 1), (2, 3), (8, 8), (17, 16), (250, 301), (16, 2)])
@pytest.mark.parametrize("subsampling", [0, 1, 2, "gray"])
def test_jpeg_decode_bit_exact_vs_pillow(cuda_dev, H, W, subsampling):
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(H * 131 + W)
    files = []
    for img in _images(H, W, rng):
        for q in (35, 90, 100):
            if subsampling == "gray":
                files.append(_encode(img[..., 0], quality=q))
            else:
                files.append(_encode(img, quality=q, subsampling=subsampling))
    out = ops.jpeg_decode(ops.jpeg_pack(files), cuda_dev)
    torch.cuda.synchronize()
    for i, (f, o) in enumerate(zip(files, out)):
        ref = _pillow(f)
        got = o.cpu().numpy()
        assert got.shape == ref.shape, (i, got.shape, ref.shape)
        bad = int((got != ref).sum())
        assert bad == 0, (i, bad, int(np.abs(got.astype(int) - ref.astype(int)).max()))


@pytest.mark.parametrize("kw", [dict(restart_marker_blocks=1), dict(restart_marker_blocks=7), dict(restart_marker_rows=1),
                                dict(optimize=True), dict(optimize=True, restart_marker_blocks=3)])
@pytest.mark.parametrize("subsampling", [0, 2, "gray"])
def test_jpeg_restart_intervals_and_optimised_tables(cuda_dev, kw, subsampling):
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(7)
    files = []
    for H, W in [(100, 130), (37, 259)]:
        for img in _images(H, W, rng):
            files.append(_encode(img[..., 1], quality=80, **kw) if subsampling == "gray"
                         else _encode(img, quality=80, subsampling=subsampling, **kw))
    packed = ops.jpeg_pack(files)
    out = ops.jpeg_decode(packed, cuda_dev)
    torch.cuda.synchronize()
    for f, o in zip(files, out):
        assert np.array_equal(o.cpu().numpy(), _pillow(f))


def test_jpeg_chest_xray_sized_batch_and_preprocess(cuda_dev):
    """A batch of 2500 x 2048 grayscale files (the dataset's image size) decoded side by side, then through
    pt_image_preprocess: `pixel_values` equal to what Pillow decode + Pillow resize + the processor give."""
    from PIL import Image
    from oracle import image_oracle as IO
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(3)
    H, W, S = 2500, 2048, 384
    yy, xx = np.mgrid[0:H, 0:W]
    base = (128 + 100 * np.sin(yy / 97.0) * np.cos(xx / 61.0)).astype(np.float32)
    files = [_encode(np.clip(base + rng.normal(0, 12 * (i + 1), (H, W)), 0, 255).astype(np.uint8), quality=92 - 10 * i)
             for i in range(3)]
    out = ops.jpeg_decode(ops.jpeg_pack(files), cuda_dev)
    torch.cuda.synchronize()
    for f, o in zip(files, out):
        ref = _pillow(f)
        assert np.array_equal(o.cpu().numpy(), ref)
        ws = torch.empty(ops.image_preprocess_workspace(H, W, S, S), dtype=torch.uint8, device=cuda_dev)
        px = torch.empty(3, S, S, device=cuda_dev)
        ops.image_preprocess(o, px, ws)
        ref_px = IO.rescale_normalize(np.asarray(Image.fromarray(ref).resize((S, S))))
        assert np.array_equal(px.cpu().numpy(), ref_px)


def test_jpeg_unsupported_files_are_rejected(cuda_dev):
    from PIL import Image
    from projectiontrainer_b200 import ops
    img = np.random.default_rng(0).integers(0, 256, (32, 32, 3), dtype=np.uint8)
    with pytest.raises(RuntimeError, match="progressive"):
        ops.jpeg_pack([_encode(img, progressive=True)])
    with pytest.raises(RuntimeError, match="not a JPEG"):
        ops.jpeg_pack([b"\x89PNG\r\n\x1a\n" + bytes(64)])
    b = io.BytesIO()
    Image.fromarray(np.dstack([img, img[..., :1]])).convert("CMYK").save(b, "JPEG")
    with pytest.raises(RuntimeError, match="components"):
        ops.jpeg_pack([b.getvalue()])
    with pytest.raises(RuntimeError, match="truncated|restart segments|bad segment"):
        ops.jpeg_pack([_encode(img, restart_marker_blocks=1)[:400]])


@pytest.mark.parametrize("subsampling", [0, 1, 2, "gray"])
@pytest.mark.parametrize("kw", [dict(), dict(restart_marker_rows=8), dict(optimize=True)])
def test_jpeg_long_segments_decode_in_parallel_subsequences(cuda_dev, subsampling, kw):
    """Entropy-coded segments of 8 KB and more take the self-synchronising parallel parse (sub-sequences of 1024 bits,
    several CTAs per segment, boundary stitching, DC running sums): same bytes as Pillow, for detailed and for flat
    images (long and short codes, few and many blocks per sub-sequence), mixed in one batch with short files."""
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(11)
    files = []
    for H, W in [(700, 900), (333, 1501)]:
        smooth, noise, edges = _images(H, W, rng)
        flat = np.full((H, W, 3), 77, dtype=np.uint8)
        flat[H // 2, W // 2] = 255
        for img, q in ((noise, 95), (smooth, 60), (edges, 85), (flat, 90), (noise // 4 + smooth // 2, 75)):
            files.append(_encode(img[..., 0], quality=q, **kw) if subsampling == "gray"
                         else _encode(img, quality=q, subsampling=subsampling, **kw))
    files.append(_encode(_images(16, 24, rng)[1], quality=90))              # a short one beside them
    packed = ops.jpeg_pack(files)
    assert sum(len(f) >= 9000 for f in files) >= 6
    out = ops.jpeg_decode(packed, cuda_dev)
    torch.cuda.synchronize()
    for i, (f, o) in enumerate(zip(files, out)):
        ref = _pillow(f)
        got = o.cpu().numpy()
        assert np.array_equal(got, ref), (i, len(f), int((got != ref).sum()))


def test_gpu_decode_loader_feeds_the_step(cuda_dev, tmp_path):
    """`gpu_decode=True` end to end: files on disk -> dataset (workers read bytes) -> collate (pack in the worker) ->
    pinned DataLoader -> `stage_batch` (JPEG decode + resize + normalize on the copy stream) -> the step.  The
    `pixel_values` the step sees equal the reference pipeline's (Pillow decode, Pillow resize, processor arithmetic)
    bit for bit, a PNG and a progressive file in the same batch included, and the losses equal those of a batch that
    carries those `pixel_values`."""
    import json
    from PIL import Image
    from oracle import image_oracle as IO
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    from projectiontrainer_b200.projector_trainer import stage1_collate
    from projectiontrainer_b200.train_projection_stage1 import XrayTextPairDataset
    vkw = dict(hidden=64, ff=128, layers=1, heads=1, image=64, patch=16)
    lkw = dict(vocab=512, hidden=128, ff=128, layers=6, hq=2, hkv=1, hd=256, window=8, q_scalar=256.0)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    Wv, Wl, P = O.init_siglip(ovc), O.init_gemma(olc), O.init_projector(ovc.hidden, olc.hidden)
    rng = np.random.default_rng(5)
    specs = [((300, 200), "JPEG", dict(quality=90, subsampling=2)), ((64, 64), "PNG", {}),
             ((129, 500), "JPEG", dict(quality=75, subsampling=0, restart_marker_rows=2)),
             ((700, 640), "JPEG", dict(quality=95)), ((90, 70), "JPEG", dict(progressive=True)),
             ((211, 97), "JPEG", dict(quality=85, subsampling=1))]
    names = []
    for i, ((h, w), fmt, kw) in enumerate(specs):
        arr = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        if i == 3:
            arr = arr[..., 0]                                   # a grayscale file, long enough for the parallel parse
        name = f"x{i}.{'png' if fmt == 'PNG' else 'jpg'}"
        Image.fromarray(arr).save(tmp_path / name, fmt, **kw)
        names.append(name)
    (tmp_path / "d.json").write_text(json.dumps([{"image": n, "normal_caption": "c"} for n in names]))
    ref = np.stack([IO.rescale_normalize(np.asarray(Image.open(tmp_path / n).convert("RGB").resize((64, 64)))) for n in names])
    _, ids, lab = O.synthetic_batch(len(names), 10, ovc, olc, left_pad=True)

    class Tok:
        pad_token_id = olc.pad_id
        def __call__(self, text, max_length, padding, truncation, return_tensors):
            Tok.i = getattr(Tok, "i", -1) + 1
            return type("Enc", (), {"input_ids": ids[Tok.i % len(names)][None]})()

    ds = XrayTextPairDataset(str(tmp_path), str(tmp_path / "d.json"), None, Tok(), 64, max_length=10, gpu_decode=True)
    loader = torch.utils.data.DataLoader(ds, batch_size=len(names), shuffle=False, num_workers=2, pin_memory=True,
                                         collate_fn=stage1_collate)
    batch = next(iter(loader))
    assert batch["jpeg"]["index"] == [0, 2, 3, 5] and batch["jpeg"]["blob"].is_pinned()
    batch["token_ids"], batch["labels"] = ids, lab              # (worker-side tokenizer call order is not the point)

    def run(b):
        eng = E.Stage1Engine(E.PackedSiglip(Wv, SiglipCfg(**vkw), cuda_dev), E.PackedGemma(Wl, GemmaCfg(**lkw), cuda_dev),
                             E.ProjectorState(P, cuda_dev), cuda_dev)
        staged = eng.stage_batch(b)
        if "_staged" in staged:
            torch.cuda.current_stream().wait_event(staged["_staged"]["event"])      # the copy stream's work
        px = staged["pixel_values"].clone()
        losses = [eng.train_step(staged, 1e-3, pad_id=olc.pad_id).item()]
        losses += [eng.train_step(eng.stage_batch(b), 1e-3, pad_id=olc.pad_id).item() for _ in range(2)]
        return px, losses

    px, l_jpeg = run(batch)
    torch.cuda.synchronize()
    assert np.array_equal(px.cpu().numpy(), ref)
    _, l_px = run({"pixel_values": torch.from_numpy(ref), "token_ids": ids, "labels": lab})
    assert np.allclose(l_jpeg, l_px, rtol=1e-6)


def test_jpeg_corrupt_entropy_data_does_not_fault(cuda_dev):
    """A file whose entropy-coded bytes are damaged (bit flips, zeroed runs, a truncated tail padded to the old length)
    still parses on the host; the device decode must stay inside its buffers — garbage pixels, no fault — and a clean
    file decoded right after it is still bit-exact."""
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(21)
    img = _images(400, 600, rng)[1]
    good = [_encode(img, quality=92, subsampling=2), _encode(img[..., 0], quality=92), _encode(img, quality=80, restart_marker_rows=1)]
    bad = []
    for f in good:
        b = bytearray(f)
        sos = b.index(b"\xff\xda")
        lo, hi = sos + 16, len(b) - 2
        for pos in rng.integers(lo, hi, 200):
            if b[pos] != 0xFF and b[pos - 1] != 0xFF:
                b[pos] ^= 1 << int(rng.integers(0, 8))
        b[lo + (hi - lo) // 2: lo + (hi - lo) // 2 + 3000] = bytes(3000)
        bad.append(bytes(b))
    out = ops.jpeg_decode(ops.jpeg_pack(bad + good), cuda_dev)
    torch.cuda.synchronize()
    for f, o in zip(good, out[3:]):
        assert np.array_equal(o.cpu().numpy(), _pillow(f))
    assert all(o.shape == (400, 600, 3) for o in out[:3])
