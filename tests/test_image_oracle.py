"""Pins oracle/image_oracle.py (numpy restatement of Pillow's antialiased bicubic resize + HF SigLIP
rescale/normalize) against the libraries the reference calls (ref:Stage1/train_projection_stage1.py:99-104,
ref:Stage2/dataset.py:90-95): Pillow itself, bit-exact, and HF's PIL-backed SiglipImageProcessor.  CPU only."""
import numpy as np
import pytest

from oracle import image_oracle as IO

PIL = pytest.importorskip("PIL.Image")


def _noise(h, w, seed=0):
    return np.random.default_rng(seed).integers(0, 256, (h, w, 3), dtype=np.uint8)


def _smooth(h, w):
    yy, xx = np.mgrid[0:h, 0:w]
    return np.stack([yy * 255 // max(h - 1, 1), xx * 255 // max(w - 1, 1), (yy + xx) % 256], -1).astype(np.uint8)


@pytest.mark.parametrize("shape", [(512, 640, 384, 384), (1000, 777, 384, 384), (200, 300, 384, 384),
                                   (384, 500, 384, 384), (500, 384, 384, 384), (384, 384, 384, 384),
                                   (7, 5, 16, 16), (33, 900, 64, 48), (3, 3, 2, 2), (1, 1, 4, 4), (1500, 1300, 384, 384)])
def test_resize_is_bit_exact_with_pillow(shape):
    h, w, oh, ow = shape
    for img in (_noise(h, w, seed=h * 7 + w), _smooth(h, w)):
        ours = IO.resize_bicubic_u8(img, oh, ow)
        ref = np.asarray(PIL.fromarray(img).resize((ow, oh)))            # Pillow default filter = BICUBIC
        assert ours.dtype == np.uint8 and np.array_equal(ours, ref)
        assert np.array_equal(ref, np.asarray(PIL.fromarray(img).resize((ow, oh), PIL.Resampling.BICUBIC)))


def test_extremes_saturate_like_pillow():
    """Bicubic overshoot on hard edges must clip to [0, 255] exactly as clip8 does."""
    img = np.zeros((640, 480, 3), dtype=np.uint8)
    img[::7] = 255
    img[:, ::5] = 255
    ours = IO.resize_bicubic_u8(img, 384, 384)
    assert np.array_equal(ours, np.asarray(PIL.fromarray(img).resize((384, 384))))
    up = IO.resize_bicubic_u8(img[:40, :30], 384, 384)
    assert np.array_equal(up, np.asarray(PIL.fromarray(img[:40, :30]).resize((384, 384))))
    assert up.min() == 0 and up.max() == 255


def test_rescale_normalize_formula_and_hf_processor():
    img = _noise(384, 384, seed=5)
    out = IO.rescale_normalize(img)
    assert out.shape == (3, 384, 384) and out.dtype == np.float32
    lut = ((np.arange(256, dtype=np.float64) * (1 / 255)).astype(np.float32) - np.float32(0.5)) / np.float32(0.5)
    assert np.array_equal(out, lut[img].transpose(2, 0, 1))
    assert out.min() >= -1.0 and out.max() <= 1.0
    tr = pytest.importorskip("transformers")
    cls = getattr(tr, "SiglipImageProcessorPil", None) or getattr(tr, "SiglipImageProcessor")
    proc = cls(size={"height": 384, "width": 384}, image_mean=[0.5] * 3, image_std=[0.5] * 3, resample=3)
    big = _noise(700, 600, seed=6)
    pil = PIL.fromarray(big).resize((384, 384))                          # the dataset's resize
    pv = proc(images=pil, return_tensors="np").pixel_values[0]
    ours = IO.preprocess(big, 384)
    # the PIL-backed processor follows rescale-then-normalize exactly; the torchvision one fuses the two
    # constants and may differ in the last bit
    assert np.abs(pv - ours).max() <= (0.0 if cls.__name__.endswith("Pil") else 6e-8)
