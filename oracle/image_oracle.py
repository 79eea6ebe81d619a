"""CPU oracle for the image preprocessing in front of the hot path.  TEST INFRASTRUCTURE ONLY.

Only tests/ may import this module; the product path (projectiontrainer_b200/) never does.

Restates what the reference's datasets do to every decoded image before the vision tower
(ref:Stage1/train_projection_stage1.py:99-104, ref:Stage2/dataset.py:90-95):

    image = Image.open(path).convert('RGB')
    image = image.resize((img_size, img_size))              # Pillow default filter: BICUBIC, antialiased
    pixel_values = processor(images=image, return_tensors="pt").pixel_values.squeeze(0)

The algorithm lives in third-party code that is not under /root/reference:

  * Pillow (requirements.txt pins pillow==11.1.0; 12.2.0 in this image), `src/libImaging/Resample.c`:
    separable two-pass resampling in 8-bit fixed point — `precompute_coeffs` (filter support scaled by the
    down-scale factor, weights normalised in double), `normalize_coeffs_8bpc` (22 fractional bits, round
    half away from zero), `ImagingResampleHorizontal_8bpc` over the rows the vertical pass needs, then
    `ImagingResampleVertical_8bpc`; each pass rounds to uint8 (`clip8`).
  * HF `SiglipImageProcessor` (transformers==4.51.0, `models/siglip/image_processing_siglip.py` +
    `image_transforms.py`): the 384x384 input is already at `size`, so its bicubic resize is the identity
    (`Image.resize` returns a copy when the size does not change); `rescale` multiplies by 1/255 in float64
    and casts to float32; `normalize` computes (x - mean) / std in float32, mean = std = 0.5; output CHW.

Parity pin: tests/test_image_oracle.py checks `resize_bicubic_u8` bit-exact against Pillow itself (imported
in the test: Pillow is part of the image on both boxes) over up- and down-scaling, odd sizes and tiny
inputs, and `rescale_normalize` against the formula above.
"""
from __future__ import annotations

import math

import numpy as np

PRECISION_BITS = 32 - 8 - 2          # Resample.c
BICUBIC_SUPPORT = 2.0


def bicubic_filter(x: np.ndarray) -> np.ndarray:
    """Resample.c `bicubic_filter`, a = -0.5 (Keys), evaluated in the source's operation order."""
    a = -0.5
    x = np.abs(x)
    near = ((a + 2.0) * x - (a + 3.0)) * x * x + 1
    far = (((x - 5) * x + 8) * x - 4) * a
    return np.where(x < 1.0, near, np.where(x < 2.0, far, 0.0))


def precompute_coeffs(in_size: int, out_size: int):
    """Resample.c `precompute_coeffs` + `normalize_coeffs_8bpc` for the full-image box (in0 = 0,
    in1 = in_size).  Returns bounds i32 [out, 2] (xmin, count) and fixed-point weights i32 [out, ksize]."""
    scale = float(np.float32(in_size) - np.float32(0.0)) / out_size
    filterscale = max(scale, 1.0)
    support = BICUBIC_SUPPORT * filterscale
    ksize = int(math.ceil(support)) * 2 + 1
    bounds = np.zeros((out_size, 2), dtype=np.int32)
    kk = np.zeros((out_size, ksize), dtype=np.int32)
    ss = 1.0 / filterscale
    for xx in range(out_size):
        center = 0.0 + (xx + 0.5) * scale
        xmin = max(int(center - support + 0.5), 0)          # C (int) cast truncates toward zero
        xmax = min(int(center + support + 0.5), in_size) - xmin
        x = np.arange(xmax, dtype=np.float64)
        w = bicubic_filter((x + xmin - center + 0.5) * ss)
        ww = 0.0
        for v in w:                                         # sequential sum, as the C loop
            ww += float(v)
        if ww != 0.0:
            w = w / ww
        fixed = np.where(w < 0, -0.5 + w * (1 << PRECISION_BITS), 0.5 + w * (1 << PRECISION_BITS))
        kk[xx, :xmax] = np.trunc(fixed).astype(np.int32)
        bounds[xx] = (xmin, xmax)
    return bounds, kk


def _clip8(acc: np.ndarray) -> np.ndarray:
    return np.clip(acc >> PRECISION_BITS, 0, 255).astype(np.uint8)


def resize_bicubic_u8(img: np.ndarray, out_h: int, out_w: int) -> np.ndarray:
    """`Image.fromarray(img).resize((out_w, out_h))` for an RGB uint8 HWC array (Resample.c
    `ImagingResampleInner`): horizontal pass over the rows the vertical pass reads, then vertical pass;
    a pass whose size does not change is skipped."""
    H, W, C = img.shape
    need_h, need_v = out_w != W, out_h != H
    bh, kh = precompute_coeffs(W, out_w)
    bv, kv = precompute_coeffs(H, out_h)
    first, last = (int(bv[0, 0]), int(bv[-1, 0] + bv[-1, 1])) if need_v else (0, H)
    cur = img
    if need_h:
        rows = img[first:last].astype(np.int64)
        tmp = np.empty((last - first, out_w, C), dtype=np.uint8)
        for xx in range(out_w):
            x0, n = bh[xx]
            acc = (rows[:, x0:x0 + n, :] * kh[xx, :n, None].astype(np.int64)).sum(axis=1) + (1 << (PRECISION_BITS - 1))
            tmp[:, xx, :] = _clip8(acc)
        cur = tmp
        bv = bv.copy()
        bv[:, 0] -= first
    if need_v:
        src = cur.astype(np.int64)
        out = np.empty((out_h, cur.shape[1], C), dtype=np.uint8)
        for yy in range(out_h):
            y0, n = bv[yy]
            acc = (src[y0:y0 + n] * kv[yy, :n, None, None].astype(np.int64)).sum(axis=0) + (1 << (PRECISION_BITS - 1))
            out[yy] = _clip8(acc)
        cur = out
    return cur.copy() if cur is img else cur


def rescale_normalize(img_u8: np.ndarray, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5),
                      rescale_factor: float = 1 / 255) -> np.ndarray:
    """HF `rescale` (float64 product cast to float32) then `normalize` ((x - mean) / std in float32),
    HWC uint8 in, CHW float32 out."""
    x = (img_u8.astype(np.float64) * rescale_factor).astype(np.float32)
    m = np.asarray(mean, dtype=np.float32)
    s = np.asarray(std, dtype=np.float32)
    return np.ascontiguousarray(((x - m) / s).transpose(2, 0, 1))


def preprocess(img_u8: np.ndarray, size: int = 384, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5)) -> np.ndarray:
    """Decoded RGB uint8 HWC image -> `pixel_values` f32 [3, size, size] as the reference's dataset builds them."""
    return rescale_normalize(resize_bicubic_u8(img_u8, size, size), mean, std)
