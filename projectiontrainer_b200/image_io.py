"""Image files -> what the GPU input pipeline takes.

The reference's datasets do `Image.open(path).convert("RGB")`, `image.resize((S, S))` and `processor(images=image)`
per item on the CPU (ref:Stage1/train_projection_stage1.py:99-104, ref:Stage2/dataset.py:86-95).  With
`gpu_decode=True` a worker only READS the file: baseline JPEG files travel as their encoded bytes and are decoded on
the GPU (`pt_jpeg_decode_batch`, bit-exact with Pillow), followed by the Pillow-exact resize / rescale / normalize
kernels (`pt_image_preprocess`).  Files the GPU decoder does not take (PNG, progressive or CMYK JPEG, ...) are decoded
by Pillow in the worker exactly as the reference does and join the batch as decoded uint8 images — that is the
reference's own loader code for those files, not a substitute for a kernel.
"""
from __future__ import annotations

import logging

import torch

logger = logging.getLogger(__name__)
_warned = set()


def load_image_item(path: str, gpu_decode: bool) -> dict:
    """{"jpeg_bytes": uint8 [n]} for a file the GPU decoder takes, else {"images_u8": uint8 [H, W, 3]} (Pillow)."""
    if gpu_decode:
        with open(path, "rb") as f:
            data = f.read()
        if data[:2] == b"\xff\xd8":
            from . import ops
            try:
                ops.jpeg_plan([data])           # host-side header parse: raises for what the kernels do not take
                return {"jpeg_bytes": torch.frombuffer(bytearray(data), dtype=torch.uint8)}
            except RuntimeError as e:
                kind = str(e).split(":")[-1].strip()[:60]
                if kind not in _warned:
                    _warned.add(kind)
                    logger.warning(f"{path}: decoded by Pillow in the worker ({kind})")
    import numpy as np
    from PIL import Image
    return {"images_u8": torch.from_numpy(np.array(Image.open(path).convert("RGB"), dtype=np.uint8))}


def collate_images(items: list) -> dict:
    """Per-item `jpeg_bytes` / `images_u8` -> batch entries: `images_u8` = list with one slot per item (a decoded
    uint8 image, or None where the image is in the JPEG pack) and, when any item carries encoded bytes, `jpeg` =
    ops.jpeg_pack of those files + `index` (their positions in the batch).  Runs in the DataLoader worker: host only."""
    out = {"images_u8": [it.get("images_u8") for it in items]}
    idx = [i for i, it in enumerate(items) if "jpeg_bytes" in it]
    if idx:
        from . import ops
        pack = ops.jpeg_pack([items[i]["jpeg_bytes"] for i in idx], pin=False)
        pack["index"] = idx
        out["jpeg"] = pack
    return out


def has_raw_images(items: list) -> bool:
    return "images_u8" in items[0] or "jpeg_bytes" in items[0]


def strip_images(item: dict) -> dict:
    return {k: v for k, v in item.items() if k not in ("images_u8", "jpeg_bytes")}
