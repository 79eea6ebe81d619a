"""Host half of the GPU JPEG decode (no GPU needed): marker parsing, table building, un-stuffing and restart cutting
(`pt_jpeg_plan` / `pt_jpeg_pack` through the C ABI), and the loader-side plumbing of `gpu_decode=True` datasets."""
import io
import struct

import numpy as np
import pytest
import torch


def _encode(arr, **kw):
    from PIL import Image
    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


def _lib_ok():
    from projectiontrainer_b200 import _lib
    return _lib.LIB_PATH.exists()


pytestmark = pytest.mark.skipif(not _lib_ok(), reason="libpt_stage1.so not built")


def test_plan_reports_geometry_and_rejects_what_the_kernels_do_not_take():
    from PIL import Image
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(0)
    img = rng.integers(0, 256, (37, 53, 3), dtype=np.uint8)
    files = [_encode(img, quality=90, subsampling=2), _encode(img[..., 0], quality=70), _encode(img[:8, :9], subsampling=0)]
    assert ops.jpeg_plan(files) == [(37, 53), (37, 53), (8, 9)]
    with pytest.raises(RuntimeError, match="progressive"):
        ops.jpeg_plan([_encode(img, progressive=True)])
    with pytest.raises(RuntimeError, match="not a JPEG"):
        ops.jpeg_plan([b"GIF89a" + bytes(32)])
    b = io.BytesIO()
    Image.fromarray(np.dstack([img, img[..., :1]])).convert("CMYK").save(b, "JPEG")
    with pytest.raises(RuntimeError, match="components"):
        ops.jpeg_plan([b.getvalue()])
    with pytest.raises(RuntimeError, match="jpeg\\[1\\]"):                 # the message names the offending file
        ops.jpeg_plan([files[0], files[0][:100]])


def test_pack_unstuffs_and_cuts_at_restart_markers():
    """The packed stream of every segment equals the file's entropy-coded bytes with FF00 -> FF and the RSTn markers
    removed; segment MCU ranges follow the restart interval."""
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(1)
    img = rng.integers(0, 256, (48, 80, 3), dtype=np.uint8)          # noise at quality 100: plenty of FF bytes
    data = _encode(img, quality=100, subsampling=2, restart_marker_blocks=2)
    packed = ops.jpeg_pack([data], pin=False)
    blob = packed["blob"].numpy().tobytes()
    magic, n, total_segs = struct.unpack_from("<Iii", blob, 0)
    assert magic == 0x4A504754 and n == 1
    mcus = (48 // 16) * (80 // 16)
    assert total_segs == (mcus + 1) // 2
    # reference un-stuffing in python
    sos = data.index(b"\xff\xda")
    start = sos + 2 + struct.unpack(">H", data[sos + 2:sos + 4])[0]
    body = data[start:data.rindex(b"\xff\xd9")]
    segs, cur, i = [], bytearray(), 0
    while i < len(body):
        if body[i] == 0xFF and i + 1 < len(body):
            if body[i + 1] == 0x00:
                cur.append(0xFF); i += 2; continue
            if 0xD0 <= body[i + 1] <= 0xD7:
                segs.append(bytes(cur)); cur = bytearray(); i += 2; continue
        cur.append(body[i]); i += 1
    segs.append(bytes(cur))
    assert len(segs) == total_segs and any(b"\xff" in s for s in segs)
    imgs_off, segs_off = struct.unpack_from("<qq", blob, 48)
    for k, ref in enumerate(segs):
        img_i, first_mcu, n_mcu, mode, off, ln = struct.unpack_from("<iiiiqq", blob, segs_off + 48 * k)
        assert (img_i, first_mcu, n_mcu, mode) == (0, 2 * k, min(2, mcus - 2 * k), 0)
        assert off % 8 == 0 and blob[off:off + ln] == ref, k
    assert packed["rgb_offsets"] == [0, 48 * 80 * 3] and packed["sizes"] == [(48, 80)]


def test_gpu_decode_dataset_items_and_collates(tmp_path):
    """gpu_decode=True: baseline JPEG files travel as bytes and are packed by the collate (in the worker), other files
    are decoded by Pillow as in the reference and share the batch; order is kept."""
    import json
    from PIL import Image
    from projectiontrainer_b200.projector_trainer import stage1_collate
    from projectiontrainer_b200.train_projection_stage1 import XrayTextPairDataset
    from projectiontrainer_b200.vqa_trainer import vqa_collate_fn, vqa_step_inputs
    rng = np.random.default_rng(2)
    names = []
    for i, (fmt, kw) in enumerate([("JPEG", dict(quality=90)), ("PNG", {}), ("JPEG", dict(progressive=True)), ("JPEG", dict(quality=60))]):
        name = f"im{i}.{'png' if fmt == 'PNG' else 'jpg'}"
        Image.fromarray(rng.integers(0, 256, (20 + i, 30, 3), dtype=np.uint8)).save(tmp_path / name, fmt, **kw)
        names.append(name)
    js = tmp_path / "d.json"
    js.write_text(json.dumps([{"image": n, "normal_caption": "clear lungs"} for n in names]))

    class Tok:
        pad_token_id, padding_side = 0, "right"
        def __call__(self, text, max_length, padding, truncation, return_tensors):
            return type("Enc", (), {"input_ids": torch.tensor([[2, 5, 6] + [0] * (max_length - 3)])})()

    ds = XrayTextPairDataset(str(tmp_path), str(js), None, Tok(), 384, max_length=8, gpu_decode=True)
    items = [ds[i] for i in range(4)]
    assert ["jpeg_bytes" in it for it in items] == [True, False, False, True]
    assert items[1]["images_u8"].shape == (21, 30, 3) and items[2]["images_u8"].shape == (22, 30, 3)
    batch = stage1_collate(items)
    assert batch["token_ids"].shape == (4, 8) and "pixel_values" not in batch
    assert [im is None for im in batch["images_u8"]] == [True, False, False, True]
    assert batch["jpeg"]["index"] == [0, 3] and batch["jpeg"]["sizes"] == [(20, 30), (23, 30)]
    vq = [{**{k: v for k, v in it.items() if k in ("jpeg_bytes", "images_u8")},
           "question_input_ids": torch.tensor([5, 6]), "answer_input_ids": torch.tensor([2, 7])} for it in items]
    cb = vqa_collate_fn(vq, Tok())
    st = vqa_step_inputs(cb, 0, 4)
    assert st["jpeg"]["index"] == [0, 3] and st["images_u8"][1] is cb["images_u8"][1] and "pixel_values" not in st


def test_host_parser_survives_damaged_files():
    """Truncated and bit-flipped files (headers included) either parse or raise RuntimeError with a message: the parser
    never reads outside the file and packing never writes outside the blob (a crash here would take a loader worker
    down)."""
    from projectiontrainer_b200 import ops
    rng = np.random.default_rng(4)
    img = rng.integers(0, 256, (64, 80, 3), dtype=np.uint8)
    seeds = [_encode(img, quality=85, subsampling=2, restart_marker_blocks=3), _encode(img[..., 0], quality=60),
             _encode(img, quality=95, subsampling=0, optimize=True)]
    ok = err = 0
    for it in range(1500):
        b = bytearray(seeds[it % 3])
        mode = it % 4
        if mode == 0:
            b = b[:int(rng.integers(2, len(b)))]
        elif mode == 1:
            for pos in rng.integers(0, min(len(b), 700), 6):
                b[pos] = int(rng.integers(0, 256))
        elif mode == 2:
            pos = int(rng.integers(2, len(b) - 4))
            b[pos:pos + 2] = b"\xff" + bytes([int(rng.integers(0xC0, 0xFF))])
        else:
            pos = int(rng.integers(2, len(b)))
            b[pos:pos] = bytes(rng.integers(0, 256, int(rng.integers(1, 40)), dtype=np.uint8))
        try:
            p = ops.jpeg_pack([bytes(b)], pin=False)
            assert p["blob"].numel() > 0 and p["rgb_offsets"][0] == 0
            ok += 1
        except RuntimeError as e:
            assert "jpeg" in str(e)
            err += 1
    assert ok > 50 and err > 200, (ok, err)
