"""GPU probe: damaged JPEG files that still pass the host parser must decode without a device fault (garbage pixels are
fine) and must not disturb a clean file in the same batch."""
import io
import sys
from pathlib import Path
sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
import numpy as np
import torch
from PIL import Image
from projectiontrainer_b200 import ops

dev = torch.device("cuda:0")
rng = np.random.default_rng(int(sys.argv[1]) if len(sys.argv) > 1 else 0)


def enc(arr, **kw):
    b = io.BytesIO()
    Image.fromarray(arr).save(b, "JPEG", **kw)
    return b.getvalue()


img = rng.integers(0, 256, (300, 420, 3), dtype=np.uint8)
seeds = [enc(img, quality=85, subsampling=2, restart_marker_blocks=5), enc(img[..., 0], quality=70), enc(img, quality=95, subsampling=0),
         enc(np.repeat(np.repeat(img, 3, 0), 3, 1), quality=90, subsampling=1)]
clean = seeds[0]
ref = np.asarray(Image.open(io.BytesIO(clean)).convert("RGB"))
done = rejected = 0
for it in range(400):
    b = bytearray(seeds[it % 4])
    for _ in range(int(rng.integers(1, 12))):
        pos = int(rng.integers(2, len(b)))
        b[pos] = int(rng.integers(0, 256))
    if it % 5 == 0:
        b = b[:int(rng.integers(len(b) // 2, len(b)))]
    try:
        sizes = ops.jpeg_plan([bytes(b)])
    except RuntimeError:
        rejected += 1
        continue
    if sizes[0][0] * sizes[0][1] > 4_000_000:
        continue
    try:
        packed = ops.jpeg_pack([bytes(b), clean])
    except RuntimeError:
        rejected += 1
        continue
    out = ops.jpeg_decode(packed, dev)
    torch.cuda.synchronize()
    assert np.array_equal(out[1].cpu().numpy(), ref), it
    done += 1
print(f"decoded {done} damaged files without a fault ({rejected} rejected by the host parser); the clean neighbour stayed bit-exact")
