"""Drop-in for ref:Stage2/dataset.py: X-ray image / question / answer triplets from a JSON list of
{"image", "problem", "normal_caption"} records.  Items (ref :116-120): `pixel_values` f32 [3,S,S] (processor
output), `question_input_ids` i64 [<=max_q_len] tokenised WITHOUT special tokens, `answer_input_ids` i64
[<=max_a_len] tokenised with them; padding happens in `vqa_collate_fn`.  Host-side data format only: no
arithmetic of the hot path lives here."""
from __future__ import annotations

import json
import logging
import os

import torch
from torch.utils.data import Dataset

logger = logging.getLogger(__name__)


class XrayVQADataset(Dataset):
    def __init__(self, image_root, json_file, processor, tokenizer, img_size, max_q_len=128, max_a_len=512,
                 image_root_2=None, gpu_preprocess=False, gpu_decode=False):
        """`gpu_preprocess=True` (an addition): items carry the decoded image as `images_u8` uint8 [H, W, 3]
        instead of `pixel_values`; resize + rescale + normalize then run on the GPU inside the trainer's
        look-ahead stage (`Stage1Engine.preprocess_images`, Pillow-exact)."""
        # gpu_decode (an addition, implies gpu_preprocess): workers only read the file; baseline JPEGs are decoded on
        # the GPU too (image_io.py)
        self.gpu_preprocess, self.gpu_decode = gpu_preprocess or gpu_decode, gpu_decode
        self.image_root, self.image_root_2, self.img_size = image_root, image_root_2, img_size
        self.processor, self.tokenizer = processor, tokenizer
        self.max_q_len, self.max_a_len = max_q_len, max_a_len
        if self.tokenizer.pad_token is None:            # ref :33-35
            self.tokenizer.pad_token = self.tokenizer.eos_token
            logger.info(f"Set tokenizer pad_token to eos_token ({self.tokenizer.eos_token})")
        with open(json_file, "r", encoding="utf-8") as f:
            self.samples = json.load(f)
        logger.info(f"Loaded {len(self.samples)} samples from {json_file}")

    def __len__(self):
        return len(self.samples)

    def _image_path(self, name: str) -> str:
        """ref :66-82: names like 'p10012261/s50349409' live under image_root_2 and may be a study
        directory holding the .jpg; everything else is relative to image_root."""
        if name.startswith("p") and "/" in name and self.image_root_2:
            path = os.path.join(self.image_root_2, name)
            if os.path.isdir(path):
                jpgs = [f for f in os.listdir(path) if f.lower().endswith(".jpg")]
                if not jpgs:
                    raise FileNotFoundError(f"No .jpg files found in directory: {path}")
                path = os.path.join(path, jpgs[0])
            return path
        return os.path.join(self.image_root, name)

    def _load(self, idx: int) -> dict:
        from PIL import Image
        s = self.samples[idx]
        name, question, answer = s.get("image"), s.get("problem"), s.get("normal_caption")
        if not all([name, question, answer]):
            raise KeyError(f"sample {idx} is missing image, problem or normal_caption")
        image = None if self.gpu_decode else Image.open(self._image_path(name)).convert("RGB")
        q = self.tokenizer(question, max_length=self.max_q_len, truncation=True, add_special_tokens=False).input_ids
        a = self.tokenizer(answer, max_length=self.max_a_len, truncation=True).input_ids
        item = {"question_input_ids": torch.tensor(q, dtype=torch.long),
                "answer_input_ids": torch.tensor(a, dtype=torch.long)}
        if self.gpu_decode:
            from .image_io import load_image_item
            item.update(load_image_item(self._image_path(name), True))
        elif self.gpu_preprocess:
            import numpy as np
            item["images_u8"] = torch.from_numpy(np.array(image, dtype=np.uint8))
        else:
            image = image.resize((self.img_size, self.img_size))
            item["pixel_values"] = self.processor(images=image, return_tensors="pt").pixel_values.squeeze(0)
        return item

    def __getitem__(self, idx):
        """A sample that cannot be read is skipped in favour of the next readable one, as the reference
        does (ref :63-65, 122-126) — iteratively here, so a run of bad records cannot exhaust the stack."""
        n = len(self)
        for k in range(n):
            j = (idx + k) % n
            try:
                return self._load(j)
            except Exception as e:
                logger.warning(f"Skipping sample {j}: {e}")
        raise RuntimeError("no readable sample in the dataset")


class SyntheticVQADataset(Dataset):
    """Ragged CXR-shaped random triplets for the boxes without data: pixels ~ U(-1,1), question ids
    ~ U{3..V-1} of random length, answers starting with BOS(2)."""

    def __init__(self, n, img_size, vocab, q_len=(8, 48), a_len=(16, 96), seed=1234):
        g = torch.Generator().manual_seed(seed)
        self.px = torch.rand(n, 3, img_size, img_size, generator=g) * 2 - 1
        self.q, self.a = [], []
        for _ in range(n):
            lq = int(torch.randint(q_len[0], q_len[1] + 1, (1,), generator=g))
            la = int(torch.randint(a_len[0], a_len[1] + 1, (1,), generator=g))
            self.q.append(torch.randint(3, vocab, (lq,), generator=g))
            a = torch.randint(3, vocab, (la,), generator=g)
            a[0] = 2
            self.a.append(a)

    def __len__(self):
        return len(self.q)

    def __getitem__(self, i):
        return {"pixel_values": self.px[i], "question_input_ids": self.q[i], "answer_input_ids": self.a[i]}
