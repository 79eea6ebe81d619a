"""Host side of the Stage-2 drop-in on CPU: collate and step-input construction bit-exact against the oracle
(which is pinned to the reference's recorded masks), accumulation bookkeeping, argument handling."""
import numpy as np
import pytest
import torch

from oracle import stage1_oracle as O
from oracle import stage2_oracle as O2
from projectiontrainer_b200.vqa_trainer import (VQATrainerStage2, accumulate_sync_flag, vqa_collate_fn,
                                                vqa_step_inputs)


class Tok:
    def __init__(self, pad, side):
        self.pad_token_id, self.eos_token_id, self.padding_side = pad, 1, side


@pytest.mark.parametrize("side", ["left", "right"])
@pytest.mark.parametrize("bucket", [1, 8, 32])
def test_collate_and_step_inputs_bit_exact_vs_oracle(side, bucket):
    vc = O.SiglipCfg(hidden=64, ff=128, layers=1, heads=1, image=64, patch=16)
    lc = O.GemmaCfg(vocab=512, hidden=64, ff=128, layers=6, hq=2, hkv=1, hd=16, window=8, q_scalar=16.0)
    px, qs, as_ = O2.synthetic_vqa_samples(5, vc, lc, seed=3, q_len=(1, 9), a_len=(1, 17))
    items = [{"pixel_values": px[i], "question_input_ids": qs[i], "answer_input_ids": as_[i]} for i in range(5)]
    cb = vqa_collate_fn(items, Tok(lc.pad_id, side))
    q_ref, a_ref = O2.vqa_collate([q.numpy() for q in qs], [a.numpy() for a in as_], lc.pad_id, side)
    assert cb["question_input_ids"].dtype == torch.int64
    assert np.array_equal(cb["question_input_ids"].numpy(), q_ref)
    assert np.array_equal(cb["answer_input_ids"].numpy(), a_ref)
    assert torch.equal(cb["pixel_values"], px)
    st = vqa_step_inputs(cb, lc.pad_id, bucket)
    n_img = 7
    attn_ref, lab_ref = O2.vqa_masks_labels(q_ref, a_ref, n_img, lc.pad_id)
    T = q_ref.shape[1] + a_ref.shape[1]
    ids, lab = st["token_ids"].numpy(), st["labels"].numpy()
    assert ids.shape[1] % bucket == 0 and ids.shape[1] - T < bucket
    assert np.array_equal(ids[:, :T], np.concatenate([q_ref, a_ref], axis=1)) and (ids[:, T:] == lc.pad_id).all()
    assert np.array_equal(lab[:, :T], lab_ref[:, n_img:]) and (lab[:, T:] == -100).all()
    assert np.array_equal((ids[:, :T] != lc.pad_id).astype(np.int64), attn_ref[:, n_img:])
    # the label span covers the answer columns, widened to bucket boundaries (extra rows carry label -100)
    Lq, La = q_ref.shape[1], a_ref.shape[1]
    off, cnt = st["label_span"]
    assert off % bucket == 0 and off <= Lq and Lq - off < bucket and off + cnt == ids.shape[1]
    assert (lab[:, off:Lq] == -100).all() and (lab[:, Lq + La:] == -100).all()
    if bucket == 1:
        assert st["label_span"] == (Lq, La)


def test_ragged_spans_bucket_to_few_graph_keys():
    """at batch_size 1 every (Lq, La) pair is distinct; the graph key (text length, label span) must not be."""
    lc = O.GemmaCfg(vocab=512, hidden=64, ff=128, layers=6, hq=2, hkv=1, hd=16, window=8, q_scalar=16.0)
    rng = np.random.default_rng(0)
    keys, exact = set(), set()
    for _ in range(400):
        Lq, La = int(rng.integers(4, 129)), int(rng.integers(4, 513))
        cb = {"question_input_ids": torch.full((1, Lq), 5, dtype=torch.int64),
              "answer_input_ids": torch.full((1, La), 7, dtype=torch.int64),
              "pixel_values": torch.zeros(1, 3, 8, 8)}
        st = vqa_step_inputs(cb, lc.pad_id, 32)
        keys.add((st["token_ids"].shape[1], st["label_span"]))
        exact.add((Lq, La))
    assert len(exact) > 390 and len(keys) <= 5 * 21, (len(exact), len(keys))


def test_accumulate_bookkeeping_matches_oracle_schedule():
    for n, ga in [(3, 2), (4, 1), (5, 4), (7, 3), (1, 8)]:
        micro, flags = 0, []
        for epoch in range(2):
            for i in range(n):
                f, micro = accumulate_sync_flag(micro, ga, i == n - 1)
                flags.append(f)
        assert flags == O2.sync_flags(n, ga) * 2, (n, ga)


def test_unsupported_stage2_modes_fail_loudly():
    class Acc:
        device = torch.device("cpu")
        is_main_process, process_index, num_processes, gradient_accumulation_steps = True, 0, 1, 1
    base = dict(accelerator=Acc(), vision_encoder=None, language_model=None, projection_layer=None,
                tokenizer=Tok(0, "right"), train_dataset=[], val_dataset=[], output_dir="/tmp/_unused_vqa",
                batch_size=1, learning_rate=1e-4, weight_decay=0.01, num_epochs=1, gradient_accumulation_steps=1,
                warmup_ratio=0.0, freeze_vision_encoder=True, freeze_projection_layer=False, freeze_llm=True,
                enable_qlora=False, train_ve_first_epoch=False, wandb_project="x")
    with pytest.raises(RuntimeError, match="CUDA"):          # adapters train on the kernels too: no CPU path
        VQATrainerStage2(**{**base, "enable_qlora": True})
    with pytest.raises(RuntimeError, match="CUDA"):          # adapters alone are a valid trainable set
        VQATrainerStage2(**{**base, "enable_qlora": True, "freeze_projection_layer": True})
    with pytest.raises(NotImplementedError):
        VQATrainerStage2(**{**base, "freeze_llm": False})
    with pytest.raises(NotImplementedError):
        VQATrainerStage2(**{**base, "train_ve_first_epoch": True})
    with pytest.raises(ValueError, match="No trainable parameters"):
        VQATrainerStage2(**{**base, "freeze_projection_layer": True})
    with pytest.raises(RuntimeError, match="CUDA"):          # no CPU path
        VQATrainerStage2(**base)


def test_stage2_cli_flags():
    from projectiontrainer_b200.train_vqa_stage2 import build_parser, main
    a = build_parser().parse_args([])
    assert a.gradient_accumulation_steps == 8 and a.batch_size == 1 and a.warmup_ratio == 0.05
    assert a.max_q_len == 128 and a.max_a_len == 512 and a.llm_name == "Qwen/Qwen3-8B"
    with pytest.raises(SystemExit):
        main(["--unfreeze_llm", "--synthetic", "4"])
    with pytest.raises(SystemExit):
        main(["--resume_qlora_adapter_path", "/nonexistent", "--synthetic", "4"])      # needs --enable_qlora


def test_qlora_adapter_directory_round_trip(tmp_path):
    """`--resume_qlora_adapter_path`: adapter_config.json + adapter_model.safetensors in peft's layout are read back
    into the LoRA fields and the tensors by peft key."""
    import json as _json
    from safetensors.torch import save_file
    from projectiontrainer_b200.lora import LoraState
    from projectiontrainer_b200.train_vqa_stage2 import load_qlora_adapter
    sd = {LoraState.key(0, "q_proj", "A"): torch.randn(8, 32), LoraState.key(0, "q_proj", "B"): torch.randn(64, 8),
          LoraState.key(1, "down_proj", "A"): torch.randn(8, 96)}
    assert LoraState.key(1, "down_proj", "A") == "base_model.model.model.layers.1.mlp.down_proj.lora_A.weight"
    assert LoraState.key(0, "q_proj", "B") == "base_model.model.model.layers.0.self_attn.q_proj.lora_B.weight"
    save_file(sd, str(tmp_path / "adapter_model.safetensors"))
    (tmp_path / "adapter_config.json").write_text(_json.dumps(
        {"peft_type": "LORA", "r": 8, "lora_alpha": 16, "target_modules": ["q_proj", "down_proj"], "lora_dropout": 0.1}))
    cfg, got = load_qlora_adapter(str(tmp_path))
    assert cfg == {"r": 8, "lora_alpha": 16, "target_modules": ["q_proj", "down_proj"], "lora_dropout": 0.1}
    assert set(got) == set(sd) and all(torch.equal(got[k], sd[k]) for k in sd)


def test_collates_keep_decoded_images_as_lists():
    """gpu_preprocess datasets hand out `images_u8` of different sizes: both collates keep them a list and
    collate everything else as before."""
    from projectiontrainer_b200.projector_trainer import stage1_collate
    a = {"images_u8": torch.zeros(5, 7, 3, dtype=torch.uint8), "token_ids": torch.arange(4), "labels": torch.arange(4)}
    b = {"images_u8": torch.zeros(9, 2, 3, dtype=torch.uint8), "token_ids": torch.arange(4), "labels": torch.arange(4)}
    out = stage1_collate([a, b])
    assert isinstance(out["images_u8"], list) and out["images_u8"][1].shape == (9, 2, 3)
    assert out["token_ids"].shape == (2, 4) and "pixel_values" not in out
    plain = stage1_collate([{"pixel_values": torch.zeros(3, 4, 4), "token_ids": torch.arange(4), "labels": torch.arange(4)}] * 2)
    assert plain["pixel_values"].shape == (2, 3, 4, 4)
    items = [{"images_u8": a["images_u8"], "question_input_ids": torch.tensor([5, 6]), "answer_input_ids": torch.tensor([2, 7, 8])},
             {"images_u8": b["images_u8"], "question_input_ids": torch.tensor([9]), "answer_input_ids": torch.tensor([2])}]
    cb = vqa_collate_fn(items, Tok(0, "left"))
    assert "pixel_values" not in cb and len(cb["images_u8"]) == 2
    st = vqa_step_inputs(cb, 0, 4)
    assert st["images_u8"] is cb["images_u8"] and st["token_ids"].tolist() == [[5, 6, 2, 7, 8, 0, 0, 0], [0, 9, 0, 0, 2, 0, 0, 0]]
    assert st["labels"].tolist() == [[-100, -100, 2, 7, 8, -100, -100, -100], [-100] * 4 + [2, -100, -100, -100]]


def test_stage1_dataset_token_cache_matches_lazy_tokenisation(tmp_path):
    """cache_tokens=True: one batched tokenizer call, same ids as the per-item path, shared with workers."""
    import json as _json
    from projectiontrainer_b200.train_projection_stage1 import XrayTextPairDataset

    class WordTok:
        pad_token_id, calls = 0, 0

        def __call__(self, text, max_length, padding, truncation, return_tensors):
            WordTok.calls += 1
            texts = [text] if isinstance(text, str) else text
            rows = []
            for t in texts:
                ids = [2] + [3 + (hash(w) % 97) for w in t.split()][: max_length - 1]
                rows.append([0] * (max_length - len(ids)) + ids)
            return type("Enc", (), {"input_ids": torch.tensor(rows)})()

    js = tmp_path / "d.json"
    js.write_text(_json.dumps([{"image": f"{i}.png", "normal_caption": "lungs are clear " * (i + 1)} for i in range(5)]))
    lazy = XrayTextPairDataset(str(tmp_path), str(js), None, WordTok(), 8, max_length=12)
    cached = XrayTextPairDataset(str(tmp_path), str(js), None, WordTok(), 8, max_length=12, cache_tokens=True)
    WordTok.calls = 0
    assert len(cached) == 5 and WordTok.calls == 1 and cached._token_matrix.is_shared()
    for i in range(5):
        assert torch.equal(cached._tokens(i), lazy._tokens(i))
    assert WordTok.calls == 1 + 5


def test_inference_vqa_stage2_flags_and_image_path_rules(tmp_path):
    """The Stage-2 inference mirror keeps the reference's flags / defaults (ref:Stage2/inference_vqa_stage2.py:300-319)
    and its image path rules (:222-241): plain entries under image_root, 'p.../s...' study directories under
    image_root_2 with the first .jpg inside."""
    from projectiontrainer_b200 import inference_vqa_stage2 as I
    a = I.build_parser().parse_args([])
    assert (a.max_new_tokens, a.do_sample, a.temperature, a.top_p, a.top_k, a.repetition_penalty, a.length_penalty,
            a.num_beams) == (512, True, 0.3, 0.9, 50, 1.8, 1.2, 3)
    assert a.vision_model_name == "StanfordAIMI/XraySigLIP__vit-l-16-siglip-384__webli" and a.image_root_2 is None
    r1, r2 = tmp_path / "a", tmp_path / "b"
    (r2 / "p10" / "s50").mkdir(parents=True)
    (r2 / "p10" / "s50" / "x.JPG").write_bytes(b"")
    (r2 / "p11" / "s51").mkdir(parents=True)
    assert I.resolve_image_path("images_002/images/1.jpg", str(r1), str(r2)) == str(r1 / "images_002/images/1.jpg")
    assert I.resolve_image_path("p10/s50", str(r1), str(r2)) == str(r2 / "p10" / "s50" / "x.JPG")
    assert I.resolve_image_path("p11/s51", str(r1), str(r2)) is None              # a study directory without a .jpg
    assert I.resolve_image_path("p10/s50", str(r1), None) == str(r1 / "p10/s50")    # no second root: the plain rule
    # a sample without image / problem is skipped before any model work
    assert I.process_sample({"image": "x.jpg"}, str(r1), None, 384, None, None, {}, 0, 1) is None


def test_inference_generation_summary_counts():
    """ref:Stage2/inference_generation.py:94,129-199: a sample is correct when its ground-truth label occurs in the
    generated text (case-insensitive); the summary counts per-label hits and candidate labels found in predictions."""
    from projectiontrainer_b200 import inference_generation as G
    a = G.build_parser().parse_args([])
    assert a.max_length == 128 and not a.verbose and G.PROMPT.startswith("Identify the diseases in this chest X-ray")
    res = [{"generated_text": "Likely ATELECTASIS with effusion", "normal_caption": "Atelectasis", "is_correct": True},
           {"generated_text": "no finding", "normal_caption": "Atelectasis", "is_correct": False},
           {"generated_text": "No Finding.", "normal_caption": "No Finding", "is_correct": True}]
    s = G.summarise(res, ["Atelectasis", "No Finding", "Edema"])
    assert (s["total"], s["correct"]) == (3, 2)
    assert s["ground_truth"] == {"Atelectasis": 2, "No Finding": 1}
    assert s["label_correct"]["Atelectasis"] == 1 and s["label_correct"]["No Finding"] == 1
    assert s["predicted"] == {"Atelectasis": 1, "No Finding": 2}
    assert G.summarise([], ["x"])["total"] == 0
