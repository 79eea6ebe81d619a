"""pt_sample_token — the next-token choice of `generate` in one launch — against a torch restatement of HF's order
(repetition penalty -> temperature -> top-k -> top-p -> softmax -> draw; HF:generation/logits_process.py, `_sample`
in HF:generation/utils.py) through the C-ABI.

The draw is an inverse CDF at a caller-supplied uniform number, so it is deterministic and checkable: the reference
builds HF's filtered distribution in float64, orders the surviving tokens as the kernel does (descending value,
ascending index) and inverts the CDF at the same u.  Tokens must be equal except where u falls within 1e-5 of a CDF
step (f32 vs f64 sums); the surviving SET must always be a superset-compatible choice (the drawn token survives HF's
filters)."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _hf_filtered(logits, prev, rp, temperature, top_k, top_p):
    """HF's processors on one row, float64: returns probabilities [V] (0 for removed tokens)."""
    sc = logits.double().clone()
    if prev is not None and rp != 1.0 and len(prev):
        g = sc[prev]
        sc[prev] = torch.where(g < 0, g * rp, g / rp)
    sc = sc / temperature
    kth = sc.topk(min(top_k, sc.numel())).values[-1]
    sc = sc.masked_fill(sc < kth, float("-inf"))
    if top_p < 1.0:
        srt, idx = sc.sort(descending=False)
        cum = srt.softmax(0).cumsum(0)
        rm = cum <= (1.0 - top_p)
        rm[-1] = False
        sc = sc.masked_fill(torch.zeros_like(rm).scatter(0, idx, rm), float("-inf"))
    return sc.softmax(0), sc


def _draw(probs, sc, u):
    keep = torch.nonzero(probs > 0).flatten()
    order = sorted(keep.tolist(), key=lambda i: (-float(sc[i]), i))
    cdf = torch.tensor([float(probs[i]) for i in order], dtype=torch.float64).cumsum(0)
    pos = int(torch.searchsorted(cdf, torch.tensor(u, dtype=torch.float64) * cdf[-1], right=True).clamp(max=len(order) - 1))
    margin = float((cdf - u * cdf[-1]).abs().min())
    return order[pos], margin


def _run(dev, logits, *, step, prev_tokens, rp, temperature, top_k, top_p, u, done=None, pad=None, eos=None):
    from projectiontrainer_b200 import ops
    B = logits.shape[0]
    out = torch.full((B, step + 2), -7, dtype=torch.int64, device=dev)
    if prev_tokens is not None:
        out[:, :step] = prev_tokens
    nxt = torch.full((B,), -1, dtype=torch.int64, device=dev)
    hold = torch.empty(B, max(logits.stride(0), logits.shape[1]), device=dev)     # a repetition penalty is applied to
    hold[:, :logits.shape[1]] = logits                                            # the logits in place: work on a copy
    logits = hold[:, :logits.shape[1]]                                            # with the caller's row stride
    ops.sample_token(logits, out=out, step=step, next_tokens=nxt, prev=out if prev_tokens is not None else None,
                     repetition_penalty=rp, temperature=temperature, top_k=top_k, top_p=top_p, u=u, done=done,
                     pad_id=pad, eos_id=eos)
    torch.cuda.synchronize()
    assert torch.equal(out[:, step], nxt) and bool((out[:, step + 1] == -7).all())
    return nxt


@pytest.mark.parametrize("V", [1000, 32003, 262144])
@pytest.mark.parametrize("cfg", [(0.7, 50, 0.9, 1.0), (1.0, 50, 1.0, 1.0), (0.3, 8, 0.5, 1.3), (1.5, 1024, 0.95, 1.1),
                                 (0.7, 1, 0.9, 1.0)], ids=["ref_call", "topk_only", "sharp", "k1024", "k1"])
def test_draw_equals_the_float64_restatement(cuda_dev, V, cfg):
    temperature, top_k, top_p, rp = cfg
    B, step = 12, 9
    g = torch.Generator(device=cuda_dev).manual_seed(V + top_k)
    ld = (V + 7) // 8 * 8 + 8
    buf = torch.randn(B, ld, generator=g, device=cuda_dev) * 3.0
    logits = buf[:, :V]
    prev = torch.randint(0, V, (B, step), generator=g, device=cuda_dev)
    prev[:, 0] = logits.argmax(1)                        # the penalty hits the top token too
    u = torch.rand(B, generator=g, device=cuda_dev)
    got = _run(cuda_dev, logits, step=step, prev_tokens=prev if rp != 1.0 else None, rp=rp, temperature=temperature,
               top_k=top_k, top_p=top_p, u=u).cpu()
    lc, pc, uc = logits.cpu(), prev.cpu(), u.cpu()
    near = 0
    for b in range(B):
        probs, sc = _hf_filtered(lc[b], pc[b] if rp != 1.0 else None, rp, temperature, top_k, top_p)
        want, margin = _draw(probs, sc, float(uc[b]))
        assert float(probs[int(got[b])]) > 0, (b, "the drawn token does not survive HF's filters")
        if int(got[b]) != want:
            assert margin < 1e-5, (b, int(got[b]), want, margin)
            near += 1
    assert near <= 1


def test_frequencies_follow_the_filtered_distribution(cuda_dev):
    """4096 rows with the SAME logits and independent u: the token frequencies match HF's filtered probabilities."""
    V, B = 5000, 4096
    g = torch.Generator(device=cuda_dev).manual_seed(5)
    row = torch.randn(V, generator=g, device=cuda_dev) * 2.0
    logits = row[None, :].expand(B, V).contiguous()
    u = torch.rand(B, generator=g, device=cuda_dev)
    got = _run(cuda_dev, logits, step=0, prev_tokens=None, rp=1.0, temperature=0.8, top_k=20, top_p=0.9, u=u).cpu()
    probs, _ = _hf_filtered(row.cpu(), None, 1.0, 0.8, 20, 0.9)
    freq = torch.bincount(got, minlength=V).double() / B
    assert float(freq[probs == 0].sum()) == 0.0
    keep = probs > 0
    # binomial standard error per token: sqrt(p (1 - p) / B) <= 0.008; 5 sigma
    assert float((freq[keep] - probs[keep]).abs().max()) < 0.04


def test_argmax_penalty_ties_and_bookkeeping(cuda_dev):
    V, B = 4099, 6
    g = torch.Generator(device=cuda_dev).manual_seed(2)
    logits = torch.randn(B, V, generator=g, device=cuda_dev)
    logits[0, 17] = logits[0, 4000] = 9.0                # a tie: the smallest index wins
    logits[1, 5] = 8.0                                   # the top token was generated before: penalised away
    logits[1, 77] = 6.0
    prev = torch.full((B, 3), 1, dtype=torch.int64, device=cuda_dev)
    prev[1, 2] = 5
    done = torch.tensor([0, 0, 1, 0, 0, 0], dtype=torch.uint8, device=cuda_dev)
    logits[3, 42] = 20.0                                 # row 3 emits eos now
    got = _run(cuda_dev, logits, step=3, prev_tokens=prev, rp=2.0, temperature=1.0, top_k=0, top_p=1.0, u=None,
               done=done, pad=0, eos=42)
    sc = logits.clone()
    gsel = sc.gather(1, prev)
    sc.scatter_(1, prev, torch.where(gsel < 0, gsel * 2.0, gsel / 2.0))
    want = sc.argmax(1)
    want[2] = 0                                          # finished row: pad
    assert int(got[0]) == 17 and int(got[1]) == 77
    assert torch.equal(got, want)
    assert done.tolist() == [0, 0, 1, 1, 0, 0]


def test_flat_rows_take_the_bisection_path(cuda_dev):
    """Rows where far more than 2048 values pass the first bound: constant rows (every token tied with the k-th), a
    two-level row (3000 tied maxima, k = 50) and a row of 4096 distinct top values among equal ones."""
    V, B = 20000, 3
    logits = torch.zeros(B, V, device=cuda_dev)
    logits[1, ::6] = 1.0                                 # 3334 tied maxima
    logits[2, :4096] = torch.linspace(1.0, 2.0, 4096, device=cuda_dev)
    u = torch.tensor([0.5, 0.999, 0.0], device=cuda_dev)
    got = _run(cuda_dev, logits, step=0, prev_tokens=None, rp=1.0, temperature=1.0, top_k=50, top_p=1.0, u=u).cpu()
    assert 0 <= int(got[0]) < V                          # all tied: any token is a valid draw
    assert int(got[1]) % 6 == 0                          # only the tied maxima survive top-k
    assert int(got[2]) == 4095                           # u = 0: the largest value
    # row 2 again at u just below 1: the 50th largest of the distinct values
    got = _run(cuda_dev, logits, step=0, prev_tokens=None, rp=1.0, temperature=1.0, top_k=50, top_p=1.0,
               u=torch.tensor([0.5, 0.5, 0.9999999], device=cuda_dev)).cpu()
    assert int(got[2]) == 4096 - 50


def test_adversarial_layout_distinct_values_through_the_bisection(cuda_dev):
    """The 4096 largest (distinct) values sit in the stride classes of 16 threads, so the bound from the thread maxima
    is far too low and ~V values pass it: the bisection has to find the exact k-th value."""
    V, B = 262144, 4
    g = torch.Generator(device=cuda_dev).manual_seed(9)
    logits = torch.randn(B, V, generator=g, device=cuda_dev)
    hot = (torch.arange(V, device=cuda_dev) % 1024) < 16
    logits[:, hot] = 50.0 + torch.rand(B, int(hot.sum()), generator=g, device=cuda_dev) * 4.0
    u = torch.rand(B, generator=g, device=cuda_dev)
    got = _run(cuda_dev, logits, step=0, prev_tokens=None, rp=1.0, temperature=0.9, top_k=50, top_p=0.95, u=u).cpu()
    lc, uc = logits.cpu(), u.cpu()
    for b in range(B):
        probs, sc = _hf_filtered(lc[b], None, 1.0, 0.9, 50, 0.95)
        want, margin = _draw(probs, sc, float(uc[b]))
        assert float(probs[int(got[b])]) > 0
        assert int(got[b]) == want or margin < 1e-5, (b, int(got[b]), want, margin)


def test_generate_with_the_kernel_and_with_the_torch_restatement_agree_on_greedy(cuda_dev):
    from oracle import stage1_oracle as O
    from projectiontrainer_b200 import engine as E
    from projectiontrainer_b200.config import GemmaCfg, SiglipCfg
    vkw = dict(hidden=64, ff=128, layers=1, heads=1, image=64, patch=16)
    lkw = dict(vocab=512, hidden=128, ff=128, layers=3, hq=4, hkv=2, hd=64, window=12, q_scalar=64.0)
    ovc, olc = O.SiglipCfg(**vkw), O.GemmaCfg(**lkw)
    eng = E.Stage1Engine(E.PackedSiglip(O.init_siglip(ovc), SiglipCfg(**vkw), cuda_dev),
                         E.PackedGemma(O.init_gemma(olc), GemmaCfg(**lkw), cuda_dev),
                         E.ProjectorState(O.init_projector(ovc.hidden, olc.hidden), cuda_dev), cuda_dev)
    px, ids, _ = O.synthetic_batch(3, 6, ovc, olc, seed=4, left_pad=False)
    kw = dict(max_new_tokens=9, do_sample=False, repetition_penalty=1.2, pad_token_id=olc.pad_id, eos_token_id=1)
    assert eng.fuse_sampler
    a = eng.generate(px, ids, **kw)
    eng.fuse_sampler = False
    b = eng.generate(px, ids, **kw)
    assert torch.equal(a, b)
    eng.fuse_sampler = True
    a = eng.generate(px, ids, max_new_tokens=9, do_sample=False, pad_token_id=olc.pad_id, eos_token_id=None)
    eng.fuse_sampler = True
    # stopping: the kernel path asks "all rows done?" only every 8th token and trims to HF's stopping point afterwards
    for row, at in ((0, 4), (1, 0), (2, 7)):
        eos = int(a[row, at])
        kw1 = dict(max_new_tokens=9, do_sample=False, pad_token_id=olc.pad_id, eos_token_id=eos)
        one = eng.generate(px[row:row + 1], ids[row:row + 1], **kw1)
        eng.fuse_sampler = False
        ref = eng.generate(px[row:row + 1], ids[row:row + 1], **kw1)
        eng.fuse_sampler = True
        assert one.shape == ref.shape and torch.equal(one, ref), (row, at, one.tolist(), ref.tolist())
        assert one.shape[1] <= 9 and int(one[0, -1]) == eos or one.shape[1] == 9
    both = eng.generate(px, ids, max_new_tokens=9, do_sample=False, pad_token_id=olc.pad_id, eos_token_id=int(a[0, 3]))
    eng.fuse_sampler = False
    both_ref = eng.generate(px, ids, max_new_tokens=9, do_sample=False, pad_token_id=olc.pad_id, eos_token_id=int(a[0, 3]))
    eng.fuse_sampler = True
    assert both.shape == both_ref.shape and torch.equal(both, both_ref)
    s = eng.generate(px, ids, max_new_tokens=9, do_sample=True, temperature=0.7, top_p=0.9, pad_token_id=olc.pad_id,
                     eos_token_id=1, generator=torch.Generator(device=cuda_dev).manual_seed(1))
    s2 = eng.generate(px, ids, max_new_tokens=9, do_sample=True, temperature=0.7, top_p=0.9, pad_token_id=olc.pad_id,
                      eos_token_id=1, generator=torch.Generator(device=cuda_dev).manual_seed(1))
    assert torch.equal(s, s2) and int(s.min()) >= 0 and int(s.max()) < olc.vocab
