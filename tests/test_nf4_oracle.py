"""oracle/nf4_oracle.py: the published known answers and the structural guarantees of bitsandbytes' double-quantised
NF4 format (the oracle is otherwise UNPINNED: bitsandbytes is not available here — see its header).  CPU only."""
import numpy as np
import torch

from oracle import nf4_oracle as N

# QLoRA (Dettmers et al. 2023), appendix E "NormalFloat 4-bit data type": the exact values, as printed
QLORA_APPENDIX_E = [-1.0, -0.6961928009986877, -0.5250730514526367, -0.39491748809814453, -0.28444138169288635,
                    -0.18477343022823334, -0.09105003625154495, 0.0, 0.07958029955625534, 0.16093020141124725,
                    0.24611230194568634, 0.33791524171829224, 0.44070982933044434, 0.5626170039176941,
                    0.7229568362236023, 1.0]
# csrc/kernels.cu dQuantizeNF4: the literals of its decision tree, in ascending order
DQUANTIZE_NF4_THRESHOLDS = [-0.8480964004993439, -0.6106329262256622, -0.4599952697753906, -0.33967943489551544,
                            -0.23460740596055984, -0.13791173323988914, -0.045525018125772476, 0.03979014977812767,
                            0.1202552504837513, 0.2035212516784668, 0.2920137718319893, 0.3893125355243683,
                            0.5016634166240692, 0.6427869200706482, 0.8614784181118011]


def _bf16(x):
    return torch.from_numpy(np.asarray(x, np.float32)).to(torch.bfloat16).float().numpy()


def test_code_books():
    assert np.array_equal(N.NF4_TABLE, np.array(QLORA_APPENDIX_E, dtype=np.float32))
    assert np.array_equal(N.NF4_MIDPOINTS, np.array(DQUANTIZE_NF4_THRESHOLDS, dtype=np.float32))
    m = N.dynamic_map()
    assert m.shape == (256,) and m.dtype == np.float32 and (np.diff(m) > 0).all()
    assert m[-1] == 1.0 and m[0] > -1.0 and (m == 0).sum() == 1          # asymmetric: +1.0 is a code, -1.0 is not
    assert np.array_equal(np.sort(-m[(m != 0) & (m != 1.0)]), m[(m != 0) & (m != 1.0)])      # otherwise sign-symmetric
    assert np.isclose(m[m > 0].min(), 0.55e-6, rtol=1e-6)                # first decade: one bucket, mean of [0.1, 1]
    for i in range(7):                                                   # 2^i codes per decade and sign
        lo, hi = 0.1 * 10.0 ** (-6 + i), 1.0 * 10.0 ** (-6 + i)
        assert ((m > lo) & (m < hi)).sum() == 2 ** i, i


def test_quantize_picks_the_nearest_code_and_packs_high_nibble_first():
    rng = np.random.default_rng(0)
    w = _bf16(rng.standard_normal(64 * 50) * 0.02)
    packed, absmax = N.quantize_nf4(w)
    assert packed.shape == (w.size // 2,) and absmax.shape == (50,)
    assert np.array_equal(absmax, np.abs(w.reshape(50, 64)).max(1))
    codes = np.stack([packed >> 4, packed & 15], 1).reshape(50, 64)
    x = w.reshape(50, 64) * (np.float32(1) / absmax)[:, None]
    nearest = np.abs(x[..., None] - N.NF4_TABLE).argmin(-1)
    assert (codes == nearest).mean() > 0.999                             # differs only on exact midpoint ties
    assert np.abs(N.NF4_TABLE[codes] - x).max() <= np.diff(N.NF4_TABLE).max() / 2 + 1e-6
    # every block holds its own extreme exactly: code 0 (-1) or 15 (+1)
    assert ((codes == 0) | (codes == 15)).any(1).all()
    # all-zero block: codes 0 and a zero absmax de-quantise to zero
    p0, a0 = N.quantize_nf4(np.zeros(64, np.float32))
    assert (p0 == 0).all() and a0[0] == 0


def test_double_quant_and_roundtrip_error_budget():
    rng = np.random.default_rng(1)
    w = _bf16(rng.standard_normal((640, 1152)) * 0.03)
    packed, absmax = N.quantize_nf4(torch.from_numpy(w).to(torch.float16).float().numpy().reshape(-1))
    q, absmax2, offset = N.double_quant(absmax)
    assert q.dtype == np.uint8 and absmax2.shape == ((absmax.size + 255) // 256,)
    assert np.isclose(offset, absmax.mean(), rtol=1e-6)
    code = N.dynamic_map()
    rec = code[q] * np.repeat(absmax2, 256)[:absmax.size] + offset
    assert np.abs(rec - absmax).max() / absmax.mean() < 0.02             # 8-bit statistics: a per-cent-level error
    r = N.roundtrip(w)
    assert r.shape == w.shape and np.array_equal(r, _bf16(r))
    rel = np.linalg.norm(r - w) / np.linalg.norm(w)
    assert 0.07 < rel < 0.11, rel                                        # Gaussian weights: about 9 %
    rr = N.roundtrip(r)                                                  # already on the grid: a second pass only
    assert np.linalg.norm(rr - r) / np.linalg.norm(r) < 0.01             # re-rounds the 8-bit statistics
    r2 = N.roundtrip(w, via_fp16=False)
    # the fp16 detour double-rounds: about 1 value in 16 lands one bf16 ulp away, nothing more
    assert 0.01 < (r2 != r).mean() < 0.1 and np.linalg.norm(r2 - r) / np.linalg.norm(r) < 2e-3


def test_dquantize8_matches_nearest_code_search():
    code = N.dynamic_map()
    rng = np.random.default_rng(2)
    x = np.concatenate([rng.uniform(-1, 1, 20000), code, [-1.0, 1.0, 0.0, 1e-9, -1e-9]]).astype(np.float32)
    got = N._dquantize8(code, x)
    err = np.abs(code[got] - x)
    best = np.abs(code[None, :] - x[:, None]).min(1)
    assert np.all(err <= best + 1e-7)
    assert np.array_equal(code[got[20000:20256]], code)                  # every code maps to itself
