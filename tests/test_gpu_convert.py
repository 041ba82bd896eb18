"""GPU parity for the device converters through the C ABI: outputs must be bit-identical to
the unmodified reference tools (golden vectors made by nvdb_convert_f16 / nvdb_quantize_i8, tests/golden/scan_*.npz) and
to the oracle's restatement on larger random inputs, including special values."""
from pathlib import Path

import numpy as np
import pytest

from _oracle import F16, F32, I8

pytestmark = pytest.mark.gpu
GOLDEN = Path(__file__).resolve().parent / "golden"


@pytest.mark.parametrize("case", ["scan_n96_d768.npz", "scan_n1024_d32.npz", "scan_n300_d384_k100.npz"])
def test_convert_matches_reference_tool_outputs(nv, oracle, case):
    """golden f16 / i8 (+scales) payloads written by the unmodified nvdb_convert_f16 / nvdb_quantize_i8
    (tests/golden/make_golden.py) from the generator's f32 rows"""
    g = np.load(GOLDEN / case)
    f32 = oracle.synth_f32(int(g["seed"]), 0, int(g["n"]), int(g["dim"]))
    h, _ = nv.capi.convert(f32, F32, F16)
    assert np.array_equal(h, g["f16"])
    q, sc = nv.capi.convert(f32, F32, I8)
    assert np.array_equal(q, g["i8"])
    assert np.array_equal(sc.view(np.uint32), g["i8_scales"].view(np.uint32))


@pytest.mark.parametrize("n,d", [(1, 1), (1000, 768), (4097, 100), (300_000, 64)])
def test_convert_matches_oracle_on_random_rows(nv, oracle, n, d):
    rng = np.random.default_rng(n + d)
    x = (rng.standard_normal((n, d)) * rng.choice([1e-6, 1e-3, 1.0, 300.0, 7e4], size=(n, 1))).astype(np.float32)
    x[rng.integers(0, n, max(1, n // 50))] = 0.0                       # all-zero rows: scale must be 1
    flat = x.reshape(-1)
    flat[rng.integers(0, flat.size, max(1, flat.size // 97))] *= -1.0
    h, _ = nv.capi.convert(x, F32, F16)
    assert np.array_equal(h, oracle.to_f16(x))
    q, sc = nv.capi.convert(x, F32, I8)
    oq, osc = oracle.quantize_i8(x)
    assert np.array_equal(q, oq) and np.array_equal(sc.view(np.uint32), osc.view(np.uint32))
    # f16 source: the reference widens exactly, then quantises the same way (apps/nvdb_quantize_i8.cpp:67-73)
    hf = h.view(np.float16).astype(np.float32)
    finite = np.isfinite(hf).all(axis=1)
    q2, sc2 = nv.capi.convert(np.ascontiguousarray(h[finite]), F16, I8)
    oq2, osc2 = oracle.quantize_i8(np.ascontiguousarray(hf[finite]))
    assert np.array_equal(q2, oq2) and np.array_equal(sc2.view(np.uint32), osc2.view(np.uint32))


def test_convert_rejects_unsupported_pairs(nv):
    x = np.zeros((4, 8), np.float32)
    with pytest.raises(nv.NvdbError):
        nv.capi.convert(x, F32, F32)
    with pytest.raises(nv.NvdbError):
        nv.capi.convert(x.astype(np.int8), I8, F16)
