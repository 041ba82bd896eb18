"""CPU tests (no GPU): pin the C restatement (oracle/nvdb_oracle.c) against
  (a) golden vectors produced by the UNMODIFIED reference binaries (tests/golden/*.npz, made by
      tests/golden/make_golden.py in the build container), and
  (b) the compiled reference itself (oracle/_ref) when it is present on this machine.
"""
import hashlib
from pathlib import Path

import numpy as np
import pytest

from _oracle import F16, F32, I8, assert_topk_equal, read_vecbin, write_raw12, write_vecbin

GOLD = Path(__file__).resolve().parent / "golden"
SCAN_CASES = sorted(p.name for p in GOLD.glob("scan_*.npz"))
DT = {"f32": F32, "f16": F16, "i8": I8}


def test_golden_files_present():
    assert len(SCAN_CASES) >= 3 and (GOLD / "dots.npz").exists()


# ------------------------------------------------------------------ (a) golden vectors
def test_dot_kernels_match_reference_golden_bits(oracle):
    """orc_dot_* == nvdb::dot_f32 / dot_f32_f16base / dot_f32_i8base, bit for bit, incl. ragged dims (tails)."""
    g = np.load(GOLD / "dots.npz")
    for d in g["dims"]:
        q, x32, x16, x8, sc = (g[f"d{d}_{s}"] for s in ("q", "x32", "x16", "x8", "scale"))
        for i in range(q.shape[0]):
            assert oracle.dot(F32, q[i], x32[i]).view(np.uint32) == g[f"d{d}_r32"][i].view(np.uint32), f"f32 d={d}"
            assert oracle.dot(F16, q[i], x16[i]).view(np.uint32) == g[f"d{d}_r16"][i].view(np.uint32), f"f16 d={d}"
            assert oracle.dot(I8, q[i], x8[i], float(sc[i])).view(np.uint32) == g[f"d{d}_r8"][i].view(np.uint32), f"i8 d={d}"


@pytest.mark.parametrize("case", SCAN_CASES)
def test_converters_match_reference_tools(oracle, case):
    """orc_convert_f16 == nvdb_convert_f16, orc_quantize_i8 == nvdb_quantize_i8 (payload and scales, bit-exact)."""
    g = np.load(GOLD / case)
    f32 = oracle.synth_f32(int(g["seed"]), 0, int(g["n"]), int(g["dim"]))
    assert hashlib.sha256(f32.tobytes()).hexdigest() == str(g["f32_sha256"]), "generator drifted"
    assert np.array_equal(oracle.to_f16(f32), g["f16"])
    i8, scales = oracle.quantize_i8(f32)
    assert np.array_equal(i8, g["i8"])
    assert np.array_equal(scales.view(np.uint32), g["i8_scales"].view(np.uint32))


@pytest.mark.parametrize("case", SCAN_CASES)
@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
def test_flat_search_matches_reference_golden(oracle, case, dt):
    """orc_flat_search{,_mt,_batch} == FlatIndex::search_topk_dot: ids == nvdb_gt_build's gtbin, score bits equal."""
    g = np.load(GOLD / case)
    n, d, k = int(g["n"]), int(g["dim"]), int(g["k"])
    f32 = oracle.synth_f32(int(g["seed"]), 0, n, d)
    base = {"f32": f32, "f16": g["f16"], "i8": g["i8"]}[dt]
    scales = g["i8_scales"] if dt == "i8" else None
    gt_ids, gt_sc = g[f"gt_ids_{dt}"], g[f"scores_{dt}"]
    bids, bsc = oracle.search_batch(DT[dt], base, scales, g["queries"], k)
    for qi, q in enumerate(g["queries"]):
        for mt in (False, True):
            ids, sc = oracle.search(DT[dt], base, scales, q, k, mt=mt)
            assert_topk_equal(ids, sc, gt_ids[qi], gt_sc[qi], what=f"{case} {dt} q{qi} mt={mt}")
        assert_topk_equal(bids[qi, :len(gt_ids[qi])], bsc[qi, :len(gt_ids[qi])], gt_ids[qi], gt_sc[qi])
        assert int(gt_ids[qi, 0]) == int(ids[0])  # queries are base rows: top-1 is the row itself (f32)


# ------------------------------------------------------------------ (b) live reference (build container)
def test_dot_kernels_match_live_reference(oracle, reference):
    rng = np.random.default_rng(3)
    for d in (768, 384, 16, 8, 5, 27, 100):
        for _ in range(200):
            q = rng.standard_normal(d).astype(np.float32)
            x = rng.standard_normal(d).astype(np.float32)
            x16 = x.astype(np.float16).view(np.uint16)
            x8 = rng.integers(-127, 128, d).astype(np.int8)
            s = float(rng.uniform(1e-3, 1e-2))
            assert oracle.dot(F32, q, x).view(np.uint32) == reference.dot(F32, q, x).view(np.uint32)
            assert oracle.dot(F16, q, x16).view(np.uint32) == reference.dot(F16, q, x16).view(np.uint32)
            assert oracle.dot(I8, q, x8, s).view(np.uint32) == reference.dot(I8, q, x8, s).view(np.uint32)


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
@pytest.mark.parametrize("mode", ["st", "omp", "async", "pool"])
def test_flat_search_matches_live_reference(oracle, reference, tmp_path, dt, mode):
    n, d, k = 6000, 768, 10
    base, scales = oracle.synth(DT[dt], 77, 0, n, d)
    p = tmp_path / "b.vecbin"
    write_vecbin(p, DT[dt], base, scales)
    h = reference.load(p)
    f = oracle.synth_f32(77, 0, n, d)
    rng = np.random.default_rng(1)
    for r in rng.integers(0, n, 6):
        q = f[r] + (0.02 * rng.standard_normal(d)).astype(np.float32)
        rid, rsc = reference.search(h, q, k, mode, threads=4)
        oid, osc = oracle.search(DT[dt], base, scales, q, k)
        assert_topk_equal(oid, osc, rid, rsc, what=f"{dt} {mode}")
    reference.free(h)


def test_vecbin_files_roundtrip_through_reference_loader(oracle, reference, tmp_path):
    """our vecbin / raw12 writers produce files the reference's VectorDataset::load accepts (strict size check)."""
    f = oracle.synth_f32(5, 0, 33, 40)
    for dt in (F32, F16, I8):
        base, scales = oracle.synth(dt, 5, 0, 33, 40)
        p = tmp_path / f"x{dt}.vecbin"
        write_vecbin(p, dt, base, scales)
        h = reference.load(p)
        L = reference.lib
        assert (L.ref_dataset_count(h), L.ref_dataset_dim(h), L.ref_dataset_dtype(h)) == (33, 40, dt)
        reference.free(h)
        dt2, b2, s2 = read_vecbin(p)
        assert dt2 == dt and np.array_equal(b2, base)
    p = tmp_path / "q.raw12"
    write_raw12(p, f)
    h = reference.load(p)
    assert reference.lib.ref_dataset_count(h) == 33 and reference.lib.ref_dataset_dtype(h) == F32
    reference.free(h)
    with open(tmp_path / "bad.vecbin", "wb") as fh:
        fh.write(open(tmp_path / "x1.vecbin", "rb").read()[:-4])
    with pytest.raises(RuntimeError):
        reference.load(tmp_path / "bad.vecbin")


def test_refine_restatements_agree(oracle):
    """CPU (double) and CUDA-order (fp32, 4 chains) restatements of the refine distance agree to ~1e-6 and rank equally."""
    n, d, Q, R, K = 3000, 768, 8, 256, 10
    base, _ = oracle.synth(F16, 9, 0, n, d)
    f = oracle.synth_f32(9, 0, n, d)
    rng = np.random.default_rng(4)
    queries = f[rng.integers(0, n, Q)] + (0.05 * rng.standard_normal((Q, d))).astype(np.float32)
    cand = np.stack([rng.permutation(n)[:R] for _ in range(Q)]).astype(np.uint32)   # no duplicate ids
    cand[:, ::9] = 0xFFFFFFFF
    i0, d0 = oracle.refine(0, F16, base, None, queries, cand, K)
    i1, d1 = oracle.refine(1, F16, base, None, queries, cand, K)
    np.testing.assert_allclose(d0, d1, rtol=2e-6)
    assert (i0 == i1).mean() > 0.95
    # brute force in numpy float64
    x = base.view(np.float16).astype(np.float64)
    for qi in range(Q):
        ids = cand[qi][cand[qi] < n]
        dist = ((queries[qi].astype(np.float64) - x[ids]) ** 2).sum(1).astype(np.float32)
        order = np.lexsort((ids, dist))[:K]
        assert np.array_equal(ids[order], i0[qi])
        assert np.array_equal(dist[order], d0[qi])


def test_limb_schemes_approximate_the_reference_scores(oracle):
    """the tensor-core path's query limb schemes (restated in the oracle) vs the reference-order FP32 scores:
    i8 two-limb <= 3e-5 relative to max|score|; f16 hi+lo limbs reconstruct the scaled query to ~2^-21."""
    n, d = 4000, 768
    base, scales = oracle.synth(I8, 41, 0, n, d)
    f = oracle.synth_f32(41, 0, n, d)
    for r in (0, 17, 3999):
        q = f[r]
        ref = oracle.score_all(I8, base, scales, q)
        sc, i1, i2 = oracle.score_all_i8_limbs(base, scales, q)
        assert np.max(np.abs(sc - ref)) <= 3e-5 * np.max(np.abs(ref))
        q1, q2, s1, s2 = oracle.query_limbs_i8(q)
        assert np.array_equal(i1, base.astype(np.int32) @ q1.astype(np.int32))
        assert np.array_equal(i2, base.astype(np.int32) @ q2.astype(np.int32))
        hi, lo, qs = oracle.query_limbs_f16(q)
        rec = (hi.view(np.float16).astype(np.float64) + lo.view(np.float16).astype(np.float64)) * float(qs)
        assert np.max(np.abs(rec - q)) <= 2.0 ** -21 * np.max(np.abs(q))
