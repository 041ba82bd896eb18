"""GPU parity for the tensor-core (tcgen05) scan through the C ABI.

  * raw accumulators (test hook): i8 limb dot products are EXACT int32 (== numpy integer matmul); f16 accumulators
    equal the real-number dot of the fp16 query limbs with the fp16 base to FP32-accumulation accuracy;
  * i8 search: ids and FP32 score bits identical to the CPU restatement of the limb scheme (oracle
    orc_score_all_i8_limbs + canonical (score desc, id asc) order), and within 5e-5 of the reference AVX2 scores;
  * f16 search: scores within RTOL_F16 of the reference-order FP32 scores, ids identical except near-ties
    (north star: 1e-3 relative; the two-limb path is held to 2e-5, the one-limb fast path to 1e-3).
"""
import numpy as np
import pytest

from _oracle import F16, F32, I8, assert_topk_equal

pytestmark = pytest.mark.gpu
RTOL_F16 = 2e-5
RTOL_F16_FAST = 1e-3
RTOL_I8_VS_REFERENCE = 5e-5


def make_queries(oracle, seed, n, d, nq, rng):
    f = oracle.synth_f32(seed, 0, n, d)
    return np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32))


def canonical_topk(scores, k):
    """top-k of one score row under (score desc, id asc)"""
    ids = np.arange(scores.size, dtype=np.uint64)
    order = np.lexsort((ids, -scores.astype(np.float64)))[:k]
    return ids[order], scores[order]


@pytest.mark.parametrize("n,d,nq", [(1000, 768, 5), (257, 64, 130), (300, 80, 3), (2000, 384, 128)])
def test_i8_limb_dot_products_are_exact(nv, oracle, n, d, nq):
    base, scales = oracle.synth(I8, 3, 0, n, d)
    rng = np.random.default_rng(0)
    queries = make_queries(oracle, 3, n, d, nq, rng)
    with nv.FlatIndexCuda.from_host(base, scales, I8) as idx:
        acc = idx.debug_mma_accumulators(queries)
    for qi in range(nq):
        q1, q2, s1, s2 = oracle.query_limbs_i8(queries[qi])
        assert np.array_equal(acc[0, qi], base.astype(np.int32) @ q1.astype(np.int32)), f"I1 q{qi}"
        assert np.array_equal(acc[1, qi], base.astype(np.int32) @ q2.astype(np.int32)), f"I2 q{qi}"


@pytest.mark.parametrize("path", ["mma", "fast"])
@pytest.mark.parametrize("n,d,nq", [(1000, 768, 5), (257, 64, 130), (300, 72, 3), (2000, 384, 128)])
def test_f16_accumulators_match_limb_dot(nv, oracle, path, n, d, nq):
    base, _ = oracle.synth(F16, 3, 0, n, d)
    rng = np.random.default_rng(0)
    queries = make_queries(oracle, 3, n, d, nq, rng)
    with nv.FlatIndexCuda.from_host(base, None, F16) as idx:
        acc = idx.debug_mma_accumulators(queries, nv.PATH_MMA if path == "mma" else nv.PATH_MMA_FAST)
    x = base.view(np.float16).astype(np.float64)
    for qi in range(nq):
        hi, lo, qs = oracle.query_limbs_f16(queries[qi])
        a = hi.view(np.float16).astype(np.float64)
        if path == "mma":
            a = a + lo.view(np.float16).astype(np.float64)
        want = x @ a
        np.testing.assert_allclose(acc[qi], want, rtol=0, atol=4e-6 * np.abs(x).max() * np.abs(a).sum(), err_msg=f"q{qi}")


@pytest.mark.parametrize("n,d", [(5000, 768), (1000, 384), (257, 64), (40000, 128), (700, 80)])
@pytest.mark.parametrize("k", [1, 10, 100])
def test_i8_search_bit_exact_vs_limb_restatement(nv, oracle, n, d, k):
    base, scales = oracle.synth(I8, 11, 0, n, d)
    rng = np.random.default_rng(5)
    with nv.FlatIndexCuda.from_host(base, scales, I8) as idx:
        idx.set_path(nv.PATH_MMA)
        for nq in (1, 4, 127, 129):
            queries = make_queries(oracle, 11, n, d, nq, rng)
            ids, sc = idx.search_topk_dot_batch(queries, k)
            kk = min(k, n)
            for qi in range(0, nq, max(1, nq // 7)):
                lsc, _, _ = oracle.score_all_i8_limbs(base, scales, queries[qi])
                oid, osc = canonical_topk(lsc, kk)
                assert np.array_equal(ids[qi, :kk], oid), f"nq={nq} q{qi}"
                assert np.array_equal(sc[qi, :kk].view(np.uint32), osc.view(np.uint32)), f"nq={nq} q{qi}"
                # and close to the reference's own (FP32 query) scores
                ref = oracle.score_all(I8, base, scales, queries[qi])
                assert np.max(np.abs(sc[qi, :kk] - ref[oid.astype(np.int64)])) <= RTOL_I8_VS_REFERENCE * np.abs(ref).max()
                if kk < k:
                    assert np.all(ids[qi, kk:] == np.uint64(0xFFFFFFFFFFFFFFFF)) and np.all(np.isneginf(sc[qi, kk:]))


@pytest.mark.parametrize("path", ["mma", "fast"])
@pytest.mark.parametrize("n,d", [(5000, 768), (1000, 384), (257, 64), (40000, 128), (700, 72)])
@pytest.mark.parametrize("k", [1, 10, 100])
def test_f16_search_within_tolerance_of_reference_scores(nv, oracle, path, n, d, k):
    base, _ = oracle.synth(F16, 11, 0, n, d)
    rng = np.random.default_rng(5)
    rtol = RTOL_F16 if path == "mma" else RTOL_F16_FAST
    near_ties = 0
    with nv.FlatIndexCuda.from_host(base, None, F16) as idx:
        idx.set_path(nv.PATH_MMA if path == "mma" else nv.PATH_MMA_FAST)
        for nq in (1, 4, 127, 129):
            queries = make_queries(oracle, 11, n, d, nq, rng)
            ids, sc = idx.search_topk_dot_batch(queries, k)
            kk = min(k, n)
            pick = list(range(0, nq, max(1, nq // 7)))
            oids, osc = oracle.search_batch(F16, base, None, queries[pick], k)
            for j, qi in enumerate(pick):
                # scores are compared relative to the largest score magnitude of the query (dot products cancel)
                scale = np.abs(osc[j, :kk]).max()
                assert np.max(np.abs(sc[qi, :kk] - osc[j, :kk])) <= rtol * scale, f"nq={nq} q{qi}"
                bad = np.nonzero(ids[qi, :kk] != oids[j, :kk])[0]
                for b in bad:   # documented near-ties: the two ids' reference scores differ by < 2*rtol*scale
                    grp = np.abs(osc[j, :kk] - osc[j, b]) <= 2 * rtol * scale
                    assert ids[qi, b] in oids[j, :kk][grp] or b == kk - 1, f"nq={nq} q{qi} pos {b}"
                    near_ties += 1
    if path == "mma":
        assert near_ties <= 2


def test_auto_path_crossover_and_timing_path(nv, oracle):
    n, d = 3000, 128
    base, _ = oracle.synth(F16, 2, 0, n, d)
    rng = np.random.default_rng(1)
    with nv.FlatIndexCuda.from_host(base, None, F16) as idx:
        t = nv.Timing()
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 1, rng), 10, t)
        assert t.path == nv.PATH_GEMV
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 2, rng), 10, t)
        assert t.path == nv.PATH_MMA
    f32 = oracle.synth_f32(2, 0, n, d)
    with nv.FlatIndexCuda.from_host(f32, None, F32) as idx:   # FP32 base: no tensor path (tf32 excluded), GEMV always
        t = nv.Timing()
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 16, rng), 10, t)
        assert t.path == nv.PATH_GEMV
        idx.set_path(nv.PATH_MMA)
        with pytest.raises(nv.NvdbError):
            idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 16, rng), 10)


def test_mma_adversarial_order_and_ties(nv, oracle):
    """scores ascending with the row id (every row beats the running threshold) and exact duplicates."""
    n, d = 6000, 64
    base = np.zeros((n, d), np.float16)
    base[:, 0] = (np.arange(n) // 4).astype(np.float16)      # groups of 4 equal scores, ascending
    q = np.zeros((5, d), np.float32)
    q[:, 0] = 1.0
    with nv.FlatIndexCuda.from_host(base.view(np.uint16), None, F16) as idx:
        idx.set_path(nv.PATH_MMA)
        ids, sc = idx.search_topk_dot_batch(q, 10)
    oids, osc = oracle.search_batch(F16, base.view(np.uint16), None, q, 10)
    assert np.array_equal(ids, oids) and np.array_equal(sc, osc)
