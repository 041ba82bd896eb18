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
        assert t.path == nv.PATH_MMA_EXACT        # two queries, small k: the tensor-core scan streams the base faster
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 2, rng), 100, t)
        assert t.path == nv.PATH_GEMV
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 3, rng), 100, t)
        assert t.path == nv.PATH_MMA_EXACT
    b8, s8 = oracle.synth(I8, 2, 0, n, d)
    with nv.FlatIndexCuda.from_host(b8, s8, I8) as idx:       # int8: the row-tile kernel from one query on (host calls)
        t = nv.Timing()
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 1, rng), 10, t)
        assert t.path == nv.PATH_MMA_EXACT
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 1, rng), 100, t)
        assert t.path == nv.PATH_GEMV
        idx.search_topk_dot_batch(make_queries(oracle, 2, n, d, 2, rng), 10, t)
        assert t.path == nv.PATH_MMA_EXACT
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


# ---- PATH_MMA_EXACT: tensor-core candidates + exact re-score + certificate + streaming-scan fallback -------------
# Bar: ids and FP32 score bits identical to the reference-order oracle (same bar as the streaming scan).

def check_exact(nv, oracle, dtype, base, scales, queries, k, row_offset=0):
    with nv.FlatIndexCuda.from_host(base, scales, dtype, row_offset=row_offset) as idx:
        idx.set_path(nv.PATH_MMA_EXACT)
        ids, sc = idx.search_topk_dot_batch(queries, k)
        stats = idx.exact_stats()
    oids, osc = oracle.search_batch(dtype, base, scales, queries, k)
    kk = min(k, base.shape[0])
    for qi in range(queries.shape[0]):
        assert_topk_equal(ids[qi, :kk], sc[qi, :kk], oids[qi, :kk] + row_offset, osc[qi, :kk], what=f"q{qi}")
        if kk < k:
            assert np.all(ids[qi, kk:] == np.uint64(0xFFFFFFFFFFFFFFFF)) and np.all(np.isneginf(sc[qi, kk:]))
    assert stats[0] == queries.shape[0]
    return stats


@pytest.mark.parametrize("dt", ["f16", "i8"])
@pytest.mark.parametrize("n,d", [(5000, 768), (1000, 384), (257, 64), (40000, 128), (700, 80)])
@pytest.mark.parametrize("k", [1, 10, 13, 100])
def test_exact_path_bit_exact_vs_oracle(nv, oracle, dt, n, d, k):
    dtype = F16 if dt == "f16" else I8
    base, scales = oracle.synth(dtype, 11, 0, n, d)
    rng = np.random.default_rng(5)
    for nq in (1, 4, 127, 129):
        queries = make_queries(oracle, 11, n, d, nq, rng)
        check_exact(nv, oracle, dtype, base, scales, queries, k)


@pytest.mark.parametrize("dt", ["f16", "i8"])
def test_exact_path_duplicates_force_fallback(nv, oracle, dt):
    """32 identical copies of every row: the candidate list cannot separate the k-th score from the excluded rows, so the
    certificate must fail and the streaming-scan fallback must still return the canonical (score desc, id asc) answer."""
    dtype = F16 if dt == "f16" else I8
    b, s = oracle.synth(dtype, 9, 0, 300, 128)
    base = np.concatenate([b] * 32)
    scales = None if s is None else np.concatenate([s] * 32)
    queries = oracle.synth_f32(9, 0, 300, 128)[[5, 77, 100, 299, 3]]
    with nv.FlatIndexCuda.from_host(base, scales, dtype) as idx:
        idx.set_path(nv.PATH_MMA_EXACT)
        ids, sc = idx.search_topk_dot_batch(queries, 10)
        nquery, nfall = idx.exact_stats()
    oids, osc = oracle.search_batch(dtype, base, scales, queries, 10)
    assert np.array_equal(ids, oids)
    assert np.array_equal(sc.view(np.uint32), osc.view(np.uint32))
    assert nquery == 5 and nfall == 5


def test_exact_path_partial_fallback_and_row_offset(nv, oracle):
    """a batch where only SOME queries hit duplicated rows: flagged ones are gathered, re-run and scattered back."""
    n, d = 4000, 128
    base, _ = oracle.synth(F16, 4, 0, n, d)
    base = base.copy()
    base[100:140] = base[100]            # 40 copies of row 100
    f = oracle.synth_f32(4, 0, n, d)
    queries = np.ascontiguousarray(f[[7, 100, 2000, 100, 3999, 50]])
    nquery, nfall = check_exact(nv, oracle, F16, base, None, queries, 10, row_offset=5_000_000_000)
    assert nfall == 2


def test_exact_path_certificate_tiers(nv, oracle):
    """40 copies of a row defeat the first certificate (k' = 16) but not the second (k' = 64); 90 copies defeat both and
    end on the streaming scan. Either way, and with the second tier switched off, the answer is the oracle's."""
    n, d, k = 4000, 128, 10
    base, _ = oracle.synth(F16, 4, 0, n, d)
    base = base.copy()
    base[100:140] = base[100]
    base[1000:1090] = base[1000]
    f = oracle.synth_f32(4, 0, n, d)
    queries = np.ascontiguousarray(f[[7, 100, 2000, 1000, 3999, 100]])
    oids, osc = oracle.search_batch(F16, base, None, queries, k)
    with nv.FlatIndexCuda.from_host(base, None, F16, row_offset=77) as idx:
        idx.set_path(nv.PATH_MMA_EXACT)
        for tier2, want in ((1, (3, 1)), (0, (0, 3))):
            idx.set_tuning(nv.capi.TUNE_EXACT_TIER2, tier2)
            ids, sc = idx.search_topk_dot_batch(queries, k)
            assert idx.exact_tiers() == want
            assert idx.exact_stats(reset=True) == (6, 3)
            assert np.array_equal(ids, oids + 77)
            assert np.array_equal(sc.view(np.uint32), osc.view(np.uint32))
        # a batch of more than one 128-query tile, half of it flagged
        idx.set_tuning(nv.capi.TUNE_EXACT_TIER2, 1)
        many = np.ascontiguousarray(np.concatenate([queries] * 30))        # 180 queries, 90 flagged
        ids, sc = idx.search_topk_dot_batch(many, k)
        assert idx.exact_tiers() == (90, 30)
        assert np.array_equal(ids, np.concatenate([oids + 77] * 30))
        assert np.array_equal(sc.view(np.uint32), np.concatenate([osc] * 30).view(np.uint32))


@pytest.mark.parametrize("k", [13, 100])
def test_exact_path_second_tier_large_k(nv, oracle, k):
    """k > 12: the second tier's lists are the widest the scan's shared memory holds (heap lists)"""
    n, d = 6000, 128
    base, _ = oracle.synth(F16, 8, 0, n, d)
    base = base.copy()
    m = k + k // 4 + 12            # more copies than the first tier's k' - k slack, fewer than the second's
    base[500:500 + m] = base[500]
    f = oracle.synth_f32(8, 0, n, d)
    base[3000] = oracle.to_f16(f[500:501])[0]       # a far-away copy as well
    queries = np.ascontiguousarray(f[[500, 10, 500, 4000]])
    oids, osc = oracle.search_batch(F16, base, None, queries, k)
    with nv.FlatIndexCuda.from_host(base, None, F16) as idx:
        idx.set_path(nv.PATH_MMA_EXACT)
        ids, sc = idx.search_topk_dot_batch(queries, k)
        nq_, nfall = idx.exact_stats()
        t2, streamed = idx.exact_tiers()
    assert np.array_equal(ids, oids)
    assert np.array_equal(sc.view(np.uint32), osc.view(np.uint32))
    assert nq_ == 4 and nfall >= 2 and t2 == nfall and streamed <= nfall


def test_exact_path_random_data_rarely_falls_back(nv, oracle):
    n, d = 200_000, 128
    with nv.FlatIndexCuda.synthetic(F16, n, d, 77) as idx:
        base, _ = idx.read_rows(0, n)
        rng = np.random.default_rng(3)
        queries = make_queries(oracle, 77, n, d, 64, rng)
        idx.set_path(nv.PATH_MMA_EXACT)
        ids, sc = idx.search_topk_dot_batch(queries, 10)
        nquery, nfall = idx.exact_stats()
        idx.set_path(nv.PATH_GEMV)
        gids, gsc = idx.search_topk_dot_batch(queries, 10)
    assert np.array_equal(ids, gids) and np.array_equal(sc.view(np.uint32), gsc.view(np.uint32))
    assert nquery == 64 and nfall <= 16, f"{nfall} of {nquery} queries fell back"


def test_exact_path_adversarial_ascending_groups(nv, oracle):
    n, d = 6000, 64
    base = np.zeros((n, d), np.float16)
    base[:, 0] = (np.arange(n) // 4).astype(np.float16)
    q = np.zeros((5, d), np.float32)
    q[:, 0] = 1.0
    check_exact(nv, oracle, F16, base.view(np.uint16), None, q, 10)


@pytest.mark.parametrize("dt", ["f16", "i8"])
@pytest.mark.parametrize("k", [10, 100])
def test_threshold_seeding_does_not_change_results(nv, oracle, dt, k):
    """the seeded two-pass scan (first ~1/16 of the rows -> admission thresholds for the rest) must return exactly what
    the single pass returns, on every tensor-core path; forced on here (knob = 2) because auto needs >= 32 tiles per CTA."""
    dtype = F16 if dt == "f16" else I8
    n, d = 100_000, 64
    rng = np.random.default_rng(8)
    with nv.FlatIndexCuda.synthetic(dtype, n, d, 5) as idx:
        base, scales = idx.read_rows(0, n)
        f = oracle.synth_f32(5, 0, n, d)
        for nq in (4, 300):
            queries = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32))
            got = {}
            for seed in (0, 2):
                idx.set_tuning(nv.capi.TUNE_MMA_SEED, seed)
                for path in (nv.PATH_MMA, nv.PATH_MMA_EXACT) + ((nv.PATH_MMA_FAST,) if dtype == F16 else ()):
                    idx.set_path(path)
                    got[(seed, path)] = idx.search_topk_dot_batch(queries, k)
            for path in (nv.PATH_MMA, nv.PATH_MMA_EXACT) + ((nv.PATH_MMA_FAST,) if dtype == F16 else ()):
                a, b = got[(0, path)], got[(2, path)]
                assert np.array_equal(a[0], b[0]), f"ids differ: path {path} nq {nq}"
                assert np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), f"scores differ: path {path} nq {nq}"
            # and the exact path still equals the oracle on a few queries
            pick = [0, nq // 2, nq - 1]
            oids, osc = oracle.search_batch(dtype, base, scales, queries[pick], k)
            for j, qi in enumerate(pick):
                assert_topk_equal(got[(2, nv.PATH_MMA_EXACT)][0][qi], got[(2, nv.PATH_MMA_EXACT)][1][qi], oids[j], osc[j],
                                  what=f"seeded exact q{qi}")


@pytest.mark.parametrize("k", [10, 100])
def test_cta_pair_kernel_matches_single_cta_kernel(nv, oracle, k):
    """cta_group::2 scan (even number of 128-query tiles) against the single-CTA scan and the oracle."""
    n, d = 60_000, 128
    rng = np.random.default_rng(12)
    with nv.FlatIndexCuda.synthetic(F16, n, d, 6) as idx:
        base, _ = idx.read_rows(0, n)
        f = oracle.synth_f32(6, 0, n, d)
        for nq in (256, 200, 1000, 300, 600):       # G = 2, 2 (padded tile), 8, 3 -> 4 and 5 -> 6 (an empty tile added)
            queries = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32))
            got = {}
            for pair in (0, 2):
                idx.set_tuning(nv.capi.TUNE_MMA_2CTA, pair)
                for path in (nv.PATH_MMA_EXACT, nv.PATH_MMA, nv.PATH_MMA_FAST):
                    idx.set_path(path)
                    got[(pair, path)] = idx.search_topk_dot_batch(queries, k)
            a, b = got[(0, nv.PATH_MMA_EXACT)], got[(2, nv.PATH_MMA_EXACT)]
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), f"exact nq={nq}"
            for path, rtol in ((nv.PATH_MMA, RTOL_F16), (nv.PATH_MMA_FAST, RTOL_F16_FAST)):
                s0, s1 = got[(0, path)][1], got[(2, path)][1]
                assert np.max(np.abs(s0 - s1)) <= rtol * np.abs(s0).max(), f"path {path} nq={nq}"
            pick = [0, nq // 3, nq - 1]
            oids, osc = oracle.search_batch(F16, base, None, queries[pick], k)
            for j, qi in enumerate(pick):
                assert_topk_equal(b[0][qi], b[1][qi], oids[j], osc[j], what=f"pair exact nq={nq} q{qi}")


def test_i8_resident_query_limb_equals_streamed_limbs(nv, oracle):
    """i8 tensor scan with query limb 0 resident in shared memory (default, k <= 16) vs both limbs streamed."""
    n, d = 50_000, 768
    rng = np.random.default_rng(4)
    with nv.FlatIndexCuda.synthetic(I8, n, d, 8) as idx:
        f = oracle.synth_f32(8, 0, n, d)
        for nq in (5, 300):
            queries = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32))
            got = {}
            for res in (0, 1):
                idx.set_tuning(nv.capi.TUNE_MMA_RESIDENT, res)
                for path in (nv.PATH_MMA, nv.PATH_MMA_EXACT):
                    idx.set_path(path)
                    got[(res, path)] = idx.search_topk_dot_batch(queries, 10)
            for path in (nv.PATH_MMA, nv.PATH_MMA_EXACT):
                a, b = got[(0, path)], got[(1, path)]
                assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), (path, nq)


@pytest.mark.parametrize("k", [10, 40])
def test_i8_cta_pair_kernel_matches_single_cta_kernel(nv, oracle, k):
    """i8 cta_group::2 scan (128-row tiles, two limb accumulators, each CTA loads 64 rows) vs the single-CTA scan:
    same integer accumulators and the same FP32 combine, so ids and score bits must be identical — with the first query
    limb resident or streamed, dims with 1 and 6 K-chunks, ragged last tile."""
    rng = np.random.default_rng(21)
    for n, d in ((60_001, 128), (20_011, 768)):
        with nv.FlatIndexCuda.synthetic(I8, n, d, 9) as idx:
            base, scales = idx.read_rows(0, n)
            f = oracle.synth_f32(9, 0, n, d)
            for nq in (256, 130, 1024, 330):   # G = 2, 2 (padded tile), 8, 3 -> 4 (an empty tile added)
                queries = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32))
                got = {}
                for pair, res in ((0, 1), (2, 1), (2, 2), (2, 0)):   # res 1: first limb resident, 2: both, 0: none
                    idx.set_tuning(nv.capi.TUNE_MMA_2CTA, pair)
                    idx.set_tuning(nv.capi.TUNE_MMA_RESIDENT, res)
                    for path in (nv.PATH_MMA, nv.PATH_MMA_EXACT):
                        idx.set_path(path)
                        got[(pair, res, path)] = idx.search_topk_dot_batch(queries, k)
                for path in (nv.PATH_MMA, nv.PATH_MMA_EXACT):
                    a = got[(0, 1, path)]
                    for key in ((2, 1, path), (2, 2, path), (2, 0, path)):
                        b = got[key]
                        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), (key, nq, n, d)
                pick = [0, nq // 2, nq - 1]
                b = got[(2, 1, nv.PATH_MMA_EXACT)]
                oids, osc = oracle.search_batch(I8, base, scales, queries[pick], k)
                for j, qi in enumerate(pick):
                    assert_topk_equal(b[0][qi], b[1][qi], oids[j], osc[j], what=f"i8 pair exact nq={nq} q{qi}")
