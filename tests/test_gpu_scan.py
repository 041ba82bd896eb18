"""GPU parity: the CUDA flat scan through the C ABI vs the CPU oracle on the same seeded inputs.
Bit-exact scores (FP32 summation order of the reference is reproduced), ids equal modulo exact-tie groups."""
import numpy as np
import pytest

from _oracle import F16, F32, I8, assert_topk_equal

pytestmark = pytest.mark.gpu
DT = {"f32": F32, "f16": F16, "i8": I8}


def make_base(oracle, dtype, n, d, seed=11):
    return oracle.synth(dtype, seed, 0, n, d)


def check(nv, oracle, dtype, base, scales, queries, k, row_offset=0):
    with nv.FlatIndexCuda.from_host(base, scales, dtype, row_offset=row_offset) as idx:
        idx.set_path(nv.PATH_GEMV)    # the bit-exact streaming scan (AUTO switches to tensor cores for batches)
        ids, sc = idx.search_topk_dot_batch(queries, k)
    oids, osc = oracle.search_batch(dtype, base, scales, queries, k)
    n = base.shape[0]
    kk = min(k, n)
    for qi in range(queries.shape[0]):
        assert_topk_equal(ids[qi, :kk], sc[qi, :kk], oids[qi, :kk] + row_offset, osc[qi, :kk], what=f"q{qi}")
        if kk < k:
            assert np.all(ids[qi, kk:] == np.uint64(0xFFFFFFFFFFFFFFFF))
            assert np.all(np.isneginf(sc[qi, kk:]))


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
@pytest.mark.parametrize("n,d", [(5000, 768), (1000, 384), (257, 64), (40000, 128)])
@pytest.mark.parametrize("k", [1, 10, 100])
def test_scan_matches_oracle(nv, oracle, dt, n, d, k):
    dtype = DT[dt]
    base, scales = make_base(oracle, dtype, n, d)
    f = oracle.synth_f32(11, 0, n, d)
    rng = np.random.default_rng(5)
    for nq in (1, 3, 8, 13):
        rows = rng.integers(0, n, nq)
        queries = f[rows] + (0.01 * rng.standard_normal((nq, d))).astype(np.float32)
        check(nv, oracle, dtype, base, scales, queries, k)


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
@pytest.mark.parametrize("d", [100, 13, 24, 7, 1, 40])
def test_scan_ragged_dims_generic_kernel(nv, oracle, dt, d):
    """dims with scalar tails / rows the tensor map cannot describe go through the generic kernel."""
    dtype = DT[dt]
    n = 3001
    base, scales = make_base(oracle, dtype, n, d, seed=3)
    rng = np.random.default_rng(6)
    queries = rng.standard_normal((5, d)).astype(np.float32)
    check(nv, oracle, dtype, base, scales, queries, 10)


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
def test_scan_ties_duplicated_rows(nv, oracle, dt):
    """4 identical copies of 2000 rows (SURVEY.md Appendix B tie probe): canonical (score desc, id asc) order."""
    dtype = DT[dt]
    b, s = make_base(oracle, dtype, 2000, 128, seed=9)
    base = np.concatenate([b] * 4)
    scales = None if s is None else np.concatenate([s] * 4)
    f = oracle.synth_f32(9, 0, 2000, 128)
    queries = f[[5, 77]]
    with nv.FlatIndexCuda.from_host(base, scales, dtype) as idx:
        idx.set_path(nv.PATH_GEMV)
        ids, sc = idx.search_topk_dot_batch(queries, 6)
        idx.set_path(nv.PATH_AUTO)   # a batch on AUTO takes the tensor-core scan for f16 / i8: tolerance-mode scores
        aids, asc = idx.search_topk_dot_batch(queries, 6)
    oids, osc = oracle.search_batch(dtype, base, scales, queries, 6)
    assert np.array_equal(ids, oids)  # canonical order is exactly reproducible, not only modulo ties
    assert np.array_equal(sc.view(np.uint32), osc.view(np.uint32))
    # identical rows get identical tensor-core scores too, so the (score desc, id asc) order is still exact
    assert np.array_equal(aids, oids)
    np.testing.assert_allclose(asc, osc, rtol=1e-4)


def test_scan_k_larger_than_n_and_row_offset(nv, oracle):
    base, scales = make_base(oracle, F16, 37, 64)
    q = oracle.synth_f32(11, 0, 37, 64)[:2]
    check(nv, oracle, F16, base, scales, q, 50, row_offset=10_000_000_000)


def test_scan_all_equal_scores(nv, oracle):
    base = np.ones((3000, 64), np.float32)
    q = np.ones((1, 64), np.float32)
    check(nv, oracle, F32, base, None, q, 10)


def test_scan_ascending_scores_worst_case_for_selector(nv, oracle):
    """Scores strictly increasing with the row id: every row passes the running threshold."""
    n, d = 20000, 16
    base = np.zeros((n, d), np.float32)
    base[:, 0] = np.arange(n, dtype=np.float32)
    q = np.zeros((1, d), np.float32)
    q[0, 0] = 1.0
    check(nv, oracle, F32, base, None, q, 10)
    check(nv, oracle, F32, base, None, q, 300)


def test_errors(nv):
    with pytest.raises(nv.NvdbError) as e:
        nv.FlatIndexCuda.from_host(np.zeros((0, 8), np.float32), None, F32)
    assert "Empty base" in str(e.value)
    with nv.FlatIndexCuda.from_host(np.zeros((4, 8), np.float32), None, F32) as idx:
        ids, sc = idx.search_topk_dot_batch(np.zeros((1, 8), np.float32), 0)
        assert ids.shape == (1, 0)
        with pytest.raises(nv.NvdbError):
            idx.search_topk_dot_batch(np.zeros((1, 9), np.float32), 1)


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
def test_device_synth_matches_oracle(nv, oracle, dt):
    dtype = DT[dt]
    n, d, seed, first = 1000, 768, 20240229, 123456789
    with nv.FlatIndexCuda.synthetic(dtype, n, d, seed, first_row=first) as idx:
        payload, scales = idx.read_rows(0, n)
    op, osc = oracle.synth(dtype, seed, first, n, d)
    assert np.array_equal(payload, op)
    if dtype == I8:
        assert np.array_equal(scales.view(np.uint32), osc.view(np.uint32))


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
@pytest.mark.parametrize("n,d,k", [(40000, 128, 10), (40000, 128, 100), (5000, 768, 1), (300, 64, 50), (70000, 64, 700)])
def test_fused_final_merge_equals_separate_merge_kernel(nv, oracle, dt, n, d, k):
    """the merge done by the scan's last CTA (default) and the separate merge kernel must agree bit for bit, for
    repeated launches on the same index (the completion counter is reset by the merging CTA)."""
    dtype = DT[dt]
    base, scales = make_base(oracle, dtype, n, d, seed=21)
    rng = np.random.default_rng(2)
    f = oracle.synth_f32(21, 0, n, d)
    with nv.FlatIndexCuda.from_host(base, scales, dtype) as idx:
        idx.set_path(nv.PATH_GEMV)
        for nq in (1, 1, 5, 13, 1):
            queries = np.ascontiguousarray(f[rng.integers(0, n, nq)])
            idx.set_tuning(nv.capi.TUNE_FUSED_MERGE, 2)      # 2 = fuse whenever the survivors' buffer fits the ring
            a = idx.search_topk_dot_batch(queries, k)
            idx.set_tuning(nv.capi.TUNE_FUSED_MERGE, 0)
            b = idx.search_topk_dot_batch(queries, k)
            assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), f"nq={nq}"


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
@pytest.mark.parametrize("n,d,k", [(6000, 64, 300), (40000, 64, 16), (60000, 32, 100), (38000, 64, 128), (5900, 128, 255)])
def test_radix_select_merge_of_many_lists(nv, oracle, dt, n, d, k):
    """> 2048 keys per query go through merge_select_kernel (radix select of the k-th key, sort of the survivors only):
    short lists with empty slots (k > rows per list), k = 100 over 148 lists, powers of two, several queries."""
    dtype = DT[dt]
    base, scales = make_base(oracle, dtype, n, d, seed=23)
    f = oracle.synth_f32(23, 0, n, d)
    rng = np.random.default_rng(8)
    for nq in (1, 5):
        queries = f[rng.integers(0, n, nq)] + (0.01 * rng.standard_normal((nq, d))).astype(np.float32)
        check(nv, oracle, dtype, base, scales, queries, k)


def test_radix_select_merge_all_equal_scores(nv, oracle):
    """Every row identical: 148 lists of equal scores, the k smallest row ids must win (keys differ only in ~row)."""
    base = np.tile(np.linspace(-1, 1, 64, dtype=np.float32), (50000, 1))
    q = np.ones((2, 64), np.float32)
    check(nv, oracle, F32, base, None, q, 100)
