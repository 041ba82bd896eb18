"""Two-stage search: INT8 coarse scan (top-R) + exact FP16 re-rank on the device. Approximate by
construction, so the bar is (a) every returned score is the bit-exact reference-order FP16 score of the returned id,
(b) results are sorted canonically, (c) recall@k against the exact FP16 scan is ~1 on the synthetic base."""
import numpy as np
import pytest

from _oracle import F16, I8

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("nq", [1, 2, 37, 300])
@pytest.mark.parametrize("k,R", [(10, 32), (10, 10), (50, 96)])
def test_two_stage_int8_scan_fp16_rerank(nv, oracle, nq, k, R):
    n, d, seed = 60_000, 128, 31
    rng = np.random.default_rng(nq + k)
    f = oracle.synth_f32(seed, 0, n, d)
    queries = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.05 * rng.standard_normal((nq, d))).astype(np.float32))
    with nv.FlatIndexCuda.synthetic(I8, n, d, seed) as coarse, nv.FlatIndexCuda.synthetic(F16, n, d, seed) as fine:
        base16, _ = fine.read_rows(0, n)
        t = nv.Timing()
        ids, sc = nv.capi.search_two_stage(coarse, fine, queries, k, R, t)
        fine.set_path(nv.PATH_GEMV)
        eids, esc = fine.search_topk_dot_batch(queries, k)
    assert t.kernel_launches >= 2
    hits = 0
    for qi in range(nq):
        # (a) scores are the exact FP16 reference scores of the returned rows
        want = np.array([oracle.dot(F16, queries[qi], base16[int(i)]) for i in ids[qi]], np.float32)
        assert np.array_equal(sc[qi].view(np.uint32), want.view(np.uint32)), f"q{qi}"
        # (b) canonical order
        order = np.lexsort((ids[qi], -sc[qi].astype(np.float64)))
        assert np.array_equal(order, np.arange(k)), f"q{qi} not sorted"
        hits += len(set(ids[qi].tolist()) & set(eids[qi].tolist()))
    recall = hits / (nq * k)
    assert recall >= (0.99 if R >= k + 16 else 0.80), f"recall@{k} = {recall:.4f} with R = {R}"


def test_two_stage_argument_checks(nv, oracle):
    with nv.FlatIndexCuda.synthetic(I8, 1000, 64, 1) as a, nv.FlatIndexCuda.synthetic(F16, 999, 64, 1) as b:
        with pytest.raises(nv.NvdbError):
            nv.capi.search_two_stage(a, b, np.zeros((1, 64), np.float32), 5, 16)
        with pytest.raises(nv.NvdbError):
            nv.capi.search_two_stage(a, a, np.zeros((1, 64), np.float32), 5, 16)
