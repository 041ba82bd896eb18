"""IVF-PQ candidate generation on the device against oracle/ivfpq_ref.py.

PARITY UNPINNED (no FAISS in the reference tree or the image, no fixtures for this stage): the oracle restates the
published IVFADC algorithm in float64. The kernels accumulate in float32 in a fixed order, so the bar is tolerance:
  * encode: list ids / code bytes equal to the oracle's on the exported parameters except at near-ties
    (relative distance gap < 1e-5), and there must be few of those;
  * search: every returned (id, distance) pair matches the oracle's float64 ADC distance of that id within
    2e-6 * (|q|^2 + |y|^2) (the kernel evaluates |q|^2 - 2 q.y + |y|^2 in float32, so its error scales with the norms,
    not with the distance), the list is ordered, and no unreturned vector of the probed lists beats the R-th returned
    distance by more than that tolerance;
  * training: deterministic for a seed; objective within 5 % of the oracle's Lloyd run from the same initial rows.
"""
import sys
from pathlib import Path

import numpy as np
import pytest

from _oracle import F16, F32, I8

sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "oracle"))
import ivfpq_ref as ref  # noqa: E402

pytestmark = pytest.mark.gpu
RTOL = 2e-5


def clustered(n, d, nclusters, seed, spread=0.35):
    rng = np.random.default_rng(seed)
    centers = rng.standard_normal((nclusters, d)).astype(np.float32)
    x = centers[rng.integers(0, nclusters, n)] + spread * rng.standard_normal((n, d)).astype(np.float32)
    return np.ascontiguousarray(x, np.float32)


def build(nv, x, dtype=F32, **kw):
    if dtype == F32:
        base = nv.FlatIndexCuda.from_host(x, None, F32)
        rows = x
    elif dtype == F16:
        h = x.astype(np.float16)
        base = nv.FlatIndexCuda.from_host(h.view(np.uint16), None, F16)
        rows = h.astype(np.float32)
    else:
        q8, sc = nv.capi.convert(x, F32, I8)
        base = nv.FlatIndexCuda.from_host(q8, sc, I8)
        rows = (q8.astype(np.float32) * sc[:, None]).astype(np.float32)
    return base, rows, nv.IvfPqIndex(base, **kw)


@pytest.mark.parametrize("dtype", [F32, F16, I8])
def test_build_layout_and_encoding_match_oracle(nv, dtype):
    n, d, nlist, M = 20_000, 48, 64, 8   # dsub = 6, the reference's m=64 / D=384 sub-vector length
    x = clustered(n, d, 40, 3)
    base, rows, ivf = build(nv, x, dtype, nlist=nlist, pq_m=M, ntrain=8000, niter=10, seed=5)
    with base, ivf:
        assert (ivf.n, ivf.dim, ivf.nlist, ivf.pq_m) == (n, d, nlist, M)
        coarse, cb, off, ids, codes = ivf.export()
    assert off[0] == 0 and off[-1] == n and np.all(np.diff(off.astype(np.int64)) >= 0)
    assert np.array_equal(np.sort(ids), np.arange(n, dtype=np.uint32))
    lists_gpu = np.empty(n, np.uint32)
    for l in range(nlist):
        seg = ids[off[l]:off[l + 1]]
        assert np.all(np.diff(seg.astype(np.int64)) > 0), "row ids ascend inside a list"
        lists_gpu[seg] = l
    codes_gpu = np.empty_like(codes)
    codes_gpu[ids] = codes
    lists_ref, codes_ref = ref.encode(rows, coarse, cb)
    # list ids: mismatches only at near-ties between two centroids
    bad = np.flatnonzero(lists_gpu != lists_ref)
    assert len(bad) <= n // 500
    d2 = ref.sq_dists(rows[bad], coarse) if len(bad) else np.zeros((0, nlist))
    for j, r in enumerate(bad):
        a, b = d2[j, lists_gpu[r]], d2[j, lists_ref[r]]
        assert abs(a - b) <= 1e-5 * max(abs(a), abs(b), 1e-12), (r, a, b)
    # codes of rows whose list agrees: mismatches only at near-ties between two codewords
    same = lists_gpu == lists_ref
    diff = (codes_gpu != codes_ref) & same[:, None]
    assert diff.sum() <= n * M // 500
    res = rows - coarse[lists_ref]
    dsub = d // M
    for r, m in zip(*np.nonzero(diff)):
        dd = ref.sq_dists(res[r:r + 1, m * dsub:(m + 1) * dsub], cb[m])[0]
        a, b = dd[codes_gpu[r, m]], dd[codes_ref[r, m]]
        assert abs(a - b) <= 1e-5 * max(abs(a), abs(b), 1e-12), (r, m, a, b)


def test_training_is_deterministic_and_as_good_as_lloyd(nv):
    n, d, nlist, M = 12_000, 32, 32, 4
    x = clustered(n, d, 50, 8)
    outs = []
    for _ in range(2):
        base, rows, ivf = build(nv, x, F32, nlist=nlist, pq_m=M, ntrain=6000, niter=12, seed=77)
        with base, ivf:
            outs.append(ivf.export())
    for a, b in zip(outs[0], outs[1]):
        assert np.array_equal(a, b), "same seed -> same index, bit for bit"
    coarse = outs[0][0]
    want = ref.kmeans(x[:6000], nlist, 12, 77)
    o_gpu, o_ref = ref.objective(x[:6000], coarse), ref.objective(x[:6000], want)
    o_init = ref.objective(x[:6000], x[ref.init_rows(6000, nlist, 77)])
    assert o_gpu <= 1.05 * o_ref and o_gpu < 0.9 * o_init, (o_gpu, o_ref, o_init)
    # first Lloyd step from the shared initial rows: means agree to float32 accuracy
    base, rows, ivf1 = build(nv, x, F32, nlist=nlist, pq_m=M, ntrain=6000, niter=1, seed=77)
    with base, ivf1:
        c1 = ivf1.export()[0]
    w1 = ref.kmeans(x[:6000], nlist, 1, 77)
    assert np.allclose(c1, w1, rtol=1e-4, atol=1e-5)


# (8, 8, 100) / (16, 8, 128): a single query's 8 probe groups x R keys = 800 / 1024 keys in the rank-counting merge, with
# groups that hold fewer than R candidates (zero padding in the middle of the key array)
@pytest.mark.parametrize("M,nprobe,R", [(8, 1, 10), (8, 8, 50), (8, 8, 100), (16, 8, 128), (16, 64, 256), (12, 5, 1000)])
def test_search_matches_oracle_adc(nv, M, nprobe, R):
    n, d, nlist = 30_000, 48, 64
    x = clustered(n, d, 60, 4)
    rng = np.random.default_rng(9)
    queries = np.ascontiguousarray(x[rng.integers(0, n, 24)] + 0.05 * rng.standard_normal((24, d)).astype(np.float32))
    base, rows, ivf = build(nv, x, F16, nlist=nlist, pq_m=M, ntrain=10_000, niter=8, seed=2)
    with base, ivf:
        coarse, cb, off, ids, codes = ivf.export()
        gi, gd = ivf.search(queries, nprobe, R)
        gi1, gd1 = ivf.search(queries[:1], nprobe, R)          # batch of one: same answer
        gi2, _ = ivf.search(queries, nprobe, R, return_dist=False)
    assert np.array_equal(gi[:1], gi1) and np.array_equal(gd[:1], gd1) and np.array_equal(gi, gi2)
    ynorm_max = float(((np.abs(coarse).max(0) + np.abs(cb).max(1).reshape(-1)) ** 2).sum())  # >= |c + reconstruction|^2
    for qi, q in enumerate(queries):
        vid, vd, _ = ref.adc_distances(q, coarse, cb, off, ids, codes, nprobe)
        truth = dict(zip(vid.tolist(), vd.tolist()))
        valid = gi[qi] != 0xFFFFFFFF
        nvalid = int(valid.sum())
        assert nvalid == min(R, len(vid)), (qi, nvalid, len(vid))
        assert np.all(valid[:nvalid]) and np.all(gd[qi][nvalid:] == np.float32(1e30))
        got_ids, got_d = gi[qi][:nvalid], gd[qi][:nvalid].astype(np.float64)
        assert len(set(got_ids.tolist())) == nvalid
        want_d = np.array([truth[int(i)] for i in got_ids])        # KeyError = a row outside the probed lists
        atol = 2e-6 * (float((q.astype(np.float64) ** 2).sum()) + ynorm_max)
        assert np.max(np.abs(got_d - want_d)) <= atol, (qi, np.max(np.abs(got_d - want_d)), atol)
        assert np.all(np.diff(gd[qi][:nvalid]) >= 0), "ascending distances"
        same = np.flatnonzero(np.diff(gd[qi][:nvalid]) == 0)
        assert np.all(got_ids[same] < got_ids[same + 1]), "id ascending inside equal distances"
        if nvalid < len(vid):   # nothing better was left behind
            rest = np.setdiff1d(vid, got_ids)
            best_rest = min(truth[int(i)] for i in rest)
            assert best_rest >= got_d[-1] - 2 * atol, qi


def test_recall_grows_with_nprobe_and_refine_recovers_exact_neighbours(nv):
    """Shape of the reference's nprobe x REFINE_K surface (Performance.md:673-687): candidate recall rises with nprobe,
    exact refine of R >= 50 candidates lifts Recall@10 far above the raw ADC ranking, and probing every list with a
    large R returns the exact neighbours."""
    n, d, nlist, M, k = 40_000, 64, 128, 16, 10
    x = clustered(n, d, 300, 6, spread=0.6)
    rng = np.random.default_rng(1)
    queries = np.ascontiguousarray(x[rng.integers(0, n, 200)] + 0.2 * rng.standard_normal((200, d)).astype(np.float32))
    base, rows, ivf = build(nv, x, F16, nlist=nlist, pq_m=M, ntrain=20_000, niter=10, seed=3)
    truth = ref.exact_l2_topk(rows, queries, k)
    with base, ivf:
        raw, refined = [], []
        for nprobe in (1, 4, 16, 64):
            ci, _ = ivf.search(queries, nprobe, k)
            raw.append(ref.recall_at_k(ci, truth))
            ri, rd = ivf.search_refine(queries, nprobe, 50, k)
            refined.append(ref.recall_at_k(ri, truth))
            assert np.all(np.diff(rd, axis=1) >= 0)
        full_i, full_d = ivf.search_refine(queries, nlist, 1000, k)
        # refine distances are the exact squared L2 of the returned rows (the refine kernel's own parity tests pin bits)
        want = ((queries[:, None, :].astype(np.float64) - rows[full_i].astype(np.float64)) ** 2).sum(2)
        assert np.allclose(full_d, want, rtol=1e-4, atol=1e-5)
    assert all(b >= a - 0.01 for a, b in zip(raw, raw[1:])), raw
    assert all(b >= a - 0.01 for a, b in zip(refined, refined[1:])), refined
    assert refined[-1] > raw[-1] + 0.05 and refined[-1] > 0.9, (raw, refined)
    assert ref.recall_at_k(full_i, truth) > 0.995


def test_small_host_searches_replayed_from_a_graph_equal_the_batch(nv):
    """Host-API searches of <= 8 queries are captured the second time a shape is seen and replayed afterwards
    (ivfpq.cu: run_or_replay). Plain, capturing and replayed calls must all return the batch's answer, also when shapes
    alternate (several cached graphs), when a workspace regrows in between (graphs dropped) and with the knob off."""
    import os
    n, d, nlist, M, k = 30_000, 64, 64, 16, 10
    x = clustered(n, d, 80, 12)
    rng = np.random.default_rng(3)
    queries = np.ascontiguousarray(x[rng.integers(0, n, 40)] + 0.1 * rng.standard_normal((40, d)).astype(np.float32))
    base, rows, ivf = build(nv, x, F16, nlist=nlist, pq_m=M, ntrain=10_000, niter=6, seed=5)
    with base, ivf:
        bi, bd = ivf.search(queries, 8, 50)               # 40 queries: never a graph
        ri, rd = ivf.search_refine(queries, 8, 50, k)
        l0 = nv.launch_count()
        for i in range(12):                               # plain, capture, then replays
            gi, gd = ivf.search(queries[i:i + 1], 8, 50)
            assert np.array_equal(gi[0], bi[i]) and np.array_equal(gd[0], bd[i]), i
        per_call = (nv.launch_count() - l0) / 12
        assert per_call == int(per_call) and per_call >= 4, "replays count the kernels they launch"
        for i in range(12):                               # two shapes alternating: capture happens on a repeat only
            gi, gd = ivf.search_refine(queries[i:i + 1], 8, 50, k)
            assert np.array_equal(gi[0], ri[i]) and np.array_equal(gd[0], rd[i]), i
            gi, gd = ivf.search_refine(queries[i:i + 1], 8, 50, k)
            assert np.array_equal(gi[0], ri[i]) and np.array_equal(gd[0], rd[i]), i
            gi, _ = ivf.search(queries[i:i + 4], 8, 50, return_dist=False)
            assert np.array_equal(gi, bi[i:i + 4]), i
            gi, _ = ivf.search(queries[i:i + 4], 8, 50, return_dist=False)
            assert np.array_equal(gi, bi[i:i + 4]), i
        big_i, _ = ivf.search(np.tile(queries, (8, 1)), 8, 200)   # regrows the workspaces: cached graphs are stale now
        assert np.array_equal(big_i[:40, :50], bi)
        for i in range(4):
            gi, gd = ivf.search_refine(queries[i:i + 1], 8, 50, k)
            assert np.array_equal(gi[0], ri[i]) and np.array_equal(gd[0], rd[i]), i
    os.environ["NVDB_B200_IVF_GRAPH"] = "0"
    try:
        base, rows, ivf = build(nv, x, F16, nlist=nlist, pq_m=M, ntrain=10_000, niter=6, seed=5)
        with base, ivf:
            for i in range(4):
                gi, gd = ivf.search_refine(queries[i:i + 1], 8, 50, k)
                assert np.array_equal(gi[0], ri[i]) and np.array_equal(gd[0], rd[i]), i
    finally:
        del os.environ["NVDB_B200_IVF_GRAPH"]


def test_argument_errors(nv):
    x = clustered(3000, 32, 10, 1)
    with nv.FlatIndexCuda.from_host(x, None, F32) as base:
        with pytest.raises(nv.NvdbError):
            nv.IvfPqIndex(base, nlist=16, pq_m=5)          # PQ_M must divide dim (nvdb_ivfpq_build.cpp:54)
        with pytest.raises(nv.NvdbError):
            nv.IvfPqIndex(base, nlist=16, pq_m=4, pq_bits=4)
        with pytest.raises(nv.NvdbError):
            nv.IvfPqIndex(base, nlist=5000, pq_m=4)        # more lists than training rows
        with nv.IvfPqIndex(base, nlist=16, pq_m=4, ntrain=0, niter=4) as ivf:
            with pytest.raises(nv.NvdbError):
                ivf.search(x[:2], 0, 10)
            with pytest.raises(nv.NvdbError):
                ivf.search_refine(x[:2], 4, 10, 20)          # k > R
            ids, dist = ivf.search(x[:2], 10_000, 5)          # nprobe clamps to nlist
            assert ids.shape == (2, 5) and np.all(ids != 0xFFFFFFFF)


def test_import_of_an_oracle_trained_index_and_export_round_trip(nv):
    """An index trained, encoded and grouped entirely by the numpy oracle, imported with nvdb_b200_ivfpq_import: the
    device search must rank exactly what the oracle ranks (this separates the search kernels from device training);
    then export -> import reproduces the same answers bit for bit (the save / load path)."""
    n, d, nlist, M, nprobe, R = 6000, 32, 16, 8, 6, 40
    x = clustered(n, d, 30, 12)
    coarse, cb = ref.train(x[:3000], nlist, M, 6, 21)
    lists, codes = ref.encode(x, coarse, cb)
    off, ids, gcodes = ref.group(lists, codes, nlist)
    rng = np.random.default_rng(2)
    queries = np.ascontiguousarray(x[rng.integers(0, n, 16)] + 0.05 * rng.standard_normal((16, d)).astype(np.float32))
    want_i, want_d = ref.search(queries, coarse, cb, off, ids, gcodes, nprobe, R)
    with nv.FlatIndexCuda.from_host(x, None, F32) as base:
        with nv.IvfPqIndex.from_arrays(base, coarse, cb, off, ids, gcodes) as ivf:
            assert (ivf.n, ivf.dim, ivf.nlist, ivf.pq_m) == (n, d, nlist, M)
            gi, gd = ivf.search(queries, nprobe, R)
            exported = ivf.export()
            ri, rd = ivf.search_refine(queries, nprobe, R, 5)
        for a, b in zip(exported, (coarse, cb, off, ids, gcodes)):
            assert np.array_equal(a, b)
        with nv.IvfPqIndex.from_arrays(base, *exported) as ivf2:
            gi2, gd2 = ivf2.search(queries, nprobe, R)
        assert np.array_equal(gi, gi2) and np.array_equal(gd, gd2)
        with pytest.raises(nv.NvdbError):
            bad = off.copy()
            bad[-1] -= 1
            nv.IvfPqIndex.from_arrays(base, coarse, cb, bad, ids, gcodes)
    ynorm_max = float(((np.abs(coarse).max(0) + np.abs(cb).max(1).reshape(-1)) ** 2).sum())
    for qi in range(len(queries)):
        atol = 2e-6 * (float((queries[qi].astype(np.float64) ** 2).sum()) + ynorm_max)
        assert np.max(np.abs(gd[qi].astype(np.float64) - want_d[qi])) <= atol
        # same ranking except where two oracle distances are closer than the tolerance
        for j in np.flatnonzero(gi[qi] != want_i[qi]):
            assert abs(want_d[qi][j] - gd[qi][j]) <= atol and gi[qi][j] in want_i[qi][max(0, j - 3):j + 4] or \
                abs(want_d[qi][min(j + 1, R - 1)] - want_d[qi][max(j - 1, 0)]) <= 4 * atol, (qi, j)
        # refine output = exact squared L2 of candidate rows, ascending, drawn from the candidates
        assert set(ri[qi].tolist()) <= set(gi[qi].tolist())
        exact = ((queries[qi][None, :].astype(np.float64) - x[ri[qi]].astype(np.float64)) ** 2).sum(1)
        assert np.allclose(rd[qi], exact, rtol=1e-5, atol=1e-6) and np.all(np.diff(rd[qi]) >= 0)


def correlated(n, d, seed):
    rng = np.random.default_rng(seed)
    mix = np.linalg.qr(rng.standard_normal((d, d)))[0]
    scales = np.geomspace(4.0, 0.05, d)
    return np.ascontiguousarray(((rng.standard_normal((n, d)) * scales) @ mix).astype(np.float32))


def test_opq_rotation_orthogonal_lower_error_and_better_recall(nv):
    """USE_OPQ=1 of the reference (faiss::OPQMatrix in front of IVF-PQ): the device rotation (tile GEMMs + Newton-Schulz
    polar factor, no SVD) must be orthogonal, must lower the PQ error of the rotated rows as the oracle's SVD-based
    OPQ does, and the OPQ index must rank better than the plain one (the reference's Table 19: 0.694 -> 0.766)."""
    n, d, nlist, M, k = 30_000, 64, 32, 8, 10
    x = correlated(n, d, 7)
    rng = np.random.default_rng(3)
    queries = np.ascontiguousarray(x[rng.integers(0, n, 150)] + 0.05 * rng.standard_normal((150, d)).astype(np.float32))
    truth = ref.exact_l2_topk(x, queries, k)
    with nv.FlatIndexCuda.from_host(x, None, F32) as base:
        with nv.IvfPqIndex(base, nlist=nlist, pq_m=M, ntrain=12_000, niter=8, seed=4) as plain:
            assert plain.rotation() is None
            p_ids, _ = plain.search(queries, nlist, k)
        with nv.IvfPqIndex(base, nlist=nlist, pq_m=M, ntrain=12_000, niter=8, seed=4, opq_niter=10) as opq:
            R = opq.rotation()
            o_ids, o_d = opq.search(queries, nlist, k)
            r_ids, r_d = opq.search_refine(queries, nlist, 200, k)
            coarse, cb, off, ids, codes = opq.export()
        with nv.IvfPqIndex(base, nlist=nlist, pq_m=M, ntrain=12_000, niter=8, seed=4, opq_niter=10) as opq2:
            assert np.array_equal(opq2.rotation(), R), "deterministic"
        # the exported arrays + rotation reproduce the index (save / load path with OPQ)
        with nv.IvfPqIndex.from_arrays(base, coarse, cb, off, ids, codes, rotation=R) as again:
            a_ids, a_d = again.search(queries, nlist, k)
    assert np.array_equal(a_ids, o_ids) and np.array_equal(a_d, o_d)
    assert R.shape == (d, d) and np.max(np.abs(R.astype(np.float64).T @ R.astype(np.float64) - np.eye(d))) < 1e-4
    e_plain, e_rot = ref.pq_error(x[:6000], M, 6, 9), ref.pq_error(x[:6000] @ R.astype(np.float64), M, 6, 9)
    R_ref, _ = ref.opq_train(x[:12_000], M, 10, 4)
    e_ref = ref.pq_error(x[:6000] @ R_ref, M, 6, 9)
    assert e_rot < 0.85 * e_plain and e_rot < 1.15 * e_ref, (e_plain, e_rot, e_ref)
    # ADC distances of the OPQ index against the oracle on the ROTATED queries
    qrot = (queries.astype(np.float64) @ R.astype(np.float64)).astype(np.float32)
    for qi in range(0, len(queries), 15):
        vid, vd, _ = ref.adc_distances(qrot[qi], coarse, cb, off, ids, codes, nlist)
        truth_d = dict(zip(vid.tolist(), vd.tolist()))
        want = np.array([truth_d[int(i)] for i in o_ids[qi]])
        assert np.allclose(o_d[qi], want, rtol=1e-4, atol=1e-3), qi
    rec_plain, rec_opq = ref.recall_at_k(p_ids, truth), ref.recall_at_k(o_ids, truth)
    assert rec_opq > rec_plain + 0.03, (rec_plain, rec_opq)
    # refine works on the ORIGINAL rows and queries: exact distances, near-perfect recall with R = 200
    want = ((queries[:, None, :].astype(np.float64) - x[r_ids].astype(np.float64)) ** 2).sum(2)
    assert np.allclose(r_d, want, rtol=1e-4, atol=1e-4) and ref.recall_at_k(r_ids, truth) > 0.97
