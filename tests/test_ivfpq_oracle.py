"""CPU checks of oracle/ivfpq_ref.py (the IVFADC restatement; parity unpinned — no FAISS, no reference fixtures):
the restatement must at least satisfy the algorithm's own invariants before the CUDA path is compared with it."""
import sys
from pathlib import Path

import numpy as np

sys.path.insert(0, str(Path(__file__).resolve().parent.parent / "oracle"))
import ivfpq_ref as ref  # noqa: E402


def clustered(n, d, nclusters, seed, spread=0.15):
    rng = np.random.default_rng(seed)
    centers = rng.standard_normal((nclusters, d)).astype(np.float32)
    x = centers[rng.integers(0, nclusters, n)] + spread * rng.standard_normal((n, d)).astype(np.float32)
    return x.astype(np.float32)


def test_init_rows_are_distinct_and_in_their_buckets():
    for n, k in ((1000, 64), (257, 256), (64, 64)):
        rows = ref.init_rows(n, k, 7)
        assert len(set(rows.tolist())) == k
        for j, r in enumerate(rows):
            assert j * n // k <= r < max((j + 1) * n // k, j * n // k + 1)


def test_kmeans_objective_does_not_increase():
    x = clustered(3000, 16, 20, 1)
    objs = [ref.objective(x, ref.kmeans(x, 32, it, 5)) for it in (0, 1, 3, 10)]
    assert all(b <= a * (1 + 1e-9) for a, b in zip(objs, objs[1:])), objs
    assert objs[-1] < 0.7 * objs[0]


def test_encode_group_search_invariants():
    n, d, nlist, M = 4000, 32, 16, 8
    x = clustered(n, d, 24, 2)
    coarse, cb = ref.train(x[:2000], nlist, M, 8, 11)
    lists, codes = ref.encode(x, coarse, cb)
    off, ids, gcodes = ref.group(lists, codes, nlist)
    assert off[0] == 0 and off[-1] == n and np.all(np.diff(off.astype(np.int64)) >= 0)
    for l in range(nlist):  # ascending row ids inside a list, all rows of the list and only those
        seg = ids[off[l]:off[l + 1]]
        assert np.all(np.diff(seg.astype(np.int64)) > 0)
        assert np.array_equal(np.sort(np.flatnonzero(lists == l)), seg)
    # PQ reconstruction is closer to the residual than the zero vector (i.e. the codebooks carry information)
    res = x - coarse[lists]
    dsub = d // M
    rec = np.concatenate([cb[m][codes[:, m]] for m in range(M)], axis=1)
    assert ((res - rec) ** 2).sum() < 0.5 * (res ** 2).sum()
    assert rec.shape == (n, M * dsub)

    q = x[:20] + 0.01
    # probing every list with R = n visits every row exactly once, sorted by (distance, id)
    i_all, d_all = ref.search(q[:3], coarse, cb, off, ids, gcodes, nlist, n)
    for qi in range(3):
        assert np.array_equal(np.sort(i_all[qi]), np.arange(n))
        assert np.all(np.diff(d_all[qi]) >= 0)
    # ADC distance of a stored vector == |q - (c + reconstruction)|^2
    full = coarse[lists] + rec
    want = ((q[0][None, :].astype(np.float64) - full.astype(np.float64)) ** 2).sum(1)
    got = np.empty(n)
    got[i_all[0]] = d_all[0]
    assert np.allclose(got, want, rtol=1e-5, atol=1e-6)
    # recall against the exact neighbours grows with nprobe and R
    truth = ref.exact_l2_topk(x, q, 10)
    rec_np = [ref.recall_at_k(ref.search(q, coarse, cb, off, ids, gcodes, p, 400)[0][:, :400], truth) for p in (1, 4, 16)]
    assert rec_np[0] <= rec_np[1] + 1e-9 <= rec_np[2] + 2e-9
    assert rec_np[-1] > 0.9


def correlated(n, d, seed):
    """rows whose energy is concentrated in a few correlated directions: plain PQ wastes codes, OPQ re-balances them"""
    rng = np.random.default_rng(seed)
    mix = np.linalg.qr(rng.standard_normal((d, d)))[0]
    scales = np.geomspace(4.0, 0.05, d)
    return ((rng.standard_normal((n, d)) * scales) @ mix).astype(np.float32)


def test_opq_rotation_is_orthogonal_and_reduces_quantisation_error():
    x = correlated(4000, 16, 3)
    R, errs = ref.opq_train(x, 4, 8, 5)
    assert np.allclose(R.T @ R, np.eye(16), atol=1e-9)
    assert errs[-1] < 0.8 * errs[0], errs          # PQ error of X R falls over the rounds
    assert ref.pq_error(x @ R, 4, 6, 9) < 0.85 * ref.pq_error(x, 4, 6, 9)
