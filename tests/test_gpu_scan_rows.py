"""GPU parity for the row-tile tensor-core scan of small batches (csrc/scan_rows.cu: base rows on M, the whole batch
resident as the B operand) behind the exact path.

The bar is the exact path's: ids and FP32 score bits identical to the streaming scan == the oracle (which is pinned to
the unmodified reference). The cases aim at what is new in this kernel: per-query candidate lists shared by 128 row
threads (atomic slots, warp compaction, pending appends retried from TMEM), the in-kernel limb conversion of the queries
into the swizzled B operand (padding columns, ragged last K chunk), and the knob that switches it off.
"""
import numpy as np
import pytest

from _oracle import F16, I8, assert_topk_equal

pytestmark = pytest.mark.gpu
TUNE_ROWS, TUNE_ROWS_SEED = 7, 8
ROWS_ALWAYS = 2            # "whenever the batch fits" (auto = int8 only)


def _search(nv, dtype, base, scales, queries, k, rows, seed=1):
    with nv.FlatIndexCuda.from_host(base, scales, dtype) as idx:
        idx.set_tuning(TUNE_ROWS, rows)
        idx.set_tuning(TUNE_ROWS_SEED, seed)
        idx.set_path(nv.PATH_MMA_EXACT)
        l0 = nv.launch_count()
        ids, sc = idx.search_topk_dot_batch(queries, k)
        launches = nv.launch_count() - l0
        stats = idx.exact_stats()
        idx.set_path(nv.PATH_GEMV)
        gids, gsc = idx.search_topk_dot_batch(queries, k)
    return ids, sc, gids, gsc, stats, launches


@pytest.mark.parametrize("dt", ["f16", "i8"])
@pytest.mark.parametrize("n,d", [(20000, 128), (5000, 768), (700, 80), (129, 64)])
@pytest.mark.parametrize("nq,k", [(3, 10), (8, 1), (16, 10), (17, 12), (64, 10), (80, 10), (5, 100), (16, 40)])
def test_rows_kernel_bit_exact_vs_oracle_and_streaming_scan(nv, oracle, dt, n, d, nq, k):
    dtype = F16 if dt == "f16" else I8
    base, scales = oracle.synth(dtype, 21, 0, n, d)
    rng = np.random.default_rng(nq * 131 + k)
    f = oracle.synth_f32(21, 0, n, d)
    queries = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32))
    ids, sc, gids, gsc, stats, _ = _search(nv, dtype, base, scales, queries, k, rows=ROWS_ALWAYS, seed=2)
    assert np.array_equal(ids, gids) and np.array_equal(sc.view(np.uint32), gsc.view(np.uint32)), "rows kernel != streaming scan"
    oids, osc = oracle.search_batch(dtype, base, scales, queries, k)
    kk = min(k, n)
    for qi in range(nq):
        assert_topk_equal(ids[qi, :kk], sc[qi, :kk], oids[qi, :kk], osc[qi, :kk], what=f"q{qi}")
    assert stats[0] == nq
    ids0, sc0, _, _, _, _ = _search(nv, dtype, base, scales, queries, k, rows=0)   # the query-tile kernels (scan_mma*.cu)
    assert np.array_equal(ids, ids0) and np.array_equal(sc.view(np.uint32), sc0.view(np.uint32))
    ids1, sc1, _, _, _, _ = _search(nv, dtype, base, scales, queries, k, rows=ROWS_ALWAYS, seed=0)   # no threshold seeding
    assert np.array_equal(ids, ids1) and np.array_equal(sc.view(np.uint32), sc1.view(np.uint32))


@pytest.mark.parametrize("dt", ["f16", "i8"])
@pytest.mark.parametrize("nq", [4, 33])
def test_rows_kernel_every_row_beats_the_threshold(nv, oracle, dt, nq):
    """Scores that GROW with the row id: every row of every tile passes the running threshold of every query, so every
    list overflows in every tile — appends pend, lists compact, appends are retried from TMEM. The worst case for the
    selector must still return exactly the streaming scan's answer."""
    dtype = F16 if dt == "f16" else I8
    n, d, k = 6000, 64, 10
    rng = np.random.default_rng(2)
    u = rng.standard_normal(d).astype(np.float32)
    u /= np.linalg.norm(u)
    ramp = (0.2 + 0.8 * np.arange(n, dtype=np.float32) / n)[:, None]
    rows = ramp * u[None, :] + 1e-3 * rng.standard_normal((n, d)).astype(np.float32)
    if dtype == F16:
        base, scales = oracle.to_f16(rows), None
    else:
        base, scales = oracle.quantize_i8(rows)
    queries = np.ascontiguousarray(u[None, :] + 0.05 * rng.standard_normal((nq, d)).astype(np.float32))
    ids, sc, gids, gsc, _, _ = _search(nv, dtype, base, scales, queries, k, rows=ROWS_ALWAYS, seed=2)
    assert np.array_equal(ids, gids) and np.array_equal(sc.view(np.uint32), gsc.view(np.uint32))
    assert int(ids.min()) > n - 400      # the best rows are the last ones


@pytest.mark.parametrize("dt", ["f16", "i8"])
def test_rows_kernel_exact_ties_and_duplicates(nv, oracle, dt):
    """40 identical copies of some rows (exact score ties inside and across tiles): the certificate fails for the queries
    that hit them and the later tiers answer; either way the canonical (score desc, id asc) result comes back."""
    dtype = F16 if dt == "f16" else I8
    n, d, k = 3000, 128, 10
    base, scales = oracle.synth(dtype, 8, 0, n, d)
    base = base.copy()
    base[500:540] = base[500]
    base[2900:2940] = base[2900]
    if scales is not None:
        scales = scales.copy()
        scales[500:540] = scales[500]
        scales[2900:2940] = scales[2900]
    f = oracle.synth_f32(8, 0, n, d)
    queries = np.ascontiguousarray(f[[7, 500, 2900, 1234, 2999, 0, 500]])
    ids, sc, gids, gsc, stats, _ = _search(nv, dtype, base, scales, queries, k, rows=ROWS_ALWAYS, seed=2)
    assert np.array_equal(ids, gids) and np.array_equal(sc.view(np.uint32), gsc.view(np.uint32))
    oids, osc = oracle.search_batch(dtype, base, scales, queries, k)
    assert np.array_equal(ids, oids) and np.array_equal(sc.view(np.uint32), osc.view(np.uint32))
    assert stats[1] >= 3                 # the three queries on duplicated rows were flagged


def test_rows_kernel_is_what_auto_runs_for_small_batches(nv, oracle):
    """AUTO, int8, 8 queries: candidate scan (scan_rows_kernel converts the queries itself) + re-score (which selects the
    candidates from the per-CTA lists itself) = 2 launches; with the knob off the query-tile kernel needs its prep and
    merge launches as well."""
    n, d, k, nq = 30000, 128, 10, 8
    base, scales = oracle.synth(I8, 3, 0, n, d)
    queries = np.ascontiguousarray(oracle.synth_f32(3, 0, n, d)[:nq])
    with nv.FlatIndexCuda.from_host(base, scales, I8) as idx:
        idx.search_topk_dot_batch(queries, k)          # row norms (once per index) + workspaces
        l0 = nv.launch_count()
        idx.search_topk_dot_batch(queries, k)
        assert nv.launch_count() - l0 == 2
        idx.set_tuning(TUNE_ROWS, 0)
        idx.search_topk_dot_batch(queries, k)
        l0 = nv.launch_count()
        idx.search_topk_dot_batch(queries, k)
        assert nv.launch_count() - l0 == 4


@pytest.mark.parametrize("dt", ["f16", "i8"])
@pytest.mark.parametrize("rounds,depth", [(1, 4), (3, 3), (2, 6)])
def test_rows_kernel_seeding_rounds_and_ring_depth_do_not_change_results(nv, dt, rounds, depth):
    """A base with enough tiles per CTA for all three seeding rounds (after tiles 0, 8 and 64) and the ring-depth
    overrides: results are those of the streaming scan bit for bit whatever NVDB_B200_ROWS_ROUNDS /
    NVDB_B200_ROWS_MAX_STAGES / NVDB_B200_GEMV_MAX_STAGES say (they are A/B aids)."""
    import os
    torch = pytest.importorskip("torch")
    dtype = F16 if dt == "f16" else I8
    n, d, k, nq = 1_300_000, 64, 10, 48            # 10157 tiles of 128 rows: 68 per CTA on 148 SMs
    dev = torch.device("cuda", 0)
    st = torch.cuda.Stream(device=dev)
    env = {"NVDB_B200_ROWS_ROUNDS": str(rounds), "NVDB_B200_ROWS_MAX_STAGES": str(depth), "NVDB_B200_GEMV_MAX_STAGES": str(depth)}
    old = {key: os.environ.get(key) for key in env}
    os.environ.update(env)
    try:
        with nv.FlatIndexCuda.synthetic(dtype, n, d, 77) as idx:
            q = torch.empty((nq, d), dtype=torch.float32, device=dev)
            for i in range(nq):
                nv.synth_fill_device(0, nv.F32, 1, d, 77, (i * 7919) % n, q[i].data_ptr(), None, st.cuda_stream)
            st.synchronize()
            q += 0.01 * torch.randn(q.shape, device=dev, generator=torch.Generator(device=dev).manual_seed(3))
            res = []
            for path, rows in ((nv.PATH_MMA_EXACT, ROWS_ALWAYS), (nv.PATH_GEMV, 1)):
                idx.set_tuning(TUNE_ROWS, rows)
                idx.set_tuning(TUNE_ROWS_SEED, 2)
                idx.set_path(path)
                ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
                sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
                for _ in range(2):                  # the second launch reuses the seeding counters (epochs)
                    idx.search_device(q.data_ptr(), nq, k, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                st.synchronize()
                res.append((ids.cpu().numpy(), sc.cpu().numpy().view(np.uint32)))
            assert idx.exact_stats()[0] == 2 * nq
        assert np.array_equal(res[0][0], res[1][0]) and np.array_equal(res[0][1], res[1][1])
    finally:
        for key, v in old.items():
            if v is None:
                os.environ.pop(key, None)
            else:
                os.environ[key] = v
