"""GPU parity for the row-sharded path on ONE device: S logical shards (each its own index with a row_offset),
packed exactly as the all-gather delivers them, merged by the library's merge kernel == the unsharded top-k
("merge of per-shard top-k == top-k of the whole")."""
import numpy as np
import pytest
import torch

from _oracle import F16, F32, I8, assert_topk_equal

pytestmark = pytest.mark.gpu
DT = {"f32": F32, "f16": F16, "i8": I8}


@pytest.mark.parametrize("dt", ["f16", "i8", "f32"])
@pytest.mark.parametrize("S,n_total,k,nq", [(2, 5001, 10, 3), (8, 20000, 100, 2), (4, 3, 5, 1), (3, 7000, 10, 16)])
def test_logical_shards_merge_equals_unsharded(nv, oracle, dt, S, n_total, k, nq):
    dtype, d = DT[dt], 128
    dev = torch.device("cuda", 0)
    base, scales = oracle.synth(dtype, 13, 0, n_total, d)
    f = oracle.synth_f32(13, 0, n_total, d)
    rng = np.random.default_rng(0)
    queries = np.ascontiguousarray(f[rng.integers(0, n_total, nq)] + (0.01 * rng.standard_normal((nq, d))).astype(np.float32))
    q_dev = torch.from_numpy(queries).to(dev)
    w = nv.packed_words(nq, k)
    gathered = torch.empty((S, w), dtype=torch.int64, device=dev)
    shards = []
    for s in range(S):
        r0, nrows = nv.shard_rows(n_total, S, s)
        ids_view = gathered[s, :nq * k].view(nq, k)
        sc_view = gathered[s, nq * k:].view(torch.float32)[:nq * k].view(nq, k)
        if nrows == 0:
            ids_view.fill_(-1)
            sc_view.fill_(float("-inf"))
            continue
        idx = nv.FlatIndexCuda.from_host(base[r0:r0 + nrows], None if scales is None else scales[r0:r0 + nrows], dtype,
                                         row_offset=r0)
        idx.set_path(nv.PATH_GEMV)
        shards.append(idx)
        idx.search_device(q_dev.data_ptr(), nq, k, ids_view.data_ptr(), sc_view.data_ptr(),
                          torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out_ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
    out_sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
    nv.merge_topk_packed_device(0, gathered.data_ptr(), S, nq, k, out_ids.data_ptr(), out_sc.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ids = out_ids.cpu().numpy().view(np.uint64)
    sc = out_sc.cpu().numpy()
    oids, osc = oracle.search_batch(dtype, base, scales, queries, k)
    kk = min(k, n_total)
    for qi in range(nq):
        assert_topk_equal(ids[qi, :kk], sc[qi, :kk], oids[qi, :kk], osc[qi, :kk], what=f"S={S} q{qi}")
        assert np.all(ids[qi, kk:] == np.uint64(0xFFFFFFFFFFFFFFFF))
    # the world==1 path of ShardedFlatIndex (same code as N ranks minus the collective)
    with nv.FlatIndexCuda.from_host(base, scales, dtype) as whole:
        whole.set_path(nv.PATH_GEMV)
        sh = nv.ShardedFlatIndex(whole, dev)
        i2, s2 = sh.search(q_dev, k)
        torch.cuda.synchronize()
        assert np.array_equal(i2.cpu().numpy().view(np.uint64)[:, :kk], ids[:, :kk])
        assert np.array_equal(s2.cpu().numpy().view(np.uint32)[:, :kk], sc.view(np.uint32)[:, :kk])
    for idx in shards:
        idx.close()
