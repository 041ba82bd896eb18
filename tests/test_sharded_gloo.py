"""CPU test of the N>1 path (world_size 2, gloo): shard plan, packed all-gather layout, merge == unsharded top-k.

The CUDA kernels cannot run here, so the two injectable steps of ShardedFlatIndex are supplied by the checker
(the oracle scans the shard; a numpy merge under (score desc, id asc)); everything else — shard_rows, the packed
buffer layout, the all-gather, padding of short / empty shards — is the product code that runs on the GPUs."""
import os
import socket
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = Path(__file__).resolve().parent.parent


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def numpy_merge(gathered, nlists, nq, k, ids_out, scores_out):
    g = gathered.numpy()
    nk = nq * k
    for q in range(nq):
        ids = np.concatenate([g[l, :nk].view(np.uint64).reshape(nq, k)[q] for l in range(nlists)])
        sc = np.concatenate([g[l, nk:].view(np.float32)[:nk].reshape(nq, k)[q] for l in range(nlists)])
        keep = ids != np.uint64(0xFFFFFFFFFFFFFFFF)
        ids, sc = ids[keep], sc[keep]
        order = np.lexsort((ids, -sc.astype(np.float64)))[:k]
        oi = np.full(k, 0xFFFFFFFFFFFFFFFF, np.uint64)
        os_ = np.full(k, -np.inf, np.float32)
        oi[:len(order)], os_[:len(order)] = ids[order], sc[order]
        ids_out[q] = torch.from_numpy(oi.view(np.int64))
        scores_out[q] = torch.from_numpy(os_)


def _worker(rank, world, port, n_total, d, k, dtype, out_dir):
    sys.path.insert(0, str(ROOT))
    sys.path.insert(0, str(ROOT / "tests"))
    import importlib
    import _oracle as O
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    nv = importlib.import_module("nano-vectordb_b200")
    orc = O.Oracle()
    r0, nrows = nv.shard_rows(n_total, world, rank)
    base, scales = orc.synth(dtype, 31, r0, max(nrows, 1), d)

    def oracle_local_search(queries, nq, kk, ids_out, scores_out):
        ids, sc = orc.search_batch(dtype, base[:nrows], None if scales is None else scales[:nrows], queries.numpy(), kk)
        ids = np.where(ids == np.uint64(0xFFFFFFFFFFFFFFFF), ids, ids + np.uint64(r0))   # row_offset
        ids_out.copy_(torch.from_numpy(ids.view(np.int64)))
        scores_out.copy_(torch.from_numpy(sc))

    sh = nv.ShardedFlatIndex(object() if nrows else None, "cpu", local_search=oracle_local_search, merge=numpy_merge)
    f = orc.synth_f32(31, 0, n_total, d)
    queries = torch.from_numpy(np.ascontiguousarray(f[[1 % n_total, n_total // 2, n_total - 1]]))
    ids, sc = sh.search(queries, k)
    np.savez(Path(out_dir) / f"r{rank}.npz", ids=ids.numpy().view(np.uint64), sc=sc.numpy())
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.parametrize("n_total,k", [(1001, 10), (5, 4), (1, 3)])
def test_world2_merge_equals_unsharded(oracle, tmp_path, n_total, k):
    from _oracle import F16
    d, world = 64, 2
    mp.spawn(_worker, args=(world, _free_port(), n_total, d, k, F16, str(tmp_path)), nprocs=world, join=True)
    base, _ = oracle.synth(F16, 31, 0, n_total, d)
    f = oracle.synth_f32(31, 0, n_total, d)
    oids, osc = oracle.search_batch(F16, base, None, f[[1 % n_total, n_total // 2, n_total - 1]], k)
    r = [np.load(tmp_path / f"r{i}.npz") for i in range(world)]
    for i in range(world):      # every rank holds the same, exact, result
        assert np.array_equal(r[i]["ids"], oids)
        assert np.array_equal(r[i]["sc"].view(np.uint32), osc.view(np.uint32))


def test_shard_rows_partition():
    import importlib
    nv = importlib.import_module("nano-vectordb_b200")
    for n in (0, 1, 7, 8, 9, 100_000_000, 2_903_054):
        for g in (1, 2, 3, 4, 8):
            parts = [nv.shard_rows(n, g, r) for r in range(g)]
            assert parts[0][0] == 0 and sum(p[1] for p in parts) == n
            for a, b in zip(parts, parts[1:]):
                assert a[0] + a[1] == b[0]
            assert max(p[1] for p in parts) == -(-n // g)
    assert nv.shard_rows(100_000_000, 8, 7) == (87_500_000, 12_500_000)
    assert nv.packed_words(1, 10) == 15 and nv.packed_words(3, 3) == 14
