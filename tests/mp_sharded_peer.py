"""Launched by torchrun with one rank per GPU (tests/test_gpu_sharded.py, scripts/gpu_round*.sh): the row-sharded index
over the NVLink peer-memory exchange against the NCCL all-gather path and the unsharded index. Prints
PEER_EXCHANGE_OK on rank 0 when every comparison was bit-exact."""
import importlib
import os
import sys
from pathlib import Path

import numpy as np
import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))


def main():
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
    dist.init_process_group("nccl", device_id=dev)
    nv = importlib.import_module("nano-vectordb_b200")
    F16, d, seed = nv.F16, 256, 77
    for n_total, k, nq in [(200_000, 10, 1), (200_001, 100, 33), (50_000, 10, 1024), (world - 1, 5, 2)]:
        r0, nrows = nv.shard_rows(n_total, world, rank)
        idx = None if nrows == 0 else nv.FlatIndexCuda.synthetic(F16, nrows, d, seed, first_row=r0, device=lr, row_offset=r0)
        qrows = [(i * 7919) % n_total for i in range(nq)]
        q = torch.empty((nq, d), dtype=torch.float32, device=dev)
        for i, r in enumerate(qrows):
            nv.synth_fill_device(lr, nv.F32, 1, d, seed, r, q[i].data_ptr(), None, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        peer = nv.ShardedFlatIndex(idx, dev, exchange="peer")
        nccl = nv.ShardedFlatIndex(idx, dev, exchange="nccl")
        res = []
        for rep in range(5):                          # several epochs through the double-buffered slots
            qq = q if rep % 2 == 0 else q.flip(0).contiguous()
            pi, ps = [t.clone() for t in peer.search(qq, k)]
            ni, ns = [t.clone() for t in nccl.search(qq, k)]
            torch.cuda.synchronize()
            assert peer.exchange == "peer" and peer.exchange_status() == 0
            assert torch.equal(pi, ni), f"rank {rank}: ids differ (n={n_total}, k={k}, nq={nq}, rep={rep})"
            assert torch.equal(ps.view(torch.int32), ns.view(torch.int32)), f"rank {rank}: scores differ"
            res.append((qq, pi, ps))
        if nq == 1:   # a later, larger call on the same object: the windows are re-created collectively
            q2 = q.repeat(1200, 1).contiguous()
            pi, ps = [t.clone() for t in peer.search(q2, k)]
            ni, ns = [t.clone() for t in nccl.search(q2, k)]
            torch.cuda.synchronize()
            assert peer.exchange_status() == 0 and torch.equal(pi, ni) and torch.equal(ps.view(torch.int32), ns.view(torch.int32))
        # every rank holds the same answer
        mine = res[0][1].clone()
        allr = [torch.empty_like(mine) for _ in range(world)]
        dist.all_gather(allr, mine)
        assert all(torch.equal(a, mine) for a in allr)
        # == the unsharded index (rank 0 builds the whole base)
        if rank == 0:
            kk = min(k, n_total)
            with nv.FlatIndexCuda.synthetic(F16, n_total, d, seed, first_row=0, device=lr, row_offset=0) as whole:
                wi = torch.empty((nq, k), dtype=torch.int64, device=dev)
                ws = torch.empty((nq, k), dtype=torch.float32, device=dev)
                whole.search_device(q.data_ptr(), nq, k, wi.data_ptr(), ws.data_ptr(), torch.cuda.current_stream().cuda_stream)
                torch.cuda.synchronize()
                a, b = res[0][2][:, :kk].cpu().numpy(), ws[:, :kk].cpu().numpy()
                assert np.array_equal(a.view(np.uint32), b.view(np.uint32)), "sharded scores != unsharded scores"
                ia, ib = res[0][1][:, :kk].cpu().numpy(), wi[:, :kk].cpu().numpy()
                for qi in range(nq):      # ids may differ only inside exact-tie groups (tie rule: id asc in both)
                    assert np.array_equal(ia[qi], ib[qi]) or len(np.unique(a[qi])) < kk, f"q{qi}: ids differ"
        peer.close()
        dist.barrier()
        if idx is not None:
            idx.close()
    if rank == 0:
        print("PEER_EXCHANGE_OK", flush=True)
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
