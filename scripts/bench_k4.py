#!/usr/bin/env python
"""K4 (BASELINE.json configs[3]): FP16 flat-scan top-100 over a row-sharded synthetic base, local top-k per GPU +
NCCL all-gather + merge. One process per GPU:

  python -m torch.distributed.run --nnodes=1 --nproc-per-node G --master-addr 127.0.0.1 --master-port P \
      scripts/bench_k4.py            # NVDB_K4_ROWS_PER_GPU (default 12,500,000 = 100 M / 8), NVDB_K4_BATCH (1024)

Every rank holds rows [rank*R, (rank+1)*R) of the global synthetic base (generated on the device, 1536 B per row:
12.5 M rows = 19.2 GB). A step = ShardedFlatIndex.search(batch queries, k=100): local exact tensor-core scan
(PATH AUTO), ONE packed all-gather, merge on every rank. Timed with CUDA events, max over ranks. Rank 0 also checks
that the top-1 of each base-row query is the row itself and prints one JSON line.
"""
import importlib
import json
import os
import sys
from pathlib import Path

import torch
import torch.distributed as dist

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")

DIM, K, SEED = 768, int(os.environ.get("NVDB_K4_K", 100)), 20240229
ROWS = int(os.environ.get("NVDB_K4_ROWS_PER_GPU", 12_500_000))
BATCH = int(os.environ.get("NVDB_K4_BATCH", 1024))
STEPS = int(os.environ.get("NVDB_K4_STEPS", 5))
rank = int(os.environ.get("RANK", "0"))
world = int(os.environ.get("WORLD_SIZE", "1"))
local_rank = int(os.environ.get("LOCAL_RANK", "0"))
torch.cuda.set_device(local_rank)
dev = torch.device("cuda", local_rank)
if world > 1:
    dist.init_process_group("nccl", device_id=dev)

idx = nv.FlatIndexCuda.synthetic(nv.F16, ROWS, DIM, SEED, first_row=rank * ROWS, device=local_rank, row_offset=rank * ROWS)
total = world * ROWS
stream = torch.cuda.Stream(device=dev)
qrows = [(i * 104729 * 31) % total for i in range(BATCH)]
q = torch.empty((BATCH, DIM), dtype=torch.float32, device=dev)
with torch.cuda.stream(stream):
    for i, r in enumerate(qrows):
        nv.synth_fill_device(local_rank, nv.F32, 1, DIM, SEED, r, q[i].data_ptr(), None, stream.cuda_stream)
stream.synchronize()
sharded = nv.ShardedFlatIndex(idx, dev, exchange=os.environ.get("NVDB_B200_EXCHANGE", "auto"))
out = {}
for batch in sorted({1, BATCH}):
    qs = q[:batch].contiguous()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for _ in range(2):
            ids, sc = sharded.search(qs, K)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize(dev)
        iters = STEPS if batch > 1 else 10 * STEPS
        e0.record(stream)
        for _ in range(iters):
            ids, sc = sharded.search(qs, K)
        e1.record(stream)
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
    ms = e0.elapsed_time(e1) / iters
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    top1 = ids[:, 0].cpu().tolist()
    ok = sum(int(a == b) for a, b in zip(top1, qrows[:batch]))
    out[f"batch{batch}"] = {"ms_per_batch": ms, "qps": batch / ms * 1e3,
                            "base_GBps_aggregate": total * DIM * 2 / (ms * 1e-3) / 1e9,
                            "self_top1_hits": ok, "queries": batch}
xq, xf = idx.exact_stats()
xstatus, xkind = sharded.exchange_status(), sharded.exchange
sharded.close()
if rank == 0:
    print(json.dumps({"config": f"K4: FP16 top-{K}, {total} x {DIM} rows row-sharded over {world} GPU(s) "
                                f"({ROWS} rows = {ROWS * DIM * 2 / 1e9:.1f} GB per GPU), local top-k + " +
                                ("peer-memory exchange + merge in one kernel (NVLink stores, no NCCL)" if xkind == "peer"
                                 else "NCCL all-gather + merge kernel"),
                      "exchange": xkind, "exchange_status": xstatus,
                      "n_gpus": world, "k": K, "results": out,
                      "rank0_exact_path_queries": xq, "rank0_certificate_fallbacks": xf}), flush=True)
if world > 1:
    dist.destroy_process_group()
