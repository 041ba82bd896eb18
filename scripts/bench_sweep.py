#!/usr/bin/env python
"""Batch sweep on one GPU: FP16 / INT8 2,903,054 x 768, top-10, batch 1..1024, per path (GEMV / MMA / MMA_FAST).
Prints one JSON line per (dtype, path, batch): ms per batch, QPS, base GB/s, tensor TFLOP/s."""
import importlib
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
N = int(os.environ.get("NVDB_BENCH_ROWS", 2_903_054))
D, K, SEED = 768, int(os.environ.get("NVDB_K", 10)), 20240229
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
PATHS = os.environ.get("NVDB_PATHS", "gemv,mma,mma_fast,mma_exact").split(",")
DTYPES = os.environ.get("NVDB_DTYPES", "f16,i8").split(",")
BATCHES = [int(b) for b in os.environ.get("NVDB_BATCHES", "1,2,4,8,16,32,64,128,256,512,1024").split(",")]


def time_search(idx, qs, nq, iters):
    out_ids = torch.empty((nq, K), dtype=torch.int64, device=dev)
    out_sc = torch.empty((nq, K), dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for _ in range(3):
            idx.search_device(qs.data_ptr(), nq, K, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        e0.record(stream)
        for _ in range(iters):
            idx.search_device(qs.data_ptr(), nq, K, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        e1.record(stream)
        stream.synchronize()
    return e0.elapsed_time(e1) / iters


for dtype, name, esz in ((nv.F16, "f16", 2), (nv.I8, "i8", 1)):
    if name not in DTYPES:
        continue
    idx = nv.FlatIndexCuda.synthetic(dtype, N, D, SEED, device=0)
    nbytes = N * D * esz + (4 * N if dtype == nv.I8 else 0)
    qall = torch.empty((1024, D), dtype=torch.float32, device=dev)
    with torch.cuda.stream(stream):
        for i in range(1024):
            nv.synth_fill_device(0, nv.F32, 1, D, SEED, (i * 104729 * 31) % N, qall[i].data_ptr(), None, stream.cuda_stream)
    stream.synchronize()
    qall += 1e-3 * torch.randn_like(qall)
    for pname, path in (("gemv", nv.PATH_GEMV), ("mma", nv.PATH_MMA), ("mma_fast", nv.PATH_MMA_FAST),
                        ("mma_exact", nv.PATH_MMA_EXACT)):
        if (dtype == nv.I8 and pname == "mma_fast") or pname not in PATHS:
            continue
        idx.set_path(path)
        for b in BATCHES:
            if pname == "gemv" and b > 64:
                continue
            iters = max(3, min(50, int(200 / max(b / 8, 1)))) if pname == "gemv" else max(5, min(50, 4096 // b))
            ms = time_search(idx, qall, b, iters)
            limbs = 2 if (pname == "mma" or dtype == nv.I8) else 1
            extra = {}
            if pname == "mma_exact":
                nqx, nfb = idx.exact_stats(reset=True)
                extra["certificate_fallback_rate"] = round(nfb / max(nqx, 1), 4)
            print(json.dumps({"dtype": name, "path": pname, "batch": b, "k": K, "ms_per_batch": round(ms, 4),
                              "qps": round(b / ms * 1e3, 1), "base_GBps_one_pass": round(nbytes / ms / 1e6, 1),
                              "useful_TFLOPs": round(2.0 * b * N * D / ms / 1e9, 1),
                              "issued_TFLOPs": round(2.0 * ((b + 127) // 128 * 128) * N * D * limbs / ms / 1e9, 1)
                              if pname != "gemv" else None, **extra}), flush=True)
    idx.close()
