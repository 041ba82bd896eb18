#!/usr/bin/env python
"""Interleaved A/B timing of search configurations on one GPU (tensor-core kernels run power-limited, so back-to-back
sweeps drift by +-15 %; here every round visits every configuration once and the report is min / median over rounds).

  NVDB_AB="f16:mma_fast:1024:10:seed=1;f16:mma_fast:1024:10:seed=0" python scripts/ab_bench.py
  fields: dtype:path:batch:k[:knob=value,...]   knobs: seed (MMA_SEED), ctas (GEMV_CTAS), pair (MMA_2CTA)
"""
import importlib
import json
import os
import statistics
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
N = int(os.environ.get("NVDB_BENCH_ROWS", 2_903_054))
D, SEED = 768, 20240229
ROUNDS = int(os.environ.get("NVDB_AB_ROUNDS", 5))
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)
PATHS = {"gemv": nv.PATH_GEMV, "mma": nv.PATH_MMA, "mma_fast": nv.PATH_MMA_FAST, "mma_exact": nv.PATH_MMA_EXACT,
         "auto": nv.PATH_AUTO}
KNOBS = {"seed": nv.capi.TUNE_MMA_SEED, "ctas": nv.capi.TUNE_GEMV_CTAS, "pair": nv.capi.TUNE_MMA_2CTA, "fuse": nv.capi.TUNE_FUSED_MERGE, "res": nv.capi.TUNE_MMA_RESIDENT}
DEFAULTS = {"seed": 1, "ctas": 1, "pair": 1, "fuse": 1, "res": 1}
DT = {"f16": nv.F16, "i8": nv.I8, "f32": nv.F32}

spec = os.environ.get("NVDB_AB", "f16:mma_fast:1024:10:seed=1;f16:mma_fast:1024:10:seed=0")
configs = []
for item in spec.split(";"):
    f = item.split(":")
    knobs = dict(DEFAULTS)
    if len(f) > 4 and f[4]:
        for kv in f[4].split(","):
            a, b = kv.split("=")
            knobs[a] = int(b)
    configs.append({"dtype": f[0], "path": f[1], "batch": int(f[2]), "k": int(f[3]), "knobs": knobs, "ms": []})

indices = {}
qall = None
for c in configs:
    if c["dtype"] not in indices:
        indices[c["dtype"]] = nv.FlatIndexCuda.synthetic(DT[c["dtype"]], N, D, SEED, device=0)
if qall is None:
    qall = torch.empty((1024, D), dtype=torch.float32, device=dev)
    with torch.cuda.stream(stream):
        for i in range(1024):
            nv.synth_fill_device(0, nv.F32, 1, D, SEED, (i * 104729 * 31) % N, qall[i].data_ptr(), None, stream.cuda_stream)
    stream.synchronize()
    qall += 1e-3 * torch.randn_like(qall)


def run(c, iters):
    idx = indices[c["dtype"]]
    idx.set_path(PATHS[c["path"]])
    for name, val in c["knobs"].items():
        idx.set_tuning(KNOBS[name], val)
    nq, k = c["batch"], c["k"]
    out_ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
    out_sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        idx.search_device(qall.data_ptr(), nq, k, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        e0.record(stream)
        for _ in range(iters):
            idx.search_device(qall.data_ptr(), nq, k, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        e1.record(stream)
        stream.synchronize()
    return e0.elapsed_time(e1) / iters


for c in configs:   # warm-up (workspaces, norms)
    run(c, 2)
for r in range(ROUNDS):
    for c in configs:
        c["ms"].append(run(c, max(3, min(30, 2048 // c["batch"]))))
for c in configs:
    ms = c.pop("ms")
    c["ms_min"] = round(min(ms), 4)
    c["ms_median"] = round(statistics.median(ms), 4)
    c["qps_median"] = round(c["batch"] / c["ms_median"] * 1e3, 1)
    if c["path"] == "mma_exact":
        nqx, nfb = indices[c["dtype"]].exact_stats(reset=False)
        c["exact_fallbacks_total"] = [nqx, nfb]
    print(json.dumps(c), flush=True)
