#!/usr/bin/env python
"""K5's refine stage beside the reference's own GPU code on the same B200: nvdb::cuda_l2_topk_batch from the UNMODIFIED
src/cuda_refine.cu (oracle/_ref/libnvdb_refcuda.so, compiled for sm_100a — measurement infrastructure, never the
product) against nvdb_b200_refine_l2 on the same base, queries and candidates. FP16 1 M x 768, R = 256, K = 10 (and the
reference's own published shape: D = 384, R = 500, Q = 10^4, Performance_CUDA.md). One JSON line per shape."""
import ctypes as C
import importlib
import json
import os
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
so = ROOT / "oracle" / "_ref" / "libnvdb_refcuda.so"
if not so.exists():
    print(json.dumps({"unavailable": "oracle/_ref/libnvdb_refcuda.so is not built"}))
    sys.exit(0)
ref = C.CDLL(str(so))
ref.refcuda_l2_topk_batch.restype = C.c_int
ref.refcuda_l2_topk_batch.argtypes = [C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p, C.c_uint32,
                                      C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p, C.c_void_p]
N = int(os.environ.get("NVDB_REFINE_ROWS", 1_000_000))
SEED, REPS = 20240229, 5
for D, R, Q, K in ((768, 256, 10_000, 10), (768, 256, 1_000, 10), (384, 500, 10_000, 10)):
    with nv.FlatIndexCuda.synthetic(nv.F16, N, D, SEED, device=0) as idx:
        base, _ = idx.read_rows(0, N)
        rng = np.random.default_rng(1)
        qrows = rng.integers(0, N, Q)
        queries = np.ascontiguousarray(base[qrows].view(np.float16).astype(np.float32))
        cand = rng.integers(0, N, (Q, R)).astype(np.uint32)
        cand[:, 0] = qrows                       # the query's own row is among its candidates
        ours, theirs = [], []
        t = nv.Timing()
        rids = np.empty((Q, K), np.uint32)
        rdist = np.empty((Q, K), np.float32)
        t4 = np.zeros(4, np.float32)
        for rep in range(REPS + 1):              # first round warms both (the reference uploads its base copy once)
            os.environ["CUDA_DBG_TIMING"] = "1"              # in-kernel phase sampling of the first 32 queries
            idx.refine_l2(queries, cand, K, timing=t)      # one chunk, four events: H2D / kernel / D2H
            os.environ.pop("CUDA_DBG_TIMING")
            phases = (t.dbg_write_cycles_avg, t.dbg_dist_cycles_avg, t.dbg_merge_cycles_avg)
            w0 = time.perf_counter()
            ids, dist = idx.refine_l2(queries, cand, K)     # the call a user makes: chunks pipelined through staging
            w1 = time.perf_counter()
            ref.refcuda_l2_topk_batch(base.ctypes.data, nv.F16, N, D, queries.ctypes.data, cand.ctypes.data, Q, R, K,
                                      rids.ctypes.data, rdist.ctypes.data, t4.ctypes.data)
            w2 = time.perf_counter()
            if rep:
                ours.append((t.kernel_ms, t.total_ms, (w1 - w0) * 1e3))
                theirs.append((float(t4[1]), float(t4[3]), (w2 - w1) * 1e3))
        # the same call with page-locked caller buffers (nvdb_b200_host_register: the CUDA_PINNED=1 analogue): no staging copy
        for a in (queries, cand):
            nv.lib.nvdb_b200_host_register(a.ctypes.data, a.nbytes)
        pinned = []
        for rep in range(REPS + 1):
            w0 = time.perf_counter()
            pids, pdist = idx.refine_l2(queries, cand, K)
            if rep:
                pinned.append((time.perf_counter() - w0) * 1e3)
        for a in (queries, cand):
            nv.lib.nvdb_b200_host_unregister(a.ctypes.data)
        assert np.array_equal(pids, ids)
        same = float(np.mean(ids == rids))
        o, r = np.median(np.array(ours), axis=0), np.median(np.array(theirs), axis=0)
        print(json.dumps({
            "shape": f"FP16 {N} x {D}, Q={Q}, R={R}, K={K}", "ids_equal_fraction": same,
            "max_rel_dist_diff": float(np.max(np.abs(dist - rdist) / np.maximum(np.abs(rdist), 1e-6))),
            "ours": {"kernel_ms": float(o[0]), "h2d_kernel_d2h_ms": float(o[1]), "wall_ms": float(o[2]),
                     "wall_ms_registered_caller_buffers": float(np.median(pinned)),
                     "phase_cycles_query_dist_select": [float(x) for x in phases],
                     "us_per_query_kernel": float(o[0]) * 1e3 / Q,
                     "gather_GBps": Q * R * D * 2 / (float(o[0]) * 1e-3) / 1e9},
            "reference_cuda_refine_sm100a": {"kernel_ms": float(r[0]), "h2d_kernel_d2h_ms": float(r[1]), "wall_ms": float(r[2]),
                                             "us_per_query_kernel": float(r[0]) * 1e3 / Q,
                                             "gather_GBps": Q * R * D * 2 / (float(r[0]) * 1e-3) / 1e9},
            "kernel_speedup": float(r[0] / o[0]), "call_speedup": float(r[2] / o[2])}), flush=True)
