#!/usr/bin/env python
"""Single-query IVF-PQ search a few times (for an ncu launch list: which kernel owns the latency)."""
import importlib, os, sys, time
from pathlib import Path
import numpy as np
ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
N, D = int(os.environ.get("NVDB_IVF_ROWS", 1_000_000)), 768
idx = nv.FlatIndexCuda.synthetic(nv.F16, N, D, 5)
ivf = nv.IvfPqIndex(idx, nlist=4096, pq_m=64, ntrain=100_000, niter=int(os.environ.get("NVDB_IVF_NITER", 4)), seed=1)
rows, _ = idx.read_rows(0, 64)
q = rows.view(np.float16).astype(np.float32)
for nprobe, R in ((64, 0), (64, 50)):
    for i in range(8):
        ivf.search(q[i:i + 1], nprobe, 10) if R == 0 else ivf.search_refine(q[i:i + 1], nprobe, R, 10)
    t0 = time.perf_counter()
    for i in range(50):
        ivf.search(q[i:i + 1], nprobe, 10) if R == 0 else ivf.search_refine(q[i:i + 1], nprobe, R, 10)
    print(nprobe, R, "avg ms", (time.perf_counter() - t0) * 1e3 / 50, flush=True)
