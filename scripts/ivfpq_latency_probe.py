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

# device-side only: queries and candidates stay on the device, CUDA events around 200 back-to-back calls
import torch
dev = torch.device("cuda", 0)
qd = torch.from_numpy(q).to(dev)
cand = torch.empty((64, 256), dtype=torch.int32, device=dev)
st = torch.cuda.Stream(device=dev)
for nprobe, R, nq in ((8, 10, 1), (64, 10, 1), (64, 50, 1), (64, 256, 1), (64, 50, 64)):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(st):
        for _ in range(5):
            ivf.search_device(qd.data_ptr(), nq, nprobe, R, cand.data_ptr(), None, st.cuda_stream)
        e0.record(st)
        for _ in range(200):
            ivf.search_device(qd.data_ptr(), nq, nprobe, R, cand.data_ptr(), None, st.cuda_stream)
        e1.record(st)
        st.synchronize()
    print("device-only nq", nq, "nprobe", nprobe, "R", R, "us per call", e0.elapsed_time(e1) * 1e3 / 200, flush=True)
