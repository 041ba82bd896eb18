#!/usr/bin/env python
"""A few exact-path searches of one batch size over the 2,903,054 x 768 synthetic base (for ncu):
   python scripts/run_rows_once.py <f16|i8> <batch> [iters]"""
import importlib
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
dt = nv.F16 if sys.argv[1] == "f16" else nv.I8
b = int(sys.argv[2])
iters = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dev = torch.device("cuda", 0)
n, d, seed = 2_903_054, 768, 20240229
st = torch.cuda.Stream(device=dev)
with nv.FlatIndexCuda.synthetic(dt, n, d, seed) as idx:
    q = torch.empty((b, d), dtype=torch.float32, device=dev)
    for i in range(b):
        nv.synth_fill_device(0, nv.F32, 1, d, seed, (i * 104729) % n, q[i].data_ptr(), None, st.cuda_stream)
    st.synchronize()
    ids = torch.empty((b, 10), dtype=torch.int64, device=dev)
    sc = torch.empty((b, 10), dtype=torch.float32, device=dev)
    idx.set_path(nv.PATH_MMA_EXACT)
    for _ in range(iters):
        idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
    st.synchronize()
    print("ok", ids[0, 0].item())
