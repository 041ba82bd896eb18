#!/usr/bin/env python
"""A few device-resident refine launches at one shape (for ncu): python scripts/run_refine_once.py <D> <Q> <R> [iters]"""
import importlib
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
D, Q, R = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
iters = int(sys.argv[4]) if len(sys.argv) > 4 else 4
dev = torch.device("cuda", 0)
n, K = 1_000_000, 10
st = torch.cuda.Stream(device=dev)
with nv.FlatIndexCuda.synthetic(nv.F16, n, D, 20240229) as idx:
    g = torch.Generator(device=dev).manual_seed(1)
    cand = torch.randint(0, n, (Q, R), dtype=torch.int32, device=dev, generator=g)
    q = torch.randn((Q, D), device=dev, generator=g)
    q /= q.norm(dim=1, keepdim=True)
    oid = torch.empty((Q, K), dtype=torch.int32, device=dev)
    od = torch.empty((Q, K), dtype=torch.float32, device=dev)
    for _ in range(iters):
        idx.refine_l2_device(q.data_ptr(), cand.data_ptr(), Q, R, K, oid.data_ptr(), od.data_ptr(), st.cuda_stream)
    st.synchronize()
    print("ok", int(oid[0, 0]))
