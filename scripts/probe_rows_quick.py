#!/usr/bin/env python
"""Scan-kernel ms of the exact path at a few batch sizes (row-tile kernel forced), for same-box A/Bs of two library builds."""
import importlib
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
dev = torch.device("cuda", 0)
n, d, seed = 2_903_054, 768, 20240229
st = torch.cuda.Stream(device=dev)
out = {}
for dt, name in ((nv.I8, "i8"), (nv.F16, "f16")):
    with nv.FlatIndexCuda.synthetic(dt, n, d, seed) as idx:
        q = torch.empty((64, d), dtype=torch.float32, device=dev)
        for i in range(64):
            nv.synth_fill_device(0, nv.F32, 1, d, seed, (i * 104729) % n, q[i].data_ptr(), None, st.cuda_stream)
        st.synchronize()
        q += 1e-3 * torch.randn_like(q)
        ids = torch.empty((64, 10), dtype=torch.int64, device=dev)
        sc = torch.empty((64, 10), dtype=torch.float32, device=dev)
        idx.set_tuning(7, 2)
        idx.set_path(nv.PATH_MMA_EXACT)
        for b in (8, 16, 32, 48, 64):
            best = 9e9
            for rep in range(3):
                for _ in range(3):
                    idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                st.synchronize()
                idx.profile(True)
                for _ in range(20):
                    idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                st.synchronize()
                scan_ms, merge_ms, calls = idx.profile_read()
                idx.profile(False)
                best = min(best, scan_ms / max(calls, 1))
            out[f"{name}_b{b}"] = round(best, 4)
print(json.dumps(out))
