#!/usr/bin/env python
"""Batch 1024 on the CTA-pair kernel against its ring depth (NVDB_B200_PAIR_MAX_STAGES), interleaved on one box."""
import importlib
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
dev = torch.device("cuda", 0)
n, d, seed = 2_903_054, 768, 20240229
st = torch.cuda.Stream(device=dev)
for dt, name in ((nv.F16, "f16"), (nv.I8, "i8")):
    with nv.FlatIndexCuda.synthetic(dt, n, d, seed) as idx:
        q = torch.empty((1024, d), dtype=torch.float32, device=dev)
        nv.synth_fill_device(0, nv.F32, 1024, d, seed, 12345, q.data_ptr(), None, st.cuda_stream)
        st.synchronize()
        q += 1e-3 * torch.randn_like(q)
        ids = torch.empty((1024, 10), dtype=torch.int64, device=dev)
        sc = torch.empty((1024, 10), dtype=torch.float32, device=dev)
        idx.set_path(nv.PATH_MMA_EXACT)
        for b in (256, 1024):
            rec = {"dtype": name, "batch": b}
            for rep in range(3):
                for depth in (8, 6, 5, 4, 3):
                    os.environ["NVDB_B200_PAIR_MAX_STAGES"] = str(depth)
                    for _ in range(2):
                        idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                    st.synchronize()
                    idx.profile(True)
                    for _ in range(6):
                        idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                    st.synchronize()
                    scan_ms, merge_ms, calls = idx.profile_read()
                    idx.profile(False)
                    k = f"max{depth}"
                    rec[k] = min(rec.get(k, 9e9), round(scan_ms / max(calls, 1), 4))
            os.environ.pop("NVDB_B200_PAIR_MAX_STAGES")
            print(json.dumps(rec), flush=True)
