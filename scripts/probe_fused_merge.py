#!/usr/bin/env python
"""Batch-1 streaming scan, k = 10 / 100: final merge fused into the last CTA (knob 4 = 1) against the separate merge
kernel (0), interleaved on one box: python scripts/probe_fused_merge.py [rows]"""
import importlib
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
dev = torch.device("cuda", 0)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2_903_054
st = torch.cuda.Stream(device=dev)
with nv.FlatIndexCuda.synthetic(nv.F16, n, 768, 20240229) as idx:
    q = torch.randn((1, 768), device=dev)
    idx.set_path(nv.PATH_GEMV)
    for k in (10, 100):
        ids = torch.empty((1, k), dtype=torch.int64, device=dev)
        sc = torch.empty((1, k), dtype=torch.float32, device=dev)
        rec = {"rows": n, "k": k}
        ref = None
        for rep in range(3):
            for fused in (1, 0):
                idx.set_tuning(4, fused)
                for _ in range(5):
                    idx.search_device(q.data_ptr(), 1, k, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                st.synchronize()
                if ref is None:
                    ref = ids.clone()
                assert torch.equal(ids, ref)
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record(st)
                for _ in range(100):
                    idx.search_device(q.data_ptr(), 1, k, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                e1.record(st)
                st.synchronize()
                key = "fused_ms" if fused else "separate_ms"
                rec[key] = min(rec.get(key, 9e9), round(e0.elapsed_time(e1) / 100, 4))
        idx.set_tuning(4, 1)
        print(json.dumps(rec), flush=True)
