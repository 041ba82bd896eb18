#!/usr/bin/env python
"""Fixed cost of the exact tensor-core path: device-timed search_device of b queries over tiny / small bases, per dtype,
with the launch list of one call (kernel names come from the library's launch counter only; use ncu for durations)."""
import importlib
import json
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
dev = torch.device("cuda", 0)
st = torch.cuda.Stream(device=dev)
out = {}
for dt, name in ((nv.I8, "i8"), (nv.F16, "f16")):
    for n in (4096, 200_000):
        with nv.FlatIndexCuda.synthetic(dt, n, 768, 1) as idx:
            q = torch.randn((64, 768), device=dev)
            ids = torch.empty((64, 10), dtype=torch.int64, device=dev)
            sc = torch.empty((64, 10), dtype=torch.float32, device=dev)
            idx.set_tuning(7, 2)
            idx.set_path(nv.PATH_MMA_EXACT)
            for b in (8, 64):
                for _ in range(20):
                    idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                st.synchronize()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                l0 = nv.launch_count()
                e0.record(st)
                for _ in range(200):
                    idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                e1.record(st)
                st.synchronize()
                idx.profile(True)
                for _ in range(20):
                    idx.search_device(q.data_ptr(), b, 10, ids.data_ptr(), sc.data_ptr(), st.cuda_stream)
                st.synchronize()
                scan_ms, merge_ms, calls = idx.profile_read()
                idx.profile(False)
                out[f"{name}_rows{n}_b{b}"] = {"us_per_call": round(e0.elapsed_time(e1) / 200 * 1e3, 1),
                                                 "scan_us": round(scan_ms / max(calls, 1) * 1e3, 1),
                                                 "after_scan_us": round(merge_ms / max(calls, 1) * 1e3, 1),
                                                 "launches_per_call": (nv.launch_count() - l0) / 220}
print(json.dumps(out, indent=1))
