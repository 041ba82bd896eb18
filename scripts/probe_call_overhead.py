#!/usr/bin/env python
"""Fixed cost of one host-buffer call: nvdb_b200_search of one query over a tiny base (the kernel itself takes ~5 us)."""
import importlib
import json
import sys
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
out = {}
for n in (2048, 200_000):
    with nv.FlatIndexCuda.synthetic(nv.F16, n, 768, 1) as idx:
        q = np.random.default_rng(0).standard_normal((1, 768)).astype(np.float32)
        t = nv.Timing()
        for _ in range(200):
            idx.search_topk_dot_batch(q, 10)
        t0 = time.perf_counter()
        for _ in range(2000):
            idx.search_topk_dot_batch(q, 10)
        wall = (time.perf_counter() - t0) / 2000 * 1e6
        idx.search_topk_dot_batch(q, 10, t)
        out[f"rows_{n}"] = {"wall_us_per_call": round(wall, 2), "timed_call_kernel_us": round(t.kernel_ms * 1e3, 2),
                            "timed_call_total_us": round(t.total_ms * 1e3, 2)}
print(json.dumps(out))
