#!/usr/bin/env python
"""Host-buffer calls (nvdb_b200_search) at a few batch sizes, INT8 and FP16, 2,903,054 x 768: ms per call."""
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
rng = np.random.default_rng(0)
for dt, name in ((nv.I8, "i8"), (nv.F16, "f16")):
    with nv.FlatIndexCuda.synthetic(dt, 2_903_054, 768, 20240229) as idx:
        for b in (8, 32, 64, 128):
            q = rng.standard_normal((b, 768)).astype(np.float32)
            for _ in range(5):
                idx.search_topk_dot_batch(q, 10)
            best = 9e9
            for rep in range(3):
                t0 = time.perf_counter()
                for _ in range(30):
                    idx.search_topk_dot_batch(q, 10)
                best = min(best, (time.perf_counter() - t0) / 30 * 1e3)
            out[f"{name}_b{b}"] = round(best, 4)
print(json.dumps(out))
