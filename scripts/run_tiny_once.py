#!/usr/bin/env python
"""A few single-query host searches over a small FP16 base (for ncu: the fixed cost of the streaming-scan kernel):
   python scripts/run_tiny_once.py <rows>"""
import importlib
import sys
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
n = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
with nv.FlatIndexCuda.synthetic(nv.F16, n, 768, 1) as idx:
    q = np.random.default_rng(0).standard_normal((1, 768)).astype(np.float32)
    for _ in range(6):
        ids, sc = idx.search_topk_dot_batch(q, 10)
    print("ok", ids[0, 0])
