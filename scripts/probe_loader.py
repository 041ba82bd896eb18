#!/usr/bin/env python
"""Loader probe: host -> device upload GB/s of nvdb_b200_index_create for a 2,903,054 x 768 FP16 base
already resident in host memory, beside torch's pageable copy (what the reference's single cudaMemcpy does)."""
import importlib
import json
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
n, d = 2_903_054, 768
x = np.empty((n, d), np.uint16)
x[:] = 0x3C00          # touch every page
out = {"bytes": x.nbytes, "upload_s": [], "pageable_s": []}
for _ in range(3):
    with nv.FlatIndexCuda.from_host(x, None, nv.F16) as idx:
        out["upload_s"].append(idx.upload_info()[0])
t = torch.from_numpy(x.view(np.int16))
for _ in range(2):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    dd = t.cuda()
    torch.cuda.synchronize()
    out["pageable_s"].append(time.perf_counter() - t0)
    del dd
out["upload_GBps_best"] = x.nbytes / min(out["upload_s"]) / 1e9
out["pageable_GBps_best"] = x.nbytes / min(out["pageable_s"]) / 1e9
print(json.dumps(out))
