#!/usr/bin/env python
"""Top-100 batch-1 searches on the 2.9 M x 768 FP16 base (streaming scan -> merge_select_kernel over 148 x 100 keys) and
batch-8 searches on the exact tensor-core path (148 x 16 keys per query): the launches ncu looks at
(scripts/gpu_round46.sh)."""
import importlib
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
N, D, SEED = 2_903_054, 768, 20240229
dev = torch.device("cuda", 0)
idx = nv.FlatIndexCuda.synthetic(nv.F16, N, D, SEED, device=0)
q = torch.empty((8, D), dtype=torch.float32, device=dev)
sp = torch.cuda.current_stream().cuda_stream
for i in range(8):
    nv.synth_fill_device(0, nv.F32, 1, D, SEED, (i * 104729 * 31) % N, q[i].data_ptr(), None, sp)
ids = torch.empty((8, 100), dtype=torch.int64, device=dev)
sc = torch.empty((8, 100), dtype=torch.float32, device=dev)
for it in range(12):
    idx.search_device(q[it % 8].data_ptr(), 1, 100, ids.data_ptr(), sc.data_ptr(), sp)
torch.cuda.synchronize()
assert int(ids[0, 0]) == ((11 % 8) * 104729 * 31) % N
for it in range(6):
    idx.search_device(q.data_ptr(), 8, 10, ids.data_ptr(), sc.data_ptr(), sp)
torch.cuda.synchronize()
print("ok")
