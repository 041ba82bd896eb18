#!/usr/bin/env python
"""IVF-PQ candidate generation + device refine on the K5 shape (BASELINE.json configs[4]: refine on 1 M x 768 FP16,
nprobe sweep) with candidates produced by this repo's own IVF-PQ (the reference uses FAISS, absent here).

Mirrors the reference's nprobe x REFINE_K grid (Performance.md:673-687: nlist=4096, m=64, Q=1000, k=10): per cell
Recall@10 against the exact L2 neighbours, per-query latency through the host API (one query per call, wall clock, like
apps/nvdb_ivf_eval.cpp:482-487) and batched throughput (all queries in one call).

Synthetic base: a noisy 24-dimensional linear manifold projected on the unit sphere, generated on the device (there is no dataset in the image); recall
numbers therefore describe THIS data, only their shape (monotone in nprobe, threshold effect of the refine depth) is
comparable with the reference's table.

  NVDB_IVF_ROWS (1,000,000)  NVDB_IVF_DIM (768)  NVDB_IVF_NLIST (4096)  NVDB_IVF_M (64)  NVDB_IVF_TRAIN (200000)
  NVDB_IVF_NITER (25)  NVDB_IVF_Q (1000)  NVDB_IVF_LATENT (24)  NVDB_IVF_OPQ (0)  NVDB_IVF_DATA (manifold | clustered)
  NVDB_IVF_CLUSTERS (2048)  NVDB_IVF_REFINE (0,20,50,256)
"""
import importlib
import json
import os
import statistics
import sys
import time
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")

N = int(os.environ.get("NVDB_IVF_ROWS", 1_000_000))
D = int(os.environ.get("NVDB_IVF_DIM", 768))
NLIST = int(os.environ.get("NVDB_IVF_NLIST", 4096))
M = int(os.environ.get("NVDB_IVF_M", 64))
NTRAIN = int(os.environ.get("NVDB_IVF_TRAIN", 200_000))
NITER = int(os.environ.get("NVDB_IVF_NITER", 25))
Q = int(os.environ.get("NVDB_IVF_Q", 1000))
OPQ = int(os.environ.get("NVDB_IVF_OPQ", 0))  # OPQ rounds (the reference's USE_OPQ=1 / OPQ_NITER=25), 0 = plain IVF-PQ
K = 10
dev = torch.device("cuda", 0)
g = torch.Generator(device=dev)
g.manual_seed(20240229)

# low intrinsic dimension (like text embeddings, unlike isotropic noise): points = normalize(z W + small noise), z in R^LAT
LAT = int(os.environ.get("NVDB_IVF_LATENT", 24))
W = torch.randn(LAT, D, device=dev, generator=g)
DATA = os.environ.get("NVDB_IVF_DATA", "manifold")
base = torch.empty(N, D, device=dev, dtype=torch.float16)
if DATA == "clustered":
    # the shape of the reference's data (L2-normalised 384-d text embeddings, Performance.md:15): topical clusters whose
    # centres lie on a low-dimensional manifold, members spread along a few local directions plus isotropic noise
    NC = int(os.environ.get("NVDB_IVF_CLUSTERS", 2048))
    SPREAD = float(os.environ.get("NVDB_IVF_SPREAD", 0.9))   # in-cluster spread along the local directions (centres: norm 1)
    NOISE = float(os.environ.get("NVDB_IVF_NOISE", 0.5))     # isotropic part
    centres = torch.nn.functional.normalize(torch.randn(NC, LAT, device=dev, generator=g) @ W, dim=1)
    W2 = torch.randn(LAT, D, device=dev, generator=g)
for r0 in range(0, N, 100_000):
    r1 = min(N, r0 + 100_000)
    z = torch.randn(r1 - r0, LAT, device=dev, generator=g)
    if DATA == "clustered":
        owner = torch.randint(0, NC, (r1 - r0,), device=dev, generator=g)
        x = centres[owner] + SPREAD * torch.nn.functional.normalize(z @ W2, dim=1) + \
            (NOISE / D ** 0.5) * torch.randn(r1 - r0, D, device=dev, generator=g)
    else:
        x = z @ W + 0.3 * torch.randn(r1 - r0, D, device=dev, generator=g)
    base[r0:r1] = torch.nn.functional.normalize(x, dim=1).half()
qrows = torch.randint(0, N, (Q,), device=dev, generator=g)
queries = torch.nn.functional.normalize(base[qrows].float() + 0.03 * torch.randn(Q, D, device=dev, generator=g), dim=1)
# exact L2 neighbours (unit vectors: same ranking as the dot product), chunked fp32 matmul — measurement plumbing only
truth = torch.empty(Q, K, dtype=torch.int64, device=dev)
best_s = torch.full((Q, K), -1e30, device=dev)
best_i = torch.zeros(Q, K, dtype=torch.int64, device=dev)
qn = (queries * queries).sum(1, keepdim=True)
for r0 in range(0, N, 200_000):
    r1 = min(N, r0 + 200_000)
    b = base[r0:r1].float()
    s = -(qn - 2.0 * queries @ b.T + (b * b).sum(1)[None, :])
    cs, ci = torch.cat([best_s, s], 1).topk(K, dim=1)
    allid = torch.cat([best_i, torch.arange(r0, r1, device=dev).expand(Q, -1)], 1)
    best_s, best_i = cs, allid.gather(1, ci)
truth = best_i.cpu().numpy()
q_host = queries.cpu().numpy()
torch.cuda.synchronize()

idx = nv.FlatIndexCuda.from_device(base.data_ptr(), None, nv.F16, N, D, keepalive=base)
t0 = time.perf_counter()
ivf = nv.IvfPqIndex(idx, nlist=NLIST, pq_m=M, ntrain=NTRAIN, niter=NITER, seed=1234, opq_niter=OPQ)
build_s = time.perf_counter() - t0
print(json.dumps({"what": "build", "rows": N, "dim": D, "nlist": NLIST, "pq_m": M, "ntrain": NTRAIN, "niter": NITER, "opq_niter": OPQ, "data": DATA,
                  "build_s": round(build_s, 2), "index_MB": round(ivf.index_bytes / 1e6, 1),
                  "base_MB": round(N * D * 2 / 1e6, 1)}), flush=True)


def recall(found):
    return float(np.mean([len(set(f.tolist()) & set(t.tolist())) / K for f, t in zip(found, truth)]))


def pct(v, p):
    v = sorted(v)
    return v[min(len(v) - 1, int(round(p / 100 * (len(v) - 1))))]


for nprobe in (8, 16, 32, 64, 128):
    for R in tuple(int(x) for x in os.environ.get("NVDB_IVF_REFINE", "0,20,50,256").split(",")):
        lat = []
        found = np.empty((Q, K), np.uint32)
        for w in range(5):  # warm-up (workspaces)
            (ivf.search(q_host[w:w + 1], nprobe, K) if R == 0 else ivf.search_refine(q_host[w:w + 1], nprobe, R, K))
        for qi in range(Q):
            t0 = time.perf_counter()
            ids, _ = ivf.search(q_host[qi:qi + 1], nprobe, K) if R == 0 else ivf.search_refine(q_host[qi:qi + 1], nprobe, R, K)
            lat.append((time.perf_counter() - t0) * 1e3)
            found[qi] = ids[0]
        ivf.search(q_host, nprobe, K) if R == 0 else ivf.search_refine(q_host, nprobe, R, K)  # workspaces of this shape
        t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            bids, _ = ivf.search(q_host, nprobe, K) if R == 0 else ivf.search_refine(q_host, nprobe, R, K)
        batch_ms = (time.perf_counter() - t0) * 1e3 / reps
        assert np.array_equal(bids, found), "batched call == one query per call"
        print(json.dumps({"what": "search", "nprobe": nprobe, "refine_R": R, "recall_at_10": round(recall(found), 4),
                          "lat_avg_ms": round(statistics.mean(lat), 4), "lat_p95_ms": round(pct(lat, 95), 4),
                          "lat_p99_ms": round(pct(lat, 99), 4), "batch_Q": Q, "batch_ms": round(batch_ms, 3),
                          "batch_qps": round(Q / batch_ms * 1e3, 1)}), flush=True)
ivf.close()
idx.close()
