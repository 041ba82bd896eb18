#!/usr/bin/env python
"""SURVEY.md 8(d) second distribution at the K2 / K3 size: 2,903,054 x 768 rows drawn around 256 Gaussian centres
(sigma 0.3, re-normalised), FP16 and INT8, top-10 (NVDB_K), batch 1 .. 1024.

For every (dtype, batch) prints one JSON line with the same measurement on the isotropic base beside it:
  * ms per batch of the streaming scan (batch <= 8) and of the exact tensor-core path (batch >= 3, what AUTO picks),
  * the exact path's certificate fallback rate (a clustered base packs a whole cluster's scores close together, which
    is what the certificate has to separate), split into queries settled by the second tier (wider candidate lists;
    NVDB_B200_EXACT_TIER2=0 switches it off for an A/B) and queries that ended on the streaming scan,
  * `equal_to_gemv`: ids and FP32 score bits of the exact path == the streaming scan on the same index, every query.
The base is generated on the device with torch (plumbing only); queries are rows of the base plus 1e-3 noise."""
import importlib
import json
import os
import sys
from pathlib import Path

import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
N = int(os.environ.get("NVDB_BENCH_ROWS", 2_903_054))
D, K, SEED = 768, int(os.environ.get("NVDB_K", 10)), 20240229
BATCHES = [int(b) for b in os.environ.get("NVDB_BATCHES", "1,8,128,1024").split(",")]
DTYPES = os.environ.get("NVDB_DTYPES", "f16,i8").split(",")
KINDS = os.environ.get("NVDB_KINDS", "isotropic,clustered").split(",")
dev = torch.device("cuda", 0)
stream = torch.cuda.Stream(device=dev)


def make_rows(kind, gen, centres, lo, hi):
    m = hi - lo
    if kind == "clustered":
        owner = torch.randint(0, centres.shape[0], (m,), device=dev, generator=gen)
        x = centres[owner] + (0.3 / D ** 0.5) * torch.randn((m, D), device=dev, generator=gen)
    else:
        x = torch.randn((m, D), device=dev, generator=gen)
    return x / x.norm(dim=1, keepdim=True)


def make_base(kind, dtype):
    gen = torch.Generator(device=dev).manual_seed(SEED)
    centres = torch.randn((256, D), device=dev, generator=gen)
    centres /= centres.norm(dim=1, keepdim=True)
    payload = torch.empty((N, D), dtype=torch.float16 if dtype == nv.F16 else torch.int8, device=dev)
    scales = torch.empty(N, dtype=torch.float32, device=dev) if dtype == nv.I8 else None
    qrows = torch.arange(1024, device=dev) * 2833 % N
    queries = torch.empty((1024, D), dtype=torch.float32, device=dev)
    step = 131072
    for lo in range(0, N, step):
        hi = min(N, lo + step)
        x = make_rows(kind, gen, centres, lo, hi)
        sel = (qrows >= lo) & (qrows < hi)
        queries[sel] = x[qrows[sel] - lo]
        if dtype == nv.F16:
            payload[lo:hi] = x.half()
        else:   # per-row symmetric scale (maxabs / 127), round to nearest: the reference quantiser's scheme
            s = x.abs().amax(dim=1).clamp_min(1e-30) / 127.0
            payload[lo:hi] = torch.round(x / s[:, None]).clamp_(-127, 127).to(torch.int8)
            scales[lo:hi] = s
    queries += 1e-3 / D ** 0.5 * torch.randn(queries.shape, device=dev, generator=gen)
    torch.cuda.synchronize()
    idx = nv.FlatIndexCuda.from_device(payload.data_ptr(), scales.data_ptr() if scales is not None else None, dtype, N,
                                       D, device=0, keepalive=(payload, scales))
    return idx, queries


def run(idx, qs, nq, iters):
    out_ids = torch.empty((nq, K), dtype=torch.int64, device=dev)
    out_sc = torch.empty((nq, K), dtype=torch.float32, device=dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for _ in range(3):
            idx.search_device(qs.data_ptr(), nq, K, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        idx.exact_stats(reset=True)
        e0.record(stream)
        for _ in range(iters):
            idx.search_device(qs.data_ptr(), nq, K, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        e1.record(stream)
        stream.synchronize()
    return e0.elapsed_time(e1) / iters, out_ids, out_sc


for dtype, name in ((nv.F16, "f16"), (nv.I8, "i8")):
    if name not in DTYPES:
        continue
    for kind in KINDS:
        idx, queries = make_base(kind, dtype)
        for b in BATCHES:
            line = {"dtype": name, "base": kind, "rows": N, "batch": b, "k": K}
            ref = None
            if b <= 8:
                idx.set_path(nv.PATH_GEMV)
                ms, gi, gs = run(idx, queries, b, 20)
                line["gemv_ms"] = round(ms, 4)
                ref = (gi.clone(), gs.clone())
            if b >= 3:
                idx.set_path(nv.PATH_MMA_EXACT)
                ms, ei, es = run(idx, queries, b, max(5, min(30, 4096 // b)))
                nqx, nfb = idx.exact_stats()
                line["exact_ms"] = round(ms, 4)
                line["exact_qps"] = round(b / ms * 1e3, 1)
                t2, streamed = idx.exact_tiers()
                line["certificate_fallback_rate"] = round(nfb / max(nqx, 1), 4)
                line["second_tier_rate"] = round(t2 / max(nqx, 1), 4)
                line["streaming_scan_rate"] = round(streamed / max(nqx, 1), 4)
                idx.exact_stats(reset=True)
                if ref is None:   # the same batch through the streaming scan, 8 queries per pass
                    idx.set_path(nv.PATH_GEMV)
                    m = b
                    ref = [], []
                    for q0 in range(0, m, 8):
                        _, gi, gs = run(idx, queries[q0:q0 + 8].contiguous(), min(8, m - q0), 1)
                        ref[0].append(gi.clone()), ref[1].append(gs.clone())
                    ref = (torch.cat(ref[0]), torch.cat(ref[1]))
                m = ref[0].shape[0]
                line["equal_to_gemv"] = bool(torch.equal(ei[:m], ref[0]) and
                                             torch.equal(es[:m].view(torch.int32), ref[1].view(torch.int32)))
            top = ref[1][0].tolist()
            line["q0_scores_1_2_10"] = [round(top[0], 4), round(top[1], 4), round(top[min(K, len(top)) - 1], 4)]
            print(json.dumps(line), flush=True)
        idx.close()
        del idx
        torch.cuda.empty_cache()
