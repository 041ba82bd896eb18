#!/usr/bin/env python
"""One small invocation of every kernel family of libnvdb_b200.so, for compute-sanitizer (memcheck / racecheck / synccheck):

    compute-sanitizer --tool <tool> --log-file gpurun_out/sanitizer_<tool>.log python scripts/sanitize_case.py

Shapes are tiny (a sanitizer run is 20-100x slower) but chosen so that every code path runs: the TMA streaming scan with
the fused last-CTA merge (f32 / f16 / i8, 1 / 2 / 8 queries per pass, k = 10 and 100 -> separate merge kernels), the generic
scan (ragged dim), the tcgen05 scans (single CTA f16 one / two limbs, i8 with and without the resident limb, CTA pairs,
shared-memory heaps for k = 100, threshold seeding), the exact path with both certificate tiers, the row-tile kernel for
small batches, refine, converters, the exchange kernel in its split and fused forms (one stream, pushes first) and a small
IVF-PQ build + search. Results are checked against each other (exact path == streaming scan), not against the oracle:
this script must run on a box where only the product library exists."""
import importlib
import os
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
F32, F16, I8 = nv.F32, nv.F16, nv.I8
dev = torch.device("cuda", 0)
SEED = 11


def queries_of(n, d, nq):
    q = torch.empty((nq, d), dtype=torch.float32, device=dev)
    s = torch.cuda.current_stream().cuda_stream
    for i in range(nq):
        nv.synth_fill_device(0, F32, 1, d, SEED, (i * 7919) % n, q[i].data_ptr(), None, s)
    torch.cuda.synchronize()
    return q.cpu().numpy() + 0.01 * np.random.default_rng(1).standard_normal((nq, d)).astype(np.float32)


def same(a, b, what):
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1].view(np.uint32), b[1].view(np.uint32)), what


def scans():
    for dt, n, d in ((F32, 3000, 128), (F16, 5000, 768), (I8, 5000, 768), (F16, 700, 72)):
        with nv.FlatIndexCuda.synthetic(dt, n, d, SEED) as idx:
            for nq, k in ((1, 10), (2, 10), (8, 10), (1, 100), (5, 100)):
                q = queries_of(n, d, nq)
                idx.set_path(nv.PATH_GEMV)
                ref = idx.search_topk_dot_batch(q, k)
                if dt != F32 and d % 16 == 0:
                    idx.set_path(nv.PATH_MMA_EXACT)
                    same(idx.search_topk_dot_batch(q, k), ref, f"exact path dt={dt} nq={nq} k={k}")
                    for path in (nv.PATH_MMA, nv.PATH_MMA_FAST):
                        if path == nv.PATH_MMA_FAST and dt == I8:
                            continue
                        idx.set_path(path)
                        ids, _ = idx.search_topk_dot_batch(q, k)
                        assert ids[0, 0] == ref[0][0, 0]
    # larger batches: CTA pairs (256 queries), seeding (knob 2 = always), heaps (k = 100), odd tile counts, tier 2
    for dt in (F16, I8):
        with nv.FlatIndexCuda.synthetic(dt, 40_000, 128, SEED) as idx:
            for nq, k in ((130, 10), (256, 10), (256, 100), (384, 10)):
                q = queries_of(40_000, 128, nq)
                idx.set_path(nv.PATH_GEMV)
                ref = idx.search_topk_dot_batch(q[:16], k)
                for seed in (1, 2):
                    idx.set_tuning(1, seed)
                    idx.set_path(nv.PATH_MMA_EXACT)
                    got = idx.search_topk_dot_batch(q, k)
                    same((got[0][:16], got[1][:16]), ref, f"exact batch dt={dt} nq={nq} k={k} seed={seed}")
                idx.set_tuning(1, 1)
    # duplicates force both certificate tiers and the streaming-scan fallback
    base = np.repeat(np.random.default_rng(2).standard_normal((64, 128)).astype(np.float16), 40, axis=0)
    with nv.FlatIndexCuda.from_host(base.view(np.uint16), None, F16) as idx:
        q = base[::40][:12].astype(np.float32)
        idx.set_path(nv.PATH_GEMV)
        ref = idx.search_topk_dot_batch(q, 10)
        idx.set_path(nv.PATH_MMA_EXACT)
        same(idx.search_topk_dot_batch(q, 10), ref, "duplicates")


def refine_convert():
    n, d = 20_000, 768
    with nv.FlatIndexCuda.synthetic(F16, n, d, SEED) as idx:
        q = queries_of(n, d, 33)
        cand = np.random.default_rng(3).integers(0, n, (33, 256)).astype(np.uint32)
        cand[:, 5] = 0xFFFFFFFF
        ids, dist = idx.refine_l2(q, cand, 10)
        assert np.all(np.diff(dist, axis=1) >= 0)
    x = np.random.default_rng(4).standard_normal((1000, 96)).astype(np.float32)
    h, _ = nv.capi.convert(x, F32, F16)
    assert np.array_equal(h.view(np.float16), x.astype(np.float16))
    nv.capi.convert(x, F32, I8)


def exchange():
    S, nq, k = 3, 5, 10
    xs = [nv.capi.PeerExchange(0, S, r, 8 * nv.packed_words(nq, k)) for r in range(S)]
    wins = [x.window() for x in xs]
    for x in xs:
        x.connect(windows=wins)
    stream = torch.cuda.Stream(device=dev)   # a real stream: handle 0 means "the index's own stream" to the C ABI
    sp = stream.cuda_stream
    n, d = 9000, 128
    shards = []
    for r in range(S):
        r0, nr = nv.shard_rows(n, S, r)
        shards.append(nv.FlatIndexCuda.synthetic(F16, nr, d, SEED, first_row=r0, row_offset=r0))
        shards[-1].set_path(nv.PATH_GEMV)
    q = torch.from_numpy(queries_of(n, d, nq)).to(dev)
    with nv.FlatIndexCuda.synthetic(F16, n, d, SEED) as whole:
        whole.set_path(nv.PATH_GEMV)
        ref = whole.search_topk_dot_batch(q.cpu().numpy(), k)
    packed = [torch.zeros(nv.packed_words(nq, k), dtype=torch.int64, device=dev) for _ in range(S)]
    for fused in range(S):     # the rank under test runs the one-call form AFTER the others' pushes (never two waiting launches)
        outs = [(torch.zeros((nq, k), dtype=torch.int64, device=dev), torch.zeros((nq, k), dtype=torch.float32, device=dev)) for _ in range(S)]
        for r in range(S):
            if r != fused:
                iv = packed[r][:nq * k].view(nq, k)
                sv = packed[r][nq * k:].view(torch.float32)[:nq * k].view(nq, k)
                shards[r].search_device(q.data_ptr(), nq, k, iv.data_ptr(), sv.data_ptr(), sp)
                xs[r].merge(packed[r].data_ptr(), nq, k, None, None, phases=nv.capi.XCHG_PUSH, stream=sp)
        xs[fused].search(shards[fused], q.data_ptr(), nq, k, outs[fused][0].data_ptr(), outs[fused][1].data_ptr(), sp)
        for r in range(S):
            if r != fused:
                xs[r].merge(None, nq, k, outs[r][0].data_ptr(), outs[r][1].data_ptr(), phases=nv.capi.XCHG_MERGE, stream=sp)
        torch.cuda.synchronize()
        for r in range(S):
            assert xs[r].status() == 0
            same((outs[r][0].cpu().numpy().view(np.uint64), outs[r][1].cpu().numpy()), ref, f"exchange fused={fused} rank={r}")
    for x in xs:
        x.close()
    for s in shards:
        s.close()


def ivfpq():
    n, d = 20_000, 64
    with nv.FlatIndexCuda.synthetic(F16, n, d, SEED) as idx:
        ivf = nv.IvfPqIndex(idx, nlist=64, pq_m=8, ntrain=5000, niter=3, seed=1)
        q = queries_of(n, d, 9)
        ivf.search(q, 8, 32)
        ids, dist = ivf.search_refine(q, 8, 32, 5)
        assert ids.shape == (9, 5)
        ivf.close()


if __name__ == "__main__":
    which = sys.argv[1:] or ["scans", "refine_convert", "exchange", "ivfpq"]
    l0 = nv.launch_count()
    for w in which:
        globals()[w]()
        torch.cuda.synchronize()
        print(f"{w}: ok", flush=True)
    print(f"sanitize_case ok: {nv.launch_count() - l0} kernel launches", flush=True)
