#!/usr/bin/env python
"""Row-tile kernel probe: parity on a small case, then ms per batch at full size (2,903,054 x 768) with the knob on / off.
  python scripts/probe_rows.py [small|perf] ..."""
import importlib
import json
import sys
from pathlib import Path

import numpy as np
import torch

ROOT = Path(__file__).resolve().parent.parent
sys.path.insert(0, str(ROOT))
nv = importlib.import_module("nano-vectordb_b200")
dev = torch.device("cuda", 0)
SEED, DIM = 20240229, 768


def queries(n, d, nq, stream):
    q = torch.empty((nq, d), dtype=torch.float32, device=dev)
    for i in range(nq):
        nv.synth_fill_device(0, nv.F32, 1, d, SEED, (i * 104729) % n, q[i].data_ptr(), None, stream.cuda_stream)
    stream.synchronize()
    return q + 1e-3 * torch.randn_like(q)


def small():
    st = torch.cuda.Stream(device=dev)
    for dt in (nv.F16, nv.I8):
        for n, d in ((20000, 128), (3000, 768)):
            with nv.FlatIndexCuda.synthetic(dt, n, d, SEED) as idx:
                for nq in (3, 8, 64):
                    q = queries(n, d, nq, st).cpu().numpy()
                    idx.set_path(nv.PATH_GEMV)
                    g = idx.search_topk_dot_batch(q, 10)
                    idx.set_path(nv.PATH_MMA_EXACT)
                    x = idx.search_topk_dot_batch(q, 10)
                    ok = np.array_equal(g[0], x[0]) and np.array_equal(g[1].view(np.uint32), x[1].view(np.uint32))
                    print(f"small dt={dt} n={n} d={d} nq={nq}: {'OK' if ok else 'MISMATCH'} stats={idx.exact_stats(reset=True)}", flush=True)
                    if not ok:
                        print(g[0][0], x[0][0], g[1][0], x[1][0])
                        sys.exit(1)


def perf():
    """interleaved A/B per batch size: rows kernel with seeding (1), without (2), query-tile kernels (0), streaming scan"""
    import time
    st = torch.cuda.Stream(device=dev)
    n = 2_903_054
    for dt, name in ((nv.F16, "f16"), (nv.I8, "i8")):
        with nv.FlatIndexCuda.synthetic(dt, n, DIM, SEED) as idx:
            q = queries(n, DIM, 128, st)
            qh = q.cpu().numpy()
            out_ids = torch.empty((128, 10), dtype=torch.int64, device=dev)
            out_sc = torch.empty((128, 10), dtype=torch.float32, device=dev)

            def run(b, iters=20):
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                with torch.cuda.stream(st):
                    for _ in range(3):
                        idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                    st.synchronize()
                    idx.profile(True)
                    e0.record(st)
                    for _ in range(iters):
                        idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                    e1.record(st)
                    st.synchronize()
                scan_ms, merge_ms, calls = idx.profile_read()
                idx.profile(False)
                return e0.elapsed_time(e1) / iters, scan_ms / max(calls, 1)

            for b in (1, 2, 3, 4, 8, 16, 32, 48, 64, 128):
                rec = {"dtype": name, "batch": b}
                for rep in range(2):
                    for label, path, rows in (("rows", nv.PATH_MMA_EXACT, 1), ("rows_noseed", nv.PATH_MMA_EXACT, 2),
                                              ("tile", nv.PATH_MMA_EXACT, 0), ("gemv", nv.PATH_GEMV, 1)):
                        if label == "gemv" and b > 8:
                            continue
                        idx.set_tuning(7, rows)
                        idx.set_path(path)
                        ms, kms = run(b)
                        key = f"{label}_ms"
                        if key not in rec or ms < rec[key]:
                            rec[key], rec[f"{label}_scan"] = round(ms, 4), round(kms, 4)
                idx.set_tuning(7, 1)
                idx.set_path(nv.PATH_MMA_EXACT)
                t0 = time.perf_counter()
                for _ in range(10):
                    idx.search_topk_dot_batch(qh[:b], 10)
                rec["rows_e2e_ms"] = round((time.perf_counter() - t0) * 100, 4)
                rec["fallbacks"] = idx.exact_stats(reset=True)[1]
                print(json.dumps(rec), flush=True)


def stages():
    """ring depth experiment (NVDB_B200_ROWS_MAX_STAGES): is a batch limited by the ring next to the resident operand
    or by its epilogue? INT8, rows kernel with seeding."""
    import os
    st = torch.cuda.Stream(device=dev)
    n = 2_903_054
    for dt, name in ((nv.I8, "i8"), (nv.F16, "f16")):
        with nv.FlatIndexCuda.synthetic(dt, n, DIM, SEED) as idx:
            q = queries(n, DIM, 128, st)
            out_ids = torch.empty((128, 10), dtype=torch.int64, device=dev)
            out_sc = torch.empty((128, 10), dtype=torch.float32, device=dev)
            idx.set_tuning(7, 2)
            idx.set_path(nv.PATH_MMA_EXACT)
            for b in (8, 16, 32, 48, 64):
                rec = {"dtype": name, "batch": b}
                for rep in range(2):
                    for ms_ in (6, 5, 4, 3):
                        os.environ["NVDB_B200_ROWS_MAX_STAGES"] = str(ms_)
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        for _ in range(3):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        st.synchronize()
                        idx.profile(True)
                        for _ in range(20):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        st.synchronize()
                        scan_ms, merge_ms, calls = idx.profile_read()
                        idx.profile(False)
                        k = f"scan_ms_max{ms_}"
                        v = round(scan_ms / max(calls, 1), 4)
                        rec[k] = min(rec.get(k, 9e9), v)
                print(json.dumps(rec), flush=True)
    os.environ.pop("NVDB_B200_ROWS_MAX_STAGES", None)


def depth1():
    """batch 1 and 2: streaming scan vs the row-tile exact path, ring depth 6 / 4 / 3, interleaved on one box"""
    import os
    st = torch.cuda.Stream(device=dev)
    n = 2_903_054
    for dt, name in ((nv.F16, "f16"), (nv.I8, "i8")):
        with nv.FlatIndexCuda.synthetic(dt, n, DIM, SEED) as idx:
            q = queries(n, DIM, 8, st)
            out_ids = torch.empty((8, 10), dtype=torch.int64, device=dev)
            out_sc = torch.empty((8, 10), dtype=torch.float32, device=dev)
            idx.set_tuning(7, 2)
            for b in (1, 2):
                rec = {"dtype": name, "batch": b}
                for rep in range(3):
                    for label, path, env in (("gemv", nv.PATH_GEMV, "NVDB_B200_GEMV_MAX_STAGES"),
                                             ("rows", nv.PATH_MMA_EXACT, "NVDB_B200_ROWS_MAX_STAGES")):
                        idx.set_path(path)
                        for ms_ in (6, 4, 3):
                            os.environ[env] = str(ms_)
                            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                            for _ in range(5):
                                idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                            st.synchronize()
                            e0.record(st)
                            for _ in range(50):
                                idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                            e1.record(st)
                            st.synchronize()
                            k = f"{label}_ms_max{ms_}"
                            rec[k] = min(rec.get(k, 9e9), round(e0.elapsed_time(e1) / 50, 4))
                            os.environ.pop(env)
                print(json.dumps(rec), flush=True)


def depth2():
    """finer ring-depth sweep, batch 1 streaming scan (all dtypes) and rows kernel b8/b32"""
    import os
    st = torch.cuda.Stream(device=dev)
    for dt, name, n in ((nv.F16, "f16", 2_903_054), (nv.I8, "i8", 2_903_054), (nv.F32, "f32", 500_000)):
        with nv.FlatIndexCuda.synthetic(dt, n, DIM, SEED) as idx:
            q = queries(n, DIM, 32, st)
            out_ids = torch.empty((32, 10), dtype=torch.int64, device=dev)
            out_sc = torch.empty((32, 10), dtype=torch.float32, device=dev)
            idx.set_tuning(7, 2)
            for label, path, env, b in (("gemv", nv.PATH_GEMV, "NVDB_B200_GEMV_MAX_STAGES", 1),
                                        ("rows", nv.PATH_MMA_EXACT, "NVDB_B200_ROWS_MAX_STAGES", 8),
                                        ("rows", nv.PATH_MMA_EXACT, "NVDB_B200_ROWS_MAX_STAGES", 32)):
                if label == "rows" and dt == nv.F32:
                    continue
                idx.set_path(path)
                rec = {"dtype": name, "kernel": label, "batch": b}
                for rep in range(5):
                    for ms_ in (6, 5, 4, 3):
                        os.environ[env] = str(ms_)
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        for _ in range(5):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        st.synchronize()
                        e0.record(st)
                        for _ in range(50):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        e1.record(st)
                        st.synchronize()
                        rec.setdefault(f"max{ms_}", []).append(round(e0.elapsed_time(e1) / 50, 4))
                        os.environ.pop(env)
                print(json.dumps(rec), flush=True)


def depth3():
    """f16 small batches: query-tile kernel (ring 6..3) against the row-tile kernel (ring 4), interleaved"""
    import os
    st = torch.cuda.Stream(device=dev)
    n = 2_903_054
    with nv.FlatIndexCuda.synthetic(nv.F16, n, DIM, SEED) as idx:
        q = queries(n, DIM, 128, st)
        out_ids = torch.empty((128, 10), dtype=torch.int64, device=dev)
        out_sc = torch.empty((128, 10), dtype=torch.float32, device=dev)
        idx.set_path(nv.PATH_MMA_EXACT)
        for b in (2, 8, 32, 64, 128):
            rec = {"dtype": "f16", "batch": b}
            for rep in range(3):
                for label, rows, env, depths in (("tile", 0, "NVDB_B200_MMA_MAX_STAGES", (6, 5, 4, 3)),
                                                 ("rows", 2, "NVDB_B200_ROWS_MAX_STAGES", (4,))):
                    if label == "rows" and b > 64:
                        continue
                    idx.set_tuning(7, rows)
                    for ms_ in depths:
                        os.environ[env] = str(ms_)
                        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                        for _ in range(5):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        st.synchronize()
                        e0.record(st)
                        for _ in range(30):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        e1.record(st)
                        st.synchronize()
                        k = f"{label}_max{ms_}"
                        rec[k] = min(rec.get(k, 9e9), round(e0.elapsed_time(e1) / 30, 4))
                        os.environ.pop(env)
            print(json.dumps(rec), flush=True)


def rounds():
    """threshold seeding rounds 0 (off) / 1 / 2 / 3 (NVDB_B200_ROWS_ROUNDS), scan kernel ms"""
    import os
    st = torch.cuda.Stream(device=dev)
    n = 2_903_054
    for dt, name in ((nv.I8, "i8"), (nv.F16, "f16")):
        with nv.FlatIndexCuda.synthetic(dt, n, DIM, SEED) as idx:
            q = queries(n, DIM, 128, st)
            out_ids = torch.empty((128, 10), dtype=torch.int64, device=dev)
            out_sc = torch.empty((128, 10), dtype=torch.float32, device=dev)
            idx.set_tuning(7, 2)
            idx.set_path(nv.PATH_MMA_EXACT)
            for b in (8, 16, 32, 48, 64):
                rec = {"dtype": name, "batch": b}
                for rep in range(3):
                    for r in (0, 1, 2, 3):
                        os.environ["NVDB_B200_ROWS_ROUNDS"] = str(max(r, 1))
                        idx.set_tuning(8, 2 if r else 0)
                        for _ in range(3):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        st.synchronize()
                        idx.profile(True)
                        for _ in range(20):
                            idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                        st.synchronize()
                        scan_ms, merge_ms, calls = idx.profile_read()
                        idx.profile(False)
                        k = f"rounds{r}"
                        rec[k] = min(rec.get(k, 9e9), round(scan_ms / max(calls, 1), 4))
                os.environ.pop("NVDB_B200_ROWS_ROUNDS")
                print(json.dumps(rec), flush=True)


def slope():
    """scan kernel ms against the number of rows: slope = steady-state cost per tile, intercept = start-up phase"""
    import os
    st = torch.cuda.Stream(device=dev)
    for dt, name in ((nv.I8, "i8"),):
        for n in (2_903_054, 1_451_527, 725_763, 362_881):
            with nv.FlatIndexCuda.synthetic(dt, n, DIM, SEED) as idx:
                q = queries(n, DIM, 128, st)
                out_ids = torch.empty((128, 10), dtype=torch.int64, device=dev)
                out_sc = torch.empty((128, 10), dtype=torch.float32, device=dev)
                idx.set_tuning(7, 2)
                idx.set_path(nv.PATH_MMA_EXACT)
                for b in (8, 32, 64):
                    rec = {"dtype": name, "rows": n, "tiles_per_cta": round(n / 128 / 148, 1), "batch": b}
                    for rep in range(3):
                        for r in (0, 1, 3):
                            os.environ["NVDB_B200_ROWS_ROUNDS"] = str(max(r, 1))
                            idx.set_tuning(8, 2 if r else 0)
                            for _ in range(3):
                                idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                            st.synchronize()
                            idx.profile(True)
                            for _ in range(20):
                                idx.search_device(q.data_ptr(), b, 10, out_ids.data_ptr(), out_sc.data_ptr(), st.cuda_stream)
                            st.synchronize()
                            scan_ms, merge_ms, calls = idx.profile_read()
                            idx.profile(False)
                            k = f"rounds{r}"
                            rec[k] = min(rec.get(k, 9e9), round(scan_ms / max(calls, 1), 4))
                    os.environ.pop("NVDB_B200_ROWS_ROUNDS")
                    print(json.dumps(rec), flush=True)


if __name__ == "__main__":
    for w in sys.argv[1:] or ["small"]:
        globals()[w]()
