#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 flat-scan path (contract: see DESIGN.md "Measurement").

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA kernels through the C ABI)
  python bench.py --impl reference --gpus N ...            the reference's own CPU flat scan on the host cores

Workload (BASELINE.json configs[1]): FP16 exact flat-scan top-10 over a 2,903,054 x 768 synthetic base,
batch 1. A "step" is one query batch scanned over the whole base. metric = the reference harness's own
`payload_equiv_bandwidth_GBps` (apps/nvdb_bench.cpp:414-418): payload bytes scanned per second; QPS is
reported beside it. At N > 1 every rank holds its own 2,903,054-row shard of an N-times larger index
(weak scaling); a step = local scan on every rank + NCCL all-gather of the local top-k + merge.
"""
from __future__ import annotations

import argparse
import importlib
import json
import os
import statistics
import sys
import tempfile
import threading
import time
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_ROWS = 2_903_054          # the reference's full set (Performance.md:195)
DIM = 768
TOPK = 10
SEED = 20240229
F32, F16, I8 = 1, 2, 3
ESZ = {F32: 4, F16: 2, I8: 1}
NQ_POOL = 64                # distinct queries cycled through the steps


def payload_bytes(n, dim, dtype):
    """reference include/nvdb/vecbin_format.h:57-59 (bytes_for_payload_and_aux)"""
    return n * dim * ESZ[dtype] + (4 * n if dtype == I8 else 0)


def measured_peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


# ----------------------------------------------------------------------------------------------------------
# clocks / throttle sampling during the timed region (NVML, ~every 25 ms)
# ----------------------------------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown", 0x4: "sw_power_cap",
               0x80: "hw_power_brake_slowdown", 0x2: "applications_clocks_setting", 0x100: "display_clock_setting"}

    def __init__(self, device_index):
        self.samples, self.reason_bits, self.max_mhz, self.err = [], 0, None, None
        self._stop = threading.Event()
        self._thread = None
        self._timed = False
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(device_index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception as e:  # noqa: BLE001
            self.nv, self.err = None, repr(e)

    def _run(self):
        while not self._stop.is_set():
            try:
                mhz = self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM)
                bits = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                self.samples.append((self._timed, mhz))
                if self._timed:
                    self.reason_bits |= int(bits)
            except Exception as e:  # noqa: BLE001
                self.err = repr(e)
                return
            time.sleep(0.025)

    def start(self):
        if self.nv:
            self._thread = threading.Thread(target=self._run, daemon=True)
            self._thread.start()

    def mark_timed(self, on):
        self._timed = on

    def stop(self):
        self._stop.set()
        if self._thread:
            self._thread.join(timeout=2)
        timed = [m for t, m in self.samples if t]
        used = timed or [m for _, m in self.samples]
        reasons = [name for bit, name in self.REASONS.items() if self.reason_bits & bit]
        out = {"sm_mhz": statistics.median(used) if used else None, "sm_max_mhz": self.max_mhz, "reasons": reasons,
               "samples_in_timed_region": len(timed), "samples_total": len(self.samples)}
        if self.err:
            out["error"] = self.err
        return out


# ----------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own OpenMP flat scan on the host cores, bounded sample
# ----------------------------------------------------------------------------------------------------------
def cpu_reference_run(sample_payload: np.ndarray, queries: np.ndarray, steps: int, warmup: int, budget_s: float):
    """Times FlatIndexOMP::search_topk_dot (compiled, unmodified reference in oracle/_ref) over a vecbin file that
    holds `sample_payload` (f16). Falls back to the C port of the oracle when oracle/_ref is absent.
    Returns dict(value GB/s, qps, ms_per_step, steps, kind, cores, sample)."""
    sys.path.insert(0, str(ROOT / "tests"))
    import _oracle as O

    n, d = sample_payload.shape
    nbytes = payload_bytes(n, d, F16)
    cores = os.cpu_count() or 1
    os.environ.setdefault("OMP_NUM_THREADS", str(cores))
    kind, run = "port", None
    tmpdir = None
    if O.REF_SO.exists() or O.REFERENCE_SRC.exists():
        try:
            ref = O.Reference()
            shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None
            tmpdir = tempfile.TemporaryDirectory(dir=shm)
            path = os.path.join(tmpdir.name, "sample_f16.vecbin")
            O.write_vecbin(path, F16, sample_payload)
            h = ref.load(path)
            kind = "reference"

            def run(q):
                return ref.search(h, q, TOPK, "omp")
        except Exception as e:  # noqa: BLE001
            print(f"[bench] oracle/_ref unusable ({e!r}); using the C port", file=sys.stderr)
    if run is None:
        orc = O.Oracle()

        def run(q):
            return orc.search(F16, sample_payload, None, q, TOPK, mt=True)

    for i in range(warmup):
        run(queries[i % len(queries)])
    t0 = time.perf_counter()
    done = 0
    for i in range(steps):
        run(queries[i % len(queries)])
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    # the reference's other multi-threaded variant, FlatIndexPool (pinned persistent workers,
    # src/flat_index_pool.cpp:94-215), on the same sample — reported beside the OMP figure
    pool = None
    if kind == "reference":
        try:
            for i in range(min(warmup, 3)):
                ref.search(h, queries[i % len(queries)], TOPK, "pool", cores)
            p0, pdone = time.perf_counter(), 0
            for i in range(min(steps, 200)):
                ref.search(h, queries[i % len(queries)], TOPK, "pool", cores)
                pdone += 1
                if time.perf_counter() - p0 > min(budget_s, 15.0):
                    break
            pdt = time.perf_counter() - p0
            pool = {"value": nbytes * pdone / pdt / 1e9, "unit": "GB/s", "threads": cores, "steps": pdone,
                    "ms_per_step": pdt / pdone * 1e3, "variant": "FlatIndexPool"}
        except Exception as e:  # noqa: BLE001
            pool = {"unavailable": repr(e)}
    if tmpdir is not None:
        tmpdir.cleanup()
    return {"value": nbytes * done / dt / 1e9, "qps_on_sample": done / dt, "ms_per_step": dt / done * 1e3, "steps": done,
            "kind": kind, "cores": cores, "pool": pool,
            "sample": f"first {n} rows x {d} f16 of the synthetic base ({nbytes / 1e6:.0f} MB), {done} single-query "
                      f"top-{TOPK} scans, FlatIndexOMP, OMP_NUM_THREADS={os.environ['OMP_NUM_THREADS']}"}


def make_cpu_sample(rows: int):
    """First `rows` rows of the synthetic base as f16 + a pool of f32 queries, generated on the CPU by the
    oracle's restatement of the generator (bit-identical to the device generator; tests/test_gpu_scan.py)."""
    sys.path.insert(0, str(ROOT / "tests"))
    import _oracle as O
    orc = O.Oracle()
    f = orc.synth_f32(SEED, 0, rows, DIM)
    payload = orc.to_f16(f)
    qrows = [(i * 104729) % rows for i in range(NQ_POOL)]
    return payload, np.ascontiguousarray(f[qrows])


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    rows = int(os.environ.get("NVDB_BENCH_CPU_ROWS", "262144"))
    payload, queries = make_cpu_sample(rows)
    r = cpu_reference_run(payload, queries, args.steps, args.warmup, budget_s=150.0)
    line = {
        "impl": "reference", "metric": "payload_equiv_bandwidth_GBps", "value": r["value"], "unit": "GB/s",
        "n_gpus": args.gpus, "steps": r["steps"], "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"FP16 exact flat-scan top-{TOPK}, {N_ROWS}x{DIM}, batch 1 (BASELINE.json configs[1])",
                   "note": "reference CPU scan is bandwidth-normalised on a bounded row sample; GB/s is size independent"},
        "qps_equiv_full_base": r["value"] * 1e9 / payload_bytes(N_ROWS, DIM, F16),
        "cpu_baseline": {"value": r["value"], "unit": "GB/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                         "pool": r.get("pool")},
        "e2e": {"value": r["value"], "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line), flush=True)
    return 0


_JSON_FD = None


def _reserve_stdout():
    """stdout carries exactly ONE JSON line: everything else a library prints to fd 1 (e.g. the NCCL version banner,
    which ignores NCCL_DEBUG_FILE) goes to stderr; emit() writes the line to the real stdout."""
    global _JSON_FD
    if _JSON_FD is None:
        sys.stdout.flush()
        _JSON_FD = os.dup(1)
        os.dup2(2, 1)


def emit(line):
    data = (json.dumps(line) + "\n").encode()
    if _JSON_FD is None:
        sys.stdout.write(data.decode())
        sys.stdout.flush()
    else:
        sys.stdout.flush()
        os.write(_JSON_FD, data)


def run_ours(args):
    import torch
    import torch.distributed as dist

    _reserve_stdout()

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; this path has no CPU fallback")
    nv = importlib.import_module("nano-vectordb_b200")
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")   # keep stdout to the one JSON line
        dist.init_process_group("nccl", device_id=dev)
    K, W = args.steps, args.warmup
    n_rows = int(os.environ.get("NVDB_BENCH_ROWS", N_ROWS))
    nbytes = payload_bytes(n_rows, DIM, F16)

    # ---- the shard of this rank, generated on the device; global rows [rank*n, (rank+1)*n)
    idx = nv.FlatIndexCuda.synthetic(F16, n_rows, DIM, SEED, first_row=rank * n_rows, device=local_rank,
                                     row_offset=rank * n_rows)
    total_rows = world * n_rows
    qrows = [(i * 104729 * 31) % total_rows for i in range(NQ_POOL)]   # queries = f32 rows of the (global) base
    stream = torch.cuda.Stream(device=dev)
    sp = stream.cuda_stream
    q_dev = torch.empty((NQ_POOL, DIM), dtype=torch.float32, device=dev)
    with torch.cuda.stream(stream):
        for i, r in enumerate(qrows):
            nv.synth_fill_device(local_rank, F32, 1, DIM, SEED, r, q_dev[i].data_ptr(), None, sp)
    stream.synchronize()
    q_host = q_dev.cpu().numpy()
    out_ids = torch.empty((1, TOPK), dtype=torch.int64, device=dev)
    out_sc = torch.empty((1, TOPK), dtype=torch.float32, device=dev)
    # local scan -> exchange + merge: ONE kernel over NVLink peer memory ("peer"), or all-gather + merge ("nccl")
    sharded = nv.ShardedFlatIndex(idx, dev, exchange=os.environ.get("NVDB_B200_EXCHANGE", "auto")) if world > 1 else None

    def step_device(i):
        if world > 1:
            sharded.search(q_dev[i % NQ_POOL:i % NQ_POOL + 1], TOPK)
        else:
            idx.search_device(q_dev[i % NQ_POOL].data_ptr(), 1, TOPK, out_ids.data_ptr(), out_sc.data_ptr(), sp)

    def sync_all():
        torch.cuda.synchronize(dev)
        if world > 1:
            dist.barrier()
            torch.cuda.synchronize(dev)

    sampler = ClockSampler(local_rank)
    sampler.start()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for i in range(W):
            step_device(i)
        sync_all()
        idx.profile(True)
        l0 = nv.launch_count()
        sampler.mark_timed(True)
        ev0.record(stream)
        for i in range(K):
            step_device(W + i)
        ev1.record(stream)
        sync_all()
        sampler.mark_timed(False)
        l1 = nv.launch_count()
    ms = ev0.elapsed_time(ev1)
    scan_ms, merge_ms, calls = idx.profile_read()
    idx.profile(False)
    if world > 1:
        t = torch.tensor([ms], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t.item())
    value = world * nbytes * K / (ms * 1e-3) / 1e9
    qps = K / (ms * 1e-3)

    # ---- end to end through the host-buffer API (H2D of the query + D2H of the result inside the timed region)
    if world == 1:
        def step_e2e(i):
            return idx.search_topk_dot(q_host[i % NQ_POOL], TOPK)
    else:
        q_pin = torch.from_numpy(q_host).pin_memory()
        r_ids = torch.empty((1, TOPK), dtype=torch.int64).pin_memory()
        r_sc = torch.empty((1, TOPK), dtype=torch.float32).pin_memory()
        q_stage = torch.empty((DIM,), dtype=torch.float32, device=dev)

        def step_e2e(i):
            with torch.cuda.stream(stream):
                q_stage.copy_(q_pin[i % NQ_POOL], non_blocking=True)
                f_ids, f_sc = sharded.search(q_stage.view(1, DIM), TOPK)
                r_ids.copy_(f_ids, non_blocking=True)
                r_sc.copy_(f_sc, non_blocking=True)
            stream.synchronize()
            return r_ids, r_sc
    for i in range(W):
        step_e2e(i)
    sync_all()
    t0 = time.perf_counter()
    for i in range(K):
        step_e2e(W + i)
    torch.cuda.synchronize(dev)
    e2e_s = time.perf_counter() - t0
    if world > 1:
        t = torch.tensor([e2e_s], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e_s = float(t.item())
    e2e_value = world * nbytes * K / e2e_s / 1e9
    clocks = sampler.stop()

    # ---- sanity: the top-1 of a base-row query is the row itself
    ids0, sc0 = (idx.search_topk_dot(q_host[0], TOPK) if world == 1 else (None, None))
    if world == 1:
        assert int(ids0[0]) == qrows[0], f"self-query check failed: {ids0[0]} vs {qrows[0]}"
    else:   # the row lives on ONE shard: every rank must find it through the exchange
        for i in (0, 1, NQ_POOL - 1):
            with torch.cuda.stream(stream):
                g_ids, _ = sharded.search(q_dev[i:i + 1], TOPK)
            stream.synchronize()
            assert int(g_ids[0, 0]) == qrows[i], f"rank {rank}: sharded self-query check failed: {int(g_ids[0, 0])} vs {qrows[i]}"
        assert sharded.exchange_status() == 0, "peer exchange timed out"

    peak, peak_src = measured_peaks()
    kern_ms = scan_ms / max(calls, 1)
    achieved = nbytes / (kern_ms * 1e-3) / 1e9 if kern_ms > 0 else None
    line = {
        "metric": "payload_equiv_bandwidth_GBps", "value": value, "unit": "GB/s", "n_gpus": world, "steps": K,
        "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": "f32", "data": "synthetic",
        "config": {"workload": f"FP16 exact flat-scan top-{TOPK}, {n_rows}x{DIM} per GPU, batch 1 (BASELINE.json configs[1])",
                   "rows_per_gpu": n_rows, "dim": DIM, "k": TOPK, "batch": 1, "base_dtype": "f16",
                   "l2": f"base ({nbytes / 1e9:.2f} GB) is larger than L2 (126 MB); queries cycle through {NQ_POOL} rows",
                   "parallelism": (f"row-shard x{world}, " + ("fused peer-memory exchange + merge kernel (NVLink stores, no NCCL)"
                                                                if sharded.exchange == "peer" else "NCCL all-gather + merge kernel"))
                   if world > 1 else "single GPU"},
        "qps": qps,
        "e2e": {"value": e2e_value, "unit": "GB/s", "qps": K / e2e_s, "ms_per_step": e2e_s / K * 1e3,
                "h2d_bytes_per_step": DIM * 4, "d2h_bytes_per_step": TOPK * 12,
                "api": "nvdb_b200_search (host buffers)" if world == 1 else f"ShardedFlatIndex.search: host query -> H2D -> local scan -> {sharded.exchange} exchange + merge -> D2H"},
        "gpu_launches": int(l1 - l0),
        "roofline": {"bound": "hbm", "kernel": "scan_gemv_tma_kernel<f16,1>", "achieved": achieved, "peak": peak,
                     "unit": "GB/s", "frac": (achieved / peak) if achieved else None, "traffic": None,
                     "peak_source": peak_src, "frac_of_nominal_8TBps": (achieved / 8000.0) if achieved else None,
                     "kernel_ms": kern_ms, "merge_kernel_ms": merge_ms / max(calls, 1),
                     "algorithmic_bytes_per_launch": nbytes},
        "clocks": clocks,
    }
    ncu_traffic = ROOT / "profiles" / "scan_f16_traffic.json"
    if ncu_traffic.exists():
        try:
            line["roofline"]["traffic"] = json.loads(ncu_traffic.read_text()).get("dram_bytes_per_launch")
        except Exception:  # noqa: BLE001
            pass
    if rank == 0 and world == 1 and not args.no_extras:
        line["extras"] = run_extras(nv, torch, idx, dev, local_rank, stream, q_dev, args)
    if rank == 0 and world == 1 and not args.no_cpu:
        rows = int(os.environ.get("NVDB_BENCH_CPU_ROWS", "262144"))
        payload, cq = make_cpu_sample(rows)
        r = cpu_reference_run(payload, cq, steps=400, warmup=3, budget_s=15.0)
        line["cpu_baseline"] = {"value": r["value"], "unit": "GB/s", "cores": r["cores"], "kind": r["kind"],
                                "sample": r["sample"], "qps_equiv_full_base": r["value"] * 1e9 / nbytes,
                                "pool": r.get("pool")}
    if sharded is not None:
        sharded.close()     # unmap the peers' windows while every rank is still alive
    idx.close()
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def _time_search(torch, idx, stream, q_dev, nq, k, iters, warm=3):
    """device-resident timing of `iters` searches of an nq-query batch; returns (ms per batch, scan kernel ms)."""
    dev = q_dev.device
    out_ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
    out_sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
    qs = q_dev
    if qs.shape[0] < nq:
        qs = qs.repeat((nq + qs.shape[0] - 1) // qs.shape[0], 1)[:nq].contiguous()
        qs += 1e-3 * torch.randn_like(qs)   # distinct queries
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with torch.cuda.stream(stream):
        for _ in range(warm):
            idx.search_device(qs.data_ptr(), nq, k, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        idx.profile(True)
        e0.record(stream)
        for _ in range(iters):
            idx.search_device(qs.data_ptr(), nq, k, out_ids.data_ptr(), out_sc.data_ptr(), stream.cuda_stream)
        e1.record(stream)
        stream.synchronize()
    scan_ms, _, calls = idx.profile_read()
    idx.profile(False)
    return e0.elapsed_time(e1) / iters, scan_ms / max(calls, 1)


def measured_tensor_peak():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return float(d["bf16_tflops"]), "measured burst (MEASURED_PEAKS.json bf16_tflops)"
    return 1590.0, "fallback (B200_PROFILING.md 1.59 PFLOP/s)"


def run_extras(nv, torch, idx16, dev, local_rank, stream, q_dev, args):
    """Other BASELINE.json configs on one GPU, short runs (explanatory numbers, not the headline)."""
    peak, _ = measured_peaks()
    tpeak, tsrc = measured_tensor_peak()
    ex = {"tensor_peak_TFLOPs": tpeak, "tensor_peak_source": tsrc}
    n = idx16.count
    nb16 = payload_bytes(n, DIM, F16)
    qbig = q_dev.repeat(1024 // NQ_POOL, 1).contiguous()
    qbig += 1e-3 * torch.randn_like(qbig)

    def sweep(idx, nbytes, plan, limbs_of):
        out = {}
        for pname, path, b in plan:
            idx.set_path(path)
            ms, kms = _time_search(torch, idx, stream, qbig, b, TOPK, iters=max(5, min(50, 2048 // b)))
            rec = {"path": pname, "ms_per_batch": ms, "qps": b / ms * 1e3, "scan_kernel_ms": kms}
            if pname == "mma_exact":
                nqx, nfb = idx.exact_stats(reset=True)
                rec["certificate_fallback_rate"] = nfb / max(nqx, 1)
            if pname == "gemv":
                rec["base_GBps"] = nbytes * ((b + 7) // 8) / (kms * 1e-3) / 1e9
            else:
                rec["base_GBps"] = nbytes / (kms * 1e-3) / 1e9
                issued = 2.0 * ((b + 127) // 128 * 128) * n * DIM * limbs_of(pname) / (kms * 1e-3) / 1e12
                rec["issued_TFLOPs"] = issued
                rec["useful_TFLOPs"] = 2.0 * b * n * DIM / (kms * 1e-3) / 1e12
                rec["tensor_frac_of_measured_bf16_peak"] = issued / tpeak
            out[f"{pname}_b{b}"] = rec
        idx.set_path(nv.PATH_AUTO)
        return out

    # Order matters on a power-capped part: the tensor-core sweeps pull the SM clock down to ~1.3 GHz for a while, and the
    # INT8 streaming scan (41 instructions per 16 bytes) is clock-sensitive. The bandwidth-bound configurations therefore
    # run first, each tensor sweep after them (profiles/README.md, "clock sensitivity").
    # FP32 500K x 768 — BASELINE.json configs[0]
    with nv.FlatIndexCuda.synthetic(F32, 500_000, DIM, SEED, device=local_rank) as idx32:
        nb32 = payload_bytes(500_000, DIM, F32)
        ms, kms = _time_search(torch, idx32, stream, q_dev, 1, TOPK, iters=50)
        ex["fp32_500k_batch1"] = {"ms": ms, "qps": 1e3 / ms, "kernel_ms": kms, "GBps": nb32 / (kms * 1e-3) / 1e9,
                                  "roofline_frac": nb32 / (kms * 1e-3) / 1e9 / peak}
    # refine: 1M x 768 f16, Q=10000, R=256, K=10 — BASELINE.json configs[4] (synthetic candidate lists)
    nref, Q, R, Kr = 1_000_000, 10_000, 256, 10
    with nv.FlatIndexCuda.synthetic(F16, nref, DIM, SEED, device=local_rank) as idxr:
        g = torch.Generator(device=dev)
        g.manual_seed(1)
        cand = torch.randint(0, nref, (Q, R), dtype=torch.int32, device=dev, generator=g)
        qs = q_dev.repeat((Q + NQ_POOL - 1) // NQ_POOL, 1)[:Q].contiguous()
        oid = torch.empty((Q, Kr), dtype=torch.int32, device=dev)
        od = torch.empty((Q, Kr), dtype=torch.float32, device=dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        with torch.cuda.stream(stream):
            for _ in range(3):
                idxr.refine_l2_device(qs.data_ptr(), cand.data_ptr(), Q, R, Kr, oid.data_ptr(), od.data_ptr(), stream.cuda_stream)
            e0.record(stream)
            for _ in range(10):
                idxr.refine_l2_device(qs.data_ptr(), cand.data_ptr(), Q, R, Kr, oid.data_ptr(), od.data_ptr(), stream.cuda_stream)
            e1.record(stream)
            stream.synchronize()
        ms = e0.elapsed_time(e1) / 10
        gb = (Q * R * DIM * 2 + Q * DIM * 4 + Q * R * 4 + Q * Kr * 8) / 1e9
        ex["refine_f16_1M_Q10000_R256_K10"] = {"kernel_ms": ms, "us_per_query": ms * 1e3 / Q, "gather_GBps": gb / (ms * 1e-3),
                                               "roofline_frac": gb / (ms * 1e-3) / peak}
        # the stage in front of it: IVF-PQ candidates from this repo's own index, refine on the
        # device; reference grid point nlist=4096, m=64, nprobe=64, REFINE_K=50 (Performance.md:673-687). The synthetic
        # base has no cluster structure, so only build time, latency and throughput are reported here — recall on
        # structured data is in profiles/r1_ivfpq_1M_*.jsonl (scripts/bench_ivfpq.py).
        if not args.no_ivfpq:
            import time as _time
            t0 = _time.perf_counter()
            ivf = nv.IvfPqIndex(idxr, nlist=4096, pq_m=64, ntrain=200_000, niter=10, seed=SEED)
            build_s = _time.perf_counter() - t0
            qh = qs[:1000].cpu().numpy()
            for _ in range(3):
                ivf.search_refine(qh[:1], 64, 50, Kr)
                ivf.search_refine(qh, 64, 50, Kr)
            t0 = _time.perf_counter()
            for i in range(200):
                ivf.search_refine(qh[i:i + 1], 64, 50, Kr)
            lat_ms = (_time.perf_counter() - t0) * 1e3 / 200
            t0 = _time.perf_counter()
            for _ in range(5):
                ivf.search_refine(qh, 64, 50, Kr)
            b_ms = (_time.perf_counter() - t0) * 1e3 / 5
            ex["ivfpq_1M_nlist4096_m64_nprobe64_R50_K10"] = {
                "build_s_niter10": build_s, "index_MB": ivf.index_bytes / 1e6, "e2e_latency_ms_per_query": lat_ms,
                "e2e_batch1000_ms": b_ms, "e2e_batch_qps": 1000 / b_ms * 1e3,
                "note": "host buffers in/out; candidates never leave the device; parity unpinned (no FAISS)"}
            ivf.close()
    # bandwidth-bound streaming scans of the other dtypes / k before any tensor sweep
    idx16.set_path(nv.PATH_GEMV)
    ms, kms = _time_search(torch, idx16, stream, q_dev, 1, 100, iters=20)
    ex["fp16_batch1_top100"] = {"ms": ms, "qps": 1e3 / ms}
    idx16.set_path(nv.PATH_AUTO)
    nb8 = payload_bytes(n, DIM, I8)
    # INT8 (+scale) — BASELINE.json configs[2]: batch 1 (bit-exact streaming scan)
    with nv.FlatIndexCuda.synthetic(I8, n, DIM, SEED, device=local_rank) as idx8:
        ex["int8_batch_sweep"] = sweep(idx8, nb8, [("gemv", nv.PATH_GEMV, 1), ("gemv", nv.PATH_GEMV, 2)], lambda p_: 2)
        r = ex["int8_batch_sweep"]["gemv_b1"]
        r["roofline_frac"] = r["base_GBps"] / peak
        # ... and batch 8 / 128 / 1024 on the tensor cores (kind::i8, two exact limbs; exact path = + re-score)
        time.sleep(0.5)
        plan8t = [("mma", nv.PATH_MMA, 8), ("mma", nv.PATH_MMA, 128), ("mma", nv.PATH_MMA, 1024),
                  ("mma_exact", nv.PATH_MMA_EXACT, 128), ("mma_exact", nv.PATH_MMA_EXACT, 1024)]
        ex["int8_batch_sweep"].update(sweep(idx8, nb8, plan8t, lambda p_: 2))
    time.sleep(0.5)
    # FP16 batch sweep (BASELINE.json configs[1]): GEMV scan -> tcgen05 GEMM crossover
    plan16 = [("gemv", nv.PATH_GEMV, 1), ("gemv", nv.PATH_GEMV, 2), ("gemv", nv.PATH_GEMV, 8),
              ("mma", nv.PATH_MMA, 2), ("mma", nv.PATH_MMA, 8), ("mma", nv.PATH_MMA, 64), ("mma", nv.PATH_MMA, 128),
              ("mma", nv.PATH_MMA, 256), ("mma", nv.PATH_MMA, 1024),
              ("mma_fast", nv.PATH_MMA_FAST, 128), ("mma_fast", nv.PATH_MMA_FAST, 1024),
              ("mma_exact", nv.PATH_MMA_EXACT, 8), ("mma_exact", nv.PATH_MMA_EXACT, 128),
              ("mma_exact", nv.PATH_MMA_EXACT, 1024)]
    ex["fp16_batch_sweep"] = sweep(idx16, nb16, plan16, lambda p_: 2 if p_ == "mma" else 1)
    # second distribution of SURVEY.md 8(d): rows around 256 Gaussian centres (sigma 0.3, re-normalised). A query's own
    # cluster packs ~11 K scores close together, which is what the exact path's certificate has to separate: the rates
    # say how many queries needed the second tier / the streaming scan (scripts/bench_clustered.py is the full sweep).
    try:
        ex["fp16_clustered_exact_b1024"] = clustered_extra(nv, torch, dev, local_rank, stream, n)
    except Exception as e:  # an explanatory extra must never cost the headline line
        ex["fp16_clustered_exact_b1024"] = {"error": repr(e)[:200]}
    return ex


def clustered_extra(nv, torch, dev, local_rank, stream, n):
    gen = torch.Generator(device=dev).manual_seed(SEED)
    centres = torch.randn((256, DIM), device=dev, generator=gen)
    centres /= centres.norm(dim=1, keepdim=True)
    payload = torch.empty((n, DIM), dtype=torch.float16, device=dev)
    qrows = torch.arange(1024, device=dev) * 2833 % n
    queries = torch.empty((1024, DIM), dtype=torch.float32, device=dev)
    for lo in range(0, n, 131072):
        hi = min(n, lo + 131072)
        owner = torch.randint(0, 256, (hi - lo,), device=dev, generator=gen)
        x = centres[owner] + (0.3 / DIM ** 0.5) * torch.randn((hi - lo, DIM), device=dev, generator=gen)
        x /= x.norm(dim=1, keepdim=True)
        sel = (qrows >= lo) & (qrows < hi)
        queries[sel] = x[qrows[sel] - lo]
        payload[lo:hi] = x.half()
    torch.cuda.synchronize(dev)
    out = {}
    with nv.FlatIndexCuda.from_device(payload.data_ptr(), None, F16, n, DIM, device=local_rank, keepalive=payload) as idx:
        idx.set_path(nv.PATH_MMA_EXACT)
        for k in (10, 100):
            ms, _ = _time_search(torch, idx, stream, queries, 1024, k, iters=5)
            nq_, nf = idx.exact_stats()
            t2, streamed = idx.exact_tiers()
            idx.exact_stats(reset=True)
            out[f"k{k}"] = {"ms_per_batch": ms, "qps": 1024 / ms * 1e3, "first_certificate_failures": nf / max(nq_, 1),
                            "second_tier": t2 / max(nq_, 1), "streaming_scan": streamed / max(nq_, 1)}
    return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=1000)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="skip the other configs (int8 / fp32 / batch sweep / refine)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-ivfpq", action="store_true", help="skip the IVF-PQ extra (index build ~2 s)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
