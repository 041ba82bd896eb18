#!/usr/bin/env python
"""bench.py — headline benchmark of the B200 flat-scan path (contract: DESIGN.md §5 "Measurement").

  python bench.py --gpus N --steps K --warmup W            our arm (CUDA kernels through the C ABI)
  python bench.py --impl reference --gpus N ...            the reference's own CPU flat scan on the host cores

metric = the reference harness's own `payload_equiv_bandwidth_GBps` (apps/nvdb_bench.cpp:414-418): payload bytes of
the index scanned per second; QPS is reported beside it. A "step" is one single-query batch scanned over the whole
index.

  N = 1   BASELINE.json configs[1]: FP16 exact flat-scan top-10 over the 2,903,054 x 768 synthetic base, batch 1.
          The other headline configurations (batch 1024, INT8, FP32 500 K, refine, small batches) are measured in the
          same run and reported INSIDE `roofline.configs` / `e2e.configs`; the fixed 100 M-row index of configs[3]
          on one GPU is `strong_scaling_n1` (what the N > 1 lines are a speed-up over).
  N > 1   BASELINE.json configs[3], STRONG scaling: FP16 exact flat-scan top-100 over ONE fixed 100,000,000 x 768
          index row-sharded over the N GPUs (rank g holds rows [g*ceil(N/G), ...)), batch 1: local scan + exchange
          over NVLink peer memory + merge, one kernel chain per rank. `configs` carries batch 1024 on the same index,
          the hard case (the 2.9 M base split N ways, batch 1) and round 1's weak-scaling line.

Timing: W untimed warm-up steps, then blocks of EXACTLY K steps, each block bracketed by a barrier +
torch.cuda.synchronize() on both sides and timed with CUDA events on the launching stream, max over ranks; blocks are
repeated until >= 0.6 s have been timed (so that the clock sampler sees the timed region) and the MEDIAN block is
reported: ms_per_step * steps is one block.
"""
from __future__ import annotations

import argparse
import ctypes
import importlib
import json
import os
import statistics
import sys
import tempfile
import threading
import time
from pathlib import Path

# Host cores this process may use, read ONCE before any OpenMP runtime is loaded: with OMP_PROC_BIND set, libgomp pins
# the initial thread to the first place when it loads, after which sched_getaffinity() reports a single core.
HOST_AFFINITY = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
HOST_CORES = len(HOST_AFFINITY) if HOST_AFFINITY else (os.cpu_count() or 1)
if "--impl" in sys.argv and "reference" in sys.argv:
    # torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm uses all host cores. Set before anything
    # loads an OpenMP runtime (force_omp_threads() below repeats it through omp_set_num_threads).
    os.environ.update({"OMP_NUM_THREADS": str(HOST_CORES), "OMP_PROC_BIND": "close", "OMP_PLACES": "cores"})

import numpy as np  # noqa: E402

ROOT = Path(__file__).resolve().parent
sys.path.insert(0, str(ROOT))

N_ROWS = 2_903_054          # the reference's full set (Performance.md:195)
K4_ROWS = 100_000_000       # BASELINE.json configs[3]
DIM = 768
TOPK = 10
K4_TOPK = 100
SEED = 20240229
F32, F16, I8 = 1, 2, 3
ESZ = {F32: 4, F16: 2, I8: 1}
NQ_POOL = 64                # distinct queries cycled through the steps
MIN_TIMED_S = 0.6           # blocks of K steps are repeated until this much has been timed


def payload_bytes(n, dim, dtype):
    """reference include/nvdb/vecbin_format.h:57-59 (bytes_for_payload_and_aux)"""
    return n * dim * ESZ[dtype] + (4 * n if dtype == I8 else 0)


def workload_name(world, total_rows=None):
    """The one string both arms print as config.workload (the driver's same_config check)."""
    if world == 1:
        return f"FP16 exact flat-scan top-{TOPK}, {N_ROWS}x{DIM}, batch 1 (BASELINE.json configs[1])"
    return (f"FP16 exact flat-scan top-{K4_TOPK}, {total_rows or K4_ROWS}x{DIM} fixed index row-sharded over {world} GPUs, "
            f"batch 1 (BASELINE.json configs[3], strong scaling)")


def peaks():
    p = ROOT / "MEASURED_PEAKS.json"
    if p.exists():
        d = json.loads(p.read_text())
        return {"hbm": float(d["hbm_gbs"]), "hbm_src": "measured (MEASURED_PEAKS.json hbm_gbs)",
                "tf": float(d.get("bf16_tflops_sustained", d["bf16_tflops"])),
                "tf_src": "measured sustained (MEASURED_PEAKS.json bf16_tflops_sustained)",
                "tf_burst": float(d["bf16_tflops"])}
    return {"hbm": 6650.0, "hbm_src": "fallback (B200_PROFILING.md 6.65 TB/s)", "tf": 1590.0,
            "tf_src": "fallback (B200_PROFILING.md 1.59 PFLOP/s)", "tf_burst": 1590.0}


# ----------------------------------------------------------------------------------------------------------
# clocks / throttle sampling during the timed region (NVML, every 10 ms)
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
            time.sleep(0.010)

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
               "samples_in_timed_region": len(timed), "samples_total": len(self.samples),
               "sm_mhz_min_timed": min(timed) if timed else None}
        if self.err:
            out["error"] = self.err
        return out


# ----------------------------------------------------------------------------------------------------------
# reference arm / cpu_baseline: the reference's own OpenMP flat scan on the host cores
# ----------------------------------------------------------------------------------------------------------
def host_cores():
    return HOST_CORES


def force_omp_threads():
    """torchrun exports OMP_NUM_THREADS=1 to every rank; the reference arm must use all host cores. The variables are
    (re)written before libgomp is loaded and, because another library may have loaded it already, the thread count is
    also set through omp_set_num_threads. Binding as the reference's REPRO.md:129-131."""
    cores = host_cores()
    os.environ["OMP_NUM_THREADS"] = str(cores)
    os.environ["OMP_PROC_BIND"] = "close"
    os.environ["OMP_PLACES"] = "cores"
    try:
        gomp = ctypes.CDLL("libgomp.so.1")
        gomp.omp_set_num_threads(cores)
        got = int(gomp.omp_get_max_threads())
    except OSError:
        got = cores
    return cores, got


def _oracle_mod():
    sys.path.insert(0, str(ROOT / "tests"))
    import _oracle as O
    return O


def make_cpu_base_file(tmpdir, rows, payload=None):
    """vecbin64 f16 file with the first `rows` rows of the synthetic base + a pool of f32 queries (rows of the base).
    payload: the same rows read back from the device-generated index (our arm); otherwise they are generated on the CPU
    by the oracle's restatement of the generator (bit-identical: tests/test_gpu_scan.py) in 128 K-row chunks."""
    O = _oracle_mod()
    import struct
    orc = O.Oracle()
    path = os.path.join(tmpdir, "base_f16.vecbin")
    qrows = [(i * 104729) % rows for i in range(NQ_POOL)]
    queries = np.stack([orc.synth_f32(SEED, r, 1, DIM)[0] for r in qrows])
    with open(path, "wb") as f:
        hdr = struct.pack("<QIIIIQ", O.VEC_MAGIC, 1, F16, DIM, 0, rows)
        f.write(hdr + b"\0" * (64 - len(hdr)))
        if payload is not None:
            f.write(memoryview(np.ascontiguousarray(payload)).cast("B"))
        else:
            step = 131072
            for r0 in range(0, rows, step):
                n = min(step, rows - r0)
                x = orc.synth_f32(SEED, r0, n, DIM)
                # IEEE round-to-nearest-even == tools/nvdb_convert_f16.cpp (SURVEY.md Appendix B probe; tests/test_oracle.py)
                f.write(memoryview(x.astype(np.float16)).cast("B"))
    return path, np.ascontiguousarray(queries, np.float32)


def cpu_reference_run(path, rows, queries, k, steps, warmup, budget_s, with_pool=True):
    """Times FlatIndexOMP::search_topk_dot (compiled, unmodified reference in oracle/_ref; the C port of the oracle only
    when _ref is absent) over the vecbin file. Returns dict(value GB/s, ms_per_step, steps, kind, cores, ...)."""
    O = _oracle_mod()
    cores, omp_threads = force_omp_threads()
    nbytes = payload_bytes(rows, DIM, F16)
    kind, run, ref, h = "port", None, None, None
    if O.REF_SO.exists() or O.REFERENCE_SRC.exists():
        try:
            ref = O.Reference()
            h = ref.load(path)
            kind = "reference"

            def run(q):
                return ref.search(h, q, k, "omp")
        except Exception as e:  # noqa: BLE001
            print(f"[bench] oracle/_ref unusable ({e!r}); using the C port", file=sys.stderr)
    if run is None:
        orc = O.Oracle()
        _, payload, _ = O.read_vecbin(path)

        def run(q):
            return orc.search(F16, payload, None, q, k, mt=True)

    for i in range(warmup):
        run(queries[i % len(queries)])
    t0 = time.perf_counter()
    done = 0
    ids0 = None
    for i in range(steps):
        r = run(queries[i % len(queries)])
        if i == 0:
            ids0 = int(r[0][0])
        done += 1
        if time.perf_counter() - t0 > budget_s:
            break
    dt = time.perf_counter() - t0
    # the reference's other multi-threaded variant, FlatIndexPool (pinned persistent workers,
    # src/flat_index_pool.cpp:94-215), on the same file — reported beside the OMP figure
    pool = None
    if kind == "reference" and with_pool:
        try:
            if HOST_AFFINITY:   # the pool's std::threads inherit the caller's mask, which OMP_PROC_BIND narrowed to one core
                os.sched_setaffinity(0, HOST_AFFINITY)
            for i in range(2):
                ref.search(h, queries[i % len(queries)], k, "pool", cores)
            p0, pdone = time.perf_counter(), 0
            for i in range(min(steps, 100)):
                ref.search(h, queries[i % len(queries)], k, "pool", cores)
                pdone += 1
                if time.perf_counter() - p0 > min(budget_s, 8.0):
                    break
            pdt = time.perf_counter() - p0
            pool = {"value": nbytes * pdone / pdt / 1e9, "unit": "GB/s", "threads": cores, "steps": pdone,
                    "ms_per_step": pdt / pdone * 1e3, "variant": "FlatIndexPool"}
        except Exception as e:  # noqa: BLE001
            pool = {"unavailable": repr(e)}
    if ref is not None and h is not None:
        ref.free(h)
    return {"value": nbytes * done / dt / 1e9, "qps": done / dt, "ms_per_step": dt / done * 1e3, "steps": done,
            "kind": kind, "cores": cores, "omp_threads": omp_threads, "pool": pool, "top1_of_query0": ids0,
            "sample": f"{done} single-query top-{k} scans of {rows} rows x {DIM} f16 ({nbytes / 1e6:.0f} MB) of the synthetic base, "
                      f"FlatIndexOMP, OMP_NUM_THREADS={omp_threads} OMP_PROC_BIND=close OMP_PLACES=cores"}


def shm_dir():
    return "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else None


def run_reference(args):
    """The reference arm: the unmodified reference's FlatIndexOMP on all host cores, on our arm's config. N = 1: the FULL
    2,903,054-row base, top-10. N > 1 (rank 0 only): a bounded sample of the fixed 100 M-row index — its first 2,903,054
    rows — top-100; GB/s is what the driver compares."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return 0
    force_omp_threads()
    world = args.gpus
    rows = int(os.environ.get("NVDB_BENCH_CPU_ROWS", N_ROWS))
    k = TOPK if world == 1 else K4_TOPK
    t0 = time.perf_counter()
    with tempfile.TemporaryDirectory(dir=shm_dir()) as tmp:
        path, queries = make_cpu_base_file(tmp, rows)
        gen_s = time.perf_counter() - t0
        r = cpu_reference_run(path, rows, queries, k, args.steps, args.warmup, budget_s=120.0)
    total_rows = int(os.environ.get("NVDB_BENCH_K4_ROWS", K4_ROWS))
    line = {
        "impl": "reference", "metric": "payload_equiv_bandwidth_GBps", "value": r["value"], "unit": "GB/s",
        "n_gpus": world, "steps": r["steps"], "warmup": args.warmup, "ms_per_step": r["ms_per_step"],
        "higher_is_better": True, "scaling": "weak" if world == 1 else "strong", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(world, total_rows), "rows_scanned_per_step": rows, "dim": DIM, "k": k, "batch": 1,
                   "base_dtype": "f16",
                   "note": ("the full base" if world == 1 and rows == N_ROWS else
                            f"bounded sample: the first {rows} rows of the index; the CPU scan is a stream, GB/s is what is compared"),
                   "base_generation_s": gen_s},
        "qps_on_scanned_rows": r["qps"],
        "cpu_baseline": {"value": r["value"], "unit": "GB/s", "cores": r["cores"], "kind": r["kind"], "sample": r["sample"],
                         "omp_threads": r["omp_threads"], "pool": r.get("pool")},
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


# ----------------------------------------------------------------------------------------------------------
# timing helpers
# ----------------------------------------------------------------------------------------------------------
class Ctx:
    """what every measurement needs: torch, the package, device, stream, process-group facts"""

    def __init__(self, torch, dist, nv, dev, local_rank, rank, world, stream):
        self.torch, self.dist, self.nv, self.dev, self.local_rank = torch, dist, nv, dev, local_rank
        self.rank, self.world, self.stream = rank, world, stream

    def sync_all(self):
        self.torch.cuda.synchronize(self.dev)
        if self.world > 1:
            self.dist.barrier()
            self.torch.cuda.synchronize(self.dev)

    def max_over_ranks(self, x):
        if self.world == 1:
            return float(x)
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item())

    def bcast_int(self, x):
        if self.world == 1:
            return int(x)
        t = self.torch.tensor([int(x)], dtype=self.torch.int64, device=self.dev)
        self.dist.broadcast(t, src=0)
        return int(t.item())


def timed_blocks(cx, step, K, W, min_total_s=MIN_TIMED_S, max_blocks=400, wall=False, on_timed=None):
    """W warm-up steps, then blocks of exactly K steps (see the module docstring). step(i) enqueues step i on cx.stream
    (wall=False: CUDA events) or runs it synchronously end to end (wall=True: perf_counter). Returns the per-block
    milliseconds (max over ranks)."""
    torch = cx.torch
    with torch.cuda.stream(cx.stream):
        for i in range(W):
            step(i)
        cx.sync_all()
        if on_timed:
            on_timed(True)
        blocks, n, i0 = [], 1, W
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        b = 0
        while b < n:
            cx.sync_all()
            t0 = time.perf_counter()
            e0.record(cx.stream)
            for i in range(K):
                step(i0 + i)
            e1.record(cx.stream)
            cx.sync_all()
            ms = (time.perf_counter() - t0) * 1e3 if wall else e0.elapsed_time(e1)
            ms = cx.max_over_ranks(ms)
            blocks.append(ms)
            i0 += K
            if b == 0:  # every rank runs the same number of blocks: rank 0 decides from the first one
                want = int(min_total_s * 1e3 / max(ms, 1e-3)) + 1
                n = cx.bcast_int(max(3, min(max_blocks, want | 1)))
            b += 1
        if on_timed:
            on_timed(False)
    return blocks


def time_search(cx, idx, q_dev, nq, k, iters, warm=3):
    """device-resident timing of `iters` searches of an nq-query batch through nvdb_b200_search_device; returns
    (ms per batch, scan-kernel ms per batch, launches per batch)."""
    torch, nv = cx.torch, cx.nv
    out_ids = torch.empty((nq, k), dtype=torch.int64, device=cx.dev)
    out_sc = torch.empty((nq, k), dtype=torch.float32, device=cx.dev)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    sp = cx.stream.cuda_stream
    with torch.cuda.stream(cx.stream):
        for _ in range(warm):
            idx.search_device(q_dev.data_ptr(), nq, k, out_ids.data_ptr(), out_sc.data_ptr(), sp)
        cx.stream.synchronize()
        idx.profile(True)
        l0 = nv.launch_count()
        e0.record(cx.stream)
        for _ in range(iters):
            idx.search_device(q_dev.data_ptr(), nq, k, out_ids.data_ptr(), out_sc.data_ptr(), sp)
        e1.record(cx.stream)
        cx.stream.synchronize()
        l1 = nv.launch_count()
    scan_ms, _, calls = idx.profile_read()
    idx.profile(False)
    return e0.elapsed_time(e1) / iters, scan_ms / max(calls, 1), (l1 - l0) / iters


def time_search_host(idx, q_host, k, iters, warm=2):
    """end to end through nvdb_b200_search with HOST buffers (H2D of the queries + D2H of ids/scores inside): wall ms"""
    for _ in range(warm):
        idx.search_topk_dot_batch(q_host, k)
    t0 = time.perf_counter()
    for _ in range(iters):
        idx.search_topk_dot_batch(q_host, k)
    return (time.perf_counter() - t0) * 1e3 / iters


def make_queries(cx, rows, total_rows):
    """f32 rows of the (global) synthetic base as queries, generated on the device"""
    torch, nv = cx.torch, cx.nv
    q = torch.empty((len(rows), DIM), dtype=torch.float32, device=cx.dev)
    sp = cx.stream.cuda_stream
    with torch.cuda.stream(cx.stream):
        for i, r in enumerate(rows):
            nv.synth_fill_device(cx.local_rank, F32, 1, DIM, SEED, r % total_rows, q[i].data_ptr(), None, sp)
    cx.stream.synchronize()
    return q


# ----------------------------------------------------------------------------------------------------------
# N = 1: the other headline configurations, each as a roofline record + its end-to-end twin
# ----------------------------------------------------------------------------------------------------------
def roofline_configs(cx, idx16, q_dev, args):
    """-> (roofline.configs, e2e.configs). Order matters on a power-capped part: tensor-core sweeps pull the SM clock to
    ~1.3 GHz for a while and the INT8 streaming scan is clock-sensitive, so bandwidth-bound configurations run first."""
    torch, nv = cx.torch, cx.nv
    pk = peaks()
    rc, ec = {}, {}
    n = idx16.count
    q1024 = q_dev.repeat(1024 // NQ_POOL, 1).contiguous()
    q1024 += 1e-3 * torch.randn_like(q1024)       # 1024 distinct queries
    q1024_host = q1024.cpu().numpy()

    def hbm_rec(name, idx, dtype, nrows, b, k, path, kernel, iters, qd=None, qh=None):
        idx.set_path(path)
        qd = q1024 if qd is None else qd
        qh = q1024_host if qh is None else qh
        ms, kms, launches = time_search(cx, idx, qd, b, k, iters)
        nbytes = payload_bytes(nrows, DIM, dtype)
        rec = {"bound": "hbm", "kernel": kernel, "kernel_ms": kms, "ms_per_batch": ms, "batch": b, "k": k,
               "algorithmic_bytes": nbytes, "achieved": nbytes / (kms * 1e-3) / 1e9, "peak": pk["hbm"], "unit": "GB/s",
               "frac": nbytes / (kms * 1e-3) / 1e9 / pk["hbm"], "frac_whole_step": nbytes / (ms * 1e-3) / 1e9 / pk["hbm"],
               "qps": b / ms * 1e3, "gpu_launches_per_batch": launches}
        if path == nv.PATH_MMA_EXACT:
            nqx, nfb = idx.exact_stats(reset=True)
            rec["certificate_fallback_rate"] = nfb / max(nqx, 1)
        rc[name] = rec
        e_ms = time_search_host(idx, qh[:b], k, max(3, iters // 2))
        ec[name] = {"value": nbytes / (e_ms * 1e-3) / 1e9, "unit": "GB/s", "ms_per_batch": e_ms, "qps": b / e_ms * 1e3,
                    "h2d_bytes_per_step": b * DIM * 4, "d2h_bytes_per_step": b * k * 12, "api": "nvdb_b200_search (host buffers)"}
        idx.set_path(nv.PATH_AUTO)

    def tensor_rec(name, idx, dtype, b, k, kernel, iters):
        idx.set_path(nv.PATH_MMA_EXACT)
        ms, kms, launches = time_search(cx, idx, q1024, b, k, iters)
        nqx, nfb = idx.exact_stats(reset=True)
        flops = 2.0 * b * n * DIM
        peak = pk["tf"] * (2.0 if dtype == I8 else 1.0)
        rc[name] = {"bound": "tensor", "kernel": kernel, "kernel_ms": kms, "ms_per_batch": ms, "batch": b, "k": k,
                    "algorithmic_flops": flops, "achieved": flops / (kms * 1e-3) / 1e12, "peak": peak, "unit": "TFLOP/s",
                    "peak_source": pk["tf_src"] + (" x 2 (kind::i8 issues twice the dense f16 rate; no measured INT8 peak in "
                                                  "MEASURED_PEAKS.json)" if dtype == I8 else ""),
                    "frac": flops / (kms * 1e-3) / 1e12 / peak, "frac_whole_step": flops / (ms * 1e-3) / 1e12 / peak,
                    "issued_limbs": 2 if dtype == I8 else 1, "qps": b / ms * 1e3, "certificate_fallback_rate": nfb / max(nqx, 1),
                    "gpu_launches_per_batch": launches,
                    "hbm_floor_ms": payload_bytes(n, DIM, dtype) / pk["hbm"] / 1e6}
        e_ms = time_search_host(idx, q1024_host[:b], k, 3)
        ec[name] = {"value": b / e_ms * 1e3, "unit": "QPS", "ms_per_batch": e_ms, "h2d_bytes_per_step": b * DIM * 4,
                    "d2h_bytes_per_step": b * k * 12, "api": "nvdb_b200_search (host buffers)"}
        idx.set_path(nv.PATH_AUTO)

    # K1: FP32 500K x 768, batch 1 (BASELINE.json configs[0])
    with nv.FlatIndexCuda.synthetic(F32, 500_000, DIM, SEED, device=cx.local_rank) as idx32:
        q32 = make_queries(cx, [(i * 7919) % 500_000 for i in range(NQ_POOL)], 500_000)
        hbm_rec("fp32_500k_b1", idx32, F32, 500_000, 1, TOPK, nv.PATH_GEMV, "scan_gemv_tma_kernel<f32,1>", 100,
                qd=q32, qh=q32.cpu().numpy())
    # K5: refine, 1M x 768 f16, Q=10000, R=256, K=10 (BASELINE.json configs[4]; synthetic candidate lists)
    nref, Q, R, Kr = 1_000_000, 10_000, 256, 10
    with nv.FlatIndexCuda.synthetic(F16, nref, DIM, SEED, device=cx.local_rank) as idxr:
        g = torch.Generator(device=cx.dev)
        g.manual_seed(1)
        cand = torch.randint(0, nref, (Q, R), dtype=torch.int32, device=cx.dev, generator=g)
        qs = q_dev.repeat((Q + NQ_POOL - 1) // NQ_POOL, 1)[:Q].contiguous()
        oid = torch.empty((Q, Kr), dtype=torch.int32, device=cx.dev)
        od = torch.empty((Q, Kr), dtype=torch.float32, device=cx.dev)
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        sp = cx.stream.cuda_stream
        with torch.cuda.stream(cx.stream):
            for _ in range(3):
                idxr.refine_l2_device(qs.data_ptr(), cand.data_ptr(), Q, R, Kr, oid.data_ptr(), od.data_ptr(), sp)
            e0.record(cx.stream)
            for _ in range(20):
                idxr.refine_l2_device(qs.data_ptr(), cand.data_ptr(), Q, R, Kr, oid.data_ptr(), od.data_ptr(), sp)
            e1.record(cx.stream)
            cx.stream.synchronize()
        ms = e0.elapsed_time(e1) / 20
        nb = Q * R * DIM * 2 + Q * DIM * 4 + Q * R * 4 + Q * Kr * 8
        rc["refine_k5"] = {"bound": "hbm", "kernel": "refine_l2_kernel<f16>", "kernel_ms": ms, "algorithmic_bytes": nb,
                           "achieved": nb / (ms * 1e-3) / 1e9, "peak": pk["hbm"], "unit": "GB/s",
                           "frac": nb / (ms * 1e-3) / 1e9 / pk["hbm"], "us_per_query": ms * 1e3 / Q,
                           "shape": f"{nref}x{DIM} f16, Q={Q}, R={R}, K={Kr}"}
        qh, ch = qs.cpu().numpy(), cand.cpu().numpy().view(np.uint32)
        idxr.refine_l2(qh, ch, Kr)
        t0 = time.perf_counter()
        for _ in range(3):
            idxr.refine_l2(qh, ch, Kr)
        e_ms = (time.perf_counter() - t0) * 1e3 / 3
        ec["refine_k5"] = {"value": Q / e_ms * 1e3, "unit": "queries/s", "ms_per_call": e_ms,
                           "h2d_bytes_per_step": Q * DIM * 4 + Q * R * 4, "d2h_bytes_per_step": Q * Kr * 8,
                           "api": "nvdb_b200_refine_l2 (host buffers)"}
        if not args.no_ivfpq:
            try:
                rc["ivfpq_k5"] = ivfpq_extra(cx, idxr, qs[:1000].cpu().numpy(), Kr)
            except Exception as e:  # noqa: BLE001 — an explanatory extra must never cost the headline line
                rc["ivfpq_k5"] = {"error": repr(e)[:200]}
    # FP16 top-100, batch 1 on the 2.9 M base (the k of configs[3])
    hbm_rec("fp16_b1_top100", idx16, F16, n, 1, K4_TOPK, nv.PATH_GEMV, "scan_gemv_tma_kernel<f16,1> + merge_select_kernel", 50)
    # K3: INT8 (+scale), 2.9M x 768: batch 1 (bit-exact streaming scan), small batches and batch 1024 (exact tensor path)
    with nv.FlatIndexCuda.synthetic(I8, n, DIM, SEED, device=cx.local_rank) as idx8:
        hbm_rec("int8_b1", idx8, I8, n, 1, TOPK, nv.PATH_GEMV, "scan_gemv_tma_kernel<i8,1>", 100)
        hbm_rec("int8_b1_auto", idx8, I8, n, 1, TOPK, nv.PATH_AUTO, "device call: scan_gemv_tma_kernel<i8,1>; host call (e2e): scan_rows_kernel<i8> + rescore_exact_kernel", 50)
        hbm_rec("int8_b2", idx8, I8, n, 2, TOPK, nv.PATH_AUTO, "scan_rows_kernel<i8> + rescore_exact_kernel (AUTO)", 50)
        for b in (8, 32, 64, 128):
            hbm_rec(f"int8_b{b}_exact", idx8, I8, n, b, TOPK, nv.PATH_MMA_EXACT, "scan_rows_kernel<i8> (b <= 64) / scan_mma2_kernel<i8> + rescore_exact_kernel", 30)
        time.sleep(0.3)
        tensor_rec("int8_b1024_exact", idx8, I8, 1024, TOPK, "scan_mma2_kernel<i8> + rescore_exact_kernel", 8)
    time.sleep(0.5)
    # K2: FP16 batch sweep (BASELINE.json configs[1]): streaming scan -> tcgen05 crossover
    hbm_rec("fp16_b2", idx16, F16, n, 2, TOPK, nv.PATH_AUTO, "scan_rows_kernel<f16> + rescore_exact_kernel (AUTO)", 50)
    for b in (8, 64, 128):
        hbm_rec(f"fp16_b{b}_exact", idx16, F16, n, b, TOPK, nv.PATH_MMA_EXACT, "scan_rows_kernel<f16> (b <= 32) / scan_mma_kernel<f16> + rescore_exact_kernel", 30)
    time.sleep(0.3)
    tensor_rec("fp16_b256_exact", idx16, F16, 256, TOPK, "scan_mma2_kernel<f16,1 limb> + rescore_exact_kernel", 10)
    tensor_rec("fp16_b1024_exact", idx16, F16, 1024, TOPK, "scan_mma2_kernel<f16,1 limb> + rescore_exact_kernel", 8)
    if not args.no_clustered:
        try:
            rc["fp16_clustered_exact_b1024"] = clustered_extra(cx, n)
        except Exception as e:  # noqa: BLE001
            rc["fp16_clustered_exact_b1024"] = {"error": repr(e)[:200]}
    return rc, ec


def ivfpq_extra(cx, idxr, qh, Kr):
    """The stage in front of the refine: IVF-PQ candidates from this repo's own index, refine on the
    device; reference grid point nlist=4096, m=64, nprobe=64, REFINE_K=50 (Performance.md:673-687). The synthetic base has
    no cluster structure, so only build time, latency and throughput are reported here; recall on structured data is in
    profiles/ (scripts/bench_ivfpq.py). Parity unpinned (no FAISS)."""
    nv = cx.nv
    t0 = time.perf_counter()
    ivf = nv.IvfPqIndex(idxr, nlist=4096, pq_m=64, ntrain=200_000, niter=10, seed=SEED)
    build_s = time.perf_counter() - t0
    for _ in range(3):
        ivf.search_refine(qh[:1], 64, 50, Kr)
        ivf.search_refine(qh, 64, 50, Kr)
    t0 = time.perf_counter()
    for i in range(200):
        ivf.search_refine(qh[i:i + 1], 64, 50, Kr)
    lat_ms = (time.perf_counter() - t0) * 1e3 / 200
    t0 = time.perf_counter()
    for _ in range(5):
        ivf.search_refine(qh, 64, 50, Kr)
    b_ms = (time.perf_counter() - t0) * 1e3 / 5
    out = {"shape": "1M x 768 f16, nlist 4096, m 64, nprobe 64, REFINE_K 50, K 10", "build_s_niter10": build_s,
           "index_MB": ivf.index_bytes / 1e6, "e2e_latency_ms_per_query": lat_ms, "e2e_batch1000_ms": b_ms,
           "e2e_batch_qps": 1000 / b_ms * 1e3,
           "note": "host buffers in/out; candidates never leave the device; parity unpinned (no FAISS)"}
    ivf.close()
    return out


def clustered_extra(cx, n):
    """second distribution of SURVEY.md 8(d): rows around 256 Gaussian centres (sigma 0.3, re-normalised). A query's own
    cluster packs ~11 K scores close together, which is what the exact path's certificate has to separate: the rates say
    how many queries needed the second tier / the streaming scan (scripts/bench_clustered.py is the full sweep)."""
    torch, nv = cx.torch, cx.nv
    gen = torch.Generator(device=cx.dev).manual_seed(SEED)
    centres = torch.randn((256, DIM), device=cx.dev, generator=gen)
    centres /= centres.norm(dim=1, keepdim=True)
    payload = torch.empty((n, DIM), dtype=torch.float16, device=cx.dev)
    qrows = torch.arange(1024, device=cx.dev) * 2833 % n
    queries = torch.empty((1024, DIM), dtype=torch.float32, device=cx.dev)
    for lo in range(0, n, 131072):
        hi = min(n, lo + 131072)
        owner = torch.randint(0, 256, (hi - lo,), device=cx.dev, generator=gen)
        x = centres[owner] + (0.3 / DIM ** 0.5) * torch.randn((hi - lo, DIM), device=cx.dev, generator=gen)
        x /= x.norm(dim=1, keepdim=True)
        sel = (qrows >= lo) & (qrows < hi)
        queries[sel] = x[qrows[sel] - lo]
        payload[lo:hi] = x.half()
    torch.cuda.synchronize(cx.dev)
    out = {}
    with nv.FlatIndexCuda.from_device(payload.data_ptr(), None, F16, n, DIM, device=cx.local_rank, keepalive=payload) as idx:
        idx.set_path(nv.PATH_MMA_EXACT)
        for k in (10, 100):
            ms, _, _ = time_search(cx, idx, queries, 1024, k, iters=5)
            nq_, nf = idx.exact_stats()
            t2, streamed = idx.exact_tiers()
            idx.exact_stats(reset=True)
            out[f"k{k}"] = {"ms_per_batch": ms, "qps": 1024 / ms * 1e3, "first_certificate_failures": nf / max(nq_, 1),
                            "second_tier": t2 / max(nq_, 1), "streaming_scan": streamed / max(nq_, 1)}
    return out


def loader_record(cx, payload):
    """host -> device upload of the 4.46 GB base through nvdb_b200_index_create (multi-threaded
    staging into two pinned buffers overlapping the DMA) beside what the reference does, one synchronous pageable
    cudaMemcpy of the whole base (src/cuda_refine.cu:179-204), here torch's pageable host -> device copy."""
    torch, nv = cx.torch, cx.nv
    nbytes = payload.nbytes
    best = None
    for _ in range(2):
        with nv.FlatIndexCuda.from_host(payload, None, F16, device=cx.local_rank) as idx:
            sec, nb = idx.upload_info()
        best = sec if best is None else min(best, sec)
    t = torch.from_numpy(payload.view(np.int16))
    torch.cuda.synchronize(cx.dev)
    pg = None
    for _ in range(2):
        t0 = time.perf_counter()
        d = t.to(cx.dev)
        torch.cuda.synchronize(cx.dev)
        dt = time.perf_counter() - t0
        pg = dt if pg is None else min(pg, dt)
        del d
    return {"bytes": nbytes, "index_create_upload_s": best, "upload_GBps": nbytes / best / 1e9,
            "pageable_cudaMemcpy_s": pg, "pageable_GBps": nbytes / pg / 1e9,
            "note": "host array already resident in page cache / RAM (read back from the device-generated base)"}


def k4_single_gpu(cx, total_rows):
    """The fixed index of BASELINE.json configs[3] (FP16, 100 M x 768 = 153.6 GB, top-100) on ONE GPU: the N = 1 point of
    the strong-scaling curve the N > 1 lines continue. Runs last, after every other index has been freed."""
    torch, nv = cx.torch, cx.nv
    torch.cuda.empty_cache()
    free_b, _ = torch.cuda.mem_get_info(cx.dev)
    need = total_rows * DIM * 2 + (3 << 30)
    if free_b < need:
        return {"skipped": f"needs {need / 1e9:.1f} GB of device memory, {free_b / 1e9:.1f} GB free"}
    t0 = time.perf_counter()
    idx = nv.FlatIndexCuda.synthetic(F16, total_rows, DIM, SEED, device=cx.local_rank)
    gen_s = time.perf_counter() - t0
    try:
        rows = [(i * 104729 * 31) % total_rows for i in range(1024)]
        q = make_queries(cx, rows, total_rows)
        nbytes = payload_bytes(total_rows, DIM, F16)
        out = {"rows": total_rows, "index_GB": nbytes / 1e9, "k": K4_TOPK, "generation_s": gen_s}
        ms1, kms1, _ = time_search(cx, idx, q, 1, K4_TOPK, iters=12, warm=2)
        out["batch1"] = {"ms_per_step": ms1, "qps": 1e3 / ms1, "value_GBps": nbytes / (ms1 * 1e-3) / 1e9, "scan_kernel_ms": kms1,
                         "frac_of_hbm_peak": nbytes / (kms1 * 1e-3) / 1e9 / peaks()["hbm"]}
        ms, kms, _ = time_search(cx, idx, q, 1024, K4_TOPK, iters=3, warm=1)
        flops = 2.0 * 1024 * total_rows * DIM
        out["batch1024"] = {"ms_per_batch": ms, "qps": 1024 / ms * 1e3, "scan_kernel_ms": kms,
                            "useful_TFLOPs": flops / (kms * 1e-3) / 1e12,
                            "frac_of_bf16_sustained": flops / (kms * 1e-3) / 1e12 / peaks()["tf"]}
        ids = torch.empty((1024, K4_TOPK), dtype=torch.int64, device=cx.dev)
        sc = torch.empty((1024, K4_TOPK), dtype=torch.float32, device=cx.dev)
        idx.search_device(q.data_ptr(), 1024, K4_TOPK, ids.data_ptr(), sc.data_ptr(), cx.stream.cuda_stream)
        cx.stream.synchronize()
        out["self_top1_hits_of_1024"] = int((ids[:, 0].cpu() == torch.tensor(rows)).sum())
        return out
    finally:
        idx.close()


# ----------------------------------------------------------------------------------------------------------
# our arm
# ----------------------------------------------------------------------------------------------------------
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
    stream = torch.cuda.Stream(device=dev)
    cx = Ctx(torch, dist, nv, dev, local_rank, rank, world, stream)
    line = run_single(cx, args) if world == 1 else run_sharded(cx, args)
    if rank == 0:
        emit(line)
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()
    return 0


def run_single(cx, args):
    torch, nv = cx.torch, cx.nv
    K, W = args.steps, args.warmup
    n_rows = int(os.environ.get("NVDB_BENCH_ROWS", N_ROWS))
    nbytes = payload_bytes(n_rows, DIM, F16)
    pk = peaks()
    idx = nv.FlatIndexCuda.synthetic(F16, n_rows, DIM, SEED, device=cx.local_rank)
    qrows = [(i * 104729 * 31) % n_rows for i in range(NQ_POOL)]
    q_dev = make_queries(cx, qrows, n_rows)
    q_host = q_dev.cpu().numpy()
    out_ids = torch.empty((1, TOPK), dtype=torch.int64, device=cx.dev)
    out_sc = torch.empty((1, TOPK), dtype=torch.float32, device=cx.dev)
    sp = cx.stream.cuda_stream

    def step_device(i):
        idx.search_device(q_dev[i % NQ_POOL].data_ptr(), 1, TOPK, out_ids.data_ptr(), out_sc.data_ptr(), sp)

    def step_e2e(i):
        return idx.search_topk_dot(q_host[i % NQ_POOL], TOPK)

    sampler = ClockSampler(cx.local_rank)
    sampler.start()
    prof = {}

    def on_timed(on):
        sampler.mark_timed(on)
        if on:
            idx.profile(True)
            prof["l0"] = nv.launch_count()
        else:
            prof["l1"] = nv.launch_count()
            prof["scan_ms"], prof["merge_ms"], prof["calls"] = idx.profile_read()
            idx.profile(False)

    blocks = timed_blocks(cx, step_device, K, W, on_timed=on_timed)
    ms = statistics.median(blocks)
    value = nbytes * K / (ms * 1e-3) / 1e9
    # ---- end to end through the host-buffer API (H2D of the query + D2H of the result inside the timed region)
    eblocks = timed_blocks(cx, step_e2e, K, W, wall=True, on_timed=sampler.mark_timed)
    e_ms = statistics.median(eblocks)
    clocks = sampler.stop()
    ids0, _ = idx.search_topk_dot(q_host[0], TOPK)
    assert int(ids0[0]) == qrows[0], f"self-query check failed: {ids0[0]} vs {qrows[0]}"

    kern_ms = prof["scan_ms"] / max(prof["calls"], 1)
    achieved = nbytes / (kern_ms * 1e-3) / 1e9
    line = {
        "metric": "payload_equiv_bandwidth_GBps", "value": value, "unit": "GB/s", "n_gpus": 1, "steps": K, "warmup": W,
        "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic",
        "config": {"workload": workload_name(1), "rows": n_rows, "dim": DIM, "k": TOPK, "batch": 1, "base_dtype": "f16",
                   "l2": f"base ({nbytes / 1e9:.2f} GB) is larger than L2 (126 MB); queries cycle through {NQ_POOL} rows",
                   "parallelism": "single GPU",
                   "timing": f"{len(blocks)} blocks of {K} steps, median block reported (min {min(blocks) / K:.4f}, max {max(blocks) / K:.4f} ms per step)"},
        "qps": K / (ms * 1e-3),
        "e2e": {"value": nbytes * K / (e_ms * 1e-3) / 1e9, "unit": "GB/s", "qps": K / (e_ms * 1e-3), "ms_per_step": e_ms / K,
                "h2d_bytes_per_step": DIM * 4, "d2h_bytes_per_step": TOPK * 12, "api": "nvdb_b200_search (host buffers)",
                "blocks": len(eblocks)},
        "gpu_launches": int(round((prof["l1"] - prof["l0"]) / len(blocks))),
        "roofline": {"bound": "hbm", "kernel": "scan_gemv_tma_kernel<f16,1>", "achieved": achieved, "peak": pk["hbm"],
                     "unit": "GB/s", "frac": achieved / pk["hbm"], "traffic": None, "peak_source": pk["hbm_src"],
                     "frac_of_nominal_8TBps": achieved / 8000.0, "kernel_ms": kern_ms,
                     "merge_kernel_ms": prof["merge_ms"] / max(prof["calls"], 1), "algorithmic_bytes_per_launch": nbytes,
                     "launches_timed": prof["calls"]},
        "clocks": clocks,
    }
    ncu_traffic = ROOT / "profiles" / "scan_f16_traffic.json"
    if ncu_traffic.exists():
        try:
            line["roofline"]["traffic"] = json.loads(ncu_traffic.read_text()).get("dram_bytes_per_launch")
            line["roofline"]["traffic_source"] = "ncu --set full capture of this kernel (profiles/scan_f16_traffic.json), not a live counter"
        except Exception:  # noqa: BLE001
            pass
    if not args.no_extras:
        rc, ec = roofline_configs(cx, idx, q_dev, args)
        line["roofline"]["configs"] = rc
        line["e2e"]["configs"] = ec
    payload = None
    if not args.no_cpu:
        payload, _ = idx.read_rows(0, n_rows)      # the CPU leg scans the very bytes the GPU scanned
    idx.close()
    if payload is not None and not args.no_extras:
        try:
            line["loader"] = loader_record(cx, payload)
        except Exception as e:  # noqa: BLE001
            line["loader"] = {"error": repr(e)[:300]}
    if not args.no_extras and not args.no_k4:
        try:
            line["strong_scaling_n1"] = k4_single_gpu(cx, int(os.environ.get("NVDB_BENCH_K4_ROWS", K4_ROWS)))
        except Exception as e:  # noqa: BLE001
            line["strong_scaling_n1"] = {"error": repr(e)[:300]}
    if payload is not None:
        with tempfile.TemporaryDirectory(dir=shm_dir()) as tmp:
            path, cq = make_cpu_base_file(tmp, n_rows, payload=payload)
            del payload
            r = cpu_reference_run(path, n_rows, cq, TOPK, steps=600, warmup=2, budget_s=15.0)
        line["cpu_baseline"] = {"value": r["value"], "unit": "GB/s", "cores": r["cores"], "kind": r["kind"],
                                "sample": r["sample"], "qps": r["qps"], "ms_per_step": r["ms_per_step"],
                                "omp_threads": r["omp_threads"], "pool": r.get("pool")}
    return line


def run_sharded(cx, args):
    """N > 1: strong scaling on the fixed 100 M-row index (module docstring)."""
    torch, nv, dist = cx.torch, cx.nv, cx.dist
    K, W = args.steps, args.warmup
    world, rank = cx.world, cx.rank
    pk = peaks()
    total_rows = int(os.environ.get("NVDB_BENCH_K4_ROWS", K4_ROWS))
    r0, nloc = nv.shard_rows(total_rows, world, rank)
    idx = nv.FlatIndexCuda.synthetic(F16, nloc, DIM, SEED, first_row=r0, device=cx.local_rank, row_offset=r0)
    nbytes_total = payload_bytes(total_rows, DIM, F16)
    qrows = [(i * 104729 * 31) % total_rows for i in range(1024)]
    q_dev = make_queries(cx, qrows, total_rows)
    q_host = q_dev[:NQ_POOL].cpu().numpy()
    exchange = os.environ.get("NVDB_B200_EXCHANGE", "auto")
    sharded = nv.ShardedFlatIndex(idx, cx.dev, exchange=exchange)

    def step_device(i):
        sharded.search(q_dev[i % NQ_POOL:i % NQ_POOL + 1], K4_TOPK)

    q_pin = torch.from_numpy(q_host).pin_memory()
    r_ids = torch.empty((1, K4_TOPK), dtype=torch.int64).pin_memory()
    r_sc = torch.empty((1, K4_TOPK), dtype=torch.float32).pin_memory()
    q_stage = torch.empty((DIM,), dtype=torch.float32, device=cx.dev)

    def step_e2e(i):
        q_stage.copy_(q_pin[i % NQ_POOL], non_blocking=True)
        f_ids, f_sc = sharded.search(q_stage.view(1, DIM), K4_TOPK)
        r_ids.copy_(f_ids, non_blocking=True)
        r_sc.copy_(f_sc, non_blocking=True)
        cx.stream.synchronize()

    sampler = ClockSampler(cx.local_rank)
    sampler.start()
    prof = {}

    def on_timed(on):
        sampler.mark_timed(on)
        if on:
            idx.profile(True)
            prof["l0"] = nv.launch_count()
        else:
            prof["l1"] = nv.launch_count()
            prof["scan_ms"], prof["merge_ms"], prof["calls"] = idx.profile_read()
            idx.profile(False)

    blocks = timed_blocks(cx, step_device, K, W, on_timed=on_timed)
    ms = statistics.median(blocks)
    eblocks = timed_blocks(cx, step_e2e, K, W, wall=True, on_timed=sampler.mark_timed)
    e_ms = statistics.median(eblocks)
    clocks = sampler.stop()
    # ---- sanity: the row a query was made from lives on ONE shard; every rank must find it through the exchange
    for i in (0, 1, NQ_POOL - 1):
        with torch.cuda.stream(cx.stream):
            g_ids, _ = sharded.search(q_dev[i:i + 1], K4_TOPK)
        cx.stream.synchronize()
        assert int(g_ids[0, 0]) == qrows[i], f"rank {rank}: sharded self-query check failed: {int(g_ids[0, 0])} vs {qrows[i]}"
    assert sharded.exchange_status() == 0, "peer exchange timed out"
    xkind = sharded.exchange

    local_bytes = payload_bytes(nloc, DIM, F16)
    kern_ms = cx.max_over_ranks(prof["scan_ms"] / max(prof["calls"], 1))
    achieved = local_bytes / (kern_ms * 1e-3) / 1e9
    configs = {}

    def timed_case(name, s_idx, s_sharded, qd, nq, k, steps, warm, rows_total, note):
        def st(i):
            s_sharded.search(qd[(i * nq) % max(1, qd.shape[0] - nq + 1):][:nq], k)
        bl = timed_blocks(cx, st, steps, warm, min_total_s=0.3)
        m = statistics.median(bl) / steps
        configs[name] = {"ms_per_step": m, "qps": nq / m * 1e3, "batch": nq, "k": k, "rows_total": rows_total,
                         "value_GBps": payload_bytes(rows_total, DIM, F16) / (m * 1e-3) / 1e9, "blocks": len(bl), "note": note}

    if not args.no_extras:
        # batch 1024 on the same fixed index (tensor-core exact path per rank + exchange kernel)
        timed_case("k4_b1024", idx, sharded, q_dev, 1024, K4_TOPK, 3, 2, total_rows,
                   "BASELINE.json configs[3] at batch 1024: QPS is the metric; the strong-scaling reference is strong_scaling_n1 of the N=1 line")
        xq, xf = idx.exact_stats(reset=True)
        configs["k4_b1024"]["rank0_certificate_fallback_rate"] = xf / max(xq, 1)
    sharded.close()
    idx.close()
    if not args.no_extras:
        # the hard case: the 2,903,054-row base of configs[1] split N ways, batch 1, top-10 — 0.68 ms / N of scan per rank
        # against the fixed cost of one exchange round
        r0s, ns = nv.shard_rows(N_ROWS, world, rank)
        with nv.FlatIndexCuda.synthetic(F16, ns, DIM, SEED, first_row=r0s, device=cx.local_rank, row_offset=r0s) as idx_s:
            sh_s = nv.ShardedFlatIndex(idx_s, cx.dev, exchange=exchange)
            qs = make_queries(cx, [(i * 104729 * 31) % N_ROWS for i in range(NQ_POOL)], N_ROWS)
            timed_case("split_2p9M_b1", idx_s, sh_s, qs, 1, TOPK, 200, 10, N_ROWS,
                       f"strong scaling of BASELINE.json configs[1]: {N_ROWS} rows split over {world} GPUs "
                       f"({payload_bytes(ns, DIM, F16) / pk['hbm'] / 1e6:.3f} ms of HBM time per rank)")
            sh_s.close()
        # round 1's line: weak scaling, every rank keeps its own 2,903,054-row shard of an N-times larger index
        with nv.FlatIndexCuda.synthetic(F16, N_ROWS, DIM, SEED, first_row=rank * N_ROWS, device=cx.local_rank,
                                        row_offset=rank * N_ROWS) as idx_w:
            sh_w = nv.ShardedFlatIndex(idx_w, cx.dev, exchange=exchange)
            qw = make_queries(cx, [(i * 104729 * 31) % (world * N_ROWS) for i in range(NQ_POOL)], world * N_ROWS)
            timed_case("weak_2p9M_per_gpu_b1", idx_w, sh_w, qw, 1, TOPK, 200, 10, world * N_ROWS,
                       "weak scaling: 2,903,054 rows per GPU, batch 1, top-10 (round 1's headline)")
            sh_w.close()

    par = (f"row-shard x{world} of one fixed index, " +
           ("fused peer-memory exchange + merge kernel (NVLink stores, no NCCL)" if xkind == "peer" else "NCCL all-gather + merge kernel"))
    line = {
        "metric": "payload_equiv_bandwidth_GBps", "value": nbytes_total * K / (ms * 1e-3) / 1e9, "unit": "GB/s",
        "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": ms / K, "higher_is_better": True, "scaling": "strong",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": {"workload": workload_name(world, total_rows), "rows_total": total_rows, "rows_per_gpu": nloc, "dim": DIM,
                   "k": K4_TOPK, "batch": 1, "base_dtype": "f16",
                   "l2": f"shard ({local_bytes / 1e9:.1f} GB) is larger than L2 (126 MB); queries cycle through {NQ_POOL} rows",
                   "parallelism": par,
                   "timing": f"{len(blocks)} blocks of {K} steps, median block reported (min {min(blocks) / K:.4f}, max {max(blocks) / K:.4f} ms per step)",
                   "strong_scaling_note": "value = bytes of the WHOLE index / step time; the N=1 point of this index is strong_scaling_n1 in the "
                                          "N=1 line (its headline is the 2.9 M-row base, same kernel, same GB/s regime)"},
        "qps": K / (ms * 1e-3),
        "e2e": {"value": nbytes_total * K / (e_ms * 1e-3) / 1e9, "unit": "GB/s", "qps": K / (e_ms * 1e-3), "ms_per_step": e_ms / K,
                "h2d_bytes_per_step": DIM * 4, "d2h_bytes_per_step": K4_TOPK * 12, "blocks": len(eblocks),
                "api": f"ShardedFlatIndex.search: pinned host query -> H2D -> local scan -> {xkind} exchange + merge -> D2H"},
        "gpu_launches": int(round((prof["l1"] - prof["l0"]) / len(blocks))),
        "roofline": {"bound": "hbm", "kernel": "scan_gemv_tma_kernel<f16,1> (per rank, its shard)", "achieved": achieved,
                     "peak": pk["hbm"], "unit": "GB/s", "frac": achieved / pk["hbm"], "traffic": None, "peak_source": pk["hbm_src"],
                     "kernel_ms": kern_ms, "algorithmic_bytes_per_launch": local_bytes,
                     "note": "slowest rank's scan kernel (includes the spin-wait for the peers when the exchange runs in its tail)"},
        "clocks": clocks,
        "configs": configs,
    }
    return line


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=20)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--no-extras", action="store_true", help="headline only (skip roofline.configs / e2e.configs / configs)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-ivfpq", action="store_true", help="skip the IVF-PQ extra (index build ~2 s)")
    ap.add_argument("--no-clustered", action="store_true", help="skip the clustered-base extra")
    ap.add_argument("--no-k4", action="store_true", help="N=1: skip the 100 M-row single-GPU point (153.6 GB)")
    args = ap.parse_args()
    if args.warmup < 3:
        args.warmup = 3
    if args.impl == "reference":
        return run_reference(args)
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
