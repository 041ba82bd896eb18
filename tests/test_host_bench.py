"""The C++ host layer (nano-vectordb_b200/host): loader behaviour on CPU, and the nvdb_bench mode=cuda harness against
golden ids written by the reference's own nvdb_gt_build (GPU)."""
import os
import re
import subprocess
from pathlib import Path

import numpy as np
import pytest

from _oracle import F16, F32, I8, REF_DIR, run_ref_tool, write_raw12, write_vecbin

ROOT = Path(__file__).resolve().parent.parent
BENCH = ROOT / "nano-vectordb_b200" / "bin" / "nvdb_bench"


@pytest.fixture(scope="module")
def bench():
    if not BENCH.exists():
        subprocess.run(["make", "-C", str(ROOT), "host"], check=True, capture_output=True)
    return BENCH


def run(bench, *args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([str(bench), *map(str, args)], capture_output=True, text=True, env=e)


def test_usage_and_loader_errors(bench, oracle, tmp_path):
    assert run(bench).returncode == 1
    f = oracle.synth_f32(1, 0, 20, 16)
    write_vecbin(tmp_path / "b.vecbin", F32, f)
    write_raw12(tmp_path / "q.raw12", f[:3])
    write_raw12(tmp_path / "q8.raw12", f[:3, :8])
    r = run(bench, tmp_path / "b.vecbin", tmp_path / "q.raw12", 5, "omp")
    assert r.returncode == 1 and "mode=cuda only" in r.stderr
    r = run(bench, tmp_path / "b.vecbin", tmp_path / "q8.raw12", 5)
    assert r.returncode == 2 and "Dim mismatch: base.dim=16, query.dim=8" in r.stderr
    r = run(bench, tmp_path / "nope.vecbin", tmp_path / "q.raw12", 5)
    assert r.returncode == 4 and "open:" in r.stderr
    raw = (tmp_path / "b.vecbin").read_bytes()
    (tmp_path / "short.vecbin").write_bytes(raw[:-4])
    r = run(bench, tmp_path / "short.vecbin", tmp_path / "q.raw12", 5)
    assert r.returncode == 4 and "VecbinHeader ok but file size mismatch" in r.stderr
    (tmp_path / "tiny").write_bytes(b"abc")
    r = run(bench, tmp_path / "tiny", tmp_path / "q.raw12", 5)
    assert r.returncode == 4 and "File too small" in r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
def test_nvdb_bench_cuda_matches_reference_gtbin(bench, oracle, tmp_path, dt):
    if not (REF_DIR / "nvdb_gt_build").exists():
        pytest.skip("oracle/_ref tools not built")
    dtype = {"f32": F32, "f16": F16, "i8": I8}[dt]
    n, d, nq, k = 30000, 768, 24, 10
    base, scales = oracle.synth(dtype, 5, 0, n, d)
    f = oracle.synth_f32(5, 0, n, d)
    rng = np.random.default_rng(0)
    q = f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32)
    write_vecbin(tmp_path / "b.vecbin", dtype, base, scales)
    write_raw12(tmp_path / "q.raw12", q)
    run_ref_tool("nvdb_gt_build", tmp_path / "b.vecbin", tmp_path / "q.raw12", k, tmp_path / "gt.gtbin",
                 env={"GT_MODE": "st", "WARMUP": "0"})
    env = {"GT_PATH": str(tmp_path / "gt.gtbin")}
    # single queries -> bit-exact streaming scan: ids identical to the reference's ground truth
    r = run(bench, tmp_path / "b.vecbin", tmp_path / "q.raw12", k, "cuda", 1, 2, env=env)
    assert r.returncode == 0, r.stderr
    out = r.stdout
    for key in ("mode=cuda", "Base count=30000 dim=768 | Query count=24 | k=10 | warmup=2", "Total:", "Avg_query:", "p50:",
                "p95:", "p99:", "sink=", "bytes_per_query=", "payload_equiv_bandwidth_GBps=", "batch_q=1"):
        assert key in out, key
    assert float(re.search(r"gt_match=([0-9.]+)", out).group(1)) == 1.0
    want = n * d * {F32: 4, F16: 2, I8: 1}[dtype] + (4 * n if dtype == I8 else 0)
    assert int(re.search(r"bytes_per_query=(\d+)", out).group(1)) == want
    # batches of 8 -> exact tensor-core path for f16 / i8 (bit-identical to the streaming scan)
    r = run(bench, tmp_path / "b.vecbin", tmp_path / "q.raw12", k, "cuda", 1, 2, 8, env=env)
    assert r.returncode == 0, r.stderr
    assert "batch_samples=3" in r.stdout and "Avg_batch:" in r.stdout and "batch_p99:" in r.stdout
    assert float(re.search(r"gt_match=([0-9.]+)", r.stdout).group(1)) >= 0.99
    assert f"path_used={1 if dtype == F32 else 4}" in r.stdout   # AUTO: exact tensor-core path for f16 / i8 batches


@pytest.mark.gpu
@pytest.mark.parametrize("dt,shards,batch", [("f16", 2, 1), ("i8", 3, 1), ("f32", 2, 4), ("f16", 2, 16)])
def test_nvdb_bench_cuda_row_shards_exchange_on_the_devices(bench, oracle, tmp_path, dt, shards, batch):
    """gpus > 1: the shards exchange and merge their lists on the devices (NVLink peer memory when they are different
    GPUs; with NVDB_BENCH_OVERSUBSCRIBE=1 several shards share one GPU) — ids identical to the reference's gtbin."""
    import torch
    if not (REF_DIR / "nvdb_gt_build").exists():
        pytest.skip("oracle/_ref tools not built")
    dtype = {"f32": F32, "f16": F16, "i8": I8}[dt]
    n, d, nq, k = 20001, 256, 16, 10
    base, scales = oracle.synth(dtype, 9, 0, n, d)
    f = oracle.synth_f32(9, 0, n, d)
    rng = np.random.default_rng(1)
    q = f[rng.integers(0, n, nq)] + (0.02 * rng.standard_normal((nq, d))).astype(np.float32)
    write_vecbin(tmp_path / "b.vecbin", dtype, base, scales)
    write_raw12(tmp_path / "q.raw12", q)
    run_ref_tool("nvdb_gt_build", tmp_path / "b.vecbin", tmp_path / "q.raw12", k, tmp_path / "gt.gtbin",
                 env={"GT_MODE": "st", "WARMUP": "0"})
    env = {"GT_PATH": str(tmp_path / "gt.gtbin"), "NVDB_B200_XCHG_TIMEOUT_MS": "10000"}
    if torch.cuda.device_count() < shards:
        env["NVDB_BENCH_OVERSUBSCRIBE"] = "1"
        # shards sharing a GPU spin for each other's results: nothing may wait for the whole device meanwhile — the
        # harness grows the workspaces beforehand, and kernels are loaded at start-up instead of at first launch
        env["CUDA_MODULE_LOADING"] = "EAGER"
    r = run(bench, tmp_path / "b.vecbin", tmp_path / "q.raw12", k, "cuda", shards, 2, batch, env=env)
    assert r.returncode == 0, r.stderr
    assert f"shards={shards}" in r.stdout
    assert float(re.search(r"gt_match=([0-9.]+)", r.stdout).group(1)) >= (1.0 if batch == 1 or dtype == F32 else 0.99)
