"""The GPU-backed host tools (nano-vectordb_b200/host: nvdb_convert_f16, nvdb_quantize_i8, nvdb_gt_build) against the
UNMODIFIED reference tools compiled into oracle/_ref: output files must be byte-identical (converters) / hold the same
ids (ground truth; tie-free data)."""
import os
import subprocess
from pathlib import Path

import numpy as np
import pytest

from _oracle import F32, REF_DIR, read_gtbin, run_ref_tool, write_raw12, write_vecbin

ROOT = Path(__file__).resolve().parent.parent
BIN = ROOT / "nano-vectordb_b200" / "bin"


@pytest.fixture(scope="module")
def tools():
    if not (BIN / "nvdb_gt_build").exists() or not (BIN / "nvdb_quantize_i8").exists():
        subprocess.run(["make", "-C", str(ROOT), "host"], check=True, capture_output=True)
    return BIN


def run(exe, *args, env=None):
    e = dict(os.environ)
    e.update(env or {})
    return subprocess.run([str(exe), *map(str, args)], capture_output=True, text=True, env=e)


def test_usage_and_input_checks_without_gpu(tools, oracle, tmp_path):
    assert run(tools / "nvdb_convert_f16").returncode == 1
    assert run(tools / "nvdb_quantize_i8").returncode == 1
    assert run(tools / "nvdb_gt_build").returncode == 1
    f = oracle.synth_f32(1, 0, 20, 16)
    h = oracle.to_f16(f)
    write_vecbin(tmp_path / "h.vecbin", 2, h)
    r = run(tools / "nvdb_convert_f16", tmp_path / "h.vecbin", tmp_path / "o.vecbin")
    assert r.returncode == 2 and "Input dtype must be Float32. Got dtype=2" in r.stderr      # reference exit code / text
    write_vecbin(tmp_path / "b.vecbin", F32, f)
    write_raw12(tmp_path / "q8.raw12", f[:3, :8])
    r = run(tools / "nvdb_gt_build", tmp_path / "b.vecbin", tmp_path / "q8.raw12", 5, tmp_path / "o.gtbin")
    assert r.returncode == 3 and "Dim mismatch: base.dim=16 query.dim=8" in r.stderr
    write_raw12(tmp_path / "q.raw12", f[:3])
    r = run(tools / "nvdb_gt_build", tmp_path / "b.vecbin", tmp_path / "q.raw12", 0, tmp_path / "o.gtbin")
    assert r.returncode == 4 and "k must be > 0" in r.stderr


@pytest.mark.gpu
def test_converters_write_the_same_bytes_as_the_reference_tools(tools, oracle, tmp_path):
    if not (REF_DIR / "nvdb_convert_f16").exists():
        pytest.skip("oracle/_ref tools not built")
    rng = np.random.default_rng(3)
    f = oracle.synth_f32(9, 0, 5003, 384) * rng.choice([1e-4, 1.0, 50.0], size=(5003, 1)).astype(np.float32)
    f[17] = 0.0
    write_vecbin(tmp_path / "f32.vecbin", F32, np.ascontiguousarray(f, np.float32))
    run_ref_tool("nvdb_convert_f16", tmp_path / "f32.vecbin", tmp_path / "ref16.vecbin")
    run_ref_tool("nvdb_quantize_i8", tmp_path / "f32.vecbin", tmp_path / "ref8.vecbin")
    run_ref_tool("nvdb_quantize_i8", tmp_path / "ref16.vecbin", tmp_path / "ref8from16.vecbin")
    for exe, src, out, ref in (("nvdb_convert_f16", "f32.vecbin", "our16.vecbin", "ref16.vecbin"),
                               ("nvdb_quantize_i8", "f32.vecbin", "our8.vecbin", "ref8.vecbin"),
                               ("nvdb_quantize_i8", "ref16.vecbin", "our8from16.vecbin", "ref8from16.vecbin")):
        r = run(tools / exe, tmp_path / src, tmp_path / out)
        assert r.returncode == 0, r.stderr
        assert (tmp_path / out).read_bytes() == (tmp_path / ref).read_bytes(), f"{exe} {src}"


@pytest.mark.gpu
@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
def test_gt_build_matches_reference_gtbin(tools, oracle, tmp_path, dt):
    if not (REF_DIR / "nvdb_gt_build").exists():
        pytest.skip("oracle/_ref tools not built")
    dtype = {"f32": 1, "f16": 2, "i8": 3}[dt]
    n, d, nq, k = 20000, 384, 37, 10
    base, scales = oracle.synth(dtype, 6, 0, n, d)
    f = oracle.synth_f32(6, 0, n, d)
    rng = np.random.default_rng(1)
    queries = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.01 * rng.standard_normal((nq, d))).astype(np.float32))
    write_vecbin(tmp_path / "b.vecbin", dtype, base, scales)
    write_raw12(tmp_path / "q.raw12", queries)
    run_ref_tool("nvdb_gt_build", tmp_path / "b.vecbin", tmp_path / "q.raw12", k, tmp_path / "ref.gtbin", env={"GT_MODE": "st"})
    r = run(tools / "nvdb_gt_build", tmp_path / "b.vecbin", tmp_path / "q.raw12", k, tmp_path / "our.gtbin", env={"GT_BATCH": "16"})
    assert r.returncode == 0, r.stderr
    assert "GT build: N=20000 Q=37 d=384 k=10 mode=cuda" in r.stdout and "Wrote GT:" in r.stdout
    ref_ids, ref_h = read_gtbin(tmp_path / "ref.gtbin")
    our_ids, our_h = read_gtbin(tmp_path / "our.gtbin")
    assert our_h == ref_h
    assert (tmp_path / "our.gtbin").stat().st_size == (tmp_path / "ref.gtbin").stat().st_size
    assert np.array_equal(our_ids, ref_ids)      # continuous random data: no exact ties
