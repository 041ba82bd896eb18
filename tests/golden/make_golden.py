#!/usr/bin/env python
"""Generates tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE (oracle/_ref, built from /root/reference by
`make -C oracle ref`). Run in the build container only:  python tests/golden/make_golden.py

The reference ships no golden vectors of its own, so these are outputs of the reference's own
binaries / library on deterministic inputs:
  * inputs  : rows of the integer-hash generator orc_synth_rows_f32 (bit-reproducible everywhere; its sha256 is
              stored so a test can prove it regenerated the same bytes)
  * f16     : reference tool nvdb_convert_f16            (tools/nvdb_convert_f16.cpp)
  * i8+scale: reference tool nvdb_quantize_i8            (apps/nvdb_quantize_i8.cpp)
  * queries : reference tool nvdb_make_query (raw12)     (tools/nvdb_make_query.cpp)
  * gt ids  : reference tool nvdb_gt_build, GT_MODE=st   (apps/nvdb_gt_build.cpp)
  * scores  : FlatIndex / FlatIndexOMP / FlatIndexPool::search_topk_dot through oracle/ref_shim.cpp
  * dots    : nvdb::dot_f32 / dot_f32_f16base / dot_f32_i8base on random vectors incl. ragged dims
"""
import hashlib
import sys
import tempfile
from pathlib import Path

import numpy as np

HERE = Path(__file__).resolve().parent
sys.path.insert(0, str(HERE.parent))
import _oracle as O  # noqa: E402

CASES = [  # name, n, dim, nq, k, seed
    ("n1024_d32", 1024, 32, 8, 10, 101),
    ("n96_d768", 96, 768, 4, 10, 102),
    ("n300_d384_k100", 300, 384, 3, 100, 103),
]


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a).tobytes()).hexdigest()


def scan_case(orc, ref, name, n, d, nq, k, seed, tmp):
    f32 = orc.synth_f32(seed, 0, n, d)
    p32, p16, p8, pq = (tmp / f"{name}.{s}" for s in ("f32", "f16", "i8", "q"))
    O.write_vecbin(p32, O.F32, f32)
    O.run_ref_tool("nvdb_convert_f16", p32, p16)
    O.run_ref_tool("nvdb_quantize_i8", p32, p8)
    O.run_ref_tool("nvdb_make_query", p32, pq, nq, 42, "random")
    _, f16, _ = O.read_vecbin(p16)
    _, i8, scales = O.read_vecbin(p8)
    _, queries, _ = O.read_vecbin(pq)
    out = dict(n=n, dim=d, nq=nq, k=k, seed=seed, f32_sha256=sha(f32), f16=f16.copy(), i8=i8.copy(),
               i8_scales=scales.copy(), queries=queries.copy())
    for tag, path in (("f32", p32), ("f16", p16), ("i8", p8)):
        gt = tmp / f"{name}.{tag}.gtbin"
        O.run_ref_tool("nvdb_gt_build", path, pq, k, gt, env={"GT_MODE": "st", "WARMUP": "0"})
        ids, meta = O.read_gtbin(gt)
        out[f"gt_ids_{tag}"] = ids.copy()
        h = ref.load(path)
        sc = np.zeros((nq, min(k, n)), np.float32)
        for mode in ("st", "omp", "pool"):
            for qi in range(nq):
                rid, rsc = ref.search(h, queries[qi], k, mode, threads=4)
                if mode == "st":
                    assert np.array_equal(rid.astype(np.uint32), ids[qi]), "gtbin vs FlatIndex"
                    sc[qi] = rsc
                else:  # the reference's variants agree bit-for-bit when scores are distinct
                    assert np.array_equal(rsc.view(np.uint32), sc[qi].view(np.uint32)), (name, tag, mode)
        ref.free(h)
        out[f"scores_{tag}"] = sc
    np.savez_compressed(HERE / f"scan_{name}.npz", **out)
    print("wrote", f"scan_{name}.npz")


def dot_cases(ref):
    rng = np.random.default_rng(7)
    dims = [1, 3, 4, 7, 8, 9, 13, 15, 16, 17, 24, 31, 33, 40, 100, 384, 768]
    rec = {}
    for d in dims:
        q = rng.standard_normal((6, d)).astype(np.float32)
        x32 = rng.standard_normal((6, d)).astype(np.float32)
        x16 = x32.astype(np.float16).view(np.uint16)
        x8 = rng.integers(-128, 128, (6, d)).astype(np.int8)
        sc = rng.uniform(1e-3, 1e-2, 6).astype(np.float32)
        r32 = np.array([ref.dot(O.F32, q[i], x32[i]) for i in range(6)], np.float32)
        r16 = np.array([ref.dot(O.F16, q[i], x16[i]) for i in range(6)], np.float32)
        r8 = np.array([ref.dot(O.I8, q[i], x8[i], float(sc[i])) for i in range(6)], np.float32)
        rec[f"d{d}_q"], rec[f"d{d}_x32"], rec[f"d{d}_x16"], rec[f"d{d}_x8"] = q, x32, x16, x8
        rec[f"d{d}_scale"], rec[f"d{d}_r32"], rec[f"d{d}_r16"], rec[f"d{d}_r8"] = sc, r32, r16, r8
    rec["dims"] = np.array(dims)
    np.savez_compressed(HERE / "dots.npz", **rec)
    print("wrote dots.npz")


def main():
    orc, ref = O.Oracle(), O.Reference()
    with tempfile.TemporaryDirectory() as t:
        tmp = Path(t)
        for c in CASES:
            scan_case(orc, ref, *c, tmp)
    dot_cases(ref)


if __name__ == "__main__":
    main()
