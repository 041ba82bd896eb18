"""ctypes bindings to the CPU oracle (oracle/nvdb_oracle.c) and to the compiled reference (oracle/_ref).

TEST INFRASTRUCTURE: imported only by tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs. The product package never imports this.
"""
from __future__ import annotations

import ctypes as C
import os
import struct
import subprocess
from pathlib import Path

import numpy as np

ROOT = Path(__file__).resolve().parent.parent
ORACLE_DIR = ROOT / "oracle"
PORT_SO = ORACLE_DIR / "_build" / "libnvdb_oracle.so"
REF_DIR = ORACLE_DIR / "_ref"
REF_SO = REF_DIR / "libnvdb_ref.so"
REFCUDA_SO = REF_DIR / "libnvdb_refcuda.so"
REFERENCE_SRC = Path("/root/reference")

F32, F16, I8 = 1, 2, 3
NP_DT = {F32: np.float32, F16: np.uint16, I8: np.int8}

_u64p = C.POINTER(C.c_uint64)
_u32p = C.POINTER(C.c_uint32)
_f32p = C.POINTER(C.c_float)


def _p(a: np.ndarray | None):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def build_port() -> Path:
    """Compile the C restatement (seconds)."""
    if not PORT_SO.exists() or PORT_SO.stat().st_mtime < (ORACLE_DIR / "nvdb_oracle.c").stat().st_mtime:
        subprocess.run(["make", "-C", str(ORACLE_DIR), "port"], check=True, capture_output=True)
    return PORT_SO


def build_ref(cuda: bool = False) -> bool:
    """Compile the unmodified reference from /root/reference when that tree is present (build container only)."""
    want = REFCUDA_SO if cuda else REF_SO
    if want.exists():
        return True
    if not REFERENCE_SRC.exists():
        return False
    subprocess.run(["make", "-C", str(ORACLE_DIR), "refcuda" if cuda else "ref"], check=True, capture_output=True)
    return want.exists()


class Oracle:
    """The C restatement."""

    def __init__(self):
        self.lib = C.CDLL(str(build_port()))
        L = self.lib
        L.orc_dot_f32.restype = C.c_float
        L.orc_dot_f16.restype = C.c_float
        L.orc_dot_i8.restype = C.c_float
        L.orc_dot_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_dot_f16.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_dot_i8.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_float]
        L.orc_score_all.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        for f in (L.orc_flat_search, L.orc_flat_search_mt):
            f.restype = C.c_uint32
            f.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p, C.c_uint32,
                          C.c_void_p, C.c_void_p]
        L.orc_flat_search_batch.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p,
                                            C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_quantize_i8.argtypes = [C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_convert_f16.argtypes = [C.c_void_p, C.c_uint64, C.c_void_p]
        L.orc_refine_l2.argtypes = [C.c_int, C.c_void_p, C.c_void_p, C.c_uint32, C.c_uint64, C.c_uint32, C.c_void_p,
                                    C.c_void_p, C.c_uint32, C.c_uint32, C.c_uint32, C.c_void_p, C.c_void_p]
        L.orc_synth_rows_f32.argtypes = [C.c_uint64, C.c_uint64, C.c_uint64, C.c_uint32, C.c_void_p]
        L.orc_score_all_i8_limbs.argtypes = [C.c_void_p, C.c_void_p, C.c_uint64, C.c_uint32, C.c_void_p, C.c_void_p,
                                             C.c_void_p, C.c_void_p]
        L.orc_query_limbs_f16.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_float)]
        L.orc_query_limbs_i8.argtypes = [C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p, C.POINTER(C.c_float),
                                         C.POINTER(C.c_float)]
        L.orc_f16_to_f32.restype = C.c_float
        L.orc_f16_to_f32.argtypes = [C.c_uint16]
        L.orc_f32_to_f16_rne.restype = C.c_uint16
        L.orc_f32_to_f16_rne.argtypes = [C.c_float]

    # --- arithmetic
    def dot(self, dtype, q, x, scale=1.0):
        q = np.ascontiguousarray(q, np.float32)
        x = np.ascontiguousarray(x, NP_DT[dtype])
        if dtype == F32:
            return np.float32(self.lib.orc_dot_f32(_p(q), _p(x), q.size))
        if dtype == F16:
            return np.float32(self.lib.orc_dot_f16(_p(q), _p(x), q.size))
        return np.float32(self.lib.orc_dot_i8(_p(q), _p(x), q.size, C.c_float(scale)))

    def score_all(self, dtype, base, scales, q):
        n, d = base.shape
        out = np.empty(n, np.float32)
        q = np.ascontiguousarray(q, np.float32)
        self.lib.orc_score_all(_p(base), _p(scales), dtype, n, d, _p(q), _p(out))
        return out

    def search(self, dtype, base, scales, q, k, mt=False):
        n, d = base.shape
        kk = min(k, n)
        ids = np.empty(max(kk, 1), np.uint64)
        sc = np.empty(max(kk, 1), np.float32)
        q = np.ascontiguousarray(q, np.float32)
        f = self.lib.orc_flat_search_mt if mt else self.lib.orc_flat_search
        r = f(_p(base), _p(scales), dtype, n, d, _p(q), k, _p(ids), _p(sc))
        return ids[:r], sc[:r]

    def search_batch(self, dtype, base, scales, queries, k):
        n, d = base.shape
        queries = np.ascontiguousarray(queries, np.float32)
        nq = queries.shape[0]
        ids = np.empty((nq, k), np.uint64)
        sc = np.empty((nq, k), np.float32)
        self.lib.orc_flat_search_batch(_p(base), _p(scales), dtype, n, d, _p(queries), nq, k, _p(ids), _p(sc))
        return ids, sc

    # --- tensor-core scan limb schemes (restatement of csrc/scan_mma.cu)
    def score_all_i8_limbs(self, base, scales, q):
        n, d = base.shape
        q = np.ascontiguousarray(q, np.float32)
        sc, i1, i2 = np.empty(n, np.float32), np.empty(n, np.int32), np.empty(n, np.int32)
        self.lib.orc_score_all_i8_limbs(_p(base), _p(scales), n, d, _p(q), _p(sc), _p(i1), _p(i2))
        return sc, i1, i2

    def query_limbs_f16(self, q):
        q = np.ascontiguousarray(q, np.float32)
        hi, lo = np.empty(q.size, np.uint16), np.empty(q.size, np.uint16)
        s = C.c_float()
        self.lib.orc_query_limbs_f16(_p(q), q.size, _p(hi), _p(lo), C.byref(s))
        return hi, lo, np.float32(s.value)

    def query_limbs_i8(self, q):
        q = np.ascontiguousarray(q, np.float32)
        q1, q2 = np.empty(q.size, np.int8), np.empty(q.size, np.int8)
        s1, s2 = C.c_float(), C.c_float()
        self.lib.orc_query_limbs_i8(_p(q), q.size, _p(q1), _p(q2), C.byref(s1), C.byref(s2))
        return q1, q2, np.float32(s1.value), np.float32(s2.value)

    # --- converters
    def to_f16(self, x):
        x = np.ascontiguousarray(x, np.float32)
        out = np.empty(x.shape, np.uint16)
        self.lib.orc_convert_f16(_p(x), x.size, _p(out))
        return out

    def quantize_i8(self, rows):
        rows = np.ascontiguousarray(rows, np.float32)
        n, d = rows.shape
        out = np.empty((n, d), np.int8)
        sc = np.empty(n, np.float32)
        self.lib.orc_quantize_i8(_p(rows), n, d, _p(out), _p(sc))
        return out, sc

    # --- refine
    def refine(self, mode, dtype, base, scales, queries, cand, K):
        n, d = base.shape
        queries = np.ascontiguousarray(queries, np.float32)
        cand = np.ascontiguousarray(cand, np.uint32)
        Q, R = cand.shape
        ids = np.empty((Q, K), np.uint32)
        dist = np.empty((Q, K), np.float32)
        self.lib.orc_refine_l2(mode, _p(base), _p(scales), dtype, n, d, _p(queries), _p(cand), Q, R, K, _p(ids),
                               _p(dist))
        return ids, dist

    # --- synthetic data
    def synth_f32(self, seed, row0, nrows, dim):
        out = np.empty((nrows, dim), np.float32)
        self.lib.orc_synth_rows_f32(seed, row0, nrows, dim, _p(out))
        return out

    def synth(self, dtype, seed, row0, nrows, dim):
        """-> (payload ndarray, scales or None) with the reference converters' semantics applied to the f32 rows."""
        f = self.synth_f32(seed, row0, nrows, dim)
        if dtype == F32:
            return f, None
        if dtype == F16:
            return self.to_f16(f), None
        return self.quantize_i8(f)


class Reference:
    """The unmodified reference, compiled by oracle/Makefile `ref` (present only if it was built)."""

    def __init__(self):
        if not build_ref():
            raise FileNotFoundError("oracle/_ref/libnvdb_ref.so not built and /root/reference absent")
        # libgomp must be loadable; same image on every box
        self.lib = C.CDLL(str(REF_SO))
        L = self.lib
        L.ref_dot_f32.restype = C.c_float
        L.ref_dot_f16.restype = C.c_float
        L.ref_dot_i8.restype = C.c_float
        L.ref_dot_f32.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.ref_dot_f16.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32]
        L.ref_dot_i8.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_float]
        L.ref_dataset_load.restype = C.c_void_p
        L.ref_dataset_load.argtypes = [C.c_char_p]
        L.ref_dataset_free.argtypes = [C.c_void_p]
        L.ref_dataset_count.restype = C.c_uint64
        L.ref_dataset_count.argtypes = [C.c_void_p]
        L.ref_dataset_dim.restype = C.c_uint32
        L.ref_dataset_dim.argtypes = [C.c_void_p]
        L.ref_dataset_dtype.restype = C.c_uint32
        L.ref_dataset_dtype.argtypes = [C.c_void_p]
        L.ref_search.restype = C.c_int64
        L.ref_search.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p, C.c_uint32, C.c_void_p, C.c_void_p]
        L.ref_last_error.restype = C.c_char_p

    def dot(self, dtype, q, x, scale=1.0):
        q = np.ascontiguousarray(q, np.float32)
        x = np.ascontiguousarray(x, NP_DT[dtype])
        if dtype == F32:
            return np.float32(self.lib.ref_dot_f32(_p(q), _p(x), q.size))
        if dtype == F16:
            return np.float32(self.lib.ref_dot_f16(_p(q), _p(x), q.size))
        return np.float32(self.lib.ref_dot_i8(_p(q), _p(x), q.size, C.c_float(scale)))

    def load(self, path):
        h = self.lib.ref_dataset_load(str(path).encode())
        if not h:
            raise RuntimeError(self.lib.ref_last_error().decode())
        return h

    def free(self, h):
        self.lib.ref_dataset_free(h)

    MODES = {"st": 0, "omp": 1, "async": 2, "pool": 3}

    def search(self, h, q, k, mode="st", threads=0):
        q = np.ascontiguousarray(q, np.float32)
        ids = np.empty(max(k, 1), np.uint64)
        sc = np.empty(max(k, 1), np.float32)
        threads = threads or os.cpu_count()
        r = self.lib.ref_search(h, self.MODES[mode], threads, _p(q), k, _p(ids), _p(sc))
        if r < 0:
            raise RuntimeError(self.lib.ref_last_error().decode())
        return ids[:r], sc[:r]


# ---------------------------------------------------------------- vecbin / raw12 / gtbin files
# byte layout = reference include/nvdb/vecbin_format.h:17-29 (same struct.pack("<QIIIIQ") as scripts/build_vecbin_chunked.py:122)
VEC_MAGIC = 0x4E56444256454331
GT_MAGIC = 0x4E56444247543031


def write_vecbin(path, dtype, payload: np.ndarray, scales: np.ndarray | None = None):
    n, d = payload.shape
    hdr = struct.pack("<QIIIIQ", VEC_MAGIC, 1, dtype, d, 0, n)
    hdr += b"\0" * (64 - len(hdr))
    with open(path, "wb") as f:
        f.write(hdr)
        f.write(np.ascontiguousarray(payload, NP_DT[dtype]).tobytes())
        if dtype == I8:
            f.write(np.ascontiguousarray(scales, np.float32).tobytes())


def write_raw12(path, payload: np.ndarray):
    n, d = payload.shape
    with open(path, "wb") as f:
        f.write(struct.pack("<III", n, 0, d))
        f.write(np.ascontiguousarray(payload, np.float32).tobytes())


def read_vecbin(path):
    raw = np.fromfile(path, np.uint8)
    magic, ver, dtype, d, _, n = struct.unpack("<QIIIIQ", raw[:32].tobytes())
    if magic == VEC_MAGIC:
        esz = {F32: 4, F16: 2, I8: 1}[dtype]
        payload = raw[64:64 + n * d * esz].view(NP_DT[dtype]).reshape(n, d)
        scales = raw[64 + n * d * esz:].view(np.float32) if dtype == I8 else None
        return dtype, payload, scales
    n, _, d = struct.unpack("<III", raw[:12].tobytes())
    return F32, raw[12:].view(np.float32).reshape(n, d), None


def read_gtbin(path):
    raw = np.fromfile(path, np.uint8)
    magic, ver, metric, k, dim, Q, N = struct.unpack("<QIIIIQQ", raw[:40].tobytes())
    assert magic == GT_MAGIC
    return raw[64:].view(np.uint32).reshape(Q, k), dict(k=k, dim=dim, Q=Q, N=N, metric=metric)


def run_ref_tool(name, *args, env=None):
    e = dict(os.environ)
    if env:
        e.update(env)
    return subprocess.run([str(REF_DIR / name), *map(str, args)], check=True, capture_output=True, text=True, env=e)


# ---------------------------------------------------------------- tie-aware comparator
def assert_topk_equal(ids, scores, ref_ids, ref_scores, *, rtol=0.0, what=""):
    """ids must match position-by-position except inside runs where the reference scores are exactly equal
    (rtol=0) or within rtol of each other (tolerance-mode paths; returns the list of such near-tie positions)."""
    ids = np.asarray(ids).astype(np.uint64)
    ref_ids = np.asarray(ref_ids).astype(np.uint64)
    scores = np.asarray(scores, np.float32)
    ref_scores = np.asarray(ref_scores, np.float32)
    assert ids.shape == ref_ids.shape, f"{what}: shape {ids.shape} vs {ref_ids.shape}"
    if rtol == 0.0:
        assert np.array_equal(scores.view(np.uint32), ref_scores.view(np.uint32)), \
            f"{what}: score bits differ\n{scores}\n{ref_scores}"
    else:
        np.testing.assert_allclose(scores, ref_scores, rtol=rtol, atol=rtol * 1e-2, err_msg=what)
    near = []
    bad = np.nonzero(ids != ref_ids)[0]
    for i in bad:
        s = ref_scores[i]
        tol = abs(s) * rtol
        grp = np.nonzero(np.abs(ref_scores - s) <= tol)[0]
        # the mismatching id must be a legitimate member of the tie group: same score (within tol)
        assert len(grp) > 1 or rtol > 0, f"{what}: id mismatch at {i} outside a tie group: {ids[i]} vs {ref_ids[i]}"
        assert abs(scores[i] - s) <= tol, f"{what}: id mismatch at {i} with different score"
        near.append(int(i))
    return near
