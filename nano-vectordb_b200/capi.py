"""ctypes binding of include/nvdb_b200.h. No compute here; every call lands in the CUDA library.

`FlatIndexCuda` mirrors the reference's index classes (reference include/nvdb/flat_index.h:9-20):
`search_topk_dot(q, k)` -> (ids, scores) sorted by score descending; plus the bench-side batched search
(reference apps/nvdb_bench.cpp:47-159) as `search_topk_dot_batch`, and `refine_l2` for
nvdb::cuda_l2_topk_batch (reference include/nvdb/cuda_refine.h:25-38).
"""
from __future__ import annotations

import ctypes as C
import os
from pathlib import Path

import numpy as np

F32, F16, I8 = 1, 2, 3
PATH_AUTO, PATH_GEMV, PATH_MMA, PATH_MMA_FAST, PATH_MMA_EXACT = 0, 1, 2, 3, 4
TUNE_MMA_SEED, TUNE_GEMV_CTAS, TUNE_MMA_2CTA, TUNE_FUSED_MERGE, TUNE_MMA_RESIDENT, TUNE_EXACT_TIER2 = 1, 2, 3, 4, 5, 6
_NP = {F32: np.float32, F16: np.uint16, I8: np.int8}

lib_path = Path(os.environ.get("NVDB_B200_LIB", Path(__file__).resolve().parent / "lib" / "libnvdb_b200.so"))
if not lib_path.exists():
    raise ImportError(f"{lib_path} is not built (run `make lib` or __graft_entry__.build()); "
                      "there is no CPU fallback for this path")
lib = C.CDLL(str(lib_path))


class Timing(C.Structure):
    _fields_ = [("h2d_ms", C.c_float), ("kernel_ms", C.c_float), ("d2h_ms", C.c_float), ("total_ms", C.c_float),
                ("kernel_launches", C.c_uint32), ("path", C.c_uint32), ("threads", C.c_uint32), ("shmem_bytes", C.c_uint32),
                ("dbg_q", C.c_uint32), ("dbg_dist_cycles_avg", C.c_float), ("dbg_write_cycles_avg", C.c_float),
                ("dbg_merge_cycles_avg", C.c_float)]


class NvdbError(RuntimeError):
    def __init__(self, status, msg):
        super().__init__(f"nvdb_b200 status {status}: {msg}")
        self.status = status
        self.msg = msg


_vp, _u32, _u64, _i = C.c_void_p, C.c_uint32, C.c_uint64, C.c_int
_sig = {
    "nvdb_b200_version": (C.c_int, []),
    "nvdb_b200_status_string": (C.c_char_p, [_i]),
    "nvdb_b200_last_error": (C.c_char_p, [_vp]),
    "nvdb_b200_device_count": (_i, [C.POINTER(_i)]),
    "nvdb_b200_index_create": (_i, [_vp, _vp, _u32, _u64, _u32, _i, _u64, C.POINTER(_vp)]),
    "nvdb_b200_index_create_device": (_i, [_vp, _vp, _u32, _u64, _u32, _i, _u64, C.POINTER(_vp)]),
    "nvdb_b200_index_create_synthetic": (_i, [_u32, _u64, _u32, _u64, _u64, _i, _u64, C.POINTER(_vp)]),
    "nvdb_b200_index_destroy": (_i, [_vp]),
    "nvdb_b200_index_info": (_i, [_vp, C.POINTER(_u64), C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_i),
                                  C.POINTER(_u64), C.POINTER(_vp), C.POINTER(_vp)]),
    "nvdb_b200_index_upload_info": (_i, [_vp, C.POINTER(C.c_double), C.POINTER(_u64)]),
    "nvdb_b200_index_set_path": (_i, [_vp, _i]),
    "nvdb_b200_index_read_rows": (_i, [_vp, _u64, _u64, _vp, _vp]),
    "nvdb_b200_search": (_i, [_vp, _vp, _u32, _u32, _vp, _vp, C.POINTER(Timing)]),
    "nvdb_b200_search_device": (_i, [_vp, _vp, _u32, _u32, _vp, _vp, _vp]),
    "nvdb_b200_search_device_static": (_i, [_vp, _vp, _u32, _u32, _vp, _vp, _vp]),
    "nvdb_b200_search_two_stage": (_i, [_vp, _vp, _vp, _u32, _u32, _u32, _vp, _vp, C.POINTER(Timing)]),
    "nvdb_b200_debug_mma_accumulators": (_i, [_vp, _vp, _u32, _i, _vp]),
    "nvdb_b200_index_set_tuning": (_i, [_vp, _i, _i]),
    "nvdb_b200_index_exact_stats": (_i, [_vp, C.POINTER(_u64), C.POINTER(_u64), _i]),
    "nvdb_b200_index_exact_tiers": (_i, [_vp, C.POINTER(_u64), C.POINTER(_u64)]),
    "nvdb_b200_merge_topk_device": (_i, [_i, _vp, _vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "nvdb_b200_merge_topk_packed_device": (_i, [_i, _vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "nvdb_b200_refine_l2": (_i, [_vp, _vp, _vp, _u32, _u32, _u32, _vp, _vp, C.POINTER(Timing)]),
    "nvdb_b200_refine_l2_device": (_i, [_vp, _vp, _vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "nvdb_b200_host_register": (_i, [_vp, C.c_size_t]),
    "nvdb_b200_host_unregister": (_i, [_vp]),
    "nvdb_b200_synth_fill_device": (_i, [_i, _u32, _u64, _u32, _u64, _u64, _vp, _vp, _vp]),
    "nvdb_b200_convert": (_i, [_i, _u32, _vp, _u32, _u64, _u32, _vp, _vp]),
    "nvdb_b200_convert_device": (_i, [_i, _u32, _vp, _u32, _u64, _u32, _vp, _vp, _vp]),
    "nvdb_b200_ivfpq_build": (_i, [_vp, _u32, _u32, _u32, _u64, _i, _u64, C.POINTER(_vp)]),
    "nvdb_b200_ivfpq_build_opq": (_i, [_vp, _u32, _u32, _u32, _u64, _i, _i, _u64, C.POINTER(_vp)]),
    "nvdb_b200_ivfpq_rotation": (_i, [_vp, _vp, C.POINTER(_i)]),
    "nvdb_b200_ivfpq_set_rotation": (_i, [_vp, _vp]),
    "nvdb_b200_ivfpq_destroy": (_i, [_vp]),
    "nvdb_b200_ivfpq_last_error": (C.c_char_p, [_vp]),
    "nvdb_b200_ivfpq_info": (_i, [_vp, C.POINTER(_u64), C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u32), C.POINTER(_u64)]),
    "nvdb_b200_ivfpq_export": (_i, [_vp, _vp, _vp, _vp, _vp, _vp]),
    "nvdb_b200_ivfpq_import": (_i, [_i, _u32, _u32, _u32, _u64, _vp, _vp, _vp, _vp, _vp, C.POINTER(_vp)]),
    "nvdb_b200_ivfpq_search": (_i, [_vp, _vp, _u32, _u32, _u32, _vp, _vp]),
    "nvdb_b200_ivfpq_search_device": (_i, [_vp, _vp, _u32, _u32, _u32, _vp, _vp, _vp]),
    "nvdb_b200_ivfpq_search_refine": (_i, [_vp, _vp, _vp, _u32, _u32, _u32, _u32, _vp, _vp]),
    "nvdb_b200_exchange_create": (_i, [_i, _u32, _u32, _u64, C.POINTER(_vp)]),
    "nvdb_b200_exchange_ipc_handle": (_i, [_vp, _vp]),
    "nvdb_b200_exchange_window": (_i, [_vp, C.POINTER(_vp), C.POINTER(_u64)]),
    "nvdb_b200_exchange_connect": (_i, [_vp, _vp, _vp]),
    "nvdb_b200_exchange_merge": (_i, [_vp, _vp, _u32, _u32, _vp, _vp, _i, _vp]),
    "nvdb_b200_exchange_status": (_i, [_vp, C.POINTER(_u32)]),
    "nvdb_b200_exchange_reset": (_i, [_vp]),
    "nvdb_b200_exchange_last_error": (C.c_char_p, [_vp]),
    "nvdb_b200_exchange_destroy": (_i, [_vp]),
    "nvdb_b200_search_exchange_device": (_i, [_vp, _vp, _vp, _u32, _u32, _vp, _vp, _vp]),
    "nvdb_b200_search_exchange": (_i, [_vp, _vp, _vp, _u32, _u32, _vp, _vp, C.POINTER(Timing)]),
    "nvdb_b200_search_exchange_split": (_i, [_vp, _vp, _vp, _u32, _u32, _vp, _vp, _i]),
    "nvdb_b200_launch_count": (_u64, []),
    "nvdb_b200_index_profile": (_i, [_vp, _i]),
    "nvdb_b200_index_profile_read": (_i, [_vp, C.POINTER(C.c_double), C.POINTER(C.c_double), C.POINTER(_u32)]),
}
for _name, (_res, _args) in _sig.items():
    _f = getattr(lib, _name)
    _f.restype = _res
    _f.argtypes = _args
EXPORTED_SYMBOLS = tuple(_sig)


def _check(status, handle=None):
    if status != 0:
        msg = lib.nvdb_b200_last_error(handle)
        raise NvdbError(status, (msg or b"").decode() or lib.nvdb_b200_status_string(status).decode())


def _ptr(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


def launch_count() -> int:
    return int(lib.nvdb_b200_launch_count())


def synth_fill_device(device, dtype, n, dim, seed, first_row, payload_ptr, scales_ptr=None, stream=None):
    _check(lib.nvdb_b200_synth_fill_device(device, dtype, n, dim, seed, first_row, payload_ptr, scales_ptr, stream))


def merge_topk_device(device, ids_ptr, scores_ptr, nlists, nq, k, out_ids_ptr, out_scores_ptr, stream=None):
    _check(lib.nvdb_b200_merge_topk_device(device, ids_ptr, scores_ptr, nlists, nq, k, out_ids_ptr, out_scores_ptr,
                                           stream))


def merge_topk_packed_device(device, packed_ptr, nlists, nq, k, out_ids_ptr, out_scores_ptr, stream=None):
    _check(lib.nvdb_b200_merge_topk_packed_device(device, packed_ptr, nlists, nq, k, out_ids_ptr, out_scores_ptr, stream))


IPC_HANDLE_BYTES = 64
XCHG_PUSH, XCHG_MERGE, XCHG_FUSED = 1, 2, 3


class PeerExchange:
    """One rank's end of the NVLink peer-memory exchange (include/nvdb_b200.h, csrc/exchange.cu): the window its
    peers store their local top-k lists into, and the fused push + wait + merge kernel. `slot_bytes` bounds the packed
    result of one call: 8 * (nq*k + (nq*k+1)//2)."""

    def __init__(self, device: int, world: int, rank: int, slot_bytes: int):
        h = _vp()
        st = lib.nvdb_b200_exchange_create(device, world, rank, slot_bytes, C.byref(h))
        if st != 0:
            raise NvdbError(st, (lib.nvdb_b200_exchange_last_error(None) or b"").decode())
        self._h = h
        self.device, self.world, self.rank, self.slot_bytes = device, world, rank, slot_bytes

    def _ck(self, st):
        if st != 0:
            raise NvdbError(st, (lib.nvdb_b200_exchange_last_error(self._h) or b"").decode())

    def ipc_handle(self) -> bytes:
        buf = (C.c_ubyte * IPC_HANDLE_BYTES)()
        self._ck(lib.nvdb_b200_exchange_ipc_handle(self._h, buf))
        return bytes(buf)

    def window(self) -> int:
        p, n = _vp(), _u64()
        self._ck(lib.nvdb_b200_exchange_window(self._h, C.byref(p), C.byref(n)))
        return int(p.value)

    def connect(self, handles: bytes | None = None, windows=None):
        """`handles`: world * 64 bytes, rank-major (peers in other processes); `windows`: world device pointers (peers
        in this process)."""
        if windows is not None:
            arr = (_vp * self.world)(*[_vp(w) for w in windows])
            self._ck(lib.nvdb_b200_exchange_connect(self._h, None, arr))
        else:
            if handles is not None and len(handles) != self.world * IPC_HANDLE_BYTES:
                raise ValueError("handles must hold world * 64 bytes")
            buf = None if handles is None else (C.c_ubyte * len(handles)).from_buffer_copy(handles)
            self._ck(lib.nvdb_b200_exchange_connect(self._h, buf, None))

    def merge(self, local_packed_ptr, nq, k, out_ids_ptr, out_scores_ptr, phases=XCHG_FUSED, stream=None):
        self._ck(lib.nvdb_b200_exchange_merge(self._h, local_packed_ptr, nq, k, out_ids_ptr, out_scores_ptr, phases,
                                              stream))

    def search(self, index, queries_ptr, nq, k, out_ids_ptr, out_scores_ptr, stream=None):
        """One rank's whole step: local scan of `index` (a FlatIndexCuda, or None for an empty shard) + exchange + merge
        (nvdb_b200_search_exchange_device); one launch for single-pass streaming scans."""
        ih = None if index is None else index._h
        st = lib.nvdb_b200_search_exchange_device(ih, self._h, queries_ptr, nq, k, out_ids_ptr, out_scores_ptr, stream)
        if st != 0:
            raise NvdbError(st, (lib.nvdb_b200_last_error(ih) or lib.nvdb_b200_exchange_last_error(self._h) or b"").decode())

    def status(self) -> int:
        s = _u32()
        self._ck(lib.nvdb_b200_exchange_status(self._h, C.byref(s)))
        return int(s.value)

    def reset(self):
        """collective recovery after a timed-out exchange (every rank, with a barrier before the next search)"""
        self._ck(lib.nvdb_b200_exchange_reset(self._h))

    def close(self):
        if self._h:
            lib.nvdb_b200_exchange_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


class FlatIndexCuda:
    """Device-resident flat index. Host arrays in, host arrays out (the *_device methods take raw pointers)."""

    def __init__(self, handle, keepalive=None):
        self._h = handle
        self._keep = keepalive
        n, dim, dt, dev, off = _u64(), _u32(), _u32(), _i(), _u64()
        pp, sp = _vp(), _vp()
        _check(lib.nvdb_b200_index_info(self._h, C.byref(n), C.byref(dim), C.byref(dt), C.byref(dev), C.byref(off),
                                        C.byref(pp), C.byref(sp)))
        self.count, self.dim, self.dtype, self.device, self.row_offset = n.value, dim.value, dt.value, dev.value, off.value
        self.payload_ptr, self.scales_ptr = pp.value, sp.value

    # ---- constructors
    @classmethod
    def from_host(cls, payload: np.ndarray, scales: np.ndarray | None, dtype: int, device=0, row_offset=0):
        payload = np.ascontiguousarray(payload, _NP[dtype])
        n, dim = payload.shape
        if scales is not None:
            scales = np.ascontiguousarray(scales, np.float32)
        h = _vp()
        _check(lib.nvdb_b200_index_create(_ptr(payload), _ptr(scales), dtype, n, dim, device, row_offset, C.byref(h)))
        return cls(h)

    @classmethod
    def from_device(cls, payload_ptr, scales_ptr, dtype, n, dim, device=0, row_offset=0, keepalive=None):
        h = _vp()
        _check(lib.nvdb_b200_index_create_device(payload_ptr, scales_ptr, dtype, n, dim, device, row_offset, C.byref(h)))
        return cls(h, keepalive)

    @classmethod
    def synthetic(cls, dtype, n, dim, seed, first_row=0, device=0, row_offset=0):
        h = _vp()
        _check(lib.nvdb_b200_index_create_synthetic(dtype, n, dim, seed, first_row, device, row_offset, C.byref(h)))
        return cls(h)

    def close(self):
        if self._h:
            lib.nvdb_b200_index_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def upload_info(self):
        """(seconds, bytes) of the host -> device upload done at creation (0, 0 for device-resident / synthetic bases)"""
        sec, nb = C.c_double(), _u64()
        _check(lib.nvdb_b200_index_upload_info(self._h, C.byref(sec), C.byref(nb)), self._h)
        return sec.value, nb.value

    def set_path(self, path):
        _check(lib.nvdb_b200_index_set_path(self._h, path), self._h)

    def set_tuning(self, knob, value):
        _check(lib.nvdb_b200_index_set_tuning(self._h, knob, value), self._h)

    def exact_stats(self, reset=False):
        """(queries answered by PATH_MMA_EXACT, how many failed the certificate and were re-run on the streaming scan)"""
        q, f = C.c_uint64(0), C.c_uint64(0)
        _check(lib.nvdb_b200_index_exact_stats(self._h, C.byref(q), C.byref(f), int(reset)), self._h)
        return q.value, f.value

    def exact_tiers(self):
        """of the certificate failures: (re-run on the tensor-core scan with wider lists, answered by the streaming scan)"""
        t, g = C.c_uint64(0), C.c_uint64(0)
        _check(lib.nvdb_b200_index_exact_tiers(self._h, C.byref(t), C.byref(g)), self._h)
        return t.value, g.value

    def read_rows(self, row0, nrows):
        payload = np.empty((nrows, self.dim), _NP[self.dtype])
        scales = np.empty(nrows, np.float32) if self.dtype == I8 else None
        _check(lib.nvdb_b200_index_read_rows(self._h, row0, nrows, _ptr(payload), _ptr(scales)), self._h)
        return payload, scales

    def debug_mma_accumulators(self, queries, path=PATH_MMA):
        """raw tensor-core accumulators: f16 -> float32 [nq, n]; i8 -> int32 [2, nq, n] (exact limb dot products)"""
        queries = np.ascontiguousarray(queries, np.float32)
        nq = queries.shape[0]
        nacc = 2 if self.dtype == I8 else 1
        out = np.zeros((nacc, nq, self.count), np.uint32)
        _check(lib.nvdb_b200_debug_mma_accumulators(self._h, _ptr(queries), nq, path, _ptr(out)), self._h)
        return out.view(np.int32) if self.dtype == I8 else out[0].view(np.float32)

    def profile(self, enable=True):
        _check(lib.nvdb_b200_index_profile(self._h, int(enable)), self._h)

    def profile_read(self):
        a, b, n = C.c_double(), C.c_double(), _u32()
        _check(lib.nvdb_b200_index_profile_read(self._h, C.byref(a), C.byref(b), C.byref(n)), self._h)
        return a.value, b.value, n.value

    # ---- search (mirrors FlatIndex::search_topk_dot / batched_scan_*)
    def search_topk_dot(self, q, k, timing: Timing | None = None):
        ids, sc = self.search_topk_dot_batch(np.asarray(q, np.float32).reshape(1, -1), k, timing)
        kk = min(k, self.count)  # the reference clamps k to n (src/flat_index.cpp:24)
        return ids[0, :kk], sc[0, :kk]

    def search_topk_dot_batch(self, queries, k, timing: Timing | None = None):
        if queries is None:
            raise NvdbError(1, "Null query")
        queries = np.ascontiguousarray(queries, np.float32)
        if queries.ndim != 2 or queries.shape[1] != self.dim:
            raise NvdbError(1, f"Dim mismatch: base.dim={self.dim}, query.dim={queries.shape[-1]}")
        nq = queries.shape[0]
        ids = np.empty((nq, k), np.uint64)
        sc = np.empty((nq, k), np.float32)
        _check(lib.nvdb_b200_search(self._h, _ptr(queries), nq, k, _ptr(ids), _ptr(sc),
                                    C.byref(timing) if timing is not None else None), self._h)
        return ids, sc

    def search_device(self, queries_ptr, nq, k, out_ids_ptr, out_scores_ptr, stream=None):
        _check(lib.nvdb_b200_search_device(self._h, queries_ptr, nq, k, out_ids_ptr, out_scores_ptr, stream), self._h)

    def search_device_static(self, queries_ptr, nq, k, out_ids_ptr, out_scores_ptr, stream=None):
        """Allocation-free streaming scan over the library's static device scratch (include/nvdb_b200.h)."""
        _check(lib.nvdb_b200_search_device_static(self._h, queries_ptr, nq, k, out_ids_ptr, out_scores_ptr, stream), self._h)

    # ---- refine (mirrors nvdb::cuda_l2_topk_batch)
    def refine_l2(self, queries, cand_ids, K, want_dist=True, timing: Timing | None = None):
        queries = np.ascontiguousarray(queries, np.float32)
        cand_ids = np.ascontiguousarray(cand_ids, np.uint32)
        Q, R = cand_ids.shape
        ids = np.full((Q, K), 0xFFFFFFFF, np.uint32)
        dist = np.full((Q, K), 1e30, np.float32) if want_dist else None
        _check(lib.nvdb_b200_refine_l2(self._h, _ptr(queries), _ptr(cand_ids), Q, R, K, _ptr(ids), _ptr(dist),
                                       C.byref(timing) if timing is not None else None), self._h)
        return ids, dist

    def refine_l2_device(self, queries_ptr, cand_ptr, Q, R, K, out_ids_ptr, out_dist_ptr, stream=None):
        _check(lib.nvdb_b200_refine_l2_device(self._h, queries_ptr, cand_ptr, Q, R, K, out_ids_ptr, out_dist_ptr,
                                              stream), self._h)


def convert(src: np.ndarray, src_dtype: int, dst_dtype: int, device: int = 0):
    """Reference converters on the device (nvdb_convert_f16 / nvdb_quantize_i8 semantics): returns (payload, scales)."""
    src = np.ascontiguousarray(src)
    n, dim = src.shape
    out = np.empty((n, dim), _NP[dst_dtype])
    scales = np.empty(n, np.float32) if dst_dtype == I8 else None
    _check(lib.nvdb_b200_convert(device, src_dtype, _ptr(src), dst_dtype, n, dim, _ptr(out), _ptr(scales)))
    return out, scales


def search_two_stage(coarse: "FlatIndexCuda", fine: "FlatIndexCuda", queries: np.ndarray, k: int, R: int, timing=None):
    """coarse top-R scan (e.g. the INT8 copy) + exact re-rank over `fine` (e.g. the FP16 copy): (ids u64, scores f32)"""
    queries = np.ascontiguousarray(queries, np.float32)
    nq = queries.shape[0]
    ids = np.empty((nq, k), np.uint64)
    sc = np.empty((nq, k), np.float32)
    _check(lib.nvdb_b200_search_two_stage(coarse._h, fine._h, _ptr(queries), nq, k, R, _ptr(ids), _ptr(sc),
                                          C.byref(timing) if timing is not None else None), coarse._h)
    return ids, sc


class IvfPqIndex:
    """IVF-PQ candidate generator over a device-resident base (reference: FAISS IndexIVFPQ as used by
    apps/nvdb_ivfpq_build.cpp:78-135 and apps/nvdb_ivf_eval.cpp:398-408,482). Parity unpinned (no FAISS here)."""

    def __init__(self, base: "FlatIndexCuda", nlist=4096, pq_m=48, pq_bits=8, ntrain=200_000, niter=25, seed=1234,
                 opq_niter=0, _handle=None):
        h = _handle
        if h is None:
            h = _vp()
            st = lib.nvdb_b200_ivfpq_build_opq(base._h, nlist, pq_m, pq_bits, ntrain, niter, opq_niter, seed, C.byref(h))
            if st != 0:
                raise NvdbError(st, (lib.nvdb_b200_ivfpq_last_error(None) or b"").decode())
        self._h = h
        self.base = base
        self._fill_info()

    @classmethod
    def from_arrays(cls, base, coarse, codebooks, list_off, ids, codes, device=0, rotation=None):
        """index from exported / externally trained parameters (nvdb_b200_ivfpq_import); `base` is only used by search_refine"""
        coarse = np.ascontiguousarray(coarse, np.float32)
        codebooks = np.ascontiguousarray(codebooks, np.float32)
        list_off = np.ascontiguousarray(list_off, np.uint32)
        ids = np.ascontiguousarray(ids, np.uint32)
        codes = np.ascontiguousarray(codes, np.uint8)
        nlist, dim = coarse.shape
        h = _vp()
        st = lib.nvdb_b200_ivfpq_import(device, dim, nlist, codebooks.shape[0], len(ids), _ptr(coarse), _ptr(codebooks),
                                        _ptr(list_off), _ptr(ids), _ptr(codes), C.byref(h))
        if st != 0:
            raise NvdbError(st, (lib.nvdb_b200_ivfpq_last_error(None) or b"").decode())
        self = cls(base, _handle=h)
        if rotation is not None:
            self.set_rotation(rotation)
        return self

    def rotation(self):
        """the OPQ rotation R [d,d] (x' = x @ R) or None"""
        has = _i()
        self._ck(lib.nvdb_b200_ivfpq_rotation(self._h, None, C.byref(has)))
        if not has.value:
            return None
        R = np.empty((self.dim, self.dim), np.float32)
        self._ck(lib.nvdb_b200_ivfpq_rotation(self._h, _ptr(R), C.byref(has)))
        return R

    def set_rotation(self, R):
        R = None if R is None else np.ascontiguousarray(R, np.float32)
        self._ck(lib.nvdb_b200_ivfpq_set_rotation(self._h, _ptr(R)))

    def _fill_info(self):
        h = self._h
        n, dim, nl, m, nbytes = _u64(), _u32(), _u32(), _u32(), _u64()
        self._ck(lib.nvdb_b200_ivfpq_info(h, C.byref(n), C.byref(dim), C.byref(nl), C.byref(m), C.byref(nbytes)))
        self.n, self.dim, self.nlist, self.pq_m, self.index_bytes = n.value, dim.value, nl.value, m.value, nbytes.value

    def _ck(self, st):
        if st != 0:
            raise NvdbError(st, (lib.nvdb_b200_ivfpq_last_error(self._h) or b"").decode())

    def close(self):
        if self._h:
            lib.nvdb_b200_ivfpq_destroy(self._h)
            self._h = None

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()

    def export(self):
        """(coarse [nlist,d], codebooks [M,256,dsub], list_off [nlist+1], ids [n], codes [n,M]) as host arrays"""
        dsub = self.dim // self.pq_m
        coarse = np.empty((self.nlist, self.dim), np.float32)
        cb = np.empty((self.pq_m, 256, dsub), np.float32)
        off = np.empty(self.nlist + 1, np.uint32)
        ids = np.empty(self.n, np.uint32)
        codes = np.empty((self.n, self.pq_m), np.uint8)
        self._ck(lib.nvdb_b200_ivfpq_export(self._h, _ptr(coarse), _ptr(cb), _ptr(off), _ptr(ids), _ptr(codes)))
        return coarse, cb, off, ids, codes

    def search(self, queries: np.ndarray, nprobe: int, R: int, return_dist=True):
        """R candidates per query by ADC distance: (ids u32 [nq,R], dist f32 [nq,R]); empty slots 0xFFFFFFFF / 1e30"""
        queries = np.ascontiguousarray(queries, np.float32).reshape(-1, self.dim)
        nq = queries.shape[0]
        ids = np.empty((nq, R), np.uint32)
        dist = np.empty((nq, R), np.float32) if return_dist else None
        self._ck(lib.nvdb_b200_ivfpq_search(self._h, _ptr(queries), nq, nprobe, R, _ptr(ids), _ptr(dist)))
        return ids, dist

    def search_device(self, queries_ptr, nq, nprobe, R, cand_ids_ptr, cand_dist_ptr=None, stream=None):
        self._ck(lib.nvdb_b200_ivfpq_search_device(self._h, queries_ptr, nq, nprobe, R, cand_ids_ptr, cand_dist_ptr, stream))

    def search_refine(self, queries: np.ndarray, nprobe: int, R: int, k: int):
        """candidates + exact L2 refine on the device: (ids u32 [nq,k], squared L2 f32 [nq,k]) ascending"""
        queries = np.ascontiguousarray(queries, np.float32).reshape(-1, self.dim)
        nq = queries.shape[0]
        ids = np.empty((nq, k), np.uint32)
        dist = np.empty((nq, k), np.float32)
        self._ck(lib.nvdb_b200_ivfpq_search_refine(self._h, self.base._h, _ptr(queries), nq, nprobe, R, k, _ptr(ids), _ptr(dist)))
        return ids, dist
