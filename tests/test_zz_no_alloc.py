"""nvdb_b200_search_device_static: the allocation-free variant of the device search (include/nvdb_b200.h; no
counterpart in the reference). Its scratch is a file-scope static
__device__ array inside the library, so the cases here are about what that changes: two calls in a row share the array
(the second must not see the first one's lists), the result is the ordinary entry point's bit for bit (integer ids,
FP32 score bits: the tolerance of every scan test in this suite is zero), a request beyond the static maximum is
refused, and device memory in use does not move across the calls.

Collected last on purpose; it passes alone in a fresh process and after the others.
"""
import numpy as np
import pytest

from _oracle import F16, F32, I8, assert_topk_equal

pytestmark = pytest.mark.gpu

UNSUPPORTED = 5     # NVDB_B200_ERR_UNSUPPORTED
STATIC_BYTES = 4 << 20


def _both(torch, idx, q_dev, nq, k, stream):
    dev = q_dev.device
    out = []
    for fn in (idx.search_device_static, idx.search_device):
        ids = torch.full((nq, k), -1, dtype=torch.int64, device=dev)
        sc = torch.full((nq, k), float("nan"), dtype=torch.float32, device=dev)
        fn(q_dev.data_ptr(), nq, k, ids.data_ptr(), sc.data_ptr(), stream.cuda_stream)
        stream.synchronize()
        out.append((ids.cpu().numpy().view(np.uint64), sc.cpu().numpy()))
    return out


@pytest.mark.parametrize("dt", ["f32", "f16", "i8"])
def test_static_scratch_matches_search_device_on_two_calls_in_a_row(nv, oracle, dt):
    torch = pytest.importorskip("torch")
    dtype = {"f32": F32, "f16": F16, "i8": I8}[dt]
    n, d = 40000, 128
    base, scales = oracle.synth(dtype, 33, 0, n, d)
    f = oracle.synth_f32(33, 0, n, d)
    rng = np.random.default_rng(7)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    with nv.FlatIndexCuda.from_host(base, scales, dtype) as idx:
        # the first call fills most of the array (8 queries x 148+ lists x 100 keys), the second reads a small corner of
        # it with other queries and another k: stale lists of call 1 must not leak into call 2
        for nq, k in [(8, 100), (1, 10), (5, 7)]:
            q = np.ascontiguousarray(f[rng.integers(0, n, nq)] + (0.05 * rng.standard_normal((nq, d))).astype(np.float32))
            q_dev = torch.from_numpy(q).to(dev)
            torch.cuda.synchronize(dev)
            (ids, sc), (rids, rsc) = _both(torch, idx, q_dev, nq, k, stream)
            assert np.array_equal(ids, rids), f"{dt} nq={nq} k={k}: ids differ from nvdb_b200_search_device"
            assert np.array_equal(sc.view(np.uint32), rsc.view(np.uint32)), f"{dt} nq={nq} k={k}: score bits differ"
            oids, osc = oracle.search_batch(dtype, base, scales, q, k)
            for qi in range(nq):
                assert_topk_equal(ids[qi], sc[qi], oids[qi], osc[qi], what=f"{dt} nq={nq} k={k} q{qi}")
            # a second identical call over the now dirty array, then the allocation check around static calls only
            ids2 = torch.empty((nq, k), dtype=torch.int64, device=dev)
            sc2 = torch.empty((nq, k), dtype=torch.float32, device=dev)
            torch.cuda.synchronize(dev)
            free1 = torch.cuda.mem_get_info(dev)[0]
            idx.search_device_static(q_dev.data_ptr(), nq, k, ids2.data_ptr(), sc2.data_ptr(), stream.cuda_stream)
            idx.search_device_static(q_dev.data_ptr(), nq, k, ids2.data_ptr(), sc2.data_ptr(), stream.cuda_stream)
            stream.synchronize()
            assert torch.cuda.mem_get_info(dev)[0] == free1, "device memory in use moved across the static calls"
            assert np.array_equal(ids2.cpu().numpy().view(np.uint64), rids)
            assert np.array_equal(sc2.cpu().numpy().view(np.uint32), rsc.view(np.uint32))


def test_static_scratch_rejects_requests_beyond_the_maximum(nv, oracle):
    torch = pytest.importorskip("torch")
    n, d = 40000, 64
    base, scales = oracle.synth(F16, 5, 0, n, d)
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)
    with nv.FlatIndexCuda.from_host(base, scales, F16) as idx:
        nq, k = 64, 100      # 64 queries x >= 148 lists x 100 keys x 8 bytes = 7.6 MB > 4 MiB
        assert nq * 148 * k * 8 > STATIC_BYTES
        q_dev = torch.randn((nq, d), dtype=torch.float32, device=dev)
        ids = torch.zeros((nq, k), dtype=torch.int64, device=dev)
        sc = torch.zeros((nq, k), dtype=torch.float32, device=dev)
        with pytest.raises(nv.NvdbError) as ei:
            idx.search_device_static(q_dev.data_ptr(), nq, k, ids.data_ptr(), sc.data_ptr(), stream.cuda_stream)
        assert ei.value.status == UNSUPPORTED
        assert "static maximum" in ei.value.msg and str(STATIC_BYTES) in ei.value.msg
        # the ordinary entry point takes the same request, and a smaller static one still works afterwards
        idx.search_device(q_dev.data_ptr(), nq, k, ids.data_ptr(), sc.data_ptr(), stream.cuda_stream)
        idx.search_device_static(q_dev.data_ptr(), 2, 10, ids.data_ptr(), sc.data_ptr(), stream.cuda_stream)
        stream.synchronize()
