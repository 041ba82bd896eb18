"""Parity at BASELINE.json's FULL sizes (K1 FP32 500,000 x 768; K2 FP16 and K3 INT8 2,903,054 x 768,
top-10). The base is generated on the device, read back once, and the CPU oracle (multi-threaded restatement of the
reference scan, bit-exact scores) scans the SAME bytes for a handful of queries; on top of that the size-independent
properties: the streaming scan and the exact tensor-core path agree bit for bit on a batch, results are sorted,
ids are unique and in range, a query that is a base row finds itself, and the merge of two row shards equals the scan
of the whole."""
import numpy as np
import pytest
import torch

from _oracle import F16, F32, I8, assert_topk_equal

pytestmark = pytest.mark.gpu
SEED, D, K = 20240229, 768, 10
CASES = {"K1_f32_500k": (F32, 500_000), "K2_f16_2.9M": (F16, 2_903_054), "K3_i8_2.9M": (I8, 2_903_054)}


def _read_all(idx, n, chunk=400_000):
    parts, sparts = [], []
    for r0 in range(0, n, chunk):
        p, s = idx.read_rows(r0, min(chunk, n - r0))
        parts.append(p)
        if s is not None:
            sparts.append(s)
    return np.concatenate(parts), (np.concatenate(sparts) if sparts else None)


@pytest.mark.parametrize("case", list(CASES))
def test_full_size_parity_and_properties(nv, oracle, case):
    dtype, n = CASES[case]
    dev = torch.device("cuda", 0)
    stream = torch.cuda.Stream(device=dev)   # a real stream: handle 0 would mean "the index's own stream" to the C ABI
    sp = stream.cuda_stream
    nq_oracle, nq_batch = 6, 64
    qrows = [(i * 104729 * 31 + 17) % n for i in range(nq_batch)]
    with nv.FlatIndexCuda.synthetic(dtype, n, D, SEED, device=0) as idx:
        q_dev = torch.empty((nq_batch, D), dtype=torch.float32, device=dev)
        for i, r in enumerate(qrows):
            nv.synth_fill_device(0, F32, 1, D, SEED, r, q_dev[i].data_ptr(), None, sp)
        torch.cuda.synchronize()
        queries = q_dev.cpu().numpy()
        queries[1::2] += (1e-3 * np.random.default_rng(3).standard_normal((nq_batch // 2, D))).astype(np.float32)

        # -- streaming scan, one query at a time (the bit-exact path), and the AUTO batch (exact tensor-core path for
        #    f16 / i8, streaming passes for f32): identical bits
        idx.set_path(nv.PATH_GEMV)
        ids1 = np.empty((nq_batch, K), np.uint64)
        sc1 = np.empty((nq_batch, K), np.float32)
        for i in range(nq_batch):
            ids1[i], sc1[i] = idx.search_topk_dot(queries[i], K)
        idx.set_path(nv.PATH_AUTO)
        ids2, sc2 = idx.search_topk_dot_batch(queries, K)
        assert np.array_equal(sc1.view(np.uint32), sc2.view(np.uint32)), "batch path scores differ from the streaming scan"
        for i in range(nq_batch):
            assert_topk_equal(ids2[i], sc2[i], ids1[i], sc1[i], what=f"{case} batch q{i}")

        # -- properties
        assert np.all(np.diff(sc1, axis=1) <= 0), "scores must be sorted descending"
        assert np.all(ids1 < n)
        assert all(len(set(row.tolist())) == K for row in ids1), "ids must be unique"
        for i in range(0, nq_batch, 2):                         # unperturbed base rows find themselves
            assert int(ids1[i, 0]) == qrows[i], f"self query {i}"

        # -- the oracle over the SAME bytes
        base, scales = _read_all(idx, n)
        for i in range(nq_oracle):
            oids, osc = oracle.search(dtype, base, scales, queries[i], K, mt=True)
            assert_topk_equal(ids1[i], sc1[i], oids, osc, what=f"{case} oracle q{i}")
        del base, scales

    # -- two row shards, merged on the device == the whole
    half = n // 2 + 1
    with nv.FlatIndexCuda.synthetic(dtype, half, D, SEED, first_row=0, device=0, row_offset=0) as a, \
            nv.FlatIndexCuda.synthetic(dtype, n - half, D, SEED, first_row=half, device=0, row_offset=half) as b:
        nq = 8
        w = nv.packed_words(nq, K)
        gathered = torch.empty((2, w), dtype=torch.int64, device=dev)
        for s, sh in enumerate((a, b)):
            sh.set_path(nv.PATH_GEMV)
            sh.search_device(q_dev.data_ptr(), nq, K, gathered[s, :nq * K].data_ptr(),
                             gathered[s, nq * K:].view(torch.float32).data_ptr(), sp)
        out_ids = torch.empty((nq, K), dtype=torch.int64, device=dev)
        out_sc = torch.empty((nq, K), dtype=torch.float32, device=dev)
        nv.merge_topk_packed_device(0, gathered.data_ptr(), 2, nq, K, out_ids.data_ptr(), out_sc.data_ptr(), sp)
        torch.cuda.synchronize()
        # q_dev holds the unperturbed rows: compare against a fresh whole-index scan of the same queries
    with nv.FlatIndexCuda.synthetic(dtype, n, D, SEED, device=0) as idx:
        idx.set_path(nv.PATH_GEMV)
        wi = torch.empty((nq, K), dtype=torch.int64, device=dev)
        ws = torch.empty((nq, K), dtype=torch.float32, device=dev)
        idx.search_device(q_dev.data_ptr(), nq, K, wi.data_ptr(), ws.data_ptr(), sp)
        torch.cuda.synchronize()
        assert torch.equal(ws.view(torch.int32), out_sc.view(torch.int32)), "shard merge scores != whole scan"
        for i in range(nq):
            assert_topk_equal(out_ids[i].cpu().numpy().view(np.uint64), out_sc[i].cpu().numpy(),
                              wi[i].cpu().numpy().view(np.uint64), ws[i].cpu().numpy(), what=f"{case} shards q{i}")
