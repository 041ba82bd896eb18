"""GPU parity for the row-sharded path on ONE device: S logical shards (each its own index with a row_offset),
packed exactly as the all-gather delivers them, merged by the library's merge kernel == the unsharded top-k
("merge of per-shard top-k == top-k of the whole")."""
import numpy as np
import pytest
import torch

from _oracle import F16, F32, I8, assert_topk_equal

pytestmark = pytest.mark.gpu
DT = {"f32": F32, "f16": F16, "i8": I8}


@pytest.mark.parametrize("dt", ["f16", "i8", "f32"])
@pytest.mark.parametrize("S,n_total,k,nq", [(2, 5001, 10, 3), (8, 20000, 100, 2), (4, 3, 5, 1), (3, 7000, 10, 16)])
def test_logical_shards_merge_equals_unsharded(nv, oracle, dt, S, n_total, k, nq):
    dtype, d = DT[dt], 128
    dev = torch.device("cuda", 0)
    base, scales = oracle.synth(dtype, 13, 0, n_total, d)
    f = oracle.synth_f32(13, 0, n_total, d)
    rng = np.random.default_rng(0)
    queries = np.ascontiguousarray(f[rng.integers(0, n_total, nq)] + (0.01 * rng.standard_normal((nq, d))).astype(np.float32))
    q_dev = torch.from_numpy(queries).to(dev)
    w = nv.packed_words(nq, k)
    gathered = torch.empty((S, w), dtype=torch.int64, device=dev)
    shards = []
    for s in range(S):
        r0, nrows = nv.shard_rows(n_total, S, s)
        ids_view = gathered[s, :nq * k].view(nq, k)
        sc_view = gathered[s, nq * k:].view(torch.float32)[:nq * k].view(nq, k)
        if nrows == 0:
            ids_view.fill_(-1)
            sc_view.fill_(float("-inf"))
            continue
        idx = nv.FlatIndexCuda.from_host(base[r0:r0 + nrows], None if scales is None else scales[r0:r0 + nrows], dtype,
                                         row_offset=r0)
        idx.set_path(nv.PATH_GEMV)
        shards.append(idx)
        idx.search_device(q_dev.data_ptr(), nq, k, ids_view.data_ptr(), sc_view.data_ptr(),
                          torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    out_ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
    out_sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
    nv.merge_topk_packed_device(0, gathered.data_ptr(), S, nq, k, out_ids.data_ptr(), out_sc.data_ptr(),
                                torch.cuda.current_stream().cuda_stream)
    torch.cuda.synchronize()
    ids = out_ids.cpu().numpy().view(np.uint64)
    sc = out_sc.cpu().numpy()
    oids, osc = oracle.search_batch(dtype, base, scales, queries, k)
    kk = min(k, n_total)
    for qi in range(nq):
        assert_topk_equal(ids[qi, :kk], sc[qi, :kk], oids[qi, :kk], osc[qi, :kk], what=f"S={S} q{qi}")
        assert np.all(ids[qi, kk:] == np.uint64(0xFFFFFFFFFFFFFFFF))
    # the world==1 path of ShardedFlatIndex (same code as N ranks minus the collective)
    with nv.FlatIndexCuda.from_host(base, scales, dtype) as whole:
        whole.set_path(nv.PATH_GEMV)
        sh = nv.ShardedFlatIndex(whole, dev)
        i2, s2 = sh.search(q_dev, k)
        torch.cuda.synchronize()
        assert np.array_equal(i2.cpu().numpy().view(np.uint64)[:, :kk], ids[:, :kk])
        assert np.array_equal(s2.cpu().numpy().view(np.uint32)[:, :kk], sc.view(np.uint32)[:, :kk])
    for idx in shards:
        idx.close()


def _packed_lists(S, nq, k, seed, dev, empty=()):
    """S packed per-shard results with global ids (descending scores per list, ties across lists included)."""
    rng = np.random.default_rng(seed)
    w = (nq * k) + (nq * k + 1) // 2
    packed = np.zeros((S, w), np.int64)
    for s in range(S):
        sc = np.sort(rng.integers(0, 50, (nq, k)).astype(np.float32) / 8.0, axis=1)[:, ::-1]   # many exact ties
        ids = (rng.permutation(nq * k).reshape(nq, k) + s * 1_000_003 + (1 << 33)).astype(np.uint64)
        if s in empty:
            ids[:] = np.uint64(0xFFFFFFFFFFFFFFFF)
            sc = np.full((nq, k), -np.inf, np.float32)
        elif k > 2:                                       # a short list: padding at the tail
            ids[:, -1] = np.uint64(0xFFFFFFFFFFFFFFFF)
            sc = sc.copy()
            sc[:, -1] = -np.inf
        packed[s, :nq * k] = ids.view(np.int64).reshape(-1)
        packed[s, nq * k:].view(np.float32)[:nq * k] = np.ascontiguousarray(sc).reshape(-1)
    return torch.from_numpy(packed).to(dev)


@pytest.mark.parametrize("S,nq,k", [(2, 1, 10), (8, 5, 100), (4, 300, 10), (3, 1024, 16), (16, 2, 7)])
def test_peer_exchange_logical_ranks_equal_packed_merge(nv, S, nq, k):
    """The peer-memory exchange kernel (csrc/exchange.cu) with S ranks living on ONE GPU (windows passed as plain
    pointers, push and merge as two launches so that no rank waits for a kernel that is not running yet) ==
    the all-gather + merge_lists path, bit for bit, on every rank, over several epochs (slot double-buffering)."""
    dev = torch.device("cuda", 0)
    slot = 8 * nv.packed_words(nq, k)
    xs = [nv.capi.PeerExchange(0, S, r, slot) for r in range(S)]
    wins = [x.window() for x in xs]
    for x in xs:
        x.connect(windows=wins)
    sp = torch.cuda.current_stream().cuda_stream
    try:
        for epoch in range(3):
            packed = _packed_lists(S, nq, k, 100 + epoch, dev, empty=(1,) if epoch == 1 else ())
            ref_ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
            ref_sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
            nv.merge_topk_packed_device(0, packed.data_ptr(), S, nq, k, ref_ids.data_ptr(), ref_sc.data_ptr(), sp)
            outs = [(torch.zeros((nq, k), dtype=torch.int64, device=dev), torch.zeros((nq, k), dtype=torch.float32, device=dev))
                    for _ in range(S)]
            for r, x in enumerate(xs):
                x.merge(packed[r].data_ptr(), nq, k, None, None, phases=nv.capi.XCHG_PUSH, stream=sp)
            for r, x in enumerate(xs):
                x.merge(None, nq, k, outs[r][0].data_ptr(), outs[r][1].data_ptr(), phases=nv.capi.XCHG_MERGE, stream=sp)
            torch.cuda.synchronize()
            for r in range(S):
                assert xs[r].status() == 0
                assert torch.equal(outs[r][0], ref_ids), f"epoch {epoch} rank {r}: ids differ"
                assert torch.equal(outs[r][1].view(torch.int32), ref_sc.view(torch.int32)), f"epoch {epoch} rank {r}: scores differ"
    finally:
        for x in xs:
            x.close()


def test_peer_exchange_fused_kernel_with_posted_peers(nv, monkeypatch):
    """The fused form (push + wait + merge in ONE launch) of one rank at a time. Kernels that wait for one another must
    never be separate launches on one GPU (nothing guarantees that they are resident together), so the other ranks of
    the epoch run the split form on the SAME stream: their pushes are enqueued before the fused launch, whose wait then
    finds every flag, and their merges after it. The rank that runs fused rotates over the epochs."""
    monkeypatch.setenv("NVDB_B200_XCHG_TIMEOUT_MS", "5000")
    dev = torch.device("cuda", 0)
    S, nq, k = 3, 64, 10
    xs = [nv.capi.PeerExchange(0, S, r, 8 * nv.packed_words(nq, k)) for r in range(S)]
    wins = [x.window() for x in xs]
    for x in xs:
        x.connect(windows=wins)
    sp = torch.cuda.current_stream().cuda_stream
    try:
        for epoch in range(6):
            fused = epoch % S
            packed = _packed_lists(S, nq, k, 7 + epoch, dev)
            ref_ids = torch.empty((nq, k), dtype=torch.int64, device=dev)
            ref_sc = torch.empty((nq, k), dtype=torch.float32, device=dev)
            nv.merge_topk_packed_device(0, packed.data_ptr(), S, nq, k, ref_ids.data_ptr(), ref_sc.data_ptr(), sp)
            outs = [(torch.zeros((nq, k), dtype=torch.int64, device=dev), torch.zeros((nq, k), dtype=torch.float32, device=dev))
                    for _ in range(S)]
            for r, x in enumerate(xs):
                if r != fused:
                    x.merge(packed[r].data_ptr(), nq, k, None, None, phases=nv.capi.XCHG_PUSH, stream=sp)
            xs[fused].merge(packed[fused].data_ptr(), nq, k, outs[fused][0].data_ptr(), outs[fused][1].data_ptr(),
                            phases=nv.capi.XCHG_FUSED, stream=sp)
            for r, x in enumerate(xs):
                if r != fused:
                    x.merge(None, nq, k, outs[r][0].data_ptr(), outs[r][1].data_ptr(), phases=nv.capi.XCHG_MERGE, stream=sp)
            torch.cuda.synchronize()
            for r in range(S):
                assert xs[r].status() == 0, "a rank timed out waiting for its peer"
                assert torch.equal(outs[r][0], ref_ids) and torch.equal(outs[r][1].view(torch.int32), ref_sc.view(torch.int32))
    finally:
        for x in xs:
            x.close()


def test_peer_exchange_timeout_pads_instead_of_hanging(nv, monkeypatch):
    """A peer that never shows up: the wait gives up after the timeout, the outputs are padding, status() reports it."""
    monkeypatch.setenv("NVDB_B200_XCHG_TIMEOUT_MS", "200")
    dev = torch.device("cuda", 0)
    S, nq, k = 2, 3, 10
    xs = [nv.capi.PeerExchange(0, S, r, 8 * nv.packed_words(nq, k)) for r in range(S)]
    wins = [x.window() for x in xs]
    for x in xs:
        x.connect(windows=wins)
    packed = _packed_lists(S, nq, k, 1, dev)
    ids = torch.zeros((nq, k), dtype=torch.int64, device=dev)
    sc = torch.zeros((nq, k), dtype=torch.float32, device=dev)
    xs[0].merge(packed[0].data_ptr(), nq, k, ids.data_ptr(), sc.data_ptr(), phases=nv.capi.XCHG_FUSED,
                stream=torch.cuda.current_stream().cuda_stream)     # rank 1 never calls
    torch.cuda.synchronize()
    assert xs[0].status() == 1
    assert bool((ids == -1).all()) and bool(torch.isinf(sc).all())
    for x in xs:
        x.close()


def test_peer_exchange_argument_errors(nv):
    with pytest.raises(nv.NvdbError):
        nv.capi.PeerExchange(0, 17, 0, 1024)           # world > 16
    with pytest.raises(nv.NvdbError):
        nv.capi.PeerExchange(0, 2, 2, 1024)            # rank >= world
    x = nv.capi.PeerExchange(0, 2, 0, 256)
    buf = torch.zeros(64, dtype=torch.int64, device="cuda")
    with pytest.raises(nv.NvdbError):                   # not connected
        x.merge(buf.data_ptr(), 1, 4, buf.data_ptr(), buf.data_ptr())
    x.close()
    x = nv.capi.PeerExchange(0, 1, 0, 256)
    with pytest.raises(nv.NvdbError):                   # packed result larger than the slot
        x.merge(buf.data_ptr(), 8, 10, buf.data_ptr(), buf.data_ptr())
    x.close()


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs")
def test_two_process_peer_exchange_equals_nccl_and_unsharded():
    """Two processes, two GPUs (torchrun): ShardedFlatIndex over the CUDA-IPC peer windows == the NCCL all-gather path
    == the unsharded index, bit for bit (tests/mp_sharded_peer.py)."""
    import socket
    import subprocess
    import sys
    from pathlib import Path
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    here = Path(__file__).resolve().parent
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr",
                        "127.0.0.1", "--master-port", str(port), str(here / "mp_sharded_peer.py")], capture_output=True,
                       text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "PEER_EXCHANGE_OK" in r.stdout


def _split_rank_push(nv, x, idx, q_dev, nq, k, packed, sp):
    """a rank in the split form: local scan into a packed buffer + push (no wait)"""
    if idx is None:
        x.merge(None, nq, k, None, None, phases=nv.capi.XCHG_PUSH, stream=sp)
        return
    ids_view = packed[:nq * k].view(nq, k)
    sc_view = packed[nq * k:].view(torch.float32)[:nq * k].view(nq, k)
    idx.search_device(q_dev.data_ptr(), nq, k, ids_view.data_ptr(), sc_view.data_ptr(), sp)
    x.merge(packed.data_ptr(), nq, k, None, None, phases=nv.capi.XCHG_PUSH, stream=sp)


@pytest.mark.parametrize("dt", ["f16", "i8", "f32"])
@pytest.mark.parametrize("n_total,k,nq", [(40_000, 10, 1), (30_001, 100, 1), (20_000, 10, 5), (9_000, 16, 40)])
def test_search_exchange_one_call_per_rank(nv, oracle, monkeypatch, dt, n_total, k, nq):
    """nvdb_b200_search_exchange_device: scan + exchange + merge of one rank (ONE launch when the streaming scan is a
    single pass: its last CTA pushes, waits and merges). Two logical ranks on one GPU and ONE stream: the rank under test
    runs the one-call form after its peer's split-form push and before the peer's merge (no kernel ever waits for a
    launch that is not in front of it); the fused rank alternates, three rounds (slot parity) == the oracle's top-k of
    the whole base, on both ranks."""
    monkeypatch.setenv("NVDB_B200_XCHG_TIMEOUT_MS", "10000")
    dtype, d, S = DT[dt], 128, 2
    dev = torch.device("cuda", 0)
    base, scales = oracle.synth(dtype, 29, 0, n_total, d)
    f = oracle.synth_f32(29, 0, n_total, d)
    rng = np.random.default_rng(3)
    xs = [nv.capi.PeerExchange(0, S, r, 8 * nv.packed_words(nq, k)) for r in range(S)]
    wins = [x.window() for x in xs]
    for x in xs:
        x.connect(windows=wins)
    shards = []
    for r in range(S):
        r0, nrows = nv.shard_rows(n_total, S, r)
        idx = nv.FlatIndexCuda.from_host(base[r0:r0 + nrows], None if scales is None else scales[r0:r0 + nrows], dtype,
                                         row_offset=r0)
        idx.set_path(nv.PATH_GEMV)
        shards.append(idx)
    stream = torch.cuda.Stream(device=dev)
    sp = stream.cuda_stream
    packed = torch.zeros(nv.packed_words(nq, k), dtype=torch.int64, device=dev)
    try:
        for rep in range(4):
            fused = rep % S
            other = 1 - fused
            queries = np.ascontiguousarray(f[rng.integers(0, n_total, nq)] + (0.01 * rng.standard_normal((nq, d))).astype(np.float32))
            q_dev = torch.from_numpy(queries).to(dev)
            outs = [(torch.zeros((nq, k), dtype=torch.int64, device=dev), torch.zeros((nq, k), dtype=torch.float32, device=dev))
                    for _ in range(S)]
            torch.cuda.synchronize()
            _split_rank_push(nv, xs[other], shards[other], q_dev, nq, k, packed, sp)
            l0 = nv.launch_count()
            xs[fused].search(shards[fused], q_dev.data_ptr(), nq, k, outs[fused][0].data_ptr(), outs[fused][1].data_ptr(), sp)
            if nq <= 4 and k <= 16:
                assert nv.launch_count() - l0 == 1, "a single-pass streaming scan must take ONE launch per rank"
            xs[other].merge(None, nq, k, outs[other][0].data_ptr(), outs[other][1].data_ptr(), phases=nv.capi.XCHG_MERGE, stream=sp)
            torch.cuda.synchronize()
            oids, osc = oracle.search_batch(dtype, base, scales, queries, k)
            for r in range(S):
                assert xs[r].status() == 0
                ids = outs[r][0].cpu().numpy().view(np.uint64)
                sc = outs[r][1].cpu().numpy()
                for qi in range(nq):
                    assert_topk_equal(ids[qi], sc[qi], oids[qi], osc[qi], what=f"rank {r} rep {rep} q{qi}")
            assert torch.equal(outs[0][0], outs[1][0]) and torch.equal(outs[0][1], outs[1][1])
    finally:
        for x in xs:
            x.close()
        for idx in shards:
            idx.close()


def test_search_exchange_empty_shard_contributes_padding(nv, oracle):
    """idx == NULL (a rank whose shard is empty) pushes padding; the other rank's result is its own top-k. Both orders:
    the empty rank in the split form around the full rank's one-call form, then the other way round."""
    dev = torch.device("cuda", 0)
    n, d, k, nq = 300, 64, 10, 2
    base, _ = oracle.synth(F16, 5, 0, n, d)
    queries = np.ascontiguousarray(oracle.synth_f32(5, 0, n, d)[:nq])
    q_dev = torch.from_numpy(queries).to(dev)
    xs = [nv.capi.PeerExchange(0, 2, r, 8 * nv.packed_words(nq, k)) for r in range(2)]
    wins = [x.window() for x in xs]
    for x in xs:
        x.connect(windows=wins)
    stream = torch.cuda.Stream(device=dev)
    sp = stream.cuda_stream
    packed = torch.zeros(nv.packed_words(nq, k), dtype=torch.int64, device=dev)
    with nv.FlatIndexCuda.from_host(base, None, F16) as idx:
        oids, osc = oracle.search_batch(F16, base, None, queries, k)
        shards = [idx, None]
        for fused in (0, 1):
            other = 1 - fused
            outs = [(torch.zeros((nq, k), dtype=torch.int64, device=dev), torch.zeros((nq, k), dtype=torch.float32, device=dev))
                    for _ in range(2)]
            _split_rank_push(nv, xs[other], shards[other], q_dev, nq, k, packed, sp)
            xs[fused].search(shards[fused], q_dev.data_ptr(), nq, k, outs[fused][0].data_ptr(), outs[fused][1].data_ptr(), sp)
            xs[other].merge(None, nq, k, outs[other][0].data_ptr(), outs[other][1].data_ptr(), phases=nv.capi.XCHG_MERGE, stream=sp)
            torch.cuda.synchronize()
            for r in range(2):
                assert xs[r].status() == 0
                for qi in range(nq):
                    assert_topk_equal(outs[r][0].cpu().numpy().view(np.uint64)[qi], outs[r][1].cpu().numpy()[qi], oids[qi], osc[qi],
                                      what=f"fused {fused} rank {r} q{qi}")
    for x in xs:
        x.close()
