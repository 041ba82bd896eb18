"""Row-sharded flat index: one process per GPU, `torch.distributed` for the one exchange step.

The reference has no multi-GPU path; its CPU variants already use this decomposition across
threads — contiguous row blocks, a top-k per block, then a merge (reference src/flat_index_pool.cpp:155-158,
212-214; src/flat_index_omp.cpp:39-75). Here a block is a GPU:

    rank g holds rows [g*ceil(N/G), min(N, (g+1)*ceil(N/G)))       (`shard_rows`)
    1. every rank scans its shard with the same queries           -> local top-k, GLOBAL row ids (row_offset)
    2. the packed local results (k*(8+4) B per query per rank) cross NVLink/NVSwitch
    3. every rank merges G*k candidates -> final top-k under (score desc, id asc)
   Steps 2 + 3 are ONE kernel per rank over peer memory (`exchange="peer"`, csrc/exchange.cu: NVLink stores into the
   peers' windows, flags, merge — no NCCL call); `exchange="nccl"` is the baseline: one all-gather + the merge kernel.
   "auto" takes the peer path when every rank could map its peers (CUDA IPC), else NCCL.

torch is plumbing only (the process group and the buffers); scan and merge are the library's CUDA kernels.
`local_search` / `merge` are injectable so that the collective logic can be exercised on CPU with the gloo
backend (tests/test_sharded_gloo.py), where the test supplies checker implementations.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_rows(n_total: int, world: int, rank: int) -> tuple[int, int]:
    """(first_row, nrows) of `rank`: contiguous ceil(N/G) blocks, the last ones possibly short or empty."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad world/rank")
    per = -(-n_total // world)
    r0 = min(n_total, rank * per)
    return r0, min(n_total, r0 + per) - r0


def packed_words(nq: int, k: int) -> int:
    """int64 words of one rank's packed result: [nq*k ids (u64)] [nq*k scores (f32), padded to 8 bytes]."""
    return nq * k + (nq * k + 1) // 2


class ShardedFlatIndex:
    """`local_index`: this rank's FlatIndexCuda (created with row_offset = first row of the shard), or None when
    the shard is empty. search() must be called by every rank of `group` with the same queries."""

    def __init__(self, local_index, device, group=None, local_search=None, merge=None, exchange="auto"):
        if exchange not in ("auto", "peer", "nccl"):
            raise ValueError("exchange must be 'auto', 'peer' or 'nccl'")
        self.idx = local_index
        self.device = torch.device(device)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._local_search = local_search or self._cuda_local_search
        self._merge = merge or self._cuda_merge
        self._bufs = {}
        # peer-memory exchange: CUDA only, and only with the library's own search / merge (the injected checker
        # implementations of the gloo tests go through the all-gather)
        injected = local_search is not None or merge is not None
        if self.device.type != "cuda" or injected or self.world == 1:
            if exchange == "peer" and (self.device.type != "cuda" or injected):
                raise ValueError("exchange='peer' needs a CUDA device and the library's own search and merge")
            exchange = "nccl"
        self.exchange = exchange          # "auto" is resolved by the first search
        self._xchg = None
        self._xchg_slot = 0
        # The C ABI reads a NULL stream as "the index's own stream", so the CUDA steps never run on torch's legacy
        # default stream (handle 0): they run on this side stream, ordered after / before the caller's stream.
        self._stream = torch.cuda.Stream(device=self.device) if self.device.type == "cuda" else None

    # ---- default (CUDA) implementations: raw pointers into the C ABI, asynchronous on the current stream
    def _cuda_local_search(self, queries, nq, k, ids_out, scores_out):
        sp = torch.cuda.current_stream(self.device).cuda_stream
        self.idx.search_device(queries.data_ptr(), nq, k, ids_out.data_ptr(), scores_out.data_ptr(), sp)

    def _cuda_merge(self, gathered, nlists, nq, k, ids_out, scores_out):
        from . import capi
        sp = torch.cuda.current_stream(self.device).cuda_stream
        capi.merge_topk_packed_device(self.device.index or 0, gathered.data_ptr(), nlists, nq, k, ids_out.data_ptr(),
                                      scores_out.data_ptr(), sp)

    def _buffers(self, nq, k):
        key = (nq, k)
        if key not in self._bufs:
            w = packed_words(nq, k)
            local = torch.empty(w, dtype=torch.int64, device=self.device)
            gathered = torch.empty((self.world, w), dtype=torch.int64, device=self.device)
            out_ids = torch.empty((nq, k), dtype=torch.int64, device=self.device)
            out_sc = torch.empty((nq, k), dtype=torch.float32, device=self.device)
            self._bufs[key] = (local, gathered, out_ids, out_sc)
        return self._bufs[key]

    # ---- peer-memory exchange (csrc/exchange.cu)
    def _peer_setup(self, slot_bytes: int) -> bool:
        """Collective: (re)create this rank's window, swap CUDA IPC handles, map the peers. False when any rank
        failed (then every rank drops to the all-gather)."""
        from . import capi
        if self._xchg is not None:
            torch.cuda.synchronize(self.device)
            dist.barrier(group=self.group)      # nobody may still be storing into a window that is about to go
            self._xchg.close()
            self._xchg = None
        dev = self.device.index or 0
        ok, x, handle = 1, None, bytes(capi.IPC_HANDLE_BYTES)
        try:
            x = capi.PeerExchange(dev, self.world, self.rank, slot_bytes)
            handle = x.ipc_handle()
        except capi.NvdbError:
            ok = 0
        mine = torch.tensor(list(handle) + [ok], dtype=torch.uint8, device=self.device)
        allh = torch.empty((self.world, capi.IPC_HANDLE_BYTES + 1), dtype=torch.uint8, device=self.device)
        dist.all_gather_into_tensor(allh.view(-1), mine, group=self.group)
        allh = allh.cpu()
        if ok and bool(allh[:, -1].all()):
            try:
                x.connect(handles=bytes(allh[:, :-1].contiguous().view(-1).tolist()))
            except capi.NvdbError:
                ok = 0
        else:
            ok = 0
        flag = torch.tensor([ok], dtype=torch.int32, device=self.device)
        dist.all_reduce(flag, op=dist.ReduceOp.MIN, group=self.group)   # also: every rank has mapped before any stores
        if int(flag.item()) != 1:
            if x is not None:
                x.close()
            return False
        self._xchg, self._xchg_slot = x, slot_bytes
        return True

    def _peer_ready(self, nq: int, k: int) -> bool:
        if self.exchange == "nccl":
            return False
        need = 8 * packed_words(nq, k)
        if self._xchg is None or need > self._xchg_slot:
            if not self._peer_setup(max(need, 1 << 16)):
                if self.exchange == "peer":
                    raise RuntimeError("exchange='peer': CUDA IPC mapping of the peers' windows failed on some rank")
                self.exchange = "nccl"
                return False
            self.exchange = "peer"
        return True

    def exchange_status(self) -> int:
        """0 = fine; 1 = a peer did not show up within the timeout (results of that call were padding)."""
        return 0 if self._xchg is None else self._xchg.status()

    def close(self):
        if self._xchg is not None:
            torch.cuda.synchronize(self.device)
            if dist.is_initialized():
                dist.barrier(group=self.group)
            self._xchg.close()
            self._xchg = None

    def search(self, queries: torch.Tensor, k: int):
        """queries: [nq, dim] f32 on `device`. Returns (ids int64 [nq,k] — u64 bit patterns, UINT64_MAX padding —,
        scores f32 [nq,k]) on every rank; buffers are reused by the next call with the same (nq, k)."""
        if self._stream is None:
            return self._search(queries, k)
        cur = torch.cuda.current_stream(self.device)
        if cur.cuda_stream != 0:        # already a real stream: no hop (two cross-stream events per call otherwise)
            return self._search(queries, k)
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream):
            out = self._search(queries, k)
        cur.wait_stream(self._stream)
        return out

    def _search(self, queries, k):
        nq = int(queries.shape[0])
        local, gathered, out_ids, out_sc = self._buffers(nq, k)
        if self.world > 1 and self._peer_ready(nq, k):
            # scan + exchange + merge in the library (ONE launch for a single-pass streaming scan)
            sp = torch.cuda.current_stream(self.device).cuda_stream
            self._xchg.search(self.idx, queries.data_ptr(), nq, k, out_ids.data_ptr(), out_sc.data_ptr(), sp)
            return out_ids, out_sc
        ids_view = local[:nq * k].view(nq, k)
        sc_view = local[nq * k:].view(torch.float32)[:nq * k].view(nq, k)
        if self.idx is None:                      # empty shard: contribute only padding
            ids_view.fill_(-1)
            sc_view.fill_(float("-inf"))
        else:
            self._local_search(queries, nq, k, ids_view, sc_view)
        if self.world == 1:
            gathered[0].copy_(local)
        else:
            dist.all_gather_into_tensor(gathered.view(-1), local, group=self.group)
        self._merge(gathered, self.world, nq, k, out_ids, out_sc)
        return out_ids, out_sc
