"""Row-sharded flat index: one process per GPU, `torch.distributed` for the one exchange step.

The reference has no multi-GPU path; its CPU variants already use this decomposition across
threads — contiguous row blocks, a top-k per block, then a merge (reference src/flat_index_pool.cpp:155-158,
212-214; src/flat_index_omp.cpp:39-75). Here a block is a GPU:

    rank g holds rows [g*ceil(N/G), min(N, (g+1)*ceil(N/G)))       (`shard_rows`)
    1. every rank scans its shard with the same queries           -> local top-k, GLOBAL row ids (row_offset)
    2. ONE all-gather of the packed local results (k*(8+4) B per query per rank) over NVLink/NVSwitch
    3. every rank merges G*k candidates -> final top-k under (score desc, id asc)   (merge kernel)

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

    def __init__(self, local_index, device, group=None, local_search=None, merge=None):
        self.idx = local_index
        self.device = torch.device(device)
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0
        self._local_search = local_search or self._cuda_local_search
        self._merge = merge or self._cuda_merge
        self._bufs = {}
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

    def search(self, queries: torch.Tensor, k: int):
        """queries: [nq, dim] f32 on `device`. Returns (ids int64 [nq,k] — u64 bit patterns, UINT64_MAX padding —,
        scores f32 [nq,k]) on every rank; buffers are reused by the next call with the same (nq, k)."""
        if self._stream is None:
            return self._search(queries, k)
        cur = torch.cuda.current_stream(self.device)
        self._stream.wait_stream(cur)
        with torch.cuda.stream(self._stream):
            out = self._search(queries, k)
        cur.wait_stream(self._stream)
        return out

    def _search(self, queries, k):
        nq = int(queries.shape[0])
        local, gathered, out_ids, out_sc = self._buffers(nq, k)
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
