"""nano-vectordb_b200 — B200-native exact flat-scan top-k + exact refine (one hot path of Nano-VectorDB).

The product is the CUDA library `lib/libnvdb_b200.so` behind the C ABI in `include/nvdb_b200.h`, plus the C++
host mirror of the reference's index API in `host/`. This Python module is only a thin ctypes door onto that C
ABI for tests and the benchmark harness; it holds no compute and no CPU fallback: importing it without the
built library raises.

The directory name contains a hyphen (it follows the reference's repo name), so import it with
`importlib.import_module("nano-vectordb_b200")`.
"""
from .capi import (F16, F32, I8, PATH_AUTO, PATH_GEMV, PATH_MMA, PATH_MMA_EXACT, PATH_MMA_FAST, FlatIndexCuda, IvfPqIndex, NvdbError, Timing, launch_count,
                   lib, lib_path, merge_topk_device, merge_topk_packed_device, synth_fill_device)
from . import capi  # noqa: F401
from .sharded import ShardedFlatIndex, packed_words, shard_rows

__all__ = ["F32", "F16", "I8", "PATH_AUTO", "PATH_GEMV", "PATH_MMA", "PATH_MMA_FAST", "PATH_MMA_EXACT", "FlatIndexCuda", "IvfPqIndex", "NvdbError", "Timing",
           "launch_count", "lib", "lib_path", "merge_topk_device", "merge_topk_packed_device", "synth_fill_device",
           "ShardedFlatIndex", "packed_words", "shard_rows"]
