"""CPU tests (no GPU): the C-ABI library loads, exports every symbol include/nvdb_b200.h declares, and fails loudly
(status code, never a CPU fallback) when no CUDA device is present."""
import ctypes as C
import re
from pathlib import Path

import numpy as np
import pytest

ROOT = Path(__file__).resolve().parent.parent
HEADER = ROOT / "include" / "nvdb_b200.h"


def declared_functions():
    text = re.sub(r"/\*.*?\*/", "", HEADER.read_text(), flags=re.S)
    return sorted(set(re.findall(r"\b(nvdb_b200_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_boundary():
    names = declared_functions()
    for must in ("nvdb_b200_index_create", "nvdb_b200_index_destroy", "nvdb_b200_search", "nvdb_b200_search_device",
                 "nvdb_b200_refine_l2", "nvdb_b200_merge_topk_device", "nvdb_b200_last_error"):
        assert must in names


def test_library_exports_every_declared_symbol(nv):
    lib = C.CDLL(str(nv.lib_path))
    missing = [n for n in declared_functions() if not hasattr(lib, n)]
    assert not missing, f"declared in include/nvdb_b200.h but not exported: {missing}"
    # and the ctypes binding covers the same surface
    assert sorted(nv.capi.EXPORTED_SYMBOLS) == declared_functions()


def test_version_and_status_strings(nv):
    assert nv.lib.nvdb_b200_version() == 101
    assert nv.lib.nvdb_b200_status_string(2) == b"Empty base"
    assert nv.lib.nvdb_b200_status_string(0) == b"ok"


def test_argument_errors_need_no_device(nv):
    h = C.c_void_p()
    x = np.zeros((4, 8), np.float32)
    p = x.ctypes.data_as(C.c_void_p)
    assert nv.lib.nvdb_b200_index_create(None, None, 1, 4, 8, 0, 0, C.byref(h)) == 1          # null payload
    assert nv.lib.nvdb_b200_index_create(p, None, 9, 4, 8, 0, 0, C.byref(h)) == 1             # bad dtype
    assert b"Unsupported base dtype" in nv.lib.nvdb_b200_last_error(None)
    assert nv.lib.nvdb_b200_index_create(p, None, 1, 0, 8, 0, 0, C.byref(h)) == 2             # "Empty base"
    assert nv.lib.nvdb_b200_index_create(p, None, 3, 4, 8, 0, 0, C.byref(h)) == 1             # i8 without scales
    assert nv.lib.nvdb_b200_search(None, p, 1, 1, p, p, None) == 1
    assert nv.lib.nvdb_b200_index_destroy(None) == 0


def test_no_cpu_fallback_without_a_device(nv):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a CUDA device is present")
    with pytest.raises(nv.NvdbError) as e:
        nv.FlatIndexCuda.from_host(np.zeros((4, 8), np.float32), None, nv.F32)
    assert e.value.status in (3, 4)   # NVDB_B200_ERR_CUDA / OOM: loud failure, no result


def test_product_package_never_imports_the_oracle():
    """no include / import / dlopen of anything under oracle/ from the product tree (comments may cite it)."""
    pkg = ROOT / "nano-vectordb_b200"
    files = [p for ext in ("*.py", "*.cu", "*.cuh", "*.cpp", "*.h") for p in pkg.rglob(ext)]
    assert files
    for p in files:
        for line in p.read_text().splitlines():
            if re.search(r"#\s*include|^\s*(import|from)\s|CDLL|dlopen|import_module", line):
                assert "oracle" not in line, f"{p}: {line}"
