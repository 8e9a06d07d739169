"""The C-ABI library loads, exports every symbol include/tmann.h declares, and fails
loudly (never falls back to a CPU path) when there is no GPU.  No compute calls here."""
import ctypes as C
import os
import re
import subprocess

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _header_symbols():
    src = open(os.path.join(ROOT, "include", "tmann.h")).read()
    return [m for m in re.findall(r"^TM_API\s+[\w\s\*]+?\**\s*\*?(\w+)\s*\(", src, re.M)]


def test_header_symbols_exported():
    from tiler_b200 import _lib
    syms = _header_symbols()
    assert len(syms) >= 25 and "ann_kdtree_create" in syms and "tm_kmodes_assign" in syms
    out = subprocess.run(["nm", "-D", "--defined-only", _lib.LIB_PATH], capture_output=True, text=True).stdout
    exported = {l.split()[-1] for l in out.splitlines() if " T " in l}
    missing = [s for s in syms if s not in exported]
    assert not missing, missing
    # the ctypes prototype table covers the header exactly
    assert sorted(_lib.PROTOTYPES) == sorted(syms)
    # nothing but the C ABI leaks out of the shared object
    extra = [s for s in exported if not (s.startswith("ann_kdtree_") or s.startswith("tm_"))]
    assert not extra, extra


def test_five_legacy_symbols_match_extern_pas():
    # extern.pas:63-67
    for s in ["ann_kdtree_create", "ann_kdtree_destroy", "ann_kdtree_search", "ann_kdtree_pri_search",
              "ann_kdtree_search_multi"]:
        assert s in _header_symbols()


def test_library_built_for_sm100a_only():
    from tiler_b200 import _lib
    out = subprocess.run(["cuobjdump", "-lelf", _lib.LIB_PATH], capture_output=True, text=True).stdout
    archs = set(re.findall(r"sm_(\d+a?)", out))
    assert archs == {"100a"}, archs


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "tiler_b200")):
        for f in files:
            if f.endswith((".py", ".cu", ".cuh", ".inc", ".cpp", ".h")):
                txt = open(os.path.join(dirpath, f)).read()
                assert "tmoracle" not in txt and "oracle/" not in txt, f


def _gpu():
    import torch
    return torch.cuda.is_available()


@pytest.mark.skipif(_gpu(), reason="checks the no-GPU failure mode")
def test_fails_loudly_without_gpu():
    import tiler_b200 as tb
    from tiler_b200._lib import TMError, lib
    rows = np.zeros((4, 192), np.float32)
    with pytest.raises(TMError, match="no CUDA device|not sm_100|CUDA"):
        tb.KDTree.from_rows(rows)
    with pytest.raises(TMError):
        tb.kmodes_assign(np.zeros((2, 80), np.uint8), np.zeros((1, 80), np.uint8))
    with pytest.raises(TMError):
        tb.psyv_queries(np.zeros((1, 64), np.int32))
    # legacy shape: NULL handle, index -1, destroy(NULL) is a no-op (extern.pas:63-65)
    assert not lib.ann_kdtree_create(None, 0, 192, 1, 0)
    lib.ann_kdtree_destroy(None)
    assert lib.ann_kdtree_search(None, None, 0.0, None) == -1
