"""tm_kmodes_assign_noalloc: the allocation-free variant of the KModes assignment pass (fixed device scratch inside the
library, no cudaMalloc / cudaFree, no synchronisation), against the existing tm_kmodes_assign_dev through the C ABI.
Integer work: the outputs must be equal bit for bit (the tolerance of every KModes test here).  Two calls in a row on
the shared scratch - the first with a centroid byte >= 16 (sets the 'wide' flag, byte-wise path), the second without
(the flag must have been cleared, bit-sliced path) - then the size limit.  Collected last on purpose (pytest -x)."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _vp(t):
    return C.c_void_p(t.data_ptr())


def _rows(rng, n, hi=16):
    X = rng.integers(0, hi, (n, 80)).astype(np.uint8)
    X[:, 64:] = rng.integers(0, 2, (n, 16))
    return X


def test_noalloc_entry_point_matches_existing_one_on_two_calls_in_a_row():
    import torch
    import tiler_b200 as tb
    from tiler_b200._lib import TMError, lib
    tb.set_device(0)
    torch.cuda.set_device(0)
    dev = torch.device("cuda", 0)
    rng = np.random.default_rng(123)
    Xh = _rows(rng, 20000)
    X = torch.from_numpy(Xh).to(dev)
    stream = torch.cuda.current_stream()
    cases = []
    c1 = _rows(rng, 1500)
    c1[7, 3] = 200                      # wide centroid: byte-wise path, flag set
    c1[11] = c1[10]                     # exact duplicate: the LAST index must win
    cases.append((c1, tb.KM_ASM))
    cases.append((_rows(rng, 4096), tb.KM_ASM))      # narrow again: stale flag / stale bit-sliced rows would show
    cases.append((Xh[::37][:333].copy(), tb.KM_GENERIC))
    for cent_h, sem in cases:
        cent = torch.from_numpy(cent_h).to(dev)
        k = cent.shape[0]
        want_c = torch.empty(len(Xh), dtype=torch.int32, device=dev)
        want_d = torch.empty(len(Xh), dtype=torch.int64, device=dev)
        rc = lib.tm_kmodes_assign_dev(_vp(X), len(Xh), _vp(cent), k, sem, _vp(want_c), _vp(want_d))
        assert rc == 0, lib.tm_last_error()
        torch.cuda.synchronize()
        got_c = torch.full((len(Xh),), -7, dtype=torch.int32, device=dev)
        got_d = torch.full((len(Xh),), -7, dtype=torch.int64, device=dev)
        torch.cuda.synchronize()
        rc = lib.tm_kmodes_assign_noalloc(_vp(X), len(Xh), _vp(cent), k, sem, _vp(got_c), _vp(got_d),
                                          C.c_void_p(stream.cuda_stream))
        assert rc == 0, lib.tm_last_error()
        torch.cuda.synchronize()
        assert torch.equal(got_c, want_c)
        assert torch.equal(got_d, want_d)
    # d_dis may be NULL, like the existing entry point
    got_c.fill_(-7)
    torch.cuda.synchronize()
    rc = lib.tm_kmodes_assign_noalloc(_vp(X), len(Xh), _vp(cent), k, sem, _vp(got_c), None, C.c_void_p(stream.cuda_stream))
    assert rc == 0
    torch.cuda.synchronize()
    assert torch.equal(got_c, want_c)
    # beyond the fixed scratch: rejected with a clear error, nothing launched
    big = torch.zeros((16385, 80), dtype=torch.uint8, device=dev)
    rc = lib.tm_kmodes_assign_noalloc(_vp(X), len(Xh), _vp(big), 16385, tb.KM_ASM, _vp(got_c), _vp(got_d), None)
    assert rc != 0 and b"TM_NOALLOC_MAX_CENTROIDS" in lib.tm_last_error()
    torch.cuda.synchronize()
    assert torch.equal(got_c, want_c)
    # exactly at the limit works
    full = torch.from_numpy(_rows(rng, 16384)).to(dev)
    lib.tm_kmodes_assign_dev(_vp(X), 4096, _vp(full), 16384, tb.KM_ASM, _vp(want_c), _vp(want_d))
    torch.cuda.synchronize()
    rc = lib.tm_kmodes_assign_noalloc(_vp(X), 4096, _vp(full), 16384, tb.KM_ASM, _vp(got_c), _vp(got_d), None)
    assert rc == 0, lib.tm_last_error()
    torch.cuda.synchronize()
    assert torch.equal(got_c[:4096], want_c[:4096]) and torch.equal(got_d[:4096], want_d[:4096])
    del TMError
