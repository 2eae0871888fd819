"""hg_csr_from_coo_static: the CSR build with its scratch in a file-scope device array (no allocation inside the
call, no counterpart in the reference).  Called through the C ABI with
caller-allocated inputs / outputs and compared with hg_csr_from_coo -- integer outputs, so bit-exact -- on two calls
in a row (the second smaller than the first: the shared scratch still holds the first call's contents).

Collected last on purpose (file name)."""
import ctypes as C

import numpy as np
import pytest
import torch

from hg_b200 import _lib, ops

pytestmark = pytest.mark.gpu

MAX_N, MAX_E = 131072, 524288          # HG_CSR_STATIC_MAX_N / HG_CSR_STATIC_MAX_EDGES of include/hg_b200.h


def _pairs(n: int, m: int, seed: int) -> torch.Tensor:
    """Sorted unique (min, max) pair list as hg_edges_fill emits it, int64 [2, E] on the GPU."""
    rng = np.random.default_rng(seed)
    a = rng.integers(0, n, size=m)
    b = rng.integers(0, n, size=m)
    keep = a != b
    lo, hi = np.minimum(a, b)[keep], np.maximum(a, b)[keep]
    key = np.unique(lo.astype(np.int64) * n + hi)
    return torch.from_numpy(np.stack([key // n, key % n])).cuda()


def _static_csr(lib, ei: torch.Tensor, n: int):
    e = ei.shape[1]
    rowptr = torch.full((n + 1,), -1, dtype=torch.int64, device="cuda")
    col = torch.full((2 * e,), -1, dtype=torch.int64, device="cuda")
    eid = torch.full((2 * e,), -1, dtype=torch.int64, device="cuda")
    rc = lib.hg_csr_from_coo_static(ei.data_ptr(), e, n, rowptr.data_ptr(), col.data_ptr(), eid.data_ptr(),
                                    torch.cuda.current_stream().cuda_stream)
    _lib.check(rc)
    return rowptr, col, eid


def test_static_scratch_csr_equals_the_workspace_entry_point_twice_in_a_row():
    lib = _lib.load_library()
    torch.cuda.set_device(0)
    cases = [(100_000, 400_000, 1), (3_000, 20_000, 2), (100_000, 400_000, 3), (2, 1, 4)]
    for n, m, seed in cases:                              # big, small (stale scratch behind it), big again, tiny
        ei = _pairs(n, m, seed)
        assert ei.shape[1] <= MAX_E
        got = _static_csr(lib, ei, n)
        want = ops.csr_from_coo(ei, n)                    # hg_csr_from_coo with a caller workspace
        torch.cuda.synchronize()
        for g, w, name in zip(got, want, ("rowptr", "col", "eid")):
            assert torch.equal(g, w), (n, name)
        assert int(got[0][-1]) == 2 * ei.shape[1]
    # no edges: rowptr all zero, nothing else touched
    empty = torch.empty((2, 0), dtype=torch.int64, device="cuda")
    rowptr, col, eid = _static_csr(lib, empty, 17)
    assert (rowptr == 0).all() and col.numel() == 0


def test_static_scratch_csr_rejects_sizes_beyond_its_limits():
    lib = _lib.load_library()
    ei = _pairs(1000, 4000, 5)
    out = torch.empty(8, dtype=torch.int64, device="cuda")
    for n, e in ((MAX_N + 1, ei.shape[1]), (1000, MAX_E + 1)):
        rc = lib.hg_csr_from_coo_static(ei.data_ptr(), e, n, out.data_ptr(), out.data_ptr(), out.data_ptr(), 0)
        assert rc == _lib.HG_ERR_UNSUPPORTED
        assert "beyond the static scratch" in _lib.last_error()
        with pytest.raises(ValueError):
            _lib.check(rc)
    assert lib.hg_csr_from_coo_static(ei.data_ptr(), -1, 1000, out.data_ptr(), out.data_ptr(), out.data_ptr(), 0) == _lib.HG_ERR_BAD_ARG
