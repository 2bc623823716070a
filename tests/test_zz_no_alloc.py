"""bf_exclusive_scan_u32_noalloc: the prefix sum for hosts that cannot give the library scratch
memory (an application whose own arena owns all of HBM, or a plugin slot that must not call
cudaMalloc while the pipeline runs).  Its tile sums live in a fixed array inside the library.

Checked through the C ABI against bf_exclusive_scan_u32 (caller-provided scratch) on two calls
in a row with different inputs -- the second call must not be disturbed by what the first left
in the shared array -- plus the size limit.  Integer arithmetic: the comparison is exact.
The file name sorts last on purpose, so a `-x` run has already executed every other test."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def test_noalloc_scan_matches_scan_with_caller_scratch_twice():
    import torch

    from beast_b200 import _cabi

    lib = _cabi.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    stream = C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    rng = np.random.default_rng(2024)
    for n, hi in ((1_000_003, 50), (70_001, 3000)):  # second call: fewer tiles than the first
        vals = torch.from_numpy(rng.integers(0, hi, size=n, dtype=np.int64).astype(np.int32)).to(dev)
        want = torch.empty(n + 1, dtype=torch.int32, device=dev)
        scratch = torch.empty(int(lib.bf_scan_scratch_elems(n)), dtype=torch.int32, device=dev)
        _cabi.call("bf_exclusive_scan_u32", n, _ptr(vals), _ptr(want), _ptr(scratch), stream)
        got = torch.full((n + 1,), -1, dtype=torch.int32, device=dev)
        torch.cuda.synchronize(dev)
        before = torch.cuda.memory_allocated(dev)
        _cabi.call("bf_exclusive_scan_u32_noalloc", n, _ptr(vals), _ptr(got), stream)
        torch.cuda.synchronize(dev)
        assert torch.cuda.memory_allocated(dev) == before
        assert torch.equal(got, want)
        ref = np.concatenate([[0], np.cumsum(vals.cpu().numpy().astype(np.int64))])
        assert np.array_equal(got.cpu().numpy().astype(np.int64) & 0xFFFFFFFF, ref & 0xFFFFFFFF)


def test_noalloc_scan_refuses_sizes_beyond_its_scratch():
    import torch

    from beast_b200 import _cabi

    lib = _cabi.load()
    dev = torch.device("cuda", 0)
    torch.cuda.set_device(dev)
    dummy = torch.zeros(8, dtype=torch.int32, device=dev)
    too_many = 65536 * 2048 + 1
    rc = lib.bf_exclusive_scan_u32_noalloc(too_many, _ptr(dummy), _ptr(dummy), None)
    assert rc == -4  # BF_ERR_TOO_LARGE
    assert b"exceeds the built-in scratch" in lib.bf_last_error()
