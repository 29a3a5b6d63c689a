"""dl_sumsq_diff_static — the squared-distance reduction for callers that cannot allocate.

Some users of the native library have nowhere to get dl_sumsq_diff's scratch buffer from: a job
whose start-up code gave all of HBM to its own pool and that treats a later cudaMalloc as an error,
a watchdog thread that must keep out of the allocator, a C program that loads the .so without
torch.  dl_sumsq_diff_static keeps its 2,560 bytes of scratch inside the library instead.  This
test drives it the way such a caller would — straight through the C ABI with raw pointers — and
checks it against the ordinary entry point on two calls in a row: different lengths, so the second
call finds the first one's partial sums still lying in the shared array.

The file name sorts last on purpose: a run that stops at the first failure has then already run
every other test.  It passes on its own in a fresh process as well as after the rest of the suite.
"""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda:0"


def _rand(n, seed):
    return torch.randn(n, generator=torch.Generator().manual_seed(seed), dtype=torch.float32).to(DEV)


def test_static_scratch_entry_point_matches_the_allocating_one():
    from torch_distlearn_b200 import _lib
    lib = _lib.cuda()
    torch.cuda.set_device(0)
    stream = torch.cuda.current_stream().cuda_stream
    assert int(lib.dl_sumsq_static_scratch_bytes()) == 2560
    assert int(lib.dl_sumsq_static_max_n()) == 1 << 31

    # the reference result comes from the existing entry point with a caller-owned scratch buffer
    scratch = torch.zeros(int(lib.dl_sumsq_scratch_bytes()), dtype=torch.uint8, device=DEV)
    for call, (n, with_b) in enumerate([(1_000_003, True), (4_099, False)]):
        a = _rand(n, 11 + call)
        b = _rand(n, 23 + call) if with_b else None
        want = torch.zeros(1, dtype=torch.float64, device=DEV)
        got = torch.full((1,), -1.0, dtype=torch.float64, device=DEV)
        b_ptr = b.data_ptr() if b is not None else None
        _lib.check(lib.dl_sumsq_diff(a.data_ptr(), b_ptr, n, want.data_ptr(), scratch.data_ptr(), stream))
        _lib.check(lib.dl_sumsq_diff_static(a.data_ptr(), b_ptr, n, got.data_ptr(), stream))
        torch.cuda.synchronize()
        # same tolerance as test_local_ops; in fact the two share kernel and block order
        torch.testing.assert_close(got, want, rtol=1e-9, atol=0)
        exact = ((a.double() - (b.double() if b is not None else 0.0)) ** 2).sum().reshape(1)
        torch.testing.assert_close(got, exact, rtol=1e-9, atol=0)


def test_static_scratch_entry_point_rejects_oversized_calls():
    from torch_distlearn_b200 import _lib
    lib = _lib.cuda()
    torch.cuda.set_device(0)
    a = _rand(16, 1)
    out = torch.zeros(1, dtype=torch.float64, device=DEV)
    rc = lib.dl_sumsq_diff_static(a.data_ptr(), None, (1 << 31) + 1, out.data_ptr(),
                                  torch.cuda.current_stream().cuda_stream)
    assert rc == -1
    msg = lib.dl_last_error().decode()
    assert "exceeds the supported maximum" in msg and str(1 << 31) in msg
    with pytest.raises(_lib.NativeError):
        _lib.check(rc)
    # and the entry point still works afterwards
    _lib.check(lib.dl_sumsq_diff_static(a.data_ptr(), None, 16, out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    torch.cuda.synchronize()
    torch.testing.assert_close(out, (a.double() ** 2).sum().reshape(1), rtol=1e-9, atol=0)
