"""The ctypes signature table binds against the built library (a misspelt type or a missing export fails here, on CPU,
instead of on the first GPU call)."""
import pytest


def test_cuda_library_loads_and_every_signature_binds():
    from torch_distlearn_b200 import _lib
    try:
        lib = _lib.cuda()
    except OSError as e:          # no CUDA runtime on this machine: nothing to bind against
        pytest.skip(f"CUDA library not loadable here: {e}")
    assert lib.dl_mnist_form(0) in (1, 2)
    assert lib.dl_bn_work_bytes(64) > 0


def test_host_library_loads():
    from torch_distlearn_b200 import _lib
    assert _lib.host() is not None
