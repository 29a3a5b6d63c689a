"""Symmetric-heap plumbing: fd passing over unix sockets (CPU) and the VMM allocate/export/import/
map path aliased by a torch tensor (single GPU)."""
import ctypes as C
import os
import threading

import pytest
import torch


def test_uds_fd_passing_roundtrip():
    from torch_distlearn_b200 import _lib
    lib = _lib.cuda()   # loads without a GPU: the CUDA runtime is linked statically and initialised lazily
    name = f"dlt-uds-{os.getpid()}".encode()
    sock = C.c_int(-1)
    _lib.check(lib.dl_uds_listen(name, C.byref(sock)))
    r, w = os.pipe()
    got = {}

    def receiver():
        fd, tag = C.c_int(-1), C.c_int(-1)
        _lib.check(lib.dl_uds_recv_fd(sock.value, C.byref(fd), C.byref(tag), 5000))
        got["fd"], got["tag"] = fd.value, tag.value

    th = threading.Thread(target=receiver)
    th.start()
    _lib.check(lib.dl_uds_send_fd(name, w, 42, 5000))
    th.join()
    lib.dl_uds_close(sock.value)
    assert got["tag"] == 42 and got["fd"] >= 0 and got["fd"] != w
    os.write(got["fd"], b"hello")      # the received descriptor is a duplicate of the pipe's write end
    assert os.read(r, 5) == b"hello"
    for fd in (r, w, got["fd"]):
        os.close(fd)


def test_uds_recv_times_out():
    from torch_distlearn_b200 import _lib
    lib = _lib.cuda()
    sock = C.c_int(-1)
    _lib.check(lib.dl_uds_listen(f"dlt-uds-to-{os.getpid()}".encode(), C.byref(sock)))
    fd, tag = C.c_int(-1), C.c_int(-1)
    assert lib.dl_uds_recv_fd(sock.value, C.byref(fd), C.byref(tag), 100) == _lib.DL_ETIMEOUT
    lib.dl_uds_close(sock.value)


@pytest.mark.gpu
def test_vmm_alloc_export_import_map_and_torch_alias():
    from torch_distlearn_b200 import _lib
    from torch_distlearn_b200.parallel.comm import _CudaMem
    lib = _lib.cuda()
    vmm, fdh, mc = C.c_int(0), C.c_int(0), C.c_int(0)
    _lib.check(lib.dl_device_caps(0, C.byref(vmm), C.byref(fdh), C.byref(mc)))
    print("caps: vmm", vmm.value, "fd", fdh.value, "multicast", mc.value)
    if not (vmm.value and fdh.value):
        pytest.skip("VMM with POSIX fd handles not supported here")
    gran = C.c_size_t(0)
    _lib.check(lib.dl_vmm_granularity(0, C.byref(gran)))
    nbytes = max(int(gran.value), 2 << 20)
    handle, fd = C.c_uint64(0), C.c_int(-1)
    _lib.check(lib.dl_vmm_alloc(0, nbytes, C.byref(handle), C.byref(fd)))
    p1 = C.c_void_p()
    _lib.check(lib.dl_vmm_map(0, handle.value, nbytes, C.byref(p1)))
    # import our own export: a second mapping of the same physical memory (what a peer would do)
    h2 = C.c_uint64(0)
    _lib.check(lib.dl_vmm_import(fd.value, C.byref(h2)))
    p2 = C.c_void_p()
    _lib.check(lib.dl_vmm_map(0, h2.value, nbytes, C.byref(p2)))
    a = torch.as_tensor(_CudaMem(p1.value, nbytes), device="cuda:0")
    b = torch.as_tensor(_CudaMem(p2.value, nbytes), device="cuda:0")
    assert a.data_ptr() == p1.value and b.data_ptr() == p2.value
    a.zero_()
    a[:1024].view(torch.float32).fill_(3.5)
    torch.cuda.synchronize()
    assert float(b[:1024].view(torch.float32).sum()) == 3.5 * 256
    del a, b
    _lib.check(lib.dl_vmm_unmap(p2.value, nbytes))
    _lib.check(lib.dl_vmm_release(h2.value))
    _lib.check(lib.dl_vmm_unmap(p1.value, nbytes))
    _lib.check(lib.dl_vmm_release(handle.value))
    lib.dl_close_fd(fd.value)


def test_checkpoint_round_trip_on_cpu(tmp_path):
    """Flat params + center + momentum checkpoint (utils/checkpoint.py) with a CPU communicator: header fields,
    shape manifest, in-place load that keeps the parameter views valid, mismatch detection."""
    import uuid
    import torch_distlearn_b200 as dl
    comm = dl.HostCommunicator(0, 1, name=f"ck-{uuid.uuid4().hex[:8]}", heap_mb=4)
    try:
        model = torch.nn.Sequential(torch.nn.Linear(5, 7), torch.nn.Linear(7, 3))
        flat = dl.FlatParams(comm, model, with_momentum=True)
        center = comm.alloc(flat.padded, torch.float32, "center")
        center.tensor.copy_(flat.params.tensor * 2)
        flat.momentum.tensor.fill_(0.25)
        path = str(tmp_path / "m.dlb200")
        dl.save_checkpoint(path, flat, center=center, momentum=flat.momentum, algo="AllReduceEA", tau=3, alpha=0.4, step=9,
                           world=4, extra={"epoch": 2})
        want_p, want_c = flat.params.tensor.clone(), center.tensor.clone()
        weight_view = model[0].weight
        flat.params.tensor.zero_(); center.tensor.zero_(); flat.momentum.tensor.zero_()
        hdr = dl.load_checkpoint(path, flat, center=center, momentum=flat.momentum)
        assert (hdr["algo"], hdr["tau"], hdr["step"], hdr["world"], hdr["extra"]["epoch"]) == ("AllReduceEA", 3, 9, 4, 2)
        assert [m["name"] for m in hdr["manifest"]] == ["0.weight", "0.bias", "1.weight", "1.bias"]
        assert torch.equal(flat.params.tensor, want_p) and torch.equal(center.tensor, want_c)
        assert float(flat.momentum.tensor[0]) == 0.25
        assert weight_view.data_ptr() == model[0].weight.data_ptr() and torch.equal(model[0].weight.flatten(), want_p[:35])
        other = dl.FlatParams(comm, torch.nn.Linear(4, 4))
        with pytest.raises(ValueError):
            dl.load_checkpoint(path, other)
    finally:
        comm.close()
