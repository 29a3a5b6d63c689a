"""ctypes bindings of the two native libraries (see include/distlearn_b200.h).

``cuda()`` returns libdistlearn_b200.so (kernels + symmetric heap), ``host()`` returns
libdistlearn_host.so (POSIX-shm CPU transport).  Both are built in-tree by
``torch_distlearn_b200.build``; a missing CUDA library on a GPU machine is an error, never a
silent fallback.
"""
from __future__ import annotations

import ctypes as C
import functools

from . import build as _build

DL_MAX_RANKS = 8
DL_MAX_BLOCKS = 160
DL_NUM_CHANNELS = 4
DL_PAD_BYTES = 256 * 1024
DL_ETIMEOUT = -5
DL_EPROTO = -6
DL_FLAG_DEBUG = 1
DL_HOST_RING = 64

VARIANT_ONESHOT, VARIANT_TWOSHOT, VARIANT_NVLS, VARIANT_LL, VARIANT_HYBRID, VARIANT_LIGHT = 0, 1, 2, 3, 4, 5
VARIANT_NAMES = {"oneshot": VARIANT_ONESHOT, "twoshot": VARIANT_TWOSHOT, "nvls": VARIANT_NVLS, "ll": VARIANT_LL,
                 "hybrid": VARIANT_HYBRID, "light": VARIANT_LIGHT}
DT_F32, DT_I32, DT_I64, DT_F64 = 0, 1, 2, 3


class DlComm(C.Structure):
    _fields_ = [
        ("rank", C.c_int32),
        ("world", C.c_int32),
        ("emulate", C.c_int32),
        ("flags", C.c_uint32),
        ("base", C.c_void_p * DL_MAX_RANKS),
        ("mc", C.c_void_p),
        ("timeout_ns", C.c_uint64),
        ("ll_off", C.c_uint64),
        ("ll_slot_bytes", C.c_uint64),
        ("host_words", C.c_void_p),
        ("ticket", C.c_uint32),
        ("nvls_permille", C.c_int32),
    ]


class _RingEntry(C.Structure):
    _fields_ = [("ticket", C.c_uint32), ("n_active", C.c_uint32)]


class DlHostWords(C.Structure):
    """What the kernels leave in pinned host memory (include/distlearn_b200.h: dl_host_words_t)."""
    _fields_ = [
        ("error", C.c_uint32),
        ("error_channel", C.c_uint32),
        ("reserved", C.c_uint32 * 2),
        ("ring", (_RingEntry * DL_HOST_RING) * DL_NUM_CHANNELS),
    ]


class DlResult(C.Structure):
    _fields_ = [
        ("payload", C.c_uint32 * DL_MAX_RANKS),
        ("n_active", C.c_uint32),
        ("error", C.c_uint32),
    ]


class NativeError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"[distlearn_b200 native error {code}] {msg}")
        self.code = code


def _sig(lib, name, restype, argtypes):
    fn = getattr(lib, name)
    fn.restype = restype
    fn.argtypes = argtypes
    return fn


u64, u32, i32, f32, vp, sz = C.c_uint64, C.c_uint32, C.c_int, C.c_float, C.c_void_p, C.c_size_t
P = C.POINTER


@functools.lru_cache(maxsize=None)
def cuda():
    path = _build.cuda_lib_path()
    if not path.exists():
        _build.build_cuda()
    lib = C.CDLL(str(path))
    cp = P(DlComm)
    _sig(lib, "dl_last_error", C.c_char_p, [])
    _sig(lib, "dl_version", i32, [])
    _sig(lib, "dl_kernel_launches", u64, [])
    _sig(lib, "dl_device_caps", i32, [i32, P(i32), P(i32), P(i32)])
    _sig(lib, "dl_vmm_granularity", i32, [i32, P(sz)])
    _sig(lib, "dl_vmm_alloc", i32, [i32, sz, P(u64), P(i32)])
    _sig(lib, "dl_vmm_import", i32, [i32, P(u64)])
    _sig(lib, "dl_vmm_map", i32, [i32, u64, sz, P(vp)])
    _sig(lib, "dl_vmm_unmap", i32, [vp, sz])
    _sig(lib, "dl_vmm_release", i32, [u64])
    _sig(lib, "dl_mc_granularity", i32, [i32, sz, P(sz)])
    _sig(lib, "dl_mc_create", i32, [i32, sz, P(u64), P(i32)])
    _sig(lib, "dl_mc_add_device", i32, [u64, i32])
    _sig(lib, "dl_mc_bind", i32, [u64, u64, sz])
    _sig(lib, "dl_mc_unbind", i32, [u64, i32, sz])
    _sig(lib, "dl_ipc_alloc", i32, [i32, sz, P(vp), vp])
    _sig(lib, "dl_ipc_open", i32, [i32, vp, P(vp)])
    _sig(lib, "dl_ipc_close", i32, [vp])
    _sig(lib, "dl_ipc_free", i32, [vp])
    _sig(lib, "dl_enable_peer", i32, [i32, i32])
    _sig(lib, "dl_can_access_peer", i32, [i32, i32, P(i32)])
    _sig(lib, "dl_memset_sync", i32, [i32, vp, i32, sz])
    _sig(lib, "dl_peak_stream", i32, [cp, i32, i32, u64, u64, u64, i32, vp])
    _sig(lib, "dl_peak_memcpy", i32, [cp, i32, i32, u64, u64, u64, vp])
    _sig(lib, "dl_litmus_scratch_bytes", sz, [i32, i32])
    _sig(lib, "dl_litmus", i32, [cp, i32, i32, u64, i32, u32, i32, vp])
    _sig(lib, "dl_host_words_alloc", i32, [P(vp), P(vp)])
    _sig(lib, "dl_host_words_free", i32, [vp])
    _sig(lib, "dl_uds_listen", i32, [C.c_char_p, P(i32)])
    _sig(lib, "dl_uds_send_fd", i32, [C.c_char_p, i32, i32, i32])
    _sig(lib, "dl_uds_recv_fd", i32, [i32, P(i32), P(i32), i32])
    _sig(lib, "dl_uds_close", i32, [i32])
    _sig(lib, "dl_close_fd", i32, [i32])
    _sig(lib, "dl_allreduce", i32, [cp, i32, u64, u64, i32, i32, u32, i32, i32, i32, vp])
    _sig(lib, "dl_fused_sgd", i32, [cp, u64, u64, u64, u64, f32, f32, f32, i32, i32, i32, u32, i32, i32, i32, vp])
    _sig(lib, "dl_fused_ea", i32, [cp, u64, u64, u64, u64, f32, i32, i32, u32, i32, i32, vp])
    _sig(lib, "dl_broadcast", i32, [cp, u64, u64, u64, i32, i32, u32, i32, i32, vp])
    _sig(lib, "dl_barrier", i32, [cp, i32, u32, i32, vp])
    _sig(lib, "dl_read_result", i32, [cp, i32, P(DlResult), vp])
    _sig(lib, "dl_reset_pad", i32, [cp])
    _sig(lib, "dl_async_ea_client", i32, [cp, i32, vp, u64, u64, f32, i32, i32, vp])
    _sig(lib, "dl_peer_pull", i32, [cp, i32, u64, vp, u64, i32, vp])
    _sig(lib, "dl_async_sync_count", i32, [cp, i32, P(u64), vp])
    _sig(lib, "dl_async_ctl_snapshot", i32, [cp, i32, P(u64), vp])
    _sig(lib, "dl_mailbox_write", i32, [cp, i32, i32, u64, vp])
    _sig(lib, "dl_mailbox_read", i32, [cp, i32, i32, P(u64), vp])
    _sig(lib, "dl_tc_gemm", i32, [vp, vp, vp, vp, i32, i32, i32, C.c_longlong, C.c_longlong, C.c_longlong, i32, i32, i32, vp])
    _sig(lib, "dl_tc_gemm_ex", i32, [vp, vp, vp, vp, i32, i32, i32, C.c_longlong, C.c_longlong, C.c_longlong, i32, i32, i32, i32, i32, vp])
    _sig(lib, "dl_tc_conv5x5", i32, [i32, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp])
    _sig(lib, "dl_tc_conv5x5_ws", i32, [i32, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, i32, i32, i32, vp])
    _sig(lib, "dl_split_tf32", i32, [vp, vp, vp, C.c_longlong, vp])
    _sig(lib, "dl_tc_gemm_streamk", i32, [i32])
    _sig(lib, "dl_tc_streamk_plan", i32, [i32, i32, i32, i32, i32, P(i32), P(i32), i32])
    _sig(lib, "dl_bn_work_bytes", sz, [i32])
    _sig(lib, "dl_bn_relu_pool_fwd", i32, [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, f32, f32, vp, vp])
    _sig(lib, "dl_bn_relu_pool_bwd", i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp])
    _sig(lib, "dl_bn_relu_pool_bwd_bias", i32, [vp, vp, vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp, vp])
    _sig(lib, "dl_colsum", i32, [vp, vp, C.c_longlong, i32, i32, vp, vp])
    _sig(lib, "dl_head_fwd", i32, [vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, vp])
    _sig(lib, "dl_head_bwd", i32, [vp, vp, vp, vp, vp, vp, vp, vp, i32, i32, i32, i32, i32, vp])
    _sig(lib, "dl_im2col5x5", i32, [vp, vp, i32, i32, i32, i32, C.c_longlong, vp])
    _sig(lib, "dl_col2im5x5", i32, [vp, vp, i32, i32, i32, i32, C.c_longlong, vp])
    _sig(lib, "dl_mnist_step", i32, [vp, vp, P(vp), P(vp), vp, vp, i32, f32, i32, vp])
    _sig(lib, "dl_mnist_form", i32, [i32])
    _sig(lib, "dl_local_sgd", i32, [vp, vp, vp, u64, f32, f32, f32, i32, f32, i32, vp])
    _sig(lib, "dl_local_sgd_from", i32, [vp, vp, vp, vp, u64, f32, f32, f32, i32, f32, i32, vp])
    _sig(lib, "dl_async_ea_client_out", i32, [cp, i32, vp, vp, u64, u64, f32, i32, i32, vp])
    _sig(lib, "dl_local_ea", i32, [vp, vp, u64, f32, vp])
    _sig(lib, "dl_sumsq_scratch_bytes", sz, [])
    _sig(lib, "dl_sumsq_diff", i32, [vp, vp, u64, vp, vp, vp])
    _sig(lib, "dl_confusion_add", i32, [vp, vp, i32, i32, vp, vp])
    _sig(lib, "dl_sumsq_static_scratch_bytes", sz, [])
    _sig(lib, "dl_sumsq_static_max_n", u64, [])
    _sig(lib, "dl_sumsq_diff_static", i32, [vp, vp, u64, vp, vp])
    return lib


@functools.lru_cache(maxsize=None)
def host():
    path = _build.host_lib_path()
    if not path.exists():
        _build.build_host()
    lib = C.CDLL(str(path))
    rp = P(DlResult)
    _sig(lib, "dlh_last_error", C.c_char_p, [])
    _sig(lib, "dlh_open", i32, [C.c_char_p, i32, i32, u64, C.c_double, P(vp)])
    _sig(lib, "dlh_close", i32, [vp, i32])
    _sig(lib, "dlh_base", vp, [vp, i32])
    _sig(lib, "dlh_set_timeout", i32, [vp, C.c_double])
    _sig(lib, "dlh_barrier", i32, [vp, i32, u32, i32, rp])
    _sig(lib, "dlh_allreduce", i32, [vp, i32, u64, u64, i32, u32, i32, i32, rp])
    _sig(lib, "dlh_fused_sgd", i32, [vp, u64, u64, u64, u64, f32, f32, f32, i32, i32, u32, i32, i32, rp])
    _sig(lib, "dlh_fused_ea", i32, [vp, u64, u64, u64, u64, f32, i32, u32, i32, rp])
    _sig(lib, "dlh_broadcast", i32, [vp, u64, u64, u64, i32, u32, i32, rp])
    _sig(lib, "dlh_async_ea_client", i32, [vp, i32, vp, u64, u64, f32, i32])
    _sig(lib, "dlh_debug_take_ticket", i32, [vp, i32, P(u32)])
    _sig(lib, "dlh_async_sync_count", i32, [vp, i32, P(u64)])
    _sig(lib, "dlh_mailbox_write", i32, [vp, i32, i32, u64])
    _sig(lib, "dlh_mailbox_read", i32, [vp, i32, i32, P(u64)])
    _sig(lib, "dlh_peer_pull", i32, [vp, i32, u64, vp, u64])
    # native batch loader (csrc/loader.cpp)
    i64 = C.c_int64
    _sig(lib, "dll_last_error", C.c_char_p, [])
    _sig(lib, "dll_create", i32, [vp, i32, vp, i64, i32, i32, vp, vp, i32, i32, i32, i32, i32, u64, i32, i32, P(vp), P(vp), P(vp)])
    _sig(lib, "dll_next", i32, [vp, P(i32), P(i32), P(i64), C.c_double])
    _sig(lib, "dll_release", i32, [vp])
    _sig(lib, "dll_num_batches", i64, [vp])
    _sig(lib, "dll_num_samples", i64, [vp])
    _sig(lib, "dll_samples_done", u64, [vp])
    _sig(lib, "dll_destroy", i32, [vp])
    _sig(lib, "dll_read_cifar_bin", i64, [C.c_char_p, vp, vp, i64])
    return lib


def check(rc: int) -> None:
    """Raise on a negative return code of libdistlearn_b200."""
    if rc < 0:
        raise NativeError(rc, cuda().dl_last_error().decode(errors="replace"))


def check_host(rc: int) -> None:
    if rc < 0:
        raise NativeError(rc, host().dlh_last_error().decode(errors="replace"))
