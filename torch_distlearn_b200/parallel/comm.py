"""Communicators: the transport layer under AllReduceSGD / AllReduceEA / AsyncEA.

The reference's transport is ``ipc.Tree`` / ``ipc.LocalhostTree`` (a k-ary TCP tree:
test/test_AllReduceSGD.lua:7, examples/mnist.lua:16).  Here a communicator owns a *symmetric
heap* — the same buffer layout on every rank, each rank's heap addressable by all the others —
and runs single-launch collectives over it:

* :class:`P2PCommunicator`   — CUDA: heaps are VMM allocations mapped across processes over
  NVLink (multicast/NVLS when the fabric exposes it, cudaIpc as fallback); collectives are the
  sm_100a kernels in ``csrc/``.
* :class:`HostCommunicator`  — CPU: heaps live in one POSIX shared-memory segment; the same
  algorithms run in C++ (``csrc/host_transport.cpp``).
* :class:`EmulatedCommunicator` — all ranks on ONE GPU, one cooperative launch plays every rank;
  used to test the multi-rank kernels on a single-GPU box.

``rank`` is 0-based; ``node_index`` (= rank + 1) is the reference's 1-based ``nodeIndex``.
"""
from __future__ import annotations

import ctypes as C
import json
import os
import threading
import uuid
from dataclasses import dataclass
from typing import List, Optional

import torch

from .. import _lib
from .._lib import (DL_PAD_BYTES, DlComm, DlHostWords, DlResult, VARIANT_HYBRID, VARIANT_LIGHT, VARIANT_LL, VARIANT_NAMES,
                    VARIANT_NVLS, VARIANT_ONESHOT, VARIANT_TWOSHOT)

VARIANT_TEXT = {VARIANT_ONESHOT: "oneshot", VARIANT_TWOSHOT: "twoshot", VARIANT_NVLS: "nvls", VARIANT_LL: "ll",
                VARIANT_HYBRID: "hybrid", VARIANT_LIGHT: "light"}
_ALIGN = 512  # byte alignment of every symmetric allocation (>= 16 for 128-bit access)

_DTYPE_CODES = {torch.float32: _lib.DT_F32, torch.int32: _lib.DT_I32, torch.int64: _lib.DT_I64,
                torch.float64: _lib.DT_F64}


class LazyCount:
    """The contributor count ``n`` of a collective (lua/AllReduceSGD.lua:20: ``_, n = allReduce(...)``) as a value that
    is read when it is USED.  The kernel leaves ``n`` in pinned host memory under this call's ticket; converting to
    ``int`` waits for the event recorded behind the launch and reads that word — no device-to-host copy, and no
    synchronisation at all for callers that never look at it.  Behaves like an int in arithmetic and comparisons."""

    __slots__ = ("_comm", "_channel", "_ticket", "_event", "_value")

    def __init__(self, comm, channel: int, ticket: int, event):
        self._comm, self._channel, self._ticket, self._event, self._value = comm, channel, ticket, event, None

    def __int__(self) -> int:
        if self._value is None:
            if self._event is not None:
                self._event.synchronize()
            self._value = self._comm._read_ring(self._channel, self._ticket)
            self._comm.poll_error()
        return self._value

    __index__ = __int__

    def __repr__(self) -> str:
        return str(int(self))

    def __eq__(self, other):
        return int(self) == other

    def __hash__(self):
        return hash(int(self))

    def __lt__(self, other): return int(self) < other
    def __le__(self, other): return int(self) <= other
    def __gt__(self, other): return int(self) > other
    def __ge__(self, other): return int(self) >= other
    def __add__(self, other): return int(self) + other
    __radd__ = __add__
    def __sub__(self, other): return int(self) - other
    def __rsub__(self, other): return other - int(self)
    def __mul__(self, other): return int(self) * other
    __rmul__ = __mul__
    def __truediv__(self, other): return int(self) / other
    def __rtruediv__(self, other): return other / int(self)
    def __bool__(self): return int(self) != 0
    def __float__(self): return float(int(self))


@dataclass
class CommResult:
    """What a collective's arrival barrier learned (lua/AllReduceSGD.lua:20: ``_, n = allReduce``)."""
    n_active: int
    payloads: List[int]


class SymBuffer:
    """A symmetric allocation: same offset in every rank's heap; ``tensor`` is the local view."""

    def __init__(self, comm: "Communicator", offset: int, numel: int, dtype: torch.dtype, tensor: torch.Tensor,
                 name: str = ""):
        self.comm, self.offset, self.numel, self.dtype, self.tensor, self.name = comm, offset, numel, dtype, tensor, name

    @property
    def nbytes(self) -> int:
        return self.numel * self.tensor.element_size()

    def sub(self, start: int, numel: int) -> "SymBuffer":
        """A view of ``numel`` elements starting at element ``start`` (must keep 16-byte alignment)."""
        es = self.tensor.element_size()
        if (start * es) % 16 or start < 0 or start + numel > self.numel:
            raise ValueError("SymBuffer.sub: range must be inside the buffer and 16-byte aligned")
        return SymBuffer(self.comm, self.offset + start * es, numel, self.dtype, self.tensor[start:start + numel], self.name)

    def __repr__(self) -> str:
        return f"SymBuffer({self.name!r}, off={self.offset}, numel={self.numel}, dtype={self.dtype})"


def _round_up(x: int, a: int) -> int:
    return (x + a - 1) // a * a


class _VariantTable:
    """Message-size -> algorithm variant, per op.  Defaults are replaced by the measured table
    that ``bench/sweep.py`` writes (``variant_table.json``); ``DL_B200_VARIANT`` forces one."""

    DEFAULT = {"allreduce": 512 * 1024, "fused_sgd": 512 * 1024, "broadcast": 256 * 1024}
    OPS = ("allreduce", "fused_sgd", "fused_ea", "broadcast")

    def __init__(self, world: int = 8):
        self.oneshot_max = dict(self.DEFAULT)
        self.nvls_min = {op: 0 for op in self.OPS}
        # ll_max: largest message that takes the push/no-barrier path (capped by the staging area the communicator
        # could afford); hybrid_min: smallest message that takes the pipelined switch+P2P kernel (broadcast has
        # none); hybrid_permille: share of every slice reduced in the switch (needs multicast; 0 = all P2P)
        self.ll_max = {op: 64 * 1024 for op in self.OPS}
        self.hybrid_min = {"allreduce": 1 << 20, "fused_sgd": 1 << 20, "fused_ea": 1 << 20}
        self.hybrid_permille = {op: (750 if world > 2 else 0) for op in self.OPS}
        self.source = "defaults"
        path = os.environ.get("DL_B200_VARIANT_TABLE") or os.path.join(os.path.dirname(__file__), "variant_table.json")
        if os.path.exists(path):
            try:
                data = json.load(open(path))
                keys = sorted(int(k) for k in data if k.isdigit())
                if keys:      # per-world tables: take the smallest measured world >= ours (else the largest)
                    pick = next((k for k in keys if k >= world), keys[-1])
                    data = data[str(pick)]
                for field in ("oneshot_max", "nvls_min", "ll_max", "hybrid_min", "hybrid_permille"):
                    entry = data.get(field, {})
                    if not isinstance(entry, dict):
                        raise ValueError(f"{field} must be an object")
                    getattr(self, field).update({k: int(v) for k, v in entry.items()})
                self.source = path
            except (OSError, ValueError, TypeError, KeyError) as e:
                # a damaged table must not pass silently for a measured one: say so once, keep the defaults
                import warnings
                warnings.warn(f"distlearn_b200: ignoring variant table {path}: {e!r}; using built-in defaults")
        # per-bucket (light) launches: all of it through the switch where that wins beside the convolutions (measured at 8
        # GPUs); at 2 GPUs the plain P2P form is the measured default.  DL_B200_LIGHT_PERMILLE overrides at any world.
        self.light_permille = int(os.environ.get("DL_B200_LIGHT_PERMILLE", "0" if world == 2 else "1000"))
        self.forced = os.environ.get("DL_B200_VARIANT")
        if os.environ.get("DL_B200_NVLS_PERMILLE"):
            self.hybrid_permille = {op: int(os.environ["DL_B200_NVLS_PERMILLE"]) for op in self.OPS}

    def pick(self, op: str, nbytes: int, world: int, has_mc: bool, requested: str = "auto", ll_cap: int = 0,
             aligned32: bool = False) -> int:
        req = self.forced or requested
        if req and req != "auto":
            v = VARIANT_NAMES[req]
            if v == VARIANT_NVLS and not has_mc:
                v = VARIANT_TWOSHOT
            if op == "fused_ea" and v == VARIANT_ONESHOT:
                v = VARIANT_TWOSHOT
            if v == VARIANT_LL and nbytes > ll_cap:
                v = VARIANT_TWOSHOT
            if v == VARIANT_HYBRID and (not aligned32 or op == "broadcast"):
                v = VARIANT_NVLS if (has_mc and world > 2) else VARIANT_TWOSHOT
            if v == VARIANT_LIGHT and (not aligned32 or op != "fused_sgd"):
                v = VARIANT_TWOSHOT
            return v
        if world > 1 and nbytes <= min(self.ll_max.get(op, 0), ll_cap):
            return VARIANT_LL
        if op != "fused_ea" and nbytes <= self.oneshot_max.get(op, 0):
            return VARIANT_ONESHOT
        if world > 1 and aligned32 and op in self.hybrid_min and nbytes >= self.hybrid_min[op]:
            return VARIANT_HYBRID
        if has_mc and world > 2 and nbytes >= self.nvls_min.get(op, 0):
            return VARIANT_NVLS
        return VARIANT_TWOSHOT

    def permille(self, op: str, has_mc: bool) -> int:
        return max(0, min(1000, self.hybrid_permille.get(op, 0))) if has_mc else 0


class Communicator:
    """Common front end; subclasses supply the heap and the native calls."""

    backend = "abstract"

    def __init__(self, rank: int, world: int, device: torch.device, heap_bytes: int):
        if not (1 <= world <= _lib.DL_MAX_RANKS):
            raise ValueError(f"world size {world} unsupported (1..{_lib.DL_MAX_RANKS}: one NVSwitch domain)")
        self.rank, self.world, self.device = rank, world, device
        self.node_index, self.num_nodes = rank + 1, world
        self.heap_bytes = heap_bytes
        self._cursor = DL_PAD_BYTES
        self._variants = _VariantTable(world)
        self.launches = 0  # native collective launches issued through this communicator
        self.max_blocks = int(os.environ.get("DL_B200_MAX_BLOCKS", "0"))

    # -- allocation -------------------------------------------------------------------------------
    def _local_bytes(self) -> torch.Tensor:  # uint8 view of the whole local heap
        raise NotImplementedError

    def alloc(self, numel: int, dtype: torch.dtype = torch.float32, name: str = "") -> SymBuffer:
        """Symmetric allocation — every rank must call alloc in the same order with the same sizes."""
        es = torch.empty((), dtype=dtype).element_size()
        nbytes = _round_up(max(numel, 1) * es, 16)
        off = _round_up(self._cursor, _ALIGN)
        if off + nbytes > self.heap_bytes:
            raise MemoryError(f"symmetric heap exhausted: need {off + nbytes} bytes, have {self.heap_bytes} "
                              f"(raise heap_mb when creating the communicator)")
        self._cursor = off + nbytes
        t = self._local_bytes()[off:off + numel * es].view(dtype)
        t.zero_()
        return SymBuffer(self, off, numel, dtype, t, name)

    def padded(self, numel: int) -> int:
        """Round an element count up so that float4 access and slicing stay aligned."""
        return _round_up(max(numel, 4), 32)

    # -- results ------------------------------------------------------------------------------------
    def _mk_result(self, res: DlResult) -> CommResult:
        return CommResult(int(res.n_active), [int(res.payload[r]) for r in range(self.world)])

    def pick_variant(self, op: str, nbytes: int, requested: str = "auto", offsets=(), n: int = 8) -> int:
        aligned32 = n % 8 == 0 and all(o % 32 == 0 for o in offsets)
        return self._variants.pick(op, nbytes, self.world, self.has_multicast, requested, self.ll_max_bytes, aligned32)

    has_multicast = False
    ll_max_bytes = 0            # largest message the small-message staging area can take (0: none)

    def ctl_snapshot(self, rank: int):
        """(sync counter, the 32 mailbox words) of ``rank``'s async-EASGD control block.  Backends that can fetch them
        in one go override this; the default asks word by word."""
        return self.sync_count(rank), [self.mailbox_read(rank, i) for i in range(32)]

    def close(self) -> None:
        pass

    def __enter__(self):
        return self

    def __exit__(self, *exc):
        self.close()


# =====================================================================================================
# CPU: POSIX shared memory
# =====================================================================================================
class HostCommunicator(Communicator):
    """Same-host CPU ranks over one shared-memory segment (``csrc/host_transport.cpp``).

    Replaces ``ipc.LocalhostTree(nodeIndex, numNodes)`` for CPU jobs (examples/mnist.lua:16)."""

    backend = "host"

    def __init__(self, rank: int, world: int, name: str, heap_mb: float = 64, timeout_s: float = 60.0):
        heap_bytes = _round_up(int(heap_mb * (1 << 20)) + DL_PAD_BYTES, 4096)
        super().__init__(rank, world, torch.device("cpu"), heap_bytes)
        self._lib = _lib.host()
        self._name = ("/" + name.strip("/")).encode()
        h = C.c_void_p()
        _lib.check_host(self._lib.dlh_open(self._name, rank, world, heap_bytes, float(timeout_s), C.byref(h)))
        self._h = h
        base = self._lib.dlh_base(h, rank)
        self._bytes = torch.frombuffer((C.c_char * heap_bytes).from_address(base), dtype=torch.uint8)
        self._res = DlResult()
        self._closed = False

    def _local_bytes(self) -> torch.Tensor:
        return self._bytes

    def set_timeout(self, seconds: float) -> None:
        self._lib.dlh_set_timeout(self._h, float(seconds))

    def barrier(self, active: int = 1, payload: int = 0, channel: int = 0, read: bool = True) -> CommResult:
        _lib.check_host(self._lib.dlh_barrier(self._h, int(active), int(payload), channel, C.byref(self._res)))
        return self._mk_result(self._res)

    def allreduce(self, buf: SymBuffer, n: Optional[int] = None, active: int = 1, payload: int = 0,
                  normalize: bool = False, variant: str = "auto", channel: int = 0, read: bool = True):
        n = buf.numel if n is None else n
        self.launches += 1
        _lib.check_host(self._lib.dlh_allreduce(self._h, _DTYPE_CODES[buf.dtype], buf.offset, n, int(active), int(payload),
                                                int(normalize), channel, C.byref(self._res)))
        return self._mk_result(self._res)

    def fused_sgd(self, grads: SymBuffer, params: SymBuffer, momentum_buf: Optional[SymBuffer], n: int, lr: float,
                  momentum: float = 0.0, weight_decay: float = 0.0, nesterov: bool = False, active: int = 1,
                  payload: int = 0, zero_grads: bool = True, variant: str = "auto", channel: int = 0, read: bool = True):
        self.launches += 1
        _lib.check_host(self._lib.dlh_fused_sgd(self._h, grads.offset, params.offset, momentum_buf.offset if momentum_buf else 0,
                                                n, lr, momentum, weight_decay, int(nesterov), int(active), int(payload),
                                                int(zero_grads), channel, C.byref(self._res)))
        return self._mk_result(self._res)

    def fused_ea(self, x: SymBuffer, c_cur: SymBuffer, c_next: SymBuffer, n: int, alpha: float, active: int = 1,
                 payload: int = 0, variant: str = "auto", channel: int = 0, read: bool = True):
        self.launches += 1
        _lib.check_host(self._lib.dlh_fused_ea(self._h, x.offset, c_cur.offset, c_next.offset, n, alpha, int(active),
                                               int(payload), channel, C.byref(self._res)))
        return self._mk_result(self._res)

    def broadcast(self, buf: SymBuffer, n: Optional[int] = None, root: int = 0, also: Optional[SymBuffer] = None,
                  payload: int = 0, variant: str = "auto", channel: int = 0, read: bool = True):
        n = buf.numel if n is None else n
        self.launches += 1
        _lib.check_host(self._lib.dlh_broadcast(self._h, buf.offset, also.offset if also else 0, n, root, int(payload),
                                                channel, C.byref(self._res)))
        return self._mk_result(self._res)

    def read_result(self, channel: int = 0) -> CommResult:
        return self._mk_result(self._res)

    # async EASGD primitives
    def async_client_sync(self, server: int, x: torch.Tensor, center: SymBuffer, n: int, alpha: float, exclusive: bool = False):
        assert x.dtype == torch.float32 and x.is_contiguous() and x.device.type == "cpu"
        self.launches += 1
        _lib.check_host(self._lib.dlh_async_ea_client(self._h, server, x.data_ptr(), center.offset, n, alpha, int(exclusive)))

    def peer_pull(self, src_rank: int, buf: SymBuffer, dst: torch.Tensor, n: Optional[int] = None):
        n = buf.numel if n is None else n
        assert dst.dtype == torch.float32 and dst.is_contiguous()
        _lib.check_host(self._lib.dlh_peer_pull(self._h, src_rank, buf.offset, dst.data_ptr(), n))

    def sync_count(self, server: int) -> int:
        v = C.c_uint64()
        _lib.check_host(self._lib.dlh_async_sync_count(self._h, server, C.byref(v)))
        return int(v.value)

    def mailbox_write(self, rank: int, index: int, value: int) -> None:
        _lib.check_host(self._lib.dlh_mailbox_write(self._h, rank, index, int(value)))

    def mailbox_read(self, rank: int, index: int) -> int:
        v = C.c_uint64()
        _lib.check_host(self._lib.dlh_mailbox_read(self._h, rank, index, C.byref(v)))
        return int(v.value)

    def synchronize(self) -> None:
        pass

    def close(self) -> None:
        if not self._closed:
            self._closed = True
            self._bytes = None
            self._lib.dlh_close(self._h, 1 if self.rank == 0 else 0)


# =====================================================================================================
# CUDA: symmetric heap over NVLink
# =====================================================================================================
class _CudaMem:
    """Exposes a raw device range through __cuda_array_interface__ so torch can alias it."""

    def __init__(self, ptr: int, nbytes: int):
        self.__cuda_array_interface__ = {"shape": (nbytes,), "typestr": "|u1", "data": (ptr, False), "version": 2,
                                         "strides": None}


_NVTX = os.environ.get("DL_B200_NVTX", "0") == "1"


class _Range:
    """NVTX range around a collective launch when DL_B200_NVTX=1 (the reference has no tracing at all);
    shows up in Nsight Systems timelines next to the kernel names."""

    def __init__(self, name: str):
        self.name = name

    def __enter__(self):
        if _NVTX:
            torch.cuda.nvtx.range_push(self.name)

    def __exit__(self, *exc):
        if _NVTX:
            torch.cuda.nvtx.range_pop()


def _stream_ptr(device: torch.device, stream: Optional[torch.cuda.Stream] = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream(device)
    return s.cuda_stream


class _CudaOps(Communicator):
    """Native-call layer shared by the real and the emulated CUDA communicators."""

    def __init__(self, rank, world, device, heap_bytes):
        super().__init__(rank, world, device, heap_bytes)
        self._lib = _lib.cuda()
        self._c = DlComm()
        self._c.rank, self._c.world, self._c.emulate = rank, world, 0
        self._c.timeout_ns = int(float(os.environ.get("DL_B200_TIMEOUT_S", "20")) * 1e9)
        if os.environ.get("DL_B200_DEBUG", "0") == "1":
            self._c.flags |= _lib.DL_FLAG_DEBUG
        self._res = DlResult()
        self._ticket = 0
        self._host_words = None
        self._host_ptr = None
        hp, dp = C.c_void_p(), C.c_void_p()
        with torch.cuda.device(device):
            _lib.check(self._lib.dl_host_words_alloc(C.byref(hp), C.byref(dp)))
        self._host_ptr = hp.value
        self._host_words = DlHostWords.from_address(hp.value)
        self._c.host_words = dp.value

    def _nb(self, nblocks: int) -> int:
        return nblocks if nblocks else self.max_blocks

    # -- small-message staging, tickets, errors ---------------------------------------------------------
    def _setup_ll(self) -> None:
        """Carve the DL_VARIANT_LL staging area out of the heap (first allocation: same offset on every rank).  It has
        to take twice the payload per (channel, parity, source); the limit shrinks with small heaps."""
        if self.world < 2:
            return
        limit = int(float(os.environ.get("DL_B200_LL_MAX_KB", "64")) * 1024)
        budget = self.heap_bytes // 8
        slots = _lib.DL_NUM_CHANNELS * 2 * self.world
        while limit >= 4096 and slots * (2 * limit + 512) > budget:
            limit //= 2
        if limit < 4096:
            return
        slot_bytes = _round_up(2 * limit + 16, 512)
        buf = self.alloc(slots * slot_bytes // 4, torch.float32, "ll.staging")
        self._c.ll_off, self._c.ll_slot_bytes = buf.offset, slot_bytes
        self.ll_max_bytes = limit

    def set_debug(self, on: bool = True) -> None:
        """Barriers check that every slot holds an epoch the protocol can produce (DL_EPROTO otherwise)."""
        self._c.flags = (self._c.flags | _lib.DL_FLAG_DEBUG) if on else (self._c.flags & ~_lib.DL_FLAG_DEBUG)

    def _begin(self, op: str) -> int:
        self._ticket = (self._ticket + 1) & 0xFFFFFFFF
        self._c.ticket = self._ticket
        self._c.nvls_permille = self._variants.permille(op, self.has_multicast)
        return self._ticket

    def _read_ring(self, channel: int, ticket: int) -> int:
        e = self._host_words.ring[channel][ticket % _lib.DL_HOST_RING]
        if e.ticket != ticket:
            raise RuntimeError("contributor count no longer available: read it within "
                               f"{_lib.DL_HOST_RING} collectives of the call that produced it")
        return int(e.n_active)

    def lazy_count(self, channel: int, ticket: int, stream=None) -> LazyCount:
        ev = torch.cuda.Event()
        ev.record(stream if stream is not None else torch.cuda.current_stream(self.device))
        return LazyCount(self, channel, ticket, ev)

    def poll_error(self) -> None:
        """Raise if any kernel of this communicator gave up waiting for a peer.  A plain read of pinned host memory —
        cheap enough for every step.  The word is sticky on the device side; reporting clears it."""
        hw = self._host_words
        if hw is not None and hw.error:
            code, ch = int(hw.error), int(hw.error_channel)
            hw.error = 0
            what = "timed out waiting for a peer" if code == -_lib.DL_ETIMEOUT else "found a barrier slot the protocol cannot produce"
            raise _lib.NativeError(-code, f"a collective on channel {ch} {what} (rank {self.rank} of {self.world}); "
                                          "replicas may have diverged: re-synchronise the parameters")

    def set_timeout(self, seconds: float) -> None:
        self._c.timeout_ns = int(seconds * 1e9)

    def barrier(self, active: int = 1, payload: int = 0, channel: int = 0, read: bool = True, stream=None):
        self.last_ticket = self._begin("barrier")
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_barrier(C.byref(self._c), int(active), int(payload), channel, _stream_ptr(self.device, stream)))
            self.launches += 1
            return self.read_result(channel, stream) if read else None

    def allreduce(self, buf: SymBuffer, n: Optional[int] = None, active: int = 1, payload: int = 0,
                  normalize: bool = False, variant: str = "auto", channel: int = 0, read: bool = False, nblocks: int = 0,
                  stream=None):
        n = buf.numel if n is None else n
        nbytes = n * buf.tensor.element_size()
        v = self.pick_variant("allreduce", nbytes, variant, (buf.offset,), n)
        if buf.dtype != torch.float32 and v not in (VARIANT_ONESHOT, VARIANT_LL):
            v = VARIANT_ONESHOT
        self.last_ticket = self._begin("allreduce")
        with torch.cuda.device(self.device), _Range(f"distlearn.allreduce[{VARIANT_TEXT[v]}]"):
            _lib.check(self._lib.dl_allreduce(C.byref(self._c), _DTYPE_CODES[buf.dtype], buf.offset, n, v, int(active),
                                              int(payload), int(normalize), self._nb(nblocks), channel,
                                              _stream_ptr(self.device, stream)))
            self.launches += 1
            return self.read_result(channel, stream) if read else None

    def fused_sgd(self, grads: SymBuffer, params: SymBuffer, momentum_buf: Optional[SymBuffer], n: int, lr: float,
                  momentum: float = 0.0, weight_decay: float = 0.0, nesterov: bool = False, active: int = 1,
                  payload: int = 0, zero_grads: bool = True, variant: str = "auto", channel: int = 0, read: bool = False,
                  nblocks: int = 0, stream=None):
        v = self.pick_variant("fused_sgd", n * 4, variant,
                              (grads.offset, params.offset, momentum_buf.offset if momentum_buf else 0), n)
        self.last_ticket = self._begin("fused_sgd")
        if v == VARIANT_LIGHT:
            # Beside the convolutions the switch path is the one that pays: a multimem load is one instruction per 16 bytes
            # of EIGHT ranks' data, where the P2P form spends eight loads and eight stores of the SMs it shares (8 GPUs,
            # per-layer buckets: 898k samples/s all-switch, 830k at 750/1000, 821k all-P2P, 879k without buckets).
            self._c.nvls_permille = self._variants.light_permille if self.has_multicast else 0
        with torch.cuda.device(self.device), _Range(f"distlearn.fused_sgd[{VARIANT_TEXT[v]}]"):
            _lib.check(self._lib.dl_fused_sgd(C.byref(self._c), grads.offset, params.offset,
                                              momentum_buf.offset if momentum_buf else 0, n, lr, momentum, weight_decay,
                                              int(nesterov), v, int(active), int(payload), int(zero_grads),
                                              self._nb(nblocks), channel, _stream_ptr(self.device, stream)))
            self.launches += 1
            return self.read_result(channel, stream) if read else None

    def fused_ea(self, x: SymBuffer, c_cur: SymBuffer, c_next: SymBuffer, n: int, alpha: float, active: int = 1,
                 payload: int = 0, variant: str = "auto", channel: int = 0, read: bool = False, nblocks: int = 0,
                 stream=None):
        v = self.pick_variant("fused_ea", n * 4, variant, (x.offset, c_cur.offset, c_next.offset), n)
        self.last_ticket = self._begin("fused_ea")
        with torch.cuda.device(self.device), _Range(f"distlearn.fused_ea[{VARIANT_TEXT[v]}]"):
            _lib.check(self._lib.dl_fused_ea(C.byref(self._c), x.offset, c_cur.offset, c_next.offset, n, alpha, v,
                                             int(active), int(payload), self._nb(nblocks), channel,
                                             _stream_ptr(self.device, stream)))
            self.launches += 1
            return self.read_result(channel, stream) if read else None

    def broadcast(self, buf: SymBuffer, n: Optional[int] = None, root: int = 0, also: Optional[SymBuffer] = None,
                  payload: int = 0, variant: str = "auto", channel: int = 0, read: bool = False, nblocks: int = 0,
                  stream=None):
        n = buf.numel if n is None else n
        v = self.pick_variant("broadcast", n * 4, variant, (buf.offset, also.offset if also else 0), n)
        self.last_ticket = self._begin("broadcast")
        with torch.cuda.device(self.device), _Range(f"distlearn.broadcast[{VARIANT_TEXT[v]}]"):
            _lib.check(self._lib.dl_broadcast(C.byref(self._c), buf.offset, also.offset if also else 0, n, root, v,
                                              int(payload), self._nb(nblocks), channel, _stream_ptr(self.device, stream)))
            self.launches += 1
            return self.read_result(channel, stream) if read else None

    def read_result(self, channel: int = 0, stream=None) -> CommResult:
        """Host-synchronising read of the last collective's result words on ``channel``."""
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_read_result(C.byref(self._c), channel, C.byref(self._res), _stream_ptr(self.device, stream)))
        return self._mk_result(self._res)

    def async_client_sync(self, server: int, x: torch.Tensor, center: SymBuffer, n: int, alpha: float,
                          exclusive: bool = False, nblocks: int = 0, stream=None, x_out: Optional[torch.Tensor] = None):
        """``x_out``: write the moved parameters there and leave ``x`` alone (so the sync can run under a forward /
        backward pass that still reads ``x``)."""
        assert x.dtype == torch.float32 and x.is_contiguous() and x.is_cuda
        out = x if x_out is None else x_out
        assert out.dtype == torch.float32 and out.is_contiguous() and out.is_cuda and out.numel() >= n
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_async_ea_client_out(C.byref(self._c), server, x.data_ptr(), out.data_ptr(), center.offset, n,
                                                        alpha, int(exclusive), self._nb(nblocks), _stream_ptr(self.device, stream)))
        self.launches += 1

    def peer_pull(self, src_rank: int, buf: SymBuffer, dst: torch.Tensor, n: Optional[int] = None, stream=None):
        n = buf.numel if n is None else n
        assert dst.dtype == torch.float32 and dst.is_contiguous() and dst.is_cuda
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_peer_pull(C.byref(self._c), src_rank, buf.offset, dst.data_ptr(), n, 0,
                                              _stream_ptr(self.device, stream)))
        self.launches += 1

    def sync_count(self, server: int) -> int:
        v = C.c_uint64()
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_async_sync_count(C.byref(self._c), server, C.byref(v), _stream_ptr(self.device)))
        return int(v.value)

    def ctl_snapshot(self, rank: int):
        """One small device-to-host copy on a private stream (never queued behind the training stream's kernels)."""
        if self.__dict__.get("_poll_stream") is None:
            self._poll_stream = torch.cuda.Stream(self.device)
            self._poll_buf = (C.c_uint64 * 33)()
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_async_ctl_snapshot(C.byref(self._c), rank, self._poll_buf, self._poll_stream.cuda_stream))
        return int(self._poll_buf[0]), [int(self._poll_buf[1 + i]) for i in range(32)]

    def mailbox_write(self, rank: int, index: int, value: int) -> None:
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_mailbox_write(C.byref(self._c), rank, index, int(value), _stream_ptr(self.device)))

    def mailbox_read(self, rank: int, index: int) -> int:
        v = C.c_uint64()
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_mailbox_read(C.byref(self._c), rank, index, C.byref(v), _stream_ptr(self.device)))
        return int(v.value)

    def synchronize(self) -> None:
        torch.cuda.synchronize(self.device)

    # -- memory-model litmus tests (csrc/litmus.cu) ---------------------------------------------------------
    def litmus(self, shape: str = "mp", fenced: bool = True, iterations: int = 100000, nblocks: int = 8, seed: int = 1):
        """Run the message-passing ("mp") or store-buffering ("sb") litmus shape across the ranks of this communicator
        (collective call).  Returns ``{"checked", "violations", "timeouts"}`` for this rank; for "sb" also ``"loaded"``,
        an int32 tensor [nblocks, iterations] of the values this rank loaded (the forbidden outcome — every rank of the
        ring loading 0 in the same iteration — is judged by the caller, who can see all ranks)."""
        which = {"mp": 0, "sb": 1}[shape]
        nbytes = int(self._lib.dl_litmus_scratch_bytes(iterations, nblocks))
        key = (iterations, nblocks)
        cache = self.__dict__.setdefault("_litmus_scratch", {})
        if key not in cache:
            cache[key] = self.alloc((nbytes + 3) // 4, torch.int32, "litmus")
        buf = cache[key]
        views = [h[buf.offset:buf.offset + buf.numel * 4].view(torch.int32) for h in self.heaps] if hasattr(self, "heaps") else [buf.tensor]
        for v in views:
            v.zero_()
        torch.cuda.synchronize(self.device)
        if hasattr(self, "_host_barrier"):
            self._host_barrier()
        with torch.cuda.device(self.device):
            _lib.check(self._lib.dl_litmus(C.byref(self._c), which, int(fenced), buf.offset, iterations, seed & 0xFFFFFFFF, nblocks,
                                           _stream_ptr(self.device)))
        torch.cuda.synchronize(self.device)
        self.launches += 1
        out = []
        for v in views:
            counts = v[:16].view(torch.int64).cpu()
            d = {"checked": int(counts[0]), "violations": int(counts[1]), "timeouts": int(counts[2])}
            if which == 1:
                per_block = (512 + 8 * iterations) // 4
                body = v[16:16 + per_block * nblocks].view(nblocks, per_block)
                d["loaded"] = body[:, 128 + iterations:128 + 2 * iterations].cpu()
            out.append(d)
        if hasattr(self, "_host_barrier"):
            self._host_barrier()
        return out if hasattr(self, "heaps") else out[0]


class P2PCommunicator(_CudaOps):
    """One process per GPU; heaps mapped across processes (VMM + fd passing, multicast if
    available, cudaIpc fallback).  ``torch.distributed`` is used ONLY to bootstrap (exchange
    handles, host barriers) — no collective on the training path goes through it."""

    backend = "p2p"

    def __init__(self, rank: Optional[int] = None, world: Optional[int] = None, device: Optional[torch.device] = None,
                 heap_mb: float = 256, mode: str = "auto", use_multicast: Optional[bool] = None, group=None,
                 self_multicast: bool = False):
        """``self_multicast`` (world 1 only): put the heap in VMM memory and bind it to a multicast object of ONE device,
        so that the switch-reduction kernels (``multimem.ld_reduce`` / ``multimem.st``: the nvls, hybrid and light variants)
        run on a single-GPU box — the driver's test machine.  Raises if the driver refuses such an object."""
        import torch.distributed as dist
        if world is None:
            world = dist.get_world_size(group) if dist.is_initialized() else 1
        if rank is None:
            rank = dist.get_rank(group) if dist.is_initialized() else 0
        if device is None:
            device = torch.device("cuda", torch.cuda.current_device())
        if not torch.cuda.is_available():
            raise RuntimeError("P2PCommunicator needs a CUDA device; use HostCommunicator on CPU")
        self._dist, self._group = dist, group
        heap_bytes = _round_up(int(heap_mb * (1 << 20)) + DL_PAD_BYTES, 2 << 20)
        super().__init__(rank, world, device, heap_bytes)
        self.mode = "local"
        self._mem_handle = None
        self._mc_handle = None
        self._mapped: List[tuple] = []
        self._ipc_ptrs: List[int] = []
        dev = device.index if device.index is not None else torch.cuda.current_device()
        self._dev = dev
        torch.cuda.set_device(dev)
        if use_multicast is None:
            use_multicast = os.environ.get("DL_B200_DISABLE_NVLS", "0") != "1"
        if world == 1 and self_multicast:
            self._bootstrap_vmm(True)
            if not self.has_multicast:
                err = getattr(self, "_mc_error", None) or "multicast is not supported on this device"
                self.close()
                raise RuntimeError(f"single-device multicast object refused: {err}")
        elif world == 1:
            self._heap_tensor = torch.zeros(heap_bytes, dtype=torch.uint8, device=device)
            self._c.base[0] = self._heap_tensor.data_ptr()
            self._bytes = self._heap_tensor
        else:
            if not dist.is_initialized():
                raise RuntimeError("torch.distributed must be initialised to bootstrap a multi-rank P2PCommunicator")
            self._bootstrap(mode, use_multicast)
        _lib.check(self._lib.dl_reset_pad(C.byref(self._c)))
        self._setup_ll()
        self._host_barrier()

    # -- bootstrap ----------------------------------------------------------------------------------
    def _host_barrier(self):
        if self.world > 1:
            torch.cuda.synchronize(self.device)
            self._dist.barrier(group=self._group)

    def _all_gather_obj(self, obj):
        if self.world == 1:
            return [obj]
        out = [None] * self.world
        self._dist.all_gather_object(out, obj, group=self._group)
        return out

    def _bootstrap(self, mode: str, use_multicast: bool):
        lib, dev = self._lib, self._dev
        vmm, fdh, mc = C.c_int(0), C.c_int(0), C.c_int(0)
        caps_ok = lib.dl_device_caps(dev, C.byref(vmm), C.byref(fdh), C.byref(mc)) == 0
        mine = {"vmm": bool(caps_ok and vmm.value and fdh.value), "mc": bool(caps_ok and mc.value), "dev": dev,
                "host": os.uname().nodename, "pid": os.getpid()}
        caps = self._all_gather_obj(mine)
        if len({c["host"] for c in caps}) != 1:
            raise RuntimeError("P2PCommunicator spans one host (one NVSwitch domain); use the dist backend across hosts")
        want_vmm = mode in ("auto", "vmm") and all(c["vmm"] for c in caps)
        want_mc = want_vmm and use_multicast and all(c["mc"] for c in caps)
        self.caps = caps
        ok = False
        if want_vmm:
            try:
                self._bootstrap_vmm(want_mc)
                ok = True
            except Exception as e:  # noqa: BLE001 - fall back collectively below
                self._vmm_error = repr(e)
                ok = False
            oks = self._all_gather_obj(ok)
            if not all(oks):
                if mode == "vmm":
                    raise RuntimeError(f"VMM heap bootstrap failed: {getattr(self, '_vmm_error', oks)}")
                self._teardown_vmm()
                ok = False
        if not ok:
            self._bootstrap_ipc()

    def _bootstrap_vmm(self, want_mc: bool):
        lib, dev = self._lib, self._dev
        gran = C.c_size_t(0)
        _lib.check(lib.dl_vmm_granularity(dev, C.byref(gran)))
        g = max(int(gran.value), 2 << 20)
        if want_mc:
            mg = C.c_size_t(0)
            if lib.dl_mc_granularity(self.world, _round_up(self.heap_bytes, g), C.byref(mg)) == 0 and mg.value:
                g = max(g, int(mg.value))
            else:
                want_mc = False
                self._mc_error = "cuMulticastGetGranularity: " + lib.dl_last_error().decode()
        self.heap_bytes = _round_up(self.heap_bytes, g)
        job = self._all_gather_obj(uuid.uuid4().hex[:12])[0]
        handle, fd = C.c_uint64(0), C.c_int(-1)
        _lib.check(lib.dl_vmm_alloc(dev, self.heap_bytes, C.byref(handle), C.byref(fd)))
        self._mem_handle = int(handle.value)
        sock = C.c_int(-1)
        name = lambda r: f"dlb200-{job}-{r}".encode()  # noqa: E731
        _lib.check(lib.dl_uds_listen(name(self.rank), C.byref(sock)))
        if self.world > 1:
            self._dist.barrier(group=self._group)  # every listener is up

        received, errors = {}, []
        expect = (self.world - 1) + (1 if (want_mc and self.rank != 0) else 0)

        def receiver():
            try:
                for _ in range(expect):
                    rfd, tag = C.c_int(-1), C.c_int(-1)
                    _lib.check(lib.dl_uds_recv_fd(sock.value, C.byref(rfd), C.byref(tag), 60000))
                    received[int(tag.value)] = int(rfd.value)
            except Exception as e:  # noqa: BLE001
                errors.append(e)

        th = threading.Thread(target=receiver, daemon=True)
        th.start()
        mc_handle, mc_fd = C.c_uint64(0), C.c_int(-1)
        mc_ok = want_mc
        if want_mc and self.rank == 0:
            if lib.dl_mc_create(self.world, self.heap_bytes, C.byref(mc_handle), C.byref(mc_fd)) != 0:
                mc_ok = False
                self._mc_error = lib.dl_last_error().decode()
        for p in range(self.world):
            if p != self.rank:
                _lib.check(lib.dl_uds_send_fd(name(p), fd.value, self.rank, 60000))
                if want_mc and self.rank == 0:
                    # tag 1000 = multicast handle (or tag 1001 with a dummy fd when creation failed)
                    _lib.check(lib.dl_uds_send_fd(name(p), mc_fd.value if mc_ok else fd.value, 1000 if mc_ok else 1001, 60000))
        th.join()
        lib.dl_uds_close(sock.value)
        if errors:
            raise errors[0]
        # map every heap
        for p in range(self.world):
            ptr = C.c_void_p()
            if p == self.rank:
                _lib.check(lib.dl_vmm_map(dev, self._mem_handle, self.heap_bytes, C.byref(ptr)))
                self._mapped.append((ptr.value, self.heap_bytes, None))
            else:
                h = C.c_uint64(0)
                _lib.check(lib.dl_vmm_import(received[p], C.byref(h)))
                lib.dl_close_fd(received[p])
                _lib.check(lib.dl_vmm_map(dev, h.value, self.heap_bytes, C.byref(ptr)))
                self._mapped.append((ptr.value, self.heap_bytes, int(h.value)))
            self._c.base[p] = ptr.value
        lib.dl_close_fd(fd.value)
        self._bytes = torch.as_tensor(_CudaMem(self._c.base[self.rank], self.heap_bytes), device=self.device)
        self.mode = "vmm"
        # multicast (NVLS)
        if want_mc:
            if self.rank == 0:
                mch = int(mc_handle.value) if mc_ok else None
            else:
                if 1000 in received:
                    h = C.c_uint64(0)
                    mc_ok = lib.dl_vmm_import(received[1000], C.byref(h)) == 0
                    lib.dl_close_fd(received[1000])
                    mch = int(h.value) if mc_ok else None
                else:
                    if 1001 in received:
                        lib.dl_close_fd(received[1001])
                    mc_ok, mch = False, None
            if mc_ok:
                mc_ok = lib.dl_mc_add_device(mch, dev) == 0
                if not mc_ok:
                    self._mc_error = lib.dl_last_error().decode()
            mc_ok = all(self._all_gather_obj(bool(mc_ok)))
            if mc_ok:
                mc_ok = lib.dl_mc_bind(mch, self._mem_handle, self.heap_bytes) == 0
                if not mc_ok:
                    self._mc_error = lib.dl_last_error().decode()
            mc_ok = all(self._all_gather_obj(bool(mc_ok)))
            if mc_ok:
                ptr = C.c_void_p()
                mc_ok = lib.dl_vmm_map(dev, mch, self.heap_bytes, C.byref(ptr)) == 0
                mc_ok = all(self._all_gather_obj(bool(mc_ok)))
                if mc_ok:
                    self._c.mc = ptr.value
                    self._mc_handle = mch
                    self.has_multicast = True
        self._bytes.zero_()
        torch.cuda.synchronize(self.device)

    def _teardown_vmm(self):
        for ptr, nbytes, h in self._mapped:
            self._lib.dl_vmm_unmap(ptr, nbytes)
            if h is not None:
                self._lib.dl_vmm_release(h)
        self._mapped = []
        if self._mem_handle is not None:
            self._lib.dl_vmm_release(self._mem_handle)
            self._mem_handle = None
        self._c.mc = None
        self.has_multicast = False

    def _bootstrap_ipc(self):
        lib, dev = self._lib, self._dev
        ptr = C.c_void_p()
        handle = (C.c_char * 64)()
        _lib.check(lib.dl_ipc_alloc(dev, self.heap_bytes, C.byref(ptr), handle))
        self._ipc_local = ptr.value
        handles = self._all_gather_obj(bytes(handle.raw))
        for p in range(self.world):
            if p == self.rank:
                self._c.base[p] = ptr.value
            else:
                q = C.c_void_p()
                buf = (C.c_char * 64).from_buffer_copy(handles[p])
                _lib.check(lib.dl_ipc_open(dev, buf, C.byref(q)))
                self._ipc_ptrs.append(q.value)
                self._c.base[p] = q.value
        self._bytes = torch.as_tensor(_CudaMem(ptr.value, self.heap_bytes), device=self.device)
        self._bytes.zero_()
        torch.cuda.synchronize(self.device)
        self.mode = "ipc"

    def _local_bytes(self) -> torch.Tensor:
        return self._bytes

    def describe(self) -> dict:
        return {"backend": self.backend, "mode": self.mode, "world": self.world, "rank": self.rank,
                "heap_bytes": self.heap_bytes, "multicast": bool(self.has_multicast),
                "mc_error": getattr(self, "_mc_error", None), "vmm_error": getattr(self, "_vmm_error", None)}

    def close(self) -> None:
        """Collective teardown: wait for the device, make sure no peer still uses our memory, then drop every mapping
        this process made — the multicast view and binding, the peers' heaps, our own heap, the cudaIpc imports — and
        release the handles, so a long-lived process that creates communicators repeatedly does not leak address space
        or physical memory.  Tensors that alias the heap (FlatParams views) must not be used afterwards."""
        if getattr(self, "_closed", False):
            return
        self._closed = True
        try:
            torch.cuda.synchronize(self.device)
        except Exception:  # noqa: BLE001
            pass
        if self.world > 1:
            try:
                self._dist.barrier(group=self._group)      # nobody is still reading or writing a peer heap
            except Exception:  # noqa: BLE001
                pass
        lib = self._lib
        self._bytes = None
        if self.mode == "vmm":
            if self._mc_handle is not None:
                if self._c.mc:
                    lib.dl_vmm_unmap(self._c.mc, self.heap_bytes)
                lib.dl_mc_unbind(self._mc_handle, self._dev, self.heap_bytes)
                lib.dl_vmm_release(self._mc_handle)
                self._mc_handle = None
            self._teardown_vmm()
        elif self.mode == "ipc":
            for q in self._ipc_ptrs:
                lib.dl_ipc_close(q)
            self._ipc_ptrs = []
            if self.world > 1:
                try:
                    self._dist.barrier(group=self._group)  # every importer has closed before the owner frees
                except Exception:  # noqa: BLE001
                    pass
            if getattr(self, "_ipc_local", None):
                lib.dl_ipc_free(self._ipc_local)
                self._ipc_local = None
        else:
            self._heap_tensor = None
        for r in range(self.world):
            self._c.base[r] = None
        self._c.mc = None
        self.has_multicast = False
        if self._host_ptr:
            self._c.host_words = None
            self._host_words = None
            lib.dl_host_words_free(self._host_ptr)
            self._host_ptr = None


class EmulatedCommunicator(_CudaOps):
    """All ``world`` ranks on ONE GPU: each rank has its own heap (a torch allocation) and every
    collective is one cooperative launch with gridDim.y = world.  ``active`` arguments are
    bitmasks over ranks, rank r publishes ``payload + r``.  Test-only."""

    backend = "emulated"

    def __init__(self, world: int, device: Optional[torch.device] = None, heap_mb: float = 16):
        device = device or torch.device("cuda", torch.cuda.current_device())
        heap_bytes = _round_up(int(heap_mb * (1 << 20)) + DL_PAD_BYTES, 4096)
        super().__init__(0, world, device, heap_bytes)
        self._c.emulate = 1
        self.heaps = [torch.zeros(heap_bytes, dtype=torch.uint8, device=device) for _ in range(world)]
        for r in range(world):
            self._c.base[r] = self.heaps[r].data_ptr()
        _lib.check(self._lib.dl_reset_pad(C.byref(self._c)))
        self._setup_ll()

    def _local_bytes(self) -> torch.Tensor:
        return self.heaps[0]

    def alloc_all(self, numel: int, dtype: torch.dtype = torch.float32, name: str = ""):
        """Returns (SymBuffer, [tensor of rank 0, tensor of rank 1, ...])."""
        buf = self.alloc(numel, dtype, name)
        es = buf.tensor.element_size()
        views = [h[buf.offset:buf.offset + numel * es].view(dtype) for h in self.heaps]
        for v in views:
            v.zero_()
        return buf, views

    def result_of(self, rank: int, channel: int = 0) -> CommResult:
        self._c.rank = rank
        try:
            return self.read_result(channel)
        finally:
            self._c.rank = 0


# =====================================================================================================
# Across hosts / mixed CPU+GPU jobs: torch.distributed collectives
# =====================================================================================================
class DistCommunicator(Communicator):
    """Same front end over ``torch.distributed`` — for jobs that do not fit one NVSwitch domain:
    several hosts (examples/client_remote.lua:31-41, the ssh lines of examples/AsyncEASGD.sh:44-46) or
    a mix of CPU and GPU nodes (examples/client_remote.sh:5-6).  The heap is a private tensor (offsets
    still match across ranks, nothing is mapped remotely) and every operation is a library collective
    (NCCL when every node is a GPU, gloo with host staging otherwise) followed by the local math — the
    reference's dataflow, kept as the portability path.  AsyncEA works here the reference's way: the rank
    that owns the center serves pull / push requests from a service thread, one client at a time."""

    backend = "dist"

    def __init__(self, rank: Optional[int] = None, world: Optional[int] = None, device: Optional[torch.device] = None,
                 heap_mb: float = 256, group=None):
        import torch.distributed as dist
        if not dist.is_initialized():
            raise RuntimeError("DistCommunicator needs an initialised torch.distributed process group")
        self._dist, self._group = dist, group
        world = dist.get_world_size(group) if world is None else world
        rank = dist.get_rank(group) if rank is None else rank
        device = device or torch.device("cpu")
        heap_bytes = _round_up(int(heap_mb * (1 << 20)) + DL_PAD_BYTES, 4096)
        super().__init__(rank, world, device, heap_bytes)
        self._bytes = torch.zeros(heap_bytes, dtype=torch.uint8, device=device)
        self._wire_cpu = dist.get_backend(group) != "nccl"       # gloo: CUDA tensors are staged through the host
        self._last: Optional[CommResult] = None
        self._start_service()

    def _local_bytes(self) -> torch.Tensor:
        return self._bytes

    # -- helpers ------------------------------------------------------------------------------------
    def _wire(self, t: torch.Tensor) -> torch.Tensor:
        return t.cpu() if (self._wire_cpu and t.is_cuda) else t

    def _meet(self, active: int, payload: int, extra: float = 0.0):
        """What the kernels' arrival barrier does: everybody learns everybody's (active, payload, extra)."""
        mine = torch.tensor([float(active), float(payload), float(extra)], dtype=torch.float64,
                            device="cpu" if self._wire_cpu else self.device)
        out = [torch.empty_like(mine) for _ in range(self.world)]
        self._dist.all_gather(out, mine, group=self._group)
        rows = [o.tolist() for o in out]
        flags = [int(r[0]) for r in rows]
        n_active = sum(1 for a in flags if a & 1)
        first = next((i for i, a in enumerate(flags) if a & 3), self.rank)
        self._last = CommResult(n_active, [int(r[1]) for r in rows])
        return flags, n_active, rows[first][2]

    def _sum(self, t: torch.Tensor) -> None:
        w = self._wire(t)
        self._dist.all_reduce(w, group=self._group)
        if w is not t:
            t.copy_(w)

    # -- collectives -----------------------------------------------------------------------------------
    def barrier(self, active: int = 1, payload: int = 0, channel: int = 0, read: bool = True, stream=None):
        self._meet(active, payload)
        self.launches += 1
        return self._last

    def allreduce(self, buf: SymBuffer, n: Optional[int] = None, active: int = 1, payload: int = 0, normalize: bool = False,
                  variant: str = "auto", channel: int = 0, read: bool = False, nblocks: int = 0, stream=None):
        n = buf.numel if n is None else n
        flags, n_active, norm = self._meet(active, payload, float(normalize))
        t = buf.tensor[:n]
        if not (active & 3):
            t.zero_()                                   # a draining node contributes zeros (lua/AllReduceSGD.lua:37)
        self._sum(t)
        if norm and n_active > 1 and t.is_floating_point():
            t.mul_(1.0 / n_active)
        self.launches += 1
        return self._last

    def fused_sgd(self, grads: SymBuffer, params: SymBuffer, momentum_buf: Optional[SymBuffer], n: int, lr: float,
                  momentum: float = 0.0, weight_decay: float = 0.0, nesterov: bool = False, active: int = 1,
                  payload: int = 0, zero_grads: bool = True, variant: str = "auto", channel: int = 0, read: bool = False,
                  nblocks: int = 0, stream=None):
        flags, n_active, lr_used = self._meet(active, payload, lr)
        g, p = grads.tensor[:n], params.tensor[:n]
        if not (active & 3):
            g.zero_()
        self._sum(g)
        if n_active > 0:
            with torch.no_grad():
                if n_active > 1:
                    g.mul_(1.0 / n_active)
                if weight_decay:
                    g.add_(p, alpha=weight_decay)
                if momentum_buf is not None:
                    m = momentum_buf.tensor[:n]
                    m.mul_(momentum).add_(g)
                    g = g.add(m, alpha=momentum) if nesterov else m
                p.add_(g, alpha=-float(lr_used))
        if zero_grads:
            grads.tensor[:n].zero_()
        self.launches += 1
        return self._last

    def fused_ea(self, x: SymBuffer, c_cur: SymBuffer, c_next: SymBuffer, n: int, alpha: float, active: int = 1,
                 payload: int = 0, variant: str = "auto", channel: int = 0, read: bool = False, nblocks: int = 0, stream=None):
        flags, n_active, _ = self._meet(active, payload)
        xv, c, cn = x.tensor[:n], c_cur.tensor[:n], c_next.tensor[:n]
        if n_active > 0:                                # lua/AllReduceEA.lua:34-45, finished nodes still move (:58-68)
            d = (xv - c) * alpha
            xv.sub_(d)
            self._sum(d)
            torch.add(c, d, out=cn)
        else:
            cn.copy_(c)
        self.launches += 1
        return self._last

    def broadcast(self, buf: SymBuffer, n: Optional[int] = None, root: int = 0, also: Optional[SymBuffer] = None,
                  payload: int = 0, variant: str = "auto", channel: int = 0, read: bool = False, nblocks: int = 0, stream=None):
        n = buf.numel if n is None else n
        self._meet(1, payload)
        t = buf.tensor[:n]
        w = self._wire(t)
        self._dist.broadcast(w, src=root if self._group is None else self._dist.get_global_rank(self._group, root), group=self._group)
        if w is not t:
            t.copy_(w)
        if also is not None:
            also.tensor[:n].copy_(t)
        self.launches += 1
        return self._last

    def read_result(self, channel: int = 0, stream=None) -> CommResult:
        return self._last if self._last is not None else CommResult(0, [0] * self.world)

    # -- asynchronous EASGD over messages ----------------------------------------------------------------
    # Without peer-mapped memory the center has to be served by its owner, which is the reference's own design
    # (lua/AsyncEA.lua:163-228): a service thread on every rank answers requests against the local heap and
    # mailbox, one at a time — so syncs are always exclusive here.  Requests travel on a private gloo group:
    # tag 0 = request header, tag 1 = server -> client payload, tag 2 = client -> server payload.
    _OP_STOP, _OP_MB_WRITE, _OP_MB_READ, _OP_SYNC_COUNT, _OP_PULL, _OP_SYNC = 0, 1, 2, 3, 4, 5

    def _start_service(self) -> None:
        import threading
        self._mailbox = [0] * 32
        self._sync_count = 0
        self._svc_lock = threading.Lock()
        self._agroup = self._dist.new_group(backend="gloo") if self.world > 1 else None
        self._svc = None
        if self.world > 1:
            self._svc = threading.Thread(target=self._serve, name="distlearn-async-service", daemon=True)
            self._svc.start()

    def _heap_f32(self, off: int, n: int) -> torch.Tensor:
        return self._bytes[off:off + 4 * n].view(torch.float32)

    def _serve(self) -> None:
        dist, g = self._dist, self._agroup
        hdr = torch.zeros(4, dtype=torch.int64)
        while True:
            try:
                src = dist.recv(hdr, group=g, tag=0)
            except Exception:  # noqa: BLE001 - the process group went away
                return
            op, a, b, c = (int(v) for v in hdr.tolist())
            if op == self._OP_STOP:
                return
            with self._svc_lock:
                if op == self._OP_MB_WRITE:
                    self._mailbox[a] = b
                    dist.send(torch.zeros(1, dtype=torch.int64), dst=src, group=g, tag=1)     # ack: the write is visible
                elif op == self._OP_MB_READ:
                    dist.send(torch.tensor([self._mailbox[a]], dtype=torch.int64), dst=src, group=g, tag=1)
                elif op == self._OP_SYNC_COUNT:
                    dist.send(torch.tensor([self._sync_count], dtype=torch.int64), dst=src, group=g, tag=1)
                elif op in (self._OP_PULL, self._OP_SYNC):
                    center = self._heap_f32(a, b)
                    dist.send(center.detach().cpu().contiguous(), dst=src, group=g, tag=1)
                    if op == self._OP_SYNC:
                        delta = torch.empty(b, dtype=torch.float32)
                        dist.recv(delta, src=src, group=g, tag=2)
                        center.add_(delta.to(center.device))                                  # lua/AsyncEA.lua:212-216
                        self._sync_count += 1

    def _request(self, server: int, op: int, a: int = 0, b: int = 0, c: int = 0) -> None:
        self._dist.send(torch.tensor([op, a, b, c], dtype=torch.int64), dst=server, group=self._agroup, tag=0)

    def _reply_i64(self, server: int) -> int:
        t = torch.zeros(1, dtype=torch.int64)
        self._dist.recv(t, src=server, group=self._agroup, tag=1)
        return int(t[0])

    def mailbox_write(self, rank: int, index: int, value: int) -> None:
        if rank == self.rank:
            with self._svc_lock:
                self._mailbox[index] = int(value)
            return
        self._request(rank, self._OP_MB_WRITE, index, int(value))
        self._reply_i64(rank)

    def mailbox_read(self, rank: int, index: int) -> int:
        if rank == self.rank:
            with self._svc_lock:
                return self._mailbox[index]
        self._request(rank, self._OP_MB_READ, index)
        return self._reply_i64(rank)

    def sync_count(self, server: int) -> int:
        if server == self.rank:
            with self._svc_lock:
                return self._sync_count
        self._request(server, self._OP_SYNC_COUNT)
        return self._reply_i64(server)

    def peer_pull(self, src_rank: int, buf: SymBuffer, dst: torch.Tensor, n: Optional[int] = None, stream=None) -> None:
        n = buf.numel if n is None else n
        if src_rank == self.rank:
            dst.view(-1)[:n].copy_(buf.tensor[:n])
            return
        self._request(src_rank, self._OP_PULL, buf.offset, n)
        t = torch.empty(n, dtype=torch.float32)
        self._dist.recv(t, src=src_rank, group=self._agroup, tag=1)
        dst.view(-1)[:n].copy_(t)

    def async_client_sync(self, server: int, x: torch.Tensor, center: SymBuffer, n: int, alpha: float, exclusive: bool = False,
                          nblocks: int = 0, stream=None) -> None:
        """d = alpha (x - c); x -= d; c += d with c on ``server`` (lua/AsyncEA.lua:95-132): pull, local move, push."""
        xv = x.view(-1)[:n]
        if server == self.rank:
            with self._svc_lock:
                c = center.tensor[:n]
                d = (xv - c) * alpha
                xv.sub_(d)
                c.add_(d)
                self._sync_count += 1
            self.launches += 1
            return
        self._request(server, self._OP_SYNC, center.offset, n)
        c = torch.empty(n, dtype=torch.float32)
        self._dist.recv(c, src=server, group=self._agroup, tag=1)
        xc = xv.detach().cpu() if xv.is_cuda else xv
        d = (xc - c) * alpha                                                   # lua/AsyncEA.lua:114
        self._dist.send(d.contiguous(), dst=server, group=self._agroup, tag=2)
        xv.sub_(d.to(xv.device))                                               # lua/AsyncEA.lua:115
        self.launches += 1

    def close(self) -> None:
        """Collective: every rank stops the service thread of its right-hand neighbour, then waits for its own."""
        if getattr(self, "_svc", None) is not None:
            try:
                self._dist.barrier(group=self._group)                         # nobody is mid-request any more
                self._request((self.rank + 1) % self.world, self._OP_STOP)
                self._svc.join(timeout=10)
            except Exception:  # noqa: BLE001
                pass
            self._svc = None

    def synchronize(self) -> None:
        if self.device.type == "cuda":
            torch.cuda.synchronize(self.device)

    def describe(self) -> dict:
        return {"backend": self.backend, "wire": "gloo+host staging" if self._wire_cpu else "nccl", "world": self.world,
                "rank": self.rank, "heap_bytes": self.heap_bytes}


def create_communicator(backend: str = "auto", **kw) -> Communicator:
    """``backend``: 'p2p' (CUDA, default when a GPU is present), 'host' (CPU shared memory),
    'dist' (torch.distributed collectives: several hosts, mixed CPU/GPU nodes)."""
    if backend == "auto":
        backend = "p2p" if torch.cuda.is_available() else "host"
    if backend == "p2p":
        return P2PCommunicator(**kw)
    if backend == "host":
        return HostCommunicator(**kw)
    if backend == "dist":
        return DistCommunicator(**kw)
    raise ValueError(f"unknown backend {backend!r}")
