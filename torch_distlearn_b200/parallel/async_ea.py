"""AsyncEA — asynchronous Elastic Averaging SGD with a center server (reference: lua/AsyncEA.lua).

Roles (one process each, as in examples/AsyncEASGD.sh:36-40): a **server** that owns the center
variable, N **clients** that train and every ``tau`` steps exchange an elastic difference with the
center, and an optional **tester** that evaluates the center on request.

The reference moves the center over TCP with a six-message handshake per sync and a server process
in the middle (lua/AsyncEA.lua:82-132 <-> :163-228).  Here the center is a buffer in the server
rank's symmetric heap; ``syncClient`` is ONE client-side kernel that loads the center over NVLink,
applies  d = alpha (x - c);  x -= d  locally and adds ``d`` into the center with vector atomics —
no host round trip, no server-side work.  The server and tester keep their API for parity:

===============  ==========================================================  =====================
reference        this library                                                lua/AsyncEA.lua
===============  ==========================================================  =====================
initServer       copy params into the heap-resident center, publish "ready"  :150-160
initClient       wait for "ready", P2P-pull the center into params           :64-78
syncClient       every tau-th call: the fused pull/update/push kernel        :134-146
syncServer       wait for >=1 new client sync, params <- center              :230-237
testNet          post a test request, (optionally) wait for the tester ack   :239-258
initTester       allocate                                                    :261-265
startTest        wait for a request, P2P-pull the center into params         :268-285
finishTest       post the ack                                                :287-292
===============  ==========================================================  =====================

``exclusive=True`` serialises client syncs with a ticket lock (the server's recvAny mutex,
:163-177); the default lock-free mode lets them overlap.  Unlike the reference there is a shutdown
protocol: clients call ``finish()``, the server's ``syncServer`` returns False once all clients
have finished (the reference server blocks forever in recvAny).
"""
from __future__ import annotations

import time
from typing import Any, Optional

import torch

from .comm import Communicator, SymBuffer
from .flat import FlatParams, walk_table

# mailbox words in the SERVER's pad
MB_READY, MB_TEST_SEQ, MB_TEST_ACK, MB_STOP = 0, 1, 2, 3
MB_DONE_BASE = 8     # words 8..15: rank r sets word 8+r when it has finished training
MB_PULLED_BASE = 16  # words 16..23: rank r sets word 16+r once it has pulled the initial center


class AsyncEA:
    def __init__(self, comm: Communicator, role: str, tau: int, alpha: float, server_rank: int = 0,
                 num_clients: Optional[int] = None, exclusive: bool = False, verbose: bool = False):
        if role not in ("server", "client", "tester"):
            raise ValueError("role must be 'server', 'client' or 'tester'")
        self.comm, self.role, self.tau, self.alpha = comm, role, int(tau), float(alpha)
        self.server_rank, self.exclusive, self.verbose = server_rank, exclusive, verbose
        self.num_clients = num_clients
        self.node = comm.node_index
        self.step = 0
        self._center: Optional[SymBuffer] = None
        self._n = 0
        self._x: Optional[torch.Tensor] = None
        self._tensors = None
        self._seen_syncs = 0
        self._seen_tests = 0
        self._poll_s = 0.0005

    # ------------------------------------------------------------------------------------------------
    def _log(self, msg: str) -> None:
        if self.verbose:
            from ..utils.color_print import print_client, print_server
            if self.role == "server":
                print_server(msg)
            else:
                print_client(self.node, msg)

    def _one_time_init(self, params: Any) -> None:
        """Every role allocates the center symmetrically (only the server's copy is live), and a flat
        local working vector for its parameters (lua/AsyncEA.lua:18-29)."""
        if self._center is not None:
            return
        if isinstance(params, FlatParams):
            n, self._x, self._tensors = params.padded, params.params.tensor, None
        else:
            tensors = walk_table(params)
            if len(tensors) == 1 and tensors[0].is_contiguous() and tensors[0].dtype == torch.float32 \
                    and tensors[0].data_ptr() % 16 == 0 and tensors[0].device.type == self.comm.device.type:
                n, self._x, self._tensors = tensors[0].numel(), tensors[0].view(-1), None
            else:
                total = sum((t.numel() + 3) // 4 * 4 for t in tensors)
                n = total
                self._x = torch.zeros(total, dtype=torch.float32, device=self.comm.device)
                self._tensors = tensors
        self._n = n
        self._center = self.comm.alloc(max(n, 4), torch.float32, "async.center")

    def _pack(self):
        if self._tensors is not None:
            off = 0
            for t in self._tensors:
                k = t.numel()
                self._x[off:off + k].copy_(t.reshape(-1))
                off += (k + 3) // 4 * 4

    def _unpack(self):
        if self._tensors is not None:
            off = 0
            for t in self._tensors:
                k = t.numel()
                t.copy_(self._x[off:off + k].view(t.shape))
                off += (k + 3) // 4 * 4

    def _wait(self, cond, timeout: Optional[float], what: str) -> bool:
        t0 = time.time()
        while not cond():
            if timeout is not None and time.time() - t0 > timeout:
                raise TimeoutError(f"AsyncEA[{self.role} {self.node}]: timed out waiting for {what}")
            time.sleep(self._poll_s)
        return True

    # ------------------------------------------------------------------------------------------------ server
    def initServer(self, params: Any) -> None:
        self._one_time_init(params)
        self._pack()
        self._center.tensor[:self._n].copy_(self._x[:self._n])
        self.comm.synchronize()
        self.comm.mailbox_write(self.server_rank, MB_READY, 1)
        self._log("center published")

    def syncServer(self, params: Any = None, timeout: Optional[float] = 60.0) -> bool:
        """Blocks until at least one more client sync has landed, then params <- center
        (lua/AsyncEA.lua:230-237, :222-226).  Returns False when every client has finished."""
        self._log("Server waiting to sync")

        last = {}

        def ready():
            # one snapshot per tick: the sync counter and every client's "done" word arrive in a single small copy
            count, words = self.comm.ctl_snapshot(self.server_rank)
            last["count"], last["done"] = count, self._count_done(words)
            return count > self._seen_syncs or last["done"]

        self._wait(ready, timeout, "a client sync")
        count = last["count"]
        if count == self._seen_syncs and last["done"]:
            return False
        self._seen_syncs = count
        self._x[:self._n].copy_(self._center.tensor[:self._n])
        self._unpack()
        return True

    def _count_done(self, words) -> bool:
        if self.num_clients is None:
            return False
        return sum(1 for r in range(self.comm.world) if words[MB_DONE_BASE + r] == 1) >= self.num_clients

    def _all_clients_done(self) -> bool:
        return self._count_done(self.comm.ctl_snapshot(self.server_rank)[1])

    @property
    def syncs(self) -> int:
        return self.comm.sync_count(self.server_rank)

    def testNet(self, block: bool = True, timeout: Optional[float] = 600.0) -> None:
        """Ask the tester to evaluate the current center (lua/AsyncEA.lua:239-258).  The reference
        blocks the server until the tester's Ack; clients are never blocked here either way."""
        seq = self.comm.mailbox_read(self.server_rank, MB_TEST_SEQ) + 1
        self.comm.mailbox_write(self.server_rank, MB_TEST_SEQ, seq)
        if block:
            self._wait(lambda: self.comm.mailbox_read(self.server_rank, MB_TEST_ACK) >= seq, timeout, "the tester's Ack")

    def stop(self) -> None:
        """Tell the tester (and anyone polling) that training is over."""
        self.comm.mailbox_write(self.server_rank, MB_STOP, 1)

    # ------------------------------------------------------------------------------------------------ client
    def initClient(self, params: Any, timeout: Optional[float] = 60.0) -> None:
        """All clients start from the server's center (lua/AsyncEA.lua:64-78)."""
        self._one_time_init(params)
        self._wait(lambda: self.comm.mailbox_read(self.server_rank, MB_READY) == 1, timeout, "the server's center")
        self.comm.peer_pull(self.server_rank, self._center, self._x, self._n)
        self._unpack()
        self.comm.synchronize()
        # The reference's initServer hands the same center to EVERY client before the first sync can happen
        # (lua/AsyncEA.lua:155-159).  Syncs need no server here, so the clients gate themselves: nobody starts
        # training (and moving the center) until all of them have taken their copy of the initial one.
        self.comm.mailbox_write(self.server_rank, MB_PULLED_BASE + self.comm.rank, 1)
        if self.num_clients is not None:
            def all_pulled():
                words = self.comm.ctl_snapshot(self.server_rank)[1]
                return sum(1 for r in range(self.comm.world) if words[MB_PULLED_BASE + r] == 1) >= self.num_clients
            self._wait(all_pulled, timeout, "every client to pull the initial center")

    def syncClient(self, params: Any = None) -> bool:
        """Call once per local step; every ``tau``-th call does the elastic exchange with the center
        (lua/AsyncEA.lua:134-146).  Returns True if it synced."""
        self.step += 1
        if self.step % self.tau != 0:
            return False
        self._log("Waiting to sync")
        self._pack()
        self.comm.async_client_sync(self.server_rank, self._x, self._center, self._n, self.alpha, self.exclusive)
        self._unpack()
        return True

    def finish(self) -> None:
        """Client-side shutdown notice (no counterpart in the reference)."""
        self.comm.synchronize()
        self.comm.mailbox_write(self.server_rank, MB_DONE_BASE + self.comm.rank, 1)

    # ------------------------------------------------------------------------------------------------ tester
    def initTester(self, params: Any) -> None:
        self._one_time_init(params)

    def startTest(self, params: Any = None, timeout: Optional[float] = None) -> bool:
        """Wait for a test request, then params <- center (lua/AsyncEA.lua:268-285).  Returns False if
        the server said stop."""
        last = {}

        def cond():
            words = self.comm.ctl_snapshot(self.server_rank)[1]
            last["seq"] = words[MB_TEST_SEQ]
            return words[MB_TEST_SEQ] > self._seen_tests or words[MB_STOP] == 1

        self._wait(cond, timeout, "a test request")
        seq = last["seq"]
        if seq == self._seen_tests:
            return False
        self._seen_tests = seq
        self.comm.peer_pull(self.server_rank, self._center, self._x, self._n)
        self._unpack()
        self.comm.synchronize()
        return True

    def finishTest(self) -> None:
        """Ack (lua/AsyncEA.lua:287-292)."""
        self.comm.mailbox_write(self.server_rank, MB_TEST_ACK, self._seen_tests)

    # snake_case aliases
    init_server, init_client, init_tester = initServer, initClient, initTester
    sync_client, sync_server, test_net = syncClient, syncServer, testNet
    start_test, finish_test = startTest, finishTest
