"""Process topology: build a communicator from reference-style flags, and spawn one process per node.

The reference starts one ``th`` process per node from a shell script and wires them with
``--nodeIndex i --numNodes N [--host h --port p]`` (examples/mnist.sh:4-10, examples/cifar10-cuda.sh:4-7,
examples/client_remote.lua:31-41).  The same flags work here::

    python examples/mnist.py --numNodes 4 --nodeIndex 1 &  ... &  wait          # like the .sh launchers
    python -m torch_distlearn_b200.launch -n 4 examples/mnist.py                # or let this module spawn them
    torchrun --nproc-per-node 8 examples/cifar10.py --cuda                      # RANK/WORLD_SIZE from the env

GPU jobs use ``torch.distributed`` (gloo) on ``--host:--port`` for bootstrap only; CPU jobs meet on a
POSIX shared-memory segment named after the port.
"""
from __future__ import annotations

import argparse
import os
import subprocess
import sys
from typing import List, Optional

import torch


def make_communicator(opt, heap_mb: Optional[float] = None, rank: Optional[int] = None, world: Optional[int] = None):
    """``opt`` carries nodeIndex/numNodes/cuda/gpu/host/port/backend/heapMB (utils.config.base_parser).
    ``rank``/``world`` override the flags (the EASGD apps add a server and a tester to the N clients)."""
    from .parallel.comm import HostCommunicator, P2PCommunicator
    rank = opt.nodeIndex - 1 if rank is None else rank
    world = opt.numNodes if world is None else world
    backend = getattr(opt, "backend", "auto")
    use_cuda = bool(getattr(opt, "cuda", False))
    if backend == "auto":
        backend = "p2p" if (use_cuda and torch.cuda.is_available()) else "host"
    heap = heap_mb if heap_mb is not None else getattr(opt, "heapMB", 256.0)
    if backend == "host":
        name = f"distlearn-{os.environ.get('DISTLEARN_JOB', os.getuid())}-{opt.port}"
        return HostCommunicator(rank, world, name=name, heap_mb=heap, timeout_s=float(os.environ.get("DL_B200_TIMEOUT_S", "120")))
    if backend == "dist":
        # several hosts and/or mixed CPU+GPU nodes: rendezvous on --host:--port like ipc.Tree(...,host,port)
        import torch.distributed as dist
        from .parallel.comm import DistCommunicator
        dev = torch.device("cpu")
        if use_cuda and torch.cuda.is_available():
            dev = torch.device("cuda", (opt.gpu - 1) % torch.cuda.device_count())
            torch.cuda.set_device(dev)
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", str(opt.host))
            os.environ.setdefault("MASTER_PORT", str(opt.port))
            dist.init_process_group(os.environ.get("DL_B200_DIST_BACKEND", "gloo"), rank=rank, world_size=world)
        return DistCommunicator(rank=rank, world=world, device=dev, heap_mb=heap)
    if not torch.cuda.is_available():
        raise RuntimeError("--backend p2p needs a CUDA device")
    ngpu = torch.cuda.device_count()
    gpu = (opt.gpu - 1) % ngpu                       # --gpu is 1-based like the reference (examples/cifar10.lua:13-17)
    dev = torch.device("cuda", gpu)
    torch.cuda.set_device(dev)
    if world > 1:
        import torch.distributed as dist
        if not dist.is_initialized():
            os.environ.setdefault("MASTER_ADDR", str(opt.host))
            os.environ.setdefault("MASTER_PORT", str(opt.port))
            dist.init_process_group("gloo", rank=rank, world_size=world)
    return P2PCommunicator(rank=rank, world=world, device=dev, heap_mb=heap)


def shutdown(comm) -> None:
    """Close the communicator and the bootstrap process group."""
    try:
        comm.synchronize()
    except Exception:  # noqa: BLE001
        pass
    comm.close()
    try:
        import torch.distributed as dist
        if dist.is_initialized():
            dist.barrier()
            dist.destroy_process_group()
    except Exception:  # noqa: BLE001
        pass


def spawn(script: str, script_args: List[str], num_nodes: int, port: int = 8080, host: str = "127.0.0.1",
          gpus: bool = False, extra_env: Optional[dict] = None, timeout: Optional[float] = None) -> int:
    """Start ``num_nodes`` copies of ``script`` (node i gets ``--nodeIndex i --numNodes N`` and, with
    ``gpus``, ``--cuda --gpu i``), wait for all; returns the worst exit code.  If one node fails the
    others are terminated (the reference's launchers just ``wait``)."""
    procs = []
    env_base = dict(os.environ)
    env_base.update(extra_env or {})
    env_base.setdefault("DISTLEARN_JOB", f"{os.getpid()}")
    for i in range(1, num_nodes + 1):
        env = dict(env_base, RANK=str(i - 1), WORLD_SIZE=str(num_nodes), LOCAL_RANK=str(i - 1), MASTER_ADDR=host,
                   MASTER_PORT=str(port), OMP_NUM_THREADS=env_base.get("OMP_NUM_THREADS", "1"))
        cmd = [sys.executable, script, "--nodeIndex", str(i), "--numNodes", str(num_nodes), "--host", host, "--port", str(port)]
        if gpus:
            cmd += ["--cuda", "--gpu", str(i)]
        procs.append(subprocess.Popen(cmd + list(script_args), env=env))
    worst = 0
    try:
        pending = list(procs)
        import time
        t0 = time.time()
        while pending:
            for p in list(pending):
                rc = p.poll()
                if rc is None:
                    continue
                pending.remove(p)
                if rc != 0:
                    worst = worst or rc
                    for q in pending:
                        q.terminate()
            if timeout is not None and time.time() - t0 > timeout:
                worst = worst or 124
                for q in pending:
                    q.terminate()
                break
            time.sleep(0.05)
    finally:
        for p in procs:
            if p.poll() is None:
                p.kill()
    return worst


def main(argv: Optional[List[str]] = None) -> int:
    ap = argparse.ArgumentParser(description="spawn one process per node (reference: examples/*.sh)")
    ap.add_argument("-n", "--numNodes", type=int, required=True)
    ap.add_argument("--port", type=int, default=8080)
    ap.add_argument("--host", default="127.0.0.1")
    ap.add_argument("--gpus", action="store_true", help="node i uses GPU i (examples/cifar10-cuda.sh:4-7)")
    ap.add_argument("script")
    ap.add_argument("script_args", nargs=argparse.REMAINDER)
    a = ap.parse_args(argv)
    return spawn(a.script, a.script_args, a.numNodes, a.port, a.host, a.gpus)


if __name__ == "__main__":
    sys.exit(main())
