"""Run a function on N ranks as N spawned processes (no GPU needed) and collect the results."""
from __future__ import annotations

import multiprocessing as mp
import os
import traceback
import uuid
from typing import Any, Callable, List


def _pack(x):
    """Tensors travel as numpy arrays: torch's fd-based tensor pickling breaks when the child exits first."""
    import torch
    if isinstance(x, torch.Tensor):
        return ("__tensor__", x.detach().cpu().numpy().copy())
    if isinstance(x, (list, tuple)):
        return type(x)(_pack(v) for v in x)
    if isinstance(x, dict):
        return {k: _pack(v) for k, v in x.items()}
    return x


def _unpack(x):
    import torch
    if isinstance(x, tuple) and len(x) == 2 and isinstance(x[0], str) and x[0] == "__tensor__":
        return torch.from_numpy(x[1])
    if isinstance(x, (list, tuple)):
        return type(x)(_unpack(v) for v in x)
    if isinstance(x, dict):
        return {k: _unpack(v) for k, v in x.items()}
    return x


def _entry(fn, rank, world, job, args, q):
    try:
        import torch
        torch.set_num_threads(1)
        out = fn(rank, world, job, *args)
        q.put((rank, "ok", _pack(out)))
    except BaseException as e:  # noqa: BLE001
        q.put((rank, "err", f"{e!r}\n{traceback.format_exc()}"))


def run_ranks(fn: Callable, world: int, *args: Any, timeout: float = 120.0) -> List[Any]:
    """fn(rank, world, job_name, *args) -> picklable result; returns results ordered by rank."""
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    job = f"dlt-{os.getpid()}-{uuid.uuid4().hex[:8]}"
    procs = [ctx.Process(target=_entry, args=(fn, r, world, job, args, q), daemon=True) for r in range(world)]
    for p in procs:
        p.start()
    results, errors = {}, []
    import queue
    import time
    deadline = time.time() + timeout
    try:
        while len(results) + len(errors) < world:
            try:
                rank, status, out = q.get(timeout=0.5)
            except queue.Empty:
                dead = [(i, p.exitcode) for i, p in enumerate(procs) if p.exitcode not in (None, 0) and i not in results]
                if dead:
                    errors.append(f"rank(s) died without a result (rank, exit code): {dead}")
                    break
                if time.time() > deadline:
                    errors.append(f"timeout after {timeout}s; finished ranks: {sorted(results)}")
                    break
                continue
            if status == "ok":
                results[rank] = _unpack(out)
            else:
                errors.append(f"rank {rank}: {out}")
    except Exception as e:  # noqa: BLE001
        errors.append(f"queue error: {e!r}")
    finally:
        for p in procs:
            p.join(timeout=5 if not errors else 1)
            if p.is_alive():
                p.kill()
    if errors:
        raise RuntimeError("\n".join(errors))
    return [results[r] for r in range(world)]
