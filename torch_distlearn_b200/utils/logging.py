"""Rank-0 printing, JSONL metrics and an ``optim.Logger`` equivalent.

The reference silences nodes > 1 by overwriting ``print`` and ``xlua.progress``
(examples/mnist.lua:20-23), logs the tester's error rates with ``optim.Logger``
(examples/EASGD_tester.lua:36-47,161) and has a dead plotting path (flag typo) —
plotting is left out here."""
from __future__ import annotations

import json
import os
import sys
import time
from typing import Any, Dict, List, Optional


class RankLogger:
    """``log.print(...)`` prints on node 1 only; ``log.metric(...)`` appends a JSON line."""

    def __init__(self, node_index: int = 1, jsonl_path: Optional[str] = None, quiet: bool = False):
        self.node_index, self.quiet = node_index, quiet
        self._f = None
        if jsonl_path and node_index == 1:
            os.makedirs(os.path.dirname(os.path.abspath(jsonl_path)), exist_ok=True)
            self._f = open(jsonl_path, "a")
        self._t0 = time.time()

    def print(self, *a, **k) -> None:
        if self.node_index == 1 and not self.quiet:
            print(*a, **k)
            sys.stdout.flush()

    def progress(self, i: int, n: int, every: int = 50) -> None:
        """xlua.progress stand-in."""
        if self.node_index == 1 and not self.quiet and (i % every == 0 or i == n):
            print(f"\r  [{i}/{n}]", end="" if i < n else "\n")
            sys.stdout.flush()

    def metric(self, **fields: Any) -> None:
        if self._f is not None:
            fields.setdefault("t", round(time.time() - self._t0, 3))
            self._f.write(json.dumps(fields) + "\n")
            self._f.flush()

    def close(self) -> None:
        if self._f is not None:
            self._f.close()
            self._f = None


class Logger:
    """``optim.Logger(path)``: ``add{name=value,...}`` appends a row to a tab-separated log
    (examples/EASGD_tester.lua:47,161).  ``style``/``plot`` are accepted and ignored."""

    def __init__(self, path: str):
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        self.path = path
        self.names: Optional[List[str]] = None
        self.rows: List[Dict[str, float]] = []

    def add(self, row: Dict[str, float]) -> None:
        if self.names is None:
            self.names = list(row)
            with open(self.path, "w") as f:
                f.write("\t".join(self.names) + "\n")
        self.rows.append(dict(row))
        with open(self.path, "a") as f:
            f.write("\t".join(f"{row.get(n, float('nan')):.6g}" for n in self.names) + "\n")

    def style(self, *_a, **_k) -> None:
        pass

    def plot(self, *_a, **_k) -> None:
        pass
