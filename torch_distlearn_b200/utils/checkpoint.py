"""Flat params+center checkpoint format (BASELINE.json north_star; SURVEY §5 "Checkpoint / resume").

The reference computes ``netFilename`` / ``optStateFilename`` but never writes them
(examples/EASGD_tester.lua:44-46): checkpointing there is vestigial.  Here one file holds the
flat vectors exactly as they sit in the symmetric heap:

    offset 0   : 8 bytes   magic  b"DLB200CK"
    offset 8   : 4 bytes   version (1), little endian
    offset 12  : 4 bytes   header_len H
    offset 16  : H bytes   JSON header {dtype, numel, padded, world, algo, tau, alpha, step,
                           steps, sections:[names], manifest:[{name, shape, offset, numel}], extra}
    then, 64-byte aligned, one little-endian float32 array of ``padded`` elements per section,
    in the order of ``sections`` ("params", then "center" and/or "momentum" when present).

Rank 0 saves after a ``synchronize*`` call (all ranks hold the same parameters then); loading
reads the file on every rank (or on rank 0 followed by ``synchronizeParameters``).
"""
from __future__ import annotations

import json
import os
import struct
from typing import Any, Dict, Optional

import numpy as np
import torch

MAGIC = b"DLB200CK"
VERSION = 1
_ALIGN = 64


def _flat_of(obj) -> torch.Tensor:
    from ..parallel.comm import SymBuffer
    if isinstance(obj, SymBuffer):
        return obj.tensor
    return obj


def save_checkpoint(path: str, flat, center: Optional[Any] = None, momentum: Optional[Any] = None,
                    algo: str = "AllReduceSGD", tau: int = 0, alpha: float = 0.0, step: int = 0,
                    steps: int = 0, world: int = 1, extra: Optional[Dict[str, Any]] = None) -> None:
    """``flat`` is a FlatParams; ``center``/``momentum`` are flat tensors or SymBuffers of the
    same padded length (``ea.center``, ``flat.momentum``)."""
    sections = {"params": flat.params.tensor}
    if center is not None:
        sections["center"] = _flat_of(center)
    if momentum is not None:
        sections["momentum"] = _flat_of(momentum)
    for k, t in sections.items():
        if t.numel() < flat.padded:
            raise ValueError(f"checkpoint section {k!r} is shorter than the parameter vector")
    header = {
        "dtype": "float32", "numel": flat.numel, "padded": flat.padded, "world": world, "algo": algo,
        "tau": tau, "alpha": alpha, "step": step, "steps": steps, "sections": list(sections),
        "manifest": flat.manifest(), "extra": extra or {},
    }
    hb = json.dumps(header).encode()
    tmp = path + ".tmp"
    with open(tmp, "wb") as f:
        f.write(MAGIC)
        f.write(struct.pack("<II", VERSION, len(hb)))
        f.write(hb)
        for t in sections.values():
            pad = (-f.tell()) % _ALIGN
            f.write(b"\0" * pad)
            f.write(t[:flat.padded].detach().to("cpu", torch.float32).contiguous().numpy().astype("<f4").tobytes())
    os.replace(tmp, path)


def read_checkpoint(path: str) -> Dict[str, Any]:
    """Returns {"header": dict, "params": tensor, "center": tensor?, "momentum": tensor?} (CPU)."""
    with open(path, "rb") as f:
        if f.read(8) != MAGIC:
            raise ValueError(f"{path}: not a distlearn_b200 checkpoint")
        version, hlen = struct.unpack("<II", f.read(8))
        if version != VERSION:
            raise ValueError(f"{path}: unsupported checkpoint version {version}")
        header = json.loads(f.read(hlen).decode())
        out: Dict[str, Any] = {"header": header}
        n = header["padded"]
        for name in header["sections"]:
            f.seek((-f.tell()) % _ALIGN, 1)
            buf = f.read(n * 4)
            if len(buf) != n * 4:
                raise ValueError(f"{path}: truncated section {name!r}")
            out[name] = torch.from_numpy(np.frombuffer(buf, dtype="<f4").copy())
    return out


def load_checkpoint(path: str, flat, center: Optional[Any] = None, momentum: Optional[Any] = None,
                    strict: bool = True) -> Dict[str, Any]:
    """Load into the live buffers (in place, so every view stays valid); returns the header."""
    ck = read_checkpoint(path)
    hdr = ck["header"]
    mine = [(m["name"], tuple(m["shape"]), m["offset"], m["numel"]) for m in flat.manifest()]
    theirs = [(m["name"], tuple(m["shape"]), m["offset"], m["numel"]) for m in hdr["manifest"]]
    same_layout = [(s, o) for _, s, o, _ in mine] == [(s, o) for _, s, o, _ in theirs]
    permuted = not same_layout and sorted((n, s) for n, s, _, _ in mine) == sorted((n, s) for n, s, _, _ in theirs)
    if strict:
        if hdr["numel"] != flat.numel or hdr["padded"] != flat.padded:
            raise ValueError(f"checkpoint has {hdr['numel']} parameters, the model has {flat.numel}")
        if not same_layout and not permuted:
            raise ValueError("checkpoint shape manifest does not match the model")

    def put(dst: torch.Tensor, src: torch.Tensor) -> None:
        if not permuted:
            dst[:flat.padded].copy_(src)
            return
        # same tensors, different layout of the flat vector (FlatParams(order=...)): move them one by one, by name
        where = {n: (o, z) for n, _, o, z in theirs}
        for n, _, o, z in mine:
            so, sz = where[n]
            dst[o:o + z].copy_(src[so:so + sz])

    put(flat.params.tensor, ck["params"])
    if center is not None and "center" in ck:
        put(_flat_of(center), ck["center"])
    if momentum is not None and "momentum" in ck:
        put(_flat_of(momentum), ck["momentum"])
    return hdr
