#!/usr/bin/env python
"""Input pipeline throughput on the host (no GPU needed): the native batch loader (csrc/loader.cpp) against the Python
batcher of utils/data.py, CIFAR-shaped samples, batch 128, label-uniform sampling (examples/cifar10.lua:53-54).

    python bench/bench_loader.py [--batches 300] [--threads 1,2,4]

Prints one JSON line per arm: samples/s delivered to the consumer (which touches nothing: the GPU step would run here)."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402

from torch_distlearn_b200.utils.data import cifar_dataset  # noqa: E402
from torch_distlearn_b200.utils.native_loader import NativeBatcher  # noqa: E402


def main() -> None:
    ap = argparse.ArgumentParser()
    ap.add_argument("--batches", type=int, default=300)
    ap.add_argument("--batch", type=int, default=128)
    ap.add_argument("--samples", type=int, default=50000)
    ap.add_argument("--threads", default="1,2,4")
    args = ap.parse_args()
    torch.set_num_threads(1)
    g = torch.Generator().manual_seed(0)
    x = torch.randint(0, 256, (args.samples, 3, 32, 32), generator=g, dtype=torch.uint8)
    y = torch.randint(0, 10, (args.samples,), generator=g)
    common = {"bench": "input_pipeline", "batch": args.batch, "sample": "3x32x32", "sampler": "label-uniform", "batches": args.batches}

    ds = cifar_dataset(num_samples=args.samples)
    get, _ = ds.sampledBatcher(args.batch, "label-uniform")
    for _ in range(5):
        get()
    t0 = time.perf_counter()
    n_py = max(args.batches // 10, 20)
    for _ in range(n_py):
        get()
    dt = time.perf_counter() - t0
    print(json.dumps({**common, "impl": "python (utils/data.py, synthesises prototype + noise per batch)", "batches": n_py,
                      "samples_per_s": round(n_py * args.batch / dt, 1), "ms_per_batch": round(dt / n_py * 1e3, 3)}))

    for threads in [int(t) for t in args.threads.split(",")]:
        nb = NativeBatcher(x, y, batch_size=args.batch, sampler="label-uniform", scale=1 / 255, mean=[0.49, 0.48, 0.45],
                           std=[0.25, 0.24, 0.26], threads=threads, depth=max(4, 2 * threads), hold=1, pin=False)
        for _ in range(10):
            nb.next()
        t0 = time.perf_counter()
        for _ in range(args.batches):
            nb.next()
        dt = time.perf_counter() - t0
        nb.close()
        print(json.dumps({**common, "impl": f"native, {threads} worker thread(s): uint8 -> float32, per-channel normalisation",
                          "samples_per_s": round(args.batches * args.batch / dt, 1), "ms_per_batch": round(dt / args.batches * 1e3, 3)}))


if __name__ == "__main__":
    main()
