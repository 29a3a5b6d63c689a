#!/usr/bin/env python
"""Turn bench/sweep.py JSON files into the markdown tables committed under profiles/.

    python bench/report.py gpurun_out/sweep_8.json [more.json ...] > profiles/sweep_8gpu.md
"""
from __future__ import annotations

import collections
import json
import sys


def human(b: int) -> str:
    for unit, div in (("GiB", 1 << 30), ("MiB", 1 << 20), ("KiB", 1 << 10)):
        if b >= div:
            return f"{b // div} {unit}"
    return f"{b} B"


def main(paths):
    for path in paths:
        d = json.load(open(path))
        rows, world = d["rows"], d["world"]
        graph = bool(rows and rows[0].get("graph"))
        print(f"## {path.split('/')[-1]} — {world} x B200, transport {d['transport']['mode']}, "
              f"multicast={d['transport']['multicast']}, timing={'CUDA-graph replay' if graph else 'eager launches'}\n")
        print(f"Per cell: time per call in us (max over ranks) / bus bandwidth GB/s / fraction of that variant's own "
              f"NVLink roofline at {d['link_gbs']:.0f} GB/s per direction ({d.get('link_gbs_source', 'recipe constant')}).  "
              f"`ll` = push path (no barrier), `hybridP` = pipelined slice pass with P/1000 of every slice reduced in the "
              f"switch and the rest over 256-bit P2P loads.  `nccl` = torch.distributed NCCL doing the same "
              f"work (all_reduce AVG + p.add_ for fused_sgd; 3 elementwise ops + all_reduce + add for fused_ea).\n")
        t = collections.defaultdict(dict)
        for r in rows:
            t[(r["op"], r["bytes"])][r["impl"]] = r
        for op in ("fused_sgd", "allreduce", "broadcast", "fused_ea"):
            sizes = sorted(b for (o, b) in t if o == op)
            if not sizes:
                continue
            seen = {i for b in sizes for i in t[(op, b)]}
            hybrids = sorted((i for i in seen if i.startswith("hybrid")), key=lambda i: int(i[6:]))
            impls = [i for i in ["ll", "oneshot", "twoshot", "nvls"] + hybrids + ["nccl"] if i in seen]
            print(f"### {op}\n")
            print("| bytes | " + " | ".join(impls) + " | best ours / nccl |")
            print("|---|" + "---|" * (len(impls) + 1))
            for b in sizes:
                cells = []
                for i in impls:
                    r = t[(op, b)].get(i)
                    cells.append(f"{r['us']:.1f} / {r['busbw_gbs']:.0f} / {r['frac_of_roofline']:.2f}" if r else "—")
                ours = [r["us"] for i, r in t[(op, b)].items() if i != "nccl"]
                nccl = t[(op, b)].get("nccl")
                speed = f"{nccl['us'] / min(ours):.2f}x" if nccl and ours else "—"
                print(f"| {human(b)} | " + " | ".join(cells) + f" | {speed} |")
            print()


if __name__ == "__main__":
    main(sys.argv[1:])
