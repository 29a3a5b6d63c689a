#!/usr/bin/env python
"""Summarise `ncu --metrics gpu__time_duration.sum --csv` output of scripts/step_once.py: the kernels of ONE training step."""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr, data = rows[hi], rows[hi + 2:]
ki, vi = hdr.index("Kernel Name"), hdr.index("Metric Value")
names = [(r[ki], float(r[vi].replace(",", "")) / 1000.0) for r in data if len(r) > vi]
idx = [i for i, (n, _) in enumerate(names) if "sgd_twoshot" in n]
step = names[idx[-2] + 1: idx[-1] + 1]
tot = sum(v for _, v in step)
ours = ("tc_gemm", "bn_", "im2col", "col2im", "colsum", "sgd_", "fused_ea", "allreduce", "head_")
agg = collections.OrderedDict()
for n, v in step:
    short = n.split("(")[0].replace("void ", "").replace("<unnamed>::", "")[:60]
    a = agg.setdefault(short, [0, 0.0])
    a[0] += 1
    a[1] += v
mine = sum(v for n, v in step if any(k in n for k in ours))
print(f"# Kernels of one training step (flagship: CIFAR convnet, batch 128, 1 x B200)\n")
print("`ncu --metrics gpu__time_duration.sum --clock-control none` over `python scripts/step_once.py 4` (eager launches,")
print("cold caches, serialised: read the SHARES, not the absolute times; the step runs in 1.17 ms as a CUDA graph).\n")
print(f"{len(step)} kernels, {tot:.0f} us in total; {100 * mine / tot:.0f} % of the time is in this library's own kernels.\n")
print("| kernel | launches | us | share |")
print("|---|---|---|---|")
for k, (c, v) in sorted(agg.items(), key=lambda kv: -kv[1][1]):
    tag = " **(ours)**" if any(o in k for o in ours) else ""
    print(f"| `{k}`{tag} | {c} | {v:.1f} | {100 * v / tot:.1f} % |")
