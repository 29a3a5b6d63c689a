#!/usr/bin/env python
"""profiles/bench_r2.md from the raw bench logs under profiles/r2/ (last JSON line of each)."""
import glob
import json
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
R2 = os.path.join(ROOT, "profiles", "r2")


def last_json(path):
    for line in reversed(open(path).read().strip().splitlines()):
        line = line.strip()
        if line.startswith("{"):
            try:
                return json.loads(line)
            except ValueError:
                continue
    return None


out = ["# Training benches of round 2 (B200 boxes of this pool; raw logs in `profiles/r2/`)\n",
       "Every row is one `bench.py` run: device-timed (CUDA events, max over ranks), 100 steps after 10 warm-up steps, synthetic "
       "CIFAR-shaped data, fp32, SM clocks 1965 MHz with no throttle reason in every run.  `vs fp32` / `vs tf32`: against the torch + "
       "NCCL arms measured in the same process right after (strict-fp32 cuDNN, the better of flat all-reduce and DDP; cuDNN with TF32 allowed).\n",
       "| log | GPUs | algo | global batch | buckets | samples/s | ms/step | e2e samples/s | vs fp32 | vs tf32 | NCCL fp32 / tf32 / DDP samples/s |",
       "|---|---|---|---|---|---|---|---|---|---|---|"]
for path in sorted(glob.glob(os.path.join(R2, "r2*_bench*.log"))):
    d = last_json(path)
    if not d or "value" not in d:
        continue
    b = d.get("baselines") or {}
    base = " / ".join(str(int(b[k]["value"])) if k in b and "value" in b[k] else "-" for k in ("nccl_fp32", "nccl_tf32", "nccl_ddp_fp32"))
    cfg = d["config"]
    out.append(f"| `{os.path.basename(path)}` | {d['n_gpus']} | {cfg.get('algo', 'sgd')} | {cfg['global_batch']} | "
               f"{cfg['bucket_mb']} MiB, {cfg['flat_order']} | **{d['value']:.0f}** | {d['ms_per_step']:.4f} | {d['e2e']['value']:.0f} | "
               f"{d.get('vs_baseline_fp32') or '-'} | {d.get('vs_baseline_tf32') or '-'} | {base if b else '(not run)'} |")
out.append("")
out.append("| log | what | result |")
out.append("|---|---|---|")
for path in sorted(glob.glob(os.path.join(R2, "r2*_async*.log")) + glob.glob(os.path.join(R2, "r2*_mnist_ea*.log"))):
    d = last_json(path)
    if not d:
        continue
    if d.get("bench") == "async_easgd_cifar":
        out.append(f"| `{os.path.basename(path)}` | async EASGD CIFAR, {d['clients']} clients, tau {d['tau']}, impl `{d['impl']}` | "
                   f"{d['client_samples_per_s_total']:.0f} samples/s over all clients, {d.get('client_samples_per_s_each', 0):.0f} per client, "
                   f"{d['syncs_per_s']:.0f} syncs/s |")
    else:
        out.append(f"| `{os.path.basename(path)}` | MNIST AllReduceEA tau {d['tau']}, batch {d['per_gpu_batch']}, {d['n_gpus']} GPUs, impl `{d['impl']}`"
                   f"{' (one-launch local step)' if d.get('fused_local_step') else ''} | {d['samples_per_s']:.0f} samples/s, {d['ms_per_step'] * 1e3:.1f} us/step |")
open(os.path.join(ROOT, "profiles", "bench_r2.md"), "w").write("\n".join(out) + "\n")
print("\n".join(out))
