#!/bin/bash
# Round 2, last GPU seconds: launch list of one training step of the final tree (ncu, one metric).
mkdir -p gpurun_out
timeout 60 python scripts/step_once.py 4 > gpurun_out/r2n_step_once_plain.log 2>&1 && \
timeout 100 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2n_launches.csv python scripts/step_once.py 4 > gpurun_out/r2n_step_once_ncu.log 2>&1
echo "rc=$? $(wc -l < gpurun_out/r2n_launches.csv) csv lines"
