#!/bin/bash
# Round 2, last 1-GPU call: the whole GPU test suite, the smoke entry point and the headline bench on the final tree.
mkdir -p gpurun_out
timeout 600 python -m pytest tests -m gpu -q -rs > gpurun_out/r2final_all.log 2>&1; echo "pytest -m gpu rc=$? $(tail -n 1 gpurun_out/r2final_all.log)"; grep -E "^FAILED|^ERROR" gpurun_out/r2final_all.log | head -10 | cut -c1-300
timeout 200 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r2final_smoke.log 2>&1; echo "smoke rc=$? $(tail -n 1 gpurun_out/r2final_smoke.log | cut -c1-200)"
timeout 300 python bench.py --steps 100 --warmup 10 > gpurun_out/r2final_bench1.log 2>&1; echo "bench rc=$? $(tail -n 1 gpurun_out/r2final_bench1.log | cut -c1-400)"
