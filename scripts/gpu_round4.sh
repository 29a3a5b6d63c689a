mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1"
timeout 200 $TR --master-port 29533 tests/run_multigpu.py --quick > gpurun_out/mg4.log 2>&1; echo "mg4 rc=$?"; grep -v "^$\|OMP\|\*\*\*" gpurun_out/mg4.log | tail -n 12
timeout 60 $TR --master-port 29545 bench.py --gpus 4 --steps 60 --warmup 10 --impl nccl > gpurun_out/bench4_nccl.log 2>&1; echo "nccl rc=$?"; tail -n 1 gpurun_out/bench4_nccl.log | cut -c1-200
timeout 60 $TR --master-port 29546 bench.py --gpus 4 --steps 60 --warmup 10 > gpurun_out/bench4_tc.log 2>&1; echo "ours rc=$?"; tail -n 1 gpurun_out/bench4_tc.log | cut -c1-200
