mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29533 tests/run_multigpu.py > gpurun_out/mg8.log 2>&1; echo "mg8 rc=$?"
timeout 600 $TR --master-port 29541 bench/sweep.py --max-mb 1024 --out gpurun_out/sweep_8.json --write-table > gpurun_out/sweep_8.log 2>&1; echo "sweep rc=$?"
timeout 600 $TR --master-port 29542 bench/sweep.py --max-mb 64 --graph --out gpurun_out/sweep_8_graph.json > gpurun_out/sweep_8_graph.log 2>&1; echo "sweepg rc=$?"
timeout 600 $TR --master-port 29543 bench.py --gpus 8 --steps 100 --warmup 10 > gpurun_out/bench8_ours.log 2>&1; echo "ours rc=$?"
timeout 600 $TR --master-port 29544 bench.py --gpus 8 --steps 100 --warmup 10 --impl nccl > gpurun_out/bench8_nccl.log 2>&1; echo "nccl rc=$?"
timeout 600 $TR --master-port 29545 bench.py --gpus 8 --steps 100 --warmup 10 --impl nccl --graph 0 > gpurun_out/bench8_nccl_nograph.log 2>&1; echo "nccl-nograph rc=$?"
tail -n 3 gpurun_out/mg8.log; tail -n 1 gpurun_out/bench8_ours.log gpurun_out/bench8_nccl.log gpurun_out/bench8_nccl_nograph.log | cut -c1-400
