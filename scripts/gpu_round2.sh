mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/t_all3.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/t_all3.log
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 120 $TR --master-port 29541 bench/sweep.py --ops fused_sgd,fused_ea --min-kb 16384 --max-mb 1024 --out gpurun_out/sweep_2b.json > gpurun_out/sweep_2b.log 2>&1; echo "sweep rc=$?"
grep '"impl": "nvls"\|"impl": "twoshot"' gpurun_out/sweep_2b.log | cut -c1-200
for bm in 0 4; do
timeout 120 $TR --master-port 2955$bm bench.py --gpus 2 --steps 60 --warmup 10 --bucket-mb $bm > gpurun_out/bench2_ours_b$bm.log 2>&1; echo "bench b$bm rc=$?"; tail -n 1 gpurun_out/bench2_ours_b$bm.log | cut -c1-330
done
timeout 120 $TR --master-port 29559 bench.py --gpus 2 --steps 60 --warmup 10 --impl nccl --graph 0 > gpurun_out/bench2_nccl.log 2>&1; echo "nccl rc=$?"; tail -n 1 gpurun_out/bench2_nccl.log | cut -c1-330
