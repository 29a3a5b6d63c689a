#!/bin/bash
# Round 2, multi-GPU call:  gpurun --gpus N --timeout 1500 -- 'bash scripts/gpu_r2_multi.sh N [full|lite|final]'
# 1 full multi-GPU check program, 2 measured NVLink peaks, 3 collective sweeps (eager + CUDA graph) and the variant
# table, 4 training benches (bucket geometry, strong scaling, EA, async) — logs under gpurun_out/, summaries in profiles/.
N=${1:-2}; MODE=${2:-full}
P=r2; [ "$MODE" = "final" ] && P=r2z   # the final run keeps the first run's logs beside its own
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
t0=$(date +%s)
timeout 400 $TR --master-port 29511 tests/run_multigpu.py > gpurun_out/${P}_mg$N.log 2>&1; echo "multigpu rc=$? ($(( $(date +%s) - t0 ))s)"; tail -n 12 gpurun_out/${P}_mg$N.log | cut -c1-220
if [ "$MODE" != "final" ]; then
timeout 120 $TR --master-port 29561 bench/peak.py --mb 256 --out gpurun_out/nvlink_peak_$N.json > gpurun_out/${P}_peak$N.log 2>&1; echo "peak rc=$?"; tail -n 22 gpurun_out/${P}_peak$N.log | tr -d '\n' | cut -c1-900; echo
mkdir -p profiles && cp gpurun_out/nvlink_peak_$N.json profiles/nvlink_peak.json 2>/dev/null
fi
MAXMB=1024; [ "$MODE" = "lite" ] && MAXMB=64
PM=0,600,750,1000; [ "$MODE" = "final" ] && PM=750,1000
timeout 500 $TR --master-port 29541 bench/sweep.py --max-mb $MAXMB --permille $PM --write-table --out gpurun_out/${P}_sweep_$N.json > gpurun_out/${P}_sweep_$N.log 2>&1; echo "sweep rc=$? ($(( $(date +%s) - t0 ))s)"
grep "variant table" gpurun_out/${P}_sweep_$N.log | cut -c1-400
timeout 300 $TR --master-port 29542 bench/sweep.py --max-mb 64 --permille 0,750,1000 --graph --write-table --out gpurun_out/${P}_sweep_${N}_graph.json > gpurun_out/${P}_sweep_${N}_graph.log 2>&1; echo "sweep graph rc=$? ($(( $(date +%s) - t0 ))s)"
grep "variant table" gpurun_out/${P}_sweep_${N}_graph.log | cut -c1-400
# install the freshly measured table for the benches below (world N entry)
python scripts/merge_variant_table.py gpurun_out/variant_table_${N}_graph.json $N; echo "table merge rc=$?"
b() { name=$1; shift; timeout 200 $TR --master-port 29551 bench.py --gpus $N --steps 100 --warmup 10 "$@" > gpurun_out/${P}_bench${N}_$name.log 2>&1; echo "bench $name rc=$? $(tail -n 1 gpurun_out/${P}_bench${N}_$name.log | cut -c1-160)"; }
[ "$MODE" != "final" ] && b b0 --baselines 0 --bucket-mb 0
[ "$MODE" != "final" ] && b exe05 --baselines 0 --bucket-mb 0.5 --flat-order execution
if [ "$MODE" = "full" ]; then
b exe05_b74 --baselines 0 --bucket-mb 0.5 --flat-order execution --bucket-blocks 74
b exe1 --baselines 0 --bucket-mb 1 --flat-order execution
fi
b default
b gb32 --global-batch 32
if [ "$MODE" = "final" ]; then b ea --algo ea --baselines 0; else b ea --algo ea; fi
echo "benches done ($(( $(date +%s) - t0 ))s)"
IMPLS="ours ours_eager hostloop"; [ "$MODE" = "lite" ] && IMPLS="ours hostloop"; [ "$MODE" = "final" ] && IMPLS="ours"
for impl in $IMPLS; do
timeout 150 $TR --master-port 29547 bench/bench_async.py --impl $impl --steps 100 > gpurun_out/${P}_async${N}_$impl.log 2>&1; echo "async $impl rc=$? $(tail -n 1 gpurun_out/${P}_async${N}_$impl.log | cut -c1-300)"
done
for impl in ours nccl; do
timeout 120 $TR --master-port 29546 bench/bench_ea.py --impl $impl --batch 1 --steps 2000 > gpurun_out/${P}_mnist_ea${N}_$impl.log 2>&1; echo "mnist ea $impl rc=$? $(tail -n 1 gpurun_out/${P}_mnist_ea${N}_$impl.log | cut -c1-300)"
done
if [ "$MODE" = "full" ]; then timeout 120 $TR --master-port 29546 bench/bench_ea.py --impl ours --fused 0 --batch 1 --steps 2000 > gpurun_out/${P}_mnist_ea${N}_ours_unfused.log 2>&1; echo "mnist ea ours unfused rc=$? $(tail -n 1 gpurun_out/${P}_mnist_ea${N}_ours_unfused.log | cut -c1-300)"; fi
echo "total $(( $(date +%s) - t0 ))s"
