#!/bin/bash
# Weak-scaling lines on N GPUs of one box: gpurun --gpus N -- "bash profiles/run_scaling.sh N [workloads...]"
N=$1; shift
WL=${@:-prune}
mkdir -p gpurun_out
for w in $WL; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus $N --steps 20 --warmup 5 --workload $w --no-cpu-baseline --no-roofline \
    > gpurun_out/r02_bench_${w}_n$N.json 2> gpurun_out/r02_bench_${w}_n$N.err
  tail -c 900 gpurun_out/r02_bench_${w}_n$N.json
done
