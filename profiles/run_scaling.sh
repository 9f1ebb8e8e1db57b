#!/bin/bash
# Weak-scaling lines on one box: gpurun --gpus G -- "bash profiles/run_scaling.sh 'N1 N2 ...' [workloads...]"
NS=$1; shift
WL=${@:-prune}
mkdir -p gpurun_out
for N in $NS; do
  for w in $WL; do
    if [ "$N" = "1" ]; then
      timeout 600 python bench.py --gpus 1 --steps 20 --warmup 5 --workload $w --no-cpu-baseline --no-roofline --loader-batches 0 --sustain 0 \
        > gpurun_out/r02_bench_${w}_n$N.json 2> gpurun_out/r02_bench_${w}_n$N.err
    else
      timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 \
        bench.py --gpus $N --steps 20 --warmup 5 --workload $w --no-cpu-baseline --no-roofline --loader-batches 0 --sustain 0 \
        > gpurun_out/r02_bench_${w}_n$N.json 2> gpurun_out/r02_bench_${w}_n$N.err
    fi
    python -c "
import json
d=json.loads(open('gpurun_out/r02_bench_${w}_n$N.json').read().strip().splitlines()[-1]); print('$w', d['n_gpus'], round(d['value'],1), round(d['ms_per_step'],3), round(d['e2e']['value'],1), d['clocks'])"
  done
done
