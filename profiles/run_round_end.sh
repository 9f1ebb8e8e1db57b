#!/bin/bash
# Round-end check on one B200: full GPU suite, smoke(), the default bench line and one line per other workload.
# Run with: gpurun --timeout 2700 -- "bash profiles/run_round_end.sh"; results come back under gpurun_out/.
set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/r02_pytest_full.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r02_pytest_full.log
tail -5 gpurun_out/r02_pytest_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r02_smoke.log 2>&1; tail -2 gpurun_out/r02_smoke.log
timeout 600 python bench.py > gpurun_out/r02_bench_prune.json 2> gpurun_out/r02_bench_prune.err; tail -c 600 gpurun_out/r02_bench_prune.json
for w in distill wavlm large final; do
  timeout 400 python bench.py --workload $w --no-cpu-baseline > gpurun_out/r02_bench_$w.json 2> gpurun_out/r02_bench_$w.err; python -c "
import json,sys
d=json.loads(open('gpurun_out/r02_bench_$w.json').read().strip().splitlines()[-1]); print('$w', d['value'], d['ms_per_step'], d['e2e']['value'], d['roofline']['frac'])"
done
# launch list of one eager step at HEAD (serialised, cold: shares only)
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv --log-file gpurun_out/r02_ncu_launches_step_head.csv python bench.py --profile-step > gpurun_out/r02_ncu_launches_head.log 2>&1
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r02_bench_reference_arm.json 2> gpurun_out/r02_bench_reference_arm.err; tail -c 400 gpurun_out/r02_bench_reference_arm.json
# per-kernel DRAM traffic of the same eager step (bench.py reads profiles/r02_step_dram_traffic.json / r02_gemm_dram_traffic.json)
timeout 900 ncu --profile-from-start off --clock-control none --page raw --csv --log-file gpurun_out/r02_step_traffic_raw.csv \
  --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum python bench.py --profile-step > gpurun_out/r02_step_traffic.log 2>&1
python profiles/step_traffic.py gpurun_out/r02_step_traffic_raw.csv gpurun_out/r02_step_dram_traffic.json gpurun_out/r02_gemm_dram_traffic.json | tail -12
