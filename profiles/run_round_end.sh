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
