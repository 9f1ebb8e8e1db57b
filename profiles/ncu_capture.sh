#!/bin/bash
# Runs ON THE GPU BOX (gpurun): launch list of one eager training step + `ncu --set full` captures of the step's
# kernels, one report per kernel family.  Usage: bash profiles/ncu_capture.sh <tag> [kernel regex ...]
# Reports land in gpurun_out/<tag>_<name>.ncu-rep; summarise them here with profiles/ncu_summary.py.
set -u
TAG=${1:-r02}
shift || true
CMD="python bench.py --profile-step"
OUT=gpurun_out
mkdir -p $OUT
$CMD > $OUT/${TAG}_plain.log 2>&1 || { echo "plain run failed"; tail -20 $OUT/${TAG}_plain.log; exit 1; }
ncu --metrics gpu__time_duration.sum --clock-control none --profile-from-start off --csv \
    --log-file $OUT/${TAG}_ncu_launches_step.csv $CMD > $OUT/${TAG}_ncu_launches.log 2>&1
echo "launch list rc=$?"
if [ $# -eq 0 ]; then
  set -- attn_fwd_kernel attn_bwd_kernel attn_bwd_prep conv0_fwd_tc conv0_bwd_fused posconv_tc_kernel \
         epilogue_bwd_kernel layernorm_fwd layernorm_bwd adamw_kernel distill_loss_fwd distill_loss_bwd sumsq_kernel \
         "gemm_bf16_kernel<.int.256, .bool.0, .bool.0, .int.2, .int.103>" \
         "gemm_bf16_kernel<.int.192, .bool.0, .bool.0, .int.2, .int.45>" \
         "gemm_bf16_kernel<.int.256, .bool.1, .bool.1, .int.2, .int.384>" \
         "gemm_bf16_kernel<.int.256, .bool.0, .bool.1, .int.2, .int.45>" \
         "gemm_bf16_kernel<.int.64, .bool.1, .bool.1, .int.1, .int.45>" \
         "gemm_bf16_kernel<.int.256, .bool.0, .bool.1, .int.2, .int.532>"
fi
for k in "$@"; do
  name=$(echo "$k" | tr -c 'A-Za-z0-9_' '_' | sed 's/__*/_/g')
  ncu --set full --clock-control none --import-source on --profile-from-start off --kernel-name-base demangled -k "regex:$k" -c 2 -f \
      -o $OUT/${TAG}_${name} $CMD > $OUT/${TAG}_${name}.log 2>&1
  echo "$k rc=$?"
  # the reports are ~35 MB each and gpurun_out/ travels back only below 64 MiB: keep the raw metric table and the
  # per-source-line page as CSV, drop the report
  if [ -f $OUT/${TAG}_${name}.ncu-rep ]; then
    ncu -i $OUT/${TAG}_${name}.ncu-rep --page raw --csv > $OUT/${TAG}_${name}.raw.csv 2>/dev/null
    ncu -i $OUT/${TAG}_${name}.ncu-rep --page source --csv 2>/dev/null | gzip > $OUT/${TAG}_${name}.source.csv.gz
    rm -f $OUT/${TAG}_${name}.ncu-rep
  fi
done
du -sh $OUT
ls -la $OUT | head -60
