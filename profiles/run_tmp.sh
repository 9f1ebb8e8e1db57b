set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gemm_gpu.py -q -x 2>&1 | tail -4
timeout 900 python -m pytest tests/test_parity_gpu.py tests/test_parity_real_gpu.py tests/test_kernels_gpu.py -q -x -k "not full_size" 2>&1 | tail -3
timeout 600 python bench.py --no-cpu-baseline --gemm-table gpurun_out/r02_gemm_table_i.txt > /dev/null 2>&1; head -4 gpurun_out/r02_gemm_table_i.txt; tail -1 gpurun_out/r02_gemm_table_i.txt
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r02_bench_w16.json 2> gpurun_out/r02_bench_w16.err; tail -c 300 gpurun_out/r02_bench_w16.json
timeout 400 python bench.py --workload large --no-cpu-baseline --no-roofline > gpurun_out/r02_bench_large_w16.json 2>/dev/null; tail -c 300 gpurun_out/r02_bench_large_w16.json
