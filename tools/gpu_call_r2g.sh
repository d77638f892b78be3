set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_gemm.py tests/test_gpu_gemm_block.py tests/test_gpu_nvfp4.py tests/test_gpu_moe.py -x -q 2>&1 | tail -4 > gpurun_out/r2g_tests.log
cat gpurun_out/r2g_tests.log
timeout 900 python tools/bench_vs_wheel.py --only gemm,fp4,moe --out gpurun_out/r2g_vs_wheel.md > gpurun_out/r2g_vs_wheel.jsonl 2>&1
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
cat gpurun_out/r2g_bench.json | cut -c1-400
