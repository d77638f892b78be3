set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_nvfp4_moe.py tests/test_gpu_nvfp4.py tests/test_gpu_moe.py -x -q 2>&1 | tail -25 > gpurun_out/r2l_tests.log
cat gpurun_out/r2l_tests.log
timeout 600 python -m pytest tests/test_gpu_vs_wheel.py -x -q -k "fp4 or moe_mm_data or moe" 2>&1 | tail -15 > gpurun_out/r2l_wheel.log
cat gpurun_out/r2l_wheel.log
