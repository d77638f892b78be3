set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_moe.py tests/test_gpu_gemm_block.py tests/test_gpu_nvfp4_moe.py tests/test_gpu_gemm.py tests/test_gpu_decoder.py -x -q 2>&1 | tail -30 > gpurun_out/r2m_tests.log
cat gpurun_out/r2m_tests.log
