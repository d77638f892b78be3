set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/r2a_smi.txt
timeout 900 python -m pytest tests/test_gpu_plugin.py tests/test_gpu_backend.py tests/test_gpu_nvfp4.py -x -q > gpurun_out/r2a_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2a_tests.log
timeout 600 python bench.py > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err; echo "bench rc=$?"
timeout 900 python tools/bench_vs_wheel.py --out gpurun_out/r2a_vs_wheel.md > gpurun_out/r2a_vs_wheel.jsonl 2> gpurun_out/r2a_vs_wheel.err; echo "vs_wheel rc=$?"
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 $NCU -k regex:gemm_fp8_v3 -s 2 -c 1 -f -o gpurun_out/r2a_gemm_v3_8192 python tools/ncu_targets.py fp8 8192 8192 8192 v3 > gpurun_out/r2a_ncu1.log 2>&1
timeout 300 $NCU -k regex:gemm_nvfp4 -s 2 -c 1 -f -o gpurun_out/r2a_nvfp4_8192 python tools/ncu_targets.py nvfp4 8192 8192 8192 > gpurun_out/r2a_ncu2.log 2>&1
timeout 300 $NCU -k regex:gemm_fp8 -s 2 -c 1 -f -o gpurun_out/r2a_gemm_o_proj python tools/ncu_targets.py fp8 256 4096 4096 > gpurun_out/r2a_ncu3.log 2>&1
ls -la gpurun_out | tail -20
