set -x
mkdir -p gpurun_out
timeout 150 python -m pytest tests/test_gpu_decoder.py tests/test_gpu_pdl.py -x -q 2>&1 | tail -3 > gpurun_out/r3h_tests.log; cat gpurun_out/r3h_tests.log
timeout 100 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r3h_bench.json 2> gpurun_out/r3h_bench.err; echo "bench rc=$?"; cut -c1-160 gpurun_out/r3h_bench.json
