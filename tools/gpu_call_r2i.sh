set -x
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_attention.py tests/test_gpu_attention_varlen.py tests/test_gpu_backend.py -x -q 2>&1 | tail -4 > gpurun_out/r2i_tests.log
cat gpurun_out/r2i_tests.log
timeout 300 python tools/bench_prefill.py > gpurun_out/r2i_prefill.txt 2>&1
cat gpurun_out/r2i_prefill.txt
for cfg in "--hq 32 --hkv 8" "--hq 16 --hkv 4" "--hq 8 --hkv 2" "--hq 4 --hkv 1"; do
  timeout 200 python tools/bench_attn.py $cfg >> gpurun_out/r2i_attn.txt 2>&1
done
cat gpurun_out/r2i_attn.txt
