# 1 GPU: attention kernel with the batched in-kernel merge: parity, ncu --set full (compare DRAM pct with round 1's
# 76.8 %), micro-benchmark at 8 / 1 kv heads, headline bench
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_pdl.py tests/test_gpu_backend.py -x -q 2>&1 | tail -3 > gpurun_out/r2y_tests.log; cat gpurun_out/r2y_tests.log
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 $NCU -k regex:paged_attn_decode -s 2 -c 1 -f -o gpurun_out/r2y_attn python tools/ncu_targets.py attn 256 4096 32 8 > gpurun_out/r2y_ncu_attn.log 2>&1
timeout 300 $NCU -k regex:paged_attn_decode -s 2 -c 1 -f -o gpurun_out/r2y_attn_tp8 python tools/ncu_targets.py attn 256 4096 4 1 > gpurun_out/r2y_ncu_attn8.log 2>&1
for cfg in "--hq 32 --hkv 8" "--hq 4 --hkv 1" "--hq 32 --hkv 8 --seqs 32 --ctx 32768"; do
  timeout 200 python tools/bench_attn.py $cfg >> gpurun_out/r2y_attn.txt 2>&1
done
cat gpurun_out/r2y_attn.txt
timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r2y_bench.json 2> gpurun_out/r2y_bench.err; cut -c1-200 gpurun_out/r2y_bench.json
