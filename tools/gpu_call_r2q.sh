# 1 GPU: re-record the round-2 evidence at HEAD (earlier r2b..r2p outputs were lost with the container):
# headline bench, v2 GEMM timelines, per-impl GEMM table, head-to-head vs the wheel, launch list, configs[3]/[4]
# lines, then the full GPU test suite
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/r2q_smi.txt
timeout 600 python bench.py > gpurun_out/r2q_bench.json 2> gpurun_out/r2q_bench.err; echo "bench rc=$?"
TRACE_FINE=1 timeout 400 python tools/trace_gemm.py "256,4096,4096;256,6144,4096;256,4096,14336;64,4096,4096" auto 64 128 256 > gpurun_out/r2q_trace.txt 2>&1
S="256,6144,4096;256,4096,4096;256,4096,14336;256,28672,4096;64,4096,4096;64,4096,14336;16,4096,4096;16,4096,14336;64,28672,4096;128,1280,8192;128,8192,1024;128,7168,8192;128,8192,3584"
for impl in v2 v3 v4 v4:2; do
  timeout 300 python tools/bench_gemm.py --impl $impl --shapes "$S" >> gpurun_out/r2q_gemm_impls.jsonl 2>&1
done
timeout 900 python tools/bench_vs_wheel.py --out gpurun_out/r2q_vs_wheel.md > gpurun_out/r2q_vs_wheel.jsonl 2> gpurun_out/r2q_vs_wheel.err; echo "vs_wheel rc=$?"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2q_launches.csv python bench.py --profile-step --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2q_ncu.log 2>&1; echo "ncu rc=$?"
python tools/summarize_launches.py gpurun_out/r2q_launches.csv > gpurun_out/r2q_launch_summary.md 2>&1
gzip -f gpurun_out/r2q_launches.csv
timeout 400 python bench.py --model mixtral-8x7b --ctx 2048 --no-cpu-baseline > gpurun_out/r2q_mixtral.json 2> gpurun_out/r2q_mixtral.err; echo "mixtral rc=$?"
timeout 400 python bench.py --quant nvfp4 --batch 32 --ctx 32768 --no-cpu-baseline > gpurun_out/r2q_nvfp4_32k.json 2> gpurun_out/r2q_nvfp4.err; echo "nvfp4 rc=$?"
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2q_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2q_tests.log
tail -3 gpurun_out/r2q_tests.log
cut -c1-300 gpurun_out/r2q_bench.json
