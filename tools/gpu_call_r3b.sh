# 1 GPU, final state: full GPU test suite, headline line, launch list, head-to-head vs the wheel, then the
# groups-per-sub-warp variant of per_token_group_fp8_quant (tools/group_quant_gps4.patch) built and tested on the box
set -x
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r3b_bench.json 2> gpurun_out/r3b_bench.err; echo "bench rc=$?"
cut -c1-200 gpurun_out/r3b_bench.json
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r3b_launches.csv python bench.py --profile-step --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r3b_ncu.log 2>&1; echo "ncu rc=$?"
python tools/summarize_launches.py gpurun_out/r3b_launches.csv "ncu launch list of ONE eager decode step (Llama-3.1-8B-shaped, FP8, batch 256, ctx 4096) at the end of round 2 (v5 GEMMs, programmatic launches), 1 x B200" > gpurun_out/r3b_launch_summary.md 2>&1
gzip -f gpurun_out/r3b_launches.csv
timeout 900 python tools/bench_vs_wheel.py --out gpurun_out/r3b_vs_wheel.md > gpurun_out/r3b_vs_wheel.jsonl 2> gpurun_out/r3b_vs_wheel.err; echo "vs_wheel rc=$?"
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/r3b_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r3b_tests.log
tail -4 gpurun_out/r3b_tests.log
timeout 200 python tools/bench_group_quant.py > gpurun_out/r3b_group_quant_base.txt 2>&1
patch -p1 < tools/group_quant_gps4.patch && make -C vllm-fork_b200 -j8 > gpurun_out/r3b_gps4_build.log 2>&1 && {
  timeout 200 python tools/bench_group_quant.py > gpurun_out/r3b_group_quant_gps4.txt 2>&1
  timeout 600 python -m pytest tests/test_gpu_fp8_quant.py tests/test_gpu_vs_wheel.py -x -q -k "group" 2>&1 | tail -3 > gpurun_out/r3b_gps4_tests.log
}
tail -5 gpurun_out/r3b_group_quant_base.txt gpurun_out/r3b_group_quant_gps4.txt gpurun_out/r3b_gps4_tests.log
