set -x
mkdir -p gpurun_out
timeout 120 tools/l2fill_probe > gpurun_out/r2p_l2fill.jsonl 2>&1; head -6 gpurun_out/r2p_l2fill.jsonl
timeout 600 python bench.py > gpurun_out/r2p_bench.json 2> gpurun_out/r2p_bench.err; echo "bench rc=$?"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2p_launches.csv python bench.py --profile-step --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2p_ncu.log 2>&1; echo "ncu rc=$?"
python tools/summarize_launches.py gpurun_out/r2p_launches.csv > gpurun_out/r2p_launch_summary.md 2>&1; head -30 gpurun_out/r2p_launch_summary.md
