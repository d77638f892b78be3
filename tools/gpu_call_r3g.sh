set -x
mkdir -p gpurun_out
timeout 120 python __graft_entry__.py smoke 2>&1 | tail -2
timeout 200 python bench.py --steps 5 --warmup 3 > gpurun_out/r3g_bench.json 2> gpurun_out/r3g_bench.err; echo "bench rc=$?"; cut -c1-160 gpurun_out/r3g_bench.json; tail -3 gpurun_out/r3g_bench.err
timeout 100 python bench.py --impl reference --model qwen2.5-0.5b --batch 8 --ctx 512 --steps 5 --warmup 3 > gpurun_out/r3g_c1_cpu.json 2> gpurun_out/r3g_c1_cpu.err; echo "ref rc=$?"; cut -c1-400 gpurun_out/r3g_c1_cpu.json
