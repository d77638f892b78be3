set -x
mkdir -p gpurun_out
timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r3i_bench_tp2.json 2> gpurun_out/r3i_bench_tp2.err; echo "bench rc=$?"
grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"tp_collective": "[a-z]*"\|"pdl_level": [0-9]\|"tiny_tokens": "[a-z]*"' gpurun_out/r3i_bench_tp2.json | head -8; tail -3 gpurun_out/r3i_bench_tp2.err
