set -x
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2k_bench_tp2.json 2> gpurun_out/r2k_bench_tp2.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/r2k_bench_tp2.json
