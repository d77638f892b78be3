# 2 GPUs: NVLink all-reduce / fused / DCP parity, bench --gpus 2 with tp_check, reference arm under torchrun
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_allreduce.py -x -q -s -k "2" > gpurun_out/r2b_allreduce.log 2>&1; echo "allreduce rc=$?" >> gpurun_out/r2b_allreduce.log
timeout 600 python -m pytest tests/test_gpu_plugin.py -x -q > gpurun_out/r2b_plugin.log 2>&1; echo "plugin rc=$?" >> gpurun_out/r2b_plugin.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2b_bench_tp2.json 2> gpurun_out/r2b_bench_tp2.err; echo "bench rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --impl reference --gpus 2 --steps 3 --warmup 1 > gpurun_out/r2b_ref_tp2.json 2> gpurun_out/r2b_ref_tp2.err; echo "ref rc=$?"
tail -3 gpurun_out/r2b_allreduce.log
