# 2 GPUs: tp_check incl. multimem; bench with comm nvlink / nvls / fused at TP2; allreduce micro-bench
set -x
mkdir -p gpurun_out
NCCL_DEBUG=WARN timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2j_bench_tp2.json 2> gpurun_out/r2j_bench_tp2.err; echo "bench rc=$?"
tail -c 2500 gpurun_out/r2j_bench_tp2.json; tail -5 gpurun_out/r2j_bench_tp2.err
for comm in nvls fused; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus 2 --steps 10 --warmup 3 --comm $comm --no-tp-check > gpurun_out/r2j_bench_tp2_$comm.json 2> gpurun_out/r2j_bench_tp2_$comm.err; echo "bench $comm rc=$?"
python -c "import json;d=json.loads(open('gpurun_out/r2j_bench_tp2_$comm.json').read().strip().splitlines()[-1]);print('$comm',d['value'],d['ms_per_step'],d['config'].get('nvswitch_multicast'))"
done
