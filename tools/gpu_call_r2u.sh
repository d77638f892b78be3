# N GPUs (argument, default 2): NVLink collectives' parity tests, headline bench with tp_check, reference arm under
# torchrun, BASELINE configs[2] (Llama-3.1-70B-shaped TP=N), all-reduce latency table
set -x
N=${1:-2}
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
if [ "$N" = "2" ]; then
timeout 600 python -m pytest tests/test_gpu_allreduce.py -x -q > gpurun_out/r2u_allreduce_tests.log 2>&1; echo "allreduce tests rc=$?" >> gpurun_out/r2u_allreduce_tests.log
tail -3 gpurun_out/r2u_allreduce_tests.log
timeout 300 $TR --master-port 29512 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > gpurun_out/r2u_ref_tp$N.json 2> gpurun_out/r2u_ref_tp$N.err; echo "ref rc=$?"
cut -c1-300 gpurun_out/r2u_ref_tp$N.json
fi
timeout 600 $TR --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/r2u_bench_tp$N.json 2> gpurun_out/r2u_bench_tp$N.err; echo "bench rc=$?"
tail -c 600 gpurun_out/r2u_bench_tp$N.err; grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"tp_collective": "[a-z]*"\|"tp_check": {[^}]*}' gpurun_out/r2u_bench_tp$N.json | head -8
B200_PDL=0 timeout 600 $TR --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --no-tp-check --no-cpu-baseline > gpurun_out/r2u_bench_tp${N}_pdl0.json 2> gpurun_out/r2u_bench_tp${N}_pdl0.err; echo "bench pdl0 rc=$?"
grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' gpurun_out/r2u_bench_tp${N}_pdl0.json | head -2
timeout 900 $TR --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 --model llama-3.1-70b --batch 128 --ctx 8192 --no-cpu-baseline > gpurun_out/r2u_c3_tp$N.json 2> gpurun_out/r2u_c3_tp$N.err; echo "c3 rc=$?"
tail -c 400 gpurun_out/r2u_c3_tp$N.err; grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"tp_collective": "[a-z]*"' gpurun_out/r2u_c3_tp$N.json | head -6
timeout 300 $TR --master-port 29541 tools/bench_allreduce.py > gpurun_out/r2u_allreduce_tp$N.jsonl 2> gpurun_out/r2u_allreduce_tp$N.err; echo "ar bench rc=$?"
tail -12 gpurun_out/r2u_allreduce_tp$N.jsonl | cut -c1-200
if [ "$N" = "8" ]; then
for comm in fused nvlink; do
timeout 600 $TR --master-port 29532 bench.py --gpus $N --steps 10 --warmup 3 --model mixtral-8x7b --ctx 2048 --comm $comm --no-tp-check --no-cpu-baseline > gpurun_out/r2u_mixtral_ep${N}_$comm.json 2> gpurun_out/r2u_mixtral_ep${N}_$comm.err; echo "rc=$?"
grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' gpurun_out/r2u_mixtral_ep${N}_$comm.json | head -2
done
timeout 600 $TR --master-port 29533 bench.py --gpus $N --steps 10 --warmup 3 --quant nvfp4 --batch 32 --ctx 32768 --no-tp-check --no-cpu-baseline > gpurun_out/r2u_nvfp4_tp$N.json 2> gpurun_out/r2u_nvfp4_tp$N.err; echo "rc=$?"
grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' gpurun_out/r2u_nvfp4_tp$N.json | head -2
fi
