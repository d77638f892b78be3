set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
timeout 900 $TR --master-port 29531 bench.py --gpus $N --steps 10 --warmup 3 --model llama-3.1-70b --batch 128 --ctx 8192 > gpurun_out/r2o_c3_tp$N.json 2> gpurun_out/r2o_c3_tp$N.err; echo "rc=$?"
tail -c 400 gpurun_out/r2o_c3_tp$N.err; grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*\|"tp_collective": "[a-z]*"' gpurun_out/r2o_c3_tp$N.json | head -6
if [ "$N" = "8" ]; then
for comm in fused nvlink; do
timeout 600 $TR --master-port 29532 bench.py --gpus $N --steps 10 --warmup 3 --model mixtral-8x7b --ctx 2048 --comm $comm --no-tp-check > gpurun_out/r2o_mixtral_ep${N}_$comm.json 2> gpurun_out/r2o_mixtral_ep${N}_$comm.err; echo "rc=$?"
grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' gpurun_out/r2o_mixtral_ep${N}_$comm.json | head -4
done
fi
