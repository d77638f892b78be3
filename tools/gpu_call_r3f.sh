# 8 GPUs: programmatic launch with plain collectives (the new default under TP) against none, configs[1] and configs[2]
set -x
mkdir -p gpurun_out
N=${1:-8}
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
run() { env $1 timeout 400 $TR --master-port 29513 bench.py --gpus $N --steps 20 --warmup 3 $2 --no-cpu-baseline > gpurun_out/r3f_$3.json 2> gpurun_out/r3f_$3.err; echo "$3 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r3f_$3.json | head -1)"; }
run "B200_PDL=0" "--no-tp-check" tp${N}_pdl0
run "B200_X=1" "" tp${N}_default
run "B200_PDL=0" "--no-tp-check" tp${N}_pdl0_again
run "B200_X=1" "--no-tp-check --model llama-3.1-70b --batch 128 --ctx 8192" c3_tp${N}_default
run "B200_PDL=0" "--no-tp-check --model llama-3.1-70b --batch 128 --ctx 8192" c3_tp${N}_pdl0
