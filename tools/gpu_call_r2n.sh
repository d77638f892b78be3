set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
timeout 600 $TR --master-port 29521 bench.py --gpus 2 --steps 10 --warmup 3 --model mixtral-8x7b --ctx 2048 --comm fused > gpurun_out/r2n_mixtral_ep2_fused.json 2> gpurun_out/r2n_mixtral_ep2_fused.err; echo "rc=$?"
tail -c 1500 gpurun_out/r2n_mixtral_ep2_fused.err; tail -c 3000 gpurun_out/r2n_mixtral_ep2_fused.json
timeout 600 $TR --master-port 29522 bench.py --gpus 2 --steps 10 --warmup 3 --model mixtral-8x7b --ctx 2048 --comm nvlink --no-tp-check > gpurun_out/r2n_mixtral_ep2_nvlink.json 2> gpurun_out/r2n_mixtral_ep2_nvlink.err; echo "rc=$?"
tail -c 600 gpurun_out/r2n_mixtral_ep2_nvlink.json
timeout 900 python -m pytest tests/test_gpu_moe.py -x -q -k "direct" 2>&1 | tail -3
