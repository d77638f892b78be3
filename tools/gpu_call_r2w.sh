# 2 GPUs: which programmatic-dependent-launch level helps which collective rung (the 4- and 8-GPU lines were slower
# with PDL on the fused multimem kernel); the NVLink collectives' pytest on both data planes
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
for comm in fused nvlink nvls; do for pdl in 0 1 2; do
  B200_PDL=$pdl timeout 300 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 --comm $comm --no-tp-check --no-cpu-baseline > gpurun_out/r2w_tp2_${comm}_pdl$pdl.json 2> gpurun_out/r2w_tp2_${comm}_pdl$pdl.err
  echo "$comm pdl=$pdl $(grep -o '"value": [0-9.]*\|"ms_per_step": [0-9.]*' gpurun_out/r2w_tp2_${comm}_pdl$pdl.json | head -2 | tr '\n' ' ')"
done; done
timeout 900 python -m pytest tests/test_gpu_allreduce.py -x -q > gpurun_out/r2w_allreduce_tests.log 2>&1; echo "allreduce tests rc=$?" >> gpurun_out/r2w_allreduce_tests.log
tail -4 gpurun_out/r2w_allreduce_tests.log
