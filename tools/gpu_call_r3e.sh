# 2 GPUs: does programmatic launch help under TP once the collectives stop triggering their dependents early?
# B200_PDL_AR: 0 = collective triggers at entry, 1 = after its last barrier, 2 = collective is a plain kernel
set -x
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
run() { env $1 timeout 300 $TR --master-port 29513 bench.py --gpus 2 --steps 20 --warmup 3 --comm $2 --no-tp-check --no-cpu-baseline > gpurun_out/r3e_$3.json 2> gpurun_out/r3e_$3.err; echo "$3 $(grep -o '"ms_per_step": [0-9.]*' gpurun_out/r3e_$3.json | head -1)"; }
run "B200_PDL=0" fused fused_pdl0
run "B200_PDL=2 B200_PDL_AR=0" fused fused_pdl2_ar0
run "B200_PDL=2 B200_PDL_AR=1" fused fused_pdl2_ar1
run "B200_PDL=2 B200_PDL_AR=2" fused fused_pdl2_ar2
run "B200_PDL=1 B200_PDL_AR=2" fused fused_pdl1_ar2
run "B200_PDL=0" fused fused_pdl0_again
run "B200_PDL=2 B200_PDL_AR=2" nvlink nvlink_pdl2_ar2
run "B200_PDL=0" nvlink nvlink_pdl0
