# 1 GPU: evidence at the v5 / PDL / folded-reduction state: headline bench, launch list, ncu --set full of the
# attention kernel (traffic) and of the v5 GEMM, head-to-head vs the wheel, configs[0] on both arms
set -x
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/r2x_bench.json 2> gpurun_out/r2x_bench.err; echo "bench rc=$?"
cut -c1-200 gpurun_out/r2x_bench.json
timeout 300 python bench.py --model qwen2.5-0.5b --batch 8 --ctx 512 > gpurun_out/r2x_c1_gpu.json 2> gpurun_out/r2x_c1_gpu.err; echo "c1 gpu rc=$?"
timeout 300 python bench.py --impl reference --model qwen2.5-0.5b --batch 8 --ctx 512 --steps 20 --warmup 3 > gpurun_out/r2x_c1_cpu.json 2> gpurun_out/r2x_c1_cpu.err; echo "c1 cpu rc=$?"
cut -c1-300 gpurun_out/r2x_c1_cpu.json
timeout 600 python bench.py --model mixtral-8x7b --ctx 2048 --no-cpu-baseline > gpurun_out/r2x_mixtral.json 2> gpurun_out/r2x_mixtral.err
timeout 600 python bench.py --quant nvfp4 --batch 32 --ctx 32768 --no-cpu-baseline > gpurun_out/r2x_nvfp4_32k.json 2> gpurun_out/r2x_nvfp4.err
timeout 900 python tools/bench_vs_wheel.py --out gpurun_out/r2x_vs_wheel.md > gpurun_out/r2x_vs_wheel.jsonl 2> gpurun_out/r2x_vs_wheel.err; echo "vs_wheel rc=$?"
timeout 600 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2x_launches.csv python bench.py --profile-step --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/r2x_ncu.log 2>&1; echo "ncu rc=$?"
python tools/summarize_launches.py gpurun_out/r2x_launches.csv "ncu launch list of ONE eager decode step (Llama-3.1-8B-shaped, FP8, batch 256, ctx 4096) with the folded split-KV reduction and the v5 GEMMs, 1 x B200" > gpurun_out/r2x_launch_summary.md 2>&1
gzip -f gpurun_out/r2x_launches.csv
NCU="ncu --set full --clock-control none --import-source on"
timeout 300 $NCU -k regex:paged_attn_decode -s 2 -c 1 -f -o gpurun_out/r2x_attn python tools/ncu_targets.py attn 256 4096 32 8 > gpurun_out/r2x_ncu_attn.log 2>&1
timeout 300 $NCU -k regex:gemm_fp8_v5 -s 2 -c 1 -f -o gpurun_out/r2x_gemm_v5_o python tools/ncu_targets.py fp8 256 4096 4096 v5 > gpurun_out/r2x_ncu_v5.log 2>&1
timeout 300 $NCU -k regex:gemm_fp8_v5 -s 2 -c 1 -f -o gpurun_out/r2x_gemm_v5_down python tools/ncu_targets.py fp8 256 4096 14336 v5 > gpurun_out/r2x_ncu_v5b.log 2>&1
timeout 300 $NCU -k regex:paged_attn_decode -s 2 -c 1 -f -o gpurun_out/r2x_attn_tp8 python tools/ncu_targets.py attn 256 4096 4 1 > gpurun_out/r2x_ncu_attn8.log 2>&1
ls -la gpurun_out | tail -12
