set -x
mkdir -p gpurun_out
timeout 300 python tools/ab_attn.py > gpurun_out/r2z_ab_attn.jsonl 2>&1; cat gpurun_out/r2z_ab_attn.jsonl | cut -c1-300
S2="256,4096,14336;128,4096,14336;128,1280,8192;128,8192,1024;256,4096,4096"
for impl in v5 auto; do for pdl in 0 1 2; do
  echo "## impl=$impl pdl=$pdl" >> gpurun_out/r2z_gemm.jsonl
  if [ "$impl" = "auto" ]; then B200_PDL=$pdl timeout 200 python tools/bench_gemm.py --shapes "$S2" >> gpurun_out/r2z_gemm.jsonl 2>&1
  else B200_PDL=$pdl timeout 200 python tools/bench_gemm.py --impl $impl --shapes "$S2" >> gpurun_out/r2z_gemm.jsonl 2>&1; fi
done; done
cut -c1-110 gpurun_out/r2z_gemm.jsonl
