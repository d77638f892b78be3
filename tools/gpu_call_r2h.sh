set -x
mkdir -p gpurun_out
S="256,6144,4096;256,4096,4096;256,4096,14336;256,1280,8192;64,4096,4096;64,4096,14336;16,4096,4096;16,4096,14336;256,28672,4096;64,28672,4096"
for impl in v2 v3 v4 v4:2; do
  timeout 300 python tools/bench_gemm.py --impl $impl --shapes "$S" >> gpurun_out/r2h_gemm_impls.jsonl 2>&1
done
cat gpurun_out/r2h_gemm_impls.jsonl
