# 1 GPU: first run of the CTA-pair (UMMA M = 128) decode GEMM: parity, then timing against v2 / v4
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q -k "pair" > gpurun_out/r2t_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2t_tests.log
tail -15 gpurun_out/r2t_tests.log
S2="256,4096,4096;256,6144,4096;256,4096,14336;256,28672,4096;128,4096,4096;128,6144,4096;128,4096,14336;64,4096,4096;64,4096,14336;16,4096,4096;16,4096,14336;128,1280,8192;128,8192,1024;128,7168,8192;128,8192,3584;256,8192,3584"
for impl in v5 v5:32 v5:64 v5:128 v5:256 v2 v4; do
  for pdl in 0 2; do
    echo "## impl=$impl pdl=$pdl" >> gpurun_out/r2t_gemm.jsonl
    B200_PDL=$pdl timeout 200 python tools/bench_gemm.py --impl $impl --shapes "$S2" >> gpurun_out/r2t_gemm.jsonl 2>&1
  done
done
grep -v Traceback gpurun_out/r2t_gemm.jsonl | cut -c1-120 | head -150
