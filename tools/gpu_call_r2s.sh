# 1 GPU: v2 GEMM tuning knobs (ring depth, L2 promotion, cache policy), SM-fill probe, PDL on a launch-bound step
set -x
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_pdl.py tests/test_gpu_gemm.py -x -q 2>&1 | tail -3 > gpurun_out/r2s_tests.log; cat gpurun_out/r2s_tests.log
timeout 150 tools/l2fill_probe > gpurun_out/r2s_l2fill.jsonl 2>&1; head -40 gpurun_out/r2s_l2fill.jsonl
S2="256,4096,4096;256,6144,4096;256,4096,14336;128,4096,4096;64,4096,4096;128,8192,1024"
run() { echo "## $1" >> gpurun_out/r2s_gemm.jsonl; env $1 timeout 200 python tools/bench_gemm.py --impl v2 --shapes "$S2" >> gpurun_out/r2s_gemm.jsonl 2>&1; }
run "B200_PDL=0"
run "B200_PDL=2"
run "B200_PDL=0 B200_GEMM_STAGES=4"
run "B200_PDL=0 B200_GEMM_STAGES=6"
run "B200_PDL=0 B200_GEMM_L2PROMO=2"
run "B200_PDL=0 B200_GEMM_L2PROMO=0"
run "B200_PDL=0 B200_GEMM_POL=1"
run "B200_PDL=0 B200_GEMM_POL=1 B200_GEMM_L2PROMO=2"
run "B200_PDL=0 B200_GEMM_BN=128"
run "B200_PDL=0 B200_GEMM_BN=128 B200_GEMM_POL=1 B200_GEMM_L2PROMO=2"
run "B200_PDL=0 B200_GEMM_BN=256"
cat gpurun_out/r2s_gemm.jsonl | cut -c1-160
B200_PDL=0 TRACE_FINE=1 timeout 200 python tools/trace_gemm.py "256,4096,4096;256,4096,14336" 64 > gpurun_out/r2s_trace.txt 2>&1
grep -E "^==|graph-timed|   [a-z]" gpurun_out/r2s_trace.txt | grep -v "kb "
for pdl in 0 2; do
  B200_PDL=$pdl timeout 300 python bench.py --no-cpu-baseline --batch 32 --ctx 512 > gpurun_out/r2s_small_pdl$pdl.json 2> gpurun_out/r2s_small_pdl$pdl.err; cut -c1-180 gpurun_out/r2s_small_pdl$pdl.json
done
timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r2s_bench.json 2> gpurun_out/r2s_bench.err; cut -c1-200 gpurun_out/r2s_bench.json
