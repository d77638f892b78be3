# 1 GPU: PDL + in-kernel split-KV reduction + cluster-multicast v2 GEMM: parity, timelines, A/B bench lines
set -x
mkdir -p gpurun_out
timeout 1200 python -m pytest tests/test_gpu_pdl.py tests/test_gpu_gemm.py tests/test_gpu_attention.py tests/test_gpu_backend.py tests/test_gpu_decoder.py -x -q > gpurun_out/r2r_tests.log 2>&1; echo "tests rc=$?" >> gpurun_out/r2r_tests.log
tail -5 gpurun_out/r2r_tests.log
S="256,4096,4096;256,6144,4096;256,4096,14336;128,4096,4096"
for cl in 1 0; do
  B200_PDL=0 B200_GEMM_CLUSTER=$cl TRACE_FINE=1 timeout 300 python tools/trace_gemm.py "$S" 64 128 > gpurun_out/r2r_trace_cl$cl.txt 2>&1
done
grep -E "^==|graph-timed" gpurun_out/r2r_trace_cl*.txt
S2="256,6144,4096;256,4096,4096;256,4096,14336;128,6144,4096;128,4096,4096;128,4096,14336;128,1280,8192;128,8192,1024;128,7168,8192;128,8192,3584;256,1280,8192;256,8192,3584"
for cl in 1 0; do for bn in 0 32 64 128; do for pdl in 0 2; do
  echo "## cluster=$cl bn=$bn pdl=$pdl" >> gpurun_out/r2r_gemm.jsonl
  B200_PDL=$pdl B200_GEMM_CLUSTER=$cl B200_GEMM_BN=$bn timeout 200 python tools/bench_gemm.py --impl v2 --shapes "$S2" >> gpurun_out/r2r_gemm.jsonl 2>&1
done; done; done
for pdl in 0 1 2; do
  B200_PDL=$pdl timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r2r_bench_pdl$pdl.json 2> gpurun_out/r2r_bench_pdl$pdl.err; echo "bench pdl=$pdl rc=$?"
  cut -c1-200 gpurun_out/r2r_bench_pdl$pdl.json
done
B200_GEMM_CLUSTER=0 timeout 400 python bench.py --no-cpu-baseline > gpurun_out/r2r_bench_nocluster.json 2> gpurun_out/r2r_bench_nocluster.err
cut -c1-200 gpurun_out/r2r_bench_nocluster.json
