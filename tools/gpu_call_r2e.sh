set -x
mkdir -p gpurun_out
for nacc in 0 1; do
B200_GEMM_NACC=$nacc timeout 600 python tools/trace_gemm.py "256,4096,4096;256,6144,4096;256,4096,14336;64,4096,4096;16,4096,14336" auto 64 128 > gpurun_out/r2e_trace_nacc$nacc.txt 2>&1
done
grep -E "^==|graph-timed" gpurun_out/r2e_trace_nacc*.txt
timeout 600 python -m pytest tests/test_gpu_gemm.py -x -q 2>&1 | tail -3
