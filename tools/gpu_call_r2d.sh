# 1 GPU: v2 FP8 GEMM timelines at the decode shapes for several batch tiles / ring depths
set -x
mkdir -p gpurun_out
timeout 600 python tools/trace_gemm.py "256,4096,4096;256,6144,4096;256,4096,14336" auto 64 128 256 128:4 256:4 > gpurun_out/r2d_trace.txt 2>&1
tail -5 gpurun_out/r2d_trace.txt
