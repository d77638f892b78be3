set -x
mkdir -p gpurun_out
timeout 600 python tools/bench_vs_wheel.py --only gemm --out gpurun_out/r3d_vs_wheel_gemm.md > gpurun_out/r3d_vs_wheel.jsonl 2> gpurun_out/r3d_vs_wheel.err
grep "down M=\|tp8 qkv" gpurun_out/r3d_vs_wheel_gemm.md | cut -c1-200
