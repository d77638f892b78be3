set -x
mkdir -p gpurun_out
timeout 200 tools/l2fill_probe > gpurun_out/r2v_l2fill.jsonl 2>&1; grep -i "row-major\|pure HBM" gpurun_out/r2v_l2fill.jsonl | cut -c1-330
timeout 300 python -m pytest tests/test_gpu_gemm.py -x -q -k "pair" 2>&1 | tail -3
