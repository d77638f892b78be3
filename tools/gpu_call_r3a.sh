set -x
mkdir -p gpurun_out
for pdl in 0 1 2; do echo "## current library at PDL level $pdl" >> gpurun_out/r3a_ab_attn.jsonl; B200_PDL=$pdl timeout 300 python tools/ab_attn.py >> gpurun_out/r3a_ab_attn.jsonl 2>&1; done
cut -c1-260 gpurun_out/r3a_ab_attn.jsonl
timeout 900 python -m pytest tests/test_gpu_attention.py tests/test_gpu_pdl.py tests/test_gpu_backend.py tests/test_gpu_decoder.py -x -q 2>&1 | tail -3 > gpurun_out/r3a_tests.log; cat gpurun_out/r3a_tests.log
for pdl in 0 2; do B200_PDL=$pdl timeout 600 python bench.py --no-cpu-baseline > gpurun_out/r3a_bench_pdl$pdl.json 2> gpurun_out/r3a_bench_pdl$pdl.err; cut -c1-200 gpurun_out/r3a_bench_pdl$pdl.json; done
