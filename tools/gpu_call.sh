#!/bin/bash
# One parametrised GPU-box runner (replaces the per-call gpu_call_r*.sh copies).
#   tools/gpu_call.sh <tag> <step> [<step> ...]
# steps: noalloc  = tests/test_zz_no_alloc.py alone in a fresh process
#        tests    = the whole -m gpu suite (-x)
#        bench    = python bench.py (default N=1 line)
#        launches = ncu launch list of a short bench run (after the plain run exited 0)
#        ncu_attn = ncu --set full of paged_attn_decode_kernel at the configs[1] shape and its TP2/4/8 shards
#        tp2      = (gpurun --gpus 2) the NVLink collectives' pytest at world 2 + the TP=2 bench line with tp_check
#        file:<path> = pytest -m gpu on one test file
# Every step writes gpurun_out/<tag>_<step>.log; the exit code is the first failing step's.
set -u
tag=$1; shift
mkdir -p gpurun_out
rc_all=0
run() {  # name, command...
  local name=$1; shift
  local t0=$(date +%s)
  "$@" > "gpurun_out/${tag}_${name}.log" 2>&1
  local rc=$?
  echo "[$tag] $name rc=$rc $(( $(date +%s) - t0 ))s: $(tail -n 1 "gpurun_out/${tag}_${name}.log")"
  [ $rc -ne 0 ] && [ $rc_all -eq 0 ] && rc_all=$rc
  return $rc
}
for step in "$@"; do
  case "$step" in
    noalloc)  run noalloc timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q ;;
    tests)    run tests timeout 1500 python -m pytest tests -m gpu -x -q ;;
    bench)    run bench timeout 600 python bench.py ;;
    launches) # launch list of ONE eager decode step (bench.py --profile-step opens the profiler range around it)
              run bench_short timeout 600 python bench.py --steps 2 --warmup 3 && \
              run launches timeout 900 ncu --profile-from-start off --metrics gpu__time_duration.sum --clock-control none \
                  --csv --log-file "gpurun_out/${tag}_launches.csv" python bench.py --profile-step --steps 1 --warmup 1 \
                  --no-cpu-baseline && \
              python tools/summarize_launches.py "gpurun_out/${tag}_launches.csv" "ncu launch list of one eager decode step (${tag})" \
                  > "gpurun_out/${tag}_launch_summary.md" ;;
    ncu_attn) # one --set full capture of the shipped decode kernel per TP shard of configs[1] (Hq/Hkv = 32/8, 16/4, 8/2, 4/1)
              for shard in "32 8" "16 4" "8 2" "4 1"; do
                n=attn_$(echo $shard | tr ' ' '_')
                run ncu_$n timeout 600 ncu --set full --clock-control none --import-source on -k regex:paged_attn_decode \
                    -s 2 -c 1 -f -o "gpurun_out/${tag}_${n}" python tools/ncu_targets.py attn 256 4096 $shard && \
                python tools/ncu_extract.py "gpurun_out/${tag}_${n}.ncu-rep" > "gpurun_out/${tag}_${n}_ncu.md" 2>&1
              done ;;
    tp2)      run allreduce timeout 900 python -m pytest tests/test_gpu_allreduce.py -m gpu -x -q -k "2-" ; \
              run bench_tp2 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 \
                  --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 ;;
    file:*)   f=${step#file:}; run "$(basename "$f" .py)" timeout 900 python -m pytest "$f" -m gpu -x -q ;;
    *) echo "unknown step $step"; rc_all=64 ;;
  esac
done
exit $rc_all
