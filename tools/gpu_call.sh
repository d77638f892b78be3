#!/bin/bash
# One parametrised GPU-box runner (replaces the per-call gpu_call_r*.sh copies).
#   tools/gpu_call.sh <tag> <step> [<step> ...]
# steps: noalloc  = tests/test_zz_no_alloc.py alone in a fresh process
#        tests    = the whole -m gpu suite (-x)
#        bench    = python bench.py (default N=1 line)
#        launches = ncu launch list of a short bench run (after the plain run exited 0)
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
    launches) run bench_short timeout 600 python bench.py --steps 2 --warmup 3 && \
              run launches timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 2000 --csv \
                  --log-file "gpurun_out/${tag}_launches.csv" python bench.py --steps 2 --warmup 3 ;;
    file:*)   f=${step#file:}; run "$(basename "$f" .py)" timeout 900 python -m pytest "$f" -m gpu -x -q ;;
    *) echo "unknown step $step"; rc_all=64 ;;
  esac
done
exit $rc_all
