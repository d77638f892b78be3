#!/usr/bin/env bash
# TEST INFRASTRUCTURE ONLY.  Builds the UNMODIFIED reference CPU paged attention
# (/root/reference/csrc/cpu/cpu_attn.cpp + utils.cpp, compiled where they lie) into
# oracle/_ref/refcpu.so (git-ignored, travels to the GPU box with the snapshot).
# Recipe = SURVEY.md F6.  The reference's own build system (cmake + oneDNN FetchContent) is NOT run.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
REF="${B200_REFERENCE_ROOT:-/root/reference}"
OUT="$HERE/_ref"
if [ ! -d "$REF/csrc/cpu" ]; then
  echo "reference tree not present at $REF -- using prebuilt $OUT/refcpu.so if any"; exit 0
fi
mkdir -p "$OUT/gen"
if [ -f "$OUT/refcpu.so" ] && [ "$OUT/refcpu.so" -nt "$REF/csrc/cpu/cpu_attn.cpp" ] && \
   [ "$OUT/refcpu.so" -nt "$HERE/ref_cpu_attn_shim.cpp" ]; then
  echo "oracle/_ref/refcpu.so up to date"; exit 0
fi
PY="${PYTHON:-python}"
# the generator writes next to its own __file__ (read-only tree) -> call its function instead
"$PY" - "$REF/csrc/cpu/generate_cpu_attn_dispatch.py" "$OUT/gen/cpu_attn_dispatch_generated.h" <<'PYEOF'
import importlib.util, sys
spec = importlib.util.spec_from_file_location("gen", sys.argv[1])
m = importlib.util.module_from_spec(spec); spec.loader.exec_module(m)
open(sys.argv[2], "w").write(m.generate_header_file())
PYEOF
TORCH_INC=$("$PY" -c "import torch.utils.cpp_extension as c; print(' '.join('-I'+p for p in c.include_paths()))")
TORCH_LIB=$("$PY" -c "import torch,os; print(os.path.join(os.path.dirname(torch.__file__),'lib'))")
PY_INC=$("$PY" -c "import sysconfig; print('-I'+sysconfig.get_paths()['include'])")
ABI=$("$PY" -c "import torch; print(int(torch._C._GLIBCXX_USE_CXX11_ABI))")
ISA_FLAGS="-mavx512f -mavx512vl -mavx512bw -mavx512dq -mf16c -mfma"
DEFS="-DVLLM_NUMA_DISABLED"
if grep -q avx512_bf16 /proc/cpuinfo; then ISA_FLAGS="$ISA_FLAGS -mavx512bf16"; fi
if grep -q avx512_vnni /proc/cpuinfo; then ISA_FLAGS="$ISA_FLAGS -mavx512vnni"; fi
if grep -q amx_bf16 /proc/cpuinfo; then ISA_FLAGS="$ISA_FLAGS -mamx-bf16 -mamx-tile"; DEFS="$DEFS -DCPU_CAPABILITY_AMXBF16"; fi
CXX=/usr/bin/g++
COMMON="-O3 -std=c++17 -fPIC -fopenmp -D_GLIBCXX_USE_CXX11_ABI=$ABI $DEFS $ISA_FLAGS -I$OUT/gen -I$REF/csrc/cpu -I$REF/csrc $TORCH_INC $PY_INC -Wno-attributes -Wno-unused-variable"
echo "[ref-build] compiling cpu_attn.cpp (takes a few minutes)"
$CXX $COMMON -c "$REF/csrc/cpu/cpu_attn.cpp" -o "$OUT/cpu_attn.o" &
$CXX $COMMON -c "$REF/csrc/cpu/utils.cpp" -o "$OUT/utils.o" &
$CXX $COMMON -c "$HERE/ref_cpu_attn_shim.cpp" -o "$OUT/shim.o" &
wait
$CXX -shared -fopenmp -o "$OUT/refcpu.so" "$OUT/cpu_attn.o" "$OUT/utils.o" "$OUT/shim.o" \
  -L"$TORCH_LIB" -Wl,-rpath,"$TORCH_LIB" -ltorch -ltorch_cpu -lc10
rm -f "$OUT"/*.o
echo "[ref-build] wrote $OUT/refcpu.so"
