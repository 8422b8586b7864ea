#!/bin/bash
# Builds lib/libmpmk.so for sm_100a (nvcc cross-compiles without a GPU).
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
NCCL_INC=$(ls -d "$(python -c 'import sysconfig;print(sysconfig.get_paths()["purelib"])')"/nvidia/nccl/include 2>/dev/null || echo /usr/include)
FLAGS="-I$NCCL_INC -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
OUT=${OUT:-lib/libmpmk.so}
BUILD=${BUILD:-build}
mkdir -p "$BUILD" "$(dirname "$OUT")"
pids=()
for f in csrc/*.cu; do
  o=$BUILD/$(basename "${f%.cu}").o
  if [ ! -f "$o" ] || [ "$f" -nt "$o" ] || [ -n "$(find csrc ../include -name '*.cuh' -newer "$o" -o -name '*.h' -newer "$o")" ]; then
    $NVCC $FLAGS ${EXTRA_DEFS:-} ${PTXAS_V:+-Xptxas -v} -c "$f" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o "$OUT" "$BUILD"/*.o -ldl
echo "built $OUT"
