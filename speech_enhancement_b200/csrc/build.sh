#!/usr/bin/env bash
# Builds libse_b200.so in-tree for sm_100a.  Usage: build.sh [extra nvcc flags]
set -euo pipefail
cd "$(dirname "$0")"
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=hidden"
mkdir -p build
objs=()
pids=()
for f in *.cu; do
  o=build/${f%.cu}.o
  objs+=("$o")
  if [[ ! -f $o || $f -nt $o || -n $(find . -maxdepth 1 \( -name '*.cuh' -o -name 'build.sh' \) -newer "$o") || ../../include/se_b200.h -nt $o ]]; then
    $NVCC $FLAGS "$@" -c "$f" -o "$o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [[ -n $p ]] && wait "$p"; done
$NVCC -shared -gencode arch=compute_100a,code=sm_100a -o libse_b200.so "${objs[@]}"
echo "built $(pwd)/libse_b200.so"
