#!/usr/bin/env bash
# Build libarbius_b200.so in-tree for sm_100a. Usage: csrc/build.sh [extra nvcc flags]
set -euo pipefail
here="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
out="$here/../arbius_b200/_native"
mkdir -p "$out/obj"
NVCC="${NVCC:-/usr/local/cuda/bin/nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC --expt-relaxed-constexpr "$@")
pids=()
objs=()
for src in "$here"/*.cu; do
  obj="$out/obj/$(basename "${src%.cu}").o"
  objs+=("$obj")
  # gemm_tc_bf16.cu is a three-line wrapper that #includes gemm_tc.cu: it is stale whenever gemm_tc.cu is newer
  dep="$src"
  [[ "$(basename "$src")" == "gemm_tc_bf16.cu" ]] && dep="$here/gemm_tc.cu"
  if [[ ! -f "$obj" || "$src" -nt "$obj" || "$dep" -nt "$obj" || -n "$(find "$here" -name "*.cuh" -newer "$obj")" || "$here/../include/arbius_b200.h" -nt "$obj" ]]; then
    "$NVCC" "${FLAGS[@]}" -Xptxas -v -c "$src" -o "$obj" 2> "$obj.log" &
    pids+=($!)
  fi
done
fail=0
for p in "${pids[@]:-}"; do [[ -z "$p" ]] || wait "$p" || fail=1; done
if [[ $fail -ne 0 ]]; then
  for l in "$out"/obj/*.log; do grep -E "error|Error" "$l" && echo "  ^ in $l"; done
  exit 1
fi
"$NVCC" -shared -o "$out/libarbius_b200.so" "${objs[@]}" -lcudart
echo "built $out/libarbius_b200.so"
