#!/usr/bin/env bash
# Build the sm_100a kernel library (and optionally the standalone self-test) in-tree.
set -euo pipefail
ROOT="$(cd "$(dirname "${BASH_SOURCE[0]}")/.." && pwd)"
SRC="$ROOT/unet_zoo_b200/csrc"
OUT="$ROOT/unet_zoo_b200/libunet_zoo_b200.so"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC)
mkdir -p "$ROOT/build"
objs=()
for f in "$SRC"/*.cu; do
  o="$ROOT/build/$(basename "${f%.cu}").o"
  if [[ ! -f "$o" || "$f" -nt "$o" || "$SRC/ptx.cuh" -nt "$o" || "$SRC/igemm.cuh" -nt "$o" || "$ROOT/include/unet_zoo_b200.h" -nt "$o" ]]; then
    nvcc "${FLAGS[@]}" -c "$f" -o "$o" &
  fi
  objs+=("$o")
done
wait
nvcc -shared -o "$OUT" "${objs[@]}" -cudart static
if [[ "${1:-}" == "selftest" ]]; then
  nvcc "${FLAGS[@]}" -o "$ROOT/build/igemm_selftest" "$ROOT/tools/igemm_selftest.cu" "${objs[@]}" -cudart static
fi
echo "built $OUT"
