#!/usr/bin/env bash
# Builds libvda_sm100a.so in-tree (next to the sources' package) for sm_100a only.
set -euo pipefail
HERE="$(cd "$(dirname "${BASH_SOURCE[0]}")" && pwd)"
OUT="$HERE/../libvda_sm100a.so"
OBJ="$HERE/build"
mkdir -p "$OBJ"
NVCC="${NVCC:-nvcc}"
FLAGS=(-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -Xcompiler -fPIC -Xcompiler -fvisibility=default)
# experiments: VDA_EXTRA_NVCC_FLAGS="-DFOO=1" VDA_OUT=/path/to/variant.so VDA_OBJ=/tmp/objdir build.sh
if [ -n "${VDA_EXTRA_NVCC_FLAGS:-}" ]; then read -r -a EXTRA <<< "$VDA_EXTRA_NVCC_FLAGS"; FLAGS+=("${EXTRA[@]}"); fi
OUT="${VDA_OUT:-$OUT}"
OBJ="${VDA_OBJ:-$OBJ}"
mkdir -p "$OBJ"
pids=()
for f in api gemm_simt gemm_tc gemm_tc2 conv_tc attn_simt attn_tc head_tail norms resize driver_kernels vis; do
  if [ ! -f "$OBJ/$f.o" ] || [ "$HERE/$f.cu" -nt "$OBJ/$f.o" ] || [ -n "$(find "$HERE" -maxdepth 1 \( -name '*.cuh' -o -name '*.h' \) -newer "$OBJ/$f.o" 2>/dev/null)" ] || [ "$HERE/../../include/vda.h" -nt "$OBJ/$f.o" ]; then
    "$NVCC" "${FLAGS[@]}" -c "$HERE/$f.cu" -o "$OBJ/$f.o" &
    pids+=($!)
  fi
done
for p in "${pids[@]:-}"; do [ -n "$p" ] && wait "$p"; done
"$NVCC" -arch=sm_100a -shared -o "$OUT" "$OBJ"/*.o -lcudart
echo "built $OUT"
