#!/bin/bash
# Development aid: builds the library from a git revision's csrc into tools/_ab/<name>.so (for same-box A/B timing
# through B200KM_LIB=...).  usage: tools/ab_build.sh <rev> <name>
set -e
rev=${1:-HEAD}; name=${2:-base}
root=$(cd "$(dirname "$0")/.." && pwd)
tmp=$(mktemp -d)
mkdir -p "$root/tools/_ab" "$tmp/csrc" "$tmp/include"
git -C "$root" archive "$rev" spark_rapids_ml_nai_b200/csrc include | tar -x -C "$tmp"
src="$tmp/spark_rapids_ml_nai_b200/csrc"
objs=""
for f in "$src"/*.cu; do
  o="$tmp/$(basename "$f" .cu).o"
  nvcc -gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-O2 --expt-relaxed-constexpr ${AB_FLAGS} -c "$f" -o "$o" &
  objs="$objs $o"
done
wait
nvcc -shared -gencode arch=compute_100a,code=sm_100a -o "$root/tools/_ab/$name.so" $objs -ldl
rm -rf "$tmp"
echo "$root/tools/_ab/$name.so"
