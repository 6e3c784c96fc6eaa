#!/bin/bash
# Builds the current csrc tree into build/lib<name>.so (A/B measurements on one GPU box:
#   python tools/quick.py c2 THEIA_B200_LIB=/root/repo/build/libA.so,/root/repo/build/libB.so)
set -e
name=$1
root=$(cd "$(dirname "$0")/.." && pwd)
mkdir -p "$root/build/obj_$name"
cd "$root/theia_b200/csrc"
for f in api bvh_build; do
  nvcc -gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true \
       -ftz=false -Xcompiler -fPIC -Xcudafe --diag_suppress=177 $EXTRA -c $f.cu -o "$root/build/obj_$name/$f.o" &
done
wait
nvcc -shared -o "$root/build/lib$name.so" "$root/build/obj_$name/api.o" "$root/build/obj_$name/bvh_build.o" \
     -gencode arch=compute_100a,code=sm_100a -lcudart_static -Xcompiler -fPIC
echo "$root/build/lib$name.so"
