#!/bin/bash
# Builds libballista_b200.so for sm_100a (in-tree, next to the package).  Usage: build.sh [-j N]
set -e
cd "$(dirname "$0")"
OUT=../libballista_b200.so
OBJ=build
mkdir -p $OBJ
NVCC=${NVCC:-/usr/local/cuda/bin/nvcc}
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -Xcompiler -fPIC,-ffp-contract=off -diag-suppress 550"
SRCS="core.cu validity.cu tpch_gen.cu agg.cu agg_registry.cu eval.cu take.cu join.cu repartition.cu comm.cu shuffle.cu csv.cu ops_basic.cu ops_agg.cu ops_join.cu capi.cu expr.cpp proto.cpp"
pids=()
for f in $SRCS; do
  o=$OBJ/${f%.*}.o
  if [ ! -f $o ] || [ -n "$(find . -maxdepth 1 \( -name '*.cu' -o -name '*.cuh' -o -name '*.h' -o -name '*.cpp' \) -newer $o | head -1)" ] || [ ../../include/ballista_b200.h -nt $o ]; then
    ( $NVCC $FLAGS -x cu -c $f -o $o 2> $OBJ/${f%.*}.log || { cat $OBJ/${f%.*}.log; exit 1; } ) &
    pids+=($!)
  fi
done
for p in "${pids[@]}"; do wait $p; done
NCCL_LIB=$(python - <<'PY'
import os
try:
    import nvidia.nccl, os
    print(os.path.join(os.path.dirname(nvidia.nccl.__file__), "lib"))
except Exception:
    print("")
PY
)
OBJS=$(for f in $SRCS; do echo $OBJ/${f%.*}.o; done)
# link against libnccl.so.2 by soname: at run time the copy torch already loaded (or the system one) is used
$NVCC -shared -o $OUT $OBJS -L/usr/lib/x86_64-linux-gnu -l:libnccl.so.2 -lcudart
echo "built $(realpath $OUT)"
