#!/bin/bash
# tools/build_variant.sh <name> <extra nvcc flags...> : builds arcanefem_b200/lib/libb200fem_<name>.so (tuning experiments).
# Recompiles the assembly kernels (assemble_*.cu) with the flags; B2F_VARIANT_FILES="sparsity convert" adds other units.
set -e
NAME=$1; shift
NVCC=/usr/local/cuda/bin/nvcc
ARCH="-gencode arch=compute_100a,code=sm_100a"
mkdir -p build/var_$NAME arcanefem_b200/lib
for f in arcanefem_b200/csrc/*.cu; do
  b=$(basename $f .cu)
  if [[ "$b" == assemble_* ]] || [[ " ${B2F_VARIANT_FILES:-} " == *" $b "* ]]; then
    $NVCC $ARCH -O3 -std=c++17 -lineinfo -ccbin /usr/bin/g++ -Xcompiler -fPIC -Iinclude -Iarcanefem_b200/csrc "$@" -c $f -o build/var_$NAME/$b.o &
  else
    [ -f build/$b.o ] || { echo "run make first"; exit 1; }
    cp build/$b.o build/var_$NAME/$b.o
  fi
done
wait
$NVCC $ARCH -shared -ccbin /usr/bin/g++ -o arcanefem_b200/lib/libb200fem_$NAME.so build/var_$NAME/*.o -ldl
echo built arcanefem_b200/lib/libb200fem_$NAME.so
