#!/bin/bash
# development aid: build parsip_b200/libpsgpu_<name>.so with extra nvcc flags for an A/B run (PSGPU_LIB=<path> selects it in api.py)
#   tools/build_variant.sh noz "-DPSGPU_NO_ZEROPAIR" cull.cu
set -e
name=$1; flags=$2; shift 2
cd "$(dirname "$0")/../parsip_b200/csrc"
mkdir -p build/var_$name
objs=""
for f in cull.cu field.cu emit.cu context.cu; do
  o=build/${f%.cu}.o
  for v in "$@"; do if [ "$v" = "$f" ]; then
    o=build/var_$name/${f%.cu}.o
    /usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -lineinfo -Xcompiler -fPIC --expt-relaxed-constexpr $flags -c $f -o $o
  fi; done
  objs="$objs $o"
done
/usr/local/cuda/bin/nvcc -gencode arch=compute_100a,code=sm_100a -shared -o ../libpsgpu_$name.so $objs build/linearize.o build/compat.o build/treeops.o build/sceneops.o
echo built parsip_b200/libpsgpu_$name.so
