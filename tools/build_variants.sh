#!/bin/bash
# Scratch: link variants of libsph_b200.so that differ in -D flags of one translation unit.
# usage: tools/build_variants.sh <file.cu> name1:"-DX=1 -DY=2" name2:"..."   -> smoothed-particle-hydrodynamics_b200/build/variants/libsph_<name>.so
set -e
PKG=/root/repo/smoothed-particle-hydrodynamics_b200
FLAGS="-gencode arch=compute_100a,code=sm_100a -O3 -lineinfo -std=c++17 -fmad=false -prec-div=true -prec-sqrt=true -ftz=false -Xcompiler -fPIC,-ffp-contract=off,-fno-fast-math -I $PKG/csrc -I $PKG/../include"
SRC=$1; shift
BASE=$(basename $SRC .cu)
mkdir -p $PKG/build/variants
for spec in "$@"; do
  name=${spec%%:*}; defs=${spec#*:}
  nvcc $FLAGS $defs -c $PKG/csrc/$SRC -o $PKG/build/variants/${BASE}_$name.o
  objs=""
  for f in sph_sort sph_neighbors sph_pcisph sph_membranes sph_slab sph_capi; do
    if [ $f == $BASE ]; then objs="$objs $PKG/build/variants/${BASE}_$name.o"; else objs="$objs $PKG/build/$f.o"; fi
  done
  nvcc -shared -o $PKG/build/variants/libsph_$name.so $objs -gencode arch=compute_100a,code=sm_100a -lcudart_static -ldl -lrt -lpthread
  echo built $name
done
