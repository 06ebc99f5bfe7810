#!/bin/bash
# usage: tools/build_variants.sh name1:"-DPDA_CHUNK_MAX=64" name2:"-DPDA_BLOCK=128 -DPDA_MIN_BLOCKS=7" ...
# builds variants/libgsmfret_<name>.so (git-ignored, travels with gpurun) and restores the default build
set -e
cd "$(dirname "$0")/.."
mkdir -p variants
for spec in "$@"; do
  name="${spec%%:*}"; flags="${spec#*:}"
  make -s -C gsmfretda_b200/csrc clean
  make -s -j8 -C gsmfretda_b200/csrc EXTRA="$flags" > /dev/null
  cp gsmfretda_b200/libgsmfret.so variants/libgsmfret_$name.so
  grep -A2 "pda_eval_kernelILi2ELb0ELi0E" gsmfretda_b200/csrc/pda_kernels.ptxas.log | grep -E "Used|spill" | sed "s/^/$name: /"
done
make -s -C gsmfretda_b200/csrc clean
make -s -j8 -C gsmfretda_b200/csrc > /dev/null
