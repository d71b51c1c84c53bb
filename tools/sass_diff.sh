#!/bin/bash
# Are the back-projection device kernels at two commits the same machine code?  (No GPU needed.)  Compiles the kernel
# translation unit at commit $1 (default: HEAD~1) and in the work tree for sm_100a and compares the SASS of each kernel
# family.  Works across the round-2 split (bp_b200.cu -> bp_kernels.cu + bp_types.h + bp_geom.h).
set -e
OLD=${1:-HEAD~1}
T=$(mktemp -d)
C=plant-3d-vision_b200/csrc
for f in bp_kernels.cu bp_kernels.h bp_types.h bp_geom.h bp_internal.h bp_b200.cu; do
  git show $OLD:$C/$f > $T/$f 2>/dev/null || rm -f $T/$f
done
git show $OLD:include/bp_b200.h > $T/bp_b200.h
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17"
OLDSRC=bp_kernels.cu; [ -f $T/bp_kernels.cu ] || OLDSRC=bp_b200.cu
(cd $T && nvcc $F -I$T -cubin -o old.cubin $OLDSRC)
nvcc $F -Iinclude -I$C -cubin -o $T/new.cubin $C/bp_kernels.cu
for k in carve_kernel carve_items_kernel cull_blocks_kernel average_kernel expand_state tail_kernel pack_bits_kernel pack_bits_u8x16 selftest fill_f32 unpad; do
  for w in old new; do
    cuobjdump -sass $T/$w.cubin | awk -v k=$k '/Function :/{f=($0~k)} f' | grep -v "Function :" \
      | sed 's#/\* 0x[0-9a-f]* \*/##' | md5sum | cut -c1-12 > $T/$w.$k
  done
  if cmp -s $T/old.$k $T/new.$k; then echo "$k: identical"; else echo "$k: DIFFERENT"; fi
done
rm -rf $T
