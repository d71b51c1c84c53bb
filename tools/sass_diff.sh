#!/bin/bash
# Are the device kernels of bp_b200.cu at two commits the same machine code?  (No GPU needed.)  Compiles the file at
# commit $1 (default: HEAD~1) and in the work tree for sm_100a and compares the SASS of each kernel family.
# Used at the end of round 1 to show that the host-side / experimental changes made after the last GPU run left
# carve_kernel, average_kernel and the read-back kernels bit-identical to the measured build.
set -e
OLD=${1:-HEAD~1}
T=$(mktemp -d)
git show $OLD:plant-3d-vision_b200/csrc/bp_b200.cu > $T/bp_b200.cu
git show $OLD:plant-3d-vision_b200/csrc/bp_internal.h > $T/bp_internal.h
git show $OLD:include/bp_b200.h > $T/bp_b200.h
F="-gencode arch=compute_100a,code=sm_100a -lineinfo -O3 -std=c++17"
(cd $T && nvcc $F -I$T -cubin -o old.cubin bp_b200.cu)
nvcc $F -Iinclude -cubin -o $T/new.cubin plant-3d-vision_b200/csrc/bp_b200.cu
for k in carve_kernel average_kernel expand_state tail_kernel pack_bits_kernel pack_bits_u8x16 selftest fill_f32 unpad; do
  for w in old new; do
    cuobjdump -sass $T/$w.cubin | awk -v k=$k '/Function :/{f=($0~k)} f' | grep -v "Function :" \
      | sed 's#/\* 0x[0-9a-f]* \*/##' | md5sum | cut -c1-12 > $T/$w.$k
  done
  if cmp -s $T/old.$k $T/new.$k; then echo "$k: identical"; else echo "$k: DIFFERENT"; fi
done
rm -rf $T
