#!/bin/bash
# SASS listings of the hot kernels of libbp_b200.so for profiles/ (no GPU needed): instruction text only (the hex
# encodings are stripped), one file per kernel, plus the loop summary of tools/sass_loops.py.
# Usage: tools/sass_dump.sh r02     -> profiles/r02_sass_<kernel>.txt, profiles/r02_sass_loops.txt
set -e
R=${1:-r02}
SO=plant-3d-vision_b200/libbp_b200.so
for k in carve_items_kernelILb0 carve_segments_kernelILb0 classify_items_kernel decode_items_kernel cull_blocks_kernel average_kernelIh carve_kernelILb0 pack_bits_u8x16_kernelILi0 tile_any_kernel tile_sat_kernel expand_state_kernel; do
  out=profiles/${R}_sass_$(echo $k | sed 's/ILb0//; s/ILi0//; s/Ih/_u8/').txt
  cuobjdump -sass $SO | awk -v k=$k '/Function :/{f=($0~k)} f' | sed 's#/\* 0x[0-9a-f]* \*/##; s/[ \t]*$//' | grep -v '^$' > $out
  echo "$out: $(grep -c ';' $out) instructions"
done
{ for k in carve_items_kernelILb0 carve_segments_kernelILb0 average_kernelIh cull_blocks_kernel; do python tools/sass_loops.py $k 30; done; } > profiles/${R}_sass_loops.txt
cat profiles/${R}_sass_loops.txt | cut -c1-220
