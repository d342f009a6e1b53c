#!/bin/bash
# SASS listings of the hot-path kernels (main instantiations) from the built objects -> profiles/sass_r01/
set -e
cd "$(dirname "$0")/.."
C=3d-gaussian-splatting_b200/csrc
OUT=profiles/sass_r01
mkdir -p $OUT
dump() {  # object, name filter (mangled substring), output
  for f in $(cuobjdump -elf $C/$1 2>/dev/null | grep -o "\.text\.[A-Za-z0-9_]*" | sort -u | sed 's/^\.text\.//' | grep -E "$2"); do
    cuobjdump -sass -fun "$f" $C/$1 | grep -E "^\s+/\*[0-9a-f]{4}\*/|Function" | sed -E 's#/\* 0x[0-9a-f]+ \*/##; s/[[:space:]]+$//' 
  done > $OUT/$3
  echo "$3: $(wc -l < $OUT/$3) lines"
}
dump raster_fwd.o 'raster_fwd_kernelILi0ELi4' raster_fwd_rgb.sass
dump raster_bwd.o 'raster_bwd_kernelILi0ELi4' raster_bwd_rgb.sass
dump project_sh.o 'project_sh_fwd_kernelILi3' project_sh_fwd_deg3.sass
dump project_sh.o 'project_sh_bwd_kernelILi3' project_sh_bwd_deg3.sass
dump radix_sort.o 'rs_onesweep_kernelItLi8' rs_onesweep_u16.sass
dump radix_sort.o 'rs_onesweep_kernelIjLi8' rs_onesweep_u32.sass
dump binning.o 'bin_emit_sorted|tile_prep' binning_emit_prep.sass
