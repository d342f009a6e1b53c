#!/bin/bash
# SASS listings of EVERY kernel family of libb200splat.so (the instantiation the config-A / default path runs) from the
# built objects -> profiles/sass_r02/ (BASELINE.json north_star: "each kernel ... ships with a committed SASS listing").
#   bash profiles/dump_sass.sh [outdir]
set -e
cd "$(dirname "$0")/.."
C=3d-gaussian-splatting_b200/csrc
OUT=${1:-profiles/sass_r02}
mkdir -p $OUT
dump() {  # object, name filter (mangled substring, egrep), output
  : > $OUT/$3
  for f in $(cuobjdump -elf $C/$1 2>/dev/null | grep -o "\.text\.[A-Za-z0-9_]*" | sort -u | sed 's/^\.text\.//' | grep -E "$2"); do
    cuobjdump -sass -fun "$f" $C/$1 | grep -E "^\s+/\*[0-9a-f]{4}\*/|Function" | sed -E 's#/\* 0x[0-9a-f]+ \*/##; s/[[:space:]]+$//' >> $OUT/$3
  done
  echo "$3: $(grep -c Function $OUT/$3) function(s), $(wc -l < $OUT/$3) lines"
}
# stage 1 / 4b
dump project_sh.o 'project_sh_fwd_kernelILi3ELb0ELb1' project_sh_fwd_deg3.sass
dump project_sh.o 'project_sh_fwd_kernelILi3ELb1ELb1' project_sh_fwd_deg3_raw.sass
dump project_sh.o 'project_sh_bwd_kernelILi3ELi0' project_sh_bwd_deg3.sass
dump project_sh.o 'project_sh_bwd_kernelILi3ELi2' project_sh_bwd_deg3_raw_accumulate.sass
dump project_sh.o 'cam_pack_kernel' cam_pack.sass
# stage 2: corner grid, tile ranges, depth sort, two-level binning (default), comparators / fall-backs
dump binning.o 'tile_count_kernel|tile_prep_kernel' tile_count_prep.sass
dump radix_sort.o 'rs_hist_kernelIj|rs_scan_kernel' rs_hist_scan.sass
dump radix_sort.o 'rs_onesweep_kernelIjLi8' rs_onesweep_u32.sass
dump radix_sort.o 'rs_onesweep_kernelIjLi5' rs_onesweep_u32_5bit.sass
dump binning.o 'coarse_count_kernel|coarse_count_stream|emit_coarse_kernel|coarse_prep_kernel|expand_kernel' bin_coarse.sass
dump binning.o 'isect_ids_from_offsets|isect_ids_kernel' isect_ids.sass
dump binning.o 'bin_scan_kernel|bin_emit_kernel|bin_count_sorted|bin_emit_sorted|tile_ranges' binning_split_and_keys64.sass
dump radix_sort.o 'rs_onesweep_kernelItLi8' rs_onesweep_u16.sass
dump radix_sort.o 'rs_onesweep_kernelImLi8|rs_hist_kernelIm' rs_onesweep_u64.sass
# stage 3 / 4a
dump raster_fwd.o 'raster_fwd_kernelILi0ELi4' raster_fwd_rgb.sass
dump raster_bwd.o 'raster_bwd_kernelILi0ELi4' raster_bwd_rgb.sass
# stage 5 and the widened path
dump api.o 'densify_stats_kernel' densify_stats.sass
dump adam.o 'adam_kernel' adam.sass
dump loss.o 'l1_ssim' l1_ssim.sass
dump densify.o 'store_gather' store_gather.sass
# multi-GPU: the library's own collectives over NVSwitch multicast memory
dump collective.o 'multimem_allreduce_kernelILi4|multimem_adam_kernel|multimem_share_slice' collective_multimem.sass
