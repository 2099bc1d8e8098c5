#!/bin/bash
# Source-level ncu captures of the heaviest kernels of one 16-frame engine call (one small report per kernel family):
#   bash tools/capture_top_kernels.sh <tag>
set -u
TAG=${1:-r02c}
export UWIE_HOST_ONE_LANE=1 UWIE_NO_GRAPH=1
ONE="python bench.py --batch 16 --chunk 16 --lanes 1 --steps 1 --warmup 1 --e2e-frames 16 --e2e-chunk 16 --no-cpu-baseline --no-p5 --probe-mb 0"
$ONE > gpurun_out/${TAG}_plain.log 2>&1 || { tail -20 gpurun_out/${TAG}_plain.log; exit 1; }
cap() {  # name, regex, count
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$2" -c $3 \
      -o gpurun_out/${TAG}_$1 $ONE > gpurun_out/${TAG}_$1.log 2>&1
}
cap gf "gf_ab_kernel|gf_recover_kernel|gf_prep" 3
cap apply "apply_kernel" 5
cap canny1 "canny_tile_kernel<.int.1>" 1
cap clahe "clahe_lut_kernel" 1
cap sel "sel_hist_kernel|sel_small" 2
ls -la gpurun_out/ | grep ncu-rep
du -sh gpurun_out
