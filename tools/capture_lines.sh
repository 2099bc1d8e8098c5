#!/bin/bash
# Per-source-line instruction / stall tables of single kernels of one 16-frame engine call, as text:
#   bash tools/capture_lines.sh <tag>     (run on the GPU box through gpurun; leaves gpurun_out/<tag>_lines_*.txt)
set -u
TAG=${1:-lines}
export UWIE_HOST_ONE_LANE=1 UWIE_NO_GRAPH=1
ONE="python bench.py --batch 16 --chunk 16 --lanes 1 --steps 1 --warmup 1 --e2e-frames 16 --e2e-chunk 16 --no-cpu-baseline --no-p5 --probe-mb 0"
$ONE > gpurun_out/${TAG}_plain.log 2>&1 || { tail -20 gpurun_out/${TAG}_plain.log; exit 1; }
cap() {  # name, kernel regex at capture time, plain name for the report (which holds that one launch)
  ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k regex:"$2" -c 1 \
      -o /tmp/${TAG}_$1 $ONE > gpurun_out/${TAG}_$1.log 2>&1
  python tools/ncu_lines.py /tmp/${TAG}_$1.ncu-rep "$3" 45 > gpurun_out/${TAG}_lines_$1.txt 2>&1
}
cap apply110 "apply_kernel<float, .bool.1, .bool.1, .bool.0>" apply_kernel
cap canny1 "canny_tile_kernel<.int.1>" canny_tile_kernel
[ "${2:-}" = "all" ] && cap lut "clahe_lut_kernel" clahe_lut_kernel
ls -la gpurun_out | grep ${TAG}
