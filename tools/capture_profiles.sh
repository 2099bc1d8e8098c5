#!/bin/bash
# Round-end measurement set, run ON THE GPU BOX through gpurun:  bash tools/capture_profiles.sh r01d
# Writes gpurun_out/<tag>_{bench,bench_reference,entry_points}.json, <tag>_launches.csv, <tag>_full.ncu-rep.
# Numbers printed under ncu are never bench values: the two bench runs come first, unprofiled.
set -u
TAG=${1:?tag}
mkdir -p gpurun_out
timeout 600 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
timeout 600 python bench.py --impl reference > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
timeout 600 python bench.py --entry-points --out gpurun_out/${TAG}_entry_points.json > /dev/null 2> gpurun_out/${TAG}_entry_points.err
ONE="python bench.py --batch 16 --chunk 16 --lanes 1 --steps 1 --warmup 1 --e2e-frames 16 --e2e-chunk 16 --no-cpu-baseline"
UWIE_HOST_ONE_LANE=1 timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $ONE > gpurun_out/${TAG}_launches.log 2>&1
UWIE_HOST_ONE_LANE=1 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:"gf_ab_kernel|gf_recover_kernel|sel_hist_kernel|sel_small_kernel|apply_kernel|clahe_lut_kernel|canny_tile_kernel<.int.1>|canny_rim_kernel<.bool.1>|canny_list_kernel|score_kernel|gf_prep" \
    -c 30 -o gpurun_out/${TAG}_full $ONE > gpurun_out/${TAG}_full.log 2>&1
tail -c 600 gpurun_out/${TAG}_bench.json
