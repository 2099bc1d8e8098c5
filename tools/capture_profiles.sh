#!/bin/bash
# Round measurement set, run ON THE GPU BOX through gpurun:  bash tools/capture_profiles.sh r02a [skip-bench]
# Writes gpurun_out/<tag>_{bench,bench_reference,entry_points}.json, <tag>_launches.csv, <tag>_full_{six,es,prims}.ncu-rep.
# Numbers printed under ncu are never bench values: the bench runs come first, unprofiled; every ncu command runs
# right after the same command line has exited 0 without ncu.
set -u
TAG=${1:?tag}
mkdir -p gpurun_out
if [ "${2:-}" != "skip-bench" ]; then
  timeout 900 python bench.py > gpurun_out/${TAG}_bench.json 2> gpurun_out/${TAG}_bench.err
  timeout 600 python bench.py --impl reference > gpurun_out/${TAG}_bench_reference.json 2> gpurun_out/${TAG}_bench_reference.err
  timeout 900 python bench.py --entry-points --out gpurun_out/${TAG}_entry_points.json > /dev/null 2> gpurun_out/${TAG}_entry_points.err
fi
export UWIE_HOST_ONE_LANE=1
ONE="python bench.py --batch 16 --chunk 16 --lanes 1 --steps 1 --warmup 1 --e2e-frames 16 --e2e-chunk 16 --no-cpu-baseline --no-p5 --probe-mb 0"
KERN="gf_|sel_hist_kernel|sel_small_kernel|apply_kernel|clahe_lut_kernel|canny_tile_kernel|canny_rim_kernel|canny_list_kernel|score_kernel|quad_|frame_hist|setup_"
timeout 600 $ONE > gpurun_out/${TAG}_plain_six.log 2>&1 && \
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv \
    --log-file gpurun_out/${TAG}_launches.csv $ONE > gpurun_out/${TAG}_launches.log 2>&1
# one engine call (the warm-up step's) of each library with the full set: per-kernel DRAM traffic, stalls, source
timeout 600 $ONE > gpurun_out/${TAG}_plain_six2.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:"$KERN" -c 75 -o gpurun_out/${TAG}_full_six $ONE > gpurun_out/${TAG}_full_six.log 2>&1
timeout 600 $ONE --library es > gpurun_out/${TAG}_plain_es.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:"$KERN" -c 75 -o gpurun_out/${TAG}_full_es $ONE --library es > gpurun_out/${TAG}_full_es.log 2>&1
timeout 300 python tools/prims_driver.py 2 > gpurun_out/${TAG}_plain_prims.log 2>&1 && \
timeout 600 ncu --set full --clock-control none --import-source on --kernel-name-base demangled \
    -k regex:"thead_kernel|clahe_l_" -s 2 -c 3 -o gpurun_out/${TAG}_full_prims python tools/prims_driver.py 2 > gpurun_out/${TAG}_full_prims.log 2>&1
# gpurun brings back at most 64 MiB: keep the raw-page CSVs (every metric of every captured launch) and drop the big reports
for X in six es; do
  if [ -f gpurun_out/${TAG}_full_${X}.ncu-rep ]; then
    ncu -i gpurun_out/${TAG}_full_${X}.ncu-rep --page raw --csv > gpurun_out/${TAG}_full_${X}_raw.csv 2>/dev/null
    rm -f gpurun_out/${TAG}_full_${X}.ncu-rep
  fi
done
tail -c 400 gpurun_out/${TAG}_bench.json 2>/dev/null
ls -la gpurun_out | tail -20
