#!/bin/bash
# Profile capture for one round (run on the GPU box through gpurun, one GPU):
#   bash profiles/capture.sh launches   -> gpurun_out/launches.csv   (per-launch durations of the bench command)
#   bash profiles/capture.sh full       -> gpurun_out/prof.ncu-rep   (ncu --set full of the two hot kernels)
# Each ncu pass only starts after the identical command has exited 0 without ncu.
set -e
CMD="python bench.py --steps 2 --warmup 1 --rows 200 --no-cpu-baseline --no-e2e"
mkdir -p gpurun_out
$CMD > gpurun_out/plain.json
if [ "$1" = "launches" ]; then
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv $CMD > gpurun_out/ncu_launches.log 2>&1
else
  ncu --set full --clock-control none --import-source on -k regex:'gp_linearize_kernel|solve_bands_kernel' -s 2 -c 3 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
fi
