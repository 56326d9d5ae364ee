#!/bin/bash
# Profile capture for one round (run on the GPU box through gpurun, one GPU):
#   bash profiles/capture.sh launches   -> gpurun_out/launches_default.csv (per-launch durations of the DEFAULT bench command)
#   bash profiles/capture.sh full       -> gpurun_out/prof.ncu-rep   (ncu --set full of the hot kernels, 200-row shard)
#   bash profiles/capture.sh lanes      -> gpurun_out/lanes_4_*.csv  (the two-lanes-per-pixel solve kernel, P=16)
# Each ncu pass only starts after the identical command has exited 0 without ncu.  The reports are turned into
# CSV pages on the box (raw metrics + per-instruction source page) and removed: gpurun brings back 64 MiB at most.
set -e
export_rep() {   # export_rep <report without extension>
  ncu -i $1.ncu-rep --page raw --csv > $1_raw.csv 2>/dev/null
  ncu -i $1.ncu-rep --page source --csv > $1_source.csv 2>/dev/null
  gzip -f $1_source.csv
  rm -f $1.ncu-rep
}
mkdir -p gpurun_out
if [ "$1" = "launches" ]; then
  CMD="python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-e2e"
  $CMD > gpurun_out/plain_default.json
  ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_default.csv $CMD > gpurun_out/ncu_launches_default.log 2>&1
elif [ "$1" = "lanes" ]; then
  export KAFKA_C5_PARAMS=16 KAFKA_C5_PIXELS=1e6
  for f in 4; do
    KAFKA_C5_FLAGS=$f python bench.py --workload c5 --no-cpu-baseline > gpurun_out/plain_lanes_$f.json
    KAFKA_C5_FLAGS=$f ncu --set full --clock-control none --import-source on -k regex:'solve_bands_pair_kernel' -s 12 -c 1 -f -o gpurun_out/lanes_$f \
      python bench.py --workload c5 --no-cpu-baseline > gpurun_out/ncu_lanes_$f.log 2>&1
    export_rep gpurun_out/lanes_$f
  done
else
  CMD="python bench.py --steps 2 --warmup 3 --rows 200 --no-cpu-baseline --no-e2e"
  $CMD > gpurun_out/plain.json
  ncu --set full --clock-control none --import-source on -k regex:'gp_linearize_kernel|solve_bands_kernel|finish_solve_kernel' -s 6 -c 6 -f -o gpurun_out/prof $CMD > gpurun_out/ncu_full.log 2>&1
  export_rep gpurun_out/prof
fi
