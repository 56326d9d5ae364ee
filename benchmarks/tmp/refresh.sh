set -x
python bench.py --workload c2 --steps 20 > gpurun_out/r01_c2.json 2>gpurun_out/r01_c2.err
python bench.py --workload c3 > gpurun_out/r01_c3.json 2>gpurun_out/r01_c3.err
python bench.py --workload c5 > gpurun_out/r01_c5.json 2>gpurun_out/r01_c5.err
python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/plain_default.json 2>gpurun_out/plain_default.err && \
ncu --metrics gpu__time_duration.sum --clock-control none -c 1500 --csv --log-file gpurun_out/launches_default.csv python bench.py --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_launches_default.log 2>&1
tail -c 300 gpurun_out/ncu_launches_default.log
wc -l gpurun_out/launches_default.csv
