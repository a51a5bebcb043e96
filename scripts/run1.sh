set -x
python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/b2.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -s 120 -c 44 --csv --log-file gpurun_out/launches_fk.csv python bench.py --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu2.log 2>&1
echo rc=$?
