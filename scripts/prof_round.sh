# usage: bash scripts/prof_round.sh <step> <tag>
#   bench   : plain bench (tuned, with cpu baseline) -> gpurun_out/bench_<tag>.json
#   launches: ncu launch list of `bench.py --steps 2 --warmup 3` (ISG_FK_NO_TUNE=1: the plan's ~1600 autotune
#             launches would otherwise fill the list; same kernels, heuristic tiles) -> gpurun_out/launches_<tag>.csv
STEP=$1; TAG=${2:-r1d}
case $STEP in
bench)
  timeout 900 python bench.py > gpurun_out/bench_$TAG.json 2> gpurun_out/bench_$TAG.err; echo rc=$?
  timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref_$TAG.json 2>> gpurun_out/bench_$TAG.err; echo rc=$?
  B=64 timeout 100 python scripts/optime.py > gpurun_out/optime_$TAG.log 2>&1
  ISG_FK_VERBOSE=1 timeout 100 python scripts/fwd_time.py > gpurun_out/tune_$TAG.log 2>&1
  timeout 200 python scripts/prof_fk.py > gpurun_out/phase_$TAG.log 2>&1
  python scripts/dec_time.py > gpurun_out/dec_$TAG.log 2>&1
  ;;
launches)
  export ISG_FK_NO_TUNE=1
  timeout 300 python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_$TAG.json 2> gpurun_out/plain_$TAG.err && \
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 800 --csv --log-file gpurun_out/launches_$TAG.csv \
     python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$TAG.log 2>&1
  echo rc=$?
  ;;
esac
