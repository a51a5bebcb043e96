# usage: bash scripts/ncu_k.sh <kernel-regex> <skip> <count> <tag>
set -x
K=$1; SKIP=${2:-0}; CNT=${3:-1}; TAG=${4:-k}
python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:$K -s $SKIP -c $CNT -f -o gpurun_out/prof_$TAG python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_$TAG.log 2>&1
echo rc=$?
