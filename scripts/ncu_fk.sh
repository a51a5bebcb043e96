# usage: bash scripts/ncu_fk.sh <skip> <count> <tag>   -- ncu --set full of fk_kernel launches in one forward at B=64
set -x
SKIP=${1:-10}; CNT=${2:-1}; TAG=${3:-fk}
python scripts/fwd_once.py > gpurun_out/plain_$TAG.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:fk_kernel -s $SKIP -c $CNT -f -o gpurun_out/prof_$TAG python scripts/fwd_once.py > gpurun_out/ncu_$TAG.log 2>&1
echo rc=$?
