# usage: bash scripts/ncu_fwd.sh <kernel-regex> <skip> <count> <tag>   (one tuned forward of batch 64; plan creation is outside the capture)
K=$1; SKIP=${2:-0}; CNT=${3:-1}; TAG=${4:-k}
timeout 120 python scripts/fwd_once.py > gpurun_out/plain_$TAG.log 2>&1 && \
timeout 400 ncu --set full --clock-control none --import-source on --profile-from-start off -k regex:$K -s $SKIP -c $CNT -f -o gpurun_out/prof_$TAG python scripts/fwd_once.py > gpurun_out/ncu_$TAG.log 2>&1
echo rc=$?
