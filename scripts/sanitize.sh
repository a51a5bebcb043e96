#!/bin/bash
# compute-sanitizer over every kernel family of libisg.  Run on the GPU box:
#   bash scripts/sanitize.sh            -> gpurun_out/sanitize_<tool>.log + gpurun_out/sanitize_summary.txt
# memcheck runs the whole workload; racecheck / synccheck run the forward + decode parts (the training part drives
# library kernels through autograd, which the tools instrument very slowly).
set -u
export ISG_FK_NO_TUNE=1
mkdir -p gpurun_out
SUM=gpurun_out/sanitize_summary.txt
: > $SUM
run() {  # tool, parts, extra flags
    local tool=$1 parts=$2; shift 2
    local log=gpurun_out/sanitize_${tool}.log
    timeout 1500 compute-sanitizer --tool $tool "$@" python scripts/sanitize_workload.py $parts > $log 2>&1
    echo "== $tool ($parts): exit $?" >> $SUM
    grep -E "ERROR SUMMARY|RACECHECK SUMMARY|sanitize workload done|Error:|hazard" $log | sort | uniq -c | head -20 >> $SUM
}
run memcheck fwd,decode,train
run racecheck fwd,decode
run synccheck fwd,decode
cat $SUM
