for v in "ISG_FK_NBUF=3" "ISG_FK_NBUF=2" "ISG_FK_NO_PIPE=1"; do
  env $v B=64 timeout 100 python scripts/optime.py > gpurun_out/optime_$v.log 2>&1
  echo "== $v"; sed -n '/pass 2/,$p' gpurun_out/optime_$v.log | awk '{printf "%s ", $2} END{print ""}'
done
