# mega (cluster-resident bottleneck runs) vs per-block launches: parity + time
ISG_FK_VERBOSE=1 timeout 120 python scripts/dbg_fk.py 2x512x512 > gpurun_out/dbg_mega.log 2>&1; echo dbg rc=$?
grep "logits\|^mega\|bottle1_5\|bottle2_8\|bottle3_8" gpurun_out/dbg_mega.log | cut -c1-110 | head
timeout 100 python scripts/fwd_time.py
ISG_FK_NO_MEGA8=1 timeout 100 python scripts/fwd_time.py
ISG_FK_NO_MEGA=1 timeout 100 python scripts/fwd_time.py
B=64 timeout 100 python scripts/optime.py > gpurun_out/optime_mega.log 2>&1
sed -n '/pass 2/,$p' gpurun_out/optime_mega.log | awk '{printf "%s ", $2} END{print ""}'
