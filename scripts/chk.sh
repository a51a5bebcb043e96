# quick regression: layerwise parity on four shapes + whole-forward time + per-op times
timeout 250 python scripts/dbg_fk.py 2x96x160 2x512x512 1x16x16 3x208x48 > gpurun_out/dbg.log 2>&1; echo dbg rc=$?
grep "logits\|border" gpurun_out/dbg.log | cut -c1-70
ISG_FK_VERBOSE=1 timeout 100 python scripts/fwd_time.py 2> gpurun_out/tune.log
ISG_FK_NO_TUNE=1 timeout 100 python scripts/fwd_time.py
B=64 timeout 100 python scripts/optime.py > gpurun_out/optime.log 2>&1
sed -n '/pass 2/,$p' gpurun_out/optime.log | awk '{printf "%s ", $2} END{print ""}'
