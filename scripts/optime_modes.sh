# per-op timing of the forward under several CTA / load-buffer modes (ISG_FK_CTAS, ISG_FK_LBUF)
for v in "ISG_FK_CTAS=1 ISG_FK_LBUF=2" "ISG_FK_CTAS=2 ISG_FK_LBUF=1" "ISG_FK_CTAS=2 ISG_FK_LBUF=2" "ISG_FK_CTAS=1 ISG_FK_LBUF=1"; do
  tag=$(echo $v | tr ' =' '__')
  env $v ISG_FK_VERBOSE=1 B=64 timeout 100 python scripts/optime.py > gpurun_out/optime_$tag.log 2>&1
  echo "== $v rc=$?"; sed -n '/pass 2/,$p' gpurun_out/optime_$tag.log | awk '{printf "%s ", $2} END{print ""}'
done
env ISG_FK_CTAS=2 ISG_FK_LBUF=1 timeout 150 python scripts/dbg_fk.py 2x96x160 2x512x512 1x16x16 3x208x48 > gpurun_out/dbg_c2.log 2>&1
grep logits gpurun_out/dbg_c2.log | cut -c1-60
