# whole-forward time (CUDA events around 20 forwards) with and without programmatic dependent launch
for v in "ISG_X=0" "ISG_FK_NO_PDL=1" "ISG_FK_CTAS=2 ISG_FK_LBUF=1" ; do
  echo "== $v"; env $v timeout 100 python scripts/fwd_time.py
done
timeout 150 python scripts/dbg_fk.py 2x96x160 2x512x512 1x16x16 3x208x48 2>&1 | grep logits | cut -c1-60
