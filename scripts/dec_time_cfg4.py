"""Decode time at cfg4 (640 instance maps of 1024x1024, boxes U(96,640), 32 canvases): tiled decode alone,
tiled decode + fused id map, and the separate gather id-map kernel; CUDA events, 2.7 GB of logits >> L2."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import synth, _lib
from instancesegmentation_b200.decode import DecodePlan, decode_masks
n_img, per, S = int(os.environ.get("NIMG", 32)), 20, 1024
boxes, first = synth.boxes(n_img, per, S, S, seed=7, lo=96, hi=640)
base = torch.from_numpy(synth.logits(per, S, S, seed=3)).cuda()
L = base.repeat(n_img, 1, 1).contiguous()
plan = DecodePlan(boxes, "cuda"); out = torch.empty(plan.total, dtype=torch.uint8, device="cuda")
nbytes = int(n_img * per * S * S * 4 + (boxes[:, 2].astype(np.int64) * boxes[:, 3]).sum() + 20 * len(boxes))
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
def timed(fn, reps=10):
    for _ in range(3): fn()
    torch.cuda.synchronize(); e0.record()
    for _ in range(reps): fn()
    e1.record(); torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps
ms = timed(lambda: decode_masks(L, out=out, plan=plan))
print(f"tiled decode            {ms*1e3:8.1f} us  {nbytes/ms/1e6:7.0f} GB/s algorithmic ({nbytes/1e6:.0f} MB)")
ms2 = timed(lambda: decode_masks(L, out=out, plan=plan, img_first=first, canvas=(S, S)))
idb = n_img * S * S * 4
print(f"tiled decode + id map   {ms2*1e3:8.1f} us  {(nbytes+idb)/ms2/1e6:7.0f} GB/s algorithmic (+{idb/1e6:.0f} MB of ids)")
l = _lib.lib(); idm = torch.empty((n_img, S, S), dtype=torch.int32, device="cuda")
fd = torch.from_numpy(first).cuda(); st = torch.cuda.current_stream().cuda_stream
ms3 = timed(lambda: _lib.check(l.isg_idmap(L.data_ptr(), 0, S, S, plan.boxes_d.data_ptr(), fd.data_ptr(), n_img, S, S, 0.5, idm.data_ptr(), st)), reps=5)
print(f"gather id-map kernel    {ms3*1e3:8.1f} us")
ms4 = timed(lambda: _lib.check(l.isg_decode(L.data_ptr(), 0, len(boxes), S, S, plan.boxes_d.data_ptr(), plan.off_d.data_ptr(), plan.max_w, plan.max_h, 0.5, out.data_ptr(), None, plan.area.data_ptr(), plan.bbox.data_ptr(), st)), reps=5)
print(f"previous decode kernel  {ms4*1e3:8.1f} us  {nbytes/ms4/1e6:7.0f} GB/s algorithmic")
r = decode_masks(L, out=out, plan=plan, img_first=first, canvas=(S, S))
print("fused id map equals gather kernel:", bool(torch.equal(r.idmap, idm)), " set fraction", float(out[:plan.total].float().mean()))
