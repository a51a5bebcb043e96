"""One pass over every kernel family of libisg at small sizes (target of scripts/sanitize.sh).
ISG_FK_NO_TUNE=1 keeps the launch count low: compute-sanitizer slows every kernel 10-100x."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment, synth
from instancesegmentation_b200.decode import decode_masks, mask_iou
from instancesegmentation_b200.preprocess import crop_resize
from instancesegmentation_b200.train_step import bce_sigmoid_loss

which = set((sys.argv[1] if len(sys.argv) > 1 else "fwd,decode,train").split(","))
z = np.load("tests/golden/segment3_state.npz")
sd = {k: torch.from_numpy(z[k]) for k in z.files}
dev = torch.device("cuda:0")


def model(prec, B):
    m = Segment(3); m.load_state_dict(sd); m.precision, m.max_batch = prec, B
    return m.to(dev).eval()


if "fwd" in which:
    # 512x512: the bottleneck runs go through the cluster-resident mega_kernel (bf16); bf16x3: split instantiations
    for prec, B, H, W in (("bf16", 2, 512, 512), ("bf16x3", 1, 512, 512), ("bf16x3", 2, 64, 96), ("fp32", 1, 64, 96)):
        x = torch.from_numpy(synth.batch(B, H, W, seed=3)).to(dev)
        with torch.no_grad():
            y = model(prec, B)(x)
        torch.cuda.synchronize()
        print("forward", prec, B, H, W, float(y.abs().mean()))
    u8 = torch.from_numpy(np.random.default_rng(0).integers(0, 256, (2, 64, 96, 3), dtype=np.uint8)).to(dev)
    with torch.no_grad():
        y = model("bf16x3", 2).forward_u8(u8)
    torch.cuda.synchronize()
    print("forward_u8", float(y.abs().mean()))
if "decode" in which:
    logits = torch.from_numpy(synth.logits(6, 96, 128, seed=5)).to(dev)
    r = decode_masks(logits)                                                     # identity-size kernel
    r2 = decode_masks(logits, want_soft=True)
    boxes, img_first = synth.boxes(2, 3, 160, 200, seed=7)
    boxes = np.minimum(boxes, 150).astype(np.int32); boxes[:, 2:] = np.maximum(boxes[:, 2:], 5)
    r3 = decode_masks(logits, boxes, img_first=img_first, canvas=(160, 200))     # tiled kernel + fused id map + finalize
    r4 = decode_masks(logits, boxes, want_soft=True, img_first=img_first, canvas=(160, 200))   # flat kernel + gather id map
    print("iou", mask_iou(r.masks, r2.masks), int(r3.area.sum()), int(r4.area.sum()))
    src = torch.from_numpy(np.random.default_rng(1).integers(0, 256, (2, 160, 200, 3), dtype=np.uint8)).to(dev)
    win = torch.from_numpy(boxes).to(dev)
    inst = torch.from_numpy(np.repeat(np.arange(2, dtype=np.int32), 3)).to(dev)
    c = crop_resize(src, win, inst, (64, 96))
    torch.cuda.synchronize()
    print("crop", tuple(c.shape))
if "train" in which:
    torch.manual_seed(0)
    m = Segment(3).to(dev).train()
    x = torch.from_numpy(synth.batch(2, 64, 64, seed=9)).to(dev)
    y = (torch.rand(2, 1, 64, 64, device=dev) < 0.3).float()
    loss = bce_sigmoid_loss(m(x), y)          # BN cluster kernels, depthwise stencil / wgrad, ConvT k8s4, fused BCE
    loss.backward()
    torch.cuda.synchronize()
    print("train loss", float(loss))
print("sanitize workload done")
