"""Inference CLI with the reference's call surface (reference infer.py:12-36): flags
``-i/--test-image-dir``, ``-o/--output-dir``, ``--continue-test`` and ``path_decompose``.
The reference's loop body stops after ``path_decompose``; here it is completed: load -> uint8 image to
the device -> resize to the network size (libisg, the crop/resize kernel of the training data path,
train_instance.py:160-196, with the whole image as the window) -> ToTensor + Normalize(0.5, 0.5)
(train_instance.py:80-85) inside the first kernel -> Segment forward (libisg) -> mask decode back to
the source resolution (libisg) -> ``<name>.png``.
"""
import argparse
import glob
import os


def parse_args(argv=None):
    ap = argparse.ArgumentParser(description="inference image")
    ap.add_argument("-i", "--test-image-dir", required=True, help="image test dir")
    ap.add_argument("-o", "--output-dir", required=True, help="image save dir")
    ap.add_argument("--continue-test", action="store_true", help="skip existing file.")
    # additions (the reference never loads a model, so it has no such flags)
    ap.add_argument("--checkpoint", default=None, help="*.pth written by train_instance.py:497-507")
    ap.add_argument("--size", type=int, default=480, help="network input size (multiple of 16)")
    ap.add_argument("--precision", choices=("fp32", "bf16x3", "bf16"), default="fp32",
                    help="fp32: CUDA cores, <= 1e-4 of the reference; bf16x3: tensor cores, logits within 2e-2 / 2e-3; "
                         "bf16: tensor cores, fastest, outside that tolerance")
    ap.add_argument("--batch", type=int, default=16)
    ap.add_argument("--soft", action="store_true", help="also write the 0-255 soft mask")
    return ap.parse_args(argv)


def path_decompose(path):
    """-> (dirname, basename without extension, extension without dot)"""
    head, tail = os.path.split(path)
    stem, ext = os.path.splitext(tail)
    return head, stem, os.path.splitext(path)[-1][1:]


def list_images(directory):
    # same pattern as the reference (infer.py:35); note it is a character class, not a list
    return sorted(glob.glob(os.path.join(directory, "*[jpg,png,jpgerr]")))


def run(args):
    import cv2
    import numpy as np
    import torch

    from instancesegmentation_b200.decode import decode_masks
    from instancesegmentation_b200.preprocess import crop_resize
    from model.segment import Segment

    dev = torch.device("cuda", torch.cuda.current_device())
    model = Segment(3)
    if args.checkpoint:
        ck = torch.load(args.checkpoint, map_location="cpu")
        model.load_state_dict(ck["state_dict"] if "state_dict" in ck else ck)
    model.precision, model.max_batch = args.precision, args.batch
    model.to(dev).eval()
    os.makedirs(args.output_dir, exist_ok=True)
    todo = []
    for path in list_images(args.test_image_dir):
        _, stem, _ = path_decompose(path)
        dst = os.path.join(args.output_dir, stem + ".png")
        if not (args.continue_test and os.path.exists(dst)):
            todo.append((path, dst))
    S = args.size
    for i in range(0, len(todo), args.batch):
        chunk, boxes = todo[i:i + args.batch], []
        x = torch.empty((len(chunk), S, S, 3), dtype=torch.uint8, device=dev)     # network inputs, uint8 NHWC
        for n, (path, _) in enumerate(chunk):
            bgr = cv2.imread(path, cv2.IMREAD_COLOR)
            if bgr is None:
                raise IOError(f"cannot read {path}")
            h, w = bgr.shape[:2]
            boxes.append((0, 0, w, h))
            rgb = torch.from_numpy(np.ascontiguousarray(bgr[:, :, ::-1])).pin_memory().to(dev, non_blocking=True)
            crop_resize(rgb[None], np.array([[0, 0, w, h]], np.int32), np.zeros(1, np.int32), (S, S), out=x[n:n + 1])
        with torch.no_grad():
            res = decode_masks(model.forward_u8(x, (0.5, 0.5, 0.5), (0.5, 0.5, 0.5)), np.array(boxes, np.int32),
                               want_soft=args.soft)
        for n, (_, dst) in enumerate(chunk):
            cv2.imwrite(dst, res.mask(n).cpu().numpy() * 255)
            if args.soft:
                cv2.imwrite(dst[:-4] + "_soft.png", res.soft_mask(n).cpu().numpy())
    return len(todo)


if __name__ == "__main__":
    print(f"{run(parse_args())} images written")
