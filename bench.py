"""bench.py -- images/s of the hot path (Segment forward + fused mask decode) on B200.

  python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference]
  torchrun ... bench.py --gpus N ...          (one rank per GPU, batch-sharded, no collective)

A step = forward of one synthetic batch (BASELINE.json configs[1]: 64 x 3x512x512 per GPU)
followed by the mask decode of its logits.  Prints ONE JSON line (contract in the task brief).
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOP_PER_IMAGE_512 = 1.1876e9      # 2 x 593.8 M MACs of the reference graph
DECODE_BYTES_PER_PIXEL = 5         # fp32 logit in + 1 mask byte out (identity size) + 20 B stats/instance
HEADLINE_PRECISION = "bf16x3"      # the tensor-core mode inside the north_star tolerance; plain bf16 is an extra key


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=100)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=("ours", "reference"), default="ours")
    ap.add_argument("--precision", choices=("bf16x3", "bf16", "fp32"), default=HEADLINE_PRECISION)
    ap.add_argument("--batch", type=int, default=64, help="images per GPU per step")
    ap.add_argument("--size", type=int, default=512)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-extras", action="store_true",
                    help="skip the extra keys of the default line (library baseline on the GPU, 1024x1024, cfg4, other modes)")
    ap.add_argument("--workload", choices=("cfg2", "cfg4", "cfg5"), default="cfg2",
                    help="cfg2 = BASELINE configs[1] (the metric's configuration); cfg4 = configs[3]: 32 source "
                         "images at 1024x1024 with 20 instances each (640 network crops), resize decode + id map")
    return ap.parse_args()


def peaks():
    p = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(p):
        d = json.load(open(p))
        return dict(hbm=d["hbm_gbs"], tf_burst=d["bf16_tflops"], tf_sust=d["bf16_tflops_sustained"], src="measured")
    return dict(hbm=6650.0, tf_burst=1590.0, tf_sust=1400.0, src="fallback")


def bind_host(index):
    """Pin this rank to the CPUs NVML reports as local to its GPU (NUMA node), BEFORE pinned buffers are allocated
    (first touch places them on that node).  Returns a short description for the JSON line."""
    try:
        import pynvml as nv
        nv.nvmlInit()
        h = nv.nvmlDeviceGetHandleByIndex(index)
        ncpu = os.cpu_count() or 1
        words = nv.nvmlDeviceGetCpuAffinity(h, (ncpu + 63) // 64)
        cpus = {64 * w + b for w, m in enumerate(words) for b in range(64) if (m >> b) & 1}
        allowed = os.sched_getaffinity(0)
        cpus = (cpus & allowed) or allowed
        os.sched_setaffinity(0, cpus)
        return {"cpus_bound": len(cpus), "cpus_total": ncpu}
    except Exception as e:
        return {"cpus_bound": None, "error": type(e).__name__}


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU with NVML during the timed region.  NVML is initialised and
    the handle fetched in the constructor (main thread, before any barrier), the thread only polls, every 100 ms."""

    def __init__(self, index):
        super().__init__(daemon=True)
        self.stop_flag, self.mhz, self.reasons, self.max_mhz, self.nv, self.h = False, [], set(), None, None, None
        self.first = threading.Event()
        try:
            import pynvml as nv
            nv.nvmlInit()
            self.nv, self.h = nv, nv.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(self.h, nv.NVML_CLOCK_SM)
        except Exception as e:  # NVML missing: record that instead of guessing
            self.reasons.add(f"nvml_unavailable:{type(e).__name__}")

    def run(self):
        nv = self.nv
        if nv is None:
            self.first.set()
            return
        names = {nv.nvmlClocksThrottleReasonHwSlowdown: "hw_slowdown",
                 nv.nvmlClocksThrottleReasonHwThermalSlowdown: "hw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwThermalSlowdown: "sw_thermal_slowdown",
                 nv.nvmlClocksThrottleReasonSwPowerCap: "sw_power_cap"}
        try:
            while not self.stop_flag:
                self.mhz.append(nv.nvmlDeviceGetClockInfo(self.h, nv.NVML_CLOCK_SM))
                r = nv.nvmlDeviceGetCurrentClocksThrottleReasons(self.h)
                self.reasons |= {n for bit, n in names.items() if r & bit}
                self.first.set()
                time.sleep(0.1)
        except Exception as e:
            self.reasons.add(f"nvml_error:{type(e).__name__}")
            self.first.set()

    def begin(self):
        """start polling and return once the first sample is in (so the thread's start-up is outside the timed region)"""
        self.start()
        self.first.wait(5.0)

    def end(self):
        self.stop_flag = True
        self.join()

    def summary(self):
        m = sorted(self.mhz)
        return {"sm_mhz": m[len(m) // 2] if m else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons)}


REF_PATH = os.environ.get("REF_PATH", "/root/reference")


def workload_config(B, S, world, precision):
    """`config` of the JSON line; the reference arm prints the same dict (same workload, bounded sample)."""
    return {"workload": f"configs[1]: Segment(3) inference batch {B}/GPU at 3x{S}x{S} + fused mask decode "
                        f"(identity size, 0/1 masks+area+bbox)",
            "global_batch": world * B, "parallelism": f"batch-shard x{world}, no collective",
            "l2": "inputs rotate over 3 batches (uint8 NHWC, 150 MB) and the activation workspace (> 700 MB) exceeds the 126 MB L2"}


def cpu_reference_arm(B, size, steps, warmup, threads=None):
    """The reference's CPU implementation of the path on all host threads: the unmodified module imported from
    REF_PATH when that tree exists on this machine (kind "reference"), else its torch-CPU restatement
    oracle/segment_oracle.py (bit-equal to the module, tests/test_oracle.py; kind "port"), + the C decode restatement."""
    import numpy as np
    import torch

    from instancesegmentation_b200 import synth
    from oracle import decode_oracle
    from oracle.segment_oracle import segment_forward
    # under torchrun OMP_NUM_THREADS=1 is inherited: ask for every CPU this process may run on
    torch.set_num_threads(threads or len(os.sched_getaffinity(0)))
    z = np.load(os.path.join(ROOT, "tests", "golden", "segment3_state.npz"))
    sd = {k: torch.from_numpy(z[k]) for k in z.files}
    kind, fwd = "port", (lambda t: segment_forward(sd, t))
    ref_file = os.path.join(REF_PATH, "model", "segment.py")
    if os.path.exists(ref_file):
        try:
            import importlib.util
            spec = importlib.util.spec_from_file_location("ref_segment_for_bench", ref_file)
            mod = importlib.util.module_from_spec(spec)
            spec.loader.exec_module(mod)
            ref = mod.Segment(3)
            ref.load_state_dict(sd, strict=True)
            ref.eval()
            kind, fwd = "reference", (lambda t: ref(t))
        except Exception:
            pass
    x = torch.from_numpy(synth.batch(B, size, size, seed=1234))
    boxes = np.tile(np.array([0, 0, size, size], np.int32), (B, 1))
    decode_oracle.build()
    times = []
    with torch.no_grad():
        for i in range(warmup + steps):
            t0 = time.perf_counter()
            y = fwd(x)
            decode_oracle.decode(y[:, 0].numpy(), boxes)
            if i >= warmup:
                times.append(time.perf_counter() - t0)
    t = sorted(times)[len(times) // 2]
    return B / t, t, torch.get_num_threads(), kind


def decode_oracle_vs_torch():
    """Part of the cpu_baseline leg: the decode has no reference code to pin against (parity unpinned), so the line reports
    how often the C restatement's mask differs from torch's own sigmoid -> F.interpolate(bilinear, align_corners=False) ->
    > 0.5 on seeded logits (identity size, a x2 upsample and two non-integer ratios), per 1e7 pixels."""
    import numpy as np
    import torch
    import torch.nn.functional as F

    from instancesegmentation_b200 import synth
    from oracle import decode_oracle
    decode_oracle.build()
    L = synth.logits(4, 256, 256, seed=5)
    boxes = np.array([[0, 0, 256, 256], [0, 0, 512, 512], [3, 5, 367, 201], [1, 1, 97, 143]], np.int32)
    r = decode_oracle.decode(L, boxes)
    mism = total = 0
    for n, (_, _, wk, hk) in enumerate(boxes):
        p = F.interpolate(torch.sigmoid(torch.from_numpy(L[n:n + 1, None])), size=(int(hk), int(wk)), mode="bilinear",
                          align_corners=False)[0, 0].numpy()
        mine = r["masks"][r["off"][n]:r["off"][n + 1]].reshape(hk, wk)
        mism += int((mine != (p > 0.5)).sum())
        total += int(hk) * int(wk)
    return {"torch_mismatch_per_1e7": mism * 1e7 / total, "pixels": total, "mismatches": mism}


def run_reference(args, rank, world):
    """--impl reference: the reference's own CPU path, all host threads, bounded sample."""
    if rank != 0:
        return
    B = 16  # bounded sample of the 64-image workload per step (the CPU path is ~3 orders slower)
    steps = max(min(args.steps, 30), 1)
    ips, t, threads, kind = cpu_reference_arm(B, args.size, steps, min(args.warmup, 2))
    sample = (f"{B} of {args.batch} images per step x {steps} timed steps (median), fp32, torch CPU conv + scalar C decode, "
              f"{threads} threads")
    emit(({
        "impl": "reference", "metric": "images/sec", "value": ips, "unit": "images/s", "n_gpus": args.gpus,
        "steps": args.steps, "warmup": args.warmup, "ms_per_step": t * 1e3 * args.batch / B, "higher_is_better": True,
        "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
        "config": workload_config(args.batch, args.size, max(world, 1), "fp32"),
        "cpu_baseline": {"value": ips, "unit": "images/s", "cores": threads, "kind": kind, "sample": sample},
        "e2e": {"value": ips, "unit": "images/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }))


N_WINDOWS = 7      # the timed region is repeated: N_WINDOWS windows of exactly K steps, each bracketed by barrier + sync


def median(v):
    v = sorted(v)
    return v[len(v) // 2]


def gather_windows(local_ms, world, dev):
    """local_ms: this rank's window times.  Returns (ms of the median window = max over ranks, per-window maxima,
    per-rank times of that window).  One stalled window on one rank cannot move the median."""
    import torch
    import torch.distributed as dist
    t = torch.tensor(local_ms, dtype=torch.float64, device=dev)
    if world > 1:
        allt = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(allt, t)
        m = torch.stack(allt).cpu()           # [world, windows]
    else:
        m = t.cpu()[None]
    wmax = m.max(0).values.tolist()
    order = sorted(range(len(wmax)), key=lambda i: wmax[i])
    w = order[len(order) // 2]
    col = m[:, w].tolist()
    return wmax[w], wmax, col


def rank_stats(col):
    return {"min": min(col), "median": median(col), "max": max(col), "argmax_rank": col.index(max(col))}


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    host_info = bind_host(local_rank)     # before any pinned allocation
    from instancesegmentation_b200 import Segment, synth
    from instancesegmentation_b200.decode import DecodePlan, decode_masks
    from instancesegmentation_b200.pipeline import GraphedStep, Pipeline

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, S, K, W = args.batch, args.size, args.steps, max(args.warmup, 3)
    z = np.load(os.path.join(ROOT, "tests", "golden", "segment3_state.npz"))
    sd = {k: torch.from_numpy(z[k]) for k in z.files}

    def make_model(precision):
        m = Segment(3)
        m.load_state_dict(sd, strict=True)
        m.precision, m.max_batch = precision, B
        return m.to(dev).eval()

    model = make_model(args.precision)

    # inputs: rank r owns images [r*B, (r+1)*B) of the global batch (contiguous shard);
    # NBUF distinct batches rotate so that consecutive steps never re-read L2-resident inputs
    NBUF = 3
    base = synth.batch(8, S, S, seed=1234 + 8 * rank)
    host = [torch.from_numpy(np.concatenate([np.roll(base, i + 1 + j, axis=0) for i in range(B // 8 or 1)])[:B]
                             .copy()).pin_memory() for j in range(NBUF)]
    xs = [h.to(dev) for h in host]
    # the headline feeds what an image decoder produces and infer.py uploads: uint8 NHWC, resident in HBM (SURVEY 8d:
    # cfg2's input is a compact device tensor); ToTensor + Normalize run inside the first kernel
    rng = np.random.default_rng(99 + rank)
    host_u8 = [torch.from_numpy(rng.integers(0, 256, size=(B, S, S, 3), dtype=np.uint8)).pin_memory()
               for _ in range(NBUF)]
    xs_u8 = [h.to(dev) for h in host_u8]
    boxes = np.tile(np.array([0, 0, S, S], np.int32), (B, 1))
    masks_buf = torch.empty(B * S * S, dtype=torch.uint8, device=dev)
    dplan = DecodePlan(boxes, dev)      # box table / offsets uploaded once: a decode is one launch

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def windows(fn, n, K):
        ms = []
        for _ in range(n):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record()
            for i in range(K):
                fn(i)
            e1.record()
            barrier()
            ms.append(e0.elapsed_time(e1))
        return ms

    def fwd(model, x):
        return model.forward_u8(x) if x.dtype == torch.uint8 else model(x)

    def measure(model, K, decode_alone=True, xs=xs_u8):
        """device-resident inputs: (a) the step = forward + decode replayed from one CUDA graph per input buffer,
        (b) the forward alone (graph replays back to back), (c) the step enqueued launch by launch, (d) the decode alone"""
        steps = [GraphedStep(model, x, plan=dplan, masks=masks_buf) for x in xs]
        fwds = [GraphedStep(model, x, decode=False) for x in xs]
        out = {"eng": next(iter(model._engines.values()))}
        for i in range(W):
            steps[i % NBUF].replay()
        out["step"] = windows(lambda i: steps[i % NBUF].replay(), N_WINDOWS, K)
        out["fwd"] = windows(lambda i: fwds[i % NBUF].replay(), 3, K)

        def eager(i):
            with torch.no_grad():
                decode_masks(fwd(model, xs[i % NBUF]), out=masks_buf, plan=dplan)
        out["eager"] = windows(eager, 3, K)
        if decode_alone:
            # K decode launches back to back over logit batches that rotate beyond L2 (4 x 67 MB + 67 MB of masks)
            with torch.no_grad():
                ys = [fwd(model, x).clone() for x in xs]
            ys.append(ys[0].clone())
            for y in ys:
                decode_masks(y, out=masks_buf, plan=dplan)
            dms = []
            for _ in range(3):
                barrier()
                for q in range(4):              # a few ms of queued work so that the K launches below run gap-free
                    fwds[q % NBUF].replay()
                e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                e0.record()
                for i in range(K):
                    decode_masks(ys[i % len(ys)], out=masks_buf, plan=dplan)
                e1.record()
                barrier()
                dms.append(e0.elapsed_time(e1))
            out["dec"] = dms
            del ys
        return out

    sampler = ClockSampler(local_rank)
    sampler.begin()                       # first NVML sample is in before any barrier of the timed windows
    r = measure(model, K)
    sampler.end()
    eng = r["eng"]
    step_ms, step_windows, step_ranks = gather_windows(r["step"], world, dev)
    fwd_ms, _, _ = gather_windows(r["fwd"], world, dev)
    eager_ms, _, _ = gather_windows(r["eager"], world, dev)
    # the same step fed the Segment.forward tensor itself (fp32 NCHW, already normalised, resident in HBM)
    K32 = min(K, 20)
    steps32 = [GraphedStep(model, x, plan=dplan, masks=masks_buf) for x in xs]
    step32_ms, _, _ = gather_windows(windows(lambda i: steps32[i % NBUF].replay(), 3, K32), world, dev)
    del steps32
    dk_ms, _, _ = gather_windows(r["dec"], world, dev)

    # ---- end-to-end through the public API with host buffers ----
    # instancesegmentation_b200.pipeline.Pipeline: pinned host batch -> H2D -> forward + decode (one graph replay per
    # slot) -> D2H of masks + area + bbox, every step, on three streams (copies of neighbouring steps overlap compute).
    out_host = [torch.empty(B * S * S, dtype=torch.uint8).pin_memory() for _ in range(2)]
    area_host = [torch.empty(B, dtype=torch.int32).pin_memory() for _ in range(2)]
    bbox_host = [torch.empty((B, 4), dtype=torch.int32).pin_memory() for _ in range(2)]

    def e2e_run(batches, K2, packed=False):
        pipe = Pipeline(model, boxes, depth=2, graphs=True, packed=packed)
        for i in range(4):
            pipe.submit(batches[i % NBUF], out_host[i % 2], (area_host[i % 2], bbox_host[i % 2]))
        pipe.finish()

        def one(i):
            pipe.submit(batches[i % NBUF], out_host[i % 2], (area_host[i % 2], bbox_host[i % 2]))
            if i == K2 - 1:
                pipe.finish()
        return windows(one, 3, K2)

    e2e_u8_ms, _, _ = gather_windows(e2e_run(host_u8, K), world, dev)
    e2e_ms, _, _ = gather_windows(e2e_run(host, K32), world, dev)
    e2e_pk_ms, _, e2e_ranks = gather_windows(e2e_run(host_u8, K, packed=True), world, dev)

    # what the host link gives this rank while every rank copies at once: pinned H2D of one uint8 batch and D2H of one
    # mask batch, concurrently on two streams
    s1, s2 = torch.cuda.Stream(dev), torch.cuda.Stream(dev)
    dst_u8 = torch.empty_like(host_u8[0], device=dev)

    def copies(i):
        with torch.cuda.stream(s1):
            dst_u8.copy_(host_u8[i % NBUF], non_blocking=True)
        with torch.cuda.stream(s2):
            out_host[i % 2].copy_(masks_buf, non_blocking=True)
        if i == 9:
            torch.cuda.current_stream().wait_stream(s1)
            torch.cuda.current_stream().wait_stream(s2)
    pcie_ms, _, _ = gather_windows(windows(copies, 3, 10), world, dev)

    extras = {}
    if world == 1 and not args.no_extras:
        extras = run_extras(args, model, make_model, measure, xs, dev, B, S)

    if rank == 0:
        pk = peaks()
        imgs = world * B * K
        flop_img = FLOP_PER_IMAGE_512 * (S / 512) ** 2
        conv_tf = (B * K * flop_img) / (fwd_ms * 1e-3) / 1e12           # per GPU
        dec_bytes = B * (S * S * DECODE_BYTES_PER_PIXEL + 20)
        dec_gbs = dec_bytes * K / (dk_ms * 1e-3) / 1e9
        dec_in_step_ms = max(step_ms - fwd_ms, 0.0) / K
        # dram__bytes_read.sum + dram__bytes_write.sum from ncu at exactly this workload (bytes per forward of 64
        # images summed over its launches / per decode launch): profiles/r1i_forward_inst_per_launch.txt,
        # profiles/r1f_ncu_decode_kernel.txt.  Any other workload or mode has no capture -> null.
        cfg2 = (B == 64 and S == 512 and args.precision == "bf16")
        line = {
            "metric": "images/sec", "value": imgs / (step_ms * 1e-3), "unit": "images/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": step_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f32" if args.precision == "fp32" else "bf16", "data": "synthetic",
            "config": workload_config(B, S, world, args.precision), "mode": args.precision,
            "timing": {"windows": N_WINDOWS, "policy": "each window = exactly K steps (one CUDA-graph replay of forward + "
                       "decode per step) between barrier + synchronize; per window the MAX over ranks; value = the median window",
                       "window_ms": step_windows, "per_rank_ms": rank_stats(step_ranks),
                       "ms_per_step_launch_by_launch": eager_ms / K, "host": host_info,
                       "input": "uint8 NHWC resident in HBM (Normalize inside the first kernel)",
                       "images_per_s_fp32_nchw_input": world * B * K32 / (step32_ms * 1e-3)},
            "roofline": {"bound": "tensor", "achieved": conv_tf, "peak": pk["tf_sust"] / 1e0, "unit": "TFLOP/s",
                         "frac": conv_tf / pk["tf_sust"],
                         # dram__bytes_read.sum + dram__bytes_write.sum of one forward of 64 images summed over its launches:
                         # bf16x3 profiles/r2c_forward_bf16x3_per_launch.txt (generic first layer), bf16 profiles/r1i_forward_inst_per_launch.txt
                         "traffic": (2829.0e6 if args.precision == "bf16x3" else 1373.8e6 if cfg2 else None) if (B == 64 and S == 512) else None,
                         "what": f"conv stack ({eng.launches_per_forward} launches/forward, mode {args.precision}), algorithmic "
                                 f"{flop_img / 1e9:.4f} GFLOP/image, peak = sustained bf16 of {pk['src']}",
                         "ms_per_forward": fwd_ms / K},
            "roofline_decode": {"bound": "hbm", "achieved": dec_gbs, "peak": pk["hbm"], "unit": "GB/s",
                                "frac": dec_gbs / pk["hbm"], "traffic": 69.74e6 if (B == 64 and S == 512) else None, "bytes_per_launch": dec_bytes,
                                "ms_per_decode": dk_ms / K, "ms_per_decode_in_step": dec_in_step_ms,
                                "frac_in_step": (dec_bytes / (dec_in_step_ms * 1e-3) / 1e9 / pk["hbm"]) if dec_in_step_ms > 0 else None,
                                "what": "decode_kernel, average over K launches issued back to back on rotating logit "
                                        "batches; in_step = graph-replayed step minus graph-replayed forward"},
            # headline end-to-end = the infer.py call surface: host uint8 NHWC images (what cv2.imread gives,
            # infer.py:run) -> H2D -> Normalize + forward + decode -> D2H masks + area + bbox, every step
            "e2e": {"value": imgs / (e2e_pk_ms * 1e-3), "unit": "images/s",
                    "h2d_bytes_per_step": B * 3 * S * S, "d2h_bytes_per_step": (B * S * S + 7) // 8 + B * 20,
                    "per_rank_ms": rank_stats(e2e_ranks),
                    "what": "Pipeline(graphs=True, packed=True).submit per step (infer.py path): pinned uint8 NHWC images -> H2D -> "
                            "ToTensor+Normalize inside the first kernel -> forward + decode (graph replay) -> masks bit-packed on the "
                            "device (isg_pack_mask_bits, 1 bit per pixel, bit-exact) -> D2H masks+area+bbox; 3 streams"},
            "e2e_byte_masks": {"value": imgs / (e2e_u8_ms * 1e-3), "unit": "images/s", "h2d_bytes_per_step": B * 3 * S * S,
                               "d2h_bytes_per_step": B * S * S + B * 20,
                               "what": "the same pipeline returning one byte per mask pixel (8x the D2H bytes)"},
            "e2e_fp32_nchw": {"value": world * B * K32 / (e2e_ms * 1e-3), "unit": "images/s",
                              "h2d_bytes_per_step": B * 3 * S * S * 4, "d2h_bytes_per_step": B * S * S + B * 20,
                              "what": "same pipeline fed the Segment.forward tensor itself (pinned fp32 NCHW, already "
                                      "normalised on the host): 4x the PCIe bytes, PCIe-bound"},
            "host_link": {"h2d_gbs_per_gpu": 10 * B * 3 * S * S / (pcie_ms * 1e-3) / 1e9,
                          "d2h_gbs_per_gpu": 10 * B * S * S / (pcie_ms * 1e-3) / 1e9,
                          "what": "pinned cudaMemcpyAsync of one uint8 batch H2D and one mask batch D2H, concurrently, "
                                  "on every rank at once (slowest rank): the ceiling of e2e"},
            "gpu_launches": K * (eng.launches_per_forward + 1),   # forward launches + one decode launch per step (graph nodes)
            "clocks": sampler.summary(),
        }
        line.update(extras)
        if world == 1 and not args.no_cpu_baseline:
            ips, tt, threads, kind = cpu_reference_arm(16, S, 10, 2)
            line["cpu_baseline"] = {"value": ips, "unit": "images/s", "cores": threads, "kind": kind,
                                    "sample": f"16 of {B} images per step x 10 steps (median), fp32 torch-CPU forward "
                                              f"+ scalar C decode"}
            line["cpu_baseline"]["decode_oracle_vs_torch"] = decode_oracle_vs_torch()
        emit(line)
    if world > 1:
        dist.destroy_process_group()


def sub_bench(extra_args, timeout=900):
    """one more bench.py run in a fresh process on the same GPU (its own plan and workspace); returns its JSON line"""
    cmd = [sys.executable, os.path.abspath(__file__), "--no-extras", "--no-cpu-baseline"] + extra_args
    try:
        p = subprocess.run(cmd, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, timeout=timeout)
        lines = [l for l in p.stdout.decode().splitlines() if l.startswith("{")]
        return json.loads(lines[-1]) if lines else {"error": f"rc {p.returncode}, no JSON line"}
    except Exception as e:
        return {"error": f"{type(e).__name__}: {e}"}


def run_extras(args, model, make_model, measure, xs, dev, B, S):
    """Extra keys of the default line (1 GPU): the same graph through stock PyTorch/cuDNN on this GPU (the library bar of
    SURVEY 2.4), the other precision modes of the engine, and BASELINE's 1024x1024 / configs[3] workloads."""
    import torch
    pk = peaks()
    flop = B * FLOP_PER_IMAGE_512 * (S / 512) ** 2
    out = {}

    # ---- stock PyTorch / cuDNN: the drop-in module's torch route in eval mode is plain nn.Conv2d / nn.BatchNorm2d /
    # nn.ConvTranspose2d / PReLU (the libisg train-mode kernels are only taken when module.training) ----
    def lib_time(fn, n=5):
        with torch.no_grad():
            for _ in range(2):
                fn()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for i in range(n):
                fn()
            e1.record()
            torch.cuda.synchronize()
        return e0.elapsed_time(e1) / n

    lib = {}
    try:
        x = xs[0]
        old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
        torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = False
        lib["eager_fp32_tf32_off_ms"] = lib_time(lambda: model._forward_autograd(x))
        torch.backends.cudnn.allow_tf32 = torch.backends.cuda.matmul.allow_tf32 = True
        lib["eager_fp32_tf32_on_ms"] = lib_time(lambda: model._forward_autograd(x))
        torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old
        xcl = x.to(memory_format=torch.channels_last)
        mcl = make_model("fp32").to(memory_format=torch.channels_last)

        def bf16_cl():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                return mcl._forward_autograd(xcl)
        lib["eager_bf16_channels_last_ms"] = lib_time(bf16_cl)
        del mcl
        best = min(v for v in lib.values())
        lib.update({"images_per_s_best": B / (best * 1e-3), "roofline_frac_best": flop / (best * 1e-3) / 1e12 / pk["tf_sust"],
                    "what": f"forward only, batch {B} at {S}x{S}, stock PyTorch {torch.__version__} / cuDNN kernels of the same "
                            f"graph on this GPU (Segment._forward_autograd in eval mode), CUDA events, 5 forwards"})
    except Exception as e:
        lib["error"] = f"{type(e).__name__}: {e}"
    out["gpu_library_baseline"] = lib
    torch.cuda.empty_cache()

    # ---- the engine's other precision modes at the same workload ----
    modes = {}
    for prec in ("bf16x3", "bf16", "fp32"):
        if prec == args.precision:
            continue
        try:
            m = make_model(prec)
            k = 20 if prec != "fp32" else 5
            r = measure(m, k, decode_alone=False)
            step, fwd = median(r["step"]), median(r["fwd"])
            modes[prec] = {"images_per_s": B * k / (step * 1e-3), "ms_per_step": step / k, "ms_per_forward": fwd / k,
                           "roofline_frac": flop * k / (fwd * 1e-3) / 1e12 / pk["tf_sust"]}
            del m, r
        except Exception as e:
            modes[prec] = {"error": f"{type(e).__name__}: {e}"}
        torch.cuda.empty_cache()
    for prec, note in (("bf16", "plain bf16 operands: OUT of the north_star logits tolerance (max-abs 0.3-0.8)"),
                       ("bf16x3", "split-precision bf16 hi+lo operands, 3 MMAs: within 2e-2 / 2e-3"),
                       ("fp32", "CUDA-core fp32 kernels: within 1e-4")):
        if prec in modes:
            modes[prec]["note"] = note
    out["modes"] = modes

    # ---- BASELINE's other single-GPU workloads, each in a fresh process ----
    ex = {}
    if S == 512 and B == 64:
        j = sub_bench(["--batch", "32", "--size", "1024", "--steps", "20", "--precision", args.precision])
        ex["b32_1024"] = {k: j.get(k) for k in ("value", "ms_per_step", "error") if k in j}
        if "roofline" in j:
            ex["b32_1024"].update(roofline_frac=j["roofline"]["frac"], ms_per_forward=j["roofline"]["ms_per_forward"],
                                  decode_frac=j["roofline_decode"]["frac"], e2e=j["e2e"]["value"])
        j = sub_bench(["--workload", "cfg4", "--steps", "3", "--precision", args.precision])
        ex["cfg4"] = {k: j.get(k) for k in ("value", "ms_per_step", "network_images_per_s", "error") if k in j}
        if "roofline" in j:
            ex["cfg4"].update(roofline_frac=j["roofline"]["frac"], decode_frac=j["roofline_decode"]["frac"],
                              ms_per_decode=j["roofline_decode"]["ms_per_decode"],
                              ms_decode_plus_idmap=j["roofline_decode"]["ms_decode_plus_idmap"],
                              crop_resize_ms=j["crop_resize"]["ms"])
        j = sub_bench(["--workload", "cfg5", "--steps", "5"])
        ex["cfg5"] = {k: j.get(k) for k in ("value", "ms_per_step", "eager_ms_per_step", "exchange", "loss", "error") if k in j}
        ex["cfg5"]["what"] = "configs[4] on ONE GPU: train_instance.py step, 16 images at 512x512, bf16 autocast; graph replay"
    out["extra"] = ex
    return out


def run_cfg4(args, rank, world, local_rank):
    """BASELINE configs[3] / SURVEY 8(d) cfg4: ``n_img`` source images of 1024x1024 with 20 instances each.
    One network sample = one object crop (train_instance.py:141-196), so a step is the forward of
    n_img*20 crops at the network size (uint8 NHWC, groups of 32) followed by ONE tiled decode launch
    (isg_decode_tiled) that resizes every logit map to its instance box (U(96,640) px, seed 7),
    thresholds it, reduces area and bbox and ORs every set pixel's instance bit into its image's
    canvas, and one finalize launch that turns the bit sets into the id map."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from instancesegmentation_b200 import Segment, synth
    from instancesegmentation_b200.decode import DecodePlan, decode_masks

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    S, K, W = 1024, args.steps, max(args.warmup, 3)
    n_img, per_img, group = 32, 20, 32
    N = n_img * per_img
    z = np.load(os.path.join(ROOT, "tests", "golden", "segment3_state.npz"))
    model = Segment(3)
    model.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}, strict=True)
    model.precision, model.max_batch = args.precision, group
    model = model.to(dev).eval()
    from instancesegmentation_b200.preprocess import crop_resize
    base = torch.from_numpy(np.stack([synth.image_u8(S, S, 1234 + 8 * rank + i) for i in range(4)])).to(dev)
    sources = base[torch.arange(n_img, device=dev) % 4].contiguous()  # the 32 source images, uint8 HWC
    sources = torch.roll(sources, shifts=1, dims=2) if rank % 2 else sources
    boxes, img_first = synth.boxes(n_img, per_img, S, S, seed=7 + rank)
    # one network sample = one instance window of its source image resized to the network size
    # (train_instance.py:160-196); the decode pastes every mask back into the same window
    win_d = torch.from_numpy(boxes).to(dev)
    inst_img_d = torch.from_numpy(np.repeat(np.arange(n_img, dtype=np.int32), per_img)).to(dev)
    crops = torch.empty((N, S, S, 3), dtype=torch.uint8, device=dev)  # [640,1024,1024,3] uint8, 2.0 GB
    dplan = DecodePlan(boxes, dev)
    logits = torch.empty((N, 1, S, S), dtype=torch.float32, device=dev)  # 2.7 GB: every step rewrites > L2
    masks = torch.empty(dplan.total, dtype=torch.uint8, device=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(3 * (K + W) + 1)]
    sampler = ClockSampler(local_rank)
    barrier()
    for i in range(W + K):
        if i == W:
            sampler.begin()
            barrier()
            ev[0].record()
        with torch.no_grad():
            crop_resize(sources, win_d, inst_img_d, (S, S), out=crops)
            if i >= W:
                ev[3 * (i - W) + 1].record()
            model.forward_u8(crops, out=logits)
            if i >= W:
                ev[3 * (i - W) + 2].record()
            res = decode_masks(logits, out=masks, plan=dplan, img_first=img_first, canvas=(S, S))
        if i >= W:
            ev[3 * (i - W) + 3].record()
    barrier()
    sampler.end()
    total_ms = ev[0].elapsed_time(ev[3 * K])
    crop_ms = sum(ev[3 * i].elapsed_time(ev[3 * i + 1]) for i in range(K))
    fwd_ms = sum(ev[3 * i + 1].elapsed_time(ev[3 * i + 2]) for i in range(K))
    dec_ms = sum(ev[3 * i + 2].elapsed_time(ev[3 * i + 3]) for i in range(K))
    # decode alone (mask/area/bbox launch without the id-map launch), same buffers
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for i in range(K):
        decode_masks(logits, out=masks, plan=dplan)
    e1.record()
    barrier()
    dk_ms = e0.elapsed_time(e1)
    t = torch.tensor([total_ms, fwd_ms, dec_ms, dk_ms, crop_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, fwd_ms, dec_ms, dk_ms, crop_ms = t.tolist()
    if rank == 0:
        pk = peaks()
        eng = next(iter(model._engines.values()))
        flop_img = FLOP_PER_IMAGE_512 * (S / 512) ** 2
        conv_tf = N * K * flop_img / (fwd_ms * 1e-3) / 1e12
        out_px = int((boxes[:, 2].astype(np.int64) * boxes[:, 3]).sum())
        dec_bytes = N * (S * S * 4 + 20) + out_px                      # full logit map in + mask out
        dec_gbs = dec_bytes * K / (dk_ms * 1e-3) / 1e9
        # dram__bytes_read.sum + dram__bytes_write.sum of one tiled decode launch at exactly this geometry (boxes
        # seed 7): profiles/r1h_decode_cfg4_launches.txt -- below the algorithmic bytes because a downsampling
        # box never touches the source rows between its taps
        dec_traffic = 1981.2e6 if (rank == 0 and world == 1) else None
        emit({
            "metric": "images/sec", "value": world * n_img * K / (total_ms * 1e-3), "unit": "images/s",
            "n_gpus": world, "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16" if args.precision == "bf16" else "f32",
            "data": "synthetic",
            "config": {"workload": f"configs[3]: {n_img} source images/GPU at 3x{S}x{S}, {per_img} instances each = {N} "
                                   f"network crops ({args.precision}; crop/pad/resize of the U(96,640) instance windows to uint8 "
                                   f"NHWC on the device, forward in groups of {group}) + resize decode back into the "
                                   f"windows + id map",
                       "global_batch": world * n_img, "network_crops_per_step": world * N,
                       "parallelism": f"image-shard x{world}, no collective",
                       "l2": f"every step streams {crops.numel() / 1e9:.1f} GB of crops and {logits.numel() * 4 / 1e9:.1f} GB "
                             f"of logits, far beyond the 126 MB L2"},
            "network_images_per_s": world * N * K / (fwd_ms * 1e-3),
            "crop_resize": {"ms": crop_ms / K, "bytes_written": N * S * S * 3,
                            "gbs": N * S * S * 3 * K / (crop_ms * 1e-3) / 1e9,
                            "what": "isg_crop_resize_u8: 640 windows -> 640 x 1024x1024x3 uint8 network inputs"},
            "roofline": {"bound": "tensor", "achieved": conv_tf, "peak": pk["tf_sust"], "unit": "TFLOP/s",
                         "frac": conv_tf / pk["tf_sust"], "traffic": None,
                         "what": f"conv stack ({eng.launches_per_forward} launches per group of {group}), algorithmic "
                                 f"{flop_img / 1e9:.4f} GFLOP/crop, peak = sustained bf16 of {pk['src']}",
                         "ms_per_forward_group": fwd_ms / K / (N / group)},
            "roofline_decode": {"bound": "hbm", "achieved": dec_gbs, "peak": pk["hbm"], "unit": "GB/s",
                                "frac": dec_gbs / pk["hbm"], "traffic": dec_traffic, "bytes_per_launch": dec_bytes,
                                "ms_per_decode": dk_ms / K, "ms_decode_plus_idmap": dec_ms / K,
                                "what": f"{N} instance maps of {S}x{S} fp32 resized to their boxes ({out_px / N:.0f} px mean)"},
            "gpu_launches": K * (eng.launches_per_forward * (N // group) + 3),   # + crop, tiled decode, id-map finalize
            "clocks": sampler.summary(),
        })
    if world > 1:
        dist.destroy_process_group()


def run_cfg5(args, rank, world, local_rank):
    """BASELINE configs[4] / SURVEY 8(e): the train_instance.py step (:372-380), 16 images per GPU at 512x512,
    bf16 autocast.  Forward/backward run through torch autograd on the drop-in module's parameters (library
    kernels -- native backward kernels are SURVEY 8f-1, "next"); what this repo owns in the step is the exchange:
    gradients accumulate straight into ONE flat fp32 bucket (views), one NCCL all-reduce over NVLink, and the
    1/world scale + Adam update in one libisg launch (isg_adam_step).  The line reports the whole step and the
    exchange separately."""
    import numpy as np
    import torch
    import torch.distributed as dist

    from instancesegmentation_b200 import Segment, synth
    from instancesegmentation_b200.train_step import FlatBucket, GraphedTrainStep

    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    B, S, K, W = 16, 512, args.steps, max(args.warmup, 3)
    torch.manual_seed(0)
    model = Segment(3).to(dev).train()
    bucket = FlatBucket(model, grads_as_views=True)
    xs = [torch.from_numpy(synth.batch(B, S, S, seed=1234 + 16 * rank + j)).to(dev) for j in range(2)]
    ys = [(torch.from_numpy(synth.logits(B, S, S, seed=77 + rank + j)) > 0.8).float()[:, None].to(dev) for j in range(2)]
    crit = torch.nn.BCELoss()

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def eager_step(i):
        bucket.zero_grads()
        with torch.autocast("cuda", dtype=torch.bfloat16):
            p = model.train_batch(xs[i % 2])
        l = crit(p.float(), ys[i % 2])
        l.backward()
        return l.detach()

    # the reference's eager sequence (sigmoid -> BCELoss -> backward), timed first for the comparison field
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    for i in range(W + K):
        if i == W:
            barrier()
            e0.record()
        eager_step(i)
        bucket.adam_step(bucket.allreduce())
        model.invalidate()
    e1.record()
    barrier()
    eager_ms = e0.elapsed_time(e1) / K

    # the measured step: forward + isg_bce_sigmoid + backward replayed from one CUDA graph
    step = GraphedTrainStep(model, bucket, xs[0], ys[0])
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(2 * K + 1)]
    sampler = ClockSampler(local_rank)
    loss = None
    for i in range(W + K):
        if i == W:
            sampler.begin()
            barrier()
            ev[0].record()
        step.x.copy_(xs[i % 2], non_blocking=True)
        step.y.copy_(ys[i % 2], non_blocking=True)
        step.graph.replay()
        loss = step.loss
        if i >= W:
            ev[2 * (i - W) + 1].record()
        bucket.adam_step(bucket.allreduce())
        model.invalidate()
        if i >= W:
            ev[2 * (i - W) + 2].record()
    barrier()
    sampler.end()
    total_ms = ev[0].elapsed_time(ev[2 * K])
    ex_ms = sum(ev[2 * i + 1].elapsed_time(ev[2 * i + 2]) for i in range(K))
    t = torch.tensor([total_ms, ex_ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    total_ms, ex_ms = t.tolist()
    if rank == 0:
        emit({
            "metric": "images/sec", "value": world * B * K / (total_ms * 1e-3), "unit": "images/s", "n_gpus": world,
            "steps": K, "warmup": W, "ms_per_step": total_ms / K, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
            "config": {"workload": f"configs[4]: train_instance.py step, {B} images/GPU at 3x{S}x{S}, bf16 autocast forward/"
                                   f"backward by torch autograd (library kernels) + fused sigmoid/BCE loss and gradient "
                                   f"(isg_bce_sigmoid) replayed from one CUDA graph, flat gradient bucket -> NCCL "
                                   f"all-reduce -> fused scale + Adam in libisg",
                       "global_batch": world * B, "parallelism": f"data-parallel x{world}, one all-reduce of "
                                                                  f"{bucket.numel * 4} bytes per step"},
            "exchange": {"ms": ex_ms / K, "bucket_bytes": bucket.numel * 4,
                         "what": "all-reduce of the flat bucket + isg_adam_step (1/world scale fused)"},
            "eager_ms_per_step": eager_ms,
            "loss": float(loss.detach()), "gpu_launches": 2 * K, "clocks": sampler.summary(),
        })
    if world > 1:
        dist.destroy_process_group()


_RESULT_FD = None


def emit(obj):
    """The one JSON line goes to the process's ORIGINAL stdout; everything libraries print at fd 1
    (e.g. NCCL's version banner) was redirected to stderr in main()."""
    data = (json.dumps(obj) + "\n").encode()
    if _RESULT_FD is None:
        sys.stdout.write(data.decode()); sys.stdout.flush()
    else:
        os.write(_RESULT_FD, data)


def main():
    global _RESULT_FD
    args = parse()
    sys.stdout.flush()
    _RESULT_FD = os.dup(1)
    os.dup2(2, 1)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    if args.impl == "reference":
        return run_reference(args, rank, world)
    if world == 1 and args.gpus > 1:
        # not launched by torchrun: re-launch ourselves one rank per GPU
        import subprocess
        cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={args.gpus}",
               "--master-addr", "127.0.0.1", "--master-port", "29511", os.path.abspath(__file__)] + sys.argv[1:]
        sys.exit(subprocess.call(cmd, stdout=_RESULT_FD))
    if args.workload == "cfg4":
        return run_cfg4(args, rank, world, local_rank)
    if args.workload == "cfg5":
        return run_cfg5(args, rank, world, local_rank)
    run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
