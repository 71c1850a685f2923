#!/usr/bin/env python
"""bench.py — images/sec of the detection hot path (decode + target assignment + loss fwd/bwd +
NMS) at YOLOv4-608 on N B200s, with the HBM roofline of the dominant kernel, the end-to-end
number through host buffers and the reference's CPU path timed beside it.

  python bench.py --gpus N --steps K --warmup W            (N>1: launched by torchrun)
  python bench.py --impl reference --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1], "YOLOv4 training loss ... batch 64 at 608x608"): per GPU one
batch of 64 synthetic YOLOv4-608 head tensors (3 maps, 255 channels, f32; 495 MB — larger than
the 126 MB L2, so every step streams from HBM) and 1..30 GT boxes per image; CIoU loss with the
reference's default weights, Rect4 matching, NMS at the shipped thresholds (conf .4 / IoU .5,
train.json5:153-154) per class.  `value` is WEAK scaling (every rank owns its own 64 images; the
only exchange is the 5 weighted loss scalars); the `strong` record beside it is the literal
configs[1] — ONE global batch of 64 split over the N GPUs (64/N images each).
One step = yh_decode_merged + yh_assign + yh_loss_dense + yh_loss_rows_compute + yh_loss_finish (full
head gradient) + yh_nms = 8 kernels, captured once as a CUDA graph over four forked/joined streams
(yolo_dl_b200/pipeline.py, schedule "overlapped"; the "grouped" / "layered" schedules that cut the loss
into L2-sized pieces were measured and are slower) and replayed per step.  The timed
region is `--steps` replays between barrier + synchronize; it is repeated `--windows` times back to
back and `value` is the median window (all windows are in the line).
"""
import argparse
import ctypes
import json
import os
import statistics
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

METRIC = "images/sec decode+loss+NMS @YOLOv4-608"
UNIT = "images/s"
SEED = 20261104 + 1
NMS_CONF, NMS_IOU, BY_CLASS = 0.4, 0.5, True
GLOBAL_BATCH = 64   # BASELINE configs[1]


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=1000)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--windows", type=int, default=5, help="timed windows of --steps replays each")
    p.add_argument("--impl", default="native", choices=["native", "reference"])
    p.add_argument("--batch", type=int, default=64, help="images per GPU (weak-scaling arm)")
    p.add_argument("--cpu-sample", type=int, default=64, help="images per CPU-baseline pass")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-extras", action="store_true", help="skip the side numbers of BASELINE configs 0, 2-4")
    p.add_argument("--no-strong", action="store_true", help="skip the strong-scaling record")
    p.add_argument("--loss-exchange", default="peer", choices=["peer", "nccl"],
                   help="N>1: in-graph peer-memory stores over NVLink (default) or a NCCL all-reduce per step")
    return p.parse_args()


def workload_config(args, world):
    return {"workload": "YOLOv4-608 head decode + target assignment + CIoU/BCE loss fwd+bwd + per-class NMS",
            "spec": "yolov4-608 (76x76,38x38,19x19; A=3; C=80; N=22743)",
            "batch_per_gpu": args.batch, "global_batch": args.batch * world,
            "gt_per_image": "U{1..30}", "box_metric": "ciou", "nms": f"conf {NMS_CONF} iou {NMS_IOU} per-class",
            "parallelism": f"images sharded over {world} GPU(s); only exchange: 5 weighted loss scalars per step",
            "l2": "inputs (495 MB/GPU/step) larger than the 126 MB L2; no flush needed"}


def git_commit():
    try:
        return subprocess.run(["git", "-C", ROOT, "rev-parse", "--short", "HEAD"], capture_output=True,
                              text=True, timeout=10).stdout.strip() or None
    except Exception:
        return None


# ----------------------------------------------------------------------------------
# clocks (sampled DURING the timed region)
# ----------------------------------------------------------------------------------
class ClockSampler:
    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake", 0x2: "applications_clocks_setting"}

    def __init__(self, index):
        self.samples, self.reasons, self.max_mhz = [], set(), None
        self._stop = threading.Event()
        self._t = None
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def _loop(self):
        while not self._stop.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                fn = getattr(self.nv, 'nvmlDeviceGetCurrentClocksEventReasons', None) or \
                    self.nv.nvmlDeviceGetCurrentClocksThrottleReasons
                r = fn(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop.wait(0.002)

    def __enter__(self):
        if self.nv:
            self._t = threading.Thread(target=self._loop, daemon=True)
            self._t.start()
        return self

    def __exit__(self, *a):
        self._stop.set()
        if self._t:
            self._t.join()

    def summary(self):
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz,
                "reasons": sorted(self.reasons), "samples": len(s)}


# ----------------------------------------------------------------------------------
# reference arm / CPU baseline: the oracle restatement of the reference's tch-on-CPU path
# ----------------------------------------------------------------------------------
def cpu_reference_step(heads, infos, C, targets):
    """Detect2D x L -> merge -> CyCxHWMatcher (C, single thread like the reference) -> YoloLoss
    + autograd backward -> NonMaxSuppression (torchvision's C++ batched_nms as tch-nms)."""
    import torch
    from oracle import detect as odet, loss as oloss, nms as onms
    r = oloss.yolo_loss(heads, infos, C, targets, oloss.LossConfig(box_metric="ciou"), use_c_matcher=True)
    with torch.no_grad():
        merged = odet.forward_merged(heads, infos, C)
        k = onms.nms_forward(merged, NMS_CONF, NMS_IOU, BY_CLASS, use_torchvision=True)
    return float(r["losses"]["total"]), len(k["batches"])


def cpu_reference_inputs(sample, device="cpu"):
    from oracle import detect as odet
    from yolo_dl_b200 import synth
    spec = synth.yolov4_608()
    heads = synth.make_heads(spec, sample, SEED, device=device)
    targets = synth.make_targets(spec, sample, SEED)
    return spec, heads, odet.make_infos(spec.shapes, spec.anchors), targets


def time_cpu_reference(sample, steps, warmup, threads=None):
    import torch
    cores = threads or os.cpu_count() or 1
    torch.set_num_threads(cores)
    spec, heads, infos, targets = cpu_reference_inputs(sample)
    for _ in range(warmup):
        cpu_reference_step(heads, infos, spec.num_classes, targets)
    t0 = time.perf_counter()
    for _ in range(steps):
        cpu_reference_step(heads, infos, spec.num_classes, targets)
    dt = time.perf_counter() - t0
    return sample * steps / dt, dt / steps, cores


def run_reference(args):
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = args.gpus
    steps, warmup = max(1, min(args.steps, 20)), max(1, min(args.warmup, 3))
    v, per_step, cores = time_cpu_reference(args.cpu_sample, steps, warmup)
    sample = f"{args.cpu_sample} images of the same workload per step, {steps} steps (bounded sample)"
    line = {"impl": "reference", "metric": METRIC, "value": v, "unit": UNIT, "n_gpus": args.gpus,
            "steps": steps, "warmup": warmup, "ms_per_step": per_step * 1e3, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": workload_config(args, world),
            "cpu_baseline": {"value": v, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
            "e2e": {"value": v, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "note": "oracle restatement of the reference's tch CPU path (same ATen kernels via PyTorch, "
                    "C matcher, torchvision batched_nms for tch-nms); the Rust reference cannot be built here; "
                    "one CPU process on rank 0 whatever N is"}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------
# helpers of the native arm
# ----------------------------------------------------------------------------------
def graph_us(dev, fn, reps=200):
    """median-free mean time of `fn(stream)` captured as a CUDA graph and replayed `reps` times"""
    import torch
    fn()
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    cap = torch.cuda.Stream(dev)
    with torch.cuda.graph(g, stream=cap):
        fn(cap.cuda_stream)
    for _ in range(5):
        g.replay()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(reps):
        g.replay()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / reps * 1e3


def in_graph_trace(step, heads, gt_d, off_d, replays=12):
    """Where every kernel of the replayed step graph runs: yh_trace_enable puts a %globaltimer
    stamp before and after every launch INSIDE a second capture of the same step.  Returns per
    launch (kernel, start_us, end_us) relative to the first start, median over the replays, and
    the span; the stamps add two one-thread nodes per kernel, so the span is a few us longer
    than the untraced step."""
    import numpy as np
    import torch
    from yolo_dl_b200 import _lib
    lib = _lib.load()
    cap_n = 128
    buf = torch.zeros(2 * cap_n, dtype=torch.int64, device=step.dev)
    lib.yh_trace_enable(buf.data_ptr(), cap_n)
    step.capture(heads, gt_d, off_d)
    ids = (ctypes.c_int32 * cap_n)()
    n = min(cap_n, lib.yh_trace_ids(ids, cap_n))
    lib.yh_trace_enable(None, 0)
    k = step.kernels_in_graph
    first = n - k          # the warm-up run inside capture() was stamped as well
    names = [_lib.KERNEL_NAMES[ids[i]] for i in range(first, n)]
    rows = []
    for _ in range(replays):
        step.replay()
        torch.cuda.synchronize()
        rows.append(buf.cpu().numpy().astype(np.int64).reshape(cap_n, 2)[first:n])
    t = np.stack(rows[2:])
    t0 = t[:, :, 0].min(axis=1)[:, None, None]
    rel = np.median((t - t0) / 1000.0, axis=0)
    launches = [{"kernel": names[i], "start_us": round(float(rel[i, 0]), 2), "end_us": round(float(rel[i, 1]), 2)}
                for i in sorted(range(k), key=lambda i: rel[i, 0])]
    per_kernel = {}
    for i in range(k):
        d = per_kernel.setdefault(names[i], {"launches": 0, "us_sum": 0.0})
        d["launches"] += 1
        d["us_sum"] = round(d["us_sum"] + float(rel[i, 1] - rel[i, 0]), 2)
    return {"span_us": round(float(rel[:, 1].max()), 2), "per_kernel": per_kernel, "launches": launches,
            "method": "%globaltimer stamps inside a traced capture of the same step graph (yh_trace_enable), "
                      f"median of {replays - 2} replays"}


def time_windows(one_step, K, windows, barrier, world, dev):
    """`windows` timed regions of exactly K steps, each bracketed by barrier + synchronize; device
    time by CUDA events, max over ranks.  Returns the list of window times in ms."""
    import torch
    import torch.distributed as dist
    out = []
    for _ in range(windows):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        barrier()
        e0.record()
        for _ in range(K):
            one_step()
        e1.record()
        barrier()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t)
        out.append(ms)
    return out


def step_us_at_batch(dev, spec, B, K, seed):
    """device-resident step time of a batch of B images on this GPU (CUDA graph, median of 3 windows)"""
    import torch
    import yolo_dl_b200 as yh
    from yolo_dl_b200 import synth
    from yolo_dl_b200.pipeline import DeviceStep
    heads = synth.make_heads(spec, B, seed, device=dev)
    gt_h, off_h = yh.pack_labels(synth.make_targets(spec, B, seed))
    gt_d, off_d = gt_h.to(dev), off_h.to(dev)
    st = DeviceStep(spec, B, dev, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=BY_CLASS)
    st.capture(heads, gt_d, off_d)
    for _ in range(5):
        st.replay()
    st.counts()
    ws = []
    for _ in range(3):
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        a.record()
        for _ in range(K):
            st.replay()
        b.record()
        torch.cuda.synchronize()
        ws.append(a.elapsed_time(b) / K * 1e3)
    return statistics.median(ws), st.kernels_in_graph


# ----------------------------------------------------------------------------------
# side numbers (BASELINE.json configs[0], [2], [3], [4]; parity cases in tests/, reported here,
# not the metric) and the reference's own ATen op sequence run eagerly on this GPU
# ----------------------------------------------------------------------------------
def aten_cuda_eager(dev, sample=64, reps=5):
    """The competitor a yolo-dl user actually runs on this box: the reference's op sequence through
    libtorch's CUDA kernels, eagerly (tch has no graphs), host matcher in C like the reference's
    single-threaded one, torchvision's CUDA batched_nms standing in for tch-nms.  Informational:
    the stated baseline of this tier stays the CPU port."""
    import torch
    spec, heads, infos, targets = cpu_reference_inputs(sample, device=dev)
    cpu_reference_step(heads, infos, spec.num_classes, targets)
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    for _ in range(reps):
        cpu_reference_step(heads, infos, spec.num_classes, targets)
    torch.cuda.synchronize()
    dt = (time.perf_counter() - t0) / reps
    return {"images_per_s": sample / dt, "ms_per_step": dt * 1e3, "batch": sample,
            "what": "oracle/{detect,loss,nms}.py with CUDA tensors (ATen CUDA kernels, eager, one stream), "
                    "host f64 matcher in C, torchvision CUDA batched_nms; wall clock incl. its host syncs"}


def side_configs(dev):
    import torch
    import yolo_dl_b200 as yh
    from yolo_dl_b200 import synth
    from yolo_dl_b200.pipeline import DeviceStep

    def cur(s):
        return s if s is not None else torch.cuda.current_stream(dev).cuda_stream

    out = {}
    # configs[0]: YOLOv4-608 head decode + NMS, batch 8 (the reference's own CPU-runnable case)
    spec = synth.yolov4_608()
    B = 8
    heads = synth.make_heads(spec, B, SEED - 1, device=dev)
    hp = yh._lib.ptr_array(heads)
    cfg0 = {}
    for conf, iou in ((0.4, 0.5), (0.1, 0.6)):
        for by_class in (False, True):
            st = DeviceStep(spec, B, dev, nms_conf=conf, nms_iou=iou, by_class=by_class, with_grad=False,
                            cand_per_image=16384, max_detections=B * 16384)

            def dn(s=None, st=st):
                st._decode(hp, cur(s))
                st._nms(hp, cur(s))
            us = graph_us(dev, dn)
            _, K = st.counts()
            cfg0[f"conf{conf}_iou{iou}_{'per_class' if by_class else 'per_image'}"] = {
                "us_per_batch": us, "images_per_s": B / (us * 1e-6), "kept": K}
            del st
    out["yolov4_608_decode_nms_b8"] = cfg0
    del heads
    # configs[2]: YOLOv4-tiny 416 decode + NMS, batch 256 (latency-bound small-head path)
    spec = synth.yolov4_tiny_416()
    B = 256
    heads = synth.make_heads(spec, B, SEED + 2, device=dev)
    hp = yh._lib.ptr_array(heads)
    cfg2 = {}
    for by_class in (False, True):
        st = DeviceStep(spec, B, dev, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=by_class, with_grad=False)

        def tiny(s=None, st=st):
            st._decode(hp, cur(s))
            st._nms(hp, cur(s))
        us = graph_us(dev, tiny)
        n0 = int(st.lib.yh_launch_count(0))
        tiny()
        cfg2["per_class" if by_class else "per_image"] = {"us_per_batch": us,
                                                          "launches_per_batch": int(st.lib.yh_launch_count(0)) - n0,
                                                          "images_per_s": B / (us * 1e-6)}
        del st
    out["yolov4_tiny_416_decode_nms_b256"] = cfg2
    del heads
    # configs[3]: YOLOv4-CSP 640, 300 GT boxes per image (30 % clustered): target assignment alone
    # and the whole loss step (assignment + loss fwd+bwd with the full head gradient)
    spec = synth.yolov4_csp_640()
    B = 64
    targets = synth.make_targets(spec, B, SEED + 3, n_min=300, n_max=300, clustered=0.3)
    gt_h, off_h = yh.pack_labels(targets)
    gt_d, off_d = gt_h.to(dev), off_h.to(dev)
    heads = synth.make_heads(spec, B, SEED + 3, device=dev)
    hp = yh._lib.ptr_array(heads)
    st = DeviceStep(spec, B, dev, max_gt=300 * B)

    def assign(s=None):
        st._assign(gt_d, off_d, cur(s))
    us = graph_us(dev, assign)
    M, _ = st.counts()

    def loss_step(s=None):
        st._assign(gt_d, off_d, cur(s))
        st._loss(st.lib.yh_loss_fwd_bwd, hp, cur(s))
    us_l = graph_us(dev, loss_step, reps=50)
    n_inst = st.N * B
    out["yolov4_csp_640_300gt_b64"] = {
        "assign_us_per_batch": us, "matches": M, "gt_boxes": 300 * B, "gt_boxes_per_s": 300 * B / (us * 1e-6),
        "assign_plus_loss_fwd_bwd_us_per_batch": us_l, "images_per_s": B / (us_l * 1e-6),
        "algorithmic_GBps": (344 * n_inst + 680 * M) / (us_l * 1e-6) / 1e9}
    del st, heads
    # configs[4]: batched NMS sweep, B = 8 images x K candidates per image, 80 classes, per image
    # (the mode every reference binary uses) and per class, through the tch_nms::nms_by_scores drop-in.
    # Above 65 536 rows in total this is ONE global NMS per group, not the reference's chunked result.
    sweep = {}
    Bn = 8
    for Kc in (1000, 4000, 16000, 65536, 100000, 200000):
        import numpy as np
        parts = [synth.make_nms_boxes(Kc, SEED + Kc + b, num_groups=80) for b in range(Bn)]
        ltrb = np.concatenate([p[0] for p in parts])
        scores = np.concatenate([p[1] for p in parts])
        cls = np.concatenate([p[2] for p in parts])
        img = np.repeat(np.arange(Bn, dtype=np.int64), Kc)
        row = {}
        for mode, grp in (("per_image", img), ("per_class", img * 80 + cls)):
            args_d = [torch.from_numpy(x).to(dev) for x in (ltrb, scores, grp)]
            yh.nms_by_scores(*args_d, 0.5)
            torch.cuda.synchronize()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            reps = 5
            a.record()
            for _ in range(reps):
                kept = yh.nms_by_scores(*args_d, 0.5)
            b.record()
            torch.cuda.synchronize()
            ms = a.elapsed_time(b) / reps
            row[mode] = {"ms": ms, "kept": int(kept.numel()), "candidates_per_s": Bn * Kc / (ms * 1e-3)}
        sweep[str(Kc)] = row
    out["nms_sweep_b8_x_candidates_per_image"] = sweep
    try:
        out["aten_cuda_eager"] = aten_cuda_eager(dev)
    except Exception as e:  # informational only
        out["aten_cuda_eager"] = {"unavailable": repr(e)[:200]}
    return out


# ----------------------------------------------------------------------------------
# native arm
# ----------------------------------------------------------------------------------
def run_native(args):
    import torch
    import torch.distributed as dist
    import yolo_dl_b200 as yh
    from yolo_dl_b200 import _lib, synth
    from yolo_dl_b200 import dist as ydist
    from yolo_dl_b200.pipeline import DeviceStep, HostStep

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a CUDA device: yolohot has no CPU fallback")
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    lib = yh.load()
    spec = synth.yolov4_608()
    B, K, W = args.batch, args.steps, max(args.warmup, 3)
    nwin = max(1, args.windows)
    # rank-specific synthetic data, generated on the device (resident in HBM before timing)
    heads = synth.make_heads(spec, B, SEED + rank, device=dev)
    targets = synth.make_targets(spec, B, SEED + rank)
    gt_h, off_h = yh.pack_labels(targets)
    gt_d, off_d = gt_h.to(dev), off_h.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=BY_CLASS)

    # the loss-scalar exchange data-parallel training needs, N > 1: the step's last
    # kernel stores this rank's five weighted scalars into every peer's mailbox over NVLink
    # (CUDA IPC peer memory, yolo_dl_b200/csrc/peer.cu) — part of the graph, no collective kernel
    # and no host launch per step.  --loss-exchange nccl: one 5-float NCCL all-reduce per step on
    # its own stream instead (the baseline this replaces).
    peer = reducer = None
    if world > 1 and args.loss_exchange == "peer":
        peer = ydist.PeerLossExchange(dev, B)
        step.capture(heads, gt_d, off_d, publisher=peer)
    else:
        step.capture(heads, gt_d, off_d, loss_slots=2 if world > 1 else 1)
        reducer = ydist.LossAllReduce(dev, B) if world > 1 else None

    def one_step():
        step.replay()
        if reducer is not None:
            reducer.launch_weighted(step.weighted5)

    for _ in range(W):
        one_step()
    M, Kdet = step.counts()
    barrier()
    with ClockSampler(local) as clk:
        win_ms = time_windows(one_step, K, nwin, barrier, world, dev)
    step.counts()   # raises if any replay of the timed region overflowed a buffer (sticky flag)
    replays = W + K * nwin
    launches = step.kernels_in_graph * K
    ms = statistics.median(win_ms)
    value = world * B * K / (ms * 1e-3)
    per_step = sorted(w / K for w in win_ms)
    value_windows = {"windows": nwin, "steps_per_window": K, "ms_per_step": [w / K for w in win_ms],
                     "median": ms / K, "min": per_step[0], "max": per_step[-1],
                     "spread_pct": 100.0 * (per_step[-1] - per_step[0]) / (ms / K)}

    # the combined loss of the last step: what the peer mailboxes hold vs a NCCL all-reduce of the
    # same weighted scalars (outside the timed region; a check, not part of the step)
    exchange = None
    if world > 1:
        chk = ydist.weighted_mean_losses(step.losses4, B)
        torch.cuda.synchronize()
        if peer is not None:
            peer.publish(step.losses4)   # the graph publishes one step behind: flush the last one
            torch.cuda.synchronize()
            dist.barrier()
            comb, seq = peer.collect()
            exchange = {"kind": "peer-memory stores over NVLink, in the step graph (csrc/peer.cu)",
                        "combined_total": float(comb[0]), "nccl_allreduce_check": float(chk[0]),
                        "published_steps": seq}
            assert seq == replays + 1, (seq, replays + 1)
            assert abs(float(comb[0]) - float(chk[0])) <= 1e-6 * abs(float(chk[0])) + 1e-7, (comb, chk)
        else:
            exchange = {"kind": "NCCL all-reduce per step on its own stream",
                        "combined_total": float(chk[0])}
    loss_value = float(step.losses4[0])

    # the same kernels launched one after another on one stream (no graph), for comparison
    Ks = max(3, min(K, 50))
    for _ in range(3):
        step.run(heads, gt_d, off_d)
    barrier()
    q0, q1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    q0.record()
    for _ in range(Ks):
        step.run(heads, gt_d, off_d)
    q1.record()
    barrier()
    seq_ms = q0.elapsed_time(q1) / Ks

    # per-kernel device times, each kernel ALONE on the stream, measured live with CUDA events on
    # the launch stream (whole-batch launches of the one-sequence form of the step)
    lib.yh_timing_enable(1)
    reps = max(5, min(K, 20))
    for _ in range(reps):
        step.run(heads, gt_d, off_d)
    nk = len(_lib.KERNEL_NAMES)
    sums = (ctypes.c_double * nk)()
    cnts = (ctypes.c_int64 * nk)()
    _lib.check(lib.yh_timing_read(sums, cnts, nk))
    lib.yh_timing_enable(0)
    kern_ms = {n: sums[i] / cnts[i] for i, n in enumerate(_lib.KERNEL_NAMES) if cnts[i]}

    # instances that pass the objectness test of the candidate filter (its class-logit traffic)
    n_inst = step.N * B
    obj_pass = 0
    for h, (H, W_) in zip(heads, spec.shapes):
        o = h.view(B, 5 + spec.num_classes, -1, H, W_)[:, 4]
        obj_pass += int((torch.sigmoid(o) >= NMS_CONF).sum())
    alg = step.algorithmic_bytes(matches=M, obj_pass=obj_pass, detections=Kdet)
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    # DRAM bytes of the dominant kernel from an `ncu --set full` capture: a committed record
    # (profiles/traffic.json) that names the commit and the command that produced it
    traffic = traffic_src = None
    try:
        tj = json.load(open(os.path.join(ROOT, "profiles", "traffic.json")))
        traffic = tj.get("loss_dense_kernel")
        traffic_src = tj.get("_source")
    except Exception:
        pass
    dom = "loss_dense_kernel"
    achieved = alg[dom] / (kern_ms[dom] * 1e-3) / 1e9
    roofline = {"kernel": dom, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "traffic_source": traffic_src,
                "dram_frac": (traffic / (kern_ms[dom] * 1e-3) / 1e9 / peak) if traffic else None,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": alg[dom], "kernel_ms": kern_ms[dom],
                "note": "frac is on ALGORITHMIC bytes of the kernel timed alone; dram_frac = ncu DRAM bytes of the "
                        "same launch / kernel_ms / peak (part of the gradient is still dirty in the L2 when the "
                        "kernel ends and is written back under whatever runs next: the last YH_KEEP_MB = 64 MB of "
                        "the stream are stored evict_last on purpose, for the scatter behind it).  frac may exceed "
                        "1.0 by a few percent: the peak is MEASURED_PEAKS' COPY bandwidth (read + write), this "
                        "kernel is a pure store stream (bare fills reach 6.9-7.0 TB/s on this GPU, "
                        "profiles/r1f_fill_bench.txt) and does not have to drain the L2 before it ends",
                "other_kernels": {n: {"ms": kern_ms[n], "algorithmic_bytes": alg.get(n),
                                      "achieved_GBps": (alg[n] / (kern_ms[n] * 1e-3) / 1e9) if alg.get(n) else None}
                                  for n in kern_ms if n != dom}}
    # whole-step view: SURVEY.md §8d algorithmic bytes of everything the step does over the step time
    step_bytes = sum(alg.values())
    roofline["step"] = {"algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms / K * 1e-3) / 1e9,
                        "frac": step_bytes / (ms / K * 1e-3) / 1e9 / peak, "unit": "GB/s",
                        "terms": {"matches": M, "objectness_passing_instances": obj_pass, "detections": Kdet,
                                  "instances": n_inst, "formula": "sum of algorithmic_bytes of all kernels: "
                                  "32NB + 344NB + 340M + 340M + (4NB + 32·C·P) + 24K + 32K"},
                        "note": "all kernels of the step over the median step time"}
    # where the kernels run inside the replayed graph (second, traced capture of the same step)
    try:
        roofline["in_graph"] = in_graph_trace(step, heads, gt_d, off_d)
        d = roofline["in_graph"]["per_kernel"].get(dom)
        if d:
            roofline["in_graph"]["dominant_achieved_GBps"] = alg[dom] / (d["us_sum"] * 1e-6) / 1e9
    except Exception as e:
        roofline["in_graph"] = {"unavailable": repr(e)[:200]}

    # ---- f3: the data-parallel gradient all-reduce either side of the path (N > 1) ---------------
    # What replaces sync_weights + the to_device fan-in through GPU 0 (multi_gpu.rs:503-530,656-726):
    # a bucketed NCCL all-reduce of a YOLOv4-sized parameter set (64 M f32 = 256 MB), alone and on
    # its own stream under the replayed step graph — the one real collective of a training step.
    dp_allreduce = None
    if world > 1:
        from yolo_dl_b200 import dp
        params = [torch.zeros(4 << 20, device=dev, requires_grad=True) for _ in range(16)]
        gar = dp.GradientAllReduce(params, bucket_bytes=64 << 20)
        nbytes = gar.nbytes()
        for _ in range(3):
            gar.reduce(B, B * world)
        barrier()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps_ar = 10
        a0.record()
        for _ in range(reps_ar):
            gar.reduce(B, B * world)
        a1.record()
        barrier()
        ar_ms = a0.elapsed_time(a1) / reps_ar
        # under the step: all-reduce on a side stream, the step graph on the current one
        side = torch.cuda.Stream(dev)
        Ko = max(10, min(K, 100))
        barrier()
        o0, o1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        o0.record()
        for _ in range(Ko):
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):
                gar.reduce(B, B * world)
            step.replay()
            torch.cuda.current_stream(dev).wait_stream(side)
        o1.record()
        barrier()
        both_ms = o0.elapsed_time(o1) / Ko
        t = torch.tensor([ar_ms, both_ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ar_ms, both_ms = float(t[0]), float(t[1])
        dp_allreduce = {"bytes": nbytes, "buckets": len(gar.buckets), "ms_alone": ar_ms,
                        "busbw_GBps": 2.0 * (world - 1) / world * nbytes / (ar_ms * 1e-3) / 1e9,
                        "ms_step_plus_allreduce_overlapped": both_ms, "ms_step_alone": ms / K,
                        "nvlink5_per_direction_GBps": 900,
                        "what": "dp.GradientAllReduce (NCCL, PreMulSum B_local/B) of 64 M f32 in 64 MB buckets"}
        del gar, params

    # ---- strong scaling: the literal BASELINE configs[1], ONE global batch of 64 over the N GPUs ----
    strong = None
    if not args.no_strong:
        Bs = GLOBAL_BATCH // world
        if world == 1:
            # one GPU: the per-GPU share of N = 1/2/4/8 run here, i.e. the compute side of the strong curve
            pts = {}
            for n_ in (1, 2, 4, 8):
                b_ = GLOBAL_BATCH // n_
                if b_ == B:
                    pts[str(n_)] = {"batch_per_gpu": b_, "us_per_step": ms / K * 1e3,
                                    "images_per_s_if_n_gpus": GLOBAL_BATCH / (ms / K * 1e-3)}
                    continue
                us, nk_ = step_us_at_batch(dev, spec, b_, max(20, min(K, 200)), SEED + 100 + n_)
                pts[str(n_)] = {"batch_per_gpu": b_, "us_per_step": us, "kernels": nk_,
                                "images_per_s_if_n_gpus": GLOBAL_BATCH / (us * 1e-6),
                                "traffic_floor_us": 380 * step.N * b_ / (peak * 1e9) * 1e6}
            base = pts["1"]["us_per_step"]
            for n_ in pts:
                pts[n_]["efficiency"] = base / (int(n_) * pts[n_]["us_per_step"])
            strong = {"global_batch": GLOBAL_BATCH, "measured_on": "one GPU running the per-GPU share of N GPUs",
                      "per_n": pts}
        elif Bs >= 1:
            heads_s = [h[:Bs].contiguous() for h in heads]
            gt_s, off_s = yh.pack_labels(targets[:Bs])
            st_s = DeviceStep(spec, Bs, dev, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=BY_CLASS)
            gts_d, offs_d = gt_s.to(dev), off_s.to(dev)
            st_s.capture(heads_s, gts_d, offs_d)
            for _ in range(W):
                st_s.replay()
            st_s.counts()
            w_s = time_windows(st_s.replay, K, nwin, barrier, world, dev)
            ms_s = statistics.median(w_s)
            strong = {"global_batch": GLOBAL_BATCH, "batch_per_gpu": Bs, "n_gpus": world,
                      "value": GLOBAL_BATCH * K / (ms_s * 1e-3), "unit": UNIT, "ms_per_step": ms_s / K,
                      "ms_per_step_windows": [w / K for w in w_s], "scaling": "strong",
                      "note": "one global batch of 64 split over the GPUs; no loss exchange in this graph "
                              "(its 32-thread publish kernel runs one step behind, off the critical path)"}
            del st_s, heads_s

    # end to end through host buffers (pinned): every step pays H2D of its heads+labels, the
    # kernels, and D2H of losses + matches + detections + head gradients.  HostPipeline overlaps
    # step i's compute with step i+1's H2D and step i-1's D2H (2 slots, 3 streams); HostStep (one
    # stream, nothing overlapped) is timed beside it.
    e2e = None
    if not args.no_e2e:
        from yolo_dl_b200.pipeline import HostPipeline
        host_heads = [h.cpu().pin_memory() for h in heads]
        Ke = max(4, min(K, 20))

        def timed(fn, finish):
            barrier()
            a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            t0 = time.perf_counter()
            a0.record()
            for _ in range(Ke):
                fn()
            finish()
            a1.record()
            torch.cuda.synchronize()
            wall = (time.perf_counter() - t0) * 1e3
            barrier()
            m = max(wall, a0.elapsed_time(a1))
            if world > 1:
                t = torch.tensor([m], device=dev)
                dist.all_reduce(t, op=dist.ReduceOp.MAX)
                m = float(t)
            return m

        hs = HostStep(spec, B, dev, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=BY_CLASS)
        for _ in range(2):
            hs.run(host_heads, gt_h, off_h)
        serial_ms = timed(lambda: hs.run(host_heads, gt_h, off_h), lambda: None)
        del hs
        pipe = HostPipeline(spec, B, dev, depth=2, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=BY_CLASS)
        last = {}

        def pstep():
            k = pipe.submit(host_heads, gt_h, off_h)
            if "k" in last:          # read the previous step's loss while this one is in flight
                last["r"] = pipe.result(last["k"])
                last["loss"] = float(last["r"][0][0])
            last["k"] = k

        def pfinish():
            last["r"] = pipe.result(last["k"])
            last["loss"] = float(last["r"][0][0])
            pipe.drain()

        for _ in range(3):
            pstep()
        pfinish()
        last.clear()
        ems = timed(pstep, pfinish)
        hl, hM, hK = last["r"]
        each_way = [pipe.h2d_bytes / (ems / Ke) / 1e6, pipe.d2h_bytes / (ems / Ke) / 1e6]
        e2e = {"value": world * B * Ke / (ems * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
               "d2h_bytes_per_step": pipe.d2h_bytes, "steps": Ke, "ms_per_step": ems / Ke,
               "includes": "per step: H2D heads+labels, the step graph, D2H losses+matches+detections+head "
                           "gradients; 2-slot pipeline over copy-in/compute/copy-out streams; timed by "
                           "max(wall clock, CUDA events) around the loop incl. the final drain",
               "unpipelined_ms_per_step": serial_ms / Ke, "pcie_GBps_each_way": each_way,
               "loss": float(hl[0]), "matches": hM, "detections": hK}
        # the box's own ceiling for this copy pattern (bare pinned cudaMemcpyAsync both ways from
        # the same number of concurrent processes, tools/pcie_sweep.sh -> profiles/r2_pcie.json)
        try:
            pj = json.load(open(os.path.join(ROOT, "profiles", "r2_pcie.json")))
            ceil = pj["duplex_GBps_per_gpu"].get(str(world))
            if ceil:
                e2e["pcie_ceiling_GBps_each_way"] = ceil
                e2e["pcie_frac"] = min(each_way) / min(ceil)
                e2e["pcie_ceiling_source"] = pj.get("_source")
        except Exception:
            pass
        del pipe, host_heads

    extras = None
    if rank == 0 and world == 1 and not args.no_extras:
        extras = side_configs(dev)

    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        v, per, cores = time_cpu_reference(args.cpu_sample, 20, 1)
        cpu = {"value": v, "unit": UNIT, "cores": cores, "kind": "port",
               "sample": f"{args.cpu_sample} images of the same workload, 20 timed passes after 1 warm-up "
                         f"({per:.2f} s/pass)"}
        # the same path on ONE host thread (the reference's matcher is single-threaded
        # anyway; this pins the ATen part to one core too): 2 passes of a quarter of the sample
        q = max(1, args.cpu_sample // 4)
        v1, per1, _ = time_cpu_reference(q, 2, 1, threads=1)
        cpu["single_thread"] = {"value": v1, "unit": UNIT, "cores": 1,
                                "sample": f"{q} images, 2 timed passes after 1 warm-up ({per1:.2f} s/pass)"}

    if rank == 0:
        line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": W,
                "ms_per_step": ms / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
                "dtype": "f32", "data": "synthetic", "config": workload_config(args, world),
                "value_windows": value_windows, "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "strong": strong, "dp_gradient_allreduce": dp_allreduce, "gpu_launches": launches,
                "execution": f"CUDA graph of {step.kernels_in_graph} kernels on forked/joined streams "
                             f"(schedule: {step.schedule}), replayed once per step",
                "sequential_ms_per_step": seq_ms, "loss_exchange": exchange, "other_configs": extras,
                "clocks": clk.summary(), "matches_per_step": M, "detections_per_step": Kdet,
                "loss": loss_value, "commit": git_commit()}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)
