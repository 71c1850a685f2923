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
train.json5:153-154) per class.  Weak scaling: every rank owns its own 64 images; the only
collective is the 5-float loss all-reduce.  One step = yh_decode_merged + yh_assign +
yh_loss_dense + yh_loss_rows_compute + yh_loss_finish (full head gradient) + yh_nms = 9 kernels, captured once as
a CUDA graph over four forked/joined streams (yolo_dl_b200/pipeline.py) and replayed per step.
"""
import argparse
import json
import os
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


def parse():
    p = argparse.ArgumentParser()
    p.add_argument("--gpus", type=int, default=1)
    p.add_argument("--steps", type=int, default=1000)
    p.add_argument("--warmup", type=int, default=10)
    p.add_argument("--impl", default="native", choices=["native", "reference"])
    p.add_argument("--batch", type=int, default=64, help="images per GPU")
    p.add_argument("--cpu-sample", type=int, default=64, help="images per CPU-baseline pass")
    p.add_argument("--no-cpu-baseline", action="store_true")
    p.add_argument("--no-e2e", action="store_true")
    p.add_argument("--no-extras", action="store_true", help="skip the side numbers of BASELINE configs 2-4")
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


def cpu_reference_inputs(sample):
    from oracle import detect as odet
    from yolo_dl_b200 import synth
    spec = synth.yolov4_608()
    heads = synth.make_heads(spec, sample, SEED)
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
                    "C matcher, torchvision batched_nms for tch-nms); the Rust reference cannot be built here"}
    print(json.dumps(line), flush=True)


# ----------------------------------------------------------------------------------
# side numbers (BASELINE.json configs[2..4]; parity cases in tests/, reported here, not the metric)
# ----------------------------------------------------------------------------------
def side_configs(dev):
    import torch
    import yolo_dl_b200 as yh
    from yolo_dl_b200 import synth
    from yolo_dl_b200.pipeline import DeviceStep

    def graph_us(fn, reps=200):
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

    out = {}
    # configs[2]: YOLOv4-tiny 416 decode + NMS, batch 256 (latency-bound small-head path)
    spec = synth.yolov4_tiny_416()
    B = 256
    heads = synth.make_heads(spec, B, SEED + 2, device=dev)
    st = DeviceStep(spec, B, dev, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=BY_CLASS, with_grad=False)
    hp = yh._lib.ptr_array(heads)

    def tiny(s=None):
        s = s if s is not None else torch.cuda.current_stream(dev).cuda_stream
        st._decode(hp, s)
        st._nms(hp, s)
    us = graph_us(tiny)
    out["yolov4_tiny_416_decode_nms_b256"] = {"us_per_batch": us, "launches_per_batch": 4,
                                              "images_per_s": B / (us * 1e-6)}
    del st, heads
    # configs[3]: YOLOv4-CSP 640, 300 GT boxes per image (30 % clustered), target assignment only
    spec = synth.yolov4_csp_640()
    B = 64
    targets = synth.make_targets(spec, B, SEED + 3, n_min=300, n_max=300, clustered=0.3)
    gt_h, off_h = yh.pack_labels(targets)
    gt_d, off_d = gt_h.to(dev), off_h.to(dev)
    st = DeviceStep(spec, B, dev, with_grad=False, max_gt=300 * B)

    def assign(s=None):
        st._assign(gt_d, off_d, s if s is not None else torch.cuda.current_stream(dev).cuda_stream)
    us = graph_us(assign)
    M, _ = st.counts()
    out["yolov4_csp_640_assign_300gt_b64"] = {"us_per_batch": us, "matches": M, "gt_boxes": 300 * B,
                                              "gt_boxes_per_s": 300 * B / (us * 1e-6)}
    del st
    # configs[4]: grouped NMS on explicit candidates (the tch_nms::nms_by_scores drop-in), 80 classes
    sweep = {}
    for Kc in (1000, 4000, 16000, 65536, 100000, 200000):
        ltrb, scores, grp = synth.make_nms_boxes(Kc, SEED + Kc, num_groups=80)
        args_d = [torch.from_numpy(x).to(dev) for x in (ltrb, scores, grp)]
        yh.nms_by_scores(*args_d, 0.5)
        torch.cuda.synchronize()
        a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        reps = 10
        a.record()
        for _ in range(reps):
            kept = yh.nms_by_scores(*args_d, 0.5)
        b.record()
        torch.cuda.synchronize()
        ms = a.elapsed_time(b) / reps
        sweep[str(Kc)] = {"ms": ms, "kept": int(kept.numel()), "candidates_per_s": Kc / (ms * 1e-3)}
    out["nms_by_scores_80_classes"] = sweep
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
    # rank-specific synthetic data, generated on the device (resident in HBM before timing)
    heads = synth.make_heads(spec, B, SEED + rank, device=dev)
    targets = synth.make_targets(spec, B, SEED + rank)
    gt_h, off_h = yh.pack_labels(targets)
    gt_d, off_d = gt_h.to(dev), off_h.to(dev)
    step = DeviceStep(spec, B, dev, nms_conf=NMS_CONF, nms_iou=NMS_IOU, by_class=BY_CLASS)

    # the step as a CUDA graph: 9 kernels on 4 forked/joined streams (pipeline.run_overlapped)
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
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    with ClockSampler(local) as clk:
        barrier()
        e0.record()
        for _ in range(K):
            one_step()
        e1.record()
        barrier()
    launches = step.kernels_in_graph * K
    ms = e0.elapsed_time(e1)
    if world > 1:
        t = torch.tensor([ms], device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        ms = float(t)
    value = world * B * K / (ms * 1e-3)

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
            assert seq == W + K + 1, (seq, W + K + 1)
            assert abs(float(comb[0]) - float(chk[0])) <= 1e-6 * abs(float(chk[0])) + 1e-7, (comb, chk)
        else:
            exchange = {"kind": "NCCL all-reduce per step on its own stream",
                        "combined_total": float(chk[0])}

    # the same 9 kernels launched one after another on one stream (no graph), for comparison
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

    # per-kernel device times, measured live with CUDA events on the launch stream
    import ctypes
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
    alg = step.algorithmic_bytes()
    peaks = {}
    try:
        peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
    except Exception:
        pass
    peak = float(peaks.get("hbm_gbs", 6650.0))
    peak_src = "measured (MEASURED_PEAKS.json hbm_gbs)" if "hbm_gbs" in peaks else "fallback 6650 GB/s"
    traffic = None
    try:
        traffic = json.load(open(os.path.join(ROOT, "profiles", "traffic.json"))).get("loss_dense_kernel")
    except Exception:
        pass
    dom = "loss_dense_kernel"
    achieved = alg[dom] / (kern_ms[dom] * 1e-3) / 1e9
    roofline = {"kernel": dom, "bound": "hbm", "achieved": achieved, "peak": peak, "unit": "GB/s",
                "frac": achieved / peak, "traffic": traffic, "peak_source": peak_src,
                "algorithmic_bytes_per_launch": alg[dom], "kernel_ms": kern_ms[dom],
                "other_kernels": {n: {"ms": kern_ms[n],
                                      "achieved_GBps": (alg[n] / (kern_ms[n] * 1e-3) / 1e9) if n in alg else None}
                                  for n in kern_ms if n != dom}}

    # whole-step view: SURVEY.md §8d algorithmic bytes of everything the step does (decode 32·N,
    # loss fwd+bwd 344·N + 340·M, filter 4·N objectness plane + 32-byte sectors of the C class
    # logits of the P instances that pass the objectness test, NMS 32·K) over the step time
    n_inst = step.N * B
    obj_pass = 0
    for h, (H, W_) in zip(heads, spec.shapes):
        o = h.view(B, 5 + spec.num_classes, -1, H, W_)[:, 4]
        obj_pass += int((torch.sigmoid(o) >= NMS_CONF).sum())
    step_bytes = 32 * n_inst + 344 * n_inst + 340 * M + 4 * n_inst + 32 * spec.num_classes * obj_pass + 32 * Kdet
    roofline["step"] = {"algorithmic_bytes": step_bytes, "achieved": step_bytes / (ms / K * 1e-3) / 1e9,
                        "frac": step_bytes / (ms / K * 1e-3) / 1e9 / peak, "unit": "GB/s",
                        "objectness_passing_instances": obj_pass,
                        "note": "all kernels of the step over the step time; the random 32-byte sector "
                                "reads/writes of the matched rows and of the NMS candidates cost about "
                                "twice their byte count in DRAM time (profiles/README.md)"}

    # end to end through host buffers (pinned): every step pays H2D of its heads+labels, the 8
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
        e2e = {"value": world * B * Ke / (ems * 1e-3), "unit": UNIT, "h2d_bytes_per_step": pipe.h2d_bytes,
               "d2h_bytes_per_step": pipe.d2h_bytes, "steps": Ke, "ms_per_step": ems / Ke,
               "includes": "per step: H2D heads+labels, 9 kernels (graph), D2H losses+matches+detections+head "
                           "gradients; 2-slot pipeline over copy-in/compute/copy-out streams; timed by "
                           "max(wall clock, CUDA events) around the loop incl. the final drain",
               "unpipelined_ms_per_step": serial_ms / Ke,
               "pcie_GBps_each_way": [pipe.h2d_bytes / (ems / Ke) / 1e6, pipe.d2h_bytes / (ems / Ke) / 1e6],
               "loss": float(hl[0]), "matches": hM, "detections": hK}
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
                "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e, "gpu_launches": launches,
                "execution": "CUDA graph of 9 kernels on 4 forked/joined streams, replayed once per step",
                "sequential_ms_per_step": seq_ms, "loss_exchange": exchange, "other_configs": extras,
                "clocks": clk.summary(), "matches_per_step": M, "detections_per_step": Kdet,
                "loss": float(step.losses4[0])}
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_native(a)
