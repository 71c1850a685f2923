"""Whole hot-path step (decode -> assign -> loss fwd+bwd -> NMS) on preallocated buffers.

`DeviceStep` is what a training worker does per batch after the backbone
(train/src/train/single_gpu.rs:152-186) with every buffer allocated once: four C-ABI calls, no
host synchronisation.  `HostStep` is the same step for callers whose tensors live in HOST memory
(the end-to-end form measured by bench.py): pinned staging, H2D of the head tensors and labels,
the step, D2H of losses / matches / detections and of the head gradients.
"""
import ctypes
import os
from typing import List, Optional, Sequence

import torch

from . import _lib
from .loss import YoloLossInit, pack_labels
from .tensors import DetectionInfo
from .detect import MergedDenseDetection
from .workspace import Workspace


class DeviceStep:
    def __init__(self, spec, batch: int, device: torch.device, loss: YoloLossInit = None,
                 nms_conf=0.4, nms_iou=0.5, by_class=True, max_gt=None, cand_per_image=1024,
                 max_detections=None, with_grad=True):
        self.lib = _lib.load()
        self.spec, self.B, self.dev = spec, batch, device
        self.C = spec.num_classes
        self.info: List[DetectionInfo] = MergedDenseDetection.make_info(spec.shapes, spec.anchors)
        self.layers = _lib.make_layers(self.info)
        self.L = len(self.info)
        self.N = sum(i.size for i in self.info)
        init = loss or YoloLossInit(box_metric="ciou")
        self.yolo_loss = init.build()  # validated config
        self.cfg = self.yolo_loss._cfg
        self.rect4 = 1 if init.match_grid_method == "Rect4" else 0
        self.thresh = float(init.anchor_scale_thresh)
        self.nms_conf, self.nms_iou, self.by_class = float(nms_conf), float(nms_iou), 1 if by_class else 0
        self.ws = Workspace()
        self.ws.ensure(device, batch, self.N, cand_per_image)
        f32, i64, i32 = torch.float32, torch.int64, torch.int32
        self.boxes = torch.empty((4, batch, 1, self.N), dtype=f32, device=device)
        max_gt = max_gt or 32 * batch
        self.mcap = max_gt * sum(i.A for i in self.info) * (5 if self.rect4 else 3)
        self.m_batch = torch.empty(self.mcap, dtype=i64, device=device)
        self.m_flat = torch.empty(self.mcap, dtype=i64, device=device)
        self.m_gt = torch.empty(self.mcap, dtype=i32, device=device)
        self.m_target = torch.empty((self.mcap, 4), dtype=f32, device=device)
        self.m_class = torch.empty(self.mcap, dtype=i64, device=device)
        self.M = torch.zeros(1, dtype=i32, device=device)
        self.losses4 = torch.zeros(4, dtype=f32, device=device)
        self.iou_score = torch.empty(self.mcap, dtype=f32, device=device)
        self.pred_rows = torch.empty((self.mcap, self.C + 5), dtype=f32, device=device)
        self.grads = [torch.empty(spec.head_shape(l, batch), dtype=f32, device=device)
                      for l in range(self.L)] if with_grad else None
        self.kcap = max_detections or batch * 1024
        self.kb = torch.empty(self.kcap, dtype=i64, device=device)
        self.kc = torch.empty(self.kcap, dtype=i64, device=device)
        self.ki = torch.empty(self.kcap, dtype=i64, device=device)
        self.tlbr = torch.empty((self.kcap, 4), dtype=f32, device=device)
        self.conf = torch.empty(self.kcap, dtype=f32, device=device)
        self.K = torch.zeros(1, dtype=i32, device=device)
        self._grad_ptrs = _lib.ptr_array(self.grads) if with_grad else None
        # staging of the matched-row gradients: reserved now, graphs cannot allocate
        self.ws.reserve_rows(self.mcap, self.C)
        self._side = None
        self._pub_stream = None
        self._graph = None
        # "overlapped": fork/join of whole-batch kernels (run_overlapped), the default; "grouped":
        # the loss in L2-sized image ranges (run_grouped) — measured and NOT faster on B200 (r2:
        # 318 / 230 us per step with groups of 8 / 16 images against 178: every ranged dense launch
        # pays ~6 us of ramp-up and drain, and the rows kernel is latency-bound at ~19 us per launch
        # however few rows it has), kept for A/B
        self.schedule = os.environ.get("YH_SCHEDULE", "overlapped")
        self.kernels_in_graph = 0

    def algorithmic_bytes(self, matches=0, obj_pass=0, detections=0):
        """SURVEY.md §8d algorithmic bytes of one launch of every kernel of the step (whole batch):
        decode 32·N·B; dense loss 344·N·B (4 read + 340 written); matched rows 340·M read and
        340·M written (scatter); candidate filter 4·N·B (objectness plane) + 32-byte sectors of
        the C class logits of the `obj_pass` instances that pass the objectness test — NOT the
        340·N·B of the reference's traffic; NMS 24·K' in + 8·K out."""
        n = self.N * self.B
        return {"decode_kernel": 32 * n, "loss_dense_kernel": 344 * n,
                "loss_rows_kernel": 340 * matches, "loss_finish_kernel": 340 * matches,
                "nms_filter_kernel": 4 * n + 32 * self.C * obj_pass,
                "nms_segment_kernel": 24 * detections, "nms_emit_kernel": 32 * detections}

    def expected_kernels(self, grouped=None):
        """kernel launches of one step: decode, 2 assignment, 2 NMS (4 with the large-segment NMS
        kernel, launched whenever the candidate capacity admits a segment above YH_BIG_MIN), and
        the loss: dense + rows + finish, or per image range dense + rows, then one finish."""
        grouped = (self.schedule == "grouped" and self._grad_ptrs is not None and self.B >= 8) \
            if grouped is None else grouped
        # NMS: filter + segment (which emits by itself) — or, when the candidate capacity admits a
        # segment above YH_BIG_MIN, filter + segment + large-segment kernel + emit
        fused = os.environ.get("YH_NMS_FUSE_EMIT", "1") != "0" and \
            self.B <= torch.cuda.get_device_properties(self.dev).multi_processor_count
        base = 3 + (4 if self.ws.key[3] > 1024 else (2 if fused else 3))
        if self.schedule == "layered" and self._grad_ptrs is not None and self.L >= 2 and not grouped:
            return base + 1 + 2 * self.L
        if not grouped:
            return base + 3
        group = max(4, (int(os.environ.get("YH_GROUP", "8")) // 4) * 4)
        R = (self.B + group - 1) // group
        if R > 16:
            group = ((self.B + 15) // 16 + 3) // 4 * 4
            R = (self.B + group - 1) // group
        return base + 2 * R + 1

    def run(self, heads: Sequence[torch.Tensor], gt_dev: torch.Tensor, off_dev: torch.Tensor, stream=None):
        """heads: L contiguous f32 CUDA tensors [B, A*(5+C), H, W]; gt_dev [G,5] f64; off_dev [B+1] i32.
        Enqueues 8 kernels on ONE stream, in the order the reference executes them; nothing
        synchronises."""
        s = stream if stream is not None else torch.cuda.current_stream(self.dev).cuda_stream
        self._decode(_lib.ptr_array(heads), s)
        self._assign(gt_dev, off_dev, s)
        self._loss(self.lib.yh_loss_fwd_bwd, _lib.ptr_array(heads), s)
        self._nms(_lib.ptr_array(heads), s)

    # -- the four C-ABI calls ------------------------------------------------------------
    def _decode(self, hp, s):
        b = self.boxes
        _lib.check(self.lib.yh_decode_merged(hp, self.layers, self.L, self.B, self.C, b[0].data_ptr(),
                                             b[1].data_ptr(), b[2].data_ptr(), b[3].data_ptr(), self.N, s))

    def _assign(self, gt_dev, off_dev, s):
        _lib.check(self.lib.yh_assign(self.ws.handle, gt_dev.data_ptr(), off_dev.data_ptr(), self.B,
                                      self.layers, self.L, self.rect4, self.thresh, self.m_batch.data_ptr(),
                                      self.m_flat.data_ptr(), self.m_gt.data_ptr(), self.m_target.data_ptr(),
                                      self.m_class.data_ptr(), self.M.data_ptr(), self.mcap, s))

    def _loss(self, fn, hp, s):
        _lib.check(fn(self.ws.handle, hp, self.layers, self.L, self.B, self.C, self.m_batch.data_ptr(),
                      self.m_flat.data_ptr(), self.m_target.data_ptr(), self.m_class.data_ptr(),
                      self.M.data_ptr(), self.mcap, self.cfg, self.losses4.data_ptr(),
                      self.iou_score.data_ptr(), self.pred_rows.data_ptr(), self._grad_ptrs, s))

    def _nms(self, hp, s):
        _lib.check(self.lib.yh_nms(self.ws.handle, hp, self.layers, self.L, self.B, self.C, self.nms_conf,
                                   self.nms_iou, self.by_class, self.kb.data_ptr(), self.kc.data_ptr(),
                                   self.ki.data_ptr(), self.tlbr.data_ptr(), self.conf.data_ptr(),
                                   self.K.data_ptr(), self.kcap, s))

    def _side_streams(self):
        """[decode, assignment + rows, NMS] streams of the fork/join.  The store stream of the dense
        loss runs on the caller's (default-priority) stream; what the block scheduler does with the
        other branches is decided by their stream priorities — CTAs of a higher priority are handed
        out first, equal priorities in launch order — and is worth 11 us of the 171 us step (r2g,
        profiles/r2g_prio*.txt, same box):

          * more than 16 images: decode and the assignment chain HIGH, the NMS chain DEFAULT.  The
            candidate filter has thousands of CTAs; on a high-priority stream they took every SM
            slot for the first 23 us, decode and the dense stream started behind them, decode then
            crawled under the stream for 115 us and the step was 171 us.  With the filter at the
            stream's own priority, launched before it, decode (20 us alone) and the filter share
            the first 40 us, the dense stream then runs nearly undisturbed (93 us in-graph instead
            of 121), and nms_segment / nms_emit land in the shadow of the scatter at the tail:
            160 us.  (Filter low but nms_segment / nms_emit high through a per-launch priority
            attribute: 170 us — they then run UNDER the stream and slow it.  Dense high: 175-181.)
          * up to 16 images every kernel is latency-bound and the filter is small: all chains HIGH,
            decode DEFAULT (35.5 us at 8 images against 40 with the setting above).

        YH_DEC_PRIO / YH_ASG_PRIO / YH_NMS_PRIO override (0 default, -1.. higher)."""
        big = self.B > 16
        prio = {"YH_DEC_PRIO": -1 if big else 0, "YH_ASG_PRIO": -1, "YH_NMS_PRIO": 0 if big else -1}
        return [torch.cuda.Stream(self.dev, priority=int(os.environ.get(k, v))) for k, v in prio.items()]

    # -- the same 8 kernels as a fork/join over four streams --------------------------------
    def run_overlapped(self, heads, gt_dev, off_dev, publisher=None, lagged_publish=True):
        """The step as the dependency graph it really is (the reference runs it as one sequence,
        train/src/train/single_gpu.rs:152-186):

            decode                                            (Detect2D x L + merge)
            assign -> loss_rows (staged row gradients) --\
            loss_dense ----------------------------------+--> loss_finish (scatter + reduce)
            nms_filter -> nms_segment (emits the kept lists itself)

        (YoloLoss fwd+bwd: neither the dense objectness pass nor the matched rows need each
        other; only the few-microsecond scatter of the staged row gradients over the dense
        gradient is ordered after both.)
        The four branches are forked from the current stream onto side streams and joined back
        into it, so the latency-bound kernels (one CTA per image: assignment, greedy NMS) run
        under the HBM-bound ones (decode, dense loss, candidate filter).  Capturable in a CUDA
        graph (`capture`)."""
        main = torch.cuda.current_stream(self.dev)
        if self._side is None:
            self._side = self._side_streams()
        s_dec, s_asg, s_nms = self._side
        hp = _lib.ptr_array(heads)
        dense_last = os.environ.get("YH_DENSE_LAST", "1") == "1"
        fork = torch.cuda.Event()
        fork.record(main)
        for st in self._side:
            st.wait_event(fork)
        pub_done = None
        if publisher is not None and lagged_publish:
            # the data-parallel loss exchange (dist.PeerLossExchange) one step behind: the stores
            # of step i-1's scalars into the peers' mailboxes (two system-scope fences over
            # NVLink) run on their own branch under step i's kernels instead of at its tail;
            # loss_finish, which overwrites losses4, waits for them
            if self._pub_stream is None:
                self._pub_stream = torch.cuda.Stream(self.dev)
            self._pub_stream.wait_event(fork)
            publisher.publish(self.losses4, self._pub_stream.cuda_stream)
            pub_done = torch.cuda.Event()
            pub_done.record(self._pub_stream)
        # enqueue the store stream last: the short latency-bound kernels of the other branches then
        # get their CTAs placed before loss_dense's ~15 400 one-tile CTAs start to stream (tools/timeline.py;
        # r1f: dense first 195.8 us per step, dense last 193.6)
        if not dense_last:
            _lib.check(self.lib.yh_loss_dense(self.ws.handle, hp, self.layers, self.L, self.B, self.C,
                                              self.cfg, self._grad_ptrs, main.cuda_stream))
        self._assign(gt_dev, off_dev, s_asg.cuda_stream)
        _lib.check(self.lib.yh_loss_rows_compute(
            self.ws.handle, hp, self.layers, self.L, self.B, self.C, self.m_batch.data_ptr(),
            self.m_flat.data_ptr(), self.m_target.data_ptr(), self.m_class.data_ptr(),
            self.M.data_ptr(), self.mcap, self.cfg, self.iou_score.data_ptr(),
            self.pred_rows.data_ptr(), 1 if self._grad_ptrs is not None else 0, s_asg.cuda_stream))
        self._nms(hp, s_nms.cuda_stream)
        self._decode(hp, s_dec.cuda_stream)
        # Above 16 images the dense stream is ordered behind decode.  It costs nothing — the stream
        # starts at ~40 us anyway, once the filter's CTAs (same priority, launched before it) have
        # all been handed out, and decode is done at ~31 — and it pins that order: the filter and
        # the dense pass are both root nodes of the graph, and WHICH of two equal-priority roots the
        # instantiated graph launches first is not ours to say.  The same step re-captured eight
        # times in one process: 158.3 / 163.0 us alternating without the edge, 158.1 / 159.6 with
        # it (profiles/r2g_recapture.txt); dense first is the 176 us schedule.
        if os.environ.get("YH_DENSE_AFTER_DECODE", "1" if self.B > 16 else "0") == "1":
            dec_done = torch.cuda.Event()
            dec_done.record(s_dec)
            main.wait_event(dec_done)
        if dense_last:
            _lib.check(self.lib.yh_loss_dense(self.ws.handle, hp, self.layers, self.L, self.B, self.C,
                                              self.cfg, self._grad_ptrs, main.cuda_stream))
        joins = []
        for st in self._side:
            e = torch.cuda.Event()
            e.record(st)
            joins.append(e)
        main.wait_event(joins[1])
        if pub_done is not None:
            main.wait_event(pub_done)
        # YH_FINISH_EARLY=1: the scatter launched into the drain of the dense pass (PDL), its rows
        # loaded before it waits (yh_loss_finish_after_dense).  Measured, not faster (r2e: 174.2 vs
        # 174.2-174.9 us per step; with a warp per row instead of one resident wave 191), so off.
        early = os.environ.get("YH_FINISH_EARLY", "0") == "1" and dense_last \
            and self._grad_ptrs is not None and pub_done is None
        fin = self.lib.yh_loss_finish_after_dense if early else self.lib.yh_loss_finish
        _lib.check(fin(
            self.ws.handle, self.layers, self.L, self.B, self.C, self.m_batch.data_ptr(),
            self.m_flat.data_ptr(), self.M.data_ptr(), self.mcap, self.cfg, self.losses4.data_ptr(),
            self._grad_ptrs, main.cuda_stream))
        if publisher is not None and not lagged_publish:
            publisher.publish(self.losses4, main.cuda_stream)
        main.wait_event(joins[0])
        main.wait_event(joins[2])

    # -- the same step with the loss cut into L2-sized image groups ----------------------------
    def run_grouped(self, heads, gt_dev, off_dev, publisher=None, group=None, dense_after=None):
        r"""run_overlapped with the loss in RANGES of `group` images (yh_loss_*_range):

            decode
            nms_filter -> nms_segment -> nms_emit
            assign ------------------\
            dense(g0) -> dense(g1) -> ...        (main stream, the store stream)
                 \-> rows(g0)  \-> rows(g1) ...  (side stream: rows of a group right behind its
                                                  dense pass, gradients stored straight into the
                                                  still-L2-resident lines)
            ... -> finish_ranges (fixed-order reduction of the partial sums)

        What this removes from the step: the staging buffer of the row gradients and the scatter
        kernel that ran alone at the tail, re-reading from DRAM every line it touched
        (loss_finish: 29 us, 18.6 MB read + 14.7 MB written for 4 MB of gradients, r1f)."""
        main = torch.cuda.current_stream(self.dev)
        if self._side is None:
            self._side = [torch.cuda.Stream(self.dev), torch.cuda.Stream(self.dev, priority=-1),
                          torch.cuda.Stream(self.dev, priority=-1)]
        if getattr(self, "_rows_stream", None) is None:
            self._rows_stream = torch.cuda.Stream(self.dev, priority=-1)
        s_dec, s_asg, s_nms = self._side
        s_rows = self._rows_stream
        group = group or int(os.environ.get("YH_GROUP", "8"))
        dense_after = dense_after or os.environ.get("YH_DENSE_AFTER", "assign")
        group = max(4, (group // 4) * 4)
        bounds = list(range(0, self.B, group)) + [self.B]
        if len(bounds) - 1 > 16:
            group = ((self.B + 15) // 16 + 3) // 4 * 4
            bounds = list(range(0, self.B, group)) + [self.B]
        R = len(bounds) - 1
        hp = _lib.ptr_array(heads)
        fork = torch.cuda.Event()
        fork.record(main)
        for st in self._side + [s_rows]:
            st.wait_event(fork)
        pub_done = None
        if publisher is not None:
            if self._pub_stream is None:
                self._pub_stream = torch.cuda.Stream(self.dev)
            self._pub_stream.wait_event(fork)
            publisher.publish(self.losses4, self._pub_stream.cuda_stream)
            pub_done = torch.cuda.Event()
            pub_done.record(self._pub_stream)
        self._assign(gt_dev, off_dev, s_asg.cuda_stream)
        asg_done = torch.cuda.Event()
        asg_done.record(s_asg)
        self._nms(hp, s_nms.cuda_stream)
        self._decode(hp, s_dec.cuda_stream)
        s_rows.wait_event(asg_done)
        if dense_after == "assign":
            main.wait_event(asg_done)
        img_off = self.lib.yh_assign_image_offsets(self.ws.handle)
        for r in range(R):
            lo, hi = bounds[r], bounds[r + 1]
            _lib.check(self.lib.yh_loss_dense_range(self.ws.handle, hp, self.layers, self.L, self.B, self.C,
                                                    lo, hi, r, R, self.cfg, self._grad_ptrs, main.cuda_stream))
            ev = torch.cuda.Event()
            ev.record(main)
            s_rows.wait_event(ev)
            _lib.check(self.lib.yh_loss_rows_range(
                self.ws.handle, hp, self.layers, self.L, self.B, self.C, lo, hi, r, R,
                self.m_batch.data_ptr(), self.m_flat.data_ptr(), self.m_target.data_ptr(),
                self.m_class.data_ptr(), self.M.data_ptr(), self.mcap, img_off, self.cfg,
                self.iou_score.data_ptr(), self.pred_rows.data_ptr(), self._grad_ptrs, s_rows.cuda_stream))
        joins = []
        for st in self._side + [s_rows]:
            e = torch.cuda.Event()
            e.record(st)
            joins.append(e)
        main.wait_event(joins[3])
        if pub_done is not None:
            main.wait_event(pub_done)
        _lib.check(self.lib.yh_loss_finish_ranges(self.ws.handle, self.layers, self.L, self.B, self.C, R,
                                                  self.M.data_ptr(), self.mcap, self.cfg,
                                                  self.losses4.data_ptr(), main.cuda_stream))
        for k in (0, 1, 2):
            main.wait_event(joins[k])

    # -- the same step with the dense pass and the gradient scatter cut by layer ----------------
    def run_layered(self, heads, gt_dev, off_dev, publisher=None):
        r"""run_overlapped with the loss cut by LAYER (yh_loss_dense_layers / yh_loss_finish_layers):

            decode
            nms_filter -> nms_segment -> nms_emit
            assign -> loss_rows (staged row gradients) ------+-----------+----------+
            dense(layer 0) -> dense(layer 1) -> dense(layer 2)           |          |
                     \-> finish(layer 0)   \-> finish(layer 1)   \-> finish(layer 2) + reduction

        In the whole-batch form the scatter of the matched rows runs alone at the tail of the step
        (27 us at 1.2 TB/s) and pays a DRAM read-modify-write for every line the 0.5 GB store
        stream has evicted.  71 % of the matched rows of a YOLOv4-608 batch lie in layers 1 and 2,
        whose gradient tensors (94 + 24 MB at 64 images) are written LAST and are still in the L2
        when their scatter runs right behind them; the scatter of layer 0 (29 % of the rows) runs
        under the dense passes of layers 1-2 instead of after everything."""
        main = torch.cuda.current_stream(self.dev)
        if self._side is None:
            self._side = self._side_streams()
        if getattr(self, "_fin_streams", None) is None:
            self._fin_streams = [torch.cuda.Stream(self.dev, priority=-1) for _ in range(max(1, self.L - 1))]
        s_dec, s_asg, s_nms = self._side
        hp = _lib.ptr_array(heads)
        fork = torch.cuda.Event()
        fork.record(main)
        for st in self._side + self._fin_streams:
            st.wait_event(fork)
        pub_done = None
        if publisher is not None:
            if self._pub_stream is None:
                self._pub_stream = torch.cuda.Stream(self.dev)
            self._pub_stream.wait_event(fork)
            publisher.publish(self.losses4, self._pub_stream.cuda_stream)
            pub_done = torch.cuda.Event()
            pub_done.record(self._pub_stream)
        self._assign(gt_dev, off_dev, s_asg.cuda_stream)
        _lib.check(self.lib.yh_loss_rows_compute(
            self.ws.handle, hp, self.layers, self.L, self.B, self.C, self.m_batch.data_ptr(),
            self.m_flat.data_ptr(), self.m_target.data_ptr(), self.m_class.data_ptr(),
            self.M.data_ptr(), self.mcap, self.cfg, self.iou_score.data_ptr(),
            self.pred_rows.data_ptr(), 1, s_asg.cuda_stream))
        rows_done = torch.cuda.Event()
        rows_done.record(s_asg)
        self._nms(hp, s_nms.cuda_stream)
        self._decode(hp, s_dec.cuda_stream)
        L = self.L
        fin_events = []
        for l in range(L):
            _lib.check(self.lib.yh_loss_dense_layers(self.ws.handle, hp, self.layers, L, self.B, self.C,
                                                     l, l + 1, l, L, self.cfg, self._grad_ptrs, main.cuda_stream))
            if l < L - 1:
                ev = torch.cuda.Event()
                ev.record(main)
                st = self._fin_streams[l]
                st.wait_event(ev)
                st.wait_event(rows_done)
                _lib.check(self.lib.yh_loss_finish_layers(
                    self.ws.handle, self.layers, L, self.B, self.C, l, l + 1, 0, L, self.m_batch.data_ptr(),
                    self.m_flat.data_ptr(), self.M.data_ptr(), self.mcap, self.cfg, self.losses4.data_ptr(),
                    self._grad_ptrs, st.cuda_stream))
                e2 = torch.cuda.Event()
                e2.record(st)
                fin_events.append(e2)
        main.wait_event(rows_done)
        if pub_done is not None:
            main.wait_event(pub_done)
        _lib.check(self.lib.yh_loss_finish_layers(
            self.ws.handle, self.layers, L, self.B, self.C, L - 1, L, 1, L, self.m_batch.data_ptr(),
            self.m_flat.data_ptr(), self.M.data_ptr(), self.mcap, self.cfg, self.losses4.data_ptr(),
            self._grad_ptrs, main.cuda_stream))
        for e in fin_events:
            main.wait_event(e)
        for st in (s_dec, s_nms):
            e = torch.cuda.Event()
            e.record(st)
            main.wait_event(e)

    def _step(self, heads, gt_dev, off_dev, publisher=None):
        if self.schedule == "grouped" and self._grad_ptrs is not None and self.B >= 8:
            self.run_grouped(heads, gt_dev, off_dev, publisher=publisher)
        elif self.schedule == "layered" and self._grad_ptrs is not None and self.L >= 2:
            self.run_layered(heads, gt_dev, off_dev, publisher=publisher)
        else:
            self.run_overlapped(heads, gt_dev, off_dev, publisher=publisher)

    def capture(self, heads, gt_dev, off_dev, loss_slots: int = 1, publisher=None):
        """Captures `run_overlapped` on these buffers into a CUDA graph; `replay()` launches it.
        The captured pointers are those of `heads`/`gt_dev`/`off_dev`: refill them in place.
        `loss_slots` > 1 captures that many copies of the graph that differ only in the
        `losses4` buffer they write and replays them round-robin, so that a consumer on another
        stream (the data-parallel loss all-reduce, dist.LossAllReduce) can still read step i's
        scalars while step i+1 runs.
        `publisher` (dist.PeerLossExchange): its 32-thread kernel becomes a node of the graph
        and writes this rank's weighted loss scalars into every peer's mailbox over NVLink — the
        data-parallel loss exchange with no collective and no host work per step.  It runs ONE
        STEP BEHIND (replay i publishes the scalars of replay i-1, under replay i's kernels):
        call `publisher.publish(step.losses4)` once after the last replay to flush."""
        self._keep = (list(heads), gt_dev, off_dev)
        self._step(heads, gt_dev, off_dev)  # warm: module attributes, streams
        torch.cuda.synchronize(self.dev)
        # [total, iou, cls, obj, 1] per slot; `weighted5` = that times the local batch size, the
        # operand of the data-parallel weighted mean (multi_gpu.rs:188-195), produced in-graph
        self._l5_ring = [torch.zeros(5, dtype=torch.float32, device=self.dev) for _ in range(loss_slots)]
        self._w5_ring = [torch.zeros(5, dtype=torch.float32, device=self.dev) for _ in range(loss_slots)]
        for l5 in self._l5_ring:
            l5[4] = 1.0
        self._loss_ring = [l5[:4] for l5 in self._l5_ring]
        self._graphs = []
        cap = torch.cuda.Stream(self.dev, priority=int(os.environ.get("YH_MAIN_PRIO", "0")))
        for k in range(loss_slots):
            self.losses4 = self._loss_ring[k]
            g = torch.cuda.CUDAGraph()
            before = int(self.lib.yh_launch_count(0))
            with torch.cuda.graph(g, stream=cap):
                self._step(heads, gt_dev, off_dev, publisher=publisher)
                if publisher is None and loss_slots > 1:
                    torch.mul(self._l5_ring[k], float(self.B), out=self._w5_ring[k])
            self.kernels_in_graph = int(self.lib.yh_launch_count(0)) - before
            self._graphs.append(g)
        self._replays = 0
        self._graph = self._graphs[0]
        self.losses4 = self._loss_ring[0]
        self.weighted5 = self._w5_ring[0]
        return self._graph

    def replay(self):
        """Launches the captured step; afterwards `self.losses4` is the buffer this replay writes."""
        k = self._replays % len(self._graphs)
        self._replays += 1
        self.losses4 = self._loss_ring[k]
        self.weighted5 = self._w5_ring[k]
        self._graphs[k].replay()

    def counts(self):
        """synchronising read of (M, K); raises on overflow."""
        return self.ws.read_counts(torch.cuda.current_stream(self.dev).cuda_stream, self.M, self.K)


def _check_flag(h_counts, what):
    f = int(h_counts[2])
    if f:
        raise _lib.YoloHotError(
            _lib.YH_ERR_INVALID if f & 8 else _lib.YH_ERR_CAPACITY,
            f"{what}: the step's sticky flag is {f} (1=matches, 2=nms candidates, 4=nms outputs overflowed; "
            "8=match list not sorted): outputs are truncated — enlarge the capacities")


class HostStep:
    """The step for HOST-resident inputs and outputs.  Per call: H2D of the L head tensors and
    the packed labels from pinned memory, DeviceStep, D2H of [losses4, M, K], the match list, the
    kept detections and (if `with_grad`) the L head-gradient tensors into pinned memory."""

    def __init__(self, spec, batch, device, with_grad=True, **kw):
        self.step = DeviceStep(spec, batch, device, with_grad=with_grad, **kw)
        st = self.step
        self.d_heads = [torch.empty(spec.head_shape(l, batch), dtype=torch.float32, device=device)
                        for l in range(st.L)]
        pin = dict(pin_memory=True)
        self.h_losses = torch.empty(4, dtype=torch.float32, **pin)
        self.h_counts = torch.empty(3, dtype=torch.int32, **pin)   # M, K, sticky overflow flag
        self.h_counts.zero_()
        # the workspace's sticky flag travels out with the results (yh_copy_error_flag_async in
        # copy_out): a replayed step must not truncate matches / candidates / detections silently
        self.h_match = torch.empty((2, st.mcap), dtype=torch.int64, **pin)
        self.h_det = torch.empty((3, st.kcap), dtype=torch.int64, **pin)
        self.h_tlbr = torch.empty((st.kcap, 4), dtype=torch.float32, **pin)
        self.h_conf = torch.empty(st.kcap, dtype=torch.float32, **pin)
        self.h_grads = [torch.empty(spec.head_shape(l, batch), dtype=torch.float32, **pin)
                        for l in range(st.L)] if with_grad else None
        self.max_gt = st.mcap // (sum(i.A for i in st.info) * (5 if st.rect4 else 3))
        self.d_gt = torch.empty((self.max_gt, 5), dtype=torch.float64, device=device)
        self.d_off = torch.empty(batch + 1, dtype=torch.int32, device=device)
        self.h2d_bytes = 0
        self.d2h_bytes = 0

    def copy_out(self):
        """D2H of every result of the step on the current stream; returns the byte count."""
        st = self.step
        # results: scalars first (small), then the bulk
        self.h_losses.copy_(st.losses4, non_blocking=True)
        self.h_counts[0:1].copy_(st.M, non_blocking=True)
        self.h_counts[1:2].copy_(st.K, non_blocking=True)
        _lib.check(st.lib.yh_copy_error_flag_async(st.ws.handle, self.h_counts[2:3].data_ptr(),
                                                   torch.cuda.current_stream(st.dev).cuda_stream))
        self.h_match[0].copy_(st.m_batch, non_blocking=True)
        self.h_match[1].copy_(st.m_flat, non_blocking=True)
        self.h_det[0].copy_(st.kb, non_blocking=True)
        self.h_det[1].copy_(st.kc, non_blocking=True)
        self.h_det[2].copy_(st.ki, non_blocking=True)
        self.h_tlbr.copy_(st.tlbr, non_blocking=True)
        self.h_conf.copy_(st.conf, non_blocking=True)
        d2h = 16 + 12 + self.h_match.numel() * 8 + self.h_det.numel() * 8 + self.h_tlbr.numel() * 4 \
            + self.h_conf.numel() * 4
        if self.h_grads is not None:
            for h, d in zip(self.h_grads, st.grads):
                h.copy_(d, non_blocking=True)
                d2h += d.numel() * 4
        return d2h

    def run(self, host_heads: Sequence[torch.Tensor], gt_host: torch.Tensor, off_host: torch.Tensor):
        """host_heads / gt_host / off_host: pinned host tensors.  Returns (losses4, M, K) on the
        host after one stream synchronisation; detections / matches / gradients are in the
        pinned h_* buffers."""
        st = self.step
        G = gt_host.shape[0]
        if G > self.max_gt:
            raise ValueError(f"{G} labels exceed the capacity {self.max_gt}")
        h2d = 0
        for d, h in zip(self.d_heads, host_heads):
            d.copy_(h, non_blocking=True)
            h2d += h.numel() * 4
        self.d_gt[:G].copy_(gt_host, non_blocking=True)
        self.d_off.copy_(off_host, non_blocking=True)
        h2d += G * 40 + off_host.numel() * 4
        st.run(self.d_heads, self.d_gt[:G], self.d_off)
        d2h = self.copy_out()
        torch.cuda.current_stream(st.dev).synchronize()
        self.h2d_bytes, self.d2h_bytes = h2d, d2h
        _check_flag(self.h_counts, "HostStep.run")
        return self.h_losses, int(self.h_counts[0]), int(self.h_counts[1])


class HostPipeline:
    """HostStep with the transfers overlapped: `depth` slots, each with its own device buffers,
    pinned result buffers and captured step graph; three streams (copy-in, compute, copy-out)
    chained by events, so that while step i computes, step i+1's heads stream in over PCIe and
    step i-1's gradients stream out (PCIe is full duplex; the single-stream HostStep serialises
    the two directions).  Every step still pays its own H2D of the inputs and D2H of all results.

        pipe.submit(host_heads, gt_host, off_host)   -> slot index (enqueues, never blocks unless
                                                        the slot's previous results are unread)
        pipe.result(slot)                            -> (losses4, M, K) after that step finished
    """

    def __init__(self, spec, batch, device, depth=2, with_grad=True, **kw):
        self.dev, self.depth = device, depth
        self.slots = [HostStep(spec, batch, device, with_grad=with_grad, **kw) for _ in range(depth)]
        self.s_in = torch.cuda.Stream(device)
        self.s_run = torch.cuda.Stream(device)
        self.s_out = torch.cuda.Stream(device)
        self.ev_in = [torch.cuda.Event() for _ in range(depth)]
        self.ev_run = [torch.cuda.Event() for _ in range(depth)]
        self.ev_out = [torch.cuda.Event() for _ in range(depth)]
        self.used = [False] * depth
        self.i = 0
        self.h2d_bytes = self.d2h_bytes = 0
        # capture each slot's step graph on its own buffers
        for hs in self.slots:
            st = hs.step
            hs.d_gt.zero_()
            hs.d_off.zero_()
            for d in hs.d_heads:
                d.zero_()
            with torch.cuda.stream(self.s_run):
                st.capture(hs.d_heads, hs.d_gt, hs.d_off)
        torch.cuda.synchronize(device)

    def submit(self, host_heads, gt_host, off_host):
        k = self.i % self.depth
        self.i += 1
        hs = self.slots[k]
        st = hs.step
        G = gt_host.shape[0]
        if G > hs.max_gt:
            raise ValueError(f"{G} labels exceed the capacity {hs.max_gt}")
        if self.used[k]:
            self.ev_out[k].synchronize()  # the slot's previous results must have left the device
        # copy-in: device input buffers of the slot are free once its previous compute finished
        with torch.cuda.stream(self.s_in):
            if self.used[k]:
                self.s_in.wait_event(self.ev_run[k])
            h2d = 0
            for d, h in zip(hs.d_heads, host_heads):
                d.copy_(h, non_blocking=True)
                h2d += h.numel() * 4
            hs.d_gt[:G].copy_(gt_host, non_blocking=True)
            hs.d_off.copy_(off_host, non_blocking=True)
            h2d += G * 40 + off_host.numel() * 4
            self.ev_in[k].record(self.s_in)
        with torch.cuda.stream(self.s_run):
            self.s_run.wait_event(self.ev_in[k])
            if self.used[k]:
                self.s_run.wait_event(self.ev_out[k])  # results of the previous use were read out
            st.replay()
            self.ev_run[k].record(self.s_run)
        with torch.cuda.stream(self.s_out):
            self.s_out.wait_event(self.ev_run[k])
            d2h = hs.copy_out()
            self.ev_out[k].record(self.s_out)
        self.used[k] = True
        self.h2d_bytes, self.d2h_bytes = h2d, d2h
        return k

    def result(self, k):
        self.ev_out[k].synchronize()
        hs = self.slots[k]
        _check_flag(hs.h_counts, "HostPipeline.result")
        return hs.h_losses, int(hs.h_counts[0]), int(hs.h_counts[1])

    def drain(self):
        for k in range(self.depth):
            if self.used[k]:
                self.ev_out[k].synchronize()
