"""ctypes binding of libyolohot.so (the C ABI of include/yolohot.h).

There is NO CPU fallback: if the shared library is missing, `load()` raises; if there is no CUDA
device, `yh_workspace_create` fails and the error text of `yh_last_error()` is raised.
"""
import ctypes
import os
from ctypes import POINTER, c_char_p, c_double, c_float, c_int, c_int32, c_int64, c_void_p

YH_MAX_LAYERS = 8
YH_MAX_ANCHORS = 8

YH_OK, YH_ERR_INVALID, YH_ERR_CUDA, YH_ERR_CAPACITY, YH_ERR_UNSUPPORTED = range(5)

METRICS = {"iou": 0, "giou": 1, "diou": 2, "ciou": 3, "hausdorff": 4}
REDUCTIONS = {"mean": 0, "sum": 1}
LOSS_KINDS = {"bce": 0, "focal": 1, "l2": 2, "crossentropy": 3}


class YhLayer(ctypes.Structure):
    _fields_ = [("H", c_int32), ("W", c_int32), ("A", c_int32), ("_pad", c_int32),
                ("flat_start", c_int64),
                ("anchor_h", c_double * YH_MAX_ANCHORS), ("anchor_w", c_double * YH_MAX_ANCHORS)]


class YhLossCfg(ctypes.Structure):
    _fields_ = [("box_metric", c_int32), ("reduction", c_int32),
                ("iou_loss_weight", c_double), ("classification_loss_weight", c_double),
                ("objectness_loss_weight", c_double), ("smooth_classification_coef", c_double),
                ("smooth_objectness_coef", c_double), ("has_objectness_pos_weight", c_int32),
                ("_pad", c_int32), ("objectness_pos_weight", c_double),
                ("objectness_loss_kind", c_int32), ("classification_loss_kind", c_int32),
                ("focal_loss_gamma", c_double)]


class YoloHotError(RuntimeError):
    def __init__(self, status, text):
        super().__init__(f"yolohot status {status}: {text}")
        self.status = status
        self.text = text


LIB_PATH = os.environ.get("YH_LIB") or os.path.join(os.path.dirname(os.path.abspath(__file__)), "libyolohot.so")
# (YH_LIB: A/B of two builds of the library on the same box, tools/ab_libs.sh)

# every symbol include/yolohot.h declares: name -> (restype, argtypes)
_PF = POINTER(c_void_p)
SYMBOLS = {
    "yh_workspace_create": (c_int, [c_int, c_int, c_int64, c_int, POINTER(c_void_p)]),
    "yh_workspace_destroy": (c_int, [c_void_p]),
    "yh_last_error": (c_char_p, []),
    "yh_version": (c_char_p, []),
    "yh_decode": (c_int, [c_void_p, c_int, c_int, POINTER(YhLayer), c_void_p, c_void_p, c_void_p,
                          c_void_p, c_int64, c_void_p]),
    "yh_decode_merged": (c_int, [_PF, POINTER(YhLayer), c_int, c_int, c_int, c_void_p, c_void_p,
                                 c_void_p, c_void_p, c_int64, c_void_p]),
    "yh_assign": (c_int, [c_void_p, c_void_p, c_void_p, c_int, POINTER(YhLayer), c_int, c_int,
                          c_double, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                          c_int, c_void_p]),
    "yh_loss_fwd_bwd": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_void_p,
                                c_void_p, c_void_p, c_void_p, c_void_p, c_int, POINTER(YhLossCfg),
                                c_void_p, c_void_p, c_void_p, _PF, c_void_p]),
    "yh_loss_dense": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, POINTER(YhLossCfg),
                              _PF, c_void_p]),
    "yh_loss_rows": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_void_p,
                             c_void_p, c_void_p, c_void_p, c_void_p, c_int, POINTER(YhLossCfg),
                             c_void_p, c_void_p, c_void_p, _PF, c_void_p]),
    "yh_loss_rows_compute": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_void_p,
                                     c_void_p, c_void_p, c_void_p, c_void_p, c_int, POINTER(YhLossCfg),
                                     c_void_p, c_void_p, c_int, c_void_p]),
    "yh_loss_finish": (c_int, [c_void_p, POINTER(YhLayer), c_int, c_int, c_int, c_void_p, c_void_p,
                               c_void_p, c_int, POINTER(YhLossCfg), c_void_p, _PF, c_void_p]),
    "yh_loss_finish_after_dense": (c_int, [c_void_p, POINTER(YhLayer), c_int, c_int, c_int, c_void_p, c_void_p,
                                           c_void_p, c_int, POINTER(YhLossCfg), c_void_p, _PF, c_void_p]),
    "yh_workspace_reserve_rows": (c_int, [c_void_p, c_int, c_int]),
    "yh_loss_dense_range": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_int, c_int,
                                    c_int, c_int, POINTER(YhLossCfg), _PF, c_void_p]),
    "yh_loss_rows_range": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_int, c_int,
                                   c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                                   c_void_p, POINTER(YhLossCfg), c_void_p, c_void_p, _PF, c_void_p]),
    "yh_loss_finish_ranges": (c_int, [c_void_p, POINTER(YhLayer), c_int, c_int, c_int, c_int, c_void_p,
                                      c_int, POINTER(YhLossCfg), c_void_p, c_void_p]),
    "yh_loss_dense_layers": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_int, c_int,
                                     c_int, c_int, POINTER(YhLossCfg), _PF, c_void_p]),
    "yh_loss_finish_layers": (c_int, [c_void_p, POINTER(YhLayer), c_int, c_int, c_int, c_int, c_int, c_int,
                                      c_int, c_void_p, c_void_p, c_void_p, c_int, POINTER(YhLossCfg),
                                      c_void_p, _PF, c_void_p]),
    "yh_assign_image_offsets": (c_void_p, [c_void_p]),
    "yh_nms": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_double, c_double,
                       c_int, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                       c_void_p]),
    "yh_nms_by_scores": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_double, c_void_p,
                                 c_void_p, c_void_p]),
    "yh_inference_pick": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p,
                                  c_void_p, c_int, c_void_p]),
    "yh_benchmark": (c_int, [c_void_p, _PF, POINTER(YhLayer), c_int, c_int, c_int, c_void_p, c_void_p,
                             c_void_p, c_int, c_double, c_void_p, c_void_p]),
    "yh_match_det_gt": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p, c_void_p,
                                c_void_p]),
    "yh_average_precision": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                     c_void_p, c_int, c_int, c_void_p, c_void_p, ctypes.c_size_t, c_void_p]),
    "yh_average_precision_scratch_bytes": (ctypes.c_size_t, [c_int, c_int]),
    "yh_average_precision_static": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int,
                                            c_void_p, c_int, c_int, c_void_p, c_void_p]),
    "yh_peer_mailbox_create": (c_int, [c_int, c_int, c_int, POINTER(c_void_p), POINTER(c_void_p),
                                       c_void_p]),
    "yh_peer_mailbox_open": (c_int, [c_int, c_void_p, POINTER(c_void_p)]),
    "yh_peer_mailbox_close": (c_int, [c_int, c_void_p]),
    "yh_peer_mailbox_destroy": (c_int, [c_void_p]),
    "yh_loss_publish": (c_int, [c_void_p, c_void_p, c_float, POINTER(c_void_p), c_int, c_void_p]),
    "yh_loss_collect": (c_int, [c_void_p, POINTER(c_float), POINTER(ctypes.c_uint32), c_void_p]),
    "yh_read_counts": (c_int, [c_void_p, c_void_p, POINTER(c_int32), c_void_p, POINTER(c_int32),
                               c_void_p]),
    "yh_workspace_candidates_needed": (c_int, [c_void_p]),
    "yh_workspace_error_flag": (c_void_p, [c_void_p]),
    "yh_copy_error_flag_async": (c_int, [c_void_p, c_void_p, c_void_p]),
    "yh_launch_count": (c_int64, [c_int]),
    "yh_timing_enable": (c_int, [c_int]),
    "yh_trace_enable": (c_int, [c_void_p, c_int]),
    "yh_trace_ids": (c_int, [POINTER(c_int32), c_int]),
    "yh_timing_read": (c_int, [POINTER(c_double), POINTER(c_int64), c_int]),
}

KERNEL_NAMES = ["decode_kernel", "assign_slots_kernel", "assign_emit_kernel", "loss_dense_kernel",
                "loss_rows_kernel", "nms_filter_kernel", "nms_segment_kernel", "nms_emit_kernel",
                "nms_pack_kernel", "pick_kernel", "benchmark_kernel",
                "match_det_gt_kernel", "ap_kernel", "loss_finish_kernel", "loss_publish_kernel",
                "nms_big_kernel"]

_lib = None


def load():
    """Loads libyolohot.so and types every entry point.  Raises if it is missing."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; "
            "g.build()'` (or `make -C yolo_dl_b200/csrc`). yolohot has no CPU fallback.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(status):
    if status != YH_OK:
        raise YoloHotError(status, load().yh_last_error().decode())


def make_layers(infos):
    """infos: sequence with .H .W .anchors [(h,w) ratio f64] .flat_start -> (YhLayer * L)."""
    arr = (YhLayer * len(infos))()
    for i, info in enumerate(infos):
        if len(info.anchors) > YH_MAX_ANCHORS:
            raise ValueError(f"layer {i}: {len(info.anchors)} anchors > {YH_MAX_ANCHORS}")
        arr[i].H, arr[i].W, arr[i].A = info.H, info.W, len(info.anchors)
        arr[i].flat_start = info.flat_start
        for a, (h, w) in enumerate(info.anchors):
            arr[i].anchor_h[a] = float(h)
            arr[i].anchor_w[a] = float(w)
    return arr


def ptr_array(tensors):
    """device pointers of a list of tensors as a `const float* const*` argument."""
    arr = (c_void_p * len(tensors))()
    for i, t in enumerate(tensors):
        arr[i] = t.data_ptr()
    return arr
