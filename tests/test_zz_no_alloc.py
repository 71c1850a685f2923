"""yh_average_precision_static — the entry point that allocates nothing (include/yolohot.h).

No counterpart in the reference: it exists for callers that cannot call cudaMalloc.  Its scratch is
a 4 MiB `static __device__` array inside libyolohot.so, shared by every call, so the checks are
  * same bits as yh_average_precision (caller scratch) through the C ABI, on two calls in a row
    whose problems differ in size (the second call sees the first call's leftovers in the array);
  * sizes beyond the array are refused with YH_ERR_CAPACITY and a message naming both sizes.
AP values are f64 computed by the same kernel in the same order: the tolerance is equality, as in
tests/test_gpu_eval.py.  Collected last on purpose.
"""
import numpy as np
import pytest
import torch

from yolo_dl_b200 import _lib

pytestmark = pytest.mark.gpu

AP_STATIC_BYTES = 4 << 20   # YH_AP_STATIC_BYTES, include/yolohot.h
COCO_THRESHOLDS = [0.5 + 0.05 * i for i in range(10)]


def _problems(dev, seed, sizes):
    """random detections of len(sizes) problems: (det_gt i32, det_iou f64, det_conf f64, offsets, num_gt)."""
    rng = np.random.default_rng(seed)
    gt, iou, conf, offs, ngt = [], [], [], [0], []
    for nd in sizes:
        g = int(rng.integers(5, 60))
        s = rng.integers(-1, g, nd)
        gt.append(s.astype(np.int32))
        iou.append(np.where(s < 0, 0.0, np.round(rng.uniform(0.05, 1.0, nd), 2)))
        conf.append(np.round(rng.uniform(0.0, 1.0, nd), 3))   # many equal confidences
        offs.append(offs[-1] + nd)
        ngt.append(g + int(rng.integers(0, 5)))
    cat = lambda parts, dt: torch.tensor(np.concatenate(parts), dtype=dt, device=dev)
    return (cat(gt, torch.int32), cat(iou, torch.float64), cat(conf, torch.float64),
            torch.tensor(offs, dtype=torch.int32, device=dev), torch.tensor(ngt, dtype=torch.int32, device=dev))


def _both(lib, dev, seed, sizes, n_points=101):
    """(AP from yh_average_precision with caller scratch, AP from yh_average_precision_static)."""
    gt, iou, conf, offs, ngt = _problems(dev, seed, sizes)
    P, dmax, T = len(sizes), max(sizes), len(COCO_THRESHOLDS)
    thr = torch.tensor(COCO_THRESHOLDS, dtype=torch.float64, device=dev)
    need = int(lib.yh_average_precision_scratch_bytes(P, dmax))
    scratch = torch.empty(max(need, 16), dtype=torch.uint8, device=dev)
    want = torch.full((P, T), -1.0, dtype=torch.float64, device=dev)
    got = torch.full((P, T), -2.0, dtype=torch.float64, device=dev)
    torch.cuda.synchronize(dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    _lib.check(lib.yh_average_precision(gt.data_ptr(), iou.data_ptr(), conf.data_ptr(), offs.data_ptr(),
                                        ngt.data_ptr(), P, dmax, thr.data_ptr(), T, n_points,
                                        want.data_ptr(), scratch.data_ptr() if need else None, need, stream))
    # everything the static call touches exists by now: it gets pointers and sizes only
    _lib.check(lib.yh_average_precision_static(gt.data_ptr(), iou.data_ptr(), conf.data_ptr(),
                                               offs.data_ptr(), ngt.data_ptr(), P, dmax, thr.data_ptr(),
                                               T, n_points, got.data_ptr(), stream))
    torch.cuda.synchronize(dev)
    return need, want.cpu(), got.cpu()


def test_static_scratch_equals_caller_scratch_two_calls_in_a_row(dev):
    lib = _lib.load()
    # call 1: three problems, the largest of 40 000 detections (P = 65 536 slots per problem)
    need1, want1, got1 = _both(lib, dev, 11, [40000, 9000, 700])
    assert 0 < need1 <= AP_STATIC_BYTES
    assert torch.equal(got1, want1), (got1 - want1).abs().max()
    assert float(want1.min()) >= 0.0 and float(want1.max()) <= 1.0 and float(want1.max()) > 0.0
    # call 2, straight after: other data, other slice sizes (P = 16 384), the array still holds call 1
    need2, want2, got2 = _both(lib, dev, 12, [8193, 12000, 3, 16000], n_points=11)
    assert 0 < need2 <= AP_STATIC_BYTES and need2 != need1
    assert torch.equal(got2, want2), (got2 - want2).abs().max()
    # call 3: the first inputs again give the first answer again
    _, want3, got3 = _both(lib, dev, 11, [40000, 9000, 700])
    assert torch.equal(got3, got1) and torch.equal(want3, want1)


def test_static_small_problems_do_not_need_the_array(dev):
    lib = _lib.load()
    need, want, got = _both(lib, dev, 13, [8192, 1, 500, 77])
    assert need == 0
    assert torch.equal(got, want)


def test_static_rejects_sizes_beyond_the_array(dev):
    lib = _lib.load()
    # the largest single problem that fits, and the first that does not
    assert lib.yh_average_precision_scratch_bytes(1, 131072) <= AP_STATIC_BYTES
    assert lib.yh_average_precision_scratch_bytes(1, 131073) > AP_STATIC_BYTES
    assert lib.yh_average_precision_scratch_bytes(4, 65536) > AP_STATIC_BYTES
    one = torch.zeros(1, dtype=torch.float64, device=dev)
    i32 = torch.zeros(2, dtype=torch.int32, device=dev)
    out = torch.full((4,), 7.0, dtype=torch.float64, device=dev)
    for P, dmax in ((1, 131073), (4, 65536)):
        # refused on the host from the sizes alone: no pointer is dereferenced, no kernel launched
        rc = lib.yh_average_precision_static(i32.data_ptr(), one.data_ptr(), one.data_ptr(), i32.data_ptr(),
                                             i32.data_ptr(), P, dmax, one.data_ptr(), 1, 101,
                                             out.data_ptr(), None)
        msg = lib.yh_last_error()
        assert rc == _lib.YH_ERR_CAPACITY, (rc, msg)
        assert b"static" in msg and str(AP_STATIC_BYTES).encode() in msg, msg
    torch.cuda.synchronize(dev)
    assert out.tolist() == [7.0] * 4
    # the same validation as the allocating entry point
    rc = lib.yh_average_precision_static(i32.data_ptr(), one.data_ptr(), one.data_ptr(), i32.data_ptr(),
                                         i32.data_ptr(), 1, 10, one.data_ptr(), 1, 1, out.data_ptr(), None)
    assert rc == _lib.YH_ERR_INVALID and b"interpolated points" in lib.yh_last_error()
