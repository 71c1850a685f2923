"""Known-answer test of the DenseDetectionTensor container operations, taken from the reference's
own `cat_test` (tch-goodies/src/detection/dense_detection_tensor.rs:582-728): same input (20 x 32
map, 80 classes, one 1x1 anchor), same EXACT f32 expectations (`==`, not a tolerance) for
cat_batch / cat_height / cat_width / slice_ratio.  Runs on the CPU: these are host-side tensor
glue, not kernels."""
import numpy as np
import pytest
import torch

from yolo_dl_b200.tensors import DenseDetectionTensor

HEIGHT, WIDTH, NUM_CLASSES = 20, 32, 80
f32 = np.float32


def _input():
    rows = torch.arange(HEIGHT, dtype=torch.float32).view(HEIGHT, 1).expand(HEIGHT, WIDTH)
    cols = torch.arange(WIDTH, dtype=torch.float32).view(1, WIDTH).expand(HEIGHT, WIDTH)
    shape = (1, 1, 1, HEIGHT, WIDTH)
    cy = (rows / f32(HEIGHT)).reshape(shape).contiguous()       # row as f32 / HEIGHT as f32
    cx = (cols / f32(WIDTH)).reshape(shape).contiguous()
    h = torch.ones(shape)
    w = torch.ones(shape)
    obj = torch.full(shape, 0.5)
    cls = torch.full((1, NUM_CLASSES, 1, HEIGHT, WIDTH), 0.1)
    return DenseDetectionTensor(cy, cx, h, w, obj, cls, [(1.0, 1.0)]).check()


def _expect(rows, cols, fn):
    """fn(row, col) evaluated in f32 like the reference's `as f32` arithmetic"""
    out = np.empty((rows, cols), dtype=np.float32)
    for r in range(rows):
        for c in range(cols):
            out[r, c] = fn(f32(r), f32(c))
    return out


def test_cat_batch():
    x = _input()
    out = DenseDetectionTensor.cat_batch([x, x])
    assert tuple(out.cy.shape) == (2, 1, 1, HEIGHT, WIDTH)          # :653-656
    assert tuple(out.class_logit.shape) == (2, NUM_CLASSES, 1, HEIGHT, WIDTH)
    assert out.anchors == [(1.0, 1.0)]


def test_cat_height_exact():
    x = _input()
    out = DenseDetectionTensor.cat_height([x, x]).check()            # :658-682
    cy, cx, h, w = (t[0, 0, 0].numpy() for t in (out.cy, out.cx, out.h, out.w))
    assert cy.shape == (2 * HEIGHT, WIDTH)
    top = _expect(HEIGHT, WIDTH, lambda r, c: r / f32(HEIGHT) / f32(2.0))
    bot = _expect(HEIGHT, WIDTH, lambda r, c: r / f32(HEIGHT) / f32(2.0) + f32(0.5))
    ex = _expect(HEIGHT, WIDTH, lambda r, c: c / f32(WIDTH))
    assert np.array_equal(cy[:HEIGHT], top) and np.array_equal(cy[HEIGHT:], bot)
    assert np.array_equal(cx[:HEIGHT], ex) and np.array_equal(cx[HEIGHT:], ex)
    assert (h == f32(0.5)).all() and (w == f32(1.0)).all()
    assert out.anchors == [(0.5, 1.0)]
    assert tuple(out.obj_logit.shape) == (1, 1, 1, 2 * HEIGHT, WIDTH)


def test_cat_width_exact():
    x = _input()
    out = DenseDetectionTensor.cat_width([x, x]).check()             # :684-708
    cy, cx, h, w = (t[0, 0, 0].numpy() for t in (out.cy, out.cx, out.h, out.w))
    assert cx.shape == (HEIGHT, 2 * WIDTH)
    ey = _expect(HEIGHT, WIDTH, lambda r, c: r / f32(HEIGHT))
    left = _expect(HEIGHT, WIDTH, lambda r, c: c / f32(WIDTH) / f32(2.0))
    right = _expect(HEIGHT, WIDTH, lambda r, c: c / f32(WIDTH) / f32(2.0) + f32(0.5))
    assert np.array_equal(cy[:, :WIDTH], ey) and np.array_equal(cy[:, WIDTH:], ey)
    assert np.array_equal(cx[:, :WIDTH], left) and np.array_equal(cx[:, WIDTH:], right)
    assert (h == f32(1.0)).all() and (w == f32(0.5)).all()
    assert out.anchors == [(1.0, 0.5)]


def test_slice_ratio_exact():
    x = _input()
    out = x.slice_ratio((0.4, 0.6), (0.5, 1.0))                     # :710-727
    NEW_H, NEW_W = HEIGHT // 5, WIDTH // 2
    cy, cx, h, w = (t[0, 0, 0].numpy() for t in (out.cy, out.cx, out.h, out.w))
    assert cy.shape == (NEW_H, NEW_W)
    assert np.array_equal(cy, _expect(NEW_H, NEW_W, lambda r, c: r / f32(NEW_H)))
    assert np.array_equal(cx, _expect(NEW_H, NEW_W, lambda r, c: c / f32(NEW_W)))
    assert (h == f32(5.0)).all() and (w == f32(2.0)).all()
    assert out.anchors == [(5.0, 2.0)]
    assert tuple(out.class_logit.shape) == (1, NUM_CLASSES, 1, NEW_H, NEW_W)


def test_validation_like_the_reference():
    x = _input()
    with pytest.raises(ValueError):
        x.slice_ratio((0.5, 0.5), (0.0, 1.0))                        # empty range (:111-118)
    with pytest.raises(ValueError):
        x.slice((0, HEIGHT + 1), (0, WIDTH))                         # :137-144
    y = DenseDetectionTensor.cat_height([x, x])
    with pytest.raises(ValueError):
        DenseDetectionTensor.cat_batch([x, y])                       # feature_h_set.len() == 1
    with pytest.raises(ValueError):
        x.index_select_batch(torch.tensor([1]))                      # :77-78
    sel = DenseDetectionTensor.cat_batch([x, x]).index_select_batch(torch.tensor([1, 0, 1]))
    assert sel.batch_size() == 3
    # unbiased_cy / unbiased_cx remove the cell offsets (:55-71): this input is offsets only
    assert float(x.unbiased_cy().abs().max()) == 0.0 and float(x.unbiased_cx().abs().max()) == 0.0
    assert tuple(x.to_tensor().shape) == (1, 5 + NUM_CLASSES, 1, HEIGHT, WIDTH)
    assert torch.equal(x.confidence(), x.obj_logit.sigmoid() * x.class_logit.sigmoid())
