"""shared helpers of the parity tests"""
import numpy as np
import torch

from oracle import detect as odet
from yolo_dl_b200 import synth
import yolo_dl_b200 as yh

# Tolerances (BASELINE.json north_star: "within 1e-5 relative in fp32").
# Decoded boxes and loss scalars: |a-b| <= 1e-5*|b| + 1e-7.
# Gradients: |a-b| <= 1e-5*|b| + 4e-6*max|b|.  Element-wise relative error is unbounded where a
# gradient cancels to ~0 — the reference's own two executors (ATen-CPU, ATen-CUDA) differ by
# 2.5e-4 relative on elements >= 1e-3*max and by 2.2e-6*max|b| absolutely (printed by
# tests/test_gpu_aten.py); this library is within 2.3e-6*max|b| of both, the bound is 4e-6.
RTOL = 1e-5
ATOL = 1e-7
GRAD_ATOL_SCALE = 4e-6


def close(a, b, rtol=RTOL, atol=ATOL, what=""):
    a = torch.as_tensor(a).detach().cpu().double()
    b = torch.as_tensor(b).detach().cpu().double()
    assert a.shape == b.shape, f"{what}: shape {tuple(a.shape)} vs {tuple(b.shape)}"
    err = (a - b).abs()
    bound = rtol * b.abs() + atol
    bad = err > bound
    if bad.any():
        i = int((err - bound).argmax())
        raise AssertionError(f"{what}: {int(bad.sum())}/{a.numel()} outside tolerance; worst at {i}: "
                             f"{a.flatten()[i].item()!r} vs {b.flatten()[i].item()!r}")


def grad_close(a, b, what=""):
    b_ = torch.as_tensor(b).detach().cpu()
    close(a, b_, rtol=RTOL, atol=GRAD_ATOL_SCALE * float(b_.abs().max()) + 1e-12, what=what)


def grad_report(a, b, floor=1e-3):
    """(worst element-wise relative error over the elements with |b| >= floor*max|b|,
    max|a-b| / max|b|) of a gradient tensor against its reference."""
    a = torch.as_tensor(a).detach().cpu().double().flatten()
    b = torch.as_tensor(b).detach().cpu().double().flatten()
    mx = float(b.abs().max())
    if mx == 0.0:
        return 0.0, float((a - b).abs().max())
    sig = b.abs() >= floor * mx
    err = (a - b).abs()
    return float((err[sig] / b.abs()[sig]).max()), float(err.max()) / mx


def build(spec, heads_cpu, dev, requires_grad=False):
    heads = [h.to(dev).requires_grad_(requires_grad) for h in heads_cpu]
    detects = [yh.Detect2DInit(spec.num_classes, a).build() for a in spec.anchors]
    merged = yh.MergedDenseDetection.from_heads(heads, detects)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    return heads, detects, merged, infos


def rows_set(b, c, n):
    return set(zip(np.asarray(b).tolist(), np.asarray(c).tolist(), np.asarray(n).tolist()))
