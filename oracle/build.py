"""Builds the C part of the oracle (oracle/_build/liboracle.so).  Test infrastructure only.

The reference's hot path is Rust over tch/libtorch plus the un-vendored tch-nms crate; neither
cargo nor rustc exists in this image, so none of ITS files can be compiled here (DESIGN.md
§oracle).  What the reference repository does hold in plain C is the vendored darknet box
arithmetic (darknet-test/darknet/src/box.c): `build_ref` compiles that file where it lies into
oracle/_ref/libdarknet_box.so as a SECONDARY cross-check of the box metrics (IoU / GIoU / DIoU
are the same formulas; CIoU differs by the sign yolo-dl gives the alpha*v term).
"""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))


def build(force: bool = False) -> str:
    out_dir = os.path.join(HERE, "_build")
    os.makedirs(out_dir, exist_ok=True)
    out = os.path.join(out_dir, "liboracle.so")
    srcs = [os.path.join(HERE, "matcher.c")]
    if not force and os.path.exists(out) and all(
            os.path.getmtime(out) >= os.path.getmtime(s) for s in srcs):
        return out
    cmd = ["gcc", "-O2", "-std=c11", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-o", out,
           *srcs, "-lm"]
    subprocess.run(cmd, check=True)
    return out


REFERENCE_BOX_C = "/root/reference/darknet-test/darknet/src/box.c"


def build_ref(force: bool = False):
    """oracle/_ref/libdarknet_box.so from the reference tree's own box.c (+ oracle/ref_shim.c).
    Returns the path, or None where /root/reference does not exist and nothing was prebuilt."""
    out_dir = os.path.join(HERE, "_ref")
    out = os.path.join(out_dir, "libdarknet_box.so")
    if not os.path.exists(REFERENCE_BOX_C):
        return out if os.path.exists(out) else None
    shim = os.path.join(HERE, "ref_shim.c")
    if not force and os.path.exists(out) and os.path.getmtime(out) >= os.path.getmtime(shim):
        return out
    os.makedirs(out_dir, exist_ok=True)
    dk = os.path.dirname(os.path.dirname(REFERENCE_BOX_C))
    cmd = ["gcc", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-w", "-I", os.path.join(dk, "include"),
           "-I", os.path.join(dk, "src"), "-o", out, REFERENCE_BOX_C, shim, "-lm"]
    subprocess.run(cmd, check=True)
    return out


def darknet_box_metrics(a_xywh, b_xywh):
    """runs oracle/_ref on float32 arrays [n,4] (x, y, w, h); returns (iou, giou, diou, ciou)"""
    import ctypes

    import numpy as np
    path = build_ref()
    if path is None:
        raise FileNotFoundError("oracle/_ref/libdarknet_box.so is not built and /root/reference is absent")
    lib = ctypes.CDLL(path)
    a = np.ascontiguousarray(a_xywh, dtype=np.float32)
    b = np.ascontiguousarray(b_xywh, dtype=np.float32)
    n = a.shape[0]
    outs = [np.empty(n, dtype=np.float32) for _ in range(4)]
    fp = ctypes.POINTER(ctypes.c_float)
    lib.dk_box_metrics.argtypes = [fp, fp, ctypes.c_int, fp, fp, fp, fp]
    lib.dk_box_metrics.restype = None
    lib.dk_box_metrics(a.ctypes.data_as(fp), b.ctypes.data_as(fp), n, *[o.ctypes.data_as(fp) for o in outs])
    return tuple(outs)


if __name__ == "__main__":
    print(build(force=True))
    print(build_ref(force=True))
