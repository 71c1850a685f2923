"""CPU tests of the boundary: the C-ABI library loads and exports every symbol include/yolohot.h
declares, struct layouts agree with the C compiler, validation and the no-GPU failure mode,
host-side packing / sharding logic.  No compute call is made here."""
import ctypes
import os
import re
import subprocess
import sys

import numpy as np
import pytest
import torch

import yolo_dl_b200 as yh
from yolo_dl_b200 import _lib, dist as ydist, synth

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
HEADER = os.path.join(ROOT, "include", "yolohot.h")


def declared_symbols():
    src = open(HEADER).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(yh_[a-z_0-9]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = yh.load()
    names = declared_symbols()
    assert len(names) >= 13
    for n in names:
        assert hasattr(lib, n), f"{n} declared in yolohot.h but not exported"
    assert set(names) == set(_lib.SYMBOLS), "ctypes table and header disagree"
    assert b"sm_100a" in lib.yh_version()


def test_struct_layout_matches_c(tmp_path):
    c = tmp_path / "sz.c"
    c.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "yolohot.h"\n'
                 'int main(){printf("%zu %zu %zu %zu %zu\\n", sizeof(YhLayer), offsetof(YhLayer, flat_start),'
                 ' offsetof(YhLayer, anchor_w), sizeof(YhLossCfg), offsetof(YhLossCfg, objectness_pos_weight));}')
    exe = tmp_path / "sz"
    subprocess.run(["gcc", "-I", os.path.join(ROOT, "include"), str(c), "-o", str(exe)], check=True)
    got = list(map(int, subprocess.run([str(exe)], capture_output=True, text=True, check=True).stdout.split()))
    assert got == [ctypes.sizeof(_lib.YhLayer), _lib.YhLayer.flat_start.offset, _lib.YhLayer.anchor_w.offset,
                   ctypes.sizeof(_lib.YhLossCfg), _lib.YhLossCfg.objectness_pos_weight.offset]


def test_header_is_plain_c(tmp_path):
    c = tmp_path / "h.c"
    c.write_text('#include "yolohot.h"\nint main(void){return YH_OK;}\n')
    subprocess.run(["gcc", "-std=c99", "-Wall", "-Werror", "-pedantic", "-I", os.path.join(ROOT, "include"),
                    "-c", str(c), "-o", str(tmp_path / "h.o")], check=True)


@pytest.mark.skipif(torch.cuda.is_available(), reason="checks the no-GPU failure mode")
def test_no_gpu_fails_loudly():
    """there is no CPU fallback: the workspace cannot be created and modules refuse CPU tensors."""
    lib = yh.load()
    h = ctypes.c_void_p()
    st = lib.yh_workspace_create(0, 1, 10, 1, ctypes.byref(h))
    assert st == _lib.YH_ERR_CUDA
    assert b"no CPU fallback" in lib.yh_last_error()
    spec = synth.small_spec()
    det = yh.Detect2DInit(spec.num_classes, spec.anchors[0]).build()
    with pytest.raises(yh.YoloHotError, match="no CPU fallback"):
        det.forward(torch.zeros(spec.head_shape(0, 1)))
    with pytest.raises(yh.YoloHotError, match="no CPU fallback"):
        yh.nms_by_scores(torch.zeros(2, 4), torch.zeros(2), torch.zeros(2, dtype=torch.long), 0.5)


def test_argument_validation_without_device():
    lib = yh.load()
    h = ctypes.c_void_p()
    assert lib.yh_workspace_create(0, 0, 10, 1, ctypes.byref(h)) == _lib.YH_ERR_INVALID
    assert b"max_batch" in lib.yh_last_error()
    assert lib.yh_workspace_create(0, 1, 10, 1, None) == _lib.YH_ERR_INVALID
    # NULL workspace / pointers are rejected before any CUDA call
    assert lib.yh_nms(None, None, None, 0, 1, 1, 0.1, 0.5, 0, None, None, None, None, None, None, 0, None) \
        == _lib.YH_ERR_INVALID
    assert lib.yh_assign(None, None, None, 1, None, 1, 1, 4.0, None, None, None, None, None, None, 0, None) \
        == _lib.YH_ERR_INVALID
    assert lib.yh_decode(None, 1, 1, None, None, None, None, None, 0, None) == _lib.YH_ERR_INVALID
    assert lib.yh_workspace_destroy(None) == _lib.YH_OK
    assert lib.yh_launch_count(1) == 0


def test_init_validation_mirrors_reference_messages():
    with pytest.raises(ValueError, match="iou_threshold must be non-negative"):
        yh.NonMaxSuppressionInit(iou_threshold=-0.1).build()
    with pytest.raises(ValueError, match="confidence_threshold must be non-negative"):
        yh.NonMaxSuppressionInit(confidence_threshold=-1).build()
    with pytest.raises(ValueError, match="anchor_scale_thresh must be greater than or equal to 1"):
        yh.YoloLossInit(anchor_scale_thresh=0.9).build()
    with pytest.raises(ValueError, match=r"smooth_objectness_coef must be in range \[0, 1\]"):
        yh.YoloLossInit(smooth_objectness_coef=-0.1).build()
    d = yh.YoloLossInit()
    assert (d.box_metric, d.iou_loss_weight, d.classification_loss_weight, d.objectness_loss_weight,
            d.smooth_classification_coef, d.smooth_objectness_coef, d.match_grid_method,
            d.anchor_scale_thresh, d.reduction) == ("diou", 0.05, 0.58, 1.0, 0.01, 0.0, "Rect4", 4.0, "mean")
    n = yh.NonMaxSuppressionInit()
    assert (n.iou_threshold, n.confidence_threshold, n.suppress_by_class) == (0.6, 0.1, False)


def test_pack_labels_and_layers():
    t = [[yh.RatioLabel(0.5, 0.4, 0.1, 0.2, 3)], [], [(0.1, 0.2, 0.3, 0.4, 7), (0.9, 0.8, 0.1, 0.1, 0)]]
    gt, offs = yh.pack_labels(t)
    assert offs.tolist() == [0, 1, 1, 3] and gt.dtype == torch.float64 and tuple(gt.shape) == (3, 5)
    assert gt[1].tolist() == [0.1, 0.2, 0.3, 0.4, 7.0]
    spec = synth.yolov4_608()
    infos = yh.MergedDenseDetection.make_info(spec.shapes, spec.anchors)
    assert [i.flat_start for i in infos] == [0, 3 * 76 * 76, 3 * 76 * 76 + 3 * 38 * 38]
    assert sum(i.size for i in infos) == 22743 == spec.num_instances
    lay = _lib.make_layers(infos)
    assert lay[2].H == 19 and lay[2].flat_start == infos[2].flat_start
    assert lay[0].anchor_h[0] == 16 / 608 and lay[0].anchor_w[0] == 12 / 608
    assert synth.yolov4_tiny_416().num_instances == 2535
    assert synth.yolov4_csp_640().num_instances == 25200


def test_shard_range_partitions_the_batch():
    for B in (1, 7, 8, 64, 65):
        for W in (1, 2, 3, 8):
            r = [ydist.shard_range(B, k, W) for k in range(W)]
            assert r[0][0] == 0 and r[-1][1] == B
            assert all(a[1] == b[0] for a, b in zip(r, r[1:]))
            sizes = [e - s for s, e in r]
            assert max(sizes) - min(sizes) <= 1
    assert ydist.gradient_scale(8, 64) == 0.125


def _gloo_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, ROOT)
    from oracle import detect as odet, loss as oloss
    spec = synth.small_spec()
    B = 5
    heads = synth.make_heads(spec, B, 3)
    targets = synth.make_targets(spec, B, 3, n_max=5)
    infos = odet.make_infos(spec.shapes, spec.anchors)
    s, e = ydist.shard_range(B, rank, world)
    r = oloss.yolo_loss([h[s:e] for h in heads], infos, spec.num_classes, targets[s:e], oloss.LossConfig())
    l4 = torch.stack([r["losses"][k] for k in ("total", "iou", "classification", "objectness")])
    out = ydist.weighted_mean_losses(l4, e - s)
    red = ydist.LossAllReduce(torch.device("cpu"), e - s)   # the preallocated form bench.py uses
    red.launch(l4)
    assert torch.equal(red.result(), out)
    q.put((rank, s, e, l4.tolist(), out.tolist()))
    dist.destroy_process_group()


def test_loss_scalar_all_reduce_gloo_world2():
    """world_size-2 gloo run of the N>1 path: images sharded by contiguous range, per-rank LOCAL
    means combined as sum(loss_g*B_g)/sum(B_g) (multi_gpu.rs:188-195) by one all-reduce."""
    import torch.multiprocessing as mp
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = 29500 + (os.getpid() % 2000)
    procs = [ctx.Process(target=_gloo_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=180) for _ in procs)
    for p in procs:
        p.join(60)
        assert p.exitcode == 0
    (_, s0, e0, l0, o0), (_, s1, e1, l1, o1) = res
    assert (s0, e0, s1, e1) == (0, 3, 3, 5)
    expect = (np.array(l0) * 3 + np.array(l1) * 2) / 5
    np.testing.assert_allclose(o0, expect, rtol=1e-6)
    np.testing.assert_allclose(o1, expect, rtol=1e-6)


def test_benchmark_ratios_follow_reference_expressions():
    """benchmark.rs:47-92 on integer counts (the host half of YoloBenchmark)."""
    from yolo_dl_b200.benchmark import ratios_from_counts
    r = ratios_from_counts(1000, 40, 20, 10, 5)
    assert r.obj_accuracy == (10 + (960 - 10)) / 1000 and r.obj_recall == 0.5
    assert r.obj_precision == 0.25 and r.class_accuracy == 0.25
    r = ratios_from_counts(1000, 0, 0, 0, 0)
    assert r.obj_recall == 1.0 and r.obj_precision == 1.0 and r.class_accuracy == 1.0


def _c_struct_fields(header, name):
    """field names of `typedef struct NAME { ... } NAME;` in declaration order"""
    import re
    body = re.search(r"typedef struct %s \{(.*?)\} %s;" % (name, name), header, re.S).group(1)
    body = re.sub(r"/\*.*?\*/", "", body, flags=re.S)
    out = []
    for decl in body.split(";"):
        decl = decl.strip()
        if not decl:
            continue
        names = decl.split(None, 1)[1]
        out += [re.sub(r"\[.*?\]", "", n).strip() for n in names.split(",")]
    return out


def test_integration_md_rust_block_matches_header():
    """INTEGRATION.md's Rust `#[repr(C)]` structs and `extern "C"` block are what a maintainer
    copies: every struct must list the header's fields in order (a stale struct makes the library
    read past it) and every function the header declares must be bound, with the same number of
    arguments."""
    import re
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    header = open(os.path.join(root, "include", "yolohot.h")).read()
    doc = open(os.path.join(root, "INTEGRATION.md")).read()
    rust = doc[doc.index("```rust"):doc.index("pub fn check(rc: c_int)")]
    for name in ("YhLayer", "YhLossCfg"):
        body = re.search(r"pub struct %s \{(.*?)\n\}" % name, rust, re.S).group(1)
        body = re.sub(r"//.*", "", body)
        rust_fields = [m.lower() for m in re.findall(r"pub (\w+)\s*:", body)]
        c_fields = [f.lower() for f in _c_struct_fields(header, name)]
        assert rust_fields == c_fields, (name, rust_fields, c_fields)
    hdr = re.sub(r"/\*.*?\*/", "", header, flags=re.S)
    c_fns = {m.group(1): m.group(2) for m in
             re.finditer(r"\b(yh_\w+)\s*\(([^;{]*?)\)\s*;", hdr, re.S)}
    rs = re.sub(r"//.*", "", rust)
    rust_fns = {m.group(1): m.group(2) for m in re.finditer(r"pub fn (yh_\w+)\s*\((.*?)\)", rs, re.S)}
    assert set(c_fns) == set(_lib.SYMBOLS), set(c_fns) ^ set(_lib.SYMBOLS)
    assert set(rust_fns) == set(c_fns), set(rust_fns) ^ set(c_fns)

    def nargs(a):
        a = a.strip()
        return 0 if a in ("", "void") else a.count(",") + 1
    for f in c_fns:
        assert nargs(c_fns[f]) == nargs(rust_fns[f]) == len(_lib.SYMBOLS[f][1]), f


def test_c_program_fails_loudly_without_a_gpu():
    """tests/c/abi_smoke.c is a plain-C caller of the ABI (gcc -std=c99, no Python, no
    torch).  On a machine without a CUDA device its first compute-side call must fail with the
    library's own message (exit code 2); on the GPU box tests/test_gpu_abi_c.py runs it to the end."""
    exe = os.path.join(ROOT, "tests", "c", "abi_smoke")
    assert os.path.exists(exe), "build it: make -C yolo_dl_b200/csrc abi_smoke"
    if torch.cuda.is_available():
        pytest.skip("a GPU is present: covered by the gpu-marked run")
    r = subprocess.run([exe], capture_output=True, text=True, timeout=120)
    assert r.returncode == 2, (r.returncode, r.stdout, r.stderr)
    assert "no CPU fallback" in r.stderr
