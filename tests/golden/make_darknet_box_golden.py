"""Generates tests/golden/darknet_box.npz: box pairs and what the REFERENCE REPOSITORY'S vendored C
code (darknet-test/darknet/src/box.c, compiled in place by oracle/build.py into oracle/_ref)
returns for them: box_iou, box_giou, box_diounms(beta1 = 1), box_ciou.  Run in the build
container, where /root/reference exists: `python tests/golden/make_darknet_box_golden.py`.
Unlike hotpath_small.npz these ARE outputs of reference-tree code (not of yolo-dl's own tch path,
which cannot be built here): a secondary pin of the box-metric family."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)

from oracle import build as obuild  # noqa: E402


def main():
    rng = np.random.default_rng(20261104)
    n = 4096
    # a: random boxes; b: a mix of jittered copies (overlapping), independent boxes (mostly
    # disjoint), nested boxes and boxes sharing an edge
    a = np.stack([rng.uniform(0.1, 0.9, n), rng.uniform(0.1, 0.9, n),
                  np.exp(rng.uniform(np.log(0.02), np.log(0.6), n)),
                  np.exp(rng.uniform(np.log(0.02), np.log(0.6), n))], 1)
    b = a.copy()
    k = n // 4
    b[:k, :2] += rng.uniform(-0.5, 0.5, (k, 2)) * a[:k, 2:]
    b[:k, 2:] *= np.exp(rng.uniform(-0.7, 0.7, (k, 2)))
    b[k:2 * k] = np.stack([rng.uniform(0.1, 0.9, k), rng.uniform(0.1, 0.9, k),
                           np.exp(rng.uniform(np.log(0.02), np.log(0.6), k)),
                           np.exp(rng.uniform(np.log(0.02), np.log(0.6), k))], 1)
    b[2 * k:3 * k, 2:] *= rng.uniform(0.2, 0.9, (k, 2))          # nested, same centre or shifted
    b[2 * k:3 * k, :2] += rng.uniform(-0.04, 0.04, (k, 2)) * a[2 * k:3 * k, 2:]
    b[3 * k:, 0] = a[3 * k:, 0] + a[3 * k:, 2]                     # touching along x: I = 0
    b[3 * k:, 2:] *= np.exp(rng.uniform(-0.5, 0.5, (n - 3 * k, 2)))
    a = a.astype(np.float32)
    b = b.astype(np.float32)
    iou, giou, diou, ciou = obuild.darknet_box_metrics(a, b)
    out = os.path.join(os.path.dirname(os.path.abspath(__file__)), "darknet_box.npz")
    np.savez_compressed(out, a_xywh=a, b_xywh=b, iou=iou, giou=giou, diou=diou, ciou=ciou)
    print(out, n, "pairs; iou>0:", int((iou > 0).sum()))


if __name__ == "__main__":
    main()
