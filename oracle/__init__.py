"""oracle — CPU restatement of the reference's detection hot path.  TEST INFRASTRUCTURE ONLY.

Only `tests/`, `__graft_entry__.smoke()` and the `cpu_baseline` / `--impl reference` legs of
`bench.py` may import this package.  The product (`yolo_dl_b200`) never does: it fails loudly
when the CUDA library is missing.

What it is: the reference (jerry73204/yolo-dl) is Rust over `tch` (libtorch); neither cargo
nor rustc exists in this image, so the reference itself cannot run here.  Every arithmetic
step of the path is a 1:1 ATen call in the reference, so this restatement issues the SAME
ATen op sequence through PyTorch on the CPU (each function cites the reference file:line it
follows); host-side f64 logic (the target matcher) is restated in plain Python and in C
(`oracle/matcher.c`).

PARITY UNPINNED: the reference holds no golden vectors, known-answer tests or fixtures for
Detect2D, CyCxHWMatcher, YoloLoss or NonMaxSuppression, and the NMS
arithmetic lives in the un-vendored git dependency `tch-nms` @
138b5e40b3c6a55bc93f35f9efc2790705fe6384 (yolo-dl/Cargo.toml:48).  What pins this oracle:
  * the reference's inline `debug-assert` restatements, turned into property tests
    (tests/test_oracle.py): detect_2d.rs:74-119, loss_.rs:342-368, loss_.rs:418-459,
    pred_target_matching.rs:161-174, merged_dense_detection.rs:105-116;
  * independent second implementations inside the oracle (Python vs C matcher; own greedy
    NMS vs torchvision.ops.batched_nms; closed-form BCE/IoU gradients vs autograd);
  * golden vectors generated from this oracle in this container (tests/golden/, script
    tests/golden/make_golden.py) so that any later drift is caught.
PINNED by code or known answers of the reference tree itself:
  * evaluation metrics (oracle/evaluation.py): the expected values of the reference's own tests
    pred_gt_matching.rs:149-240 and average_precision.rs:560-705 (tests/test_oracle_eval.py);
  * the DenseDetectionTensor container operations: the reference's `cat_test`,
    dense_detection_tensor.rs:582-728 (tests/test_tensors_known_answers.py);
  * the box-metric family (oracle/box.py), SECONDARY: outputs of the vendored darknet C code
    (darknet-test/darknet/src/box.c:135-256) compiled in place into oracle/_ref by
    oracle/build.py, committed as tests/golden/darknet_box.npz (tests/test_oracle_ref.py).
    IoU / GIoU / DIoU agree to 2e-6; CIoU's alpha*v term agrees in magnitude with the opposite
    sign (yolo-dl's cycxhw.rs:119 adds it).
"""
