"""yolo_dl_b200 — B200-native (sm_100a) detection hot path of jerry73204/yolo-dl:
YOLO head decode, training target assignment + box/objectness/class loss (fwd+bwd), batched NMS.

The product is `libyolohot.so` (C ABI: include/yolohot.h, kernels: yolo_dl_b200/csrc/).  This
package is the host-side mirror of the reference's module interface over that ABI; it has no
CPU fallback and never imports `oracle/`.
"""
from . import _lib
from ._lib import YoloHotError, load
from .benchmark import YoloBenchmark, YoloBenchmarkInit, YoloBenchmarkOutput
from .detect import Detect2D, Detect2DInit, MergeDetect2D, MergedDenseDetection
from .loss import CyCxHWMatcher, CyCxHWMatcherInit, YoloLoss, YoloLossInit, pack_labels
from .nms import (NonMaxSuppression, NonMaxSuppressionInit, YoloInference, YoloInferenceInit,
                  nms_by_scores)
from .tensors import (DenseDetectionTensor, DenseDetectionTensorList, DetectionInfo,
                      FlatIndexTensor, MatchingOutput, NmsOutput, RatioLabel, YoloInferenceOutput,
                      YoloLossAuxiliary, YoloLossOutput)
from .workspace import Workspace

__all__ = [n for n in dir() if not n.startswith("_")]
