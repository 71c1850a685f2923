// eval.cu — evaluation metrics on the device:
//   match_det_gt                    yolo-dl/src/loss/pred_gt_matching.rs:109-142
//   bbox RectFloat::iou_with        bbox/src/rect.rs:113-140  (scalar f64, epsilon 1e-8)
//   ApCalculator::compute_by_detections / compute_by_prec_rec
//                                   yolo-dl/src/loss/average_precision.rs:87-219
//   interpolate_stepwise_values     yolo-dl/src/utils/algorithm.rs:38-91
// The reference does all of this on the host in f64 (R64) with Vec / HashMap / sort; these
// kernels keep f64 and the reference's operation order (the library is built -fmad=false), so the
// integers (matched ground truth) are bit-exact and the AP values agree to the last bit with the
// plain-Python restatement in oracle/evaluation.py, which is pinned by the reference's own
// known-answer tests (tests/test_oracle_eval.py).
//
// match_det_gt_kernel  one warp per detection, lanes across the ground truths of its segment
//                      (segment = image, or image x class: the caller decides); arg-max with
//                      Iterator::max_by_key's "last maximum" rule.
// ap_kernel            one CTA per problem (a class of an evaluation set).  Two bitonic sorts of
//                      detection indices in shared memory — by (ground truth, -IoU, index) to
//                      mark the best detection of every ground truth, by (-confidence, ground
//                      truth, -IoU, index) for the precision/recall sweep — then per IoU
//                      threshold a block scan (running TP), a reverse block scan (precision
//                      envelope) and one binary search per interpolation point.
#include <math.h>

#include "common.cuh"

namespace yh {

// ---------------------------------------------------------------------------------
// scalar IoU, f64 (rect.rs:113-140); boxes are (t, l, b, r)
// ---------------------------------------------------------------------------------
__device__ __forceinline__ double iou_f64(const double4 a, const double4 b, const double eps) {
  const double t = fmax(a.x, b.x), l = fmax(a.y, b.y);
  const double bb = fmin(a.z, b.z), r = fmin(a.w, b.w);
  const double inter = (bb > t && r > l) ? (bb - t) * (r - l) : 0.0;
  const double area_a = (a.z - a.x) * (a.w - a.y);
  const double area_b = (b.z - b.x) * (b.w - b.y);
  const double uni = area_a + area_b - inter + eps;
  return inter / uni;
}

constexpr int kMatchThreads = 256;

__global__ void __launch_bounds__(kMatchThreads)
match_det_gt_kernel(const double* __restrict__ dets, const int* __restrict__ det_off,
                    const double* __restrict__ gts, const int* __restrict__ gt_off, const int S,
                    const int D, const double eps, int* __restrict__ out_gt,
                    double* __restrict__ out_iou) {
  const int lane = threadIdx.x & 31;
  const int d = blockIdx.x * (kMatchThreads / 32) + (threadIdx.x >> 5);
  if (d >= D) return;
  // segment of detection d: last s with det_off[s] <= d
  int lo = 0, hi = S;
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (det_off[mid] <= d) lo = mid; else hi = mid;
  }
  const int g0 = gt_off[lo], g1 = gt_off[lo + 1];
  const double4 db = *reinterpret_cast<const double4*>(dets + (size_t)d * 4);
  double best = -1.0;
  int best_g = -1;
  for (int g = g0 + lane; g < g1; g += 32) {
    const double4 gb = *reinterpret_cast<const double4*>(gts + (size_t)g * 4);
    const double v = iou_f64(db, gb, eps);
    if (v >= best) {  // ascending g within a lane: '>=' keeps the last maximum
      best = v;
      best_g = g;
    }
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    const double ob = __shfl_xor_sync(0xffffffffu, best, o);
    const int og = __shfl_xor_sync(0xffffffffu, best_g, o);
    if (og >= 0 && (best_g < 0 || ob > best || (ob == best && og > best_g))) {
      best = ob;
      best_g = og;
    }
  }
  if (lane == 0) {
    // (!abs_diff_eq!(iou, 0.0)).then(...)  (pred_gt_matching.rs:131), f64::EPSILON
    const bool none = best_g < 0 || fabs(best) <= 2.220446049250313e-16;
    out_gt[d] = none ? -1 : best_g;
    double v = 0.0;
    if (!none) {
      // DetBoxStruct::iou recomputes ground_truth.iou_with(detection) (:96-106)
      const double4 gb = *reinterpret_cast<const double4*>(gts + (size_t)best_g * 4);
      v = iou_f64(gb, db, eps);
    }
    out_iou[d] = v;
  }
}

// ---------------------------------------------------------------------------------
// AP
// ---------------------------------------------------------------------------------
constexpr int kApThreads = 1024;
constexpr int kApMax = 8192;  // detections per problem held in shared memory

struct ApView {
  const int* gt;        // matched ground truth per detection, -1 = none
  const double* iou;
  const double* conf;
};

// order 1: (ground truth asc [-1 first], IoU desc, index asc)
struct ByGroup {
  ApView v;
  __device__ __forceinline__ bool operator()(int a, int b) const {
    if (a == b) return false;
    if (a == 0x7fffffff || b == 0x7fffffff) return b == 0x7fffffff;
    const int ga = v.gt[a], gb = v.gt[b];
    if (ga != gb) return ga < gb;
    const double ia = v.iou[a], ib = v.iou[b];
    if (ia != ib) return ia > ib;
    return a < b;
  }
};
// order 2: (confidence desc, then order 1): the stable sort_by_key(-confidence) of :194
struct ByConf {
  ApView v;
  __device__ __forceinline__ bool operator()(int a, int b) const {
    if (a == b) return false;
    if (a == 0x7fffffff || b == 0x7fffffff) return b == 0x7fffffff;
    const double ca = v.conf[a], cb = v.conf[b];
    if (ca != cb) return ca > cb;
    return ByGroup{v}(a, b);
  }
};

template <class Less>
__device__ void bitonic_sort_idx(int* s, const int P, const Less less) {
  for (int k = 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j >= 1; j >>= 1) {
      for (int p = threadIdx.x; p < P / 2; p += kApThreads) {
        const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
        const int l = i | j;
        const bool asc = (i & k) == 0;
        const int a = s[i], c = s[l];
        if (less(c, a) == asc) {
          s[i] = c;
          s[l] = a;
        }
      }
      __syncthreads();
    }
  }
}
