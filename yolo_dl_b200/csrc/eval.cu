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

#include <atomic>

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

// block-wide inclusive scans over `n` ints / doubles in shared memory, chunked: thread t owns the
// contiguous elements [t*per, (t+1)*per); s_part[kApThreads] is scratch
__device__ void block_inclusive_sum(int* a, const int n, int* s_part) {
  const int per = (n + kApThreads - 1) / kApThreads;
  const int lo = threadIdx.x * per, hi = min(n, lo + per);
  int run = 0;
  for (int i = lo; i < hi; ++i) {
    run += a[i];
    a[i] = run;
  }
  s_part[threadIdx.x] = run;
  __syncthreads();
  // Hillis-Steele over the kApThreads partial sums
  for (int o = 1; o < kApThreads; o <<= 1) {
    const int v = (threadIdx.x >= o) ? s_part[threadIdx.x - o] : 0;
    __syncthreads();
    s_part[threadIdx.x] += v;
    __syncthreads();
  }
  const int before = threadIdx.x ? s_part[threadIdx.x - 1] : 0;
  for (int i = lo; i < hi; ++i) a[i] += before;
  __syncthreads();
}

// suffix maximum in place: a[i] = max(a[i..n))
__device__ void block_suffix_max(double* a, const int n, double* s_part) {
  const int per = (n + kApThreads - 1) / kApThreads;
  const int lo = threadIdx.x * per, hi = min(n, lo + per);
  double run = -INFINITY;
  for (int i = hi - 1; i >= lo; --i) {
    run = fmax(run, a[i]);
    a[i] = run;
  }
  s_part[threadIdx.x] = run;
  __syncthreads();
  for (int o = 1; o < kApThreads; o <<= 1) {
    const double v = (threadIdx.x + o < kApThreads) ? s_part[threadIdx.x + o] : -INFINITY;
    __syncthreads();
    s_part[threadIdx.x] = fmax(s_part[threadIdx.x], v);
    __syncthreads();
  }
  const double after = (threadIdx.x + 1 < kApThreads) ? s_part[threadIdx.x + 1] : -INFINITY;
  for (int i = lo; i < hi; ++i) a[i] = fmax(a[i], after);
  __syncthreads();
}

// dynamic shared memory layout (P = next pow2 of D, D <= kApMax):
//   int    idx[P]      sorted detection indices (order 2)
//   int    tp[P]       per position: is TP -> running TP
//   double env[P]      per position: precision -> suffix max
//   u8     best[P]     detection (by LOCAL index) is the best of its ground truth
//   scratch for the scans / the point values
__global__ void __launch_bounds__(kApThreads)
ap_kernel(const int* __restrict__ det_gt, const double* __restrict__ det_iou,
          const double* __restrict__ det_conf, const int* __restrict__ prob_off,
          const int* __restrict__ num_gt, const double* __restrict__ thresholds, const int T,
          const int n_points, double* __restrict__ out_ap, unsigned char* __restrict__ scratch,
          const size_t scratch_per_problem) {
  extern __shared__ __align__(16) unsigned char smem_raw[];
  const int p = blockIdx.x;
  const int d0 = prob_off[p], D = prob_off[p + 1] - d0;
  if (D <= 0) {  // the reference panics on an empty curve (average_precision.rs:90); 0 here
    for (int k = threadIdx.x; k < T; k += kApThreads) out_ap[(size_t)p * T + k] = 0.0;
    return;
  }
  int P = 2;
  while (P < D) P <<= 1;
  // the scan partials always live in shared memory; the four per-detection arrays too when every
  // problem of the launch fits (<= kApMax detections), else in this problem's slice of the caller's scratch
  // (L2-resident; the same code walks either through generic pointers)
  double* s_dpart = reinterpret_cast<double*>(smem_raw);  // kApThreads doubles (also the point values)
  int* s_ipart = reinterpret_cast<int*>(s_dpart + kApThreads);  // kApThreads ints
  // (one decision per LAUNCH: with scratch the shared-memory request holds the partials only)
  unsigned char* arrays = (scratch == nullptr) ? reinterpret_cast<unsigned char*>(s_ipart + kApThreads)
                                               : scratch + (size_t)p * scratch_per_problem;
  double* env = reinterpret_cast<double*>(arrays);
  int* idx = reinterpret_cast<int*>(env + P);
  int* tp = idx + P;
  unsigned char* best = reinterpret_cast<unsigned char*>(tp + P);

  ApView v;
  v.gt = det_gt + d0;
  v.iou = det_iou + d0;
  v.conf = det_conf + d0;

  // order 1 -> best-of-ground-truth flags (:171-189)
  for (int i = threadIdx.x; i < P; i += kApThreads) idx[i] = i < D ? i : 0x7fffffff;
  __syncthreads();
  bitonic_sort_idx(idx, P, ByGroup{v});
  for (int i = threadIdx.x; i < D; i += kApThreads) {
    const int me = idx[i];
    best[me] = (i == 0 || v.gt[idx[i - 1]] != v.gt[me]) ? 1 : 0;
  }
  __syncthreads();
  // order 2 (:194); the list is already in order 1, any correct sort gives the same result
  bitonic_sort_idx(idx, P, ByConf{v});

  const double ngt = (double)num_gt[p];
  for (int k = 0; k < T; ++k) {
    const double thr = thresholds[k];
    for (int i = threadIdx.x; i < D; i += kApThreads) {
      const int me = idx[i];
      tp[i] = (best[me] && v.iou[me] >= thr) ? 1 : 0;
    }
    __syncthreads();
    block_inclusive_sum(tp, D, s_ipart);
    // precision = tp / (tp + fp), tp + fp = position + 1  (:199-213)
    for (int i = threadIdx.x; i < D; i += kApThreads) {
      const double a = (double)tp[i], b = (double)(i + 1 - tp[i]);
      env[i] = a / (a + b);
    }
    __syncthreads();
    block_suffix_max(env, D, s_dpart);  // precision envelope (:102-135)

    // sentinel `last` (:97-100): recall min(max_recall + 1e-3, 1), precision 0
    const double max_recall = (double)tp[D - 1] / ngt;
    const double r_last = fmin(max_recall + 1e-3, 1.0);
    // interpolation points (:141-149, algorithm.rs:38-91): the value at x is the envelope at the
    // first curve entry with recall > x; past the last entry only the point that exhausts the
    // curve may contribute (its precision if that recall is exactly 1)
    double val = 0.0;
    if ((int)threadIdx.x < n_points) {
      const double x = (double)threadIdx.x / (double)(n_points - 1);
      // entries j = 0..D-1 real (recall tp[j]/ngt), j = D the sentinel
      int lo = 0, hi = D + 1;  // first j with recall_j > x, D+1 = none
      while (lo < hi) {
        const int mid = (lo + hi) >> 1;
        const double r = mid < D ? (double)tp[mid] / ngt : r_last;
        if (r > x) hi = mid; else lo = mid + 1;
      }
      if (lo <= D) {
        val = lo < D ? env[lo] : 0.0;
      } else {
        // x >= r_last: this point contributes only if it is the FIRST such point and r_last == 1
        const bool first_past = threadIdx.x == 0 ||
                                !((double)(threadIdx.x - 1) / (double)(n_points - 1) >= r_last);
        if (first_past && r_last == 1.0) {
          int a = 0, b = D;  // first j with recall_j >= 1.0 (sentinel at D qualifies)
          while (a < b) {
            const int mid = (a + b) >> 1;
            if ((double)tp[mid] / ngt >= 1.0) b = mid; else a = mid + 1;
          }
          val = a < D ? env[a] : 0.0;
        }
      }
    }
    __syncthreads();
    if ((int)threadIdx.x < n_points) s_dpart[threadIdx.x] = val;
    __syncthreads();
    if (threadIdx.x == 0) {
      double s = 0.0;
      for (int i = 0; i < n_points; ++i) s += s_dpart[i];  // Iterator::sum order (:148)
      out_ap[(size_t)p * T + k] = s / (double)n_points;
    }
    __syncthreads();
  }
}

// env f64[P] | idx i32[P] | tp i32[P] | best u8[P], rounded to 16 bytes
static size_t ap_array_bytes(int P) {
  return (((size_t)P * (sizeof(double) + 2 * sizeof(int) + 1)) + 15) & ~(size_t)15;
}
static size_t ap_smem_bytes(int P) {
  return kApThreads * (sizeof(double) + sizeof(int)) + (P <= kApMax ? ap_array_bytes(P) : 0);
}
static int ap_pow2(int d) {
  int P = 2;
  while (P < d) P <<= 1;
  return P;
}

}  // namespace yh

extern "C" {

int yh_match_det_gt(const double* dets_tlbr_dev, const int32_t* det_offsets_dev, int D,
                    const double* gts_tlbr_dev, const int32_t* gt_offsets_dev, int S,
                    int32_t* out_gt_dev, double* out_iou_dev, void* stream) {
  using namespace yh;
  YH_REQUIRE(D >= 0 && S >= 1, "D=%d S=%d", D, S);
  YH_REQUIRE(det_offsets_dev && gt_offsets_dev, "offsets are NULL");
  if (D == 0) return YH_OK;
  // gts may be NULL when no segment has a ground truth (it is then never dereferenced)
  YH_REQUIRE(dets_tlbr_dev && out_gt_dev && out_iou_dev, "NULL argument");
  YH_REQUIRE((((uintptr_t)dets_tlbr_dev | (uintptr_t)gts_tlbr_dev) & 31) == 0,
             "boxes must be 32-byte aligned");
  cudaStream_t s = (cudaStream_t)stream;
  const int warps = kMatchThreads / 32;
  YH_TIME(YH_K_MATCH_DET_GT, s);
  match_det_gt_kernel<<<(D + warps - 1) / warps, kMatchThreads, 0, s>>>(
      dets_tlbr_dev, det_offsets_dev, gts_tlbr_dev, gt_offsets_dev, S, D, 1e-8, out_gt_dev,
      out_iou_dev);
  YH_LAUNCHED(s);
  return YH_OK;
}

size_t yh_average_precision_scratch_bytes(int num_problems, int max_detections_per_problem) {
  if (num_problems < 1 || max_detections_per_problem <= yh::kApMax) return 0;
  return (size_t)num_problems * yh::ap_array_bytes(yh::ap_pow2(max_detections_per_problem));
}

int yh_average_precision(const int32_t* det_gt_dev, const double* det_iou_dev,
                         const double* det_conf_dev, const int32_t* problem_offsets_dev,
                         const int32_t* num_ground_truth_dev, int num_problems,
                         int max_detections_per_problem, const double* iou_thresholds_dev,
                         int num_thresholds, int n_points, double* out_ap_dev, void* scratch_dev,
                         size_t scratch_bytes, void* stream) {
  using namespace yh;
  YH_REQUIRE(num_problems >= 1 && num_thresholds >= 1, "num_problems=%d num_thresholds=%d",
             num_problems, num_thresholds);
  // ApCalculator::new (:72-84); a single point divides by n_points-1 = 0 in the reference
  YH_REQUIRE(n_points >= 2 && n_points <= kApThreads, "invalid number of interpolated points %d",
             n_points);
  YH_REQUIRE(max_detections_per_problem >= 0 && max_detections_per_problem <= (1 << 30),
             "max_detections_per_problem=%d", max_detections_per_problem);
  YH_REQUIRE(problem_offsets_dev && num_ground_truth_dev && iou_thresholds_dev && out_ap_dev,
             "NULL argument");
  const int P = ap_pow2(max_detections_per_problem);
  const size_t need = yh_average_precision_scratch_bytes(num_problems, max_detections_per_problem);
  if (need > 0 && (scratch_dev == nullptr || scratch_bytes < need)) {
    set_error("%d detections in one problem exceed the %d held in shared memory: pass %zu bytes of "
              "device scratch (yh_average_precision_scratch_bytes), got %zu",
              max_detections_per_problem, kApMax, need, scratch_dev ? scratch_bytes : (size_t)0);
    return YH_ERR_CAPACITY;
  }
  YH_REQUIRE(need == 0 || ((uintptr_t)scratch_dev & 15) == 0, "scratch not 16-byte aligned");
  const size_t smem = ap_smem_bytes(P);
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaFuncSetAttribute(ap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  YH_TIME(YH_K_AP, s);
  ap_kernel<<<num_problems, kApThreads, smem, s>>>(det_gt_dev, det_iou_dev, det_conf_dev,
                                                   problem_offsets_dev, num_ground_truth_dev,
                                                   iou_thresholds_dev, num_thresholds, n_points,
                                                   out_ap_dev,
                                                   need ? (unsigned char*)scratch_dev : nullptr,
                                                   need ? ap_array_bytes(P) : (size_t)0);
  YH_LAUNCHED(s);
  return YH_OK;
}

}  // extern "C"

// ---- yh_average_precision_static: the same kernel with its scratch in a file-scope device array.
// No counterpart in the reference; for callers that cannot allocate (include/yolohot.h).
namespace yh {

// Every GPU that loads the library holds these YH_AP_STATIC_BYTES until the process exits.
static __device__ __align__(256) unsigned char g_ap_static[YH_AP_STATIC_BYTES];

constexpr int kApMaxDevices = 64;
// device address of g_ap_static per device.  In host code `g_ap_static` is only the host-side
// stand-in of the symbol: the address a kernel may use comes from cudaGetSymbolAddress, asked
// once per device with that device current.
static std::atomic<unsigned char*> g_ap_static_ptr[kApMaxDevices];

static int ap_static_address(unsigned char** out) {
  int dev = 0;
  YH_CUDA(cudaGetDevice(&dev));
  YH_REQUIRE(dev >= 0 && dev < kApMaxDevices, "device %d beyond the %d this entry point tracks", dev,
             kApMaxDevices);
  unsigned char* p = g_ap_static_ptr[dev].load(std::memory_order_acquire);
  if (p == nullptr) {
    void* sym = nullptr;
    YH_CUDA(cudaGetSymbolAddress(&sym, g_ap_static));
    YH_REQUIRE(sym != nullptr, "cudaGetSymbolAddress returned NULL for the static scratch");
    p = static_cast<unsigned char*>(sym);
    g_ap_static_ptr[dev].store(p, std::memory_order_release);  // racing threads store the same value
  }
  *out = p;
  return YH_OK;
}

}  // namespace yh

extern "C" {

int yh_average_precision_static(const int32_t* det_gt_dev, const double* det_iou_dev,
                                const double* det_conf_dev, const int32_t* problem_offsets_dev,
                                const int32_t* num_ground_truth_dev, int num_problems,
                                int max_detections_per_problem, const double* iou_thresholds_dev,
                                int num_thresholds, int n_points, double* out_ap_dev, void* stream) {
  using namespace yh;
  YH_REQUIRE(num_problems >= 1 && num_thresholds >= 1, "num_problems=%d num_thresholds=%d",
             num_problems, num_thresholds);
  YH_REQUIRE(n_points >= 2 && n_points <= kApThreads, "invalid number of interpolated points %d",
             n_points);
  YH_REQUIRE(max_detections_per_problem >= 0 && max_detections_per_problem <= (1 << 30),
             "max_detections_per_problem=%d", max_detections_per_problem);
  YH_REQUIRE(problem_offsets_dev && num_ground_truth_dev && iou_thresholds_dev && out_ap_dev,
             "NULL argument");
  const int P = ap_pow2(max_detections_per_problem);
  // same rule as yh_average_precision: up to kApMax detections per problem everything is in
  // shared memory and the static array is not touched
  const size_t need = yh_average_precision_scratch_bytes(num_problems, max_detections_per_problem);
  if (need > (size_t)YH_AP_STATIC_BYTES) {
    set_error("yh_average_precision_static: %d problems of up to %d detections need %zu bytes of "
              "scratch, the static array holds %zu (YH_AP_STATIC_BYTES); use yh_average_precision "
              "with caller scratch",
              num_problems, max_detections_per_problem, need, (size_t)YH_AP_STATIC_BYTES);
    return YH_ERR_CAPACITY;
  }
  unsigned char* scratch = nullptr;
  if (need > 0) {
    const int rc = ap_static_address(&scratch);
    if (rc != YH_OK) return rc;
  }
  // nothing in the array has to start at zero: the kernel writes idx[0,P), best/tp/env[0,D) of a
  // problem's slice before it reads them, so the previous call's contents are never observed
  const size_t smem = ap_smem_bytes(P);
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaFuncSetAttribute(ap_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem));
  YH_TIME(YH_K_AP, s);
  ap_kernel<<<num_problems, kApThreads, smem, s>>>(det_gt_dev, det_iou_dev, det_conf_dev,
                                                   problem_offsets_dev, num_ground_truth_dev,
                                                   iou_thresholds_dev, num_thresholds, n_points,
                                                   out_ap_dev, scratch,
                                                   need ? ap_array_bytes(P) : (size_t)0);
  YH_LAUNCHED(s);
  return YH_OK;
}

}  // extern "C"
