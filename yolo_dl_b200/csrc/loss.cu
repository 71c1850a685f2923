// loss.cu — K4 (dense objectness BCE forward+backward, writes the whole head gradient) and
// K3 (matched rows: box metric + class BCE forward+backward, objectness correction), i.e.
// YoloLoss::forward (yolo-dl/src/loss/loss_.rs:227-468) plus what total_loss.backward()
// delivers to the three head tensors (train/src/train/single_gpu.rs:161-168).
//
// Decomposition.  The objectness target is zero everywhere except at the M matched cells
// (loss_.rs:393-415), so the dense pass treats EVERY cell as target 0:
//     BCE(x, 0) = softplus(x),   d/dx = sigmoid(x)
// and the row pass, which runs after it on the same stream, corrects the M matched cells:
//     BCE(x, t) - BCE(x, 0) = -t*x + (lw-1)*softplus(-x),  lw = 1 + (pos_weight-1)*t
// and overwrites their objectness gradient.  No dense target plane is ever materialised
// (the reference builds zeros_like + index_put_, then reads both planes).
//
// K4 traffic: reads the obj plane (4 B/instance), writes the head gradient (4*(5+C)
// B/instance, zeros outside the obj channels) = the 344*N bytes/image of SURVEY.md §8d.
// It is a pure streaming kernel: float4 stores over the flat gradient tensor of each layer
// (the tensor is contiguous, so the flat view is 16-byte aligned even for 19x19 maps whose
// channel rows are not).
// K3: a warp per 8 matched rows (scalar box part one row per lane, class part lanes across the
// C channels); the 5+C logits of a row are (5+C) scattered 4-byte reads, M is 10..15 per GT.
//
// Reductions are deterministic: per-CTA partials in fixed slots, summed in a fixed order by
// the last CTA of K3 (ticket counter), in f64.
#include <math.h>
#include <stdlib.h>

#include "common.cuh"

namespace yh {

constexpr int kDenseThreads = 256;
constexpr int kRowThreads = 128;
constexpr int kPartHalf = YH_PARTIALS_MAX / 2;  // K4 partials in [0,kPartHalf), K3 after

struct GradTab {
  float* grad[YH_MAX_LAYERS];
};

struct LossParams {
  int B;
  int metric;
  int reduction;       // YhReduction
  float w_iou, w_cls, w_obj;
  float cls_pos, cls_neg;   // 1-0.5*s, 1-pos (loss_.rs:326-327), cast to f32 by Tensor::full
  float obj_base, obj_coef; // 1-s_o, s_o (loss_.rs:397-406)
  int has_pw;
  float pw;
  int obj_kind, cls_kind;   // YhLossKind
  float gamma;              // focal_loss_gamma
};

__device__ __forceinline__ float softplus_f(float x) {
  // log(1 + exp(x)) = max(x,0) + log1p(exp(-|x|))
  return fmaxf(x, 0.0f) + log1pf(expf(-fabsf(x)));
}

__device__ __forceinline__ float sigmoid_l(float x) { return 1.0f / (1.0f + expf(-x)); }

// The non-default loss kinds, per logit: loss(x, t) and d loss / d x.
//   L2     mse on the logit (l2_loss.rs:13-15): (x-t)^2
//   Focal  (focal_loss.rs:90-98) around BceWithLogits (reduction None, optional pos_weight):
//          bce * (t*a + (1-t)*(1-a)) * (1 - p_t)^gamma,  p_t = t*p + (1-t)*(1-p),  a = 0.25;
//          reverse mode of the same chain; pow(., 0) has a zero derivative like ATen's.
__device__ __forceinline__ float kind_loss(const int kind, const float x, const float t,
                                           const bool has_pw, const float pw, const float gamma,
                                           float& dx) {
  if (kind == YH_LOSS_L2) {
    const float d = x - t;
    dx = 2.0f * d;
    return d * d;
  }
  const float kAlpha = 0.25f;  // FocalLossInit::default (focal_loss.rs:27-34)
  const float lw = has_pw ? (pw - 1.0f) * t + 1.0f : 1.0f;
  const float b = (1.0f - t) * x + lw * softplus_f(-x);
  const float p = sigmoid_l(x);
  float db;
  if (has_pw) {
    const float tp = pw * t;
    db = ((tp + 1.0f) - t) * p - tp;
  } else {
    db = p - t;
  }
  const float pt = t * p + (1.0f - t) * (1.0f - p);
  const float af = t * kAlpha + (1.0f - t) * (1.0f - kAlpha);
  const float q = -pt + 1.0f;
  float mf = 1.0f, dmf = 0.0f;
  if (gamma != 0.0f) {
    mf = powf(q, gamma);
    const float dpt = (t - (1.0f - t)) * ((1.0f - p) * p);
    dmf = -(gamma * powf(q, gamma - 1.0f)) * dpt;
  }
  dx = (db * mf + b * dmf) * af;
  return (b * af) * mf;
}

// ---------------------------------------------------------------------------------
// K4
// ---------------------------------------------------------------------------------
// GRAD: the flat gradient tensor of every layer is cut into tiles of kTileQuads float4 quads
// (32 KB); a tile never straddles two layers.  One 64-bit modulo per TILE (not per store)
// tells whether the tile intersects an objectness plane (floats [4*run, 5*run) of an image):
//   * it does not (~97 % of the tiles): kTileUnroll unconditional 16-byte streaming zero
//     stores per thread — no loads, no address arithmetic beyond an immediate offset.
//     tools/fill_bench.cu measured this form at 6.0-6.3 TB/s on B200 (cudaMemset: 6.6-6.8),
//     level with TMA bulk stores from a zeroed shared buffer, so the simpler form is used;
//   * it does: per-quad test with a single conditional wrap instead of a modulo; the quads
//     that hold an objectness element are NOT written here but by pass 1, which spreads all
//     B*N/4 of them evenly over the grid (left in the tiles they serialise a few CTAs on
//     load latency + transcendental math while the others have long finished: 97 -> ~85 us).
// per-layer prefixes of loss_dense_kernel's two passes, computed on the host
struct DenseTab {
  int L;
  long long run[YH_MAX_LAYERS];              // A*HW: floats per entry per image
  long long tile_prefix[YH_MAX_LAYERS + 1];  // pass 2, in tiles
  long long qprefix[YH_MAX_LAYERS + 1];      // pass 1, in quads
  // pass 2: tiles from this index on are stored with the L2 evict_last priority instead of as
  // streaming stores — the END of the store stream (the small layers, where most matched rows
  // lie) is then still in the L2 when the scatter of the staged rows runs behind it
  long long keep_from;
};

constexpr int kTileUnroll = 8;
constexpr int kTileQuads = kDenseThreads * kTileUnroll;  // 2048 quads = 8192 floats = 32 KB
constexpr int kTileFloats = kTileQuads * 4;
#ifndef YH_KEEP_MB_DEFAULT
#define YH_KEEP_MB_DEFAULT 64
#endif

// BCE: the default objectness kind keeps its own instantiation (40 registers, 6 CTAs/SM)
template <bool GRAD, bool BCE>
__global__ void __launch_bounds__(kDenseThreads)
loss_dense_kernel(const LayerTab t, const GradTab g, const int B, const float numel,
                  const float w_obj, const int reduction, const int kind, const float gamma,
                  double* __restrict__ partials, const int n_obj, const DenseTab d) {
  // PDL: this kernel reads nothing an earlier kernel of its stream wrote, so it works first and
  // waits for its predecessor only before it EXITS (completion stays transitive along the
  // stream).  Consecutive ranged launches (yh_loss_dense_range per image group) then overlap the
  // drain of one with the ramp-up of the next instead of losing ~1.5 us of bandwidth per launch.
  griddep_launch();
  double acc = 0.0;
  if (GRAD) {
    // NOT a persistent kernel: the first n_obj CTAs do pass 1, every later CTA writes exactly one
    // tile of pass 2 and exits.  The hardware hands CTAs out in index order, so the tiles being
    // written at any moment are a small contiguous window of the tensor: 6.9-7.0 TB/s on B200
    // where a grid-stride loop over the same tiles reaches 6.0-6.3 and cudaMemset 6.65
    // (tools/fill_bench.cu, profiles/r1f_fill_bench.txt).  The objectness CTAs go first so that
    // their load latency and transcendental math are hidden under the store stream.
    // A one-tile CTA cannot afford a prologue: the per-layer prefixes come precomputed in `d`
    // and the layer's fields are picked with compile-time indices (a run-time index into a
    // kernel-parameter array makes every thread copy that array to local memory first).
    const int E = t.E;
    if ((int)blockIdx.x >= n_obj) {
      // ---- pass 2: one 32 KB tile of zeros -------------------------------------------------
      const long long tile = (long long)blockIdx.x - n_obj;
      long long run = d.run[0], tp = 0;
      float* __restrict__ gl = g.grad[0];
#pragma unroll
      for (int k = 1; k < YH_MAX_LAYERS; ++k)
        if (k < d.L && tile >= d.tile_prefix[k]) {
          run = d.run[k];
          tp = d.tile_prefix[k];
          gl = g.grad[k];
        }
      const long long per_img = run * E;                 // floats per image
      const long long len = (long long)B * per_img;
      const long long idx0 = (tile - tp) * kTileFloats;  // first float of the tile
      // its offset inside the image (32-bit modulo whenever the layer tensor allows it)
      const long long r0 = (len < (1LL << 31)) ? (long long)((unsigned)idx0 % (unsigned)per_img)
                                               : idx0 % per_img;
      const long long o0 = 4 * run, o1 = 5 * run;        // obj plane of an image
      float* __restrict__ gbase = gl + idx0;
      const bool whole = idx0 + kTileFloats <= len;
      const bool one_wrap = per_img >= kTileFloats + 4;
      const bool touches = !one_wrap || (r0 + kTileFloats > o0 && r0 < o1) ||
                           (r0 + kTileFloats > per_img + o0);
      const float4 zero4 = make_float4(0.f, 0.f, 0.f, 0.f);
      if (whole && !touches) {
        float4* q = reinterpret_cast<float4*>(gbase) + threadIdx.x;
        if (tile >= d.keep_from) {
          const unsigned long long pol = l2_policy_evict_last();
#pragma unroll
          for (int u = 0; u < kTileUnroll; ++u) st_keep4(q + u * kDenseThreads, zero4, pol);
        } else {
#pragma unroll
          for (int u = 0; u < kTileUnroll; ++u) st_stream4(q + u * kDenseThreads, zero4);
        }
        griddep_wait();
        return;
      }
      // a tile that meets an objectness plane or the end of the tensor: per-quad test, quads
      // with an objectness element belong to pass 1
#pragma unroll
      for (int u = 0; u < kTileUnroll; ++u) {
        const int qi = u * kDenseThreads + threadIdx.x;
        const long long idx = idx0 + 4LL * qi;
        if (idx >= len) break;
        long long r = r0 + 4LL * qi;
        if (one_wrap) {
          if (r >= per_img) r -= per_img;
        } else {
          r %= per_img;
        }
        const bool hit = (r + 3 >= o0 && r < o1) || (r + 3 >= per_img + o0);
        if (hit) continue;
        float* gp = gbase + 4 * qi;
        if (idx + 3 < len) {
          st_stream4(reinterpret_cast<float4*>(gp), zero4);
        } else {
#pragma unroll
          for (int k = 0; k < 4; ++k)
            if (idx + k < len) st_stream1(gp + k, 0.0f);
        }
      }
      griddep_wait();
      return;
    }
    // ---- pass 1: the quads that hold at least one objectness element ---------------------
    // (B*N/4 of them, spread evenly over the n_obj leading CTAs).  Slot j of image b of layer l
    // is quad floor((b*per_img + 4*run)/4) + j, if it still starts before the plane's end.
    const long long total = d.qprefix[YH_MAX_LAYERS];
    const long long stride = (long long)n_obj * kDenseThreads;
    for (long long i = (long long)blockIdx.x * kDenseThreads + threadIdx.x; i < total; i += stride) {
      int run = (int)d.run[0];
      long long qp = 0;
      const float* __restrict__ hl = t.head[0];
      float* __restrict__ gl = g.grad[0];
#pragma unroll
      for (int k = 1; k < YH_MAX_LAYERS; ++k)
        if (k < d.L && i >= d.qprefix[k]) {
          run = (int)d.run[k];
          qp = d.qprefix[k];
          hl = t.head[k];
          gl = g.grad[k];
        }
      const int slots = run / 4 + 2;
      const long long li = i - qp;
      const long long b = li / slots;
      const int j = (int)(li - b * slots);
      const long long per_img = (long long)run * E;
      const long long len = (long long)B * per_img;
      const long long o0 = b * per_img + 4LL * run, o1 = o0 + run;  // obj plane of image b
      const long long idx = (o0 / 4 + j) * 4;                        // first float of the quad
      if (idx >= o1) continue;
      const float* __restrict__ hp = hl + idx;
      float* __restrict__ gp = gl + idx;
      const bool full = idx + 3 < len;
      float x[4] = {0.f, 0.f, 0.f, 0.f};
      if (idx >= o0 && idx + 3 < o1) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(hp));
        x[0] = v.x; x[1] = v.y; x[2] = v.z; x[3] = v.w;
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (idx + k >= o0 && idx + k < o1) x[k] = __ldcs(hp + k);
      }
      float v[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
      for (int k = 0; k < 4; ++k) {
        if (idx + k >= o0 && idx + k < o1) {
          float gi;
          if (BCE) {
            acc += (double)softplus_f(x[k]);
            // BCE backward, target 0: sigmoid(x) [/numel for Mean] * upstream w_obj
            gi = sigmoid_f(x[k]);
          } else {
            acc += (double)kind_loss(kind, x[k], 0.0f, false, 0.0f, gamma, gi);
          }
          if (reduction == YH_REDUCTION_MEAN) gi = gi / numel;
          v[k] = gi * w_obj;
        }
      }
      if (full) {
        st_stream4(reinterpret_cast<float4*>(gp), make_float4(v[0], v[1], v[2], v[3]));
      } else {
#pragma unroll
        for (int k = 0; k < 4; ++k)
          if (idx + k < len) st_stream1(gp + k, v[k]);
      }
    }
  } else {
    // forward only: just the objectness planes
    long long prefix[YH_MAX_LAYERS + 1];
    prefix[0] = 0;
#pragma unroll
    for (int l = 0; l < YH_MAX_LAYERS; ++l)
      prefix[l + 1] = prefix[l] + ((l < t.L) ? (long long)B * t.A[l] * t.HW[l] : 0LL);
    const long long total = prefix[YH_MAX_LAYERS];
    const long long stride = (long long)gridDim.x * kDenseThreads;
    for (long long i = (long long)blockIdx.x * kDenseThreads + threadIdx.x; i < total; i += stride) {
      int l = 0;
#pragma unroll
      for (int k = 1; k < YH_MAX_LAYERS; ++k)
        if (k < t.L && i >= prefix[k]) l = k;
      const long long li = i - prefix[l];
      const int run = t.A[l] * t.HW[l];
      const long long per_img = (long long)run * t.E;
      const long long b = li / run;
      const long long j = li - b * run;
      const float x = __ldcs(t.head[l] + b * per_img + 4LL * run + j);
      float unused;
      acc += BCE ? (double)softplus_f(x)
                                   : (double)kind_loss(kind, x, 0.0f, false, 0.0f, gamma, unused);
    }
  }

  // CTA partial, fixed slot
  __shared__ double s_part[kDenseThreads / 32];
  acc = warp_sum_d(acc);
  if ((threadIdx.x & 31) == 0) s_part[threadIdx.x >> 5] = acc;
  __syncthreads();
  if (threadIdx.x == 0) {
    double tot = 0.0;
#pragma unroll
    for (int k = 0; k < kDenseThreads / 32; ++k) tot += s_part[k];
    partials[(size_t)blockIdx.x * 4] = tot;
  }
  griddep_wait();
}

// ---------------------------------------------------------------------------------
// box metric, forward + reverse-mode gradient of the same op sequence
// (tch-goodies/src/compound_tensor/cycxhw.rs:67-120, tlbr.rs:81-134,214-225)
// ---------------------------------------------------------------------------------
// ATen's maximum/minimum backward: all of the gradient to the selected operand, half on ties
__device__ __forceinline__ float w_gt(float a, float b) { return a > b ? 1.f : (a == b ? 0.5f : 0.f); }
__device__ __forceinline__ float w_lt(float a, float b) { return a < b ? 1.f : (a == b ? 0.5f : 0.f); }

struct Box4 {
  float cy, cx, h, w;
};

// Returns the metric; if want_grad, d receives d(metric)/d(pred) * g.
__device__ __forceinline__ float box_metric(const int metric, const Box4 p, const Box4 q,
                                            const float g, const bool want_grad, Box4& d) {
  const float EPS = 1e-16f;  // utils.rs:5 (f64 1e-16 added to an f32 tensor)
  // TLBR (tlbr.rs:214-225)
  const float pt = p.cy - p.h / 2.0f, pb = p.cy + p.h / 2.0f;
  const float pl = p.cx - p.w / 2.0f, pr = p.cx + p.w / 2.0f;
  const float qt = q.cy - q.h / 2.0f, qb = q.cy + q.h / 2.0f;
  const float ql = q.cx - q.w / 2.0f, qr = q.cx + q.w / 2.0f;
  // intersection (tlbr.rs:81-108)
  const float max_t = fmaxf(pt, qt), max_l = fmaxf(pl, ql);
  const float min_b = fminf(pb, qb), min_r = fminf(pr, qr);
  const float dhh = min_b - max_t, dww = min_r - max_l;
  const float ih = fmaxf(dhh, 0.0f), iw = fmaxf(dww, 0.0f);
  const float inter = ih * iw;
  const float ap = p.h * p.w, aq = q.h * q.w;
  const float outer = ap + aq - inter + EPS;
  const float iou = inter / outer;

  float m = iou;
  // closure (tlbr.rs:111-134)
  const float ct = fminf(pt, qt), cl = fminf(pl, ql), cb = fmaxf(pb, qb), cr = fmaxf(pr, qr);
  const float ch = cb - ct, cw = cr - cl;
  float carea = 0.f, gu = 0.f, gv = 0.f, diag = 0.f, cdist = 0.f, qd = 0.f;
  float dy = 0.f, dx = 0.f, ang_p = 0.f, ang_q = 0.f, shape = 0.f, alpha = 0.f;
  const float PI2 = (float)(3.14159265358979323846 * 3.14159265358979323846);
  if (metric == YH_METRIC_GIOU) {
    carea = ch * cw;
    gu = carea - outer;
    gv = carea + EPS;
    m = iou - gu / gv;
  } else if (metric == YH_METRIC_DIOU || metric == YH_METRIC_CIOU) {
    diag = ch * ch + cw * cw + EPS;
    dy = p.cy - q.cy;
    dx = p.cx - q.cx;
    cdist = dy * dy + dx * dx;
    qd = cdist / diag;
    m = iou - qd;
    if (metric == YH_METRIC_CIOU) {
      ang_p = atan2f(p.h, p.w);
      ang_q = atan2f(q.h, q.w);
      const float da = ang_p - ang_q;
      shape = (da * da) * 4.0f / PI2;
      alpha = shape / (1.0f - iou + shape);  // no_grad (cycxhw.rs:117)
      m = iou - qd + alpha * shape;          // '+' as in the reference (cycxhw.rs:119)
    }
  }
  if (!want_grad) return m;

  float d_pcy = 0.f, d_pcx = 0.f, d_ph = 0.f, d_pw = 0.f;
  float d_pt = 0.f, d_pb = 0.f, d_pl = 0.f, d_pr = 0.f;
  float d_iou = g;
  float d_outer = 0.f;
  float d_ch = 0.f, d_cw = 0.f;
  if (metric == YH_METRIC_GIOU) {
    // m = iou - gu/gv
    const float d_gu = -g / gv;
    const float d_gv = g * (gu / gv) / gv;
    const float d_carea = d_gu + d_gv;
    d_outer += -d_gu;
    d_ch += d_carea * cw;
    d_cw += d_carea * ch;
  } else if (metric == YH_METRIC_DIOU || metric == YH_METRIC_CIOU) {
    const float d_cdist = -g / diag;
    const float d_diag = g * (cdist / diag) / diag;
    d_ch += d_diag * (2.0f * ch);
    d_cw += d_diag * (2.0f * cw);
    d_pcy += d_cdist * (2.0f * dy);
    d_pcx += d_cdist * (2.0f * dx);
    if (metric == YH_METRIC_CIOU) {
      const float d_shape = g * alpha;
      const float d_ang = d_shape / PI2 * 4.0f * (2.0f * (ang_p - ang_q));
      const float rec = 1.0f / (p.h * p.h + p.w * p.w);  // atan2 backward
      d_ph += d_ang * p.w * rec;
      d_pw += d_ang * (-p.h) * rec;
    }
  }
  // closure edges
  d_pb += d_ch * w_gt(pb, qb);
  d_pt += -d_ch * w_lt(pt, qt);
  d_pr += d_cw * w_gt(pr, qr);
  d_pl += -d_cw * w_lt(pl, ql);
  // iou = inter / outer
  float d_inter = d_iou / outer;
  d_outer += -d_iou * (inter / outer) / outer;
  // outer = ap + aq - inter + eps
  const float d_ap = d_outer;
  d_inter += -d_outer;
  d_ph += d_ap * p.w;
  d_pw += d_ap * p.h;
  // inter = ih * iw, clamp_min backward passes where the input >= 0
  const float d_ih = d_inter * iw, d_iw = d_inter * ih;
  const float d_dhh = dhh >= 0.0f ? d_ih : 0.0f;
  const float d_dww = dww >= 0.0f ? d_iw : 0.0f;
  d_pb += d_dhh * w_lt(pb, qb);
  d_pt += -d_dhh * w_gt(pt, qt);
  d_pr += d_dww * w_lt(pr, qr);
  d_pl += -d_dww * w_gt(pl, ql);
  // pt = cy - h/2 ...
  d_pcy += d_pt + d_pb;
  d_ph += (d_pb - d_pt) / 2.0f;
  d_pcx += d_pl + d_pr;
  d_pw += (d_pr - d_pl) / 2.0f;
  d.cy = d_pcy;
  d.cx = d_pcx;
  d.h = d_ph;
  d.w = d_pw;
  return m;
}

// Hausdorff distance between two boxes (tlbr.rs:137-177 via cycxhw.rs:122-125), forward and the
// reverse mode of the same ATen chain: clamp_min passes the gradient where its input >= 0,
// pow(2) gives 2x, maximum gives everything to the larger operand and half to each on a tie
// (ties are the common case here: most clamped differences are 0), sqrt gives g/(2*sqrt).
// Returns the distance; if want_grad, d receives d(distance)/d(pred) * g.
__device__ __forceinline__ float hausdorff_metric(const Box4 p, const Box4 q, const float g,
                                                  const bool want_grad, Box4& d) {
  const float pt = p.cy - p.h / 2.0f, pb = p.cy + p.h / 2.0f;
  const float pl = p.cx - p.w / 2.0f, pr = p.cx + p.w / 2.0f;
  const float qt = q.cy - q.h / 2.0f, qb = q.cy + q.h / 2.0f;
  const float ql = q.cx - q.w / 2.0f, qr = q.cx + q.w / 2.0f;
  // self = pred, other = target: dt = tr - tl, dl = lr - ll, db = bl - br, dr = rl - rr
  const float dv[4] = {qt - pt, ql - pl, pb - qb, pr - qr};  // t, l, b, r
  float cl[4], cr[4];                                        // clamp_min(d,0), clamp_min(-d,0)
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    cl[k] = fmaxf(dv[k], 0.0f);
    cr[k] = fmaxf(-dv[k], 0.0f);
  }
  // the eight corner terms in the reference's order: (t,l) (t,r) (b,l) (b,r), _l then _r
  const int ia[4] = {0, 0, 2, 2}, ib[4] = {1, 3, 1, 3};
  float T[8];
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    T[k] = cl[ia[k]] * cl[ia[k]] + cl[ib[k]] * cl[ib[k]];
    T[4 + k] = cr[ia[k]] * cr[ia[k]] + cr[ib[k]] * cr[ib[k]];
  }
  float run[8];
  run[0] = T[0];
#pragma unroll
  for (int k = 1; k < 8; ++k) run[k] = fmaxf(run[k - 1], T[k]);
  const float dist = sqrtf(run[7]);
  if (!want_grad) return dist;

  float gT[8];
  float gr = g / (2.0f * dist);  // sqrt backward
#pragma unroll
  for (int k = 7; k >= 1; --k) {
    gT[k] = gr * w_gt(T[k], run[k - 1]);
    gr = gr * w_gt(run[k - 1], T[k]);
  }
  gT[0] = gr;
  float gcl[4] = {0.f, 0.f, 0.f, 0.f}, gcr[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    gcl[ia[k]] += gT[k] * (2.0f * cl[ia[k]]);
    gcl[ib[k]] += gT[k] * (2.0f * cl[ib[k]]);
    gcr[ia[k]] += gT[4 + k] * (2.0f * cr[ia[k]]);
    gcr[ib[k]] += gT[4 + k] * (2.0f * cr[ib[k]]);
  }
  float gd[4];
#pragma unroll
  for (int k = 0; k < 4; ++k)
    gd[k] = (dv[k] >= 0.0f ? gcl[k] : 0.0f) - (-dv[k] >= 0.0f ? gcr[k] : 0.0f);
  const float d_pt = -gd[0], d_pl = -gd[1], d_pb = gd[2], d_pr = gd[3];
  d.cy = d_pt + d_pb;
  d.h = (d_pb - d_pt) / 2.0f;
  d.cx = d_pl + d_pr;
  d.w = (d_pr - d_pl) / 2.0f;
  return dist;
}

// ---------------------------------------------------------------------------------
// K3
// ---------------------------------------------------------------------------------
// A warp owns kRowsPerWarp consecutive matched rows.
//   phase A  lane r < kRowsPerWarp owns row m0+r: index loads, cell location, the five leading
//            logits, decode, box metric forward + reverse, objectness correction — the ~900
//            instruction scalar part is paid once per kRowsPerWarp rows, not once per row;
//   phase B  the whole warp walks the rows: lanes across the C class channels (three per lane
//            up to C = 96, the rest in a tail loop), class BCE forward + backward.
// Every load of both phases is issued before the first value is consumed.
// (the template parameter RPW of loss_rows_kernel: 2, 4 or 8 rows per warp)
constexpr int kClsRegs = 3;

// An image range [b_lo, b_hi) of a batch (yh_loss_*_range): which rows of the (batch, flat)-sorted
// match list belong to it.  img_off (the per-image row offsets yh_assign leaves in its workspace)
// answers with two loads; without it every CTA binary-searches the batch column.
struct RowRange {
  int b_lo, b_hi;      // b_hi < 0: all rows
  const int* img_off;  // [B+1] or null
  int part_off;        // first partial-sum slot of this launch
};

__device__ __forceinline__ int first_row_of_image(const long long* __restrict__ m_batch, const int M,
                                                  const int b) {
  int lo = 0, hi = M;
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldg(m_batch + mid) < (long long)b) lo = mid + 1; else hi = mid;
  }
  return lo;
}

// DIRECT (g.grad[0] != null): the gradients of a row are stored straight over the dense gradient
// instead of being staged for loss_finish_kernel.  Meant for a launch that FOLLOWS the dense pass
// of the same few images while their gradient lines are still in the 126 MB L2: the ~85 4-byte
// stores of a row then land in resident lines (no read-modify-write through DRAM, no staging
// buffer, no scatter kernel).
template <int kRowsPerWarp>
__global__ void __launch_bounds__(kRowThreads)
loss_rows_kernel(const LayerTab t, const GradTab g, const LossParams p, const RowRange rr,
                 const long long* __restrict__ m_batch, const long long* __restrict__ m_flat,
                 const float* __restrict__ m_target, const long long* __restrict__ m_class,
                 const int* __restrict__ M_dev, const int capacity,
                 float* __restrict__ iou_score, float* __restrict__ pred_rows,
                 float* __restrict__ row_grad, double* __restrict__ partials) {
  griddep_launch();
  griddep_wait();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int warps = kRowThreads / 32;
  // The match row of the warp's FIRST pass is requested together with the match count instead
  // of after it (the lists hold `capacity` rows, so the read is in bounds whatever M is; it is
  // used only once m < M is known): one dependent memory round trip less on the step's critical
  // chain assign -> rows -> finish.
  const int m_first = (blockIdx.x * warps + wid) * kRowsPerWarp + lane;
  const bool pre = rr.b_hi < 0 && lane < kRowsPerWarp && m_first < capacity;
  long long pre_b = 0, pre_f = 0, pre_c = 0;
  float4 pre_t = make_float4(0.f, 0.f, 0.f, 0.f);
  if (pre) {
    pre_b = m_batch[m_first];
    pre_f = m_flat[m_first];
    pre_c = m_class[m_first];
    pre_t = *reinterpret_cast<const float4*>(m_target + (size_t)m_first * 4);
  }
  const int M = min(*M_dev, capacity);
  const bool direct = g.grad[0] != nullptr;
  const bool want_grad = row_grad != nullptr || direct;
  int m_begin = 0, m_end = M;
  if (rr.b_hi >= 0) {
    if (rr.img_off != nullptr) {
      m_begin = min(__ldg(rr.img_off + rr.b_lo), M);
      m_end = min(__ldg(rr.img_off + rr.b_hi), M);
    } else {
      __shared__ int s_rng[2];
      if (threadIdx.x < 2) s_rng[threadIdx.x] = first_row_of_image(m_batch, M, threadIdx.x ? rr.b_hi : rr.b_lo);
      __syncthreads();
      m_begin = s_rng[0];
      m_end = s_rng[1];
    }
  }
  const bool haus = p.metric == YH_METRIC_HAUSDORFF;
  const float Mf = (float)M;
  const float MCf = (float)((long long)M * t.C);
  const float numel_obj = (float)((long long)p.B * t.N);

  double a_iou = 0.0, a_cls = 0.0, a_obj = 0.0;

  for (int m0 = m_begin + (blockIdx.x * warps + wid) * kRowsPerWarp; m0 < m_end;
       m0 += gridDim.x * warps * kRowsPerWarp) {
    // ---- phase A, part 1: where is my row -------------------------------------------
    const int m = m0 + lane;
    const bool own = lane < kRowsPerWarp && m < m_end;
    long long off = 0;     // offset of entry 0 of the row inside its head tensor
    int es = 0, l = 0, cls = 0, anchor = 0, row = 0, col = 0;
    float4 tg = make_float4(0.f, 0.f, 0.f, 0.f);
    if (own) {
      const bool first = pre && m == m_first;
      const int b = (int)(first ? pre_b : m_batch[m]);
      const int flat = (int)(first ? pre_f : m_flat[m]);
      cls = (int)(first ? pre_c : m_class[m]);
      tg = first ? pre_t : *reinterpret_cast<const float4*>(m_target + (size_t)m * 4);
      YH_CHECK(b >= 0 && b < p.B && flat >= 0 && flat < t.N && cls >= 0 && cls < t.C);
      const Cell c = locate(t, flat);
      l = c.layer;
      anchor = c.anchor;
      row = c.row;
      col = c.col;
      es = t.A[l] * t.HW[l];
      off = (long long)head_offset(t, l, b, c.anchor, c.cell);
    }
    // ---- all loads ----------------------------------------------------------------------
    float xb[5] = {0.f, 0.f, 0.f, 0.f, 0.f};
    if (own) {
      const float* hp = t.head[l] + off;
#pragma unroll
      for (int e = 0; e < 5; ++e) xb[e] = ld_scatter(hp + (size_t)e * es, t.l2hint);
    }
    float xc[kRowsPerWarp][kClsRegs];
#pragma unroll
    for (int r = 0; r < kRowsPerWarp; ++r) {
      const long long off_r = __shfl_sync(0xffffffffu, off, r);
      const int es_r = __shfl_sync(0xffffffffu, es, r);
      const int l_r = __shfl_sync(0xffffffffu, l, r);
      const float* hp = t.head[l_r] + off_r;
#pragma unroll
      for (int j = 0; j < kClsRegs; ++j) {
        const int k = lane + 32 * j;
        xc[r][j] = (m0 + r < m_end && k < t.C) ? ld_scatter(hp + (size_t)(5 + k) * es_r, t.l2hint) : 0.0f;
      }
    }

    // ---- phase A, part 2: box metric, objectness correction, their gradients ----------
    if (own) {
      const float s0 = sigmoid_f(xb[0]), s1 = sigmoid_f(xb[1]);
      const float s2 = sigmoid_f(xb[2]), s3 = sigmoid_f(xb[3]);
      const float xo = xb[4];
      Box4 pb;
      pb.cy = decode_center(s0, row, (float)t.H[l]);
      pb.cx = decode_center(s1, col, (float)t.W[l]);
      pb.h = decode_size(s2, t.ah[l][anchor]);
      pb.w = decode_size(s3, t.aw[l][anchor]);
      Box4 tb;
      tb.cy = tg.x; tb.cx = tg.y; tb.h = tg.z; tb.w = tg.w;
      // d total / d metric: total = w_iou * mean(1 - metric); Hausdorff: w_iou * mean(distance)
      float gm = p.w_iou;
      if (p.reduction == YH_REDUCTION_MEAN) gm = gm / Mf;
      if (!haus) gm = -gm;
      Box4 d = {0.f, 0.f, 0.f, 0.f};
      const float metric = haus ? hausdorff_metric(pb, tb, gm, want_grad, d)
                                : box_metric(p.metric, pb, tb, gm, want_grad, d);
      a_iou += haus ? (double)metric : (double)(1.0f - metric);
      // objectness target at this cell (loss_.rs:397-406); Hausdorff has no score (loss_.rs:285-288)
      const float to = haus ? p.obj_base
                            : p.obj_base + fminf(fmaxf(metric, 0.0f), 1.0f) * p.obj_coef;
      float dxo = 0.0f;  // d objectness-loss / d xo for the non-BCE kinds
      if (p.obj_kind == YH_LOSS_BCE) {
        const float lwm1 = p.has_pw ? (p.pw - 1.0f) * to : 0.0f;
        a_obj += (double)(-to * xo) + (double)(lwm1 * softplus_f(-xo));
      } else {
        // K4 summed loss(x, 0) for this cell: replace it by loss(x, t)
        float d0;
        const float l0 = kind_loss(p.obj_kind, xo, 0.0f, false, 0.0f, p.gamma, d0);
        const float lt = kind_loss(p.obj_kind, xo, to, p.has_pw != 0, p.pw, p.gamma, dxo);
        a_obj += (double)lt - (double)l0;
      }
      if (iou_score) iou_score[m] = metric;
      if (pred_rows) {
        float* r = pred_rows + (size_t)m * t.E;
        r[0] = pb.cy; r[1] = pb.cx; r[2] = pb.h; r[3] = pb.w; r[4] = xo;
      }
      if (want_grad) {
        // staged per row ([M, 5+C], coalesced); loss_finish_kernel scatters it into the heads —
        // or, DIRECT, entry e of the row goes to its place in the dense gradient (stride es)
        float* gp = direct ? g.grad[l] + off : row_grad + (size_t)m * t.E;
        const size_t gs = direct ? (size_t)es : (size_t)1;
        // chain through the decode (detect_2d.rs:69-72,122-133):
        //   cy = (2s-0.5)/H + off  -> dcy/ds = 2/H ; s' = s(1-s)
        //   h  = (2s)^2 * a        -> dh/ds  = a * 2*(2s) * 2
        const float Hf = (float)t.H[l], Wf = (float)t.W[l];
        gp[0] = (d.cy / Hf * 2.0f) * ((1.0f - s0) * s0);
        gp[gs] = (d.cx / Wf * 2.0f) * ((1.0f - s1) * s1);
        gp[2 * gs] = ((d.h * t.ah[l][anchor]) * (2.0f * (s2 * 2.0f)) * 2.0f) * ((1.0f - s2) * s2);
        gp[3 * gs] = ((d.w * t.aw[l][anchor]) * (2.0f * (s3 * 2.0f)) * 2.0f) * ((1.0f - s3) * s3);
        // objectness gradient with the real target (replaces K4's target-0 value)
        const float so = sigmoid_f(xo);
        float gi;
        if (p.obj_kind != YH_LOSS_BCE) {
          gi = dxo;
        } else if (p.has_pw) {
          const float tp = p.pw * to;
          gi = ((tp + 1.0f) - to) * so - tp;
        } else {
          gi = so - to;
        }
        if (p.reduction == YH_REDUCTION_MEAN) gi = gi / numel_obj;
        gp[4 * gs] = gi * p.w_obj;
      }
    }

    // ---- phase B: class BCE over C channels, lanes strided ------------------------------
#pragma unroll
    for (int r = 0; r < kRowsPerWarp; ++r) {
      const long long off_r = __shfl_sync(0xffffffffu, off, r);
      const int es_r = __shfl_sync(0xffffffffu, es, r);
      const int l_r = __shfl_sync(0xffffffffu, l, r);
      const int cls_r = __shfl_sync(0xffffffffu, cls, r);
      if (m0 + r >= m_end) break;  // warp-uniform
      const float* hp = t.head[l_r] + off_r;
      float* gp = !want_grad ? nullptr
                             : (direct ? g.grad[l_r] + off_r : row_grad + (size_t)(m0 + r) * t.E);
      const size_t gs = direct ? (size_t)es_r : (size_t)1;
      float bce = 0.f;
      // CrossEntropyLoss (cross_entropy.rs:19-58): the dense target's arg-max (first maximum:
      // the class, or 0 when pos == neg), log_softmax + nll per ROW
      const int ce_t = (p.cls_pos > p.cls_neg) ? cls_r : 0;
      float lse = 0.f;
      if (p.cls_kind == YH_LOSS_CROSS_ENTROPY) {
        float mx = -INFINITY;
        for (int k = lane, j = 0; k < t.C; k += 32, ++j) {
          const float x = j == 0 ? xc[r][0] : j == 1 ? xc[r][1] : j == 2 ? xc[r][2]
                                                                        : hp[(size_t)(5 + k) * es_r];
          mx = fmaxf(mx, x);
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) mx = fmaxf(mx, __shfl_xor_sync(0xffffffffu, mx, o));
        float se = 0.f;
        for (int k = lane, j = 0; k < t.C; k += 32, ++j) {
          const float x = j == 0 ? xc[r][0] : j == 1 ? xc[r][1] : j == 2 ? xc[r][2]
                                                                        : hp[(size_t)(5 + k) * es_r];
          se += expf(x - mx);
        }
        se = warp_sum(se);
        lse = mx + logf(se);
      }
      for (int k = lane, j = 0; k < t.C; k += 32, ++j) {
        float x;
        if (j == 0) x = xc[r][0];
        else if (j == 1) x = xc[r][1];
        else if (j == 2) x = xc[r][2];
        else x = hp[(size_t)(5 + k) * es_r];
        const float tt = (k == cls_r) ? p.cls_pos : p.cls_neg;
        float gi;
        if (p.cls_kind == YH_LOSS_BCE) {
          bce += (1.0f - tt) * x + softplus_f(-x);
          gi = sigmoid_f(x) - tt;
          if (p.reduction == YH_REDUCTION_MEAN) gi = gi / MCf;
        } else if (p.cls_kind == YH_LOSS_CROSS_ENTROPY) {
          if (k == ce_t) bce += lse - x;
          gi = expf(x - lse) - (k == ce_t ? 1.0f : 0.0f);
          // tch 0.7's cross_entropy_for_logits is log_softmax + nll_loss with nll_loss's DEFAULT
          // Mean reduction: already a scalar mean over the M rows, which cross_entropy.rs:47-51's
          // .sum() / .mean() leave unchanged — 1/M for BOTH reductions
          gi = gi / Mf;
        } else {
          bce += kind_loss(p.cls_kind, x, tt, false, 0.0f, p.gamma, gi);
          if (p.reduction == YH_REDUCTION_MEAN) gi = gi / MCf;
        }
        if (want_grad) gp[(size_t)(5 + k) * gs] = gi * p.w_cls;
        if (pred_rows) pred_rows[(size_t)(m0 + r) * t.E + 5 + k] = x;
      }
      bce = warp_sum(bce);
      a_cls += (double)bce;  // identical in every lane; lane 0's copy is used
    }
  }
  // rows were owned by lanes 0..kRowsPerWarp-1: fold them into lane 0 in a fixed order
  a_iou = warp_sum_d(a_iou);
  a_obj = warp_sum_d(a_obj);

  // CTA partials -> fixed slot (summed in a fixed order by loss_finish_kernel)
  __shared__ double s_acc[kRowThreads / 32][3];
  if (lane == 0) {
    s_acc[wid][0] = a_iou;
    s_acc[wid][1] = a_cls;
    s_acc[wid][2] = a_obj;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    double s[3] = {0.0, 0.0, 0.0};
    for (int k = 0; k < warps; ++k) {
      s[0] += s_acc[k][0];
      s[1] += s_acc[k][1];
      s[2] += s_acc[k][2];
    }
    double* dst = partials + (size_t)(kPartHalf + rr.part_off + blockIdx.x) * 4;
    dst[0] = s[0];
    dst[1] = s[1];
    dst[2] = s[2];
  }
}

// ---------------------------------------------------------------------------------
// K3b: ordered after K4 AND K3.  Scatters the staged row gradients into the head gradient
// (overwriting K4's values at the matched cells: pure fire-and-forget 4-byte stores, a warp per
// row, lanes across the 5+C entries) and, in CTA 0, sums all partials in a fixed order into the
// four loss scalars.  Splitting this off K3 lets K3 (latency-bound gathers) run UNDER K4
// (HBM-bound stream) instead of after it.
// ---------------------------------------------------------------------------------
constexpr int kFinishThreads = 256;

__global__ void __launch_bounds__(kFinishThreads)
loss_finish_kernel(const LayerTab t, const GradTab g, const LossParams p,
                   const long long* __restrict__ m_batch, const long long* __restrict__ m_flat,
                   const int* __restrict__ M_dev, const int capacity,
                   const float* __restrict__ row_grad, const double* __restrict__ partials,
                   const int dense_per_range, const int rows_per_range, const int num_ranges,
                   const int num_row_ranges, const int range_stride, const int layer_lo,
                   const int layer_hi, const int do_reduce, const int early, const int split_layer,
                   float* __restrict__ losses4, int* __restrict__ err_flag) {
  // early (yh_loss_finish_after_dense): the stream's previous kernel is the dense pass and
  // everything this kernel READS (match list, staged row gradients) was complete before it was
  // enqueued, so with the PDL attribute its CTAs come in while the last wave of the dense pass
  // drains, load their rows, and only then wait for the dense pass: the stores go out the moment
  // it has finished instead of a launch + three dependent round trips later (this kernel is the
  // ~24 us tail of the overlapped step).
  griddep_launch();
  if (!early) griddep_wait();
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int warps = kFinishThreads / 32;
  const int M = min(*M_dev, capacity);
  if (g.grad[0] != nullptr && row_grad != nullptr) {
    // The scatter is bound by the NUMBER of store requests (one per matched element, each to its
    // own sector), not by bytes.  Matched cells come in clusters (a label takes a cell and its
    // left/right/top/bottom neighbours) and the match list is sorted by (image, flat), so rows
    // m-3..m+3 are the only ones that can share an aligned 4-float quad with row m: the row with
    // the lowest flat of a quad stores the whole quad (its members' values, zeros elsewhere —
    // what loss_dense wrote there) with ONE 16-byte store and the other members store nothing.
    // Not for the objectness entry (its neighbours hold sigmoid values, not zeros) and not for a
    // quad that crosses the end of its entry run: those elements are stored one by one.
    // split_layer (yh_loss_dense kept the end of its store stream in the L2): two sweeps over the
    // list, first the rows of the layers >= split_layer — written last, still resident — then
    // the others, whose line fills would otherwise push the resident lines out before their turn
    const bool two = split_layer > layer_lo && split_layer < layer_hi && split_layer < t.L;
    for (int sweep = 0; sweep < (two ? 2 : 1); ++sweep) {
    const int s_lo = two ? (sweep == 0 ? split_layer : layer_lo) : layer_lo;
    const int s_hi = two ? (sweep == 0 ? layer_hi : split_layer) : layer_hi;
    for (int m = blockIdx.x * warps + wid; m < M; m += gridDim.x * warps) {
      const int kk = lane - 3;  // lanes 0..6 look at rows m-3 .. m+3
      long long nb = -1;
      int nflat = 0;
      if (lane < 7 && m + kk >= 0 && m + kk < M) {
        nb = __ldg(m_batch + m + kk);
        nflat = (int)__ldg(m_flat + m + kk);
      }
      const int b = (int)__shfl_sync(0xffffffffu, nb, 3);
      const int flat = __shfl_sync(0xffffffffu, nflat, 3);
      YH_CHECK(b >= 0 && b < p.B && flat >= 0 && flat < t.N);
      const Cell c = locate(t, flat);
      const int l = c.layer;
      if (l < s_lo || l >= s_hi) continue;  // another sweep's / another launch's layers (yh_loss_finish_layers)
      const int esi = t.A[l] * t.HW[l];
      const size_t es = (size_t)esi;
      const int j = flat - t.start[l];  // position inside the entry run
      const int dk = nflat - flat;
      const bool near = lane < 7 && nb == (long long)b && dk >= -3 && dk <= 3 &&
                        nflat >= t.start[l] && nflat < t.start[l] + esi;
      const unsigned near_mask = __ballot_sync(0xffffffffu, near);  // bit 3 = this row
      // the merge below relies on the order yh_assign emits: report anything else loudly
      // (yh_read_counts -> YH_ERR_INVALID) instead of scattering wrong values quietly
      if (lane == 2 && nb >= 0 && (nb > (long long)b || (nb == (long long)b && nflat >= flat)))
        atomicOr(err_flag, 8);
      int d[7];  // flat distance of rows m-3 .. m+3 (meaningful where near_mask has the bit)
#pragma unroll
      for (int k = 0; k < 7; ++k) d[k] = __shfl_sync(0xffffffffu, dk, k);
      float* gp = g.grad[l] + head_offset(t, l, b, c.anchor, c.cell);
      const float* src = row_grad + (size_t)m * t.E;
      // All loads of a row first, then the stores: the staged values of this row at its (up to)
      // three entries per lane AND those of rows m+1..m+3 wherever they are neighbours (needed
      // only where they share a quad, fetched regardless: 12 independent loads instead of three
      // rounds of load -> test -> dependent load).  This kernel is latency-bound (ncu: DRAM 16 %,
      // SM 20 % busy, r1f) and runs ALONE at the tail of the overlapped step.
      for (int e0 = 0; e0 < t.E; e0 += 96) {
        float v[3], vn[3][3];
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const int e = e0 + lane + 32 * i;
          const bool ok = e < t.E;
          v[i] = ok ? __ldcs(src + e) : 0.0f;
#pragma unroll
          for (int k = 0; k < 3; ++k)
            vn[i][k] = (ok && ((near_mask >> (4 + k)) & 1u)) ? __ldcs(src + (size_t)(k + 1) * t.E + e) : 0.0f;
        }
        if (early) griddep_wait();  // from here on: stores over what the dense pass wrote
#pragma unroll
        for (int i = 0; i < 3; ++i) {
          const int e = e0 + lane + 32 * i;
          if (e >= t.E) continue;
          const long long q = ((long long)b * t.E + e) * (long long)es + j;  // float index in the tensor
          const int a = (int)(q & 3);
          if (e == 4 || j - a < 0 || j - a + 3 >= esi) {
            YH_CHECK(q >= 0 && q < (long long)p.B * t.E * esi);
            gp[(size_t)e * es] = v[i];
            continue;
          }
          // members of this element's quad among the neighbouring rows: 0 <= a + d <= 3
          float o[4] = {0.f, 0.f, 0.f, 0.f};
#pragma unroll
          for (int c4 = 0; c4 < 4; ++c4)
            if (c4 == a) o[c4] = v[i];
          bool lead = true;
#pragma unroll
          for (int k = 0; k < 7; ++k) {
            if (k == 3) continue;
            const int pos = a + d[k];
            const bool in = ((near_mask >> k) & 1u) && pos >= 0 && pos <= 3;
            if (in && k < 3) lead = false;  // a row with a lower flat stores the quad
            if (in && k > 3) {
              const float vk = vn[i][k > 3 ? k - 4 : 0];
#pragma unroll
              for (int c4 = 0; c4 < 4; ++c4)
                if (c4 == pos) o[c4] = vk;
            }
          }
          if (lead) {
            YH_CHECK(q - a >= 0 && q - a + 3 < (long long)p.B * t.E * esi && ((q - a) & 3) == 0);
            *reinterpret_cast<float4*>(g.grad[l] + (q - a)) = make_float4(o[0], o[1], o[2], o[3]);
          }
        }
      }
    }
    }
  }
  if (blockIdx.x != 0 || !do_reduce) return;
  if (early) griddep_wait();  // the dense pass's partial sums

  // CTA 0: fixed-order final sums
  __shared__ double s_red[kFinishThreads][4];
  // every thread sums a strided subset; the loads of a thread are independent of its sums,
  // so they are issued kFinalUnroll at a time (the association is still a fixed one)
  constexpr int kFinalUnroll = 8;
  // launch r of a ranged loss (yh_loss_*_range) wrote its partials from slot r * range_stride
  const int n_dense_partials = dense_per_range * num_ranges;
  const int n_row_partials = rows_per_range * num_row_ranges;
  double r[4] = {0.0, 0.0, 0.0, 0.0};
  for (int i0 = threadIdx.x; i0 < n_dense_partials; i0 += kFinishThreads * kFinalUnroll) {
    double v[kFinalUnroll];
#pragma unroll
    for (int u = 0; u < kFinalUnroll; ++u) {
      const int i = i0 + u * kFinishThreads;
      const int slot = (i / dense_per_range) * range_stride + i % dense_per_range;
      v[u] = (i < n_dense_partials) ? __ldcg(&partials[(size_t)slot * 4]) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < kFinalUnroll; ++u) r[3] += v[u];
  }
  for (int i0 = threadIdx.x; i0 < n_row_partials; i0 += kFinishThreads * kFinalUnroll) {
    double v[kFinalUnroll][3];
#pragma unroll
    for (int u = 0; u < kFinalUnroll; ++u) {
      const int i = i0 + u * kFinishThreads;
      const int slot = (i / rows_per_range) * range_stride + i % rows_per_range;
      const double* src = partials + (size_t)(kPartHalf + slot) * 4;
#pragma unroll
      for (int k = 0; k < 3; ++k) v[u][k] = (i < n_row_partials) ? __ldcg(&src[k]) : 0.0;
    }
#pragma unroll
    for (int u = 0; u < kFinalUnroll; ++u) {
      r[0] += v[u][0];
      r[1] += v[u][1];
      r[2] += v[u][2];
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) s_red[threadIdx.x][k] = r[k];
  __syncthreads();
  for (int o = kFinishThreads / 2; o > 0; o >>= 1) {
    if (threadIdx.x < o) {
#pragma unroll
      for (int k = 0; k < 4; ++k) s_red[threadIdx.x][k] += s_red[threadIdx.x + o][k];
    }
    __syncthreads();
  }
  if (threadIdx.x == 0) {
    double iou_sum = s_red[0][0], cls_sum = s_red[0][1];
    double obj_sum = s_red[0][3] + s_red[0][2];
    float iou_l, cls_l, obj_l;
    if (p.reduction == YH_REDUCTION_MEAN) {
      // empty inputs give 0 (loss_.rs:309-314, bce_with_logits_loss.rs:74-77)
      iou_l = M > 0 ? (float)(iou_sum / (double)M) : 0.0f;
      const double cls_n = (p.cls_kind == YH_LOSS_CROSS_ENTROPY) ? (double)M
                                                                 : (double)M * (double)t.C;
      cls_l = M > 0 ? (float)(cls_sum / cls_n) : 0.0f;
      obj_l = (float)(obj_sum / ((double)p.B * (double)t.N));
    } else {
      iou_l = (float)iou_sum;
      // CrossEntropy: nll_loss's own Mean reduction has already happened (see loss_rows_kernel);
      // M = 0 gives the mean of an empty tensor, NaN, exactly as cross_entropy.rs:41-51 does
      cls_l = (p.cls_kind == YH_LOSS_CROSS_ENTROPY) ? (float)(cls_sum / (double)M) : (float)cls_sum;
      obj_l = (float)obj_sum;
    }
    // loss_.rs:258-260, f32 tensor arithmetic with f32-cast scalars
    const float total = p.w_iou * iou_l + p.w_cls * cls_l + p.w_obj * obj_l;
    losses4[0] = total;
    losses4[1] = iou_l;
    losses4[2] = cls_l;
    losses4[3] = obj_l;
  }
}

}  // namespace yh

namespace yh {

enum { kPhaseDense = 1, kPhaseRows = 2, kPhaseFinish = 4 };

// One launch of a RANGED loss: images [b_lo, b_hi) of a batch of B, launch `index` of `count`.
// count == 0: the plain whole-batch call.
struct LossRange {
  int b_lo, b_hi, index, count;
  const int32_t* img_off;  // per-image row offsets of the match list ([B+1], device) or null
  int direct;              // rows phase: scatter straight into grad_heads (no staging, no finish scatter)
  int layer_lo, layer_hi;  // dense / finish phase: only these layers (layer_hi <= 0: all)
  int finish_reduce;       // finish phase: also reduce the partial sums (the LAST finish launch)
  int row_ranges;          // finish phase: launches of the rows phase (1: one whole-batch launch)
  int finish_early;        // finish phase: yh_loss_finish_after_dense (loads before the PDL wait)
};
constexpr int kMaxRanges = 16;
constexpr int kRangeSlots = kPartHalf / kMaxRanges;  // partial-sum slots per ranged launch

// How much of the END of the dense store stream is written with the L2 evict_last priority
// (YH_KEEP_MB, default below): the tiles of the small layers, written last, where most matched
// rows lie (71 % in layers 1-2 of YOLOv4-608, 118 MB at 64 images against 126 MB of L2).
// tools/gather_bench.cu on B200: the scatter of 11 788 rows behind a 495 MB fill takes 33 us
// after streaming stores, 23 us when the last 118 MB were stored evict_last and are scattered
// first; 1 M stores into 64 MB just written: 10.8 us (93 G/s against 24 G/s into evicted lines).
static long long keep_bytes() {
  static const long long v = []() {
    const char* e = getenv("YH_KEEP_MB");
    const long long mb = e ? atoll(e) : YH_KEEP_MB_DEFAULT;
    return (mb > 0 ? mb : 0) << 20;
  }();
  return v;
}
// Only when the gradient is several times what is kept: a gradient that fits the L2 anyway (up to
// ~16 images of YOLOv4-608) gains nothing, and its latency-bound step loses ~3 us to the slower
// stores and the second sweep of the scatter (35.2 -> 38.4 us at 8 images, r2g)
static bool keep_active(const long long dense_bytes) {
  return keep_bytes() > 0 && dense_bytes >= 3 * keep_bytes();
}
static long long keep_tiles(const long long dense_tiles) {
  if (!keep_active(dense_tiles * kTileFloats * 4)) return 0;
  return keep_bytes() / ((long long)kTileFloats * 4);
}
// first layer of which at least a quarter lies in the kept tail: the scatter takes the rows of
// the layers from there on first (YH_MAX_LAYERS: no split)
static int keep_split_layer(const LayerTab& tab, const int B) {
  long long left = keep_bytes();
  int split = YH_MAX_LAYERS;
  long long total = 0;
  for (int l = 0; l < tab.L; ++l) total += (long long)B * tab.A[l] * tab.HW[l] * tab.E * 4;
  if (!keep_active(total)) return split;
  for (int l = tab.L - 1; l >= 1 && left > 0; --l) {
    const long long bytes = (long long)B * tab.A[l] * tab.HW[l] * tab.E * 4;
    if (left * 4 >= bytes) split = l;
    left -= bytes;
  }
  return split;
}

// The three phases of YoloLoss::forward+backward.
//   kPhaseDense   needs only the heads;
//   kPhaseRows    needs the heads and the matches (yh_assign); stages the row gradients in the
//                 workspace, so it may run concurrently with kPhaseDense on another stream — or,
//                 ranged + direct, stores them over the dense gradient of the same images (then
//                 it is ordered after that range's kPhaseDense);
//   kPhaseFinish  ordered after both: scatters the staged rows over the dense gradient (not
//                 after direct rows) and reduces the partial sums into losses4.
static int loss_phases(int phases, YhWorkspace* ws, const float* const* heads,
                       const YhLayer* layers, int L, int B, int C, const int64_t* m_batch,
                       const int64_t* m_flat, const float* m_target, const int64_t* m_class,
                       const int32_t* M_dev, int capacity, const YhLossCfg* cfg,
                       float* losses4_dev, float* iou_score_dev, float* pred_rows_dev,
                       float* const* grad_heads, int want_grad_rows, void* stream,
                       const LossRange* range = nullptr) {
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  YH_REQUIRE(cfg != nullptr, "NULL argument");
  if (phases & kPhaseRows) {
    YH_REQUIRE(M_dev != nullptr, "NULL argument");
    YH_REQUIRE(capacity == 0 || (m_batch && m_flat && m_target && m_class), "match list is NULL");
    YH_REQUIRE(((uintptr_t)m_target & 15) == 0, "m_target not 16-byte aligned");
  }
  if (phases & kPhaseFinish) {
    YH_REQUIRE(losses4_dev != nullptr && M_dev != nullptr, "NULL argument");
    YH_REQUIRE(capacity == 0 || grad_heads == nullptr || (m_batch && m_flat), "match list is NULL");
  }
  YH_REQUIRE(capacity >= 0, "capacity=%d", capacity);
  YH_REQUIRE(B >= 1 && B <= ws->max_batch, "B=%d outside 1..max_batch=%d", B, ws->max_batch);
  const bool ranged = range != nullptr && range->count > 0;
  const int num_ranges = ranged ? range->count : 1;
  if (ranged) {
    YH_REQUIRE(range->count <= kMaxRanges, "num_ranges=%d outside 1..%d", range->count, kMaxRanges);
    if (!(phases & kPhaseFinish)) {
      YH_REQUIRE(range->index >= 0 && range->index < range->count, "range_index=%d outside 0..%d",
                 range->index, range->count - 1);
      YH_REQUIRE(range->b_lo >= 0 && range->b_lo < range->b_hi && range->b_hi <= B,
                 "image range [%d,%d) outside the batch of %d", range->b_lo, range->b_hi, B);
    }
  }
  LayerTab tab;
  int rc = build_layer_tab(layers, L, C, heads, &tab);
  if (rc) return rc;
  // YoloLossInit::build validation (loss_.rs:64-87)
  YH_REQUIRE(cfg->smooth_classification_coef >= 0.0 && cfg->smooth_classification_coef <= 1.0,
             "smooth_classification_coef must be in range [0, 1]");
  YH_REQUIRE(cfg->smooth_objectness_coef >= 0.0 && cfg->smooth_objectness_coef <= 1.0,
             "smooth_objectness_coef must be in range [0, 1]");
  YH_REQUIRE(cfg->iou_loss_weight >= 0.0, "iou_loss_weight must be non-negative");
  YH_REQUIRE(cfg->objectness_loss_weight >= 0.0, "objectness_loss_weight must be non-negative");
  YH_REQUIRE(cfg->classification_loss_weight >= 0.0,
             "classification_loss_weight must be non-negative");
  YH_REQUIRE(cfg->focal_loss_gamma >= 0.0, "focal_loss_gamma must be non-negative");
  YH_REQUIRE(cfg->objectness_loss_kind >= YH_LOSS_BCE && cfg->objectness_loss_kind <= YH_LOSS_L2,
             "objectness_loss_kind=%d", cfg->objectness_loss_kind);
  YH_REQUIRE(cfg->classification_loss_kind >= YH_LOSS_BCE &&
                 cfg->classification_loss_kind <= YH_LOSS_CROSS_ENTROPY,
             "classification_loss_kind=%d", cfg->classification_loss_kind);
  YH_REQUIRE(cfg->reduction == YH_REDUCTION_MEAN || cfg->reduction == YH_REDUCTION_SUM,
             "reduction %d is not supported", cfg->reduction);
  YH_REQUIRE(cfg->box_metric >= YH_METRIC_IOU && cfg->box_metric <= YH_METRIC_HAUSDORFF,
             "box_metric=%d", cfg->box_metric);

  GradTab g;
  memset(&g, 0, sizeof(g));
  if (grad_heads) {
    for (int l = 0; l < L; ++l) {
      YH_REQUIRE(grad_heads[l] != nullptr, "grad_heads[%d] is NULL", l);
      YH_REQUIRE(((uintptr_t)grad_heads[l] & 15) == 0, "grad_heads[%d] not 16-byte aligned", l);
      YH_REQUIRE(!heads || ((uintptr_t)heads[l] & 15) == 0, "heads[%d] not 16-byte aligned", l);
      g.grad[l] = grad_heads[l];
    }
  }
  LossParams p;
  p.B = B;
  p.metric = cfg->box_metric;
  p.reduction = cfg->reduction;
  p.w_iou = (float)cfg->iou_loss_weight;
  p.w_cls = (float)cfg->classification_loss_weight;
  p.w_obj = (float)cfg->objectness_loss_weight;
  const double pos = 1.0 - 0.5 * cfg->smooth_classification_coef;
  p.cls_pos = (float)pos;
  p.cls_neg = (float)(1.0 - pos);
  p.obj_base = (float)(1.0 - cfg->smooth_objectness_coef);
  p.obj_coef = (float)cfg->smooth_objectness_coef;
  p.has_pw = cfg->has_objectness_pos_weight ? 1 : 0;
  p.pw = (float)cfg->objectness_pos_weight;
  p.obj_kind = cfg->objectness_loss_kind;
  p.cls_kind = cfg->classification_loss_kind;
  p.gamma = (float)cfg->focal_loss_gamma;

  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(ws->device));
  note_batch(B);
  const float numel = (float)((long long)B * tab.N);
  // objectness CTAs (and the whole grid of the forward-only form) per SM.  Sweep on B200
  // (profiles/README.md, r1f): 1 -> 78.4 us, 2 -> 77.1, 3 -> 77.9, 4 -> 79.1, 8 -> 80.6; tiles of
  // 16 KB 105 us (the CTA prologue shows), 32 KB 77.7, 64 KB 77.1.
  static const int dense_cps = []() {  // tuning knob for sweeps
    const char* e = getenv("YH_DENSE_CTAS_PER_SM");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : 2;
  }();
  const int slot_room = ranged ? kRangeSlots : kPartHalf;
  int dense_grid = ws->num_sms * dense_cps;
  if (dense_grid > slot_room) dense_grid = slot_room;
  int row_grid = ws->num_sms * 4;  // resident in one wave; rows beyond that grid-stride
  // rows per warp.  The scalar box part runs one row per lane, the class part lanes across C:
  // more rows per warp amortise the first, fewer put more warps on the same rows (the kernel is
  // latency-bound).  Inside the overlapped step at 64 images (r2c, same box): 8 -> 177.0 us per
  // step, 4 -> 167.1, 2 -> 176.5 (alone: 37.7 / 37.1 / 35.1 us); a batch of <= 16 images has
  // ~200 rows per image and wants every warp it can get: 2 (18.3 -> 9.4 us at 8 images).
  static const int rpw_env = []() {  // sweep knob
    const char* e = getenv("YH_ROWS_PER_WARP");
    const int v = e ? atoi(e) : 0;
    return (v == 2 || v == 4 || v == 8) ? v : 0;
  }();
  const int rpw = rpw_env ? rpw_env : (B <= 16 ? 2 : 4);
  const int rows_per_cta = (kRowThreads / 32) * rpw;
  const int need = (capacity + rows_per_cta - 1) / rows_per_cta;
  if (row_grid > need) row_grid = need;
  if (row_grid < 1) row_grid = 1;
  const bool rows_whole = ranged && range->row_ranges == 1;  // layered finish after a plain rows launch
  if (row_grid > (rows_whole ? kPartHalf : slot_room)) row_grid = rows_whole ? kPartHalf : slot_room;
  const int part_off = ranged ? range->index * kRangeSlots : 0;

  if (phases & kPhaseDense) {
    // the images of this launch: the whole batch, or [b_lo, b_hi) of it — the same kernel on
    // pointers advanced by b_lo images (the loss is normalised by the WHOLE batch: numel)
    const int b_lo = ranged ? range->b_lo : 0;
    const int Bd = ranged ? range->b_hi - range->b_lo : B;
    // ... and/or only the layers [layer_lo, layer_hi) (yh_loss_dense_layers): the kernel walks
    // compacted tables, so a subset is just a shorter layer list
    const int l_lo = (ranged && range->layer_hi > 0) ? range->layer_lo : 0;
    const int l_hi = (ranged && range->layer_hi > 0) ? range->layer_hi : L;
    YH_REQUIRE(l_lo >= 0 && l_lo < l_hi && l_hi <= L, "layers [%d,%d) outside 0..%d", l_lo, l_hi, L);
    LayerTab dt_tab = tab;
    GradTab dg;
    memset(&dg, 0, sizeof(dg));
    DenseTab dt;
    memset(&dt, 0, sizeof(dt));
    dt.L = l_hi - l_lo;
    dt_tab.L = l_hi - l_lo;
    for (int k = 0; k < YH_MAX_LAYERS; ++k) {
      long long tiles = 0, oq = 0;
      const int l = l_lo + k;
      if (l < l_hi) {
        dt.run[k] = (long long)tab.A[l] * tab.HW[l];
        const long long shift = (long long)b_lo * dt.run[k] * tab.E;  // floats
        YH_REQUIRE(shift % 4 == 0,
                   "b_lo=%d: layer %d has %lld floats per image, an image range must start at a "
                   "16-byte boundary of every head tensor (use a multiple of 4 images)",
                   b_lo, l, dt.run[k] * tab.E);
        dt_tab.H[k] = tab.H[l];
        dt_tab.W[k] = tab.W[l];
        dt_tab.A[k] = tab.A[l];
        dt_tab.HW[k] = tab.HW[l];
        dt_tab.head[k] = tab.head[l] + shift;
        if (grad_heads) dg.grad[k] = g.grad[l] + shift;
        const long long quads = ((long long)Bd * dt.run[k] * tab.E + 3) / 4;
        tiles = (quads + kTileQuads - 1) / kTileQuads;
        oq = (long long)Bd * (dt.run[k] / 4 + 2);
      } else {
        dt_tab.head[k] = nullptr;
      }
      dt.tile_prefix[k + 1] = dt.tile_prefix[k] + tiles;
      dt.qprefix[k + 1] = dt.qprefix[k] + oq;
    }
    // with a gradient: dense_grid objectness CTAs followed by one CTA per 32 KB tile
    const long long dense_tiles = dt.tile_prefix[YH_MAX_LAYERS];
    dt.keep_from = dense_tiles - ((ranged || !grad_heads) ? 0 : keep_tiles(dense_tiles));
    YH_REQUIRE(dense_tiles + dense_grid < (1LL << 31), "head tensors too large for one launch");
    const bool bce = p.obj_kind == YH_LOSS_BCE;
    double* part = ws->partials + (size_t)part_off * 4;
    YH_TIME(YH_K_LOSS_DENSE, s);
    // PDL only behind another dense launch of the same step (a later range / layer part).  The
    // kernel works BEFORE it waits, so as the first kernel of a call it must not be allowed to
    // overlap whatever the caller ran last: a captured step whose root node carried the attribute
    // read labels the preceding copy had not made visible yet (r2c: replay on refilled inputs).
    const bool dense_pdl = ranged && range->index > 0;
#define YH_DENSE(G, K)                                                                          \
  YH_CUDA(launch_pdl_if(dense_pdl, loss_dense_kernel<G, K>,                                      \
                        dim3((unsigned)(dense_grid + (G ? dense_tiles : 0))), dim3(kDenseThreads), 0, \
                        s, dt_tab, dg, Bd, numel, p.w_obj, p.reduction, p.obj_kind, p.gamma, part, \
                        dense_grid, dt))
    if (grad_heads) {
      if (bce) YH_DENSE(true, true); else YH_DENSE(true, false);
    } else {
      if (bce) YH_DENSE(false, true); else YH_DENSE(false, false);
    }
#undef YH_DENSE
    YH_LAUNCHED(s);
  }
  const bool direct = ranged && range->direct && grad_heads != nullptr;
  const bool want_grad = grad_heads != nullptr || want_grad_rows != 0;
  const bool staged = want_grad && !direct && !(ranged && (phases & kPhaseFinish) && !grad_heads);
  if (staged && (phases & (kPhaseRows | kPhaseFinish)) && capacity > 0 &&
      (size_t)capacity * (size_t)tab.E > ws->row_grad_floats) {
    // no allocation and no synchronisation inside a compute entry point (yolohot.h conventions)
    set_error("row-gradient staging holds %zu floats, capacity=%d rows of %d need %zu: call "
              "yh_workspace_reserve_rows(ws, capacity, C) first",
              ws->row_grad_floats, capacity, tab.E, (size_t)capacity * (size_t)tab.E);
    return YH_ERR_CAPACITY;
  }
  if (phases & kPhaseRows) {
    RowRange rr;
    rr.b_lo = ranged ? range->b_lo : 0;
    rr.b_hi = ranged ? range->b_hi : -1;
    rr.img_off = ranged ? range->img_off : nullptr;
    rr.part_off = part_off;
    GradTab rg;
    memset(&rg, 0, sizeof(rg));
    if (direct) rg = g;
    YH_TIME(YH_K_LOSS_ROWS, s);
#define YH_ROWS(RPW)                                                                            \
  YH_CUDA(launch_pdl_if((phases & kPhaseDense) != 0, loss_rows_kernel<RPW>, dim3(row_grid),     \
                     dim3(kRowThreads), 0, s, tab, rg, p, rr,                                  \
                     (const long long*)m_batch, (const long long*)m_flat, m_target,             \
                     (const long long*)m_class, M_dev, capacity, iou_score_dev, pred_rows_dev,  \
                     (want_grad && !direct) ? ws->row_grad : nullptr, ws->partials))
    if (rpw == 2) YH_ROWS(2); else if (rpw == 4) YH_ROWS(4); else YH_ROWS(8);
#undef YH_ROWS
    YH_LAUNCHED(s);
  }
  if (phases & kPhaseFinish) {
    const bool fin_early = range != nullptr && range->finish_early;
    int fin_grid = 1;
    if (grad_heads && capacity > 0) {
      fin_grid = (capacity + kFinishThreads / 32 - 1) / (kFinishThreads / 32);
      // early: a warp per row at most (CTAs beyond the match count exit at once), so that every
      // row's loads are in flight before the wait; otherwise one resident wave, grid-striding
      static const bool early_uncapped = []() {
        const char* e = getenv("YH_FINISH_EARLY_UNCAPPED");
        return e && e[0] == '1';
      }();
      if (!(fin_early && early_uncapped) && fin_grid > ws->num_sms * 8) fin_grid = ws->num_sms * 8;
    }
    YH_TIME(YH_K_LOSS_FINISH, s);
    YH_CUDA(launch_pdl_if((phases & kPhaseRows) != 0 || fin_early, loss_finish_kernel, dim3(fin_grid),
                          dim3(kFinishThreads), 0, s, tab, g, p,
                       (const long long*)m_batch, (const long long*)m_flat, M_dev, capacity,
                       grad_heads ? (const float*)ws->row_grad : (const float*)nullptr,
                       (const double*)ws->partials, dense_grid, row_grid, num_ranges,
                       (ranged && range->row_ranges > 0) ? range->row_ranges : num_ranges, kRangeSlots,
                       (ranged && range->layer_hi > 0) ? range->layer_lo : 0,
                       (ranged && range->layer_hi > 0) ? range->layer_hi : YH_MAX_LAYERS,
                       (ranged && range->layer_hi > 0) ? range->finish_reduce : 1, fin_early ? 1 : 0,
                       (ranged && !fin_early) ? YH_MAX_LAYERS : keep_split_layer(tab, B),
                       losses4_dev, ws->err_flag));
    YH_LAUNCHED(s);
  }
  return YH_OK;
}

}  // namespace yh

extern "C" {

int yh_loss_fwd_bwd(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                    int B, int C, const int64_t* m_batch, const int64_t* m_flat,
                    const float* m_target, const int64_t* m_class, const int32_t* M_dev,
                    int capacity, const YhLossCfg* cfg, float* losses4_dev, float* iou_score_dev,
                    float* pred_rows_dev, float* const* grad_heads, void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  return yh::loss_phases(yh::kPhaseDense | yh::kPhaseRows | yh::kPhaseFinish, ws, heads, layers, L,
                         B, C, m_batch, m_flat, m_target, m_class, M_dev, capacity, cfg,
                         losses4_dev, iou_score_dev, pred_rows_dev, grad_heads, 0, stream);
}

int yh_loss_dense(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L, int B,
                  int C, const YhLossCfg* cfg, float* const* grad_heads, void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  return yh::loss_phases(yh::kPhaseDense, ws, heads, layers, L, B, C, nullptr, nullptr, nullptr,
                         nullptr, nullptr, 0, cfg, nullptr, nullptr, nullptr, grad_heads, 0,
                         stream);
}

int yh_loss_rows(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L, int B,
                 int C, const int64_t* m_batch, const int64_t* m_flat, const float* m_target,
                 const int64_t* m_class, const int32_t* M_dev, int capacity, const YhLossCfg* cfg,
                 float* losses4_dev, float* iou_score_dev, float* pred_rows_dev,
                 float* const* grad_heads, void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  return yh::loss_phases(yh::kPhaseRows | yh::kPhaseFinish, ws, heads, layers, L, B, C, m_batch,
                         m_flat, m_target, m_class, M_dev, capacity, cfg, losses4_dev,
                         iou_score_dev, pred_rows_dev, grad_heads, 0, stream);
}

int yh_loss_rows_compute(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                         int B, int C, const int64_t* m_batch, const int64_t* m_flat,
                         const float* m_target, const int64_t* m_class, const int32_t* M_dev,
                         int capacity, const YhLossCfg* cfg, float* iou_score_dev,
                         float* pred_rows_dev, int want_grad, void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  return yh::loss_phases(yh::kPhaseRows, ws, heads, layers, L, B, C, m_batch, m_flat, m_target,
                         m_class, M_dev, capacity, cfg, nullptr, iou_score_dev, pred_rows_dev,
                         nullptr, want_grad, stream);
}

int yh_loss_finish(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C,
                   const int64_t* m_batch, const int64_t* m_flat, const int32_t* M_dev,
                   int capacity, const YhLossCfg* cfg, float* losses4_dev,
                   float* const* grad_heads, void* stream) {
  return yh::loss_phases(yh::kPhaseFinish, ws, nullptr, layers, L, B, C, m_batch, m_flat, nullptr,
                         nullptr, M_dev, capacity, cfg, losses4_dev, nullptr, nullptr, grad_heads,
                         0, stream);
}

// ---- ranged forms: the batch in groups of images small enough for the L2 --------------------
int yh_loss_dense_range(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                        int B, int C, int b_lo, int b_hi, int range_index, int num_ranges,
                        const YhLossCfg* cfg, float* const* grad_heads, void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  yh::LossRange r = {b_lo, b_hi, range_index, num_ranges, nullptr, 0, 0, 0, 1, 0};
  YH_REQUIRE(num_ranges >= 1, "num_ranges=%d", num_ranges);
  return yh::loss_phases(yh::kPhaseDense, ws, heads, layers, L, B, C, nullptr, nullptr, nullptr,
                         nullptr, nullptr, 0, cfg, nullptr, nullptr, nullptr, grad_heads, 0, stream,
                         &r);
}

int yh_loss_rows_range(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                       int B, int C, int b_lo, int b_hi, int range_index, int num_ranges,
                       const int64_t* m_batch, const int64_t* m_flat, const float* m_target,
                       const int64_t* m_class, const int32_t* M_dev, int capacity,
                       const int32_t* image_row_offsets_dev, const YhLossCfg* cfg,
                       float* iou_score_dev, float* pred_rows_dev, float* const* grad_heads,
                       void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  YH_REQUIRE(num_ranges >= 1, "num_ranges=%d", num_ranges);
  yh::LossRange r = {b_lo, b_hi, range_index, num_ranges, image_row_offsets_dev, 1, 0, 0, 1, 0};
  return yh::loss_phases(yh::kPhaseRows, ws, heads, layers, L, B, C, m_batch, m_flat, m_target,
                         m_class, M_dev, capacity, cfg, nullptr, iou_score_dev, pred_rows_dev,
                         grad_heads, 0, stream, &r);
}

int yh_loss_finish_ranges(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C,
                          int num_ranges, const int32_t* M_dev, int capacity, const YhLossCfg* cfg,
                          float* losses4_dev, void* stream) {
  YH_REQUIRE(num_ranges >= 1, "num_ranges=%d", num_ranges);
  yh::LossRange r = {0, B, 0, num_ranges, nullptr, 1, 0, 0, 1, 0};
  return yh::loss_phases(yh::kPhaseFinish, ws, nullptr, layers, L, B, C, nullptr, nullptr, nullptr,
                         nullptr, M_dev, capacity, cfg, losses4_dev, nullptr, nullptr, nullptr, 0,
                         stream, &r);
}

// ---- layered forms: the dense pass cut by LAYER, the scatter of a layer group behind it ------
int yh_loss_dense_layers(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                         int B, int C, int layer_lo, int layer_hi, int part_index, int num_parts,
                         const YhLossCfg* cfg, float* const* grad_heads, void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  YH_REQUIRE(num_parts >= 1 && layer_hi > 0, "num_parts=%d layer_hi=%d", num_parts, layer_hi);
  yh::LossRange r = {0, B, part_index, num_parts, nullptr, 0, layer_lo, layer_hi, 0, 0};
  return yh::loss_phases(yh::kPhaseDense, ws, heads, layers, L, B, C, nullptr, nullptr, nullptr,
                         nullptr, nullptr, 0, cfg, nullptr, nullptr, nullptr, grad_heads, 0, stream,
                         &r);
}

int yh_loss_finish_layers(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C, int layer_lo,
                          int layer_hi, int reduce, int num_dense_parts, const int64_t* m_batch,
                          const int64_t* m_flat, const int32_t* M_dev, int capacity,
                          const YhLossCfg* cfg, float* losses4_dev, float* const* grad_heads,
                          void* stream) {
  YH_REQUIRE(num_dense_parts >= 1 && layer_hi > 0 && layer_lo >= 0 && layer_lo < layer_hi && layer_hi <= L,
             "num_dense_parts=%d layers [%d,%d)", num_dense_parts, layer_lo, layer_hi);
  yh::LossRange r = {0, B, 0, num_dense_parts, nullptr, 0, layer_lo, layer_hi, reduce ? 1 : 0, 1};
  return yh::loss_phases(yh::kPhaseFinish, ws, nullptr, layers, L, B, C, m_batch, m_flat, nullptr,
                         nullptr, M_dev, capacity, cfg, losses4_dev, nullptr, nullptr, grad_heads, 0,
                         stream, &r);
}

int yh_loss_finish_after_dense(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C,
                               const int64_t* m_batch, const int64_t* m_flat, const int32_t* M_dev,
                               int capacity, const YhLossCfg* cfg, float* losses4_dev,
                               float* const* grad_heads, void* stream) {
  // the plain whole-batch finish (count 0: not ranged, same partial-sum slots as yh_loss_dense and
  // yh_loss_rows_compute) with finish_early set
  yh::LossRange r = {0, B, 0, 0, nullptr, 0, 0, 0, 1, 1, 1};
  return yh::loss_phases(yh::kPhaseFinish, ws, nullptr, layers, L, B, C, m_batch, m_flat, nullptr,
                         nullptr, M_dev, capacity, cfg, losses4_dev, nullptr, nullptr, grad_heads, 0,
                         stream, &r);
}

}  // extern "C"
