// assign.cu — K2: training target assignment on the device
// (CyCxHWMatcher::match_bboxes, yolo-dl/src/loss/pred_target_matching.rs:51-276).
//
// The reference folds (GT x layer x anchor x neighbour-cell) entries into a HashMap keyed by
// (batch, layer, anchor, row, col); on a key collision the GT whose centre is closer to the
// cell centre wins, strict '<', so the earlier GT keeps ties (:180-217).  Collisions only
// happen between GTs of the SAME image, so one CTA owns one image: its GT boxes are staged in
// shared memory, its N slots live in an L2-resident slice of the workspace and the fold
// becomes two order-independent atomicMin passes:
//   pass A: slot_dist[flat] = min over entries of bits(dist^2)        (f64, >= 0 => monotone)
//   pass B: slot_gt[flat]   = min GT index among entries with dist^2 == slot_dist[flat]
// = "smallest distance, then earliest GT", exactly the fold's result for any entry order.
// All decisions use f64 and the reference's operation order; products and sums are written
// with __dmul_rn/__dadd_rn so no FMA can change a floor(), a '< 0.5' or a distance tie.
// A second kernel turns the occupied slots into the (batch, flat)-sorted match list.
#include <math.h>

#include "common.cuh"

namespace yh {

constexpr int kAssignThreads = 512;
constexpr int kMaxGtSmem = 1024;  // GT rows staged per pass

struct AssignTab {
  int L, N, rect4;
  double thresh;
  int H[YH_MAX_LAYERS], W[YH_MAX_LAYERS], A[YH_MAX_LAYERS], start[YH_MAX_LAYERS];
  double ah[YH_MAX_LAYERS][YH_MAX_ANCHORS], aw[YH_MAX_LAYERS][YH_MAX_ANCHORS];
};

// One (GT, layer, anchor) entry -> up to 5 cells (pred_target_matching.rs:87-150): centre, top,
// left, bottom, right.  Fills flat[k] / d2[k] (f64 bits of the squared distance between the GT
// centre and the cell centre) for the cells that exist and returns their bit mask; static slot
// indices, so the ten values stay in registers.
__device__ __forceinline__ unsigned cells_of(const AssignTab& t, int l, int a, double gcy, double gcx,
                                             double gh, double gw, int flat[5],
                                             unsigned long long d2[5]) {
  // anchor-ratio filter (:139-150)
  const double ah = t.ah[l][a], aw = t.aw[l][a];
  const bool bounded = (gh / ah <= t.thresh) && (ah / gh <= t.thresh) && (gw / aw <= t.thresh) &&
                       (aw / gw <= t.thresh);
  if (!bounded) return 0u;
  // Transform::from_sizes_exact(unit, feature): sy = H/1, ty = 0 - 0*sy; cy*sy + ty (:88-94,
  // bbox/src/transform.rs:16-37, bbox/src/cycxhw.rs:40-47)
  const double Hd = (double)t.H[l], Wd = (double)t.W[l];
  const double sy = Hd / 1.0, sx = Wd / 1.0;
  const double ty = __dsub_rn(0.0, __dmul_rn(0.0, sy)), tx = __dsub_rn(0.0, __dmul_rn(0.0, sx));
  const double cy = __dadd_rn(__dmul_rn(gcy, sy), ty);
  const double cx = __dadd_rn(__dmul_rn(gcx, sx), tx);
  const double fy = floor(cy), fx = floor(cx);
  const long long row = (long long)fy, col = (long long)fx;
  const double cy_fract = cy - trunc(cy), cx_fract = cx - trunc(cx);  // f64::fract
  const bool want[5] = {true, cy_fract < 0.5, cx_fract < 0.5, t.rect4 && cy_fract > 0.5,
                        t.rect4 && cx_fract > 0.5};
  const long long rr[5] = {row, row - 1, row, row + 1, row};
  const long long cc[5] = {col, col, col - 1, col, col + 1};
  unsigned mask = 0u;
#pragma unroll
  for (int i = 0; i < 5; ++i) {
    flat[i] = 0;
    d2[i] = 0ull;
    if (!want[i] || rr[i] < 0 || rr[i] >= t.H[l] || cc[i] < 0 || cc[i] >= t.W[l]) continue;  // (:112-116)
    // distance of the GT centre to the cell centre, ratio units (:193-204)
    const double pcy = __dadd_rn((double)rr[i], 0.5) / Hd;
    const double pcx = __dadd_rn((double)cc[i], 0.5) / Wd;
    const double dy = __dsub_rn(gcy, pcy), dx = __dsub_rn(gcx, pcx);
    flat[i] = t.start[l] + (int)cc[i] + t.W[l] * ((int)rr[i] + t.H[l] * a);
    YH_CHECK(flat[i] >= 0 && flat[i] < t.N);
    d2[i] = (unsigned long long)__double_as_longlong(__dadd_rn(__dmul_rn(dy, dy), __dmul_rn(dx, dx)));
    mask |= 1u << i;
  }
  return mask;
}

// Slot-table invariant: between two yh_assign calls every slot of the workspace is EMPTY
// (dist = ~0, gt = kEmptyGt; set once by yh_workspace_create).  A call touches only the slots
// its entries fall into (~10..15 per GT), records them in a per-image occupancy bitmap and the
// emit kernel puts them back to EMPTY after reading them — no O(N) initialisation or scan of
// the table per image, which is what made the first version of this path 58 us per step.
constexpr int kEmptyGt = 0x7f7f7f7f;  // cudaMemset(0x7f)
constexpr int kBitsSmem = 2048;       // occupancy words kept in shared memory (65536 instances)

__global__ void __launch_bounds__(kAssignThreads)
assign_slots_kernel(const AssignTab t, const double* __restrict__ gt,
                    const int* __restrict__ gt_off, unsigned long long* __restrict__ slot_dist,
                    int* __restrict__ slot_gt, unsigned* __restrict__ slot_bits, const int words,
                    int* __restrict__ img_count) {
  __shared__ double s_gt[kMaxGtSmem][4];
  __shared__ unsigned s_bits[kBitsSmem];
  __shared__ int s_cnt[kAssignThreads / 32];
  griddep_launch();
  griddep_wait();
  const int b = blockIdx.x;
  const int g0 = gt_off[b], g1 = gt_off[b + 1];
  const int ng = g1 - g0;
  unsigned long long* sd = slot_dist + (size_t)b * (size_t)t.N;
  int* sg = slot_gt + (size_t)b * (size_t)t.N;
  unsigned* gbits = slot_bits + (size_t)b * (size_t)words;
  const bool in_smem = words <= kBitsSmem;
  unsigned* bits = in_smem ? s_bits : gbits;  // generic pointer: atomicOr works on either

  for (int i = threadIdx.x; i < words; i += kAssignThreads) bits[i] = 0u;

  // total anchors over layers, and prefix for (layer, anchor) enumeration
  int la_total = 0;
  for (int l = 0; l < t.L; ++l) la_total += t.A[l];

  if (ng <= kMaxGtSmem && ng * la_total <= kAssignThreads) {
    // The usual case (<= 56 labels per image with 9 anchors): one (GT, layer, anchor) entry per
    // thread.  The cells are computed ONCE (the f64 divisions are the arithmetic of this kernel)
    // and pass B issues its <= 5 slot reads together instead of one dependent L2 round trip per
    // cell: 14.4 -> (r2) us per call at 8 images, where this kernel heads the critical path.
    __syncthreads();
    for (int i = threadIdx.x; i < ng * 4; i += kAssignThreads) {
      const int g = i >> 2, k = i & 3;
      s_gt[g][k] = gt[(size_t)(g0 + g) * 5 + k];
    }
    __syncthreads();
    int fl[5];
    unsigned long long dd[5];
    unsigned cm = 0u;
    const int wi = threadIdx.x;
    int gidx = 0;
    if (wi < ng * la_total) {
      const int g = wi / la_total;
      int r = wi - g * la_total;
      int l = 0;
      while (r >= t.A[l]) { r -= t.A[l]; ++l; }
      gidx = g;
      const double gh = s_gt[g][2], gw = s_gt[g][3];
      if (!(fabs(gh) <= 2.220446049250313e-16 || fabs(gw) <= 2.220446049250313e-16))
        cm = cells_of(t, l, r, s_gt[g][0], s_gt[g][1], gh, gw, fl, dd);
    }
#pragma unroll
    for (int i = 0; i < 5; ++i)
      if ((cm >> i) & 1u) {
        atomicMin(&sd[fl[i]], dd[i]);
        atomicOr(&bits[fl[i] >> 5], 1u << (fl[i] & 31));
      }
    __syncthreads();
    unsigned long long cur[5];
#pragma unroll
    for (int i = 0; i < 5; ++i) cur[i] = ((cm >> i) & 1u) ? __ldcg(&sd[fl[i]]) : 0ull;
#pragma unroll
    for (int i = 0; i < 5; ++i)
      if (((cm >> i) & 1u) && cur[i] == dd[i]) atomicMin(&sg[fl[i]], gidx);
    __syncthreads();
  } else {
    for (int pass = 0; pass < 2; ++pass) {
      for (int base = 0; base < ng; base += kMaxGtSmem) {
        const int chunk = min(kMaxGtSmem, ng - base);
        __syncthreads();
        for (int i = threadIdx.x; i < chunk * 4; i += kAssignThreads) {
          const int g = i >> 2, k = i & 3;
          s_gt[g][k] = gt[(size_t)(g0 + base + g) * 5 + k];
        }
        __syncthreads();
        const int work = chunk * la_total;
        for (int wi = threadIdx.x; wi < work; wi += kAssignThreads) {
          const int g = wi / la_total;
          int r = wi - g * la_total;
          int l = 0;
          while (r >= t.A[l]) { r -= t.A[l]; ++l; }
          const double gh = s_gt[g][2], gw = s_gt[g][3];
          // abs_diff_eq!(h, 0.0) with the default f64 epsilon (:65-68)
          if (fabs(gh) <= 2.220446049250313e-16 || fabs(gw) <= 2.220446049250313e-16) continue;
          const int gidx = base + g;
          int fl[5];
          unsigned long long dd[5];
          const unsigned cm = cells_of(t, l, r, s_gt[g][0], s_gt[g][1], gh, gw, fl, dd);
#pragma unroll
          for (int i = 0; i < 5; ++i) {
            if (!((cm >> i) & 1u)) continue;
            if (pass == 0) {
              atomicMin(&sd[fl[i]], dd[i]);
              atomicOr(&bits[fl[i] >> 5], 1u << (fl[i] & 31));
            } else if (__ldcg(&sd[fl[i]]) == dd[i]) {
              atomicMin(&sg[fl[i]], gidx);
            }
          }
        }
      }
      __syncthreads();
    }
  }

  // matches of this image = occupied slots; publish the bitmap for the emit kernel
  int local = 0;
  for (int i = threadIdx.x; i < words; i += kAssignThreads) {
    const unsigned w = bits[i];
    local += __popc(w);
    if (in_smem) gbits[i] = w;
  }
  for (int o = 16; o > 0; o >>= 1) local += __shfl_xor_sync(0xffffffffu, local, o);
  if ((threadIdx.x & 31) == 0) s_cnt[threadIdx.x >> 5] = local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int i = 0; i < kAssignThreads / 32; ++i) tot += s_cnt[i];
    img_count[b] = tot;
  }
}

// Ordered compaction of the occupied slots into the match list, sorted by (batch, flat);
// puts every slot it reads back to EMPTY.
// 1024 threads: one tile covers 32 768 instances, so every YOLOv4 head is ONE tile and the kernel
// is three dependent memory round trips (counts + bitmap -> slot -> label row) instead of one
// such chain per 8192 instances; under the dense store stream of the overlapped step a round
// trip costs 2-3 us and this kernel sits on the step's critical path (r2: 38 us in-graph).
constexpr int kEmitThreads = 1024;
constexpr int kEmitList = 4096;  // matches of one tile (1024 words = 32 768 instances) kept in smem

__global__ void __launch_bounds__(kEmitThreads)
assign_emit_kernel(int N, int B, const int words, const double* __restrict__ gt,
                   const int* __restrict__ gt_off, unsigned long long* __restrict__ slot_dist,
                   int* __restrict__ slot_gt, const unsigned* __restrict__ slot_bits,
                   const int* __restrict__ img_count, long long* __restrict__ out_batch,
                   long long* __restrict__ out_flat, int* __restrict__ out_gt,
                   float* __restrict__ out_target, long long* __restrict__ out_class,
                   int* __restrict__ out_M, int capacity, int* __restrict__ err_flag,
                   int* __restrict__ img_off) {
  __shared__ int s_warp[kEmitThreads / 32];
  __shared__ int s_base;
  __shared__ int s_list[kEmitList];
  __shared__ int s_pref[kEmitThreads];       // exclusive prefix of the tile's words (overflow path)
  __shared__ unsigned s_word[kEmitThreads];  // the tile's words (overflow path)
  griddep_launch();
  griddep_wait();
  const int b = blockIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const unsigned* bits = slot_bits + (size_t)b * (size_t)words;

  // everything that depends only on the slot kernel is requested together: the counts of the
  // earlier images, this image's label offset and its first tile of bitmap words
  int part = 0;
  for (int i = threadIdx.x; i < b; i += kEmitThreads) part += img_count[i];
  const unsigned w_first = ((int)threadIdx.x < words) ? bits[threadIdx.x] : 0u;
  const int g0 = gt_off[b];
  const int mine = img_count[b];
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (lane == 0) s_warp[wid] = part;
  __syncthreads();
  if (threadIdx.x == 0) {
    int tot = 0;
    for (int i = 0; i < kEmitThreads / 32; ++i) tot += s_warp[i];
    s_base = tot;
    img_off[b] = tot;  // first row of image b in the match list (yh_assign_image_offsets)
    if (b == B - 1) {
      const int M = tot + mine;
      img_off[B] = M;
      *out_M = M;
      if (M > capacity) atomicOr(err_flag, 1);
    }
  }
  __syncthreads();
  int base = s_base;
  if (mine == 0) return;  // block-uniform
  unsigned long long* sd = slot_dist + (size_t)b * (size_t)N;
  int* sg = slot_gt + (size_t)b * (size_t)N;

  for (int tile = 0; tile < words; tile += kEmitThreads) {
    const int wi = tile + threadIdx.x;
    unsigned w = tile == 0 ? w_first : ((wi < words) ? bits[wi] : 0u);
    const int c = __popc(w);
    // exclusive scan of c over the CTA
    int incl = c;
#pragma unroll
    for (int o = 1; o < 32; o <<= 1) {
      const int v = __shfl_up_sync(0xffffffffu, incl, o);
      if (lane >= o) incl += v;
    }
    __syncthreads();  // s_warp reuse
    if (lane == 31) s_warp[wid] = incl;
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int k = 0; k < kEmitThreads / 32; ++k) {
      const int v = s_warp[k];
      if (k < wid) before += v;
      total += v;
    }
    // publish the flats of this tile in order, then one thread per match (no serial walk over
    // the bits of a word: every match costs two dependent L2 round trips, slot -> GT row)
    int lp = before + incl - c;
    s_pref[threadIdx.x] = lp;
    s_word[threadIdx.x] = w;
    if (total <= kEmitList) {
      while (w) {
        const int bit = __ffs(w) - 1;
        w &= w - 1;
        YH_CHECK(lp < kEmitList);
        s_list[lp++] = wi * 32 + bit;
      }
    }
    __syncthreads();
    for (int j = threadIdx.x; j < total; j += kEmitThreads) {
      int flat;
      if (total <= kEmitList) {
        flat = s_list[j];
      } else {
        // more matches in the tile than the list holds (CSP-640 with 300 labels per image: ~20 000):
        // the word of match j by binary search over the words' prefix, then its (j - prefix)-th bit
        int lo = 0, hi = kEmitThreads - 1;  // last word whose exclusive prefix is <= j
        while (lo < hi) {
          const int mid = (lo + hi + 1) >> 1;
          if (s_pref[mid] <= j) lo = mid; else hi = mid - 1;
        }
        unsigned ww = s_word[lo];
        for (int r = j - s_pref[lo]; r > 0; --r) ww &= ww - 1;
        flat = (tile + lo) * 32 + __ffs(ww) - 1;
      }
      const int pos = base + j;
      YH_CHECK(flat >= 0 && flat < N);
      const int g = sg[flat];
      YH_CHECK(g >= 0 && g < gt_off[b + 1] - g0);  // an occupied slot names a label of this image
      sg[flat] = kEmptyGt;
      sd[flat] = ~0ull;
      if (pos < capacity) {
        const int grow = g0 + g;
        out_batch[pos] = b;
        out_flat[pos] = flat;
        out_gt[pos] = grow;
        const double* r = gt + (size_t)grow * 5;
        // rect.cast::<f32>() (:247)
        *reinterpret_cast<float4*>(out_target + (size_t)pos * 4) =
            make_float4((float)r[0], (float)r[1], (float)r[2], (float)r[3]);
        out_class[pos] = (long long)r[4];
      }
    }
    base += total;
  }
}

}  // namespace yh

extern "C" int yh_assign(YhWorkspace* ws, const double* gt_dev, const int32_t* gt_offsets_dev,
                         int B, const YhLayer* layers, int L, int rect4,
                         double anchor_scale_thresh, int64_t* out_batch, int64_t* out_flat,
                         int32_t* out_gt, float* out_target, int64_t* out_class,
                         int32_t* out_M_dev, int capacity, void* stream) {
  using namespace yh;
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  YH_REQUIRE(gt_offsets_dev != nullptr, "gt_offsets is NULL");
  YH_REQUIRE(out_batch && out_flat && out_gt && out_target && out_class && out_M_dev,
             "output pointer is NULL");
  YH_REQUIRE(B >= 1 && B <= ws->max_batch, "B=%d outside 1..max_batch=%d", B, ws->max_batch);
  // CyCxHWMatcherInit::build (:32-35)
  YH_REQUIRE(anchor_scale_thresh >= 1.0, "anchor_scale_thresh must be greater than or equal to 1");
  YH_REQUIRE(capacity >= 0, "capacity=%d", capacity);
  YH_REQUIRE(((uintptr_t)out_target & 15) == 0, "out_target not 16-byte aligned");
  LayerTab lt;
  int rc = build_layer_tab(layers, L, 1, nullptr, &lt);
  if (rc) return rc;
  YH_REQUIRE(lt.N <= ws->max_instances, "instances %d > workspace max %lld", lt.N,
             (long long)ws->max_instances);
  AssignTab t;
  t.L = L;
  t.N = lt.N;
  t.rect4 = rect4 ? 1 : 0;
  t.thresh = anchor_scale_thresh;
  for (int l = 0; l < L; ++l) {
    t.H[l] = lt.H[l];
    t.W[l] = lt.W[l];
    t.A[l] = lt.A[l];
    t.start[l] = lt.start[l];
    for (int a = 0; a < lt.A[l]; ++a) {
      t.ah[l][a] = layers[l].anchor_h[a];
      t.aw[l][a] = layers[l].anchor_w[a];
    }
  }
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(ws->device));
  note_batch(B);
  YH_TIME(YH_K_ASSIGN_SLOTS, s);
  const int words = (t.N + 31) / 32;
  assign_slots_kernel<<<B, kAssignThreads, 0, s>>>(t, gt_dev, gt_offsets_dev, ws->slot_dist,
                                                   ws->slot_gt, ws->slot_bits, words,
                                                   ws->img_count);
  YH_LAUNCHED(s);
  YH_TIME(YH_K_ASSIGN_EMIT, s);
  YH_CUDA(launch_pdl(assign_emit_kernel, dim3(B), dim3(kEmitThreads), 0, s, t.N, B, words, gt_dev,
                     gt_offsets_dev, ws->slot_dist, ws->slot_gt, ws->slot_bits, ws->img_count,
                     (long long*)out_batch, (long long*)out_flat, out_gt, out_target,
                     (long long*)out_class, out_M_dev, capacity, ws->err_flag, ws->img_off));
  YH_LAUNCHED(s);
  return YH_OK;
}

extern "C" const int32_t* yh_assign_image_offsets(const YhWorkspace* ws) {
  return ws ? ws->img_off : nullptr;
}
