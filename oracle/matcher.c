/* oracle/matcher.c — C restatement of CyCxHWMatcher::match_bboxes
 * (yolo-dl/src/loss/pred_target_matching.rs:51-276): same fold order (GT x layer x anchor x
 * neighbour cell), f64, a hash map keyed by the instance, strict '<' on the squared distance so
 * the first-seen GT keeps ties (:180-217).  Single thread like the reference.
 * TEST INFRASTRUCTURE ONLY: checker for tests/ and the timed CPU baseline of bench.py.
 * Compiled with -ffp-contract=off so that a*b+c is never fused (rustc does not fuse either). */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
  int64_t key; /* batch * N + flat, -1 = empty */
  int64_t gt;  /* index inside the image */
  int32_t layer, row, col, _pad;
} Slot;

static uint64_t mix(uint64_t x) {
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  x *= 0xc4ceb9fe1a85ec53ULL;
  x ^= x >> 33;
  return x;
}

long oracle_match_bboxes(const double* gt, const int32_t* offs, int B, int L, const int32_t* Hs,
                         const int32_t* Ws, const int32_t* As, const int64_t* starts,
                         const double* anchors /* [L][amax][2] h,w */, int amax, int rect4,
                         double thresh, int64_t* out_b, int64_t* out_flat, int64_t* out_gt,
                         long cap) {
  const double snap = 0.5;
  int64_t N = 0;
  for (int l = 0; l < L; ++l) N += (int64_t)As[l] * Hs[l] * Ws[l];
  size_t tab = 64;
  while ((long)tab < 2 * cap + 16) tab <<= 1;
  Slot* slots = (Slot*)malloc(tab * sizeof(Slot));
  if (!slots) return -1;
  for (size_t i = 0; i < tab; ++i) slots[i].key = -1;
  long used = 0;

  for (int b = 0; b < B; ++b) {
    for (int g = offs[b]; g < offs[b + 1]; ++g) {
      const double gcy = gt[5 * g], gcx = gt[5 * g + 1], gh = gt[5 * g + 2], gw = gt[5 * g + 3];
      if (fabs(gh - 0.0) <= 2.220446049250313e-16 || fabs(gw - 0.0) <= 2.220446049250313e-16)
        continue; /* (:62-73) */
      for (int l = 0; l < L; ++l) {
        const int H = Hs[l], W = Ws[l];
        /* Transform::from_sizes_exact (bbox/src/transform.rs:16-37) */
        const double sy = ((0.0 + (double)H) - 0.0) / ((0.0 + 1.0) - 0.0);
        const double sx = ((0.0 + (double)W) - 0.0) / ((0.0 + 1.0) - 0.0);
        const double ty = 0.0 - 0.0 * sy, tx = 0.0 - 0.0 * sx;
        const double cy = gcy * sy + ty, cx = gcx * sx + tx;
        const int64_t row = (int64_t)floor(cy), col = (int64_t)floor(cx);
        const double cyf = cy - trunc(cy), cxf = cx - trunc(cx);
        int64_t rr[5], cc[5];
        int n = 0;
        rr[n] = row; cc[n] = col; ++n;
        if (cyf < snap) { rr[n] = row - 1; cc[n] = col; ++n; }
        if (cxf < snap) { rr[n] = row; cc[n] = col - 1; ++n; }
        if (rect4 && cyf > 1.0 - snap) { rr[n] = row + 1; cc[n] = col; ++n; }
        if (rect4 && cxf > 1.0 - snap) { rr[n] = row; cc[n] = col + 1; ++n; }
        for (int a = 0; a < As[l]; ++a) {
          const double ah = anchors[((size_t)l * amax + a) * 2], aw = anchors[((size_t)l * amax + a) * 2 + 1];
          if (!(gh / ah <= thresh && ah / gh <= thresh && gw / aw <= thresh && aw / gw <= thresh))
            continue; /* (:139-150) */
          for (int i = 0; i < n; ++i) {
            if (rr[i] < 0 || rr[i] >= H || cc[i] < 0 || cc[i] >= W) continue; /* (:112-116) */
            const int64_t flat = starts[l] + cc[i] + (int64_t)W * (rr[i] + (int64_t)H * a);
            const int64_t key = (int64_t)b * N + flat;
            size_t h = (size_t)mix((uint64_t)key) & (tab - 1);
            while (slots[h].key != -1 && slots[h].key != key) h = (h + 1) & (tab - 1);
            if (slots[h].key == -1) { /* Vacant (:212-214) */
              if (used >= cap) { free(slots); return -2; }
              slots[h].key = key;
              slots[h].gt = g - offs[b];
              slots[h].layer = l;
              slots[h].row = (int32_t)rr[i];
              slots[h].col = (int32_t)cc[i];
              ++used;
            } else { /* Occupied (:186-211) */
              const double* o = gt + 5 * (size_t)(offs[b] + slots[h].gt);
              const double pcy = ((double)rr[i] + 0.5) / (double)H;
              const double pcx = ((double)cc[i] + 0.5) / (double)W;
              const double d_orig = (o[0] - pcy) * (o[0] - pcy) + (o[1] - pcx) * (o[1] - pcx);
              const double d_new = (gcy - pcy) * (gcy - pcy) + (gcx - pcx) * (gcx - pcx);
              if (d_new < d_orig) slots[h].gt = g - offs[b];
            }
          }
        }
      }
    }
  }
  long m = 0;
  for (size_t i = 0; i < tab; ++i) {
    if (slots[i].key == -1) continue;
    out_b[m] = slots[i].key / N;
    out_flat[m] = slots[i].key % N;
    out_gt[m] = slots[i].gt;
    ++m;
  }
  free(slots);
  return m;
}
