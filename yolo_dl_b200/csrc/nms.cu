// nms.cu — K5: candidate filter + grouped greedy NMS + post-NMS class pick.
//   NonMaxSuppression::forward   yolo-dl/src/loss/nms.rs:72-170
//   tch_nms::nms_by_scores       external (call site nms.rs:142-148); semantics restated in
//                                oracle/nms.py: greedy, score-descending, equal scores keep the
//                                lower index, suppress iff IoU > thr, IoU = inter/(a+b-inter)
//   YoloInference::forward       yolo-dl/src/loss/inference.rs:50-77 (class pick)
//
// Pipeline of yh_nms (3 launches; 4 + a memset when the candidate capacity admits large segments):
//   nms_filter_kernel   one warp per 32 instances.  Reads the objectness plane (4 B/instance);
//                       only where sigmoid(obj) >= thr — conf = obj*cls <= obj, so nothing else
//                       can pass (nms.rs:97-99) — the warp reads that instance's C class logits
//                       together (lane = class), ballots conf >= thr and appends 64-bit keys
//                       (~ordered(score) << 32 | class*N + instance) to the image's candidate
//                       list with ONE atomicAdd per ballot.  The [B,C,N] confidence tensor of
//                       the reference is never materialised.
//   nms_segment_kernel  one CTA per segment of <= YH_BIG_MIN candidates: sorts the keys (rank sort
//                       up to 1024 keys; bitonic in shared memory up to 8192, hybrid above), so
//                       the list is in score-descending, then nonzero()-index-ascending order,
//                       and runs the greedy pass in tiles of 512 candidates: test against the
//                       kept list (shared-memory cache), 512 x 512 suppression bit matrix, one
//                       warp walks the tile in order.  No K x K mask exists anywhere.
//                       Per-class suppression only adds "same class" to the predicate, so one
//                       sorted list per image serves both modes (nms.rs:129-133).
//   nms_emit_kernel     ordered compaction of the per-image kept lists into the outputs — behind
//                       the large-segment kernel only: otherwise nms_segment_kernel emits by
//                       itself (SegEmit: kept counts published per image, look-back for the base).
//   nms_big_kernel      segments (an image; a group bucket of yh_nms_by_scores) with more than
//                       kBigMin candidates: the same greedy pass spread over ALL SMs — merge sort,
//                       box gather, and per 512-candidate tile the test against the kept list on
//                       every SM but one, which resolves the previous tile meanwhile.  Per-image
//                       groups (`suppress_by_class: false`) are the only mode the reference's
//                       binaries use (train/src/train/multi_gpu.rs:72, single_gpu.rs:59,
//                       detect/src/main.rs:84), so one big group must not sit on one SM.
// yh_nms_by_scores exposes the same engine on explicit [K,4] ltrb boxes / scores / int64 groups,
// i.e. the drop-in for tch_nms::nms_by_scores itself.
#include <math.h>
#include <stdlib.h>
#include <string.h>

#ifdef YH_SEG_PROF
#include <cstdio>
#endif
#include "common.cuh"

namespace yh {

constexpr int kFilterThreads = 256;
constexpr int kNmsThreads = 1024;
constexpr int kSortTile = 8192;    // keys sorted in shared memory at once (64 KB)
constexpr int kKeptCache = 2048;   // kept boxes cached in shared memory (aliases the sort tile)

// ---------------------------------------------------------------------------------
// keys
// ---------------------------------------------------------------------------------
// order-preserving float -> uint (ascending), inverted so that ascending keys = descending score
__device__ __forceinline__ unsigned score_to_bits(float s) {
  unsigned u = __float_as_uint(s);
  u = (u & 0x80000000u) ? ~u : (u | 0x80000000u);
  return ~u;
}
__device__ __forceinline__ float bits_to_score(unsigned k) {
  unsigned u = ~k;
  u = (u & 0x80000000u) ? (u & 0x7fffffffu) : ~u;
  return __uint_as_float(u);
}
__device__ __forceinline__ unsigned long long make_key(float score, unsigned idx) {
  return ((unsigned long long)score_to_bits(score) << 32) | (unsigned long long)idx;
}

// ---------------------------------------------------------------------------------
// box providers: key index -> (t,l,b,r) box and suppression group
// ---------------------------------------------------------------------------------
struct HeadBoxes {
  LayerTab t;
  int by_class;
  // idx = class*N + instance (nonzero() order inside an image, nms.rs:100-103)
  __device__ __forceinline__ void get(int seg, unsigned idx, float4& box, long long& grp) const {
    const int cls = (int)(idx / (unsigned)t.N);
    const int n = (int)(idx - (unsigned)cls * (unsigned)t.N);
    YH_CHECK(cls < t.C);
    const Cell c = locate(t, n);
    const int l = c.layer;
    const size_t es = (size_t)t.A[l] * (size_t)t.HW[l];
    const float* hp = t.head[l] + head_offset(t, l, seg, c.anchor, c.cell);
    const float cy = decode_center(sigmoid_f(__ldg(hp)), c.row, (float)t.H[l]);
    const float cx = decode_center(sigmoid_f(__ldg(hp + es)), c.col, (float)t.W[l]);
    const float h = decode_size(sigmoid_f(__ldg(hp + 2 * es)), t.ah[l][c.anchor]);
    const float w = decode_size(sigmoid_f(__ldg(hp + 3 * es)), t.aw[l][c.anchor]);
    // nms.rs:91-94
    box.x = cy - h / 2.0f;  // t
    box.y = cx - w / 2.0f;  // l
    box.z = cy + h / 2.0f;  // b
    box.w = cx + w / 2.0f;  // r
    grp = by_class ? (long long)cls : 0ll;
  }
};

struct ArrayBoxes {
  const float* ltrb;        // [K,4] l,t,r,b (nms.rs:128)
  const long long* groups;  // [K]
  __device__ __forceinline__ void get(int, unsigned idx, float4& box, long long& grp) const {
    const float4 v = __ldg(reinterpret_cast<const float4*>(ltrb) + idx);
    box.x = v.y;
    box.y = v.x;
    box.z = v.w;
    box.w = v.z;
    grp = __ldg(groups + idx);
  }
};

// IoU > thr, arithmetic of the greedy reference (areas first, inter/(a+b-inter), f32; the
// comparison promotes to the f64 threshold the caller passes, nms.rs:146)
__device__ __forceinline__ float box_area(const float4 a) { return (a.w - a.y) * (a.z - a.x); }

__device__ __forceinline__ bool suppresses(const float4 a, const float area_a, const float4 c,
                                           const float area_c, const double thr) {
  const float xx1 = fmaxf(a.y, c.y), yy1 = fmaxf(a.x, c.x);
  const float xx2 = fminf(a.w, c.w), yy2 = fminf(a.z, c.z);
  const float w = fmaxf(0.0f, xx2 - xx1), h = fmaxf(0.0f, yy2 - yy1);
  // disjoint boxes (most pairs): inter = 0, so the quotient is 0 or 0/0 = NaN and neither is
  // > thr for a non-negative threshold (a negative one takes the full path) — same answer, no division
  // (the IEEE division is a third of this function; the kept-list test of the large-segment
  // kernel is bound by exactly this arithmetic on all SMs)
  if (!(w > 0.0f && h > 0.0f) && thr >= 0.0) return false;
  const float inter = w * h;
  const float ovr = inter / (area_a + area_c - inter);
  return (double)ovr > thr;
}

// ---------------------------------------------------------------------------------
// candidate filter (heads -> per-image key lists)
// ---------------------------------------------------------------------------------
// Every thread owns kFilterPer instances (strided by the CTA width) and issues all of their
// objectness loads before the first use; a passing instance is expanded by its whole warp with
// all ceil(C/32) class loads in flight together.  (The first version — one instance per thread,
// class loads one at a time — ran as five waves of warps each waiting on four dependent DRAM
// round trips: 33 us for 5.8 MB of compulsory reads.)
constexpr int kFilterClsRegs = 3;  // class loads in flight per expansion (C <= 96), tail loops

template <int kFilterPer>
__global__ void __launch_bounds__(kFilterThreads)
nms_filter_kernel(const LayerTab t, const float thr, unsigned long long* __restrict__ keys,
                  const int stride, const int cap, int* __restrict__ count,
                  int* __restrict__ err_flag, int* __restrict__ seg_ready) {
  griddep_launch();  // nms_segment_kernel's CTAs may take their places while the filter runs
  const int b = blockIdx.y;
  // the fused emit of nms_segment_kernel (SegEmit) starts from zeroed "kept count published" words
  if (blockIdx.x == 0 && threadIdx.x == 0) seg_ready[b] = 0;
  const int lane = threadIdx.x & 31;
  const float* hp[kFilterPer];
  unsigned es[kFilterPer];
  float xo[kFilterPer];
  int nn[kFilterPer];
#pragma unroll
  for (int j = 0; j < kFilterPer; ++j) {
    const int n = (blockIdx.x * kFilterPer + j) * kFilterThreads + threadIdx.x;
    nn[j] = n;
    hp[j] = nullptr;
    es[j] = 0;
    xo[j] = -INFINITY;
    if (n < t.N) {
      const Cell c = locate(t, n);
      es[j] = (unsigned)(t.A[c.layer] * t.HW[c.layer]);
      hp[j] = t.head[c.layer] + head_offset(t, c.layer, b, c.anchor, c.cell);
      xo[j] = __ldg(hp[j] + 4 * (size_t)es[j]);
    }
  }
  unsigned long long* out = keys + (size_t)b * (size_t)stride;
#pragma unroll
  for (int j = 0; j < kFilterPer; ++j) {
    const float so = sigmoid_f(xo[j]);  // sigmoid(-inf) = 0 never passes a positive threshold
    unsigned m = __ballot_sync(0xffffffffu, nn[j] < t.N && so >= thr);
    while (m) {
      const int src = __ffs(m) - 1;
      m &= m - 1;
      const float* hp_s = (const float*)__shfl_sync(0xffffffffu, (unsigned long long)hp[j], src);
      const unsigned es_s = __shfl_sync(0xffffffffu, es[j], src);
      const float so_s = __shfl_sync(0xffffffffu, so, src);
      const int n_s = __shfl_sync(0xffffffffu, nn[j], src);
      float xc[kFilterClsRegs];
#pragma unroll
      for (int r = 0; r < kFilterClsRegs; ++r) {
        const int k = r * 32 + lane;
        xc[r] = (k < t.C) ? ld_scatter(hp_s + (size_t)(5 + k) * es_s, t.l2hint) : 0.0f;
      }
      for (int k0 = 0, r = 0; k0 < t.C; k0 += 32, ++r) {
        const int k = k0 + lane;
        bool hit = false;
        float conf = 0.0f;
        if (k < t.C) {
          float x;
          if (r == 0) x = xc[0];
          else if (r == 1) x = xc[1];
          else if (r == 2) x = xc[2];
          else x = __ldg(hp_s + (size_t)(5 + k) * es_s);
          conf = so_s * sigmoid_f(x);
          hit = conf >= thr;
        }
        const unsigned hm = __ballot_sync(0xffffffffu, hit);
        if (hm) {
          int base = 0;
          if (lane == 0) base = atomicAdd(&count[b], __popc(hm));
          base = __shfl_sync(0xffffffffu, base, 0);
          if (hit) {
            const int pos = base + __popc(hm & ((1u << lane) - 1u));
            if (pos < cap)
              out[pos] = make_key(conf, (unsigned)k * (unsigned)t.N + (unsigned)n_s);
            else
              atomicOr(err_flag, 2);
          }
        }
      }
    }
  }
}

// explicit scores -> keys of one segment
__global__ void nms_pack_kernel(const float* __restrict__ scores, const int K,
                                unsigned long long* __restrict__ keys, int* __restrict__ count) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < K) keys[i] = make_key(scores[i], (unsigned)i);
  if (i == 0) count[0] = K;
}

// ---------------------------------------------------------------------------------
// bitonic sort, one CTA
// ---------------------------------------------------------------------------------
// compare-exchange passes j = j_hi .. 1 of merge size k over n (pow2) keys in shared memory;
// gbase = global index of s[0] (the direction of a pair depends on the global index)
__device__ __forceinline__ void smem_passes(unsigned long long* s, const int n, const int gbase,
                                            const int k, const int j_hi) {
  for (int j = j_hi; j >= 1; j >>= 1) {
    for (int p = threadIdx.x; p < n / 2; p += kNmsThreads) {
      const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
      const int l = i | j;
      const bool asc = ((gbase + i) & k) == 0;
      const unsigned long long a = s[i], c = s[l];
      if ((a > c) == asc) {
        s[i] = c;
        s[l] = a;
      }
    }
    __syncthreads();
  }
}

// sorts keys[0..K) ascending in place; keys has room for next_pow2(K) entries
__device__ void sort_keys(unsigned long long* __restrict__ keys, const int K,
                          unsigned long long* s) {
  if (K <= 1) return;
#ifndef YH_SEG_NO_RANKSORT
  if (K <= kNmsThreads) {
    // rank sort: keys are unique (the index is part of the key), so the sorted position of a key
    // is the number of smaller keys — K broadcast reads of shared memory per thread and two
    // barriers, where the bitonic network below needs log2(P)*(log2(P)+1)/2 block-wide passes
    // (36 for the ~200 candidates per image of a training step: ~5 us of barriers on one SM)
    unsigned long long my = 0;
    if ((int)threadIdx.x < K) {
      my = keys[threadIdx.x];
      s[threadIdx.x] = my;
    }
    __syncthreads();
    if ((int)threadIdx.x < K) {
      int rank = 0;
#pragma unroll 8
      for (int i = 0; i < K; ++i) rank += s[i] < my;
      keys[rank] = my;
    }
    __syncthreads();
    return;
  }
#endif
  int P = 2;
  while (P < K) P <<= 1;
  if (P <= kSortTile) {
    for (int i = threadIdx.x; i < P; i += kNmsThreads) s[i] = (i < K) ? keys[i] : ~0ull;
    __syncthreads();
    for (int k = 2; k <= P; k <<= 1) smem_passes(s, P, 0, k, k >> 1);
    for (int i = threadIdx.x; i < K; i += kNmsThreads) keys[i] = s[i];
    __syncthreads();
    return;
  }
  // hybrid: tiles of kSortTile in shared memory, wider strides through global memory (L2)
  for (int i = K + threadIdx.x; i < P; i += kNmsThreads) keys[i] = ~0ull;
  __syncthreads();
  for (int base = 0; base < P; base += kSortTile) {
    for (int i = threadIdx.x; i < kSortTile; i += kNmsThreads) s[i] = keys[base + i];
    __syncthreads();
    for (int k = 2; k <= kSortTile; k <<= 1) smem_passes(s, kSortTile, base, k, k >> 1);
    for (int i = threadIdx.x; i < kSortTile; i += kNmsThreads) keys[base + i] = s[i];
    __syncthreads();
  }
  for (int k = kSortTile * 2; k <= P; k <<= 1) {
    for (int j = k >> 1; j >= kSortTile; j >>= 1) {
      for (int p = threadIdx.x; p < P / 2; p += kNmsThreads) {
        const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
        const int l = i | j;
        const bool asc = (i & k) == 0;
        const unsigned long long a = keys[i], c = keys[l];
        if ((a > c) == asc) {
          keys[i] = c;
          keys[l] = a;
        }
      }
      __syncthreads();
    }
    for (int base = 0; base < P; base += kSortTile) {
      for (int i = threadIdx.x; i < kSortTile; i += kNmsThreads) s[i] = keys[base + i];
      __syncthreads();
      smem_passes(s, kSortTile, base, k, kSortTile >> 1);
      for (int i = threadIdx.x; i < kSortTile; i += kNmsThreads) keys[base + i] = s[i];
      __syncthreads();
    }
  }
}

// ---------------------------------------------------------------------------------
// sort + greedy NMS, one CTA per segment
// ---------------------------------------------------------------------------------
struct SegView {
  unsigned long long* keys;      // [S * stride]
  unsigned long long* kept_key;  // [S * stride]
  float4* kept_box;              // [S * stride]
  long long* kept_grp;           // [S * stride]
  int* count;                    // [S] candidates (may exceed cap: clamped)
  int* kept_count;               // [S]
  long long stride;
  int cap;
  const int* seg_off;            // optional: segment s starts at seg_off[s] instead of s*stride
  int* need;                     // atomicMax of the candidate counts that exceeded cap
  int big_min;                   // segments with more candidates belong to nms_big_kernel (0: none do)
};

// Fused emit (yh_nms without large segments): the segment kernel itself compacts the kept lists
// into the outputs — the CTA of image s learns its first output row from the kept counts its
// predecessors publish (ready[i] = kept_i + 1; CTAs are handed out in index order, so the ones a
// CTA waits for are always resident or done) — instead of a second kernel behind it: one launch
// and one dependent-kernel gap (~8 us of the NMS chain) less.
struct SegEmit {
  int on;             // 0: only the kept lists are produced (yh_nms_by_scores; large segments)
  int N;              // instances per image (index = class * N + instance)
  long long* kb;
  long long* kc;
  long long* ki;
  float* tlbr;
  float* conf;
  int* out_K;
  int capacity;
  int* err_flag;
  int* ready;         // [S]: kept + 1 of every finished image; zeroed by the filter kernel of the call
};

// Greedy pass, in tiles of kTile candidates (score order):
//   a. every candidate of the tile is tested against the boxes kept by earlier tiles (two
//      threads per candidate, each walking half of the kept list: shared-memory cache first);
//   b. the survivors' pairwise suppression bits are computed for the whole tile at once — a warp
//      per (row, 32-column word), one ballot each — into a kTile x kTile bit matrix in shared
//      memory (upper triangle, rows of survivors only);
//   c. ONE warp walks the tile in order with the removed-set spread over its lanes (lane w =
//      word w): per KEPT candidate one shared-memory row read and an OR, nothing per suppressed
//      candidate — the only serial part of greedy NMS, ~50 cycles per kept box;
//   d. the kept candidates are appended in order (global lists + the shared-memory cache).
// (The first version resolved 32 candidates at a time with a shuffle/ballot loop per candidate
// and re-tested the later warps after every round: 43 us for ~80 candidates per image.)
constexpr int kTile = 512;
constexpr int kTileWords = kTile / 32;
constexpr size_t kSegSmem = (size_t)kSortTile * sizeof(unsigned long long) +
                            (size_t)kTile * kTileWords * sizeof(unsigned);

// In-order walk of one tile by ONE warp.  rem: lane w holds word w of the removed set (bit set =
// not a survivor); mask: the tile's suppression bits in shared memory, row r at
// mask[r*kTileWords + word], upper triangle.  Per 32-candidate word the 32 in-word rows are read
// into registers (broadcast loads), the word is resolved by a 32-step chain of pure ALU
// operations, and the rows of what it kept are OR-ed into the later words with independent loads
// (lane = word): ~300 cycles per word whatever is kept, where the first version paid a dependent
// shared-memory load plus a shuffle (~120 cycles) per KEPT box — 30 us for a tile of survivors.
// Returns the keep bits, lane w holding word w.
__device__ __forceinline__ unsigned walk_tile(const unsigned* mask, unsigned rem, const int nwords,
                                              const int lane) {
  unsigned keep = 0u;
  for (int w = 0; w < nwords; ++w) {
    unsigned cur = __shfl_sync(0xffffffffu, rem, w);
    unsigned m[32];
#pragma unroll
    for (int j = 0; j < 32; ++j) m[j] = mask[(w * 32 + j) * kTileWords + w];
    unsigned kw = 0u;
#pragma unroll
    for (int j = 0; j < 32; ++j) {
      const bool k = !((cur >> j) & 1u);
      kw |= k ? (1u << j) : 0u;
      cur |= k ? m[j] : 0u;  // rows of suppressed candidates may hold anything: never used
    }
    unsigned add = 0u, t = kw;
    const bool later = lane > w && lane < nwords;
    while (t) {
      const int j = __ffs(t) - 1;
      t &= t - 1;
      if (later) add |= mask[(w * 32 + j) * kTileWords + lane];
    }
    rem |= add;
    if (lane == w) keep = kw;
  }
  return keep;
}

template <class P>
__global__ void __launch_bounds__(kNmsThreads)
nms_segment_kernel(const P prov, const SegView v, const double iou_thr, const SegEmit em) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  __shared__ float4 t_box[kTile];
  __shared__ long long t_grp[kTile];
  __shared__ unsigned s_al[2][kTileWords];
  __shared__ unsigned s_keep[32];
  __shared__ unsigned s_nz[kTileWords];  // bit r: row r of the bit matrix has a bit set
  static_assert(kNmsThreads == 2 * kTile, "two threads per candidate of a tile");

#ifdef YH_SEG_PROF  // `make prof`: %globaltimer stamps at the phase boundaries, printed by segment 0
  unsigned long long T[10];
  int ti_ = 0;
#define YH_STAMP() do { if (threadIdx.x == 0 && ti_ < 10) asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(T[ti_++])); } while (0)
#else
#define YH_STAMP() do { } while (0)
#endif
  YH_STAMP();
  griddep_launch();
  griddep_wait();
  YH_STAMP();
  const int seg = blockIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int found = v.count[seg];  // reset to 0 by the emit / rank kernel that follows
  const int K = min(found, v.cap);
  // the filter keeps counting past the capacity: tell the host how much room the caller needs
  if (threadIdx.x == 0 && found > v.cap) atomicMax(v.need, found);
  if (v.big_min > 0 && K > v.big_min) return;  // nms_big_kernel's (all threads: block-uniform)
  const size_t seg_base = v.seg_off ? (size_t)v.seg_off[seg] : (size_t)seg * (size_t)v.stride;
  unsigned long long* keys = v.keys + seg_base;
  unsigned long long* kept_key = v.kept_key + seg_base;
  float4* kept_box = v.kept_box + seg_base;
  long long* kept_grp = v.kept_grp + seg_base;

  YH_STAMP();
  sort_keys(keys, K, s_keys);
  __syncthreads();
  YH_STAMP();

  // the sort tile becomes the kept-box cache; the bit matrix lives behind it
  float4* c_box = reinterpret_cast<float4*>(s_keys);
  long long* c_grp = reinterpret_cast<long long*>(c_box + kKeptCache);
  unsigned* mask = reinterpret_cast<unsigned*>(s_keys + kSortTile);
  int nkept = 0;

  const int ti = threadIdx.x & (kTile - 1);
  const int half = threadIdx.x >> 9;  // kTile == 512
  for (int base = 0; base < K; base += kTile) {
    const int nt = min(K - base, kTile);
    const int nwords = (nt + 31) >> 5;
    // ---- a. load, test against earlier tiles' kept boxes ------------------------------------
    const int i = base + ti;
    bool alive = ti < nt;
    unsigned long long key = 0;
    float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
    long long grp = 0;
    if (alive) {
      key = keys[i];
      prov.get(seg, (unsigned)(key & 0xffffffffull), box, grp);
      const float area = box_area(box);
      const int nc = min(nkept, kKeptCache);
      for (int q = half; q < nc; q += 2) {
        const float4 kb = c_box[q];
        if (c_grp[q] == grp && suppresses(kb, box_area(kb), box, area, iou_thr)) {
          alive = false;
          break;
        }
      }
      if (alive) {
        for (int q = kKeptCache + half; q < nkept; q += 2) {
          const float4 kb = kept_box[q];
          if (kept_grp[q] == grp && suppresses(kb, box_area(kb), box, area, iou_thr)) {
            alive = false;
            break;
          }
        }
      }
    }
    if (half == 0) {
      t_box[ti] = box;
      t_grp[ti] = grp;
    }
    if (threadIdx.x < kTileWords) s_nz[threadIdx.x] = 0u;
    {
      const unsigned am = __ballot_sync(0xffffffffu, alive);
      if (lane == 0) s_al[half][wid & (kTileWords - 1)] = am;
    }
    __syncthreads();
    YH_STAMP();

    // ---- b. suppression bits among the survivors (upper triangle) ---------------------------
    for (int item = wid; item < nt * nwords; item += kNmsThreads / 32) {
      const int r = item / nwords, w = item - r * nwords;
      if (w < (r >> 5)) continue;
      const unsigned al_r = s_al[0][r >> 5] & s_al[1][r >> 5];
      if (!((al_r >> (r & 31)) & 1u)) continue;
      const int j = w * 32 + lane;
      const unsigned al_w = s_al[0][w] & s_al[1][w];
      bool hit = false;
      if (j > r && ((al_w >> lane) & 1u)) {
        const float4 br = t_box[r], bj = t_box[j];
        hit = t_grp[r] == t_grp[j] && suppresses(br, box_area(br), bj, box_area(bj), iou_thr);
      }
      const unsigned m = __ballot_sync(0xffffffffu, hit);
      if (lane == 0) {
        mask[r * kTileWords + w] = m;
        if (m) atomicOr(&s_nz[r >> 5], 1u << (r & 31));
      }
    }
    __syncthreads();
    YH_STAMP();

    // ---- c. in-order walk, one warp ----------------------------------------------------------
    // (the register walk of nms_big_kernel, walk_tile, is NOT used here: inside the overlapped
    // training step it made this kernel 42 us instead of 33 and the step 193 us instead of 177-180
    // — same-box A/B of the two libraries, profiles/r2_ab_libs.txt — although alone both take 22.7 us)
    if (wid == 0) {
      unsigned rem = (lane < nwords) ? ~(s_al[0][lane] & s_al[1][lane]) : ~0u;
      unsigned keep = 0u;
      // Only a survivor that suppresses somebody (a row with a bit set: s_nz) needs a step of the
      // serial chain — a dependent shared-memory load and a shuffle, ~90 ns; the survivors with
      // empty rows are kept or not by what the others did and cost nothing.  Per-class NMS of a
      // training step (~50-200 candidates per image over 80 classes: hardly any pair overlaps):
      // phase c 4.9 -> 1.2 us of the kernel's 12 -> 8.5 (make prof, profiles/r2g_segprof.txt)
      for (int w = 0; w < nwords; ++w) {
        unsigned cur = __shfl_sync(0xffffffffu, rem, w);
        const unsigned nz = s_nz[w];
        unsigned t = ~cur & nz;  // alive suppressors of this word, visited in order
        while (t) {
          const int j = __ffs(t) - 1;
          const int r = w * 32 + j;
          const unsigned mrow = (lane >= w && lane < nwords) ? mask[r * kTileWords + lane] : 0u;
          rem |= mrow;
          cur |= __shfl_sync(0xffffffffu, mrow, w);
          t &= ~cur & ~((2u << j) - 1u);  // suppressed by j, or at / before j: done
        }
        if (lane == w) keep = ~cur;
      }
      s_keep[lane] = keep;
    }
    __syncthreads();
    YH_STAMP();

    // ---- d. append the kept candidates in order ----------------------------------------------
    int before = 0, total = 0;
#pragma unroll
    for (int w = 0; w < kTileWords; ++w) {
      const int c = __popc(s_keep[w]);
      if (w < (ti >> 5)) before += c;
      total += c;
    }
    const unsigned kwm = s_keep[ti >> 5];
    if (half == 0 && ((kwm >> (ti & 31)) & 1u)) {
      const int pos = nkept + before + __popc(kwm & ((1u << (ti & 31)) - 1u));
      YH_CHECK(pos >= 0 && pos < K);
      kept_key[pos] = key;
      kept_box[pos] = box;
      kept_grp[pos] = grp;
      if (pos < kKeptCache) {
        c_box[pos] = box;
        c_grp[pos] = grp;
      }
    }
    nkept += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) v.kept_count[seg] = nkept;
  if (em.on) {  // kernel-uniform
    const int S = (int)gridDim.x;
    __shared__ int s_esum[kNmsThreads / 32];
    if (threadIdx.x == 0) {
      v.count[seg] = 0;  // leave the candidate count 0 for the next filter pass
      // relaxed: the word itself is the whole message (every CTA emits its OWN lists)
      asm volatile("st.relaxed.gpu.global.s32 [%0], %1;" ::"l"(em.ready + seg), "r"(nkept + 1) : "memory");
    }
    int part = 0;
    for (int i = threadIdx.x; i < seg; i += kNmsThreads) {
      int r;
      for (;;) {
        asm volatile("ld.relaxed.gpu.global.s32 %0, [%1];" : "=r"(r) : "l"(em.ready + i) : "memory");
        if (r != 0) break;
        __nanosleep(256);  // 256 images: 32 k pollers on 256 words would crowd the CTAs still working
      }
      part += r - 1;
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
    if (lane == 0) s_esum[wid] = part;
    __syncthreads();  // also: this CTA's kept lists (phase d) are visible to all its threads
    int out_base = 0;
#pragma unroll
    for (int k = 0; k < kNmsThreads / 32; ++k) out_base += s_esum[k];
    if (threadIdx.x == 0 && seg == S - 1) {
      const int total = out_base + nkept;
      *em.out_K = min(total, em.capacity);
      if (total > em.capacity) atomicOr(em.err_flag, 4);
    }
    for (int r = threadIdx.x; r < nkept; r += kNmsThreads) {
      const int pos = out_base + r;
      if (pos >= em.capacity) break;
      const unsigned long long kkey = kept_key[r];
      const unsigned idx = (unsigned)(kkey & 0xffffffffull);
      const unsigned cls = idx / (unsigned)em.N;
      em.kb[pos] = seg;
      em.kc[pos] = (long long)cls;
      em.ki[pos] = (long long)(idx - cls * (unsigned)em.N);
      reinterpret_cast<float4*>(em.tlbr)[pos] = kept_box[r];
      em.conf[pos] = bits_to_score((unsigned)(kkey >> 32));
    }
    // (the ready words are put back to zero by the NEXT call's filter kernel, nms_filter_kernel)
  }
#ifdef YH_SEG_PROF
  YH_STAMP();
  if (threadIdx.x == 0 && seg == 0) {
    printf("SEGPROF seg=%d K=%d kept=%d total=%llu:", seg, K, nkept, T[ti_ - 1] - T[0]);
    for (int i = 1; i < ti_; ++i) printf(" %llu", T[i] - T[i - 1]);
    printf(" ns (wait | count | sort | a | b | c | d+end)\n");
  }
#endif
#undef YH_STAMP
}


// ---------------------------------------------------------------------------------
// sort + greedy NMS of LARGE segments on all SMs
// ---------------------------------------------------------------------------------
// One persistent cooperative launch (one CTA per SM) handles every segment with more than
// kBigMin candidates.  The grid splits into up to kBigGroups groups of >= kBigMinCtas CTAs, each
// with its own barrier, and the big segments are dealt round-robin to the groups (one segment:
// all 148 SMs; 8 images of 200 000 candidates: 8 groups of 18 SMs; 256 buckets of 6 000 rows:
// 37 groups of 4).  Per segment:
//   1. merge sort: every CTA sorts tiles of 8192 keys in shared memory, then log2(tiles) merge
//      passes, each output chunk of 8192 located by a merge-path binary search, loaded as a
//      bitonic sequence (A ascending ++ B descending) and merged in shared memory;
//   2. gather: the (t,l,b,r) box and group of every candidate in sorted order, once, into a
//      buffer laid out like the key lists (the head-reading provider decodes them; every CTA
//      re-decoding every tile would cost more than the tests);
//   3. greedy pass in tiles of 512, software-pipelined over the group barrier: in phase p
//        * CTAs 1.. of the group prepare tile p+1 — its test against the boxes kept by tiles < p
//          (kept list split evenly over the CTAs and staged through shared memory, two threads
//          per candidate, results OR-ed into a 512-bit mask with one atomicOr per warp), its
//          512 x 512 in-tile suppression bit matrix and the 512 x 512 matrix "candidate of tile p
//          suppresses candidate of tile p+1" (both pure functions of the boxes);
//        * CTA 0 resolves tile p with no IoU arithmetic: alive = ~mask & ~OR(rows of what tile
//          p-1 kept), register walk over the in-tile matrix (walk_tile), append — the only
//          serial part, ~10 us per tile.
//      (The first version did the previous-tile test, the matrix and a shuffle walk on the
//      resolver: ~170 us per tile of survivors — one SM's worth of IoU arithmetic.)
// ONE group of 65 536 / 200 000 candidates: 617 ms / - on one CTA (round 1), 21.8 / 50.2 ms with
// the first version, 3.86 / 13.7 ms now (the kept-list test on all SMs is what remains: ncu SM
// 39 % busy, 1.66e9 warp instructions, profiles/r2_nms_big_ncu.txt).
constexpr int kBigThreads = 1024;
constexpr int kBigMin = YH_BIG_MIN;
constexpr int kBigGroups = 64;   // most CTA groups one launch splits into (YH_BIG_GROUPS in common.cuh)
constexpr int kBigMinCtas = 4;   // fewest CTAs of a group: one resolver, three testers
constexpr int kStage = 1024;  // kept boxes a tester CTA stages in shared memory per round (24 KB)

struct BigView {
  float4* box;        // boxes of the candidates in sorted order, laid out like the key lists
  long long* grp;     // (segment base + position): no per-group copy, any number of groups
  unsigned* sync;     // [kBigGroups * 32] barrier counters, zero at launch
  unsigned* dead;     // [kBigGroups * 2 * kTileWords] cross-test masks, zero at launch
  int* nkept;         // [kBigGroups * 2] kept count as of the start of a phase (double-buffered)
  unsigned* mat;      // [kBigGroups][2 parities][in-tile | previous-tile][kTile * kTileWords] bit matrices
};

// barrier of the n CTAs of a group: a monotonically increasing arrival counter
__device__ __forceinline__ void group_sync(unsigned* ctr, unsigned& target, const unsigned n) {
  __syncthreads();
  if (threadIdx.x == 0) {
    target += n;
    __threadfence();
    atomicAdd(ctr, 1u);
    unsigned seen;
    do {
      asm volatile("ld.acquire.gpu.u32 %0, [%1];" : "=r"(seen) : "l"(ctr) : "memory");
    } while (seen < target);
  }
  __syncthreads();
}

// ascending bitonic merge of the bitonic sequence s[0..kSortTile)
__device__ __forceinline__ void smem_merge_asc(unsigned long long* s) {
  for (int j = kSortTile >> 1; j >= 1; j >>= 1) {
    for (int p = threadIdx.x; p < kSortTile / 2; p += kBigThreads) {
      const int i = ((p & ~(j - 1)) << 1) | (p & (j - 1));
      const int l = i | j;
      const unsigned long long a = s[i], c = s[l];
      if (a > c) {
        s[i] = c;
        s[l] = a;
      }
    }
    __syncthreads();
  }
}

// merge path: how many of the first `diag` outputs of merge(A[0..nA), B[0..nB)) come from A
__device__ __forceinline__ int merge_path(const unsigned long long* __restrict__ A, const int nA,
                                          const unsigned long long* __restrict__ B, const int nB,
                                          const int diag) {
  int lo = max(0, diag - nB), hi = min(diag, nA);
  while (lo < hi) {
    const int mid = (lo + hi) >> 1;
    if (__ldcg(A + mid) < __ldcg(B + (diag - 1 - mid))) lo = mid + 1; else hi = mid;
  }
  return lo;
}

template <class P>
__global__ void __launch_bounds__(kBigThreads)
nms_big_kernel(const P prov, const SegView v, const int S, const BigView big,
               const double iou_thr) {
  extern __shared__ __align__(16) unsigned long long s_keys[];
  __shared__ float4 t_box[kTile];
  __shared__ long long t_grp[kTile];
  __shared__ unsigned s_al[2][kTileWords];
  __shared__ unsigned s_keep[32];
  __shared__ int s_list[kBigThreads];
  __shared__ int s_scan[kBigThreads / 32];
  __shared__ int s_split[2];
  static_assert(kBigThreads == 2 * kTile, "two threads per candidate of a tile");
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;

  // ---- which segments are big, and how the grid splits -------------------------------------
  int nbig = 0;
  for (int i = threadIdx.x; i < S; i += kBigThreads) nbig += min(v.count[i], v.cap) > v.big_min;
  for (int o = 16; o > 0; o >>= 1) nbig += __shfl_xor_sync(0xffffffffu, nbig, o);
  if (lane == 0) s_scan[wid] = nbig;
  __syncthreads();
  nbig = 0;
  for (int k = 0; k < kBigThreads / 32; ++k) nbig += s_scan[k];
  __syncthreads();
  if (nbig == 0) return;
  // few big segments: all SMs on each; many: up to kBigGroups groups of >= kBigMinCtas CTAs side
  // by side (256 buckets of 6 000 rows ran as 32 rounds of 8 groups of 18 SMs in the first
  // version: the testers idled, the 8 resolvers set the pace)
  const int G = (int)gridDim.x;
  int ngroups = min(min(nbig, kBigGroups), max(G / kBigMinCtas, 1));
  const int csz = G / ngroups;              // CTAs per group
  const int g = (int)blockIdx.x / csz;      // my group
  const int r = (int)blockIdx.x - g * csz;  // my rank inside it
  if (g >= ngroups) return;                 // left-over CTAs
  YH_CHECK(ngroups <= kBigGroups && csz >= 1);
  unsigned* ctr = big.sync + g * 32;
  unsigned target = 0;
  unsigned* dead = big.dead + g * 2 * kTileWords;
  int* nk_pub = big.nkept + g * 2;

  int seen = 0;  // big segments before the current chunk
  for (int chunk = 0; chunk < S; chunk += kBigThreads) {
    // big segments of this chunk, in order; mine are those whose rank % ngroups == g
    const int si = chunk + threadIdx.x;
    const bool isbig = si < S && min(v.count[si], v.cap) > v.big_min;
    const unsigned bm = __ballot_sync(0xffffffffu, isbig);
    if (lane == 0) s_scan[wid] = __popc(bm);
    __syncthreads();
    int before = 0, total = 0;
    for (int k = 0; k < kBigThreads / 32; ++k) {
      if (k < wid) before += s_scan[k];
      total += s_scan[k];
    }
    const int rank = seen + before + __popc(bm & ((1u << lane) - 1u));
    // compact my group's segments: position = (rank - first rank of mine in this chunk) / ngroups
    const int first = ((g - seen) % ngroups + ngroups) % ngroups;  // chunk-local rank of my first
    if (isbig && rank % ngroups == g) s_list[(rank - seen - first) / ngroups] = si;
    const int nmine = (total > first) ? (total - first + ngroups - 1) / ngroups : 0;
    __syncthreads();
    seen += total;

    for (int li = 0; li < nmine; ++li) {
      const int seg = s_list[li];
      const int K = min(v.count[seg], v.cap);
      const size_t seg_base = v.seg_off ? (size_t)v.seg_off[seg] : (size_t)seg * (size_t)v.stride;
      unsigned long long* keys = v.keys + seg_base;
      unsigned long long* kept_key = v.kept_key + seg_base;
      float4* kept_box = v.kept_box + seg_base;
      long long* kept_grp = v.kept_grp + seg_base;
      float4* bbox = big.box + seg_base;
      long long* bgrp = big.grp + seg_base;

      // ---- 1. merge sort (result in `keys`) -------------------------------------------------
      const int ntile = (K + kSortTile - 1) / kSortTile;
#ifdef YH_BOUNDS
      {  // the segment's room: its stride, or the distance to the next bucket's offset
        const long long room = v.seg_off ? (long long)(v.seg_off[seg + 1] - v.seg_off[seg]) : v.stride;
        YH_CHECK((ntile > 1 ? (long long)ntile * kSortTile : (long long)K) <= room);
      }
#endif
      const int Kp = ntile * kSortTile;  // merge passes only when ntile > 1: then Kp <= next_pow2(K), the segment's room
      int npass = 0;
      for (int run = 1; run < ntile; run <<= 1) ++npass;
      unsigned long long* src = (npass & 1) ? kept_key : keys;
      unsigned long long* dst = (npass & 1) ? keys : kept_key;
      for (int t = r; t < ntile; t += csz) {
        const int base = t * kSortTile;
        for (int i = threadIdx.x; i < kSortTile; i += kBigThreads)
          s_keys[i] = (base + i < K) ? keys[base + i] : ~0ull;
        __syncthreads();
        for (int k = 2; k <= kSortTile; k <<= 1) smem_passes(s_keys, kSortTile, 0, k, k >> 1);
        // several tiles: the padding keys of the last tile are written too (the merges read them;
        // Kp <= the segment's power-of-two room once K > kSortTile).  ONE tile: only the K real
        // keys go back — a segment of 1025..8191 candidates has room for next_pow2(K) < 8192 only
        const int nwrite = ntile > 1 ? kSortTile : K;
        for (int i = threadIdx.x; i < nwrite; i += kBigThreads) src[base + i] = s_keys[i];
        __syncthreads();
      }
      group_sync(ctr, target, csz);
      for (int run = kSortTile; run < Kp; run <<= 1) {
        for (int c = r; c < ntile; c += csz) {
          const int obase = c * kSortTile;
          const int pbase = obase / (2 * run) * (2 * run);
          const unsigned long long* A = src + pbase;
          const int nA = min(run, Kp - pbase);
          const unsigned long long* Bq = A + nA;
          const int nB = min(run, Kp - pbase - nA);
          const int diag = obase - pbase;
          if (threadIdx.x < 2) s_split[threadIdx.x] = merge_path(A, nA, Bq, nB, diag + threadIdx.x * kSortTile);
          __syncthreads();
          const int a0 = s_split[0], a1 = s_split[1];
          const int b0 = diag - a0, na = a1 - a0;
          for (int i = threadIdx.x; i < kSortTile; i += kBigThreads)
            s_keys[i] = (i < na) ? __ldcg(A + a0 + i) : __ldcg(Bq + b0 + (kSortTile - 1 - i));
          __syncthreads();
          smem_merge_asc(s_keys);
          for (int i = threadIdx.x; i < kSortTile; i += kBigThreads) dst[obase + i] = s_keys[i];
          __syncthreads();
        }
        group_sync(ctr, target, csz);
        unsigned long long* tmp = src;
        src = dst;
        dst = tmp;
      }
      // ---- 2. gather the boxes in sorted order ----------------------------------------------
      for (int i = r * kBigThreads + threadIdx.x; i < K; i += csz * kBigThreads) {
        float4 box;
        long long grp;
        prov.get(seg, (unsigned)(__ldcg(keys + i) & 0xffffffffull), box, grp);
        bbox[i] = box;
        bgrp[i] = grp;
      }
      if (r == 0 && threadIdx.x < 2) nk_pub[threadIdx.x] = 0;
      group_sync(ctr, target, csz);

      // ---- 3. greedy pass, pipelined over tiles ---------------------------------------------
      const int ntiles = (K + kTile - 1) / kTile;
      const int ti = threadIdx.x & (kTile - 1);
      const int half = threadIdx.x >> 9;
      unsigned* mask = reinterpret_cast<unsigned*>(s_keys);  // resolver: 32 KB in-tile bit matrix
      int nkept = 0;
      if (csz == 1) {
        // a group of one CTA (a GPU with fewer than two SMs per group): everything on this CTA,
        // unpipelined — the same passes as nms_segment_kernel, kept list read from global memory
        for (int p = 0; p < ntiles; ++p) {
          const int base = p * kTile;
          const int nt = min(K - base, kTile);
          const int nwords = (nt + 31) >> 5;
          bool alive = ti < nt;
          unsigned long long key = 0;
          float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
          long long grp = 0;
          if (alive) {
            key = __ldcg(keys + base + ti);
            box = __ldcg(bbox + base + ti);
            grp = __ldcg(bgrp + base + ti);
            const float area = box_area(box);
            for (int q = half; alive && q < nkept; q += 2) {
              const float4 kb = __ldcg(kept_box + q);
              if (__ldcg(kept_grp + q) == grp && suppresses(kb, box_area(kb), box, area, iou_thr))
                alive = false;
            }
          }
          if (half == 0) {
            t_box[ti] = box;
            t_grp[ti] = grp;
          }
          {
            const unsigned am = __ballot_sync(0xffffffffu, alive);
            if (lane == 0) s_al[half][wid & (kTileWords - 1)] = am;
          }
          __syncthreads();
          for (int item = wid; item < nt * nwords; item += kBigThreads / 32) {
            const int rr = item / nwords, w = item - rr * nwords;
            if (w < (rr >> 5)) continue;
            const int j = w * 32 + lane;
            bool hit = false;
            if (j > rr && j < nt) {
              const float4 br = t_box[rr], bj = t_box[j];
              hit = t_grp[rr] == t_grp[j] && suppresses(br, box_area(br), bj, box_area(bj), iou_thr);
            }
            const unsigned m = __ballot_sync(0xffffffffu, hit);
            if (lane == 0) mask[rr * kTileWords + w] = m;
          }
          __syncthreads();
          if (wid == 0) {
            const unsigned rem = (lane < nwords) ? ~(s_al[0][lane] & s_al[1][lane]) : ~0u;
            s_keep[lane] = walk_tile(mask, rem, nwords, lane);
          }
          __syncthreads();
          int before_k = 0, total_k = 0;
#pragma unroll
          for (int w = 0; w < kTileWords; ++w) {
            const int c = __popc(s_keep[w]);
            if (w < (ti >> 5)) before_k += c;
            total_k += c;
          }
          const unsigned kwm = s_keep[ti >> 5];
          if (half == 0 && ((kwm >> (ti & 31)) & 1u)) {
            const int lp = nkept + before_k + __popc(kwm & ((1u << (ti & 31)) - 1u));
            kept_key[lp] = key;
            kept_box[lp] = box;
            kept_grp[lp] = grp;
          }
          nkept += total_k;
          __threadfence();  // the next tile reads the kept list through the L2
          __syncthreads();
        }
      } else {
        // Phase p (p = -1 .. ntiles-1), one group barrier each:
        //   CTAs 1..  prepare tile p+1 — everything that is a pure function of the boxes plus the
        //             test against the kept list as it stood before tile p:
        //               dead    (512 bits)   box of tile p+1 suppressed by a box kept by tiles < p
        //               m_in    (512 x 512)  in-tile suppression bits, upper triangle
        //               m_prev  (512 x 512)  bit (i, j): candidate i of tile p suppresses j of tile p+1
        //   CTA 0     resolves tile p from what phase p-1 prepared: alive = ~dead & ~OR of the
        //             m_prev rows of the candidates tile p-1 KEPT, then the in-order walk over
        //             m_in — no IoU arithmetic at all on the serial path.
        unsigned* s_prev_keep = reinterpret_cast<unsigned*>(t_grp);  // resolver: keep words of tile p-1
        unsigned* s_or = s_prev_keep + 32;                           // resolver: OR of m_prev rows
        if (r == 0 && threadIdx.x < 32) s_prev_keep[threadIdx.x] = 0u;
        for (int p = -1; p < ntiles; ++p) {
          if (r == 0) {
            if (p >= 0) {
              const int base = p * kTile;
              const int nt = min(K - base, kTile);
              const int nwords = (nt + 31) >> 5;
              const unsigned* m_in = big.mat + ((size_t)(g * 2 + (p & 1)) * 2 + 0) * (kTile * kTileWords);
              const unsigned* m_prev = big.mat + ((size_t)(g * 2 + (p & 1)) * 2 + 1) * (kTile * kTileWords);
              // the in-tile matrix into shared memory (32 KB, two 16-byte loads per thread)
              {
                const uint4* src4 = reinterpret_cast<const uint4*>(m_in);
                uint4* dst4 = reinterpret_cast<uint4*>(mask);
                const uint4 v0 = __ldcg(src4 + threadIdx.x), v1 = __ldcg(src4 + kBigThreads + threadIdx.x);
                dst4[threadIdx.x] = v0;
                dst4[kBigThreads + threadIdx.x] = v1;
              }
              // my candidate (threads 0..511) and its cross-test bit
              bool alive = half == 0 && ti < nt;
              unsigned long long key = 0;
              float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
              long long grp = 0;
              if (alive) {
                key = __ldcg(keys + base + ti);
                box = __ldcg(bbox + base + ti);
                grp = __ldcg(bgrp + base + ti);
                unsigned dw;
                asm volatile("ld.relaxed.gpu.u32 %0, [%1];" : "=r"(dw) : "l"(dead + (p & 1) * kTileWords + (ti >> 5)) : "memory");
                alive = !((dw >> (ti & 31)) & 1u);
              }
              // OR of the m_prev rows of what tile p-1 kept: a warp per 16 rows, lane = word
              if (threadIdx.x < kTileWords) s_or[threadIdx.x] = 0u;
              __syncthreads();
              if (p > 0) {
                unsigned acc = 0u;
                const unsigned pk = s_prev_keep[wid >> 1];          // rows wid*16 .. wid*16+15
                const unsigned bits = (pk >> ((wid & 1) * 16)) & 0xffffu;
                if (lane < kTileWords) {
                  unsigned t = bits;
                  while (t) {
                    const int j = __ffs(t) - 1;
                    t &= t - 1;
                    acc |= __ldcg(m_prev + (wid * 16 + j) * kTileWords + lane);
                  }
                  if (acc) atomicOr(&s_or[lane], acc);
                }
              }
              __syncthreads();
              if (alive) alive = !((s_or[ti >> 5] >> (ti & 31)) & 1u);
              if (half == 0) {
                const unsigned am = __ballot_sync(0xffffffffu, alive);
                if (lane == 0) s_al[0][wid] = am;
              }
              __syncthreads();
              if (threadIdx.x < kTileWords) dead[(p & 1) * kTileWords + threadIdx.x] = 0u;
              if (wid == 0) {
                const unsigned rem = (lane < nwords) ? ~s_al[0][lane] : ~0u;
                s_keep[lane] = walk_tile(mask, rem, nwords, lane);
              }
              __syncthreads();
              int before_k = 0, total_k = 0;
#pragma unroll
              for (int w = 0; w < kTileWords; ++w) {
                const int c = __popc(s_keep[w]);
                if (w < (ti >> 5)) before_k += c;
                total_k += c;
              }
              const unsigned kwm = s_keep[ti >> 5];
              if (half == 0 && ((kwm >> (ti & 31)) & 1u)) {
                const int lp = nkept + before_k + __popc(kwm & ((1u << (ti & 31)) - 1u));
                YH_CHECK(lp >= 0 && lp < K);
                kept_key[lp] = key;
                kept_box[lp] = box;
                kept_grp[lp] = grp;
              }
              nkept += total_k;
              if (threadIdx.x < 32) s_prev_keep[threadIdx.x] = threadIdx.x < kTileWords ? s_keep[threadIdx.x] : 0u;
              // what the testers of phase p+1 work against: everything kept by tiles <= p
              if (threadIdx.x == 0) nk_pub[(p + 1) & 1] = nkept;
            }
          } else if (p + 1 < ntiles) {
            const int q = p + 1;  // the tile this phase prepares
            const int T = csz - 1, tr = r - 1;
            const int base = q * kTile;
            const int nt = min(K - base, kTile);
            // boxes of tile q and of tile q-1 in shared memory
            float4* q_box = reinterpret_cast<float4*>(s_keys);             // 8 KB
            long long* q_grp = reinterpret_cast<long long*>(q_box + kTile);  // 4 KB
            float4* o_box = reinterpret_cast<float4*>(q_grp + kTile);       // 8 KB (tile q-1)
            long long* o_grp = reinterpret_cast<long long*>(o_box + kTile);  // 4 KB
            float4* st_box = reinterpret_cast<float4*>(o_grp + kTile);      // kStage x 16 B
            long long* st_grp = reinterpret_cast<long long*>(st_box + kStage);
            if (half == 0) {
              const bool in = ti < nt;
              q_box[ti] = in ? __ldcg(bbox + base + ti) : make_float4(0.f, 0.f, 0.f, 0.f);
              q_grp[ti] = in ? __ldcg(bgrp + base + ti) : -1ll;
            } else if (q > 0) {
              o_box[ti] = __ldcg(bbox + base - kTile + ti);
              o_grp[ti] = __ldcg(bgrp + base - kTile + ti);
            }
            __syncthreads();
            // (a) my share of the kept list (as of the start of this phase) against tile q
            if (p >= 0) {
              int n0;
              asm volatile("ld.relaxed.gpu.s32 %0, [%1];" : "=r"(n0) : "l"(nk_pub + (p & 1)) : "memory");
              const int lo = (int)((long long)n0 * tr / T), hi = (int)((long long)n0 * (tr + 1) / T);
              const float4 box = q_box[ti];
              const long long grp = q_grp[ti];
              const float area = box_area(box);
              bool hit = false;
              for (int c0 = lo; c0 < hi; c0 += kStage) {
                const int cnt = min(kStage, hi - c0);
                for (int i = threadIdx.x; i < cnt; i += kBigThreads) {
                  st_box[i] = __ldcg(kept_box + c0 + i);
                  st_grp[i] = __ldcg(kept_grp + c0 + i);
                }
                __syncthreads();
                if (ti < nt && !hit) {
                  for (int k = half; k < cnt; k += 2) {
                    const float4 kb = st_box[k];
                    if (st_grp[k] == grp && suppresses(kb, box_area(kb), box, area, iou_thr)) {
                      hit = true;
                      break;
                    }
                  }
                }
                __syncthreads();
              }
              const unsigned m = __ballot_sync(0xffffffffu, hit);
              if (lane == 0 && m) atomicOr(dead + (q & 1) * kTileWords + (wid & (kTileWords - 1)), m);
            }
            // (b) my share of the two bit matrices of tile q: one ballot per (row, 32-column word)
            unsigned* m_in = big.mat + ((size_t)(g * 2 + (q & 1)) * 2 + 0) * (kTile * kTileWords);
            unsigned* m_prev = big.mat + ((size_t)(g * 2 + (q & 1)) * 2 + 1) * (kTile * kTileWords);
            const int per_mat = kTile * kTileWords;
            const int nitems = (q > 0 ? 2 : 1) * per_mat;
            for (int item = tr * (kBigThreads / 32) + wid; item < nitems; item += T * (kBigThreads / 32)) {
              const int which = item / per_mat;
              const int rr = (item - which * per_mat) / kTileWords;
              const int w = item & (kTileWords - 1);
              const int j = w * 32 + lane;
              bool hit = false;
              if (which == 0) {
                if (w < (rr >> 5) || rr >= nt) continue;  // lower triangle / past the end: never read
                if (j > rr && j < nt) {
                  const float4 br = q_box[rr], bj = q_box[j];
                  hit = q_grp[rr] == q_grp[j] && suppresses(br, box_area(br), bj, box_area(bj), iou_thr);
                }
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                YH_CHECK(rr >= 0 && rr < kTile && w >= 0 && w < kTileWords);
                if (lane == 0) m_in[rr * kTileWords + w] = m;
              } else {
                if (j < nt) {
                  const float4 br = o_box[rr], bj = q_box[j];
                  hit = o_grp[rr] == q_grp[j] && suppresses(br, box_area(br), bj, box_area(bj), iou_thr);
                }
                const unsigned m = __ballot_sync(0xffffffffu, hit);
                if (lane == 0) m_prev[rr * kTileWords + w] = m;
              }
            }
          }
          group_sync(ctr, target, csz);
        }
      }
      if (r == 0 && threadIdx.x == 0) v.kept_count[seg] = nkept;
    }
  }
}

// ---------------------------------------------------------------------------------
// ordered compaction of the kept lists
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(256)
nms_emit_kernel(const SegView v, const int S, const int N, long long* __restrict__ kb,
                long long* __restrict__ kc, long long* __restrict__ ki, float* __restrict__ tlbr,
                float* __restrict__ conf, long long* __restrict__ keep_idx,
                int* __restrict__ out_K, const int capacity, int* __restrict__ err_flag) {
  __shared__ int s_warp[8];
  __shared__ int s_base;
  griddep_launch();
  griddep_wait();
  const int seg = blockIdx.x;
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  int part = 0;
  for (int i = threadIdx.x; i < seg; i += 256) part += v.kept_count[i];
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (lane == 0) s_warp[wid] = part;
  __syncthreads();
  const int mine = v.kept_count[seg];
  if (threadIdx.x == 0) {
    // both greedy kernels have read the candidate count: leave it 0 for the next filter pass
    // (a per-call memset node would sit behind the dense store stream in the overlapped step)
    v.count[seg] = 0;
    int tot = 0;
    for (int i = 0; i < 8; ++i) tot += s_warp[i];
    s_base = tot;
    if (seg == S - 1) {
      const int total = tot + mine;
      *out_K = min(total, capacity);
      if (total > capacity) atomicOr(err_flag, 4);
    }
  }
  __syncthreads();
  const int base = s_base;
  const size_t seg_base = v.seg_off ? (size_t)v.seg_off[seg] : (size_t)seg * (size_t)v.stride;
  const unsigned long long* kk = v.kept_key + seg_base;
  const float4* bx = v.kept_box + seg_base;
  for (int r = threadIdx.x; r < mine; r += 256) {
    const int pos = base + r;
    if (pos >= capacity) break;
    const unsigned long long key = kk[r];
    const unsigned idx = (unsigned)(key & 0xffffffffull);
    if (keep_idx) {
      keep_idx[pos] = (long long)idx;
    } else {
      const unsigned cls = idx / (unsigned)N;
      kb[pos] = seg;
      kc[pos] = (long long)cls;
      ki[pos] = (long long)(idx - cls * (unsigned)N);
      reinterpret_cast<float4*>(tlbr)[pos] = bx[r];
      conf[pos] = bits_to_score((unsigned)(key >> 32));
    }
  }
}

// ---------------------------------------------------------------------------------
// YoloInference class pick (inference.rs:50-77): among kept rows of one (batch, instance)
// keep the max-confidence one; max_by_key keeps the LAST maximum, i.e. the highest row.
// One CTA: slot table init -> atomicMax -> ordered in-place compaction.
// ---------------------------------------------------------------------------------
__global__ void __launch_bounds__(kNmsThreads)
pick_kernel(long long* __restrict__ kb, long long* __restrict__ kc, long long* __restrict__ ki,
            float* __restrict__ tlbr, float* __restrict__ conf, int* __restrict__ io_K,
            const int capacity, unsigned long long* __restrict__ slot, const int max_batch,
            const long long max_instances) {
  __shared__ int s_warp[kNmsThreads / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int K = min(*io_K, capacity);
  for (int i = threadIdx.x; i < K; i += kNmsThreads) {
    const long long b = kb[i], n = ki[i];
    if (b >= 0 && b < max_batch && n >= 0 && n < max_instances) slot[b * max_instances + n] = 0ull;
  }
  __syncthreads();
  for (int i = threadIdx.x; i < K; i += kNmsThreads) {
    const long long b = kb[i], n = ki[i];
    if (b >= 0 && b < max_batch && n >= 0 && n < max_instances)
      atomicMax(&slot[b * max_instances + n],
                ((unsigned long long)__float_as_uint(conf[i]) << 32) | (unsigned)i);
  }
  __syncthreads();
  int base = 0;
  for (int tile = 0; tile < K; tile += kNmsThreads) {
    const int i = tile + threadIdx.x;
    bool keep = false;
    long long b = 0, c = 0, n = 0;
    float4 box = make_float4(0.f, 0.f, 0.f, 0.f);
    float cf = 0.f;
    if (i < K) {
      b = kb[i];
      c = kc[i];
      n = ki[i];
      box = reinterpret_cast<const float4*>(tlbr)[i];
      cf = conf[i];
      const bool in = b >= 0 && b < max_batch && n >= 0 && n < max_instances;
      keep = !in || slot[b * max_instances + n] ==
                        (((unsigned long long)__float_as_uint(cf) << 32) | (unsigned)i);
    }
    const unsigned m = __ballot_sync(0xffffffffu, keep);
    if (lane == 0) s_warp[wid] = __popc(m);
    __syncthreads();
    int before = 0, total = 0;
#pragma unroll
    for (int k = 0; k < kNmsThreads / 32; ++k) {
      const int cnt = s_warp[k];
      if (k < wid) before += cnt;
      total += cnt;
    }
    if (keep) {
      const int pos = base + before + __popc(m & ((1u << lane) - 1u));
      kb[pos] = b;
      kc[pos] = c;
      ki[pos] = n;
      reinterpret_cast<float4*>(tlbr)[pos] = box;
      conf[pos] = cf;
    }
    base += total;
    __syncthreads();
  }
  if (threadIdx.x == 0) *io_K = base;
}

// ---------------------------------------------------------------------------------
// yh_nms_by_scores on many rows: rows of different groups never interact, so the rows are
// hashed by group into kBySegs buckets (count -> offsets -> fill), every bucket is one segment
// of nms_segment_kernel (a CTA each, in parallel, and a candidate only ever meets the kept boxes
// of its own bucket), and the kept rows of all buckets are merged back into one score-ordered
// list by a final sort.  (One segment for everything — the first version — tests every
// candidate against every kept box of every group on ONE SM: 417 ms for 65 536 rows of 80 classes.)
// ---------------------------------------------------------------------------------
constexpr int kBySegs = 256;

// Group ids are usually small dense integers (batch*C + class, nms.rs:129-133): the low 8 bits
// taken as they are put up to 256 consecutive ids into 256 DIFFERENT buckets (a random hash of 80
// ids puts two of them into one bucket for ~11 buckets, and the fullest bucket sets the time);
// the upper bits are hashed in so that ids that differ by multiples of 256 still spread.
__device__ __forceinline__ int group_bucket(long long g) {
  static_assert(kBySegs == 256, "low 8 bits");
  unsigned long long x = (unsigned long long)g >> 8;
  x ^= x >> 33;
  x *= 0xff51afd7ed558ccdULL;
  x ^= x >> 33;
  return (int)(((unsigned long long)g ^ x) & 255ull);
}

__global__ void __launch_bounds__(256)
by_count_kernel(const long long* __restrict__ groups, const int K, int* __restrict__ count) {
  __shared__ int s_hist[kBySegs];
  for (int i = threadIdx.x; i < kBySegs; i += 256) s_hist[i] = 0;
  __syncthreads();
  for (int i = blockIdx.x * 256 + threadIdx.x; i < K; i += gridDim.x * 256)
    atomicAdd(&s_hist[group_bucket(groups[i])], 1);
  __syncthreads();
  for (int i = threadIdx.x; i < kBySegs; i += 256)
    if (s_hist[i]) atomicAdd(&count[i], s_hist[i]);
}

// offsets with power-of-two room per bucket (the sort pads), cursors to 0
__global__ void __launch_bounds__(kBySegs)
by_scan_kernel(const int* __restrict__ count, int* __restrict__ off, int* __restrict__ cursor) {
  __shared__ int s_room[kBySegs];
  const int t = threadIdx.x;
  int room = 2;
  while (room < count[t]) room <<= 1;
  s_room[t] = room;
  cursor[t] = 0;
  __syncthreads();
  if (t == 0) {
    int acc = 0;
    for (int i = 0; i < kBySegs; ++i) {
      off[i] = acc;
      acc += s_room[i];
    }
    off[kBySegs] = acc;
  }
}

__global__ void __launch_bounds__(256)
by_fill_kernel(const float* __restrict__ scores, const long long* __restrict__ groups, const int K,
               unsigned long long* __restrict__ keys, const int* __restrict__ off,
               int* __restrict__ cursor) {
  const int i = blockIdx.x * 256 + threadIdx.x;
  if (i >= K) return;
  const int b = group_bucket(groups[i]);
  const int pos = atomicAdd(&cursor[b], 1);
  keys[(size_t)off[b] + pos] = make_key(scores[i], (unsigned)i);
}

// Global score order of the kept rows of all buckets WITHOUT a sort: every bucket's kept list is
// already sorted (keys are unique), so the rank of a kept row is its position in its own list
// plus, for every other bucket, the number of keys below it (one binary search each, the lists
// are L2-resident).  All SMs work; the single-CTA sort this replaces took 4.4 ms for the 182 k
// kept rows of the 200 000-candidate case, 0.9 ms for 65 536.  The bucket counters go back to 0.
constexpr int kRankThreads = 256;

// a warp per kept row (lanes across the buckets); grid sized for the upper bound K kept rows
__global__ void __launch_bounds__(kRankThreads)
by_rank_kernel(const SegView v, long long* __restrict__ keep, int* __restrict__ out_K,
               int* __restrict__ count) {
  static_assert(kBySegs == kRankThreads, "one thread per bucket");
  __shared__ int s_off[kBySegs];
  __shared__ int s_cnt[kBySegs];
  __shared__ int s_pre[kBySegs + 1];  // exclusive prefix of the kept counts
  __shared__ int s_warp[kRankThreads / 32];
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  const int c = v.kept_count[threadIdx.x];
  s_off[threadIdx.x] = v.seg_off[threadIdx.x];
  s_cnt[threadIdx.x] = c;
  int incl = c;
#pragma unroll
  for (int o = 1; o < 32; o <<= 1) {
    const int u = __shfl_up_sync(0xffffffffu, incl, o);
    if (lane >= o) incl += u;
  }
  if (lane == 31) s_warp[wid] = incl;
  __syncthreads();
  int before = 0;
#pragma unroll
  for (int k = 0; k < kRankThreads / 32; ++k)
    if (k < wid) before += s_warp[k];
  s_pre[threadIdx.x] = before + incl - c;
  if (threadIdx.x == kRankThreads - 1) s_pre[kBySegs] = before + incl;
  __syncthreads();
  const int total = s_pre[kBySegs];
  if (blockIdx.x == 0) {
    count[threadIdx.x] = 0;  // candidate counters for the next call
    if (threadIdx.x == 0) *out_K = total;
  }
  const int g = blockIdx.x * (kRankThreads / 32) + wid;  // kept row, in bucket order
  if (g >= total) return;
  int lo = 0, hi = kBySegs;  // the bucket: last one whose prefix is <= g
  while (hi - lo > 1) {
    const int mid = (lo + hi) >> 1;
    if (s_pre[mid] <= g) lo = mid; else hi = mid;
  }
  const int seg = lo, r = g - s_pre[seg];
  const unsigned long long key = v.kept_key[(size_t)s_off[seg] + r];
  int part = 0;
  for (int t = lane; t < kBySegs; t += 32) {
    const int n = s_cnt[t];
    if (t == seg || n == 0) continue;
    const unsigned long long* o = v.kept_key + (size_t)s_off[t];
    int a = 0, b = n;  // first position whose key is not below `key`
    while (a < b) {
      const int mid = (a + b) >> 1;
      if (__ldg(o + mid) < key) a = mid + 1; else b = mid;
    }
    part += a;
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) part += __shfl_xor_sync(0xffffffffu, part, o);
  if (lane == 0) keep[r + part] = (long long)(key & 0xffffffffull);
}


static int next_pow2(long long x) {
  long long p = 1;
  while (p < x) p <<= 1;
  return (int)p;
}

// cudaFuncSetAttribute is per device: one flag per (kernel instantiation, device).  The
// reference calls this path from one OS thread per device inside one process
// (train/src/train/multi_gpu.rs:578-646); a race between threads only repeats the call.
constexpr int kMaxDevices = 64;

template <class P>
static int launch_segments(YhWorkspace* ws, const P& prov, SegView v, int S, double iou_thr,
                           cudaStream_t s, SegEmit em = SegEmit{}) {
  static bool seg_attr[kMaxDevices] = {};
  static bool big_attr[kMaxDevices] = {};
  const int dev = ws->device;
  const size_t smem = kSegSmem;
  if (dev >= kMaxDevices || !seg_attr[dev]) {
    YH_CUDA(cudaFuncSetAttribute(nms_segment_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)smem));
    if (dev < kMaxDevices) seg_attr[dev] = true;
  }
  // segments above kBigMin candidates need the big buffers; a workspace without them (candidate
  // capacity <= kBigMin) cannot hold such a segment in the first place
  const bool big = ws->big_box != nullptr && v.cap > kBigMin;
  v.big_min = big ? kBigMin : 0;
  YH_TIME(YH_K_NMS_SEGMENT, s);
  if (big) em.on = 0;  // the large-segment kernel finishes some of the lists: nms_emit_kernel behind it
  YH_CUDA(launch_pdl(nms_segment_kernel<P>, dim3(S), dim3(kNmsThreads), smem, s, prov, v, iou_thr, em));
  YH_LAUNCHED(s);
  if (!big) return YH_OK;
  // the multi-CTA kernel: one CTA per SM, co-resident (cooperative launch: its groups spin on
  // their own barriers); exits at once when no segment is big
  const size_t big_smem = (size_t)kSortTile * sizeof(unsigned long long);
  if (dev >= kMaxDevices || !big_attr[dev]) {
    YH_CUDA(cudaFuncSetAttribute(nms_big_kernel<P>, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)big_smem));
    if (dev < kMaxDevices) big_attr[dev] = true;
  }
  BigView bv;
  bv.box = ws->big_box;
  bv.grp = ws->big_grp;
  bv.sync = ws->big_sync;
  bv.dead = ws->big_sync + kBigGroups * 32;
  bv.nkept = (int*)(ws->big_sync + kBigGroups * 32 + kBigGroups * 2 * kTileWords);
  bv.mat = ws->big_mat;
  YH_CUDA(cudaMemsetAsync(ws->big_sync, 0, YH_BIG_SYNC_WORDS * sizeof(unsigned), s));
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3((unsigned)ws->num_sms);
  cfg.blockDim = dim3(kBigThreads);
  cfg.dynamicSmemBytes = big_smem;
  cfg.stream = s;
  cudaLaunchAttribute attr[1];
  attr[0].id = cudaLaunchAttributeCooperative;
  attr[0].val.cooperative = 1;
  cfg.attrs = attr;
  cfg.numAttrs = 1;
  YH_TIME(YH_K_NMS_BIG, s);
  YH_CUDA(cudaLaunchKernelEx(&cfg, nms_big_kernel<P>, prov, v, S, bv, iou_thr));
  YH_LAUNCHED(s);
  return YH_OK;
}

static SegView make_view(YhWorkspace* ws) {
  SegView v;
  v.keys = ws->cand_key;
  v.kept_key = ws->kept_key;
  v.kept_box = ws->kept_box;
  v.kept_grp = ws->kept_grp;
  v.count = ws->cand_count;
  v.kept_count = ws->kept_count;
  v.stride = ws->cand_stride;
  v.cap = ws->cand_per_image;
  v.seg_off = nullptr;
  v.need = ws->cand_need;
  v.big_min = 0;
  return v;
}

}  // namespace yh

extern "C" {

int yh_nms(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L, int B, int C,
           double confidence_threshold, double iou_threshold, int suppress_by_class, int64_t* kb,
           int64_t* kc, int64_t* ki, float* tlbr, float* conf, int32_t* out_K_dev, int capacity,
           void* stream) {
  using namespace yh;
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  YH_REQUIRE(kb && kc && ki && tlbr && conf && out_K_dev, "output pointer is NULL");
  YH_REQUIRE(((uintptr_t)tlbr & 15) == 0, "tlbr not 16-byte aligned");
  YH_REQUIRE(B >= 1 && B <= ws->max_batch && B <= 65535, "B=%d outside 1..max_batch=%d", B,
             ws->max_batch);
  YH_REQUIRE(capacity >= 0, "capacity=%d", capacity);
  // NonMaxSuppressionInit::build (nms.rs:35-39)
  YH_REQUIRE(iou_threshold >= 0.0, "iou_threshold must be non-negative");
  YH_REQUIRE(confidence_threshold >= 0.0, "confidence_threshold must be non-negative");
  HeadBoxes prov;
  int rc = build_layer_tab(layers, L, C, heads, &prov.t);
  if (rc) return rc;
  prov.by_class = suppress_by_class ? 1 : 0;
  YH_REQUIRE((long long)prov.t.N * C < (1ll << 32), "C*N=%lld does not fit the 32-bit key index",
             (long long)prov.t.N * C);
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(ws->device));
  note_batch(B);
  SegView v = make_view(ws);
  // instances per thread.  Sweep inside the overlapped step on B200 (r1f): 8 -> 193.1 us per step,
  // 4 -> 191.3, 2 -> 191.0, 1 -> 184.4..189.9 (kernel alone 46 / 38 / 36 / 35.5 us): many short
  // CTAs interleave better with the other kernels of the step than few long ones.
  static const int filter_per = []() {
    const char* e = getenv("YH_FILTER_PER");
    const int v = e ? atoi(e) : 0;
    return (v == 1 || v == 2 || v == 4 || v == 8) ? v : 1;
  }();
  const int per_cta = kFilterThreads * filter_per;
  dim3 grid((prov.t.N + per_cta - 1) / per_cta, B);
  YH_TIME(YH_K_NMS_FILTER, s);
#define YH_FILTER(P)                                                                          \
  nms_filter_kernel<P><<<grid, kFilterThreads, 0, s>>>(prov.t, (float)confidence_threshold,   \
                                                       v.keys, (int)v.stride, v.cap, v.count, \
                                                       ws->err_flag, ws->seg_ready)
  if (filter_per == 2) YH_FILTER(2);
  else if (filter_per == 4) YH_FILTER(4);
  else if (filter_per == 8) YH_FILTER(8);
  else YH_FILTER(1);
#undef YH_FILTER
  YH_LAUNCHED(s);
  // without large segments the segment kernel emits by itself (SegEmit); YH_NMS_FUSE_EMIT=0: the
  // separate emit kernel as before (A/B)
  static const bool fuse_emit = []() {
    const char* e = getenv("YH_NMS_FUSE_EMIT");
    return !(e && e[0] == '0');
  }();
  const bool big_path = ws->big_box != nullptr && v.cap > kBigMin;
  SegEmit em;
  memset(&em, 0, sizeof(em));
  // ... and only while every image's CTA is resident at once (one per SM: 57 registers x 1024
  // threads): with a second wave a CTA that waits for its predecessors also keeps its SM from the
  // next image (tiny-416 at 256 images: 43 us instead of 33, profiles/r2g_tiny.txt)
  if (fuse_emit && !big_path && B <= ws->num_sms) {
    em.on = 1;
    em.N = prov.t.N;
    em.kb = (long long*)kb;
    em.kc = (long long*)kc;
    em.ki = (long long*)ki;
    em.tlbr = tlbr;
    em.conf = conf;
    em.out_K = out_K_dev;
    em.capacity = capacity;
    em.err_flag = ws->err_flag;
    em.ready = ws->seg_ready;
  }
  rc = launch_segments(ws, prov, v, B, iou_threshold, s, em);
  if (rc) return rc;
  if (em.on) return YH_OK;
  YH_TIME(YH_K_NMS_EMIT, s);
  if (ws->big_box != nullptr && v.cap > kBigMin) {
    // behind the cooperative large-segment kernel: plain launch
    nms_emit_kernel<<<B, 256, 0, s>>>(v, B, prov.t.N, (long long*)kb, (long long*)kc,
                                      (long long*)ki, tlbr, conf, nullptr, out_K_dev, capacity,
                                      ws->err_flag);
  } else {
    YH_CUDA(launch_pdl(nms_emit_kernel, dim3(B), dim3(256), 0, s, v, B, prov.t.N, (long long*)kb,
                       (long long*)kc, (long long*)ki, tlbr, conf, (long long*)nullptr, out_K_dev,
                       capacity, ws->err_flag));
  }
  YH_LAUNCHED(s);
  return YH_OK;
}

int yh_nms_by_scores(YhWorkspace* ws, const float* ltrb, const float* scores,
                     const int64_t* groups, int K, double iou_threshold, int64_t* keep,
                     int32_t* out_K_dev, void* stream) {
  using namespace yh;
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  YH_REQUIRE(K >= 0, "K=%d", K);
  YH_REQUIRE(out_K_dev != nullptr, "out_K is NULL");
  YH_REQUIRE(K == 0 || (ltrb && scores && groups && keep), "NULL argument");
  YH_REQUIRE(((uintptr_t)ltrb & 15) == 0, "ltrb not 16-byte aligned");
  const long long room = (long long)ws->max_batch * ws->cand_stride;
  YH_REQUIRE(K == 0 || next_pow2(K) <= room, "K=%d exceeds the workspace (%lld keys)", K, room);
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(ws->device));
  note_batch(1);
  if (K == 0) {
    YH_CUDA(cudaMemsetAsync(out_K_dev, 0, sizeof(int32_t), s));
    return YH_OK;
  }
  SegView v = make_view(ws);
  ArrayBoxes prov;
  prov.ltrb = ltrb;
  prov.groups = (const long long*)groups;
  // rows above which the bucketed path is used.  B200, 80 groups (tools/nms_small_sweep.py):
  // 1000 rows 0.42 ms as one segment vs 0.10 ms bucketed, 2000 rows 0.92 vs 0.115; with a single
  // group the buckets cost 0.02 ms extra; below ~256 rows both are at the launch floor.
  static const int by_min = []() {
    const char* e = getenv("YH_BY_MIN");
    const int v = e ? atoi(e) : 0;
    return v > 0 ? v : 256;
  }();
  if (K > by_min && 2LL * K + 2LL * kBySegs <= room) {
    // many rows: one segment per group bucket, all buckets in parallel
    int* by_count = ws->by_ints;
    int* by_off = by_count + kBySegs;
    int* by_cursor = by_off + kBySegs + 1;
    int* by_kept = by_cursor + kBySegs;
    v.count = by_count;
    v.kept_count = by_kept;
    v.seg_off = by_off;
    v.cap = K;
    const int blocks = (K + 255) / 256;
    YH_TIME(YH_K_NMS_PACK, s);
    by_count_kernel<<<blocks < 592 ? blocks : 592, 256, 0, s>>>((const long long*)groups, K, by_count);
    YH_LAUNCHED(s);
    YH_TIME(YH_K_NMS_PACK, s);
    by_scan_kernel<<<1, kBySegs, 0, s>>>(by_count, by_off, by_cursor);
    YH_LAUNCHED(s);
    YH_TIME(YH_K_NMS_PACK, s);
    by_fill_kernel<<<blocks, 256, 0, s>>>(scores, (const long long*)groups, K, v.keys, by_off,
                                          by_cursor);
    YH_LAUNCHED(s);
    int rc = launch_segments(ws, prov, v, kBySegs, iou_threshold, s);
    if (rc) return rc;
    YH_TIME(YH_K_NMS_EMIT, s);
    by_rank_kernel<<<(K + kRankThreads / 32 - 1) / (kRankThreads / 32), kRankThreads, 0, s>>>(
        v, (long long*)keep, out_K_dev, by_count);
    YH_LAUNCHED(s);
    return YH_OK;
  }
  v.stride = room;  // one segment over the whole region
  v.cap = K;
  YH_TIME(YH_K_NMS_PACK, s);
  nms_pack_kernel<<<(K + 255) / 256, 256, 0, s>>>(scores, K, v.keys, v.count);
  YH_LAUNCHED(s);
  int rc = launch_segments(ws, prov, v, 1, iou_threshold, s);
  if (rc) return rc;
  YH_TIME(YH_K_NMS_EMIT, s);
  nms_emit_kernel<<<1, 256, 0, s>>>(v, 1, 1, nullptr, nullptr, nullptr, nullptr, nullptr,
                                    (long long*)keep, out_K_dev, K, ws->err_flag);
  YH_LAUNCHED(s);
  return YH_OK;
}

int yh_inference_pick(YhWorkspace* ws, int64_t* kb, int64_t* kc, int64_t* ki, float* tlbr,
                      float* conf, int32_t* io_K_dev, int capacity, void* stream) {
  using namespace yh;
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  YH_REQUIRE(kb && kc && ki && tlbr && conf && io_K_dev, "NULL argument");
  YH_REQUIRE(((uintptr_t)tlbr & 15) == 0, "tlbr not 16-byte aligned");
  YH_REQUIRE(capacity >= 0, "capacity=%d", capacity);
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(ws->device));
  YH_TIME(YH_K_PICK, s);
  pick_kernel<<<1, kNmsThreads, 0, s>>>((long long*)kb, (long long*)kc, (long long*)ki, tlbr, conf,
                                        io_K_dev, capacity, ws->pick_slot, ws->max_batch,
                                        (long long)ws->max_instances);
  YH_LAUNCHED(s);
  return YH_OK;
}

}  // extern "C"
