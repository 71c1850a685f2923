// decode.cu — K1: YOLO head decode (Detect2D::forward, tch-modules/src/detect_2d.rs:42-152),
// optionally fused with the layer merge of MergedDenseDetection::from_detection_tensors
// (tch-goodies/src/detection/merged_dense_detection.rs:60-91).
//
// Streaming kernel: per instance it reads the 4 box logits (entries 0..3 of the entry-major
// head, each a unit-stride run along the cell index) and writes cy,cx,h,w.  Algorithmic traffic
// 32 B per instance.  The grid is (chunks, B) where a chunk is kDecodeThreads*PER
// consecutive cells of the A*HW run of ONE layer (it may cross anchor planes, never layers): no
// per-thread search for the layer, and the grid terms arange(size)/size of detect_2d.rs:122-133
// are two <= 1024-entry tables built once per CTA in shared memory (the reference caches the
// same tables per (H,W), detect_2d.rs:162-238).  A thread: 16-byte loads of its quads of all four
// box entries (all in flight before the tables are built), the sigmoids, 16-byte stores.
// 19x19 / 13x13 planes and odd batch strides are not 16-byte aligned: those use scalar accesses,
// still coalesced.
#include <stdlib.h>

#include "common.cuh"

namespace yh {

constexpr int kDecodeThreads = 256;
constexpr int kGridTabMax = 1024;

struct ChunkTab {
  int cpi[YH_MAX_LAYERS];     // chunks per image of the layer (over its A*HW run)
  int cstart[YH_MAX_LAYERS];  // first chunk of the layer
};

// One CTA = kDecodeThreads * PER consecutive cells of the A*HW RUN of one (image, layer), for ALL
// FOUR box entries.  In the entry-major head, entry e of all A anchors of an image is one
// contiguous run of A*HW floats (channel = e*A + anchor), and the merged output index of cell c of
// that run is flat_start + c — so a chunk may cross anchor planes: 24 chunks of 1024 cells per
// YOLOv4-608 image at 92 % lane use, where chunks confined to one anchor plane needed 27 at 82 %
// (the 19x19 plane filled 35 % of its chunk).  Every thread has 4 * PER/4 16-byte loads in flight
// (cy, cx, h, w logits of its PER cells) while the CTA fills the two grid-offset tables.
template <int PER>
__global__ void __launch_bounds__(kDecodeThreads)
decode_kernel(const LayerTab t, const ChunkTab ct, float* __restrict__ cy, float* __restrict__ cx,
              float* __restrict__ h, float* __restrict__ w, long long out_stride) {
  static_assert(PER % 4 == 0, "whole quads per thread");
  constexpr int Q = PER / 4;
  constexpr int kChunk = kDecodeThreads * PER;
  __shared__ float s_row[kGridTabMax];
  __shared__ float s_col[kGridTabMax];
  const int b = blockIdx.y;
  int l = 0;
#pragma unroll
  for (int i = 1; i < YH_MAX_LAYERS; ++i)
    if (i < t.L && (int)blockIdx.x >= ct.cstart[i]) l = i;
  const int chunk = blockIdx.x - ct.cstart[l];
  const int H = t.H[l], W = t.W[l], HW = t.HW[l], A = t.A[l];
  const int run = A * HW;
  const float Hf = (float)H, Wf = (float)W;
  const bool tab_h = H <= kGridTabMax, tab_w = W <= kGridTabMax;

  // entry e of image b starts e*run floats after entry 0
  const float* __restrict__ p0 = t.head[l] + (size_t)b * (size_t)(run * t.E);
  float a[4][PER];
  int c0[Q], cnt[Q];
#pragma unroll
  for (int q = 0; q < Q; ++q) {
    // quad q of this thread: quads of a warp are adjacent (coalesced 512-byte requests)
    c0[q] = chunk * kChunk + (q * kDecodeThreads + threadIdx.x) * 4;
    cnt[q] = c0[q] < run ? min(4, run - c0[q]) : 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float* p = p0 + (size_t)e * run + c0[q];
      if (cnt[q] == 4 && ((uintptr_t)p & 15) == 0) {
        const float4 v = __ldcs(reinterpret_cast<const float4*>(p));
        a[e][4 * q + 0] = v.x; a[e][4 * q + 1] = v.y; a[e][4 * q + 2] = v.z; a[e][4 * q + 3] = v.w;
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i) a[e][4 * q + i] = (i < cnt[q]) ? __ldcs(p + i) : 0.0f;
      }
    }
  }
  // arange(size)[idx]/size, one IEEE division per table entry instead of one per element
  if (tab_h)
    for (int i = threadIdx.x; i < H; i += kDecodeThreads) s_row[i] = (float)i / Hf;
  if (tab_w)
    for (int i = threadIdx.x; i < W; i += kDecodeThreads) s_col[i] = (float)i / Wf;
  __syncthreads();

#pragma unroll
  for (int q = 0; q < Q; ++q) {
    if (cnt[q] == 0) continue;
    int anchor = c0[q] / HW;
    const int rem = c0[q] - anchor * HW;
    int row = rem / W, col = rem - row * W;
    float o[4][4];
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      // (sigmoid*2 - 0.5)/size + arange(size)[idx]/size   == decode_center()
      const int an = min(anchor, A - 1);
      const float gr = tab_h ? s_row[row] : (float)row / Hf;
      const float gc = tab_w ? s_col[col] : (float)col / Wf;
      o[0][i] = (sigmoid_f(a[0][4 * q + i]) * 2.0f - 0.5f) / Hf + gr;
      o[1][i] = (sigmoid_f(a[1][4 * q + i]) * 2.0f - 0.5f) / Wf + gc;
      o[2][i] = decode_size(sigmoid_f(a[2][4 * q + i]), t.ah[l][an]);
      o[3][i] = decode_size(sigmoid_f(a[3][4 * q + i]), t.aw[l][an]);
      if (++col == W) {
        col = 0;
        if (++row == H) { row = 0; ++anchor; }
      }
    }
    YH_CHECK(c0[q] >= 0 && c0[q] + cnt[q] <= run && t.start[l] + c0[q] + cnt[q] <= out_stride);
    const size_t off = (size_t)b * (size_t)out_stride + (size_t)t.start[l] + (size_t)c0[q];
    float* outs[4] = {cy + off, cx + off, h + off, w + off};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      float* out = outs[e];
      if (cnt[q] == 4 && ((uintptr_t)out & 15) == 0) {
        st_stream4(reinterpret_cast<float4*>(out), make_float4(o[e][0], o[e][1], o[e][2], o[e][3]));
      } else {
#pragma unroll
        for (int i = 0; i < 4; ++i)
          if (i < cnt[q]) st_stream1(out + i, o[e][i]);
      }
    }
  }
}

static int launch_decode(const LayerTab& tab, int B, float* cy, float* cx, float* h, float* w,
                         int64_t out_stride, cudaStream_t s) {
  YH_REQUIRE(B >= 1 && B <= 65535, "B=%d outside 1..65535", B);
  YH_REQUIRE(cy && cx && h && w, "output pointer is NULL");
  YH_REQUIRE(out_stride >= tab.N, "out_batch_stride=%lld < instances %d", (long long)out_stride,
             tab.N);
  // cells per thread: 4 (1024-cell chunks) or 8 (2048); sweep knob
  static const int per = []() {
    const char* e = getenv("YH_DECODE_PER");
    const int v = e ? atoi(e) : 0;
    return v == 8 ? 8 : 4;
  }();
  const int chunk_cells = kDecodeThreads * per;
  ChunkTab ct;
  int chunks = 0;
  for (int l = 0; l < YH_MAX_LAYERS; ++l) {
    ct.cpi[l] = l < tab.L ? (tab.A[l] * tab.HW[l] + chunk_cells - 1) / chunk_cells : 1;
    ct.cstart[l] = chunks;
    if (l < tab.L) chunks += ct.cpi[l];
  }
  dim3 grid(chunks, B);
  YH_TIME(YH_K_DECODE, s);
  if (per == 8)
    decode_kernel<8><<<grid, kDecodeThreads, 0, s>>>(tab, ct, cy, cx, h, w, (long long)out_stride);
  else
    decode_kernel<4><<<grid, kDecodeThreads, 0, s>>>(tab, ct, cy, cx, h, w, (long long)out_stride);
  YH_LAUNCHED(s);
  return YH_OK;
}

}  // namespace yh

extern "C" {

int yh_decode(const float* head, int B, int C, const YhLayer* layer, float* cy, float* cx,
              float* h, float* w, int64_t out_batch_stride, void* stream) {
  YH_REQUIRE(layer != nullptr && head != nullptr, "head/layer is NULL");
  YhLayer one = *layer;
  one.flat_start = 0;  // outputs are addressed relative to the layer
  yh::LayerTab tab;
  const float* heads[1] = {head};
  int rc = yh::build_layer_tab(&one, 1, C, heads, &tab);
  if (rc) return rc;
  return yh::launch_decode(tab, B, cy, cx, h, w, out_batch_stride, (cudaStream_t)stream);
}

int yh_decode_merged(const float* const* heads, const YhLayer* layers, int L, int B, int C,
                     float* cy, float* cx, float* h, float* w, int64_t out_batch_stride,
                     void* stream) {
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  yh::LayerTab tab;
  int rc = yh::build_layer_tab(layers, L, C, heads, &tab);
  if (rc) return rc;
  return yh::launch_decode(tab, B, cy, cx, h, w, out_batch_stride, (cudaStream_t)stream);
}

}  // extern "C"
