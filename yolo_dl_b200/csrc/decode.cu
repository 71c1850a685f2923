// decode.cu — K1: YOLO head decode (Detect2D::forward, tch-modules/src/detect_2d.rs:42-152),
// optionally fused with the layer merge of MergedDenseDetection::from_detection_tensors
// (tch-goodies/src/detection/merged_dense_detection.rs:60-91).
//
// Streaming kernel: per instance it reads the 4 box logits (entries 0..3 of the entry-major
// head, each a unit-stride run along the cell index) and writes cy,cx,h,w.  Algorithmic traffic
// 32 B per instance.  The grid is (chunks, B) where a chunk is kDecodeThreads*4
// consecutive cells of ONE (layer, anchor): a CTA never straddles planes, so there is no
// per-thread search for the layer, and the grid terms arange(size)/size of detect_2d.rs:122-133
// are two <= 1024-entry tables built once per CTA in shared memory (the reference caches the
// same tables per (H,W), detect_2d.rs:162-238).  A thread: four 16-byte loads (one per box
// entry, all in flight before the tables are built), sixteen sigmoids, four 16-byte stores.
// 19x19 / 13x13 planes and odd batch strides are not 16-byte aligned: those use scalar accesses,
// still coalesced.
#include "common.cuh"

namespace yh {

constexpr int kDecodeThreads = 256;
constexpr int kDecodeChunk = kDecodeThreads * 4;
constexpr int kGridTabMax = 1024;

struct ChunkTab {
  int cpp[YH_MAX_LAYERS];     // chunks per (layer, anchor) plane
  int cstart[YH_MAX_LAYERS];  // first chunk of the layer
};

// One CTA = 1024 cells of one (image, layer, anchor) for ALL FOUR box entries: every thread has
// four 16-byte loads in flight (cy, cx, h, w logits of its 4 cells) while the CTA fills the two
// grid-offset tables, instead of one load per thread and four times as many CTAs (r1e: 6912 CTAs
// of one load each, 25 us for 46 MB — bound by per-CTA latency, not by bandwidth or math).
__global__ void __launch_bounds__(kDecodeThreads)
decode_kernel(const LayerTab t, const ChunkTab ct, float* __restrict__ cy, float* __restrict__ cx,
              float* __restrict__ h, float* __restrict__ w, long long out_stride) {
  __shared__ float s_row[kGridTabMax];
  __shared__ float s_col[kGridTabMax];
  const int b = blockIdx.y;
  int l = 0;
#pragma unroll
  for (int i = 1; i < YH_MAX_LAYERS; ++i)
    if (i < t.L && (int)blockIdx.x >= ct.cstart[i]) l = i;
  const int rel = blockIdx.x - ct.cstart[l];
  const int anchor = rel / ct.cpp[l];
  const int chunk = rel - anchor * ct.cpp[l];
  const int H = t.H[l], W = t.W[l], HW = t.HW[l], A = t.A[l];
  const float Hf = (float)H, Wf = (float)W;
  const bool tab_h = H <= kGridTabMax, tab_w = W <= kGridTabMax;

  const int cell0 = chunk * kDecodeChunk + threadIdx.x * 4;
  const bool live = cell0 < HW;
  const int cnt = live ? min(4, HW - cell0) : 0;
  // entry e of (b, anchor) starts e*A*HW floats after entry 0 (channel = e*A + anchor)
  const size_t es = (size_t)A * (size_t)HW;
  const float* __restrict__ p0 =
      t.head[l] + ((size_t)b * (size_t)(A * t.E) + (size_t)anchor) * (size_t)HW + cell0;
  float a[4][4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const float* p = p0 + (size_t)e * es;
    if (cnt == 4 && ((uintptr_t)p & 15) == 0) {
      const float4 v = __ldcs(reinterpret_cast<const float4*>(p));
      a[e][0] = v.x; a[e][1] = v.y; a[e][2] = v.z; a[e][3] = v.w;
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) a[e][i] = (i < cnt) ? __ldcs(p + i) : 0.0f;
    }
  }
  // arange(size)[idx]/size, one IEEE division per table entry instead of one per element
  if (tab_h)
    for (int i = threadIdx.x; i < H; i += kDecodeThreads) s_row[i] = (float)i / Hf;
  if (tab_w)
    for (int i = threadIdx.x; i < W; i += kDecodeThreads) s_col[i] = (float)i / Wf;
  __syncthreads();
  if (!live) return;

  const float ah = t.ah[l][anchor], aw = t.aw[l][anchor];
  int row = cell0 / W, col = cell0 - row * W;
  float o[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    // (sigmoid*2 - 0.5)/size + arange(size)[idx]/size   == decode_center()
    const float gr = tab_h ? s_row[row] : (float)row / Hf;
    const float gc = tab_w ? s_col[col] : (float)col / Wf;
    o[0][i] = (sigmoid_f(a[0][i]) * 2.0f - 0.5f) / Hf + gr;
    o[1][i] = (sigmoid_f(a[1][i]) * 2.0f - 0.5f) / Wf + gc;
    o[2][i] = decode_size(sigmoid_f(a[2][i]), ah);
    o[3][i] = decode_size(sigmoid_f(a[3][i]), aw);
    if (++col == W) { col = 0; ++row; if (row >= H) row = H - 1; }
  }
  const size_t off = (size_t)b * (size_t)out_stride + (size_t)t.start[l] + (size_t)anchor * HW + cell0;
  float* outs[4] = {cy + off, cx + off, h + off, w + off};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    float* out = outs[e];
    if (cnt == 4 && ((uintptr_t)out & 15) == 0) {
      st_stream4(reinterpret_cast<float4*>(out), make_float4(o[e][0], o[e][1], o[e][2], o[e][3]));
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i)
        if (i < cnt) st_stream1(out + i, o[e][i]);
    }
  }
}

static int launch_decode(const LayerTab& tab, int B, float* cy, float* cx, float* h, float* w,
                         int64_t out_stride, cudaStream_t s) {
  YH_REQUIRE(B >= 1 && B <= 65535, "B=%d outside 1..65535", B);
  YH_REQUIRE(cy && cx && h && w, "output pointer is NULL");
  YH_REQUIRE(out_stride >= tab.N, "out_batch_stride=%lld < instances %d", (long long)out_stride,
             tab.N);
  ChunkTab ct;
  int chunks = 0;
  for (int l = 0; l < YH_MAX_LAYERS; ++l) {
    ct.cpp[l] = l < tab.L ? (tab.HW[l] + kDecodeChunk - 1) / kDecodeChunk : 1;
    ct.cstart[l] = chunks;
    if (l < tab.L) chunks += ct.cpp[l] * tab.A[l];
  }
  dim3 grid(chunks, B);
  YH_TIME(YH_K_DECODE, s);
  decode_kernel<<<grid, kDecodeThreads, 0, s>>>(tab, ct, cy, cx, h, w, (long long)out_stride);
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
