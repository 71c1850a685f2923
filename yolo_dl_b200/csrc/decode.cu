// decode.cu — K1: YOLO head decode (Detect2D::forward, tch-modules/src/detect_2d.rs:42-152),
// optionally fused with the layer merge of MergedDenseDetection::from_detection_tensors
// (tch-goodies/src/detection/merged_dense_detection.rs:60-91).
//
// HBM-bound streaming kernel: per instance it reads the 4 box logits (entries 0..3 of the
// entry-major head, each a unit-stride run along the cell index) and writes cy,cx,h,w.
// Algorithmic traffic 32 B per instance.  The four outputs are independent
// functions of one logit each (cy of entry 0, cx of 1, h of 2, w of 3), so the grid is
// (chunks, B, 4 entries) where a chunk is kDecodeThreads*4 consecutive cells of ONE
// (layer, anchor) plane: a CTA never straddles planes, so there is no per-thread search for
// the layer, the entry is block-uniform (one instantiation per entry), and the grid term
// arange(size)/size of detect_2d.rs:122-133 is a <= 1024-entry table built once per CTA in
// shared memory (the reference caches the same table per (H,W), detect_2d.rs:162-238).
// A thread: one 16-byte load, four sigmoids, one 16-byte store.  19x19 / 13x13 planes and odd
// batch strides are not 16-byte aligned: those use scalar accesses, still coalesced.
#include "common.cuh"

namespace yh {

constexpr int kDecodeThreads = 256;
constexpr int kDecodeChunk = kDecodeThreads * 4;
constexpr int kGridTabMax = 1024;

struct ChunkTab {
  int cpp[YH_MAX_LAYERS];     // chunks per (layer, anchor) plane
  int cstart[YH_MAX_LAYERS];  // first chunk of the layer
};

template <int E>
__device__ __forceinline__ void decode_plane(const LayerTab& t, const ChunkTab& ct,
                                             float* __restrict__ outp, long long out_stride,
                                             float* s_tab) {
  const int b = blockIdx.y;
  int l = 0;
#pragma unroll
  for (int i = 1; i < YH_MAX_LAYERS; ++i)
    if (i < t.L && (int)blockIdx.x >= ct.cstart[i]) l = i;
  const int rel = blockIdx.x - ct.cstart[l];
  const int anchor = rel / ct.cpp[l];
  const int chunk = rel - anchor * ct.cpp[l];
  const int H = t.H[l], W = t.W[l], HW = t.HW[l];
  const float size = (E == 0) ? (float)H : (float)W;
  const int tab_n = (E == 0) ? H : W;
  const bool use_tab = E < 2 && tab_n <= kGridTabMax;
  if (use_tab) {
    for (int i = threadIdx.x; i < tab_n; i += kDecodeThreads) s_tab[i] = (float)i / size;
    __syncthreads();
  }
  const int cell0 = chunk * kDecodeChunk + threadIdx.x * 4;
  if (cell0 >= HW) return;
  const float anc = (E == 2) ? t.ah[l][anchor] : t.aw[l][anchor];
  const float* __restrict__ p =
      t.head[l] + ((size_t)b * (size_t)(t.A[l] * t.E) + (size_t)(E * t.A[l] + anchor)) * (size_t)HW + cell0;
  float* __restrict__ out =
      outp + (size_t)b * (size_t)out_stride + (size_t)t.start[l] + (size_t)anchor * HW + cell0;
  int row = cell0 / W, col = cell0 - row * W;
  const int cnt = min(4, HW - cell0);

  float a[4] = {0.f, 0.f, 0.f, 0.f};
  if (cnt == 4 && ((uintptr_t)p & 15) == 0) {
    const float4 v = __ldcs(reinterpret_cast<const float4*>(p));
    a[0] = v.x; a[1] = v.y; a[2] = v.z; a[3] = v.w;
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (i < cnt) a[i] = __ldcs(p + i);
  }
  float o[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const float s = sigmoid_f(a[i]);
    if (E < 2) {
      const int idx = (E == 0) ? row : col;
      // (sigmoid*2 - 0.5)/size + arange(size)[idx]/size   == decode_center()
      const float g = use_tab ? s_tab[idx] : (float)idx / size;
      o[i] = (s * 2.0f - 0.5f) / size + g;
      if (++col == W) { col = 0; ++row; }
    } else {
      o[i] = decode_size(s, anc);
    }
  }
  if (cnt == 4 && ((uintptr_t)out & 15) == 0) {
    st_stream4(reinterpret_cast<float4*>(out), make_float4(o[0], o[1], o[2], o[3]));
  } else {
#pragma unroll
    for (int i = 0; i < 4; ++i)
      if (i < cnt) st_stream1(out + i, o[i]);
  }
}

__global__ void __launch_bounds__(kDecodeThreads)
decode_kernel(const LayerTab t, const ChunkTab ct, float* __restrict__ cy, float* __restrict__ cx,
              float* __restrict__ h, float* __restrict__ w, long long out_stride) {
  __shared__ float s_tab[kGridTabMax];
  switch (blockIdx.z) {  // block-uniform
    case 0: decode_plane<0>(t, ct, cy, out_stride, s_tab); break;
    case 1: decode_plane<1>(t, ct, cx, out_stride, s_tab); break;
    case 2: decode_plane<2>(t, ct, h, out_stride, s_tab); break;
    default: decode_plane<3>(t, ct, w, out_stride, s_tab); break;
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
  dim3 grid(chunks, B, 4);
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
