// decode.cu — K1: YOLO head decode (Detect2D::forward, tch-modules/src/detect_2d.rs:42-152),
// optionally fused with the layer merge of MergedDenseDetection::from_detection_tensors
// (tch-goodies/src/detection/merged_dense_detection.rs:60-91).
//
// HBM-bound streaming kernel: per instance it reads the 4 box logits (entries 0..3 of the
// entry-major head, each a unit-stride run along the cell index) and writes cy,cx,h,w.
// Algorithmic traffic 32 B per instance.  One thread handles QUAD=4
// consecutive flat indices: where the four share (layer, anchor) and the addresses are
// 16-byte aligned it uses one float4 load per entry and one float4 store per output,
// otherwise (19x19 / 13x13 maps, odd batch strides) it falls back to scalar accesses that
// are still coalesced across the warp.
#include "common.cuh"

namespace yh {

constexpr int kDecodeThreads = 256;

__device__ __forceinline__ void decode_one(const LayerTab& t, int l, int anchor, int row,
                                           int col, float x0, float x1, float x2, float x3,
                                           float& cy, float& cx, float& h, float& w) {
  cy = decode_center(sigmoid_f(x0), row, (float)t.H[l]);
  cx = decode_center(sigmoid_f(x1), col, (float)t.W[l]);
  h = decode_size(sigmoid_f(x2), t.ah[l][anchor]);
  w = decode_size(sigmoid_f(x3), t.aw[l][anchor]);
}

__global__ void __launch_bounds__(kDecodeThreads)
decode_kernel(const LayerTab t, float* __restrict__ cy, float* __restrict__ cx,
              float* __restrict__ h, float* __restrict__ w, long long out_stride) {
  const int b = blockIdx.y;
  const int n0 = (blockIdx.x * kDecodeThreads + threadIdx.x) * 4;
  if (n0 >= t.N) return;
  const size_t ob = (size_t)b * (size_t)out_stride + (size_t)n0;

  const Cell c0 = locate(t, n0);
  const int l = c0.layer;
  const size_t estride = (size_t)t.A[l] * (size_t)t.HW[l];
  const float* p = t.head[l] + head_offset(t, l, b, c0.anchor, c0.cell);
  const bool same = (n0 + 3 < t.N) && (c0.cell + 3 < t.HW[l]);
  const bool in_al = (((uintptr_t)p | (uintptr_t)(estride * sizeof(float))) & 15) == 0;

  if (same && in_al) {
    const float4 v0 = __ldcs(reinterpret_cast<const float4*>(p));
    const float4 v1 = __ldcs(reinterpret_cast<const float4*>(p + estride));
    const float4 v2 = __ldcs(reinterpret_cast<const float4*>(p + 2 * estride));
    const float4 v3 = __ldcs(reinterpret_cast<const float4*>(p + 3 * estride));
    const float a0[4] = {v0.x, v0.y, v0.z, v0.w};
    const float a1[4] = {v1.x, v1.y, v1.z, v1.w};
    const float a2[4] = {v2.x, v2.y, v2.z, v2.w};
    const float a3[4] = {v3.x, v3.y, v3.z, v3.w};
    float ocy[4], ocx[4], oh[4], ow[4];
    int row = c0.row, col = c0.col;
#pragma unroll
    for (int i = 0; i < 4; ++i) {
      decode_one(t, l, c0.anchor, row, col, a0[i], a1[i], a2[i], a3[i], ocy[i], ocx[i], oh[i],
                 ow[i]);
      if (++col == t.W[l]) { col = 0; ++row; }
    }
    const bool out_al = (((uintptr_t)(cy + ob) | (uintptr_t)(cx + ob) | (uintptr_t)(h + ob) |
                          (uintptr_t)(w + ob)) & 15) == 0;
    if (out_al) {
      st_stream4(reinterpret_cast<float4*>(cy + ob), make_float4(ocy[0], ocy[1], ocy[2], ocy[3]));
      st_stream4(reinterpret_cast<float4*>(cx + ob), make_float4(ocx[0], ocx[1], ocx[2], ocx[3]));
      st_stream4(reinterpret_cast<float4*>(h + ob), make_float4(oh[0], oh[1], oh[2], oh[3]));
      st_stream4(reinterpret_cast<float4*>(w + ob), make_float4(ow[0], ow[1], ow[2], ow[3]));
    } else {
#pragma unroll
      for (int i = 0; i < 4; ++i) {
        st_stream1(cy + ob + i, ocy[i]);
        st_stream1(cx + ob + i, ocx[i]);
        st_stream1(h + ob + i, oh[i]);
        st_stream1(w + ob + i, ow[i]);
      }
    }
    return;
  }

  // generic path: each of the (up to) four instances located on its own
  float x[4][4];
  Cell cs[4];
  int cnt = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (n0 + i < t.N) {
      cs[i] = locate(t, n0 + i);
      const int li = cs[i].layer;
      const size_t es = (size_t)t.A[li] * (size_t)t.HW[li];
      const float* q = t.head[li] + head_offset(t, li, b, cs[i].anchor, cs[i].cell);
      x[i][0] = __ldcs(q);
      x[i][1] = __ldcs(q + es);
      x[i][2] = __ldcs(q + 2 * es);
      x[i][3] = __ldcs(q + 3 * es);
      cnt = i + 1;
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    if (i < cnt) {
      float ocy, ocx, oh, ow;
      decode_one(t, cs[i].layer, cs[i].anchor, cs[i].row, cs[i].col, x[i][0], x[i][1], x[i][2],
                 x[i][3], ocy, ocx, oh, ow);
      st_stream1(cy + ob + i, ocy);
      st_stream1(cx + ob + i, ocx);
      st_stream1(h + ob + i, oh);
      st_stream1(w + ob + i, ow);
    }
  }
}

static int launch_decode(const LayerTab& tab, int B, float* cy, float* cx, float* h, float* w,
                         int64_t out_stride, cudaStream_t s) {
  YH_REQUIRE(B >= 1 && B <= 65535, "B=%d outside 1..65535", B);
  YH_REQUIRE(cy && cx && h && w, "output pointer is NULL");
  YH_REQUIRE(out_stride >= tab.N, "out_batch_stride=%lld < instances %d", (long long)out_stride,
             tab.N);
  const int quads = (tab.N + 3) / 4;
  dim3 grid((quads + kDecodeThreads - 1) / kDecodeThreads, B);
  YH_TIME(YH_K_DECODE, s);
  decode_kernel<<<grid, kDecodeThreads, 0, s>>>(tab, cy, cx, h, w, (long long)out_stride);
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
