// benchmark.cu — the counters of YoloBenchmark::forward (yolo-dl/src/loss/benchmark.rs:34-100) on
// the device.  The reference builds two boolean masks over [B,1,N] and [M,1],
// a [M,C] confidence tensor, a max_dim and five count_nonzero calls, each followed by an
// i64::from(tensor) device->host copy.  Here ONE launch produces the five integers:
//   counts[0] all_count     = B*N                                   (:47)
//   counts[1] all_pos       = #{ sigmoid(obj) >= thr } over B*N     (:44,48)
//   counts[2] matched_count = M                                     (:52)
//   counts[3] matched_pos   = #{ sigmoid(obj_m) >= thr } over M     (:45,53)
//   counts[4] class_hits    = #{ any_c(sigmoid(obj_m)*sigmoid(cls_m,c) >= thr)
//                                && argmax_c cls_m,c == target class }   (:83-88)
// and the host forms the four f64 ratios exactly as benchmark.rs:59-92 does.
// CTAs [0, plane_ctas) stream the objectness planes of the head tensors (4 B per instance, the
// same addressing as loss_dense_kernel<false>); the remaining CTAs take the matched rows, one
// warp per row of the `pred_rows` buffer yh_loss_rows already gathered ([cap, 5+C], contiguous).
// Integer atomics only: the result does not depend on the order of execution.
#include "common.cuh"

namespace yh {

constexpr int kBenchThreads = 256;

__global__ void __launch_bounds__(kBenchThreads)
benchmark_kernel(const LayerTab t, const int B, const float thr, const int plane_ctas,
                 const float* __restrict__ pred_rows, const long long* __restrict__ m_class,
                 const int* __restrict__ M_dev, const int capacity,
                 unsigned long long* __restrict__ counts) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  __shared__ unsigned s_cnt[2][kBenchThreads / 32];
  unsigned c0 = 0, c1 = 0;  // plane CTAs: c0 = all_pos.  row CTAs: c0 = matched_pos, c1 = class hits

  if ((int)blockIdx.x < plane_ctas) {
    long long prefix[YH_MAX_LAYERS + 1];
    prefix[0] = 0;
#pragma unroll
    for (int l = 0; l < YH_MAX_LAYERS; ++l)
      prefix[l + 1] = prefix[l] + ((l < t.L) ? (long long)B * t.A[l] * t.HW[l] : 0LL);
    const long long total = prefix[YH_MAX_LAYERS];
    const long long stride = (long long)plane_ctas * kBenchThreads;
    for (long long i = (long long)blockIdx.x * kBenchThreads + threadIdx.x; i < total; i += stride) {
      int l = 0;
#pragma unroll
      for (int k = 1; k < YH_MAX_LAYERS; ++k)
        if (k < t.L && i >= prefix[k]) l = k;
      const long long li = i - prefix[l];
      const int run = t.A[l] * t.HW[l];
      const long long b = li / run;
      const long long j = li - b * run;
      const float x = __ldcs(t.head[l] + b * ((long long)run * t.E) + 4LL * run + j);
      c0 += sigmoid_f(x) >= thr ? 1u : 0u;
    }
  } else {
    const int M = min(*M_dev, capacity);
    const int warps = kBenchThreads / 32;
    const int row_ctas = gridDim.x - plane_ctas;
    for (int m = ((int)blockIdx.x - plane_ctas) * warps + wid; m < M; m += row_ctas * warps) {
      const float* r = pred_rows + (size_t)m * t.E;
      const float so = sigmoid_f(r[4]);
      // first maximum of the class logits (max_dim) and any(conf >= thr)
      float best = -INFINITY;
      int best_k = 0x7fffffff;
      bool any = false;
      for (int k = lane; k < t.C; k += 32) {
        const float x = r[5 + k];
        any = any || (so * sigmoid_f(x) >= thr);
        if (x > best || (x == best && k < best_k)) {
          best = x;
          best_k = k;
        }
      }
#pragma unroll
      for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int ok = __shfl_xor_sync(0xffffffffu, best_k, o);
        if (ob > best || (ob == best && ok < best_k)) {
          best = ob;
          best_k = ok;
        }
      }
      any = __any_sync(0xffffffffu, any);
      if (lane == 0) {
        c0 += so >= thr ? 1u : 0u;
        c1 += (any && (long long)best_k == m_class[m]) ? 1u : 0u;
      }
    }
  }

#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    c0 += __shfl_xor_sync(0xffffffffu, c0, o);
    c1 += __shfl_xor_sync(0xffffffffu, c1, o);
  }
  if (lane == 0) {
    s_cnt[0][wid] = c0;
    s_cnt[1][wid] = c1;
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    unsigned a = 0, b = 0;
#pragma unroll
    for (int k = 0; k < kBenchThreads / 32; ++k) {
      a += s_cnt[0][k];
      b += s_cnt[1][k];
    }
    if ((int)blockIdx.x < plane_ctas) {
      if (a) atomicAdd(&counts[1], (unsigned long long)a);
    } else {
      if (a) atomicAdd(&counts[3], (unsigned long long)a);
      if (b) atomicAdd(&counts[4], (unsigned long long)b);
    }
    if (blockIdx.x == 0) {
      counts[0] = (unsigned long long)B * (unsigned long long)t.N;
      counts[2] = (unsigned long long)min(*M_dev, capacity);
    }
  }
}

}  // namespace yh

extern "C" int yh_benchmark(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                            int B, int C, const float* pred_rows_dev, const int64_t* m_class,
                            const int32_t* M_dev, int capacity, double confidence_threshold,
                            int64_t* counts5_dev, void* stream) {
  using namespace yh;
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  YH_REQUIRE(counts5_dev != nullptr && M_dev != nullptr, "NULL argument");
  YH_REQUIRE(capacity == 0 || (pred_rows_dev && m_class), "matched rows are NULL");
  YH_REQUIRE(heads != nullptr, "heads is NULL");
  YH_REQUIRE(capacity >= 0, "capacity=%d", capacity);
  YH_REQUIRE(B >= 1 && B <= ws->max_batch, "B=%d outside 1..max_batch=%d", B, ws->max_batch);
  LayerTab tab;
  int rc = build_layer_tab(layers, L, C, heads, &tab);
  if (rc) return rc;
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(ws->device));
  YH_CUDA(cudaMemsetAsync(counts5_dev, 0, 5 * sizeof(int64_t), s));
  const int plane_ctas = ws->num_sms * 4;
  int row_ctas = (capacity + (kBenchThreads / 32) - 1) / (kBenchThreads / 32);
  if (row_ctas > ws->num_sms * 4) row_ctas = ws->num_sms * 4;
  if (row_ctas < 1) row_ctas = 1;
  YH_TIME(YH_K_BENCHMARK, s);
  benchmark_kernel<<<plane_ctas + row_ctas, kBenchThreads, 0, s>>>(
      tab, B, (float)confidence_threshold, plane_ctas, pred_rows_dev, (const long long*)m_class,
      M_dev, capacity, (unsigned long long*)counts5_dev);
  YH_LAUNCHED(s);
  return YH_OK;
}
