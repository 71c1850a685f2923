// gather_bench.cu — scratch microbenchmark (not part of the library): what is the CEILING of the
// three random-access passes of the step (nms_filter's class logits, loss_rows' matched logits,
// loss_finish's scatter)?  Each moves ~1 M isolated 4-byte elements of a 495 MB tensor, one per
// 128-byte line, and takes 24-35 us alone, i.e. 30-40 G accesses/s whatever the kernel.  If bare
// random loads / stores with unlimited memory-level parallelism reach no more than that, the
// passes are at the DRAM's random-access rate and only fewer touched lines can help; if they
// reach a multiple, the kernels are short of loads in flight.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o gather_bench gather_bench.cu && ./gather_bench
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__device__ __forceinline__ unsigned long long mix(unsigned long long x) {
  x ^= x >> 33; x *= 0xff51afd7ed558ccdULL; x ^= x >> 33; x *= 0xc4ceb9fe1a85ec53ULL; x ^= x >> 33;
  return x;
}

template <int HINT>
__device__ __forceinline__ float ld1(const float* p) {
  float v;
  if (HINT == 64) asm volatile("ld.global.nc.L2::64B.f32 %0, [%1];" : "=f"(v) : "l"(p));
  else if (HINT == 128) asm volatile("ld.global.nc.L2::128B.f32 %0, [%1];" : "=f"(v) : "l"(p));
  else if (HINT == 256) asm volatile("ld.global.nc.L2::256B.f32 %0, [%1];" : "=f"(v) : "l"(p));
  else v = __ldg(p);
  return v;
}

__global__ void __launch_bounds__(256) fill_st4(float4* __restrict__ p, long long nquads) {
  const long long i0 = ((long long)blockIdx.x * 256 + threadIdx.x);
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long i = i0; i < nquads; i += (long long)gridDim.x * 256) __stcs(p + i, z);
}

// K independent random loads, U per thread, all issued before the first use
template <int HINT, int U>
__global__ void __launch_bounds__(256) gather_flat(const float* __restrict__ src, long long n,
                                                   long long K, unsigned long long seed,
                                                   float* __restrict__ sink) {
  const long long t = (long long)blockIdx.x * 256 + threadIdx.x;
  float v[U];
#pragma unroll
  for (int u = 0; u < U; ++u) {
    const long long k = t * U + u;
    v[u] = k < K ? ld1<HINT>(src + mix(seed + k) % n) : 0.f;
  }
  float s = 0.f;
#pragma unroll
  for (int u = 0; u < U; ++u) s += v[u];
  if (s == 1234.5f) sink[0] = s;
}

// the shape of the filter expansion / the rows kernel: R instances at random (image, cell), a warp
// each, 80 class logits `es` floats apart (3 loads per lane in flight)
template <int HINT>
__global__ void __launch_bounds__(128) gather_rows(const float* __restrict__ src, int B, int es,
                                                   int E, int R, unsigned long long seed,
                                                   float* __restrict__ sink) {
  const int r = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= R) return;
  const unsigned long long h = mix(seed + r);
  const long long b = h % B, c = (h >> 32) % es;
  const float* p = src + b * (long long)E * es + c;
  float v[3];
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int k = lane + 32 * j;
    v[j] = k < E - 5 ? ld1<HINT>(p + (long long)(5 + k) * es) : 0.f;
  }
  const float s = v[0] + v[1] + v[2];
  if (s == 1234.5f) sink[0] = s;
}

// K random stores of W floats (W = 1, 4, 8: 4 B, one aligned 16 B quad, one aligned 32 B sector)
template <int W>
__global__ void __launch_bounds__(256) scatter_flat(float* __restrict__ dst, long long n, long long K,
                                                    unsigned long long seed) {
  const long long k = (long long)blockIdx.x * 256 + threadIdx.x;
  if (k >= K) return;
  const long long i = (mix(seed + k) % (n - 8)) & ~(long long)(W - 1);
  if (W == 1) dst[i] = 1.f;
  if (W >= 4) *reinterpret_cast<float4*>(dst + i) = make_float4(1.f, 2.f, 3.f, 4.f);
  if (W == 8) *reinterpret_cast<float4*>(dst + i + 4) = make_float4(1.f, 2.f, 3.f, 4.f);
}

// the shape of the scatter: a warp per row, 85 stores `es` floats apart
__global__ void __launch_bounds__(128) scatter_rows(float* __restrict__ dst, int B, int es, int E,
                                                    int R, unsigned long long seed) {
  const int r = blockIdx.x * 4 + (threadIdx.x >> 5), lane = threadIdx.x & 31;
  if (r >= R) return;
  const unsigned long long h = mix(seed + r);
  const long long b = h % B, c = (h >> 32) % es;
  float* p = dst + b * (long long)E * es + c;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int k = lane + 32 * j;
    if (k < E) p[(long long)k * es] = 1.f;
  }
}


// ---- the dense fill with a per-region store policy, then the scatter of the matched rows -----
// YOLOv4-608 at 64 images: layer 0 = 377 MB, layer 1 = 94 MB, layer 2 = 24 MB, written in this
// order; 29 / 49 / 22 % of the matched rows lie in layers 0 / 1 / 2.  POLICY of the tail
// (layers 1-2, 118 MB <= the 126 MB L2): 0 = st.global.cs like layer 0, 1 = plain st.global,
// 2 = st.global with an L2::evict_last cache policy.
template <int POLICY>
__global__ void __launch_bounds__(256) fill_tiles(float4* __restrict__ p, long long ntiles,
                                                  long long tail_from) {
  const long long tile = blockIdx.x;
  if (tile >= ntiles) return;
  float4* q = p + tile * 2048 + threadIdx.x;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  if (POLICY == 0 || tile < tail_from) {
#pragma unroll
    for (int u = 0; u < 8; ++u) __stcs(q + u * 256, z);
  } else if (POLICY == 1) {
#pragma unroll
    for (int u = 0; u < 8; ++u) q[u * 256] = z;
  } else {
    unsigned long long pol;
    asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
#pragma unroll
    for (int u = 0; u < 8; ++u)
      asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %1, %1, %1}, %2;" ::"l"(q + u * 256), "f"(0.f), "l"(pol) : "memory");
  }
}

// a warp per matched row; rows [0, R0) in layer 0, [R0, R1) in layer 1, the rest in layer 2
__global__ void __launch_bounds__(128) scatter_rows3(float* __restrict__ dst, int B, int E, int R,
                                                     int R0, int R1, unsigned long long seed, int order) {
  int r = blockIdx.x * 4 + (threadIdx.x >> 5);
  const int lane = threadIdx.x & 31;
  if (r >= R) return;
  if (order) r = R - 1 - r;  // tail layers first
  const int es0 = 3 * 76 * 76, es1 = 3 * 38 * 38, es2 = 3 * 19 * 19;
  const long long base1 = (long long)B * E * es0, base2 = base1 + (long long)B * E * es1;
  const int es = r < R0 ? es0 : (r < R1 ? es1 : es2);
  const long long base = r < R0 ? 0 : (r < R1 ? base1 : base2);
  const unsigned long long h = mix(seed + r);
  const long long b = h % B, c = (h >> 32) % es;
  float* p = dst + base + b * (long long)E * es + c;
#pragma unroll
  for (int j = 0; j < 3; ++j) {
    const int k = lane + 32 * j;
    if (k < E) p[(long long)k * es] = 1.f;
  }
}

int main(int argc, char** argv) {
  const bool only_policy = argc > 1;  // any argument: just the fill-policy x scatter part
  // layer 0 of YOLOv4-608 at 64 images: [64, 85, 3*76*76] = 377 MB; the whole head is 495 MB
  const int B = 64, E = 85, es = 3 * 76 * 76;
  const long long n = (long long)B * E * es;
  float *head, *grad, *sink;
  CK(cudaMalloc(&head, n * 4));
  CK(cudaMalloc(&grad, n * 4));
  CK(cudaMalloc(&sink, 256));
  CK(cudaMemset(head, 0, n * 4));
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  unsigned long long seed = 1;

  auto timeit = [&](const char* name, long long accesses, auto launch, bool refill) {
    float best = 1e9f, sum = 0.f;
    const int R = 6;
    for (int i = 0; i < R + 2; ++i) {
      seed += 0x9e3779b97f4a7c15ULL;
      if (refill) fill_st4<<<148 * 16, 256>>>((float4*)grad, n / 4);  // the state loss_dense leaves
      else fill_st4<<<148 * 16, 256>>>((float4*)grad, 160LL * 1024 * 1024 / 16);  // evicts `head` lines
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(a));
      launch();
      CK(cudaEventRecord(b));
      CK(cudaDeviceSynchronize());
      CK(cudaGetLastError());
      float ms;
      CK(cudaEventElapsedTime(&ms, a, b));
      if (i >= 2) { best = ms < best ? ms : best; sum += ms; }
    }
    printf("%-52s %8lld acc  best %7.1f us  mean %7.1f us  %6.1f G acc/s\n", name, accesses,
           best * 1e3, sum / R * 1e3, accesses / (best * 1e-3) / 1e9);
  };

  if (!only_policy)
  for (long long K : {250000LL, 1000000LL, 4000000LL}) {
    const int g1 = (int)((K + 255) / 256), g4 = (int)((K / 4 + 255) / 256);
    timeit("gather_flat ldg           1/thread", K, [&] { gather_flat<0, 1><<<g1, 256>>>(head, n, K, seed, sink); }, false);
    timeit("gather_flat L2::64B       1/thread", K, [&] { gather_flat<64, 1><<<g1, 256>>>(head, n, K, seed, sink); }, false);
    timeit("gather_flat L2::64B       4/thread", K, [&] { gather_flat<64, 4><<<g4, 256>>>(head, n, K, seed, sink); }, false);
    timeit("gather_flat L2::128B      4/thread", K, [&] { gather_flat<128, 4><<<g4, 256>>>(head, n, K, seed, sink); }, false);
    timeit("scatter 4 B  (after a full dense fill)", K, [&] { scatter_flat<1><<<g1, 256>>>(grad, n, K, seed); }, true);
    timeit("scatter 16 B (after a full dense fill)", K, [&] { scatter_flat<4><<<g1, 256>>>(grad, n, K, seed); }, true);
    timeit("scatter 32 B sector (after a full dense fill)", K, [&] { scatter_flat<8><<<g1, 256>>>(grad, n, K, seed); }, true);
  }
  if (!only_policy)
  for (int R : {2972, 11885, 47540}) {
    const long long acc = (long long)R * 80;
    timeit("gather_rows  80 x stride es, ldg", acc, [&] { gather_rows<0><<<(R + 3) / 4, 128>>>(head, B, es, E, R, seed, sink); }, false);
    timeit("gather_rows  80 x stride es, L2::64B", acc, [&] { gather_rows<64><<<(R + 3) / 4, 128>>>(head, B, es, E, R, seed, sink); }, false);
    timeit("scatter_rows 85 x stride es (after dense fill)", (long long)R * 85, [&] { scatter_rows<<<(R + 3) / 4, 128>>>(grad, B, es, E, R, seed); }, true);
  }
  // scatter into lines that ARE resident: fill only 64 MB, scatter inside it
  if (!only_policy) {
    const long long n64 = 64LL * 1024 * 1024 / 4, K = 1000000;
    float best = 1e9f;
    for (int i = 0; i < 6; ++i) {
      seed += 77;
      fill_st4<<<148 * 16, 256>>>((float4*)grad, n64 / 4);
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(a));
      scatter_flat<1><<<(int)((K + 255) / 256), 256>>>(grad, n64, K, seed);
      CK(cudaEventRecord(b));
      CK(cudaDeviceSynchronize());
      float ms;
      CK(cudaEventElapsedTime(&ms, a, b));
      best = ms < best ? ms : best;
    }
    printf("%-52s %8lld acc  best %7.1f us  %6.1f G acc/s\n", "scatter 4 B into a 64 MB region just written", K,
           best * 1e3, K / (best * 1e-3) / 1e9);
  }
  // ---- fill policy of the tail layers x scatter ------------------------------------------------
  {
    const long long nall = (long long)B * E * (3 * 76 * 76 + 3 * 38 * 38 + 3 * 19 * 19);
    float* g3;
    CK(cudaMalloc(&g3, nall * 4 + 65536));
    const long long ntiles = (nall + 8191) / 8192;
    const long long tail_from = ((long long)B * E * es) / 8192;
    const int R = 11788, R0 = (int)(R * 0.29), R1 = (int)(R * 0.78);
    for (int policy = 0; policy < 3; ++policy)
      for (int order = 0; order < 2; ++order) {
        float fbest = 1e9f, sbest = 1e9f, ssum = 0.f;
        for (int i = 0; i < 8; ++i) {
          seed += 12345;
          CK(cudaDeviceSynchronize());
          CK(cudaEventRecord(a));
          if (policy == 0) fill_tiles<0><<<(int)ntiles, 256>>>((float4*)g3, ntiles, tail_from);
          if (policy == 1) fill_tiles<1><<<(int)ntiles, 256>>>((float4*)g3, ntiles, tail_from);
          if (policy == 2) fill_tiles<2><<<(int)ntiles, 256>>>((float4*)g3, ntiles, tail_from);
          CK(cudaEventRecord(b));
          cudaEvent_t c2, d2;
          CK(cudaEventCreate(&c2));
          CK(cudaEventCreate(&d2));
          CK(cudaEventRecord(c2));
          scatter_rows3<<<(R + 3) / 4, 128>>>(g3, B, E, R, R0, R1, seed, order);
          CK(cudaEventRecord(d2));
          CK(cudaDeviceSynchronize());
          CK(cudaGetLastError());
          float f, sc;
          CK(cudaEventElapsedTime(&f, a, b));
          CK(cudaEventElapsedTime(&sc, c2, d2));
          if (i >= 2) { fbest = f < fbest ? f : fbest; sbest = sc < sbest ? sc : sbest; ssum += sc; }
          CK(cudaEventDestroy(c2));
          CK(cudaEventDestroy(d2));
        }
        printf("tail policy %d (0 .cs, 1 plain, 2 evict_last) order %d: fill %6.1f us (%5.0f GB/s)  scatter of %d rows best %5.1f mean %5.1f us\n",
               policy, order, fbest * 1e3, nall * 4 / fbest / 1e6, R, sbest * 1e3, ssum / 6 * 1e3);
      }
  }
  return 0;
}
