// fill_bench.cu — scratch microbenchmark (not part of the library): which way of writing a large
// zero-mostly tensor gets closest to the HBM write roofline on B200?  Decides the design of
// loss_dense_kernel's zero-fill fast path.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o fill_bench fill_bench.cu && ./fill_bench
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

// B: grid-stride tiles, float4 streaming stores, UNROLL per thread
template <int UNROLL, int MODE>
__global__ void __launch_bounds__(256) fill_st4(float4* __restrict__ p, long long nquads) {
  const long long tile = 256LL * UNROLL;
  const long long ntiles = (nquads + tile - 1) / tile;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    float4* q = p + t * tile + threadIdx.x;
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      if (t * tile + threadIdx.x + u * 256 < nquads) {
        if (MODE == 0) __stcs(q + u * 256, z);
        else if (MODE == 1) q[u * 256] = z;
        else __stwt(q + u * 256, z);
      }
    }
  }
}

// D: 256-bit stores
template <int UNROLL>
__global__ void __launch_bounds__(256) fill_st8(float* __restrict__ p, long long n8) {
  const long long tile = 256LL * UNROLL;
  const long long ntiles = (n8 + tile - 1) / tile;
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
#pragma unroll
    for (int u = 0; u < UNROLL; ++u) {
      const long long i = t * tile + threadIdx.x + u * 256;
      if (i < n8) {
        float* q = p + i * 8;
        asm volatile("st.global.v8.f32 [%0], {%1,%1,%1,%1,%1,%1,%1,%1};" ::"l"(q), "f"(0.f) : "memory");
      }
    }
  }
}

// C: TMA bulk shared->global from a zeroed shared buffer
template <int BYTES>
__global__ void __launch_bounds__(128) fill_tma(char* __restrict__ p, long long nbytes) {
  extern __shared__ __align__(128) char zbuf[];
  for (int i = threadIdx.x; i < BYTES / 16; i += blockDim.x)
    reinterpret_cast<float4*>(zbuf)[i] = make_float4(0.f, 0.f, 0.f, 0.f);
  asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
  __syncthreads();
  if (threadIdx.x == 0) {
    const long long ntiles = (nbytes + BYTES - 1) / BYTES;
    const uint32_t s = (uint32_t)__cvta_generic_to_shared(zbuf);
    for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
      long long off = t * BYTES;
      uint32_t sz = (uint32_t)((nbytes - off < BYTES) ? (nbytes - off) : BYTES);
      asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(p + off), "r"(s), "r"(sz) : "memory");
      asm volatile("cp.async.bulk.commit_group;" ::: "memory");
    }
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}


// E: one tile per CTA, no loop (the hardware's in-order CTA dispatch keeps the set of tiles
// being written at any moment small and contiguous)
template <int UNROLL>
__global__ void __launch_bounds__(256) fill_once(float4* __restrict__ p, long long nquads) {
  const long long tile = 256LL * UNROLL;
  const long long t = blockIdx.x;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  float4* q = p + t * tile + threadIdx.x;
#pragma unroll
  for (int u = 0; u < UNROLL; ++u)
    if (t * tile + threadIdx.x + u * 256 < nquads) __stcs(q + u * 256, z);
}

// F: persistent, but every CTA owns a CONTIGUOUS range of tiles
template <int UNROLL>
__global__ void __launch_bounds__(256) fill_chunk(float4* __restrict__ p, long long nquads) {
  const long long tile = 256LL * UNROLL;
  const long long ntiles = (nquads + tile - 1) / tile;
  const long long per = (ntiles + gridDim.x - 1) / gridDim.x;
  const long long t0 = blockIdx.x * per, t1 = min(t0 + per, ntiles);
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long t = t0; t < t1; ++t) {
    float4* q = p + t * tile + threadIdx.x;
#pragma unroll
    for (int u = 0; u < UNROLL; ++u)
      if (t * tile + threadIdx.x + u * 256 < nquads) __stcs(q + u * 256, z);
  }
}

// G: grid-stride tiles, but a WARP writes 4 KB contiguous (8 x 512 B back to back)
__global__ void __launch_bounds__(256) fill_warpc(float4* __restrict__ p, long long nquads) {
  const long long tile = 256LL * 8;
  const long long ntiles = (nquads + tile - 1) / tile;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    float4* q = p + t * tile + wid * 256 + lane;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (t * tile + wid * 256 + lane + u * 32 < nquads) __stcs(q + u * 32, z);
  }
}

template <typename F>
static float timeit(const char* name, double bytes, F f) {
  cudaEvent_t a, b;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  for (int i = 0; i < 3; ++i) f();
  CK(cudaDeviceSynchronize());
  float best = 1e9f, tot = 0.f;
  const int R = 10;
  for (int i = 0; i < R; ++i) {
    CK(cudaEventRecord(a));
    f();
    CK(cudaEventRecord(b));
    CK(cudaEventSynchronize(b));
    float ms;
    CK(cudaEventElapsedTime(&ms, a, b));
    best = ms < best ? ms : best;
    tot += ms;
  }
  CK(cudaGetLastError());
  printf("%-34s best %8.1f us  %7.1f GB/s   mean %8.1f us %7.1f GB/s\n", name, best * 1e3, bytes / best / 1e6,
         tot / R * 1e3, bytes / (tot / R) / 1e6);
  return best;
}

int main() {
  const long long nbytes = 494927680LL / 16 * 16;  // the YOLOv4-608 B=64 head gradient
  char* p;
  CK(cudaMalloc(&p, nbytes + 256));
  int sms;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  printf("SMs %d, bytes %lld\n", sms, nbytes);
  const long long nq = nbytes / 16;
  timeit("cudaMemsetAsync", nbytes, [&] { cudaMemsetAsync(p, 0, nbytes); });
  for (int k : {2, 4, 8, 16}) {
    char nm[64];
    snprintf(nm, 64, "st4.cs unroll8 grid %dxSM", k);
    timeit(nm, nbytes, [&] { fill_st4<8, 0><<<sms * k, 256>>>((float4*)p, nq); });
  }
  {
    const long long t8 = (nq + 2047) / 2048, t4 = (nq + 1023) / 1024, t2 = (nq + 511) / 512, t1 = (nq + 255) / 256;
    timeit("once 32KB/CTA grid=ntiles", nbytes, [&] { fill_once<8><<<(unsigned)t8, 256>>>((float4*)p, nq); });
    timeit("once 16KB/CTA grid=ntiles", nbytes, [&] { fill_once<4><<<(unsigned)t4, 256>>>((float4*)p, nq); });
    timeit("once 8KB/CTA grid=ntiles", nbytes, [&] { fill_once<2><<<(unsigned)t2, 256>>>((float4*)p, nq); });
    timeit("once 4KB/CTA grid=ntiles", nbytes, [&] { fill_once<1><<<(unsigned)t1, 256>>>((float4*)p, nq); });
    for (int k : {4, 8, 16}) {
      char nm[64];
      snprintf(nm, 64, "chunk-contiguous unroll8 grid %dxSM", k);
      timeit(nm, nbytes, [&] { fill_chunk<8><<<sms * k, 256>>>((float4*)p, nq); });
      snprintf(nm, 64, "warp-contiguous 4KB grid %dxSM", k);
      timeit(nm, nbytes, [&] { fill_warpc<<<sms * k, 256>>>((float4*)p, nq); });
    }
  }
  timeit("st4.cs unroll4 grid 8xSM", nbytes, [&] { fill_st4<4, 0><<<sms * 8, 256>>>((float4*)p, nq); });
  timeit("st4.cs unroll16 grid 8xSM", nbytes, [&] { fill_st4<16, 0><<<sms * 8, 256>>>((float4*)p, nq); });
  timeit("st4.cs unroll1 grid 8xSM", nbytes, [&] { fill_st4<1, 0><<<sms * 8, 256>>>((float4*)p, nq); });
  timeit("st4 default unroll8 grid 8xSM", nbytes, [&] { fill_st4<8, 1><<<sms * 8, 256>>>((float4*)p, nq); });
  timeit("st4.wt unroll8 grid 8xSM", nbytes, [&] { fill_st4<8, 2><<<sms * 8, 256>>>((float4*)p, nq); });
  timeit("st8 unroll4 grid 8xSM", nbytes, [&] { fill_st8<4><<<sms * 8, 256>>>((float*)p, nbytes / 32); });
  timeit("st8 unroll8 grid 4xSM", nbytes, [&] { fill_st8<8><<<sms * 4, 256>>>((float*)p, nbytes / 32); });
  CK(cudaFuncSetAttribute(fill_tma<65536>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  for (int k : {1, 2, 3}) {
    char nm[64];
    snprintf(nm, 64, "tma bulk 16KB grid %dxSM", k);
    timeit(nm, nbytes, [&] { fill_tma<16384><<<sms * k, 128, 16384>>>(p, nbytes); });
    snprintf(nm, 64, "tma bulk 32KB grid %dxSM", k);
    timeit(nm, nbytes, [&] { fill_tma<32768><<<sms * k, 128, 32768>>>(p, nbytes); });
    snprintf(nm, 64, "tma bulk 64KB grid %dxSM", k);
    timeit(nm, nbytes, [&] { fill_tma<65536><<<sms * k, 128, 65536>>>(p, nbytes); });
  }
  timeit("tma bulk 4KB grid 4xSM", nbytes, [&] { fill_tma<4096><<<sms * 4, 128, 4096>>>(p, nbytes); });
  // copy for reference (read+write), same size
  char* q;
  CK(cudaMalloc(&q, nbytes));
  timeit("cudaMemcpyAsync D2D (r+w bytes)", 2.0 * nbytes, [&] { cudaMemcpyAsync(q, p, nbytes, cudaMemcpyDeviceToDevice); });
  return 0;
}
