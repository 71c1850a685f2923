// corun_bench.cu — scratch microbenchmark (not part of the library): how much does a
// latency-bound kernel (dependent L2 round trips, one CTA per image — the shape of
// assign_slots / nms_segment) slow down next to the HBM store stream of loss_dense, and does
// it matter whether that stream is issued as LSU stores (st.global.cs) or as TMA bulk stores?
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o corun_bench corun_bench.cu && ./corun_bench
#include <cstdio>
#include <cstdlib>
#include <cstdint>
#include <cuda_runtime.h>

#define CK(x) do { cudaError_t e = (x); if (e != cudaSuccess) { printf("%s: %s\n", #x, cudaGetErrorString(e)); exit(1); } } while (0)

__global__ void __launch_bounds__(256) fill_st4(float4* __restrict__ p, long long nquads) {
  const long long tile = 256LL * 8;
  const long long ntiles = (nquads + tile - 1) / tile;
  const float4 z = make_float4(0.f, 0.f, 0.f, 0.f);
  for (long long t = blockIdx.x; t < ntiles; t += gridDim.x) {
    float4* q = p + t * tile + threadIdx.x;
#pragma unroll
    for (int u = 0; u < 8; ++u)
      if (t * tile + threadIdx.x + u * 256 < nquads) __stcs(q + u * 256, z);
  }
}

// TMA bulk stores, at most INFLIGHT groups outstanding per CTA
template <int BYTES, int INFLIGHT>
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
      asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(INFLIGHT - 1) : "memory");
    }
    asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
  }
}

// latency-bound: ROUNDS dependent rounds of (load -> atomicMin -> __syncthreads) on an L2-resident table
__global__ void __launch_bounds__(512) chaser(unsigned* __restrict__ table, int n, int rounds,
                                              unsigned long long* __restrict__ stamps) {
  unsigned long long t0;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t0));
  unsigned* tb = table + (size_t)blockIdx.x * n;
  unsigned idx = threadIdx.x * 37u % n;
  for (int r = 0; r < rounds; ++r) {
    const unsigned v = __ldcg(tb + idx);
    atomicMin(tb + (idx * 13u + 7u) % n, v + r);
    idx = (v + idx * 31u + 1u) % n;
    __syncthreads();
  }
  unsigned long long t1;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t1));
  if (threadIdx.x == 0) {
    atomicMin(&stamps[0], t0);
    atomicMax(&stamps[1], t1);
  }
}

int main() {
  const long long nbytes = 494927680LL / 16 * 16;
  char* p;
  CK(cudaMalloc(&p, nbytes + 256));
  int sms;
  CK(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0));
  const int n = 22743, imgs = 64, rounds = 8;
  unsigned* table;
  CK(cudaMalloc(&table, (size_t)imgs * n * 4));
  CK(cudaMemset(table, 0x11, (size_t)imgs * n * 4));
  unsigned long long* stamps;
  CK(cudaMallocManaged(&stamps, 16));
  cudaStream_t s1, s2, s3;
  CK(cudaStreamCreate(&s1));
  int lo, hi;
  CK(cudaDeviceGetStreamPriorityRange(&lo, &hi));
  CK(cudaStreamCreateWithPriority(&s2, cudaStreamNonBlocking, hi));
  CK(cudaStreamCreateWithPriority(&s3, cudaStreamNonBlocking, hi));
  cudaEvent_t a, b, go;
  CK(cudaEventCreate(&a));
  CK(cudaEventCreate(&b));
  CK(cudaEventCreate(&go));
  CK(cudaFuncSetAttribute(fill_tma<32768, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  CK(cudaFuncSetAttribute(fill_tma<32768, 1>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  CK(cudaFuncSetAttribute(fill_tma<32768, 4>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));
  CK(cudaFuncSetAttribute(fill_tma<16384, 2>, cudaFuncAttributeMaxDynamicSharedMemorySize, 65536));

  auto run = [&](const char* name, int mode, int k, bool chase = true) {
    float fill_best = 1e9f;
    double chase_sum = 0;
    const int R = 8;
    for (int i = 0; i < R + 2; ++i) {
      stamps[0] = ~0ull;
      stamps[1] = 0;
      CK(cudaDeviceSynchronize());
      CK(cudaEventRecord(a, s1));
      const long long nq = nbytes / 16;
      if (mode == 1) fill_st4<<<sms * k, 256, 0, s1>>>((float4*)p, nq);
      if (mode == 2) fill_tma<32768, 2><<<sms * k, 128, 32768, s1>>>(p, nbytes);
      if (mode == 3) fill_tma<32768, 1><<<sms * k, 128, 32768, s1>>>(p, nbytes);
      if (mode == 4) fill_tma<32768, 4><<<sms * k, 128, 32768, s1>>>(p, nbytes);
      if (mode == 5) fill_tma<16384, 2><<<sms * k, 128, 16384, s1>>>(p, nbytes);
      CK(cudaEventRecord(b, s1));
      if (chase) chaser<<<imgs, 512, 0, s2>>>(table, n, rounds, stamps);
      CK(cudaDeviceSynchronize());
      CK(cudaGetLastError());
      float ms = 0;
      CK(cudaEventElapsedTime(&ms, a, b));
      if (i >= 2) {
        if (ms < fill_best) fill_best = ms;
        chase_sum += (double)(stamps[1] - stamps[0]) / 1000.0;
      }
    }
    printf("%-40s fill %7.1f us (%6.0f GB/s)   chaser %7.1f us\n", name, fill_best * 1e3,
           mode ? nbytes / fill_best / 1e6 : 0.0, chase_sum / R);
  };
  run("chaser alone", 0, 0);
  run("st4.cs grid 8xSM alone", 1, 8, false);
  run("st4.cs grid 16xSM alone", 1, 16, false);
  run("tma 32KB inflight2 grid 2xSM alone", 2, 2, false);
  run("tma 32KB inflight4 grid 2xSM alone", 4, 2, false);
  for (int k : {2, 8, 16}) {
    char nm[64];
    snprintf(nm, 64, "st4.cs grid %dxSM", k);
    run(nm, 1, k);
  }
  for (int k : {1, 2, 4}) {
    char nm[64];
    snprintf(nm, 64, "tma 32KB inflight2 grid %dxSM", k);
    run(nm, 2, k);
    snprintf(nm, 64, "tma 32KB inflight1 grid %dxSM", k);
    run(nm, 3, k);
    snprintf(nm, 64, "tma 32KB inflight4 grid %dxSM", k);
    run(nm, 4, k);
    snprintf(nm, 64, "tma 16KB inflight2 grid %dxSM", k);
    run(nm, 5, k);
  }
  return 0;
}
