// pcie_bench.cu — what the box's host<->device path gives N GPUs at once, with nothing of the
// hot path in the way: bare cudaMemcpyAsync of pinned buffers, H2D and D2H concurrently on two
// streams (the e2e arm of bench.py moves 495 MB each way per step per GPU).
//
//   pcie_bench <device> <MB> <reps> <start_epoch_s> <bind>
//     bind = 0  pinned memory from cudaHostAlloc, thread wherever the OS puts it
//     bind = 1  thread pinned to the CPUs of the GPU's NUMA node (sysfs numa_node of its PCI
//               function), buffers malloc'ed + first-touched there, then cudaHostRegister'ed
//   all processes of a run are given the same start_epoch_s and spin until then, so that their
//   timed regions overlap (tools/pcie_sweep.sh launches 1/2/4/8 of them).
// Prints one line: device, NUMA node, GB/s H2D alone, D2H alone, H2D and D2H while both run.
#ifndef _GNU_SOURCE
#define _GNU_SOURCE
#endif
#include <cuda_runtime.h>
#include <sched.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <sys/time.h>
#include <unistd.h>

#define CK(x)                                                                         \
  do {                                                                                \
    cudaError_t e = (x);                                                              \
    if (e != cudaSuccess) {                                                           \
      fprintf(stderr, "%s: %s\n", #x, cudaGetErrorString(e));                         \
      return 1;                                                                       \
    }                                                                                 \
  } while (0)

static double now_s() {
  struct timeval tv;
  gettimeofday(&tv, nullptr);
  return tv.tv_sec + tv.tv_usec * 1e-6;
}

static int numa_node_of(int device) {
  char bus[64] = "", path[256];
  if (cudaDeviceGetPCIBusId(bus, sizeof(bus), device) != cudaSuccess) return -1;
  for (char* p = bus; *p; ++p)
    if (*p >= 'A' && *p <= 'Z') *p += 32;
  snprintf(path, sizeof(path), "/sys/bus/pci/devices/%s/numa_node", bus);
  FILE* f = fopen(path, "r");
  if (!f) return -1;
  int n = -1;
  if (fscanf(f, "%d", &n) != 1) n = -1;
  fclose(f);
  return n;
}

// pins the calling thread to the CPUs listed in /sys/devices/system/node/node<k>/cpulist
static int bind_to_node(int node) {
  char path[128], buf[4096];
  snprintf(path, sizeof(path), "/sys/devices/system/node/node%d/cpulist", node);
  FILE* f = fopen(path, "r");
  if (!f) return -1;
  if (!fgets(buf, sizeof(buf), f)) {
    fclose(f);
    return -1;
  }
  fclose(f);
  cpu_set_t set;
  CPU_ZERO(&set);
  int count = 0;
  for (char* tok = strtok(buf, ",\n"); tok; tok = strtok(nullptr, ",\n")) {
    int a, b;
    if (sscanf(tok, "%d-%d", &a, &b) == 2) {
      for (int c = a; c <= b; ++c) CPU_SET(c, &set), ++count;
    } else if (sscanf(tok, "%d", &a) == 1) {
      CPU_SET(a, &set), ++count;
    }
  }
  if (!count || sched_setaffinity(0, sizeof(set), &set) != 0) return -1;
  return count;
}

int main(int argc, char** argv) {
  if (argc < 6) {
    fprintf(stderr, "usage: pcie_bench <device> <MB> <reps> <start_epoch_s> <bind>\n");
    return 2;
  }
  const int device = atoi(argv[1]);
  const size_t bytes = (size_t)atol(argv[2]) << 20;
  const int reps = atoi(argv[3]);
  const double start = atof(argv[4]);
  const int bind = atoi(argv[5]);
  CK(cudaSetDevice(device));
  const int node = numa_node_of(device);
  int cpus = 0;
  void *h_in = nullptr, *h_out = nullptr;
  if (bind) {
    if (node >= 0) cpus = bind_to_node(node);
    h_in = aligned_alloc(4096, bytes);
    h_out = aligned_alloc(4096, bytes);
    if (!h_in || !h_out) return 1;
    memset(h_in, 1, bytes);  // first touch on this thread's node
    memset(h_out, 0, bytes);
    CK(cudaHostRegister(h_in, bytes, cudaHostRegisterDefault));
    CK(cudaHostRegister(h_out, bytes, cudaHostRegisterDefault));
  } else {
    CK(cudaHostAlloc(&h_in, bytes, cudaHostAllocDefault));
    CK(cudaHostAlloc(&h_out, bytes, cudaHostAllocDefault));
    memset(h_in, 1, bytes);
    memset(h_out, 0, bytes);
  }
  void *d_in, *d_out;
  CK(cudaMalloc(&d_in, bytes));
  CK(cudaMalloc(&d_out, bytes));
  CK(cudaMemset(d_out, 2, bytes));
  cudaStream_t s_in, s_out;
  CK(cudaStreamCreateWithFlags(&s_in, cudaStreamNonBlocking));
  CK(cudaStreamCreateWithFlags(&s_out, cudaStreamNonBlocking));
  cudaEvent_t a0, a1, b0, b1;
  CK(cudaEventCreate(&a0));
  CK(cudaEventCreate(&a1));
  CK(cudaEventCreate(&b0));
  CK(cudaEventCreate(&b1));
  // warm-up
  CK(cudaMemcpyAsync(d_in, h_in, bytes, cudaMemcpyHostToDevice, s_in));
  CK(cudaMemcpyAsync(h_out, d_out, bytes, cudaMemcpyDeviceToHost, s_out));
  CK(cudaDeviceSynchronize());
  while (now_s() < start) {
  }
  float ms;
  double gb = bytes * (double)reps / 1e9;
  // H2D alone
  CK(cudaEventRecord(a0, s_in));
  for (int i = 0; i < reps; ++i) CK(cudaMemcpyAsync(d_in, h_in, bytes, cudaMemcpyHostToDevice, s_in));
  CK(cudaEventRecord(a1, s_in));
  CK(cudaDeviceSynchronize());
  CK(cudaEventElapsedTime(&ms, a0, a1));
  const double h2d = gb / (ms * 1e-3);
  // D2H alone
  CK(cudaEventRecord(b0, s_out));
  for (int i = 0; i < reps; ++i) CK(cudaMemcpyAsync(h_out, d_out, bytes, cudaMemcpyDeviceToHost, s_out));
  CK(cudaEventRecord(b1, s_out));
  CK(cudaDeviceSynchronize());
  CK(cudaEventElapsedTime(&ms, b0, b1));
  const double d2h = gb / (ms * 1e-3);
  // both directions at once
  CK(cudaEventRecord(a0, s_in));
  CK(cudaEventRecord(b0, s_out));
  for (int i = 0; i < reps; ++i) {
    CK(cudaMemcpyAsync(d_in, h_in, bytes, cudaMemcpyHostToDevice, s_in));
    CK(cudaMemcpyAsync(h_out, d_out, bytes, cudaMemcpyDeviceToHost, s_out));
  }
  CK(cudaEventRecord(a1, s_in));
  CK(cudaEventRecord(b1, s_out));
  CK(cudaDeviceSynchronize());
  float ms_in, ms_out;
  CK(cudaEventElapsedTime(&ms_in, a0, a1));
  CK(cudaEventElapsedTime(&ms_out, b0, b1));
  printf("device %d numa_node %d bind %d cpus %d  MB %zu reps %d  h2d_alone %.1f GB/s  d2h_alone %.1f GB/s  "
         "duplex h2d %.1f d2h %.1f GB/s\n",
         device, node, bind, cpus, bytes >> 20, reps, h2d, d2h, gb / (ms_in * 1e-3), gb / (ms_out * 1e-3));
  return 0;
}
