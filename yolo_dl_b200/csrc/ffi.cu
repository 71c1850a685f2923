// ffi.cu — workspace, error string, counters: the non-kernel part of the C ABI.
#include <cstdarg>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <vector>

#include "common.cuh"

namespace yh {

static thread_local char g_err[512] = "";
static thread_local int64_t g_launches = 0;

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_err, sizeof(g_err), fmt, ap);
  va_end(ap);
}

int64_t& launch_counter() { return g_launches; }

// PDL pays where the step is launch-latency-bound (8 images per GPU: 53.7 -> 50.4 us, r2) and
// costs where it is bandwidth-bound (64 images: 178 -> 189 us — early-launched dependents sit on
// SM resources the streaming kernels want), so by default it follows the batch size of the call.
static thread_local int g_pdl_batch = 1;
void note_batch(int B) { g_pdl_batch = B; }

bool pdl_enabled() {
  static const int forced = []() {
    const char* e = getenv("YH_PDL");
    return e ? (e[0] == '0' ? 0 : 1) : -1;
  }();
  if (forced >= 0) return forced != 0;
  return g_pdl_batch <= 16;
}

// ---- optional per-launch timing: one CUDA event pair per launch, on the launch stream ----
struct TimedLaunch {
  int id;
  cudaEvent_t a, b;
};
static thread_local bool g_timing = false;
static thread_local std::vector<TimedLaunch>* g_timed = nullptr;
static thread_local std::vector<cudaEvent_t>* g_pool = nullptr;

static cudaEvent_t get_event() {
  if (g_pool && !g_pool->empty()) {
    cudaEvent_t e = g_pool->back();
    g_pool->pop_back();
    return e;
  }
  cudaEvent_t e = nullptr;
  cudaEventCreate(&e);
  return e;
}

// ---- optional in-graph trace: a one-thread kernel before and after every launch writes
// %globaltimer into trace[2*id], trace[2*id+1]; unlike events it can be captured into a CUDA
// graph, so it shows where each kernel really runs inside the overlapped, replayed step
// (tools/timeline.py).  Debug aid only.
static thread_local unsigned long long* g_trace = nullptr;
static thread_local int g_trace_cap = 0;
static thread_local std::vector<int>* g_trace_ids = nullptr;  // kernel id of every traced launch

__global__ void stamp_kernel(unsigned long long* slot) {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  *slot = t;
}

void timing_begin(int id, cudaStream_t s) {
  if (g_trace) {
    // one slot per LAUNCH, in issue order (a ranged loss launches the same kernel several times)
    if (!g_trace_ids) g_trace_ids = new std::vector<int>();
    if ((int)g_trace_ids->size() < g_trace_cap) {
      stamp_kernel<<<1, 1, 0, s>>>(g_trace + 2 * g_trace_ids->size());
      g_trace_ids->push_back(id);
    } else {
      g_trace_ids->push_back(-1 - id);  // over capacity: counted, not stamped
    }
    return;
  }
  if (!g_timing) return;
  if (!g_timed) g_timed = new std::vector<TimedLaunch>();
  TimedLaunch t;
  t.id = id;
  t.a = get_event();
  t.b = get_event();
  cudaEventRecord(t.a, s);
  g_timed->push_back(t);
}

void timing_end(cudaStream_t s) {
  if (g_trace) {
    if (g_trace_ids && !g_trace_ids->empty() && g_trace_ids->back() >= 0)
      stamp_kernel<<<1, 1, 0, s>>>(g_trace + 2 * (g_trace_ids->size() - 1) + 1);
    return;
  }
  if (!g_timing || !g_timed || g_timed->empty()) return;
  cudaEventRecord(g_timed->back().b, s);
}

int build_layer_tab(const YhLayer* layers, int L, int C, const float* const* heads,
                    LayerTab* tab) {
  YH_REQUIRE(layers != nullptr, "layers is NULL");
  YH_REQUIRE(L >= 1 && L <= YH_MAX_LAYERS, "L=%d outside 1..%d", L, YH_MAX_LAYERS);
  YH_REQUIRE(C >= 1, "num_classes=%d must be >= 1", C);
  memset(tab, 0, sizeof(*tab));
  tab->L = L;
  tab->C = C;
  tab->E = 5 + C;
  int64_t expect = 0;
  for (int l = 0; l < L; ++l) {
    const YhLayer& y = layers[l];
    YH_REQUIRE(y.H > 0 && y.W > 0, "layer %d: feature size %dx%d", l, y.H, y.W);
    YH_REQUIRE(y.A >= 1 && y.A <= YH_MAX_ANCHORS, "layer %d: A=%d outside 1..%d", l, y.A,
               YH_MAX_ANCHORS);
    // flat ranges are consecutive (merged_dense_detection.rs:68-72)
    YH_REQUIRE(y.flat_start == expect, "layer %d: flat_start=%lld, expected %lld", l,
               (long long)y.flat_start, (long long)expect);
    expect += (int64_t)y.A * y.H * y.W;
    YH_REQUIRE(expect < (int64_t)1 << 31, "too many instances");
    tab->H[l] = y.H;
    tab->W[l] = y.W;
    tab->A[l] = y.A;
    tab->HW[l] = y.H * y.W;
    tab->start[l] = (int)y.flat_start;
    for (int a = 0; a < y.A; ++a) {
      tab->ah[l][a] = (float)y.anchor_h[a];
      tab->aw[l][a] = (float)y.anchor_w[a];
    }
    if (heads) {
      YH_REQUIRE(heads[l] != nullptr, "heads[%d] is NULL", l);
      tab->head[l] = heads[l];
    }
  }
  tab->N = (int)expect;
  static const int l2hint = []() {
    const char* e = getenv("YH_L2_HINT");
    return e ? atoi(e) : 1;  // 64-byte fills for the isolated logit loads (filter alone 35.3 -> 32.9 us, r2)
  }();
  tab->l2hint = l2hint;
  return YH_OK;
}

int reserve_rows(YhWorkspace* ws, size_t floats) {
  if (floats <= ws->row_grad_floats) return YH_OK;
  YH_CUDA(cudaSetDevice(ws->device));
  // the old buffer may still be in use by work already enqueued on this device
  YH_CUDA(cudaDeviceSynchronize());
  if (ws->row_grad) YH_CUDA(cudaFree(ws->row_grad));
  ws->row_grad = nullptr;
  ws->row_grad_floats = 0;
  YH_CUDA(cudaMalloc((void**)&ws->row_grad, floats * sizeof(float)));
  ws->row_grad_floats = floats;
  return YH_OK;
}

}  // namespace yh

extern "C" {

int yh_workspace_reserve_rows(YhWorkspace* ws, int capacity, int C) {
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  YH_REQUIRE(capacity >= 0 && C >= 1, "capacity=%d C=%d", capacity, C);
  return yh::reserve_rows(ws, (size_t)capacity * (size_t)(5 + C));
}

const char* yh_last_error(void) { return yh::g_err; }

const char* yh_version(void) {
#ifdef YH_BOUNDS
  return "yolohot 0.2 (sm_100a, YH_BOUNDS: device asserts on)";
#else
  return "yolohot 0.2 (sm_100a)";
#endif
}

int64_t yh_launch_count(int reset) {
  int64_t v = yh::g_launches;
  if (reset) yh::g_launches = 0;
  return v;
}

int yh_trace_enable(unsigned long long* dev_buf, int capacity) {
  yh::g_trace = dev_buf;
  yh::g_trace_cap = dev_buf ? capacity : 0;
  if (dev_buf && yh::g_trace_ids) yh::g_trace_ids->clear();
  return YH_OK;
}

int yh_trace_ids(int32_t* ids, int n) {
  const int have = yh::g_trace_ids ? (int)yh::g_trace_ids->size() : 0;
  for (int i = 0; i < n && i < have; ++i) ids[i] = (*yh::g_trace_ids)[i];
  return have;
}

int yh_timing_enable(int on) {
  yh::g_timing = on != 0;
  return YH_OK;
}

int yh_timing_read(double* ms_sum, int64_t* launches, int n) {
  using namespace yh;
  for (int i = 0; i < n; ++i) {
    if (ms_sum) ms_sum[i] = 0.0;
    if (launches) launches[i] = 0;
  }
  if (!g_timed) return YH_OK;
  if (!g_pool) g_pool = new std::vector<cudaEvent_t>();
  for (const TimedLaunch& t : *g_timed) {
    float ms = 0.f;
    YH_CUDA(cudaEventSynchronize(t.b));
    YH_CUDA(cudaEventElapsedTime(&ms, t.a, t.b));
    if (t.id < n) {
      if (ms_sum) ms_sum[t.id] += ms;
      if (launches) launches[t.id] += 1;
    }
    g_pool->push_back(t.a);
    g_pool->push_back(t.b);
  }
  g_timed->clear();
  return YH_OK;
}

int yh_workspace_destroy(YhWorkspace* ws) {
  if (!ws) return YH_OK;
  cudaSetDevice(ws->device);
  cudaFree(ws->slot_dist);
  cudaFree(ws->slot_gt);
  cudaFree(ws->slot_bits);
  cudaFree(ws->img_count);
  cudaFree(ws->img_off);
  cudaFree(ws->pick_slot);
  cudaFree(ws->partials);
  cudaFree(ws->ticket);
  cudaFree(ws->row_grad);
  cudaFree(ws->cand_key);
  cudaFree(ws->cand_count);
  cudaFree(ws->kept_key);
  cudaFree(ws->kept_box);
  cudaFree(ws->kept_count);
  cudaFree(ws->seg_ready);
  cudaFree(ws->kept_grp);
  cudaFree(ws->by_ints);
  cudaFree(ws->big_box);
  cudaFree(ws->big_grp);
  cudaFree(ws->big_sync);
  cudaFree(ws->big_mat);
  cudaFree(ws->cand_need);
  cudaFree(ws->err_flag);
  if (ws->host_counts) cudaFreeHost(ws->host_counts);
  delete ws;
  return YH_OK;
}

int yh_workspace_create(int device, int max_batch, int64_t max_instances, int cand_per_image,
                        YhWorkspace** out) {
  YH_REQUIRE(out != nullptr, "out is NULL");
  *out = nullptr;
  YH_REQUIRE(max_batch >= 1, "max_batch=%d", max_batch);
  YH_REQUIRE(max_instances >= 1 && max_instances < ((int64_t)1 << 31), "max_instances=%lld",
             (long long)max_instances);
  YH_REQUIRE(cand_per_image >= 1, "cand_per_image=%d", cand_per_image);
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e != cudaSuccess || count == 0) {
    yh::set_error("no CUDA device (%s); yolohot has no CPU fallback",
                  e == cudaSuccess ? "device count 0" : cudaGetErrorString(e));
    cudaGetLastError();
    return YH_ERR_CUDA;
  }
  YH_REQUIRE(device >= 0 && device < count, "device=%d outside 0..%d", device, count - 1);
  YH_CUDA(cudaSetDevice(device));

  YhWorkspace* ws = new YhWorkspace();
  memset(ws, 0, sizeof(*ws));
  ws->device = device;
  ws->max_batch = max_batch;
  ws->max_instances = max_instances;
  ws->cand_per_image = cand_per_image;
  cudaDeviceProp prop;
  if (cudaGetDeviceProperties(&prop, device) != cudaSuccess) {
    delete ws;
    yh::set_error("cudaGetDeviceProperties failed");
    return YH_ERR_CUDA;
  }
  ws->num_sms = prop.multiProcessorCount;

  size_t slots = (size_t)max_batch * (size_t)max_instances;
  ws->cand_stride = 2;
  while (ws->cand_stride < cand_per_image) ws->cand_stride <<= 1;
  size_t cands = (size_t)max_batch * (size_t)ws->cand_stride;
#define YH_ALLOC(ptr, bytes)                                                            \
  do {                                                                                  \
    cudaError_t _e = cudaMalloc((void**)&(ptr), (bytes));                               \
    if (_e != cudaSuccess) {                                                            \
      yh::set_error("cudaMalloc(%zu) for %s failed: %s", (size_t)(bytes), #ptr,         \
                    cudaGetErrorString(_e));                                            \
      yh_workspace_destroy(ws);                                                         \
      return YH_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)
  YH_ALLOC(ws->slot_dist, slots * sizeof(unsigned long long));
  YH_ALLOC(ws->slot_gt, slots * sizeof(int));
  YH_ALLOC(ws->slot_bits, (size_t)max_batch * (size_t)((max_instances + 31) / 32) * sizeof(unsigned));
  YH_ALLOC(ws->img_count, (size_t)max_batch * sizeof(int));
  YH_ALLOC(ws->img_off, ((size_t)max_batch + 1) * sizeof(int));
  YH_ALLOC(ws->pick_slot, slots * sizeof(unsigned long long));
  YH_ALLOC(ws->partials, (size_t)YH_PARTIALS_MAX * 4 * sizeof(double));
  YH_ALLOC(ws->ticket, 64);
  YH_ALLOC(ws->cand_key, cands * sizeof(unsigned long long));
  YH_ALLOC(ws->cand_count, (size_t)max_batch * sizeof(int));
  YH_ALLOC(ws->kept_key, cands * sizeof(unsigned long long));
  YH_ALLOC(ws->kept_box, cands * sizeof(float4));
  YH_ALLOC(ws->kept_count, (size_t)max_batch * sizeof(int));
  YH_ALLOC(ws->seg_ready, ((size_t)max_batch + 2) * sizeof(int));
  YH_ALLOC(ws->kept_grp, cands * sizeof(long long));
  YH_ALLOC(ws->by_ints, 2048 * sizeof(int));
  YH_ALLOC(ws->cand_need, 64);
  if (cand_per_image > YH_BIG_MIN) {
    // the multi-CTA NMS kernel: boxes of the candidates in sorted order (same layout as the key
    // lists), barrier words and the bit matrices of the tiles in flight
    YH_ALLOC(ws->big_box, cands * sizeof(float4));
    YH_ALLOC(ws->big_grp, cands * sizeof(long long));
    YH_ALLOC(ws->big_sync, YH_BIG_SYNC_WORDS * sizeof(unsigned));
    YH_ALLOC(ws->big_mat, (size_t)YH_BIG_GROUPS * 2 * 2 * 512 * 16 * sizeof(unsigned));
  }
  YH_ALLOC(ws->err_flag, 64);
#undef YH_ALLOC
  // the matcher's slot table is EMPTY between calls (assign.cu)
  if (cudaMemset(ws->slot_dist, 0xff, slots * sizeof(unsigned long long)) != cudaSuccess ||
      cudaMemset(ws->slot_gt, 0x7f, slots * sizeof(int)) != cudaSuccess ||
      cudaMemset(ws->cand_count, 0, (size_t)max_batch * sizeof(int)) != cudaSuccess ||
      cudaMemset(ws->seg_ready, 0, ((size_t)max_batch + 2) * sizeof(int)) != cudaSuccess ||
      cudaMemset(ws->by_ints, 0, 2048 * sizeof(int)) != cudaSuccess ||
      cudaMemset(ws->ticket, 0, 64) != cudaSuccess ||
      cudaMemset(ws->cand_need, 0, 64) != cudaSuccess ||
      cudaMemset(ws->err_flag, 0, 64) != cudaSuccess ||
      cudaMallocHost((void**)&ws->host_counts, ((size_t)max_batch + 8) * sizeof(int32_t)) !=
          cudaSuccess) {
    yh::set_error("workspace init failed: %s", cudaGetErrorString(cudaGetLastError()));
    yh_workspace_destroy(ws);
    return YH_ERR_CUDA;
  }
  *out = ws;
  return YH_OK;
}

const int32_t* yh_workspace_error_flag(const YhWorkspace* ws) { return ws ? ws->err_flag : nullptr; }

int yh_copy_error_flag_async(YhWorkspace* ws, int32_t* flag_host_pinned, void* stream) {
  YH_REQUIRE(ws != nullptr && flag_host_pinned != nullptr, "NULL argument");
  YH_CUDA(cudaSetDevice(ws->device));
  YH_CUDA(cudaMemcpyAsync(flag_host_pinned, ws->err_flag, sizeof(int32_t), cudaMemcpyDeviceToHost,
                          (cudaStream_t)stream));
  return YH_OK;
}

int yh_workspace_candidates_needed(const YhWorkspace* ws) {
  return ws ? ws->cand_need_host : 0;
}

int yh_read_counts(YhWorkspace* ws, const int32_t* M_dev, int32_t* M_host, const int32_t* K_dev,
                   int32_t* K_host, void* stream) {
  YH_REQUIRE(ws != nullptr, "workspace is NULL");
  cudaStream_t s = (cudaStream_t)stream;
  YH_CUDA(cudaSetDevice(ws->device));
  int32_t* h = ws->host_counts;
  h[0] = h[1] = h[2] = h[3] = 0;
  if (M_dev) YH_CUDA(cudaMemcpyAsync(&h[0], M_dev, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  if (K_dev) YH_CUDA(cudaMemcpyAsync(&h[1], K_dev, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  YH_CUDA(cudaMemcpyAsync(&h[2], ws->err_flag, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  YH_CUDA(cudaMemsetAsync(ws->err_flag, 0, sizeof(int32_t), s));
  YH_CUDA(cudaMemcpyAsync(&h[3], ws->cand_need, sizeof(int32_t), cudaMemcpyDeviceToHost, s));
  YH_CUDA(cudaMemsetAsync(ws->cand_need, 0, sizeof(int32_t), s));
  YH_CUDA(cudaStreamSynchronize(s));
  ws->cand_need_host = h[3];
  if (M_host) *M_host = h[0];
  if (K_host) *K_host = h[1];
  if (h[2] & 8) {
    yh::set_error("the match list given to the loss is not sorted by (batch, flat) with unique "
                  "entries (yh_assign emits it that way; flag=%d)", h[2]);
    return YH_ERR_INVALID;
  }
  if (h[2] != 0) {
    yh::set_error("a device buffer overflowed (flag=%d): 1=matches, 2=nms candidates (largest image "
                  "had %d, cand_per_image=%d: recreate the workspace with "
                  "yh_workspace_candidates_needed), 4=nms outputs",
                  h[2], h[3], ws->cand_per_image);
    return YH_ERR_CAPACITY;
  }
  return YH_OK;
}

}  // extern "C"
