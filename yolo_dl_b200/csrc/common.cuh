// common.cuh — shared definitions of the yolohot kernels (sm_100a).
//
// Everything here is device-side plumbing: the layer table handed to kernels by value,
// the workspace, the one decode routine every kernel shares (so that boxes decoded by
// yh_decode, by the loss and by the NMS candidate filter are bit-identical), and
// warp/block reduction helpers.  The library is compiled with -fmad=false: the
// reference's arithmetic is a chain of separate ATen ops (no fused multiply-add), and
// index decisions (NMS keeps, matcher winners) must not depend on contraction.
#pragma once

#include <cuda_runtime.h>
#include <stdint.h>

// compute-sanitizer is closed on this pool (operators' decision, B200_PROFILING.md), so the
// bounds discipline of the kernels with hand-rolled indexing is checked by the kernels themselves:
// a build with -DYH_BOUNDS (make bounds -> libyolohot_bounds.so) turns every YH_CHECK into a device
// assert (message + trap -> cudaErrorAssert at the next call), and the whole GPU suite is run once
// against that library (YH_LIB=..., profiles/r2_bounds_pytest.log).  No cost in the shipped build.
#ifdef YH_BOUNDS
#include <assert.h>
#define YH_CHECK(cond) assert(cond)
#else
#define YH_CHECK(cond) ((void)0)
#endif

#include "yolohot.h"

namespace yh {

// ---------------------------------------------------------------------------------
// errors / launch accounting
// ---------------------------------------------------------------------------------
void set_error(const char* fmt, ...);
int64_t& launch_counter();

#define YH_CUDA(expr)                                                                   \
  do {                                                                                  \
    cudaError_t _e = (expr);                                                            \
    if (_e != cudaSuccess) {                                                            \
      yh::set_error("%s failed at %s:%d: %s", #expr, __FILE__, __LINE__,                \
                    cudaGetErrorString(_e));                                            \
      return YH_ERR_CUDA;                                                               \
    }                                                                                   \
  } while (0)

#define YH_REQUIRE(cond, ...)                                                           \
  do {                                                                                  \
    if (!(cond)) {                                                                      \
      yh::set_error(__VA_ARGS__);                                                       \
      return YH_ERR_INVALID;                                                            \
    }                                                                                   \
  } while (0)

// kernel ids of the optional per-launch CUDA-event timing (yh_timing_enable / yh_timing_read)
enum KernelId {
  YH_K_DECODE = 0, YH_K_ASSIGN_SLOTS, YH_K_ASSIGN_EMIT, YH_K_LOSS_DENSE, YH_K_LOSS_ROWS,
  YH_K_NMS_FILTER, YH_K_NMS_SEGMENT, YH_K_NMS_EMIT, YH_K_NMS_PACK, YH_K_PICK, YH_K_BENCHMARK,
  YH_K_MATCH_DET_GT, YH_K_AP, YH_K_LOSS_FINISH, YH_K_LOSS_PUBLISH, YH_K_NMS_BIG, YH_K_COUNT
};
void timing_begin(int id, cudaStream_t s);  // no-ops unless timing is enabled on this thread
void timing_end(cudaStream_t s);

// before every <<<>>>
#define YH_TIME(id, s) yh::timing_begin(yh::id, s)
// after every <<<>>>: count it and surface launch-configuration errors
#define YH_LAUNCHED(s)                                                                  \
  do {                                                                                  \
    yh::timing_end(s);                                                                  \
    yh::launch_counter()++;                                                             \
    YH_CUDA(cudaGetLastError());                                                        \
  } while (0)

// ---------------------------------------------------------------------------------
// programmatic dependent launch (PDL) along the two dependent chains of the step
//   assign_slots -> assign_emit -> loss_rows -> loss_finish     nms_filter -> nms_segment -> nms_emit
// Every kernel of a chain calls griddep_launch() first thing (the NEXT kernel of the stream may then
// be scheduled as soon as all CTAs of this one have started) and griddep_wait() before it touches
// global memory (blocks until the PREVIOUS kernel of the stream has completed and flushed).  With
// the launch attribute of launch_pdl the ~2-4 us launch + drain gap between dependent kernels — most
// of the step at 8 images per GPU — overlaps the predecessor's execution; without the attribute both
// instructions are no-ops and the stream serialises as usual.  Capturable into CUDA graphs.
// ---------------------------------------------------------------------------------
__device__ __forceinline__ void griddep_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }
__device__ __forceinline__ void griddep_launch() { asm volatile("griddepcontrol.launch_dependents;" ::: "memory"); }

bool pdl_enabled();   // YH_PDL=0/1 forces; default: on for calls on batches of <= 16 images
void note_batch(int B);  // the entry points tell the batch size of the call on this thread

template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl_if(bool allow, void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                                 cudaStream_t s, Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = grid;
  cfg.blockDim = block;
  cfg.dynamicSmemBytes = smem;
  cfg.stream = s;
  cudaLaunchAttribute at[1];
  at[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
  at[0].val.programmaticStreamSerializationAllowed = 1;
  cfg.attrs = at;
  cfg.numAttrs = (allow && pdl_enabled()) ? 1 : 0;
  return cudaLaunchKernelEx(&cfg, kern, KArgs(args)...);
}

// Only for a kernel whose stream predecessor is a kernel of THIS library launched by the same
// entry point (assign_slots -> assign_emit, dense -> rows -> finish, filter -> segment -> emit).
template <typename... KArgs, typename... Args>
inline cudaError_t launch_pdl(void (*kern)(KArgs...), dim3 grid, dim3 block, size_t smem,
                              cudaStream_t s, Args&&... args) {
  return launch_pdl_if(true, kern, grid, block, smem, s, static_cast<Args&&>(args)...);
}

// ---------------------------------------------------------------------------------
// layer table (kernel parameter, by value)
// ---------------------------------------------------------------------------------
struct LayerTab {
  int L;
  int N;                               // sum_l A*H*W
  int C;                               // classes
  int E;                               // 5 + C
  int H[YH_MAX_LAYERS], W[YH_MAX_LAYERS], A[YH_MAX_LAYERS], HW[YH_MAX_LAYERS];
  int start[YH_MAX_LAYERS];            // flat_start
  float ah[YH_MAX_LAYERS][YH_MAX_ANCHORS];  // anchors cast to f32 (detect_2d.rs:203-209)
  float aw[YH_MAX_LAYERS][YH_MAX_ANCHORS];
  const float* head[YH_MAX_LAYERS];    // [B, A*E, H, W]
  int l2hint;                          // YH_L2_HINT: prefetch-size hint of the scattered logit loads
};

// Validates the layers and fills tab (host). heads may be NULL (then tab.head is NULL).
int build_layer_tab(const YhLayer* layers, int L, int C, const float* const* heads,
                    LayerTab* tab);

// Grows ws->row_grad to `floats` (no-op when large enough).  Only yh_workspace_reserve_rows calls
// it: it synchronises the device and reallocates, which no compute entry point may do.
int reserve_rows(YhWorkspace* ws, size_t floats);

// Position of a flat instance index inside its layer.
struct Cell {
  int layer, anchor, row, col, cell;  // cell = row*W + col
};

__device__ __forceinline__ Cell locate(const LayerTab& t, int flat) {
  Cell c;
  int l = 0;
#pragma unroll
  for (int i = 1; i < YH_MAX_LAYERS; ++i)
    if (i < t.L && flat >= t.start[i]) l = i;
  int rem = flat - t.start[l];
  c.layer = l;
  c.anchor = rem / t.HW[l];
  c.cell = rem - c.anchor * t.HW[l];
  c.row = c.cell / t.W[l];
  c.col = c.cell - c.row * t.W[l];
  return c;
}

// Offset (in floats) of entry 0 of (b, anchor, cell) inside head tensor of layer l;
// entry e lives at + e * A * HW  (channel = e*A + anchor, detect_2d.rs:66).
__device__ __forceinline__ size_t head_offset(const LayerTab& t, int l, int b, int anchor,
                                              int cell) {
  return ((size_t)b * (size_t)(t.A[l] * t.E) + (size_t)anchor) * (size_t)t.HW[l] + (size_t)cell;
}

// ---------------------------------------------------------------------------------
// decode (detect_2d.rs:69-72,122-133), f32, op order of the ATen chain
// ---------------------------------------------------------------------------------
__device__ __forceinline__ float sigmoid_f(float x) { return 1.0f / (1.0f + expf(-x)); }

__device__ __forceinline__ float decode_center(float sig, int idx, float size) {
  // (sigmoid*2 - 0.5)/size + arange(size)[idx]/size
  return (sig * 2.0f - 0.5f) / size + (float)idx / size;
}

__device__ __forceinline__ float decode_size(float sig, float anchor) {
  float t = sig * 2.0f;
  return (t * t) * anchor;
}

// One isolated 4-byte logit (class logits of an NMS candidate / a matched row: every element in
// its own 128-byte line, 69 KB apart).  r1 measured 3.6 sectors of DRAM traffic per such request:
// the L2 fills whole lines.  hint 1 asks for 64-byte fills (PTX .L2::64B prefetch size).
__device__ __forceinline__ float ld_scatter(const float* p, const int hint) {
  float v;
  if (hint == 1) {
    asm volatile("ld.global.nc.L2::64B.f32 %0, [%1];" : "=f"(v) : "l"(p));
  } else {
    v = __ldg(p);
  }
  return v;
}

// ---------------------------------------------------------------------------------
// reductions
// ---------------------------------------------------------------------------------
__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

__device__ __forceinline__ double warp_sum_d(double v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
  return v;
}

// streaming stores / loads: the head gradient is written once and never re-read by us
__device__ __forceinline__ void st_stream4(float4* p, float4 v) { __stcs(p, v); }
__device__ __forceinline__ void st_stream1(float* p, float v) { __stcs(p, v); }
// a store whose line should OUTLIVE the stream around it in the L2 (evict_last priority)
__device__ __forceinline__ unsigned long long l2_policy_evict_last() {
  unsigned long long pol;
  asm volatile("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(pol));
  return pol;
}
__device__ __forceinline__ void st_keep4(float4* p, float4 v, unsigned long long pol) {
  asm volatile("st.global.L2::cache_hint.v4.f32 [%0], {%1, %2, %3, %4}, %5;" ::"l"(p), "f"(v.x),
               "f"(v.y), "f"(v.z), "f"(v.w), "l"(pol)
               : "memory");
}

}  // namespace yh

// ---------------------------------------------------------------------------------
// workspace
// ---------------------------------------------------------------------------------
struct YhWorkspace {
  int device;
  int max_batch;
  int64_t max_instances;
  int cand_per_image;
  int num_sms;

  // assignment (assign.cu): per (image, flat) slot; EMPTY (~0 / 0x7f7f7f7f) between calls
  unsigned long long* slot_dist;  // [max_batch * max_instances] f64 bits of best dist^2
  int* slot_gt;                   // [max_batch * max_instances] winning GT (local index)
  unsigned* slot_bits;            // [max_batch * ceil(max_instances/32)] occupancy bitmap
  int* img_count;                 // [max_batch] matches per image
  int* img_off;                   // [max_batch + 1] first match row of every image (assign_emit)
  unsigned long long* pick_slot;  // [max_batch * max_instances] scratch of yh_inference_pick

  // loss (loss.cu)
  double* partials;               // [YH_PARTIALS_MAX * 4]
  unsigned int* ticket;           // last-block-done counter
  float* row_grad;                // [row_grad_floats] staged gradients of the matched rows [M, 5+C]
  size_t row_grad_floats;

  // nms (nms.cu): per-image regions of cand_stride (= next pow2 of cand_per_image) entries
  long long cand_stride;
  unsigned long long* cand_key;   // [max_batch * cand_stride] (~score bits << 32 | index)
  int* cand_count;                // [max_batch]
  unsigned long long* kept_key;   // [max_batch * cand_stride]
  float4* kept_box;               // [max_batch * cand_stride] (t,l,b,r)
  long long* kept_grp;            // [max_batch * cand_stride] suppression group of a kept box
  int* kept_count;                // [max_batch]
  int* seg_ready;                 // [max_batch + 2] nms_segment_kernel's fused emit: kept count + 1 of
                                  // every finished image (0 between calls) | finished-CTA ticket

  int* by_ints;                   // yh_nms_by_scores buckets: count | off | cursor | kept | total

  // nms_big_kernel (segments above YH_BIG_MIN candidates); allocated when cand_per_image > YH_BIG_MIN
  float4* big_box;                // [max_batch * cand_stride] sorted candidates' boxes, laid out like cand_key
  long long* big_grp;             // [max_batch * cand_stride]
  unsigned* big_sync;             // barrier counters | cross-test masks | kept counts (YH_BIG_SYNC_WORDS)
  unsigned* big_mat;              // [YH_BIG_GROUPS][2][2][512 * 16] suppression bit matrices of the tiles in flight
  int* cand_need;                 // device: largest candidate count that exceeded the capacity
  int cand_need_host;             // its value at the last yh_read_counts

  int* err_flag;                  // device: sticky overflow flag
  int32_t* host_counts;           // pinned, [max_batch + 8]
};

#define YH_PARTIALS_MAX 8192
#define YH_BIG_MIN 1024          // candidates per segment above which the multi-CTA NMS kernel takes it
                                 // (one group of 1000 rows: 0.59 ms on one CTA; 5000 rows on all SMs: 0.60 ms)
#define YH_BIG_GROUPS 64         // CTA groups nms_big_kernel may split into
#define YH_BIG_SYNC_WORDS (YH_BIG_GROUPS * (32 + 32 + 2))
