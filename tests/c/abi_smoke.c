/* abi_smoke.c — calls the C ABI of include/yolohot.h from plain C, no Python, no torch:
 * what a foreign-language binding (the Rust `extern "C"` block of INTEGRATION.md) does.
 *
 *   decode -> assign -> loss fwd+bwd -> NMS -> count read, on a small two-layer head whose
 *   values are generated here; results are checked against scalar C restatements of the
 *   reference formulas (detect_2d.rs:69-72,122-133; bce_with_logits_loss.rs:60-85).
 *
 * Build: make -C yolo_dl_b200/csrc abi_smoke     (gcc -std=c99, links libyolohot.so + cudart)
 * Exit code 0 = all checks passed; 2 = no CUDA device (the library has no CPU fallback and says
 * so); 1 = a check failed.
 */
#include <math.h>
#include <stdint.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>

#include <cuda_runtime_api.h>

#include "yolohot.h"

#define CHECK_YH(call)                                                              \
  do {                                                                              \
    int rc_ = (call);                                                               \
    if (rc_ != YH_OK) {                                                             \
      fprintf(stderr, "%s -> %d: %s\n", #call, rc_, yh_last_error());               \
      return rc_ == YH_ERR_CUDA && !have_gpu ? 2 : 1;                               \
    }                                                                               \
  } while (0)
#define CHECK_CU(call)                                                              \
  do {                                                                              \
    cudaError_t e_ = (call);                                                        \
    if (e_ != cudaSuccess) {                                                        \
      fprintf(stderr, "%s: %s\n", #call, cudaGetErrorString(e_));                   \
      return 1;                                                                     \
    }                                                                               \
  } while (0)
#define EXPECT(cond, ...)                                                           \
  do {                                                                              \
    if (!(cond)) {                                                                  \
      fprintf(stderr, "FAILED %s:%d: ", __FILE__, __LINE__);                        \
      fprintf(stderr, __VA_ARGS__);                                                 \
      fprintf(stderr, "\n");                                                        \
      return 1;                                                                     \
    }                                                                               \
  } while (0)

enum { B = 4, C = 3, E = 5 + C, L = 2, A = 2 };
static const int HH[L] = {8, 4}, WW[L] = {12, 6};

static uint32_t rng_state = 20261104u;
static float frand(void) { /* xorshift, uniform in (-1, 1) */
  rng_state ^= rng_state << 13;
  rng_state ^= rng_state >> 17;
  rng_state ^= rng_state << 5;
  return (float)(rng_state >> 8) / 8388608.0f - 1.0f;
}

static float sigmoidf_(float x) { return 1.0f / (1.0f + expf(-x)); }

int main(void) {
  int have_gpu = 0, ndev = 0;
  if (cudaGetDeviceCount(&ndev) == cudaSuccess && ndev > 0) have_gpu = 1;
  printf("%s, %d CUDA device(s)\n", yh_version(), ndev);

  YhLayer layers[L];
  size_t head_floats[L];
  int64_t N = 0;
  memset(layers, 0, sizeof(layers));
  for (int l = 0; l < L; ++l) {
    layers[l].H = HH[l];
    layers[l].W = WW[l];
    layers[l].A = A;
    layers[l].flat_start = N;
    for (int a = 0; a < A; ++a) {
      layers[l].anchor_h[a] = 0.1 * (l + 1) * (a + 1);
      layers[l].anchor_w[a] = 0.15 * (l + 1) * (a + 1);
    }
    N += (int64_t)A * HH[l] * WW[l];
    head_floats[l] = (size_t)B * A * E * HH[l] * WW[l];
  }

  YhWorkspace* ws = NULL;
  CHECK_YH(yh_workspace_create(0, B, N, 1024, &ws)); /* fails loudly without a GPU */

  /* heads: box logits ~U(-1,1), objectness ~U(-3,-1) except a few hot cells, classes ~U(-2,0) */
  float* h_head[L];
  float* d_head[L];
  float* d_grad[L];
  for (int l = 0; l < L; ++l) {
    const size_t hw = (size_t)HH[l] * WW[l];
    h_head[l] = (float*)malloc(head_floats[l] * sizeof(float));
    for (int b = 0; b < B; ++b)
      for (int e = 0; e < E; ++e)
        for (int a = 0; a < A; ++a)
          for (size_t c = 0; c < hw; ++c) {
            float v = frand();
            if (e == 4) v = v - 2.0f + ((c % 17 == 3) ? 4.0f : 0.0f);
            if (e >= 5) v = v - 1.0f + ((c % 17 == 3 && e == 5 + (int)(c % C)) ? 4.0f : 0.0f);
            h_head[l][(((size_t)b * E + e) * A + a) * hw + c] = v;
          }
    CHECK_CU(cudaMalloc((void**)&d_head[l], head_floats[l] * sizeof(float)));
    CHECK_CU(cudaMalloc((void**)&d_grad[l], head_floats[l] * sizeof(float)));
    CHECK_CU(cudaMemcpy(d_head[l], h_head[l], head_floats[l] * sizeof(float), cudaMemcpyHostToDevice));
  }

  /* ---- decode (Detect2D x L + merge) ---- */
  float* d_box;
  CHECK_CU(cudaMalloc((void**)&d_box, 4 * (size_t)B * N * sizeof(float)));
  const float* heads_c[L] = {d_head[0], d_head[1]};
  CHECK_YH(yh_decode_merged(heads_c, layers, L, B, C, d_box, d_box + (size_t)B * N,
                            d_box + 2 * (size_t)B * N, d_box + 3 * (size_t)B * N, N, NULL));
  float* h_box = (float*)malloc(4 * (size_t)B * N * sizeof(float));
  CHECK_CU(cudaMemcpy(h_box, d_box, 4 * (size_t)B * N * sizeof(float), cudaMemcpyDeviceToHost));
  for (int l = 0; l < L; ++l) {
    const int H = HH[l], W = WW[l];
    const size_t hw = (size_t)H * W;
    for (int b = 0; b < B; b += B - 1)
      for (int a = 0; a < A; ++a)
        for (int r = 0; r < H; r += 3)
          for (int c = 0; c < W; c += 5) {
            const size_t cell = (size_t)r * W + c;
            const float* x = h_head[l] + ((size_t)b * E * A + a) * hw + cell;
            const size_t es = (size_t)A * hw;
            const float cy = (sigmoidf_(x[0]) * 2.0f - 0.5f) / (float)H + (float)r / (float)H;
            const float t2 = sigmoidf_(x[3 * es]) * 2.0f;
            const float w = t2 * t2 * (float)layers[l].anchor_w[a];
            const size_t flat = (size_t)layers[l].flat_start + c + (size_t)W * (r + (size_t)H * a);
            const float gcy = h_box[(size_t)b * N + flat];
            const float gw = h_box[3 * (size_t)B * N + (size_t)b * N + flat];
            EXPECT(fabsf(gcy - cy) <= 1e-5f * fabsf(cy) + 1e-7f, "cy %g vs %g", gcy, cy);
            EXPECT(fabsf(gw - w) <= 1e-5f * fabsf(w) + 1e-7f, "w %g vs %g", gw, w);
          }
  }

  /* ---- assign: two labels per image ---- */
  double h_gt[B * 2][5];
  int32_t h_off[B + 1];
  for (int b = 0; b <= B; ++b) h_off[b] = 2 * b;
  for (int i = 0; i < B * 2; ++i) {
    h_gt[i][0] = 0.2 + 0.07 * i;  /* cy */
    h_gt[i][1] = 0.8 - 0.06 * i;  /* cx */
    h_gt[i][2] = 0.15 + 0.02 * i; /* h */
    h_gt[i][3] = 0.2 + 0.01 * i;  /* w */
    h_gt[i][4] = (double)(i % C);
  }
  double* d_gt;
  int32_t* d_off;
  CHECK_CU(cudaMalloc((void**)&d_gt, sizeof(h_gt)));
  CHECK_CU(cudaMalloc((void**)&d_off, sizeof(h_off)));
  CHECK_CU(cudaMemcpy(d_gt, h_gt, sizeof(h_gt), cudaMemcpyHostToDevice));
  CHECK_CU(cudaMemcpy(d_off, h_off, sizeof(h_off), cudaMemcpyHostToDevice));
  const int cap = B * 2 * L * A * 5;
  int64_t *m_batch, *m_flat, *m_class;
  int32_t *m_gt, *d_M, *d_K;
  float* m_target;
  CHECK_CU(cudaMalloc((void**)&m_batch, cap * sizeof(int64_t)));
  CHECK_CU(cudaMalloc((void**)&m_flat, cap * sizeof(int64_t)));
  CHECK_CU(cudaMalloc((void**)&m_class, cap * sizeof(int64_t)));
  CHECK_CU(cudaMalloc((void**)&m_gt, cap * sizeof(int32_t)));
  CHECK_CU(cudaMalloc((void**)&m_target, cap * 4 * sizeof(float)));
  CHECK_CU(cudaMalloc((void**)&d_M, sizeof(int32_t)));
  CHECK_CU(cudaMalloc((void**)&d_K, sizeof(int32_t)));
  CHECK_YH(yh_assign(ws, d_gt, d_off, B, layers, L, 1, 4.0, m_batch, m_flat, m_gt, m_target, m_class,
                     d_M, cap, NULL));

  /* ---- loss forward + backward ---- */
  YhLossCfg cfg;
  memset(&cfg, 0, sizeof(cfg));
  cfg.box_metric = YH_METRIC_CIOU;
  cfg.reduction = YH_REDUCTION_MEAN;
  cfg.iou_loss_weight = 0.05;
  cfg.classification_loss_weight = 0.58;
  cfg.objectness_loss_weight = 1.0;
  cfg.smooth_classification_coef = 0.01;
  float *d_losses, *d_iou, *d_rows;
  CHECK_CU(cudaMalloc((void**)&d_losses, 4 * sizeof(float)));
  CHECK_CU(cudaMalloc((void**)&d_iou, cap * sizeof(float)));
  CHECK_CU(cudaMalloc((void**)&d_rows, (size_t)cap * E * sizeof(float)));
  float* grads_c[L] = {d_grad[0], d_grad[1]};
  /* the loss entry points never allocate: without the reserve they must refuse */
  int rc = yh_loss_fwd_bwd(ws, heads_c, layers, L, B, C, m_batch, m_flat, m_target, m_class, d_M, cap,
                           &cfg, d_losses, d_iou, d_rows, grads_c, NULL);
  EXPECT(rc == YH_ERR_CAPACITY, "expected YH_ERR_CAPACITY before yh_workspace_reserve_rows, got %d", rc);
  CHECK_YH(yh_workspace_reserve_rows(ws, cap, C));
  CHECK_YH(yh_loss_fwd_bwd(ws, heads_c, layers, L, B, C, m_batch, m_flat, m_target, m_class, d_M, cap,
                           &cfg, d_losses, d_iou, d_rows, grads_c, NULL));

  /* ---- NMS ---- */
  const int kcap = B * 1024;
  int64_t *kb, *kc, *ki;
  float *tlbr, *conf;
  CHECK_CU(cudaMalloc((void**)&kb, kcap * sizeof(int64_t)));
  CHECK_CU(cudaMalloc((void**)&kc, kcap * sizeof(int64_t)));
  CHECK_CU(cudaMalloc((void**)&ki, kcap * sizeof(int64_t)));
  CHECK_CU(cudaMalloc((void**)&tlbr, (size_t)kcap * 4 * sizeof(float)));
  CHECK_CU(cudaMalloc((void**)&conf, kcap * sizeof(float)));
  CHECK_YH(yh_nms(ws, heads_c, layers, L, B, C, 0.3, 0.5, 0, kb, kc, ki, tlbr, conf, d_K, kcap, NULL));

  int32_t M = -1, K = -1;
  CHECK_YH(yh_read_counts(ws, d_M, &M, d_K, &K, NULL)); /* the one synchronising call */
  EXPECT(M > 0 && M <= cap, "M=%d", M);
  EXPECT(K > 0 && K <= kcap, "K=%d", K);

  float losses[4];
  CHECK_CU(cudaMemcpy(losses, d_losses, sizeof(losses), cudaMemcpyDeviceToHost));
  for (int i = 0; i < 4; ++i) EXPECT(isfinite(losses[i]) && losses[i] > 0.0f, "loss[%d]=%g", i, losses[i]);
  const float total = 0.05f * losses[1] + 0.58f * losses[2] + 1.0f * losses[3];
  EXPECT(fabsf(total - losses[0]) <= 1e-6f * fabsf(total), "total %g vs %g", losses[0], total);

  /* objectness gradient of an unmatched cell: sigmoid(x) / (B*N) (target 0, weight 1) */
  int64_t* h_mb = (int64_t*)malloc(M * sizeof(int64_t));
  int64_t* h_mf = (int64_t*)malloc(M * sizeof(int64_t));
  CHECK_CU(cudaMemcpy(h_mb, m_batch, M * sizeof(int64_t), cudaMemcpyDeviceToHost));
  CHECK_CU(cudaMemcpy(h_mf, m_flat, M * sizeof(int64_t), cudaMemcpyDeviceToHost));
  for (int i = 1; i < M; ++i)
    EXPECT(h_mb[i] > h_mb[i - 1] || (h_mb[i] == h_mb[i - 1] && h_mf[i] > h_mf[i - 1]),
           "match list not sorted at %d", i);
  float* h_grad0 = (float*)malloc(head_floats[0] * sizeof(float));
  CHECK_CU(cudaMemcpy(h_grad0, d_grad[0], head_floats[0] * sizeof(float), cudaMemcpyDeviceToHost));
  {
    const size_t hw = (size_t)HH[0] * WW[0];
    int checked = 0;
    for (size_t cell = 0; cell < hw && checked < 8; ++cell) {
      int matched = 0;
      for (int i = 0; i < M; ++i)
        if (h_mb[i] == 0 && h_mf[i] == (int64_t)cell) matched = 1; /* image 0, anchor 0, layer 0 */
      if (matched) continue;
      const size_t idx = ((size_t)4 * A + 0) * hw + cell; /* image 0, entry 4, anchor 0 */
      const float want = sigmoidf_(h_head[0][idx]) / (float)(B * N);
      EXPECT(fabsf(h_grad0[idx] - want) <= 1e-5f * want + 1e-12f, "obj grad %g vs %g", h_grad0[idx], want);
      EXPECT(h_grad0[cell] == 0.0f, "box gradient of an unmatched cell is %g", h_grad0[cell]);
      ++checked;
    }
    EXPECT(checked == 8, "checked %d cells", checked);
  }

  /* kept detections: ordered by (batch, confidence descending), confidences above the threshold */
  int64_t* h_kb = (int64_t*)malloc(K * sizeof(int64_t));
  float* h_conf = (float*)malloc(K * sizeof(float));
  CHECK_CU(cudaMemcpy(h_kb, kb, K * sizeof(int64_t), cudaMemcpyDeviceToHost));
  CHECK_CU(cudaMemcpy(h_conf, conf, K * sizeof(float), cudaMemcpyDeviceToHost));
  for (int i = 0; i < K; ++i) {
    EXPECT(h_conf[i] >= 0.3f && h_kb[i] >= 0 && h_kb[i] < B, "row %d: conf %g batch %lld", i, h_conf[i],
           (long long)h_kb[i]);
    if (i) EXPECT(h_kb[i] > h_kb[i - 1] || h_conf[i] <= h_conf[i - 1], "order at row %d", i);
  }

  /* argument validation speaks the reference's messages */
  rc = yh_assign(ws, d_gt, d_off, B, layers, L, 1, 0.5, m_batch, m_flat, m_gt, m_target, m_class, d_M,
                 cap, NULL);
  EXPECT(rc == YH_ERR_INVALID && strstr(yh_last_error(), "anchor_scale_thresh") != NULL, "rc=%d %s", rc,
         yh_last_error());

  printf("abi_smoke OK: M=%d K=%d losses total=%.6f iou=%.6f cls=%.6f obj=%.6f launches=%lld\n", M, K,
         losses[0], losses[1], losses[2], losses[3], (long long)yh_launch_count(0));
  yh_workspace_destroy(ws);
  return 0;
}
