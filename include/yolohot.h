/*
 * yolohot.h — C ABI of the B200-native (sm_100a) detection hot path for yolo-dl.
 *
 * This is the drop-in boundary: every entry point below replaces the body of one
 * Rust method of the reference (jerry73204/yolo-dl).  The Rust side keeps its
 * signatures and only marshals `Tensor::data_ptr()`s and shapes; see
 * INTEGRATION.md for the `extern "C"` block a maintainer would add.
 *
 * Conventions:
 *   - every `*_dev` / head / output pointer is caller-owned DEVICE memory on the
 *     device the workspace was created for; the library never frees caller memory;
 *   - `stream` is a `cudaStream_t` passed as `void*` (NULL = legacy default stream,
 *     which is what tch 0.7 runs on);
 *   - no entry point synchronises the device unless its comment says so; counts
 *     (`M`, `K`) are produced in device memory, `yh_read_counts` is the one
 *     synchronising read (the reference already syncs there: `nonzero`, host matcher);
 *   - return value: 0 (YH_OK) or a YhStatus; `yh_last_error()` gives the text
 *     (thread-local), the Rust wrapper maps it to `anyhow::Error` / panic as today;
 *   - there is NO CPU fallback: without a CUDA device every compute call fails.
 *
 * Layout of one head tensor (input of Detect2D, `tch-modules/src/detect_2d.rs:66`):
 *   contiguous f32 `[B, A*(5+C), H, W]` VIEWED as `[B, 5+C, A, H, W]`, i.e.
 *   channel = entry*A + anchor, entries 0=cy 1=cx 2=h 3=w 4=obj 5..=class.
 * Flat instance index inside an image (`merged_dense_detection.rs:374-376`):
 *   flat = layer.flat_start + col + W*(row + H*anchor).
 */
#ifndef YOLOHOT_H
#define YOLOHOT_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define YH_MAX_LAYERS 8
#define YH_MAX_ANCHORS 8

typedef enum YhStatus {
  YH_OK = 0,
  YH_ERR_INVALID = 1,     /* bad argument (shape, null pointer, unsupported option) */
  YH_ERR_CUDA = 2,        /* a CUDA runtime call failed (text in yh_last_error) */
  YH_ERR_CAPACITY = 3,    /* an output or workspace buffer was too small */
  YH_ERR_UNSUPPORTED = 4  /* valid in the reference but outside this path */
} YhStatus;

/* `yolo-dl/src/loss/misc.rs:14-21` BoxMetric */
typedef enum YhBoxMetric {
  YH_METRIC_IOU = 0,
  YH_METRIC_GIOU = 1,
  YH_METRIC_DIOU = 2,
  YH_METRIC_CIOU = 3,
  YH_METRIC_HAUSDORFF = 4
} YhBoxMetric;

/* tch `Reduction` as used by `loss_.rs:301-319` (None returns per-row tensors and is
 * not part of the scalar-loss ABI). */
typedef enum YhReduction { YH_REDUCTION_MEAN = 0, YH_REDUCTION_SUM = 1 } YhReduction;

/* `ObjectnessLossKind` / `ClassificationLossKind` (`loss_.rs:191-205`): BceWithLogitsLoss
 * (`bce_with_logits_loss.rs:60-85`), FocalLoss wrapping it (`focal_loss.rs:66-108`), L2Loss = mse on
 * the LOGITS (`l2_loss.rs:13-15`), CrossEntropyLoss with the dense target's arg-max
 * (`cross_entropy.rs:19-58`, classification only). */
typedef enum YhLossKind {
  YH_LOSS_BCE = 0,
  YH_LOSS_FOCAL = 1,
  YH_LOSS_L2 = 2,
  YH_LOSS_CROSS_ENTROPY = 3
} YhLossKind;

/* One detection layer = one `DetectionInfo` (`merged_dense_detection.rs:497-507`)
 * plus what `Detect2D` holds (`detect_2d.rs:33-39`). Anchors are RATIO units (h, w)
 * in f64 exactly as the reference stores them (`Ratio<HW<R64>>`); the decode casts
 * them to f32 (`detect_2d.rs:203-209`), the matcher uses them as f64
 * (`pred_target_matching.rs:141`). */
typedef struct YhLayer {
  int32_t H, W, A, _pad;
  int64_t flat_start; /* DetectionInfo.flat_index_range.start */
  double anchor_h[YH_MAX_ANCHORS];
  double anchor_w[YH_MAX_ANCHORS];
} YhLayer;

/* `YoloLossInit` (`yolo-dl/src/loss/loss_.rs:21-36`, defaults `:169-189`), BCE kinds only. */
typedef struct YhLossCfg {
  int32_t box_metric;   /* YhBoxMetric; default DIoU */
  int32_t reduction;    /* YhReduction; default Mean */
  double iou_loss_weight;            /* 0.05 */
  double classification_loss_weight; /* 0.58 */
  double objectness_loss_weight;     /* 1.0 */
  double smooth_classification_coef; /* 0.01 */
  double smooth_objectness_coef;     /* 0.0 */
  int32_t has_objectness_pos_weight; /* 0 */
  int32_t _pad;
  double objectness_pos_weight;
  int32_t objectness_loss_kind;      /* YhLossKind: Bce (default), Focal, L2        (loss_.rs:115-145) */
  int32_t classification_loss_kind;  /* YhLossKind: Bce (default), Focal, L2, CrossEntropy (:89-113) */
  double focal_loss_gamma;           /* 0.0 (YoloLossInit default, loss_.rs:176); alpha is FocalLossInit's
                                        fixed 0.25 (`tch-modules/src/focal_loss.rs:27-34`) */
} YhLossCfg;

typedef struct YhWorkspace YhWorkspace;

/* Sizes the per-module scratch (slot tables, candidate lists, partial sums).
 * Replaces nothing in the reference; created by the module constructors
 * (`Detect2DInit::build` detect_2d.rs:12, `YoloLossInit::build` loss_.rs:39,
 * `NonMaxSuppressionInit::build` nms.rs:28).
 *   max_batch        largest B of any call
 *   max_instances    largest N = sum_l A*H*W
 *   cand_per_image   capacity of the per-image NMS candidate list */
int yh_workspace_create(int device, int max_batch, int64_t max_instances,
                        int cand_per_image, YhWorkspace** out);
int yh_workspace_destroy(YhWorkspace* ws);
/* Reserves the staging buffer of the matched-row gradients ([capacity, 5+C] f32) that the loss
 * phases use when they produce gradients.  This is the ONLY call that grows it (it synchronises
 * the device and reallocates): the loss entry points never allocate or synchronise, they return
 * YH_ERR_CAPACITY when `capacity` rows do not fit what was reserved.  Call it from the module
 * constructor with the largest match capacity (G_max * sum_l A_l * 5), or whenever that grows. */
int yh_workspace_reserve_rows(YhWorkspace* ws, int capacity, int C);
const char* yh_last_error(void);
const char* yh_version(void);

/* ---- a1: Detect2D::forward (`tch-modules/src/detect_2d.rs:42-152`) -------------------
 * cy = (2*sigmoid(x0) - 0.5)/H + row/H ; cx likewise with W ;
 * h  = (2*sigmoid(x2))^2 * anchor_h   ; w likewise.  obj/class logits stay views of `head`.
 * Output element (b, a, row, col) is written at out[b*out_batch_stride + (a*H+row)*W + col]:
 * pass stride A*H*W for the four `[B,1,A,H,W]` tensors of DenseDetectionTensor, or the
 * merged stride N with pointers pre-offset by flat_start to write straight into
 * MergedDenseDetection's `[B,1,N]` planes (`merged_dense_detection.rs:60-91`). */
int yh_decode(const float* head, int B, int C, const YhLayer* layer,
              float* cy, float* cx, float* h, float* w, int64_t out_batch_stride,
              void* stream);

/* a1+a3 fused: all layers in ONE launch, straight into the merged
 * `[B,1,N]` planes (batch stride = out_batch_stride >= N). */
int yh_decode_merged(const float* const* heads, const YhLayer* layers, int L, int B, int C,
                     float* cy, float* cx, float* h, float* w, int64_t out_batch_stride,
                     void* stream);

/* ---- a4: CyCxHWMatcher::match_bboxes (`yolo-dl/src/loss/pred_target_matching.rs:51-276`)
 * gt_dev: [G,5] f64 rows (cy, cx, h, w, class) in ratio units, images concatenated;
 * gt_offsets_dev: [B+1] int32 prefix offsets into gt_dev.
 * Outputs (capacity rows each), sorted by (batch, flat):
 *   out_batch/out_flat  FlatIndexTensor {batches, flats}       (:238-241)
 *   out_gt              row of gt_dev that won the slot
 *   out_target          [cap,4] f32 cy,cx,h,w of that GT        (:247)
 *   out_class           i64 class                               (:248)
 *   out_M_dev           int32 count M (device); M > capacity sets YH_ERR_CAPACITY at
 *                       the next yh_read_counts.
 * Decisions are taken in f64 with the reference's operation order (no FMA). */
int yh_assign(YhWorkspace* ws, const double* gt_dev, const int32_t* gt_offsets_dev, int B,
              const YhLayer* layers, int L, int rect4, double anchor_scale_thresh,
              int64_t* out_batch, int64_t* out_flat, int32_t* out_gt, float* out_target,
              int64_t* out_class, int32_t* out_M_dev, int capacity, void* stream);

/* ---- a5..a11: YoloLoss::forward + autograd backward (`yolo-dl/src/loss/loss_.rs:227-277`)
 * Reads the L head tensors and the matches of yh_assign; writes
 *   losses4_dev   [total, iou, classification, objectness]            (:258-273)
 *   iou_score_dev [cap] box metric per matched row (nullable)          (:296-299)
 *   pred_rows_dev [cap, 5+C] gathered cy,cx,h,w,obj_logit,class_logit — the
 *                 `MatchingOutput.pred` rows (`merged_dense_detection.rs:280-307`), nullable
 * The match list must be in the order yh_assign emits it — ascending (batch, flat), no
 * duplicates (what `MatchingOutput::cat_mini_batches` of per-image lists also gives): the
 * gradient scatter merges neighbouring matched cells into one store.  Anything else is reported
 * by the next yh_read_counts as YH_ERR_INVALID.
 *   grad_heads    L device pointers, each `[B, A*(5+C), H, W]`: d total_loss / d head,
 *                 i.e. what `total_loss.backward()` delivers to the backbone
 *                 (`train/src/train/single_gpu.rs:161-168`). NULL = forward only. */
int yh_loss_fwd_bwd(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                    int B, int C, const int64_t* m_batch, const int64_t* m_flat,
                    const float* m_target, const int64_t* m_class, const int32_t* M_dev,
                    int capacity, const YhLossCfg* cfg, float* losses4_dev,
                    float* iou_score_dev, float* pred_rows_dev, float* const* grad_heads,
                    void* stream);

/* The phases of yh_loss_fwd_bwd as separate calls, so that a caller with streams can overlap
 * them with the rest of the step (YoloLoss::forward does assign -> loss strictly in sequence,
 * `loss_.rs:233-256`, although the dense objectness pass needs no matches):
 *   yh_loss_dense         the objectness BCE over all B*N cells against target 0 and, if
 *                         grad_heads, the whole head gradient (zeros + sigmoid(x)*w/(B*N)).
 *                         Needs only the heads: may run concurrently with yh_assign / yh_decode /
 *                         yh_nms / yh_loss_rows_compute on other streams.
 *   yh_loss_rows_compute  the M matched rows (box metric, class BCE, objectness correction);
 *                         with want_grad their gradients are staged in the workspace
 *                         ([M, 5+C]).  Ordered after yh_assign only.
 *   yh_loss_finish        ordered after BOTH of the above on the same workspace (same capacity
 *                         and cfg): scatters the staged row gradients over the dense gradient
 *                         (grad_heads, NULL = forward only) and reduces all partial sums, in a
 *                         fixed order, into losses4_dev.
 *   yh_loss_rows          == yh_loss_rows_compute; yh_loss_finish on one stream (after dense).
 * yh_loss_fwd_bwd == yh_loss_dense; yh_loss_rows on one stream. */
int yh_loss_dense(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L, int B,
                  int C, const YhLossCfg* cfg, float* const* grad_heads, void* stream);
int yh_loss_rows(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L, int B,
                 int C, const int64_t* m_batch, const int64_t* m_flat, const float* m_target,
                 const int64_t* m_class, const int32_t* M_dev, int capacity, const YhLossCfg* cfg,
                 float* losses4_dev, float* iou_score_dev, float* pred_rows_dev,
                 float* const* grad_heads, void* stream);
int yh_loss_rows_compute(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                         int B, int C, const int64_t* m_batch, const int64_t* m_flat,
                         const float* m_target, const int64_t* m_class, const int32_t* M_dev,
                         int capacity, const YhLossCfg* cfg, float* iou_score_dev,
                         float* pred_rows_dev, int want_grad, void* stream);
int yh_loss_finish(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C,
                   const int64_t* m_batch, const int64_t* m_flat, const int32_t* M_dev,
                   int capacity, const YhLossCfg* cfg, float* losses4_dev,
                   float* const* grad_heads, void* stream);

/* yh_loss_finish for a caller that has arranged its streams so that
 *   (1) the kernel enqueued on `stream` right before this call is yh_loss_dense of the same step, and
 *   (2) yh_assign and yh_loss_rows_compute of the step had COMPLETED (joined by an event) before this
 *       call was enqueued (yolo_dl_b200/pipeline.py run_overlapped).
 * The scatter kernel is then launched with programmatic dependent launch: its CTAs arrive while the
 * dense pass drains, load the match list and the staged gradients, and wait for the dense pass only
 * before their first store.  Same results as yh_loss_finish.  Violating (2) reads stale rows. */
int yh_loss_finish_after_dense(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C,
                               const int64_t* m_batch, const int64_t* m_flat, const int32_t* M_dev,
                               int capacity, const YhLossCfg* cfg, float* losses4_dev,
                               float* const* grad_heads, void* stream);

/* The same loss in RANGES of images, for callers that order their streams around the 126 MB L2
 * (yolo_dl_b200/pipeline.py; YoloLoss::forward itself is one sequence, loss_.rs:233-256).  The
 * batch is cut into num_ranges <= 16 contiguous image ranges [b_lo, b_hi) (b_lo a multiple of 4:
 * every head tensor must start an image range on a 16-byte boundary); for each range
 *   yh_loss_dense_range   the dense objectness pass over those images (B stays the WHOLE batch:
 *                         the losses are normalised by it), then
 *   yh_loss_rows_range    ordered after it: the matched rows of those images, their gradients
 *                         stored straight over the dense gradient while its lines are still in
 *                         the L2 (no staging buffer, no scatter pass; grad_heads NULL = forward
 *                         only).  image_row_offsets_dev = yh_assign_image_offsets(ws) of the
 *                         workspace whose yh_assign produced the match list, or NULL (the kernel
 *                         then binary-searches the batch column);
 * and, after all ranges, yh_loss_finish_ranges reduces the partial sums into losses4_dev.
 * Results equal yh_loss_fwd_bwd's: gradients bit for bit, the four scalars to f32 rounding of a
 * differently associated f64 sum. */
int yh_loss_dense_range(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                        int B, int C, int b_lo, int b_hi, int range_index, int num_ranges,
                        const YhLossCfg* cfg, float* const* grad_heads, void* stream);
int yh_loss_rows_range(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                       int B, int C, int b_lo, int b_hi, int range_index, int num_ranges,
                       const int64_t* m_batch, const int64_t* m_flat, const float* m_target,
                       const int64_t* m_class, const int32_t* M_dev, int capacity,
                       const int32_t* image_row_offsets_dev, const YhLossCfg* cfg,
                       float* iou_score_dev, float* pred_rows_dev, float* const* grad_heads,
                       void* stream);
int yh_loss_finish_ranges(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C,
                          int num_ranges, const int32_t* M_dev, int capacity, const YhLossCfg* cfg,
                          float* losses4_dev, void* stream);
/* The same loss cut by LAYER (yolo_dl_b200/pipeline.py, schedule "layered"): the gradient tensors
 * of the small layers fit the 126 MB L2, the big first layer does not, and the gradient scatter
 * of the matched rows (yh_loss_finish) costs a DRAM read-modify-write per line once the dense
 * stream has evicted it.  So
 *   yh_loss_dense_layers   the dense pass over layers [layer_lo, layer_hi) only (launch
 *                          part_index of num_parts <= 16: its partial sums go to their own slots);
 *   yh_loss_finish_layers  ordered after yh_loss_rows_compute (ONE whole-batch launch, same
 *                          capacity / cfg) and after the dense launches of those layers: scatters
 *                          the staged gradients of the matched rows that lie in layers
 *                          [layer_lo, layer_hi); the launch with reduce != 0 must come after ALL
 *                          dense launches and also writes losses4_dev.
 * A caller runs finish(layer l) right behind dense(layer l) while dense(l+1 ..) streams on.
 * Results are yh_loss_fwd_bwd's: gradients bit for bit, the scalars to f32 rounding. */
int yh_loss_dense_layers(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L,
                         int B, int C, int layer_lo, int layer_hi, int part_index, int num_parts,
                         const YhLossCfg* cfg, float* const* grad_heads, void* stream);
int yh_loss_finish_layers(YhWorkspace* ws, const YhLayer* layers, int L, int B, int C, int layer_lo,
                          int layer_hi, int reduce, int num_dense_parts, const int64_t* m_batch,
                          const int64_t* m_flat, const int32_t* M_dev, int capacity,
                          const YhLossCfg* cfg, float* losses4_dev, float* const* grad_heads,
                          void* stream);
/* Device pointer to the [B+1] int32 first-row-of-image offsets of the match list the last yh_assign
 * on this workspace produced (valid once that call has run on the stream). */
const int32_t* yh_assign_image_offsets(const YhWorkspace* ws);

/* ---- a12+a13: NonMaxSuppression::forward (`yolo-dl/src/loss/nms.rs:72-170`) incl. the
 * grouped greedy NMS of tch_nms::nms_by_scores (call site nms.rs:142-148).
 * conf = sigmoid(obj)*sigmoid(cls); candidates: sigmoid(obj) >= thr && conf >= thr (f32
 * compare, nms.rs:97-99); group = batch (by_class=0) or batch*C+class (nms.rs:129-133).
 * Greedy rule (oracle/nms.py, the stand-in for the un-vendored tch-nms): candidates in
 * score-descending order, equal scores in nonzero() order (class*N+instance ascending); a
 * candidate is dropped iff a kept box of its group has IoU > iou_threshold.
 * DEVIATION (SURVEY.md §8b lists a `compat_chunk65536` flag for this export; there is none): above
 * 65 536 candidates per batch the reference cuts the (batch, class, instance)-ordered rows into
 * independent 65 536-row chunks, runs tch-nms on each and then gathers with the CHUNK-LOCAL indices
 * (nms.rs:135-160: the chunk offset is never added) — the rows it returns past the first chunk are
 * not the rows NMS kept.  That result is not reproduced: yh_nms runs ONE global NMS per image (per
 * image and class with suppress_by_class) for any number of candidates, which is what the chunk
 * loop approximates and equals it exactly whenever the batch has <= 65 536 candidates.  The
 * chunk-wise variant (with the offset repaired) exists only in the oracle (oracle/nms.py,
 * `compat_chunk65536`) to document the difference.
 * Kept rows are emitted ordered by (batch, score desc, class*N+instance asc):
 *   kb/kc/ki  i64 batches/classes/instances ; tlbr [cap,4] f32 (t,l,b,r) ; conf [cap]
 *   out_K_dev int32 count (device), clamped to capacity (overflow -> yh_read_counts error).
 * Does not synchronise. */
int yh_nms(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L, int B,
           int C, double confidence_threshold, double iou_threshold, int suppress_by_class,
           int64_t* kb, int64_t* kc, int64_t* ki, float* tlbr, float* conf,
           int32_t* out_K_dev, int capacity, void* stream);

/* ---- a13 alone: drop-in for `tch_nms::nms_by_scores(ltrb, scores, groups, iou_thr)`
 * (external crate tch-nms @138b5e40, call site `yolo-dl/src/loss/nms.rs:142-148`).
 *   ltrb [K,4] f32 (l,t,r,b) 16-byte aligned ; scores [K] f32 ; groups [K] i64
 *   keep [K] i64: kept row indices, score-descending (ties: lower row first)
 *   out_K_dev int32 count (device).
 * K is limited by the workspace: next_pow2(K) <= max_batch * next_pow2(cand_per_image). */
int yh_nms_by_scores(YhWorkspace* ws, const float* ltrb, const float* scores,
                     const int64_t* groups, int K, double iou_threshold, int64_t* keep,
                     int32_t* out_K_dev, void* stream);

/* ---- f2: the counters of YoloBenchmark::forward (`yolo-dl/src/loss/benchmark.rs:34-100`).
 * One launch; counts5_dev (device int64[5]) receives
 *   [all_count = B*N, all_pos = #{sigmoid(obj) >= thr}, matched_count = M,
 *    matched_pos = #{sigmoid(obj of matched rows) >= thr},
 *    class_hits = #{any_c(conf_c >= thr) && argmax_c(class logit) == target class}]
 * from which the caller forms obj_accuracy / obj_recall / obj_precision / class_accuracy in f64
 * exactly as benchmark.rs:59-92.  pred_rows_dev / m_class / M_dev are the outputs of
 * yh_loss_fwd_bwd and yh_assign.  The threshold is compared in f32 like `Tensor::ge(f64)`. */
int yh_benchmark(YhWorkspace* ws, const float* const* heads, const YhLayer* layers, int L, int B,
                 int C, const float* pred_rows_dev, const int64_t* m_class, const int32_t* M_dev,
                 int capacity, double confidence_threshold, int64_t* counts5_dev, void* stream);

/* ---- f4: evaluation metrics, f64 like the reference's R64 host code.
 * yh_match_det_gt = match_det_gt (`yolo-dl/src/loss/pred_gt_matching.rs:109-142`) with the scalar
 * `bbox::RectFloat::iou_with(.., 1e-8)` (`bbox/src/rect.rs:113-140`), batched over S segments
 * (an image, or an image x class: the caller's grouping).  Boxes are f64 (t,l,b,r) rows;
 * det_offsets_dev / gt_offsets_dev are [S+1] int32 prefix offsets.  Per detection:
 *   out_gt_dev   row (into gts) of the ground truth of maximum IoU among ALL ground truths of its
 *                segment — the LAST of equal maxima (Iterator::max_by_key) — or -1 when that IoU
 *                is 0 within f64 epsilon (:131) or the segment has no ground truth (the
 *                reference would panic on the empty max, :128);
 *   out_iou_dev  `DetBoxStruct::iou` (:96-106): IoU with that ground truth, 0 for none. */
int yh_match_det_gt(const double* dets_tlbr_dev, const int32_t* det_offsets_dev, int D,
                    const double* gts_tlbr_dev, const int32_t* gt_offsets_dev, int S,
                    int32_t* out_gt_dev, double* out_iou_dev, void* stream);

/* yh_average_precision = ApCalculator::compute_by_detections for every (problem, IoU threshold)
 * (`yolo-dl/src/loss/average_precision.rs:157-219`, envelope + n-point interpolation :87-155,
 * `utils/algorithm.rs:38-91`).  A problem is a contiguous range of detections
 * [problem_offsets[p], problem_offsets[p+1]) (typically one class) with num_ground_truth[p];
 * det_gt / det_iou / det_conf are yh_match_det_gt's outputs and the detections' confidences.
 * out_ap_dev is [num_problems, num_thresholds] f64; MeanApCalculator::compute_mean_ap (:252-275)
 * is the mean of a row.  Detections of equal confidence are ordered by (ground truth ascending,
 * none first; IoU descending; index) — the reference iterates a HashMap there (unspecified).
 * n_points in 2..1024 (COCO: 101).  An empty problem yields 0 (the reference panics, :90).
 * Any number of detections per problem (the reference's Vec code has no bound, :157-219): up to
 * 8192 the per-detection arrays live in shared memory; above, in `scratch_dev` (device memory,
 * 16-byte aligned, yh_average_precision_scratch_bytes(num_problems, max_detections_per_problem)
 * bytes, which is 0 when no scratch is needed; too little scratch -> YH_ERR_CAPACITY). */
size_t yh_average_precision_scratch_bytes(int num_problems, int max_detections_per_problem);
int yh_average_precision(const int32_t* det_gt_dev, const double* det_iou_dev,
                         const double* det_conf_dev, const int32_t* problem_offsets_dev,
                         const int32_t* num_ground_truth_dev, int num_problems,
                         int max_detections_per_problem, const double* iou_thresholds_dev,
                         int num_thresholds, int n_points, double* out_ap_dev, void* scratch_dev,
                         size_t scratch_bytes, void* stream);

/* yh_average_precision_static: NO COUNTERPART IN THE REFERENCE — added
 * for callers that cannot allocate device memory (code under CUDA graph capture,
 * tight serving loops, harnesses that forbid cudaMalloc).  Same kernel, arguments, results and
 * error behaviour as yh_average_precision, but the per-detection arrays of problems above 8192
 * detections live in a file-scope `static __device__` array of YH_AP_STATIC_BYTES inside the
 * library instead of caller scratch; the call allocates and frees nothing.
 *   limit   yh_average_precision_scratch_bytes(num_problems, max_detections_per_problem) must be
 *           <= YH_AP_STATIC_BYTES (e.g. one problem of 131 072 detections, or three of 65 536);
 *           beyond that -> YH_ERR_CAPACITY with a message naming both sizes (use
 *           yh_average_precision with caller scratch).  Up to 8192 detections per problem the
 *           array is not touched at all.
 *   cost    every GPU that loads the library holds the YH_AP_STATIC_BYTES until the process exits.
 *   sharing every call on a device uses the SAME array, which keeps the last call's contents.
 *           Nothing in it has to start at zero (the kernel writes each element it reads first),
 *           but two calls that overlap on the GPU (different streams, or two host threads on one
 *           device) RACE on it: serialise them, or use yh_average_precision with own scratch.
 *   set-up  the first call on each device asks cudaGetSymbolAddress for the array's device
 *           address with that device current (the calling thread's current device must be the one
 *           `stream` belongs to); later calls reuse it. */
#define YH_AP_STATIC_BYTES (4u << 20)
int yh_average_precision_static(const int32_t* det_gt_dev, const double* det_iou_dev,
                                const double* det_conf_dev, const int32_t* problem_offsets_dev,
                                const int32_t* num_ground_truth_dev, int num_problems,
                                int max_detections_per_problem, const double* iou_thresholds_dev,
                                int num_thresholds, int n_points, double* out_ap_dev, void* stream);

/* ---- §8e: the data-parallel loss exchange as peer-memory stores over NVLink / NVSwitch.
 * Replaces the per-step master-device averaging of the loss scalars
 * (`train/src/train/multi_gpu.rs:188-195`, `yolo-dl/src/loss/loss_.rs:527-577`): sum_g loss_g*B_g /
 * sum_g B_g.  Every rank (one process per GPU of ONE node) creates a mailbox
 * [slots][world][8 f32], exchanges the 64-byte CUDA IPC handle with its peers (any transport:
 * torch.distributed all_gather here) and opens theirs.  yh_loss_publish — one 32-thread kernel,
 * capturable in a CUDA graph, enqueued after the loss of a step — writes this rank's
 * [B_g*total, B_g*iou, B_g*cls, B_g*obj, B_g, -, -, seq] into entry [seq % slots][rank] of every
 * peer's mailbox (peer_mails[r], r = 0..world-1, own mailbox at [rank]); the sequence number lives
 * on the device and advances with every launch.  yh_loss_collect (synchronises `stream`) returns
 * the combined [total, iou, cls, obj] of the newest step all ranks have published, and its
 * sequence number (0 = none complete yet).  No collective kernel, no host work per step. */
typedef struct YhMailbox YhMailbox;
int yh_peer_mailbox_create(int device, int world, int slots, YhMailbox** out, void** mail_dev,
                           unsigned char* handle64);
int yh_peer_mailbox_open(int device, const unsigned char* handle64, void** peer_mail);
int yh_peer_mailbox_close(int device, void* peer_mail);
int yh_peer_mailbox_destroy(YhMailbox* m);
int yh_loss_publish(YhMailbox* m, const float* losses4_dev, float weight, void* const* peer_mails,
                    int rank, void* stream);
int yh_loss_collect(YhMailbox* m, float* out4, unsigned* out_seq, void* stream);

/* ---- a14: YoloInference::forward post-pass (`yolo-dl/src/loss/inference.rs:50-77`):
 * among kept rows of the same (batch, instance) keep the one with max confidence.
 * max_by_key keeps the LAST maximum: equal confidences keep the later row.
 * In-place, order-preserving compaction of the yh_nms outputs; *io_K_dev updated. */
int yh_inference_pick(YhWorkspace* ws, int64_t* kb, int64_t* kc, int64_t* ki, float* tlbr,
                      float* conf, int32_t* io_K_dev, int capacity, void* stream);

/* Synchronising read of device-side counters; returns YH_ERR_CAPACITY if any call since
 * the last read overflowed a buffer. Any pointer may be NULL. */
int yh_read_counts(YhWorkspace* ws, const int32_t* M_dev, int32_t* M_host,
                   const int32_t* K_dev, int32_t* K_host, void* stream);

/* The reference's candidate filter is `nonzero()`: it takes however many rows pass (nms.rs:97-103).
 * yh_nms keeps per-image candidate lists of `cand_per_image` entries; the filter keeps COUNTING past
 * that, so when yh_read_counts reports YH_ERR_CAPACITY (flag 2) this returns the largest per-image
 * candidate count seen since the previous read (0 if none overflowed): destroy the workspace, create
 * one with that cand_per_image and call yh_nms again — the count -> allocate -> fill form of
 * nonzero(), paid only when a batch outgrows the workspace.  Host value, no synchronisation. */
int yh_workspace_candidates_needed(const YhWorkspace* ws);

/* Device pointer to the workspace's sticky int32 overflow / order flag (bits: 1 matches, 2 NMS
 * candidates, 4 NMS outputs, 8 unsorted match list; cleared by yh_read_counts).  A caller that
 * replays a captured step and copies its results out asynchronously copies this word with them
 * instead of paying yh_read_counts' synchronisation: non-zero = the step's outputs are truncated. */
const int32_t* yh_workspace_error_flag(const YhWorkspace* ws);
/* The same word copied to pinned host memory on `stream` (asynchronous, does not clear it). */
int yh_copy_error_flag_async(YhWorkspace* ws, int32_t* flag_host_pinned, void* stream);

/* Number of kernels this library has launched on the calling thread since the last
 * reset (bench.py's `gpu_launches`). */
int64_t yh_launch_count(int reset);

/* Optional per-launch timing for benchmarks: while enabled (thread-local) every kernel launch of
 * this library is bracketed by a CUDA event pair on its launch stream.  yh_timing_read
 * synchronises on those events and returns, per kernel id (order: decode, assign_slots,
 * assign_emit, loss_dense, loss_rows, nms_filter, nms_segment, nms_emit, nms_pack, pick,
 * benchmark, match_det_gt, ap, loss_finish, loss_publish, nms_big), the summed device time in ms
 * and the launch count since the last read. */
#define YH_NUM_KERNELS 16
int yh_timing_enable(int on);
/* Debug aid: while dev_buf (2 * capacity uint64, device) is set on this thread, launch number i
 * since the call (in issue order, i < capacity) is bracketed by one-thread kernels that write
 * %globaltimer (ns) to dev_buf[2*i] and dev_buf[2*i+1]; capturable into a CUDA graph, so it shows
 * where every kernel really runs inside an overlapped, replayed step.  NULL disables.
 * yh_trace_ids copies the kernel id (order of yh_timing_read) of the first n traced launches and
 * returns how many launches were issued while tracing was on. */
int yh_trace_enable(unsigned long long* dev_buf, int capacity);
int yh_trace_ids(int32_t* ids, int n);
int yh_timing_read(double* ms_sum, int64_t* launches, int n);

#ifdef __cplusplus
}
#endif
#endif /* YOLOHOT_H */
