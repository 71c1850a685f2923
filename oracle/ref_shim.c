/* ref_shim.c — test infrastructure only.  Batch entry point around the reference repository's
 * vendored darknet box metrics (darknet-test/darknet/src/box.c:135-256: box_iou, box_giou,
 * box_diounms, box_ciou), which oracle/build.py compiles FROM WHERE THEY LIE under /root/reference
 * into oracle/_ref/libdarknet_box.so together with this file.  Nothing of box.c is copied here:
 * only the struct layout (`include/darknet.h:867-869`: four floats x, y, w, h) and the four
 * prototypes are re-declared.  xcalloc is the one external symbol box.c's (unused) NMS helpers
 * reference; it is stubbed so that the library loads with RTLD_NOW. */
#include <stdlib.h>

typedef struct { float x, y, w, h; } box;
float box_iou(box a, box b);
float box_giou(box a, box b);
float box_diounms(box a, box b, float beta1);
float box_ciou(box a, box b);

void* xcalloc(size_t n, size_t s) { return calloc(n, s); }

/* a, b: n rows of (x = centre x, y = centre y, w, h) */
void dk_box_metrics(const float* a, const float* b, int n, float* iou, float* giou, float* diou,
                    float* ciou) {
  for (int i = 0; i < n; ++i) {
    box p = {a[4 * i], a[4 * i + 1], a[4 * i + 2], a[4 * i + 3]};
    box q = {b[4 * i], b[4 * i + 1], b[4 * i + 2], b[4 * i + 3]};
    iou[i] = box_iou(p, q);
    giou[i] = box_giou(p, q);
    diou[i] = box_diounms(p, q, 1.0f); /* iou - (rho^2/c^2)^1: the DIoU of the paper */
    ciou[i] = box_ciou(p, q);
  }
}
