/* oracle/nms_c.c -- TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).
 *
 * Plain-C restatement of the reference's custom_nms (README.md:62-83 =
 * tech.md:72-93) with torchvision.ops.box_iou's fp32 arithmetic
 * (torchvision/ops/boxes.py:297,321-339,369; dependency unpinned in
 * requirements.txt:7, 0.26 restated here).  Exists so the full-size NMS
 * stress config (30k boxes x 32 images x 5 thresholds) can be checked in
 * seconds; it is itself validated against oracle.hotpath.custom_nms (the
 * verbatim torch snippet) in tests/test_oracle.py.  PARITY UNPINNED: the
 * reference ships no golden vectors for this path.
 *
 * Build: gcc -O2 -ffp-contract=off -shared -fPIC  (no FMA contraction: every
 * fp32 operation rounds once, like torch's elementwise kernels).
 */
#include <stdint.h>
#include <stdlib.h>

typedef struct { float s; int i; } yv_key;

/* D5: descending score, ties -> lower original index first
 * (torch.argsort(descending=True, stable=True), README.md:64). */
static int cmp_desc(const void *a, const void *b) {
    int ia = ((const yv_key *)a)->i, ib = ((const yv_key *)b)->i;
    float sa = ((const yv_key *)a)->s, sb = ((const yv_key *)b)->s;
    if (sa > sb) return -1;
    if (sa < sb) return 1;
    return (ia > ib) - (ia < ib);
}

static inline float max_t(float a, float b) {   /* torch.max: NaN propagates */
    if (a != a) return a;
    if (b != b) return b;
    return a > b ? a : b;
}
static inline float min_t(float a, float b) {
    if (a != a) return a;
    if (b != b) return b;
    return a < b ? a : b;
}
static inline float clamp0(float d) { return d < 0.0f ? 0.0f : d; }  /* NaN stays NaN */

float yv_oracle_iou(const float *p, const float *q) {
    float area1 = (p[2] - p[0]) * (p[3] - p[1]);
    float area2 = (q[2] - q[0]) * (q[3] - q[1]);
    float w = clamp0(min_t(p[2], q[2]) - max_t(p[0], q[0]));
    float h = clamp0(min_t(p[3], q[3]) - max_t(p[1], q[1]));
    float inter = w * h;
    float uni = (area1 + area2) - inter;
    return inter / uni;
}

/* boxes[n][4] xyxy f32, scores[n] f32 (no NaN scores), thr already cast to
 * fp32 (torch compares the fp32 tensor against the scalar in fp32).  keep[]
 * must hold n ints; returns the number kept, in descending-score order. */
int yv_oracle_nms(const float *boxes, const float *scores, int n, float thr, int32_t *keep) {
    if (n <= 0) return 0;
    yv_key *keys = (yv_key *)malloc(sizeof(yv_key) * (size_t)n);
    int *order = (int *)malloc(sizeof(int) * (size_t)n);
    unsigned char *dead = (unsigned char *)calloc((size_t)n, 1);
    for (int i = 0; i < n; ++i) { keys[i].s = scores[i]; keys[i].i = i; }
    qsort(keys, (size_t)n, sizeof(yv_key), cmp_desc);
    for (int i = 0; i < n; ++i) order[i] = keys[i].i;
    free(keys);
    int nk = 0;
    for (int a = 0; a < n; ++a) {
        if (dead[a]) continue;
        int i = order[a];
        keep[nk++] = i;
        const float *p = boxes + 4 * (size_t)i;
        for (int b = a + 1; b < n; ++b) {
            if (dead[b]) continue;
            float iou = yv_oracle_iou(p, boxes + 4 * (size_t)order[b]);
            if (!(iou < thr)) dead[b] = 1;      /* survivors: iou < thr (strict); NaN suppressed */
        }
    }
    free(order);
    free(dead);
    return nk;
}

/* Batched convenience: offsets[B+1] into concatenated boxes/scores; keep is
 * written at the same offsets, counts[B]. */
void yv_oracle_nms_batched(const float *boxes, const float *scores, const int64_t *offsets,
                           int B, float thr, int32_t *keep, int32_t *counts) {
    for (int b = 0; b < B; ++b) {
        int64_t o = offsets[b];
        counts[b] = yv_oracle_nms(boxes + 4 * o, scores + o, (int)(offsets[b + 1] - o), thr, keep + o);
    }
}
