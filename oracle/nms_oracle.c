/* ORACLE (test infrastructure only): plain-C restatement of the reference's BEV NMS arithmetic.
 *   rotated BEV overlap / IoU : mmdet3d/ops/pcdet_nms/src/iou3d_nms_kernel.cu:36-234 (same maths as iou3d_cpu.cpp:59-229)
 *   axis-aligned BEV IoU      : mmdet3d/ops/pcdet_nms/src/iou3d_nms_kernel.cu:314-325
 *   greedy keep loop          : mmdet3d/ops/pcdet_nms/src/iou3d_nms.cpp:116-132
 * Pinned against the reference's own iou3d_cpu.cpp (oracle/_ref, CPU) in tests/test_oracle.py and against the
 * reference's own CUDA kernels recompiled for sm_100a (oracle/_ref, GPU box) in tests/test_nms_gpu.py.
 * Compiled with -ffp-contract=off: single-precision operations in source order. */
#include <math.h>
#include <stdint.h>
#include <string.h>

#define EPS 1e-8f

typedef struct { float x, y; } pt;

static float cross3(pt p1, pt p2, pt p0) { return (p1.x - p0.x) * (p2.y - p0.y) - (p2.x - p0.x) * (p1.y - p0.y); }
static float cross2(pt a, pt b) { return a.x * b.y - a.y * b.x; }

static int rect_cross(pt p1, pt p2, pt q1, pt q2) {
  return fminf(p1.x, p2.x) <= fmaxf(q1.x, q2.x) && fminf(q1.x, q2.x) <= fmaxf(p1.x, p2.x) &&
         fminf(p1.y, p2.y) <= fmaxf(q1.y, q2.y) && fminf(q1.y, q2.y) <= fmaxf(p1.y, p2.y);
}

static int in_box(const float* box, pt p) {
  const float margin = 1e-2f;
  float cx = box[0], cy = box[1];
  float c = cosf(-box[6]), s = sinf(-box[6]);
  float rx = (p.x - cx) * c + (p.y - cy) * (-s);
  float ry = (p.x - cx) * s + (p.y - cy) * c;
  return fabsf(rx) < box[3] / 2 + margin && fabsf(ry) < box[4] / 2 + margin;
}

static int seg_x(pt p1, pt p0, pt q1, pt q0, pt* ans) {
  if (!rect_cross(p0, p1, q0, q1)) return 0;
  float s1 = cross3(q0, p1, p0), s2 = cross3(p1, q1, p0), s3 = cross3(p0, q1, q0), s4 = cross3(q1, p1, q0);
  if (!(s1 * s2 > 0 && s3 * s4 > 0)) return 0;
  float s5 = cross3(q1, p1, p0);
  if (fabsf(s5 - s1) > EPS) {
    ans->x = (s5 * q0.x - s1 * q1.x) / (s5 - s1);
    ans->y = (s5 * q0.y - s1 * q1.y) / (s5 - s1);
  } else {
    float a0 = p0.y - p1.y, b0 = p1.x - p0.x, c0 = p0.x * p1.y - p1.x * p0.y;
    float a1 = q0.y - q1.y, b1 = q1.x - q0.x, c1 = q0.x * q1.y - q1.x * q0.y;
    float D = a0 * b1 - a1 * b0;
    ans->x = (b0 * c1 - b1 * c0) / D;
    ans->y = (a1 * c0 - a0 * c1) / D;
  }
  return 1;
}

static void corners(const float* b, pt* c) {
  float hx = b[3] / 2, hy = b[4] / 2;
  float x1 = b[0] - hx, y1 = b[1] - hy, x2 = b[0] + hx, y2 = b[1] + hy;
  pt raw[4] = {{x1, y1}, {x2, y1}, {x2, y2}, {x1, y2}};
  float ca = cosf(b[6]), sa = sinf(b[6]);
  for (int k = 0; k < 4; ++k) {
    c[k].x = (raw[k].x - b[0]) * ca + (raw[k].y - b[1]) * (-sa) + b[0];
    c[k].y = (raw[k].x - b[0]) * sa + (raw[k].y - b[1]) * ca + b[1];
  }
  c[4] = c[0];
}

float orc_bev_overlap(const float* a, const float* b) {
  pt ca[5], cb[5], pts[16], ctr = {0.f, 0.f};
  int cnt = 0;
  corners(a, ca);
  corners(b, cb);
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j)
      if (seg_x(ca[i + 1], ca[i], cb[j + 1], cb[j], &pts[cnt])) {
        ctr.x += pts[cnt].x;
        ctr.y += pts[cnt].y;
        cnt++;
      }
  for (int k = 0; k < 4; ++k) {
    if (in_box(a, cb[k])) { ctr.x += cb[k].x; ctr.y += cb[k].y; pts[cnt++] = cb[k]; }
    if (in_box(b, ca[k])) { ctr.x += ca[k].x; ctr.y += ca[k].y; pts[cnt++] = ca[k]; }
  }
  ctr.x /= cnt;
  ctr.y /= cnt;
  for (int j = 0; j < cnt - 1; ++j)
    for (int i = 0; i < cnt - j - 1; ++i)
      if (atan2f(pts[i].y - ctr.y, pts[i].x - ctr.x) > atan2f(pts[i + 1].y - ctr.y, pts[i + 1].x - ctr.x)) {
        pt t = pts[i]; pts[i] = pts[i + 1]; pts[i + 1] = t;
      }
  float area = 0.f;
  for (int k = 0; k < cnt - 1; ++k) {
    pt u = {pts[k].x - pts[0].x, pts[k].y - pts[0].y}, v = {pts[k + 1].x - pts[0].x, pts[k + 1].y - pts[0].y};
    area += cross2(u, v);
  }
  return fabsf(area) / 2.0f;
}

float orc_iou_bev(const float* a, const float* b) {
  float sa = a[3] * a[4], sb = b[3] * b[4], s = orc_bev_overlap(a, b);
  return s / fmaxf(sa + sb - s, EPS);
}

float orc_iou_normal(const float* a, const float* b) {
  float left = fmaxf(a[0] - a[3] / 2, b[0] - b[3] / 2), right = fminf(a[0] + a[3] / 2, b[0] + b[3] / 2);
  float top = fmaxf(a[1] - a[4] / 2, b[1] - b[4] / 2), bottom = fminf(a[1] + a[4] / 2, b[1] + b[4] / 2);
  float w = fmaxf(right - left, 0.f), h = fmaxf(bottom - top, 0.f);
  float inter = w * h, Sa = a[3] * a[4], Sb = b[3] * b[4];
  return inter / fmaxf(Sa + Sb - inter, EPS);
}

/* mode 0 overlap, 1 rotated IoU, 2 axis-aligned IoU */
void orc_pair_matrix(const float* a, int na, const float* b, int nb, int mode, float* out) {
  for (int i = 0; i < na; ++i)
    for (int j = 0; j < nb; ++j) {
      const float *p = a + 7 * i, *q = b + 7 * j;
      out[(long)i * nb + j] = mode == 0 ? orc_bev_overlap(p, q) : mode == 1 ? orc_iou_bev(p, q) : orc_iou_normal(p, q);
    }
}

/* suppression mask exactly as the reference lays it out: mask[i * col_blocks + c], bit j = IoU(i, 64c + j) > thr,
 * only j > i inside the diagonal block; all (row, col) blocks computed like the reference kernel does */
void orc_nms_mask(const float* boxes, int n, float thr, int rotated, uint64_t* mask) {
  int cb = (n + 63) / 64;
  memset(mask, 0, sizeof(uint64_t) * (size_t)n * cb);
  for (int i = 0; i < n; ++i)
    for (int c = 0; c < cb; ++c) {
      uint64_t t = 0;
      int lim = n - c * 64 < 64 ? n - c * 64 : 64;
      int start = (i / 64 == c) ? (i % 64) + 1 : 0;
      for (int j = start; j < lim; ++j) {
        const float *p = boxes + 7 * i, *q = boxes + 7 * (c * 64 + j);
        float v = rotated ? orc_iou_bev(p, q) : orc_iou_normal(p, q);
        if (v > thr) t |= 1ull << j;
      }
      mask[(size_t)i * cb + c] = t;
    }
}

/* iou3d_nms.cpp:116-132 */
int orc_nms_greedy(const uint64_t* mask, int n, long* keep) {
  int cb = (n + 63) / 64, kept = 0;
  uint64_t remv[cb > 0 ? cb : 1];
  memset(remv, 0, sizeof(remv));
  for (int i = 0; i < n; ++i) {
    int nblock = i / 64, inblock = i % 64;
    if (!(remv[nblock] & (1ull << inblock))) {
      keep[kept++] = i;
      const uint64_t* p = mask + (size_t)i * cb;
      for (int j = nblock; j < cb; ++j) remv[j] |= p[j];
    }
  }
  return kept;
}
