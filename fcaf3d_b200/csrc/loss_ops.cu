// Box-regression and centerness losses of the FCAF3D head, value and gradient in ONE launch each (sm_100a).
//
// Reference (fp32 torch graphs, ~40 small kernels + the un-vendored `sort_v` CUDA op for the rotated case):
//   IoU3DLoss                mmdet3d/models/losses/iou3d_loss.py:11-13,21-35,46-75
//   cal_iou_3d / cal_iou     mmdet3d/ops/rotated_iou/oriented_iou_loss.py:86-109, :38-58, box2corners_th :6-35
//   oriented_box_intersection_2d  mmdet3d/ops/rotated_iou/box_intersection_2d.py:13-183
//   axis-aligned IoU         mmdet3d/core/bbox/iou_calculators/iou3d_calculator.py:279-292,310-312
//   _bbox_pred_to_bbox       mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py:281-330
//   centerness BCE           fcaf3d_neck_with_head.py:191-193 -> mmdet CrossEntropyLoss(use_sigmoid=True)
//
// One thread per box pair.  The gradient is carried by forward-mode dual numbers (value + one tangent per input of the
// pair) through exactly the reference's expression graph, so it equals what torch autograd returns for that graph: masked
// intersection points have zero gradient, the angular sort is not differentiated, |shoelace| / 2 takes the sign of the sum.
// The rotated intersection needs the polygon's vertex ORDER before it can be differentiated; a first float pass finds the
// (at most 8) vertices and sorts them, the dual pass recomputes only those.
#include "common.cuh"

namespace {

template <int N>
struct Dual {
  float v;
  float d[N];
  __device__ Dual() {}
  __device__ Dual(float x) : v(x) {
#pragma unroll
    for (int i = 0; i < N; ++i) d[i] = 0.f;
  }
};

template <int N>
__device__ __forceinline__ Dual<N> seed(float x, int k) {
  Dual<N> r(x);
  r.d[k] = 1.f;
  return r;
}
#define DUAL_BIN(op, val, da, db)                                                      \
  template <int N>                                                                     \
  __device__ __forceinline__ Dual<N> operator op(const Dual<N>& a, const Dual<N>& b) { \
    Dual<N> r;                                                                         \
    r.v = val;                                                                         \
    _Pragma("unroll") for (int i = 0; i < N; ++i) r.d[i] = (da)*a.d[i] + (db)*b.d[i];  \
    return r;                                                                          \
  }
DUAL_BIN(+, a.v + b.v, 1.f, 1.f)
DUAL_BIN(-, a.v - b.v, 1.f, -1.f)
DUAL_BIN(*, a.v * b.v, b.v, a.v)
DUAL_BIN(/, a.v / b.v, 1.f / b.v, -(a.v / b.v) / b.v)
#undef DUAL_BIN
template <int N>
__device__ __forceinline__ Dual<N> operator+(const Dual<N>& a, float b) { Dual<N> r = a; r.v += b; return r; }
template <int N>
__device__ __forceinline__ Dual<N> operator-(const Dual<N>& a, float b) { Dual<N> r = a; r.v -= b; return r; }
template <int N>
__device__ __forceinline__ Dual<N> operator-(float a, const Dual<N>& b) {
  Dual<N> r;
  r.v = a - b.v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = -b.d[i];
  return r;
}
template <int N>
__device__ __forceinline__ Dual<N> operator*(const Dual<N>& a, float b) {
  Dual<N> r;
  r.v = a.v * b;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = a.d[i] * b;
  return r;
}
template <int N>
__device__ __forceinline__ Dual<N> operator*(float b, const Dual<N>& a) { return a * b; }
template <int N>
__device__ __forceinline__ Dual<N> operator/(const Dual<N>& a, float b) { return a * (1.f / b); }
template <int N>
__device__ __forceinline__ Dual<N> chain(const Dual<N>& a, float v, float dv) {   // f(a) with f = v, f' = dv at a.v
  Dual<N> r;
  r.v = v;
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = dv * a.d[i];
  return r;
}
template <int N> __device__ __forceinline__ Dual<N> dsin(const Dual<N>& a) { return chain(a, sinf(a.v), cosf(a.v)); }
template <int N> __device__ __forceinline__ Dual<N> dcos(const Dual<N>& a) { return chain(a, cosf(a.v), -sinf(a.v)); }
template <int N> __device__ __forceinline__ Dual<N> dexp(const Dual<N>& a) { const float e = expf(a.v); return chain(a, e, e); }
template <int N> __device__ __forceinline__ Dual<N> dsqrt(const Dual<N>& a) { const float s = sqrtf(a.v); return chain(a, s, 0.5f / s); }
template <int N> __device__ __forceinline__ Dual<N> dabs(const Dual<N>& a) {
  return chain(a, fabsf(a.v), a.v > 0.f ? 1.f : (a.v < 0.f ? -1.f : 0.f));
}
template <int N>
__device__ __forceinline__ Dual<N> datan2(const Dual<N>& y, const Dual<N>& x) {
  const float r2 = x.v * x.v + y.v * y.v;
  Dual<N> r;
  r.v = atan2f(y.v, x.v);
#pragma unroll
  for (int i = 0; i < N; ++i) r.d[i] = (x.v * y.d[i] - y.v * x.d[i]) / r2;
  return r;
}
// torch.min / torch.max of two tensors: the gradient goes to the smaller / larger argument, half to each on a tie
template <int N>
__device__ __forceinline__ Dual<N> dmin(const Dual<N>& a, float b) {
  if (a.v < b) return a;
  if (a.v > b) return Dual<N>(b);
  return chain(a, a.v, 0.5f);
}
template <int N>
__device__ __forceinline__ Dual<N> dmax(const Dual<N>& a, float b) {
  if (a.v > b) return a;
  if (a.v < b) return Dual<N>(b);
  return chain(a, a.v, 0.5f);
}
template <int N>
__device__ __forceinline__ Dual<N> dclamp_min(const Dual<N>& a, float lo) {   // clamp_min: gradient 1 where a >= lo
  return a.v >= lo ? a : Dual<N>(lo);
}

// float twins, so that the geometry below is written once
__device__ __forceinline__ float val(float a) { return a; }
template <int N> __device__ __forceinline__ float val(const Dual<N>& a) { return a.v; }
__device__ __forceinline__ float dsin(float a) { return sinf(a); }
__device__ __forceinline__ float dcos(float a) { return cosf(a); }

// corners of (x, y, w, h, alpha): local (+-w/2, +-h/2) in the order (+,+), (-,+), (-,-), (+,-), rotated by alpha, shifted
template <class T>
__device__ __forceinline__ void box_corners(const T& x, const T& y, const T& w, const T& h, const T& alpha, T (*c)[2]) {
  const float sx[4] = {0.5f, -0.5f, -0.5f, 0.5f};
  const float sy[4] = {0.5f, 0.5f, -0.5f, -0.5f};
  const T s = dsin(alpha), co = dcos(alpha);
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    const T lx = w * sx[k], ly = h * sy[k];
    c[k][0] = lx * co - ly * s + x;
    c[k][1] = lx * s + ly * co + y;
  }
}

// intersection point of edge i of box 1 (corners a) with edge j of box 2 (corners b); only called for valid pairs
template <class T>
__device__ __forceinline__ void edge_point(const T (*a)[2], const float (*b)[2], int i, int j, T* px, T* py) {
  const T x1 = a[i][0], y1 = a[i][1], x2 = a[(i + 1) & 3][0], y2 = a[(i + 1) & 3][1];
  const float x3 = b[j][0], y3 = b[j][1], x4 = b[(j + 1) & 3][0], y4 = b[(j + 1) & 3][1];
  const T num = (x1 - x2) * (y3 - y4) - (y1 - y2) * (x3 - x4);
  const T den_t = (x1 - x3) * (y3 - y4) - (y1 - y3) * (x3 - x4);
  const T t = den_t / (num + 1e-8f);
  *px = x1 + t * (x2 - x1);
  *py = y1 + t * (y2 - y1);
}

// corners of `ca` inside the rectangle `cb` (box1_in_box2), edges included with a 1e-6 margin
__device__ __forceinline__ bool corner_inside(const float* m, const float (*cb)[2]) {
  const float abx = cb[1][0] - cb[0][0], aby = cb[1][1] - cb[0][1];
  const float adx = cb[3][0] - cb[0][0], ady = cb[3][1] - cb[0][1];
  const float amx = m[0] - cb[0][0], amy = m[1] - cb[0][1];
  const float p_ab = (abx * amx + aby * amy) / (abx * abx + aby * aby);
  const float p_ad = (adx * amx + ady * amy) / (adx * adx + ady * ady);
  return p_ab > -1e-6f && p_ab < 1.f + 1e-6f && p_ad > -1e-6f && p_ad < 1.f + 1e-6f;
}

// Vertex indices (0-3 corners of box 1, 4-7 corners of box 2, 8 + 4 i + j edge intersections) of the intersection
// polygon in counter-clockwise order; returns their number (0 when fewer than 3 are valid: no area, no gradient).
__device__ int polygon_order(const float (*c1)[2], const float (*c2)[2], int* order) {
  float vx[24], vy[24];
  unsigned int mask = 0;
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    vx[k] = c1[k][0], vy[k] = c1[k][1];
    vx[4 + k] = c2[k][0], vy[4 + k] = c2[k][1];
    if (corner_inside(c1[k], c2)) mask |= 1u << k;
    if (corner_inside(c2[k], c1)) mask |= 1u << (4 + k);
  }
  for (int i = 0; i < 4; ++i)
    for (int j = 0; j < 4; ++j) {
      const float x1 = c1[i][0], y1 = c1[i][1], x2 = c1[(i + 1) & 3][0], y2 = c1[(i + 1) & 3][1];
      const float x3 = c2[j][0], y3 = c2[j][1], x4 = c2[(j + 1) & 3][0], y4 = c2[(j + 1) & 3][1];
      const float num = (x1 - x2) * (y3 - y4) - (y1 - y2) * (x3 - x4);
      const float den_t = (x1 - x3) * (y3 - y4) - (y1 - y3) * (x3 - x4);
      const float den_u = (x1 - x2) * (y1 - y3) - (y1 - y2) * (x1 - x3);
      const float t = num == 0.f ? -1.f : den_t / num;
      const float u = num == 0.f ? -1.f : -den_u / num;
      const bool ok = t > 0.f && t < 1.f && u > 0.f && u < 1.f;
      const int k = 8 + 4 * i + j;
      vx[k] = vy[k] = 0.f;
      if (ok) {
        float px, py;
        edge_point<float>(c1, c2, i, j, &px, &py);
        vx[k] = px, vy[k] = py;
        mask |= 1u << k;
      }
    }
  const int nv = __popc(mask);
  if (nv < 3) return 0;
  float mx = 0.f, my = 0.f;
  for (int k = 0; k < 24; ++k)
    if ((mask >> k) & 1u) mx += vx[k], my += vy[k];
  mx /= (float)nv, my /= (float)nv;
  float ang[24];
  for (int k = 0; k < 24; ++k) ang[k] = atan2f(vy[k] - my, vx[k] - mx);
  unsigned int used = 0;
  int cnt = 0;
  for (; cnt < nv && cnt < 8; ++cnt) {          // selection sort, ascending angle, ties: lower index first (sort_v)
    int best = -1;
    for (int k = 0; k < 24; ++k)
      if (((mask & ~used) >> k) & 1u)
        if (best < 0 || ang[k] < ang[best]) best = k;
    used |= 1u << best;
    order[cnt] = best;
  }
  return cnt;
}

// rotated 3-D IoU of box b1 = (x, y, z, w, l, h, alpha) (T: float or dual) with the constant box b2
template <class T>
__device__ T rot_iou3d(const T* b1, const float* b2) {
  float c1f[4][2], c2[4][2];
  box_corners<float>(b2[0], b2[1], b2[3], b2[4], b2[6], c2);
  T c1[4][2];
  box_corners<T>(b1[0], b1[1], b1[3], b1[4], b1[6], c1);
#pragma unroll
  for (int k = 0; k < 4; ++k) c1f[k][0] = val(c1[k][0]), c1f[k][1] = val(c1[k][1]);
  int order[8];
  const int cnt = polygon_order(c1f, c2, order);
  T total(0.f);
  if (cnt) {
    T px[8], py[8];
    for (int k = 0; k < cnt; ++k) {
      const int id = order[k];
      if (id < 4) {
        px[k] = c1[id][0], py[k] = c1[id][1];
      } else if (id < 8) {
        px[k] = T(c2[id - 4][0]), py[k] = T(c2[id - 4][1]);
      } else {
        edge_point<T>(c1, c2, (id - 8) >> 2, (id - 8) & 3, &px[k], &py[k]);
      }
    }
    for (int k = 0; k < cnt; ++k) {
      const int k1 = k + 1 == cnt ? 0 : k + 1;
      total = total + (px[k] * py[k1] - py[k] * px[k1]);
    }
  }
  const T inter = dabs(total) / 2.f;
  const T u = b1[3] * b1[4] + (b2[3] * b2[4]) - inter;
  const T iou2d = inter / u;
  const T zmax1 = b1[2] + b1[5] * 0.5f, zmin1 = b1[2] - b1[5] * 0.5f;
  const float zmax2 = b2[2] + b2[5] * 0.5f, zmin2 = b2[2] - b2[5] * 0.5f;
  const T zov = dclamp_min(dmin(zmax1, zmax2) - dmax(zmin1, zmin2), 0.f);
  const T inter3d = iou2d * u * zov;
  const T v1 = b1[3] * b1[4] * b1[5];
  const float v2 = b2[3] * b2[4] * b2[5];
  return inter3d / (v1 + v2 - inter3d);
}

// axis-aligned 3-D IoU of (centre, size) boxes: overlap / max(vol1 + vol2 - overlap, 1e-6)
template <class T>
__device__ T aa_iou3d(const T* b1, const float* b2) {
  T ov(1.f), vol(1.f);
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const T lo = b1[a] - b1[3 + a] / 2.f, hi = b1[a] + b1[3 + a] / 2.f;
    const float lo2 = b2[a] - b2[3 + a] / 2.f, hi2 = b2[a] + b2[3 + a] / 2.f;
    ov = ov * dclamp_min(dmin(hi, hi2) - dmax(lo, lo2), 0.f);
    vol = vol * (hi - lo);
  }
  const float vol2 = (b2[3]) * (b2[4]) * (b2[5]);
  const T uni = vol + vol2 - ov;
  return ov / dmax(uni, 1e-6f);
}

// distances -> box in dual numbers (the same expressions as decode_box in head_ops.cu)
template <class T>
__device__ void decode_dual(const float* p, const T* bp, int n_reg, int yaw_mode, T* o) {
  o[0] = (bp[1] - bp[0]) / 2.f + p[0];
  o[1] = (bp[3] - bp[2]) / 2.f + p[1];
  o[2] = (bp[5] - bp[4]) / 2.f + p[2];
  o[3] = bp[0] + bp[1];
  o[4] = bp[2] + bp[3];
  o[5] = bp[4] + bp[5];
  o[6] = T(0.f);
  if (n_reg == 6 || yaw_mode == 0) return;
  if (yaw_mode == 1) {
    o[6] = bp[6];
  } else if (yaw_mode == 2) {
    const T norm = dsqrt(bp[6] * bp[6] + bp[7] * bp[7]);
    o[6] = datan2(bp[6] / norm, bp[7] / norm);
  } else {
    const T scale = bp[0] + bp[1] + bp[2] + bp[3];
    const T q = dexp(dsqrt(bp[6] * bp[6] + bp[7] * bp[7]));
    o[3] = scale / (q + 1.f);
    o[4] = scale / (q + 1.f) * q;
    o[6] = datan2(bp[6], bp[7]) * 0.5f;
  }
}

__global__ void __launch_bounds__(64) k_rot_iou3d(const float* __restrict__ pred, const float* __restrict__ tgt, int n,
                                                  float* __restrict__ iou, float* __restrict__ dpred) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  float b2[7];
  Dual<7> b1[7];
#pragma unroll
  for (int k = 0; k < 7; ++k) b1[k] = seed<7>(pred[(long long)i * 7 + k], k), b2[k] = tgt[(long long)i * 7 + k];
  const Dual<7> r = rot_iou3d<Dual<7>>(b1, b2);
  iou[i] = r.v;
  if (dpred)
#pragma unroll
    for (int k = 0; k < 7; ++k) dpred[(long long)i * 7 + k] = r.d[k];
}

template <int R>
__global__ void __launch_bounds__(64) k_box_iou_decoded(const float* __restrict__ points, const float* __restrict__ bbox_pred,
                                                        int yaw_mode, const float* __restrict__ tgt, int rotated, int n,
                                                        float* __restrict__ iou, float* __restrict__ dbbox) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  Dual<R> bp[R], box[7];
  float b2[7];
#pragma unroll
  for (int k = 0; k < R; ++k) bp[k] = seed<R>(bbox_pred[(long long)i * R + k], k);
#pragma unroll
  for (int k = 0; k < 7; ++k) b2[k] = tgt[(long long)i * 7 + k];
  decode_dual<Dual<R>>(points + (long long)i * 3, bp, R, yaw_mode, box);
  const Dual<R> r = rotated ? rot_iou3d<Dual<R>>(box, b2) : aa_iou3d<Dual<R>>(box, b2);
  iou[i] = r.v;
#pragma unroll
  for (int k = 0; k < R; ++k) dbbox[(long long)i * R + k] = r.d[k];
}

// binary cross entropy with logits, element-wise, numerically stable: max(x, 0) - x t + log(1 + exp(-|x|))
__global__ void k_bce_logits(const float* __restrict__ x, const float* __restrict__ t, long long n,
                             float* __restrict__ loss, float* __restrict__ grad) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float xv = x[i], tv = t[i];
  loss[i] = fmaxf(xv, 0.f) - xv * tv + log1pf(expf(-fabsf(xv)));
  grad[i] = 1.f / (1.f + expf(-xv)) - tv;
}

}  // namespace

extern "C" {

int b2_rot_iou3d(const float* pred, const float* tgt, long long n, float* iou, float* dpred, void* stream) {
  B2_CHECK_ARG(n >= 0 && n < (1ll << 31), "b2_rot_iou3d: bad n");
  if (n == 0) return B2_OK;
  B2_CHECK_ARG(pred && tgt && iou, "b2_rot_iou3d: null pointer");
  k_rot_iou3d<<<b2_div_up(n, 64), 64, 0, (cudaStream_t)stream>>>(pred, tgt, (int)n, iou, dpred);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_box_iou_decoded(const float* points, const float* bbox_pred, int n_reg, int yaw_mode, const float* tgt,
                       int rotated, long long n, float* iou, float* dbbox_pred, void* stream) {
  B2_CHECK_ARG(n >= 0 && n < (1ll << 31), "b2_box_iou_decoded: bad n");
  B2_CHECK_ARG(n_reg == 6 || n_reg == 7 || n_reg == 8, "b2_box_iou_decoded: n_reg must be 6, 7 or 8");
  B2_CHECK_ARG(yaw_mode >= 0 && yaw_mode <= 3 && (yaw_mode < 2 || n_reg == 8) && (yaw_mode != 1 || n_reg >= 7),
               "b2_box_iou_decoded: yaw_mode %d does not fit n_reg %d", yaw_mode, n_reg);
  if (n == 0) return B2_OK;
  B2_CHECK_ARG(points && bbox_pred && tgt && iou && dbbox_pred, "b2_box_iou_decoded: null pointer");
  const int grid = b2_div_up(n, 64);
  cudaStream_t st = (cudaStream_t)stream;
  if (n_reg == 6)
    k_box_iou_decoded<6><<<grid, 64, 0, st>>>(points, bbox_pred, yaw_mode, tgt, rotated, (int)n, iou, dbbox_pred);
  else if (n_reg == 7)
    k_box_iou_decoded<7><<<grid, 64, 0, st>>>(points, bbox_pred, yaw_mode, tgt, rotated, (int)n, iou, dbbox_pred);
  else
    k_box_iou_decoded<8><<<grid, 64, 0, st>>>(points, bbox_pred, yaw_mode, tgt, rotated, (int)n, iou, dbbox_pred);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_bce_logits(const float* x, const float* t, long long n, float* loss, float* grad, void* stream) {
  B2_CHECK_ARG(n >= 0, "b2_bce_logits: bad n");
  if (n == 0) return B2_OK;
  B2_CHECK_ARG(x && t && loss && grad, "b2_bce_logits: null pointer");
  k_bce_logits<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(x, t, n, loss, grad);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
