// Device-side point-cloud augmentation of the FCAF3D training pipeline: one fused pass per scene for
//   IndoorPointSample   -> row gather by the sampled indices          (transforms_3d.py:838-880)
//   RandomFlip3D        -> sign of x / y                              (transforms_3d.py:144-160, depth_points.py:28-33)
//   GlobalRotScaleTrans -> rotation about z, uniform scale, translation (transforms_3d.py:541-648, base_points.py:139-269)
// and the same maps for the ground-truth boxes (depth_box3d.py:119-208, base_box3d.py:149-222).  The random numbers are
// drawn by the caller (host RNG in the reference's draw order, or a device permutation), so the kernels are deterministic
// functions of their arguments and can be checked against the reference's own containers.
// HBM-bound: dim * 4 bytes read (scattered rows) + dim * 4 bytes written per output point.
#include "common.cuh"
#include <math.h>

__global__ void k_augment_points(const float* __restrict__ in, long long n_in, int dim,
                                 const long long* __restrict__ choices, long long n_out, float sgn_x, float sgn_y,
                                 float c, float s, float scale, float tx, float ty, float tz, float* __restrict__ out,
                                 int* __restrict__ err) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_out) return;
  const long long src = choices ? choices[i] : i;
  if (src < 0 || src >= n_in) {
    atomicExch(err, 1);
    return;
  }
  const float* p = in + src * dim;
  float* o = out + i * dim;
  const float x = sgn_x * p[0], y = sgn_y * p[1], z = p[2];
  // xyz @ rot_mat_T with rot_mat_T = [[c, s, 0], [-s, c, 0], [0, 0, 1]]
  const float xr = x * c + y * (-s);
  const float yr = x * s + y * c;
  o[0] = xr * scale + tx;
  o[1] = yr * scale + ty;
  o[2] = z * scale + tz;
  for (int d = 3; d < dim; ++d) o[d] = p[d];
}

// boxes (cx, cy, cz, dx, dy, dz, yaw), gravity centre
__global__ void k_augment_boxes(const float* __restrict__ in, int n, int with_yaw, int flip_h, int flip_v, float angle,
                                float c, float s, float scale, float tx, float ty, float tz, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* b = in + 7ll * i;
  float x = b[0], y = b[1], z = b[2], dx = b[3], dy = b[4], dz = b[5], yaw = b[6];
  if (flip_h) {
    x = -x;
    if (with_yaw) yaw = -yaw + 3.14159265358979323846f;
  }
  if (flip_v) {
    y = -y;
    if (with_yaw) yaw = -yaw;
  }
  const float xr = x * c + y * (-s), yr = x * s + y * c;
  if (with_yaw) {
    yaw -= angle;
  } else {  // axis-aligned boxes stay axis-aligned: sizes = extents of the rotated footprint
    const float ndx = fabsf(dx * c) + fabsf(dy * s), ndy = fabsf(dx * s) + fabsf(dy * c);
    dx = ndx;
    dy = ndy;
  }
  float* o = out + 7ll * i;
  o[0] = xr * scale + tx;
  o[1] = yr * scale + ty;
  o[2] = z * scale + tz;
  o[3] = dx * scale;
  o[4] = dy * scale;
  o[5] = dz * scale;
  o[6] = yaw;
}

extern "C" {

// out [n_out, dim] = augment(in[choices[i]]) (choices int64 [n_out] on the device, NULL = identity with n_out == n_in);
// err int32 [1] device flag (zeroed by the caller): 1 = an index outside [0, n_in)
int b2_augment_points(const float* in, long long n_in, int dim, const long long* choices, long long n_out, int flip_h,
                      int flip_v, float angle, float scale, float tx, float ty, float tz, float* out, int* err,
                      void* stream) {
  B2_CHECK_ARG(dim >= 3, "points need at least xyz (dim = %d)", dim);
  B2_CHECK_ARG(choices != nullptr || n_out == n_in, "without sampled indices n_out must equal n_in");
  if (n_out == 0) return B2_OK;
  k_augment_points<<<b2_div_up(n_out, 256), 256, 0, (cudaStream_t)stream>>>(
      in, n_in, dim, choices, n_out, flip_h ? -1.f : 1.f, flip_v ? -1.f : 1.f, cosf(angle), sinf(angle), scale, tx, ty, tz,
      out, err);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_augment_boxes(const float* boxes, int n, int with_yaw, int flip_h, int flip_v, float angle, float scale, float tx,
                     float ty, float tz, float* out, void* stream) {
  if (n == 0) return B2_OK;
  k_augment_boxes<<<b2_div_up(n, 128), 128, 0, (cudaStream_t)stream>>>(boxes, n, with_yaw, flip_h, flip_v, angle,
                                                                     cosf(angle), sinf(angle), scale, tx, ty, tz, out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
