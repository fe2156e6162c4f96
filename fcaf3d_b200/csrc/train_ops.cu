// Training-path kernels: weight gradient of the sparse convolution, max-pool backward, the FCAF3D target
// assigner, sigmoid focal loss and the axis-aligned IoU loss (forward + backward), angular vertex sort for the
// rotated IoU loss.
//
// Reference behaviour:
//   conv backward            MinkowskiEngine v0.5.4 ConvolutionBackwardKernelGPU (not vendored)
//   Fcaf3DAssigner.assign    mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py:387-466, compute_centerness :377-384
//   sigmoid focal loss       mmcv-full 1.3.8 sigmoid_focal_loss CUDA op (label -1 = background)
//   axis-aligned IoU loss    mmdet3d/models/losses/iou3d_loss.py:21-35 + core/bbox/iou_calculators/iou3d_calculator.py:279-312
//   sort_v                   lilanxiao/Rotated_IoU@3bdca6b cuda_op/sort_vert_kernel.cu (not vendored)
#include "common.cuh"
#include <mma.h>
#include <stdlib.h>

template <typename T>
__device__ __forceinline__ float ldf(const T* p);
template <>
__device__ __forceinline__ float ldf<float>(const float* p) { return *p; }
template <>
__device__ __forceinline__ float ldf<__nv_bfloat16>(const __nv_bfloat16* p) { return __bfloat162float(*p); }

// ---------------------------------------------------------------------------
// wgrad: dW[k][ci][co] = sum_o in[nbr[k][o]][ci] * dout[o][co]   (fp32 accumulate)
// CTA = (64 x 64 tile of dW[k], one chunk of output rows); partial sums are combined with
// fp32 atomics into a zero-initialised dW (split-K over row chunks).
// ---------------------------------------------------------------------------
#define WG_ROWS 16

template <typename TA, typename TG>
__global__ void __launch_bounds__(256) k_wgrad(const TA* __restrict__ in, const TG* __restrict__ dout,
                                               const int* __restrict__ nbr, int ld, int n_out, int cin, int cout,
                                               int rows_per_cta, float* __restrict__ dw) {
  __shared__ float As[WG_ROWS][64 + 4];
  __shared__ float Gs[WG_ROWS][64 + 4];
  __shared__ int s_idx[WG_ROWS];
  const int k = blockIdx.z;
  const int tiles_co = (cout + 63) / 64;
  const int ci0 = (blockIdx.y / tiles_co) * 64, co0 = (blockIdx.y % tiles_co) * 64;
  const int r_begin = blockIdx.x * rows_per_cta, r_end = min(r_begin + rows_per_cta, n_out);
  const int tid = threadIdx.x, tx = tid & 15, ty = tid >> 4;  // 4 x 4 outputs per thread
  float acc[4][4];
#pragma unroll
  for (int i = 0; i < 4; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  for (int r0 = r_begin; r0 < r_end; r0 += WG_ROWS) {
    __syncthreads();
    if (tid < WG_ROWS) {
      const int o = r0 + tid;
      s_idx[tid] = (o < r_end) ? (nbr ? nbr[(long long)k * ld + o] : o) : -1;
    }
    __syncthreads();
    // 16 rows x 64 channels of A and G: 1024 elements each, 4 per thread
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = tid + u * 256, rr = e >> 6, c = e & 63;
      const int src = s_idx[rr];
      const int o = r0 + rr;
      As[rr][c] = (src >= 0 && ci0 + c < cin) ? ldf<TA>(in + (long long)src * cin + ci0 + c) : 0.f;
      Gs[rr][c] = (src >= 0 && o < r_end && co0 + c < cout) ? ldf<TG>(dout + (long long)o * cout + co0 + c) : 0.f;
    }
    __syncthreads();
#pragma unroll
    for (int rr = 0; rr < WG_ROWS; ++rr) {
      const float4 a = *reinterpret_cast<const float4*>(&As[rr][ty * 4]);
      const float4 g = *reinterpret_cast<const float4*>(&Gs[rr][tx * 4]);
      const float av[4] = {a.x, a.y, a.z, a.w}, gv[4] = {g.x, g.y, g.z, g.w};
#pragma unroll
      for (int i = 0; i < 4; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(av[i], gv[j], acc[i][j]);
    }
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const int ci = ci0 + ty * 4 + i;
    if (ci >= cin) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int co = co0 + tx * 4 + j;
      if (co < cout && acc[i][j] != 0.f) atomicAdd(dw + ((long long)k * cin + ci) * cout + co, acc[i][j]);
    }
  }
}

// bf16 tensor-core variant (legacy mma.sync path through nvcuda::wmma; A is consumed transposed straight from the
// row-major gathered tile: element (m = ci, k = row) of A^T lives at As[row][ci] = column-major with ld = WGT_LD).
// CTA = 4 warps, 64 x 64 tile of dW[k]; each warp 32 x 32 (2 x 2 fragments); 64 gathered rows per iteration.
#define WGT_ROWS 64
#define WGT_LD 72  // 64 + 8 bf16 of padding: ldmatrix rows land in different banks

__global__ void __launch_bounds__(128) k_wgrad_wmma(const __nv_bfloat16* __restrict__ in,
                                                    const __nv_bfloat16* __restrict__ dout, const int* __restrict__ nbr,
                                                    int ld, int n_out, int cin, int cout, int rows_per_cta,
                                                    float* __restrict__ dw) {
  using namespace nvcuda;
  __shared__ __align__(32) __nv_bfloat16 As[WGT_ROWS * WGT_LD];
  __shared__ __align__(32) __nv_bfloat16 Gs[WGT_ROWS * WGT_LD];
  __shared__ __align__(32) float Cs[4][16 * 16];
  __shared__ int s_idx[WGT_ROWS];
  const int k = blockIdx.z;
  const int tiles_co = cout / 64;
  const int ci0 = (blockIdx.y / tiles_co) * 64, co0 = (blockIdx.y % tiles_co) * 64;
  const int r_begin = blockIdx.x * rows_per_cta, r_end = min(r_begin + rows_per_cta, n_out);
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int wm = (warp >> 1) * 32, wn = (warp & 1) * 32;
  wmma::fragment<wmma::accumulator, 16, 16, 16, float> acc[2][2];
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) wmma::fill_fragment(acc[i][j], 0.f);
  for (int r0 = r_begin; r0 < r_end; r0 += WGT_ROWS) {
    __syncthreads();
    if (tid < WGT_ROWS) {
      const int o = r0 + tid;
      s_idx[tid] = (o < r_end) ? (nbr ? nbr[(long long)k * ld + o] : o) : -1;
    }
    __syncthreads();
    // 64 rows x 8 chunks of 16 B for each operand: 512 chunks, 4 per thread
#pragma unroll
    for (int u = 0; u < 4; ++u) {
      const int e = tid + u * 128, rr = e >> 3, ch = e & 7;
      const int src = s_idx[rr];
      uint4 a = make_uint4(0u, 0u, 0u, 0u), g = make_uint4(0u, 0u, 0u, 0u);
      if (src >= 0) {
        a = __ldg(reinterpret_cast<const uint4*>(in + (long long)src * cin + ci0) + ch);
        g = __ldg(reinterpret_cast<const uint4*>(dout + (long long)(r0 + rr) * cout + co0) + ch);
      }
      *reinterpret_cast<uint4*>(&As[rr * WGT_LD + ch * 8]) = a;
      *reinterpret_cast<uint4*>(&Gs[rr * WGT_LD + ch * 8]) = g;
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < WGT_ROWS; kk += 16) {
      wmma::fragment<wmma::matrix_a, 16, 16, 16, __nv_bfloat16, wmma::col_major> fa[2];
      wmma::fragment<wmma::matrix_b, 16, 16, 16, __nv_bfloat16, wmma::row_major> fb[2];
#pragma unroll
      for (int i = 0; i < 2; ++i) wmma::load_matrix_sync(fa[i], &As[kk * WGT_LD + wm + i * 16], WGT_LD);
#pragma unroll
      for (int j = 0; j < 2; ++j) wmma::load_matrix_sync(fb[j], &Gs[kk * WGT_LD + wn + j * 16], WGT_LD);
#pragma unroll
      for (int i = 0; i < 2; ++i)
#pragma unroll
        for (int j = 0; j < 2; ++j) wmma::mma_sync(acc[i][j], fa[i], fb[j], acc[i][j]);
    }
  }
  // fragments -> smem -> fp32 atomics into dW[k]
#pragma unroll
  for (int i = 0; i < 2; ++i)
#pragma unroll
    for (int j = 0; j < 2; ++j) {
      wmma::store_matrix_sync(Cs[warp], acc[i][j], 16, wmma::mem_row_major);
      __syncwarp();
      for (int e = lane; e < 256; e += 32) {
        const int ci = ci0 + wm + i * 16 + (e >> 4), co = co0 + wn + j * 16 + (e & 15);
        const float v = Cs[warp][e];
        if (v != 0.f) atomicAdd(dw + ((long long)k * cin + ci) * cout + co, v);
      }
      __syncwarp();
    }
}

// Narrow-input weight gradient (the stem: cin = 3, cout = 64, 27 offsets over ~92 k output rows per scene).
// CTA = 256 threads = 64 output channels x 4 groups of kernel offsets (k = grp, grp + 4, ...: up to 7 per group);
// a thread keeps its 7 x 4 partial sums in registers.  Per batch of 32 output rows the CTA stages the gathered
// inputs as s_a[row][grp][7][4] (zero where the neighbour is absent) and dout as s_g[row][64]; the inner loop is
// 7 broadcast LDS.128 + 1 LDS per 28 FMAs.  One fp32 atomic per (k, ci, co) and CTA at the end.
#define WN_ROWS 32
template <typename TA, typename TG>
__global__ void __launch_bounds__(256) k_wgrad_narrow(const TA* __restrict__ in, const TG* __restrict__ dout,
                                                      const int* __restrict__ nbr, int ld, int n_out, int cin, int cout,
                                                      int kvol, int rows_per_cta, float* __restrict__ dw) {
  __shared__ __align__(16) float s_a[WN_ROWS][4][7][4];
  __shared__ float s_g[WN_ROWS][64];
  const int tid = threadIdx.x, co = tid & 63, grp = tid >> 6;
  const int r_begin = blockIdx.x * rows_per_cta, r_end = min(r_begin + rows_per_cta, n_out);
  float acc[7][4];
#pragma unroll
  for (int i = 0; i < 7; ++i)
#pragma unroll
    for (int c = 0; c < 4; ++c) acc[i][c] = 0.f;
  for (int e = tid; e < WN_ROWS * 4 * 7 * 4; e += 256) (&s_a[0][0][0][0])[e] = 0.f;  // slots of unused (k, ci) stay zero
  for (int r0 = r_begin; r0 < r_end; r0 += WN_ROWS) {
    __syncthreads();
    for (int p = tid; p < kvol * WN_ROWS; p += 256) {
      const int k = p / WN_ROWS, rr = p % WN_ROWS, o = r0 + rr;
      const int src = (o < r_end) ? (nbr ? __ldg(nbr + (long long)k * ld + o) : o) : -1;
      float* dst = &s_a[rr][k & 3][k >> 2][0];
      for (int c = 0; c < cin; ++c) dst[c] = (src >= 0) ? ldf<TA>(in + (long long)src * cin + c) : 0.f;
    }
    for (int e = tid; e < WN_ROWS * 64; e += 256) {
      const int rr = e >> 6, c = e & 63, o = r0 + rr;
      s_g[rr][c] = (o < r_end && c < cout) ? ldf<TG>(dout + (long long)o * cout + c) : 0.f;
    }
    __syncthreads();
#pragma unroll 4
    for (int rr = 0; rr < WN_ROWS; ++rr) {
      const float g = s_g[rr][co];
#pragma unroll
      for (int i = 0; i < 7; ++i) {
        const float4 a = *reinterpret_cast<const float4*>(&s_a[rr][grp][i][0]);
        acc[i][0] = fmaf(a.x, g, acc[i][0]);
        acc[i][1] = fmaf(a.y, g, acc[i][1]);
        acc[i][2] = fmaf(a.z, g, acc[i][2]);
        acc[i][3] = fmaf(a.w, g, acc[i][3]);
      }
    }
  }
  if (co < cout) {
#pragma unroll
    for (int i = 0; i < 7; ++i) {
      const int k = grp + 4 * i;
      if (k >= kvol) continue;
#pragma unroll
      for (int c = 0; c < 4; ++c)
        if (c < cin && acc[i][c] != 0.f) atomicAdd(dw + ((long long)k * cin + c) * cout + co, acc[i][c]);
    }
  }
}

// bf16 operand layouts of a convolution kernel w [K, Cin, Cout] fp32, both in one pass:
//   wt_fwd [K, Cout, Cin] = w[k]^T          (forward: "W_k^T" rows of the tcgen05 B operand)
//   wt_bwd [K, Cin, Cout] = w[K-1-k]        (input gradient: the transposed map pairs offset k with K-1-k)
__global__ void __launch_bounds__(256) k_pack_conv_weights(const float* __restrict__ w, int kvol, int cin, int cout,
                                                           __nv_bfloat16* __restrict__ wt_fwd,
                                                           __nv_bfloat16* __restrict__ wt_bwd) {
  __shared__ float tile[32][33];
  const int k = blockIdx.z, ci0 = blockIdx.y * 32, co0 = blockIdx.x * 32;
  const int tx = threadIdx.x & 31, ty = threadIdx.x >> 5;
  const float* wk = w + (long long)k * cin * cout;
  __nv_bfloat16* bk = wt_bwd ? wt_bwd + (long long)(kvol - 1 - k) * cin * cout : nullptr;
  for (int r = ty; r < 32; r += 8) {
    const int ci = ci0 + r, co = co0 + tx;
    float v = 0.f;
    if (ci < cin && co < cout) {
      v = wk[(long long)ci * cout + co];
      if (bk) bk[(long long)ci * cout + co] = __float2bfloat16(v);
    }
    tile[r][tx] = v;
  }
  __syncthreads();
  if (wt_fwd) {
    __nv_bfloat16* fk = wt_fwd + (long long)k * cin * cout;
    for (int r = ty; r < 32; r += 8) {
      const int co = co0 + r, ci = ci0 + tx;
      if (ci < cin && co < cout) fk[(long long)co * cin + ci] = __float2bfloat16(tile[tx][r]);
    }
  }
}

// max-pool backward: dx[arg[o, c], c] += dy[o, c]
template <typename T>
__global__ void k_maxpool_bwd(const T* __restrict__ dy, const int* __restrict__ arg, long long n_out, int c,
                              float* __restrict__ dx) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n_out * c) return;
  const int r = arg[t];
  if (r >= 0) atomicAdd(dx + (long long)r * c + (int)(t % c), ldf<T>(dy + t));
}

// ---------------------------------------------------------------------------
// Fcaf3DAssigner (one scene per call).  points [P,3] (all levels concatenated, level_off[L+1]),
// boxes [B,7] = (gravity centre, size, yaw), labels_in [B].
// ---------------------------------------------------------------------------
struct FaceDist {
  float d[6];
};

// face distances of point p to box b (fcaf3d_neck_with_head.py:417-431), same operation order
__device__ __forceinline__ FaceDist face_distances(const float* __restrict__ p, const float* __restrict__ b) {
  const float sx = p[0] - b[0], sy = p[1] - b[1], sz = p[2] - b[2];
  // rotation_3d_in_axis(shift, -yaw, axis=2): [x', y'] = [x cos + y sin', ...] with angle = -yaw
  const float ang = -b[6];
  const float rs = sinf(ang), rc = cosf(ang);
  const float rx = sx * rc + sy * rs;       // einsum with rot_mat_T = [[c, -s, 0], [s, c, 0], [0, 0, 1]]
  const float ry = sx * (-rs) + sy * rc;
  const float cx = b[0] + rx, cy = b[1] + ry, cz = b[2] + sz;
  FaceDist f;
  f.d[0] = cx - b[0] + b[3] / 2;
  f.d[1] = b[0] + b[3] / 2 - cx;
  f.d[2] = cy - b[1] + b[4] / 2;
  f.d[3] = b[1] + b[4] / 2 - cy;
  f.d[4] = cz - b[2] + b[5] / 2;
  f.d[5] = b[2] + b[5] / 2 - cz;
  return f;
}
__device__ __forceinline__ bool inside_box(const FaceDist& f) {
  float m = f.d[0];
#pragma unroll
  for (int i = 1; i < 6; ++i) m = fminf(m, f.d[i]);
  return m > 0.f;
}
// compute_centerness (:377-384)
__device__ __forceinline__ float centerness_of(const FaceDist& f) {
  const float x = fminf(f.d[0], f.d[1]) / fmaxf(f.d[0], f.d[1]);
  const float y = fminf(f.d[2], f.d[3]) / fmaxf(f.d[2], f.d[3]);
  const float z = fminf(f.d[4], f.d[5]) / fmaxf(f.d[4], f.d[5]);
  return sqrtf(x * y * z);
}

// pass 1: positives per (level, box)
__global__ void k_assign_count(const float* __restrict__ points, int P, const int* __restrict__ level_off, int L,
                               const float* __restrict__ boxes, int B, int* __restrict__ counts) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int lvl = 0;
  while (lvl + 1 < L && p >= level_off[lvl + 1]) ++lvl;
  for (int b = 0; b < B; ++b) {
    const FaceDist f = face_distances(points + 3ll * p, boxes + 7ll * b);
    if (inside_box(f)) atomicAdd(&counts[lvl * B + b], 1);
  }
}

// pass 2: best level per box (:437-446) and the (topk+1)-th largest centerness among its valid points (:452-453).
// One CTA per box; exact selection by a 3-pass radix select over the non-negative centerness bit patterns.
__global__ void __launch_bounds__(256) k_assign_select(const float* __restrict__ points, int P,
                                                       const int* __restrict__ level_off, int L,
                                                       const float* __restrict__ boxes, int B,
                                                       const int* __restrict__ counts, int limit, int topk,
                                                       int* __restrict__ best_level, float* __restrict__ thr) {
  __shared__ unsigned int hist[2048];
  __shared__ unsigned int s_prefix, s_need;
  __shared__ int s_best;
  const int b = blockIdx.x;
  if (threadIdx.x == 0) {
    int first_low = -1;
    for (int l = 0; l < L; ++l)
      if (counts[l * B + b] < limit) { first_low = l; break; }
    int best = (first_low < 0) ? L - 1 : max(first_low - 1, 0);
    s_best = best;
    best_level[b] = best;
  }
  __syncthreads();
  const int lvl = s_best;
  const int beg = level_off[lvl], end = level_off[lvl + 1];
  // torch.topk(centerness, min(topk + 1, P), dim=0).values[-1] over ALL P points, invalid entries being -1
  const int kk = min(topk + 1, P);
  const int n_valid = counts[lvl * B + b];
  if (n_valid < kk) {  // the kk-th largest is an invalid entry
    if (threadIdx.x == 0) thr[b] = -1.f;
    return;
  }
  unsigned int prefix = 0, need = (unsigned int)kk;
  const int shifts[3] = {21, 10, 0}, widths[3] = {11, 11, 10};
  for (int pass = 0; pass < 3; ++pass) {
    const int shift = shifts[pass], bins = 1 << widths[pass];
    const unsigned int hi_mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + widths[pass]));
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int p = beg + threadIdx.x; p < end; p += blockDim.x) {
      const FaceDist f = face_distances(points + 3ll * p, boxes + 7ll * b);
      if (!inside_box(f)) continue;
      const unsigned int u = __float_as_uint(centerness_of(f));  // >= 0: bit pattern order == value order
      if ((u & hi_mask) == prefix) atomicAdd(&hist[(u >> shift) & (bins - 1)], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned int acc = 0;
      int d = bins - 1;
      for (; d > 0; --d) {
        if (acc + hist[d] >= need) break;
        acc += hist[d];
      }
      s_prefix = prefix | ((unsigned int)d << shift);
      s_need = need - acc;
    }
    __syncthreads();
    prefix = s_prefix;
    need = s_need;
    __syncthreads();
  }
  if (threadIdx.x == 0) thr[b] = __uint_as_float(prefix);
}

// pass 3: per point, the valid box of minimal volume (first index on ties) (:455-466)
__global__ void k_assign_final(const float* __restrict__ points, int P, const int* __restrict__ level_off, int L,
                               const float* __restrict__ boxes, int B, const long long* __restrict__ labels_in,
                               const int* __restrict__ best_level, const float* __restrict__ thr,
                               float* __restrict__ centerness_t, float* __restrict__ box_t,
                               long long* __restrict__ labels_out) {
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  int lvl = 0;
  while (lvl + 1 < L && p >= level_off[lvl + 1]) ++lvl;
  const float float_max = 1e8f;
  float best_vol = float_max;
  int best = 0;
  for (int b = 0; b < B; ++b) {
    const float* bx = boxes + 7ll * b;
    const FaceDist f = face_distances(points + 3ll * p, bx);
    float vol = float_max;
    if (inside_box(f) && best_level[b] == lvl && centerness_of(f) > thr[b]) vol = bx[3] * bx[4] * bx[5];
    if (vol < best_vol) {
      best_vol = vol;
      best = b;
    }
  }
  const float* bx = boxes + 7ll * best;
  const FaceDist f = face_distances(points + 3ll * p, bx);
  centerness_t[p] = centerness_of(f);
#pragma unroll
  for (int u = 0; u < 7; ++u) box_t[7ll * p + u] = bx[u];
  labels_out[p] = (best_vol == float_max) ? -1 : labels_in[best];
}

// ---------------------------------------------------------------------------
// Batched Fcaf3DAssigner: all scenes of the batch in three launches.  Points are LEVEL-major (the rows of each level's
// coordinate map, scenes grouped inside a level): segment j = level * S + scene covers rows [seg[2j], seg[2j+1]).
// Boxes of all scenes are concatenated, box_off[S+1].  Same arithmetic per (point, box) as the per-scene kernels.
// ---------------------------------------------------------------------------
#define AB_MAX_SEG 256
__device__ __forceinline__ int find_segment(const int* s_seg, int nseg, int p) {
  for (int j = 0; j < nseg; ++j)
    if (p >= s_seg[2 * j] && p < s_seg[2 * j + 1]) return j;
  return -1;
}

__global__ void k_assignb_count(const float* __restrict__ points, int P, const int* __restrict__ seg, int S, int L,
                                const float* __restrict__ boxes, const int* __restrict__ box_off, int Btot,
                                int* __restrict__ counts, int* __restrict__ scene_of) {
  __shared__ int s_seg[2 * AB_MAX_SEG];
  for (int i = threadIdx.x; i < 2 * S * L; i += blockDim.x) s_seg[i] = seg[i];
  __syncthreads();
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int j = find_segment(s_seg, S * L, p);
  if (j < 0) {
    scene_of[p] = -1;
    return;
  }
  const int lvl = j / S, sc = j - lvl * S;
  scene_of[p] = sc;
  for (int b = box_off[sc]; b < box_off[sc + 1]; ++b) {
    const FaceDist f = face_distances(points + 3ll * p, boxes + 7ll * b);
    if (inside_box(f)) atomicAdd(&counts[lvl * Btot + b], 1);
  }
}

__global__ void __launch_bounds__(256) k_assignb_select(const float* __restrict__ points, const int* __restrict__ seg,
                                                        int S, int L, const float* __restrict__ boxes,
                                                        const int* __restrict__ box_off, int Btot,
                                                        const int* __restrict__ counts, int limit, int topk,
                                                        int* __restrict__ best_level, float* __restrict__ thr) {
  __shared__ unsigned int hist[2048];
  __shared__ unsigned int s_prefix, s_need;
  __shared__ int s_best, s_scene, s_np;
  const int b = blockIdx.x;
  if (threadIdx.x == 0) {
    int sc = 0;
    while (sc + 1 < S && b >= box_off[sc + 1]) ++sc;
    int first_low = -1;
    for (int l = 0; l < L; ++l)
      if (counts[l * Btot + b] < limit) { first_low = l; break; }
    const int best = (first_low < 0) ? L - 1 : max(first_low - 1, 0);
    int np = 0;
    for (int l = 0; l < L; ++l) np += seg[2 * (l * S + sc) + 1] - seg[2 * (l * S + sc)];
    s_best = best;
    s_scene = sc;
    s_np = np;
    best_level[b] = best;
  }
  __syncthreads();
  const int lvl = s_best, sc = s_scene;
  const int beg = seg[2 * (lvl * S + sc)], end = seg[2 * (lvl * S + sc) + 1];
  const int kk = min(topk + 1, s_np);  // torch.topk(..., min(topk + 1, len(points))) over the scene's points
  const int n_valid = counts[lvl * Btot + b];
  if (n_valid < kk) {
    if (threadIdx.x == 0) thr[b] = -1.f;
    return;
  }
  unsigned int prefix = 0, need = (unsigned int)kk;
  const int shifts[3] = {21, 10, 0}, widths[3] = {11, 11, 10};
  for (int pass = 0; pass < 3; ++pass) {
    const int shift = shifts[pass], bins = 1 << widths[pass];
    const unsigned int hi_mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + widths[pass]));
    for (int i = threadIdx.x; i < 2048; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int p = beg + threadIdx.x; p < end; p += blockDim.x) {
      const FaceDist f = face_distances(points + 3ll * p, boxes + 7ll * b);
      if (!inside_box(f)) continue;
      const unsigned int u = __float_as_uint(centerness_of(f));
      if ((u & hi_mask) == prefix) atomicAdd(&hist[(u >> shift) & (bins - 1)], 1u);
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned int acc = 0;
      int d = bins - 1;
      for (; d > 0; --d) {
        if (acc + hist[d] >= need) break;
        acc += hist[d];
      }
      s_prefix = prefix | ((unsigned int)d << shift);
      s_need = need - acc;
    }
    __syncthreads();
    prefix = s_prefix;
    need = s_need;
    __syncthreads();
  }
  if (threadIdx.x == 0) thr[b] = __uint_as_float(prefix);
}

__global__ void k_assignb_final(const float* __restrict__ points, int P, const int* __restrict__ seg, int S, int L,
                                const float* __restrict__ boxes, const int* __restrict__ box_off,
                                const long long* __restrict__ labels_in, const int* __restrict__ best_level,
                                const float* __restrict__ thr, float* __restrict__ centerness_t,
                                float* __restrict__ box_t, long long* __restrict__ labels_out,
                                int* __restrict__ n_pos) {
  __shared__ int s_seg[2 * AB_MAX_SEG];
  for (int i = threadIdx.x; i < 2 * S * L; i += blockDim.x) s_seg[i] = seg[i];
  __syncthreads();
  const int p = blockIdx.x * blockDim.x + threadIdx.x;
  if (p >= P) return;
  const int j = find_segment(s_seg, S * L, p);
  const int lvl = j < 0 ? 0 : j / S, sc = j < 0 ? 0 : j - lvl * S;
  const int b0 = j < 0 ? 0 : box_off[sc], b1 = j < 0 ? 0 : box_off[sc + 1];
  const float float_max = 1e8f;
  float best_vol = float_max;
  int best = b0;
  for (int b = b0; b < b1; ++b) {
    const float* bx = boxes + 7ll * b;
    const FaceDist f = face_distances(points + 3ll * p, bx);
    float vol = float_max;
    if (inside_box(f) && best_level[b] == lvl && centerness_of(f) > thr[b]) vol = bx[3] * bx[4] * bx[5];
    if (vol < best_vol) {
      best_vol = vol;
      best = b;
    }
  }
  if (b1 == b0) {  // scene without boxes: everything is background
    centerness_t[p] = 0.f;
#pragma unroll
    for (int u = 0; u < 7; ++u) box_t[7ll * p + u] = 0.f;
    labels_out[p] = -1;
    return;
  }
  const float* bx = boxes + 7ll * best;
  const FaceDist f = face_distances(points + 3ll * p, bx);
  centerness_t[p] = centerness_of(f);
#pragma unroll
  for (int u = 0; u < 7; ++u) box_t[7ll * p + u] = bx[u];
  const bool pos = best_vol != float_max;
  labels_out[p] = pos ? labels_in[best] : -1;
  if (pos) atomicAdd(&n_pos[sc], 1);
}

// ---------------------------------------------------------------------------
// Sigmoid focal loss, mmcv semantics: label outside [0, C) (e.g. -1) marks an all-negative row.
// loss_elem[n, c] and grad[n, c] (d loss_elem / d logit); the caller reduces / scales.
// ---------------------------------------------------------------------------
__global__ void k_focal(const float* __restrict__ logits, const long long* __restrict__ labels, long long n, int c,
                        float gamma, float alpha, float* __restrict__ loss, float* __restrict__ grad) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c) return;
  const long long row = t / c;
  const int cls = (int)(t - row * c);
  const float x = logits[t];
  const float p = 1.f / (1.f + expf(-x));
  const bool pos = labels[row] == cls;
  const float FLT_MIN_ = 1.17549435e-38f;
  if (pos) {
    const float w = alpha * powf(1.f - p, gamma);
    if (loss) loss[t] = -w * logf(fmaxf(p, FLT_MIN_));
    // d/dx [-a (1-p)^g log p] = a (1-p)^g (g p log p - (1 - p))
    if (grad) grad[t] = -alpha * powf(1.f - p, gamma) * ((1.f - p) - gamma * p * logf(fmaxf(p, FLT_MIN_)));
  } else {
    const float w = (1.f - alpha) * powf(p, gamma);
    if (loss) loss[t] = -w * logf(fmaxf(1.f - p, FLT_MIN_));
    if (grad) grad[t] = (1.f - alpha) * powf(p, gamma) * (p - gamma * (1.f - p) * logf(fmaxf(1.f - p, FLT_MIN_)));
  }
}

// ---------------------------------------------------------------------------
// Axis-aligned 3-D IoU of aligned pairs (centre, size) boxes + gradient w.r.t. the prediction.
// iou = overlap / max(area1 + area2 - overlap, 1e-6);  loss_i = 1 - iou_i.
// ---------------------------------------------------------------------------
__global__ void k_aa_iou(const float* __restrict__ pred, const float* __restrict__ tgt, int n, int stride,
                         float* __restrict__ iou, float* __restrict__ dpred) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = pred + (long long)i * stride;
  const float* t = tgt + (long long)i * stride;
  float lo1[3], hi1[3], lo2[3], hi2[3], wh[3];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    lo1[a] = p[a] - p[3 + a] / 2;
    hi1[a] = p[a] + p[3 + a] / 2;
    lo2[a] = t[a] - t[3 + a] / 2;
    hi2[a] = t[a] + t[3 + a] / 2;
    wh[a] = fmaxf(fminf(hi1[a], hi2[a]) - fmaxf(lo1[a], lo2[a]), 0.f);
  }
  const float area1 = (hi1[0] - lo1[0]) * (hi1[1] - lo1[1]) * (hi1[2] - lo1[2]);
  const float area2 = (hi2[0] - lo2[0]) * (hi2[1] - lo2[1]) * (hi2[2] - lo2[2]);
  const float overlap = wh[0] * wh[1] * wh[2];
  const float uni = area1 + area2 - overlap;
  const float den = fmaxf(uni, 1e-6f);
  iou[i] = overlap / den;
  if (!dpred) return;
  // gradients w.r.t. lo1/hi1 then chain to (centre, size)
  float g_lo[3], g_hi[3];
  const float d_ov = (uni > 1e-6f) ? (den + overlap) / (den * den) : 1.f / den;  // d iou / d overlap (union depends on it)
  const float d_a1 = (uni > 1e-6f) ? -overlap / (den * den) : 0.f;               // d iou / d area1
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const int b = (a + 1) % 3, c = (a + 2) % 3;
    const float dov_dwh = wh[b] * wh[c];
    float dwh_dhi = 0.f, dwh_dlo = 0.f;
    if (fminf(hi1[a], hi2[a]) - fmaxf(lo1[a], lo2[a]) > 0.f) {
      dwh_dhi = (hi1[a] <= hi2[a]) ? 1.f : 0.f;   // torch.min picks the first argument on ties
      dwh_dlo = (lo1[a] >= lo2[a]) ? -1.f : 0.f;  // torch.max likewise
    }
    const float da1_dext = (hi1[b] - lo1[b]) * (hi1[c] - lo1[c]);
    g_hi[a] = d_ov * dov_dwh * dwh_dhi + d_a1 * da1_dext;
    g_lo[a] = d_ov * dov_dwh * dwh_dlo - d_a1 * da1_dext;
  }
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    dpred[(long long)i * stride + a] = g_lo[a] + g_hi[a];
    dpred[(long long)i * stride + 3 + a] = 0.5f * (g_hi[a] - g_lo[a]);
  }
  for (int a = 6; a < stride; ++a) dpred[(long long)i * stride + a] = 0.f;
}

// ---------------------------------------------------------------------------
// sort_v: indices of the valid polygon vertices (of 24 candidates, mean-centred) in counter-clockwise angular
// order, the first index repeated, then padded with an invalid intersection slot (index >= 8).  One thread per box.
// ---------------------------------------------------------------------------
__global__ void k_sort_vertices(const float* __restrict__ vertices, const uint8_t* __restrict__ mask,
                                const int* __restrict__ num_valid, int n, int* __restrict__ idx) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* v = vertices + (long long)i * 48;
  const uint8_t* m = mask + (long long)i * 24;
  int pad = 8;
  for (int j = 8; j < 24; ++j)
    if (!m[j]) { pad = j; break; }
  int* out = idx + (long long)i * 9;
  const int nv = num_valid[i];
  if (nv < 3) {
    for (int j = 0; j < 9; ++j) out[j] = pad;
    return;
  }
  float ang[24];
  for (int j = 0; j < 24; ++j) ang[j] = m[j] ? atan2f(v[2 * j + 1], v[2 * j]) : 1e9f;
  // selection sort of at most 8 valid vertices by ascending angle (ties: lower index first)
  unsigned int used = 0;
  int cnt = 0;
  for (; cnt < nv && cnt < 8; ++cnt) {
    int best = -1;
    for (int j = 0; j < 24; ++j)
      if (m[j] && !((used >> j) & 1u) && (best < 0 || ang[j] < ang[best])) best = j;
    used |= 1u << best;
    out[cnt] = best;
  }
  out[cnt] = out[0];
  for (int j = cnt + 1; j < 9; ++j) out[j] = pad;
}

int b2_wgrad_tc_dispatch(const void* in, const void* dout, const int* nbr, long long ld, long long n_out, int cin,
                         int cout, int kvol, float* dw, cudaStream_t st);

extern "C" {

// dw [kvol, cin, cout] fp32 must be zero-initialised by the caller; a_dtype / g_dtype: 0 = f32, 1 = bf16
int b2_conv_wgrad(const void* in, int a_dtype, const void* dout, int g_dtype, const int* nbr, long long ld,
                  long long n_out, int cin, int cout, int kvol, float* dw, void* stream) {
  B2_CHECK_ARG(nbr != nullptr || kvol == 1, "identity map requires kernel volume 1");
  if (n_out == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int tiles = ((cin + 63) / 64) * ((cout + 63) / 64);
  // enough row chunks to fill the GPU, at least 256 rows each
  long long chunks = (148ll * 4 + (long long)tiles * kvol - 1) / ((long long)tiles * kvol);
  if (chunks < 1) chunks = 1;
  long long rows_per = (n_out + chunks - 1) / chunks;
  if (rows_per < 256) rows_per = 256;
  rows_per = (rows_per + WG_ROWS - 1) / WG_ROWS * WG_ROWS;
  dim3 grid((unsigned)((n_out + rows_per - 1) / rows_per), tiles, kvol);
  if (cin <= 4 && cout <= 64 && kvol <= 28) {
    long long rp = b2_div_up(b2_div_up(n_out, 148ll * 4), (long long)WN_ROWS) * WN_ROWS;
    if (rp < 8 * WN_ROWS) rp = 8 * WN_ROWS;
    const unsigned gx = (unsigned)b2_div_up(n_out, rp);
#define WN_GO(TA, TG) \
  k_wgrad_narrow<TA, TG><<<gx, 256, 0, st>>>((const TA*)in, (const TG*)dout, nbr, (int)ld, (int)n_out, cin, cout, kvol, (int)rp, dw)
    if (a_dtype == 0 && g_dtype == 0) WN_GO(float, float);
    else if (a_dtype == 1 && g_dtype == 1) WN_GO(__nv_bfloat16, __nv_bfloat16);
    else if (a_dtype == 0) WN_GO(float, __nv_bfloat16);
    else WN_GO(__nv_bfloat16, float);
#undef WN_GO
    B2_LAUNCH_CHECK();
    return B2_OK;
  }
  // tcgen05 path (wgrad_tc.cu); B2_WGRAD_WMMA=1 keeps the legacy mma.sync kernel below for A/B runs
  static const bool legacy_wmma = (getenv("B2_WGRAD_WMMA") && atoi(getenv("B2_WGRAD_WMMA")) == 1);
  const int c64n = cin / 64;
  if (!legacy_wmma && a_dtype == 1 && g_dtype == 1 && cin % 64 == 0 && cout % 64 == 0 && (c64n & (c64n - 1)) == 0 &&
      (nbr == nullptr || ld % 4 == 0))
    return b2_wgrad_tc_dispatch(in, dout, nbr, ld, n_out, cin, cout, kvol, dw, st);
  if (a_dtype == 1 && g_dtype == 1 && cin % 64 == 0 && cout % 64 == 0) {
    long long c2 = (148ll * 8 + (long long)tiles * kvol - 1) / ((long long)tiles * kvol);
    long long rp = (n_out + (c2 < 1 ? 1 : c2) - 1) / (c2 < 1 ? 1 : c2);
    if (rp < 1024) rp = 1024;
    rp = (rp + WGT_ROWS - 1) / WGT_ROWS * WGT_ROWS;
    dim3 g2((unsigned)((n_out + rp - 1) / rp), tiles, kvol);
    k_wgrad_wmma<<<g2, 128, 0, st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)dout, nbr, (int)ld, (int)n_out, cin,
                                     cout, (int)rp, dw);
    B2_LAUNCH_CHECK();
    return B2_OK;
  }
#define WG_GO(TA, TG) \
  k_wgrad<TA, TG><<<grid, 256, 0, st>>>((const TA*)in, (const TG*)dout, nbr, (int)ld, (int)n_out, cin, cout, (int)rows_per, dw)
  if (a_dtype == 0 && g_dtype == 0) WG_GO(float, float);
  else if (a_dtype == 1 && g_dtype == 1) WG_GO(__nv_bfloat16, __nv_bfloat16);
  else if (a_dtype == 0) WG_GO(float, __nv_bfloat16);
  else WG_GO(__nv_bfloat16, float);
#undef WG_GO
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// bf16 tcgen05 operand layouts of w [kvol, cin, cout] fp32: wt_fwd [kvol, cout, cin] (= w[k]^T) and
// wt_bwd [kvol, cin, cout] (= w[kvol-1-k]); either output may be NULL
int b2_pack_conv_weights(const float* w, int kvol, int cin, int cout, void* wt_fwd, void* wt_bwd, void* stream) {
  B2_CHECK_ARG(kvol > 0 && cin > 0 && cout > 0, "empty kernel");
  dim3 grid(b2_div_up(cout, 32), b2_div_up(cin, 32), kvol);
  k_pack_conv_weights<<<grid, 256, 0, (cudaStream_t)stream>>>(w, kvol, cin, cout, (__nv_bfloat16*)wt_fwd,
                                                              (__nv_bfloat16*)wt_bwd);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// dx [n_in, c] fp32, zero-initialised by the caller
int b2_maxpool_bwd(const void* dy, int dtype, const int* argmax, long long n_out, int c, float* dx, void* stream) {
  if (n_out == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int G = b2_div_up(n_out * c, 256);
  if (dtype == 0) k_maxpool_bwd<float><<<G, 256, 0, st>>>((const float*)dy, argmax, n_out, c, dx);
  else k_maxpool_bwd<__nv_bfloat16><<<G, 256, 0, st>>>((const __nv_bfloat16*)dy, argmax, n_out, c, dx);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// scratch: counts int32 [L*B] (zeroed inside), best_level int32 [B], thr f32 [B]
int b2_fcaf3d_assign(const float* points, long long P, const int* level_off, int L, const float* boxes, int B,
                     const long long* labels_in, int limit, int topk, int* counts, int* best_level, float* thr,
                     float* centerness_t, float* box_t, long long* labels_out, void* stream) {
  B2_CHECK_ARG(L >= 1 && B >= 0, "bad assigner shape");
  cudaStream_t st = (cudaStream_t)stream;
  if (P == 0) return B2_OK;
  if (B == 0) {
    B2_CUDA(cudaMemsetAsync(labels_out, 0xFF, (size_t)P * 8, st));  // -1
    B2_CUDA(cudaMemsetAsync(centerness_t, 0, (size_t)P * 4, st));
    B2_CUDA(cudaMemsetAsync(box_t, 0, (size_t)P * 28, st));
    return B2_OK;
  }
  B2_CUDA(cudaMemsetAsync(counts, 0, (size_t)L * B * 4, st));
  k_assign_count<<<b2_div_up(P, 128), 128, 0, st>>>(points, (int)P, level_off, L, boxes, B, counts);
  B2_LAUNCH_CHECK();
  k_assign_select<<<B, 256, 0, st>>>(points, (int)P, level_off, L, boxes, B, counts, limit, topk, best_level, thr);
  B2_LAUNCH_CHECK();
  k_assign_final<<<b2_div_up(P, 128), 128, 0, st>>>(points, (int)P, level_off, L, boxes, B, labels_in, best_level, thr,
                                                   centerness_t, box_t, labels_out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// batched assigner: points level-major, seg int32 [L*S][2] (row ranges, index level*S+scene), boxes of all scenes
// concatenated with box_off int32 [S+1].  Scratch: counts int32 [L*Btot] (zeroed inside), best_level [Btot], thr [Btot].
// Outputs as b2_fcaf3d_assign plus scene_of int32 [P] and n_pos int32 [S] (positives per scene, zeroed inside).
int b2_fcaf3d_assign_batched(const float* points, long long P, const int* seg, int S, int L, const float* boxes,
                             const int* box_off, int Btot, const long long* labels_in, int limit, int topk, int* counts,
                             int* best_level, float* thr, float* centerness_t, float* box_t, long long* labels_out,
                             int* scene_of, int* n_pos, void* stream) {
  B2_CHECK_ARG(L >= 1 && S >= 1 && S * L <= AB_MAX_SEG && Btot >= 0, "bad batched assigner shape (scenes x levels <= 256)");
  cudaStream_t st = (cudaStream_t)stream;
  B2_CUDA(cudaMemsetAsync(n_pos, 0, (size_t)S * 4, st));
  if (P == 0) return B2_OK;
  if (Btot > 0) B2_CUDA(cudaMemsetAsync(counts, 0, (size_t)L * Btot * 4, st));
  k_assignb_count<<<b2_div_up(P, 128), 128, 0, st>>>(points, (int)P, seg, S, L, boxes, box_off, Btot, counts, scene_of);
  B2_LAUNCH_CHECK();
  if (Btot > 0) {
    k_assignb_select<<<Btot, 256, 0, st>>>(points, seg, S, L, boxes, box_off, Btot, counts, limit, topk, best_level, thr);
    B2_LAUNCH_CHECK();
  }
  k_assignb_final<<<b2_div_up(P, 128), 128, 0, st>>>(points, (int)P, seg, S, L, boxes, box_off, labels_in, best_level, thr,
                                                    centerness_t, box_t, labels_out, n_pos);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// loss / grad: [n, c] element-wise (either may be NULL)
int b2_sigmoid_focal(const float* logits, const long long* labels, long long n, int c, float gamma, float alpha,
                     float* loss, float* grad, void* stream) {
  if (n == 0) return B2_OK;
  k_focal<<<b2_div_up(n * c, 256), 256, 0, (cudaStream_t)stream>>>(logits, labels, n, c, gamma, alpha, loss, grad);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// pred / tgt rows of `stride` floats starting with (cx, cy, cz, dx, dy, dz); iou [n]; dpred [n, stride] or NULL
int b2_aa_iou3d(const float* pred, const float* tgt, long long n, int stride, float* iou, float* dpred, void* stream) {
  B2_CHECK_ARG(stride >= 6, "boxes need at least 6 values");
  if (n == 0) return B2_OK;
  k_aa_iou<<<b2_div_up(n, 128), 128, 0, (cudaStream_t)stream>>>(pred, tgt, (int)n, stride, iou, dpred);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// vertices [n, 24, 2] fp32 (mean-centred), mask [n, 24] uint8, num_valid [n] int32 -> idx [n, 9] int32
int b2_sort_vertices(const float* vertices, const uint8_t* mask, const int* num_valid, long long n, int* idx,
                     void* stream) {
  if (n == 0) return B2_OK;
  k_sort_vertices<<<b2_div_up(n, 128), 128, 0, (cudaStream_t)stream>>>(vertices, mask, num_valid, (int)n, idx);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
