// Kernel-map (neighbour table) construction, CUDA-core sparse convolution (fp32 accumulate),
// max pooling, per-segment channel statistics and the fused affine/activation pass.
//
// Behavioural counterparts in the reference: MinkowskiConvolution / MinkowskiMaxPooling /
// MinkowskiInstanceNorm / MinkowskiBatchNorm / MinkowskiReLU / MinkowskiELU as called from
// mmdet3d/models/backbones/me_resnet.py:18-25,55-78 and
// mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py:49-72,83-85 (MinkowskiEngine v0.5.4, not vendored).
#include "common.cuh"

int b2_optin_smem(const void* func, int bytes, unsigned long long* done);   // tc_map.cu
#include <stdlib.h>

// ---------------------------------------------------------------------------
// Neighbour table: nbr[k * ld + o] = input row at (coord_o + offset_k) or -1.
// Offset order follows ME: x fastest (k = ix + KS*iy + KS*KS*iz); odd kernels are
// centred, even kernels start at the voxel itself.  One CTA = one 128-row output
// tile; tile_mask[tile] bit k says whether any row of the tile has neighbour k.
// ---------------------------------------------------------------------------
template <int KS>
__global__ void __launch_bounds__(128) k_kernel_map(const B2Slot* __restrict__ table, unsigned int cap_mask,
                                                    const int* __restrict__ out_coords, int n_out, int step,
                                                    int* __restrict__ nbr, int ld, unsigned int* __restrict__ tile_mask) {
  constexpr int K = KS * KS * KS;
  const int o = blockIdx.x * 128 + threadIdx.x;
  int res[K];
  if (o < n_out) {
    const int4 c = __ldg(reinterpret_cast<const int4*>(out_coords) + o);
#pragma unroll
    for (int k = 0; k < K; ++k) {
      const int ix = k % KS, iy = (k / KS) % KS, iz = k / (KS * KS);
      const int base = (KS & 1) ? -(KS / 2) : 0;
      const int x = c.y + (ix + base) * step, y = c.z + (iy + base) * step, z = c.w + (iz + base) * step;
      int r = -1;
      if (b2_coord_in_range(c.x, x, y, z)) r = b2_table_find(table, cap_mask, b2_pack_key(c.x, x, y, z));
      res[k] = r;
    }
  } else {
#pragma unroll
    for (int k = 0; k < K; ++k) res[k] = -1;
  }
  unsigned int m = 0;
#pragma unroll
  for (int k = 0; k < K; ++k) {
    nbr[(long long)k * ld + o] = res[k];
    if (__syncthreads_or(res[k] >= 0)) m |= (1u << k);
  }
  if (threadIdx.x == 0) tile_mask[blockIdx.x] = m;
}

// The same table with the offsets spread over blockIdx.y (KPB offsets per CTA): a thread does KPB independent probes
// instead of KS^3 sequential ones, so the small maps of the deep levels (21 - 423 tiles), which were pure probe latency
// (27 dependent L2 round trips, ~21 - 35 us whatever the size), finish in a few microseconds.  tile_mask is zeroed by the
// caller; a CTA ORs in the bits of its offsets (one atomic per CTA).
template <int KS, int KPB>
__global__ void __launch_bounds__(128) k_kernel_map_par(const B2Slot* __restrict__ table, unsigned int cap_mask,
                                                        const int* __restrict__ out_coords, int n_out, int step,
                                                        int* __restrict__ nbr, int ld, unsigned int* __restrict__ tile_mask) {
  const int o = blockIdx.x * 128 + threadIdx.x;
  const int k0 = blockIdx.y * KPB;
  int res[KPB];
#pragma unroll
  for (int j = 0; j < KPB; ++j) res[j] = -1;
  if (o < n_out) {
    const int4 c = __ldg(reinterpret_cast<const int4*>(out_coords) + o);
#pragma unroll
    for (int j = 0; j < KPB; ++j) {
      const int k = k0 + j;
      const int ix = k % KS, iy = (k / KS) % KS, iz = k / (KS * KS);
      const int base = (KS & 1) ? -(KS / 2) : 0;
      const int x = c.y + (ix + base) * step, y = c.z + (iy + base) * step, z = c.w + (iz + base) * step;
      if (b2_coord_in_range(c.x, x, y, z)) res[j] = b2_table_find(table, cap_mask, b2_pack_key(c.x, x, y, z));
    }
  }
  unsigned int m = 0;
#pragma unroll
  for (int j = 0; j < KPB; ++j) {
    nbr[(long long)(k0 + j) * ld + o] = res[j];
    if (__syncthreads_or(res[j] >= 0)) m |= (1u << (k0 + j));
  }
  if (threadIdx.x == 0 && m) atomicOr(tile_mask + blockIdx.x, m);
}

// Stride-2 table (out map = in coordinates floored to 2*step) built from the INPUT side: an input voxel at u*step is the
// offset-k neighbour of the output voxels at u*step - d_k*step that are multiples of 2*step - per axis one candidate
// when u is even (d = 0), two when it is odd (d = -1, +1; even kernels: exactly one) - so ~3.4 probes of the OUT table
// per input row instead of 27 probes of the input table per output row.  Every (k, o) entry has at most one writer
// (input coordinates are unique), so the table equals k_kernel_map's bit for bit.  nbr is pre-filled with -1 and
// tile_mask with 0 by the caller (b2_kernel_map_strided).
template <int KS>
__global__ void __launch_bounds__(256) k_kernel_map_strided_in(const B2Slot* __restrict__ out_table, unsigned int cap_mask,
                                                               const int* __restrict__ in_coords, int n_in, int step,
                                                               int* __restrict__ nbr, int ld,
                                                               unsigned int* __restrict__ tile_mask) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n_in) return;
  const int4 c = __ldg(reinterpret_cast<const int4*>(in_coords) + i);
  const int v[3] = {c.y, c.z, c.w};
  int cnt[3], idx[3][2], q[3][2];
#pragma unroll
  for (int a = 0; a < 3; ++a) {
    const bool odd = b2_floor_div(v[a], step) & 1;
    if (KS & 1) {                 // centred: offset index = d + KS/2, d in -KS/2 .. KS/2 (KS = 3: -1, 0, 1)
      if (!odd) {
        cnt[a] = 1; idx[a][0] = KS / 2; q[a][0] = v[a];
        idx[a][1] = 0; q[a][1] = 0;
      } else {
        cnt[a] = 2;
        idx[a][0] = KS / 2 - 1; q[a][0] = v[a] + step;    // d = -1
        idx[a][1] = KS / 2 + 1; q[a][1] = v[a] - step;    // d = +1
      }
    } else {                      // even kernels start at the voxel itself: d in 0 .. KS-1 (KS = 2: 0, 1)
      cnt[a] = 1; idx[a][1] = 0; q[a][1] = 0;
      idx[a][0] = odd ? 1 : 0;
      q[a][0] = odd ? v[a] - step : v[a];
    }
  }
  for (int az = 0; az < cnt[2]; ++az)
    for (int ay = 0; ay < cnt[1]; ++ay)
      for (int ax = 0; ax < cnt[0]; ++ax) {
        const int x = q[0][ax], y = q[1][ay], z = q[2][az];
        if (!b2_coord_in_range(c.x, x, y, z)) continue;
        const int o = b2_table_find(out_table, cap_mask, b2_pack_key(c.x, x, y, z));
        if (o < 0) continue;
        const int k = idx[0][ax] + KS * idx[1][ay] + KS * KS * idx[2][az];
        nbr[(long long)k * ld + o] = i;
        // a tile's bit is set by its first writer only: ~440 rows x offsets land on every word, same-address atomics
        // serialise in L2, a plain (L1-bypassing) read of the word does not
        unsigned int* word = tile_mask + (o >> 7);
        if (!((*reinterpret_cast<volatile unsigned int*>(word) >> k) & 1u)) atomicOr(word, 1u << k);
      }
}

// Neighbour table of a CHILDREN map (rows 8*i + c, c = cx + 2*cy + 4*cz, all eight children of every parent present)
// derived from the parent's 3^3 table without touching the hash: the neighbour of child c at offset d lies in the parent
// at offset e = floor((c + d) / 2) and is that parent's child c' = (c + d) - 2e.
__global__ void k_kernel_map_children(const int* __restrict__ pnbr, int pld, int n_parent, int* __restrict__ nbr, int ld,
                                      unsigned int* __restrict__ tile_mask) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  const int k = blockIdx.y;  // child offset index, x fastest
  if (t >= ld) return;
  int out = -1;
  if (t < 8ll * n_parent) {
    const int i = (int)(t >> 3), c = (int)(t & 7);
    const int d[3] = {k % 3 - 1, (k / 3) % 3 - 1, k / 9 - 1};
    const int cc[3] = {c & 1, (c >> 1) & 1, (c >> 2) & 1};
    int ke = 0, cnew = 0;
#pragma unroll
    for (int a = 0; a < 3; ++a) {
      const int s = cc[a] + d[a];              // -1 .. 2
      const int e = (s < 0) ? -1 : (s > 1 ? 1 : 0);
      const int cn = s - 2 * e;
      ke += (e + 1) * (a == 0 ? 1 : (a == 1 ? 3 : 9));
      cnew += cn << a;
    }
    const int p = __ldg(pnbr + (long long)ke * pld + i);
    out = (p >= 0) ? 8 * p + cnew : -1;
  }
  nbr[(long long)k * ld + t] = out;
  if (k == 0 && (t & 127) == 0) tile_mask[t >> 7] = 0x7FFFFFFu;  // dense blocks: every offset is in use (conservative)
}

// Pruning right behind a children map (fcaf3d_neck_with_head.py:99-102): the rows that survive the prune are known from
// the coarser level's scores before the 3^3 convolution of the up-block runs, and in eval mode that convolution is
// followed by per-row work only, so it is enough to evaluate it at the surviving rows.  Step 1: where every row of the
// children map ("gen" rows 8 i + c) and every backbone row goes in the pruned map (-1 = dropped), and which gen row
// feeds each pruned row.  dst_b / dst_a are the union's row maps, mask / pos the prune mask and its exclusive scan.
__global__ void k_prune_plan(const int* __restrict__ dst_b, int n_gen, const int* __restrict__ dst_a, int n_a,
                             const uint8_t* __restrict__ mask, const int* __restrict__ pos, int* __restrict__ gen_to_kept,
                             int* __restrict__ a_to_kept, int* __restrict__ kept_src) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n_gen) {
    const int u = dst_b[t];
    const int r = mask[u] ? pos[u] : -1;
    gen_to_kept[t] = r;
    if (r >= 0) kept_src[r] = t;
  } else if (t < n_gen + n_a) {
    const int u = dst_a[t - n_gen];
    a_to_kept[t - n_gen] = mask[u] ? pos[u] : -1;
  }
}

// Step 2: both 3^3 tables of the pruned rows from the PARENT's table (same index arithmetic as k_kernel_map_children):
// nbr_up[k][r] = gen row next to pruned row r (input of the up-block convolution = all gen rows),
// nbr_out[k][r] = the same neighbour as a pruned row, -1 if it was dropped (table of the pruned map itself).
// One 128-row tile per CTA; exact tile masks.
__global__ void __launch_bounds__(128) k_kernel_map_children_pruned(
    const int* __restrict__ pnbr, int pld, const int* __restrict__ kept_src, const int* __restrict__ gen_to_kept,
    int n_keep, int ld, int* __restrict__ nbr_up, unsigned int* __restrict__ mask_up, int* __restrict__ nbr_out,
    unsigned int* __restrict__ mask_out) {
  __shared__ unsigned int s_up, s_out;
  if (threadIdx.x == 0) s_up = s_out = 0u;
  __syncthreads();
  const int r = blockIdx.x * 128 + threadIdx.x;
  const int t = (r < n_keep) ? kept_src[r] : -1;
  const int i = t >> 3, c = t & 7;
  unsigned int bits_up = 0u, bits_out = 0u;
#pragma unroll
  for (int k = 0; k < 27; ++k) {
    int g = -1, o = -1;
    if (t >= 0) {
      const int d[3] = {k % 3 - 1, (k / 3) % 3 - 1, k / 9 - 1};
      int ke = 0, cnew = 0;
#pragma unroll
      for (int a = 0; a < 3; ++a) {
        const int s = ((c >> a) & 1) + d[a];     // -1 .. 2
        const int e = (s < 0) ? -1 : (s > 1 ? 1 : 0);
        ke += (e + 1) * (a == 0 ? 1 : (a == 1 ? 3 : 9));
        cnew += (s - 2 * e) << a;
      }
      const int p = __ldg(pnbr + (long long)ke * pld + i);
      if (p >= 0) {
        g = 8 * p + cnew;
        o = __ldg(gen_to_kept + g);
      }
    }
    nbr_up[(long long)k * ld + r] = g;
    nbr_out[(long long)k * ld + r] = o;
    if (__any_sync(0xffffffffu, g >= 0)) bits_up |= 1u << k;
    if (__any_sync(0xffffffffu, o >= 0)) bits_out |= 1u << k;
  }
  if ((threadIdx.x & 31) == 0) {
    atomicOr(&s_up, bits_up);
    atomicOr(&s_out, bits_out);
  }
  __syncthreads();
  if (threadIdx.x == 0) {
    mask_up[blockIdx.x] = s_up;
    mask_out[blockIdx.x] = s_out;
  }
}

// ---------------------------------------------------------------------------
// CUDA-core implicit-GEMM sparse convolution (the fp32-parity path and the path
// for narrow layers: Cin = 3 stem, Cout <= 32 heads).  Output-stationary:
// a CTA owns a 128 x 64 output tile and accumulates over all kernel offsets, so
// there are no atomics and no scatter.
//   out[o, :] = act( (sum_k in[nbr[k][o], :] @ W[k] + bias) * scale + shift + residual[o, :] )
// ---------------------------------------------------------------------------
#define CV_TM 128
#define CV_TN 64
#define CV_TK 16

template <typename T>
__device__ __forceinline__ float to_f32(T v);
template <>
__device__ __forceinline__ float to_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ float to_f32<__nv_bfloat16>(__nv_bfloat16 v) { return __bfloat162float(v); }
template <typename T>
__device__ __forceinline__ T from_f32(float v);
template <>
__device__ __forceinline__ float from_f32<float>(float v) { return v; }
template <>
__device__ __forceinline__ __nv_bfloat16 from_f32<__nv_bfloat16>(float v) { return __float2bfloat16(v); }

template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256) k_conv_cc(const TIn* __restrict__ in, const float* __restrict__ w,
                                                 const int* __restrict__ nbr, int ld,
                                                 const unsigned int* __restrict__ tile_mask, int n_out, int cin, int cout,
                                                 int kvol, const float* __restrict__ bias, const float* __restrict__ scale,
                                                 const float* __restrict__ shift, const TOut* __restrict__ residual,
                                                 int act, int ch_mod, TOut* __restrict__ out) {
  __shared__ __align__(16) float As[CV_TK][CV_TM + 4];
  __shared__ __align__(16) float Bs[CV_TK][CV_TN + 4];
  __shared__ int s_idx[CV_TM];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;  // 16 x 16 threads; thread tile 8 rows x 4 cols
  const int row0 = blockIdx.x * CV_TM, col0 = blockIdx.y * CV_TN;
  const unsigned int tmask = (tile_mask != nullptr && nbr != nullptr) ? tile_mask[blockIdx.x] : 0xFFFFFFFFu;

  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;

  for (int k = 0; k < kvol; ++k) {
    if (kvol <= 32 && !((tmask >> k) & 1u)) continue;
    __syncthreads();
    if (tid < CV_TM) {
      const int o = row0 + tid;
      int r = -1;
      if (o < n_out) r = nbr ? nbr[(long long)k * ld + o] : o;
      s_idx[tid] = r;
    }
    __syncthreads();
    const float* wk = w + (long long)k * cin * cout;
    for (int c0 = 0; c0 < cin; c0 += CV_TK) {
      // A tile: 128 rows x 16 channels; thread -> (row = tid/2, 8 channels)
      {
        const int r = tid >> 1, cb = (tid & 1) * 8;
        const int src = s_idx[r];
#pragma unroll
        for (int u = 0; u < 8; ++u) {
          const int c = c0 + cb + u;
          float v = 0.f;
          if (src >= 0 && c < cin) v = to_f32<TIn>(in[(long long)src * cin + c]);
          As[cb + u][r] = v;
        }
      }
      // B tile: 16 x 64; thread -> (kk = tid/16, 4 cols)
      {
        const int kk = tid >> 4, nb = (tid & 15) * 4;
        const int c = c0 + kk;
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          const int n = col0 + nb + u;
          float v = 0.f;
          if (c < cin && n < cout) v = __ldg(wk + (long long)c * cout + n);
          Bs[kk][nb + u] = v;
        }
      }
      __syncthreads();
#pragma unroll
      for (int kk = 0; kk < CV_TK; ++kk) {
        const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 8]);
        const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][ty * 8 + 4]);
        const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
        const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
        const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
        for (int i = 0; i < 8; ++i)
#pragma unroll
          for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
      }
      __syncthreads();
    }
  }

#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int o = row0 + ty * 8 + i;
    if (o >= n_out) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = col0 + tx * 4 + j;
      if (n >= cout) continue;
      const int ch = ch_mod > 0 ? (n % ch_mod) : n;
      float v = acc[i][j];
      if (bias) v += bias[ch];
      if (scale) v = v * scale[ch] + shift[ch];
      if (residual) v += to_f32<TOut>(residual[(long long)o * cout + n]);
      v = b2_apply_act(v, act);
      out[(long long)o * cout + n] = from_f32<TOut>(v);
    }
  }
}

// Cin <= 8 (the stem: Cin = 3, K = 27): flatten (offset, channel) into one inner dimension of K*Cin so that the
// 16-wide inner tiles are full (81 -> 6 tiles instead of 27 tiles with 3 of 16 lanes used).
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(256) k_conv_flat(const TIn* __restrict__ in, const float* __restrict__ w,
                                                   const int* __restrict__ nbr, int ld, int n_out, int cin, int cout,
                                                   int kvol, const float* __restrict__ bias,
                                                   const float* __restrict__ scale, const float* __restrict__ shift,
                                                   const TOut* __restrict__ residual, int act, int ch_mod,
                                                   TOut* __restrict__ out) {
  __shared__ __align__(16) float As[CV_TK][CV_TM + 4];
  __shared__ __align__(16) float Bs[CV_TK][CV_TN + 4];
  __shared__ int s_nbr[27 * CV_TM];
  const int tid = threadIdx.x;
  const int tx = tid & 15, ty = tid >> 4;
  const int row0 = blockIdx.x * CV_TM, col0 = blockIdx.y * CV_TN;
  for (int i = tid; i < kvol * CV_TM; i += 256) {
    const int k = i / CV_TM, r = i - k * CV_TM;
    s_nbr[i] = (row0 + r < n_out) ? nbr[(long long)k * ld + row0 + r] : -1;
  }
  float acc[8][4];
#pragma unroll
  for (int i = 0; i < 8; ++i)
#pragma unroll
    for (int j = 0; j < 4; ++j) acc[i][j] = 0.f;
  const int J = kvol * cin;
  __syncthreads();
  for (int j0 = 0; j0 < J; j0 += CV_TK) {
    {
      const int r = tid >> 1, cb = (tid & 1) * 8;
#pragma unroll
      for (int u = 0; u < 8; ++u) {
        const int j = j0 + cb + u;
        float v = 0.f;
        if (j < J) {
          const int k = j / cin, c = j - k * cin;
          const int src = s_nbr[k * CV_TM + r];
          if (src >= 0) v = to_f32<TIn>(in[(long long)src * cin + c]);
        }
        As[cb + u][r] = v;
      }
    }
    {
      const int kk = tid >> 4, nb = (tid & 15) * 4;
      const int j = j0 + kk;
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int n = col0 + nb + u;
        Bs[kk][nb + u] = (j < J && n < cout) ? __ldg(w + (long long)j * cout + n) : 0.f;
      }
    }
    __syncthreads();
#pragma unroll
    for (int kk = 0; kk < CV_TK; ++kk) {
      const float4 a0 = *reinterpret_cast<const float4*>(&As[kk][ty * 8]);
      const float4 a1 = *reinterpret_cast<const float4*>(&As[kk][ty * 8 + 4]);
      const float4 b = *reinterpret_cast<const float4*>(&Bs[kk][tx * 4]);
      const float a[8] = {a0.x, a0.y, a0.z, a0.w, a1.x, a1.y, a1.z, a1.w};
      const float bb[4] = {b.x, b.y, b.z, b.w};
#pragma unroll
      for (int i = 0; i < 8; ++i)
#pragma unroll
        for (int j = 0; j < 4; ++j) acc[i][j] = fmaf(a[i], bb[j], acc[i][j]);
    }
    __syncthreads();
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int o = row0 + ty * 8 + i;
    if (o >= n_out) continue;
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      const int n = col0 + tx * 4 + j;
      if (n >= cout) continue;
      const int ch = ch_mod > 0 ? (n % ch_mod) : n;
      float v = acc[i][j];
      if (bias) v += bias[ch];
      if (scale) v = v * scale[ch] + shift[ch];
      if (residual) v += to_f32<TOut>(residual[(long long)o * cout + n]);
      out[(long long)o * cout + n] = from_f32<TOut>(b2_apply_act(v, act));
    }
  }
}

// Narrow-input convolution that only touches the neighbours that exist (the stem: Cin = 3, 27 offsets, but a voxel of the
// stride-2 map has ~3 of its 27 input neighbours: a dense 81-wide inner product multiplies 90 % zeros).
// Every WARP works on its own 32 output rows without block-wide barriers: phase 1, lane = row, reads the row's 27 table
// entries (coalesced over the lanes, all loads independent) and keeps the present ones as (k << 26 | input row) in the
// warp's shared list; phase 2, two rows at a time, 16 lanes x 4 output channels per row, walks the list (input loads of
// four neighbours issued back to back, weights [K][Cin][64] from shared memory) and writes one 128-byte output row per
// half warp.  Accumulation order per output element: k ascending, then channel (as in k_conv_flat).
#define CN_WARPS 8
template <typename TIn, typename TOut>
__global__ void __launch_bounds__(CN_WARPS * 32) k_conv_narrow(const TIn* __restrict__ in, const float* __restrict__ w,
                                                               const int* __restrict__ nbr, int ld, int n_out, int cin,
                                                               int cout, int kvol, const float* __restrict__ bias,
                                                               const float* __restrict__ scale,
                                                               const float* __restrict__ shift,
                                                               const TOut* __restrict__ residual, int act, int ch_mod,
                                                               TOut* __restrict__ out) {
  extern __shared__ __align__(16) unsigned char cn_smem[];
  int* s_list = reinterpret_cast<int*>(cn_smem);                                    // [CN_WARPS][27][32]
  float* s_w = reinterpret_cast<float*>(cn_smem + CN_WARPS * 27 * 32 * 4);          // [kvol][cin][64]
  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  for (int i = tid; i < kvol * cin * 64; i += CN_WARPS * 32) {
    const int n = i & 63, kc = i >> 6;
    s_w[i] = (n < cout) ? __ldg(w + (long long)kc * cout + n) : 0.f;
  }
  __syncthreads();
  int* list = s_list + warp * 27 * 32;
  const int half = lane >> 4, tx = lane & 15;
  for (long long row0 = ((long long)blockIdx.x * CN_WARPS + warp) * 32; row0 < n_out;
       row0 += (long long)gridDim.x * CN_WARPS * 32) {
    int cnt = 0;
    {
      const long long o = row0 + lane;
      if (o < n_out) {
        int v[27];
#pragma unroll
        for (int k = 0; k < 27; ++k) v[k] = (k < kvol) ? __ldg(nbr + (long long)k * ld + o) : -1;
#pragma unroll
        for (int k = 0; k < 27; ++k)
          if (v[k] >= 0) list[(cnt++) * 32 + lane] = (k << 26) | v[k];
      }
    }
    __syncwarp();
#pragma unroll 1
    for (int rr = 0; rr < 16; ++rr) {
      const int r = rr * 2 + half;
      const long long o = row0 + r;
      const int n_e = __shfl_sync(0xffffffffu, cnt, r);
      float acc[4] = {0.f, 0.f, 0.f, 0.f};
      for (int j0 = 0; j0 < n_e; j0 += 4) {
        int e[4];
        float a[4][4];
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          e[u] = (j0 + u < n_e) ? list[(j0 + u) * 32 + r] : -1;
          const TIn* p = in + (long long)(e[u] & 0x3FFFFFF) * cin;
#pragma unroll
          for (int c = 0; c < 4; ++c) a[u][c] = (e[u] >= 0 && c < cin) ? to_f32<TIn>(p[c]) : 0.f;
        }
#pragma unroll
        for (int u = 0; u < 4; ++u) {
          if (e[u] < 0) continue;
          const int k = e[u] >> 26;
#pragma unroll
          for (int c = 0; c < 4; ++c) {
            if (c >= cin) continue;
            const float4 wv = *reinterpret_cast<const float4*>(&s_w[((k * cin + c) << 6) + tx * 4]);
            acc[0] = fmaf(a[u][c], wv.x, acc[0]);
            acc[1] = fmaf(a[u][c], wv.y, acc[1]);
            acc[2] = fmaf(a[u][c], wv.z, acc[2]);
            acc[3] = fmaf(a[u][c], wv.w, acc[3]);
          }
        }
      }
      if (o < n_out) {
#pragma unroll
        for (int j = 0; j < 4; ++j) {
          const int n = tx * 4 + j;
          if (n >= cout) continue;
          const int ch = ch_mod > 0 ? (n % ch_mod) : n;
          float v = acc[j];
          if (bias) v += bias[ch];
          if (scale) v = v * scale[ch] + shift[ch];
          if (residual) v += to_f32<TOut>(residual[o * cout + n]);
          out[o * cout + n] = from_f32<TOut>(b2_apply_act(v, act));
        }
      }
    }
    __syncwarp();
  }
}

// ---------------------------------------------------------------------------
// Max pooling over a neighbour table (2^3, stride 2): out[o, c] = max_k in[nbr[k][o], c]
// ---------------------------------------------------------------------------
template <typename T>
__global__ void k_maxpool(const T* __restrict__ in, const int* __restrict__ nbr, int ld, int kvol, int n_out, int c,
                          T* __restrict__ out, int* __restrict__ arg) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n_out * c) return;
  const int o = (int)(t / c), ch = (int)(t - (long long)o * c);
  float best = -INFINITY;
  int besti = -1;
  for (int k = 0; k < kvol; ++k) {
    const int r = nbr[(long long)k * ld + o];
    if (r >= 0) {
      const float v = to_f32<T>(in[(long long)r * c + ch]);
      if (besti < 0 || v > best) {
        best = v;
        besti = r;
      }
    }
  }
  out[t] = from_f32<T>(besti >= 0 ? best : 0.f);
  if (arg) arg[t] = besti;
}

// 8 bf16 channels per thread (16-byte loads); same max / first-wins rule as k_maxpool
__global__ void k_maxpool_bf16x8(const uint4* __restrict__ in, const int* __restrict__ nbr, int ld, int kvol, int n_out,
                                 int vpr, uint4* __restrict__ out) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n_out * vpr) return;
  const int o = (int)(t / vpr), v = (int)(t - (long long)o * vpr);
  float best[8];
  bool any = false;
  for (int k = 0; k < kvol; ++k) {
    const int r = nbr[(long long)k * ld + o];
    if (r < 0) continue;
    const uint4 u = __ldg(in + (long long)r * vpr + v);
    const unsigned w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float lo = __uint_as_float(w4[e] << 16), hi = __uint_as_float(w4[e] & 0xFFFF0000u);
      best[2 * e] = (!any || lo > best[2 * e]) ? lo : best[2 * e];
      best[2 * e + 1] = (!any || hi > best[2 * e + 1]) ? hi : best[2 * e + 1];
    }
    any = true;
  }
  unsigned r4[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(any ? best[2 * e] : 0.f, any ? best[2 * e + 1] : 0.f);
    r4[e] = *reinterpret_cast<const unsigned*>(&h);
  }
  out[t] = make_uint4(r4[0], r4[1], r4[2], r4[3]);
}

// Max pool over act(x * scale[s] + shift[s]) (s = scene of the output row) without materialising the normalised rows:
// every gathered value is normalised, activated and rounded to bf16 exactly as k_affine_act_bf16x8 would store it, then
// the maximum is taken, so the result equals b2_affine_act followed by b2_maxpool bit for bit.
__global__ void k_maxpool_affine_bf16x8(const uint4* __restrict__ in, const int* __restrict__ nbr, int ld, int kvol,
                                        int n_out, int vpr, const int* __restrict__ out_coords,
                                        const float* __restrict__ scale, const float* __restrict__ shift, int act,
                                        uint4* __restrict__ out) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (long long)n_out * vpr) return;
  const int o = (int)(t / vpr), v = (int)(t - (long long)o * vpr);
  const int s = out_coords ? __ldg(out_coords + 4ll * o) : 0;
  const float* sc = scale + ((long long)s * vpr + v) * 8;
  const float* sh = shift + ((long long)s * vpr + v) * 8;
  float a[8], b[8];
  {  // 32-byte aligned rows of 8 floats: two 16-byte loads each instead of eight scalar ones
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(sc)), a1 = __ldg(reinterpret_cast<const float4*>(sc) + 1);
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(sh)), b1 = __ldg(reinterpret_cast<const float4*>(sh) + 1);
    a[0] = a0.x; a[1] = a0.y; a[2] = a0.z; a[3] = a0.w; a[4] = a1.x; a[5] = a1.y; a[6] = a1.z; a[7] = a1.w;
    b[0] = b0.x; b[1] = b0.y; b[2] = b0.z; b[3] = b0.w; b[4] = b1.x; b[5] = b1.y; b[6] = b1.z; b[7] = b1.w;
  }
  float best[8];
  bool any = false;
  // kvol <= 8 (2^3 pooling): all neighbour indices first, then all row loads, so that up to 8 loads are in flight
  int rr[8];
  uint4 uu[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) rr[k] = (k < kvol) ? __ldg(nbr + (long long)k * ld + o) : -1;
#pragma unroll
  for (int k = 0; k < 8; ++k)
    uu[k] = (rr[k] >= 0) ? __ldg(in + (long long)rr[k] * vpr + v) : make_uint4(0u, 0u, 0u, 0u);
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    if (rr[k] < 0) continue;
    const uint4 u = uu[k];
    const unsigned w4[4] = {u.x, u.y, u.z, u.w};
    float f[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      f[2 * e] = __uint_as_float(w4[e] << 16);
      f[2 * e + 1] = __uint_as_float(w4[e] & 0xFFFF0000u);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = f[e] * a[e] + b[e];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(b2_apply_act_bf16(f[2 * e], act), b2_apply_act_bf16(f[2 * e + 1], act));
      const unsigned w = *reinterpret_cast<const unsigned*>(&h);
      const float lo = __uint_as_float(w << 16), hi = __uint_as_float(w & 0xFFFF0000u);
      best[2 * e] = (!any || lo > best[2 * e]) ? lo : best[2 * e];
      best[2 * e + 1] = (!any || hi > best[2 * e + 1]) ? hi : best[2 * e + 1];
    }
    any = true;
  }
  unsigned r4[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(any ? best[2 * e] : 0.f, any ? best[2 * e + 1] : 0.f);
    r4[e] = *reinterpret_cast<const unsigned*>(&h);
  }
  out[t] = make_uint4(r4[0], r4[1], r4[2], r4[3]);
}

__global__ void k_affine_act_bf16x8(const uint4* __restrict__ x, const int* __restrict__ coords, long long n, int c,
                                    const float* __restrict__ scale, const float* __restrict__ shift,
                                    const uint4* __restrict__ residual, int act, uint4* __restrict__ y) {
  const int vpr = c / 8;
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * vpr) return;
  const long long r = t / vpr;
  const int ch0 = (int)(t - r * vpr) * 8;
  const int s = coords ? coords[4 * r] : 0;
  const uint4 u = x[t];
  const unsigned w4[4] = {u.x, u.y, u.z, u.w};
  float f[8];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    f[2 * e] = __uint_as_float(w4[e] << 16);
    f[2 * e + 1] = __uint_as_float(w4[e] & 0xFFFF0000u);
  }
  if (scale) {
    const float* sc = scale + (long long)s * c + ch0;
    const float* sh = shift + (long long)s * c + ch0;
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = f[e] * __ldg(sc + e) + __ldg(sh + e);
  }
  if (residual) {
    const uint4 q = residual[t];
    const unsigned q4[4] = {q.x, q.y, q.z, q.w};
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      f[2 * e] += __uint_as_float(q4[e] << 16);
      f[2 * e + 1] += __uint_as_float(q4[e] & 0xFFFF0000u);
    }
  }
  unsigned r4[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(b2_apply_act_bf16(f[2 * e], act), b2_apply_act_bf16(f[2 * e + 1], act));
    r4[e] = *reinterpret_cast<const unsigned*>(&h);
  }
  y[t] = make_uint4(r4[0], r4[1], r4[2], r4[3]);
}

// The BatchNorm form of the above (one scale / shift row, scale != NULL): four vectors per thread, 256 vectors apart, which
// share their channels because 256 % (c / 8) == 0 - scale and shift are read once, the four row loads are in flight
// together (one vector per thread reached ~3.2 TB/s of the 6.5 TB/s HBM copy rate; same arithmetic, same bits).
__global__ void __launch_bounds__(256) k_affine_act_bf16x8_v4(const uint4* __restrict__ x, long long total, int c,
                                                              const float* __restrict__ scale,
                                                              const float* __restrict__ shift,
                                                              const uint4* __restrict__ residual, int act,
                                                              uint4* __restrict__ y) {
  const int vpr = c / 8;
  const long long t0 = (long long)blockIdx.x * 1024 + threadIdx.x;
  if (t0 >= total) return;
  const int ch0 = (int)(t0 % vpr) * 8;
  uint4 u[4], q[4];
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long t = t0 + i * 256;
    if (t < total) {
      u[i] = x[t];
      if (residual) q[i] = residual[t];
    }
  }
  float sc[8], sh[8];
  {
    const float4 a0 = __ldg(reinterpret_cast<const float4*>(scale + ch0)), a1 = __ldg(reinterpret_cast<const float4*>(scale + ch0) + 1);
    const float4 b0 = __ldg(reinterpret_cast<const float4*>(shift + ch0)), b1 = __ldg(reinterpret_cast<const float4*>(shift + ch0) + 1);
    sc[0] = a0.x; sc[1] = a0.y; sc[2] = a0.z; sc[3] = a0.w; sc[4] = a1.x; sc[5] = a1.y; sc[6] = a1.z; sc[7] = a1.w;
    sh[0] = b0.x; sh[1] = b0.y; sh[2] = b0.z; sh[3] = b0.w; sh[4] = b1.x; sh[5] = b1.y; sh[6] = b1.z; sh[7] = b1.w;
  }
#pragma unroll
  for (int i = 0; i < 4; ++i) {
    const long long t = t0 + i * 256;
    if (t >= total) break;
    const unsigned w4[4] = {u[i].x, u[i].y, u[i].z, u[i].w};
    float f[8];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      f[2 * e] = __uint_as_float(w4[e] << 16);
      f[2 * e + 1] = __uint_as_float(w4[e] & 0xFFFF0000u);
    }
#pragma unroll
    for (int e = 0; e < 8; ++e) f[e] = f[e] * sc[e] + sh[e];
    if (residual) {
      const unsigned q4[4] = {q[i].x, q[i].y, q[i].z, q[i].w};
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        f[2 * e] += __uint_as_float(q4[e] << 16);
        f[2 * e + 1] += __uint_as_float(q4[e] & 0xFFFF0000u);
      }
    }
    unsigned r4[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const __nv_bfloat162 h = __floats2bfloat162_rn(b2_apply_act_bf16(f[2 * e], act), b2_apply_act_bf16(f[2 * e + 1], act));
      r4[e] = *reinterpret_cast<const unsigned*>(&h);
    }
    y[t] = make_uint4(r4[0], r4[1], r4[2], r4[3]);
  }
}

// ---------------------------------------------------------------------------
// Per-segment channel statistics (segments = contiguous row ranges: scenes for
// InstanceNorm, the whole tensor for BatchNorm).  Deterministic two-level sum:
// partial[s][chunk][c] then a final reduce; second pass centres on the mean.
// ---------------------------------------------------------------------------
#define ST_CHUNKS 64

template <typename T, int POW>
__global__ void __launch_bounds__(256) k_colsum_partial(const T* __restrict__ x, const int* __restrict__ seg_off, int c,
                                                        const float* __restrict__ center, float* __restrict__ partial) {
  const int s = blockIdx.y, chunk = blockIdx.x;
  const int r0 = seg_off[s], r1 = seg_off[s + 1];
  const int n = r1 - r0;
  const int per = (n + ST_CHUNKS - 1) / ST_CHUNKS;
  const int a = r0 + chunk * per;
  const int b = min(a + per, r1);
  // thread -> channel lane; 256 threads cover channels in a strided way, rows in y when c < 256
  const int lanes_c = c < 256 ? c : 256;
  const int rows_par = 256 / lanes_c;  // >= 1
  const int cl = threadIdx.x % lanes_c, rl = threadIdx.x / lanes_c;
  __shared__ float red[256];
  for (int cb = 0; cb < c; cb += lanes_c) {
    const int ch = cb + cl;
    float acc = 0.f;
    if (ch < c && rl < rows_par) {
      const float m = center ? center[(long long)s * c + ch] : 0.f;
      for (int r = a + rl; r < b; r += rows_par) {
        float v = to_f32<T>(x[(long long)r * c + ch]) - m;
        acc += (POW == 2) ? v * v : v;
      }
    }
    red[threadIdx.x] = acc;
    __syncthreads();
    if (rl == 0 && ch < c) {
      float t = 0.f;
      for (int j = 0; j < rows_par; ++j) t += red[j * lanes_c + cl];
      partial[((long long)s * ST_CHUNKS + chunk) * c + ch] = t;
    }
    __syncthreads();
  }
}

__global__ void k_colsum_final(const float* __restrict__ partial, const int* __restrict__ seg_off, int c, int nseg,
                               float* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nseg * c) return;
  const int s = t / c, ch = t - s * c;
  float acc = 0.f;
  for (int j = 0; j < ST_CHUNKS; ++j) acc += partial[((long long)s * ST_CHUNKS + j) * c + ch];
  const int n = seg_off[s + 1] - seg_off[s];
  out[t] = n > 0 ? acc / (float)n : 0.f;
}

// InstanceNorm statistics in ONE pass over the rows (the two-pass kernels above read the largest tensor of the network
// twice with 2-byte loads): per scene and channel, sums of (x - pivot) and (x - pivot)^2 with pivot = the scene's first
// row (shifted sums: no cancellation for the magnitudes a convolution output has), 16-byte loads, IN_CHUNKS x nseg CTAs,
// partials added in a fixed order (deterministic).  k_instnorm_final turns them into the folded (scale, shift) of
// MinkowskiInstanceNorm directly: biased variance, eps inside the square root (ME App. A.6; me_resnet.py:22).
#define IN_CHUNKS 74
template <typename T>
__global__ void __launch_bounds__(256) k_instnorm_partial(const T* __restrict__ x, const int* __restrict__ seg_off, int c,
                                                          float* __restrict__ partial) {
  constexpr int E = 16 / (int)sizeof(T);
  __shared__ float red[256 * 2 * E];
  const int s = blockIdx.y, chunk = blockIdx.x;
  const int r0 = seg_off[s], r1 = seg_off[s + 1];
  const int n = r1 - r0;
  const int V = c / E, rows_par = 256 / V;
  const int vl = threadIdx.x % V, rl = threadIdx.x / V;
  const int per = (n + IN_CHUNKS - 1) / IN_CHUNKS;
  const int a = r0 + chunk * per, b = min(a + per, r1);
  float piv[E], s1[E], s2[E];
#pragma unroll
  for (int e = 0; e < E; ++e) piv[e] = s1[e] = s2[e] = 0.f;
  if (n > 0) {
#pragma unroll
    for (int e = 0; e < E; ++e) piv[e] = to_f32<T>(x[(long long)r0 * c + vl * E + e]);
  }
  for (int r = a + rl; r < b; r += rows_par) {
    const uint4 u = __ldg(reinterpret_cast<const uint4*>(x + (long long)r * c) + vl);
    const T* v = reinterpret_cast<const T*>(&u);
#pragma unroll
    for (int e = 0; e < E; ++e) {
      const float d = to_f32<T>(v[e]) - piv[e];
      s1[e] += d;
      s2[e] += d * d;
    }
  }
#pragma unroll
  for (int e = 0; e < E; ++e) {
    red[(threadIdx.x * 2) * E + e] = s1[e];
    red[(threadIdx.x * 2 + 1) * E + e] = s2[e];
  }
  __syncthreads();
  if (rl == 0) {
    float* out = partial + ((long long)s * IN_CHUNKS + chunk) * 2 * c;
#pragma unroll
    for (int e = 0; e < E; ++e) {
      float t1 = 0.f, t2 = 0.f;
      for (int j = 0; j < rows_par; ++j) {
        t1 += red[((j * V + vl) * 2) * E + e];
        t2 += red[((j * V + vl) * 2 + 1) * E + e];
      }
      out[vl * E + e] = t1;
      out[c + vl * E + e] = t2;
    }
  }
}

template <typename T>
__global__ void k_instnorm_final(const T* __restrict__ x, const float* __restrict__ partial, const int* __restrict__ seg_off,
                                 int c, int nseg, const float* __restrict__ weight, const float* __restrict__ bias, float eps,
                                 float* __restrict__ scale, float* __restrict__ shift) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= nseg * c) return;
  const int s = t / c, ch = t - s * c;
  const int r0 = seg_off[s], n = seg_off[s + 1] - r0;
  double mean = 0.0, var = 0.0;
  if (n > 0) {
    double a1 = 0.0, a2 = 0.0;
    // all 2 x 74 loads are issued before the (ordered) additions: the loop used to be one dependent load latency per trip
    float p1[IN_CHUNKS], p2[IN_CHUNKS];
#pragma unroll
    for (int j = 0; j < IN_CHUNKS; ++j) {
      p1[j] = __ldg(partial + ((long long)s * IN_CHUNKS + j) * 2 * c + ch);
      p2[j] = __ldg(partial + ((long long)s * IN_CHUNKS + j) * 2 * c + c + ch);
    }
#pragma unroll
    for (int j = 0; j < IN_CHUNKS; ++j) {
      a1 += (double)p1[j];
      a2 += (double)p2[j];
    }
    const double m = a1 / n;
    mean = (double)to_f32<T>(x[(long long)r0 * c + ch]) + m;
    var = a2 / n - m * m;
    if (var < 0.0) var = 0.0;
  }
  const float sc = (weight ? weight[ch] : 1.f) / sqrtf((float)var + eps);
  scale[t] = sc;
  shift[t] = (bias ? bias[ch] : 0.f) - (float)mean * sc;
}

// y = act(x * scale[s, c] + shift[s, c] + residual); s = batch index of the row (coords) or 0
template <typename T>
__global__ void k_affine_act(const T* __restrict__ x, const int* __restrict__ coords, long long n, int c,
                             const float* __restrict__ scale, const float* __restrict__ shift,
                             const T* __restrict__ residual, int act, T* __restrict__ y) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c) return;
  const long long r = t / c;
  const int ch = (int)(t - r * c);
  const int s = coords ? coords[4 * r] : 0;
  float v = to_f32<T>(x[t]);
  if (scale) v = v * scale[(long long)s * c + ch] + shift[(long long)s * c + ch];
  if (residual) v += to_f32<T>(residual[t]);
  y[t] = from_f32<T>(b2_apply_act(v, act));
}

__global__ void k_f32_to_bf16(const float* __restrict__ x, long long n, __nv_bfloat16* __restrict__ y) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) y[t] = __float2bfloat16(x[t]);
}
__global__ void k_bf16_to_f32(const __nv_bfloat16* __restrict__ x, long long n, float* __restrict__ y) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t < n) y[t] = __bfloat162float(x[t]);
}

// points [n, 3+F] fp32 -> coords (b, floor(xyz / voxel)) int32 [n, 4] and feats = rest / feat_div
// (single_stage_sparse.py:34-36 + ME.utils.batch_sparse_collate)
__global__ void k_quantize_points(const float* __restrict__ pts, int n, int dim, int batch, float voxel_size,
                                  float feat_div, int row0, int* __restrict__ coords, float* __restrict__ feats) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* p = pts + (long long)i * dim;
  int4 c;
  c.x = batch;
  c.y = (int)floorf(p[0] / voxel_size);
  c.z = (int)floorf(p[1] / voxel_size);
  c.w = (int)floorf(p[2] / voxel_size);
  reinterpret_cast<int4*>(coords)[row0 + i] = c;
  for (int f = 3; f < dim; ++f) feats[(long long)(row0 + i) * (dim - 3) + (f - 3)] = p[f] / feat_div;
}

extern "C" {

// ksize in {1,2,3}; step = input tensor stride * dilation; ld = n_out rounded up to 128
int b2_kernel_map(const void* in_table, long long cap, const int* out_coords, long long n_out, int ksize, int step,
                  int* nbr, long long ld, unsigned int* tile_mask, void* stream) {
  B2_CHECK_ARG(ld % 128 == 0 && ld >= n_out, "ld must be n_out rounded up to 128");
  B2_CHECK_ARG(ksize >= 1 && ksize <= 3, "kernel size %d unsupported (1..3)", ksize);
  if (ld == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int G = (int)(ld / 128);
  const B2Slot* t = (const B2Slot*)in_table;
  const unsigned m = (unsigned)(cap - 1);
  static const bool par = !(getenv("B2_KMAP_PAR") && atoi(getenv("B2_KMAP_PAR")) == 0);   // A/B switch
  if (ksize == 1) k_kernel_map<1><<<G, 128, 0, st>>>(t, m, out_coords, (int)n_out, step, nbr, (int)ld, tile_mask);
  if (ksize >= 2 && par) {
    B2_CUDA(cudaMemsetAsync(tile_mask, 0, sizeof(unsigned int) * (size_t)G, st));
    if (ksize == 2)
      k_kernel_map_par<2, 2><<<dim3(G, 4), 128, 0, st>>>(t, m, out_coords, (int)n_out, step, nbr, (int)ld, tile_mask);
    else
      k_kernel_map_par<3, 3><<<dim3(G, 9), 128, 0, st>>>(t, m, out_coords, (int)n_out, step, nbr, (int)ld, tile_mask);
  } else {
    if (ksize == 2) k_kernel_map<2><<<G, 128, 0, st>>>(t, m, out_coords, (int)n_out, step, nbr, (int)ld, tile_mask);
    if (ksize == 3) k_kernel_map<3><<<G, 128, 0, st>>>(t, m, out_coords, (int)n_out, step, nbr, (int)ld, tile_mask);
  }
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_kernel_map_strided(const void* out_table, long long cap, const int* in_coords, long long n_in, int ksize, int step,
                          int* nbr, long long ld, unsigned int* tile_mask, void* stream) {
  B2_CHECK_ARG(ld % 128 == 0, "ld must be n_out rounded up to 128");
  B2_CHECK_ARG(ksize == 2 || ksize == 3, "kernel size %d unsupported (2, 3)", ksize);
  B2_CHECK_ARG(step >= 1, "step must be positive");
  if (ld == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int kvol = ksize * ksize * ksize;
  B2_CUDA(cudaMemsetAsync(nbr, 0xFF, sizeof(int) * (size_t)kvol * (size_t)ld, st));
  B2_CUDA(cudaMemsetAsync(tile_mask, 0, sizeof(unsigned int) * (size_t)(ld / 128), st));
  if (n_in == 0) return B2_OK;
  const B2Slot* t = (const B2Slot*)out_table;
  const unsigned m = (unsigned)(cap - 1);
  const int G = (int)b2_div_up(n_in, 256);
  if (ksize == 2) k_kernel_map_strided_in<2><<<G, 256, 0, st>>>(t, m, in_coords, (int)n_in, step, nbr, (int)ld, tile_mask);
  else k_kernel_map_strided_in<3><<<G, 256, 0, st>>>(t, m, in_coords, (int)n_in, step, nbr, (int)ld, tile_mask);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// 3^3 stride-1 table of a children map (b2_map_children order) from the parent map's 3^3 stride-1 table
int b2_kernel_map_children(const int* parent_nbr, long long parent_ld, long long n_parent, int* nbr, long long ld,
                           unsigned int* tile_mask, void* stream) {
  B2_CHECK_ARG(ld % 128 == 0 && ld >= 8 * n_parent, "ld must be 8 * n_parent rounded up to 128");
  if (ld == 0) return B2_OK;
  dim3 grid(b2_div_up(ld, 256), 27);
  k_kernel_map_children<<<grid, 256, 0, (cudaStream_t)stream>>>(parent_nbr, (int)parent_ld, (int)n_parent, nbr, (int)ld,
                                                               tile_mask);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_prune_plan(const int* dst_b, long long n_gen, const int* dst_a, long long n_a, const uint8_t* mask, const int* pos,
                  int* gen_to_kept, int* a_to_kept, int* kept_src, void* stream) {
  B2_CHECK_ARG(n_gen >= 0 && n_a >= 0 && n_gen + n_a < (1ll << 31), "row counts out of range");
  if (n_gen + n_a == 0) return B2_OK;
  k_prune_plan<<<b2_div_up(n_gen + n_a, 256), 256, 0, (cudaStream_t)stream>>>(dst_b, (int)n_gen, dst_a, (int)n_a, mask, pos,
                                                                             gen_to_kept, a_to_kept, kept_src);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_kernel_map_children_pruned(const int* parent_nbr, long long parent_ld, const int* kept_src, const int* gen_to_kept,
                                  long long n_keep, long long ld, int* nbr_up, unsigned int* mask_up, int* nbr_out,
                                  unsigned int* mask_out, void* stream) {
  B2_CHECK_ARG(ld % 128 == 0 && ld >= n_keep, "ld must be n_keep rounded up to 128");
  if (ld == 0) return B2_OK;
  k_kernel_map_children_pruned<<<(int)(ld / 128), 128, 0, (cudaStream_t)stream>>>(
      parent_nbr, (int)parent_ld, kept_src, gen_to_kept, (int)n_keep, (int)ld, nbr_up, mask_up, nbr_out, mask_out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// in_dtype/out_dtype: 0 = f32, 1 = bf16.  nbr == NULL means identity map (kvol must be 1).
// ch_mod > 0: epilogue channel = col % ch_mod (generative transpose packs 8 kernels along columns).
int b2_conv_cc(const void* in, int in_dtype, const float* w, const int* nbr, long long ld, const unsigned int* tile_mask,
               long long n_out, int cin, int cout, int kvol, const float* bias, const float* scale, const float* shift,
               const void* residual, int act, int ch_mod, void* out, int out_dtype, void* stream) {
  B2_CHECK_ARG(nbr != nullptr || kvol == 1, "identity map requires kernel volume 1");
  B2_CHECK_ARG((scale == nullptr) == (shift == nullptr), "scale and shift come together");
  if (n_out == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  dim3 grid(b2_div_up(n_out, CV_TM), b2_div_up(cout, CV_TN));
  const bool flat = (nbr != nullptr && cin <= 8 && kvol <= 27);
  static const bool narrow_ok = !(getenv("B2_CONV_NARROW") && atoi(getenv("B2_CONV_NARROW")) == 0);
  const bool narrow = narrow_ok && nbr != nullptr && cin <= 4 && cout <= 64 && kvol <= 27;
  const int cn_rows = CN_WARPS * 32;
  const int gn = (int)(b2_div_up(n_out, cn_rows) < 148 * 4 ? b2_div_up(n_out, cn_rows) : 148 * 4);
  const int cn_smem_bytes = CN_WARPS * 27 * 32 * 4 + kvol * cin * 64 * 4;
#define CC_LAUNCH(TI, TO)                                                                                          \
  do {                                                                                                             \
    if (narrow) {                                                                                                  \
      static unsigned long long cn_done = 0;                                                                       \
      if (cn_smem_bytes > 48 * 1024)                                                                               \
        if (int rc0 = b2_optin_smem((const void*)k_conv_narrow<TI, TO>, 56 * 1024, &cn_done)) return rc0;          \
      k_conv_narrow<TI, TO><<<gn, CN_WARPS * 32, cn_smem_bytes, st>>>((const TI*)in, w, nbr, (int)ld, (int)n_out, cin, \
                                                cout, kvol, bias, scale, shift, (const TO*)residual, act, ch_mod, (TO*)out); \
    } else if (flat)                                                                                                 \
      k_conv_flat<TI, TO><<<grid, 256, 0, st>>>((const TI*)in, w, nbr, (int)ld, (int)n_out, cin, cout, kvol, bias,  \
                                                scale, shift, (const TO*)residual, act, ch_mod, (TO*)out);         \
    else                                                                                                           \
      k_conv_cc<TI, TO><<<grid, 256, 0, st>>>((const TI*)in, w, nbr, (int)ld, tile_mask, (int)n_out, cin, cout,     \
                                              kvol, bias, scale, shift, (const TO*)residual, act, ch_mod, (TO*)out); \
  } while (0)
  if (in_dtype == 0 && out_dtype == 0) CC_LAUNCH(float, float);
  else if (in_dtype == 0 && out_dtype == 1) CC_LAUNCH(float, __nv_bfloat16);
  else if (in_dtype == 1 && out_dtype == 0) CC_LAUNCH(__nv_bfloat16, float);
  else CC_LAUNCH(__nv_bfloat16, __nv_bfloat16);
#undef CC_LAUNCH
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_maxpool(const void* in, int dtype, const int* nbr, long long ld, int kvol, long long n_out, int c, void* out,
               int* argmax, void* stream) {
  if (n_out == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int G = b2_div_up(n_out * c, 256);
  if (dtype == 0)
    k_maxpool<float><<<G, 256, 0, st>>>((const float*)in, nbr, (int)ld, kvol, (int)n_out, c, (float*)out, argmax);
  else if (c % 8 == 0 && argmax == nullptr)
    k_maxpool_bf16x8<<<b2_div_up(n_out * (c / 8), 256), 256, 0, st>>>((const uint4*)in, nbr, (int)ld, kvol, (int)n_out,
                                                                    c / 8, (uint4*)out);
  else
    k_maxpool<__nv_bfloat16><<<G, 256, 0, st>>>((const __nv_bfloat16*)in, nbr, (int)ld, kvol, (int)n_out, c,
                                                (__nv_bfloat16*)out, argmax);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_maxpool_affine(const void* in, const int* nbr, long long ld, int kvol, long long n_out, int c, const int* out_coords,
                      const float* scale, const float* shift, int act, void* out, void* stream) {
  B2_CHECK_ARG(c % 8 == 0, "channel count must be a multiple of 8");
  B2_CHECK_ARG(kvol >= 1 && kvol <= 8, "kernel volume %d unsupported (1..8)", kvol);
  B2_CHECK_ARG(scale != nullptr && shift != nullptr, "scale and shift are required");
  if (n_out == 0) return B2_OK;
  k_maxpool_affine_bf16x8<<<b2_div_up(n_out * (c / 8), 256), 256, 0, (cudaStream_t)stream>>>(
      (const uint4*)in, nbr, (int)ld, kvol, (int)n_out, c / 8, out_coords, scale, shift, act, (uint4*)out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

size_t b2_segment_stats_ws_bytes(int nseg, int c) { return (size_t)nseg * ST_CHUNKS * c * sizeof(float); }

// mean[s, c], var[s, c] (biased) over rows seg_off[s]..seg_off[s+1]
int b2_segment_stats(const void* x, int dtype, const int* seg_off, int nseg, int c, float* mean, float* var, void* ws,
                     size_t ws_bytes, void* stream) {
  B2_CHECK_ARG(nseg > 0 && c > 0, "bad shape");
  if (ws_bytes < b2_segment_stats_ws_bytes(nseg, c)) {
    b2_set_error("b2_segment_stats workspace too small");
    return B2_ERR_WS;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float* partial = (float*)ws;
  dim3 grid(ST_CHUNKS, nseg);
  const int GF = b2_div_up((long long)nseg * c, 256);
  if (dtype == 0) {
    k_colsum_partial<float, 1><<<grid, 256, 0, st>>>((const float*)x, seg_off, c, nullptr, partial);
    k_colsum_final<<<GF, 256, 0, st>>>(partial, seg_off, c, nseg, mean);
    k_colsum_partial<float, 2><<<grid, 256, 0, st>>>((const float*)x, seg_off, c, mean, partial);
    k_colsum_final<<<GF, 256, 0, st>>>(partial, seg_off, c, nseg, var);
  } else {
    k_colsum_partial<__nv_bfloat16, 1><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, seg_off, c, nullptr, partial);
    k_colsum_final<<<GF, 256, 0, st>>>(partial, seg_off, c, nseg, mean);
    k_colsum_partial<__nv_bfloat16, 2><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, seg_off, c, mean, partial);
    k_colsum_final<<<GF, 256, 0, st>>>(partial, seg_off, c, nseg, var);
  }
  b2_count_launches(3);  // four kernels behind one check
  B2_LAUNCH_CHECK();
  return B2_OK;
}

size_t b2_instnorm_ws_bytes(int nseg, int c) { return (size_t)nseg * IN_CHUNKS * 2 * c * sizeof(float); }

int b2_instnorm_scale_shift(const void* x, int dtype, const int* seg_off, int nseg, int c, const float* weight,
                            const float* bias, float eps, float* scale, float* shift, void* ws, size_t ws_bytes,
                            void* stream) {
  const int E = dtype == 0 ? 4 : 8;
  B2_CHECK_ARG(nseg > 0 && c > 0 && c % E == 0 && c / E <= 256 && 256 % (c / E) == 0,
               "channel count %d unsupported (vectors per row must divide 256)", c);
  if (ws_bytes < b2_instnorm_ws_bytes(nseg, c)) {
    b2_set_error("b2_instnorm_scale_shift workspace too small");
    return B2_ERR_WS;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float* partial = (float*)ws;
  dim3 grid(IN_CHUNKS, nseg);
  const int GF = b2_div_up((long long)nseg * c, 256);
  if (dtype == 0) {
    k_instnorm_partial<float><<<grid, 256, 0, st>>>((const float*)x, seg_off, c, partial);
    k_instnorm_final<float><<<GF, 256, 0, st>>>((const float*)x, partial, seg_off, c, nseg, weight, bias, eps, scale, shift);
  } else {
    k_instnorm_partial<__nv_bfloat16><<<grid, 256, 0, st>>>((const __nv_bfloat16*)x, seg_off, c, partial);
    k_instnorm_final<__nv_bfloat16><<<GF, 256, 0, st>>>((const __nv_bfloat16*)x, partial, seg_off, c, nseg, weight, bias, eps,
                                                       scale, shift);
  }
  b2_count_launches(1);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// coords == NULL -> one (scale, shift) row for all; else row batch index selects scale[s, :]
int b2_affine_act(const void* x, int dtype, const int* coords, long long n, int c, const float* scale, const float* shift,
                  const void* residual, int act, void* y, void* stream) {
  if (n == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int G = b2_div_up(n * c, 256);
  if (dtype == 0)
    k_affine_act<float><<<G, 256, 0, st>>>((const float*)x, coords, n, c, scale, shift, (const float*)residual, act,
                                           (float*)y);
  else if (c % 8 == 0 && coords == nullptr && scale != nullptr && 256 % (c / 8) == 0 && c % 32 == 0)
    k_affine_act_bf16x8_v4<<<b2_div_up(n * (c / 8), 1024), 256, 0, st>>>((const uint4*)x, n * (c / 8), c, scale, shift,
                                                                       (const uint4*)residual, act, (uint4*)y);
  else if (c % 8 == 0)
    k_affine_act_bf16x8<<<b2_div_up(n * (c / 8), 256), 256, 0, st>>>((const uint4*)x, coords, n, c, scale, shift,
                                                                   (const uint4*)residual, act, (uint4*)y);
  else
    k_affine_act<__nv_bfloat16><<<G, 256, 0, st>>>((const __nv_bfloat16*)x, coords, n, c, scale, shift,
                                                   (const __nv_bfloat16*)residual, act, (__nv_bfloat16*)y);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_cast(const void* x, int src_dtype, long long n, void* y, void* stream) {
  if (n == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int G = b2_div_up(n, 256);
  if (src_dtype == 0)
    k_f32_to_bf16<<<G, 256, 0, st>>>((const float*)x, n, (__nv_bfloat16*)y);
  else
    k_bf16_to_f32<<<G, 256, 0, st>>>((const __nv_bfloat16*)x, n, (float*)y);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_quantize_points(const float* pts, long long n, int dim, int batch, float voxel_size, float feat_div,
                       long long row0, int* coords, float* feats, void* stream) {
  B2_CHECK_ARG(dim >= 3, "points need at least xyz");
  if (n == 0) return B2_OK;
  k_quantize_points<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(pts, (int)n, dim, batch, voxel_size, feat_div,
                                                                       (int)row0, coords, feats);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
