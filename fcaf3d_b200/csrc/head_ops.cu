// Head tail of the detector: trilinear score interpolation, per-scene top-k masks (pruning and
// nms_pre), head post-processing (exp / scale / row max), box decoding and candidate assembly.
//
// Reference behaviour: mmdet3d/models/dense_heads/fcaf3d_neck_with_head.py
//   _prune :110-126 (features_at_coordinates + per-scene topk mask), forward_single :256-279,
//   _get_bboxes_single :227-253, _bbox_pred_to_bbox :281-330, _nms :332-374 (score threshold).
#include "common.cuh"
#include <cooperative_groups.h>
#include <stdlib.h>
namespace cg = cooperative_groups;

// ---------------------------------------------------------------------------
// features_at_coordinates: 8-corner trilinear interpolation on the source grid
// (ME v0.5.4 semantics): absent corners contribute 0.
// ---------------------------------------------------------------------------
__global__ void k_interp(const B2Slot* __restrict__ table, unsigned int cap_mask, const float* __restrict__ feats, int c,
                         int ts, const float* __restrict__ query, int m, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const float4 q = __ldg(reinterpret_cast<const float4*>(query) + i);
  const int b = (int)q.x;
  const float fts = (float)ts;
  const int bx = (int)floorf(q.y / fts) * ts, by = (int)floorf(q.z / fts) * ts, bz = (int)floorf(q.w / fts) * ts;
  int rows[8];
  float wts[8];
#pragma unroll
  for (int k = 0; k < 8; ++k) {
    const int x = bx + ((k & 1) ? ts : 0), y = by + ((k & 2) ? ts : 0), z = bz + ((k & 4) ? ts : 0);
    const float wx = 1.f - fabsf(q.y - (float)x) / fts;
    const float wy = 1.f - fabsf(q.z - (float)y) / fts;
    const float wz = 1.f - fabsf(q.w - (float)z) / fts;
    wts[k] = wx * wy * wz;
    int r = -1;
    if (wts[k] != 0.f && b2_coord_in_range(b, x, y, z)) r = b2_table_find(table, cap_mask, b2_pack_key(b, x, y, z));
    rows[k] = r;
  }
  for (int ch = 0; ch < c; ++ch) {
    float acc = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k)
      if (rows[k] >= 0) acc += wts[k] * feats[(long long)rows[k] * c + ch];
    out[(long long)i * c + ch] = acc;
  }
}

// ---------------------------------------------------------------------------
// Per-segment top-k mask by radix select.  One CTA per segment (scene).
// mask[i] = 1 for the k largest scores of the segment; ties at the threshold are
// resolved towards the lowest row (torch.topk leaves this unspecified).
// ---------------------------------------------------------------------------
__device__ __forceinline__ unsigned int f32_ordered(float v) {
  unsigned int u = __float_as_uint(v);
  if (u == 0x80000000u) u = 0u;  // -0.0 == +0.0 (torch.topk compares values, not bit patterns)
  return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

#define TOPK_BINS 2048

__global__ void __launch_bounds__(1024) k_topk_mask(const float* __restrict__ scores, const int* __restrict__ seg_off,
                                                    int k, uint8_t* __restrict__ mask) {
  __shared__ unsigned int hist[TOPK_BINS];
  __shared__ unsigned int s_prefix, s_need;
  __shared__ int s_warp[32];
  __shared__ int s_run;
  const int beg = seg_off[blockIdx.x], n = seg_off[blockIdx.x + 1] - beg;
  if (n <= 0) return;
  if (k >= n) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) mask[beg + i] = 1;
    return;
  }
  if (k <= 0) {
    for (int i = threadIdx.x; i < n; i += blockDim.x) mask[beg + i] = 0;
    return;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  // radix select on the order-preserving key, 11 + 11 + 10 bits, most significant digit first
  unsigned int prefix = 0, need = (unsigned int)k;
  const int shifts[3] = {21, 10, 0}, widths[3] = {11, 11, 10};
  for (int pass = 0; pass < 3; ++pass) {
    const int shift = shifts[pass], bins = 1 << widths[pass];
    const unsigned int hi_mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + widths[pass]));
    for (int i = threadIdx.x; i < TOPK_BINS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    // 4 independent loads per thread per trip (the single CTA of a segment is latency bound otherwise)
    for (int i0 = threadIdx.x; i0 - lane < n; i0 += blockDim.x * 4) {  // trip count is warp-uniform
      unsigned int u4[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = i0 + e * (int)blockDim.x;
        u4[e] = (i < n) ? f32_ordered(scores[beg + i]) : 0u;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = i0 + e * (int)blockDim.x;
        const bool valid = (i < n) && ((u4[e] & hi_mask) == prefix);
        // warp-aggregated histogram update: scores cluster in a few bins, plain shared atomics would serialise
        const unsigned int vm = __ballot_sync(0xffffffffu, valid);
        if (valid) {
          const unsigned int d = (u4[e] >> shift) & (unsigned int)(bins - 1);
          const unsigned int peers = __match_any_sync(vm, d);
          if (lane == __ffs(peers) - 1) atomicAdd(&hist[d], (unsigned int)__popc(peers));
        }
      }
    }
    __syncthreads();
    if (warp == 0) {
      // lane l owns bins [l*per, (l+1)*per); find the bin that holds the need-th largest element
      const int per = bins / 32;
      unsigned int mine = 0;
      for (int j = 0; j < per; ++j) mine += hist[lane * per + j];
      unsigned int above = 0;  // elements in lanes with higher bins
      for (int l = 31; l >= 0; --l) {
        const unsigned int v = __shfl_sync(0xffffffffu, mine, l);
        if (l > lane) above += v;
      }
      if (above < need && above + mine >= need) {
        unsigned int acc = above;
        int d = lane * per + per - 1;
        for (; d > lane * per; --d) {
          if (acc + hist[d] >= need) break;
          acc += hist[d];
        }
        s_prefix = prefix | ((unsigned int)d << shift);
        s_need = need - acc;
      }
    }
    __syncthreads();
    prefix = s_prefix;
    need = s_need;
    __syncthreads();
  }
  // prefix == ordered bits of the k-th largest score; `need` of the rows equal to it are taken, lowest rows first
  if (threadIdx.x == 0) s_run = 0;
  __syncthreads();
  // ordered pass: every thread owns 4 consecutive rows, ties are ranked in row order
  for (int base = 0; base < n; base += blockDim.x * 4) {
    const int i = base + threadIdx.x * 4;
    unsigned int u4[4];
    int eqc = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      u4[e] = (i + e < n) ? f32_ordered(scores[beg + i + e]) : 0u;
      eqc += (i + e < n && u4[e] == prefix) ? 1 : 0;
    }
    int incl = eqc;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int rank = s_run + incl - eqc;
    for (int w = 0; w < warp; ++w) rank += s_warp[w];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (i + e < n) {
        const bool eq = u4[e] == prefix;
        mask[beg + i + e] = (u4[e] > prefix || (eq && rank < (int)need)) ? 1 : 0;
        rank += eq ? 1 : 0;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_warp[w];
      s_run += t;
    }
    __syncthreads();
  }
}

// Same selection for LONG segments (the prune step: 8 scenes x 180 k scores): a thread-block cluster of CL CTAs shares one
// segment.  Every CTA histograms its slice of the rows, the slices' histograms are added into CTA 0's shared memory
// through distributed shared memory, CTA 0 picks the digit, everybody reads it back; for the ties at the threshold the
// CTAs exchange their counts of equal rows so that the lowest rows win exactly as in the one-CTA kernel (bit-identical
// masks).  Eight cluster barriers per call instead of one CTA walking the whole segment four times.
template <int CL>
__global__ void __launch_bounds__(1024) k_topk_mask_cl(const float* __restrict__ scores, const int* __restrict__ seg_off,
                                                       int k, uint8_t* __restrict__ mask) {
  cg::cluster_group cluster = cg::this_cluster();
  __shared__ unsigned int hist[TOPK_BINS];
  __shared__ unsigned int merged[3][TOPK_BINS];  // used in CTA 0 only: cluster-wide histogram of each pass
  __shared__ unsigned int s_prefix, s_need;
  __shared__ int s_warp[32];
  __shared__ int s_run, s_eq;
  const int crank = (int)cluster.block_rank();
  const int seg = blockIdx.x / CL;
  const int beg = seg_off[seg], n = seg_off[seg + 1] - beg;
  const int per = ((n + CL - 1) / CL + 3) & ~3;
  const int lo = min(crank * per, n), hi = min(lo + per, n);
  if (n <= 0) return;
  if (k >= n || k <= 0) {  // cluster-uniform
    for (int i = lo + threadIdx.x; i < hi; i += blockDim.x) mask[beg + i] = (k >= n) ? 1 : 0;
    return;
  }
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  if (crank == 0)
    for (int i = threadIdx.x; i < 3 * TOPK_BINS; i += blockDim.x) (&merged[0][0])[i] = 0;
  cluster.sync();
  unsigned int* merged0 = cluster.map_shared_rank(&merged[0][0], 0);
  const unsigned int* prefix0 = cluster.map_shared_rank(&s_prefix, 0);
  const unsigned int* need0 = cluster.map_shared_rank(&s_need, 0);
  unsigned int prefix = 0, need = (unsigned int)k;
  const int shifts[3] = {21, 10, 0}, widths[3] = {11, 11, 10};
  const int m = hi - lo;
  for (int pass = 0; pass < 3; ++pass) {
    const int shift = shifts[pass], bins = 1 << widths[pass];
    const unsigned int hi_mask = pass == 0 ? 0u : (0xFFFFFFFFu << (shift + widths[pass]));
    for (int i = threadIdx.x; i < TOPK_BINS; i += blockDim.x) hist[i] = 0;
    __syncthreads();
    for (int i0 = threadIdx.x; i0 - lane < m; i0 += blockDim.x * 4) {
      unsigned int u4[4];
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = i0 + e * (int)blockDim.x;
        u4[e] = (i < m) ? f32_ordered(scores[beg + lo + i]) : 0u;
      }
#pragma unroll
      for (int e = 0; e < 4; ++e) {
        const int i = i0 + e * (int)blockDim.x;
        const bool valid = (i < m) && ((u4[e] & hi_mask) == prefix);
        const unsigned int vm = __ballot_sync(0xffffffffu, valid);
        if (valid) {
          const unsigned int d = (u4[e] >> shift) & (unsigned int)(bins - 1);
          const unsigned int peers = __match_any_sync(vm, d);
          if (lane == __ffs(peers) - 1) atomicAdd(&hist[d], (unsigned int)__popc(peers));
        }
      }
    }
    __syncthreads();
    for (int i = threadIdx.x; i < bins; i += blockDim.x) {
      const unsigned int v = hist[i];
      if (v) atomicAdd(merged0 + pass * TOPK_BINS + i, v);
    }
    cluster.sync();
    if (crank == 0 && warp == 0) {
      const unsigned int* h = &merged[pass][0];
      const int per_l = bins / 32;
      unsigned int mine = 0;
      for (int j = 0; j < per_l; ++j) mine += h[lane * per_l + j];
      unsigned int above = 0;
      for (int l = 31; l >= 0; --l) {
        const unsigned int v = __shfl_sync(0xffffffffu, mine, l);
        if (l > lane) above += v;
      }
      if (above < need && above + mine >= need) {
        unsigned int acc = above;
        int d = lane * per_l + per_l - 1;
        for (; d > lane * per_l; --d) {
          if (acc + h[d] >= need) break;
          acc += h[d];
        }
        s_prefix = prefix | ((unsigned int)d << shift);
        s_need = need - acc;
      }
    }
    cluster.sync();
    prefix = *prefix0;
    need = *need0;
  }
  // ties: rows equal to the threshold are taken in row order; the slices are in row order, so a CTA starts its count at
  // the number of equal rows in the slices before it
  int eq_local = 0;
  for (int i = threadIdx.x; i < m; i += blockDim.x) eq_local += (f32_ordered(scores[beg + lo + i]) == prefix) ? 1 : 0;
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) eq_local += __shfl_xor_sync(0xffffffffu, eq_local, o);
  if (lane == 0) s_warp[warp] = eq_local;
  __syncthreads();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_warp[w];
    s_eq = t;
  }
  cluster.sync();
  if (threadIdx.x == 0) {
    int t = 0;
    for (int r = 0; r < crank; ++r) t += *cluster.map_shared_rank(&s_eq, r);
    s_run = t;
  }
  __syncthreads();
  for (int base = 0; base < m; base += blockDim.x * 4) {
    const int i = base + threadIdx.x * 4;
    unsigned int u4[4];
    int eqc = 0;
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      u4[e] = (i + e < m) ? f32_ordered(scores[beg + lo + i + e]) : 0u;
      eqc += (i + e < m && u4[e] == prefix) ? 1 : 0;
    }
    int incl = eqc;
#pragma unroll
    for (int d = 1; d < 32; d <<= 1) {
      const int t = __shfl_up_sync(0xffffffffu, incl, d);
      if (lane >= d) incl += t;
    }
    if (lane == 31) s_warp[warp] = incl;
    __syncthreads();
    int rank = s_run + incl - eqc;
    for (int w = 0; w < warp; ++w) rank += s_warp[w];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      if (i + e < m) {
        const bool eq = u4[e] == prefix;
        mask[beg + lo + i + e] = (u4[e] > prefix || (eq && rank < (int)need)) ? 1 : 0;
        rank += eq ? 1 : 0;
      }
    }
    __syncthreads();
    if (threadIdx.x == 0) {
      int t = 0;
      for (int w = 0; w < (int)(blockDim.x >> 5); ++w) t += s_warp[w];
      s_run += t;
    }
    __syncthreads();
  }
  cluster.sync();  // nobody leaves while a peer may still read its shared memory
}

// ---------------------------------------------------------------------------
// Head post-processing on the fused head output feat[n, 1 + n_reg + n_cls]
// (centerness | regression | class logits).  forward_single :257-269.
// ---------------------------------------------------------------------------
__global__ void k_head_post(const float* __restrict__ feat, int n, int feat_ld, int n_reg, int n_cls,
                            const float* __restrict__ scale_ptr,
                            float* __restrict__ centerness, float* __restrict__ bbox_pred, float* __restrict__ cls,
                            float* __restrict__ prune_score, float* __restrict__ max_score) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const float* f = feat + (long long)i * feat_ld;
  const float ctr = f[0];
  const float scale = scale_ptr ? __ldg(scale_ptr) : 1.f;
  centerness[i] = ctr;
  for (int r = 0; r < n_reg; ++r) {
    const float v = f[1 + r];
    bbox_pred[(long long)i * n_reg + r] = (r < 6) ? expf(scale * v) : v;
  }
  const float sc = 1.f / (1.f + expf(-ctr));
  float mlogit = -INFINITY, mscore = -INFINITY;
  for (int k = 0; k < n_cls; ++k) {
    const float v = f[1 + n_reg + k];
    cls[(long long)i * n_cls + k] = v;
    mlogit = fmaxf(mlogit, v);
    mscore = fmaxf(mscore, (1.f / (1.f + expf(-v))) * sc);
  }
  prune_score[i] = mlogit;
  if (max_score) max_score[i] = mscore;
}

// Same arithmetic with coalesced traffic: a CTA stages 256 rows of feat in shared memory (row pitch feat_ld + 1 words),
// one thread per row does the per-row scalars from there, and the [rows, n_reg] / [rows, n_cls] blocks - contiguous in
// the outputs - are written with consecutive threads on consecutive words.
#define HP_ROWS 256
__global__ void __launch_bounds__(256) k_head_post_tiled(const float* __restrict__ feat, int n, int feat_ld, int n_reg,
                                                         int n_cls, const float* __restrict__ scale_ptr,
                                                         float* __restrict__ centerness, float* __restrict__ bbox_pred,
                                                         float* __restrict__ cls, float* __restrict__ prune_score,
                                                         float* __restrict__ max_score) {
  extern __shared__ float s_feat[];
  const int row0 = blockIdx.x * HP_ROWS;
  const int rows = min(HP_ROWS, n - row0);
  const int ldp = feat_ld + 1;
  const float* src = feat + (long long)row0 * feat_ld;
  for (int t = threadIdx.x; t < rows * feat_ld; t += 256) {
    const int r = t / feat_ld;
    s_feat[r * ldp + (t - r * feat_ld)] = __ldg(src + t);
  }
  __syncthreads();
  const float scale = scale_ptr ? __ldg(scale_ptr) : 1.f;
  if (threadIdx.x < rows) {
    const float* f = s_feat + threadIdx.x * ldp;
    const int i = row0 + threadIdx.x;
    const float ctr = f[0];
    centerness[i] = ctr;
    const float sc = 1.f / (1.f + expf(-ctr));
    float mlogit = -INFINITY, mscore = -INFINITY;
    for (int k = 0; k < n_cls; ++k) {
      const float v = f[1 + n_reg + k];
      mlogit = fmaxf(mlogit, v);
      mscore = fmaxf(mscore, (1.f / (1.f + expf(-v))) * sc);
    }
    prune_score[i] = mlogit;
    if (max_score) max_score[i] = mscore;
  }
  float* bp = bbox_pred + (long long)row0 * n_reg;
  for (int t = threadIdx.x; t < rows * n_reg; t += 256) {
    const int r = t / n_reg, k = t - r * n_reg;
    const float v = s_feat[r * ldp + 1 + k];
    bp[t] = (k < 6) ? expf(scale * v) : v;
  }
  float* cp = cls + (long long)row0 * n_cls;
  for (int t = threadIdx.x; t < rows * n_cls; t += 256) {
    const int r = t / n_cls, k = t - r * n_cls;
    cp[t] = s_feat[r * ldp + 1 + n_reg + k];
  }
}

// distances -> box (_bbox_pred_to_bbox :281-330).  yaw_mode: 0 none(6), 1 naive, 2 sin-cos, 3 fcaf3d (Mobius)
__device__ __forceinline__ void decode_box(const float* __restrict__ p, const float* __restrict__ bp, int n_reg,
                                           int yaw_mode, float* __restrict__ o) {
  const float xc = p[0] + (bp[1] - bp[0]) / 2;
  const float yc = p[1] + (bp[3] - bp[2]) / 2;
  const float zc = p[2] + (bp[5] - bp[4]) / 2;
  o[0] = xc;
  o[1] = yc;
  o[2] = zc;
  if (n_reg == 6 || yaw_mode == 0) {
    o[3] = bp[0] + bp[1];
    o[4] = bp[2] + bp[3];
    o[5] = bp[4] + bp[5];
    o[6] = 0.f;
  } else if (yaw_mode == 1) {
    o[3] = bp[0] + bp[1];
    o[4] = bp[2] + bp[3];
    o[5] = bp[4] + bp[5];
    o[6] = bp[6];
  } else if (yaw_mode == 2) {
    const float norm = powf(powf(bp[6], 2.f) + powf(bp[7], 2.f), 0.5f);
    o[3] = bp[0] + bp[1];
    o[4] = bp[2] + bp[3];
    o[5] = bp[4] + bp[5];
    o[6] = atan2f(bp[6] / norm, bp[7] / norm);
  } else {
    const float scale = bp[0] + bp[1] + bp[2] + bp[3];
    const float q = expf(sqrtf(bp[6] * bp[6] + bp[7] * bp[7]));
    o[3] = scale / (1 + q);
    o[4] = scale / (1 + q) * q;
    o[5] = bp[5] + bp[4];
    o[6] = 0.5f * atan2f(bp[6], bp[7]);
  }
}

__global__ void k_decode_boxes(const float* __restrict__ points, const float* __restrict__ bbox_pred, int n, int n_reg,
                               int yaw_mode, float* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  decode_box(points + (long long)i * 3, bbox_pred + (long long)i * n_reg, n_reg, yaw_mode, out + (long long)i * 7);
}

// Selected rows of one level -> the scene-major candidate arrays.
// Row i (scene s) lands at dst_base[s] + (pos[i] - pos[seg_off[s]]).
__global__ void k_emit_candidates(const int* __restrict__ coords, const float* __restrict__ centerness,
                                  const float* __restrict__ bbox_pred, const float* __restrict__ cls,
                                  const uint8_t* __restrict__ mask, const int* __restrict__ pos,
                                  const int* __restrict__ seg_off, const int* __restrict__ dst_base, int n, int n_reg,
                                  int n_cls, int yaw_mode, float voxel_size, float* __restrict__ cand_boxes,
                                  float* __restrict__ cand_scores) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !mask[i]) return;
  const int4 c = __ldg(reinterpret_cast<const int4*>(coords) + i);
  const int s = c.x;
  const int j = dst_base[s] + pos[i] - pos[seg_off[s]];
  float p[3] = {(float)c.y * voxel_size, (float)c.z * voxel_size, (float)c.w * voxel_size};
  decode_box(p, bbox_pred + (long long)i * n_reg, n_reg, yaw_mode, cand_boxes + (long long)j * 7);
  const float sc = 1.f / (1.f + expf(-centerness[i]));
  for (int k = 0; k < n_cls; ++k)
    cand_scores[(long long)j * n_cls + k] = (1.f / (1.f + expf(-cls[(long long)i * n_cls + k]))) * sc;
}

// flags in (scene, class, candidate) order: flag[n_cls*cand_off[s] + c*M_s + j] = score[cand_off[s]+j][c] > thr
__global__ void k_class_flags(const float* __restrict__ cand_scores, const int* __restrict__ cand_off, int nscene,
                              int n_cls, float thr, int total, uint8_t* __restrict__ flags) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total * n_cls) return;
  // find scene: n_cls*cand_off[s] <= t
  int lo = 0, hi = nscene - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if ((long long)n_cls * cand_off[mid] <= t) lo = mid; else hi = mid - 1;
  }
  const int s = lo;
  const int ms = cand_off[s + 1] - cand_off[s];
  const int local = t - n_cls * cand_off[s];
  const int c = local / ms, j = local - c * ms;
  flags[t] = cand_scores[(long long)(cand_off[s] + j) * n_cls + c] > thr ? 1 : 0;
}

// compact the flagged (scene, class, candidate) triples into NMS segment arrays
__global__ void k_class_compact(const float* __restrict__ cand_boxes, const float* __restrict__ cand_scores,
                                const int* __restrict__ cand_off, int nscene, int n_cls, int total,
                                const uint8_t* __restrict__ flags, const int* __restrict__ pos,
                                float* __restrict__ seg_boxes, float* __restrict__ seg_scores, int* __restrict__ seg_src) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= total * n_cls || !flags[t]) return;
  int lo = 0, hi = nscene - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if ((long long)n_cls * cand_off[mid] <= t) lo = mid; else hi = mid - 1;
  }
  const int s = lo;
  const int ms = cand_off[s + 1] - cand_off[s];
  const int local = t - n_cls * cand_off[s];
  const int c = local / ms, j = local - c * ms;
  const int src = cand_off[s] + j;
  const int d = pos[t];
  const float* b = cand_boxes + (long long)src * 7;
  float* o = seg_boxes + (long long)d * 7;
#pragma unroll
  for (int u = 0; u < 7; ++u) o[u] = b[u];
  seg_scores[d] = cand_scores[(long long)src * n_cls + c];
  seg_src[d] = src;
}

// segment offsets: seg_off[q] = pos[start_of_segment q] (q = s*n_cls + c), seg_off[nseg] = total kept
__global__ void k_segment_offsets(const int* __restrict__ pos, const int* __restrict__ total_flagged,
                                  const int* __restrict__ cand_off, int nscene, int n_cls, int total_flags,
                                  int* __restrict__ seg_off) {
  const int q = blockIdx.x * blockDim.x + threadIdx.x;
  const int nseg = nscene * n_cls;
  if (q > nseg) return;
  if (q == nseg) {
    seg_off[q] = *total_flagged;
    return;
  }
  const int s = q / n_cls, c = q - s * n_cls;
  const int ms = cand_off[s + 1] - cand_off[s];
  const long long start = (long long)n_cls * cand_off[s] + (long long)c * ms;
  seg_off[q] = (start < total_flags) ? pos[start] : *total_flagged;
}

// kept entries (sorted order inside segments) -> final detections
__global__ void k_collect_dets(const float* __restrict__ sorted_boxes, const float* __restrict__ seg_scores,
                               const int* __restrict__ order, const int* __restrict__ seg_off, int nseg, int n_cls,
                               const uint8_t* __restrict__ keep, const int* __restrict__ pos, int n,
                               float* __restrict__ out_boxes, float* __restrict__ out_scores, int* __restrict__ out_labels,
                               int* __restrict__ out_scene) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n || !keep[i]) return;
  int lo = 0, hi = nseg - 1;
  while (lo < hi) {
    const int mid = (lo + hi + 1) >> 1;
    if (seg_off[mid] <= i) lo = mid; else hi = mid - 1;
  }
  const int q = lo;
  const int d = pos[i];
  const float* b = sorted_boxes + (long long)i * 7;
  float* o = out_boxes + (long long)d * 7;
#pragma unroll
  for (int u = 0; u < 7; ++u) o[u] = b[u];
  out_scores[d] = seg_scores[seg_off[q] + order[i]];
  out_labels[d] = q % n_cls;
  out_scene[d] = q / n_cls;
}

extern "C" {

int b2_interp(const void* table, long long cap, const float* feats, int c, int tensor_stride, const float* query,
              long long m, float* out, void* stream) {
  B2_CHECK_ARG(tensor_stride > 0 && c > 0, "bad interpolation arguments");
  if (m == 0) return B2_OK;
  k_interp<<<b2_div_up(m, 128), 128, 0, (cudaStream_t)stream>>>((const B2Slot*)table, (unsigned)(cap - 1), feats, c,
                                                              tensor_stride, query, (int)m, out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// n_total: rows over all segments (host-known); long segments take the 8-CTA cluster kernel
int b2_topk_mask_n(const float* scores, const int* seg_off, int nseg, int k, long long n_total, uint8_t* mask,
                   void* stream) {
  if (nseg == 0) return B2_OK;
  static const int cl_env = getenv("B2_TOPK_CLUSTER") ? atoi(getenv("B2_TOPK_CLUSTER")) : -1;
  const bool use_cl = cl_env >= 0 ? cl_env > 0 : (n_total / nseg >= 32768);
  if (use_cl) {
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)nseg * 8);
    cfg.blockDim = dim3(1024);
    cfg.stream = (cudaStream_t)stream;
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeClusterDimension;
    attr[0].val.clusterDim.x = 8;
    attr[0].val.clusterDim.y = 1;
    attr[0].val.clusterDim.z = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
    B2_CUDA(cudaLaunchKernelEx(&cfg, k_topk_mask_cl<8>, scores, seg_off, k, mask));
  } else {
    k_topk_mask<<<nseg, 1024, 0, (cudaStream_t)stream>>>(scores, seg_off, k, mask);
  }
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_topk_mask(const float* scores, const int* seg_off, int nseg, int k, uint8_t* mask, void* stream) {
  return b2_topk_mask_n(scores, seg_off, nseg, k, 0, mask, stream);
}

int b2_head_post(const float* feat, long long n, int feat_ld, int n_reg, int n_cls, const float* scale,
                 float* centerness,
                 float* bbox_pred, float* cls, float* prune_score, float* max_score, void* stream) {
  if (n == 0) return B2_OK;
  B2_CHECK_ARG(feat_ld >= 1 + n_reg + n_cls, "feat_ld smaller than 1 + n_reg + n_cls");
  const size_t smem = (size_t)HP_ROWS * (feat_ld + 1) * sizeof(float);
  if (smem <= 48 * 1024)
    k_head_post_tiled<<<b2_div_up(n, HP_ROWS), 256, smem, (cudaStream_t)stream>>>(feat, (int)n, feat_ld, n_reg, n_cls, scale,
                                                                                centerness, bbox_pred, cls, prune_score,
                                                                                max_score);
  else
    k_head_post<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(feat, (int)n, feat_ld, n_reg, n_cls, scale, centerness,
                                                                   bbox_pred, cls, prune_score, max_score);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_decode_boxes(const float* points, const float* bbox_pred, long long n, int n_reg, int yaw_mode, float* out,
                    void* stream) {
  B2_CHECK_ARG(n_reg == 6 || n_reg == 7 || n_reg == 8, "n_reg must be 6, 7 or 8");
  if (n == 0) return B2_OK;
  k_decode_boxes<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(points, bbox_pred, (int)n, n_reg, yaw_mode, out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_emit_candidates(const int* coords, const float* centerness, const float* bbox_pred, const float* cls,
                       const uint8_t* mask, const int* pos, const int* seg_off, const int* dst_base, long long n,
                       int n_reg, int n_cls, int yaw_mode, float voxel_size, float* cand_boxes, float* cand_scores,
                       void* stream) {
  if (n == 0) return B2_OK;
  k_emit_candidates<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(coords, centerness, bbox_pred, cls, mask, pos,
                                                                       seg_off, dst_base, (int)n, n_reg, n_cls, yaw_mode,
                                                                       voxel_size, cand_boxes, cand_scores);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// Builds the (scene, class) NMS segments from the scene-major candidate arrays.
// flags/pos: scratch of total*n_cls entries; seg_* outputs sized total*n_cls; seg_off[nscene*n_cls + 1].
int b2_class_segments(const float* cand_boxes, const float* cand_scores, const int* cand_off, int nscene, int n_cls,
                      long long total, float score_thr, uint8_t* flags, int* pos, int* n_flagged, float* seg_boxes,
                      float* seg_scores, int* seg_src, int* seg_off, void* ws, size_t ws_bytes, void* stream);

int b2_collect_dets(const float* sorted_boxes, const float* seg_scores, const int* order, const int* seg_off, int nseg,
                    int n_cls, const uint8_t* keep, const int* pos, long long n, float* out_boxes, float* out_scores,
                    int* out_labels, int* out_scene, void* stream) {
  if (n == 0) return B2_OK;
  k_collect_dets<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(sorted_boxes, seg_scores, order, seg_off, nseg,
                                                                    n_cls, keep, pos, (int)n, out_boxes, out_scores,
                                                                    out_labels, out_scene);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"

int b2_scan_flags_launch(const uint8_t* flags, long long n, int* pos, int* total, void* ws, size_t ws_bytes,
                         cudaStream_t st);

extern "C" int b2_class_segments(const float* cand_boxes, const float* cand_scores, const int* cand_off, int nscene,
                                 int n_cls, long long total, float score_thr, uint8_t* flags, int* pos, int* n_flagged,
                                 float* seg_boxes, float* seg_scores, int* seg_src, int* seg_off, void* ws,
                                 size_t ws_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  const long long nf = total * n_cls;
  const int nseg = nscene * n_cls;
  if (nf > 0) {
    k_class_flags<<<b2_div_up(nf, 256), 256, 0, st>>>(cand_scores, cand_off, nscene, n_cls, score_thr, (int)total, flags);
    B2_LAUNCH_CHECK();
  }
  int rc = b2_scan_flags_launch(flags, nf, pos, n_flagged, ws, ws_bytes, st);
  if (rc) return rc;
  if (nf > 0) {
    k_class_compact<<<b2_div_up(nf, 256), 256, 0, st>>>(cand_boxes, cand_scores, cand_off, nscene, n_cls, (int)total,
                                                       flags, pos, seg_boxes, seg_scores, seg_src);
    B2_LAUNCH_CHECK();
  }
  k_segment_offsets<<<b2_div_up(nseg + 1, 256), 256, 0, st>>>(pos, n_flagged, cand_off, nscene, n_cls, (int)nf, seg_off);
  B2_LAUNCH_CHECK();
  return B2_OK;
}
