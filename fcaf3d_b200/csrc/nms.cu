// Batched BEV NMS (rotated "nms_gpu" and axis-aligned "nms_normal_gpu" flavours).
//
// Same per-pair arithmetic as the reference's pcdet_nms extension
// (mmdet3d/ops/pcdet_nms/src/iou3d_nms_kernel.cu:104-234 rotated BEV overlap / IoU,
//  :314-325 axis-aligned BEV IoU, :267-372 64x64 suppression tiles,
//  mmdet3d/ops/pcdet_nms/src/iou3d_nms.cpp:116-132 greedy reduce on the host),
// restructured for B200: every (scene, class) segment of a batch goes through ONE launch per
// stage, only upper-triangular tiles are evaluated, per-box trigonometry is hoisted out of the
// pair loop, far-apart pairs are rejected by a bounding-circle test, and the greedy reduce runs
// on the device (no cudaMalloc / D2H / host loop per class).
#include "common.cuh"

#define NMS_EPS 1e-8f
#define NMS_TILE 64

struct P2 {
  float x, y;
};

__device__ __forceinline__ float cross3(const P2& p1, const P2& p2, const P2& p0) {
  return (p1.x - p0.x) * (p2.y - p0.y) - (p2.x - p0.x) * (p1.y - p0.y);
}
__device__ __forceinline__ float cross2(const P2& a, const P2& b) { return a.x * b.y - a.y * b.x; }

// per-box quantities that the reference recomputes for every pair
struct BoxPrep {
  float cx, cy, dx, dy;     // centre and full sizes
  float cosm, sinm;         // cos(-heading), sin(-heading) used by the corner-in-box test
  P2 c[4];                  // rotated corners
  float rad;                // circumscribed radius (for the conservative far-apart rejection)
};

__device__ __forceinline__ void rotate_about(const P2& ctr, float ca, float sa, P2& p) {
  const float nx = (p.x - ctr.x) * ca + (p.y - ctr.y) * (-sa) + ctr.x;
  const float ny = (p.x - ctr.x) * sa + (p.y - ctr.y) * ca + ctr.y;
  p.x = nx;
  p.y = ny;
}

__device__ __forceinline__ void prep_box(const float* __restrict__ b, BoxPrep& o) {
  const float ang = b[6];
  const float hx = b[3] / 2, hy = b[4] / 2;
  const float x1 = b[0] - hx, y1 = b[1] - hy, x2 = b[0] + hx, y2 = b[1] + hy;
  P2 ctr;
  ctr.x = b[0];
  ctr.y = b[1];
  o.cx = b[0];
  o.cy = b[1];
  o.dx = b[3];
  o.dy = b[4];
  o.c[0].x = x1; o.c[0].y = y1;
  o.c[1].x = x2; o.c[1].y = y1;
  o.c[2].x = x2; o.c[2].y = y2;
  o.c[3].x = x1; o.c[3].y = y2;
  const float ca = cosf(ang), sa = sinf(ang);
#pragma unroll
  for (int k = 0; k < 4; ++k) rotate_about(ctr, ca, sa, o.c[k]);
  o.cosm = cosf(-ang);
  o.sinm = sinf(-ang);
  o.rad = sqrtf(hx * hx + hy * hy);
}

__device__ __forceinline__ int seg_intersect(const P2& p1, const P2& p0, const P2& q1, const P2& q0, P2& ans) {
  // bounding-rectangle rejection of the two segments
  const int hit = fminf(p0.x, p1.x) <= fmaxf(q0.x, q1.x) && fminf(q0.x, q1.x) <= fmaxf(p0.x, p1.x) &&
                  fminf(p0.y, p1.y) <= fmaxf(q0.y, q1.y) && fminf(q0.y, q1.y) <= fmaxf(p0.y, p1.y);
  if (!hit) return 0;
  const float s1 = cross3(q0, p1, p0);
  const float s2 = cross3(p1, q1, p0);
  const float s3 = cross3(p0, q1, q0);
  const float s4 = cross3(q1, p1, q0);
  if (!(s1 * s2 > 0 && s3 * s4 > 0)) return 0;
  const float s5 = cross3(q1, p1, p0);
  if (fabsf(s5 - s1) > NMS_EPS) {
    ans.x = (s5 * q0.x - s1 * q1.x) / (s5 - s1);
    ans.y = (s5 * q0.y - s1 * q1.y) / (s5 - s1);
  } else {
    const float a0 = p0.y - p1.y, b0 = p1.x - p0.x, c0 = p0.x * p1.y - p1.x * p0.y;
    const float a1 = q0.y - q1.y, b1 = q1.x - q0.x, c1 = q0.x * q1.y - q1.x * q0.y;
    const float D = a0 * b1 - a1 * b0;
    ans.x = (b0 * c1 - b1 * c0) / D;
    ans.y = (a1 * c0 - a0 * c1) / D;
  }
  return 1;
}

__device__ __forceinline__ int corner_in_box(const BoxPrep& box, const P2& p) {
  const float MARGIN = 1e-2f;
  const float rx = (p.x - box.cx) * box.cosm + (p.y - box.cy) * (-box.sinm);
  const float ry = (p.x - box.cx) * box.sinm + (p.y - box.cy) * box.cosm;
  return (fabsf(rx) < box.dx / 2 + MARGIN && fabsf(ry) < box.dy / 2 + MARGIN);
}

__device__ float bev_overlap(const BoxPrep& A, const BoxPrep& B) {
  P2 pts[16];
  P2 ctr;
  ctr.x = 0.f;
  ctr.y = 0.f;
  int cnt = 0;
#pragma unroll
  for (int i = 0; i < 4; ++i) {
#pragma unroll
    for (int j = 0; j < 4; ++j) {
      P2 hit;
      if (seg_intersect(A.c[(i + 1) & 3], A.c[i], B.c[(j + 1) & 3], B.c[j], hit)) {
        pts[cnt] = hit;
        ctr.x = ctr.x + hit.x;
        ctr.y = ctr.y + hit.y;
        cnt++;
      }
    }
  }
#pragma unroll
  for (int k = 0; k < 4; ++k) {
    if (corner_in_box(A, B.c[k])) {
      ctr.x = ctr.x + B.c[k].x;
      ctr.y = ctr.y + B.c[k].y;
      pts[cnt] = B.c[k];
      cnt++;
    }
    if (corner_in_box(B, A.c[k])) {
      ctr.x = ctr.x + A.c[k].x;
      ctr.y = ctr.y + A.c[k].y;
      pts[cnt] = A.c[k];
      cnt++;
    }
  }
  ctr.x /= cnt;
  ctr.y /= cnt;
  // order the polygon vertices by angle around their mean (same exchange sort, same comparison).
  // NOTE: kept in exactly this form on purpose.  Hoisting the atan2f calls out of the comparisons is ~35 % faster on
  // the rotated workload, but it changes how nvcc contracts the area accumulation below into FMAs: 2.6 % of the IoU
  // values then differ from the reference kernel in the last bits (measured, 60 000 pairs), which can flip a
  // keep decision when an IoU sits on the threshold.  Bit-identical suppression masks win over speed here.
  for (int j = 0; j < cnt - 1; ++j) {
    for (int i = 0; i < cnt - j - 1; ++i) {
      if (atan2f(pts[i].y - ctr.y, pts[i].x - ctr.x) > atan2f(pts[i + 1].y - ctr.y, pts[i + 1].x - ctr.x)) {
        const P2 t = pts[i];
        pts[i] = pts[i + 1];
        pts[i + 1] = t;
      }
    }
  }
  float area = 0.f;
  for (int k = 0; k < cnt - 1; ++k) {
    P2 u, v;
    u.x = pts[k].x - pts[0].x;
    u.y = pts[k].y - pts[0].y;
    v.x = pts[k + 1].x - pts[0].x;
    v.y = pts[k + 1].y - pts[0].y;
    area += cross2(u, v);
  }
  return fabsf(area) / 2.0f;
}

__device__ __forceinline__ float iou_rotated(const BoxPrep& A, const BoxPrep& B) {
  const float sa = A.dx * A.dy;
  const float sb = B.dx * B.dy;
  const float s = bev_overlap(A, B);
  return s / fmaxf(sa + sb - s, NMS_EPS);
}

__device__ __forceinline__ float iou_aligned(const float* a, const float* b) {
  const float left = fmaxf(a[0] - a[3] / 2, b[0] - b[3] / 2), right = fminf(a[0] + a[3] / 2, b[0] + b[3] / 2);
  const float top = fmaxf(a[1] - a[4] / 2, b[1] - b[4] / 2), bottom = fminf(a[1] + a[4] / 2, b[1] + b[4] / 2);
  const float width = fmaxf(right - left, 0.f), height = fmaxf(bottom - top, 0.f);
  const float inter = width * height;
  const float Sa = a[3] * a[4];
  const float Sb = b[3] * b[4];
  return inter / fmaxf(Sa + Sb - inter, NMS_EPS);
}

// ---------------------------------------------------------------------------
// Stage 1: per-segment sort by score (descending; ties -> lower index first).
// One CTA per segment, bitonic network in shared memory, <= 4096 entries.
// ---------------------------------------------------------------------------
#define SORT_MAX 4096

__global__ void __launch_bounds__(1024) k_nms_sort(const float* __restrict__ boxes, const float* __restrict__ scores,
                                                   const int* __restrict__ seg_off, float* __restrict__ sorted_boxes,
                                                   int* __restrict__ order) {
  __shared__ float s_key[SORT_MAX];
  __shared__ int s_idx[SORT_MAX];
  const int q = blockIdx.x;
  const int beg = seg_off[q], n = seg_off[q + 1] - beg;
  if (n <= 0) return;
  int m = 1;
  while (m < n) m <<= 1;
  for (int i = threadIdx.x; i < m; i += blockDim.x) {
    s_key[i] = (i < n) ? scores[beg + i] : -INFINITY;
    s_idx[i] = (i < n) ? i : 0x7FFFFFFF;
  }
  __syncthreads();
  for (int k = 2; k <= m; k <<= 1) {
    for (int j = k >> 1; j > 0; j >>= 1) {
      for (int i = threadIdx.x; i < m; i += blockDim.x) {
        const int ixj = i ^ j;
        if (ixj > i) {
          const float ka = s_key[i], kb = s_key[ixj];
          const int ia = s_idx[i], ib = s_idx[ixj];
          // "a before b" in the final order: higher score first, then lower index
          const bool a_first = (ka > kb) || (ka == kb && ia < ib) || (kb != kb && ka == ka);
          const bool up = ((i & k) == 0);
          if (up ? !a_first : a_first) {
            s_key[i] = kb;
            s_key[ixj] = ka;
            s_idx[i] = ib;
            s_idx[ixj] = ia;
          }
        }
      }
      __syncthreads();
    }
  }
  for (int i = threadIdx.x; i < n; i += blockDim.x) {
    const int src = s_idx[i];
    order[beg + i] = src;
    const float* b = boxes + (long long)(beg + src) * 7;
    float* o = sorted_boxes + (long long)(beg + i) * 7;
#pragma unroll
    for (int u = 0; u < 7; ++u) o[u] = b[u];
  }
}

// ---------------------------------------------------------------------------
// Stage 2: suppression bit masks, upper-triangular 64x64 tiles of every segment.
// tile_off[q] = first global tile id of segment q; mask_off[q] = first word of its
// mask (n_q rows x T_q words).  256 threads: 4 threads per row, 16 columns each.
// ---------------------------------------------------------------------------
template <bool ROTATED>
__global__ void __launch_bounds__(256) k_nms_mask(const float* __restrict__ boxes, const int* __restrict__ seg_off,
                                                  const int* __restrict__ tile_off, const long long* __restrict__ mask_off,
                                                  int nseg, int total_tiles, float thresh,
                                                  unsigned long long* __restrict__ mask) {
  __shared__ BoxPrep s_col[NMS_TILE];
  __shared__ BoxPrep s_row[NMS_TILE];
  // axis-aligned flavour: per-box edges (x1, x2, y1, y2) are hoisted out of the pair loop; they are the very
  // expressions of iou_aligned (a[0] -+ a[3] / 2 ...), so every pair sees bit-identical operands
  __shared__ float4 s_coledge[NMS_TILE];
  __shared__ float2 s_colsize[NMS_TILE];
  __shared__ float4 s_rowedge[NMS_TILE];
  __shared__ float2 s_rowsize[NMS_TILE];
  __shared__ unsigned long long s_bits[NMS_TILE];
  for (int tile = blockIdx.x; tile < total_tiles; tile += gridDim.x) {
    // locate the segment (binary search) and the (row, col) tile inside its upper triangle
    int lo = 0, hi = nseg - 1;
    while (lo < hi) {
      const int mid = (lo + hi + 1) >> 1;
      if (tile_off[mid] <= tile) lo = mid; else hi = mid - 1;
    }
    const int q = lo;
    const int beg = seg_off[q], n = seg_off[q + 1] - beg;
    const int T = (n + NMS_TILE - 1) / NMS_TILE;
    int local = tile - tile_off[q];
    int tr = 0;
    while (local >= T - tr) {
      local -= T - tr;
      ++tr;
    }
    const int tc = tr + local;
    const int row_n = min(n - tr * NMS_TILE, NMS_TILE), col_n = min(n - tc * NMS_TILE, NMS_TILE);

    __syncthreads();
    const int t = threadIdx.x;
    if (t < NMS_TILE) {
      s_bits[t] = 0ull;
      if (t < col_n) {
        const float* b = boxes + (long long)(beg + tc * NMS_TILE + t) * 7;
        if (ROTATED) {
          prep_box(b, s_col[t]);
        } else {
          s_coledge[t] = make_float4(b[0] - b[3] / 2, b[0] + b[3] / 2, b[1] - b[4] / 2, b[1] + b[4] / 2);
          s_colsize[t] = make_float2(b[3], b[4]);
        }
      }
    } else if (t < 2 * NMS_TILE) {
      const int r = t - NMS_TILE;
      if (r < row_n) {
        const float* b = boxes + (long long)(beg + tr * NMS_TILE + r) * 7;
        if (ROTATED) {
          prep_box(b, s_row[r]);
        } else {
          s_rowedge[r] = make_float4(b[0] - b[3] / 2, b[0] + b[3] / 2, b[1] - b[4] / 2, b[1] + b[4] / 2);
          s_rowsize[r] = make_float2(b[3], b[4]);
        }
      }
    }
    __syncthreads();
    const int r = t & 63, quarter = t >> 6;
    if (r < row_n) {
      unsigned long long bits = 0ull;
      int c0 = quarter * 16, c1 = min(c0 + 16, col_n);
      if (tr == tc) c0 = max(c0, r + 1);
      float4 ae = make_float4(0.f, 0.f, 0.f, 0.f);
      float2 as = make_float2(0.f, 0.f);
      if (!ROTATED) {
        ae = s_rowedge[r];
        as = s_rowsize[r];
      }
      for (int c = c0; c < c1; ++c) {
        float iou;
        if (ROTATED) {
          const BoxPrep& A = s_row[r];
          const BoxPrep& B = s_col[c];
          const float ddx = A.cx - B.cx, ddy = A.cy - B.cy;
          const float reach = A.rad + B.rad + 0.05f;
          // far apart => the reference finds no vertex, overlap 0, IoU 0: bit stays clear (needs thresh >= 0)
          if (thresh >= 0.f && ddx * ddx + ddy * ddy > reach * reach) continue;
          iou = iou_rotated(A, B);
        } else {
          const float4 be = s_coledge[c];
          const float2 bs = s_colsize[c];
          const float left = fmaxf(ae.x, be.x), right = fminf(ae.y, be.y);
          const float top = fmaxf(ae.z, be.z), bottom = fminf(ae.w, be.w);
          const float width = fmaxf(right - left, 0.f), height = fmaxf(bottom - top, 0.f);
          // disjoint footprints: IoU is 0 (or NaN for degenerate boxes), never above a threshold >= 0
          if (thresh >= 0.f && !(width > 0.f && height > 0.f)) continue;
          const float inter = width * height;
          const float Sa = as.x * as.y;
          const float Sb = bs.x * bs.y;
          iou = inter / fmaxf(Sa + Sb - inter, NMS_EPS);
        }
        if (iou > thresh) bits |= 1ull << c;
      }
      if (bits) atomicOr(&s_bits[r], bits);
    }
    __syncthreads();
    if (t < row_n) mask[mask_off[q] + (long long)(tr * NMS_TILE + t) * T + tc] = s_bits[t];
  }
}

// Axis-aligned flavour, wide work items: one CTA = 64 rows x up to 256 columns (4 column tiles) of one segment's
// upper triangle; thread (row r, column tile j) evaluates 64 pairs and writes its whole 64-bit word, so there are no
// shared atomics and one barrier per 16384 pairs.  grid = (items of the longest segment, segments).
// Same per-pair arithmetic as above (hoisted edges, inline Sa/Sb products, IEEE division).
#define NMS_WIDE 4
__device__ __forceinline__ int nms_items_in_row(int T, int tr) { return (T - tr + NMS_WIDE - 1) / NMS_WIDE; }

__global__ void __launch_bounds__(256) k_nms_mask_aligned(const float* __restrict__ boxes, const int* __restrict__ seg_off,
                                                          const long long* __restrict__ mask_off, float thresh,
                                                          unsigned long long* __restrict__ mask) {
  __shared__ float4 s_coledge[NMS_WIDE * NMS_TILE];
  __shared__ float2 s_colsize[NMS_WIDE * NMS_TILE];
  const int q = blockIdx.y;
  const int beg = seg_off[q], n = seg_off[q + 1] - beg;
  const int T = (n + NMS_TILE - 1) / NMS_TILE;
  int local = blockIdx.x, tr = 0;
  while (tr < T && local >= nms_items_in_row(T, tr)) {
    local -= nms_items_in_row(T, tr);
    ++tr;
  }
  if (tr >= T) return;  // this segment has fewer work items than the longest one
  const int tc0 = tr + local * NMS_WIDE;
  const int t = threadIdx.x;
  {
    const int c = tc0 * NMS_TILE + t;  // one column box per thread
    if (c < n) {
      const float* b = boxes + (long long)(beg + c) * 7;
      s_coledge[t] = make_float4(b[0] - b[3] / 2, b[0] + b[3] / 2, b[1] - b[4] / 2, b[1] + b[4] / 2);
      s_colsize[t] = make_float2(b[3], b[4]);
    }
  }
  __syncthreads();
  const int r = t & 63, j = t >> 6;
  const int tc = tc0 + j;
  const int row = tr * NMS_TILE + r;
  if (row >= n || tc >= T) return;
  const float* a = boxes + (long long)(beg + row) * 7;
  const float4 ae = make_float4(a[0] - a[3] / 2, a[0] + a[3] / 2, a[1] - a[4] / 2, a[1] + a[4] / 2);
  const float2 as = make_float2(a[3], a[4]);
  const int col_n = min(n - tc * NMS_TILE, NMS_TILE);
  unsigned long long bits = 0ull;
  for (int c = (tr == tc) ? r + 1 : 0; c < col_n; ++c) {
    const float4 be = s_coledge[j * NMS_TILE + c];
    const float2 bs = s_colsize[j * NMS_TILE + c];
    const float left = fmaxf(ae.x, be.x), right = fminf(ae.y, be.y);
    const float top = fmaxf(ae.z, be.z), bottom = fminf(ae.w, be.w);
    const float width = fmaxf(right - left, 0.f), height = fmaxf(bottom - top, 0.f);
    // disjoint footprints: IoU is 0 (or NaN for degenerate boxes), never above a threshold >= 0
    if (thresh >= 0.f && !(width > 0.f && height > 0.f)) continue;
    const float inter = width * height;
    const float Sa = as.x * as.y;
    const float Sb = bs.x * bs.y;
    const float iou = inter / fmaxf(Sa + Sb - inter, NMS_EPS);
    if (iou > thresh) bits |= 1ull << c;
  }
  mask[mask_off[q] + (long long)row * T + tc] = bits;
}

// ---------------------------------------------------------------------------
// Stage 3: greedy reduce on the device, one CTA per segment (same visiting order
// as the reference's host loop).  keep[i] in sorted order; counts[q] = kept boxes.
// ---------------------------------------------------------------------------
#define RED_MAX_WORDS 1024

__global__ void __launch_bounds__(128) k_nms_reduce(const unsigned long long* __restrict__ mask,
                                                    const int* __restrict__ seg_off,
                                                    const long long* __restrict__ mask_off, uint8_t* __restrict__ keep,
                                                    int* __restrict__ counts) {
  __shared__ unsigned long long s_remv[RED_MAX_WORDS];
  __shared__ unsigned long long s_diag[NMS_TILE];
  __shared__ unsigned long long s_keepw;
  const int q = blockIdx.x;
  const int beg = seg_off[q], n = seg_off[q + 1] - beg;
  const int T = (n + NMS_TILE - 1) / NMS_TILE;
  const unsigned long long* m = mask + mask_off[q];
  for (int j = threadIdx.x; j < T; j += blockDim.x) s_remv[j] = 0ull;
  int kept = 0;
  __syncthreads();
  for (int blk = 0; blk < T; ++blk) {
    const int rows = min(n - blk * NMS_TILE, NMS_TILE);
    if (threadIdx.x < NMS_TILE)
      s_diag[threadIdx.x] = (threadIdx.x < rows) ? m[(long long)(blk * NMS_TILE + threadIdx.x) * T + blk] : 0ull;
    __syncthreads();
    if (threadIdx.x == 0) {
      unsigned long long cur = s_remv[blk], kw = 0ull;
      for (int i = 0; i < rows; ++i) {
        if (!((cur >> i) & 1ull)) {
          kw |= 1ull << i;
          cur |= s_diag[i];
        }
      }
      s_keepw = kw;
    }
    __syncthreads();
    const unsigned long long kw = s_keepw;
    kept += __popcll(kw);
    for (int j = blk + 1 + threadIdx.x; j < T; j += blockDim.x) {
      unsigned long long acc = 0ull, w = kw;
      while (w) {
        const int i = __ffsll((long long)w) - 1;
        w &= w - 1;
        acc |= m[(long long)(blk * NMS_TILE + i) * T + j];
      }
      s_remv[j] |= acc;
    }
    if (threadIdx.x < rows) keep[beg + blk * NMS_TILE + threadIdx.x] = (uint8_t)((kw >> threadIdx.x) & 1ull);
    __syncthreads();
  }
  if (threadIdx.x == 0 && counts) counts[q] = kept;
}

// ---------------------------------------------------------------------------
// Blocked greedy NMS: one CTA per segment walks its boxes in blocks of 64 (score order) and evaluates
//   (1) the 64 x 64 diagonal tile of the block, (2) a serial resolve of the block against the boxes kept so far,
//   (3) the rows of the boxes that were just KEPT against all later columns.
// The reference's host loop only ever reads mask rows of kept boxes (iou3d_nms.cpp:121-131), so rows of suppressed
// boxes are never needed: the keep-set is identical while the pair count drops from n^2/2 to ~n_kept * n / 2, and no
// mask is materialised in global memory at all.  Per-pair arithmetic = the functions above.
// ---------------------------------------------------------------------------
__global__ void k_nms_prep(const float* __restrict__ boxes, int n, BoxPrep* __restrict__ prep) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n) prep_box(boxes + (long long)i * 7, prep[i]);
}

// NT threads per segment: 256 for the cheap axis-aligned pairs, 512 for the rotated ones (their update phase is the
// long pole: a few kept rows x thousands of columns of polygon clipping per round)
// STAGED (axis-aligned only): the edges and sizes of ALL boxes of the segment are put in dynamic shared memory once
// (24 bytes per box), so neither the row staging nor the per-column walk of a block waits for global loads.
template <bool ROTATED, int NT, bool STAGED>
__global__ void __launch_bounds__(NT) k_nms_greedy(const float* __restrict__ boxes, const BoxPrep* __restrict__ prep,
                                                    const int* __restrict__ seg_off, float thresh,
                                                    uint8_t* __restrict__ keep, int* __restrict__ counts, int cap) {
  extern __shared__ __align__(16) unsigned char s_dyn[];
  __shared__ unsigned long long s_remv[RED_MAX_WORDS];
  __shared__ unsigned long long s_diag[NMS_TILE];
  __shared__ BoxPrep s_row[ROTATED ? NMS_TILE : 1];
  __shared__ float4 s_rowedge_blk[STAGED ? 1 : NMS_TILE];
  __shared__ float2 s_rowsize_blk[STAGED ? 1 : NMS_TILE];
  __shared__ int s_klist[NMS_TILE];
  __shared__ int s_alive[NMS_TILE];
  __shared__ int s_nk;
  float4* s_alledge = reinterpret_cast<float4*>(s_dyn);                 // [cap]
  float2* s_allsize = reinterpret_cast<float2*>(s_dyn + (size_t)cap * sizeof(float4));
  const int q = blockIdx.x;
  const int beg = seg_off[q], n = seg_off[q + 1] - beg;
  const int T = (n + NMS_TILE - 1) / NMS_TILE;
  const int t = threadIdx.x;
  for (int j = t; j < T; j += blockDim.x) s_remv[j] = 0ull;
  if (STAGED) {
    for (int j = t; j < n; j += NT) {
      const float* b = boxes + (long long)(beg + j) * 7;
      const float b0 = b[0], b1 = b[1], b3 = b[3], b4 = b[4];
      s_alledge[j] = make_float4(b0 - b3 / 2, b0 + b3 / 2, b1 - b4 / 2, b1 + b4 / 2);
      s_allsize[j] = make_float2(b3, b4);
    }
  }
  int kept = 0;
  __syncthreads();
  for (int blk = 0; blk < T; ++blk) {
    const int rows = min(n - blk * NMS_TILE, NMS_TILE);
    const float4* s_rowedge = STAGED ? s_alledge + blk * NMS_TILE : s_rowedge_blk;
    const float2* s_rowsize = STAGED ? s_allsize + blk * NMS_TILE : s_rowsize_blk;
    // boxes of this block that no earlier block suppressed, in visiting order (every thread knows the mask)
    const unsigned long long alive =
        ~s_remv[blk] & (rows == NMS_TILE ? ~0ull : ((1ull << rows) - 1ull));
    const int na = __popcll(alive);
    if (t < NMS_TILE) {
      s_diag[t] = 0ull;
      if ((alive >> t) & 1ull) s_alive[__popcll(alive & ((1ull << t) - 1ull))] = t;
      if (t < rows) {
        const int gi = beg + blk * NMS_TILE + t;
        if (ROTATED) {
          s_row[t] = prep[gi];
        } else if (!STAGED) {
          const float* b = boxes + (long long)gi * 7;
          s_rowedge_blk[t] = make_float4(b[0] - b[3] / 2, b[0] + b[3] / 2, b[1] - b[4] / 2, b[1] + b[4] / 2);
          s_rowsize_blk[t] = make_float2(b[3], b[4]);
        }
      }
    }
    __syncthreads();
    // (1) diagonal tile.  Only pairs of boxes that are still alive at the start of the block can matter: the resolve
    // below never reads the row of a box that an earlier block suppressed, and a bit for an already suppressed column
    // changes nothing.  With a few dozen survivors per thousands of boxes that is a handful of pairs per block instead
    // of 2016 - and on the rotated path one pair is a polygon clip with an angular sort (tens of thousands of
    // instructions), which made this phase ~90 % of the kernel.  The alive pairs are spread evenly over the threads.
#pragma unroll 1
    for (int p = t; p < na * na; p += NT) {
      const int ia = p / na, ib = p - ia * na;
      if (ia >= ib) continue;
      const int r = s_alive[ia], c = s_alive[ib];
      float iou;
      if (ROTATED) {
        const BoxPrep& A = s_row[r];
        const BoxPrep& B = s_row[c];
        const float ddx = A.cx - B.cx, ddy = A.cy - B.cy;
        const float reach = A.rad + B.rad + 0.05f;
        if (thresh >= 0.f && ddx * ddx + ddy * ddy > reach * reach) continue;
        iou = iou_rotated(A, B);
      } else {
        const float4 ae = s_rowedge[r], be = s_rowedge[c];
        const float2 as = s_rowsize[r], bs = s_rowsize[c];
        const float left = fmaxf(ae.x, be.x), right = fminf(ae.y, be.y);
        const float top = fmaxf(ae.z, be.z), bottom = fminf(ae.w, be.w);
        const float width = fmaxf(right - left, 0.f), height = fmaxf(bottom - top, 0.f);
        if (thresh >= 0.f && !(width > 0.f && height > 0.f)) continue;
        const float inter = width * height;
        const float Sa = as.x * as.y;
        const float Sb = bs.x * bs.y;
        iou = inter / fmaxf(Sa + Sb - inter, NMS_EPS);
      }
      if (iou > thresh) atomicOr(&s_diag[r], 1ull << c);
    }
    __syncthreads();
    // (2) serial resolve of this block (same visiting order as the reference's host loop)
    if (t == 0) {
      unsigned long long cur = s_remv[blk], kw = 0ull;
      int nk = 0;
      for (int ia = 0; ia < na; ++ia) {   // rows dead at the start of the block cannot be kept: walk the alive ones
        const int i = s_alive[ia];
        if (!((cur >> i) & 1ull)) {
          kw |= 1ull << i;
          cur |= s_diag[i];
          s_klist[nk++] = i;
        }
      }
      s_remv[blk] = ~kw;  // reused below as "not kept" bits of this block
      s_nk = nk;
    }
    __syncthreads();
    const int nk = s_nk;
    kept += nk;
    if (t < rows) keep[beg + blk * NMS_TILE + t] = (uint8_t)(((~s_remv[blk]) >> t) & 1ull);
    // (3) rows of the boxes kept in this block against every later column.  A thread owns columns (col0 + t, + NT, ...):
    // the column's box is loaded once per block, the kept rows come from shared memory (broadcast), and the walk over
    // the kept rows stops at the first one that suppresses the column - further bits cannot change the outcome.
    const int col0 = (blk + 1) * NMS_TILE;
    if (nk > 0) {
      for (int j = col0 + t; j < n; j += NT) {
        if ((s_remv[j >> 6] >> (j & 63)) & 1ull) continue;  // suppressed by an earlier block
        bool sup = false;
        if (ROTATED) {
          BoxPrep B = prep[beg + j];
          // Any kept row that overlaps enough suppresses the column, so the visiting order is free: the kept row whose
          // centre is nearest goes first - it is the likeliest suppressor, and a column that this block suppresses then
          // costs one polygon clip instead of one per kept row up to the first hit.
          int first = 0;
          float dmin = 3.0e38f;
          for (int ai = 0; ai < nk; ++ai) {
            const BoxPrep& A = s_row[s_klist[ai]];
            const float ex = A.cx - B.cx, ey = A.cy - B.cy;
            const float d = ex * ex + ey * ey;
            if (d < dmin) {
              dmin = d;
              first = ai;
            }
          }
#pragma unroll 1
          for (int it = 0; it < nk; ++it) {
            const int ai = (it == 0) ? first : (it <= first ? it - 1 : it);
            // B stays in registers across the walk, but the compiler must not hoist B-only sub-expressions out of the
            // loop: that would change how the pair arithmetic is contracted into FMAs (tools/check_nms_codegen.py)
            asm volatile("" : "+f"(B.cx), "+f"(B.cy), "+f"(B.dx), "+f"(B.dy), "+f"(B.cosm), "+f"(B.sinm), "+f"(B.c[0].x),
                         "+f"(B.c[0].y), "+f"(B.c[1].x), "+f"(B.c[1].y), "+f"(B.c[2].x), "+f"(B.c[2].y), "+f"(B.c[3].x),
                         "+f"(B.c[3].y), "+f"(B.rad));
            const BoxPrep& A = s_row[s_klist[ai]];
            const float ddx = A.cx - B.cx, ddy = A.cy - B.cy;
            const float reach = A.rad + B.rad + 0.05f;
            if (thresh >= 0.f && ddx * ddx + ddy * ddy > reach * reach) continue;
            if (iou_rotated(A, B) > thresh) {
              sup = true;
              break;
            }
          }
        } else {
          float4 be;
          float Sb;
          if (STAGED) {
            be = s_alledge[j];
            const float2 bs = s_allsize[j];
            Sb = bs.x * bs.y;
          } else {
            const float* b = boxes + (long long)(beg + j) * 7;
            const float b3 = b[3], b4 = b[4];
            be = make_float4(b[0] - b3 / 2, b[0] + b3 / 2, b[1] - b4 / 2, b[1] + b4 / 2);
            Sb = b3 * b4;
          }
          for (int ai = 0; ai < nk && !sup; ++ai) {
            const int a = s_klist[ai];
            const float4 ae = s_rowedge[a];
            const float left = fmaxf(ae.x, be.x), right = fminf(ae.y, be.y);
            const float top = fmaxf(ae.z, be.z), bottom = fminf(ae.w, be.w);
            const float width = fmaxf(right - left, 0.f), height = fmaxf(bottom - top, 0.f);
            if (thresh >= 0.f && !(width > 0.f && height > 0.f)) continue;
            const float2 as = s_rowsize[a];
            const float inter = width * height;
            const float Sa = as.x * as.y;
            sup = inter / fmaxf(Sa + Sb - inter, NMS_EPS) > thresh;
          }
        }
        if (sup) atomicOr(&s_remv[j >> 6], 1ull << (j & 63));
      }
    }
    __syncthreads();
  }
  if (t == 0 && counts) counts[q] = kept;
}

// pairwise BEV IoU / overlap matrices (reference: boxes_iou_bev_gpu / boxes_overlap_bev_gpu,
// iou3d_nms_kernel.cu:236-265)
__global__ void k_pair_matrix(const float* __restrict__ a, int na, const float* __restrict__ b, int nb, int mode,
                              float* __restrict__ out) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  const int i = blockIdx.y * blockDim.y + threadIdx.y;
  if (i >= na || j >= nb) return;
  float v;
  if (mode == 2) {
    v = iou_aligned(a + (long long)i * 7, b + (long long)j * 7);
  } else {
    BoxPrep A, B;
    prep_box(a + (long long)i * 7, A);
    prep_box(b + (long long)j * 7, B);
    v = (mode == 0) ? bev_overlap(A, B) : iou_rotated(A, B);
  }
  out[(long long)i * nb + j] = v;
}

// kept indices in order (compat with the reference's `keep` list)
__global__ void k_keep_list(const uint8_t* __restrict__ keep, const int* __restrict__ pos, int n,
                            long long* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n && keep[i]) out[pos[i]] = i;
}

extern "C" {

// Host-side helper: fills tile_off[nseg+1] (int32) and mask_off[nseg+1] (int64, in 8-byte words) from
// host segment offsets; returns total mask words through *mask_words.
int b2_nms_plan(const int* seg_off_host, int nseg, int* tile_off_host, long long* mask_off_host, long long* mask_words,
                int* max_seg) {
  long long words = 0;
  int tiles = 0, mx = 0;
  for (int q = 0; q < nseg; ++q) {
    const int n = seg_off_host[q + 1] - seg_off_host[q];
    B2_CHECK_ARG(n >= 0, "segment offsets must be non-decreasing");
    const int T = (n + NMS_TILE - 1) / NMS_TILE;
    tile_off_host[q] = tiles;
    mask_off_host[q] = words;
    tiles += T * (T + 1) / 2;
    words += (long long)n * T;
    if (n > mx) mx = n;
  }
  tile_off_host[nseg] = tiles;
  mask_off_host[nseg] = words;
  *mask_words = words;
  *max_seg = mx;
  return B2_OK;
}

// boxes [n,7] / scores [n] grouped in nseg segments (device seg_off[nseg+1]).
// sort != 0: each segment is first ordered by descending score -> sorted_boxes, order (segment-local source index).
// sort == 0: boxes are taken as already ordered (sorted_boxes may alias boxes, order unused).
// keep[n] (uint8, in sorted order), counts[nseg].  mask: scratch of mask_words 8-byte words.
int b2_nms_batched(const float* boxes, const float* scores, const int* seg_off, const int* tile_off,
                   const long long* mask_off, int nseg, int total_tiles, int max_seg, float thresh, int rotated, int sort,
                   float* sorted_boxes, int* order, unsigned long long* mask, uint8_t* keep, int* counts, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (nseg == 0) return B2_OK;
  B2_CHECK_ARG(max_seg <= RED_MAX_WORDS * NMS_TILE, "segment too long for the device reduce (%d)", max_seg);
  const float* sb = boxes;
  if (sort) {
    B2_CHECK_ARG(max_seg <= SORT_MAX, "segment longer than %d cannot be sorted in-kernel (%d)", SORT_MAX, max_seg);
    k_nms_sort<<<nseg, 1024, 0, st>>>(boxes, scores, seg_off, sorted_boxes, order);
    B2_LAUNCH_CHECK();
    sb = sorted_boxes;
  }
  if (total_tiles > 0) {
    const int grid = total_tiles < 148 * 8 ? total_tiles : 148 * 8;
    if (rotated) {
      k_nms_mask<true><<<grid, 256, 0, st>>>(sb, seg_off, tile_off, mask_off, nseg, total_tiles, thresh, mask);
    } else if (nseg > 65535) {
      k_nms_mask<false><<<grid, 256, 0, st>>>(sb, seg_off, tile_off, mask_off, nseg, total_tiles, thresh, mask);
    } else {
      const int Tmax = (max_seg + NMS_TILE - 1) / NMS_TILE;
      int items = 0;
      for (int tr = 0; tr < Tmax; ++tr) items += (Tmax - tr + NMS_WIDE - 1) / NMS_WIDE;
      k_nms_mask_aligned<<<dim3(items, nseg), 256, 0, st>>>(sb, seg_off, mask_off, thresh, mask);
    }
    B2_LAUNCH_CHECK();
  }
  k_nms_reduce<<<nseg, 128, 0, st>>>(mask, seg_off, mask_off, keep, counts);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

size_t b2_nms_greedy_ws_bytes(long long n, int rotated) { return rotated ? (size_t)n * sizeof(BoxPrep) + 256 : 256; }

// Same contract as b2_nms_batched (keep[n] in per-segment sorted order, counts[nseg]) without a suppression mask:
// ws (b2_nms_greedy_ws_bytes) holds the per-box corner / trig data of the rotated flavour.
int b2_nms_greedy_batched(const float* boxes, const float* scores, const int* seg_off, int nseg, long long n,
                          int max_seg, float thresh, int rotated, int sort, float* sorted_boxes, int* order, void* ws,
                          size_t ws_bytes, uint8_t* keep, int* counts, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (nseg == 0) return B2_OK;
  B2_CHECK_ARG(max_seg <= RED_MAX_WORDS * NMS_TILE, "segment too long for the device reduce (%d)", max_seg);
  if (ws_bytes < b2_nms_greedy_ws_bytes(n, rotated)) {
    b2_set_error("b2_nms_greedy_batched workspace too small");
    return B2_ERR_WS;
  }
  const float* sb = boxes;
  if (sort) {
    B2_CHECK_ARG(max_seg <= SORT_MAX, "segment longer than %d cannot be sorted in-kernel (%d)", SORT_MAX, max_seg);
    k_nms_sort<<<nseg, 1024, 0, st>>>(boxes, scores, seg_off, sorted_boxes, order);
    B2_LAUNCH_CHECK();
    sb = sorted_boxes;
  }
  BoxPrep* prep = reinterpret_cast<BoxPrep*>(ws);
  if (rotated) {
    if (n > 0) {
      k_nms_prep<<<b2_div_up(n, 128), 128, 0, st>>>(sb, (int)n, prep);
      B2_LAUNCH_CHECK();
    }
    // 57 registers per thread: 1024 threads per segment fit, and the per-column walk of the update phase uses them
    // (B2_NMS_NT_ROT=512 for A/B runs)
    static const int ntr_env = getenv("B2_NMS_NT_ROT") ? atoi(getenv("B2_NMS_NT_ROT")) : 0;
    const int ntr = ntr_env ? ntr_env : (max_seg > 1024 ? 1024 : 512);
    if (ntr >= 1024) k_nms_greedy<true, 1024, false><<<nseg, 1024, 0, st>>>(sb, prep, seg_off, thresh, keep, counts, 0);
    else k_nms_greedy<true, 512, false><<<nseg, 512, 0, st>>>(sb, prep, seg_off, thresh, keep, counts, 0);
  } else {
    // threads per segment: the update phase is one independent walk per column, so a segment of thousands of boxes
    // feeds 1024 threads (B2_NMS_NT=256 / 512 / 1024 for A/B runs)
    static const int nt_env = getenv("B2_NMS_NT") ? atoi(getenv("B2_NMS_NT")) : 0;
    const int nt = nt_env ? nt_env : (max_seg > 1024 ? 1024 : (max_seg > 256 ? 512 : 256));
    // all boxes of a segment staged in shared memory (24 B per box) when the longest segment fits (B2_NMS_STAGED=0: off)
    static const bool staged_ok = !(getenv("B2_NMS_STAGED") && atoi(getenv("B2_NMS_STAGED")) == 0);
    const size_t dyn = (size_t)max_seg * (sizeof(float4) + sizeof(float2));
    if (nt >= 1024 && staged_ok && max_seg > 0 && dyn <= 160 * 1024) {
      auto kern = k_nms_greedy<false, 1024, true>;
      static unsigned long long configured = 0;   // one bit per device: the attribute is per device
      {
        int dev = 0;
        B2_CUDA(cudaGetDevice(&dev));
        if (dev >= 64 || !((configured >> dev) & 1ull)) {
          B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, 160 * 1024));
          if (dev < 64) configured |= 1ull << dev;
        }
      }
      kern<<<nseg, 1024, dyn, st>>>(sb, prep, seg_off, thresh, keep, counts, max_seg);
    } else if (nt >= 1024) {
      k_nms_greedy<false, 1024, false><<<nseg, 1024, 0, st>>>(sb, prep, seg_off, thresh, keep, counts, 0);
    } else if (nt >= 512) {
      k_nms_greedy<false, 512, false><<<nseg, 512, 0, st>>>(sb, prep, seg_off, thresh, keep, counts, 0);
    } else {
      k_nms_greedy<false, 256, false><<<nseg, 256, 0, st>>>(sb, prep, seg_off, thresh, keep, counts, 0);
    }
  }
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// mode 0: rotated BEV overlap area, 1: rotated BEV IoU, 2: axis-aligned BEV IoU
int b2_boxes_pair_matrix(const float* a, long long na, const float* b, long long nb, int mode, float* out, void* stream) {
  if (na == 0 || nb == 0) return B2_OK;
  dim3 block(16, 16), grid(b2_div_up(nb, 16), b2_div_up(na, 16));
  k_pair_matrix<<<grid, block, 0, (cudaStream_t)stream>>>(a, (int)na, b, (int)nb, mode, out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_keep_list(const uint8_t* keep, const int* pos, long long n, long long* out, void* stream) {
  if (n == 0) return B2_OK;
  k_keep_list<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(keep, pos, (int)n, out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"

// ===========================================================================
// Allocation-free single-segment NMS.
//
// For callers that are not allowed to allocate device memory when they call us: a perception loop that
// reserved its whole GPU budget at start-up and bans cudaMalloc afterwards (cudaMalloc can stall the device
// and can fail late, once memory is fragmented or taken by other tenants), or a host runtime that owns the
// GPU heap and gives this library no allocator.  b2_nms_batched needs a caller-provided mask buffer of
// n * ceil(n/64) words; this variant keeps that scratch (and the three tiny plan arrays the kernels read)
// in a fixed array inside the library image instead, so the caller passes only boxes in and keep/count out.
//
// The price: B2_NOALLOC_MAX_BOXES boxes at most, and ~0.5 MB of device memory held by every GPU that loads
// the library, until the process exits.
// ===========================================================================
#define B2_NOALLOC_MAX_BOXES 2048
#define B2_NOALLOC_HDR_WORDS 32  // 256 bytes: seg_off[2] | tile_off[2] | mask_off[2]
#define B2_NOALLOC_WORDS (B2_NOALLOC_HDR_WORDS + B2_NOALLOC_MAX_BOXES * (B2_NOALLOC_MAX_BOXES / NMS_TILE))

static __device__ __align__(256) unsigned long long g_noalloc_scratch[B2_NOALLOC_WORDS];

// plan of a one-segment problem, written on the device so that no host->device copy is needed
__global__ void k_noalloc_plan(unsigned long long* hdr, int n) {
  int* seg_off = reinterpret_cast<int*>(hdr);           // words 0
  int* tile_off = reinterpret_cast<int*>(hdr + 1);      // word 1
  long long* mask_off = reinterpret_cast<long long*>(hdr + 2);  // words 2..3
  const int T = (n + NMS_TILE - 1) / NMS_TILE;
  seg_off[0] = 0;
  seg_off[1] = n;
  tile_off[0] = 0;
  tile_off[1] = T * (T + 1) / 2;
  mask_off[0] = 0;
  mask_off[1] = (long long)n * T;
}

#include <mutex>
#define B2_MAX_DEVICES 64
static std::mutex g_noalloc_mu;
static unsigned long long* g_noalloc_ptr[B2_MAX_DEVICES] = {nullptr};

// Device address of the scratch array on the CURRENT device.  In host code `g_noalloc_scratch` is only a
// host-side handle of the symbol, never a device address, so it is resolved once per device through
// cudaGetSymbolAddress (with that device current) and the resolved pointer is what the kernels receive.
static int noalloc_scratch(unsigned long long** out) {
  int dev = -1;
  B2_CUDA(cudaGetDevice(&dev));
  B2_CHECK_ARG(dev >= 0 && dev < B2_MAX_DEVICES, "device ordinal %d out of range", dev);
  std::lock_guard<std::mutex> lk(g_noalloc_mu);
  if (g_noalloc_ptr[dev] == nullptr) {
    void* p = nullptr;
    B2_CUDA(cudaGetSymbolAddress(&p, g_noalloc_scratch));
    g_noalloc_ptr[dev] = static_cast<unsigned long long*>(p);
  }
  *out = g_noalloc_ptr[dev];
  return B2_OK;
}

extern "C" int b2_nms_single_noalloc_max_boxes(void) { return B2_NOALLOC_MAX_BOXES; }

// Greedy BEV NMS of ONE list of boxes that is already sorted by descending score (the contract of the
// reference's nms_gpu / nms_normal_gpu), with no allocation of any kind inside the call.
//   boxes_sorted [n,7] device, keep [n] device (1 = kept), count [1] device (number kept); n <= 2048.
// Notes for the caller:
//  * the first call on each device resolves the scratch address (cudaGetSymbolAddress); make that first
//    call during start-up, before any stream capture is begun on that device;
//  * all calls on one device share the one scratch array: two calls whose kernels may overlap on the GPU
//    (different streams, no dependency) race on it — serialise them;
//  * the array keeps the previous call's contents; every mask word that a call reads is written earlier in
//    that same call (upper-triangular tiles) and the plan words are rewritten each time, so nothing has to
//    be cleared and stale data is never observed.
extern "C" int b2_nms_single_noalloc(const float* boxes_sorted, long long n, float thresh, int rotated, uint8_t* keep,
                                     int* count, void* stream) {
  B2_CHECK_ARG(n >= 0, "negative box count");
  B2_CHECK_ARG(n <= B2_NOALLOC_MAX_BOXES,
               "b2_nms_single_noalloc handles at most %d boxes (got %lld): use b2_nms_batched with caller scratch",
               B2_NOALLOC_MAX_BOXES, n);
  cudaStream_t st = (cudaStream_t)stream;
  unsigned long long* scratch = nullptr;
  int rc = noalloc_scratch(&scratch);
  if (rc) return rc;
  const int* seg_off = reinterpret_cast<const int*>(scratch);
  const int* tile_off = reinterpret_cast<const int*>(scratch + 1);
  const long long* mask_off = reinterpret_cast<const long long*>(scratch + 2);
  unsigned long long* mask = scratch + B2_NOALLOC_HDR_WORDS;
  k_noalloc_plan<<<1, 1, 0, st>>>(scratch, (int)n);
  B2_LAUNCH_CHECK();
  const int T = (int)((n + NMS_TILE - 1) / NMS_TILE);
  const int total_tiles = T * (T + 1) / 2;
  if (total_tiles > 0) {
    const int grid = total_tiles < 148 * 8 ? total_tiles : 148 * 8;
    if (rotated)
      k_nms_mask<true><<<grid, 256, 0, st>>>(boxes_sorted, seg_off, tile_off, mask_off, 1, total_tiles, thresh, mask);
    else
      k_nms_mask<false><<<grid, 256, 0, st>>>(boxes_sorted, seg_off, tile_off, mask_off, 1, total_tiles, thresh, mask);
    B2_LAUNCH_CHECK();
  }
  k_nms_reduce<<<1, 128, 0, st>>>(mask, seg_off, mask_off, keep, count);
  B2_LAUNCH_CHECK();
  return B2_OK;
}
