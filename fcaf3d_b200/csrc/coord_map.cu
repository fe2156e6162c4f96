// Coordinate manager kernels: packed-key open-addressing hash, dedup in first-occurrence
// order, stride / children / union maps, order-preserving compaction (chained scan).
//
// Replaces (behaviourally) MinkowskiEngine v0.5.4's CoordinateMapManager::insert_and_map /
// stride / stride_region / union_map / prune, which the reference reaches from
// mmdet3d/models/detectors/single_stage_sparse.py:34-37, backbones/me_resnet.py:19-24,66-75 and
// dense_heads/fcaf3d_neck_with_head.py:60-66,101,125 (ME itself is not vendored in the reference).
#include "common.cuh"

// ---------------------------------------------------------------------------
// Single-pass chained scan (decoupled look-back) over uint8 flags -> int32
// exclusive positions.  4096 flags per CTA; dynamic CTA ids (ticket) make the
// look-back deadlock-free regardless of hardware scheduling order.
// ---------------------------------------------------------------------------
#define SCAN_THREADS 256
#define SCAN_ITEMS 16
#define SCAN_TILE (SCAN_THREADS * SCAN_ITEMS)
#define ST_INVALID 0ull
#define ST_AGG 1ull
#define ST_PREFIX 2ull

// WINNER >= 1: the flags are not read from memory but computed in place - row i is flagged when it is the winning (lowest)
// input row of its hash slot (b2_map_build: saves the pass that wrote the flags and the one that read them back).
// WINNER == 2 (map builds whose caller wants no inverse map): the scan also does the finalize step for the winners - it
// writes their compacted coordinate row (floored to `quant`), optionally their input row (unique_index) and the compacted
// row id into their hash slot - so the build is two kernels (insert, scan) and the positions never go to memory.
template <int WINNER>
__global__ void __launch_bounds__(SCAN_THREADS) k_scan_flags(const uint8_t* __restrict__ flags,
                                                             const int* __restrict__ row_slot,
                                                             B2Slot* __restrict__ table, int n,
                                                             int* __restrict__ pos, int* __restrict__ total,
                                                             unsigned long long* status, unsigned int* ticket,
                                                             int nblocks, const int* __restrict__ coords = nullptr,
                                                             int quant = 0, int* __restrict__ out_coords = nullptr,
                                                             int* __restrict__ unique_index = nullptr) {
  __shared__ int s_bid;
  __shared__ int s_warp[SCAN_THREADS / 32];
  __shared__ int s_prefix;
  if (threadIdx.x == 0) s_bid = (int)atomicAdd(ticket, 1u);
  __syncthreads();
  const int bid = s_bid;
  const int lane = threadIdx.x & 31, warp = threadIdx.x >> 5;
  const long long base = (long long)bid * SCAN_TILE + (long long)threadIdx.x * SCAN_ITEMS;

  int f[SCAN_ITEMS];
  int sl[WINNER ? SCAN_ITEMS : 1];
  if (WINNER) {
    if (base + SCAN_ITEMS <= n) {
#pragma unroll
      for (int i = 0; i < SCAN_ITEMS / 4; ++i) {
        const int4 v = __ldg(reinterpret_cast<const int4*>(row_slot + base) + i);
        sl[4 * i] = v.x;
        sl[4 * i + 1] = v.y;
        sl[4 * i + 2] = v.z;
        sl[4 * i + 3] = v.w;
      }
    } else {
#pragma unroll
      for (int i = 0; i < SCAN_ITEMS; ++i) sl[i] = (base + i < n) ? row_slot[base + i] : -1;
    }
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i)
      f[i] = (sl[i] >= 0 && table[sl[i]].pad == (unsigned int)(base + i)) ? 1 : 0;
  } else if (base + SCAN_ITEMS <= n) {
    const uint4 v = *reinterpret_cast<const uint4*>(flags + base);
    const unsigned w[4] = {v.x, v.y, v.z, v.w};
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) f[i] = ((w[i >> 2] >> ((i & 3) * 8)) & 0xFF) ? 1 : 0;
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) f[i] = (base + i < n) ? (flags[base + i] ? 1 : 0) : 0;
  }
  int local = 0;
#pragma unroll
  for (int i = 0; i < SCAN_ITEMS; ++i) local += f[i];

  int incl = local;
#pragma unroll
  for (int d = 1; d < 32; d <<= 1) {
    int t = __shfl_up_sync(0xffffffffu, incl, d);
    if (lane >= d) incl += t;
  }
  if (lane == 31) s_warp[warp] = incl;
  __syncthreads();
  int warp_off = 0, block_sum = 0;
#pragma unroll
  for (int w = 0; w < SCAN_THREADS / 32; ++w) {
    int s = s_warp[w];
    if (w < warp) warp_off += s;
    block_sum += s;
  }

  if (warp == 0) {
    if (lane == 0) {
      unsigned long long st = ((bid == 0 ? ST_PREFIX : ST_AGG) << 62) | (unsigned int)block_sum;
      atomicExch(&status[bid], st);
    }
    int excl = 0;
    if (bid > 0) {
      int look = bid - 1;
      for (;;) {
        const int idx = look - lane;
        unsigned long long s;
        if (idx >= 0) {
          do {
            s = *reinterpret_cast<volatile unsigned long long*>(&status[idx]);
          } while ((s >> 62) == ST_INVALID);
        } else {
          s = (ST_PREFIX << 62);
        }
        const unsigned pmask = __ballot_sync(0xffffffffu, (s >> 62) == ST_PREFIX);
        int val = (int)(unsigned int)(s & 0xFFFFFFFFull);
        const int first = __ffs(pmask) - 1;  // nearest predecessor holding an inclusive prefix
        if (pmask && lane > first) val = 0;
#pragma unroll
        for (int d = 16; d > 0; d >>= 1) val += __shfl_xor_sync(0xffffffffu, val, d);
        excl += val;
        if (pmask) break;
        look -= 32;
      }
      if (lane == 0) {
        unsigned long long st = (ST_PREFIX << 62) | (unsigned int)(excl + block_sum);
        atomicExch(&status[bid], st);
      }
    }
    if (lane == 0) {
      s_prefix = excl;
      if (bid == nblocks - 1) *total = excl + block_sum;
    }
  }
  __syncthreads();
  int run = s_prefix + warp_off + (incl - local);
  if (WINNER == 2) {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
      if (f[i]) {
        const long long row = base + i;
        int4 c = __ldg(reinterpret_cast<const int4*>(coords) + row);
        if (quant > 0) {
          c.y = b2_floor_div(c.y, quant) * quant;
          c.z = b2_floor_div(c.z, quant) * quant;
          c.w = b2_floor_div(c.w, quant) * quant;
        }
        reinterpret_cast<int4*>(out_coords)[run] = c;
        if (unique_index) unique_index[run] = (int)row;
        table[sl[i]].val = (unsigned int)run;
        ++run;
      }
    }
    return;
  }
  if (base + SCAN_ITEMS <= n) {
    int o[SCAN_ITEMS];
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
      o[i] = run;
      run += f[i];
    }
    int4* dst = reinterpret_cast<int4*>(pos + base);
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS / 4; ++i) dst[i] = make_int4(o[4 * i], o[4 * i + 1], o[4 * i + 2], o[4 * i + 3]);
  } else {
#pragma unroll
    for (int i = 0; i < SCAN_ITEMS; ++i) {
      if (base + i < n) pos[base + i] = run;
      run += f[i];
    }
  }
}

size_t b2_scan_ws_bytes_impl(long long n) {
  long long nblocks = (n + SCAN_TILE - 1) / SCAN_TILE;
  return (size_t)(nblocks + 2) * 8;
}

// flags/pos must be 16-byte aligned (torch allocations are).
int b2_scan_flags_launch(const uint8_t* flags, long long n, int* pos, int* total, void* ws, size_t ws_bytes,
                         cudaStream_t st) {
  if (n <= 0) {
    B2_CUDA(cudaMemsetAsync(total, 0, sizeof(int), st));
    return B2_OK;
  }
  const size_t need = b2_scan_ws_bytes_impl(n);
  if (ws_bytes < need) {
    b2_set_error("scan workspace too small: %zu < %zu", ws_bytes, need);
    return B2_ERR_WS;
  }
  const int nblocks = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
  B2_CUDA(cudaMemsetAsync(ws, 0, need, st));
  unsigned int* ticket = reinterpret_cast<unsigned int*>(ws);
  unsigned long long* status = reinterpret_cast<unsigned long long*>(ws) + 1;
  k_scan_flags<0><<<nblocks, SCAN_THREADS, 0, st>>>(flags, nullptr, nullptr, (int)n, pos, total, status, ticket, nblocks);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// exclusive positions of the winning rows of a hash build (see k_scan_flags<true>); row_slot must be 16-byte aligned
static int b2_scan_winners_launch(const int* row_slot, B2Slot* table, long long n, int* pos, int* total, void* ws,
                                  size_t ws_bytes, cudaStream_t st, const int* coords = nullptr, int quant = 0,
                                  int* out_coords = nullptr, int* unique_index = nullptr) {
  const size_t need = b2_scan_ws_bytes_impl(n);
  if (ws_bytes < need) {
    b2_set_error("scan workspace too small: %zu < %zu", ws_bytes, need);
    return B2_ERR_WS;
  }
  const int nblocks = (int)((n + SCAN_TILE - 1) / SCAN_TILE);
  B2_CUDA(cudaMemsetAsync(ws, 0, need, st));
  unsigned int* ticket = reinterpret_cast<unsigned int*>(ws);
  unsigned long long* status = reinterpret_cast<unsigned long long*>(ws) + 1;
  if (out_coords != nullptr)
    k_scan_flags<2><<<nblocks, SCAN_THREADS, 0, st>>>(nullptr, row_slot, table, (int)n, nullptr, total, status, ticket, nblocks,
                                                      coords, quant, out_coords, unique_index);
  else
    k_scan_flags<1><<<nblocks, SCAN_THREADS, 0, st>>>(nullptr, row_slot, table, (int)n, pos, total, status, ticket, nblocks);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// ---------------------------------------------------------------------------
// Hash insert / finalize
// ---------------------------------------------------------------------------
__device__ __forceinline__ int4 load_coord(const int* __restrict__ coords, long long i) {
  return __ldg(reinterpret_cast<const int4*>(coords) + i);
}

// quant > 0: spatial coords are floored to multiples of quant (stride map, ME A.3).
__global__ void k_hash_insert(const int* __restrict__ coords, int n, int quant, B2Slot* table, unsigned int cap_mask,
                              int* __restrict__ row_slot, int* err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  int4 c = load_coord(coords, i);
  if (quant > 0) {
    c.y = b2_floor_div(c.y, quant) * quant;
    c.z = b2_floor_div(c.z, quant) * quant;
    c.w = b2_floor_div(c.w, quant) * quant;
  }
  if (!b2_coord_in_range(c.x, c.y, c.z, c.w)) {
    atomicExch(err, 1);
    row_slot[i] = -1;
    return;
  }
  const unsigned long long key = b2_pack_key(c.x, c.y, c.z, c.w);
  unsigned int h = b2_hash_key(key) & cap_mask;
#pragma unroll 1
  for (;;) {
    unsigned long long old = atomicCAS(&table[h].key, B2_EMPTY_KEY, key);
    if (old == B2_EMPTY_KEY || old == key) {
      // lowest input row wins (ME CPU first-occurrence rule).  The input row lives in the slot's second word; the
      // first one (what lookups return) receives the compacted row id from the winner in k_build_finalize
      atomicMin(&table[h].pad, (unsigned int)i);
      row_slot[i] = (int)h;
      return;
    }
    h = (h + 1) & cap_mask;
  }
}

// winners write the compacted coordinate row + unique_index and turn their slot's value into the compacted row id;
// everybody records the winner row of its voxel (read from the slot's second word, which nobody writes here)
__global__ void k_build_finalize(const int* __restrict__ coords, int n, int quant, B2Slot* __restrict__ table,
                                 const int* __restrict__ row_slot,
                                 const int* __restrict__ pos, int* __restrict__ out_coords,
                                 int* __restrict__ unique_index, int* __restrict__ inverse_map) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int s = row_slot[i];
  if (s < 0) {
    inverse_map[i] = -1;
    return;
  }
  const int w = (int)table[s].pad;  // original row of the representative
  inverse_map[i] = pos[w];
  if (w == i) {  // this row is the representative of its voxel
    int4 c = load_coord(coords, i);
    if (quant > 0) {
      c.y = b2_floor_div(c.y, quant) * quant;
      c.z = b2_floor_div(c.z, quant) * quant;
      c.w = b2_floor_div(c.w, quant) * quant;
    }
    const int p = pos[i];
    reinterpret_cast<int4*>(out_coords)[p] = c;
    unique_index[p] = i;
    table[s].val = (unsigned int)p;
  }
}

// plain insert of already-unique coordinates (value = row)
__global__ void k_hash_insert_unique(const int* __restrict__ coords, int n, B2Slot* table, unsigned int cap_mask,
                                     int* err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  const int4 c = load_coord(coords, i);
  if (!b2_coord_in_range(c.x, c.y, c.z, c.w)) {
    atomicExch(err, 1);
    return;
  }
  const unsigned long long key = b2_pack_key(c.x, c.y, c.z, c.w);
  unsigned int h = b2_hash_key(key) & cap_mask;
#pragma unroll 1
  for (;;) {
    unsigned long long old = atomicCAS(&table[h].key, B2_EMPTY_KEY, key);
    if (old == B2_EMPTY_KEY || old == key) {
      atomicMin(&table[h].val, (unsigned int)i);
      return;
    }
    h = (h + 1) & cap_mask;
  }
}

__global__ void k_lookup(const B2Slot* __restrict__ table, unsigned int cap_mask, const int* __restrict__ q, int m,
                         int* __restrict__ rows) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= m) return;
  const int4 c = load_coord(q, i);
  int r = -1;
  if (b2_coord_in_range(c.x, c.y, c.z, c.w)) r = b2_table_find(table, cap_mask, b2_pack_key(c.x, c.y, c.z, c.w));
  rows[i] = r;
}

// children of every voxel: row 8*i+k, k = ix + 2*iy + 4*iz (x fastest), offsets {0, half}
__global__ void k_children(const int* __restrict__ coords, int n, int half, int* __restrict__ out) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * 8) return;
  const int i = t >> 3, k = t & 7;
  int4 c = load_coord(coords, i);
  c.y += (k & 1) ? half : 0;
  c.z += (k & 2) ? half : 0;
  c.w += (k & 4) ? half : 0;
  reinterpret_cast<int4*>(out)[t] = c;
}

// boundaries of batch-sorted rows: offsets[b] = first row with batch >= b, offsets[B] = n
__global__ void k_batch_offsets(const int* __restrict__ coords, int n, int nb, int* __restrict__ offsets, int* err) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  const int prev = (i == 0) ? -1 : coords[4 * (long long)(i - 1)];
  const int cur = (i == n) ? nb : coords[4 * (long long)i];
  if (i < n && (cur < prev || cur >= nb || cur < 0)) {
    atomicExch(err, 2);  // rows are not grouped by ascending batch index
    return;
  }
  for (int b = prev + 1; b <= cur && b <= nb; ++b) offsets[b] = i;
}

// same, with the row count read from device memory (so the host needs no round trip before launching it)
__global__ void k_batch_offsets_dev(const int* __restrict__ coords, const int* __restrict__ n_ptr, int nb,
                                    int* __restrict__ offsets, int* err) {
  const int n = *n_ptr;
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i > n) return;
  const int prev = (i == 0) ? -1 : coords[4 * (long long)(i - 1)];
  const int cur = (i == n) ? nb : coords[4 * (long long)i];
  if (i < n && (cur < prev || cur >= nb || cur < 0)) {
    atomicExch(err, 2);
    return;
  }
  for (int b = prev + 1; b <= cur && b <= nb; ++b) offsets[b] = i;
}

// union: flags of b-rows missing from a
__global__ void k_union_probe(const B2Slot* __restrict__ a_table, unsigned int cap_mask, const int* __restrict__ b_coords,
                              int nb_rows, int* __restrict__ found, uint8_t* __restrict__ missing) {
  const int j = blockIdx.x * blockDim.x + threadIdx.x;
  if (j >= nb_rows) return;
  const int4 c = load_coord(b_coords, j);
  int r = -1;
  if (b2_coord_in_range(c.x, c.y, c.z, c.w)) r = b2_table_find(a_table, cap_mask, b2_pack_key(c.x, c.y, c.z, c.w));
  found[j] = r;
  missing[j] = (r < 0) ? 1 : 0;
}

// union rows stay grouped by batch: scene s = [a rows of s][b-only rows of s]
__global__ void k_union_assign(const int* __restrict__ a_coords, int na, const int* __restrict__ a_off,
                               const int* __restrict__ b_coords, int nb_rows, const int* __restrict__ b_off,
                               const int* __restrict__ found, const int* __restrict__ pos, const int* __restrict__ total,
                               int* __restrict__ out_coords, int* __restrict__ dst_a, int* __restrict__ dst_b) {
  const int t = blockIdx.x * blockDim.x + threadIdx.x;
  if (t < na) {
    const int4 c = load_coord(a_coords, t);
    const int bo = b_off[c.x];
    const int shift = (bo < nb_rows) ? pos[bo] : *total;  // b-only rows of earlier scenes
    const int d = t + shift;
    dst_a[t] = d;
    reinterpret_cast<int4*>(out_coords)[d] = c;
  } else if (t < na + nb_rows) {
    const int j = t - na;
    const int4 c = load_coord(b_coords, j);
    const int f = found[j];
    if (f >= 0) {
      const int bo = b_off[c.x];
      const int shift = (bo < nb_rows) ? pos[bo] : *total;
      dst_b[j] = f + shift;
    } else {
      const int d = a_off[c.x + 1] + pos[j];
      dst_b[j] = d;
      reinterpret_cast<int4*>(out_coords)[d] = c;
    }
  }
}

__global__ void k_add_total(const int* __restrict__ total, int add, int* __restrict__ out) { *out = *total + add; }

// order-preserving row compaction
__global__ void k_gather_coords(const int* __restrict__ coords, const uint8_t* __restrict__ mask,
                                const int* __restrict__ pos, int n, int* __restrict__ out) {
  const int i = blockIdx.x * blockDim.x + threadIdx.x;
  if (i >= n) return;
  if (mask[i]) reinterpret_cast<int4*>(out)[pos[i]] = load_coord(coords, i);
}

// out[pos[i]] = in[i] for masked rows; 16-byte vectors (row_bytes % 16 == 0) or 4-byte words
template <typename V>
__global__ void k_gather_rows(const V* __restrict__ in, const uint8_t* __restrict__ mask, const int* __restrict__ pos,
                              long long n, int vec_per_row, V* __restrict__ out) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * vec_per_row) return;
  const long long i = t / vec_per_row;
  const int v = (int)(t - i * vec_per_row);
  if (mask[i]) out[(long long)pos[i] * vec_per_row + v] = in[t];
}

// out[i] = in[idx[i]] (idx < 0 -> zeros)
template <typename V>
__global__ void k_index_rows(const V* __restrict__ in, const int* __restrict__ idx, long long n, int vec_per_row,
                             V* __restrict__ out) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * vec_per_row) return;
  const long long i = t / vec_per_row;
  const int v = (int)(t - i * vec_per_row);
  const int r = idx[i];
  V z;
  memset(&z, 0, sizeof(V));
  out[t] = (r >= 0) ? in[(long long)r * vec_per_row + v] : z;
}

// union feature assembly: out[dst_a[i]] = a[i]; out[dst_b[j]] += b[j]  (out pre-zeroed; a first, then b)
__global__ void k_scatter_rows_f32(const float* __restrict__ in, const int* __restrict__ dst, long long n, int c,
                                   float* __restrict__ out, int accumulate) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c) return;
  const long long i = t / c;
  const int ch = (int)(t - i * c);
  if (dst[i] < 0) return;
  const long long o = (long long)dst[i] * c + ch;
  out[o] = accumulate ? out[o] + in[t] : in[t];
}

__global__ void k_scatter_rows_bf16(const __nv_bfloat16* __restrict__ in, const int* __restrict__ dst, long long n,
                                    int c, __nv_bfloat16* __restrict__ out, int accumulate) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * c) return;
  const long long i = t / c;
  const int ch = (int)(t - i * c);
  if (dst[i] < 0) return;
  const long long o = (long long)dst[i] * c + ch;
  out[o] = accumulate ? __float2bfloat16(__bfloat162float(out[o]) + __bfloat162float(in[t])) : in[t];
}

// 16-byte vectorised variant (8 bf16 per thread): one row = c/8 vectors
__global__ void k_scatter_rows_bf16x8(const uint4* __restrict__ in, const int* __restrict__ dst, long long n, int vpr,
                                      uint4* __restrict__ out, int accumulate) {
  const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= n * vpr) return;
  const long long i = t / vpr;
  const int v = (int)(t - i * vpr);
  if (dst[i] < 0) return;
  const long long o = (long long)dst[i] * vpr + v;
  uint4 a = __ldg(in + t);
  if (accumulate) {
    const uint4 b = out[o];
    const unsigned aw[4] = {a.x, a.y, a.z, a.w}, bw[4] = {b.x, b.y, b.z, b.w};
    unsigned r[4];
#pragma unroll
    for (int e = 0; e < 4; ++e) {
      const float lo = __uint_as_float(aw[e] << 16) + __uint_as_float(bw[e] << 16);
      const float hi = __uint_as_float(aw[e] & 0xFFFF0000u) + __uint_as_float(bw[e] & 0xFFFF0000u);
      const __nv_bfloat162 h = __floats2bfloat162_rn(lo, hi);
      r[e] = *reinterpret_cast<const unsigned*>(&h);
    }
    a = make_uint4(r[0], r[1], r[2], r[3]);
  }
  out[o] = a;
}

// ---------------------------------------------------------------------------
// extern "C" entry points
// ---------------------------------------------------------------------------
static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

extern "C" {

size_t b2_scan_ws_bytes(long long n) { return b2_scan_ws_bytes_impl(n); }

size_t b2_map_build_ws_bytes(long long n) {
  // row_slot[n] i32 | pos[n] i32 | flags[n] u8 | scan ws
  return align_up((size_t)n * 4, 256) * 2 + align_up((size_t)n, 256) + align_up(b2_scan_ws_bytes_impl(n), 256) + 256;
}

int b2_scan_mask(const uint8_t* mask, long long n, int* pos, int* total, void* ws, size_t ws_bytes, void* stream) {
  return b2_scan_flags_launch(mask, n, pos, total, ws, ws_bytes, (cudaStream_t)stream);
}

int b2_table_clear(void* table, long long cap, void* stream) {
  B2_CHECK_ARG(cap > 0 && (cap & (cap - 1)) == 0, "hash capacity must be a power of two, got %lld", cap);
  B2_CUDA(cudaMemsetAsync(table, 0xFF, (size_t)cap * sizeof(B2Slot), (cudaStream_t)stream));
  return B2_OK;
}

int b2_map_build(const int* coords, long long n, int quant, void* table, long long cap, int* out_coords,
                 int* unique_index, int* inverse_map, int* n_out, int* err, void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2_CHECK_ARG(cap > 0 && (cap & (cap - 1)) == 0 && cap >= 2 * n, "hash capacity %lld must be pow2 >= 2n (n=%lld)", cap, n);
  B2_CHECK_ARG(n < (1ll << 31), "too many rows");
  if (b2_map_build_ws_bytes(n) > ws_bytes) {
    b2_set_error("b2_map_build workspace too small");
    return B2_ERR_WS;
  }
  B2_CUDA(cudaMemsetAsync(table, 0xFF, (size_t)cap * sizeof(B2Slot), st));
  if (n == 0) {
    B2_CUDA(cudaMemsetAsync(n_out, 0, sizeof(int), st));
    return B2_OK;
  }
  char* p = reinterpret_cast<char*>(ws);
  int* row_slot = reinterpret_cast<int*>(p);
  p += align_up((size_t)n * 4, 256);
  int* pos = reinterpret_cast<int*>(p);
  p += align_up((size_t)n * 4, 256);
  p += align_up((size_t)n, 256);   // (formerly one flag byte per row; the workspace layout is unchanged)
  void* scan_ws = p;
  const size_t scan_bytes = ws_bytes - (size_t)(p - reinterpret_cast<char*>(ws));
  const int T = 256, G = b2_div_up(n, T);
  k_hash_insert<<<G, T, 0, st>>>(coords, (int)n, quant, (B2Slot*)table, (unsigned)(cap - 1), row_slot, err);
  B2_LAUNCH_CHECK();
  if (inverse_map == nullptr)   // nobody asked for the inverse map: insert + scan-with-finalize, two kernels
    return b2_scan_winners_launch(row_slot, (B2Slot*)table, n, pos, n_out, scan_ws, scan_bytes, st, coords, quant, out_coords,
                                  unique_index);
  // the scan computes "row i won its slot" itself (k_scan_flags<1>), and so does the finalize kernel
  int rc = b2_scan_winners_launch(row_slot, (B2Slot*)table, n, pos, n_out, scan_ws, scan_bytes, st);
  if (rc) return rc;
  k_build_finalize<<<G, T, 0, st>>>(coords, (int)n, quant, (B2Slot*)table, row_slot, pos, out_coords, unique_index,
                                    inverse_map);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_map_table(const int* coords, long long n, void* table, long long cap, int* err, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  B2_CHECK_ARG(cap > 0 && (cap & (cap - 1)) == 0 && cap >= 2 * n, "hash capacity %lld must be pow2 >= 2n (n=%lld)", cap, n);
  B2_CUDA(cudaMemsetAsync(table, 0xFF, (size_t)cap * sizeof(B2Slot), st));
  if (n == 0) return B2_OK;
  k_hash_insert_unique<<<b2_div_up(n, 256), 256, 0, st>>>(coords, (int)n, (B2Slot*)table, (unsigned)(cap - 1), err);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_map_lookup(const void* table, long long cap, const int* query, long long m, int* rows, void* stream) {
  if (m == 0) return B2_OK;
  k_lookup<<<b2_div_up(m, 256), 256, 0, (cudaStream_t)stream>>>((const B2Slot*)table, (unsigned)(cap - 1), query, (int)m,
                                                              rows);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_map_children(const int* coords, long long n, int half, int* out_coords, void* stream) {
  B2_CHECK_ARG(half > 0, "generative transpose needs tensor_stride >= 2");
  if (n == 0) return B2_OK;
  k_children<<<b2_div_up(n * 8, 256), 256, 0, (cudaStream_t)stream>>>(coords, (int)n, half, out_coords);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_batch_offsets(const int* coords, long long n, int num_batches, int* offsets, int* err, void* stream) {
  B2_CHECK_ARG(num_batches > 0 && num_batches < 1023, "num_batches out of range");
  k_batch_offsets<<<b2_div_up(n + 1, 256), 256, 0, (cudaStream_t)stream>>>(coords, (int)n, num_batches, offsets, err);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// row count taken from device memory (*n_dev <= max_n): lets the caller fetch count and offsets with ONE host read
int b2_batch_offsets_dev(const int* coords, const int* n_dev, long long max_n, int num_batches, int* offsets, int* err,
                         void* stream) {
  B2_CHECK_ARG(num_batches > 0 && num_batches < 1023, "num_batches out of range");
  k_batch_offsets_dev<<<b2_div_up(max_n + 1, 256), 256, 0, (cudaStream_t)stream>>>(coords, n_dev, num_batches, offsets,
                                                                                 err);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

size_t b2_map_union_ws_bytes(long long nb) {
  // found[nb] i32 | pos[nb] i32 | missing[nb] u8 | total | scan
  return align_up((size_t)nb * 4, 256) * 2 + align_up((size_t)nb, 256) + 256 + align_up(b2_scan_ws_bytes_impl(nb), 256) + 256;
}

// a_off / b_off: int32[B+1] batch offsets of both maps.  n_out = na + #(b rows absent from a).
int b2_map_union(const void* a_table, long long a_cap, const int* a_coords, long long na, const int* a_off,
                 const int* b_coords, long long nb_rows, const int* b_off, int* out_coords, int* dst_a, int* dst_b,
                 int* n_out, void* ws, size_t ws_bytes, void* stream) {
  cudaStream_t st = (cudaStream_t)stream;
  if (b2_map_union_ws_bytes(nb_rows) > ws_bytes) {
    b2_set_error("b2_map_union workspace too small");
    return B2_ERR_WS;
  }
  char* p = reinterpret_cast<char*>(ws);
  int* found = reinterpret_cast<int*>(p);
  p += align_up((size_t)nb_rows * 4, 256);
  int* pos = reinterpret_cast<int*>(p);
  p += align_up((size_t)nb_rows * 4, 256);
  uint8_t* missing = reinterpret_cast<uint8_t*>(p);
  p += align_up((size_t)nb_rows, 256);
  int* total = reinterpret_cast<int*>(p);
  p += 256;
  void* scan_ws = p;
  const size_t scan_bytes = ws_bytes - (size_t)(p - reinterpret_cast<char*>(ws));
  if (nb_rows > 0) {
    k_union_probe<<<b2_div_up(nb_rows, 256), 256, 0, st>>>((const B2Slot*)a_table, (unsigned)(a_cap - 1), b_coords,
                                                          (int)nb_rows, found, missing);
    B2_LAUNCH_CHECK();
  }
  int rc = b2_scan_flags_launch(missing, nb_rows, pos, total, scan_ws, scan_bytes, st);
  if (rc) return rc;
  if (na + nb_rows > 0) {
    k_union_assign<<<b2_div_up(na + nb_rows, 256), 256, 0, st>>>(a_coords, (int)na, a_off, b_coords, (int)nb_rows, b_off,
                                                                found, pos, total, out_coords, dst_a, dst_b);
    B2_LAUNCH_CHECK();
  }
  k_add_total<<<1, 1, 0, st>>>(total, (int)na, n_out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_gather_coords(const int* coords, const uint8_t* mask, const int* pos, long long n, int* out, void* stream) {
  if (n == 0) return B2_OK;
  k_gather_coords<<<b2_div_up(n, 256), 256, 0, (cudaStream_t)stream>>>(coords, mask, pos, (int)n, out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// row_bytes must be a multiple of 4; rows are copied as 16-byte vectors when possible
int b2_gather_rows(const void* in, const uint8_t* mask, const int* pos, long long n, int row_bytes, void* out,
                   void* stream) {
  B2_CHECK_ARG(row_bytes > 0 && row_bytes % 2 == 0, "row_bytes must be even");
  if (n == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (row_bytes % 16 == 0) {
    const int v = row_bytes / 16;
    k_gather_rows<uint4><<<b2_div_up(n * v, 256), 256, 0, st>>>((const uint4*)in, mask, pos, n, v, (uint4*)out);
  } else if (row_bytes % 4 == 0) {
    const int v = row_bytes / 4;
    k_gather_rows<unsigned><<<b2_div_up(n * v, 256), 256, 0, st>>>((const unsigned*)in, mask, pos, n, v, (unsigned*)out);
  } else {
    const int v = row_bytes / 2;
    k_gather_rows<unsigned short><<<b2_div_up(n * v, 256), 256, 0, st>>>((const unsigned short*)in, mask, pos, n, v,
                                                                         (unsigned short*)out);
  }
  B2_LAUNCH_CHECK();
  return B2_OK;
}

int b2_index_rows(const void* in, const int* idx, long long n, int row_bytes, void* out, void* stream) {
  B2_CHECK_ARG(row_bytes > 0 && row_bytes % 2 == 0, "row_bytes must be even");
  if (n == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (row_bytes % 16 == 0) {
    const int v = row_bytes / 16;
    k_index_rows<uint4><<<b2_div_up(n * v, 256), 256, 0, st>>>((const uint4*)in, idx, n, v, (uint4*)out);
  } else if (row_bytes % 4 == 0) {
    const int v = row_bytes / 4;
    k_index_rows<unsigned><<<b2_div_up(n * v, 256), 256, 0, st>>>((const unsigned*)in, idx, n, v, (unsigned*)out);
  } else {
    const int v = row_bytes / 2;
    k_index_rows<unsigned short><<<b2_div_up(n * v, 256), 256, 0, st>>>((const unsigned short*)in, idx, n, v,
                                                                        (unsigned short*)out);
  }
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// dtype: 0 = f32, 1 = bf16.  accumulate=1 adds into out (rows of dst must be unique within one call).
int b2_scatter_rows(const void* in, const int* dst, long long n, int c, int dtype, int accumulate, void* out,
                    void* stream) {
  if (n == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (dtype == 0)
    k_scatter_rows_f32<<<b2_div_up(n * c, 256), 256, 0, st>>>((const float*)in, dst, n, c, (float*)out, accumulate);
  else if (c % 8 == 0)
    k_scatter_rows_bf16x8<<<b2_div_up(n * (c / 8), 256), 256, 0, st>>>((const uint4*)in, dst, n, c / 8, (uint4*)out,
                                                                       accumulate);
  else
    k_scatter_rows_bf16<<<b2_div_up(n * c, 256), 256, 0, st>>>((const __nv_bfloat16*)in, dst, n, c, (__nv_bfloat16*)out,
                                                              accumulate);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
