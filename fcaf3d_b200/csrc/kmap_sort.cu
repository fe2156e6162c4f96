// Sparsity-aware row grouping of a 3^3 neighbour table (the "sorted implicit GEMM" idea of TorchSparse++, MLSys'23):
// output rows that use the same SLABS of the kernel are made adjacent, so that a 128-row tile of the tcgen05 convolution
// walks only the offsets its rows actually use.  FCAF3D's maps are surfaces about two voxels thick: a row has ~13 of the 27
// neighbours, but the union over 128 arbitrary rows is all 27.  The grouping key is 6 bits - "has a neighbour in the
// x-1 / x+1 / y-1 / y+1 / z-1 / z+1 slab of 9 offsets" - which separates floor-like, wall-like and edge rows: measured on the
// synthetic ScanNet-shape scenes 27 -> 15.8 offsets per tile at tensor stride 8 (a full lexicographic sort by the 27-bit
// mask gives 17.6, sorting by the key and then by the mask 15.2), 17.7 at stride 16.  With 64 buckets the sort is a
// counting sort: one pass for keys + per-block histograms, a scan of the histograms, one pass for the stable positions and
// one for the permuted table (which reads only the PRESENT entries: absent ones are known from the row mask).
// The coordinate map keeps its MinkowskiEngine row order; only the convolution walks the rows through `perm`:
//   perm[j]          = output row handled at position j (-1 in the padding)
//   nbr_sorted[k][j] = nbr[k][perm[j]],  tile_mask_sorted[t] = offsets used by positions 128 t .. 128 t + 127
// Per-row arithmetic (the order of the k accumulation) is unchanged, so results are bit-identical to the unsorted walk.
#include "common.cuh"

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

#define KS_BLOCK 1024
#define KS_KEYS 64

// bit masks over the 27 offsets (k = ix + 3 iy + 9 iz, x fastest) of the six outer slabs
__device__ __forceinline__ unsigned int slab_key(unsigned int m) {
  const unsigned int XM = 0x1249249u, XP = XM << 2;               // ix == 0 / ix == 2
  const unsigned int YM = 0x01C0E07u, YP = YM << 6;               // iy == 0 / iy == 2
  const unsigned int ZM = 0x00001FFu, ZP = ZM << 18;              // iz == 0 / iz == 2
  return ((m & XM) ? 1u : 0u) | ((m & XP) ? 2u : 0u) | ((m & YM) ? 4u : 0u) | ((m & YP) ? 8u : 0u) |
         ((m & ZM) ? 16u : 0u) | ((m & ZP) ? 32u : 0u);
}

// rowinfo[o] = 27-bit row mask (the key is recomputed from it); hist[key][block] = rows of the block with that key
__global__ void __launch_bounds__(KS_BLOCK) k_row_keys(const int* __restrict__ nbr, int ld, int n_out, int kvol,
                                                       unsigned int* __restrict__ rowinfo, int* __restrict__ hist,
                                                       int nblk) {
  __shared__ int s_hist[KS_KEYS];
  if (threadIdx.x < KS_KEYS) s_hist[threadIdx.x] = 0;
  __syncthreads();
  const int o = blockIdx.x * KS_BLOCK + threadIdx.x;
  if (o < n_out) {
    unsigned int m = 0;
    for (int k = 0; k < kvol; ++k) m |= (__ldg(nbr + (long long)k * ld + o) >= 0 ? 1u : 0u) << k;
    const unsigned int key = kvol == 27 ? slab_key(m) : 0u;
    rowinfo[o] = m;
    atomicAdd(&s_hist[key], 1);
  }
  __syncthreads();
  if (threadIdx.x < KS_KEYS) hist[threadIdx.x * nblk + blockIdx.x] = s_hist[threadIdx.x];
}

// exclusive scan of the key-major histogram array (one block; total = 64 * nblk entries) in place
__global__ void __launch_bounds__(1024) k_scan_hist(int* __restrict__ hist, int total) {
  __shared__ int s_sum[1024];
  const int per = (total + 1023) / 1024;
  const int lo = min(threadIdx.x * per, total), hi = min(lo + per, total);
  int s = 0;
  for (int i = lo; i < hi; ++i) s += hist[i];
  s_sum[threadIdx.x] = s;
  __syncthreads();
  for (int d = 1; d < 1024; d <<= 1) {   // Hillis-Steele inclusive scan
    const int v = threadIdx.x >= d ? s_sum[threadIdx.x - d] : 0;
    __syncthreads();
    s_sum[threadIdx.x] += v;
    __syncthreads();
  }
  int run = s_sum[threadIdx.x] - s;
  for (int i = lo; i < hi; ++i) {
    const int v = hist[i];
    hist[i] = run;
    run += v;
  }
}

// stable position of every row: base of (key, block) + rows of earlier warps with the key + rank inside the warp
__global__ void __launch_bounds__(KS_BLOCK) k_positions(const unsigned int* __restrict__ rowinfo, int n_out, int kvol,
                                                        const int* __restrict__ base, int nblk, int* __restrict__ perm,
                                                        unsigned int* __restrict__ mask_sorted) {
  __shared__ int s_cnt[KS_BLOCK / 32][KS_KEYS];
  const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
  for (int i = threadIdx.x; i < (KS_BLOCK / 32) * KS_KEYS; i += KS_BLOCK) (&s_cnt[0][0])[i] = 0;
  __syncthreads();
  const int o = blockIdx.x * KS_BLOCK + threadIdx.x;
  const bool live = o < n_out;
  const unsigned int info = live ? rowinfo[o] : 0u;
  const unsigned int key = live ? (kvol == 27 ? slab_key(info) : 0u) : 0xFFu;
  const unsigned int peers = __match_any_sync(0xffffffffu, key);
  const int rank = __popc(peers & ((1u << lane) - 1u));
  if (live && rank == 0) s_cnt[warp][key] = __popc(peers);
  __syncthreads();
  if (threadIdx.x < KS_KEYS) {
    int run = 0;
    for (int w = 0; w < KS_BLOCK / 32; ++w) {
      const int v = s_cnt[w][threadIdx.x];
      s_cnt[w][threadIdx.x] = run;
      run += v;
    }
  }
  __syncthreads();
  if (live) {
    const int p = base[key * nblk + blockIdx.x] + s_cnt[warp][key] + rank;
    perm[p] = o;
    mask_sorted[p] = info;
  }
}

// one CTA of 128 threads = one tile of sorted positions: tile mask, then the permuted table (coalesced writes; only present
// entries are read from the unsorted table)
__global__ void __launch_bounds__(128) k_permute_sorted(const int* __restrict__ nbr, int ld, int n_out, int kvol,
                                                        int* __restrict__ perm,
                                                        const unsigned int* __restrict__ mask_sorted,
                                                        int* __restrict__ nbr_sorted,
                                                        unsigned int* __restrict__ tile_mask_sorted) {
  __shared__ unsigned int s_or[4];
  const int j = blockIdx.x * 128 + threadIdx.x;
  const bool live = j < n_out;
  const int src = live ? perm[j] : -1;
  const unsigned int m = live ? mask_sorted[j] : 0u;
  if (!live) perm[j] = -1;
  unsigned int t = m;
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) t |= __shfl_xor_sync(0xffffffffu, t, d);
  if ((threadIdx.x & 31) == 0) s_or[threadIdx.x >> 5] = t;
  __syncthreads();
  const unsigned int tmask = s_or[0] | s_or[1] | s_or[2] | s_or[3];
  if (threadIdx.x == 0) tile_mask_sorted[blockIdx.x] = tmask;
  for (int k = 0; k < kvol; ++k) {
    int v = -1;
    if ((m >> k) & 1u) v = __ldg(nbr + (long long)k * ld + src);
    nbr_sorted[(long long)k * ld + j] = v;
  }
}

extern "C" {

size_t b2_kmap_sort_ws_bytes(long long n_out, int kvol) {
  (void)kvol;
  const long long nblk = (n_out + KS_BLOCK - 1) / KS_BLOCK;
  return 2 * align_up((size_t)(n_out > 0 ? n_out : 1) * 4, 256) + align_up((size_t)(nblk > 0 ? nblk : 1) * KS_KEYS * 4, 256) + 256;
}

// nbr [kvol, ld] (-1 = absent, padding rows -1) -> perm int32 [ld], nbr_sorted [kvol, ld], tile_mask_sorted [ld / 128]
int b2_kmap_sort(const int* nbr, long long ld, long long n_out, int kvol, int* perm, int* nbr_sorted,
                 unsigned int* tile_mask_sorted, void* ws, size_t ws_bytes, void* stream) {
  B2_CHECK_ARG(kvol >= 1 && kvol <= 27 && ld % 128 == 0 && n_out <= ld, "bad neighbour table shape");
  B2_CHECK_ARG(n_out < (1ll << 31), "too many rows");
  B2_CHECK_ARG(ws_bytes >= b2_kmap_sort_ws_bytes(n_out, kvol), "workspace too small");
  if (ld == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const int nblk = (int)((n_out + KS_BLOCK - 1) / KS_BLOCK);
  const size_t seg = align_up((size_t)(n_out > 0 ? n_out : 1) * 4, 256);
  uint8_t* w = (uint8_t*)ws;
  unsigned int* rowinfo = (unsigned int*)w;
  unsigned int* mask_s = (unsigned int*)(w + seg);
  int* hist = (int*)(w + 2 * seg);
  if (n_out > 0) {
    k_row_keys<<<nblk, KS_BLOCK, 0, st>>>(nbr, (int)ld, (int)n_out, kvol, rowinfo, hist, nblk);
    B2_LAUNCH_CHECK();
    k_scan_hist<<<1, 1024, 0, st>>>(hist, nblk * KS_KEYS);
    B2_LAUNCH_CHECK();
    k_positions<<<nblk, KS_BLOCK, 0, st>>>(rowinfo, (int)n_out, kvol, hist, nblk, perm, mask_s);
    B2_LAUNCH_CHECK();
  }
  k_permute_sorted<<<(unsigned)(ld / 128), 128, 0, st>>>(nbr, (int)ld, (int)n_out, kvol, perm, mask_s, nbr_sorted,
                                                        tile_mask_sorted);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
