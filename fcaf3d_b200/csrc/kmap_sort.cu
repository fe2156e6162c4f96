// Sparsity-aware row grouping of a neighbour table (the "sorted implicit GEMM" idea of TorchSparse++, MLSys'23):
// output rows with the same set of present kernel offsets are made adjacent, so that a 128-row tile of the tcgen05
// convolution touches only the offsets its rows actually use (on FCAF3D's surface-like maps the union over a random
// tile is all 27 offsets while a row has ~13: measured 27 -> 15.7 offsets per tile at tensor stride 8).
// The coordinate map keeps its MinkowskiEngine row order; only the convolution walks the rows through `perm`:
//   perm[j]          = output row handled at position j (-1 in the padding)
//   nbr_sorted[k][j] = nbr[k][perm[j]],  tile_mask_sorted[t] = offsets used by positions 128 t .. 128 t + 127
// Per-row arithmetic (the order of the k accumulation) is unchanged, so results are bit-identical to the unsorted walk.
#include "common.cuh"
#include <cub/device/device_radix_sort.cuh>

static inline size_t align_up(size_t v, size_t a) { return (v + a - 1) / a * a; }

__global__ void k_row_masks(const int* __restrict__ nbr, int ld, int n_out, int kvol, unsigned int* __restrict__ mask,
                            int* __restrict__ rows) {
  const int o = blockIdx.x * blockDim.x + threadIdx.x;
  if (o >= n_out) return;
  unsigned int m = 0;
  for (int k = 0; k < kvol; ++k) m |= (__ldg(nbr + (long long)k * ld + o) >= 0 ? 1u : 0u) << k;
  mask[o] = m;
  rows[o] = o;
}

// one CTA of 128 threads = one tile position block, blockIdx.y = kernel offset
__global__ void __launch_bounds__(128) k_permute_table(const int* __restrict__ nbr, int ld, int n_out,
                                                       const int* __restrict__ rows_sorted, int* __restrict__ perm,
                                                       int* __restrict__ nbr_sorted,
                                                       unsigned int* __restrict__ tile_mask_sorted) {
  const int j = blockIdx.x * 128 + threadIdx.x, k = blockIdx.y;
  const int src = (j < n_out) ? rows_sorted[j] : -1;
  const int v = (src >= 0) ? __ldg(nbr + (long long)k * ld + src) : -1;
  nbr_sorted[(long long)k * ld + j] = v;
  if (k == 0) perm[j] = src;
  const int any = __syncthreads_or(v >= 0);
  if (threadIdx.x == 0 && any) atomicOr(&tile_mask_sorted[blockIdx.x], 1u << k);
}

static size_t sort_temp_bytes(long long n, int kvol) {
  size_t bytes = 0;
  cub::DeviceRadixSort::SortPairs(nullptr, bytes, (const unsigned int*)nullptr, (unsigned int*)nullptr,
                                  (const int*)nullptr, (int*)nullptr, (int)n, 0, kvol);
  return bytes;
}

extern "C" {

size_t b2_kmap_sort_ws_bytes(long long n_out, int kvol) {
  return 4 * align_up((size_t)n_out * 4, 256) + align_up(sort_temp_bytes(n_out, kvol), 256) + 256;
}

// nbr [kvol, ld] (-1 = absent, padding rows -1) -> perm int32 [ld], nbr_sorted [kvol, ld], tile_mask_sorted [ld / 128]
int b2_kmap_sort(const int* nbr, long long ld, long long n_out, int kvol, int* perm, int* nbr_sorted,
                 unsigned int* tile_mask_sorted, void* ws, size_t ws_bytes, void* stream) {
  B2_CHECK_ARG(kvol >= 1 && kvol <= 32 && ld % 128 == 0 && n_out <= ld, "bad neighbour table shape");
  B2_CHECK_ARG(ws_bytes >= b2_kmap_sort_ws_bytes(n_out, kvol), "workspace too small");
  if (ld == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t seg = align_up((size_t)n_out * 4, 256);
  uint8_t* w = (uint8_t*)ws;
  unsigned int* mask = (unsigned int*)w;
  int* rows = (int*)(w + seg);
  unsigned int* mask_s = (unsigned int*)(w + 2 * seg);
  int* rows_s = (int*)(w + 3 * seg);
  void* temp = w + 4 * seg;
  size_t temp_bytes = sort_temp_bytes(n_out, kvol);
  B2_CUDA(cudaMemsetAsync(tile_mask_sorted, 0, (size_t)(ld / 128) * 4, st));
  if (n_out > 0) {
    k_row_masks<<<b2_div_up(n_out, 256), 256, 0, st>>>(nbr, (int)ld, (int)n_out, kvol, mask, rows);
    B2_LAUNCH_CHECK();
    B2_CUDA(cub::DeviceRadixSort::SortPairs(temp, temp_bytes, mask, mask_s, rows, rows_s, (int)n_out, 0, kvol, st));
    b2_count_launches(1);
  }
  dim3 grid((unsigned)(ld / 128), kvol);
  k_permute_table<<<grid, 128, 0, st>>>(nbr, (int)ld, (int)n_out, rows_s, perm, nbr_sorted, tile_mask_sorted);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
