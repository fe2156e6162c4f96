// Shared device/host helpers for the fcaf3d_b200 C-ABI library (sm_100a only).
#pragma once
#include <cuda_runtime.h>
#include <cuda_bf16.h>
#include <stdint.h>
#include <stdio.h>

#define B2_OK 0
#define B2_ERR_ARG -1
#define B2_ERR_CUDA -2
#define B2_ERR_WS -3

void b2_set_error(const char* fmt, ...);

#define B2_CHECK_ARG(cond, ...)                 \
  do {                                          \
    if (!(cond)) {                              \
      b2_set_error(__VA_ARGS__);                \
      return B2_ERR_ARG;                        \
    }                                           \
  } while (0)

#define B2_CUDA(call)                                                              \
  do {                                                                             \
    cudaError_t e_ = (call);                                                       \
    if (e_ != cudaSuccess) {                                                       \
      b2_set_error("%s:%d %s -> %s", __FILE__, __LINE__, #call, cudaGetErrorString(e_)); \
      return B2_ERR_CUDA;                                                          \
    }                                                                              \
  } while (0)

// every kernel launch of the library is counted (bench.py reports the number launched inside its timed region)
void b2_count_launches(int n);
#define B2_LAUNCH_CHECK()          \
  do {                             \
    b2_count_launches(1);          \
    B2_CUDA(cudaGetLastError());   \
  } while (0)

static inline int b2_div_up(long long a, long long b) { return (int)((a + b - 1) / b); }

// ---------------------------------------------------------------------------
// Packed voxel key: (batch:10 | x:18 | y:18 | z:18), biased so that unsigned
// key order == lexicographic (b, x, y, z) order.  Valid ranges: 0 <= b < 1023,
// -131072 <= x,y,z < 131072.
// ---------------------------------------------------------------------------
#define B2_COORD_BIAS 131072
#define B2_COORD_MASK 0x3FFFFull
#define B2_EMPTY_KEY 0xFFFFFFFFFFFFFFFFull

__host__ __device__ __forceinline__ bool b2_coord_in_range(int b, int x, int y, int z) {
  return (unsigned)b < 1023u && (unsigned)(x + B2_COORD_BIAS) < 262144u &&
         (unsigned)(y + B2_COORD_BIAS) < 262144u && (unsigned)(z + B2_COORD_BIAS) < 262144u;
}

__host__ __device__ __forceinline__ unsigned long long b2_pack_key(int b, int x, int y, int z) {
  return ((unsigned long long)(unsigned)b << 54) |
         ((unsigned long long)(unsigned)(x + B2_COORD_BIAS) << 36) |
         ((unsigned long long)(unsigned)(y + B2_COORD_BIAS) << 18) |
         ((unsigned long long)(unsigned)(z + B2_COORD_BIAS));
}

__host__ __device__ __forceinline__ unsigned int b2_hash_key(unsigned long long k) {
  k ^= k >> 33;
  k *= 0xff51afd7ed558ccdull;
  k ^= k >> 33;
  k *= 0xc4ceb9fe1a85ec53ull;
  k ^= k >> 33;
  return (unsigned int)k;
}

// One hash slot = 16 bytes so that a probe is a single 128-bit load.
struct __align__(16) B2Slot {
  unsigned long long key;
  unsigned int val;
  unsigned int pad;
};

__device__ __forceinline__ int b2_table_find(const B2Slot* __restrict__ table, unsigned int cap_mask,
                                             unsigned long long key) {
  unsigned int h = b2_hash_key(key) & cap_mask;
#pragma unroll 1
  for (;;) {
    const uint4 s = __ldg(reinterpret_cast<const uint4*>(table + h));
    const unsigned long long k = ((unsigned long long)s.y << 32) | s.x;
    if (k == key) return (int)s.z;
    if (k == B2_EMPTY_KEY) return -1;
    h = (h + 1) & cap_mask;
  }
}

__device__ __forceinline__ int b2_floor_div(int a, int b) {
  int q = a / b;
  int r = a % b;
  return (r != 0 && ((r < 0) != (b < 0))) ? q - 1 : q;
}

// activation codes shared by every fused epilogue
#define B2_ACT_NONE 0
#define B2_ACT_RELU 1
#define B2_ACT_ELU 2

__device__ __forceinline__ float b2_apply_act(float v, int act) {
  if (act == B2_ACT_RELU) return v > 0.f ? v : 0.f;
  if (act == B2_ACT_ELU) return v > 0.f ? v : expm1f(v);
  return v;
}

// ELU of a value that is stored as bf16 right afterwards.  expm1f costs ~40 instructions per element, which made the
// epilogues of the neck's short-K tcgen05 layers (a 128 x 256 tile has 32 k elements for 8 warps: ~17 us per tile
// whatever K was) and the bf16 normalisation passes arithmetic-bound.  exp(v) - 1 through ex2.approx away from zero and
// the cubic Taylor polynomial near zero stays within 1e-5 relative of expm1f, three orders of magnitude below the bf16
// rounding that follows.  fp32 outputs keep expm1f (b2_apply_act).
__device__ __forceinline__ float b2_elu_bf16(float v) {
  const float e = __expf(v) - 1.f;
  const float p = v * (1.f + v * (0.5f + v * (1.f / 6.f)));
  const float neg = (v > -0.03f) ? p : e;
  return v > 0.f ? v : neg;
}

__device__ __forceinline__ float b2_apply_act_bf16(float v, int act) {
  if (act == B2_ACT_RELU) return v > 0.f ? v : 0.f;
  if (act == B2_ACT_ELU) return b2_elu_bf16(v);
  return v;
}
