// Training-mode BatchNorm over the feature rows [N, C] fused with the activation and the residual add
// (MinkowskiBatchNorm = nn.BatchNorm1d on .F, ME App. A.6; me_resnet.py:63, BasicBlock; fcaf3d_neck_with_head.py:53,67,70).
// torch's channels-last batch-norm kernels are built for NHWC images; on [N, 64..512] row matrices with N up to 1.4 M they
// reach a fraction of HBM bandwidth (measured 8 ms of a 42 ms training step for 46 layers).  Here:
//   forward : k_bn_stats (one pass, shifted sums, BN_PARTS partial CTAs) -> k_bn_finalize (mean / invstd / folded scale and
//             shift in double, running statistics updated like nn.BatchNorm1d) -> b2_affine_act (normalise + residual + act)
//   backward: k_bn_bwd_reduce (sum dz, sum dz * xhat with dz = dy * act'(y)) -> k_bn_bwd_finalize -> k_bn_bwd_apply
//             dx = w * invstd * (dz - mean(dz) - xhat * mean(dz * xhat)), optional d(residual) = dz
// Rows are read 8 channels (16 or 32 bytes) per thread; C % 8 == 0, C <= 2048.  HBM-bound: forward 2 reads + 1 write of
// [N, C], backward 6 reads + 1..2 writes.
#include "common.cuh"
#include <algorithm>

#define BN_PARTS 592  // partial CTAs = 4 per SM
// (fewer partial CTAs for small layers - at least 128 rows each - was measured: no change of the training step)
static inline int bn_parts(long long n) { return (int)(n < BN_PARTS ? n : BN_PARTS); }

template <typename T>
__device__ __forceinline__ void load8(const T* p, float v[8]);
template <>
__device__ __forceinline__ void load8<float>(const float* p, float v[8]) {
  const float4 a = *reinterpret_cast<const float4*>(p), b = *reinterpret_cast<const float4*>(p + 4);
  v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
}
template <>
__device__ __forceinline__ void load8<__nv_bfloat16>(const __nv_bfloat16* p, float v[8]) {
  const uint4 u = *reinterpret_cast<const uint4*>(p);
  const uint32_t w[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    v[2 * e] = __uint_as_float(w[e] << 16);
    v[2 * e + 1] = __uint_as_float(w[e] & 0xFFFF0000u);
  }
}
template <typename T>
__device__ __forceinline__ void store8(T* p, const float v[8]);
template <>
__device__ __forceinline__ void store8<float>(float* p, const float v[8]) {
  *reinterpret_cast<float4*>(p) = make_float4(v[0], v[1], v[2], v[3]);
  *reinterpret_cast<float4*>(p + 4) = make_float4(v[4], v[5], v[6], v[7]);
}
template <>
__device__ __forceinline__ void store8<__nv_bfloat16>(__nv_bfloat16* p, const float v[8]) {
  uint32_t w[4];
#pragma unroll
  for (int e = 0; e < 4; ++e) {
    const __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * e], v[2 * e + 1]);
    w[e] = *reinterpret_cast<const uint32_t*>(&h);
  }
  *reinterpret_cast<uint4*>(p) = make_uint4(w[0], w[1], w[2], w[3]);
}

// derivative of the activation expressed through its output y (ReLU: y > 0; ELU(alpha = 1): y > 0 ? 1 : y + 1)
__device__ __forceinline__ float act_grad_from_output(float y, int act) {
  if (act == 1) return y > 0.f ? 1.f : 0.f;
  if (act == 2) return y > 0.f ? 1.f : y + 1.f;
  return 1.f;
}

// CTA -> rows [a, b); thread -> 8 channels (tl) of every rows_par-th row (rl); two per-channel sums per thread, reduced
// over rl in shared memory and written to partial[cta][2][c]
template <typename T, int MODE>  // MODE 0: sums of (x - k), (x - k)^2; MODE 1: sums of dz, dz * xhat
__global__ void __launch_bounds__(256) k_bn_partial(const T* __restrict__ x, const T* __restrict__ dy,
                                                    const T* __restrict__ y, long long n, int c, int act,
                                                    const float* __restrict__ mean, const float* __restrict__ invstd,
                                                    float* __restrict__ partial) {
  __shared__ float red[256 * 16];
  const int lanes = c >> 3, rows_par = 256 / lanes;
  const int tl = threadIdx.x % lanes, rl = threadIdx.x / lanes;
  const long long per = (n + gridDim.x - 1) / gridDim.x;
  const long long a = (long long)blockIdx.x * per, b = min(a + per, n);
  float s1[8], s2[8], k8[8], m8[8], i8[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) s1[i] = s2[i] = 0.f;
  if (rl < rows_par) {
    if (MODE == 0) {
      load8<T>(x + tl * 8, k8);  // shift = first row: keeps the one-pass variance well conditioned
    } else {
#pragma unroll
      for (int i = 0; i < 8; ++i) {
        m8[i] = mean[tl * 8 + i];
        i8[i] = invstd[tl * 8 + i];
      }
    }
    // U rows per trip with all their loads issued first (one row per trip left ~2 MB in flight on the whole GPU: latency
    // bound at ~40 % of HBM); the sums are still accumulated row by row in the same order
    constexpr int U = MODE == 0 ? 4 : 1;   // (two rows per trip in the backward sums cost occupancy: 1.26 -> 1.45 ms per step)
    long long r = a + rl;
    for (; r + (long long)(U - 1) * rows_par < b; r += (long long)U * rows_par) {
      float v[U][8], g[U][8], o[U][8];
#pragma unroll
      for (int u = 0; u < U; ++u) {
        const long long off = (r + (long long)u * rows_par) * c + tl * 8;
        load8<T>(x + off, v[u]);
        if (MODE == 1) {
          load8<T>(dy + off, g[u]);
          if (act != 0) load8<T>(y + off, o[u]);
        }
      }
#pragma unroll
      for (int u = 0; u < U; ++u) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          if (MODE == 0) {
            const float d = v[u][i] - k8[i];
            s1[i] += d;
            s2[i] += d * d;
          } else {
            const float dz = act != 0 ? g[u][i] * act_grad_from_output(o[u][i], act) : g[u][i];
            s1[i] += dz;
            s2[i] += dz * ((v[u][i] - m8[i]) * i8[i]);
          }
        }
      }
    }
    for (; r < b; r += rows_par) {
      float v[8];
      load8<T>(x + r * c + tl * 8, v);
      if (MODE == 0) {
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float d = v[i] - k8[i];
          s1[i] += d;
          s2[i] += d * d;
        }
      } else {
        float g[8], o[8];
        load8<T>(dy + r * c + tl * 8, g);
        if (act != 0) load8<T>(y + r * c + tl * 8, o);
#pragma unroll
        for (int i = 0; i < 8; ++i) {
          const float dz = act != 0 ? g[i] * act_grad_from_output(o[i], act) : g[i];
          s1[i] += dz;
          s2[i] += dz * ((v[i] - m8[i]) * i8[i]);
        }
      }
    }
  }
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    red[threadIdx.x * 16 + i] = s1[i];
    red[threadIdx.x * 16 + 8 + i] = s2[i];
  }
  __syncthreads();
  if (rl == 0) {
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      float t1 = 0.f, t2 = 0.f;
      for (int j = 0; j < rows_par; ++j) {
        t1 += red[(j * lanes + tl) * 16 + i];
        t2 += red[(j * lanes + tl) * 16 + 8 + i];
      }
      // layout [stat][channel][part]: the final reduction reads a channel's partials contiguously
      partial[((long long)(tl * 8 + i)) * gridDim.x + blockIdx.x] = t1;
      partial[((long long)(c + tl * 8 + i)) * gridDim.x + blockIdx.x] = t2;
    }
  }
}

// Final reduction of the partial sums: one warp per channel adds its contiguous run of partials (both statistics) in
// double and combines the lanes by shuffles (one thread per channel walking all partials was latency bound at 90 us per
// call, a 1024-thread CTA per 32 channels still 18 us: too few CTAs for 64-channel layers).
__device__ __forceinline__ void bn_reduce_parts(const float* __restrict__ partial, int parts, int c, int ch, double& s1,
                                                double& s2) {
  const int lane = threadIdx.x & 31;
  double a1 = 0.0, a2 = 0.0;
  const float* p1 = partial + (long long)ch * parts;
  const float* p2 = partial + (long long)(c + ch) * parts;
  for (int p = lane; p < parts; p += 32) {
    a1 += (double)p1[p];
    a2 += (double)p2[p];
  }
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) {
    a1 += __shfl_xor_sync(0xFFFFFFFFu, a1, o);
    a2 += __shfl_xor_sync(0xFFFFFFFFu, a2, o);
  }
  s1 = a1;
  s2 = a2;
}

// partial sums -> mean / invstd / folded scale, shift; running statistics like nn.BatchNorm1d
template <typename T>
__global__ void __launch_bounds__(256) k_bn_finalize(const float* __restrict__ partial, int parts, long long n, int c,
                                                     const T* __restrict__ x, const float* __restrict__ weight,
                                                     const float* __restrict__ bias, float* running_mean,
                                                     float* running_var, long long* num_batches_tracked, float momentum,
                                                     float eps, float* __restrict__ mean, float* __restrict__ invstd,
                                                     float* __restrict__ scale, float* __restrict__ shift) {
  const int ch = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (blockIdx.x == 0 && threadIdx.x == 0 && num_batches_tracked) *num_batches_tracked += 1;
  if (ch >= c) return;
  double s1, s2;
  bn_reduce_parts(partial, parts, c, ch, s1, s2);
  if ((threadIdx.x & 31) != 0) return;
  float k;
  {
    float k8[8];
    load8<T>(x + (ch & ~7), k8);
    k = k8[ch & 7];
  }
  const double m1 = s1 / (double)n;
  const double var = fmax(s2 / (double)n - m1 * m1, 0.0);
  const float mu = (float)((double)k + m1);
  const float is = (float)(1.0 / sqrt(var + (double)eps));
  mean[ch] = mu;
  invstd[ch] = is;
  const float w = weight ? weight[ch] : 1.f, b = bias ? bias[ch] : 0.f;
  scale[ch] = w * is;
  shift[ch] = b - mu * (w * is);
  if (running_mean) {
    const float unbiased = (float)(var * ((double)n / (double)(n > 1 ? n - 1 : 1)));
    running_mean[ch] = (1.f - momentum) * running_mean[ch] + momentum * mu;
    running_var[ch] = (1.f - momentum) * running_var[ch] + momentum * unbiased;
  }
}

__global__ void __launch_bounds__(256) k_bn_bwd_finalize(const float* __restrict__ partial, int parts, int c,
                                                         float* __restrict__ sums) {
  const int ch = blockIdx.x * 8 + (threadIdx.x >> 5);
  if (ch >= c) return;
  double s1, s2;
  bn_reduce_parts(partial, parts, c, ch, s1, s2);
  if ((threadIdx.x & 31) != 0) return;
  sums[ch] = (float)s1;      // d bias
  sums[c + ch] = (float)s2;  // d weight
}

template <typename T, int VPT>
__global__ void __launch_bounds__(256) k_bn_bwd_apply(const T* __restrict__ x, const T* __restrict__ dy,
                                                      const T* __restrict__ y, long long n, int c, int act,
                                                      const float* __restrict__ mean, const float* __restrict__ invstd,
                                                      const float* __restrict__ weight, const float* __restrict__ sums,
                                                      T* __restrict__ dx, T* __restrict__ dres) {
  // VPT vectors of 8 channels per thread, 256 vectors apart (coalesced per pass; the same channels when 256 % lanes == 0,
  // which the launcher guarantees for VPT > 1), all loads issued before the arithmetic
  const int lanes = c >> 3;
  const long long total = n * lanes;
  const long long t0 = (long long)blockIdx.x * (256 * VPT) + threadIdx.x;
  if (t0 >= total) return;
  const int ch0 = (int)(t0 % lanes) * 8;
  float v[VPT][8], g[VPT][8], o[VPT][8];
#pragma unroll
  for (int u = 0; u < VPT; ++u) {
    const long long t = t0 + u * 256;
    if (t < total) {
      load8<T>(x + t * 8, v[u]);
      load8<T>(dy + t * 8, g[u]);
      if (act != 0) load8<T>(y + t * 8, o[u]);
    }
  }
  const float inv_n = 1.f / (float)n;
  float is8[8], mean8[8], w8[8], sb8[8], sw8[8];
#pragma unroll
  for (int i = 0; i < 8; ++i) {
    const int ch = ch0 + i;
    is8[i] = invstd[ch];
    mean8[i] = mean[ch];
    w8[i] = weight ? weight[ch] : 1.f;
    sb8[i] = sums[ch] * inv_n;
    sw8[i] = sums[c + ch] * inv_n;
  }
#pragma unroll
  for (int u = 0; u < VPT; ++u) {
    const long long t = t0 + u * 256;
    if (t >= total) break;
    float dz[8], out[8];
#pragma unroll
    for (int i = 0; i < 8; ++i) {
      dz[i] = act != 0 ? g[u][i] * act_grad_from_output(o[u][i], act) : g[u][i];
      const float xhat = (v[u][i] - mean8[i]) * is8[i];
      out[i] = (w8[i] * is8[i]) * (dz[i] - sb8[i] - xhat * sw8[i]);
    }
    store8<T>(dx + t * 8, out);
    if (dres) store8<T>(dres + t * 8, dz);
  }
}

extern "C" {

size_t b2_bn_ws_bytes(int c) { return (size_t)BN_PARTS * 2 * c * sizeof(float) + 256; }

// x [n, c] (dtype 0 = f32, 1 = bf16) -> mean, invstd, scale = w * invstd, shift = b - mean * scale (all f32 [c]);
// running_mean / running_var (f32 [c], may be NULL) and num_batches_tracked (int64 [1], may be NULL) are updated in place
int b2_bn_train_stats(const void* x, int dtype, long long n, int c, const float* weight, const float* bias,
                      float* running_mean, float* running_var, long long* num_batches_tracked, float momentum, float eps,
                      float* mean, float* invstd, float* scale, float* shift, void* ws, size_t ws_bytes, void* stream) {
  B2_CHECK_ARG(c % 8 == 0 && c >= 8 && c <= 2048, "BatchNorm kernels need C %% 8 == 0 and C <= 2048 (got %d)", c);
  B2_CHECK_ARG(n > 0, "BatchNorm over an empty tensor");
  B2_CHECK_ARG((running_mean == nullptr) == (running_var == nullptr), "running statistics come together");
  if (ws_bytes < b2_bn_ws_bytes(c)) {
    b2_set_error("b2_bn_train_stats workspace too small");
    return B2_ERR_WS;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float* partial = (float*)ws;
  const int parts = bn_parts(n);
  if (dtype == 0) {
    k_bn_partial<float, 0><<<parts, 256, 0, st>>>((const float*)x, nullptr, nullptr, n, c, 0, nullptr, nullptr, partial);
    k_bn_finalize<float><<<b2_div_up(c, 8), 256, 0, st>>>(partial, parts, n, c, (const float*)x, weight, bias,
                                                            running_mean, running_var, num_batches_tracked, momentum, eps,
                                                            mean, invstd, scale, shift);
  } else {
    k_bn_partial<__nv_bfloat16, 0><<<parts, 256, 0, st>>>((const __nv_bfloat16*)x, nullptr, nullptr, n, c, 0, nullptr,
                                                         nullptr, partial);
    k_bn_finalize<__nv_bfloat16><<<b2_div_up(c, 8), 256, 0, st>>>(partial, parts, n, c, (const __nv_bfloat16*)x, weight,
                                                                    bias, running_mean, running_var, num_batches_tracked,
                                                                    momentum, eps, mean, invstd, scale, shift);
  }
  b2_count_launches(1);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// backward of y = act(bn(x) [+ residual]): sums f32 [2, c] = (d bias, d weight), dx [n, c], dres [n, c] or NULL
int b2_bn_act_bwd(const void* x, const void* dy, const void* y, int dtype, long long n, int c, int act, const float* mean,
                  const float* invstd, const float* weight, float* sums, void* dx, void* dres, void* ws, size_t ws_bytes,
                  void* stream) {
  B2_CHECK_ARG(c % 8 == 0 && c >= 8 && c <= 2048, "BatchNorm kernels need C %% 8 == 0 and C <= 2048 (got %d)", c);
  B2_CHECK_ARG(n > 0, "BatchNorm over an empty tensor");
  B2_CHECK_ARG(act == 0 || y != nullptr, "the activation's derivative needs the forward output");
  if (ws_bytes < b2_bn_ws_bytes(c)) {
    b2_set_error("b2_bn_act_bwd workspace too small");
    return B2_ERR_WS;
  }
  cudaStream_t st = (cudaStream_t)stream;
  float* partial = (float*)ws;
  const int parts = bn_parts(n);
  const int G = b2_div_up(n * (c / 8), 256);
  const bool vpt2 = 256 % (c / 8) == 0;   // two vectors per thread share their channels
  const int G2 = b2_div_up(n * (c / 8), 512);
  if (dtype == 0) {
    k_bn_partial<float, 1><<<parts, 256, 0, st>>>((const float*)x, (const float*)dy, (const float*)y, n, c, act, mean,
                                                  invstd, partial);
    k_bn_bwd_finalize<<<b2_div_up(c, 8), 256, 0, st>>>(partial, parts, c, sums);
    k_bn_bwd_apply<float, 1><<<G, 256, 0, st>>>((const float*)x, (const float*)dy, (const float*)y, n, c, act, mean, invstd,
                                                weight, sums, (float*)dx, (float*)dres);
  } else {
    k_bn_partial<__nv_bfloat16, 1><<<parts, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                         (const __nv_bfloat16*)y, n, c, act, mean, invstd, partial);
    k_bn_bwd_finalize<<<b2_div_up(c, 8), 256, 0, st>>>(partial, parts, c, sums);
    if (vpt2)
      k_bn_bwd_apply<__nv_bfloat16, 2><<<G2, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                          (const __nv_bfloat16*)y, n, c, act, mean, invstd, weight, sums,
                                                          (__nv_bfloat16*)dx, (__nv_bfloat16*)dres);
    else
      k_bn_bwd_apply<__nv_bfloat16, 1><<<G, 256, 0, st>>>((const __nv_bfloat16*)x, (const __nv_bfloat16*)dy,
                                                          (const __nv_bfloat16*)y, n, c, act, mean, invstd, weight, sums,
                                                          (__nv_bfloat16*)dx, (__nv_bfloat16*)dres);
  }
  b2_count_launches(2);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
