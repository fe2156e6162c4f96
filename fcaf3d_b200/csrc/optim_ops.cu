// Optimiser step of the training recipe as two launches over ALL parameter tensors (sm_100a).
//
// Reference recipe (configs/fcaf3d/fcaf3d.py:30-31): AdamW(lr 1e-3, weight_decay 1e-4) with gradient clipping at L2 norm 10
// (mmcv OptimizerHook -> torch.nn.utils.clip_grad_norm_ then torch.optim.AdamW.step).  Here:
//   b2_grad_sqnorm  squared L2 norm of every gradient element, deterministic (per-chunk partial sums, then one block adds
//                   them in chunk order)
//   b2_adamw_step   g' = g * min(1, max_norm / (sqrt(sqnorm) + 1e-6)); decoupled weight decay; first / second moment
//                   update; bias-corrected step - torch.optim.AdamW's arithmetic, element for element
// The tensors are addressed through a device table of (param, grad, exp_avg, exp_avg_sq) pointers + sizes and a chunk list
// (tensor index, element offset): one CTA per chunk of OPT_CHUNK elements, so 300 tensors of 27 elements .. 7 M elements
// load-balance without one launch per tensor.
#include "common.cuh"

#define OPT_CHUNK 16384
#define OPT_THREADS 256

struct B2OptTensor {
  float* p;
  const float* g;
  float* m;
  float* v;
  long long n;
};

__global__ void __launch_bounds__(OPT_THREADS) k_grad_sqnorm(const B2OptTensor* __restrict__ tab, const int2* __restrict__ chunks,
                                                             float* __restrict__ partial) {
  const int2 c = chunks[blockIdx.x];
  const B2OptTensor t = tab[c.x];
  const long long lo = (long long)c.y * OPT_CHUNK;
  const long long hi = min(lo + OPT_CHUNK, t.n);
  float s = 0.f;
  for (long long i = lo + threadIdx.x; i < hi; i += OPT_THREADS) {
    const float g = t.g[i];
    s += g * g;
  }
  __shared__ float red[OPT_THREADS / 32];
#pragma unroll
  for (int d = 16; d > 0; d >>= 1) s += __shfl_xor_sync(0xffffffffu, s, d);
  if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = s;
  __syncthreads();
  if (threadIdx.x == 0) {
    float tsum = 0.f;
    for (int w = 0; w < OPT_THREADS / 32; ++w) tsum += red[w];
    partial[blockIdx.x] = tsum;
  }
}

__global__ void __launch_bounds__(1024) k_sum_partials(const float* __restrict__ partial, int n, float* __restrict__ out) {
  __shared__ double red[1024];
  double s = 0.0;
  for (int i = threadIdx.x; i < n; i += 1024) s += (double)partial[i];
  red[threadIdx.x] = s;
  __syncthreads();
  for (int d = 512; d > 0; d >>= 1) {
    if (threadIdx.x < d) red[threadIdx.x] += red[threadIdx.x + d];
    __syncthreads();
  }
  if (threadIdx.x == 0) out[0] = (float)red[0];
}

__global__ void __launch_bounds__(OPT_THREADS) k_adamw(const B2OptTensor* __restrict__ tab, const int2* __restrict__ chunks,
                                                       float decay, float beta1, float beta2, float one_m_b1,
                                                       float one_m_b2, float eps, float step_size,
                                                       float bias_c2_sqrt, float max_norm,
                                                       const float* __restrict__ sqnorm) {
  const int2 c = chunks[blockIdx.x];
  const B2OptTensor t = tab[c.x];
  const long long lo = (long long)c.y * OPT_CHUNK;
  const long long hi = min(lo + OPT_CHUNK, t.n);
  float coef = 1.f;
  if (max_norm > 0.f && sqnorm != nullptr) coef = fminf(max_norm / (sqrtf(*sqnorm) + 1e-6f), 1.f);
  for (long long i = lo + threadIdx.x; i < hi; i += OPT_THREADS) {
    const float g = t.g[i] * coef;
    float p = t.p[i] * decay;
    const float m = beta1 * t.m[i] + one_m_b1 * g;
    const float v = beta2 * t.v[i] + one_m_b2 * g * g;
    const float denom = sqrtf(v) / bias_c2_sqrt + eps;
    p -= step_size * (m / denom);
    t.p[i] = p;
    t.m[i] = m;
    t.v[i] = v;
  }
}

extern "C" {

int b2_opt_chunk_elems(void) { return OPT_CHUNK; }

// table: n_tensors records of 5 x 8 bytes (param, grad, exp_avg, exp_avg_sq pointers, element count); chunks: n_chunks
// int32 pairs (tensor index, chunk index inside the tensor); partial: n_chunks floats of scratch
int b2_grad_sqnorm(const void* table, const int* chunks, int n_chunks, float* partial, float* sqnorm_out, void* stream) {
  B2_CHECK_ARG(n_chunks >= 0, "b2_grad_sqnorm: bad chunk count");
  cudaStream_t st = (cudaStream_t)stream;
  if (n_chunks == 0) {
    B2_CUDA(cudaMemsetAsync(sqnorm_out, 0, 4, st));
    return B2_OK;
  }
  B2_CHECK_ARG(table && chunks && partial && sqnorm_out, "b2_grad_sqnorm: null pointer");
  k_grad_sqnorm<<<n_chunks, OPT_THREADS, 0, st>>>((const B2OptTensor*)table, (const int2*)chunks, partial);
  B2_LAUNCH_CHECK();
  k_sum_partials<<<1, 1024, 0, st>>>(partial, n_chunks, sqnorm_out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// step: 1-based number of this update; max_norm <= 0 or sqnorm == NULL: no clipping
int b2_adamw_step(const void* table, const int* chunks, int n_chunks, double lr, double beta1, double beta2, double eps,
                  double wd, long long step, double max_norm, const float* sqnorm, void* stream) {
  B2_CHECK_ARG(n_chunks >= 0 && step >= 1, "b2_adamw_step: bad chunk count / step");
  if (n_chunks == 0) return B2_OK;
  B2_CHECK_ARG(table && chunks, "b2_adamw_step: null pointer");
  const double bc1 = 1.0 - pow(beta1, (double)step);
  const double bc2 = 1.0 - pow(beta2, (double)step);
  // the scalar factors are formed in double like torch forms them in python, then rounded once
  k_adamw<<<n_chunks, OPT_THREADS, 0, (cudaStream_t)stream>>>(
      (const B2OptTensor*)table, (const int2*)chunks, (float)(1.0 - lr * wd), (float)beta1, (float)beta2,
      (float)(1.0 - beta1), (float)(1.0 - beta2), (float)eps, (float)(lr / bc1), (float)sqrt(bc2), (float)max_norm, sqnorm);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

}  // extern "C"
