// tcgen05 implicit-GEMM sparse convolution for sm_100a (bf16 operands, fp32 accumulators in TMEM).
//
// Output-stationary: one CTA owns a 128-row x BN-column output tile and walks the kernel offsets
// that the tile actually uses (tile_mask).  For every (offset k, 64-channel slice) stage, one of two
// groups of four producer warps gathers the 128 neighbour rows (cp.async, 16 B per request, zero-fill
// for missing neighbours) into 128B-swizzled shared memory - the groups take alternate stages -, lane 0
// of warp 9 fetches the matching [BN x 64] slice of W_k^T with one TMA tile load, and a single thread
// of warp 8 issues tcgen05.mma (M=128, N=BN, K=16) and releases the stage with tcgen05.commit
// (3, 4 or 6 stages).  The accumulator never leaves TMEM until the fused epilogue
// (bias, folded BatchNorm scale/shift, residual, ReLU/ELU), so there are no atomics, no scatter and
// no intermediate [pairs x C] buffers as in MinkowskiEngine's gather-GEMM-scatter
// (reference call sites: me_resnet.py:66-78 BasicBlock convs, fcaf3d_neck_with_head.py:52,60-69).
#include "common.cuh"

#include "tc_common.cuh"
#include <stdlib.h>

// 8 producer / epilogue warps, the MMA warp, and the warp whose lane 0 issues the weight-tile TMA loads
#define CT_THREADS 320

template <int BN, int ST>
struct TcCfg {
  static constexpr int STAGES = ST;
  static constexpr int A_BYTES = TC_BM * TC_BK * 2;
  static constexpr int B_BYTES = BN * TC_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int NBR_BYTES = 27 * TC_BM * 4;  // the tile's neighbour table (<= 27 offsets x 128 rows)
  static constexpr int EPI_BYTES = 3 * BN * 4;      // bias | scale | shift of the tile's columns
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + NBR_BYTES + EPI_BYTES;
};

// Warp roles: warps 0..7 = producers (cp.async gathers), then epilogue; warp 8 = TMEM owner + MMA issuer.
// Producer thread t: 16-byte chunk (t & 7) of rows (t >> 3) + 32 p; everything that does not depend on the stage is
// hoisted (the producers are instruction-latency bound: a stage must cost them as few dependent instructions as possible).
// GRP == 2: the producer warps work as two groups of four that fill alternate pipeline stages (each thread copies 8 rows'
// chunks of a stage instead of 4, every other stage).  A producer warp needs ~120 dependent instructions per stage it
// takes part in (wait, address arithmetic, copies, arrive) at ~8 cycles each, and that chain - not the L2 fabric or the
// tensor pipe - paced the C = 64 layers (ncu source view: 2/3 of the stall samples in the producer loop and its per-offset
// set-up); with two groups the per-stage fixed part of the chain is paid half as often per warp.  The weight tile's TMA
// load moves from thread 0 of the producers to lane 0 of a warp of its own.
// Head epilogue (EPI == 1, BN == 32: the fused 128 -> 1 + n_reg + n_cls head GEMM): instead of storing the [rows, 32] logits
// and running b2_head_post over them, every epilogue thread - it holds all 32 columns of its row - writes the head's outputs
// directly (forward_single, fcaf3d_neck_with_head.py:256-279: centerness, exp(scale * distances), class logits, row max for
// the pruning score; and max_c sigmoid(cls) * sigmoid(centerness) for the candidate top-k, :237-238).  Same expressions as
// k_head_post_tiled, hence bit-identical values.

#ifdef B2_TC_DBG
// timing experiments only (results are wrong with any bit set): 1 = no gathers, 2 = weight tiles only for the first S stages,
// 4 = no MMA, 8 = every neighbour absent (zero-fill copies only), 16 = no proxy fence before the MMAs, 32 = one arrival per
// producer warp instead of one per thread, 64 = no neighbour-table staging, 128 = no epilogue loads / stores, 256 = no TMEM read
__device__ int g_tc_dbg = 0;
extern "C" int b2_tc_debug(int v) { return cudaMemcpyToSymbol(g_tc_dbg, &v, sizeof(int)) == cudaSuccess ? 0 : -1; }
#define TC_DBG(bit) (dbg & (bit))
// cycle counters summed over CTAs: 0 CTAs, 1 stages, 2 prologue, 3 MMA thread waiting for full, 4 MMA thread issuing,
// 5 producer thread 0 waiting for empty, 6 producer thread 0 issuing, 7 weight thread waiting for empty,
// 8 waiting for the accumulator (producer thread 0, after its last stage), 9 epilogue, 10 CTA lifetime
__device__ unsigned long long g_tc_stat[16];
extern "C" int b2_tc_debug_stats(unsigned long long* out16, int reset) {
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(out16, g_tc_stat, sizeof(g_tc_stat)) != cudaSuccess) return -1;
  if (reset) {
    unsigned long long z[16] = {0};
    if (cudaMemcpyToSymbol(g_tc_stat, z, sizeof(z)) != cudaSuccess) return -1;
  }
  return 0;
}
#define TC_CLK() clock64()
#define TC_STAT(i, v) atomicAdd(&g_tc_stat[i], (unsigned long long)(v))
#else
#define TC_CLK() 0ll
#define TC_STAT(i, v)
#define TC_DBG(bit) false
#endif

// TIn = __nv_bfloat16 (kind::f16, 64 channels per 128-byte stage row) or float (kind::tf32: fp32 rows and weights read as
// TF32 operands, 32 channels per stage row, fp32 accumulation - the fp32-storage mode of the tensor-core path).
template <int BN, int ST, typename TOut, int GRP, typename TIn = __nv_bfloat16, int EPI = 0>
__global__ void __launch_bounds__(CT_THREADS)
k_conv_tc(const __grid_constant__ CUtensorMap map_b, const TIn* __restrict__ in, const int* __restrict__ nbr,
          int ld,
          const unsigned int* __restrict__ tile_mask, const int* __restrict__ out_rows, int n_out, int cin, int cout,
          int kvol, const float* __restrict__ bias, const float* __restrict__ scale, const float* __restrict__ shift,
          const TOut* __restrict__ residual, int act, int ch_mod, TOut* __restrict__ out, const TcHeadEpi hp = TcHeadEpi()) {
  using Cfg = TcCfg<BN, ST>;
  constexpr int S = Cfg::STAGES;
  constexpr int BKE = 128 / (int)sizeof(TIn);   // channels per stage row (128 bytes)
  constexpr int CH16 = 16 / (int)sizeof(TIn);   // channels per 16-byte copy
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);  // stays a shared-space pointer
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* accum_bar = empty_bar + S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  int* s_nbr = reinterpret_cast<int*>(smem + S * Cfg::STAGE_BYTES + 256);
  float* s_epi = reinterpret_cast<float*>(smem + S * Cfg::STAGE_BYTES + 256 + Cfg::NBR_BYTES);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
#ifdef B2_TC_DBG
  const int dbg = g_tc_dbg;
#endif
  const long long t_start = TC_CLK();
  // sorted tables put the tiles with the most offsets last: walk them first so that the heavy tiles do not form the tail
  const int tile = (out_rows != nullptr) ? (int)(gridDim.x - 1 - blockIdx.x) : (int)blockIdx.x;
  const int row0 = tile * TC_BM, col0 = blockIdx.y * BN;
  const unsigned int kmask = (nbr != nullptr) ? tile_mask[tile] : 1u;
  const int chunks = cin / BKE;
  // the staged neighbour table holds 16-byte units (row * row bytes / 16, -1 = absent), see tc_produce
  const int row_units = cin * (int)sizeof(TIn) / 16;
  const int nstage = __popc(kmask & ((1u << kvol) - 1u)) * chunks;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], TC_PRODUCERS / 2 + 1);  // one group's cp.async arrivals + the weight tile's expect_tx
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == TC_PRODUCERS / 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // stage the tile's neighbour rows in shared memory once (coalesced, independent loads in flight)
  if (nbr != nullptr && !TC_DBG(64)) {
    const int total = kvol * TC_BM;
    for (int i0 = tid; i0 < total; i0 += CT_THREADS * 4) {
      int v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * CT_THREADS;
        const int k = i >> 7, r = i & (TC_BM - 1);
        v[u] = (i < total && ((kmask >> k) & 1u)) ? __ldg(nbr + (long long)k * ld + row0 + r) : -1;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * CT_THREADS < total) s_nbr[i0 + u * CT_THREADS] = v[u] >= 0 ? v[u] * row_units : -1;
    }
  } else {
    for (int r = tid; r < TC_BM; r += CT_THREADS) s_nbr[r] = (row0 + r < n_out) ? (row0 + r) * row_units : -1;
  }
  for (int j = tid; j < BN; j += CT_THREADS) {
    const int n = col0 + j;
    const int ch = ch_mod > 0 ? (n % ch_mod) : n;
    s_epi[j] = bias ? __ldg(bias + ch) : 0.f;
    s_epi[BN + j] = scale ? __ldg(scale + ch) : 1.f;
    s_epi[2 * BN + j] = shift ? __ldg(shift + ch) : 0.f;
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  if (tid == 0) {
    TC_STAT(0, 1);
    TC_STAT(1, nstage);
    TC_STAT(2, TC_CLK() - t_start);
  }

  if (warp < TC_PRODUCERS / 32) {
    // ===================== producers ======================================================
    {
      const int group = warp >> 2, gt = tid & 127;
      const int chunk16 = gt & 7, rsub = gt >> 3;  // rsub 0..15; rows rsub + 16 p keep (row & 7) == (rsub & 7)
      const uint32_t dst_off = (uint32_t)(rsub * 128 + ((chunk16 ^ (rsub & 7)) << 4));  // + 2048 per pass (16 rows)
      const uint4* in_thr = reinterpret_cast<const uint4*>(in) + chunk16;
      const unsigned int km = kmask & ((1u << kvol) - 1u);
      int it = 0;
      if (chunks == 1)
        tc_produce<S, Cfg::STAGE_BYTES, 1>(km, 1, group, s_nbr + rsub, in_thr, smem_u32(smem) + dst_off, full_bar, empty_bar, it);
      else if (chunks == 2)
        tc_produce<S, Cfg::STAGE_BYTES, 2>(km, 2, group, s_nbr + rsub, in_thr, smem_u32(smem) + dst_off, full_bar, empty_bar, it);
      else if (chunks == 4)
        tc_produce<S, Cfg::STAGE_BYTES, 4>(km, 4, group, s_nbr + rsub, in_thr, smem_u32(smem) + dst_off, full_bar, empty_bar, it);
      else
        tc_produce<S, Cfg::STAGE_BYTES, 0>(km, chunks, group, s_nbr + rsub, in_thr, smem_u32(smem) + dst_off, full_bar,
                                           empty_bar, it);
    }

    // ===================== epilogue: TMEM -> registers -> global ===========================
    const long long t_e0 = TC_CLK();
    if (nstage > 0) {
      mbar_wait(accum_bar, 0);
      tc_fence_after();
    }
    const long long t_e1 = TC_CLK();
    if (tid == 0) TC_STAT(8, t_e1 - t_e0);
    const int q = warp & 3, half = warp >> 2;          // TMEM lane quarter, column half
    constexpr int CPW = (BN >= 64) ? BN / 2 : BN;       // columns per warp
    // tile position -> output row (identity, or the permutation of a sparsity-sorted neighbour table, kmap_sort.cu)
    int o = row0 + q * 32 + lane;
    if (out_rows != nullptr) {
      o = __ldg(out_rows + o);
      if (o < 0) o = n_out;
    }
    if (EPI == 1) {
      // (staging the rows through shared memory for coalesced block writes was measured slower: 212 against 199 us per step)
      if (half == 0) {
        uint32_t v[32];
        if (nstage > 0) {
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16), v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if (o < n_out) {
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = (__uint_as_float(v[j]) + s_epi[j]) * s_epi[BN + j] + s_epi[2 * BN + j];
          tc_head_row(f, (long long)o, hp);
        }
      }
    } else if ((BN >= 64 || half == 0) && !TC_DBG(128)) {
      for (int cc = half * CPW; cc < half * CPW + CPW; cc += 32) {
        uint32_t v[32];
        if (nstage > 0 && !TC_DBG(256)) {
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)cc, v);
        } else {
#pragma unroll
          for (int j = 0; j < 32; ++j) v[j] = 0u;
        }
        if (o < n_out) {
          const long long base = (long long)o * cout + col0 + cc;
          float f[32];
#pragma unroll
          for (int j = 0; j < 32; ++j)
            f[j] = (__uint_as_float(v[j]) + s_epi[cc + j]) * s_epi[BN + cc + j] + s_epi[2 * BN + cc + j];
          if (residual) {
            if (sizeof(TOut) == 2) {
              const uint4* rp = reinterpret_cast<const uint4*>(residual + base);
#pragma unroll
              for (int qq = 0; qq < 4; ++qq) {
                const uint4 u = __ldg(rp + qq);
                const uint32_t w4[4] = {u.x, u.y, u.z, u.w};
#pragma unroll
                for (int e = 0; e < 4; ++e) {
                  f[qq * 8 + e * 2] += __uint_as_float(w4[e] << 16);
                  f[qq * 8 + e * 2 + 1] += __uint_as_float(w4[e] & 0xFFFF0000u);
                }
              }
            } else {
              const float4* rp = reinterpret_cast<const float4*>(residual + base);
#pragma unroll
              for (int qq = 0; qq < 8; ++qq) {
                const float4 u = __ldg(rp + qq);
                f[qq * 4] += u.x;
                f[qq * 4 + 1] += u.y;
                f[qq * 4 + 2] += u.z;
                f[qq * 4 + 3] += u.w;
              }
            }
          }
          if (sizeof(TOut) == 2 && act == B2_ACT_ELU) {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = b2_elu_bf16(f[j]);
          } else {
#pragma unroll
            for (int j = 0; j < 32; ++j) f[j] = b2_apply_act(f[j], act);
          }
          if (sizeof(TOut) == 2) {
            uint4* op = reinterpret_cast<uint4*>(out + base);
#pragma unroll
            for (int qq = 0; qq < 4; ++qq) {
              uint32_t w4[4];
#pragma unroll
              for (int e = 0; e < 4; ++e) {
                const __nv_bfloat162 h = __floats2bfloat162_rn(f[qq * 8 + e * 2], f[qq * 8 + e * 2 + 1]);
                w4[e] = *reinterpret_cast<const uint32_t*>(&h);
              }
              op[qq] = make_uint4(w4[0], w4[1], w4[2], w4[3]);
            }
          } else {
            float4* op = reinterpret_cast<float4*>(out + base);
#pragma unroll
            for (int qq = 0; qq < 8; ++qq) op[qq] = make_float4(f[qq * 4], f[qq * 4 + 1], f[qq * 4 + 2], f[qq * 4 + 3]);
          }
        }
      }
    }
    tc_fence_before();
    if (tid == 0) TC_STAT(9, TC_CLK() - t_e1);
  } else if (warp == TC_PRODUCERS / 32 + 1) {
    // ===================== weight tiles (GRP == 2): one thread, one TMA tile load per stage ===
    if (lane == 0) {
      const uint32_t smem_base = smem_u32(smem);
      int it = 0;
      for (int k = 0; k < kvol; ++k) {
        if (!((kmask >> k) & 1u)) continue;
        for (int c = 0; c < chunks; ++c, ++it) {
          const int s = it % S;
          const long long t0 = TC_CLK();
          mbar_wait(&empty_bar[s], ((uint32_t)(it / S) & 1u) ^ 1u);
          TC_STAT(7, TC_CLK() - t0);
          if (TC_DBG(2) && it >= S) {
            mbar_arrive(&full_bar[s]);
            continue;
          }
          mbar_expect_tx(&full_bar[s], (uint32_t)Cfg::B_BYTES);
          tma_tile_2d(smem_base + s * Cfg::STAGE_BYTES + Cfg::A_BYTES, &map_b, &full_bar[s], c * BKE, k * cout + col0);
        }
      }
    }
    __syncwarp();
  } else {
    // ===================== MMA issuer (one elected thread; the warp waits together) ==========
    if (nstage > 0) {
      // instruction descriptor: D = F32; A / B format 1 = BF16 (kind::f16) or 2 = TF32 (kind::tf32); N >> 3; M >> 4
      constexpr uint32_t fmt = sizeof(TIn) == 2 ? 1u : 2u;
      constexpr uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((128u >> 4) << 24);
      // shared-memory descriptors differ only in their 14-bit (address >> 4) field: one descriptor, then additions
      // (no carry out of the field: shared addresses stay below 256 KB)
      const uint64_t desc0 = umma_desc_sw128(smem_u32(smem));
      const bool leader = elect_one();
      long long t_mw = 0, t_mi = 0;
      int s = 0;
      uint32_t phase = 0;
      for (int it = 0; it < nstage; ++it) {
        const long long t0 = TC_CLK();
        mbar_wait(&full_bar[s], phase);
        const long long t1 = TC_CLK();
        t_mw += t1 - t0;
        if (!TC_DBG(16)) fence_async_smem();  // cp.async wrote through the generic proxy, tcgen05.mma reads through the async proxy
        tc_fence_after();
        if (leader) {
          const uint64_t da = desc0 + (uint64_t)((s * Cfg::STAGE_BYTES) >> 4);
          const uint64_t db = da + (uint64_t)(Cfg::A_BYTES >> 4);
          if (!TC_DBG(4) || it == 0) {
#pragma unroll
            for (int kk = 0; kk < TC_BK / 16; ++kk) {
              if (sizeof(TIn) == 2)
                umma_bf16(tmem_base, da + 2 * kk, db + 2 * kk, idesc, (it > 0 || kk > 0) ? 1u : 0u);
              else
                umma_tf32(tmem_base, da + 2 * kk, db + 2 * kk, idesc, (it > 0 || kk > 0) ? 1u : 0u);
            }
          }
          umma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
        }
        __syncwarp();
        t_mi += TC_CLK() - t1;
        if (++s == S) {
          s = 0;
          phase ^= 1u;
        }
      }
      if (leader) {
        umma_commit(accum_bar);
        TC_STAT(3, t_mw);
        TC_STAT(4, t_mi);
      }
    }
    __syncwarp();
  }
  __syncthreads();
  if (tid == 0) TC_STAT(10, TC_CLK() - t_start);
  if (warp == TC_PRODUCERS / 32) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
  }
}

template <int BN, int ST, typename TOut, int GRP>
static int launch_tc_g(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                     const int* out_rows, long long n_out, int cin, int cout, int kvol, const float* bias, const float* scale,
                     const float* shift, const void* residual, int act, int ch_mod, void* out, cudaStream_t st) {
  using Cfg = TcCfg<BN, ST>;
  auto kern = k_conv_tc<BN, ST, TOut, GRP>;
  static unsigned long long configured = 0;  // per template instance, one bit per device
  if (int rc0 = b2_optin_smem((const void*)kern, Cfg::SMEM, &configured)) return rc0;
  CUtensorMap map_b;
  const int rc = b2_make_bf16_map(&map_b, wt, (long long)kvol * cout, cin, BN);
  if (rc) return rc;
  dim3 grid(b2_div_up(n_out, TC_BM), cout / BN);
  kern<<<grid, CT_THREADS, Cfg::SMEM, st>>>(map_b, (const __nv_bfloat16*)in, nbr, (int)ld, tile_mask, out_rows,
                                            (int)n_out, cin, cout, kvol, bias, scale, shift, (const TOut*)residual, act,
                                            ch_mod, (TOut*)out, TcHeadEpi());
  B2_LAUNCH_CHECK();
  return B2_OK;
}

template <int BN, int ST, typename TOut>
static int launch_tc(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                     const int* out_rows, long long n_out, int cin, int cout, int kvol, const float* bias, const float* scale,
                     const float* shift, const void* residual, int act, int ch_mod, void* out, cudaStream_t st) {
  return launch_tc_g<BN, ST, TOut, 2>(in, wt, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                                      residual, act, ch_mod, out, st);
}

template <int BN, int ST>
static int launch_tf32(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                       const int* out_rows, long long n_out, int cin, int cout, int kvol, const float* bias,
                       const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
                       cudaStream_t st) {
  using Cfg = TcCfg<BN, ST>;
  auto kern = k_conv_tc<BN, ST, float, 2, float>;
  static unsigned long long configured = 0;
  if (int rc0 = b2_optin_smem((const void*)kern, Cfg::SMEM, &configured)) return rc0;
  CUtensorMap map_b;
  const int rc = b2_make_tf32_map(&map_b, wt, (long long)kvol * cout, cin, BN);
  if (rc) return rc;
  dim3 grid(b2_div_up(n_out, TC_BM), cout / BN);
  kern<<<grid, CT_THREADS, Cfg::SMEM, st>>>(map_b, (const float*)in, nbr, (int)ld, tile_mask, out_rows, (int)n_out, cin, cout,
                                            kvol, bias, scale, shift, (const float*)residual, act, ch_mod, (float*)out, TcHeadEpi());
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// The three 1x1 heads as one 128 -> 32-column GEMM whose epilogue is the head post-processing (see TcHeadEpi)
extern "C" int b2_head_fused(const void* in_bf16, long long n, const void* wt_bf16, int cin, int cols, const float* bias,
                             int n_reg, int n_cls, const float* level_scale, float* centerness, float* bbox_pred, float* cls,
                             float* prune_score, float* max_score, void* stream) {
  B2_CHECK_ARG(cin % TC_BK == 0 && cin > 0, "b2_head_fused needs cin %% 64 == 0 (got %d)", cin);
  B2_CHECK_ARG(cols == 32 && 1 + n_reg + n_cls <= 32 && n_reg >= 6 && n_cls >= 1,
               "b2_head_fused: 1 + n_reg + n_cls (%d) must fit the 32 padded head columns", 1 + n_reg + n_cls);
  if (n == 0) return B2_OK;
  B2_CHECK_ARG(in_bf16 && wt_bf16 && centerness && bbox_pred && cls && prune_score, "b2_head_fused: null pointer");
  using Cfg = TcCfg<32, 4>;
  auto kern = k_conv_tc<32, 4, float, 2, __nv_bfloat16, 1>;
  static unsigned long long configured = 0;
  if (int rc0 = b2_optin_smem((const void*)kern, Cfg::SMEM, &configured)) return rc0;
  CUtensorMap map_b;
  const int rc = b2_make_bf16_map(&map_b, wt_bf16, cols, cin, 32);
  if (rc) return rc;
  TcHeadEpi hp;
  hp.centerness = centerness, hp.bbox_pred = bbox_pred, hp.cls = cls, hp.prune_score = prune_score, hp.max_score = max_score;
  hp.scale = level_scale, hp.n_reg = n_reg, hp.n_cls = n_cls;
  {
    int taken = 0;
    const int rcp = b2_head_tcp_dispatch(in_bf16, n, wt_bf16, cin, bias, hp, (cudaStream_t)stream, &taken);
    if (rcp != B2_OK || taken) return rcp;
  }
  kern<<<dim3(b2_div_up(n, TC_BM), 1), CT_THREADS, Cfg::SMEM, (cudaStream_t)stream>>>(
      map_b, (const __nv_bfloat16*)in_bf16, nullptr, 0, nullptr, nullptr, (int)n, cin, cols, 1, bias, nullptr, nullptr, nullptr,
      B2_ACT_NONE, 0, nullptr, hp);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// fp32 rows and fp32 weights [kvol, cout, cin] through the tensor cores as TF32 operands (10-bit mantissa products, fp32
// accumulation in TMEM), fp32 output: the fp32-storage mode of b2_conv_tc.  cin % 32 == 0, cout % 32 == 0.
extern "C" int b2_conv_tc_tf32(const void* in_f32, long long n_in, const void* wt_f32, const int* nbr, long long ld,
                               const unsigned int* tile_mask, const int* out_rows, long long n_out, int cin, int cout,
                               int kvol, const float* bias, const float* scale, const float* shift, const void* residual,
                               int act, int ch_mod, void* out, void* stream) {
  B2_CHECK_ARG(n_in * (cin / 4) < (1ll << 31), "b2_conv_tc_tf32: input of %lld rows x %d channels exceeds 2^31 16-byte units", n_in, cin);
  B2_CHECK_ARG(cin % 32 == 0 && cin > 0, "b2_conv_tc_tf32 needs cin %% 32 == 0 (got %d)", cin);
  B2_CHECK_ARG(cout % 32 == 0 && cout > 0, "b2_conv_tc_tf32 needs cout %% 32 == 0 (got %d)", cout);
  B2_CHECK_ARG(nbr != nullptr || kvol == 1, "identity map requires kernel volume 1");
  B2_CHECK_ARG(nbr == nullptr || (tile_mask != nullptr && kvol <= 27 && ld % TC_BM == 0),
               "neighbour table needs a tile mask, kvol <= 27 and ld %% 128 == 0");
  B2_CHECK_ARG((scale == nullptr) == (shift == nullptr), "scale and shift come together");
  B2_CHECK_ARG(out_rows == nullptr || nbr != nullptr, "a row permutation needs a neighbour table");
  if (n_out == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  {
    int taken = 0;
    const int rc = b2_conv_tcp_dispatch(in_f32, wt_f32, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                                        residual, act, ch_mod, out, 1, st, &taken);
    if (rc != B2_OK || taken) return rc;
  }
  if (cout % 128 == 0)
    return launch_tf32<128, 3>(in_f32, wt_f32, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                               residual, act, ch_mod, out, st);
  if (cout % 64 == 0)
    return launch_tf32<64, 4>(in_f32, wt_f32, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                              residual, act, ch_mod, out, st);
  return launch_tf32<32, 4>(in_f32, wt_f32, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                            residual, act, ch_mod, out, st);
}

extern "C" int b2_conv_tc(const void* in_bf16, long long n_in, const void* wt_bf16, const int* nbr, long long ld,
                          const unsigned int* tile_mask, const int* out_rows, long long n_out, int cin, int cout, int kvol, const float* bias,
                          const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
                          int out_dtype, void* stream) {
  B2_CHECK_ARG(cin % TC_BK == 0 && cin > 0, "b2_conv_tc needs cin %% 64 == 0 (got %d)", cin);
  B2_CHECK_ARG(n_in * (cin / 8) < (1ll << 31), "b2_conv_tc: input of %lld rows x %d channels exceeds 2^31 16-byte units", n_in, cin);
  B2_CHECK_ARG(cout % 32 == 0 && cout > 0, "b2_conv_tc needs cout %% 32 == 0 (got %d)", cout);
  B2_CHECK_ARG(nbr != nullptr || kvol == 1, "identity map requires kernel volume 1");
  B2_CHECK_ARG(nbr == nullptr || (tile_mask != nullptr && kvol <= 27 && ld % TC_BM == 0),
               "neighbour table needs a tile mask, kvol <= 27 and ld %% 128 == 0");
  B2_CHECK_ARG((scale == nullptr) == (shift == nullptr), "scale and shift come together");
  B2_CHECK_ARG(out_rows == nullptr || nbr != nullptr, "a row permutation needs a neighbour table");
  if (n_out == 0) return B2_OK;
  cudaStream_t st = (cudaStream_t)stream;
  if (out_dtype == 1) {
    int taken = 0;
    const int rc = b2_conv_tcp_dispatch(in_bf16, wt_bf16, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale,
                                        shift, residual, act, ch_mod, out, 0, st, &taken);
    if (rc != B2_OK || taken) return rc;
  }
  const bool wide = (cout % 128 == 0);
  const bool mid = (cout % 64 == 0);
#define TC_GO(BN, T)                                                                                               \
  return launch_tc<BN, (BN <= 64 ? 4 : 3), T>(in_bf16, wt_bf16, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, \
                                              shift, residual, act, ch_mod, out, st)
  // 256 output columns per CTA (one CTA per SM, 48 KB stages): half as many MMA instructions and gathered bytes per flop.
  // Measured: 256->256 on 12.5 k rows 63 -> 53 us; but 512->512 on 2.7 k rows (42 CTAs instead of 84) 78 -> 88 us, so it
  // is used only when it still yields at least half an SM wave of CTAs.  B2_TC_BN256=0 / 1 forces it off / on.
  static const int bn256_env = getenv("B2_TC_BN256") ? atoi(getenv("B2_TC_BN256")) : -1;
  const long long cta256 = (long long)b2_div_up(n_out, TC_BM) * (cout / 256);
  // dense GEMMs (identity map: generative transposes, heads) have a main loop of cin / 64 stages only and are bound by
  // the epilogue, which two resident CTAs per SM overlap better than one: 128-column tiles by default (generative
  // transposes of one step 249 -> 193 us against 256-column tiles; 64 columns are no better).  B2_TC_DENSE_BN = 64 /
  // 128 / 256 picks the tile width for A/B runs.
  static const int dense_bn = getenv("B2_TC_DENSE_BN") ? atoi(getenv("B2_TC_DENSE_BN")) : 128;
  const bool dense = (nbr == nullptr);
  if (dense && dense_bn == 64 && out_dtype == 1 && cout % 64 == 0)
    return launch_tc<64, 4, __nv_bfloat16>(in_bf16, wt_bf16, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias,
                                           scale, shift, residual, act, ch_mod, out, st);
  const bool no256 = dense && (dense_bn == 64 || dense_bn == 128);
  // Deep pipelines for grids of at most one CTA per SM (the deep levels: 84 - 99 CTAs): such a CTA has the whole shared
  // memory of its SM, and the layer is bound by latency x bytes in flight, so it takes 6 stages of 32 KB (128 columns) or
  // 4 of 48 KB (256 columns) instead of 3.  B2_TC_DEEP=0 keeps 3 stages.
  static const bool deep_ok = !(getenv("B2_TC_DEEP") && atoi(getenv("B2_TC_DEEP")) == 0);
  const long long tiles = b2_div_up(n_out, TC_BM);
  if (deep_ok && nbr != nullptr && out_dtype == 1 && kvol * (cin / TC_BK) >= 24) {
    const bool use256 = cout % 256 == 0 && bn256_env != 0 && (bn256_env == 1 || cta256 >= 74);
    if (use256 && cta256 <= 148)
      return launch_tc<256, 4, __nv_bfloat16>(in_bf16, wt_bf16, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias,
                                              scale, shift, residual, act, ch_mod, out, st);
    if (!use256 && wide && tiles * (cout / 128) <= 148)
      return launch_tc<128, 6, __nv_bfloat16>(in_bf16, wt_bf16, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias,
                                              scale, shift, residual, act, ch_mod, out, st);
  }
  if (cout % 256 == 0 && bn256_env != 0 && !no256 && (bn256_env == 1 || cta256 >= 74)) {
    if (out_dtype == 1)
      return launch_tc<256, 3, __nv_bfloat16>(in_bf16, wt_bf16, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias,
                                              scale, shift, residual, act, ch_mod, out, st);
    return launch_tc<256, 3, float>(in_bf16, wt_bf16, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale,
                                    shift, residual, act, ch_mod, out, st);
  }
  if (out_dtype == 1) {
    if (wide) TC_GO(128, __nv_bfloat16);
    if (mid) TC_GO(64, __nv_bfloat16);
    TC_GO(32, __nv_bfloat16);
  } else {
    if (wide) TC_GO(128, float);
    if (mid) TC_GO(64, float);
    TC_GO(32, float);
  }
#undef TC_GO
}
