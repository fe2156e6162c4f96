// tcgen05 implicit-GEMM sparse convolution, two 128-row output tiles per CTA sharing every weight tile.
//
// Why: with one tile per CTA the kernel sits on the L2->SM fabric limit and a large share of that traffic is the
// per-offset weight slice W_k^T that every 128-row tile re-reads (C=64 layers: 221 KB of weights vs ~210 KB of
// gathered rows per tile; C>=256 layers: >75 % weights).  Here a CTA owns MT=2 tiles and two accumulators in TMEM
// (2 x BN columns); the weight slice of a (offset, 64-channel chunk) is staged ONCE in its own ring and consumed by the
// MMAs of both tiles, the gathered A tiles flow through a second ring.  Same producer / issuer / epilogue roles,
// same asynchronous cp.async->mbarrier completion and same fused epilogue as conv_tc.cu.
#include "tc_common.cuh"

template <int BN, int MT>
struct Tc2Cfg {
  static constexpr int SA = (BN <= 64) ? 4 : 3;          // A ring (16 KB stages)
  static constexpr int SB = 2;                           // B ring
  static constexpr int A_BYTES = TC_BM * TC_BK * 2;
  static constexpr int B_BYTES = BN * TC_BK * 2;
  static constexpr int RING_BYTES = SA * A_BYTES + SB * B_BYTES;
  static constexpr int NBR_BYTES = 27 * TC_BM * MT * 4;
  static constexpr int EPI_BYTES = 3 * BN * 4;
  static constexpr int SMEM = RING_BYTES + 1024 /*align*/ + 256 /*barriers*/ + NBR_BYTES + EPI_BYTES;
};

template <int BN, int MT, typename TOut>
__global__ void __launch_bounds__(TC_THREADS)
k_conv_tc2(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ wt, const int* __restrict__ nbr, int ld,
           const unsigned int* __restrict__ tile_mask, int n_out, int cin, int cout, int kvol,
           const float* __restrict__ bias, const float* __restrict__ scale, const float* __restrict__ shift,
           const TOut* __restrict__ residual, int act, int ch_mod, TOut* __restrict__ out) {
  using Cfg = Tc2Cfg<BN, MT>;
  constexpr int SA = Cfg::SA, SB = Cfg::SB;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint8_t* smem_b = smem + SA * Cfg::A_BYTES;
  uint64_t* a_full = reinterpret_cast<uint64_t*>(smem + Cfg::RING_BYTES);
  uint64_t* a_empty = a_full + SA;
  uint64_t* b_full = a_empty + SA;
  uint64_t* b_empty = b_full + SB;
  uint64_t* accum_bar = b_empty + SB;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  int* s_nbr = reinterpret_cast<int*>(smem + Cfg::RING_BYTES + 256);
  float* s_epi = reinterpret_cast<float*>(smem + Cfg::RING_BYTES + 256 + Cfg::NBR_BYTES);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int n_tiles = (n_out + TC_BM - 1) / TC_BM;
  const int tile0 = blockIdx.x * MT;
  const int row0 = tile0 * TC_BM, col0 = blockIdx.y * BN;
  const int chunks = cin / TC_BK;
  const unsigned int kall = (kvol >= 32) ? 0xFFFFFFFFu : ((1u << kvol) - 1u);
  unsigned int tmask[MT];
  unsigned int umask = 0;
#pragma unroll
  for (int t = 0; t < MT; ++t) {
    tmask[t] = (tile0 + t < n_tiles) ? ((nbr != nullptr) ? (tile_mask[tile0 + t] & kall) : 1u) : 0u;
    umask |= tmask[t];
  }

  if (tid == 0) {
    for (int s = 0; s < SA; ++s) {
      mbar_init(&a_full[s], TC_PRODUCERS);
      mbar_init(&a_empty[s], 1);
    }
    for (int s = 0; s < SB; ++s) {
      mbar_init(&b_full[s], TC_PRODUCERS);
      mbar_init(&b_empty[s], 1);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == TC_PRODUCERS / 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)(BN * MT))
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  // neighbour rows of the CTA's MT*128 output rows: s_nbr[k][r], r < MT*128
  constexpr int ROWS = TC_BM * MT;
  if (nbr != nullptr) {
    const int total = kvol * ROWS;
    for (int i0 = tid; i0 < total; i0 += TC_THREADS * 4) {
      int v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * TC_THREADS;
        const int k = i / ROWS, r = i - k * ROWS;
        v[u] = (i < total && ((umask >> k) & 1u) && row0 + r < ld) ? __ldg(nbr + (long long)k * ld + row0 + r) : -1;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * TC_THREADS < total) s_nbr[i0 + u * TC_THREADS] = v[u];
    }
  } else {
    for (int r = tid; r < ROWS; r += TC_THREADS) s_nbr[r] = (row0 + r < n_out) ? row0 + r : -1;
  }
  for (int j = tid; j < BN; j += TC_THREADS) {
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
  const uint32_t smem_a0 = smem_u32(smem), smem_b0 = smem_u32(smem_b);

  if (warp < TC_PRODUCERS / 32) {
    // ===================== producers ======================================================
    const int chunk16 = tid & 7, rsub = tid >> 3;
    const uint32_t dst_off = (uint32_t)(rsub * 128 + ((chunk16 ^ (rsub & 7)) << 4));
    const __nv_bfloat16* in_thr = in + chunk16 * 8;
    const __nv_bfloat16* w_thr = wt + (long long)(col0 + rsub) * cin + chunk16 * 8;
    const long long w_pass = 32ll * cin;
    int a_it = 0, b_it = 0;
    for (int k = 0; k < kvol; ++k) {
      if (!((umask >> k) & 1u)) continue;
      const __nv_bfloat16* src[MT][TC_BM / 32];
      uint32_t nbytes[MT][TC_BM / 32];
#pragma unroll
      for (int t = 0; t < MT; ++t) {
#pragma unroll
        for (int p = 0; p < TC_BM / 32; ++p) {
          const int r = s_nbr[k * ROWS + t * TC_BM + p * 32 + rsub];
          src[t][p] = in_thr + (r >= 0 ? (long long)r * cin : 0ll);
          nbytes[t][p] = r >= 0 ? 16u : 0u;
        }
      }
      const __nv_bfloat16* wk = w_thr + (long long)k * cout * cin;
      for (int c = 0; c < chunks; ++c) {
        {
          const int s = b_it % SB;
          mbar_wait(&b_empty[s], ((uint32_t)(b_it / SB) & 1u) ^ 1u);
          const uint32_t b_dst = smem_b0 + s * Cfg::B_BYTES + dst_off;
#pragma unroll
          for (int p = 0; p < BN / 32; ++p) cp_async16(b_dst + p * 4096, wk + p * w_pass + c * TC_BK, 16u);
          cp_async_arrive_noinc(&b_full[s]);
          ++b_it;
        }
#pragma unroll
        for (int t = 0; t < MT; ++t) {
          if (!((tmask[t] >> k) & 1u)) continue;
          const int s = a_it % SA;
          mbar_wait(&a_empty[s], ((uint32_t)(a_it / SA) & 1u) ^ 1u);
          const uint32_t a_dst = smem_a0 + s * Cfg::A_BYTES + dst_off;
#pragma unroll
          for (int p = 0; p < TC_BM / 32; ++p) cp_async16(a_dst + p * 4096, src[t][p] + c * TC_BK, nbytes[t][p]);
          cp_async_arrive_noinc(&a_full[s]);
          ++a_it;
        }
      }
    }

    // ===================== epilogue ========================================================
    if (umask != 0u) {
      mbar_wait(accum_bar, 0);
      tc_fence_after();
    }
    const int q = warp & 3;
    // MT == 2: warps 0-3 -> tile 0, warps 4-7 -> tile 1 (all BN columns); MT == 1: the two halves split the columns
    const int t_epi = (MT == 2) ? (warp >> 2) : 0;
    constexpr int CPW = (MT == 2) ? BN : ((BN >= 64) ? BN / 2 : BN);
    const int cbeg = (MT == 2) ? 0 : (warp >> 2) * CPW;
    const bool has_work = (MT == 2) ? true : (BN >= 64 || (warp >> 2) == 0);
    const int o = row0 + t_epi * TC_BM + q * 32 + lane;
    if (has_work) {
      for (int cc = cbeg; cc < cbeg + CPW; cc += 32) {
        uint32_t v[32];
        if (tmask[t_epi] != 0u) {
          tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(t_epi * BN + cc), v);
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
#pragma unroll
          for (int j = 0; j < 32; ++j) f[j] = b2_apply_act(f[j], act);
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
  } else {
    // ===================== MMA issuer (one thread) ==========================================
    if (lane == 0 && umask != 0u) {
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((128u >> 4) << 24);
      int a_it = 0, b_it = 0;
      uint32_t started = 0;
      for (int k = 0; k < kvol; ++k) {
        if (!((umask >> k) & 1u)) continue;
        for (int c = 0; c < chunks; ++c) {
          const int sb = b_it % SB;
          mbar_wait(&b_full[sb], (uint32_t)(b_it / SB) & 1u);
          const uint32_t b_base = smem_b0 + sb * Cfg::B_BYTES;
#pragma unroll
          for (int t = 0; t < MT; ++t) {
            if (!((tmask[t] >> k) & 1u)) continue;
            const int sa = a_it % SA;
            mbar_wait(&a_full[sa], (uint32_t)(a_it / SA) & 1u);
            fence_async_smem();  // cp.async (generic proxy) -> tcgen05.mma (async proxy)
            tc_fence_after();
            const uint32_t a_base = smem_a0 + sa * Cfg::A_BYTES;
#pragma unroll
            for (int kk = 0; kk < TC_BK / 16; ++kk) {
              const uint64_t da = umma_desc_sw128(a_base + kk * 32);
              const uint64_t db = umma_desc_sw128(b_base + kk * 32);
              umma_bf16(tmem_base + (uint32_t)(t * BN), da, db, idesc, (((started >> t) & 1u) || kk > 0) ? 1u : 0u);
            }
            started |= 1u << t;
            umma_commit(&a_empty[sa]);
            ++a_it;
          }
          umma_commit(&b_empty[sb]);  // the weight stage is free once the MMAs of every tile have read it
          ++b_it;
        }
      }
      umma_commit(accum_bar);
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == TC_PRODUCERS / 32) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)(BN * MT))
                 : "memory");
  }
}

template <int BN, int MT, typename TOut>
static int launch_tc2(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                      long long n_out, int cin, int cout, int kvol, const float* bias, const float* scale,
                      const float* shift, const void* residual, int act, int ch_mod, void* out, cudaStream_t st) {
  using Cfg = Tc2Cfg<BN, MT>;
  auto kern = k_conv_tc2<BN, MT, TOut>;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    configured = true;
  }
  const long long tiles = (n_out + TC_BM - 1) / TC_BM;
  dim3 grid((unsigned)((tiles + MT - 1) / MT), cout / BN);
  kern<<<grid, TC_THREADS, Cfg::SMEM, st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)wt, nbr, (int)ld, tile_mask,
                                            (int)n_out, cin, cout, kvol, bias, scale, shift, (const TOut*)residual, act,
                                            ch_mod, (TOut*)out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// Called by b2_conv_tc for layers with at least two row tiles and cout % 64 == 0.
int b2_conv_tc2_dispatch(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                         long long n_out, int cin, int cout, int kvol, const float* bias, const float* scale,
                         const float* shift, const void* residual, int act, int ch_mod, void* out, int out_dtype,
                         cudaStream_t st) {
#define TC2_GO(BN, T) \
  return launch_tc2<BN, 2, T>(in, wt, nbr, ld, tile_mask, n_out, cin, cout, kvol, bias, scale, shift, residual, act, ch_mod, out, st)
  if (out_dtype == 1) {
    if (cout % 128 == 0) TC2_GO(128, __nv_bfloat16);
    TC2_GO(64, __nv_bfloat16);
  } else {
    if (cout % 128 == 0) TC2_GO(128, float);
    TC2_GO(64, float);
  }
#undef TC2_GO
}
