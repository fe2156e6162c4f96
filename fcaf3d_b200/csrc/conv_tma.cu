// tcgen05 implicit-GEMM sparse convolution with TMA operand staging (sm_100a).
//
// Same tiling, pipeline and fused epilogue as conv_tc.cu, but the operands are moved by the Tensor Memory Accelerator:
//   A (gathered neighbour rows)  cp.async.bulk.tensor.2d ... tile::gather4 : one instruction fetches FOUR arbitrary rows
//                                (64 bf16 = 128 B each) of the feature matrix into the 128B-swizzled stage; a missing
//                                neighbour is row index -1, which is out of bounds and therefore zero-filled by TMA;
//   B (W_k^T slice)              one cp.async.bulk.tensor.2d tile load of [BN x 64].
// One warp issues all copies of a stage (lane l: rows 4l..4l+3), completion is counted in bytes on the stage's
// mbarrier (expect_tx / complete_tx), so no thread ever waits on a load and no generic->async proxy fence is needed.
// The per-thread cp.async version is bound by LSU issue (a zero-filled row costs as much as a real one); here a stage
// costs the SM 33 TMA instructions instead of ~1500 thread-level LDGSTS.
#include "tc_common.cuh"
#include <mutex>

#define TMA_THREADS 192  // warp 0: TMA producer, warp 1: TMEM + MMA issuer, warps 2-5: epilogue

template <int BN>
struct TmaCfg {
  static constexpr int STAGES = (BN <= 64) ? 4 : 3;
  static constexpr int A_BYTES = TC_BM * TC_BK * 2;
  static constexpr int B_BYTES = BN * TC_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int NBR_BYTES = 27 * TC_BM * 4;
  static constexpr int EPI_BYTES = 3 * BN * 4;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 + 256 + NBR_BYTES + EPI_BYTES;
};

template <int BN, typename TOut>
__global__ void __launch_bounds__(TMA_THREADS)
k_conv_tma(const __grid_constant__ CUtensorMap map_a, const __grid_constant__ CUtensorMap map_b,
           const int* __restrict__ nbr, int ld, const unsigned int* __restrict__ tile_mask, int n_out, int cin, int cout,
           int kvol, const float* __restrict__ bias, const float* __restrict__ scale, const float* __restrict__ shift,
           const TOut* __restrict__ residual, int act, int ch_mod, TOut* __restrict__ out) {
  using Cfg = TmaCfg<BN>;
  constexpr int S = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* accum_bar = empty_bar + S;
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(accum_bar + 1);
  int* s_nbr = reinterpret_cast<int*>(smem + S * Cfg::STAGE_BYTES + 256);
  float* s_epi = reinterpret_cast<float*>(smem + S * Cfg::STAGE_BYTES + 256 + Cfg::NBR_BYTES);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int row0 = blockIdx.x * TC_BM, col0 = blockIdx.y * BN;
  const unsigned int kmask = ((nbr != nullptr) ? tile_mask[blockIdx.x] : 1u) & ((1u << kvol) - 1u);
  const int chunks = cin / TC_BK;
  const int nstage = __popc(kmask) * chunks;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], 1);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_a)) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(reinterpret_cast<uint64_t>(&map_b)) : "memory");
  }
  if (warp == 1) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)BN)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  if (nbr != nullptr) {
    const int total = kvol * TC_BM;
    for (int i0 = tid; i0 < total; i0 += TMA_THREADS * 4) {
      int v[4];
#pragma unroll
      for (int u = 0; u < 4; ++u) {
        const int i = i0 + u * TMA_THREADS;
        const int k = i >> 7, r = i & (TC_BM - 1);
        v[u] = (i < total && ((kmask >> k) & 1u)) ? __ldg(nbr + (long long)k * ld + row0 + r) : -1;
      }
#pragma unroll
      for (int u = 0; u < 4; ++u)
        if (i0 + u * TMA_THREADS < total) s_nbr[i0 + u * TMA_THREADS] = v[u];
    }
  } else {
    for (int r = tid; r < TC_BM; r += TMA_THREADS) s_nbr[r] = (row0 + r < n_out) ? row0 + r : -1;
  }
  for (int j = tid; j < BN; j += TMA_THREADS) {
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
  const uint32_t smem_base = smem_u32(smem);

  if (warp == 0) {
    // ===================== TMA producer warp ==============================================
    int it = 0;
    for (int k = 0; k < kvol; ++k) {
      if (!((kmask >> k) & 1u)) continue;
      const int4 rows = *reinterpret_cast<const int4*>(&s_nbr[k * TC_BM + lane * 4]);
      for (int c = 0; c < chunks; ++c, ++it) {
        const int s = it % S;
        if (lane == 0) {
          mbar_wait(&empty_bar[s], ((uint32_t)(it / S) & 1u) ^ 1u);
          mbar_expect_tx(&full_bar[s], (uint32_t)Cfg::STAGE_BYTES);
        }
        __syncwarp();
        const uint32_t a_dst = smem_base + s * Cfg::STAGE_BYTES;
        tma_gather4(a_dst + lane * 512, &map_a, &full_bar[s], c * TC_BK, rows.x, rows.y, rows.z, rows.w);
        if (lane == 0) tma_tile_2d(a_dst + Cfg::A_BYTES, &map_b, &full_bar[s], c * TC_BK, k * cout + col0);
      }
    }
  } else if (warp == 1) {
    // ===================== MMA issuer =====================================================
    if (lane == 0 && nstage > 0) {
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | ((uint32_t)(BN >> 3) << 17) | ((128u >> 4) << 24);
      for (int it = 0; it < nstage; ++it) {
        const int s = it % S;
        mbar_wait(&full_bar[s], (uint32_t)(it / S) & 1u);
        tc_fence_after();
        const uint32_t a_base = smem_base + s * Cfg::STAGE_BYTES;
        const uint32_t b_base = a_base + Cfg::A_BYTES;
#pragma unroll
        for (int kk = 0; kk < TC_BK / 16; ++kk) {
          const uint64_t da = umma_desc_sw128(a_base + kk * 32);
          const uint64_t db = umma_desc_sw128(b_base + kk * 32);
          umma_bf16(tmem_base, da, db, idesc, (it > 0 || kk > 0) ? 1u : 0u);
        }
        umma_commit(&empty_bar[s]);
      }
      umma_commit(accum_bar);
    }
    __syncwarp();
  } else {
    // ===================== epilogue warps 2..5 (TMEM lane quarter = warp % 4) ================
    if (nstage > 0) {
      mbar_wait(accum_bar, 0);
      tc_fence_after();
    }
    const int q = warp & 3;
    const int o = row0 + q * 32 + lane;
    for (int cc = 0; cc < BN; cc += 32) {
      uint32_t v[32];
      if (nstage > 0) {
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
    tc_fence_before();
  }
  __syncthreads();
  if (warp == 1) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)BN) : "memory");
  }
}

// ---------------------------------------------------------------------------
// host side: tensor maps through the driver entry point (the library does not link libcuda)
// ---------------------------------------------------------------------------
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult qres;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &qres) == cudaSuccess &&
        qres == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

// 2-D bf16 tensor [rows, cols] (cols contiguous), box = [box_rows x 64 columns], 128-byte swizzle, zero fill out of bounds
int b2_make_bf16_map(CUtensorMap* map, const void* base, long long rows, int cols, int box_rows) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    b2_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return B2_ERR_CUDA;
  }
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * 2};
  const cuuint32_t box[2] = {(cuuint32_t)TC_BK, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, const_cast<void*>(base), dims, strides, box, estr,
                  CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                  CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b2_set_error("cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%d box_rows=%d", (int)r, rows, cols, box_rows);
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

template <int BN, typename TOut>
static int launch_tma(const void* in, long long n_in, const void* wt, const int* nbr, long long ld,
                      const unsigned int* tile_mask, long long n_out, int cin, int cout, int kvol, const float* bias,
                      const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
                      cudaStream_t st) {
  using Cfg = TmaCfg<BN>;
  auto kern = k_conv_tma<BN, TOut>;
  static bool configured = false;
  if (!configured) {
    B2_CUDA(cudaFuncSetAttribute(kern, cudaFuncAttributeMaxDynamicSharedMemorySize, Cfg::SMEM));
    configured = true;
  }
  CUtensorMap map_a, map_b;
  int rc = b2_make_bf16_map(&map_a, in, n_in, cin, 1);
  if (rc) return rc;
  rc = b2_make_bf16_map(&map_b, wt, (long long)kvol * cout, cin, BN);
  if (rc) return rc;
  dim3 grid(b2_div_up(n_out, TC_BM), cout / BN);
  kern<<<grid, TMA_THREADS, Cfg::SMEM, st>>>(map_a, map_b, nbr, (int)ld, tile_mask, (int)n_out, cin, cout, kvol, bias,
                                             scale, shift, (const TOut*)residual, act, ch_mod, (TOut*)out);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// n_in = number of rows of the input feature matrix (bounds of the gather; rows outside are zero-filled)
int b2_conv_tma_dispatch(const void* in, long long n_in, const void* wt, const int* nbr, long long ld,
                         const unsigned int* tile_mask, long long n_out, int cin, int cout, int kvol, const float* bias,
                         const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
                         int out_dtype, cudaStream_t st) {
#define TMA_GO(BN, T) \
  return launch_tma<BN, T>(in, n_in, wt, nbr, ld, tile_mask, n_out, cin, cout, kvol, bias, scale, shift, residual, act, ch_mod, out, st)
  if (out_dtype == 1) {
    if (cout % 128 == 0) TMA_GO(128, __nv_bfloat16);
    if (cout % 64 == 0) TMA_GO(64, __nv_bfloat16);
    TMA_GO(32, __nv_bfloat16);
  } else {
    if (cout % 128 == 0) TMA_GO(128, float);
    if (cout % 64 == 0) TMA_GO(64, float);
    TMA_GO(32, float);
  }
#undef TMA_GO
}
