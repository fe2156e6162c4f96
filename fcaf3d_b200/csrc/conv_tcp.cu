// Persistent variant of the tcgen05 implicit-GEMM sparse convolution (conv_tc.cu) for layers with many row tiles.
//
// Cycle counters in k_conv_tc (tools/bench_conv.py, -DB2_TC_DBG) on the 64-channel levels: a CTA lives ~20 k cycles for a
// 16-stage main loop, of which ~3.6 k are its prologue (tile mask -> neighbour rows -> barrier set-up, TMEM allocation),
// ~3 k the accumulator drain and epilogue, and another ~25 % of the layer's time lies between CTAs (launch, tail); inside the
// main loop the two resident CTAs already pull ~60 B/clk, the fill bandwidth of one SM.  So the main loop has to run all
// the time: here a CTA stays resident and walks tiles blockIdx.x, blockIdx.x + gridDim.x, ...
//   * barriers, TMEM (two accumulators of BN columns) and the epilogue constants are set up once;
//   * the pipeline stages run on across tile boundaries (one global stage counter);
//   * four epilogue warps drain accumulator (j & 1) of tile j while the MMA warp already fills the other one with tile j + 1;
//   * the producers re-stage the neighbour rows of the next tile while the MMA warp still consumes the stages in flight.
// Warp roles: 0-7 producers (two groups of four warps on alternate stages, 8 rows x 16 B per thread and stage; with four
// producer warps - one per scheduler - the ~200 instructions a warp needs per stage were latency-bound), 8 MMA issuer,
// 9 weight-tile TMA, 10-13 epilogue (TMEM lane quarter = warp & 3).  Same arithmetic and accumulation order per output row as
// k_conv_tc: results are bit-identical.
#include "common.cuh"

#include "tc_common.cuh"
#include <stdlib.h>
#include <algorithm>

#define TCP_THREADS 448
#define TCP_PRODUCERS 256

template <int BN, int ST>
struct TcpCfg {
  static constexpr int STAGES = ST;
  static constexpr int A_BYTES = TC_BM * TC_BK * 2;
  static constexpr int B_BYTES = BN * TC_BK * 2;
  static constexpr int STAGE_BYTES = A_BYTES + B_BYTES;
  static constexpr int NBR_BYTES = 27 * TC_BM * 4;
  static constexpr int EPI_BYTES = 3 * BN * 4;
  static constexpr int SMEM = STAGES * STAGE_BYTES + 1024 /*align*/ + 256 /*barriers*/ + NBR_BYTES + EPI_BYTES;
  static constexpr int TMEM_COLS = 2 * BN;   // 128 or 256: a power of two >= 32
};

#ifdef B2_TC_DBG
// cycle counters summed over CTAs (tools/bench_conv.py): 0 CTAs, 1 stages, 2 tiles, 3 MMA waiting for full, 4 MMA issuing,
// 5 MMA waiting for a drained accumulator, 6 producer 0 staging neighbour rows, 7 producer 0 waiting for empty, 8 producer 0
// issuing, 9 epilogue thread waiting for the accumulator, 10 epilogue, 11 CTA lifetime, 12 prologue
__device__ unsigned long long g_tcp_stat[16];
extern "C" int b2_tcp_debug_stats(unsigned long long* out16, int reset) {
  if (cudaDeviceSynchronize() != cudaSuccess) return -1;
  if (cudaMemcpyFromSymbol(out16, g_tcp_stat, sizeof(g_tcp_stat)) != cudaSuccess) return -1;
  if (reset) {
    unsigned long long z[16] = {0};
    if (cudaMemcpyToSymbol(g_tcp_stat, z, sizeof(z)) != cudaSuccess) return -1;
  }
  return 0;
}
#define TCP_CLK() clock64()
__device__ __forceinline__ unsigned long long tcp_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
  return t;
}
#define TCP_STAT(i, v) atomicAdd(&g_tcp_stat[i], (unsigned long long)(v))
#else
#define TCP_CLK() 0ll
#define TCP_STAT(i, v)
#endif

__device__ __forceinline__ void tcp_bar_producers() { asm volatile("bar.sync 1, 256;" ::: "memory"); }

// one 32-column chunk of a row: (acc + bias) * scale + shift (+ residual), activation, store (bf16, or fp32 in the TF32 mode)
template <typename T>
__device__ __forceinline__ void tcp_store_chunk(const uint32_t* v, const float* s_epi, int bn, int cc,
                                                const T* __restrict__ residual, T* __restrict__ out, long long base,
                                                int act) {
  float f[32];
#pragma unroll
  for (int j = 0; j < 32; ++j) f[j] = (__uint_as_float(v[j]) + s_epi[cc + j]) * s_epi[bn + cc + j] + s_epi[2 * bn + cc + j];
  if (residual) {
    if (sizeof(T) == 2) {
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
  if (sizeof(T) == 2 && act == B2_ACT_ELU) {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = b2_elu_bf16(f[j]);
  } else {
#pragma unroll
    for (int j = 0; j < 32; ++j) f[j] = b2_apply_act(f[j], act);
  }
  if (sizeof(T) == 2) {
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

// Tile walked by CTA c of G in its round i (-1: none left).  Rounds alternate direction: tiles are ordered by their number
// of offsets (sorted tables), and a plain round-robin would give CTA 0 the heaviest tile of every round.
__device__ __forceinline__ int tcp_tile(int i, int c, int G, int n_tiles, bool rev) {
  const int idx = i * G + ((i & 1) ? G - 1 - c : c);
  return idx < n_tiles ? (rev ? n_tiles - 1 - idx : idx) : -1;
}

// EPI == 1 (BN == 32): the fused head GEMM, whose epilogue writes the head outputs (TcHeadEpi) instead of `out`.
// nbr == nullptr: identity map (dense GEMM over consecutive rows: generative transposes, heads), one stage per 64 channels.
// T = __nv_bfloat16 (kind::f16) or float (kind::tf32: fp32 rows and weights as TF32 operands, 32 channels per stage row,
// fp32 output - the persistent form of b2_conv_tc_tf32).
template <int BN, int ST, int EPI = 0, typename T = __nv_bfloat16>
__global__ void __launch_bounds__(TCP_THREADS, 2)
k_conv_tcp(const __grid_constant__ CUtensorMap map_b, const T* __restrict__ in, const int* __restrict__ nbr,
           int ld, const unsigned int* __restrict__ tile_mask, const int* __restrict__ out_rows, int n_out, int n_tiles,
           int cin, int cout, int kvol, const float* __restrict__ bias, const float* __restrict__ scale,
           const float* __restrict__ shift, const T* __restrict__ residual, int act, int ch_mod,
           T* __restrict__ out, const TcHeadEpi hp) {
  using Cfg = TcpCfg<BN, ST>;
  constexpr int S = Cfg::STAGES;
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem + S * Cfg::STAGE_BYTES);
  uint64_t* empty_bar = full_bar + S;
  uint64_t* acc_full = empty_bar + S;    // [2] accumulator b holds a finished tile
  uint64_t* acc_empty = acc_full + 2;    // [2] accumulator b has been drained
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);
  int* s_nbr = reinterpret_cast<int*>(smem + S * Cfg::STAGE_BYTES + 256);
  float* s_epi = reinterpret_cast<float*>(smem + S * Cfg::STAGE_BYTES + 256 + Cfg::NBR_BYTES);

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const long long t_start = TCP_CLK();
#ifdef B2_TC_DBG
  const unsigned long long ns_start = tcp_ns();
#endif
  const int col0 = blockIdx.y * BN;
  constexpr int BKE = 128 / (int)sizeof(T);   // channels per stage row (128 bytes)
  const int chunks = cin / BKE;
  const unsigned int kvalid = (1u << kvol) - 1u;
  const bool dense = nbr == nullptr;   // every tile has the one offset of the identity map
  const bool rev = out_rows != nullptr;   // sorted tables put the tiles with the most offsets last: walk them first
  const int step = gridDim.x;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], TCP_PRODUCERS / 2 + 1);  // one producer group's cp.async arrivals + the weight tile's expect_tx
      mbar_init(&empty_bar[s], 1);
    }
    for (int b = 0; b < 2; ++b) {
      mbar_init(&acc_full[b], 1);
      mbar_init(&acc_empty[b], 128);
    }
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == 8) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  for (int j = tid; j < BN; j += TCP_THREADS) {
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
    TCP_STAT(0, 1);
    TCP_STAT(12, TCP_CLK() - t_start);
  }

  int t = tcp_tile(0, blockIdx.x, step, n_tiles, rev);   // (every CTA has one: the grid is at most n_tiles wide)
  unsigned int kmask = t >= 0 ? (dense ? 1u : (__ldg(tile_mask + t) & kvalid)) : 0u;

  if (warp < 8) {
    // ===================== producers: neighbour rows of the tile, then cp.async gathers ===================
    const int group = warp >> 2, gt = tid & 127;
    const int chunk16 = gt & 7, rsub = gt >> 3;  // rsub 0..15; rows rsub + 16 p keep (row & 7) == (rsub & 7)
    const uint32_t dst_off = (uint32_t)(rsub * 128 + ((chunk16 ^ (rsub & 7)) << 4));  // + 2048 per pass (16 rows)
    // the staged neighbour table holds 16-byte units (row * cin / 8, -1 = absent): the multiplication is done once per entry
    // at staging time and a producer needs three instructions per row (load, sign test, 64-bit address)
    const uint4* in_thr = reinterpret_cast<const uint4*>(in) + chunk16;
    const int row_units = cin * (int)sizeof(T) / 16;
    const uint32_t smem_base = smem_u32(smem);
    int it = 0;
    long long t_ps = 0;
    for (int i = 0; t >= 0; ++i) {
      const int row0 = t * TC_BM;
      const int tn = tcp_tile(i + 1, blockIdx.x, step, n_tiles, rev);
      const unsigned int knext = tn >= 0 ? (dense ? 1u : (__ldg(tile_mask + tn) & kvalid)) : 0u;
      const long long t_s0 = TCP_CLK();
      tcp_bar_producers();  // every producer has finished with the previous tile's rows
      if (dense) {
        if (tid < TC_BM) s_nbr[tid] = (row0 + tid < n_out) ? (row0 + tid) * row_units : -1;
      } else {
        const int r = tid & (TC_BM - 1), kh = tid >> 7;   // offsets kh, kh + 2, ...
#pragma unroll
        for (int k0 = 0; k0 < 28; k0 += 14) {
          int v[7];
#pragma unroll
          for (int u = 0; u < 7; ++u) {
            const int k = k0 + 2 * u + kh;
            v[u] = (k < 27 && ((kmask >> k) & 1u)) ? __ldg(nbr + (long long)k * ld + row0 + r) : -1;
          }
#pragma unroll
          for (int u = 0; u < 7; ++u) {
            const int k = k0 + 2 * u + kh;
            if (k < 27) s_nbr[k * TC_BM + r] = v[u] >= 0 ? v[u] * row_units : -1;
          }
        }
      }
      tcp_bar_producers();
      t_ps += TCP_CLK() - t_s0;
      if (chunks == 1)
        tc_produce<S, Cfg::STAGE_BYTES, 1>(kmask, 1, group, s_nbr + rsub, in_thr, smem_base + dst_off, full_bar, empty_bar, it);
      else if (chunks == 2)
        tc_produce<S, Cfg::STAGE_BYTES, 2>(kmask, 2, group, s_nbr + rsub, in_thr, smem_base + dst_off, full_bar, empty_bar, it);
      else if (chunks == 4)
        tc_produce<S, Cfg::STAGE_BYTES, 4>(kmask, 4, group, s_nbr + rsub, in_thr, smem_base + dst_off, full_bar, empty_bar, it);
      else
        tc_produce<S, Cfg::STAGE_BYTES, 0>(kmask, chunks, group, s_nbr + rsub, in_thr, smem_base + dst_off, full_bar, empty_bar,
                                            it);
      kmask = knext;
      t = tn;
    }
    if (tid == 0) TCP_STAT(6, t_ps);
  } else if (warp == 8) {
    // ===================== MMA issuer (one elected thread; the warp waits together) =======================
    constexpr uint32_t fmt = sizeof(T) == 2 ? 1u : 2u;   // A / B format: BF16 or TF32
    constexpr uint32_t idesc = (1u << 4) | (fmt << 7) | (fmt << 10) | ((uint32_t)(BN >> 3) << 17) | ((128u >> 4) << 24);
    const uint64_t desc0 = umma_desc_sw128(smem_u32(smem));
    const bool leader = elect_one();
    int s = 0;
    uint32_t phase = 0;
    long long t_mw = 0, t_mi = 0, t_ma = 0;
    for (int j = 0; t >= 0; ++j) {
      const int tn = tcp_tile(j + 1, blockIdx.x, step, n_tiles, rev);
      const unsigned int knext = tn >= 0 ? (dense ? 1u : (__ldg(tile_mask + tn) & kvalid)) : 0u;
      const int nstage = __popc(kmask) * chunks;
      const int b = j & 1;
      const long long t_a0 = TCP_CLK();
      mbar_wait(&acc_empty[b], (((uint32_t)j >> 1) & 1u) ^ 1u);  // the epilogue has drained this accumulator (tile j - 2)
      t_ma += TCP_CLK() - t_a0;
      if (lane == 0) {
        TCP_STAT(1, nstage);
        TCP_STAT(2, 1);
      }
      tc_fence_after();
      const uint32_t d_tmem = tmem_base + (uint32_t)(b * BN);
      for (int i = 0; i < nstage; ++i) {
        const long long t0 = TCP_CLK();
        mbar_wait(&full_bar[s], phase);
        const long long t1 = TCP_CLK();
        t_mw += t1 - t0;
        fence_async_smem();  // cp.async wrote through the generic proxy, tcgen05.mma reads through the async proxy
        tc_fence_after();
        if (leader) {
          const uint64_t da = desc0 + (uint64_t)((s * Cfg::STAGE_BYTES) >> 4);
          const uint64_t db = da + (uint64_t)(Cfg::A_BYTES >> 4);
#pragma unroll
          for (int kk = 0; kk < TC_BK / 16; ++kk) {
            if (sizeof(T) == 2)
              umma_bf16(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (i > 0 || kk > 0) ? 1u : 0u);
            else
              umma_tf32(d_tmem, da + 2 * kk, db + 2 * kk, idesc, (i > 0 || kk > 0) ? 1u : 0u);
          }
          umma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
        }
        __syncwarp();
        t_mi += TCP_CLK() - t1;
        if (++s == S) {
          s = 0;
          phase ^= 1u;
        }
      }
      if (leader) umma_commit(&acc_full[b]);
      __syncwarp();
      kmask = knext;
      t = tn;
    }
    if (lane == 0) {
      TCP_STAT(3, t_mw);
      TCP_STAT(4, t_mi);
      TCP_STAT(5, t_ma);
    }
  } else if (warp == 9) {
    // ===================== weight tiles: one TMA tile load per stage =======================================
    const bool leader = elect_one();
    const uint32_t smem_base = smem_u32(smem);
    int s = 0;
    uint32_t phase = 0;
    for (int i = 0; t >= 0; ++i) {
      const int tn = tcp_tile(i + 1, blockIdx.x, step, n_tiles, rev);
      const unsigned int knext = tn >= 0 ? (dense ? 1u : (__ldg(tile_mask + tn) & kvalid)) : 0u;
      for (int k = 0; k < kvol; ++k) {
        if (!((kmask >> k) & 1u)) continue;
        for (int c = 0; c < chunks; ++c) {
          mbar_wait(&empty_bar[s], phase ^ 1u);
          if (leader) {
            mbar_expect_tx(&full_bar[s], (uint32_t)Cfg::B_BYTES);
            tma_tile_2d(smem_base + s * Cfg::STAGE_BYTES + Cfg::A_BYTES, &map_b, &full_bar[s], c * BKE, k * cout + col0);
          }
          __syncwarp();
          if (++s == S) {
            s = 0;
            phase ^= 1u;
          }
        }
      }
      kmask = knext;
      t = tn;
    }
  } else {
    // ===================== epilogue warps: TMEM -> registers -> global, one tile behind the MMA warp =======
    const int q = warp & 3;
    for (int j = 0; t >= 0; ++j) {
      const int tn = tcp_tile(j + 1, blockIdx.x, step, n_tiles, rev);
      const unsigned int knext = tn >= 0 ? (dense ? 1u : (__ldg(tile_mask + tn) & kvalid)) : 0u;
      const int row0 = t * TC_BM;
      int o = row0 + q * 32 + lane;
      if (out_rows != nullptr) {
        o = __ldg(out_rows + o);
        if (o < 0) o = n_out;
      }
      const int b = j & 1;
      const long long t_e0 = TCP_CLK();
      mbar_wait(&acc_full[b], ((uint32_t)j >> 1) & 1u);
      const long long t_e1 = TCP_CLK();
      tc_fence_after();
      const uint32_t taddr = tmem_base + (uint32_t)(b * BN) + ((uint32_t)(q * 32) << 16);
#pragma unroll 1
      for (int cc = 0; cc < BN; cc += 32) {
        uint32_t v[32];
        if (kmask != 0u) {
          tmem_ld32(taddr + (uint32_t)cc, v);
        } else {
#pragma unroll
          for (int e = 0; e < 32; ++e) v[e] = 0u;
        }
        if (EPI == 1) {
          if (o < n_out) {
            float f[32];
#pragma unroll
            for (int e = 0; e < 32; ++e) f[e] = (__uint_as_float(v[e]) + s_epi[e]) * s_epi[BN + e] + s_epi[2 * BN + e];
            tc_head_row(f, (long long)o, hp);
          }
        } else {
          if (o < n_out) tcp_store_chunk<T>(v, s_epi, BN, cc, residual, out, (long long)o * cout + col0 + cc, act);
        }
      }
      tc_fence_before();
      mbar_arrive(&acc_empty[b]);
      if (tid == 10 * 32) {
        TCP_STAT(9, t_e1 - t_e0);
        TCP_STAT(10, TCP_CLK() - t_e1);
      }
      kmask = knext;
      t = tn;
    }
  }
  tc_fence_before();
  __syncthreads();
  if (tid == 0) TCP_STAT(11, TCP_CLK() - t_start);
#ifdef B2_TC_DBG
  if (tid == 0) {
    const unsigned long long ns_end = tcp_ns();
    TCP_STAT(13, ns_end - ns_start);
    atomicMax(&g_tcp_stat[14], ns_end - ns_start);   // longest CTA (ns)
    atomicMax(&g_tcp_stat[15], (unsigned long long)(TCP_CLK() - t_start));   // longest CTA (cycles)
  }
#endif
  if (warp == 8) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base), "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
  }
}

template <int BN, int ST, typename T>
static int launch_tcp(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                      const int* out_rows, long long n_out, int cin, int cout, int kvol, const float* bias,
                      const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
                      int ctas, cudaStream_t st) {
  using Cfg = TcpCfg<BN, ST>;
  auto kern = k_conv_tcp<BN, ST, 0, T>;
  static unsigned long long configured = 0;  // per template instance, one bit per device
  if (int rc0 = b2_optin_smem((const void*)kern, Cfg::SMEM, &configured)) return rc0;
  CUtensorMap map_b;
  const int rc = sizeof(T) == 2 ? b2_make_bf16_map(&map_b, wt, (long long)kvol * cout, cin, BN)
                                : b2_make_tf32_map(&map_b, wt, (long long)kvol * cout, cin, BN);
  if (rc) return rc;
  const int n_tiles = (int)b2_div_up(n_out, TC_BM);
  const int gy = cout / BN;
  dim3 grid((unsigned)std::min(n_tiles, std::max(1, ctas / gy)), (unsigned)gy);
  kern<<<grid, TCP_THREADS, Cfg::SMEM, st>>>(map_b, (const T*)in, nbr, (int)ld, tile_mask, out_rows, (int)n_out, n_tiles, cin,
                                             cout, kvol, bias, scale, shift, (const T*)residual, act, ch_mod, (T*)out,
                                             TcHeadEpi());
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// 0 = by size (default), 1 = never, 2 = whenever the layer shape allows it (tests compare the two kernels bit for bit)
static int g_persist_mode = 0;
extern "C" int b2_tc_persist_mode(int mode) {
  const int old = g_persist_mode;
  if (mode >= 0 && mode <= 2) g_persist_mode = mode;
  return old;
}

// Called by b2_conv_tc for bf16 -> bf16 layers with a neighbour table.  *taken = 0: not a layer for the persistent kernel
// (few tiles, or more than 128 output columns) - the caller goes on to k_conv_tc.  B2_TC_PERSIST=0 switches it off,
// B2_TC_PERSIST_MIN_TILES (default: 2 resident CTAs x SMs x 2) is the smallest number of row tiles it takes.
int b2_conv_tcp_dispatch(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                         const int* out_rows, long long n_out, int cin, int cout, int kvol, const float* bias,
                         const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
                         int tf32, cudaStream_t st, int* taken) {
  *taken = 0;
  static const int enabled = getenv("B2_TC_PERSIST") ? atoi(getenv("B2_TC_PERSIST")) : 1;
  static const int min_env = getenv("B2_TC_PERSIST_MIN_TILES") ? atoi(getenv("B2_TC_PERSIST_MIN_TILES")) : 0;
  const bool dense = nbr == nullptr;   // identity map: any multiple of 128 output columns (column tiles on grid.y)
  if (!enabled || g_persist_mode == 1 || (dense ? (cout % 128 != 0 || kvol != 1) : (cout != 64 && cout != 128)))
    return B2_OK;
  static int sms[64] = {0};
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 64) return B2_OK;
  if (sms[dev] == 0) B2_CUDA(cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev));
  static const int ctas_env = getenv("B2_TC_PERSIST_CTAS") ? atoi(getenv("B2_TC_PERSIST_CTAS")) : 0;
  const int ctas = ctas_env > 0 ? ctas_env : 2 * sms[dev];
  const long long tiles = b2_div_up(n_out, TC_BM) * (dense ? cout / 128 : 1);
  if (g_persist_mode != 2 && tiles < (min_env > 0 ? min_env : 4 * sms[dev])) return B2_OK;
  *taken = 1;
  if (tf32) {   // fp32 rows / weights / output (b2_conv_tc_tf32)
    if (cout == 64)
      return launch_tcp<64, 4, float>(in, wt, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                                      residual, act, ch_mod, out, ctas, st);
    return launch_tcp<128, 3, float>(in, wt, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                                     residual, act, ch_mod, out, ctas, st);
  }
  if (cout == 64)
    return launch_tcp<64, 4, __nv_bfloat16>(in, wt, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                                            residual, act, ch_mod, out, ctas, st);
  return launch_tcp<128, 3, __nv_bfloat16>(in, wt, nbr, ld, tile_mask, out_rows, n_out, cin, cout, kvol, bias, scale, shift,
                                           residual, act, ch_mod, out, ctas, st);
}

// The fused head GEMM (b2_head_fused) on the persistent kernel: 32 head columns, epilogue = head post-processing.
int b2_head_tcp_dispatch(const void* in, long long n, const void* wt, int cin, const float* bias, const TcHeadEpi& hp,
                         cudaStream_t st, int* taken) {
  *taken = 0;
  static const int enabled = getenv("B2_TC_PERSIST") ? atoi(getenv("B2_TC_PERSIST")) : 1;
  static const int min_env = getenv("B2_TC_PERSIST_MIN_TILES") ? atoi(getenv("B2_TC_PERSIST_MIN_TILES")) : 0;
  if (!enabled || g_persist_mode == 1) return B2_OK;
  static int sms[64] = {0};
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  if (dev >= 64) return B2_OK;
  if (sms[dev] == 0) B2_CUDA(cudaDeviceGetAttribute(&sms[dev], cudaDevAttrMultiProcessorCount, dev));
  const int n_tiles = (int)b2_div_up(n, TC_BM);
  if (g_persist_mode != 2 && n_tiles < (min_env > 0 ? min_env : 4 * sms[dev])) return B2_OK;
  *taken = 1;
  using Cfg = TcpCfg<32, 4>;
  auto kern = k_conv_tcp<32, 4, 1>;
  static unsigned long long configured = 0;
  if (int rc0 = b2_optin_smem((const void*)kern, Cfg::SMEM, &configured)) return rc0;
  CUtensorMap map_b;
  const int rc = b2_make_bf16_map(&map_b, wt, 32, cin, 32);
  if (rc) return rc;
  kern<<<dim3((unsigned)std::min(n_tiles, 2 * sms[dev]), 1), TCP_THREADS, Cfg::SMEM, st>>>(
      map_b, (const __nv_bfloat16*)in, nullptr, 0, nullptr, nullptr, (int)n, n_tiles, cin, 32, 1, bias, nullptr, nullptr,
      nullptr, B2_ACT_NONE, 0, nullptr, hp);
  B2_LAUNCH_CHECK();
  return B2_OK;
}
