// tcgen05 weight gradient of the sparse convolution (sm_100a, bf16 operands, fp32 accumulators in TMEM).
//
//   dW[k][ci][co] = sum_o in[nbr[k][o]][ci] * dout[o][co]
//
// is ONE GEMM  D[(k, ci), co] = A^T G  with  M = kvol * cin,  N = cout  and the reduction over the OUTPUT ROWS o:
// A[o][(k, ci)] = in[nbr[k][o]][ci] (gathered, zero where the neighbour is absent) and G = dout.  Both operands are
// stored row by row with the channels contiguous, i.e. they are "MN-major" for the tensor core: a gathered piece of
// 64 channels is a 128-byte row, 8 consecutive rows form one 1024-byte SWIZZLE_128B atom, and the UMMA shared-memory
// descriptor walks the 64-channel groups with the leading byte offset and the 8-row groups with the stride byte offset
// (canonical layout ((8,n),(8,k)):((1,LBO),(8,SBO)) in 16-byte units).  No transposes anywhere.
//
// CTA = WG_MT (4) M tiles of 128 dW rows (= 8 "sub-tiles" of [32 rows x 64 channels], a sub-tile = one kernel offset k
// and one 64-channel slice of cin) x BN columns x one chunk of output rows.  Per stage of 32 output rows the 8 producer
// warps gather 8 sub-tiles of A and BN/64 sub-tiles of G with cp.async (completion on the stage's mbarrier), one thread
// issues WG_MT x 2 tcgen05.mma (M=128, N=BN, K=16) that share the G tile, tcgen05.commit releases the stage.  The
// WG_MT accumulators live side by side in TMEM (WG_MT * BN <= 512 columns) until the row chunk is exhausted; partial
// sums of the row chunks are combined with 16-byte fp32 reductions into the zero-initialised dW (split-K).
//
// Reference behaviour: MinkowskiEngine v0.5.4 ConvolutionBackwardKernelGPU (grad_kernel; SURVEY App. A.4, not vendored).
#include "tc_common.cuh"
#include <stdlib.h>

#define WG_KR 32                    // output rows per stage (two K=16 MMA steps)
#define WG_SUB_BYTES (WG_KR * 128)  // one [32 rows x 64 channels] bf16 sub-tile, four swizzle atoms
#define WG_MT 4                     // M tiles per CTA
#define WG_NSUB (2 * WG_MT)         // A sub-tiles per CTA and stage
#define WG_IB 8                     // stages per neighbour-index block
#define WG_MAX_STAGES 8

template <int BN>
struct WgCfg {
  static constexpr int A_BYTES = WG_NSUB * WG_SUB_BYTES;    // 32 KB
  static constexpr int G_BYTES = (BN / 64) * WG_SUB_BYTES;  // 4 or 8 KB
  static constexpr int STAGE_BYTES = A_BYTES + G_BYTES;
  static constexpr int IDX_BYTES = 2 * WG_NSUB * WG_IB * WG_KR * 4;  // double-buffered index blocks
  static constexpr int MAX_SMEM = 227 * 1024;
  static constexpr int STAGES = (MAX_SMEM - 2048 - IDX_BYTES) / STAGE_BYTES;  // 5 for both tile widths
  static constexpr int smem_bytes(int stages) { return stages * STAGE_BYTES + 1024 /*align*/ + 1024 /*barriers*/ + IDX_BYTES; }
  static constexpr int TMEM_COLS = WG_MT * BN;  // 256 or 512
};

// MN-major operand, 128-byte swizzle: 64-element groups `lbo` bytes apart, 8-row (K) groups 1024 bytes apart
__device__ __forceinline__ uint64_t umma_desc_mn_sw128(uint32_t saddr, uint32_t lbo) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)((lbo >> 4) & 0x3FFF) << 16;
  d |= (uint64_t)(1024 >> 4) << 32;
  d |= (uint64_t)1 << 46;  // descriptor version (sm_100)
  d |= (uint64_t)2 << 61;  // SWIZZLE_128B
  return d;
}

template <int BN, int SHARE>
__global__ void __launch_bounds__(TC_THREADS)
k_wgrad_tc(const __nv_bfloat16* __restrict__ in, const __nv_bfloat16* __restrict__ dout, const int* __restrict__ nbr,
           int ld, int n_out, int cin, int cout, int kvol, int rows_per_cta, float* __restrict__ dw, int dbg) {
  using Cfg = WgCfg<BN>;
  constexpr int S = Cfg::STAGES;
  // shared memory: [0, 1 KB) barriers | index blocks | S stages (1024-byte aligned)
  extern __shared__ uint8_t smem_raw[];
  uint8_t* smem0 = smem_raw + ((1024u - (smem_u32(smem_raw) & 1023u)) & 1023u);
  uint64_t* full_bar = reinterpret_cast<uint64_t*>(smem0);
  uint64_t* empty_bar = full_bar + WG_MAX_STAGES;
  uint64_t* accum_bar = empty_bar + WG_MAX_STAGES;
  uint64_t* idx_bar = accum_bar + 1;  // [2]
  uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(idx_bar + 2);
  int* s_idx = reinterpret_cast<int*>(smem0 + 1024);
  uint8_t* smem = smem0 + 1024 + Cfg::IDX_BYTES;

  const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
  const int c64n = cin >> 6;                       // 64-channel slices of cin
  const int nsub = kvol * c64n;                    // sub-tiles along M: s = k * c64n + slice
  const int s0 = blockIdx.y * WG_NSUB;
  const int nsub_cta = min(WG_NSUB, nsub - s0);    // >= 1
  const int nmt = (nsub_cta + 1) >> 1;             // M tiles in use (the last may have an all-zero upper half)
  const int n0 = blockIdx.z * BN;
  const int r_begin = blockIdx.x * rows_per_cta, r_end = min(r_begin + rows_per_cta, n_out);
  const int nstage = (r_end - r_begin + WG_KR - 1) / WG_KR;

  if (tid == 0) {
    for (int s = 0; s < S; ++s) {
      mbar_init(&full_bar[s], TC_PRODUCERS);
      mbar_init(&empty_bar[s], 1);
    }
    mbar_init(accum_bar, 1);
    mbar_init(&idx_bar[0], TC_PRODUCERS);
    mbar_init(&idx_bar[1], TC_PRODUCERS);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
  }
  if (warp == TC_PRODUCERS / 32) {
    asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(tmem_slot)),
                 "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
    asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
  }
  tc_fence_before();
  __syncthreads();
  tc_fence_after();
  const uint32_t tmem_base = *tmem_slot;
  const uint32_t smem_base = smem_u32(smem);

  if (warp < TC_PRODUCERS / 32) {
    // ===================== producers ======================================================
    // thread t: 16-byte chunk (t & 7) of row (t >> 3) of every sub-tile of the stage
    const int chunk16 = tid & 7, r = tid >> 3;
    const uint32_t dst_off = (uint32_t)(r * 128 + ((chunk16 ^ (r & 7)) << 4));
    // Neighbour indices are staged through shared memory in blocks of WG_IB stages (WG_IB * 32 rows x the CTA's
    // distinct kernel offsets), one block ahead, with coalesced 16-byte cp.async.  The stage loop itself is kept to a
    // few dozen instructions (the producers are instruction-issue bound): SHARE is a template parameter, pointers
    // advance incrementally, rows past the end and unused sub-tiles are -1 in the staged indices (the neighbour table
    // is padded with -1 up to ld) so the loop needs no bounds logic for A.
    constexpr int NK = WG_NSUB / SHARE;    // distinct kernel offsets of this CTA
    constexpr int IBR = WG_IB * WG_KR;     // rows per index block
    const int k0 = s0 / c64n;
    const __nv_bfloat16* in_thr = in + (s0 % c64n) * 64 + chunk16 * 8;
    auto issue_idx_block = [&](int b) {
      int* dstb = s_idx + (b & 1) * (WG_NSUB * IBR);
      const int o0 = r_begin + b * IBR;
      for (int e = tid; e < NK * (IBR / 4); e += TC_PRODUCERS) {
        const int dk = e / (IBR / 4), c4 = (e % (IBR / 4)) * 4;
        int* dst = dstb + dk * IBR + c4;
        if (s0 + dk * SHARE >= nsub) {
          *reinterpret_cast<int4*>(dst) = make_int4(-1, -1, -1, -1);
        } else if (nbr != nullptr) {
          const bool ok = o0 + c4 + 3 < ld;
          cp_async16(smem_u32(dst), nbr + (long long)(k0 + dk) * ld + (ok ? o0 + c4 : 0), ok ? 16u : 0u);
        } else {
          const int o = o0 + c4;
          *reinterpret_cast<int4*>(dst) = make_int4(o < r_end ? o : -1, o + 1 < r_end ? o + 1 : -1,
                                                    o + 2 < r_end ? o + 2 : -1, o + 3 < r_end ? o + 3 : -1);
        }
      }
      cp_async_arrive_noinc(&idx_bar[b & 1]);
    };
    issue_idx_block(0);
    const int* my_idx = s_idx;
    int o = r_begin + r;
    const __nv_bfloat16* g_ptr = dout + (long long)o * cout + n0 + chunk16 * 8;
    const long long g_step = (long long)WG_KR * cout;
    int s = 0;
    uint32_t ph = 1;  // parity to wait for on the stage's empty barrier
    for (int it = 0; it < nstage; ++it) {
      if ((it & (WG_IB - 1)) == 0) {
        const int b = it / WG_IB;
        // every producer is past the last stage of block b-1, so its buffer may be overwritten with block b+1
        asm volatile("bar.sync 1, %0;" ::"n"(TC_PRODUCERS) : "memory");
        if ((b + 1) * WG_IB < nstage) issue_idx_block(b + 1);
        mbar_wait(&idx_bar[b & 1], (uint32_t)(b >> 1) & 1u);
        my_idx = s_idx + (b & 1) * (WG_NSUB * IBR) + r;
      }
      int v[NK];
#pragma unroll
      for (int dk = 0; dk < NK; ++dk) v[dk] = my_idx[dk * IBR];
      my_idx += WG_KR;
      mbar_wait(&empty_bar[s], ph);
      const uint32_t a_dst = smem_base + s * Cfg::STAGE_BYTES + dst_off;
      if (!(dbg & 2)) {
#pragma unroll
        for (int dk = 0; dk < NK; ++dk) {
          const bool ok = v[dk] >= 0;
          const __nv_bfloat16* src = in_thr + (ok ? (long long)v[dk] * cin : 0ll);
          const uint32_t nb = ok ? 16u : 0u;
#pragma unroll
          for (int j = 0; j < SHARE; ++j) cp_async16(a_dst + (dk * SHARE + j) * WG_SUB_BYTES, src + j * 64, nb);
        }
      }
      const bool okg = o < r_end;
#pragma unroll
      for (int g = 0; g < BN / 64; ++g)
        cp_async16(a_dst + Cfg::A_BYTES + g * WG_SUB_BYTES, okg ? g_ptr + g * 64 : dout, okg ? 16u : 0u);
      cp_async_arrive_noinc(&full_bar[s]);
      o += WG_KR;
      g_ptr += g_step;
      if (++s == S) {
        s = 0;
        ph ^= 1u;
      }
    }

    // ===================== epilogue: TMEM -> fp32 reductions into dW ========================
    mbar_wait(accum_bar, 0);
    tc_fence_after();
    const int q = warp & 3, half = warp >> 2;  // TMEM lane quarter; the two warps of a quarter alternate over M tiles
    const int m = q * 32 + lane;               // row of the M tile = (sub-tile parity, channel)
    for (int j = half; j < nmt; j += 2) {
      const int s = s0 + 2 * j + (m >> 6);     // warp-uniform
      if (s >= nsub || (dbg & 4)) continue;
      const int k = s / c64n, ci = (s % c64n) * 64 + (m & 63);
      float* dst = dw + ((long long)k * cin + ci) * cout + n0;
      for (int cc = 0; cc < BN; cc += 32) {
        uint32_t v[32];
        tmem_ld32(tmem_base + ((uint32_t)(q * 32) << 16) + (uint32_t)(j * BN + cc), v);
#pragma unroll
        for (int e = 0; e < 8; ++e) {
          const float4 f = make_float4(__uint_as_float(v[4 * e]), __uint_as_float(v[4 * e + 1]),
                                       __uint_as_float(v[4 * e + 2]), __uint_as_float(v[4 * e + 3]));
          if (f.x != 0.f || f.y != 0.f || f.z != 0.f || f.w != 0.f)
            atomicAdd(reinterpret_cast<float4*>(dst + cc + 4 * e), f);
        }
      }
    }
    tc_fence_before();
  } else {
    // ===================== MMA issuer (one elected thread; the warp waits together) ==========
    {
      // bf16 x bf16 -> f32, A and B both MN-major (bits 15, 16), N = BN, M = 128
      constexpr uint32_t idesc = (1u << 4) | (1u << 7) | (1u << 10) | (1u << 15) | (1u << 16) |
                                 ((uint32_t)(BN >> 3) << 17) | ((128u >> 4) << 24);
      // descriptors differ from stage to stage only in the 14-bit start-address field: add (byte offset >> 4)
      const uint64_t da0 = umma_desc_mn_sw128(smem_base, WG_SUB_BYTES);
      const uint64_t db0 = umma_desc_mn_sw128(smem_base + Cfg::A_BYTES, WG_SUB_BYTES);
      const bool leader = elect_one();  // (tcgen05.mma under `lane == 0` compiles to a per-lane serialisation loop)
      int s = 0;
      uint32_t ph = 0;
      for (int it = 0; it < nstage; ++it) {
        mbar_wait(&full_bar[s], ph);
        fence_async_smem();  // cp.async wrote through the generic proxy, tcgen05.mma reads through the async proxy
        tc_fence_after();
        if (leader) {
          const uint64_t so = (uint64_t)((s * Cfg::STAGE_BYTES) >> 4);
#pragma unroll
          for (int kk = 0; kk < WG_KR / 16; ++kk) {
            if (kk == 1 && (dbg & 8)) break;
            const uint64_t db = db0 + so + (uint64_t)((kk * 2048) >> 4);
            for (int j = 0; j < ((dbg & 1) ? 0 : nmt); ++j) {
              const uint64_t da = da0 + so + (uint64_t)((j * 2 * WG_SUB_BYTES + kk * 2048) >> 4);
              umma_bf16(tmem_base + (uint32_t)(j * BN), da, db, idesc, (it > 0 || kk > 0) ? 1u : 0u);
            }
          }
          umma_commit(&empty_bar[s]);  // frees the smem stage once these MMAs have read it
        }
        __syncwarp();
        if (++s == S) {
          s = 0;
          ph ^= 1u;
        }
      }
      if (leader) umma_commit(accum_bar);
    }
    __syncwarp();
  }
  __syncthreads();
  if (warp == TC_PRODUCERS / 32) {
    tc_fence_after();
    asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem_base),
                 "r"((uint32_t)Cfg::TMEM_COLS)
                 : "memory");
  }
}

template <int BN, int SHARE>
static int launch_wgrad_tc(const void* in, const void* dout, const int* nbr, long long ld, long long n_out, int cin,
                           int cout, int kvol, float* dw, cudaStream_t st) {
  using Cfg = WgCfg<BN>;
  auto kern = k_wgrad_tc<BN, SHARE>;
  static unsigned long long configured = 0;   // one bit per device
  if (int rc0 = b2_optin_smem((const void*)kern, Cfg::smem_bytes(Cfg::STAGES), &configured)) return rc0;
  constexpr int stages = Cfg::STAGES;  // measured: 3, 4 and 5 stages give the same time (not latency bound)
  static const int dbg = getenv("B2_WG_DEBUG") ? atoi(getenv("B2_WG_DEBUG")) : 0;  // timing experiments only
  const int nsub = kvol * (cin / 64);
  const int gy = b2_div_up(nsub, WG_NSUB), gz = cout / BN;
  // row chunks: about two CTAs per SM in total (one is resident at a time), at least 16 stages each
  long long chunks = (2ll * 148) / ((long long)gy * gz);  // rounded down: no third wave
  const long long max_chunks = b2_div_up(n_out, 16ll * WG_KR);
  if (chunks > max_chunks) chunks = max_chunks;
  if (chunks < 1) chunks = 1;
  long long rows_per = b2_div_up(b2_div_up(n_out, chunks), (long long)WG_KR) * WG_KR;
  dim3 grid((unsigned)b2_div_up(n_out, rows_per), gy, gz);
  kern<<<grid, TC_THREADS, Cfg::smem_bytes(stages), st>>>((const __nv_bfloat16*)in, (const __nv_bfloat16*)dout, nbr, (int)ld,
                                            (int)n_out, cin, cout, kvol, (int)rows_per, dw, dbg);
  B2_LAUNCH_CHECK();
  return B2_OK;
}

// bf16 x bf16 weight gradient on tcgen05; eligibility (cin / 64 a power of two, cout % 64 == 0) is checked by the caller
int b2_wgrad_tc_dispatch(const void* in, const void* dout, const int* nbr, long long ld, long long n_out, int cin,
                         int cout, int kvol, float* dw, cudaStream_t st) {
  const int c64n = cin / 64;
#define WG_GO(BN)                                                                                      \
  do {                                                                                                 \
    if (c64n == 1) return launch_wgrad_tc<BN, 1>(in, dout, nbr, ld, n_out, cin, cout, kvol, dw, st);   \
    if (c64n == 2) return launch_wgrad_tc<BN, 2>(in, dout, nbr, ld, n_out, cin, cout, kvol, dw, st);   \
    if (c64n == 4) return launch_wgrad_tc<BN, 4>(in, dout, nbr, ld, n_out, cin, cout, kvol, dw, st);   \
    return launch_wgrad_tc<BN, 8>(in, dout, nbr, ld, n_out, cin, cout, kvol, dw, st);                  \
  } while (0)
  if (cout % 128 == 0) WG_GO(128);
  WG_GO(64);
#undef WG_GO
}
