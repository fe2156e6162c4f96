// PTX helpers shared by the tcgen05 sparse-convolution kernels (mbarrier, cp.async, tcgen05 MMA / TMEM access).
#pragma once
#include "common.cuh"
#include <cuda.h>

#define TC_BM 128
#define TC_BK 64
#define TC_PRODUCERS 256
#define TC_THREADS 288

__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_arrive(uint64_t* bar) {
  asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
  const uint32_t addr = smem_u32(bar);
  uint32_t ok;
  do {
    asm volatile(
        "{\n\t.reg .pred p;\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n\t"
        "selp.u32 %0, 1, 0, p;\n\t}"
        : "=r"(ok)
        : "r"(addr), "r"(parity)
        : "memory");
  } while (!ok);
}
__device__ __forceinline__ void cp_async16(uint32_t dst, const void* src, uint32_t src_bytes) {
  asm volatile("cp.async.cg.shared.global [%0], [%1], 16, %2;" ::"r"(dst), "l"(src), "r"(src_bytes) : "memory");
}
// the mbarrier receives one arrival from this thread once all of its earlier cp.async copies have landed
__device__ __forceinline__ void cp_async_arrive_noinc(uint64_t* bar) {
  asm volatile("cp.async.mbarrier.arrive.noinc.shared::cta.b64 [%0];" ::"r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void cp_async_commit() { asm volatile("cp.async.commit_group;" ::: "memory"); }
template <int N>
__device__ __forceinline__ void cp_async_wait() {
  asm volatile("cp.async.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ void fence_async_smem() { asm volatile("fence.proxy.async.shared::cta;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// true in exactly one lane of a converged warp (the compiler then knows the guarded region runs in one thread: tcgen05.mma /
// TMA instructions inside it are issued without a per-lane serialisation loop)
__device__ __forceinline__ bool elect_one() {
  uint32_t pred;
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "elect.sync _|p, 0xffffffff;\n\t"
      "selp.u32 %0, 1, 0, p;\n\t}"
      : "=r"(pred));
  return pred != 0;
}
__device__ __forceinline__ void umma_commit(uint64_t* bar) {
  asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];" ::"r"(smem_u32(bar))
               : "memory");
}
__device__ __forceinline__ void umma_bf16(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
__device__ __forceinline__ void umma_tf32(uint32_t tmem_d, uint64_t desc_a, uint64_t desc_b, uint32_t idesc,
                                          uint32_t accumulate) {
  asm volatile(
      "{\n\t.reg .pred p;\n\t"
      "setp.ne.b32 p, %4, 0;\n\t"
      "tcgen05.mma.cta_group::1.kind::tf32 [%0], %1, %2, %3, p;\n\t}"
      ::"r"(tmem_d), "l"(desc_a), "l"(desc_b), "r"(idesc), "r"(accumulate)
      : "memory");
}
// K-major, 128-byte swizzle, 8-row atoms 1024 B apart (SBO); descriptor version 1 (sm_100)
__device__ __forceinline__ uint64_t umma_desc_sw128(uint32_t saddr) {
  uint64_t d = 0;
  d |= (uint64_t)((saddr >> 4) & 0x3FFF);
  d |= (uint64_t)1 << 16;             // leading byte offset (unused for swizzled K-major)
  d |= (uint64_t)(1024 >> 4) << 32;   // stride byte offset
  d |= (uint64_t)1 << 46;             // version
  d |= (uint64_t)2 << 61;             // SWIZZLE_128B
  return d;
}
__device__ __forceinline__ void tmem_ld32(uint32_t taddr, uint32_t* r) {
  asm volatile(
      "tcgen05.ld.sync.aligned.32x32b.x32.b32 "
      "{%0,%1,%2,%3,%4,%5,%6,%7,%8,%9,%10,%11,%12,%13,%14,%15,"
      "%16,%17,%18,%19,%20,%21,%22,%23,%24,%25,%26,%27,%28,%29,%30,%31}, [%32];"
      : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]), "=r"(r[8]),
        "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15]), "=r"(r[16]),
        "=r"(r[17]), "=r"(r[18]), "=r"(r[19]), "=r"(r[20]), "=r"(r[21]), "=r"(r[22]), "=r"(r[23]), "=r"(r[24]),
        "=r"(r[25]), "=r"(r[26]), "=r"(r[27]), "=r"(r[28]), "=r"(r[29]), "=r"(r[30]), "=r"(r[31])
      : "r"(taddr)
      : "memory");
  asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory");
}


// outputs of the fused head epilogue (EPI == 1 of k_conv_tc / k_conv_tcp; see conv_tc.cu)
struct TcHeadEpi {
  float* centerness = nullptr;
  float* bbox_pred = nullptr;
  float* cls = nullptr;
  float* prune_score = nullptr;
  float* max_score = nullptr;
  const float* scale = nullptr;
  int n_reg = 0, n_cls = 0;
};

// Head epilogue of one row (forward_single, fcaf3d_neck_with_head.py:256-279, :237-238): f = the row's 32 head columns after
// bias / scale / shift.  Same expressions as k_head_post_tiled, hence bit-identical values.
__device__ __forceinline__ void tc_head_row(const float* f, long long o, const TcHeadEpi& hp) {
  const float hscale = hp.scale ? __ldg(hp.scale) : 1.f;
  const float ctr = f[0];
  hp.centerness[o] = ctr;
  const float sc = 1.f / (1.f + expf(-ctr));
  float mlogit = -INFINITY, mscore = -INFINITY;
  const int c0 = 1 + hp.n_reg, c1 = c0 + hp.n_cls;
#pragma unroll
  for (int j = 1; j < 32; ++j) {
    if (j < c0) {
      const int k = j - 1;
      hp.bbox_pred[o * hp.n_reg + k] = (k < 6) ? expf(hscale * f[j]) : f[j];
    } else if (j < c1) {
      mlogit = fmaxf(mlogit, f[j]);
      mscore = fmaxf(mscore, (1.f / (1.f + expf(-f[j]))) * sc);
      hp.cls[o * hp.n_cls + (j - c0)] = f[j];
    }
  }
  hp.prune_score[o] = mlogit;
  if (hp.max_score) hp.max_score[o] = mscore;
}

// ---- producer side of the gathers (shared by k_conv_tc and k_conv_tcp) ------------------------------------------
// 64-bit address of 16-byte unit `unit` behind `base` in one instruction (kept opaque so that the compiler holds the eight
// row addresses of an offset in registers instead of rebuilding them for every stage)
__device__ __forceinline__ const uint4* tc_unit_ptr(const uint4* base, int unit) {
  unsigned long long a;
  asm("mad.wide.u32 %0, %1, 16, %2;" : "=l"(a) : "r"((unsigned int)unit), "l"((unsigned long long)base));
  return reinterpret_cast<const uint4*>(a);
}

// One producer thread's part of a tile: for every used offset the group's stages (alternate stages of the global counter
// `it`), 8 rows x 16 bytes per stage.  CH = channel slices per offset (cin / 64) when known at compile time, 0 = `chunks`.
// Per row and offset: shared load, clamp, address; per row and stage: sign test + cp.async.
template <int S, int STAGE_BYTES, int CH>
__device__ __forceinline__ void tc_produce(unsigned int kmask, int chunks, int group, const int* nb0, const uint4* in_thr,
                                            uint32_t a_dst0, uint64_t* full_bar, uint64_t* empty_bar, int& it) {
  const int nch = CH > 0 ? CH : chunks;
  for (unsigned int m = kmask; m != 0u; m &= m - 1u) {
    const int k = __ffs((int)m) - 1;
    if (nch == 1 && (it & 1) != group) {  // the other group's offset
      ++it;
      continue;
    }
    const int* nb = nb0 + k * TC_BM;
    int r[TC_BM / 16];
    const uint4* src[TC_BM / 16];
#pragma unroll
    for (int p = 0; p < TC_BM / 16; ++p) r[p] = nb[p * 16];
#pragma unroll
    for (int p = 0; p < TC_BM / 16; ++p) src[p] = tc_unit_ptr(in_thr, max(r[p], 0));
#pragma unroll(CH > 0 ? CH : 1)
    for (int c = 0; c < nch; ++c, ++it) {
      if ((it & 1) != group) continue;
      const int s = it % S;
      mbar_wait(&empty_bar[s], ((uint32_t)(it / S) & 1u) ^ 1u);
      const uint32_t a_dst = a_dst0 + s * STAGE_BYTES;
#pragma unroll
      for (int p = 0; p < TC_BM / 16; ++p) cp_async16(a_dst + p * 2048, src[p] + c * 8, r[p] >= 0 ? 16u : 0u);
      cp_async_arrive_noinc(&full_bar[s]);
    }
  }
}

// ---- TMA (cp.async.bulk.tensor) helpers --------------------------------------------------------------------
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void tma_gather4(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int col, int r0, int r1,
                                            int r2, int r3) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile::gather4.mbarrier::complete_tx::bytes"
      " [%0], [%1, {%3, %4, %5, %6, %7}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(col), "r"(r0), "r"(r1), "r"(r2), "r"(r3)
      : "memory");
}
__device__ __forceinline__ void tma_tile_2d(uint32_t dst, const CUtensorMap* map, uint64_t* bar, int col, int row) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1, {%3, %4}], [%2];"
      ::"r"(dst), "l"(reinterpret_cast<uint64_t>(map)), "r"(smem_u32(bar)), "r"(col), "r"(row)
      : "memory");
}


// host (tc_map.cu): 2-D tensor maps [rows, cols], box [box_rows x 128 bytes], 128B swizzle, zero fill
int b2_make_bf16_map(CUtensorMap* map, const void* base, long long rows, int cols, int box_rows);
int b2_make_tf32_map(CUtensorMap* map, const void* base, long long rows, int cols, int box_rows);
// host (conv_tcp.cu): the persistent kernel for bf16 layers with many row tiles; *taken = 0 when the layer is not one of its
int b2_conv_tcp_dispatch(const void* in, const void* wt, const int* nbr, long long ld, const unsigned int* tile_mask,
                         const int* out_rows, long long n_out, int cin, int cout, int kvol, const float* bias,
                         const float* scale, const float* shift, const void* residual, int act, int ch_mod, void* out,
                         int tf32, cudaStream_t st, int* taken);
int b2_head_tcp_dispatch(const void* in, long long n, const void* wt, int cin, const float* bias, const TcHeadEpi& hp,
                         cudaStream_t st, int* taken);
// host: opt-in of one kernel to `bytes` of dynamic shared memory on the CURRENT device, once per device
int b2_optin_smem(const void* func, int bytes, unsigned long long* done);
