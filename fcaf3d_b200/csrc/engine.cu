// Native executor of the fused FCAF3D inference step (sm_100a).
//
// The Python host of this package enqueues one step as ~200 separate C-ABI calls; at ~25 us of interpreter time per call the
// host needs as long as the GPU (6.2 ms against 6.4 ms per batch of 8 ScanNet-shaped scenes), so a faster kernel shows up
// nowhere.  b2_engine_infer walks the SAME sequence of entry points of this library in C++ - quantise, coordinate maps one
// level ahead of the features, kernel maps, fused convolutions, neck with prune-before-convolution, head, batched top-k /
// decode / NMS tail - with the same arguments, so its results are bit-identical to detector.simple_test_fused (tests pin
// that), while the host side of a step shrinks to the launches themselves plus the row-count reads.
//
// What it executes is the reference's call chain SingleStageSparse3DDetector.simple_test -> extract_feat -> MEResNet3D.forward
// -> Fcaf3DNeckWithHead.forward / get_bboxes (single_stage_sparse.py:32-59, me_resnet.py:66-99,
// fcaf3d_neck_with_head.py:94-126,205-253,256-279,332-374) in eval mode with BatchNorm folded.
//
// Memory: nothing is allocated per call.  All intermediates are bump-allocated from the caller's arena (reset every call);
// the small pinned staging ring and the events of the row-count reads are created once in b2_engine_create.
#include "common.cuh"
#include "../../include/fcaf3d_b200.h"

#include <stdlib.h>
#include <thread>
#include <string.h>

#include <vector>

namespace {

struct ConvW {
  const void* wt;      // bf16 [kvol, cout, cin]
  const float* w32;    // fp32 [kvol, cin, cout]
  const float* scale;  // folded BatchNorm (NULL = none)
  const float* shift;
  const float* bias;
  int kvol, cin, cout;
};

constexpr int MAX_LEVELS = 4;
constexpr int RING = 1024;      // pinned staging slots / events
constexpr int SLOT_INTS = 128;  // ints per slot

struct Engine {
  int n_outs, n_cls, n_reg, yaw_mode, pts_threshold, nms_pre, dim, head_cols;
  int blocks[MAX_LEVELS];
  float voxel_size, score_thr, iou_thr, inorm_eps;
  ConvW stem;
  const float *inorm_w, *inorm_b;
  std::vector<ConvW> conv1[MAX_LEVELS], conv2[MAX_LEVELS];
  ConvW down[MAX_LEVELS];
  ConvW out_block[MAX_LEVELS], up_gen[MAX_LEVELS], up_conv[MAX_LEVELS];
  ConvW head;
  const float* level_scale[MAX_LEVELS];
  int* pinned = nullptr;  // RING * SLOT_INTS ints
  cudaEvent_t events[RING];
  int next_slot = 0;
  int device = -1;
  long long sort_min_rows = 32768;   // B2_SORT_MIN_ROWS (0 = never sort)
  cudaStream_t coord_stream = nullptr;   // coordinate work (hash builds, kernel maps) runs here, ahead of the features
  bool split = true;                     // B2_ENGINE_SPLIT=0: everything on the caller's stream
};

struct Map {
  bool on_F = false;   // coordinates produced on the feature stream (a pruned map): whatever consumes it runs there too
  int* coords = nullptr;
  long long n = 0;
  int ts = 1;
  void* table = nullptr;
  long long cap = 0;
  int* off_dev = nullptr;
  std::vector<int> off;  // nb + 1 host copies of the batch offsets
  int children_of = -1;
};

struct KMap {
  int in, out, ks;
  int* nbr;
  long long ld;
  unsigned int* tmask;
  long long n_out;
  int kvol;
  // sparsity-grouped copy for the tcgen05 convolutions (b2_kmap_sort), NULL = walk the table as built
  int* perm = nullptr;
  int* nbr_sorted = nullptr;
  unsigned int* tmask_sorted = nullptr;
};

struct Pending {  // a map build / union whose row count is still in flight
  bool valid = false;
  bool on_F = false;
  int slot = 0;
  int* status = nullptr;
  int* out_coords = nullptr;
  int* uidx = nullptr;
  int* inverse = nullptr;
  int* dst_a = nullptr;
  int* dst_b = nullptr;
  void* table = nullptr;
  long long cap = 0;
  int ts = 1;
};

static long long next_pow2(long long v) {
  long long p = 1;
  while (p < v) p <<= 1;
  return p;
}
static long long table_cap(long long n) { return next_pow2(2 * n) < 1024 ? 1024 : next_pow2(2 * n); }

__global__ void k_coords_to_float(const int* __restrict__ c, long long n4, float* __restrict__ out) {
  const long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x;
  if (i < n4) out[i] = (float)c[i];
}

struct Run {
  Engine& e;
  char* base;
  size_t cap, off = 0;
  cudaStream_t st;   // the stream ops are enqueued on right now (stF or stC)
  void* stv;
  cudaStream_t stF, stC;
  bool split = false;    // coordinate work on stC, features on stF, F waits for C by events
  bool c_dirty = false;  // C work enqueued since F last waited for C
  int nb = 0;
  int dt = 1;  // activation dtype: 1 bf16, 0 fp32
  bool tf32 = false;   // fp32 rows, wide convolutions as TF32 tensor-core operands (ConvW::wt then holds fp32 [kvol, cout, cin])
  std::vector<Map> maps;
  std::vector<KMap> kmaps;

  Run(Engine& eng, void* arena, size_t bytes, void* stream)
      : e(eng), base((char*)arena), cap(bytes), st((cudaStream_t)stream), stv(stream), stF((cudaStream_t)stream),
        stC(eng.coord_stream) {
    split = eng.split && eng.coord_stream != nullptr;
    maps.reserve(256);
    kmaps.reserve(256);   // pointers into these vectors are handed out: they must never reallocate (checked below)
  }

  void* take(size_t bytes) {
    const size_t a = (off + 255) & ~(size_t)255;
    if (a + bytes > cap) {
      b2_set_error("b2_engine_infer: arena too small (%zu bytes needed so far, %zu given)", a + bytes, cap);
      return nullptr;
    }
    off = a + bytes;
    return base + a;
  }
  size_t esize() const { return dt == 1 ? 2 : 4; }

  // Two streams: everything that depends on COORDINATES only (hash builds, kernel maps, table grouping, children / union
  // maps) is enqueued on stC, the feature kernels on stF; before a feature op F waits (by event) for the coordinate work
  // enqueued so far.  Nothing on C ever waits for F inside a call, and the host blocks on C events only, so while the GPU
  // still runs the convolutions of this call on F the host is already enqueueing - and the GPU executing - the small,
  // latency-bound coordinate kernels of the NEXT call on C.
  void to_F() { st = stF, stv = (void*)stF; }
  int fsync() {
    if (!split || !c_dirty) return B2_OK;
    const int s = slot();
    B2_CUDA(cudaEventRecord(e.events[s], stC));
    B2_CUDA(cudaStreamWaitEvent(stF, e.events[s], 0));
    c_dirty = false;
    return B2_OK;
  }
  struct CoordScope {   // ops of this scope go to C (or to F when they touch a map that was made on F)
    Run& r;
    int rc = B2_OK;
    cudaStream_t old_st;
    void* old_stv;
    CoordScope(Run& r_, bool on_f) : r(r_), old_st(r_.st), old_stv(r_.stv) {
      if (r.split && r.st == r.stC && !on_f) return;   // nested inside a coordinate scope: stay on C
      if (on_f || !r.split) {
        rc = r.fsync();
        r.to_F();
      } else {
        r.st = r.stC, r.stv = (void*)r.stC;
        r.c_dirty = true;
      }
    }
    ~CoordScope() { r.st = old_st, r.stv = old_stv; }
  };

  int slot() {
    const int s = e.next_slot;
    e.next_slot = (e.next_slot + 1) % RING;
    return s;
  }
  // host ints -> device (pinned staging slot, asynchronous copy)
  int upload(const int* vals, int n, int** dev) {
    B2_CHECK_ARG(n <= SLOT_INTS, "engine: upload of %d ints exceeds a staging slot", n);
    *dev = (int*)take((size_t)(n > 0 ? n : 1) * 4);
    if (!*dev) return B2_ERR_WS;
    if (n == 0) return B2_OK;
    const int s = slot();
    int* h = e.pinned + (size_t)s * SLOT_INTS;
    memcpy(h, vals, (size_t)n * 4);
    B2_CUDA(cudaMemcpyAsync(*dev, h, (size_t)n * 4, cudaMemcpyHostToDevice, st));
    return B2_OK;
  }
  // device ints -> pinned slot, event recorded; read_wait returns the host pointer once the copy has landed
  int read_start(const int* dev, int n, int* slot_out) {
    B2_CHECK_ARG(n <= SLOT_INTS, "engine: read of %d ints exceeds a staging slot", n);
    const int s = slot();
    B2_CUDA(cudaMemcpyAsync(e.pinned + (size_t)s * SLOT_INTS, dev, (size_t)n * 4, cudaMemcpyDeviceToHost, st));
    B2_CUDA(cudaEventRecord(e.events[s], st));
    *slot_out = s;
    return B2_OK;
  }
  int read_wait(int s, const int** host) {
    B2_CUDA(cudaEventSynchronize(e.events[s]));
    *host = e.pinned + (size_t)s * SLOT_INTS;
    return B2_OK;
  }

  // ---- coordinate maps -------------------------------------------------------------------------------------------
  // want_uidx: the input map (features are picked at the representatives' rows); nobody reads an inverse map in this pass,
  // so every build is the two-kernel variant of b2_map_build
  int build_start(const int* coords, long long n, int quant, int ts_out, Pending* p, bool want_uidx = false) {
    CoordScope sc(*this, false);
    if (sc.rc) return sc.rc;
    p->cap = table_cap(n);
    p->ts = ts_out;
    p->table = take((size_t)p->cap * 16);
    p->out_coords = (int*)take((size_t)(n > 0 ? n : 1) * 16);
    p->uidx = want_uidx ? (int*)take((size_t)(n > 0 ? n : 1) * 4) : nullptr;
    p->inverse = nullptr;
    p->status = (int*)take((size_t)(3 + nb) * 4);
    const size_t wsb = b2_map_build_ws_bytes(n);
    void* ws = take(wsb);
    if (!p->table || !p->out_coords || (want_uidx && !p->uidx) || !p->status || !ws) return B2_ERR_WS;
    B2_CUDA(cudaMemsetAsync(p->status, 0, (size_t)(3 + nb) * 4, st));
    int rc = b2_map_build(coords, n, quant, p->table, p->cap, p->out_coords, p->uidx, p->inverse, p->status, p->status + 1,
                          ws, wsb, stv);
    if (rc) return rc;
    rc = b2_batch_offsets_dev(p->out_coords, p->status, n, nb, p->status + 2, p->status + 1, stv);
    if (rc) return rc;
    rc = read_start(p->status, 3 + nb, &p->slot);
    if (rc) return rc;
    p->valid = true;
    return B2_OK;
  }
  int build_finish(Pending* p, int* map_index) {
    B2_CHECK_ARG(p->valid, "engine: map build was never started");
    const int* h;
    int rc = read_wait(p->slot, &h);
    if (rc) return rc;
    B2_CHECK_ARG(h[1] != 1, "coordinate outside the packed key range (batch < 1023, |xyz| < 131072)");
    B2_CHECK_ARG(h[1] != 2, "coordinate rows must be grouped by ascending batch index");
    Map m;
    m.coords = p->out_coords;
    m.n = h[0];
    m.ts = p->ts;
    m.table = p->table;
    m.cap = p->cap;
    m.off_dev = p->status + 2;
    m.off.assign(h + 2, h + 3 + nb);
    maps.push_back(m);
    *map_index = (int)maps.size() - 1;
    p->valid = false;
    return B2_OK;
  }
  int ensure_table(int mi) {
    Map& m = maps[mi];
    if (m.table) return B2_OK;
    CoordScope sc(*this, m.on_F);
    if (sc.rc) return sc.rc;
    m.cap = table_cap(m.n);
    m.table = take((size_t)m.cap * 16);
    int* err = (int*)take(4);
    if (!m.table || !err) return B2_ERR_WS;
    B2_CUDA(cudaMemsetAsync(err, 0, 4, st));
    return b2_map_table(m.coords, m.n, m.table, m.cap, err, stv);
  }
  int children(int mi, int* out_index) {
    const Map src = maps[mi];
    B2_CHECK_ARG(src.ts % 2 == 0, "generative transpose needs an even tensor stride");
    CoordScope sc(*this, src.on_F);
    if (sc.rc) return sc.rc;
    Map m;
    m.on_F = src.on_F;
    m.n = src.n * 8;
    m.ts = src.ts / 2;
    m.coords = (int*)take((size_t)(m.n > 0 ? m.n : 1) * 16);
    if (!m.coords) return B2_ERR_WS;
    int rc = b2_map_children(src.coords, src.n, m.ts, m.coords, stv);
    if (rc) return rc;
    m.off.resize(nb + 1);
    for (int b = 0; b <= nb; ++b) m.off[b] = 8 * src.off[b];
    rc = upload(m.off.data(), nb + 1, &m.off_dev);
    if (rc) return rc;
    m.children_of = mi;
    maps.push_back(m);
    *out_index = (int)maps.size() - 1;
    return B2_OK;
  }
  int union_start(int ai, int bi, Pending* p) {
    int rc = ensure_table(ai);
    if (rc) return rc;
    const Map &a = maps[ai], &b = maps[bi];
    B2_CHECK_ARG(a.ts == b.ts, "union of maps with different tensor strides");
    p->on_F = a.on_F || b.on_F;
    CoordScope sc(*this, p->on_F);
    if (sc.rc) return sc.rc;
    p->ts = a.ts;
    p->out_coords = (int*)take((size_t)(a.n + b.n > 0 ? a.n + b.n : 1) * 16);
    p->dst_a = (int*)take((size_t)(a.n > 0 ? a.n : 1) * 4);
    p->dst_b = (int*)take((size_t)(b.n > 0 ? b.n : 1) * 4);
    p->status = (int*)take((size_t)(3 + nb) * 4);
    const size_t wsb = b2_map_union_ws_bytes(b.n);
    void* ws = take(wsb);
    if (!p->out_coords || !p->dst_a || !p->dst_b || !p->status || !ws) return B2_ERR_WS;
    B2_CUDA(cudaMemsetAsync(p->status, 0, (size_t)(3 + nb) * 4, st));
    rc = b2_map_union(a.table, a.cap, a.coords, a.n, a.off_dev, b.coords, b.n, b.off_dev, p->out_coords, p->dst_a, p->dst_b,
                      p->status, ws, wsb, stv);
    if (rc) return rc;
    rc = b2_batch_offsets_dev(p->out_coords, p->status, a.n + b.n, nb, p->status + 2, p->status + 1, stv);
    if (rc) return rc;
    rc = read_start(p->status, 3 + nb, &p->slot);
    if (rc) return rc;
    p->valid = true;
    return B2_OK;
  }
  int union_finish(Pending* p, int* map_index) {
    const int* h;
    int rc = read_wait(p->slot, &h);
    if (rc) return rc;
    Map m;
    m.on_F = p->on_F;
    m.coords = p->out_coords;
    m.n = h[0];
    m.ts = p->ts;
    m.off_dev = p->status + 2;
    m.off.assign(h + 2, h + 3 + nb);
    maps.push_back(m);
    *map_index = (int)maps.size() - 1;
    p->valid = false;
    return B2_OK;
  }

  // ---- kernel maps -----------------------------------------------------------------------------------------------
  KMap* find_kmap(int in, int out, int ks) {
    for (auto& k : kmaps)
      if (k.in == in && k.out == out && k.ks == ks) return &k;
    return nullptr;
  }
  int alloc_kmap(int in, int out, int ks, long long n_out, KMap* k) {
    k->in = in, k->out = out, k->ks = ks, k->kvol = ks * ks * ks, k->n_out = n_out;
    k->ld = (n_out + 127) / 128 * 128;
    k->nbr = (int*)take((size_t)(k->kvol * k->ld > 0 ? k->kvol * k->ld : 1) * 4);
    k->tmask = (unsigned int*)take((size_t)(k->ld / 128 > 0 ? k->ld / 128 : 1) * 4);
    return (k->nbr && k->tmask) ? B2_OK : B2_ERR_WS;
  }
  int kernel_map(int in, int out, int ks, KMap** res) {
    if (KMap* k = find_kmap(in, out, ks)) {
      *res = k;
      return B2_OK;
    }
    KMap k;
    int rc = alloc_kmap(in, out, ks, maps[out].n, &k);
    if (rc) return rc;
    CoordScope sc(*this, maps[in].on_F || maps[out].on_F);
    if (sc.rc) return sc.rc;
    if (in == out && maps[in].children_of >= 0 && ks == 3) {
      const int parent = maps[in].children_of;
      KMap* pk;
      rc = kernel_map(parent, parent, 3, &pk);
      if (rc) return rc;
      rc = b2_kernel_map_children(pk->nbr, pk->ld, maps[parent].n, k.nbr, k.ld, k.tmask, stv);
    } else if (in != out && maps[out].ts == 2 * maps[in].ts && ((ks == 3 && 2 * maps[in].n < 3 * maps[out].n) || ks == 2)) {
      rc = ensure_table(out);
      if (rc) return rc;
      rc = b2_kernel_map_strided(maps[out].table, maps[out].cap, maps[in].coords, maps[in].n, ks, maps[in].ts, k.nbr, k.ld,
                                 k.tmask, stv);
    } else {
      rc = ensure_table(in);
      if (rc) return rc;
      rc = b2_kernel_map(maps[in].table, maps[in].cap, maps[out].coords, maps[out].n, ks, maps[in].ts, k.nbr, k.ld, k.tmask,
                         stv);
    }
    if (rc) return rc;
    B2_CHECK_ARG(kmaps.size() < kmaps.capacity(), "engine: too many kernel maps in one step");
    kmaps.push_back(k);
    *res = &kmaps.back();
    return B2_OK;
  }

  // rows grouped by the kernel slabs they use (kmap_sort.cu): pays for tables that several tcgen05 convolutions walk
  int sort_kmap(KMap* k) {
    if (k->perm || k->kvol != 27 || k->ld == 0) return B2_OK;
    CoordScope sc(*this, maps[k->in].on_F || maps[k->out].on_F);
    if (sc.rc) return sc.rc;
    k->perm = (int*)take((size_t)k->ld * 4);
    k->nbr_sorted = (int*)take((size_t)k->kvol * k->ld * 4);
    k->tmask_sorted = (unsigned int*)take((size_t)(k->ld / 128) * 4);
    const size_t wsb = b2_kmap_sort_ws_bytes(k->n_out, k->kvol);
    void* ws = take(wsb);
    if (!k->perm || !k->nbr_sorted || !k->tmask_sorted || !ws) return B2_ERR_WS;
    return b2_kmap_sort(k->nbr, k->ld, k->n_out, k->kvol, k->perm, k->nbr_sorted, k->tmask_sorted, ws, wsb, stv);
  }

  // ---- convolution (the dispatch of ops.sparse_conv) ---------------------------------------------------------------
  int conv(const ConvW& w, const void* x, int x_dt, long long n_in, const KMap* km, long long n_out, const void* residual,
           int act, int ch_mod, int out_dt, void** out) {
    *out = take((size_t)(n_out > 0 ? n_out : 1) * w.cout * (out_dt == 1 ? 2 : 4));
    if (!*out) return B2_ERR_WS;
    if (int rcs = fsync()) return rcs;
    const int* nbr = km ? km->nbr : nullptr;
    const long long ld = km ? km->ld : 0;
    const unsigned int* tm = km ? km->tmask : nullptr;
    if (tf32 && x_dt == 0 && out_dt == 0 && w.wt && w.cin % 32 == 0 && w.cout % 32 == 0) {
      const bool sorted = km && km->perm && w.kvol == 27;
      return b2_conv_tc_tf32(x, n_in, w.wt, sorted ? km->nbr_sorted : nbr, ld, sorted ? km->tmask_sorted : tm,
                             sorted ? km->perm : nullptr, n_out, w.cin, w.cout, w.kvol, w.bias, w.scale, w.shift, residual,
                             act, ch_mod, *out, stv);
    }
    if (x_dt == 1 && w.wt && w.cin % 64 == 0 && w.cout % 32 == 0) {
      if (km && km->perm && w.kvol == 27)
        return b2_conv_tc(x, n_in, w.wt, km->nbr_sorted, ld, km->tmask_sorted, km->perm, n_out, w.cin, w.cout, w.kvol,
                          w.bias, w.scale, w.shift, residual, act, ch_mod, *out, out_dt, stv);
      return b2_conv_tc(x, n_in, w.wt, nbr, ld, tm, nullptr, n_out, w.cin, w.cout, w.kvol, w.bias, w.scale, w.shift,
                        residual, act, ch_mod, *out, out_dt, stv);
    }
    return b2_conv_cc(x, x_dt, w.w32, nbr, ld, tm, n_out, w.cin, w.cout, w.kvol, w.bias, w.scale, w.shift, residual, act,
                      ch_mod, *out, out_dt, stv);
  }
};

struct Level {
  int map = -1;
  float *centerness = nullptr, *bbox = nullptr, *cls = nullptr, *prune = nullptr, *max_score = nullptr;
};

#define CK(call)            \
  do {                      \
    int rc_ = (call);       \
    if (rc_) return rc_;    \
  } while (0)
#define FK(call)            \
  do {                      \
    CK(r.fsync());          \
    CK(call);               \
  } while (0)
#define TAKE(var, type, count)                                      \
  type* var = (type*)r.take((size_t)((count) > 0 ? (count) : 1) * sizeof(type)); \
  if (!var) return B2_ERR_WS

static int infer(Engine& e, Run& r, const float* const* points, const long long* n_points, int n_scenes, int dtype,
                 float* out_boxes, float* out_scores, int* out_labels, int* out_scene, int* out_count, long long out_cap,
                 long long* n_rows_out, void* tail_stream, void* inputs_ready_event, int inputs_ready,
                 void* arena_free_event) {
  if (r.split) {
    // C starts as soon as the inputs are there and the arena is free - NOT behind the feature work still queued on F
    if (inputs_ready_event != nullptr) {
      B2_CUDA(cudaStreamWaitEvent(r.stC, (cudaEvent_t)inputs_ready_event, 0));
    } else if (!inputs_ready) {   // unknown producer: the inputs are ordered on F only, so C has to wait for F's queue
      const int s = r.slot();
      B2_CUDA(cudaEventRecord(e.events[s], r.stF));
      B2_CUDA(cudaStreamWaitEvent(r.stC, e.events[s], 0));
    }
    if (arena_free_event != nullptr) B2_CUDA(cudaStreamWaitEvent(r.stC, (cudaEvent_t)arena_free_event, 0));
  }
  r.nb = n_scenes;
  r.dt = dtype == 1 ? 1 : 0;
  r.tf32 = dtype == 2;
  const int nb = n_scenes, L = e.n_outs, dim = e.dim, fdim = e.dim - 3;
  B2_CHECK_ARG(nb >= 1 && nb + 3 <= SLOT_INTS && nb * (L + 1) + 1 <= SLOT_INTS, "engine: too many scenes per call (%d)", nb);
  long long total = 0;
  for (int b = 0; b < nb; ++b) total += n_points[b];
  B2_CHECK_ARG(total < (1ll << 31), "engine: too many points");

  // ---- input: quantise, dedup, the first strided map from the raw rows (both row counts come back behind one wait)
  TAKE(coords, int, total * 4);
  TAKE(feats, float, total * fdim);
  {
    Run::CoordScope sc(r, false);
    if (sc.rc) return sc.rc;
    long long row0 = 0;
    for (int b = 0; b < nb; ++b) {
      CK(b2_quantize_points(points[b], n_points[b], dim, b, e.voxel_size, 255.f, row0, coords, feats, r.stv));
      row0 += n_points[b];
    }
  }
  Pending p0, p1;
  CK(r.build_start(coords, total, 0, 1, &p0, true));
  CK(r.build_start(coords, total, 2, 2, &p1));
  int* uidx = p0.uidx;
  int m1;
  CK(r.build_finish(&p0, &m1));
  TAKE(x0, float, r.maps[m1].n * fdim);
  FK(b2_index_rows(feats, uidx, r.maps[m1].n, fdim * 4, x0, r.stv));

  // ---- stem: conv 3^3 s2 -> InstanceNorm + ReLU -> max-pool 2^3 s2 (me_resnet.py:18-25)
  int m2;
  CK(r.build_finish(&p1, &m2));
  KMap* km;
  CK(r.kernel_map(m1, m2, 3, &km));
  Pending pp;
  CK(r.build_start(r.maps[m2].coords, r.maps[m2].n, 4, 4, &pp));
  void* f0;
  CK(r.conv(e.stem, x0, 0, r.maps[m1].n, km, r.maps[m2].n, nullptr, B2_ACT_NONE, 0, r.dt, &f0));
  const int c0 = e.stem.cout;
  TAKE(in_scale, float, nb * c0);
  TAKE(in_shift, float, nb * c0);
  {
    const size_t wsb = b2_instnorm_ws_bytes(nb, c0);
    void* ws = r.take(wsb);
    if (!ws) return B2_ERR_WS;
    FK(b2_instnorm_scale_shift(f0, r.dt, r.maps[m2].off_dev, nb, c0, e.inorm_w, e.inorm_b, e.inorm_eps, in_scale, in_shift,
                               ws, wsb, r.stv));
  }
  int m4;
  CK(r.build_finish(&pp, &m4));
  Pending pn;
  CK(r.build_start(r.maps[m4].coords, r.maps[m4].n, r.maps[m4].ts * 2, r.maps[m4].ts * 2, &pn));
  CK(r.kernel_map(m2, m4, 2, &km));
  void* f = r.take((size_t)(r.maps[m4].n > 0 ? r.maps[m4].n : 1) * c0 * r.esize());
  if (!f) return B2_ERR_WS;
  if (r.dt == 1) {
    FK(b2_maxpool_affine(f0, km->nbr, km->ld, km->kvol, km->n_out, c0, r.maps[m4].coords, in_scale, in_shift, B2_ACT_RELU,
                         f, r.stv));
  } else {
    FK(b2_affine_act(f0, r.dt, r.maps[m2].coords, r.maps[m2].n, c0, in_scale, in_shift, nullptr, B2_ACT_RELU, f0, r.stv));
    FK(b2_maxpool(f0, r.dt, km->nbr, km->ld, km->kvol, km->n_out, c0, f, nullptr, r.stv));
  }

  // ---- residual stages (me_resnet.py:27-41,66-78): conv + folded BatchNorm + ReLU (+ residual) = one launch each
  int key = m4;
  void* stage_f[MAX_LEVELS];
  int stage_map[MAX_LEVELS];
  for (int i = 0; i < L; ++i) {
    for (int j = 0; j < e.blocks[i]; ++j) {
      int out = key;
      if (j == 0) CK(r.build_finish(&pn, &out));
      const long long n_in = r.maps[key].n, n_out = r.maps[out].n;
      KMap* km1;
      CK(r.kernel_map(key, out, 3, &km1));
      void* h;
      CK(r.conv(e.conv1[i][j], f, r.dt, n_in, km1, n_out, nullptr, B2_ACT_RELU, 0, r.dt, &h));
      void* res = f;
      if (j == 0) {
        KMap* kmd;
        CK(r.kernel_map(key, out, 1, &kmd));
        CK(r.conv(e.down[i], f, r.dt, n_in, kmd, n_out, nullptr, B2_ACT_NONE, 0, r.dt, &res));
      }
      KMap* km2;
      CK(r.kernel_map(out, out, 3, &km2));
      // the stage's stride-1 table is walked by 2 * blocks - 1 convolutions: group its rows once (bit-identical results)
      if ((r.dt == 1 || r.tf32) && e.sort_min_rows > 0 && e.blocks[i] >= 2 && n_out >= e.sort_min_rows) CK(r.sort_kmap(km2));
      void* o;
      CK(r.conv(e.conv2[i][j], h, r.dt, n_out, km2, n_out, res, B2_ACT_RELU, 0, r.dt, &o));
      f = o;
      key = out;
      if (j == 0 && i + 1 < L)
        CK(r.build_start(r.maps[key].coords, r.maps[key].n, r.maps[key].ts * 2, r.maps[key].ts * 2, &pn));
    }
    stage_f[i] = f;
    stage_map[i] = key;
  }

  // ---- neck + heads, coarse -> fine (fcaf3d_neck_with_head.py:94-126,256-279)
  Level lv[MAX_LEVELS];
  Pending pu;
  int pre_gen = -1;
  f = stage_f[L - 1];
  key = stage_map[L - 1];
  for (int i = L - 1; i >= 0; --i) {
    if (i < L - 1) {
      const int gkey = pre_gen, parent = r.maps[gkey].children_of;
      const long long n_par = r.maps[parent].n, n_gen = r.maps[gkey].n;
      const int cg = e.up_gen[i + 1].cout / 8;  // channels of the finer level
      // The prune chain of this level (interpolation, top-k, scan, row maps, tables of the pruned rows: ~0.25 ms of small
      // launches) needs the previous level's scores only, the generative transpose needs its features only: with a tail
      // stream the chain runs THERE, next to the transpose on the feature stream, and the feature stream waits for it
      // right before the up-block convolution.
      const bool can_aux = tail_stream != nullptr && r.split && tail_stream != (void*)r.stF;
      int aux_ev = -1;
      if (can_aux) {
        aux_ev = r.slot();
        B2_CUDA(cudaEventRecord(e.events[aux_ev], r.stF));
      }
      bool in_aux = false;
      cudaStream_t f_orig = r.stF;
      auto aux_begin = [&]() -> int {
        B2_CUDA(cudaStreamWaitEvent((cudaStream_t)tail_stream, e.events[aux_ev], 0));
        r.stF = (cudaStream_t)tail_stream;
        r.to_F();
        r.c_dirty = true;      // the aux stream has not seen any coordinate work yet
        in_aux = true;
        return r.fsync();
      };
      auto aux_end = [&]() -> int {
        if (!in_aux) return B2_OK;
        int rc_ = r.fsync();
        if (rc_) return rc_;
        const int s = r.slot();
        B2_CUDA(cudaEventRecord(e.events[s], r.stF));
        r.stF = f_orig;
        r.to_F();
        B2_CUDA(cudaStreamWaitEvent(f_orig, e.events[s], 0));
        in_aux = false;
        return B2_OK;
      };
      void* g;
      CK(r.conv(e.up_gen[i + 1], f, r.dt, n_par, nullptr, n_par, nullptr, B2_ACT_ELU, cg, r.dt, &g));  // [n_par, 8 cg] = [n_gen, cg]
      int ukey;
      int *dst_a = pu.dst_a, *dst_b = pu.dst_b;
      CK(r.union_finish(&pu, &ukey));
      const long long nu = r.maps[ukey].n;
      int max_rows = 0;
      for (int b = 0; b < nb; ++b) max_rows = max(max_rows, r.maps[ukey].off[b + 1] - r.maps[ukey].off[b]);
      const bool need_prune = e.pts_threshold >= 0 && max_rows > e.pts_threshold;
      uint8_t* mask = nullptr;
      int* pos = nullptr;
      long long n_keep = nu;
      int pkey = ukey;
      if (need_prune) {
        // prune by the coarser level's interpolated max-class score: ready before the up-block's 3^3 convolution runs
        if (can_aux) CK(aux_begin());
        const Level& prev = lv[i + 1];
        CK(r.ensure_table(prev.map));
        TAKE(q, float, nu * 4);
        CK(r.fsync());
        k_coords_to_float<<<b2_div_up(nu * 4, 256), 256, 0, r.st>>>(r.maps[ukey].coords, nu * 4, q);
        B2_LAUNCH_CHECK();
        TAKE(sc, float, nu);
        FK(b2_interp(r.maps[prev.map].table, r.maps[prev.map].cap, prev.prune, 1, r.maps[prev.map].ts, q, nu, sc, r.stv));
        mask = (uint8_t*)r.take((size_t)nu);
        pos = (int*)r.take((size_t)nu * 4);
        int* tot = (int*)r.take(4);
        if (!mask || !pos || !tot) return B2_ERR_WS;
        FK(b2_topk_mask_n(sc, r.maps[ukey].off_dev, nb, e.pts_threshold, nu, mask, r.stv));
        B2_CUDA(cudaMemsetAsync(tot, 0, 4, r.st));
        const size_t wsb = b2_scan_ws_bytes(nu);
        void* ws = r.take(wsb);
        if (!ws) return B2_ERR_WS;
        FK(b2_scan_mask(mask, nu, pos, tot, ws, wsb, r.stv));
        Map pm;  // top-k keeps exactly min(N_b, k) rows of every scene: the pruned map's offsets are known on the host
        pm.on_F = true;
        pm.ts = r.maps[ukey].ts;
        pm.off.resize(nb + 1);
        pm.off[0] = 0;
        for (int b = 0; b < nb; ++b)
          pm.off[b + 1] = pm.off[b] + min(r.maps[ukey].off[b + 1] - r.maps[ukey].off[b], e.pts_threshold);
        pm.n = n_keep = pm.off[nb];
        CK(r.upload(pm.off.data(), nb + 1, &pm.off_dev));
        pm.coords = (int*)r.take((size_t)(n_keep > 0 ? n_keep : 1) * 16);
        if (!pm.coords) return B2_ERR_WS;
        FK(b2_gather_coords(r.maps[ukey].coords, mask, pos, nu, pm.coords, r.stv));
        r.maps.push_back(pm);
        pkey = (int)r.maps.size() - 1;
      }
      if (need_prune && nu == n_gen) {
        // behind the convolution there is per-row work only (folded BatchNorm, ELU, + backbone row): evaluate it at the
        // surviving rows; its table and the pruned map's own 3^3 table come from the parent's table by index arithmetic
        TAKE(g2k, int, n_gen);
        TAKE(a2k, int, r.maps[stage_map[i]].n);
        TAKE(ksrc, int, n_keep);
        FK(b2_prune_plan(dst_b, n_gen, dst_a, r.maps[stage_map[i]].n, mask, pos, g2k, a2k, ksrc, r.stv));
        KMap* pk;
        CK(r.kernel_map(parent, parent, 3, &pk));
        KMap up, ko;
        CK(r.alloc_kmap(gkey, pkey, 3, n_keep, &up));
        CK(r.alloc_kmap(pkey, pkey, 3, n_keep, &ko));
        FK(b2_kernel_map_children_pruned(pk->nbr, pk->ld, ksrc, g2k, n_keep, up.ld, up.nbr, up.tmask, ko.nbr, ko.tmask, r.stv));
        CK(aux_end());
        r.kmaps.push_back(ko);
        void* o;
        CK(r.conv(e.up_conv[i + 1], g, r.dt, n_gen, &up, n_keep, nullptr, B2_ACT_ELU, 0, r.dt, &o));
        FK(b2_scatter_rows(stage_f[i], a2k, r.maps[stage_map[i]].n, cg, r.dt, 1, o, r.stv));
        f = o;
      } else {
        CK(aux_end());
        KMap* kg;
        CK(r.kernel_map(gkey, gkey, 3, &kg));
        void* o;
        CK(r.conv(e.up_conv[i + 1], g, r.dt, n_gen, kg, n_gen, nullptr, B2_ACT_ELU, 0, r.dt, &o));
        void* u = r.take((size_t)(nu > 0 ? nu : 1) * cg * r.esize());
        if (!u) return B2_ERR_WS;
        B2_CUDA(cudaMemsetAsync(u, 0, (size_t)nu * cg * r.esize(), r.st));
        FK(b2_scatter_rows(stage_f[i], dst_a, r.maps[stage_map[i]].n, cg, r.dt, 0, u, r.stv));
        FK(b2_scatter_rows(o, dst_b, n_gen, cg, r.dt, 1, u, r.stv));
        f = u;
        if (need_prune) {
          void* pr = r.take((size_t)(n_keep > 0 ? n_keep : 1) * cg * r.esize());
          if (!pr) return B2_ERR_WS;
          FK(b2_gather_rows(u, mask, pos, nu, (int)(cg * r.esize()), pr, r.stv));
          f = pr;
        }
      }
      key = pkey;
    }
    if (i > 0) {
      // the next (finer) level's generated map and its union with the backbone map depend on coordinates only: enqueue
      // them before this level's out_block / heads so that the union's row count is not waited for with an idle GPU
      CK(r.children(key, &pre_gen));
      CK(r.union_start(stage_map[i - 1], pre_gen, &pu));
    }
    const long long n = r.maps[key].n;
    KMap* ko;
    CK(r.kernel_map(key, key, 3, &ko));
    void* o;
    CK(r.conv(e.out_block[i], f, r.dt, n, ko, n, nullptr, B2_ACT_ELU, 0, r.dt, &o));
    Level& l = lv[i];
    l.map = key;
    l.centerness = (float*)r.take((size_t)(n > 0 ? n : 1) * 4);
    l.bbox = (float*)r.take((size_t)(n > 0 ? n : 1) * e.n_reg * 4);
    l.cls = (float*)r.take((size_t)(n > 0 ? n : 1) * e.n_cls * 4);
    l.prune = (float*)r.take((size_t)(n > 0 ? n : 1) * 4);
    l.max_score = (float*)r.take((size_t)(n > 0 ? n : 1) * 4);
    if (!l.centerness || !l.bbox || !l.cls || !l.prune || !l.max_score) return B2_ERR_WS;
    if (r.dt == 1 && e.head.wt && e.head_cols == 32 && e.head.cin % 64 == 0 && 1 + e.n_reg + e.n_cls <= 32) {
      // head GEMM with the head post-processing as its epilogue: the [n, 32] logits never go to memory
      FK(b2_head_fused(o, n, e.head.wt, e.head.cin, e.head_cols, e.head.bias, e.n_reg, e.n_cls, e.level_scale[i], l.centerness,
                       l.bbox, l.cls, l.prune, l.max_score, r.stv));
    } else {
      void* h;
      CK(r.conv(e.head, o, r.dt, n, nullptr, n, nullptr, B2_ACT_NONE, 0, 0, &h));
      FK(b2_head_post((const float*)h, n, e.head_cols, e.n_reg, e.n_cls, e.level_scale[i], l.centerness, l.bbox, l.cls, l.prune,
                      l.max_score, r.stv));
    }
  }

  // ---- tail: per (scene, level) top-k -> decode -> per (scene, class) BEV NMS (fcaf3d_neck_with_head.py:205-253,332-374)
  // The tail is a chain of small launches (one CTA per scene or per (scene, class) segment: ~0.4 ms at a few per cent of
  // the SMs).  With a tail stream it is enqueued THERE, behind an event on the main stream, so the next step's
  // quantise / hash / stem work on the main stream fills the GPU while this tail drains.
  CK(r.fsync());
  if (tail_stream != nullptr && tail_stream != r.stv) {
    const int s = r.slot();
    B2_CUDA(cudaEventRecord(e.events[s], r.st));
    B2_CUDA(cudaStreamWaitEvent((cudaStream_t)tail_stream, e.events[s], 0));
    r.st = r.stF = (cudaStream_t)tail_stream;   // from here on "F" is the tail stream (no coordinate work follows)
    r.stv = tail_stream;
    r.split = false;
  }
  const int n_cls = e.n_cls, n_reg = e.n_reg;
  const int rotated = n_reg != 6;
  const int yaw_mode = rotated ? e.yaw_mode : 0;
  std::vector<int> meta(nb + 1 + L * nb);
  std::vector<int> cnt(nb * L);
  for (int b = 0; b < nb; ++b)
    for (int li = 0; li < L; ++li) {
      const Map& m = r.maps[lv[li].map];
      const int rows = m.off[b + 1] - m.off[b];
      cnt[b * L + li] = e.nms_pre > 0 ? min(rows, e.nms_pre) : rows;
    }
  meta[0] = 0;
  int max_seg = 0;
  for (int b = 0; b < nb; ++b) {
    int s = 0;
    for (int li = 0; li < L; ++li) s += cnt[b * L + li];
    meta[b + 1] = meta[b] + s;
    max_seg = max(max_seg, s);
  }
  const long long tot_c = meta[nb];
  for (int li = 0; li < L; ++li)
    for (int b = 0; b < nb; ++b) {
      int s = meta[b];
      for (int l2 = 0; l2 < li; ++l2) s += cnt[b * L + l2];
      meta[nb + 1 + li * nb + b] = s;
    }
  int* meta_dev;
  CK(r.upload(meta.data(), (int)meta.size(), &meta_dev));
  TAKE(cand_boxes, float, tot_c * 7);
  TAKE(cand_scores, float, tot_c * n_cls);
  for (int li = 0; li < L; ++li) {
    const Map& m = r.maps[lv[li].map];
    const long long k = e.nms_pre > 0 ? e.nms_pre : m.n;
    uint8_t* mask = (uint8_t*)r.take((size_t)(m.n > 0 ? m.n : 1));
    int* pos = (int*)r.take((size_t)(m.n > 0 ? m.n : 1) * 4);
    int* tot = (int*)r.take(4);
    const size_t wsb = b2_scan_ws_bytes(m.n);
    void* ws = r.take(wsb);
    if (!mask || !pos || !tot || !ws) return B2_ERR_WS;
    FK(b2_topk_mask_n(lv[li].max_score, m.off_dev, nb, (int)k, m.n, mask, r.stv));
    B2_CUDA(cudaMemsetAsync(tot, 0, 4, r.st));
    FK(b2_scan_mask(mask, m.n, pos, tot, ws, wsb, r.stv));
    FK(b2_emit_candidates(m.coords, lv[li].centerness, lv[li].bbox, lv[li].cls, mask, pos, m.off_dev,
                          meta_dev + nb + 1 + li * nb, m.n, n_reg, n_cls, yaw_mode, e.voxel_size, cand_boxes, cand_scores,
                          r.stv));
  }
  const long long nf = tot_c * n_cls;
  const int nseg = nb * n_cls;
  B2_CHECK_ARG(nf <= out_cap, "engine: output buffers hold %lld rows, %lld needed", out_cap, nf);
  B2_CHECK_ARG(max_seg <= 4096, "engine: %d candidates per scene exceed the in-kernel NMS sort capacity", max_seg);
  *n_rows_out = nf;
  TAKE(flags, uint8_t, nf);
  TAKE(fpos, int, nf);
  TAKE(seg_boxes, float, nf * 7);
  TAKE(seg_scores, float, nf);
  TAKE(seg_src, int, nf);
  TAKE(seg_off, int, nseg + 2);
  B2_CUDA(cudaMemsetAsync(seg_off, 0, (size_t)(nseg + 2) * 4, r.st));
  {
    const size_t wsb = b2_scan_ws_bytes(nf);
    void* ws = r.take(wsb);
    if (!ws) return B2_ERR_WS;
    FK(b2_class_segments(cand_boxes, cand_scores, meta_dev, nb, n_cls, tot_c, e.score_thr, flags, fpos, seg_off + nseg + 1,
                         seg_boxes, seg_scores, seg_src, seg_off, ws, wsb, r.stv));
  }
  TAKE(keep, uint8_t, nf);
  TAKE(counts, int, nseg);
  TAKE(sorted_boxes, float, nf * 7);
  TAKE(order, int, nf);
  B2_CUDA(cudaMemsetAsync(keep, 0, (size_t)nf, r.st));
  B2_CUDA(cudaMemsetAsync(counts, 0, (size_t)nseg * 4, r.st));
  if (nf > 0) {
    const size_t wsb = b2_nms_greedy_ws_bytes(nf, rotated);
    void* ws = r.take(wsb);
    if (!ws) return B2_ERR_WS;
    FK(b2_nms_greedy_batched(seg_boxes, seg_scores, seg_off, nseg, nf, max_seg, e.iou_thr, rotated, 1, sorted_boxes, order,
                             ws, wsb, keep, counts, r.stv));
  }
  TAKE(kpos, int, nf);
  B2_CUDA(cudaMemsetAsync(out_count, 0, 4, r.st));
  {
    const size_t wsb = b2_scan_ws_bytes(nf);
    void* ws = r.take(wsb);
    if (!ws) return B2_ERR_WS;
    FK(b2_scan_mask(keep, nf, kpos, out_count, ws, wsb, r.stv));
  }
  FK(b2_collect_dets(sorted_boxes, seg_scores, order, seg_off, nseg, n_cls, keep, kpos, nf, out_boxes, out_scores, out_labels,
                     out_scene, r.stv));
  return B2_OK;
}

}  // namespace

extern "C" {

// desc: int64 words (see fcaf3d_b200/engine.py, which is the only writer of this table):
//   [0] magic 0xB2E1  [1] n_outs  [2] n_cls  [3] n_reg  [4] yaw_mode  [5] pts_threshold  [6] nms_pre  [7] point dim
//   [8] head columns (padded)  [9..12] blocks per stage  [13] number of convolution records that follow from word 16
//   records of 8 words: wt bf16 ptr, w fp32 ptr, scale ptr, shift ptr, bias ptr, kvol, cin, cout; order: stem,
//   (InstanceNorm weight ptr, bias ptr in the first two words), per stage: per block conv1, conv2, then the stage's
//   downsample; per level out_block; per level >= 1 generative transpose (packed [1, 8 cout, cin]) and up convolution;
//   head; then one record per level whose first word is the level's Scale parameter pointer.
// fl: voxel_size, score_thr, iou_thr, InstanceNorm eps.
void* b2_engine_create(const long long* desc, int n_words, const double* fl) {
  if (!desc || n_words < 16 || desc[0] != 0xB2E1) {
    b2_set_error("b2_engine_create: bad descriptor");
    return nullptr;
  }
  Engine* e = new Engine();
  if (const char* v = getenv("B2_SORT_MIN_ROWS")) e->sort_min_rows = atoll(v);
  e->n_outs = (int)desc[1], e->n_cls = (int)desc[2], e->n_reg = (int)desc[3], e->yaw_mode = (int)desc[4];
  e->pts_threshold = (int)desc[5], e->nms_pre = (int)desc[6], e->dim = (int)desc[7], e->head_cols = (int)desc[8];
  if (e->n_outs < 1 || e->n_outs > MAX_LEVELS) {
    b2_set_error("b2_engine_create: 1..4 levels supported");
    delete e;
    return nullptr;
  }
  for (int i = 0; i < MAX_LEVELS; ++i) e->blocks[i] = (int)desc[9 + i];
  e->voxel_size = (float)fl[0], e->score_thr = (float)fl[1], e->iou_thr = (float)fl[2], e->inorm_eps = (float)fl[3];
  const long long* p = desc + 16;
  const long long* end = desc + n_words;
  bool ok = true;
  auto rec = [&](ConvW* w) {
    if (p + 8 > end) {
      ok = false;
      return;
    }
    w->wt = (const void*)p[0], w->w32 = (const float*)p[1], w->scale = (const float*)p[2], w->shift = (const float*)p[3];
    w->bias = (const float*)p[4], w->kvol = (int)p[5], w->cin = (int)p[6], w->cout = (int)p[7];
    p += 8;
  };
  ConvW tmp;
  rec(&e->stem);
  rec(&tmp);
  e->inorm_w = (const float*)tmp.wt, e->inorm_b = tmp.w32;
  for (int i = 0; i < e->n_outs; ++i) {
    e->conv1[i].resize(e->blocks[i]);
    e->conv2[i].resize(e->blocks[i]);
    for (int j = 0; j < e->blocks[i]; ++j) rec(&e->conv1[i][j]), rec(&e->conv2[i][j]);
    rec(&e->down[i]);
  }
  for (int i = 0; i < e->n_outs; ++i) rec(&e->out_block[i]);
  for (int i = 1; i < e->n_outs; ++i) rec(&e->up_gen[i]), rec(&e->up_conv[i]);
  rec(&e->head);
  for (int i = 0; i < e->n_outs; ++i) {
    rec(&tmp);
    e->level_scale[i] = (const float*)tmp.wt;
  }
  if (!ok || p != end) {
    b2_set_error("b2_engine_create: descriptor length does not match the network it describes");
    delete e;
    return nullptr;
  }
  if (cudaGetDevice(&e->device) != cudaSuccess ||
      cudaHostAlloc((void**)&e->pinned, (size_t)RING * SLOT_INTS * 4, cudaHostAllocDefault) != cudaSuccess) {
    b2_set_error("b2_engine_create: pinned staging allocation failed");
    delete e;
    return nullptr;
  }
  // Row-count reads wait on these events.  Spinning (the CUDA default) is 2 - 4 % faster per step than sleeping, but needs a
  // core per waiting thread: with more (local ranks x 3 threads: caller + two lane threads) than cores the waits sleep
  // (cudaEventBlockingSync).  B2_ENGINE_BLOCKING_SYNC=0 / 1 overrides.
  const int local_ranks = getenv("LOCAL_WORLD_SIZE") ? atoi(getenv("LOCAL_WORLD_SIZE")) : 1;
  const unsigned cores = std::thread::hardware_concurrency();
  bool blocking = cores > 0 && (unsigned)(local_ranks * 3) > cores;
  if (const char* v = getenv("B2_ENGINE_BLOCKING_SYNC")) blocking = atoi(v) != 0;
  for (int i = 0; i < RING; ++i)
    cudaEventCreateWithFlags(&e->events[i], cudaEventDisableTiming | (blocking ? cudaEventBlockingSync : 0));
  if (const char* v = getenv("B2_ENGINE_SPLIT")) e->split = atoi(v) != 0;
  if (e->split && cudaStreamCreateWithFlags(&e->coord_stream, cudaStreamNonBlocking) != cudaSuccess) {
    e->coord_stream = nullptr;
    e->split = false;
  }
  return e;
}

int b2_engine_destroy(void* engine) {
  Engine* e = (Engine*)engine;
  if (!e) return B2_OK;
  for (int i = 0; i < RING; ++i) cudaEventDestroy(e->events[i]);
  if (e->coord_stream) cudaStreamDestroy(e->coord_stream);
  if (e->pinned) cudaFreeHost(e->pinned);
  delete e;
  return B2_OK;
}

// rows the output buffers must hold for `n_scenes` scenes (every candidate of every class)
long long b2_engine_out_rows(void* engine, int n_scenes) {
  Engine* e = (Engine*)engine;
  if (!e || e->nms_pre <= 0) return -1;
  return (long long)n_scenes * e->n_outs * e->nms_pre * e->n_cls;
}

int b2_engine_infer(void* engine, const float* const* points_host, const long long* n_points_host, int n_scenes, int dtype,
                    void* arena, size_t arena_bytes, float* out_boxes, float* out_scores, int* out_labels, int* out_scene,
                    int* out_count, long long out_cap, long long* n_rows_host, size_t* arena_used_host, void* stream,
                    void* tail_stream, void* inputs_ready_event, int inputs_ready, void* arena_free_event) {
  Engine* e = (Engine*)engine;
  B2_CHECK_ARG(e && points_host && n_points_host && arena && out_boxes && out_scores && out_labels && out_scene && out_count,
               "b2_engine_infer: null argument");
  B2_CHECK_ARG(dtype >= 0 && dtype <= 2, "b2_engine_infer: dtype 0 (fp32), 1 (bf16) or 2 (fp32 rows, TF32 tensor cores)");
  int dev = -1;
  B2_CUDA(cudaGetDevice(&dev));
  B2_CHECK_ARG(dev == e->device, "b2_engine_infer: engine was created on device %d, current device is %d", e->device, dev);
  Run r(*e, arena, arena_bytes, stream);
  long long rows = 0;
  const int rc = infer(*e, r, points_host, n_points_host, n_scenes, dtype, out_boxes, out_scores, out_labels, out_scene,
                       out_count, out_cap, &rows, tail_stream, inputs_ready_event, inputs_ready, arena_free_event);
  if (n_rows_host) *n_rows_host = rows;
  if (arena_used_host) *arena_used_host = r.off;
  return rc;
}

}  // extern "C"
