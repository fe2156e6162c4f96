// Host helpers of the tcgen05 kernels: TMA tensor maps through the driver entry point (the library does not link
// libcuda) and the per-device opt-in to large dynamic shared memory.
#include "tc_common.cuh"

#include <mutex>

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

static int make_map(CUtensorMap* map, CUtensorMapDataType dt, int esize, const void* base, long long rows, int cols,
                    int box_rows, int box_cols) {
  EncodeTiledFn fn = encode_fn();
  if (!fn) {
    b2_set_error("cuTensorMapEncodeTiled is not available from this driver");
    return B2_ERR_CUDA;
  }
  const cuuint64_t dims[2] = {(cuuint64_t)cols, (cuuint64_t)rows};
  const cuuint64_t strides[1] = {(cuuint64_t)cols * esize};
  const cuuint32_t box[2] = {(cuuint32_t)box_cols, (cuuint32_t)box_rows};
  const cuuint32_t estr[2] = {1, 1};
  CUresult r = fn(map, dt, 2, const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE,
                  CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
  if (r != CUDA_SUCCESS) {
    b2_set_error("cuTensorMapEncodeTiled failed (%d) rows=%lld cols=%d box_rows=%d", (int)r, rows, cols, box_rows);
    return B2_ERR_CUDA;
  }
  return B2_OK;
}

// 2-D bf16 tensor [rows, cols] (cols contiguous), box = [box_rows x 64 columns], 128-byte swizzle, zero fill out of bounds
int b2_make_bf16_map(CUtensorMap* map, const void* base, long long rows, int cols, int box_rows) {
  return make_map(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, 2, base, rows, cols, box_rows, TC_BK);
}

// 2-D fp32 tensor read as tf32 operands: box = [box_rows x 32 columns] (128 bytes per row), 128-byte swizzle
int b2_make_tf32_map(CUtensorMap* map, const void* base, long long rows, int cols, int box_rows) {
  return make_map(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 4, base, rows, cols, box_rows, 32);
}

// cudaFuncAttributeMaxDynamicSharedMemorySize is per device: `done` is one flag word per kernel instance (bit = ordinal)
int b2_optin_smem(const void* func, int bytes, unsigned long long* done) {
  int dev = 0;
  B2_CUDA(cudaGetDevice(&dev));
  static std::mutex mu;
  std::lock_guard<std::mutex> lock(mu);
  if (dev < 64 && ((*done >> dev) & 1ull)) return B2_OK;
  B2_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  if (dev < 64) *done |= 1ull << dev;
  return B2_OK;
}
