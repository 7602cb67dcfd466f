// copy_async.cu -- launchers for the register-free copy variants:
//   BULK : cp.async.bulk  global -> shared -> global            (SASS: UBLKCP)
//   TMA  : cp.async.bulk.tensor.2d through CUtensorMap tiles    (SASS: UTMALDG / UTMASTG)
// Both move raw bytes, so one instantiation serves float and double.
#include <mutex>

#include "internal.h"
#include "copy_ring.cuh"

namespace b200 {

static int ensure_smem_attr(const void* func, int bytes, int device) {
  // cudaFuncSetAttribute is per device; remember what was granted
  static std::mutex mu;
  static int granted[2][64] = {};
  const int which = (func == (const void*)copy_bulk_kernel) ? 0 : 1;
  std::lock_guard<std::mutex> lock(mu);
  if (device < 64 && granted[which][device] >= bytes) return B200_OK;
  B200_CUDA(cudaFuncSetAttribute(func, cudaFuncAttributeMaxDynamicSharedMemorySize, bytes));
  if (device < 64) granted[which][device] = bytes;
  return B200_OK;
}

static int ring_shape(const Resolved& r, const DeviceProps& props, int* stages, unsigned* chunk) {
  *stages = r.stages > 0 ? r.stages : 4;
  *chunk = r.chunk_bytes > 0 ? (unsigned)r.chunk_bytes : 32768u;
  if (*stages < 2 || *stages > kMaxStages || *chunk % 16 != 0 || *chunk < 1024) {
    set_error("bulk copy: stages must be 2..%d and chunk_bytes a multiple of 16 >= 1024 (got %d, %u)", kMaxStages,
              *stages, *chunk);
    return B200_EINVAL;
  }
  if ((size_t)*stages * *chunk > (size_t)props.max_smem_optin - 1024) {
    set_error("bulk copy: %d stages x %u bytes exceed the %d bytes of shared memory per block", *stages, *chunk,
              props.max_smem_optin);
    return B200_EINVAL;
  }
  return B200_OK;
}

int launch_copy_bulk(const void* src, void* dst, size_t bytes, const Resolved& r, const DeviceProps& props,
                     cudaStream_t stream) {
  if (((uintptr_t)src | (uintptr_t)dst) & 15) {
    set_error("bulk copy needs 16-byte aligned pointers");
    return B200_EINVAL;
  }
  int stages; unsigned chunk;
  if (int rc = ring_shape(r, props, &stages, &chunk)) return rc;
  int device = 0;
  B200_CUDA(cudaGetDevice(&device));
  const int smem = stages * (int)chunk;
  if (int rc = ensure_smem_attr((const void*)copy_bulk_kernel, smem, device)) return rc;

  BulkCopyArgs args;
  args.src = static_cast<const char*>(src);
  args.dst = static_cast<char*>(dst);
  args.chunk_bytes = chunk;
  args.nchunks = bytes / chunk;
  args.covered_bytes = args.nchunks * chunk;
  args.total_bytes = bytes;
  args.stages = stages;
  args.evict_first = (r.cache_hint == B200_HINT_EVICT_FIRST || r.cache_hint == B200_HINT_STREAM) ? 1 : 0;
  const unsigned grid = grid_for(r.grid_policy == INT32_MAX ? -1 : r.grid_policy, args.nchunks ? args.nchunks : 1, props.num_sms);
  copy_bulk_kernel<<<grid, 128, smem, stream>>>(args);
  B200_CUDA(cudaPeekAtLastError());
  count_launch();
  return B200_OK;
}

// cuTensorMapEncodeTiled comes from the driver; fetch it through the runtime so that the library
// has no link-time dependency on libcuda (it must load on a machine without a driver).
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);

static EncodeTiledFn encode_tiled_fn() {
  static EncodeTiledFn fn = nullptr;
  static std::once_flag once;
  std::call_once(once, [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
        q == cudaDriverEntryPointSuccess)
      fn = reinterpret_cast<EncodeTiledFn>(p);
  });
  return fn;
}

int launch_copy_tma(const void* src, void* dst, size_t bytes, const Resolved& r, const DeviceProps& props,
                    cudaStream_t stream) {
  if (((uintptr_t)src | (uintptr_t)dst) & 15) {
    set_error("TMA copy needs 16-byte aligned pointers");
    return B200_EINVAL;
  }
  EncodeTiledFn encode = encode_tiled_fn();
  if (!encode) {
    set_error("cuTensorMapEncodeTiled is not available from this driver");
    return B200_EUNSUPPORTED;
  }
  int stages; unsigned chunk;
  if (int rc = ring_shape(r, props, &stages, &chunk)) return rc;
  // rows of 256 x 8 bytes = 2 KiB (the largest inner box TMA allows), boxes of chunk/2048 rows
  constexpr unsigned kRowElems = 256, kRowBytes = kRowElems * 8;
  if (chunk % kRowBytes != 0 || chunk / kRowBytes > 256) {
    set_error("TMA copy: chunk_bytes must be a multiple of %u and at most %u", kRowBytes, kRowBytes * 256);
    return B200_EINVAL;
  }
  const int rows_per_chunk = (int)(chunk / kRowBytes);
  const unsigned long long rows = bytes / kRowBytes;
  int device = 0;
  B200_CUDA(cudaGetDevice(&device));
  const int smem = stages * (int)chunk;
  if (int rc = ensure_smem_attr((const void*)copy_tma_kernel, smem, device)) return rc;

  BulkCopyArgs args;
  args.src = static_cast<const char*>(src);
  args.dst = static_cast<char*>(dst);
  args.chunk_bytes = chunk;
  args.nchunks = (rows + rows_per_chunk - 1) / rows_per_chunk;
  args.covered_bytes = rows * kRowBytes;
  args.total_bytes = bytes;
  args.stages = stages;
  args.evict_first = 0;

  CUtensorMap src_map, dst_map;
  if (rows > 0) {
    const cuuint64_t dims[2] = {kRowElems, rows};
    const cuuint64_t strides[1] = {kRowBytes};  // byte stride of dimension 1
    const cuuint32_t box[2] = {kRowElems, (cuuint32_t)rows_per_chunk};
    const cuuint32_t estr[2] = {1, 1};
    const CUtensorMapL2promotion promo = CU_TENSOR_MAP_L2_PROMOTION_L2_256B;
    CUresult e1 = encode(&src_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, const_cast<void*>(src), dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, promo,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    CUresult e2 = encode(&dst_map, CU_TENSOR_MAP_DATA_TYPE_FLOAT64, 2, dst, dims, strides, box, estr,
                         CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_NONE, promo,
                         CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (e1 != CUDA_SUCCESS || e2 != CUDA_SUCCESS) {
      set_error("cuTensorMapEncodeTiled failed (%d, %d) for %llu rows", (int)e1, (int)e2, rows);
      return B200_ECUDA;
    }
  } else {
    memset(&src_map, 0, sizeof src_map);
    memset(&dst_map, 0, sizeof dst_map);
  }
  const unsigned grid = grid_for(r.grid_policy == INT32_MAX ? -1 : r.grid_policy, args.nchunks ? args.nchunks : 1, props.num_sms);
  copy_tma_kernel<<<grid, 128, smem, stream>>>(args, src_map, dst_map, rows_per_chunk);
  B200_CUDA(cudaPeekAtLastError());
  count_launch();
  return B200_OK;
}

}  // namespace b200
