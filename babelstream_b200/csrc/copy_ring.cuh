// copy_ring.cuh -- register-free copy kernels (cp.async.bulk / TMA tensor tiles through a
// shared-memory ring).  Non-template __global__ functions: include from copy_async.cu ONLY.
#pragma once

#include "kernels.cuh"

namespace b200 {

// ------------------------------------------------------------------------------------------------
// Copy without registers: global -> shared -> global through the async (TMA) proxy.
// One thread per CTA drives a ring of STAGES shared-memory buffers:
//     load(k)  : mbarrier.expect_tx + cp.async.bulk[.tensor] global->shared, completes on full[s]
//     store(k) : cp.async.bulk[.tensor] shared->global, tracked by bulk async-groups
// A stage is refilled as soon as the store that read it has finished READING shared memory
// (cp.async.bulk.wait_group.read), so STAGES-1 loads stay in flight per CTA.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) { return (uint32_t)__cvta_generic_to_shared(p); }

__device__ __forceinline__ void mbar_init(uint64_t* bar, unsigned count) {
  asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, unsigned bytes) {
  asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes) : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, unsigned parity) {
  asm volatile(
      "{\n"
      ".reg .pred p;\n"
      "B200_WAIT_%=:\n"
      "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n"
      "@p bra B200_DONE_%=;\n"
      "bra B200_WAIT_%=;\n"
      "B200_DONE_%=:\n"
      "}" ::"r"(smem_u32(bar)), "r"(parity) : "memory");
}
__device__ __forceinline__ void bulk_load(void* smem_dst, const void* gsrc, unsigned bytes, uint64_t* bar) {
  asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                   smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void bulk_load_hint(void* smem_dst, const void* gsrc, unsigned bytes, uint64_t* bar,
                                               uint64_t policy) {
  asm volatile(
      "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes.L2::cache_hint [%0], [%1], %2, [%3], %4;" ::"r"(
          smem_u32(smem_dst)), "l"(gsrc), "r"(bytes), "r"(smem_u32(bar)), "l"(policy) : "memory");
}
__device__ __forceinline__ void bulk_store(void* gdst, const void* smem_src, unsigned bytes) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group [%0], [%1], %2;" ::"l"(gdst), "r"(smem_u32(smem_src)),
               "r"(bytes) : "memory");
}
__device__ __forceinline__ void bulk_store_hint(void* gdst, const void* smem_src, unsigned bytes, uint64_t policy) {
  asm volatile("cp.async.bulk.global.shared::cta.bulk_group.L2::cache_hint [%0], [%1], %2, %3;" ::"l"(gdst),
               "r"(smem_u32(smem_src)), "r"(bytes), "l"(policy) : "memory");
}
__device__ __forceinline__ void bulk_commit() { asm volatile("cp.async.bulk.commit_group;" ::: "memory"); }
template <int N> __device__ __forceinline__ void bulk_wait_read() {
  asm volatile("cp.async.bulk.wait_group.read %0;" ::"n"(N) : "memory");
}
template <int N> __device__ __forceinline__ void bulk_wait_all() {
  asm volatile("cp.async.bulk.wait_group %0;" ::"n"(N) : "memory");
}
__device__ __forceinline__ uint64_t policy_evict_first() {
  uint64_t p;
  asm volatile("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(p));
  return p;
}
__device__ __forceinline__ void tma_load_2d(void* smem_dst, const CUtensorMap* map, int c0, int c1, uint64_t* bar) {
  asm volatile(
      "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes [%0], [%1, {%2, %3}], [%4];" ::"r"(
          smem_u32(smem_dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_u32(bar)) : "memory");
}
__device__ __forceinline__ void tma_store_2d(const CUtensorMap* map, int c0, int c1, const void* smem_src) {
  asm volatile("cp.async.bulk.tensor.2d.global.shared::cta.tile.bulk_group [%0, {%1, %2}], [%3];" ::"l"(map), "r"(c0),
               "r"(c1), "r"(smem_u32(smem_src)) : "memory");
}

constexpr int kMaxStages = 8;

struct BulkCopyArgs {
  const char* src;
  char* dst;
  unsigned long long nchunks;       // chunks handled by the ring (the last one may be partial with TMA)
  unsigned long long covered_bytes; // bytes the ring moves; [covered_bytes, total_bytes) is the remainder
  unsigned long long total_bytes;
  unsigned int chunk_bytes;       // multiple of 16
  int stages;
  int evict_first;
};

// Generic ring, shared by the linear-bulk and the tensor-map variants.
template <bool TENSOR>
__device__ __forceinline__ void copy_ring(const BulkCopyArgs& args, const CUtensorMap* src_map,
                                          const CUtensorMap* dst_map, int rows_per_chunk, char* smem,
                                          uint64_t* full) {
  const int S = args.stages;
  const unsigned bytes = args.chunk_bytes;
  const unsigned long long first = blockIdx.x, stride = gridDim.x;
  if (first >= args.nchunks) return;
  const unsigned long long mine = (args.nchunks - first + stride - 1) / stride;  // chunks of this CTA
  const uint64_t policy = args.evict_first ? policy_evict_first() : 0;

  auto issue_load = [&](unsigned long long k) {
    const int s = (int)(k % S);
    const unsigned long long chunk = first + k * stride;
    mbar_expect_tx(&full[s], bytes);
    if constexpr (TENSOR) tma_load_2d(smem + (size_t)s * bytes, src_map, 0, (int)(chunk * rows_per_chunk), &full[s]);
    else if (args.evict_first) bulk_load_hint(smem + (size_t)s * bytes, args.src + chunk * bytes, bytes, &full[s], policy);
    else bulk_load(smem + (size_t)s * bytes, args.src + chunk * bytes, bytes, &full[s]);
  };

  const unsigned long long pre = mine < (unsigned long long)S ? mine : (unsigned long long)S;
  for (unsigned long long k = 0; k < pre; k++) issue_load(k);

  for (unsigned long long k = 0; k < mine; k++) {
    const int s = (int)(k % S);
    mbar_wait(&full[s], (unsigned)((k / S) & 1));
    const unsigned long long chunk = first + k * stride;
    if constexpr (TENSOR) tma_store_2d(dst_map, 0, (int)(chunk * rows_per_chunk), smem + (size_t)s * bytes);
    else if (args.evict_first) bulk_store_hint(args.dst + chunk * bytes, smem + (size_t)s * bytes, bytes, policy);
    else bulk_store(args.dst + chunk * bytes, smem + (size_t)s * bytes, bytes);
    bulk_commit();
    // refill the stage used by the PREVIOUS store: allow only the newest group to still read smem
    if (k >= 1 && (k - 1) + S < mine) {
      bulk_wait_read<1>();
      issue_load((k - 1) + S);
    }
  }
  bulk_wait_all<0>();
}

__device__ __forceinline__ void copy_remainder(const BulkCopyArgs& args) {
  // bytes the ring does not cover: plain 16-byte / 1-byte accesses by block 0
  if (blockIdx.x != 0) return;
  const unsigned long long done = args.covered_bytes;
  const unsigned long long rem = args.total_bytes - done;
  const unsigned long long nv = rem / 16;
  for (unsigned long long v = threadIdx.x; v < nv; v += blockDim.x) {
    const Words<16> w = ldg<16, H_DEFAULT, true>(args.src + done + v * 16);
    stg<16, H_DEFAULT>(args.dst + done + v * 16, w);
  }
  for (unsigned long long i = done + nv * 16 + threadIdx.x; i < args.total_bytes; i += blockDim.x)
    args.dst[i] = args.src[i];
}

__global__ void __launch_bounds__(128)
copy_bulk_kernel(const BulkCopyArgs args) {
  extern __shared__ __align__(1024) char smem[];
  __shared__ uint64_t full[kMaxStages];
  if (threadIdx.x == 0) {
    for (int s = 0; s < args.stages; s++) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    copy_ring<false>(args, nullptr, nullptr, 0, smem, full);
  }
  copy_remainder(args);
}

// Tensor-map variant: the array prefix is viewed as a 2-D tensor of rows of `row_bytes` bytes;
// a chunk is a box of rows_per_chunk rows.  The last box may hang over the end of the tensor:
// TMA zero-fills the out-of-bounds rows on load and drops them on store.
__global__ void __launch_bounds__(128)
copy_tma_kernel(const BulkCopyArgs args, const __grid_constant__ CUtensorMap src_map,
                const __grid_constant__ CUtensorMap dst_map, const int rows_per_chunk) {
  extern __shared__ __align__(1024) char smem[];
  __shared__ uint64_t full[kMaxStages];
  if (threadIdx.x == 0) {
    asm volatile("prefetch.tensormap [%0];" ::"l"(&src_map) : "memory");
    asm volatile("prefetch.tensormap [%0];" ::"l"(&dst_map) : "memory");
    for (int s = 0; s < args.stages; s++) mbar_init(&full[s], 1);
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    copy_ring<true>(args, &src_map, &dst_map, rows_per_chunk, smem, full);
  }
  copy_remainder(args);
}

}  // namespace b200
