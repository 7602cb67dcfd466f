// dot_static.cu -- b200_launch_dot_static: dot for callers that cannot hand the library any device memory of
// their own beyond the two vectors and the result.
//
// Who that is.  b200_launch_dot wants a workspace that the CALLER obtained from cudaMalloc and zeroed.  Some hosts of
// this library have no way of doing that: a plug-in that is given device pointers by its host application but does not
// link the CUDA runtime itself (a Go/Java/Fortran binding that only forwards addresses); a caller inside a region where
// the allocator must stay quiet -- cudaMalloc and cudaFree synchronise the whole device, which stalls every other
// stream of a latency-sensitive service; and a process that runs under a fixed memory budget in which every byte of
// device memory was handed out at start-up, so an allocation at call time can only fail.  For them the block partials,
// the tickets and the group sums live in the array below, which is part of the library's own image.
//
// What it costs.  The array belongs to the module, not to a call: every GPU on which this library's code gets loaded
// holds its own copy of it (kDotWorkspaceBytes = 139 776 bytes) from then until the process exits, whether or not this
// entry point is ever used there.
//
// What the caller must know.
//   * ONE array per GPU, shared by every call on that GPU.  Two b200_launch_dot_static calls that can overlap on the
//     device (different streams, no dependency between them) race on the partials and the tickets and both return
//     garbage.  Keep them on one stream, or order them with events.  Use b200_launch_dot with one workspace per stream
//     when concurrent dots are needed.
//   * The array keeps what the previous call left in it.  The kernel returns its tickets to zero by itself, but a call
//     that was aborted half-way (device reset, a sticky error) would not have; the entry point therefore clears the
//     ticket / group-sum region (not the partials: they are written before they are read) on the caller's stream ahead
//     of every launch.  That is a cudaMemsetAsync of 8.5 KB, not an allocation.
//   * In host code `g_dot_scratch` is only a handle for the runtime, not an address a kernel may be given.  The device
//     address is looked up with cudaGetSymbolAddress, once per GPU, while that GPU is current, and cached.  The look-up
//     happens inside the first call made on a GPU: make that first call outside any stream capture.
//   * No cudaMalloc, cudaFree, cudaMallocAsync, cudaHostAlloc or torch allocation is made here or below.
//
// No counterpart in the reference: its model allocates the block sums in its constructor (src/cuda/CUDAStream.cu:128-145).
#include <atomic>
#include <mutex>

#include "internal.h"
#include "kernels.cuh"

namespace b200 {

// The whole workspace of one dot launch (layout: kDotWs* in kernels.cuh).  Zero-initialised at module load.
static __device__ __align__(256) unsigned char g_dot_scratch[kDotWorkspaceBytes];

static_assert(B200_DOT_STATIC_MAX_BLOCKS == kDotMaxBlocks, "header and kernel disagree about the static workspace's capacity");
static_assert(B200_DOT_STATIC_SCRATCH_BYTES == kDotWorkspaceBytes, "header and kernel disagree about the static workspace's size");

constexpr int kStaticMaxDevices = 64;
static std::atomic<void*> g_scratch_addr[kStaticMaxDevices];  // device address of g_dot_scratch on GPU d, or null
static std::mutex g_scratch_mutex;

// Device address of this GPU's copy of g_dot_scratch.  First call on a GPU: cudaGetSymbolAddress with that GPU current
// (it also makes the runtime load the module there, which is why it must not happen under stream capture).
static int static_scratch_for_current_device(void** out) {
  int dev = -1;
  B200_CUDA(cudaGetDevice(&dev));
  if (dev < 0 || dev >= kStaticMaxDevices) {
    set_error("b200_launch_dot_static: device index %d is beyond the %d devices the address cache holds", dev, kStaticMaxDevices);
    return B200_EINVAL;
  }
  void* p = g_scratch_addr[dev].load(std::memory_order_acquire);
  if (!p) {
    std::lock_guard<std::mutex> lock(g_scratch_mutex);
    p = g_scratch_addr[dev].load(std::memory_order_relaxed);
    if (!p) {
      B200_CUDA(cudaGetSymbolAddress(&p, g_dot_scratch));
      if (!p || (reinterpret_cast<uintptr_t>(p) & 255)) {
        set_error("b200_launch_dot_static: cudaGetSymbolAddress gave %p for the scratch array", p);
        return B200_ECUDA;
      }
      g_scratch_addr[dev].store(p, std::memory_order_release);
    }
  }
  *out = p;
  return B200_OK;
}

}  // namespace b200

using namespace b200;

extern "C" int b200_launch_dot_static(const void* a, const void* b, size_t n, int elem_bytes, void* result_dev,
                                      void* result_host_mapped, const b200_launch_cfg* cfg, void* stream) {
  if (elem_bytes != 4 && elem_bytes != 8) {
    set_error("elem_bytes must be 4 (float) or 8 (double), got %d", elem_bytes);
    return B200_EINVAL;
  }
  if (n > 0 && (!a || !b)) { set_error("b200_launch_dot_static: null pointer"); return B200_EINVAL; }
  if (!result_dev && !result_host_mapped) {
    set_error("b200_launch_dot_static: nowhere to put the sum (result_dev and result_host_mapped are both null)");
    return B200_EINVAL;
  }
  // The limit: one partial per block, B200_DOT_STATIC_MAX_BLOCKS partials.  The automatic grid stays far below it for
  // every n; only an explicit cfg->grid on a long vector can ask for more.
  {
    const DeviceProps* props = current_device_props(nullptr);
    if (!props) return B200_ENODEVICE;
    Resolved r;
    if (int rc = resolve_cfg(B200_DOT, n, elem_bytes, props->num_sms, props->l2_bytes, cfg, &r)) return rc;
    const Geometry g = make_geometry(a, b, nullptr, n, elem_bytes, r.vec_bytes, r.block, r.unroll, r.cache_hint);
    const unsigned grid = plan_sweep(B200_DOT, r, g, props->num_sms).grid;  // what launch_dot_common will launch
    if (grid > (unsigned)B200_DOT_STATIC_MAX_BLOCKS) {
      set_error("b200_launch_dot_static: this launch shape needs %u block partials for n = %zu, the library's built-in "
                "scratch array holds %d; use a smaller grid (or the automatic one)", grid, n, B200_DOT_STATIC_MAX_BLOCKS);
      return B200_EINVAL;
    }
  }
  void* scratch = nullptr;
  if (int rc = static_scratch_for_current_device(&scratch)) return rc;
  // tickets, status word, group sums: everything behind the partials must be zero when the kernel starts
  B200_CUDA(cudaMemsetAsync(static_cast<char*>(scratch) + kDotWsTop, 0, kDotWorkspaceBytes - kDotWsTop,
                            static_cast<cudaStream_t>(stream)));
  return launch_dot_common(a, b, n, elem_bytes, scratch, result_dev, result_host_mapped, 0, 0, 0, nullptr, cfg, nullptr, stream);
}
