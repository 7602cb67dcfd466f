// internal.h -- shared declarations of libb200stream's translation units (not part of the ABI).
#pragma once

#include <cuda_runtime.h>
#include <stddef.h>
#include <stdint.h>

#include "../../include/b200_stream.h"

namespace b200 {

// thread-local error message behind b200_last_error()
void set_error(const char* fmt, ...) __attribute__((format(printf, 1, 2)));
int cuda_fail(cudaError_t e, const char* expr, const char* file, int line);  // sets the message, returns B200_ECUDA

#define B200_CUDA(expr)                                                                   \
  do {                                                                                    \
    cudaError_t e__ = (expr);                                                             \
    if (e__ != cudaSuccess) return ::b200::cuda_fail(e__, #expr, __FILE__, __LINE__);     \
  } while (0)

struct DeviceProps {
  int num_sms;
  int l2_bytes;
  int cc_major, cc_minor;
  int max_smem_optin;
};
// cached per device; returns nullptr (and sets the error) when there is no usable device
const DeviceProps* current_device_props(int* device_out);

void count_launch(unsigned n = 1);
void uncount_launch(unsigned n);   // a stream capture "launched" n kernels that never ran
uint64_t thread_launch_count();    // launches counted on the calling thread

// Every register-path kernel (elementwise, dot, collect) is launched through here.  flags & B200_LAUNCH_PDL adds the
// programmatic-stream-serialization attribute (see pdl_entry() in kernels.cuh).  The two async-proxy copy variants
// (copy_async.cu) launch with <<<>>> and have no dependent-launch entry: for them B200_LAUNCH_PDL has no effect (a plain
// launch is fully stream-ordered, the stronger guarantee).
template <typename... KArgs, typename... Args>
int launch_kernel(void (*kernel)(KArgs...), unsigned grid, int block, size_t smem, cudaStream_t stream, int flags,
                  Args&&... args) {
  cudaLaunchConfig_t cfg = {};
  cfg.gridDim = dim3(grid);
  cfg.blockDim = dim3((unsigned)block);
  cfg.dynamicSmemBytes = smem;
  cfg.stream = stream;
  cudaLaunchAttribute attr[1];
  if (flags & B200_LAUNCH_PDL) {
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = 1;
  }
  const cudaError_t e = cudaLaunchKernelEx(&cfg, kernel, static_cast<KArgs>(args)...);
  if (e != cudaSuccess) return cuda_fail(e, "cudaLaunchKernelEx", __FILE__, __LINE__);
  count_launch();  // only what was really enqueued
  return B200_OK;
}

// bound of the peer-memory all-reduce's wait for a missing peer (dot.cu)
unsigned long long p2p_timeout_ns();
void set_p2p_timeout_ns(unsigned long long ns);
unsigned long long p2p_timeout_gen();  // changes whenever the bound does (a captured dot launch has the bound baked in)

// A fully resolved launch: no "auto" left.
struct Resolved {
  int vec_bytes, unroll, block, cache_hint, variant, contract, stages, chunk_bytes;
  int flags;        // B200_LAUNCH_* bits
  int keep_mb;      // MB of traffic at the END of the sweep whose lines stay in L2 (evict-first hint only); 0 = none
  int grid_policy;  // the cfg.grid convention: >0 exact, <0 per-SM, INT32_MAX one tile per block
};

// Size class of a problem: the footprint of the three arrays against the L2 size (tuning.cu)
enum { SIZE_L2 = 0, SIZE_KNEE = 1, SIZE_HBM = 2, SIZE_HBM_XL = 3 };
int size_class(size_t n, int elem_bytes, int l2_bytes);

// Fills every auto field of `in` (may be null) for (kernel, n, elem_bytes).  Pure host logic.
int resolve_cfg(int kernel, size_t n, int elem_bytes, int num_sms, int l2_bytes, const b200_launch_cfg* in,
                Resolved* out);
// grid size for `ntiles` tiles under the policy
unsigned grid_for(int grid_policy, unsigned long long ntiles, int num_sms);
// first tile/block (dispatch order) of the keep region: the last keep_mb MB of a sweep that moves `bytes_per_unit`
// bytes per tile (elementwise) or per block (dot); ~0 when nothing is kept
unsigned long long keep_from_for(int keep_mb, int cache_hint, unsigned long long units, double bytes_per_unit);

// Contiguous shard g of ngpus with boundaries on multiples of align_bytes (128 for device memory,
// b200_shard_range; 2 MiB for managed memory so that no page has two owners)
void shard_range_aligned(size_t n, int elem_bytes, int ngpus, int g, size_t align_bytes, size_t* begin, size_t* count);

// Geometry of one vectorised pass over n elements whose (up to three) base pointers are p0..p2
// (null = unused).  Falls back to scalar accesses when the pointers are not mutually congruent
// modulo vec_bytes.
struct Geometry {
  int vec_bytes;
  int unroll;      // == the requested unroll, or 4 on the scalar fallback path
  int cache_hint;  // == the requested hint, or B200_HINT_DEFAULT on the scalar fallback path
  unsigned head;
  unsigned long long nvec, ntiles;
};
Geometry make_geometry(const void* p0, const void* p1, const void* p2, size_t n, int elem_bytes, int vec_bytes,
                       int block, int unroll, int cache_hint);

// How a grid walks its tiles (pure host logic, tuning.cu; unit-tested without a GPU through b200_plan_sweep).
//   elementwise: dispatch index p in [0, ntiles) starts at vector origin + stride*p (blocks take p = blockIdx.x,
//                blockIdx.x + grid, ...); p >= keep_from keeps its lines in L2.
//   dot:         logical block L in [0, grid) owns tiles [L*tiles_per_block, min((L+1)*tiles_per_block, ntiles))
//                (tiles_per_block = 0: grid-stride L, L+grid, ...); dispatch index p is L, or grid-1-L when reversed;
//                p >= keep_from loads with evict-normal.
constexpr unsigned kDotMaxPartials = 16384;  // == kDotMaxBlocks (kernels.cuh): block partials the dot workspace holds
struct SweepPlan {
  unsigned grid = 1;
  unsigned tiles_per_block = 0;
  bool reverse = false;
  unsigned long long keep_from = ~0ull;  // ~0: no keep region
  unsigned long long origin = 0;
  long long stride = 0;
};
SweepPlan plan_sweep(int kernel, const Resolved& r, const Geometry& g, int num_sms);

// Completion doorbell of a launch (see kernels.cuh); flag == nullptr: none.
struct Bell {
  void* ticket = nullptr;      // device word, zero between launches
  void* flag = nullptr;        // device-visible address of a mapped pinned host word
  unsigned long long seq = 0;  // value the kernel stores when its last block has finished
};

// Per-op elementwise launchers (one translation unit each, so they compile in parallel).
struct EwCall {
  void *a, *b, *c;
  size_t n;
  int elem_bytes;
  double s0, s1, s2;
  Resolved r;
  Geometry g;
  unsigned grid;
  cudaStream_t stream;
  Bell bell;
  SweepPlan sweep;                          // direction, keep region (plan_sweep)
};
int launch_ew_init(const EwCall& call);
int launch_ew_copy(const EwCall& call);
int launch_ew_mul(const EwCall& call);
int launch_ew_add(const EwCall& call);
int launch_ew_triad(const EwCall& call);
int launch_ew_nstream(const EwCall& call);

struct DotCall {
  const void *a, *b;
  size_t n;
  int elem_bytes;
  void* workspace;
  void* result_dev;
  void* result_host;
  // fused all-reduce over peer memory (p2p_world <= 1: none)
  int p2p_world = 0, p2p_rank = 0;
  unsigned long long p2p_seq = 0;
  void* const* p2p_mailboxes = nullptr;
  bool p2p_send_only = false;  // the sending half only; launch_dot_collect sums later
  void* status_host = nullptr; // mapped pinned word set to 1 when the collect times out (LEVEL 2), or null
  void* seq_counters = nullptr;// device: [0] doorbell value, [1] all-reduce sequence number, bumped by the kernel (LEVEL 2)
  Bell bell;  // only .flag and .seq are used: dot has its own ticket
  Resolved r;
  Geometry g;
  unsigned grid;
  SweepPlan sweep;                          // chunking, direction, keep region (plan_sweep)
  cudaStream_t stream;
};
int launch_dot_impl(const DotCall& call);
int launch_dot_collect(int elem_bytes, void* result_dev, void* result_host, int world, int rank, unsigned long long seq,
                       void* const* mailboxes, void* status_dev, int flags, cudaStream_t stream);

// shared bodies of the LEVEL 1 launchers (api.cu), also called by LEVEL 2 with a doorbell
int launch_ew(int kernel, void* a, void* b, void* c, size_t n, int elem_bytes, double s0, double s1, double s2,
              const b200_launch_cfg* cfg, const b200_doorbell* bell, void* stream);
int launch_dot_common(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* result_dev,
                      void* result_host_mapped, int world, int rank, unsigned long long seq, void* const* mailboxes,
                      const b200_launch_cfg* cfg, const b200_doorbell* bell, void* stream, bool send_only = false,
                      void* status_host = nullptr, void* seq_counters = nullptr);

int launch_copy_bulk(const void* src, void* dst, size_t bytes, const Resolved& r, const DeviceProps& props,
                     cudaStream_t stream);
int launch_copy_tma(const void* src, void* dst, size_t bytes, const Resolved& r, const DeviceProps& props,
                    cudaStream_t stream);

}  // namespace b200
