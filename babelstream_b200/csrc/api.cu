// api.cu -- LEVEL 0 (queries) and LEVEL 1 (raw launchers) of include/b200_stream.h.
#include <atomic>
#include <dlfcn.h>
#include <limits.h>
#include <mutex>
#include <stdio.h>
#include <string.h>

#include "internal.h"
#include "kernels.cuh"

namespace b200 {

static std::atomic<uint64_t> g_launches{0};
static thread_local uint64_t t_launches = 0;
void count_launch(unsigned n) { g_launches.fetch_add(n, std::memory_order_relaxed); t_launches += n; }
void uncount_launch(unsigned n) { g_launches.fetch_sub(n, std::memory_order_relaxed); t_launches -= n; }
uint64_t thread_launch_count() { return t_launches; }

int cuda_fail(cudaError_t e, const char* expr, const char* file, int line) {
  // same information as the reference's CU() macro prints (src/cuda/CUDAStream.cu:14-17)
  set_error("CUDA Error at %s:%d: %s (%d)\n  %s", file, line, cudaGetErrorString(e), (int)e, expr);
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver || e == cudaErrorInvalidDevice) return B200_ENODEVICE;
  if (e == cudaErrorMemoryAllocation) return B200_ENOMEM;
  return B200_ECUDA;
}

static const int kMaxDevices = 64;
static DeviceProps g_props[kMaxDevices];
static std::atomic<int> g_props_ready[kMaxDevices];
static std::mutex g_props_mu;

static int load_props(int dev) {
  std::lock_guard<std::mutex> lock(g_props_mu);
  if (g_props_ready[dev].load(std::memory_order_acquire)) return B200_OK;
  DeviceProps p;
  B200_CUDA(cudaDeviceGetAttribute(&p.num_sms, cudaDevAttrMultiProcessorCount, dev));
  B200_CUDA(cudaDeviceGetAttribute(&p.l2_bytes, cudaDevAttrL2CacheSize, dev));
  B200_CUDA(cudaDeviceGetAttribute(&p.cc_major, cudaDevAttrComputeCapabilityMajor, dev));
  B200_CUDA(cudaDeviceGetAttribute(&p.cc_minor, cudaDevAttrComputeCapabilityMinor, dev));
  B200_CUDA(cudaDeviceGetAttribute(&p.max_smem_optin, cudaDevAttrMaxSharedMemoryPerBlockOptin, dev));
  if (p.cc_major != 10) {
    // the fatbin holds sm_100a code only: fail loudly instead of hitting "no kernel image"
    set_error("device %d is compute capability %d.%d; libb200stream is built for sm_100a (B200) only", dev,
              p.cc_major, p.cc_minor);
    return B200_EUNSUPPORTED;
  }
  g_props[dev] = p;
  g_props_ready[dev].store(1, std::memory_order_release);
  return B200_OK;
}

const DeviceProps* current_device_props(int* device_out) {
  int dev = -1;
  cudaError_t e = cudaGetDevice(&dev);
  if (e != cudaSuccess) { cuda_fail(e, "cudaGetDevice(&dev)", __FILE__, __LINE__); return nullptr; }
  if (dev < 0 || dev >= kMaxDevices) { set_error("device index %d out of range", dev); return nullptr; }
  if (!g_props_ready[dev].load(std::memory_order_acquire) && load_props(dev) != B200_OK) return nullptr;
  if (device_out) *device_out = dev;
  return &g_props[dev];
}

static int check_common(size_t n, int elem_bytes) {
  (void)n;
  if (elem_bytes != 4 && elem_bytes != 8) { set_error("elem_bytes must be 4 (float) or 8 (double), got %d", elem_bytes); return B200_EINVAL; }
  return B200_OK;
}

// shared body of the six elementwise launchers
static int check_bell(const b200_doorbell* bell, bool need_ticket) {
  if (!bell) return B200_OK;
  if (!bell->flag_host_mapped || (need_ticket && !bell->ticket) || bell->seq == 0) {
    set_error("doorbell needs a flag address, a non-zero seq%s", need_ticket ? " and a ticket word" : "");
    return B200_EINVAL;
  }
  return B200_OK;
}

int launch_ew(int kernel, void* a, void* b, void* c, size_t n, int elem_bytes, double s0, double s1, double s2,
              const b200_launch_cfg* cfg, const b200_doorbell* bell, void* stream) {
  if (int rc = check_common(n, elem_bytes)) return rc;
  if (int rc = check_bell(bell, true)) return rc;
  const bool need_a = kernel != B200_MUL, need_b = kernel != B200_COPY, need_c = true;  // which arrays the op touches
  if (n > 0 && ((need_a && !a) || (need_b && !b) || (need_c && !c))) { set_error("null array pointer"); return B200_EINVAL; }
  // No-alias rule: the array a kernel writes must not overlap an array it reads through the read-only
  // (ld.global.nc) path -- an in-place call such as triad(a, a, c) would read, non-coherently, memory the same
  // grid is writing.  (nstream's a is read and written by design and is loaded coherently.)
  if (n > 0) {
    const size_t bytes = n * (size_t)elem_bytes;
    const char* out = static_cast<const char*>(kernel == B200_MUL ? b : (kernel == B200_COPY || kernel == B200_ADD) ? c : a);
    const void* ins[3] = {(kernel == B200_COPY || kernel == B200_ADD) ? a : nullptr,
                          (kernel == B200_ADD || kernel == B200_TRIAD || kernel == B200_NSTREAM) ? b : nullptr,
                          (kernel == B200_MUL || kernel == B200_TRIAD || kernel == B200_NSTREAM) ? c : nullptr};
    for (const void* in : ins) {
      const char* r = static_cast<const char*>(in);
      if (r && kernel != B200_INIT && r < out + bytes && out < r + bytes) {
        set_error("the array a kernel writes must not overlap the arrays it reads (in-place calls are not supported)");
        return B200_EINVAL;
      }
    }
    if (kernel == B200_INIT) {
      const char* p[3] = {static_cast<const char*>(a), static_cast<const char*>(b), static_cast<const char*>(c)};
      for (int i = 0; i < 3; i++)
        for (int j = i + 1; j < 3; j++)
          if (p[i] < p[j] + bytes && p[j] < p[i] + bytes) { set_error("init: the three arrays must not overlap"); return B200_EINVAL; }
    }
  }
  const DeviceProps* props = current_device_props(nullptr);
  if (!props) return B200_ENODEVICE;
  EwCall call;
  if (int rc = resolve_cfg(kernel, n, elem_bytes, props->num_sms, props->l2_bytes, cfg, &call.r)) return rc;
  if (n == 0) {
    if (bell) { set_error("a doorbell launch needs n > 0 (nothing would ring it)"); return B200_EINVAL; }
    return B200_OK;
  }
  if (bell) { call.bell.ticket = bell->ticket; call.bell.flag = bell->flag_host_mapped; call.bell.seq = bell->seq; }
  call.a = a; call.b = b; call.c = c; call.n = n; call.elem_bytes = elem_bytes;
  call.s0 = s0; call.s1 = s1; call.s2 = s2;
  call.stream = static_cast<cudaStream_t>(stream);

  if (kernel == B200_COPY && call.r.variant != B200_VARIANT_LDST && !bell) {
    const size_t bytes = n * (size_t)elem_bytes;
    if (((uintptr_t)a | (uintptr_t)c) & 15) {
      // the async-proxy paths need 16-byte alignment: fall through to the register path
    } else if (call.r.variant == B200_VARIANT_BULK) {
      return launch_copy_bulk(a, c, bytes, call.r, *props, call.stream);
    } else {
      return launch_copy_tma(a, c, bytes, call.r, *props, call.stream);
    }
  }
  call.g = make_geometry(need_a ? a : nullptr, need_b ? b : nullptr, c, n, elem_bytes, call.r.vec_bytes, call.r.block,
                         call.r.unroll, call.r.cache_hint);
  if (call.g.ntiles >> 32) {
    set_error("launch shape gives %llu tiles (2^32 or more): use a larger block or unroll", (unsigned long long)call.g.ntiles);
    return B200_EINVAL;
  }
  call.sweep = plan_sweep(kernel, call.r, call.g, props->num_sms);
  call.grid = call.sweep.grid;
  switch (kernel) {
    case B200_INIT: return launch_ew_init(call);
    case B200_COPY: return launch_ew_copy(call);
    case B200_MUL: return launch_ew_mul(call);
    case B200_ADD: return launch_ew_add(call);
    case B200_TRIAD: return launch_ew_triad(call);
    case B200_NSTREAM: return launch_ew_nstream(call);
  }
  set_error("not an elementwise kernel: %d", kernel);
  return B200_EINVAL;
}

int launch_dot_common(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* result_dev,
                      void* result_host_mapped, int world, int rank, unsigned long long seq, void* const* mailboxes,
                      const b200_launch_cfg* cfg, const b200_doorbell* bell, void* stream, bool send_only, void* status_host,
                      void* seq_counters) {
  if (int rc = check_common(n, elem_bytes)) return rc;
  if (bell && seq_counters) {
    if (!bell->flag_host_mapped) { set_error("doorbell needs a flag address"); return B200_EINVAL; }
  } else if (int rc = check_bell(bell, false)) return rc;
  if (!workspace || (n > 0 && (!a || !b))) { set_error("b200_launch_dot: null pointer"); return B200_EINVAL; }
  const DeviceProps* props = current_device_props(nullptr);
  if (!props) return B200_ENODEVICE;
  DotCall call;
  if (int rc = resolve_cfg(B200_DOT, n, elem_bytes, props->num_sms, props->l2_bytes, cfg, &call.r)) return rc;
  call.a = a; call.b = b; call.n = n; call.elem_bytes = elem_bytes;
  call.workspace = workspace; call.result_dev = result_dev; call.result_host = result_host_mapped;
  call.p2p_world = world; call.p2p_rank = rank; call.p2p_seq = seq; call.p2p_mailboxes = mailboxes;
  call.p2p_send_only = send_only;
  call.status_host = status_host;
  call.seq_counters = seq_counters;
  if (bell) { call.bell.flag = bell->flag_host_mapped; call.bell.seq = bell->seq; }
  call.stream = static_cast<cudaStream_t>(stream);
  call.g = make_geometry(a, b, nullptr, n, elem_bytes, call.r.vec_bytes, call.r.block, call.r.unroll, call.r.cache_hint);
  call.sweep = plan_sweep(B200_DOT, call.r, call.g, props->num_sms);
  call.grid = call.sweep.grid;
  return launch_dot_impl(call);
}

}  // namespace b200

using namespace b200;

extern "C" {

const char* b200_version(void) { return "b200-stream 0.1 (sm_100a; variant of BabelStream 5.0)"; }

uint64_t b200_launch_count(void) { return g_launches.load(std::memory_order_relaxed); }

int b200_device_count(void) {
  int count = 0;
  cudaError_t e = cudaGetDeviceCount(&count);
  if (e == cudaErrorNoDevice || e == cudaErrorInsufficientDriver) { (void)cudaGetLastError(); return 0; }
  if (e != cudaSuccess) return cuda_fail(e, "cudaGetDeviceCount(&count)", __FILE__, __LINE__);
  return count;
}

int b200_device_name(int device, char* buf, int buflen) {
  if (!buf || buflen <= 0) { set_error("b200_device_name: no buffer"); return B200_EINVAL; }
  cudaDeviceProp p;
  B200_CUDA(cudaGetDeviceProperties(&p, device));
  snprintf(buf, (size_t)buflen, "%s", p.name);
  return B200_OK;
}

int b200_driver_version(int device) {
  (void)device;
  int v = 0;
  B200_CUDA(cudaDriverGetVersion(&v));
  return v;
}

int b200_device_info(int device, int* num_sms, int* cc_major, int* cc_minor, size_t* total_mem, int* l2_bytes,
                     int* mem_clock_khz, int* mem_bus_bits) {
  int v;
  if (num_sms) { B200_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMultiProcessorCount, device)); *num_sms = v; }
  if (cc_major) { B200_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMajor, device)); *cc_major = v; }
  if (cc_minor) { B200_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrComputeCapabilityMinor, device)); *cc_minor = v; }
  if (l2_bytes) { B200_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrL2CacheSize, device)); *l2_bytes = v; }
  if (mem_clock_khz) { B200_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrMemoryClockRate, device)); *mem_clock_khz = v; }
  if (mem_bus_bits) { B200_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrGlobalMemoryBusWidth, device)); *mem_bus_bits = v; }
  if (total_mem) {
    cudaDeviceProp p;
    B200_CUDA(cudaGetDeviceProperties(&p, device));
    *total_mem = p.totalGlobalMem;
  }
  return B200_OK;
}

int b200_device_ecc(int device) {
  int v = 0;
  B200_CUDA(cudaDeviceGetAttribute(&v, cudaDevAttrEccEnabled, device));
  return v ? 1 : 0;
}

// The reference's banner prints NVML's CURRENT memory clock (src/cuda/CUDAStream.cu:96-108), not the attribute's
// maximum.  NVML is resolved at run time; without it the attribute value is returned.
int b200_device_mem_clock_mhz(int device) {
  int khz = 0;
  B200_CUDA(cudaDeviceGetAttribute(&khz, cudaDevAttrMemoryClockRate, device));
  int mhz = khz / 1000;
  static void* nvml = dlopen("libnvidia-ml.so.1", RTLD_NOW | RTLD_LOCAL);
  if (nvml) {
    typedef int (*InitFn)(void);
    typedef int (*ByBusFn)(const char*, void**);
    typedef int (*ClockFn)(void*, int, unsigned int*);
    static InitFn init = (InitFn)dlsym(nvml, "nvmlInit_v2");
    static ByBusFn by_bus = (ByBusFn)dlsym(nvml, "nvmlDeviceGetHandleByPciBusId_v2");
    static ClockFn clock = (ClockFn)dlsym(nvml, "nvmlDeviceGetClockInfo");
    char bus[32];
    void* dev = nullptr;
    unsigned int mem = 0;
    if (init && by_bus && clock && cudaDeviceGetPCIBusId(bus, (int)sizeof bus, device) == cudaSuccess && init() == 0 &&
        by_bus(bus, &dev) == 0 && clock(dev, /*NVML_CLOCK_MEM*/ 2, &mem) == 0 && mem > 0)
      mhz = (int)mem;
    (void)cudaGetLastError();
  }
  return mhz;
}

int b200_launch_init(void* a, void* b, void* c, size_t n, int elem_bytes, double ia, double ib, double ic,
                     const b200_launch_cfg* cfg, void* stream) {
  return launch_ew(B200_INIT, a, b, c, n, elem_bytes, ia, ib, ic, cfg, nullptr, stream);
}
int b200_launch_copy(const void* a, void* c, size_t n, int elem_bytes, const b200_launch_cfg* cfg, void* stream) {
  return launch_ew(B200_COPY, const_cast<void*>(a), nullptr, c, n, elem_bytes, 0, 0, 0, cfg, nullptr, stream);
}
int b200_launch_mul(void* b, const void* c, size_t n, int elem_bytes, double scalar, const b200_launch_cfg* cfg,
                    void* stream) {
  return launch_ew(B200_MUL, nullptr, b, const_cast<void*>(c), n, elem_bytes, scalar, 0, 0, cfg, nullptr, stream);
}
int b200_launch_add(const void* a, const void* b, void* c, size_t n, int elem_bytes, const b200_launch_cfg* cfg,
                    void* stream) {
  return launch_ew(B200_ADD, const_cast<void*>(a), const_cast<void*>(b), c, n, elem_bytes, 0, 0, 0, cfg, nullptr, stream);
}
int b200_launch_triad(void* a, const void* b, const void* c, size_t n, int elem_bytes, double scalar,
                      const b200_launch_cfg* cfg, void* stream) {
  return launch_ew(B200_TRIAD, a, const_cast<void*>(b), const_cast<void*>(c), n, elem_bytes, scalar, 0, 0, cfg, nullptr, stream);
}
int b200_launch_nstream(void* a, const void* b, const void* c, size_t n, int elem_bytes, double scalar,
                        const b200_launch_cfg* cfg, void* stream) {
  return launch_ew(B200_NSTREAM, a, const_cast<void*>(b), const_cast<void*>(c), n, elem_bytes, scalar, 0, 0, cfg, nullptr, stream);
}

size_t b200_dot_workspace_bytes(void) { return kDotWorkspaceBytes; }

int b200_launch_dot(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* result_dev,
                    void* result_host_mapped, const b200_launch_cfg* cfg, void* stream) {
  return launch_dot_common(a, b, n, elem_bytes, workspace, result_dev, result_host_mapped, 0, 0, 0, nullptr, cfg, nullptr, stream);
}

int b200_launch_doorbell(int kernel, void* a, void* b, void* c, size_t n, int elem_bytes, double s0, double s1, double s2,
                         const b200_launch_cfg* cfg, const b200_doorbell* bell, void* stream) {
  if (kernel == B200_DOT || kernel < 0 || kernel >= B200_NUM_KERNELS) {
    set_error("b200_launch_doorbell: kernel id %d is not an elementwise kernel", kernel);
    return B200_EINVAL;
  }
  if (!bell) { set_error("b200_launch_doorbell: null doorbell"); return B200_EINVAL; }
  return launch_ew(kernel, a, b, c, n, elem_bytes, s0, s1, s2, cfg, bell, stream);
}

int b200_launch_dot_doorbell(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* result_dev,
                             void* result_host_mapped, const b200_launch_cfg* cfg, const b200_doorbell* bell,
                             void* stream) {
  if (!bell) { set_error("b200_launch_dot_doorbell: null doorbell"); return B200_EINVAL; }
  return launch_dot_common(a, b, n, elem_bytes, workspace, result_dev, result_host_mapped, 0, 0, 0, nullptr, cfg, bell, stream);
}

size_t b200_dot_mailbox_bytes(void) { return kMailboxBytes; }

int b200_launch_dot_allreduce_p2p(const void* a, const void* b, size_t n, int elem_bytes, void* workspace,
                                  void* result_dev, void* result_host_mapped, int world, int rank,
                                  unsigned long long seq, void* const* mailboxes, const b200_launch_cfg* cfg,
                                  void* stream) {
  if (world < 1 || world > kMaxPeers) { set_error("dot p2p: world must be 1..%d", kMaxPeers); return B200_EINVAL; }
  return launch_dot_common(a, b, n, elem_bytes, workspace, result_dev, result_host_mapped, world, rank, seq, mailboxes,
                           cfg, nullptr, stream);
}

int b200_launch_dot_send_p2p(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* partial_dev,
                             int world, int rank, unsigned long long seq, void* const* mailboxes,
                             const b200_launch_cfg* cfg, void* stream) {
  if (world < 2 || world > kMaxPeers) { set_error("dot send: world must be 2..%d", kMaxPeers); return B200_EINVAL; }
  return launch_dot_common(a, b, n, elem_bytes, workspace, partial_dev, nullptr, world, rank, seq, mailboxes, cfg, nullptr,
                           stream, /*send_only=*/true);
}

int b200_launch_dot_collect_p2p(int elem_bytes, void* result_dev, void* result_host_mapped, int world, int rank,
                                unsigned long long seq, void* const* mailboxes, void* workspace, int flags, void* stream) {
  if (int rc = check_common(0, elem_bytes)) return rc;
  if (!current_device_props(nullptr)) return B200_ENODEVICE;
  void* status = workspace ? static_cast<char*>(workspace) + kDotWsStatus : nullptr;
  return launch_dot_collect(elem_bytes, result_dev, result_host_mapped, world, rank, seq, mailboxes, status, flags,
                            static_cast<cudaStream_t>(stream));
}

size_t b200_dot_status_offset(void) { return kDotWsStatus; }

int b200_set_p2p_timeout_ms(double ms) {
  if (!(ms > 0)) { set_error("b200_set_p2p_timeout_ms: the timeout must be positive"); return B200_EINVAL; }
  set_p2p_timeout_ns((unsigned long long)(ms * 1e6));
  return B200_OK;
}

// ---- cross-process peer memory (one process per GPU) ------------------------------------------
int b200_dot_mailbox_create(void** mailbox) {
  if (!mailbox) { set_error("b200_dot_mailbox_create: null argument"); return B200_EINVAL; }
  if (!current_device_props(nullptr)) return B200_ENODEVICE;
  void* p = nullptr;
  B200_CUDA(cudaMalloc(&p, kMailboxBytes));  // its own allocation: cudaIpcGetMemHandle exports whole allocations
  cudaError_t e = cudaMemset(p, 0, kMailboxBytes);
  if (e == cudaSuccess) e = cudaDeviceSynchronize();
  if (e != cudaSuccess) { cudaFree(p); return cuda_fail(e, "cudaMemset(mailbox)", __FILE__, __LINE__); }
  *mailbox = p;
  return B200_OK;
}

int b200_dot_mailbox_destroy(void* mailbox) {
  if (mailbox) B200_CUDA(cudaFree(mailbox));
  return B200_OK;
}

int b200_ipc_export(void* dev_ptr, void* handle_out) {
  static_assert(sizeof(cudaIpcMemHandle_t) == B200_IPC_HANDLE_BYTES, "handle size is part of the ABI");
  if (!dev_ptr || !handle_out) { set_error("b200_ipc_export: null argument"); return B200_EINVAL; }
  cudaIpcMemHandle_t h;
  B200_CUDA(cudaIpcGetMemHandle(&h, dev_ptr));
  memcpy(handle_out, &h, sizeof h);
  return B200_OK;
}

int b200_ipc_open(const void* handle, void** dev_ptr) {
  if (!handle || !dev_ptr) { set_error("b200_ipc_open: null argument"); return B200_EINVAL; }
  cudaIpcMemHandle_t h;
  memcpy(&h, handle, sizeof h);
  void* p = nullptr;
  B200_CUDA(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
  *dev_ptr = p;
  return B200_OK;
}

int b200_ipc_close(void* dev_ptr) {
  if (dev_ptr) B200_CUDA(cudaIpcCloseMemHandle(dev_ptr));
  return B200_OK;
}

}  // extern "C"
