// context.cu -- LEVEL 2 of include/b200_stream.h: the Stream<T> object.
//
// Three device-resident arrays of n elements, sharded contiguously over G GPUs of one box
// (b200_shard_range).  Every public call is blocking, as src/Stream.h:36 demands.
//   G == 1 : the caller's thread launches and waits; no helper threads at all.
//   G  > 1 : the caller's thread drives GPU 0 itself and G-1 worker threads drive the others, released by a
//            shared sequence number and joined through ONE completion counter, so that the G launches start
//            together instead of one after another from a single thread (SURVEY.md section 7 "launch skew").
//            Each worker is pinned to its own core among the CPUs local to its GPU (option pin=off: not).
// Every kernel call replays a one-node CUDA graph that was captured the first time the call was made
// with these arguments (a "plan"): the arrays, the shape and the scalar never change between the calls
// of a benchmark run, and cudaGraphLaunch + cudaStreamSynchronize returns 2.0-2.5 us sooner than
// <<<>>> + cudaStreamSynchronize on B200 at every size (profiles/r02/lab_path.txt).  dot's per-call
// sequence numbers live in device memory for the same reason.
// The elementwise kernels never communicate.  dot is the one exchange step: each GPU reduces its
// shard to one T, and the G values are combined inside the dot kernel by peer stores over NVLink
// by a rank-order sum of G mapped host scalars whose doorbells the caller polls (dot_combine=host, the
// default inside one process), inside the dot kernel by peer stores over NVLink (p2p: what process groups
// use, b200_join_group) or by ONE ncclAllReduce of count 1 (nccl).
#include <ctype.h>
#include <errno.h>
#include <limits.h>
#include <dlfcn.h>
#include <sched.h>
#include <nccl.h>  // types only: the library is dlopen'ed, so single-GPU use needs no NCCL at all

#include <atomic>
#include <chrono>
#include <condition_variable>
#include <mutex>
#include <new>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <string>
#include <thread>
#include <vector>

#if defined(__x86_64__)
#include <immintrin.h>
#define B200_PAUSE() _mm_pause()
#else
#define B200_PAUSE() do { } while (0)
#endif

#include "internal.h"
#include "kernels.cuh"

namespace b200 {

// ---------------------------------------------------------------------------------------------
// NCCL, resolved at run time
// ---------------------------------------------------------------------------------------------
struct NcclApi {
  void* handle = nullptr;
  ncclResult_t (*CommInitAll)(ncclComm_t*, int, const int*) = nullptr;
  ncclResult_t (*CommDestroy)(ncclComm_t) = nullptr;
  ncclResult_t (*AllReduce)(const void*, void*, size_t, ncclDataType_t, ncclRedOp_t, ncclComm_t, cudaStream_t) = nullptr;
  const char* (*GetErrorString)(ncclResult_t) = nullptr;
  ncclResult_t (*GetVersion)(int*) = nullptr;

  int load() {
    if (handle) return B200_OK;
    const char* env = getenv("BABELSTREAM_B200_NCCL_LIB");
    const char* names[] = {env, "libnccl.so.2", "libnccl.so"};
    for (const char* nm : names) {
      if (!nm || !*nm) continue;
      handle = dlopen(nm, RTLD_NOW | RTLD_GLOBAL);
      if (handle) break;
    }
    if (!handle) { set_error("cannot load NCCL (libnccl.so.2): %s", dlerror()); return B200_ENCCL; }
    CommInitAll = (decltype(CommInitAll))dlsym(handle, "ncclCommInitAll");
    CommDestroy = (decltype(CommDestroy))dlsym(handle, "ncclCommDestroy");
    AllReduce = (decltype(AllReduce))dlsym(handle, "ncclAllReduce");
    GetErrorString = (decltype(GetErrorString))dlsym(handle, "ncclGetErrorString");
    GetVersion = (decltype(GetVersion))dlsym(handle, "ncclGetVersion");
    if (!CommInitAll || !CommDestroy || !AllReduce || !GetErrorString) {
      set_error("libnccl is missing a required symbol");
      return B200_ENCCL;
    }
    return B200_OK;
  }
};

enum { CMD_NONE = 0, CMD_KERNEL, CMD_READ, CMD_WRITE, CMD_PREFETCH, CMD_QUIT };
enum { COMBINE_AUTO = -1, COMBINE_NCCL = 0, COMBINE_HOST = 1, COMBINE_P2P = 2 };
enum { SYNC_AUTO = -1, SYNC_STREAM = 0, SYNC_SPIN = 1, SYNC_DOORBELL = 2 };

struct Shard {
  int dev = 0;
  int pin_cpu = -1;              // CPU the shard's worker thread is pinned to (-1: unpinned)
  size_t begin = 0, count = 0;
  cudaStream_t stream = nullptr;
  void *a = nullptr, *b = nullptr, *c = nullptr;
  void* dot_ws = nullptr;        // partials + ticket
  void* dot_dev = nullptr;       // 2 x 8 bytes: [0] this GPU's partial, [1] all-reduced value
  void* dot_host = nullptr;      // mapped pinned, 64 bytes: [0,8) dot's sum, [8,16) dot's doorbell flag,
                                 // [16,24) the elementwise kernels' doorbell flag, [32,36) all-reduce timeout status
  void* dot_host_dev = nullptr;  // device-side alias of dot_host
  void* bell_ticket = nullptr;   // device word of the elementwise kernels' doorbell
  unsigned long long bell_seq = 0;      // last value the elementwise doorbell was asked to ring with
  unsigned long long dot_deferred = 0;  // dot_combine=host: doorbell value the CALLER still has to wait for (0: none)
  void* seq_counters = nullptr;  // device, 2 x 8 bytes: dot's doorbell value and all-reduce sequence number (bumped by the kernel)
  unsigned long long dot_bell_seq = 0;  // host mirror of seq_counters[0]
  struct Plan {                  // one captured launch, replayed while its key matches
    cudaGraph_t graph = nullptr;
    cudaGraphExec_t exec = nullptr;
    double s0 = 0, s1 = 0, s2 = 0;
    uint64_t gen = 0;            // ctx->tune_gen it was built under
    unsigned nlaunch = 0;        // kernel nodes (for b200_launch_count)
    bool bell = false;           // dot: built with the doorbell
    uint64_t timeout_gen = 0;    // p2p_timeout_gen() it was built under: the wait bound is a kernel parameter
  } plans[B200_NUM_KERNELS][2];  // [kernel][sweep direction]
  void* mailbox = nullptr;       // p2p combine: kMaxPeers slots in this GPU's HBM, written by the peers
  int rank = 0;
  cudaEvent_t ev0 = nullptr, ev1 = nullptr;
  ncclComm_t comm = nullptr;
  float last_ms = -1.f;
  int status = B200_OK;
  std::string error;
};

struct Command {
  int kind = CMD_NONE;
  int kernel = 0;
  double s0 = 0, s1 = 0, s2 = 0;
  int reverse = 0;                              // kernel: sweep the shard from the top down
  void* host[3] = {nullptr, nullptr, nullptr};  // read: destinations; write: sources
  int prefetch_to = 0;                          // CMD_PREFETCH: 0 = the shard's GPU, 1 = the CPU
};

}  // namespace b200

using namespace b200;

struct b200_ctx {
  int first_device = 0, ngpus = 1;
  size_t n = 0;
  int elem_bytes = 8;
  std::vector<Shard*> shards;

  // options
  b200_launch_cfg tune[B200_NUM_KERNELS];
  bool timing = false;
  bool created = false;  // b200_create finished: the allocation-time options are frozen
  int sync_mode = SYNC_AUTO;
  std::vector<int> claimed_cpus;  // CPUs this context's workers were given (process-wide registry, see choose_worker_cpus)
  int caller_device = -1;    // the device that was current when b200_create was called; restored by create and destroy
  bool use_graphs = true;    // option launch=graph|direct
  bool pin = true;           // option pin=auto|off: pin every worker thread to a core of its own near its GPU
  bool alternate = true;     // option sweep=alternate|forward: consecutive kernels sweep in opposite directions
  int next_reverse = 0;      // direction of the next kernel call
  uint64_t tune_gen = 1;     // bumped by every option change: invalidates the cached plans
  // one process per GPU (b200_join_group): this context is rank group_rank of group_world
  int group_world = 1, group_rank = 0;
  void* group_mailboxes[kMaxPeers] = {};
  int combine = COMBINE_AUTO;  // resolved by ensure_combine_ready: p2p if every pair has peer access, else nccl
  NcclApi nccl;
  bool nccl_ready = false;
  bool p2p_ready = false;
  void* mailboxes[kMaxPeers] = {};

  // mem=managed: the reference's MEM=MANAGED build (src/cuda/model.cmake:6-10, CUDAStream.cu:37-38,217-221).
  // One cudaMallocManaged range per array; every shard is a sub-range whose preferred location is its
  // GPU.  get_arrays hands out the ranges themselves (no gather, no second copy of the data).
  bool managed = false;
  void* managed_base[3] = {nullptr, nullptr, nullptr};
  bool managed_on_host = false;  // get_arrays moved the pages to the CPU; move them back before the next kernel
  // mem=pagefault: the reference's MEM=PAGEFAULT build (src/cuda/model.cmake:6-10, CUDAStream.cu:39-40,217-221):
  // the arrays are plain system allocations that the GPUs reach through the page-fault path (HMM or ATS).
  // Implies `managed` (one range per array, get_arrays returns the ranges); advice and prefetches are
  // hints only there and a driver that refuses them for system memory is not an error.
  bool pagefault = false;

  // host mirrors for get_arrays
  void* host[3] = {nullptr, nullptr, nullptr};
  bool host_pinned = false;

  // opt-in host-side profile of the blocking path (BABELSTREAM_B200_HOSTPROF=1): ns spent in
  // cudaSetDevice / enqueue / wait, printed by b200_destroy
  bool hostprof = false;
  uint64_t prof_ns[3] = {0, 0, 0};
  uint64_t prof_calls = 0;

  // workers (ngpus > 1)
  std::vector<std::thread> workers;
  Command cmd;
  std::atomic<uint64_t> seq{0};
  std::atomic<uint64_t> done_total{0};  // every worker adds 1 per finished command
  std::atomic<int> sleepers{0};
  std::mutex mu;
  std::condition_variable cv;
};

namespace b200 {

static int stream_wait(b200_ctx* ctx, Shard* s) {
  if (ctx->sync_mode == SYNC_SPIN) {
    for (;;) {
      cudaError_t e = cudaStreamQuery(s->stream);
      if (e == cudaSuccess) return B200_OK;
      if (e != cudaErrorNotReady) return cuda_fail(e, "cudaStreamQuery(stream)", __FILE__, __LINE__);
      B200_PAUSE();
    }
  }
  B200_CUDA(cudaStreamSynchronize(s->stream));
  return B200_OK;
}

// Wait for the doorbell of the launch just issued (kernels.cuh): spin on the host word the kernel's
// last block writes.  Every 64 Ki polls the stream is queried, so that a failed launch or a faulting
// kernel ends in an error instead of a hang.
static int bell_wait(Shard* s, unsigned long long seq, int flag_offset) {
  const unsigned long long* flag = reinterpret_cast<const unsigned long long*>(static_cast<char*>(s->dot_host) + flag_offset);
  for (unsigned spins = 1;; ++spins) {
    if (__atomic_load_n(flag, __ATOMIC_ACQUIRE) == seq) return B200_OK;
    if ((spins & 0xffffu) == 0) {
      const cudaError_t e = cudaStreamQuery(s->stream);
      if (e == cudaSuccess) {
        if (__atomic_load_n(flag, __ATOMIC_ACQUIRE) == seq) return B200_OK;
        set_error("kernel finished without ringing its doorbell (seq %llu)", seq);
        return B200_ECUDA;
      }
      if (e != cudaErrorNotReady) return cuda_fail(e, "cudaStreamQuery(stream)", __FILE__, __LINE__);
    }
    B200_PAUSE();
  }
}

// Does this call wait on a doorbell?  auto: dot only.  Measured on B200 (profiles/r01/wait_modes.md):
// dot's blocking call gets 1.3-1.9 us shorter at every size, because the host no longer waits for the
// end-of-kernel flush of dot's own write into host memory; the ticket is one dot takes anyway.  The
// elementwise kernels have no such write, the ring saves them 0.2 us at best, and the per-block ticket
// costs 1-3 us as soon as there are ~1000 blocks (same-address atomics drain at ~3 ns each) -- so they
// ring only when asked to (sync=doorbell).
static bool use_doorbell(const b200_ctx* ctx, const Shard* s, int kernel) {
  if (ctx->timing || s->count == 0) return false;
  if (kernel == B200_DOT && ctx->ngpus > 1 && ctx->combine == COMBINE_NCCL) return false;  // the all-reduce follows
  if (ctx->sync_mode == SYNC_DOORBELL) return true;
  return ctx->sync_mode == SYNC_AUTO && kernel == B200_DOT;
}

// Launch one kernel on one shard (device already current).  No synchronisation.
static int shard_launch(b200_ctx* ctx, Shard* s, const Command& cmd, const b200_doorbell* bell) {
  const int eb = ctx->elem_bytes;
  b200_launch_cfg cfg_store = ctx->tune[cmd.kernel];
  if (cmd.reverse) cfg_store.flags |= B200_LAUNCH_REVERSE;
  const b200_launch_cfg* cfg = &cfg_store;
  void* st = s->stream;
  if (cmd.kernel == B200_DOT) {
    const bool single = ctx->ngpus == 1;
    void* const status = static_cast<char*>(s->dot_host_dev) + 32;
    if (single && ctx->group_world > 1) {
      // one process per GPU: same fused kernel, the peers' mailboxes are IPC mappings
      return launch_dot_common(s->a, s->b, s->count, eb, s->dot_ws, s->dot_dev, s->dot_host_dev, ctx->group_world,
                               ctx->group_rank, 0, ctx->group_mailboxes, cfg, bell, st, false, status, s->seq_counters);
    }
    if (!single && ctx->combine == COMBINE_P2P) {
      // fused: partial + all-reduce over peer memory in ONE launch; an empty shard still takes part
      return launch_dot_common(s->a, s->b, s->count, eb, s->dot_ws, s->dot_dev, s->dot_host_dev, ctx->ngpus, s->rank,
                               0, ctx->mailboxes, cfg, bell, st, false, status, s->seq_counters);
    }
    // single GPU: the kernel writes the sum straight into mapped host memory (zero-copy)
    if (s->count == 0) {
      B200_CUDA(cudaMemsetAsync(s->dot_dev, 0, 8, s->stream));
      if (single || ctx->combine == COMBINE_HOST) memset(s->dot_host, 0, 8);
      return B200_OK;
    }
    const bool to_host = single || ctx->combine == COMBINE_HOST;
    return launch_dot_common(s->a, s->b, s->count, eb, s->dot_ws, s->dot_dev, to_host ? s->dot_host_dev : nullptr, 0, 0, 0,
                             nullptr, cfg, bell, st, false, nullptr, s->seq_counters);
  }
  if (s->count == 0) return B200_OK;
  if (cmd.kernel < 0 || cmd.kernel >= B200_NUM_KERNELS) { set_error("unknown kernel id %d", cmd.kernel); return B200_EINVAL; }
  // launch_ew ignores the arrays a kernel does not touch
  return launch_ew(cmd.kernel, s->a, s->b, s->c, s->count, eb, cmd.s0, cmd.s1, cmd.s2, cfg, bell, st);
}

// Is the fused peer-memory all-reduce part of this context's dot?
static bool dot_is_p2p(const b200_ctx* ctx) {
  return (ctx->ngpus == 1 && ctx->group_world > 1) || (ctx->ngpus > 1 && ctx->combine == COMBINE_P2P);
}

// Capture the launch of `cmd` on shard `s` into a one-node graph (a plan).  Nothing is enqueued.
static int build_plan(b200_ctx* ctx, Shard* s, const Command& cmd, const b200_doorbell* bell, Shard::Plan* p) {
  if (p->exec) { cudaGraphExecDestroy(p->exec); p->exec = nullptr; }
  if (p->graph) { cudaGraphDestroy(p->graph); p->graph = nullptr; }
  const uint64_t before = thread_launch_count();
  B200_CUDA(cudaStreamBeginCapture(s->stream, cudaStreamCaptureModeThreadLocal));
  const int rc = shard_launch(ctx, s, cmd, bell);
  cudaGraph_t graph = nullptr;
  const cudaError_t e = cudaStreamEndCapture(s->stream, &graph);
  const unsigned n = (unsigned)(thread_launch_count() - before);
  uncount_launch(n);  // a capture launches nothing; every replay is counted instead
  if (rc != B200_OK) { if (graph) cudaGraphDestroy(graph); (void)cudaGetLastError(); return rc; }
  B200_CUDA(e);
  if (n == 0) { cudaGraphDestroy(graph); set_error("internal: plan without a kernel"); return B200_ECUDA; }
  const cudaError_t ie = cudaGraphInstantiate(&p->exec, graph, 0);
  if (ie != cudaSuccess) { cudaGraphDestroy(graph); p->exec = nullptr; return cuda_fail(ie, "cudaGraphInstantiate", __FILE__, __LINE__); }
  p->graph = graph;
  p->s0 = cmd.s0; p->s1 = cmd.s1; p->s2 = cmd.s2;
  p->gen = ctx->tune_gen;
  p->timeout_gen = p2p_timeout_gen();
  p->nlaunch = n;
  p->bell = bell != nullptr;
  return B200_OK;
}

// Everything one GPU does for one command, including the wait.  Runs on the caller's thread for
// shard 0 and on the shard's worker thread for the others.
static int shard_execute(b200_ctx* ctx, int g, const Command& cmd) {
  Shard* s = ctx->shards[g];
  const size_t eb = (size_t)ctx->elem_bytes;
  using clk = std::chrono::steady_clock;
  clk::time_point t0, t1, t2;
  if (ctx->hostprof) t0 = clk::now();
  // Shard 0 runs on the caller's thread: make the shard's device current for the call and hand the caller's
  // device back afterwards (the workers of shards 1..G-1 own their device for life)
  struct DeviceScope {
    int saved = -1, mine = -1;
    ~DeviceScope() { if (saved >= 0 && saved != mine) cudaSetDevice(saved); }
  } scope;
  if (g == 0) {
    scope.mine = s->dev;
    if (cudaGetDevice(&scope.saved) != cudaSuccess) { (void)cudaGetLastError(); scope.saved = -1; }
    if (scope.saved != s->dev) B200_CUDA(cudaSetDevice(s->dev));
  }
  if (cmd.kind == CMD_KERNEL) {
    if (ctx->hostprof) t1 = clk::now();
    const bool is_dot = cmd.kernel == B200_DOT;
    const bool nccl_dot = is_dot && ctx->ngpus > 1 && ctx->combine == COMBINE_NCCL;
    const bool launches = is_dot ? (s->count > 0 || dot_is_p2p(ctx)) : s->count > 0;
    if (ctx->timing) B200_CUDA(cudaEventRecord(s->ev0, s->stream));
    b200_doorbell bell_store, *bell = nullptr;
    unsigned long long want = 0;
    int flag_off = 8;
    // the kernel bumps seq_counters[0] on every launch that RUNS: the host copy follows only once the launch went through
    if (is_dot && launches) want = s->dot_bell_seq + 1;
    if (use_doorbell(ctx, s, cmd.kernel)) {
      if (is_dot) {  // value: the device-side counter
        bell_store.ticket = nullptr;
        bell_store.flag_host_mapped = static_cast<char*>(s->dot_host_dev) + 8;
        bell_store.seq = 0;
      } else {
        bell_store.ticket = s->bell_ticket;
        bell_store.flag_host_mapped = static_cast<char*>(s->dot_host_dev) + 16;
        bell_store.seq = want = ++s->bell_seq;
        flag_off = 16;
      }
      bell = &bell_store;
    }
    // Replay a plan whenever nothing in the parameter block changes from call to call.
    const bool replay = ctx->use_graphs && launches && !ctx->timing && !nccl_dot && !(bell && !is_dot);
    if (replay) {
      Shard::Plan& p = s->plans[cmd.kernel][cmd.reverse ? 1 : 0];
      if (!p.exec || p.gen != ctx->tune_gen || (is_dot && p.timeout_gen != p2p_timeout_gen()) || p.s0 != cmd.s0 || p.s1 != cmd.s1 || p.s2 != cmd.s2 || p.bell != (bell != nullptr)) {
        if (int rc = build_plan(ctx, s, cmd, bell, &p)) return rc;
      }
      B200_CUDA(cudaGraphLaunch(p.exec, s->stream));
      count_launch(p.nlaunch);
    } else if (int rc = shard_launch(ctx, s, cmd, bell)) {
      return rc;
    }
    if (is_dot && launches) s->dot_bell_seq = want;
    if (ctx->hostprof) t2 = clk::now();
    if (nccl_dot) {
      // the one collective of the whole benchmark: G partial sums -> every GPU holds the total
      char* buf = static_cast<char*>(s->dot_dev);
      ncclResult_t r = ctx->nccl.AllReduce(buf, buf + 8, 1, eb == 8 ? ncclDouble : ncclFloat, ncclSum, s->comm, s->stream);
      if (r != ncclSuccess) { set_error("ncclAllReduce failed: %s", ctx->nccl.GetErrorString(r)); return B200_ENCCL; }
      if (g == 0) B200_CUDA(cudaMemcpyAsync(s->dot_host, buf + 8, eb, cudaMemcpyDeviceToHost, s->stream));
    }
    if (ctx->timing) B200_CUDA(cudaEventRecord(s->ev1, s->stream));
    // In-process fused all-reduce: every GPU ends up with the same bits and the caller reads GPU 0's copy, so
    // only worker 0 has to see the kernel finish; stream order protects whatever the others enqueue next.
    // dot_combine=host: the caller polls every GPU's doorbell itself (b200_dot), so a worker's own wake-up and the
    // join are not on the path of the last partial to arrive either.
    const bool host_defer = is_dot && g != 0 && ctx->ngpus > 1 && ctx->combine == COMBINE_HOST && bell && !ctx->timing;
    s->dot_deferred = host_defer ? want : 0;
    const bool skip_wait = host_defer || (is_dot && g != 0 && ctx->ngpus > 1 && ctx->combine == COMBINE_P2P && !ctx->timing);
    if (!skip_wait) {
      if (int rc = bell ? bell_wait(s, want, flag_off) : stream_wait(ctx, s)) return rc;
    }
    if (ctx->hostprof && g == 0) {
      const clk::time_point t3 = clk::now();
      ctx->prof_ns[0] += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(t1 - t0).count();
      ctx->prof_ns[1] += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(t2 - t1).count();
      ctx->prof_ns[2] += (uint64_t)std::chrono::duration_cast<std::chrono::nanoseconds>(t3 - t2).count();
      ctx->prof_calls++;
    }
    if (ctx->timing) B200_CUDA(cudaEventElapsedTime(&s->last_ms, s->ev0, s->ev1));
    if (is_dot && !skip_wait && dot_is_p2p(ctx)) {
      volatile unsigned int* status = reinterpret_cast<volatile unsigned int*>(static_cast<char*>(s->dot_host) + 32);
      if (*status) {
        *status = 0;
        set_error("dot: the peer-memory all-reduce timed out after %.0f ms waiting for another rank's partial sum "
                  "(rank %d of %d); the result is NaN", (double)p2p_timeout_ns() * 1e-6,
                  ctx->ngpus > 1 ? s->rank : ctx->group_rank, ctx->ngpus > 1 ? ctx->ngpus : ctx->group_world);
        return B200_ETIMEOUT;
      }
    }
    return B200_OK;
  }
  if (cmd.kind == CMD_READ || cmd.kind == CMD_WRITE) {
    void* dev[3] = {s->a, s->b, s->c};
    for (int k = 0; k < 3; k++) {
      if (!cmd.host[k] || s->count == 0) continue;
      char* h = static_cast<char*>(cmd.host[k]) + s->begin * eb;
      // cudaMemcpyDefault: the device side is a managed range under mem=managed
      if (cmd.kind == CMD_READ) B200_CUDA(cudaMemcpyAsync(h, dev[k], s->count * eb, cudaMemcpyDefault, s->stream));
      else B200_CUDA(cudaMemcpyAsync(dev[k], h, s->count * eb, cudaMemcpyDefault, s->stream));
    }
    B200_CUDA(cudaStreamSynchronize(s->stream));
    return B200_OK;
  }
  if (cmd.kind == CMD_PREFETCH) {
    void* dev[3] = {s->a, s->b, s->c};
    const int dst = cmd.prefetch_to ? cudaCpuDeviceId : s->dev;
    for (int k = 0; k < 3 && s->count; k++) {
      const cudaError_t e = cudaMemPrefetchAsync(dev[k], s->count * eb, dst, s->stream);
      if (e != cudaSuccess && ctx->pagefault) { (void)cudaGetLastError(); break; }  // a hint there: the faults move the pages
      B200_CUDA(e);
    }
    B200_CUDA(cudaStreamSynchronize(s->stream));
    return B200_OK;
  }
  return B200_OK;
}

// One CPU per worker thread: among the CPUs this process may run on, prefer those local to the worker's GPU
// (/sys/bus/pci/devices/<bdf>/local_cpulist) and a physical core no other worker -- and not the caller, who
// drives GPU 0 from whatever core it is on -- has been given.  Workers spin between the calls of a
// benchmark run; two of them on the hyper-threads of one core, or one migrating in the middle of a call,
// shows up directly in the join time.  Anything unreadable: the shard stays unpinned.
static std::vector<int> parse_cpulist(const char* path) {
  std::vector<int> cpus;
  FILE* f = fopen(path, "r");
  if (!f) return cpus;
  char buf[4096];
  if (fgets(buf, sizeof buf, f)) {
    for (char* p = buf; *p && *p != '\n';) {
      char* end;
      const long lo = strtol(p, &end, 10);
      if (end == p) break;
      long hi = lo;
      p = end;
      if (*p == '-') { hi = strtol(p + 1, &end, 10); p = end; }
      for (long c = lo; c <= hi && c < CPU_SETSIZE; c++) cpus.push_back((int)c);
      if (*p == ',') p++;
    }
  }
  fclose(f);
  return cpus;
}

static long core_key(int cpu) {
  char path[128];
  long pkg = 0, core = cpu;
  snprintf(path, sizeof path, "/sys/devices/system/cpu/cpu%d/topology/physical_package_id", cpu);
  if (FILE* f = fopen(path, "r")) { if (fscanf(f, "%ld", &pkg) != 1) pkg = 0; fclose(f); }
  snprintf(path, sizeof path, "/sys/devices/system/cpu/cpu%d/topology/core_id", cpu);
  if (FILE* f = fopen(path, "r")) { if (fscanf(f, "%ld", &core) != 1) core = cpu; fclose(f); }
  return pkg * 100000 + core;
}

// CPUs handed to worker threads by ANY context of this process: two multi-GPU contexts side by side must not stack
// their spinning workers on the same cores while others idle.  Released in b200_destroy.  (Separate PROCESSES cannot
// see each other's choice: give them disjoint taskset masks, or pin=off.)
static std::mutex g_pin_mutex;
static std::vector<int> g_pinned_cpus;

static void release_worker_cpus(b200_ctx* ctx) {
  std::lock_guard<std::mutex> lock(g_pin_mutex);
  for (int c : ctx->claimed_cpus) {
    for (size_t i = 0; i < g_pinned_cpus.size(); i++)
      if (g_pinned_cpus[i] == c) { g_pinned_cpus.erase(g_pinned_cpus.begin() + (long)i); break; }
  }
  ctx->claimed_cpus.clear();
}

static void choose_worker_cpus(b200_ctx* ctx) {
  std::lock_guard<std::mutex> lock(g_pin_mutex);
  cpu_set_t allowed;
  CPU_ZERO(&allowed);
  if (sched_getaffinity(0, sizeof allowed, &allowed) != 0) return;
  if (CPU_COUNT(&allowed) < ctx->ngpus) return;  // fewer CPUs than threads: pinning would only stack them
  std::vector<long> used_cores;
  std::vector<int> used_cpus;
  const int here = sched_getcpu();
  if (here >= 0) { used_cores.push_back(core_key(here)); used_cpus.push_back(here); }
  for (int c : g_pinned_cpus) { used_cores.push_back(core_key(c)); used_cpus.push_back(c); }  // other contexts' workers
  auto is_in = [](const auto& v, auto x) { for (auto y : v) if (y == x) return true; return false; };
  for (int g = 1; g < ctx->ngpus; g++) {
    Shard* s = ctx->shards[g];
    std::vector<int> local;
    char bdf[32] = {0};
    if (cudaDeviceGetPCIBusId(bdf, sizeof bdf, s->dev) == cudaSuccess) {
      for (char* p = bdf; *p; p++) *p = (char)tolower(*p);
      char path[128];
      snprintf(path, sizeof path, "/sys/bus/pci/devices/%s/local_cpulist", bdf);
      local = parse_cpulist(path);
    } else {
      (void)cudaGetLastError();
    }
    std::vector<int> everywhere;
    for (int c = 0; c < CPU_SETSIZE; c++) if (CPU_ISSET(c, &allowed)) everywhere.push_back(c);
    int pick = -1;
    // 1. local CPU on an untouched core  2. any CPU on an untouched core  3. a local, then any, untouched hyper-thread
    for (int pass = 0; pass < 4 && pick < 0; pass++) {
      const std::vector<int>& pool = (pass == 0 || pass == 2) ? local : everywhere;
      for (int c : pool) {
        if (!CPU_ISSET(c, &allowed) || is_in(used_cpus, c)) continue;
        if (pass < 2 && is_in(used_cores, core_key(c))) continue;
        pick = c;
        break;
      }
    }
    s->pin_cpu = pick;
    if (pick >= 0) {
      used_cpus.push_back(pick); used_cores.push_back(core_key(pick));
      g_pinned_cpus.push_back(pick); ctx->claimed_cpus.push_back(pick);
    }
  }
}

static void worker_main(b200_ctx* ctx, int g) {
  Shard* s = ctx->shards[g];
  cudaSetDevice(s->dev);
  if (s->pin_cpu >= 0) {
    cpu_set_t set;
    CPU_ZERO(&set);
    CPU_SET(s->pin_cpu, &set);
    if (sched_setaffinity(0, sizeof set, &set) != 0) s->pin_cpu = -1;  // not allowed here: run unpinned
  }
  uint64_t seen = 0;
  for (;;) {
    // spin briefly (the benchmark issues calls back to back), then sleep on the condition variable
    int spins = 0;
    while (ctx->seq.load(std::memory_order_acquire) == seen) {
      if (++spins < (1 << 16)) { B200_PAUSE(); continue; }
      std::unique_lock<std::mutex> lock(ctx->mu);
      ctx->sleepers.fetch_add(1);
      ctx->cv.wait(lock, [&] { return ctx->seq.load() != seen; });
      ctx->sleepers.fetch_sub(1);
    }
    seen++;
    const Command cmd = ctx->cmd;  // published before seq was bumped
    if (cmd.kind == CMD_QUIT) { ctx->done_total.fetch_add(1, std::memory_order_release); return; }
    const int rc = shard_execute(ctx, g, cmd);
    s->status = rc;
    if (rc != B200_OK) s->error = b200_last_error();
    ctx->done_total.fetch_add(1, std::memory_order_release);
  }
}

// Run one command on every shard and wait for all of them.
static int run_all(b200_ctx* ctx, const Command& cmd) {
  if (ctx->ngpus == 1) return shard_execute(ctx, 0, cmd);
  ctx->cmd = cmd;
  const uint64_t target = ctx->seq.fetch_add(1, std::memory_order_seq_cst) + 1;
  if (ctx->sleepers.load(std::memory_order_seq_cst) > 0) {
    { std::lock_guard<std::mutex> lock(ctx->mu); }
    ctx->cv.notify_all();
  }
  // GPU 0 is driven from here: no hand-over on the path of the shard whose result the caller reads
  Shard* s0 = ctx->shards[0];
  s0->status = shard_execute(ctx, 0, cmd);
  if (s0->status != B200_OK) s0->error = b200_last_error();
  // join: ONE counter that every worker bumps once per command (not G flags polled in series)
  const uint64_t all_done = target * (uint64_t)(ctx->ngpus - 1);
  unsigned spins = 0;
  while (ctx->done_total.load(std::memory_order_acquire) < all_done) {
    if (++spins < (1u << 22)) B200_PAUSE();
    else std::this_thread::yield();
  }
  int rc = B200_OK;
  for (Shard* s : ctx->shards)
    if (s->status != B200_OK && rc == B200_OK) { rc = s->status; set_error("GPU %d: %s", s->dev, s->error.c_str()); }
  return rc;
}

static int parse_cfg_fields(b200_launch_cfg* c, const std::string& fields) {
  // "vec=32,unroll=2,block=256,grid=-4,hint=3,variant=2,stages=4,chunk=32768,contract=1"
  // Strict: an unknown field name or a value that is not a whole number is an error, not a silent "auto".
  size_t pos = 0;
  while (pos < fields.size()) {
    size_t end = fields.find(',', pos);
    if (end == std::string::npos) end = fields.size();
    const std::string kv = fields.substr(pos, end - pos);
    pos = end + 1;
    if (kv.empty()) continue;
    const size_t eq = kv.find('=');
    if (eq == std::string::npos || eq == 0) { set_error("tune: '%s' is not field=value", kv.c_str()); return B200_EINVAL; }
    const std::string k = kv.substr(0, eq);
    const char* text = kv.c_str() + eq + 1;
    char* stop = nullptr;
    errno = 0;
    const long lv = strtol(text, &stop, 10);
    if (stop == text || *stop != '\0' || errno == ERANGE || lv < INT_MIN || lv > INT_MAX) {
      set_error("tune: value of '%s' is not an integer: '%s'", k.c_str(), text);
      return B200_EINVAL;
    }
    const int v = (int)lv;
    if (k == "vec") c->vec_bytes = v; else if (k == "unroll") c->unroll = v; else if (k == "block") c->block = v;
    else if (k == "grid") c->grid = v; else if (k == "hint") c->cache_hint = v; else if (k == "variant") c->variant = v;
    else if (k == "stages") c->stages = v; else if (k == "chunk") c->chunk_bytes = v; else if (k == "contract") c->contract = v;
    else if (k == "keep") c->keep_mb = v; else if (k == "flags") c->flags = v;
    else {
      set_error("tune: unknown field '%s' (fields: vec unroll block grid hint variant stages chunk contract keep flags)", k.c_str());
      return B200_EINVAL;
    }
  }
  return B200_OK;
}

// Does the calling thread have a CUDA context current?  Asked of the driver (cuCtxGetCurrent, fetched through the runtime:
// no link-time dependency on libcuda) because every runtime call that could tell would create one.
static bool thread_has_cuda_context() {
  using GetCurrentFn = int (*)(void**);
  static GetCurrentFn fn = [] {
    void* p = nullptr;
    cudaDriverEntryPointQueryResult q;
    if (cudaGetDriverEntryPoint("cuCtxGetCurrent", &p, cudaEnableDefault, &q) != cudaSuccess || q != cudaDriverEntryPointSuccess) {
      (void)cudaGetLastError();
      p = nullptr;
    }
    return reinterpret_cast<GetCurrentFn>(p);
  }();
  if (!fn) return true;  // cannot tell: behave as before (restore the caller's device)
  void* cur = nullptr;
  return fn(&cur) == 0 && cur != nullptr;
}

static int kernel_by_name(const std::string& s) {
  const char* names[B200_NUM_KERNELS] = {"copy", "mul", "add", "triad", "nstream", "dot", "init"};
  for (int i = 0; i < B200_NUM_KERNELS; i++) if (s == names[i]) return i;
  return -1;
}

static int apply_option(b200_ctx* ctx, const std::string& key, const std::string& val) {
  ctx->tune_gen++;  // whatever changes: the cached plans are rebuilt on their next use
  if (key == "sweep") {
    if (val != "alternate" && val != "forward") { set_error("sweep must be alternate or forward, got '%s'", val.c_str()); return B200_EINVAL; }
    ctx->alternate = val == "alternate";
    ctx->next_reverse = 0;
  } else if (key == "pin") {
    if (val != "auto" && val != "off") { set_error("pin must be auto or off, got '%s'", val.c_str()); return B200_EINVAL; }
    if (ctx->created) { set_error("pin is fixed when the worker threads start: pass it to b200_create_opts or set BABELSTREAM_B200_PIN"); return B200_EINVAL; }
    ctx->pin = val == "auto";
  } else if (key == "launch") {
    if (val != "graph" && val != "direct") { set_error("launch must be graph or direct, got '%s'", val.c_str()); return B200_EINVAL; }
    ctx->use_graphs = val == "graph";
  } else if (key == "contract") {
    const bool off = val == "off" || val == "0" || val == "unfused";
    if (!off && val != "fast" && val != "1" && val != "fused" && val != "on") {
      set_error("contract must be fast (fused, on, 1) or off (unfused, 0), got '%s'", val.c_str()); return B200_EINVAL;
    }
    for (auto& t : ctx->tune) t.contract = off ? 2 : 1;
  } else if (key == "copy_variant") {
    if (val != "auto" && val != "ldst" && val != "bulk" && val != "tma") {
      set_error("copy_variant must be auto, ldst, bulk or tma, got '%s'", val.c_str()); return B200_EINVAL;
    }
    int v = val == "ldst" ? B200_VARIANT_LDST : val == "bulk" ? B200_VARIANT_BULK : val == "tma" ? B200_VARIANT_TMA : B200_VARIANT_AUTO;
    ctx->tune[B200_COPY].variant = v;
  } else if (key == "dot_variant") {
    if (val != "auto" && val != "two_pass" && val != "ticket") {
      set_error("dot_variant must be auto, two_pass or ticket, got '%s'", val.c_str()); return B200_EINVAL;
    }
    ctx->tune[B200_DOT].variant = val == "two_pass" ? B200_VARIANT_DOT_TWO_PASS : val == "ticket" ? B200_VARIANT_DOT_TICKET : 0;
  } else if (key == "dot_combine") {
    if (val != "auto" && val != "host" && val != "p2p" && val != "nccl") {
      set_error("dot_combine must be auto, host, p2p or nccl, got '%s'", val.c_str()); return B200_EINVAL;
    }
    ctx->combine = val == "host" ? COMBINE_HOST : val == "p2p" ? COMBINE_P2P : val == "nccl" ? COMBINE_NCCL : COMBINE_AUTO;
  } else if (key == "timing") {
    const bool on = val == "1" || val == "on" || val == "true";
    if (!on && val != "0" && val != "off" && val != "false") {
      set_error("timing must be 0 (off, false) or 1 (on, true), got '%s'", val.c_str()); return B200_EINVAL;
    }
    ctx->timing = on;
  } else if (key == "mem") {
    const bool pf = val == "pagefault";
    const bool want = val == "managed" || pf;
    if (!want && val != "device" && val != "default") { set_error("mem must be device, managed or pagefault, got '%s'", val.c_str()); return B200_EINVAL; }
    if (ctx->created && (want != ctx->managed || pf != ctx->pagefault)) {
      set_error("mem is fixed when the arrays are allocated: pass it to b200_create_opts or set BABELSTREAM_B200_MEM");
      return B200_EINVAL;
    }
    ctx->managed = want;
    ctx->pagefault = pf;
  } else if (key == "sync") {
    if (val != "auto" && val != "stream" && val != "spin" && val != "doorbell") {
      set_error("sync must be auto, stream, spin or doorbell, got '%s'", val.c_str()); return B200_EINVAL;
    }
    ctx->sync_mode = val == "spin" ? SYNC_SPIN : val == "stream" ? SYNC_STREAM : val == "doorbell" ? SYNC_DOORBELL : SYNC_AUTO;
  } else if (key == "tune") {
    // "<kernel>:<fields>;<kernel>:<fields>"   kernel "all" applies to the six elementwise kernels.
    // Parsed into a copy, checked as a launch would check it, and only then adopted: a bad shape fails HERE, not at
    // the first call that uses it (a failed dot launch used to leave the doorbell count one short for good).
    b200_launch_cfg next[B200_NUM_KERNELS];
    memcpy(next, ctx->tune, sizeof next);
    bool touched[B200_NUM_KERNELS] = {};
    size_t pos = 0;
    while (pos < val.size()) {
      size_t end = val.find(';', pos);
      if (end == std::string::npos) end = val.size();
      const std::string item = val.substr(pos, end - pos);
      pos = end + 1;
      if (item.empty()) continue;
      const size_t colon = item.find(':');
      if (colon == std::string::npos) { set_error("tune: '%s' is not <kernel>:<field>=<v>,...", item.c_str()); return B200_EINVAL; }
      const std::string kn = item.substr(0, colon), fields = item.substr(colon + 1);
      if (kn == "all") {
        for (int k = 0; k < B200_NUM_KERNELS; k++)
          if (k != B200_DOT) { if (int rc = parse_cfg_fields(&next[k], fields)) return rc; touched[k] = true; }
      } else {
        const int k = kernel_by_name(kn);
        if (k < 0) { set_error("tune: unknown kernel '%s'", kn.c_str()); return B200_EINVAL; }
        if (int rc = parse_cfg_fields(&next[k], fields)) return rc;
        touched[k] = true;
      }
    }
    const size_t per_gpu = ctx->ngpus > 0 ? (ctx->n + (size_t)ctx->ngpus - 1) / (size_t)ctx->ngpus : ctx->n;
    for (int k = 0; k < B200_NUM_KERNELS; k++) {
      if (!touched[k]) continue;
      b200_sweep_plan plan;
      if (b200_plan_sweep(k, per_gpu, ctx->elem_bytes, 148, 126 << 20, &next[k], &plan)) {
        const std::string why = b200_last_error();
        const char* names[B200_NUM_KERNELS] = {"copy", "mul", "add", "triad", "nstream", "dot", "init"};
        set_error("tune: %s: %s", names[k], why.c_str());
        return B200_EINVAL;
      }
    }
    memcpy(ctx->tune, next, sizeof next);
  } else {
    set_error("unknown option '%s'", key.c_str());
    return B200_EINVAL;
  }
  return B200_OK;
}

static void free_host_mirrors(b200_ctx* ctx) {
  for (auto& h : ctx->host) {
    if (!h) continue;
    if (ctx->host_pinned) cudaFreeHost(h); else free(h);
    h = nullptr;
  }
}

static int ensure_host_mirrors(b200_ctx* ctx) {
  if (ctx->host[0]) return B200_OK;
  const size_t bytes = ctx->n * (size_t)ctx->elem_bytes;
  // pinned while it is reasonable (D2H at full PCIe speed); beyond the limit fall back to pageable
  // memory like the reference's std::vector (CUDAStream.cu:225-233) instead of pinning > 1/3 of RAM
  double limit_gb = 24.0;
  if (const char* e = getenv("BABELSTREAM_B200_PINNED_LIMIT_GB")) limit_gb = atof(e);
  bool pinned = 3.0 * (double)bytes <= limit_gb * 1e9;
  if (pinned) {
    for (auto& h : ctx->host) {
      if (cudaHostAlloc(&h, bytes ? bytes : 1, cudaHostAllocPortable) != cudaSuccess) { (void)cudaGetLastError(); h = nullptr; pinned = false; break; }
    }
    if (!pinned) { ctx->host_pinned = true; free_host_mirrors(ctx); }
  }
  if (!pinned) {
    for (auto& h : ctx->host) {
      h = aligned_alloc(2 * 1024 * 1024, ((bytes ? bytes : 1) + (2u << 20) - 1) / (2u << 20) * (2u << 20));
      if (!h) { set_error("cannot allocate %zu bytes of host memory for get_arrays", bytes); return B200_ENOMEM; }
    }
  }
  ctx->host_pinned = pinned;
  return B200_OK;
}

// Bring up whatever the selected multi-GPU dot combine needs.  Caller's thread, workers idle.
static int ensure_combine_ready(b200_ctx* ctx) {
  if (ctx->ngpus == 1) return B200_OK;
  struct DeviceScope {  // the caller's current device is put back on every way out
    int saved = -1;
    DeviceScope() { if (cudaGetDevice(&saved) != cudaSuccess) { (void)cudaGetLastError(); saved = -1; } }
    ~DeviceScope() { if (saved >= 0) cudaSetDevice(saved); }
  } scope;
  if (ctx->combine == COMBINE_AUTO) {
    // Default inside ONE process: the caller is the only consumer of the sum, so every GPU stores its partial
    // into mapped host memory and rings a doorbell there; the caller polls the G doorbells and adds the G values
    // in rank order.  8 GPUs x 2^25 fp64, blocking dot through the driver: 49.3 TB/s (0.97 of 8 x one GPU)
    // against 48.2 TB/s for the fused peer-memory all-reduce (whose result makes one more hop: peers -> GPU 0 ->
    // host) and 38.0 TB/s for ncclAllReduce (profiles/r02/inprocess_8gpu.txt).  The fused all-reduce is what
    // process groups use (b200_join_group), where every rank needs the sum.
    ctx->combine = COMBINE_HOST;
  }
  if (ctx->combine == COMBINE_HOST) return B200_OK;
  if (ctx->combine == COMBINE_NCCL) {
    if (ctx->nccl_ready) return B200_OK;
    if (int rc = ctx->nccl.load()) return rc;
    std::vector<int> devs(ctx->ngpus);
    std::vector<ncclComm_t> comms(ctx->ngpus);
    for (int g = 0; g < ctx->ngpus; g++) devs[g] = ctx->first_device + g;
    ncclResult_t r = ctx->nccl.CommInitAll(comms.data(), ctx->ngpus, devs.data());
    if (r != ncclSuccess) { set_error("ncclCommInitAll failed: %s", ctx->nccl.GetErrorString(r)); return B200_ENCCL; }
    for (int g = 0; g < ctx->ngpus; g++) ctx->shards[g]->comm = comms[g];
    ctx->nccl_ready = true;
    return B200_OK;
  }
  if (ctx->p2p_ready) return B200_OK;
  if (ctx->ngpus > kMaxPeers) { set_error("dot_combine=p2p supports at most %d GPUs", kMaxPeers); return B200_EUNSUPPORTED; }
  for (Shard* s : ctx->shards) {
    B200_CUDA(cudaSetDevice(s->dev));
    for (Shard* peer : ctx->shards) {
      if (peer == s) continue;
      int can = 0;
      B200_CUDA(cudaDeviceCanAccessPeer(&can, s->dev, peer->dev));
      if (!can) { set_error("GPU %d cannot access GPU %d: dot_combine=p2p needs NVLink/PCIe peer access", s->dev, peer->dev); return B200_EUNSUPPORTED; }
      cudaError_t e = cudaDeviceEnablePeerAccess(peer->dev, 0);
      if (e == cudaErrorPeerAccessAlreadyEnabled) (void)cudaGetLastError();
      else if (e != cudaSuccess) return cuda_fail(e, "cudaDeviceEnablePeerAccess", __FILE__, __LINE__);
    }
    if (!s->mailbox) {
      B200_CUDA(cudaMalloc(&s->mailbox, b200_dot_mailbox_bytes()));
      B200_CUDA(cudaMemset(s->mailbox, 0, b200_dot_mailbox_bytes()));
    }
    ctx->mailboxes[s->rank] = s->mailbox;
  }
  for (Shard* s : ctx->shards) { B200_CUDA(cudaSetDevice(s->dev)); B200_CUDA(cudaDeviceSynchronize()); }
  ctx->p2p_ready = true;
  return B200_OK;
}

}  // namespace b200

// ---------------------------------------------------------------------------------------------
extern "C" {

void b200_destroy(b200_ctx* ctx) {
  if (!ctx) return;
  if (ctx->hostprof && ctx->prof_calls)
    fprintf(stderr, "b200 hostprof: %llu calls; per call: set_device %.0f ns, enqueue %.0f ns, wait %.0f ns\n",
            (unsigned long long)ctx->prof_calls, (double)ctx->prof_ns[0] / ctx->prof_calls,
            (double)ctx->prof_ns[1] / ctx->prof_calls, (double)ctx->prof_ns[2] / ctx->prof_calls);
  if (!ctx->workers.empty()) {
    Command q; q.kind = CMD_QUIT;
    ctx->cmd = q;
    ctx->seq.fetch_add(1);
    { std::lock_guard<std::mutex> lock(ctx->mu); }
    ctx->cv.notify_all();
    for (auto& t : ctx->workers) t.join();
  }
  release_worker_cpus(ctx);
  for (Shard* s : ctx->shards) {
    if (!s) continue;
    cudaSetDevice(s->dev);
    if (s->comm && ctx->nccl.CommDestroy) ctx->nccl.CommDestroy(s->comm);
    if (s->stream) cudaStreamDestroy(s->stream);
    if (s->ev0) cudaEventDestroy(s->ev0);
    if (s->ev1) cudaEventDestroy(s->ev1);
    for (auto& per_kernel : s->plans)
      for (auto& p : per_kernel) {
        if (p.exec) cudaGraphExecDestroy(p.exec);
        if (p.graph) cudaGraphDestroy(p.graph);
      }
    if (!ctx->managed) { cudaFree(s->a); cudaFree(s->b); cudaFree(s->c); }
    cudaFree(s->dot_ws); cudaFree(s->dot_dev); cudaFree(s->mailbox); cudaFree(s->bell_ticket); cudaFree(s->seq_counters);
    if (s->dot_host) cudaFreeHost(s->dot_host);
    delete s;
  }
  for (void* p : ctx->managed_base) if (p) { if (ctx->pagefault) free(p); else cudaFree(p); }
  free_host_mirrors(ctx);
  if (ctx->caller_device >= 0) cudaSetDevice(ctx->caller_device);
  delete ctx;
}

b200_ctx* b200_create(int first_device, int ngpus, size_t n, int elem_bytes, int* status) {
  return b200_create_opts(first_device, ngpus, n, elem_bytes, nullptr, status);
}

b200_ctx* b200_create_opts(int first_device, int ngpus, size_t n, int elem_bytes, const char* const* options, int* status) {
  int rc = B200_OK;
  b200_ctx* ctx = nullptr;
  auto fail = [&](int code) -> b200_ctx* {
    if (status) *status = code;
    const std::string keep = b200_last_error();  // destroy may overwrite it
    b200_destroy(ctx);
    set_error("%s", keep.c_str());
    return nullptr;
  };
  if (elem_bytes != 4 && elem_bytes != 8) { set_error("elem_bytes must be 4 or 8, got %d", elem_bytes); return fail(B200_EINVAL); }
  if (ngpus == 0) {
    const char* e = getenv("BABELSTREAM_B200_NGPUS");
    ngpus = e ? atoi(e) : 1;
  }
  const int count = b200_device_count();
  if (count <= 0) { set_error("no CUDA device available: libb200stream has no CPU fallback"); return fail(B200_ENODEVICE); }
  if (ngpus < 1 || first_device < 0 || first_device + ngpus > count) {
    // the reference throws std::runtime_error("Invalid device index") here, CUDAStream.cu:87-88
    set_error("Invalid device index: devices %d..%d requested, %d present", first_device, first_device + ngpus - 1, count);
    return fail(B200_ENODEVICE);
  }
  ctx = new (std::nothrow) b200_ctx;
  if (!ctx) { set_error("out of host memory"); return fail(B200_ENOMEM); }
  ctx->first_device = first_device; ctx->ngpus = ngpus; ctx->n = n; ctx->elem_bytes = elem_bytes;
  memset(ctx->tune, 0, sizeof ctx->tune);
  if (cudaGetDevice(&ctx->caller_device) != cudaSuccess) { (void)cudaGetLastError(); ctx->caller_device = -1; }
  // A caller that has never touched CUDA on this thread (the reference's driver: main.cpp only constructs the model) has
  // no device of its own to go back to -- cudaGetDevice just says 0.  "Restoring" 0 would create a primary context on
  // GPU 0 for `--device 1` (memory and start-up time on a GPU the run does not use) and make every later blocking call
  // switch devices twice.  For such a caller the context's first device simply stays current, as the reference's
  // constructor leaves it (CUDAStream.cu:89).
  if (!thread_has_cuda_context()) ctx->caller_device = -1;

  // options from the environment: BABELSTREAM_B200_<KEY>
  const char* keys[] = {"contract", "copy_variant", "dot_variant", "dot_combine", "timing", "sync", "tune", "mem", "launch", "sweep", "pin"};
  for (const char* k : keys) {
    std::string env = "BABELSTREAM_B200_";
    for (const char* p = k; *p; p++) env += (char)toupper(*p);
    if (const char* v = getenv(env.c_str())) { rc = apply_option(ctx, k, v); if (rc) return fail(rc); }
  }
  // explicit options override the environment
  for (const char* const* o = options; o && *o; o++) {
    const std::string kv = *o;
    const size_t eq = kv.find('=');
    if (eq == std::string::npos) { set_error("option '%s' is not key=value", *o); return fail(B200_EINVAL); }
    rc = apply_option(ctx, kv.substr(0, eq), kv.substr(eq + 1));
    if (rc) return fail(rc);
  }
  ctx->created = false;

  // mem=managed: the three arrays are single managed ranges; shard boundaries move to 2 MiB multiples
  // (the managed page size) so that no page is shared by two GPUs
  // (mem=pagefault: the same, on 2 MiB huge-page boundaries)
  const size_t shard_align = (ctx->managed && ngpus > 1 && n * (size_t)elem_bytes / (size_t)ngpus >= (8u << 20)) ? (2u << 20) : 128;
  if (ctx->pagefault) {
    // "p = (T*)malloc(array_bytes)" (CUDAStream.cu:39-40).  Every GPU that holds a shard must be able to
    // touch pageable memory; 2 MiB alignment so that shard boundaries fall on (huge) page boundaries.
    for (int g = 0; g < ngpus; g++) {
      int pageable = 0;
      cudaDeviceGetAttribute(&pageable, cudaDevAttrPageableMemoryAccess, first_device + g);
      if (!pageable) {
        set_error("mem=pagefault needs GPU access to pageable system memory (HMM or ATS), which device %d does not offer", first_device + g);
        return fail(B200_EUNSUPPORTED);
      }
    }
    const size_t huge = (size_t)2 << 20;
    const size_t array_bytes = ((n ? n * (size_t)elem_bytes : 256) + huge - 1) / huge * huge;
    for (int k = 0; k < 3; k++) {
      void* p = nullptr;
      if (posix_memalign(&p, huge, array_bytes) != 0 || !p) { set_error("out of host memory: %zu bytes for a pagefault array", array_bytes); return fail(B200_ENOMEM); }
      ctx->managed_base[k] = p;
    }
  } else if (ctx->managed) {
    int dev0_managed = 0;
    cudaDeviceGetAttribute(&dev0_managed, cudaDevAttrConcurrentManagedAccess, first_device);
    if (!dev0_managed) { set_error("mem=managed needs concurrent managed access, which device %d does not offer", first_device); return fail(B200_EUNSUPPORTED); }
    cudaError_t e = cudaSetDevice(first_device);
    const size_t array_bytes = n ? n * (size_t)elem_bytes : 256;
    for (int k = 0; k < 3 && e == cudaSuccess; k++) e = cudaMallocManaged(&ctx->managed_base[k], array_bytes);
    if (e != cudaSuccess) return fail(cuda_fail(e, "cudaMallocManaged", __FILE__, __LINE__));
  }

  if (const char* hp = getenv("BABELSTREAM_B200_HOSTPROF")) ctx->hostprof = atoi(hp) != 0;

  for (int g = 0; g < ngpus; g++) {
    Shard* s = new (std::nothrow) Shard;
    if (!s) { set_error("out of host memory"); return fail(B200_ENOMEM); }
    ctx->shards.push_back(s);
    s->dev = first_device + g;
    s->rank = g;
    shard_range_aligned(n, elem_bytes, ngpus, g, shard_align, &s->begin, &s->count);
    auto cu = [&](cudaError_t e, const char* what) { if (e != cudaSuccess && rc == B200_OK) rc = cuda_fail(e, what, __FILE__, __LINE__); };
    cu(cudaSetDevice(s->dev), "cudaSetDevice");
    if (rc) return fail(rc);
    if (!current_device_props(nullptr)) return fail(B200_EUNSUPPORTED);
    size_t free_b = 0, total_b = 0;
    cu(cudaMemGetInfo(&free_b, &total_b), "cudaMemGetInfo");
    const size_t shard_bytes = s->count * (size_t)elem_bytes;
    if (rc == B200_OK && 3.0 * (double)shard_bytes + 1e6 > (double)free_b) {
      // reference: "Device does not have enough memory for all buffers", CUDAStream.cu:135-139
      set_error("Device does not have enough memory for all buffers: %.3f GB requested on GPU %d, %.3f GB free",
                3.0 * shard_bytes * 1e-9, s->dev, free_b * 1e-9);
      return fail(B200_ENOMEM);
    }
    cu(cudaStreamCreateWithFlags(&s->stream, cudaStreamNonBlocking), "cudaStreamCreateWithFlags");
    cu(cudaEventCreate(&s->ev0), "cudaEventCreate");
    cu(cudaEventCreate(&s->ev1), "cudaEventCreate");
    const size_t alloc = shard_bytes ? shard_bytes : 256;
    if (ctx->managed) {
      void** mine[3] = {&s->a, &s->b, &s->c};
      for (int k = 0; k < 3; k++) {
        *mine[k] = static_cast<char*>(ctx->managed_base[k]) + s->begin * (size_t)elem_bytes;
        if (!shard_bytes) continue;
        if (ctx->pagefault) {
          // hints: the pages are not populated yet, init_arrays first-touches them from the owning GPU
          (void)cudaMemAdvise(*mine[k], shard_bytes, cudaMemAdviseSetPreferredLocation, s->dev);
          (void)cudaGetLastError();
          continue;
        }
        cu(cudaMemAdvise(*mine[k], shard_bytes, cudaMemAdviseSetPreferredLocation, s->dev), "cudaMemAdvise(preferred location)");
        cu(cudaMemPrefetchAsync(*mine[k], shard_bytes, s->dev, s->stream), "cudaMemPrefetchAsync");
      }
    } else {
      cu(cudaMalloc(&s->a, alloc), "cudaMalloc(a)");
      cu(cudaMalloc(&s->b, alloc), "cudaMalloc(b)");
      cu(cudaMalloc(&s->c, alloc), "cudaMalloc(c)");
    }
    cu(cudaMalloc(&s->dot_ws, kDotWorkspaceBytes), "cudaMalloc(dot workspace)");
    cu(cudaMalloc(&s->dot_dev, 16), "cudaMalloc(dot scalars)");
    cu(cudaMalloc(&s->bell_ticket, 256), "cudaMalloc(doorbell ticket)");
    cu(cudaMalloc(&s->seq_counters, 256), "cudaMalloc(sequence counters)");
    cu(cudaHostAlloc(&s->dot_host, 64, cudaHostAllocMapped | cudaHostAllocPortable), "cudaHostAlloc(dot result)");
    if (rc) return fail(rc);
    memset(s->dot_host, 0, 64);
    cu(cudaHostGetDevicePointer(&s->dot_host_dev, s->dot_host, 0), "cudaHostGetDevicePointer");
    cu(cudaMemsetAsync(s->dot_ws, 0, kDotWorkspaceBytes, s->stream), "cudaMemsetAsync");
    cu(cudaMemsetAsync(s->dot_dev, 0, 16, s->stream), "cudaMemsetAsync");
    cu(cudaMemsetAsync(s->bell_ticket, 0, 256, s->stream), "cudaMemsetAsync");
    cu(cudaMemsetAsync(s->seq_counters, 0, 256, s->stream), "cudaMemsetAsync");
    cu(cudaStreamSynchronize(s->stream), "cudaStreamSynchronize");
    if (rc) return fail(rc);
  }

  if (ngpus > 1) {
    rc = ensure_combine_ready(ctx);
    if (rc) return fail(rc);
    if (ctx->pin) choose_worker_cpus(ctx);
    for (int g = 1; g < ngpus; g++) ctx->workers.emplace_back(worker_main, ctx, g);
  }
  cudaSetDevice(ctx->caller_device >= 0 ? ctx->caller_device : first_device);  // the caller's current device is left as it was
  ctx->created = true;
  if (status) *status = B200_OK;
  return ctx;
}

int b200_set_option(b200_ctx* ctx, const char* key, const char* value) {
  if (!ctx || !key || !value) { set_error("b200_set_option: null argument"); return B200_EINVAL; }
  if (int rc = apply_option(ctx, key, value)) return rc;
  return ensure_combine_ready(ctx);  // e.g. dot_combine switched to nccl or p2p: bring it up now
}

int b200_join_group(b200_ctx* ctx, int world, int rank, void* const* mailboxes) {
  if (!ctx) { set_error("null context"); return B200_EINVAL; }
  if (ctx->ngpus != 1) { set_error("b200_join_group: the context must own exactly one GPU (it has %d)", ctx->ngpus); return B200_EINVAL; }
  if (world < 1 || world > kMaxPeers || rank < 0 || rank >= world || (world > 1 && !mailboxes)) {
    set_error("b200_join_group: bad world/rank (%d/%d; at most %d ranks)", world, rank, kMaxPeers);
    return B200_EINVAL;
  }
  for (int g = 0; g < kMaxPeers; g++) {
    ctx->group_mailboxes[g] = (world > 1 && g < world) ? mailboxes[g] : nullptr;
    if (world > 1 && g < world && !mailboxes[g]) { set_error("b200_join_group: mailbox %d is null", g); ctx->group_world = 1; return B200_EINVAL; }
  }
  ctx->group_world = world;
  ctx->group_rank = rank;
  ctx->tune_gen++;  // dot's plan changes shape
  // Every rank counts its dot calls from the moment it joins: the all-reduce sequence number (device counter)
  // restarts at 0, and this rank's OWN mailbox is wiped so that no {flag, payload} word of an earlier group,
  // context or LEVEL 1 call can be mistaken for a partial of the new group.  The caller must put a barrier
  // between the join of all ranks and the first dot (a fast peer's partial must not land before the wipe).
  Shard* s = ctx->shards[0];
  int saved = -1;
  if (cudaGetDevice(&saved) != cudaSuccess) { (void)cudaGetLastError(); saved = -1; }
  cudaError_t e = cudaSetDevice(s->dev);
  if (e == cudaSuccess) e = cudaMemsetAsync(static_cast<char*>(s->seq_counters) + 8, 0, 8, s->stream);
  if (e == cudaSuccess && world > 1) e = cudaMemsetAsync(mailboxes[rank], 0, b200_dot_mailbox_bytes(), s->stream);
  if (e == cudaSuccess) e = cudaStreamSynchronize(s->stream);
  if (saved >= 0 && saved != s->dev) cudaSetDevice(saved);
  if (e != cudaSuccess) return cuda_fail(e, "b200_join_group: reset of the sequence counter / own mailbox", __FILE__, __LINE__);
  *reinterpret_cast<volatile unsigned int*>(static_cast<char*>(s->dot_host) + 32) = 0;
  return B200_OK;
}

int b200_describe(b200_ctx* ctx, char* buf, int buflen) {
  if (!ctx || !buf || buflen <= 0) { set_error("b200_describe: null argument"); return B200_EINVAL; }
  int saved = -1;
  if (cudaGetDevice(&saved) != cudaSuccess) { (void)cudaGetLastError(); saved = -1; }
  if (saved != ctx->shards[0]->dev) B200_CUDA(cudaSetDevice(ctx->shards[0]->dev));  // describe the context's GPU, not the caller's
  const DeviceProps* props = current_device_props(nullptr);
  if (saved >= 0 && saved != ctx->shards[0]->dev) cudaSetDevice(saved);
  if (!props) return B200_ENODEVICE;
  std::string out;
  char line[256];
  snprintf(line, sizeof line, "GPUs: %d (devices %d..%d), %zu elements of %d bytes, shard %zu elements\n", ctx->ngpus,
           ctx->first_device, ctx->first_device + ctx->ngpus - 1, ctx->n, ctx->elem_bytes, ctx->shards[0]->count);
  out += line;
  if (ctx->pagefault) out += "  memory: PAGEFAULT (one system allocation per array, reached by the GPUs through page faults; get_arrays returns the ranges)\n";
  else if (ctx->managed) out += "  memory: MANAGED (one cudaMallocManaged range per array, preferred location = owning GPU, prefetched; get_arrays returns the ranges)\n";
  const char* names[B200_NUM_KERNELS] = {"copy", "mul", "add", "triad", "nstream", "dot", "init"};
  for (int k = 0; k < B200_NUM_KERNELS; k++) {
    Resolved r;
    if (resolve_cfg(k, ctx->shards[0]->count, ctx->elem_bytes, props->num_sms, props->l2_bytes, &ctx->tune[k], &r)) continue;
    const Geometry g = make_geometry(nullptr, nullptr, nullptr, ctx->shards[0]->count, ctx->elem_bytes, r.vec_bytes, r.block, r.unroll, r.cache_hint);
    const SweepPlan sp = plan_sweep(k, r, g, props->num_sms);
    snprintf(line, sizeof line, "  %-8s vec %2dB unroll %d block %4d grid %-9u hint %d variant %d %s", names[k], r.vec_bytes,
             r.unroll, r.block, sp.grid, r.cache_hint, r.variant, r.contract == 2 ? "unfused" : "fused");
    out += line;
    if (sp.tiles_per_block) { snprintf(line, sizeof line, ", %u tiles per block", sp.tiles_per_block); out += line; }
    if (sp.keep_from != ~0ull) { snprintf(line, sizeof line, ", last %d MB of a sweep kept in L2", r.keep_mb); out += line; }
    out += "\n";
  }
  if (ctx->ngpus > 1) {
    int v = 0;
    if (ctx->nccl_ready && ctx->nccl.GetVersion) ctx->nccl.GetVersion(&v);
    snprintf(line, sizeof line, "  dot combine: %s", ctx->combine == COMBINE_NCCL ? "ncclAllReduce(count=1)"
             : ctx->combine == COMBINE_P2P ? "fused peer-memory all-reduce inside the dot kernel (NVLink)" : "host sum in rank order");
    out += line;
    if (v) { snprintf(line, sizeof line, " (NCCL %d)", v); out += line; }
    out += "\n";
  }
  if (ctx->group_world > 1) {
    snprintf(line, sizeof line, "  rank %d of %d processes; dot: fused peer-memory all-reduce over IPC mailboxes\n", ctx->group_rank,
             ctx->group_world);
    out += line;
  }
  const char* sync_names[] = {"auto (dot: doorbell rung by the kernel; elementwise: cudaStreamSynchronize)", "cudaStreamSynchronize",
                              "cudaStreamQuery spin", "doorbell"};
  snprintf(line, sizeof line, "  wait: %s\n", sync_names[ctx->sync_mode + 1]);
  out += line;
  out += ctx->alternate ? "  sweep: consecutive kernels walk the arrays in opposite directions (sweep=forward: always bottom-up)\n"
                        : "  sweep: forward (always bottom-up)\n";
  out += ctx->use_graphs ? "  launch: every call replays a captured one-kernel CUDA graph (launch=direct: <<<>>>)\n"
                         : "  launch: direct (<<<>>> per call)\n";
  if (ctx->ngpus > 1) {
    out += "  threads: the caller drives GPU 0; workers";
    for (int g = 1; g < ctx->ngpus; g++) {
      if (ctx->shards[g]->pin_cpu >= 0) snprintf(line, sizeof line, " %d->cpu%d", g, ctx->shards[g]->pin_cpu);
      else snprintf(line, sizeof line, " %d->unpinned", g);
      out += line;
    }
    out += "\n";
  }
  snprintf(buf, (size_t)buflen, "%s", out.c_str());
  return B200_OK;
}

// mem=managed: after get_arrays the pages live in host memory; bring every shard home before it is used
// on the GPU again (the kernels would fault them over one by one otherwise)
static int managed_home(b200_ctx* ctx) {
  if (!ctx->managed_on_host) return B200_OK;
  Command cmd;
  cmd.kind = CMD_PREFETCH; cmd.prefetch_to = 0;
  if (int rc = run_all(ctx, cmd)) return rc;
  ctx->managed_on_host = false;
  return B200_OK;
}

static int run_kernel(b200_ctx* ctx, int kernel, double s0, double s1, double s2) {
  if (!ctx) { set_error("null context"); return B200_EINVAL; }
  if (int rc = managed_home(ctx)) return rc;
  Command cmd;
  cmd.kind = CMD_KERNEL; cmd.kernel = kernel; cmd.s0 = s0; cmd.s1 = s1; cmd.s2 = s2;
  if (ctx->alternate) { cmd.reverse = ctx->next_reverse; ctx->next_reverse ^= 1; }
  return run_all(ctx, cmd);
}

int b200_init_arrays(b200_ctx* ctx, double ia, double ib, double ic) { return run_kernel(ctx, B200_INIT, ia, ib, ic); }
int b200_copy(b200_ctx* ctx) { return run_kernel(ctx, B200_COPY, 0, 0, 0); }
int b200_mul(b200_ctx* ctx, double scalar) { return run_kernel(ctx, B200_MUL, scalar, 0, 0); }
int b200_add(b200_ctx* ctx) { return run_kernel(ctx, B200_ADD, 0, 0, 0); }
int b200_triad(b200_ctx* ctx, double scalar) { return run_kernel(ctx, B200_TRIAD, scalar, 0, 0); }
int b200_nstream(b200_ctx* ctx, double scalar) { return run_kernel(ctx, B200_NSTREAM, scalar, 0, 0); }

// dot_combine=host: the doorbells of the GPUs whose workers returned right after launching
static int collect_deferred_dots(b200_ctx* ctx) {
  for (Shard* s : ctx->shards) {
    if (!s->dot_deferred) continue;
    const int rc = bell_wait(s, s->dot_deferred, 8);
    s->dot_deferred = 0;
    if (rc) return rc;
  }
  return B200_OK;
}

int b200_warmup(b200_ctx* ctx, double scalar) {
  if (!ctx) { set_error("null context"); return B200_EINVAL; }
  if (int rc = managed_home(ctx)) return rc;
  const int kernels[6] = {B200_COPY, B200_MUL, B200_ADD, B200_TRIAD, B200_DOT, B200_NSTREAM};
  for (int reverse = 0; reverse < (ctx->alternate ? 2 : 1); reverse++) {
    for (int k : kernels) {
      Command cmd;
      cmd.kind = CMD_KERNEL; cmd.kernel = k; cmd.reverse = reverse;
      cmd.s0 = (k == B200_MUL || k == B200_TRIAD || k == B200_NSTREAM) ? scalar : 0.0; cmd.s1 = 0; cmd.s2 = 0;
      if (int rc = run_all(ctx, cmd)) return rc;
      if (k == B200_DOT) { if (int rc = collect_deferred_dots(ctx)) return rc; }
    }
  }
  ctx->next_reverse = 0;
  return B200_OK;
}

int b200_dot(b200_ctx* ctx, double* result) {
  if (int rc = run_kernel(ctx, B200_DOT, 0, 0, 0)) return rc;
  double total;
  if (ctx->ngpus > 1 && ctx->combine == COMBINE_HOST) {
    if (int rc = collect_deferred_dots(ctx)) return rc;
    // G values, added in rank order in type T: deterministic
    if (ctx->elem_bytes == 8) { double t = 0; for (Shard* s : ctx->shards) t += *static_cast<volatile double*>(s->dot_host); total = t; }
    else { float t = 0; for (Shard* s : ctx->shards) t += *static_cast<volatile float*>(s->dot_host); total = t; }
  } else {
    const void* h = ctx->shards[0]->dot_host;
    total = ctx->elem_bytes == 8 ? *static_cast<const volatile double*>(h) : (double)*static_cast<const volatile float*>(h);
  }
  if (result) *result = total;
  return B200_OK;
}

int b200_run(b200_ctx* ctx, int kernel, double scalar, double* result) {
  switch (kernel) {
    case B200_COPY: return b200_copy(ctx);
    case B200_MUL: return b200_mul(ctx, scalar);
    case B200_ADD: return b200_add(ctx);
    case B200_TRIAD: return b200_triad(ctx, scalar);
    case B200_NSTREAM: return b200_nstream(ctx, scalar);
    case B200_DOT: return b200_dot(ctx, result);
  }
  set_error("b200_run: kernel id %d is not a benchmark kernel", kernel);
  return B200_EINVAL;
}

// Several blocking calls from one call: what the driver's loop over bench[] does (src/main.cpp:150-165), for hosts whose
// own per-call overhead is not negligible (an interpreter).  Every kernel is launched and waited for before the next.
int b200_run_sequence(b200_ctx* ctx, const int* kernels, int count, double scalar, double* last_dot) {
  if (!ctx || (count > 0 && !kernels) || count < 0) { set_error("b200_run_sequence: bad arguments"); return B200_EINVAL; }
  for (int i = 0; i < count; i++)
    if (int rc = b200_run(ctx, kernels[i], scalar, last_dot)) return rc;
  return B200_OK;
}

static int transfer(b200_ctx* ctx, int kind, void* a, void* b, void* c) {
  if (!ctx) { set_error("null context"); return B200_EINVAL; }
  Command cmd;
  cmd.kind = kind; cmd.host[0] = a; cmd.host[1] = b; cmd.host[2] = c;
  return run_all(ctx, cmd);
}

int b200_read_arrays(b200_ctx* ctx, void* a, void* b, void* c) { return transfer(ctx, CMD_READ, a, b, c); }
int b200_write_arrays(b200_ctx* ctx, const void* a, const void* b, const void* c) {
  return transfer(ctx, CMD_WRITE, const_cast<void*>(a), const_cast<void*>(b), const_cast<void*>(c));
}

int b200_get_arrays(b200_ctx* ctx, const void** a, const void** b, const void** c) {
  if (!ctx) { set_error("null context"); return B200_EINVAL; }
  if (ctx->managed) {
    // "Unified memory: return pointers to device memory" (CUDAStream.cu:217-221) -- after moving the pages
    // to the CPU in bulk, all shards in parallel, so that check_solution does not fault them over one by one
    if (!ctx->managed_on_host) {
      Command cmd;
      cmd.kind = CMD_PREFETCH; cmd.prefetch_to = 1;
      if (int rc = run_all(ctx, cmd)) return rc;
      ctx->managed_on_host = true;
    }
    if (a) *a = ctx->managed_base[0];
    if (b) *b = ctx->managed_base[1];
    if (c) *c = ctx->managed_base[2];
    return B200_OK;
  }
  if (int rc = ensure_host_mirrors(ctx)) return rc;
  if (int rc = transfer(ctx, CMD_READ, ctx->host[0], ctx->host[1], ctx->host[2])) return rc;
  if (a) *a = ctx->host[0];
  if (b) *b = ctx->host[1];
  if (c) *c = ctx->host[2];
  return B200_OK;
}

int b200_num_gpus(b200_ctx* ctx) { return ctx ? ctx->ngpus : B200_EINVAL; }
const char* b200_memory_mode(b200_ctx* ctx) { return !ctx ? "" : ctx->pagefault ? "PAGEFAULT" : ctx->managed ? "MANAGED" : "DEFAULT"; }
size_t b200_array_size(b200_ctx* ctx) { return ctx ? ctx->n : 0; }
int b200_elem_bytes(b200_ctx* ctx) { return ctx ? ctx->elem_bytes : B200_EINVAL; }

double b200_last_kernel_ms(b200_ctx* ctx) {
  if (!ctx || !ctx->timing) return -1.0;
  float ms = -1.f;
  for (Shard* s : ctx->shards) if (s->last_ms > ms) ms = s->last_ms;
  return (double)ms;
}

int b200_shard_info(b200_ctx* ctx, int g, int* device, void** a, void** b, void** c, size_t* begin, size_t* count,
                    void** stream) {
  if (!ctx || g < 0 || g >= ctx->ngpus) { set_error("b200_shard_info: bad shard index"); return B200_EINVAL; }
  Shard* s = ctx->shards[g];
  if (device) *device = s->dev;
  if (a) *a = s->a;
  if (b) *b = s->b;
  if (c) *c = s->c;
  if (begin) *begin = s->begin;
  if (count) *count = s->count;
  if (stream) *stream = s->stream;
  return B200_OK;
}

}  // extern "C"
