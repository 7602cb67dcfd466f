// dot.cu -- launcher for the two dot variants (see kernels.cuh for the algorithm).
#include <stdlib.h>

#include <atomic>

#include "internal.h"
#include "kernels.cuh"

namespace b200 {

// How long the collecting warp of the peer-memory all-reduce waits for a missing peer before it gives up
// and reports B200_ETIMEOUT: BABELSTREAM_B200_P2P_TIMEOUT_MS (default 2000), b200_set_p2p_timeout_ms().
// 0 = not read from the environment yet.  Atomics: the worker threads of a multi-GPU context read it concurrently.
static std::atomic<unsigned long long> g_p2p_timeout_ns{0};
static std::atomic<unsigned long long> g_p2p_timeout_gen{1};  // bumped by every change: captured dot plans carry the value
unsigned long long p2p_timeout_ns() {
  unsigned long long ns = g_p2p_timeout_ns.load(std::memory_order_acquire);
  if (ns == 0) {
    double ms = 2000.0;
    if (const char* e = getenv("BABELSTREAM_B200_P2P_TIMEOUT_MS")) { const double v = atof(e); if (v > 0) ms = v; }
    unsigned long long expected = 0;
    ns = (unsigned long long)(ms * 1e6);
    if (!g_p2p_timeout_ns.compare_exchange_strong(expected, ns, std::memory_order_acq_rel)) ns = expected;  // someone set it meanwhile
  }
  return ns;
}
void set_p2p_timeout_ns(unsigned long long ns) {
  g_p2p_timeout_ns.store(ns ? ns : 1, std::memory_order_release);
  g_p2p_timeout_gen.fetch_add(1, std::memory_order_acq_rel);
}
unsigned long long p2p_timeout_gen() { return g_p2p_timeout_gen.load(std::memory_order_acquire); }

#define B200_DOT_FOR_HINT(H)                                                                      \
  B200_DOT_CASE(8, 1, H) B200_DOT_CASE(8, 2, H) B200_DOT_CASE(8, 4, H)                            \
  B200_DOT_CASE(16, 1, H) B200_DOT_CASE(16, 2, H) B200_DOT_CASE(16, 4, H) B200_DOT_CASE(16, 8, H) \
  B200_DOT_CASE(32, 1, H) B200_DOT_CASE(32, 2, H) B200_DOT_CASE(32, 4, H)

template <typename T, bool FUSED, int MODE>
static int launch_dot_typed(const DotCall& call) {
  using Kernel = void (*)(const DotArgs);
  Kernel k = nullptr;
  const int vb = call.g.vec_bytes, u = call.g.unroll, h = call.g.cache_hint;
#define B200_DOT_CASE(VB, U, H) \
  if (vb == VB && u == U && h == H) k = dot_kernel<T, VB, U, H, FUSED, MODE>;
  B200_DOT_FOR_HINT(1) B200_DOT_FOR_HINT(2) B200_DOT_FOR_HINT(3) B200_DOT_FOR_HINT(4) B200_DOT_FOR_HINT(5)
  if constexpr (sizeof(T) == 4) { if (vb == 4 && u == 4 && h == 1) k = dot_kernel<T, 4, 4, 1, FUSED, MODE>; }
#undef B200_DOT_CASE
  if (!k) {
    set_error("no dot kernel compiled for vec_bytes=%d unroll=%d cache_hint=%d", vb, u, h);
    return B200_EUNSUPPORTED;
  }
  if (call.r.block > dot_block_bound(vb, u, MODE)) {
    set_error("dot: block %d too large for vec_bytes=%d unroll=%d%s (max %d)", call.r.block, vb, u,
              MODE >= DOT_TICKET_P2P ? " with the fused peer-memory all-reduce" : "", dot_block_bound(vb, u, MODE));
    return B200_EINVAL;
  }
  DotArgs args;
  args.a = call.a; args.b = call.b;
  args.partials = call.workspace;
  args.result_dev = call.result_dev; args.result_host = call.result_host;
  args.nvec = call.g.nvec; args.ntiles = call.g.ntiles; args.n = call.n; args.head = call.g.head;
  args.tiles_per_block = call.sweep.tiles_per_block;
  args.sweep.reverse = call.sweep.reverse ? 1u : 0u;
  args.sweep.keep_from = call.sweep.keep_from >= 0xffffffffull ? ~0u : (unsigned)call.sweep.keep_from;
  args.sweep.origin = 0; args.sweep.stride = 0;
  args.bell_flag = static_cast<unsigned long long*>(call.bell.flag); args.bell_seq = call.bell.seq;
  args.seq_counters = static_cast<unsigned long long*>(call.seq_counters);
  args.p2p.world = call.p2p_world; args.p2p.rank = call.p2p_rank; args.p2p.seq = call.p2p_seq;
  args.p2p.timeout_ns = p2p_timeout_ns();
  args.p2p.status_dev = reinterpret_cast<unsigned int*>(static_cast<char*>(call.workspace) + kDotWsStatus);
  args.p2p.status_host = static_cast<unsigned int*>(call.status_host);
  for (int g = 0; g < kMaxPeers; g++)
    args.p2p.mailbox[g] = (call.p2p_world > 1 && g < call.p2p_world) ? static_cast<MailSlot*>(call.p2p_mailboxes[g]) : nullptr;
  if (int rc = launch_kernel(k, call.grid, call.r.block, 0, call.stream, call.r.flags, args)) return rc;
  if (MODE == DOT_TWO_PASS) {
    // one block of 32 warps: warp g adds group g -- the same order as the ticket variant, whatever the grid
    if (int rc = launch_kernel(dot_final_kernel<T>, 1u, 1024, 0, call.stream, call.r.flags, static_cast<const T*>(call.workspace),
                               call.grid, static_cast<T*>(call.result_dev), static_cast<T*>(call.result_host), args.bell_flag,
                               args.bell_seq, args.seq_counters))
      return rc;
  }
  return B200_OK;
}

int launch_dot_impl(const DotCall& call) {
  const bool fused = call.r.contract != 2;
  const bool ticket = call.r.variant == B200_VARIANT_DOT_TICKET;
  if (call.p2p_world > 1) {
    if (!ticket) { set_error("the fused peer-memory all-reduce needs the ticket variant of dot"); return B200_EINVAL; }
    if (call.p2p_world > kMaxPeers || call.p2p_rank < 0 || call.p2p_rank >= call.p2p_world || !call.p2p_mailboxes ||
        (call.p2p_seq == 0 && !call.seq_counters)) {
      set_error("dot p2p: bad world/rank/seq/mailboxes (world %d, rank %d)", call.p2p_world, call.p2p_rank);
      return B200_EINVAL;
    }
  }
  static_assert(kDotMaxPartials == (unsigned)kDotMaxBlocks, "host and device disagree about the dot workspace");
  if (call.grid > (unsigned)kDotMaxBlocks) {
    set_error("dot grid %u exceeds the workspace capacity of %d blocks", call.grid, kDotMaxBlocks);
    return B200_EINVAL;
  }
  const int mode = call.p2p_world > 1 ? (call.p2p_send_only ? DOT_TICKET_SEND : DOT_TICKET_P2P) : ticket ? DOT_TICKET : DOT_TWO_PASS;
#define B200_DOT_DISPATCH(T)                                                                                   \
  switch (mode) {                                                                                              \
    case DOT_TWO_PASS: return fused ? launch_dot_typed<T, true, DOT_TWO_PASS>(call) : launch_dot_typed<T, false, DOT_TWO_PASS>(call); \
    case DOT_TICKET: return fused ? launch_dot_typed<T, true, DOT_TICKET>(call) : launch_dot_typed<T, false, DOT_TICKET>(call);       \
    case DOT_TICKET_SEND: return fused ? launch_dot_typed<T, true, DOT_TICKET_SEND>(call) : launch_dot_typed<T, false, DOT_TICKET_SEND>(call); \
    default: return fused ? launch_dot_typed<T, true, DOT_TICKET_P2P>(call) : launch_dot_typed<T, false, DOT_TICKET_P2P>(call);     \
  }
  if (call.elem_bytes == 8) { B200_DOT_DISPATCH(double) }
  B200_DOT_DISPATCH(float)
#undef B200_DOT_DISPATCH
}

int launch_dot_collect(int elem_bytes, void* result_dev, void* result_host, int world, int rank, unsigned long long seq,
                       void* const* mailboxes, void* status_dev, int flags, cudaStream_t stream) {
  if (world < 2 || world > kMaxPeers || rank < 0 || rank >= world || !mailboxes || seq == 0) {
    set_error("dot collect: bad world/rank/seq/mailboxes (world %d, rank %d)", world, rank);
    return B200_EINVAL;
  }
  P2PArgs p;
  for (int g = 0; g < kMaxPeers; g++) p.mailbox[g] = g < world ? static_cast<MailSlot*>(mailboxes[g]) : nullptr;
  p.seq = seq; p.world = world; p.rank = rank;
  p.timeout_ns = p2p_timeout_ns();
  p.status_dev = static_cast<unsigned int*>(status_dev);
  p.status_host = nullptr;
  if (!p.mailbox[rank]) { set_error("dot collect: own mailbox is null"); return B200_EINVAL; }
  unsigned long long* const no_bell = nullptr;
  if (elem_bytes == 8)
    return launch_kernel(dot_collect_kernel<double>, 1u, 32, 0, stream, flags, p, static_cast<double*>(result_dev),
                         static_cast<double*>(result_host), no_bell, 0ull);
  return launch_kernel(dot_collect_kernel<float>, 1u, 32, 0, stream, flags, p, static_cast<float*>(result_dev),
                       static_cast<float*>(result_host), no_bell, 0ull);
}

}  // namespace b200
