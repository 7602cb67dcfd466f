// launch_ew.cuh -- picks the ew_kernel instantiation for a resolved configuration and launches it.
// Included by exactly one translation unit per op (ew_<op>.cu).
#pragma once

#include "internal.h"
#include "kernels.cuh"

namespace b200 {

// The compiled variant space per (T, op, rounding):
//   8-byte: unroll 1 2 4     16-byte: unroll 1 2 4 8     32-byte: unroll 1 2 4      x  5 cache hints
//   float scalar fallback (misaligned operands): 4-byte accesses, unroll 4, default hint
//   (the double scalar fallback is the 8-byte, unroll 4, default-hint member of the table)
#define B200_EW_FOR_HINT(H)                                                                   \
  B200_EW_CASE(8, 1, H) B200_EW_CASE(8, 2, H) B200_EW_CASE(8, 4, H)                           \
  B200_EW_CASE(16, 1, H) B200_EW_CASE(16, 2, H) B200_EW_CASE(16, 4, H) B200_EW_CASE(16, 8, H) \
  B200_EW_CASE(32, 1, H) B200_EW_CASE(32, 2, H) B200_EW_CASE(32, 4, H)

template <typename T, int OP, bool FUSED>
int launch_ew_typed(const EwCall& call) {
  using Kernel = void (*)(const EwArgs, const T, const T, const T);
  Kernel k = nullptr;
  const int vb = call.g.vec_bytes, u = call.g.unroll, h = call.g.cache_hint;
#define B200_EW_CASE(VB, U, H) \
  if (vb == VB && u == U && h == H) k = ew_kernel<T, OP, VB, U, H, FUSED>;
  B200_EW_FOR_HINT(1) B200_EW_FOR_HINT(2) B200_EW_FOR_HINT(3) B200_EW_FOR_HINT(4) B200_EW_FOR_HINT(5)
  if constexpr (sizeof(T) == 4) { if (vb == 4 && u == 4 && h == 1) k = ew_kernel<T, OP, 4, 4, 1, FUSED>; }
#undef B200_EW_CASE
  if (!k) {
    set_error("no elementwise kernel compiled for vec_bytes=%d unroll=%d cache_hint=%d", vb, u, h);
    return B200_EUNSUPPORTED;
  }
  if (call.r.block > ew_max_block(vb, u)) {
    set_error("block %d too large for vec_bytes=%d unroll=%d (max %d)", call.r.block, vb, u, ew_max_block(vb, u));
    return B200_EINVAL;
  }
  EwArgs args;
  args.a = call.a; args.b = call.b; args.c = call.c;
  args.nvec = call.g.nvec; args.ntiles = call.g.ntiles; args.n = call.n; args.head = call.g.head;
  const size_t head_bytes = (size_t)call.g.head * sizeof(T);
  args.va = static_cast<char*>(call.a) + head_bytes;
  args.vb = static_cast<char*>(call.b) + head_bytes;
  args.vc = static_cast<char*>(call.c) + head_bytes;
  args.sweep.reverse = call.sweep.reverse ? 1u : 0u;
  args.sweep.origin = call.sweep.origin;
  args.sweep.stride = call.sweep.stride;
  args.sweep.keep_from = call.sweep.keep_from >= 0xffffffffull ? ~0u : (unsigned)call.sweep.keep_from;
  args.bell.ticket = static_cast<unsigned int*>(call.bell.ticket);
  args.bell.flag = static_cast<unsigned long long*>(call.bell.flag);
  args.bell.seq = call.bell.seq;
  return launch_kernel(k, call.grid, call.r.block, 0, call.stream, call.r.flags, args, (T)call.s0, (T)call.s1, (T)call.s2);
}

template <int OP>
int launch_ew_op(const EwCall& call) {
  constexpr bool HAS_AXPY = (OP == OP_TRIAD || OP == OP_NSTREAM);
  if (call.elem_bytes == 8) {
    if constexpr (HAS_AXPY) { if (call.r.contract == 2) return launch_ew_typed<double, OP, false>(call); }
    return launch_ew_typed<double, OP, true>(call);
  }
  if constexpr (HAS_AXPY) { if (call.r.contract == 2) return launch_ew_typed<float, OP, false>(call); }
  return launch_ew_typed<float, OP, true>(call);
}

}  // namespace b200
