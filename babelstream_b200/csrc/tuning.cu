// tuning.cu -- host-side launch logic: the per-size configuration table, the grid policy, the
// vector geometry for arbitrary pointer alignment, and the multi-GPU shard partition.
// Pure host code (no CUDA calls): unit-tested on a machine without a GPU.
//
// The reference fixes its shape at compile time: 256 threads, UNROLL_FACTOR 4 scalar accesses,
// grid = ceil(N/4/256), dot = 4 blocks/SM x 1024 threads (src/cuda/CUDAStream.h:18-19,
// CUDAStream.cu:126,196-198).  Here the shape is chosen per (kernel, bytes, L2 size); the
// entries below are the winners of tools/sweep.py on a B200 (profiles/r01/sweep_*_top.csv, profiles/r02/).
#include <limits.h>
#include <stdarg.h>
#include <stdio.h>
#include <string.h>

#include "internal.h"

namespace b200 {

static thread_local char g_error[512] = "";

void set_error(const char* fmt, ...) {
  va_list ap;
  va_start(ap, fmt);
  vsnprintf(g_error, sizeof g_error, fmt, ap);
  va_end(ap);
}
const char* last_error_message() { return g_error; }

int size_class(size_t n, int elem_bytes, int l2_bytes) {
  if (l2_bytes <= 0) l2_bytes = 126 << 20;
  const double f = 3.0 * (double)n * elem_bytes / (double)l2_bytes;
  return f <= 0.8 ? SIZE_L2 : f <= 3.2 ? SIZE_KNEE : f <= 16.0 ? SIZE_HBM : SIZE_HBM_XL;
}

int resolve_cfg(int kernel, size_t n, int elem_bytes, int num_sms, int l2_bytes, const b200_launch_cfg* in,
                Resolved* out) {
  if (kernel < 0 || kernel >= B200_NUM_KERNELS) { set_error("unknown kernel id %d", kernel); return B200_EINVAL; }
  if (elem_bytes != 4 && elem_bytes != 8) { set_error("elem_bytes must be 4 or 8, got %d", elem_bytes); return B200_EINVAL; }
  b200_launch_cfg z;
  memset(&z, 0, sizeof z);
  const b200_launch_cfg& c = in ? *in : z;
  if (num_sms <= 0) num_sms = 148;
  if (l2_bytes <= 0) l2_bytes = 126 << 20;

  // BabelStream keeps all three arrays live across kernels, so the size class is judged on 3 arrays
  const int cls = size_class(n, elem_bytes, l2_bytes);
  const bool l2_resident = cls == SIZE_L2;

  Resolved r;
  memset(&r, 0, sizeof r);
  r.variant = B200_VARIANT_LDST;
  r.grid_policy = INT_MAX;  // one tile per block: always beats a persistent grid-stride loop (by 4-10 %)
  // ---- the per-size table (auto) ------------------------------------------------------------------
  // Sources: round 1 tools/sweep.py (profiles/r01/sweep_f64_top.csv: 6560 shapes, back to back), round 2
  // tools/lab/ew_lab (profiles/r02/lab_*: blocking calls in the driver's kernel order, which is what src/main.cpp
  // times) and the driver itself with BABELSTREAM_B200_TUNE overrides (profiles/r02/driver_session_*.txt).
  // The table assumes that consecutive kernels sweep in opposite directions (B200_LAUNCH_REVERSE on every other
  // launch: LEVEL 2's default, see SweepArgs in kernels.cuh): what a kernel touches last is what the next one
  // reads first, so the cache policy decides how much of the 126 MB L2 that hand-over can use.
  // Classes by the footprint F of the three arrays against the L2 size:
  //
  //   SIZE_L2     F <= 0.8 L2   launch-latency bound (8-20 us per call): fat threads in small blocks, default
  //                             cache policy so that consecutive kernels hit in L2
  //   SIZE_KNEE   F <= 3.2 L2   part of every array still lives in L2 when the next kernel reads it: same shape,
  //                             default policy (no evict-first: the lines are wanted again)
  //   SIZE_HBM    F <= 16 L2    HBM-bound, 50-250 us per call: ramp-up and tail are 3-4 % of the call, and they
  //                             shrink with the tile -- 4 KB per array per block (256 threads x 16 B) runs the
  //                             blocking call 2.3 us (triad, add) faster than 8 KB at 2^25 fp64 although both
  //                             stream at the same 7.0 TB/s back to back.
  //                             F <= 8 L2 (2^25 fp64): evict-first for the body of the sweep, the last 128 MB of
  //                             traffic kept (loads evict-normal, stores evict-last).  Six-kernel step, driver-timed:
  //                             646 us forward -> 625 us alternating, 96 MB kept -> 617.6 us with 128 MB kept (one
  //                             policy for everything: 626 us default, 635 us no-allocate).
  //                             8 L2 < F (2^26 fp64): default policy for everything, nothing pinned: 1223 us against
  //                             1228 us (evict-first + keep) and 1254 us forward.
  //   SIZE_HBM_XL beyond        steady state is everything: 8 KB per array per block streams 1.2-1.4 % faster at
  //                             2^29 fp64 than 4 KB, the ~2 us of tail no longer matter.  L1 no-allocate, normal L2
  //                             priority for all lines: the step takes 2417 / 4801 / 9561 us at 2^27 / 2^28 / 2^29
  //                             against 2428 / 4818 / 9589 us with evict-first and 2445 / 4828 / 9594 us forward;
  //                             triad alone at 2^29: 1780-1783 us (7.23-7.24 TB/s) against 1787 us evict-first
  //                             alternating and 1793 us forward.
  //
  // Mixing priorities is what hurts at the large sizes: lines stored evict-last (or loaded evict-normal) among
  // evict-first traffic are never displaced by it and squat in L2 for the following kernels, whose streams then
  // thrash in what is left (add with the default policy between evict-first neighbours at 2^28: add -11 us, the
  // triad and dot after it +18 us each; a kept tail at 2^29: triad +2.6 %).
  // What did NOT help (profiles/r02/README.md): cp.async.bulk staging through shared memory with either
  // register or bulk stores (the cub::DeviceTransform ublkcp shape), L2 prefetch ahead of the loads (per-line
  // prefetch.global.L2 -- what the reference's thrust model runs -- or cp.async.bulk.prefetch.L2), persistent
  // multi-stage rings: all within +-0.5 % of plain LDG/STG at 2^25 and 1-2.5 % slower at 2^29.
  const double footprint = 3.0 * (double)n * elem_bytes / (double)l2_bytes;
  const bool keep_tail = cls == SIZE_HBM && footprint <= 8.0;
  r.cache_hint = cls == SIZE_HBM_XL ? B200_HINT_NOALLOC : keep_tail ? B200_HINT_EVICT_FIRST : B200_HINT_DEFAULT;
  switch (kernel) {
    case B200_COPY: case B200_MUL:
      // one read and one write stream: two 1024-thread blocks per SM run the alternating blocking call 0.7 % faster
      // at 2^28 than four of 512 (612 vs 617 us; profiles/r02/sweep_alt_2p28_top.csv), no difference at 2^25
      r.vec_bytes = 16; r.unroll = 2; r.block = cls == SIZE_HBM_XL ? 1024 : 512; break;
    case B200_ADD: case B200_TRIAD:
      r.vec_bytes = 16; r.unroll = 1; r.block = cls == SIZE_HBM_XL ? 512 : 256; break;
    case B200_NSTREAM:
      if (cls == SIZE_HBM_XL) { r.vec_bytes = 16; r.unroll = 2; r.block = 512; }
      else { r.vec_bytes = 32; r.unroll = 1; r.block = 256; }
      break;
    case B200_INIT: r.vec_bytes = 32; r.unroll = 1; r.block = 256; break;
    case B200_DOT:
      // read-only: a bounded grid (the partials must fit the workspace), each block adding one contiguous chunk.
      // Tuned in the driver's own situation -- blocking call, issued right after triad whose dirty lines are
      // still draining from L2 (profiles/r01/sweep_dot_blocking_after_triad_top.csv): 8 blocks of 512
      // threads per SM (two waves, the second balancing the first) with 2 x 16-byte loads per
      // operand in flight is within 1.1 % of the best shape at 2^25, 2^27 and 2^29; shapes with
      // less than full occupancy win in isolation but lose up to 10 % in sequence.  Same cache policy as the
      // elementwise kernels of the class (profiles/r02/driver_session_g_hints.txt: a dot that does not allocate
      // among default-policy neighbours is 4 us faster at 2^26 and makes the nstream after it 4.5 us slower).
      r.variant = B200_VARIANT_DOT_TICKET;
      r.vec_bytes = 16; r.unroll = 2; r.block = 512; r.grid_policy = -8;
      break;
  }
  // The tail of a sweep stays in L2 for the next kernel -- only where that pays (see SweepArgs in kernels.cuh)
  r.keep_mb = keep_tail ? 128 : 0;
  if (cls <= SIZE_KNEE) {
    // Working set inside L2 (<= 2^22 doubles) or not much larger (2^23, 2^24): one 32-byte access per thread in
    // 256-thread blocks.  Driver-timed candidates: profiles/r01/small_size_candidates.txt (2^20..2^22),
    // profiles/r02/class_calibration.txt (2^20..2^24: at the knee the six-kernel step takes 181.6 us with this
    // shape and default policy against 189.6 us with the HBM shapes + no-allocate and 229.6 us for cuda-stream
    // at 2^23; dot alone 28.3 vs 33.5 us because triad's output stays in L2 for it).  The alternating sweep takes
    // the knee further: 176.7 us at 2^23 and 330.2 us at 2^24 against 185.3 / 350.9 us forward.
    if (kernel == B200_DOT) { if (l2_resident) { r.vec_bytes = 32; r.unroll = 4; r.block = 256; r.grid_policy = -2; } }
    else { r.vec_bytes = 32; r.unroll = 1; r.block = 256; }
  }
  r.contract = 1;
  r.stages = 4;
  r.chunk_bytes = 32768;

  // ---- explicit overrides ----
  if (c.vec_bytes) r.vec_bytes = c.vec_bytes;
  if (c.unroll) r.unroll = c.unroll;
  if (c.block) r.block = c.block;
  if (c.grid) r.grid_policy = c.grid;
  if (c.cache_hint) r.cache_hint = c.cache_hint;
  if (c.variant) r.variant = c.variant;
  if (c.contract) r.contract = c.contract;
  if (c.stages) r.stages = c.stages;
  if (c.chunk_bytes) r.chunk_bytes = c.chunk_bytes;
  r.flags = c.flags;
  if (c.keep_mb) r.keep_mb = c.keep_mb < 0 ? 0 : c.keep_mb;

  // ---- validation ----
  if (r.vec_bytes != 8 && r.vec_bytes != 16 && r.vec_bytes != 32 && r.vec_bytes != elem_bytes) {
    set_error("vec_bytes must be 8, 16, 32 or elem_bytes (scalar), got %d", r.vec_bytes); return B200_EINVAL;
  }
  if (r.vec_bytes == 4) { r.unroll = 4; r.cache_hint = B200_HINT_DEFAULT; }  // float scalar path: one shape
  const bool unroll_ok = r.unroll == 1 || r.unroll == 2 || r.unroll == 4 || (r.unroll == 8 && r.vec_bytes == 16);
  if (!unroll_ok) { set_error("unroll must be 1, 2, 4 (or 8 with 16-byte vectors), got %d", r.unroll); return B200_EINVAL; }
  // register budget: see ew_max_block / dot_max_block in kernels.cuh
  const int max_block = kernel == B200_DOT ? 512
                                           : (r.vec_bytes * r.unroll <= 32 ? 1024 : 512);
  if (r.block < 32 || r.block > max_block || r.block % 32) {
    set_error("block must be a multiple of 32 in [32, %d] for vec_bytes=%d unroll=%d, got %d", max_block, r.vec_bytes,
              r.unroll, r.block);
    return B200_EINVAL;
  }
  if (r.cache_hint < B200_HINT_DEFAULT || r.cache_hint > B200_HINT_PREFETCH) {
    set_error("cache_hint out of range: %d", r.cache_hint); return B200_EINVAL;
  }
  if (r.contract != 1 && r.contract != 2) { set_error("contract must be 1 (fused) or 2 (unfused)"); return B200_EINVAL; }
  if (kernel == B200_DOT) {
    if (r.variant != B200_VARIANT_DOT_TWO_PASS && r.variant != B200_VARIANT_DOT_TICKET) {
      set_error("dot variant must be two_pass (1) or ticket (2)"); return B200_EINVAL;
    }
    if (r.grid_policy == INT_MAX) r.grid_policy = -4;  // a dot grid is always bounded (workspace)
  } else if (kernel == B200_COPY) {
    if (r.variant < B200_VARIANT_LDST || r.variant > B200_VARIANT_TMA) { set_error("copy variant out of range"); return B200_EINVAL; }
  } else if (r.variant != B200_VARIANT_LDST) {
    set_error("only copy has bulk/tma variants"); return B200_EINVAL;
  }
  *out = r;
  return B200_OK;
}

unsigned grid_for(int grid_policy, unsigned long long ntiles, int num_sms) {
  if (ntiles == 0) return 1;  // still launched: block 0 handles the ragged edges
  unsigned long long g;
  if (grid_policy == INT_MAX) g = ntiles;
  else if (grid_policy < 0) g = (unsigned long long)(-(long long)grid_policy) * (unsigned long long)num_sms;
  else g = (unsigned long long)grid_policy;
  if (g > ntiles) g = ntiles;
  if (g > 0x7fffffffull) g = 0x7fffffffull;  // gridDim.x limit; the tile loop covers the rest
  if (g == 0) g = 1;
  return (unsigned)g;
}

unsigned long long keep_from_for(int keep_mb, int cache_hint, unsigned long long units, double bytes_per_unit) {
  if (keep_mb <= 0 || cache_hint != B200_HINT_EVICT_FIRST || units == 0 || !(bytes_per_unit > 0)) return ~0ull;
  const double want = (double)keep_mb * 1048576.0 / bytes_per_unit;
  const unsigned long long keep = want >= (double)units ? units : (unsigned long long)want;
  return units - keep;
}

SweepPlan plan_sweep(int kernel, const Resolved& r, const Geometry& g, int num_sms) {
  SweepPlan sp;
  sp.grid = grid_for(r.grid_policy, g.ntiles, num_sms);
  sp.reverse = (r.flags & B200_LAUNCH_REVERSE) != 0;
  const double vec_tile_bytes = (double)r.block * g.unroll * g.vec_bytes;  // one tile of one array
  if (kernel == B200_DOT) {
    if (!(r.flags & B200_LAUNCH_DOT_GRIDSTRIDE) && g.ntiles > 0) {
      // one contiguous chunk of k tiles per logical block; no empty blocks
      const unsigned long long k = (g.ntiles + sp.grid - 1) / sp.grid;
      sp.tiles_per_block = (unsigned)k;
      sp.grid = (unsigned)((g.ntiles + k - 1) / k);
    }
    sp.keep_from = keep_from_for(r.keep_mb, g.cache_hint, sp.grid,
                                 (double)(sp.tiles_per_block ? sp.tiles_per_block : 1) * vec_tile_bytes * 2.0);
    return sp;
  }
  const int arrays = (kernel == B200_COPY || kernel == B200_MUL) ? 2 : kernel == B200_NSTREAM ? 4 : 3;
  sp.keep_from = keep_from_for(r.keep_mb, g.cache_hint, g.ntiles, vec_tile_bytes * arrays);
  const unsigned long long tile_vecs = (unsigned long long)r.block * (unsigned)g.unroll;
  sp.origin = sp.reverse && g.ntiles ? (g.ntiles - 1) * tile_vecs : 0;
  sp.stride = sp.reverse ? -(long long)tile_vecs : (long long)tile_vecs;
  return sp;
}

Geometry make_geometry(const void* p0, const void* p1, const void* p2, size_t n, int elem_bytes, int vec_bytes,
                       int block, int unroll, int cache_hint) {
  Geometry g;
  const void* ptrs[3] = {p0, p1, p2};
  // all used pointers must share their offset within a vector, otherwise no common `head` exists
  bool congruent = true;
  uintptr_t off = 0;
  bool have = false;
  for (const void* p : ptrs) {
    if (!p) continue;
    const uintptr_t o = (uintptr_t)p % (uintptr_t)vec_bytes;
    if (o % elem_bytes) congruent = false;  // not even element-aligned relative to the vector grid
    if (!have) { off = o; have = true; }
    else if (o != off) congruent = false;
  }
  if (vec_bytes == elem_bytes || !congruent) {
    g.vec_bytes = elem_bytes; g.unroll = 4; g.cache_hint = B200_HINT_DEFAULT; g.head = 0;
    g.nvec = n;
  } else {
    g.vec_bytes = vec_bytes; g.unroll = unroll; g.cache_hint = cache_hint;
    const unsigned epv = (unsigned)(vec_bytes / elem_bytes);
    unsigned head = off ? (unsigned)((vec_bytes - off) / elem_bytes) : 0u;
    if (head > n) head = (unsigned)n;
    g.head = head;
    g.nvec = (n - head) / epv;
  }
  const unsigned long long tile = (unsigned long long)block * g.unroll;
  g.ntiles = (g.nvec + tile - 1) / tile;
  return g;
}

void shard_range_aligned(size_t n, int elem_bytes, int ngpus, int g, size_t align_bytes, size_t* begin, size_t* count) {
  // ceil split rounded up to a multiple of align_bytes; trailing shards may be short or empty
  const size_t align = align_bytes / (size_t)elem_bytes;
  size_t per = (n + (size_t)ngpus - 1) / (size_t)ngpus;
  per = (per + align - 1) / align * align;
  size_t b = per * (size_t)g;
  if (b > n) b = n;
  size_t e = b + per;
  if (e > n) e = n;
  *begin = b;
  *count = e - b;
}

}  // namespace b200

// ---- exported host logic --------------------------------------------------------------------
extern "C" {

const char* b200_last_error(void) { return b200::last_error_message(); }

int b200_shard_range(size_t n, int elem_bytes, int ngpus, int g, size_t* begin, size_t* count) {
  if (ngpus < 1 || g < 0 || g >= ngpus || (elem_bytes != 4 && elem_bytes != 8) || !begin || !count) {
    b200::set_error("b200_shard_range: bad arguments (n=%zu elem_bytes=%d ngpus=%d g=%d)", n, elem_bytes, ngpus, g);
    return B200_EINVAL;
  }
  b200::shard_range_aligned(n, elem_bytes, ngpus, g, 128, begin, count);
  return B200_OK;
}

int b200_resolve_cfg(int kernel, size_t n, int elem_bytes, int num_sms, int l2_bytes, const b200_launch_cfg* in,
                     b200_launch_cfg* out, unsigned* grid_out) {
  b200::Resolved r;
  if (int rc = b200::resolve_cfg(kernel, n, elem_bytes, num_sms, l2_bytes, in, &r)) return rc;
  if (out) {
    memset(out, 0, sizeof *out);
    out->vec_bytes = r.vec_bytes; out->unroll = r.unroll; out->block = r.block; out->grid = r.grid_policy;
    out->cache_hint = r.cache_hint; out->variant = r.variant; out->contract = r.contract;
    out->stages = r.stages; out->chunk_bytes = r.chunk_bytes; out->flags = r.flags; out->keep_mb = r.keep_mb;
  }
  if (grid_out) {
    // geometry for perfectly aligned arrays
    const b200::Geometry g = b200::make_geometry(nullptr, nullptr, nullptr, n, elem_bytes, r.vec_bytes, r.block,
                                                 r.unroll, r.cache_hint);
    *grid_out = b200::plan_sweep(kernel, r, g, num_sms > 0 ? num_sms : 148).grid;
  }
  return B200_OK;
}

int b200_plan_sweep(int kernel, size_t n, int elem_bytes, int num_sms, int l2_bytes, const b200_launch_cfg* in,
                    b200_sweep_plan* out) {
  if (!out) { b200::set_error("b200_plan_sweep: null output"); return B200_EINVAL; }
  b200::Resolved r;
  if (int rc = b200::resolve_cfg(kernel, n, elem_bytes, num_sms, l2_bytes, in, &r)) return rc;
  const b200::Geometry g = b200::make_geometry(nullptr, nullptr, nullptr, n, elem_bytes, r.vec_bytes, r.block, r.unroll,
                                               r.cache_hint);
  const b200::SweepPlan sp = b200::plan_sweep(kernel, r, g, num_sms > 0 ? num_sms : 148);
  if (kernel == B200_DOT && sp.grid > b200::kDotMaxPartials) {  // the launcher refuses the same way
    b200::set_error("dot grid %u exceeds the workspace capacity of %u blocks", sp.grid, b200::kDotMaxPartials);
    return B200_EINVAL;
  }
  memset(out, 0, sizeof *out);
  out->grid = sp.grid; out->tiles_per_block = sp.tiles_per_block; out->reverse = sp.reverse ? 1 : 0;
  out->keep_from = sp.keep_from; out->origin = sp.origin; out->stride = sp.stride;
  out->ntiles = g.ntiles; out->nvec = g.nvec; out->tile_vecs = (unsigned long long)r.block * (unsigned)g.unroll;
  out->vec_bytes = g.vec_bytes; out->unroll = g.unroll; out->cache_hint = g.cache_hint;
  return B200_OK;
}

}  // extern "C"
