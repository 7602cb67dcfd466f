// kernels.cuh -- hand-written sm_100a device code for the BabelStream kernels.
//
// What is computed (reference: /root/reference/src/cuda/CUDAStream.cu:203-296, and the CPU
// models the results must match bit for bit, src/serial/SerialStream.cpp:36-113):
//   init    a[i]=ia; b[i]=ib; c[i]=ic          copy    c[i] = a[i]
//   mul     b[i] = s*c[i]                      add     c[i] = a[i] + b[i]
//   triad   a[i] = b[i] + s*c[i]               nstream a[i] = a[i] + (b[i] + s*c[i])
//   dot     sum_i a[i]*b[i]
//
// Nothing here is a contraction, so tensor cores / TMEM are not used; every kernel is bound by
// HBM bandwidth (<= 1 flop per 12 bytes).  The design levers are the ones that matter for a
// streaming kernel on B200:
//   * 128-bit and 256-bit fully coalesced accesses (LDG.E.128 / LDG.E.ENL2.256, sm_100 only),
//   * all loads of a tile issued before the first use (memory-level parallelism),
//   * explicit L1/L2 cache policy per access (no_allocate / evict_first / .cs / L2::256B),
//   * tile = blockDim * UNROLL vectors, grid either one tile per block or a grid-stride loop
//     over k blocks per SM -- chosen per size in tuning.cuh,
//   * for copy: no registers at all -- cp.async.bulk / TMA tensor tiles through a shared-memory
//     ring driven by one thread per CTA.
//
// Rounding is spelled out with the _rn intrinsics, which the compiler never contracts or
// re-associates: FUSED = true reproduces the reference's default build (GCC contracts
// b + s*c into fma(s,c,b)), FUSED = false its -ffp-contract=off build.  The file is compiled
// with -fmad=false as well, so no other FMA can appear.
#pragma once

#include <cuda.h>
#include <cuda_runtime.h>
#include <stdint.h>

namespace b200 {

enum Op : int { OP_COPY = 0, OP_MUL = 1, OP_ADD = 2, OP_TRIAD = 3, OP_NSTREAM = 4, OP_DOT = 5, OP_INIT = 6 };
enum Hint : int { H_DEFAULT = 1, H_NOALLOC = 2, H_EVICT_FIRST = 3, H_STREAM = 4, H_PREFETCH = 5 };

// ------------------------------------------------------------------------------------------------
// Exactly-rounded scalar arithmetic
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double mul_rn(double x, double y) { return __dmul_rn(x, y); }
__device__ __forceinline__ float mul_rn(float x, float y) { return __fmul_rn(x, y); }
__device__ __forceinline__ double add_rn(double x, double y) { return __dadd_rn(x, y); }
__device__ __forceinline__ float add_rn(float x, float y) { return __fadd_rn(x, y); }
__device__ __forceinline__ double fma_rn(double x, double y, double z) { return __fma_rn(x, y, z); }
__device__ __forceinline__ float fma_rn(float x, float y, float z) { return __fmaf_rn(x, y, z); }

// b + s*c with the reference's two possible roundings
template <bool FUSED, typename T>
__device__ __forceinline__ T axpy(T s, T c, T b) {
  if constexpr (FUSED) return fma_rn(s, c, b);
  else return add_rn(b, mul_rn(s, c));
}

// ------------------------------------------------------------------------------------------------
// Vector register packs and hinted global accesses.  VB = bytes per access (4, 8, 16, 32).
// Data travels as raw 32/64-bit words; 256-bit accesses (v4.b64) exist on sm_100 and later.
// ------------------------------------------------------------------------------------------------
template <int VB> struct Words;
template <> struct Words<4> { uint32_t w[1]; };
template <> struct Words<8> { uint64_t w[1]; };
template <> struct Words<16> { uint64_t w[2]; };
template <> struct Words<32> { uint64_t w[4]; };

#define B200_LD_QUAL_1_RW ""
#define B200_LD_QUAL_1_RO ".nc"
#define B200_LD_QUAL_2_RW ".L1::no_allocate"
#define B200_LD_QUAL_2_RO ".nc.L1::no_allocate"
#define B200_LD_QUAL_3_RW ".L1::no_allocate.L2::evict_first"      /* 256-bit accesses only */
#define B200_LD_QUAL_3_RO ".nc.L1::no_allocate.L2::evict_first"
#define B200_LD_QUAL_3P_RW ".L1::no_allocate.L2::cache_hint"      /* narrower: createpolicy operand */
#define B200_LD_QUAL_3P_RO ".nc.L1::no_allocate.L2::cache_hint"
#define B200_LD_QUAL_4_RW ".cs"
#define B200_LD_QUAL_5_RW ".L1::no_allocate.L2::256B"
#define B200_LD_QUAL_5_RO ".nc.L1::no_allocate.L2::256B"
#define B200_ST_QUAL_1 ""
#define B200_ST_QUAL_2 ".L1::no_allocate"
#define B200_ST_QUAL_3 ".L1::no_allocate.L2::evict_first"
#define B200_ST_QUAL_3P ".L1::no_allocate.L2::cache_hint"
#define B200_ST_QUAL_4 ".cs"
#define B200_ST_QUAL_5 ".L1::no_allocate"

#define B200_LD_BODY(Q)                                                                                         \
  if constexpr (VB == 4)                                                                                        \
    asm volatile("ld.global" Q ".b32 %0, [%1];" : "=r"(r.w[0]) : "l"(p));                                       \
  else if constexpr (VB == 8)                                                                                   \
    asm volatile("ld.global" Q ".b64 %0, [%1];" : "=l"(r.w[0]) : "l"(p));                                       \
  else if constexpr (VB == 16)                                                                                  \
    asm volatile("ld.global" Q ".v2.b64 {%0,%1}, [%2];" : "=l"(r.w[0]), "=l"(r.w[1]) : "l"(p));                 \
  else                                                                                                          \
    asm volatile("ld.global" Q ".v4.b64 {%0,%1,%2,%3}, [%4];"                                                   \
                 : "=l"(r.w[0]), "=l"(r.w[1]), "=l"(r.w[2]), "=l"(r.w[3]) : "l"(p));

// same, with an L2 cache-policy operand (a createpolicy register: the policy can change from tile to tile)
#define B200_LD_BODY_POLICY(Q)                                                                                  \
  if constexpr (VB == 4)                                                                                        \
    asm volatile("ld.global" Q ".b32 %0, [%1], %2;" : "=r"(r.w[0]) : "l"(p), "l"(policy));                      \
  else if constexpr (VB == 8)                                                                                   \
    asm volatile("ld.global" Q ".b64 %0, [%1], %2;" : "=l"(r.w[0]) : "l"(p), "l"(policy));                      \
  else if constexpr (VB == 16)                                                                                  \
    asm volatile("ld.global" Q ".v2.b64 {%0,%1}, [%2], %3;" : "=l"(r.w[0]), "=l"(r.w[1]) : "l"(p), "l"(policy)); \
  else                                                                                                          \
    asm volatile("ld.global" Q ".v4.b64 {%0,%1,%2,%3}, [%4], %5;"                                               \
                 : "=l"(r.w[0]), "=l"(r.w[1]), "=l"(r.w[2]), "=l"(r.w[3]) : "l"(p), "l"(policy));

#define B200_ST_BODY(Q)                                                                                         \
  if constexpr (VB == 4)                                                                                        \
    asm volatile("st.global" Q ".b32 [%0], %1;" ::"l"(p), "r"(r.w[0]) : "memory");                              \
  else if constexpr (VB == 8)                                                                                   \
    asm volatile("st.global" Q ".b64 [%0], %1;" ::"l"(p), "l"(r.w[0]) : "memory");                              \
  else if constexpr (VB == 16)                                                                                  \
    asm volatile("st.global" Q ".v2.b64 [%0], {%1,%2};" ::"l"(p), "l"(r.w[0]), "l"(r.w[1]) : "memory");         \
  else                                                                                                          \
    asm volatile("st.global" Q ".v4.b64 [%0], {%1,%2,%3,%4};" ::"l"(p), "l"(r.w[0]), "l"(r.w[1]), "l"(r.w[2]),  \
                 "l"(r.w[3]) : "memory");

#define B200_ST_BODY_POLICY(Q)                                                                                  \
  if constexpr (VB == 4)                                                                                        \
    asm volatile("st.global" Q ".b32 [%0], %1, %2;" ::"l"(p), "r"(r.w[0]), "l"(policy) : "memory");             \
  else if constexpr (VB == 8)                                                                                   \
    asm volatile("st.global" Q ".b64 [%0], %1, %2;" ::"l"(p), "l"(r.w[0]), "l"(policy) : "memory");             \
  else if constexpr (VB == 16)                                                                                  \
    asm volatile("st.global" Q ".v2.b64 [%0], {%1,%2}, %3;" ::"l"(p), "l"(r.w[0]), "l"(r.w[1]), "l"(policy) : "memory"); \
  else                                                                                                          \
    asm volatile("st.global" Q ".v4.b64 [%0], {%1,%2,%3,%4}, %5;" ::"l"(p), "l"(r.w[0]), "l"(r.w[1]), "l"(r.w[2]), \
                 "l"(r.w[3]), "l"(policy) : "memory");

// L2 cache-policy registers.  H_EVICT_FIRST accesses carry one as an operand, so that the policy can differ
// between the body of a sweep (evict-first: the line is not wanted again) and its last tiles (see SweepArgs).
enum L2Policy : int { L2_EVICT_FIRST = 0, L2_EVICT_NORMAL = 1, L2_EVICT_LAST = 2 };
template <int KIND>
__device__ __forceinline__ uint64_t l2_policy() {
  uint64_t policy;
  // not volatile: a pure value (ptxas folds it into a constant), free to be hoisted and selected between
  if constexpr (KIND == L2_EVICT_FIRST) asm("createpolicy.fractional.L2::evict_first.b64 %0, 1.0;" : "=l"(policy));
  else if constexpr (KIND == L2_EVICT_NORMAL) asm("createpolicy.fractional.L2::evict_normal.b64 %0, 1.0;" : "=l"(policy));
  else asm("createpolicy.fractional.L2::evict_last.b64 %0, 1.0;" : "=l"(policy));
  return policy;
}
template <int VB, int HINT>
__device__ __forceinline__ uint64_t make_policy() {
  if constexpr (HINT == H_EVICT_FIRST) return l2_policy<L2_EVICT_FIRST>();
  else return 0;
}

// The sweep: in which order a grid walks its tiles, and which of them keep their lines in L2.
//   reverse    block p (dispatch order) handles tile ntiles-1-p: the arrays are swept from the top down.  Consecutive
//              kernels of a sequence alternate the direction, so each one starts on the data its predecessor touched
//              last -- the part that is still in the 126 MB L2 (output lines not yet written back, inputs just read).
//   keep_from  tiles p >= keep_from (dispatch order, i.e. the END of the sweep) load with evict_normal and store with
//              evict_last instead of evict_first, so that the successor finds them.  Only with H_EVICT_FIRST (the other
//              hints keep everything by default); ~0 = no tile.  Measured on B200 (profiles/r02/lab_zigzag.txt,
//              driver_session_e_sweep_keep.txt): 2^25 fp64, six-kernel sequence of blocking calls: 646 us forward ->
//              638 us alternating -> 618 us with the last 128 MB kept; at 2^27 and beyond keeping HURTS (kept lines are
//              never displaced by evict-first traffic and squat in L2: +2.6 % at 2^29) -- there every line gets the
//              same priority and the alternation alone gains 0.5-1.2 % (which class does what: tuning.cu).
// The elementwise kernels take the direction as an affine map, prepared by the launcher so that the per-block prologue
// (the latency before a block's first load, which a one-tile-per-block grid pays 2^19 times at 2^29) stays short:
// dispatch index p starts at vector origin + stride*p, stride = +-(vectors per tile).  p < 2^32 (checked on the host).
struct SweepArgs {
  unsigned long long origin;  // ew: first vector of dispatch tile 0 (0, or (ntiles-1)*tile_vecs when reversed)
  long long stride;           // ew: +tile_vecs or -tile_vecs
  unsigned int keep_from;     // dispatch index (ew: tile, dot: block) from which lines are kept; ~0u: none
  unsigned int reverse;       // dot: logical block = gridDim.x-1-blockIdx.x
};

// READONLY: the array is not written by this kernel (-> .nc, the non-coherent path).
template <int VB, int HINT, bool READONLY>
__device__ __forceinline__ Words<VB> ldg(const void* p, uint64_t policy = 0) {
  Words<VB> r;
  if constexpr (HINT == H_DEFAULT) { if constexpr (READONLY) { B200_LD_BODY(B200_LD_QUAL_1_RO) } else { B200_LD_BODY(B200_LD_QUAL_1_RW) } }
  else if constexpr (HINT == H_NOALLOC) { if constexpr (READONLY) { B200_LD_BODY(B200_LD_QUAL_2_RO) } else { B200_LD_BODY(B200_LD_QUAL_2_RW) } }
  else if constexpr (HINT == H_EVICT_FIRST) { if constexpr (READONLY) { B200_LD_BODY_POLICY(B200_LD_QUAL_3P_RO) } else { B200_LD_BODY_POLICY(B200_LD_QUAL_3P_RW) } }
  else if constexpr (HINT == H_STREAM) { B200_LD_BODY(B200_LD_QUAL_4_RW) }
  else { if constexpr (READONLY) { B200_LD_BODY(B200_LD_QUAL_5_RO) } else { B200_LD_BODY(B200_LD_QUAL_5_RW) } }
  (void)policy;
  return r;
}

template <int VB, int HINT>
__device__ __forceinline__ void stg(void* p, const Words<VB>& r, uint64_t policy = 0) {
  if constexpr (HINT == H_DEFAULT) { B200_ST_BODY(B200_ST_QUAL_1) }
  else if constexpr (HINT == H_NOALLOC) { B200_ST_BODY(B200_ST_QUAL_2) }
  else if constexpr (HINT == H_EVICT_FIRST) { B200_ST_BODY_POLICY(B200_ST_QUAL_3P) }
  else if constexpr (HINT == H_STREAM) { B200_ST_BODY(B200_ST_QUAL_4) }
  else { B200_ST_BODY(B200_ST_QUAL_5) }
  (void)policy;
}

// words <-> elements (register renaming only; no instructions survive)
template <typename T, int VB> struct Elems { T v[VB / sizeof(T)]; };

template <typename T, int VB>
__device__ __forceinline__ Elems<T, VB> unpack(const Words<VB>& w) {
  Elems<T, VB> e;
  if constexpr (sizeof(T) == 8) {
#pragma unroll
    for (int i = 0; i < VB / 8; i++) e.v[i] = __longlong_as_double((long long)w.w[i]);
  } else if constexpr (VB == 4) {
    e.v[0] = __uint_as_float(w.w[0]);
  } else {
#pragma unroll
    for (int i = 0; i < VB / 8; i++) {
      e.v[2 * i] = __uint_as_float((uint32_t)(w.w[i] & 0xffffffffull));
      e.v[2 * i + 1] = __uint_as_float((uint32_t)(w.w[i] >> 32));
    }
  }
  return e;
}

template <typename T, int VB>
__device__ __forceinline__ Words<VB> pack(const Elems<T, VB>& e) {
  Words<VB> w;
  if constexpr (sizeof(T) == 8) {
#pragma unroll
    for (int i = 0; i < VB / 8; i++) w.w[i] = (uint64_t)__double_as_longlong(e.v[i]);
  } else if constexpr (VB == 4) {
    w.w[0] = __float_as_uint(e.v[0]);
  } else {
#pragma unroll
    for (int i = 0; i < VB / 8; i++)
      w.w[i] = (uint64_t)__float_as_uint(e.v[2 * i]) | ((uint64_t)__float_as_uint(e.v[2 * i + 1]) << 32);
  }
  return w;
}

// ------------------------------------------------------------------------------------------------
// Programmatic dependent launch.  Every kernel starts with: "the next kernel of this stream may be
// scheduled as soon as all my blocks have started" and "wait until the previous kernel of the stream
// has completed and its memory is visible".  Both are no-ops unless the launch carries the
// programmatic-stream-serialization attribute (cfg.flags & B200_LAUNCH_PDL), in which case the
// ~1-2 us drain-and-launch gap between consecutive kernels of an asynchronous sequence disappears:
// the successor's blocks are already resident, parked in griddepcontrol.wait, when the predecessor's
// last block retires.  A successor can only be scheduled once ALL blocks of its predecessor are
// running, so it never takes an SM slot a predecessor block is still waiting for.
// ------------------------------------------------------------------------------------------------
__device__ __forceinline__ void pdl_entry() {
  asm volatile("griddepcontrol.launch_dependents;" ::: "memory");
  asm volatile("griddepcontrol.wait;" ::: "memory");
}

// ------------------------------------------------------------------------------------------------
// Completion doorbell.  A blocking caller normally learns that a kernel has finished from
// cudaStreamSynchronize, i.e. from the semaphore the front end releases after the grid has drained
// and its caches are flushed, plus the driver's own bookkeeping: 2-3 us after the last store.  With
// a doorbell the kernel rings the host itself: every block takes a ticket once its stores are
// fenced, and the block that takes the last one writes `seq` to a word of mapped pinned host memory
// that the caller spins on.  Stream order still holds for whatever is enqueued next, so "results are
// complete when the call returns" (src/Stream.h:36) is kept.  Default for dot, which takes a ticket
// anyway and whose result is itself a write into host memory (the call gets 1.3-1.9 us shorter);
// opt-in for the elementwise kernels, where ~1000 same-address tickets cost more than the ring saves.
// ------------------------------------------------------------------------------------------------
struct Doorbell {
  unsigned int* ticket;            // device word, zero between launches
  unsigned long long* flag;        // device-visible address of mapped pinned host memory, or nullptr
  unsigned long long seq;          // value to store
};

__device__ __forceinline__ void ring_host(unsigned long long* flag, unsigned long long seq) {
  asm volatile("st.release.sys.global.u64 [%0], %1;" ::"l"(flag), "l"(seq) : "memory");
}

// Block-level: call with all threads once the block's stores have been issued.
__device__ __forceinline__ void doorbell_block_done(const Doorbell& bell) {
  __syncthreads();  // every thread's stores are issued ...
  if (threadIdx.x == 0) {
    __threadfence();  // ... and ordered before the ticket
    const unsigned t = atomicAdd(bell.ticket, 1u);
    if (t == gridDim.x - 1) {
      *bell.ticket = 0u;  // ready for the next launch (stream order separates the launches)
      __threadfence();    // acquire side: the other blocks' stores are visible before the host is told
      ring_host(bell.flag, bell.seq);
    }
  }
}

// ------------------------------------------------------------------------------------------------
// Elementwise kernels.  One template for init/copy/mul/add/triad/nstream.
//
// Geometry: the arrays are addressed from element `head` (0 <= head < VB/sizeof(T), chosen by the
// launcher so that a+head, b+head, c+head are VB-aligned), as nvec vectors of VB bytes.  A tile
// is blockDim.x*UNROLL consecutive vectors; thread t of the block owns vectors t, t+blockDim.x, ...
// of the tile, so every warp-level access covers 32*VB contiguous bytes.  Blocks walk the tiles
// with stride gridDim.x (gridDim.x == ntiles gives one tile per block).  The < 2*VB/sizeof(T)
// ragged elements before `head` and after the last full vector are done by block 0.
// ------------------------------------------------------------------------------------------------
struct EwArgs {
  void* a;
  void* b;
  void* c;
  char* va;                   // a + head, b + head, c + head: the VB-aligned bases of the vector part
  char* vb;
  char* vc;
  unsigned long long nvec;    // full vectors after head
  unsigned long long ntiles;  // ceil(nvec / (blockDim.x * UNROLL))
  unsigned long long n;       // total elements
  unsigned int head;          // leading scalar elements
  SweepArgs sweep;            // direction and L2 keep region
  Doorbell bell;              // flag == nullptr: none
};

template <typename T, int OP, bool FUSED>
__device__ __forceinline__ void ew_scalar(T* a, T* b, T* c, unsigned long long i, T s0, T s1, T s2) {
  if constexpr (OP == OP_INIT) { a[i] = s0; b[i] = s1; c[i] = s2; }
  else if constexpr (OP == OP_COPY) c[i] = a[i];
  else if constexpr (OP == OP_MUL) b[i] = mul_rn(s0, c[i]);
  else if constexpr (OP == OP_ADD) c[i] = add_rn(a[i], b[i]);
  else if constexpr (OP == OP_TRIAD) a[i] = axpy<FUSED>(s0, c[i], b[i]);
  else if constexpr (OP == OP_NSTREAM) a[i] = add_rn(a[i], axpy<FUSED>(s0, c[i], b[i]));
}

// Threads per block the register budget allows: up to 32 bytes per thread per array fit in the 64
// registers of a 1024-thread block; wider tiles run with at most 512 threads (128 registers).
constexpr int ew_max_block(int vb, int unroll) { return vb * unroll <= 32 ? 1024 : 512; }

template <typename T, int OP, int VB, int UNROLL, int HINT, bool FUSED>
__global__ void __launch_bounds__(ew_max_block(VB, UNROLL))
ew_kernel(const EwArgs args, const T s0, const T s1, const T s2) {
  constexpr int EPV = VB / (int)sizeof(T);
  constexpr bool LA = (OP == OP_COPY || OP == OP_ADD || OP == OP_NSTREAM);
  constexpr bool LB = (OP == OP_ADD || OP == OP_TRIAD || OP == OP_NSTREAM);
  constexpr bool LC = (OP == OP_MUL || OP == OP_TRIAD || OP == OP_NSTREAM);
  constexpr bool A_RO = (OP == OP_COPY || OP == OP_ADD);  // a is read AND written by nstream: no .nc
  pdl_entry();

  T* const a = static_cast<T*>(args.a);
  T* const b = static_cast<T*>(args.b);
  T* const c = static_cast<T*>(args.c);
  char* const va = args.va;
  char* const vb = args.vb;
  char* const vc = args.vc;

  const unsigned long long nvec = args.nvec;
  const unsigned long long tile_vecs = (unsigned long long)blockDim.x * UNROLL;

  const uint64_t body_policy = make_policy<VB, HINT>();
  uint64_t keep_ld_policy = 0, keep_st_policy = 0;
  if constexpr (HINT == H_EVICT_FIRST) { keep_ld_policy = l2_policy<L2_EVICT_NORMAL>(); keep_st_policy = l2_policy<L2_EVICT_LAST>(); }
  Elems<T, VB> init_a, init_b, init_c;
  if constexpr (OP == OP_INIT) {
#pragma unroll
    for (int e = 0; e < EPV; e++) { init_a.v[e] = s0; init_b.v[e] = s1; init_c.v[e] = s2; }
  }

  for (unsigned long long ptile = blockIdx.x; ptile < args.ntiles; ptile += gridDim.x) {
    const unsigned p32 = (unsigned)ptile;
    const unsigned long long vbase = args.sweep.origin + (unsigned long long)(args.sweep.stride * (long long)p32);
    const unsigned long long v0 = vbase + threadIdx.x;
    const bool full = vbase + tile_vecs <= nvec;  // block-uniform
    uint64_t policy = body_policy, st_policy = body_policy;
    if constexpr (HINT == H_EVICT_FIRST) {
      const bool keep = p32 >= args.sweep.keep_from;
      policy = keep ? keep_ld_policy : body_policy;
      st_policy = keep ? keep_st_policy : body_policy;
    }

    Words<VB> wa[UNROLL], wb[UNROLL], wc[UNROLL];
    // 1. every load of the tile first
#pragma unroll
    for (int j = 0; j < UNROLL; j++) {
      const unsigned long long v = v0 + (unsigned long long)j * blockDim.x;
      if (full || v < nvec) {
        if constexpr (LA) wa[j] = ldg<VB, HINT, A_RO>(va + v * VB, policy);
        if constexpr (LB) wb[j] = ldg<VB, HINT, true>(vb + v * VB, policy);
        if constexpr (LC) wc[j] = ldg<VB, HINT, true>(vc + v * VB, policy);
      }
    }
    // 2. arithmetic and stores
#pragma unroll
    for (int j = 0; j < UNROLL; j++) {
      const unsigned long long v = v0 + (unsigned long long)j * blockDim.x;
      if (full || v < nvec) {
        if constexpr (OP == OP_INIT) {
          stg<VB, HINT>(va + v * VB, pack<T, VB>(init_a), st_policy);
          stg<VB, HINT>(vb + v * VB, pack<T, VB>(init_b), st_policy);
          stg<VB, HINT>(vc + v * VB, pack<T, VB>(init_c), st_policy);
        } else if constexpr (OP == OP_COPY) {
          stg<VB, HINT>(vc + v * VB, wa[j], st_policy);
        } else {
          Elems<T, VB> ea, eb, ec, r;
          if constexpr (LA) ea = unpack<T, VB>(wa[j]);
          if constexpr (LB) eb = unpack<T, VB>(wb[j]);
          if constexpr (LC) ec = unpack<T, VB>(wc[j]);
#pragma unroll
          for (int e = 0; e < EPV; e++) {
            if constexpr (OP == OP_MUL) r.v[e] = mul_rn(s0, ec.v[e]);
            else if constexpr (OP == OP_ADD) r.v[e] = add_rn(ea.v[e], eb.v[e]);
            else if constexpr (OP == OP_TRIAD) r.v[e] = axpy<FUSED>(s0, ec.v[e], eb.v[e]);
            else r.v[e] = add_rn(ea.v[e], axpy<FUSED>(s0, ec.v[e], eb.v[e]));  // nstream: a + (b + s*c)
          }
          if constexpr (OP == OP_MUL) stg<VB, HINT>(vb + v * VB, pack<T, VB>(r), st_policy);
          else if constexpr (OP == OP_ADD) stg<VB, HINT>(vc + v * VB, pack<T, VB>(r), st_policy);
          else stg<VB, HINT>(va + v * VB, pack<T, VB>(r), st_policy);
        }
      }
    }
  }

  // ragged edges: [0, head) and [head + nvec*EPV, n)
  if (blockIdx.x == 0) {
    const unsigned long long tail_begin = args.head + nvec * EPV;
    const unsigned long long edge = args.head + (args.n - tail_begin);
    if (threadIdx.x < edge) {
      const unsigned long long i = threadIdx.x < args.head ? threadIdx.x : tail_begin + (threadIdx.x - args.head);
      ew_scalar<T, OP, FUSED>(a, b, c, i, s0, s1, s2);
    }
  }
  if (args.bell.flag) doorbell_block_done(args.bell);  // grid-uniform
}

// ------------------------------------------------------------------------------------------------
// Dot.  Pass 1: every thread accumulates UNROLL*EPV independent chains over its vectors
// (FUSED: acc = fma(a,b,acc); else acc = acc + round(a*b)), folds them in a fixed order, then
// warp shuffle tree -> one shared-memory slot per warp -> warp 0 -> block partial.
// Pass 2 (deterministic, two levels): the gridDim.x partials are split into NG = min(32, gridDim.x)
// interleaved groups (block b belongs to group b % NG).  ONE WARP adds the partials of a group in a fixed
// order (lane-strided chains, shuffle tree), and one warp adds the NG group sums the same way.
// It runs either as a second launch (dot_final_kernel: warp g of one block takes group g) or,
// TICKET = true, inside pass 1: the last block of each group to finish (a ticket per group) adds its
// group, and whichever of those NG blocks comes last (a second ticket) adds the group sums.  The
// order of the additions never depends on which block does them, so both variants give the same bits.
// Why two levels: with one ticket, ~1200 blocks that finish together queue ~1200 atomics on ONE L2
// address (2-3 ns each) and the winner then needs a whole-block pass over all partials; with groups
// no address sees more than ~37 atomics and the critical path after the last block's loads is two
// atomics and two one-warp sums.
// ------------------------------------------------------------------------------------------------
constexpr int kDotMaxBlocks = 16384;
constexpr int kDotGroups = 32;
constexpr int kDotTicketStride = 64;  // in 32-bit words: every ticket on its own 256-byte line
// workspace layout in bytes (zero-initialised once by the caller; every launch leaves it zeroed again,
// except the status word):
constexpr size_t kDotWsTop = (size_t)kDotMaxBlocks * 8;                         // top-level ticket
constexpr size_t kDotWsStatus = kDotWsTop + 128;                                // != 0: a peer-memory collect timed out
constexpr size_t kDotWsGroupSums = kDotWsTop + 256;                             // T[kDotGroups]
constexpr size_t kDotWsGroupTickets = kDotWsGroupSums + 256;                    // kDotGroups tickets, 256 bytes apart
constexpr size_t kDotWorkspaceBytes = kDotWsGroupTickets + (size_t)kDotGroups * kDotTicketStride * 4;

template <typename T>
__device__ __forceinline__ T warp_sum(T v) {
#pragma unroll
  for (int o = 16; o > 0; o >>= 1) v = add_rn(v, __shfl_down_sync(0xffffffffu, v, o));
  return v;
}

// result valid in thread 0.  scratch: >= 32 elements of shared memory.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* scratch) {
  const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
  v = warp_sum(v);
  if (lane == 0) scratch[wid] = v;
  __syncthreads();
  if (wid == 0) {
    v = lane < nw ? scratch[lane] : T(0);
    v = warp_sum(v);
  }
  __syncthreads();  // scratch may be reused
  return v;
}

__device__ __forceinline__ unsigned dot_num_groups(unsigned nparts) { return nparts < (unsigned)kDotGroups ? nparts : (unsigned)kDotGroups; }

// One full warp adds the partials of group g (partials g, g+ng, g+2ng, ...).  Valid in lane 0.
template <typename T>
__device__ __forceinline__ T dot_group_sum(const T* partials, unsigned nparts, unsigned ng, unsigned g) {
  const unsigned lane = threadIdx.x & 31u;
  T v = T(0);
  for (unsigned i = g + ng * lane; i < nparts; i += ng * 32u) v = add_rn(v, __ldcg(partials + i));
  return warp_sum(v);
}
// One full warp adds the ng <= 32 group sums.  Valid in lane 0.
template <typename T>
__device__ __forceinline__ T dot_top_sum(const T* gsums, unsigned ng) {
  const unsigned lane = threadIdx.x & 31u;
  return warp_sum(lane < ng ? __ldcg(gsums + lane) : T(0));
}

// ------------------------------------------------------------------------------------------------
// Fused dot + all-reduce over NVLink peer memory (no NCCL, no extra launch).
// Every GPU owns a mailbox of kMaxPeers slots in its own HBM, mapped into all peers.  The last block
// of GPU r's dot kernel  (1) stores its partial into slot r of EVERY GPU's mailbox (peer stores travel
// over NVLink / NVSwitch),  (2) waits until all G slots of its OWN mailbox carry this call's sequence
// number,  (3) adds the G values in rank order.  Every GPU ends up with the same bits.  The G kernels
// run on G different devices at the same time, so the wait cannot deadlock on scheduling; a
// globaltimer bound (P2PArgs::timeout_ns) turns a missing peer into a NaN result plus a status word
// that the host checks, instead of a hang.
//
// Wire format: every 8-byte word of a slot carries its own flag -- {payload: 32 bits, flag = low 32
// bits of seq} -- and an aligned 8-byte store is single-copy atomic, so a word is valid exactly when
// its flag matches.  The sender needs NO release fence (which would hold the kernel until every peer
// store has been acknowledged across NVLink) and the receiver no acquire: a double travels as two such
// words, a float as one, all stored and polled with relaxed system-scope accesses.  (The layout NCCL's
// low-latency protocol uses for the same reason.)
// ------------------------------------------------------------------------------------------------
constexpr int kMaxPeers = 8;
struct alignas(16) MailSlot {
  unsigned long long w[2];  // w[i] = (flag << 32) | payload_i ; payload_0 = low half of the value's bits
};
// A mailbox is TWO banks of kMaxPeers slots, used alternately (bank = seq & 1).  GPU A can only start
// call k+2 after it has finished call k+1, which needed B's partial of k+1, which B sent after it had
// read everything of call k: so a slot of bank k&1 is never overwritten while a peer may still read it,
// however far the GPUs drift apart between calls.
constexpr int kMailBanks = 2;
constexpr size_t kMailboxBytes = sizeof(MailSlot) * kMaxPeers * kMailBanks;

__device__ __forceinline__ void st_relaxed_sys(unsigned long long* p, unsigned long long v) {
  asm volatile("st.relaxed.sys.global.u64 [%0], %1;" ::"l"(p), "l"(v) : "memory");
}
__device__ __forceinline__ unsigned long long ld_relaxed_sys(const unsigned long long* p) {
  unsigned long long v;
  asm volatile("ld.relaxed.sys.global.u64 %0, [%1];" : "=l"(v) : "l"(p) : "memory");
  return v;
}
__device__ __forceinline__ unsigned long long global_timer_ns() {
  unsigned long long t;
  asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(t));
  return t;
}

struct P2PArgs {
  MailSlot* mailbox[kMaxPeers];  // mailbox[g] = GPU g's mailbox as seen from THIS device
  unsigned long long seq;        // strictly increasing per dot call, equal on all GPUs
  unsigned long long timeout_ns; // how long the collecting warp waits for a peer before it gives up
  unsigned int* status_dev;      // device word, set to 1 when the wait timed out (never cleared by a kernel)
  unsigned int* status_host;     // same, mapped pinned host memory; may be null
  int world;                     // 0 or 1: no exchange
  int rank;
};

template <typename T> __device__ __forceinline__ unsigned long long to_bits(T v);
template <> __device__ __forceinline__ unsigned long long to_bits<double>(double v) { return (unsigned long long)__double_as_longlong(v); }
template <> __device__ __forceinline__ unsigned long long to_bits<float>(float v) { return (unsigned long long)__float_as_uint(v); }
template <typename T> __device__ __forceinline__ T from_bits(unsigned long long b);
template <> __device__ __forceinline__ double from_bits<double>(unsigned long long b) { return __longlong_as_double((long long)b); }
template <> __device__ __forceinline__ float from_bits<float>(unsigned long long b) { return __uint_as_float((unsigned)b); }

// The all-reduce has two halves, usable separately (DOT_TICKET_SEND + dot_collect_kernel) by callers that
// do not need the sum at once: send early, collect late, and the wait for the slowest GPU hides under
// whatever runs in between.
// send: lanes < world of one warp store {bits, seq} into slot `rank` of every GPU's mailbox.
template <typename T>
__device__ __forceinline__ void p2p_send_warp(const P2PArgs& p, T mine, unsigned long long seq) {
  const int lane = threadIdx.x & 31;
  const int bank = (int)(seq & 1ull) * kMaxPeers;
  if (lane < p.world) {
    MailSlot* dst = p.mailbox[lane] + bank + p.rank;   // my slot in GPU `lane`'s mailbox
    const unsigned long long bits = to_bits<T>(mine), flag = (seq & 0xffffffffull) << 32;
    st_relaxed_sys(&dst->w[0], flag | (bits & 0xffffffffull));
    if constexpr (sizeof(T) == 8) st_relaxed_sys(&dst->w[1], flag | (bits >> 32));
  }
}
// collect: one full warp waits for the `world` slots of its OWN mailbox (local HBM) to carry p.seq and
// adds them in rank order; the sum (identical on every GPU) is returned in every lane.
template <typename T>
__device__ __forceinline__ T p2p_collect_warp(const P2PArgs& p, unsigned long long seq) {
  const int lane = threadIdx.x & 31;
  const int bank = (int)(seq & 1ull) * kMaxPeers;
  bool ok = true;
  T got = T(0);
  if (lane < p.world) {
    const MailSlot* src = p.mailbox[p.rank] + bank + lane;  // GPU `lane`'s slot in my mailbox
    const unsigned long long flag = seq & 0xffffffffull;
    const unsigned long long t0 = global_timer_ns();
    unsigned long long w0, w1 = flag << 32;
    for (;;) {
      w0 = ld_relaxed_sys(&src->w[0]);
      if constexpr (sizeof(T) == 8) w1 = ld_relaxed_sys(&src->w[1]);
      if ((w0 >> 32) == flag && (w1 >> 32) == flag) break;
      if (global_timer_ns() - t0 > p.timeout_ns) { ok = false; break; }
    }
    got = from_bits<T>((w0 & 0xffffffffull) | (w1 << 32));
  }
  T total = T(0);
  for (int r = 0; r < p.world; r++) total = add_rn(total, __shfl_sync(0xffffffffu, got, r));  // rank order
  if (!__all_sync(0xffffffffu, ok)) {
    // a peer never arrived: the sum becomes NaN AND the failure is reported out of band, so that the
    // host can tell it from NaN data (b200_dot returns B200_ETIMEOUT)
    total = from_bits<T>(~0ull);
    if (lane == 0) {
      if (p.status_dev) *p.status_dev = 1u;
      if (p.status_host) asm volatile("st.relaxed.sys.global.u32 [%0], %1;" ::"l"(p.status_host), "r"(1u) : "memory");
    }
  }
  return total;
}

// Called by warp 0 of the last block; `mine` is this GPU's partial.  Returns the all-reduced sum
// (identical on every GPU) in every lane.
// Only the MODE >= 2 instantiations of dot_kernel contain this code: inlined or not, it raises the
// kernel's register allocation from 38 to 42, which drops the residency of 512-thread blocks from 3
// to 2 per SM (-3.5 % at 2^29) -- so single-GPU, NCCL and host-combine runs use the lean kernel, and
// the MODE 2/3 kernels are register-capped by their launch bound instead (the spills land in this
// once-per-launch code).  `pp` points INTO the __grid_constant__ parameter block: nothing is copied.
template <typename T, bool COLLECT>
__device__ __noinline__ T p2p_allreduce_warp(const P2PArgs* pp, T mine, unsigned long long seq) {
  p2p_send_warp<T>(*pp, mine, seq);
  if constexpr (COLLECT) return p2p_collect_warp<T>(*pp, seq);
  else return mine;
}

struct DotArgs {
  const void* a;
  const void* b;
  void* partials;        // the workspace: T[gridDim.x] partials, then tickets and group sums (kDotWs* offsets)
  void* result_dev;      // may be null
  void* result_host;     // mapped pinned, may be null
  unsigned long long nvec, ntiles, n;
  unsigned int head;
  // Work assignment.  tiles_per_block = k > 0: LOGICAL block L owns the contiguous tiles [L*k, (L+1)*k) and always adds
  // them in ascending order; k = 0: grid-stride (tiles L, L+G, ...).  sweep.reverse maps dispatch index p to L = G-1-p,
  // i.e. changes WHEN a chunk is read, never what a thread adds in which order: the result has the same bits in both
  // directions.  Blocks p >= sweep.keep_from load with evict_normal (H_EVICT_FIRST only).
  unsigned int tiles_per_block;
  SweepArgs sweep;
  unsigned long long* bell_flag;  // doorbell (mapped pinned host word) rung after the result is stored, or nullptr
  unsigned long long bell_seq;
  // Sequence numbers kept ON THE DEVICE (LEVEL 2): when non-null, the doorbell value is ++seq_counters[0] and the
  // all-reduce sequence number ++seq_counters[1] instead of bell_seq / p2p.seq -- nothing in the parameter
  // block changes from call to call, so the launch can be replayed from a CUDA graph.
  unsigned long long* seq_counters;
  P2PArgs p2p;           // world <= 1: plain single-GPU dot
};

// dot blocks have at most 512 threads: the streaming shapes are held to 42 registers so that three such blocks fit an
// SM (launch bound 512 x 3), the widest tiles (2 operands x 128 bytes per thread) get the 128 registers they need.
constexpr int dot_max_block(int vb, int unroll) { return (void)vb, (void)unroll, 512; }

// MODE: how the block partials become the result
enum DotMode : int { DOT_TWO_PASS = 0,    // partials only; dot_final_kernel follows
                     DOT_TICKET = 1,      // the last block to finish runs the second pass
                     DOT_TICKET_P2P = 2,  // ... and then the all-reduce over peer memory
                     DOT_TICKET_SEND = 3 };// ... and then only SENDS its partial to the peers (dot_collect_kernel sums later)
// the P2P instantiations run with at most 512 threads and must keep 3 blocks of them resident
constexpr int dot_block_bound(int vb, int unroll, int mode) {
  return mode >= DOT_TICKET_P2P ? 512 : dot_max_block(vb, unroll);
}
// Three 512-thread blocks per SM need <= 42 registers; the once-per-launch tail (tickets, group sums, all-reduce) would
// otherwise push the streaming shapes to 44-48 and cost a third of the occupancy (-3.5 % at 2^29).
constexpr int dot_min_blocks(int vb, int unroll, int mode) { return (void)mode, vb * unroll <= 32 ? 3 : 1; }

template <typename T, int VB, int UNROLL, int HINT, bool FUSED, int MODE>
__global__ void __launch_bounds__(dot_block_bound(VB, UNROLL, MODE), dot_min_blocks(VB, UNROLL, MODE))
dot_kernel(const __grid_constant__ DotArgs args) {
  constexpr int EPV = VB / (int)sizeof(T);
  __shared__ T scratch[32];
  __shared__ bool group_last;
  pdl_entry();

  const T* const a = static_cast<const T*>(args.a);
  const T* const b = static_cast<const T*>(args.b);
  const char* const va = reinterpret_cast<const char*>(a + args.head);
  const char* const vb = reinterpret_cast<const char*>(b + args.head);
  const unsigned long long nvec = args.nvec;
  const unsigned long long tile_vecs = (unsigned long long)blockDim.x * UNROLL;

  uint64_t policy = make_policy<VB, HINT>();
  if constexpr (HINT == H_EVICT_FIRST) { if (blockIdx.x >= args.sweep.keep_from) policy = l2_policy<L2_EVICT_NORMAL>(); }
  const unsigned lb = args.sweep.reverse ? gridDim.x - 1u - blockIdx.x : blockIdx.x;  // logical block
  const unsigned long long k = args.tiles_per_block;
  const unsigned long long first = k ? (unsigned long long)lb * k : lb, step = k ? 1ull : gridDim.x;
  const unsigned long long last = k ? (first + k < args.ntiles ? first + k : args.ntiles) : args.ntiles;
  T acc[UNROLL][EPV];
#pragma unroll
  for (int j = 0; j < UNROLL; j++)
#pragma unroll
    for (int e = 0; e < EPV; e++) acc[j][e] = T(0);

  for (unsigned long long tile = first; tile < last; tile += step) {
    const unsigned long long v0 = tile * tile_vecs + threadIdx.x;
    const bool full = (tile + 1) * tile_vecs <= nvec;
    Words<VB> wa[UNROLL], wb[UNROLL];
#pragma unroll
    for (int j = 0; j < UNROLL; j++) {
      const unsigned long long v = v0 + (unsigned long long)j * blockDim.x;
      if (full || v < nvec) {
        wa[j] = ldg<VB, HINT, true>(va + v * VB, policy);
        wb[j] = ldg<VB, HINT, true>(vb + v * VB, policy);
      }
    }
#pragma unroll
    for (int j = 0; j < UNROLL; j++) {
      const unsigned long long v = v0 + (unsigned long long)j * blockDim.x;
      if (full || v < nvec) {
        const Elems<T, VB> ea = unpack<T, VB>(wa[j]), eb = unpack<T, VB>(wb[j]);
#pragma unroll
        for (int e = 0; e < EPV; e++) {
          if constexpr (FUSED) acc[j][e] = fma_rn(ea.v[e], eb.v[e], acc[j][e]);
          else acc[j][e] = add_rn(acc[j][e], mul_rn(ea.v[e], eb.v[e]));
        }
      }
    }
  }

  // fixed-order fold of the per-thread chains
  T sum = T(0);
#pragma unroll
  for (int j = 0; j < UNROLL; j++)
#pragma unroll
    for (int e = 0; e < EPV; e++) sum = add_rn(sum, acc[j][e]);

  // ragged edges, logical block 0 only
  if (lb == 0) {
    const unsigned long long tail_begin = args.head + nvec * EPV;
    const unsigned long long edge = args.head + (args.n - tail_begin);
    if (threadIdx.x < edge) {
      const unsigned long long i = threadIdx.x < args.head ? threadIdx.x : tail_begin + (threadIdx.x - args.head);
      if constexpr (FUSED) sum = fma_rn(a[i], b[i], sum);
      else sum = add_rn(sum, mul_rn(a[i], b[i]));
    }
  }

  sum = block_sum(sum, scratch);
  T* const partials = static_cast<T*>(args.partials);
  if (threadIdx.x == 0) partials[lb] = sum;

  if constexpr (MODE != DOT_TWO_PASS) {
    char* const ws = reinterpret_cast<char*>(args.partials);
    unsigned int* const top_ticket = reinterpret_cast<unsigned int*>(ws + kDotWsTop);
    T* const gsums = reinterpret_cast<T*>(ws + kDotWsGroupSums);
    const unsigned ng = dot_num_groups(gridDim.x), g = lb % ng;
    unsigned int* const my_ticket = reinterpret_cast<unsigned int*>(ws + kDotWsGroupTickets) + g * kDotTicketStride;
    const unsigned gsize = (gridDim.x - g + ng - 1) / ng;  // blocks in my group
    if (threadIdx.x == 0) {
      __threadfence();  // partial visible before the ticket is taken
      group_last = atomicAdd(my_ticket, 1u) == gsize - 1;
    }
    __syncthreads();
    if (group_last && threadIdx.x < 32) {  // block-uniform; warp 0 of the last block of group g
      __threadfence();  // acquire side: the group's partials
      const T gs = dot_group_sum<T>(partials, gridDim.x, ng, g);
      unsigned last = 0;
      if (threadIdx.x == 0) {
        gsums[g] = gs;
        *my_ticket = 0u;  // ready for the next launch (stream order separates the launches)
        __threadfence();
        last = atomicAdd(top_ticket, 1u) == ng - 1;
      }
      last = __shfl_sync(0xffffffffu, last, 0);
      if (last) {
        __threadfence();
        T v = dot_top_sum<T>(gsums, ng);
        unsigned long long bell_seq = args.bell_seq, p2p_seq = args.p2p.seq;
        if (threadIdx.x == 0) {
          *top_ticket = 0u;
          if (args.seq_counters) {  // this thread is the only one of the grid that touches the counters
            bell_seq = args.seq_counters[0] + 1; args.seq_counters[0] = bell_seq;
            if constexpr (MODE >= DOT_TICKET_P2P) { p2p_seq = args.seq_counters[1] + 1; args.seq_counters[1] = p2p_seq; }
          }
        }
        if constexpr (MODE >= DOT_TICKET_P2P) {  // fused all-reduce over peer memory (or its sending half)
          p2p_seq = __shfl_sync(0xffffffffu, p2p_seq, 0);
          v = p2p_allreduce_warp<T, MODE == DOT_TICKET_P2P>(&args.p2p, __shfl_sync(0xffffffffu, v, 0), p2p_seq);
        }
        if (threadIdx.x == 0) {
          if (args.result_dev) *static_cast<T*>(args.result_dev) = v;
          if (args.result_host) *static_cast<T*>(args.result_host) = v;
          if (args.bell_flag) ring_host(args.bell_flag, bell_seq);  // release: after the result
        }
      }
    }
  }
}

// The collecting half of the split all-reduce: one warp.
template <typename T>
__global__ void __launch_bounds__(32)
dot_collect_kernel(const __grid_constant__ P2PArgs p, T* result_dev, T* result_host, unsigned long long* bell_flag,
                   unsigned long long bell_seq) {
  pdl_entry();
  const T v = p2p_collect_warp<T>(p, p.seq);
  if (threadIdx.x == 0) {
    if (result_dev) *result_dev = v;
    if (result_host) *result_host = v;
    if (bell_flag) ring_host(bell_flag, bell_seq);
  }
}

// Second launch of the two-pass variant: warp g adds group g, then warp 0 adds the group sums -- the same
// two functions, hence the same bits, as the ticket variant.
template <typename T>
__global__ void __launch_bounds__(1024)
dot_final_kernel(const T* partials, unsigned nparts, T* result_dev, T* result_host, unsigned long long* bell_flag,
                 unsigned long long bell_seq, unsigned long long* seq_counters) {
  __shared__ T gs[kDotGroups];
  pdl_entry();
  const unsigned ng = dot_num_groups(nparts), w = threadIdx.x >> 5;
  if (w < ng) {
    const T v = dot_group_sum<T>(partials, nparts, ng, w);
    if ((threadIdx.x & 31u) == 0) gs[w] = v;
  }
  __syncthreads();
  if (w == 0) {
    const unsigned lane = threadIdx.x;
    const T v = warp_sum(lane < ng ? gs[lane] : T(0));
    if (lane == 0) {
      if (seq_counters) { bell_seq = seq_counters[0] + 1; seq_counters[0] = bell_seq; }  // see DotArgs::seq_counters
      if (result_dev) *result_dev = v;
      if (result_host) *result_host = v;
      if (bell_flag) ring_host(bell_flag, bell_seq);
    }
  }
}

}  // namespace b200
