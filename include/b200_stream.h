/*
 * b200_stream.h -- the C ABI of libb200stream.so: hand-written sm_100a BabelStream kernels.
 *
 * This is the drop-in boundary.  The C++ host class B200Stream<T> (babelstream_b200/src/b200/)
 * implements the reference's Stream<T> interface (/root/reference/src/Stream.h:29-49) purely in
 * terms of the LEVEL 2 calls below and is driven by the reference's unmodified src/main.cpp.
 * Python (babelstream_b200/stream.py, ctypes), cgo, JNI ... bind the same symbols.
 *
 * Conventions
 *   - plain pointers and sizes only; no C++ types, no exceptions cross this boundary;
 *   - every call returns B200_OK (0) or a negative B200_E* code; b200_last_error() gives the
 *     message of the calling thread's last failure;
 *   - elem_bytes is 4 (float) or 8 (double): the two types Stream<T> is instantiated for
 *     (src/cuda/CUDAStream.cu:349-350);
 *   - scalars cross as double and are rounded to T exactly once inside, which is what the
 *     reference's `const T scalar = startScalar;` does (src/serial/SerialStream.cpp:68,87,97);
 *   - there is NO CPU fallback: without a CUDA device the compute entry points fail with
 *     B200_ENODEVICE.
 */
#ifndef B200_STREAM_H
#define B200_STREAM_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define B200_OK 0
#define B200_EINVAL (-1)    /* bad argument */
#define B200_ENODEVICE (-2) /* no CUDA device / bad device index (reference: runtime_error "Invalid device index") */
#define B200_ENOMEM (-3)    /* arrays do not fit (reference: "Device does not have enough memory for all buffers") */
#define B200_ECUDA (-4)     /* a CUDA runtime call failed (reference: CU() macro, CUDAStream.cu:14-26) */
#define B200_ENCCL (-5)     /* NCCL missing or failed (multi-GPU dot only) */
#define B200_EUNSUPPORTED (-6)
#define B200_ETIMEOUT (-7)  /* the peer-memory all-reduce gave up waiting for a rank (multi-GPU dot); the sum is NaN */

/* Kernel ids: the reference's BenchId values (src/benchmark.h:18) plus init. */
enum { B200_COPY = 0, B200_MUL = 1, B200_ADD = 2, B200_TRIAD = 3, B200_NSTREAM = 4, B200_DOT = 5, B200_INIT = 6,
       B200_NUM_KERNELS = 7 };

/* cache_hint values */
enum { B200_HINT_AUTO = 0,
       B200_HINT_DEFAULT = 1,     /* ld.global / st.global */
       B200_HINT_NOALLOC = 2,     /* ld.global[.nc].L1::no_allocate / st.global.L1::no_allocate */
       B200_HINT_EVICT_FIRST = 3, /* + .L2::evict_first on loads and stores */
       B200_HINT_STREAM = 4,      /* ld.global.cs / st.global.cs */
       B200_HINT_PREFETCH = 5     /* L1::no_allocate + .L2::256B prefetch on loads */ };

/* variant values */
enum { B200_VARIANT_AUTO = 0,
       B200_VARIANT_LDST = 1,     /* register path: vector LDG/STG (all kernels) */
       B200_VARIANT_BULK = 2,     /* copy: cp.async.bulk global->shared->global ring (UBLKCP) */
       B200_VARIANT_TMA = 3,      /* copy: cp.async.bulk.tensor 2-D tiles through a CUtensorMap (UTMALDG/UTMASTG) */
       B200_VARIANT_DOT_TWO_PASS = 1, /* dot: block partials, then a 1-block second launch */
       B200_VARIANT_DOT_TICKET = 2    /* dot: same second pass run by the last block to finish, one launch */ };

/* Launch configuration.  A zeroed struct (or NULL) means "auto": the per-size table in
 * csrc/tuning.cu.  The reference's counterparts are compile-time: TBSIZE 256, UNROLL_FACTOR 4,
 * GRID_STRIDE (src/cuda/CUDAStream.h:18, model.cmake:6-22). */
typedef struct b200_launch_cfg {
  int vec_bytes;  /* 0 auto | 8 = 64-bit | 16 = 128-bit | 32 = 256-bit accesses (4: float scalar path) */
  int unroll;     /* 0 auto | 1,2,4,8 vectors per thread per tile (8 only with 16-byte vectors) */
  int block;      /* 0 auto | threads per block, multiple of 32: up to 1024 while vec_bytes*unroll <= 32, else up to
                     512; dot: up to 512 */
  int grid;       /* 0 auto | >0 exact block count (grid-stride) | <0: -k means k blocks per SM |
                     INT32_MAX: one tile per block */
  int cache_hint; /* B200_HINT_* */
  int variant;    /* B200_VARIANT_* */
  int contract;   /* 0 auto (=fused) | 1 fused: fma(s,c,b), the reference's default -ffp-contract=fast build |
                     2 unfused: round(s*c) then add, the reference built -ffp-contract=off */
  int stages;     /* bulk/tma copy: ring depth (0 auto) */
  int chunk_bytes;/* bulk/tma copy: bytes per stage (0 auto) */
  int flags;      /* B200_LAUNCH_* bits */
  int keep_mb;    /* 0 auto | -1 none | > 0: MB of traffic at the END of the sweep whose lines are kept in L2 (loads
                     evict_normal, stores evict_last instead of evict_first) for the next kernel; evict-first hint only */
  int reserved[1];
} b200_launch_cfg;

/* flags */
#define B200_LAUNCH_PDL 1 /* programmatic dependent launch: the kernel may be SCHEDULED before the previous kernel of
                             the stream has drained; it waits (griddepcontrol.wait) for that kernel's completion and
                             memory before its first access, so stream-order semantics are unchanged.  Removes the
                             ~2 us drain-and-launch gap between back-to-back kernels of an asynchronous sequence. */
#define B200_LAUNCH_REVERSE 2 /* sweep the arrays from the top down (block p handles tile ntiles-1-p).  Results are
                             bit-identical in both directions, dot included.  A caller that alternates the direction of
                             consecutive kernels lets each one start on the data its predecessor touched last, i.e. on
                             what is still in the 126 MB L2; LEVEL 2 does so by default (option sweep=alternate). */
#define B200_LAUNCH_DOT_GRIDSTRIDE 4 /* dot: round-1 work assignment (block b adds tiles b, b+G, ...) instead of one
                             contiguous chunk per block; different summation order, hence different low-order bits */

/* ------------------------------------------------------------------------------------------
 * LEVEL 0 -- library / device queries (replace listDevices / getDeviceName / getDeviceDriver,
 * src/Stream.h:51-54, src/cuda/CUDAStream.cu:311-347).  Safe to call without a GPU.
 * ------------------------------------------------------------------------------------------ */
const char* b200_version(void);
const char* b200_last_error(void);
int b200_device_count(void);                           /* >= 0, or a negative code */
int b200_device_name(int device, char* buf, int buflen);
int b200_driver_version(int device);                   /* cudaDriverGetVersion value, or negative */
int b200_device_info(int device, int* num_sms, int* cc_major, int* cc_minor, size_t* total_mem,
                     int* l2_bytes, int* mem_clock_khz, int* mem_bus_bits);
int b200_device_ecc(int device);                       /* 1 = ECC on, 0 = off (banner: CUDAStream.cu:108-109) */
int b200_device_mem_clock_mhz(int device);             /* NVML's current memory clock as the reference prints it
                                                          (CUDAStream.cu:96-108); the attribute's value without NVML */
uint64_t b200_launch_count(void);                      /* kernels launched by this library so far */

/* Host-side logic, usable (and tested) without a GPU. */
/* Contiguous shard of GPU g out of ngpus: [begin, begin+count).  Shard boundaries are multiples of
 * 128 bytes so every shard starts on a full cache line.  This is the layout of mem=device
 * contexts; mem=managed and mem=pagefault contexts with shards of 8 MiB or more put the boundaries on
 * multiples of 2 MiB (the managed page size) instead -- ask b200_shard_info for the shards of a context. */
int b200_shard_range(size_t n, int elem_bytes, int ngpus, int g, size_t* begin, size_t* count);
/* Resolve "auto" fields of `in` for (kernel, n, elem_bytes) on a GPU with num_sms SMs and l2_bytes of L2:
 * fills `out`, and the resulting grid size / tile count. */
int b200_resolve_cfg(int kernel, size_t n, int elem_bytes, int num_sms, int l2_bytes,
                     const b200_launch_cfg* in, b200_launch_cfg* out, unsigned* grid_out);
/* The walk a launch with this configuration makes over its tiles (vector-aligned arrays assumed), for inspection and
 * tests.  Elementwise kernels: dispatch index p in [0, ntiles) -- block b takes p = b, b + grid, ... -- starts at vector
 * origin + stride*p and covers tile_vecs vectors of vec_bytes bytes; p >= keep_from keeps its lines in L2
 * (keep_from = ~0: none).  dot: logical block L in [0, grid) owns tiles [L*tiles_per_block, (L+1)*tiles_per_block)
 * (tiles_per_block = 0: L, L + grid, ...); reverse only changes the order in which the logical blocks are dispatched. */
typedef struct b200_sweep_plan {
  unsigned grid, tiles_per_block;
  int reverse, vec_bytes;
  int unroll, cache_hint; /* as launched: 4 / default when the scalar path is taken (vec_bytes == elem_bytes) */
  unsigned long long keep_from, origin;
  long long stride;
  unsigned long long ntiles, nvec, tile_vecs;
} b200_sweep_plan;
int b200_plan_sweep(int kernel, size_t n, int elem_bytes, int num_sms, int l2_bytes, const b200_launch_cfg* in,
                    b200_sweep_plan* out);

/* ------------------------------------------------------------------------------------------
 * LEVEL 1 -- raw launchers: one kernel on caller-owned DEVICE pointers, asynchronous on
 * `stream` (a cudaStream_t passed as void*; NULL = the legacy default stream) of the CURRENT
 * device.  Any pointer alignment and any n >= 0 are accepted.  The array a kernel writes must not overlap the
 * arrays it reads (they are loaded through the read-only path): in-place calls return B200_EINVAL.
 * Each replaces the body of the reference model's method of the same name:
 *   init_arrays CUDAStream.cu:203-211   copy :237-243   mul :245-251   add :253-259
 *   triad :261-267   nstream :269-275   dot :277-309
 * ------------------------------------------------------------------------------------------ */
int b200_launch_init(void* a, void* b, void* c, size_t n, int elem_bytes, double init_a, double init_b,
                     double init_c, const b200_launch_cfg* cfg, void* stream);
int b200_launch_copy(const void* a, void* c, size_t n, int elem_bytes, const b200_launch_cfg* cfg, void* stream);
int b200_launch_mul(void* b, const void* c, size_t n, int elem_bytes, double scalar, const b200_launch_cfg* cfg,
                    void* stream);
int b200_launch_add(const void* a, const void* b, void* c, size_t n, int elem_bytes, const b200_launch_cfg* cfg,
                    void* stream);
int b200_launch_triad(void* a, const void* b, const void* c, size_t n, int elem_bytes, double scalar,
                      const b200_launch_cfg* cfg, void* stream);
int b200_launch_nstream(void* a, const void* b, const void* c, size_t n, int elem_bytes, double scalar,
                        const b200_launch_cfg* cfg, void* stream);
/* dot: `workspace` is device memory of at least b200_dot_workspace_bytes() bytes, zero-initialised
 * once by the caller (the kernel leaves it zeroed again).  The sum (type T) is written to
 * result_dev (device pointer, may be NULL) and to result_host_mapped (device-visible address of
 * mapped pinned host memory, may be NULL).  Deterministic: bit-identical for equal
 * (n, cfg, pointer alignment). */
size_t b200_dot_workspace_bytes(void);
int b200_launch_dot(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* result_dev,
                    void* result_host_mapped, const b200_launch_cfg* cfg, void* stream);

/* dot WITHOUT a caller-supplied workspace, and without any allocation inside the library.  This one has no counterpart
 * in the reference (its model allocates the block sums in its constructor, CUDAStream.cu:128-145); it exists for the
 * callers of this ABI that cannot give b200_launch_dot the zeroed cudaMalloc'ed workspace it asks for:
 *   - bindings that are handed device addresses by their host program but do not link the CUDA runtime themselves,
 *     so they have nothing to allocate with;
 *   - services that keep the allocator out of their steady state, because cudaMalloc/cudaFree synchronise the whole
 *     device and stall every other stream;
 *   - processes whose device memory was handed out completely at start-up, where a later allocation just fails.
 * The block partials, tickets and group sums live in an array that is part of the library image instead
 * (B200_DOT_STATIC_SCRATCH_BYTES per GPU; every GPU that loads the library's code keeps its copy until the process
 * exits).  Same kernel, same summation order, hence the same bits as b200_launch_dot for equal (n, cfg, alignment).
 * Rules:
 *   - at most B200_DOT_STATIC_MAX_BLOCKS blocks: the automatic grid never comes near it, whatever n; a cfg whose grid
 *     asks for more (an explicit block count on a long vector) is refused with B200_EINVAL;
 *   - all calls on one GPU share the one array: calls that may overlap on the GPU (different streams with no ordering
 *     between them) race and return garbage -- keep them on one stream or order them with events; concurrent dots need
 *     b200_launch_dot with a workspace each;
 *   - the first call on each GPU looks the array's device address up (cudaGetSymbolAddress with that GPU current) and
 *     caches it: make that first call before any stream capture begins;
 *   - the ticket region is cleared on `stream` ahead of every launch (cudaMemsetAsync), so a previous call that died
 *     half-way cannot poison the next one;
 *   - at least one of result_dev / result_host_mapped must be given.
 * Calls no cudaMalloc, cudaFree or other allocating routine. */
#define B200_DOT_STATIC_MAX_BLOCKS 16384
#define B200_DOT_STATIC_SCRATCH_BYTES 139776
int b200_launch_dot_static(const void* a, const void* b, size_t n, int elem_bytes, void* result_dev,
                           void* result_host_mapped, const b200_launch_cfg* cfg, void* stream);

/* Fused dot + all-reduce over NVLink peer memory: ONE launch per GPU, no NCCL.  GPU `rank` of `world`
 * (<= 8) computes its partial as b200_launch_dot does; the last block then stores it into slot `rank`
 * of every GPU's mailbox (peer stores), waits for the `world` partials of this call in its own
 * mailbox and adds them in rank order, so every GPU obtains the same bits.  mailboxes[g] is GPU g's
 * mailbox (b200_dot_mailbox_bytes() of zero-initialised device memory on GPU g, peer access enabled
 * from the current device); `seq` must be non-zero, equal on all GPUs and strictly increasing from
 * call to call.  All `world` launches must be issued (on their own devices) for any to complete;
 * a peer that never arrives turns the result into NaN after the timeout below instead of hanging. */
size_t b200_dot_mailbox_bytes(void);
int b200_launch_dot_allreduce_p2p(const void* a, const void* b, size_t n, int elem_bytes, void* workspace,
                                  void* result_dev, void* result_host_mapped, int world, int rank,
                                  unsigned long long seq, void* const* mailboxes, const b200_launch_cfg* cfg,
                                  void* stream);

/* The same all-reduce in two halves, for ASYNCHRONOUS callers that do not need the sum at once: `send`
 * is the dot kernel whose last block stores the partial into every peer's mailbox and returns without
 * waiting (partial_dev, may be NULL, receives this GPU's partial); `collect` is a one-warp kernel that
 * waits for the `world` partials of call `seq` in the GPU's own mailbox, adds them in rank order (same
 * bits as the fused call) and stores the sum.  Whatever is enqueued between the two hides the wait for
 * the slowest GPU.  Rule: on every GPU, collect(seq) must be enqueued before the next send (the
 * mailbox has two banks). */
int b200_launch_dot_send_p2p(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* partial_dev,
                             int world, int rank, unsigned long long seq, void* const* mailboxes,
                             const b200_launch_cfg* cfg, void* stream);
int b200_launch_dot_collect_p2p(int elem_bytes, void* result_dev, void* result_host_mapped, int world, int rank,
                                unsigned long long seq, void* const* mailboxes, void* workspace /* may be NULL */,
                                int flags /* B200_LAUNCH_* */, void* stream);
/* A rank that never sends turns the sum into NaN after the timeout (default 2000 ms, BABELSTREAM_B200_P2P_TIMEOUT_MS)
 * AND sets the 32-bit word at byte b200_dot_status_offset() of the dot workspace to 1 (never cleared by a kernel), so
 * that the failure can be told from NaN data; LEVEL 2's b200_dot returns B200_ETIMEOUT. */
size_t b200_dot_status_offset(void);
int b200_set_p2p_timeout_ms(double ms);

/* Completion doorbell: lets a BLOCKING caller learn that a kernel has finished without
 * cudaStreamSynchronize.  Every block takes a ticket once its stores are fenced; the block that takes
 * the last one stores `seq` (release, system scope) to a word of mapped pinned host memory, which the
 * caller spins on.  Whatever is enqueued next on the stream is still ordered after the whole kernel.
 * `ticket` is one zero-initialised device word per stream (left zero; dot ignores it and uses the
 * ticket in its workspace), `seq` must be non-zero and differ from the flag's current value.
 * This is what LEVEL 2 does for dot (always: -1.3..1.9 us per call, the host no longer waits for the
 * end-of-kernel flush of dot's result write) and, under option sync=doorbell, for every kernel (not the
 * default: an elementwise kernel gains 0.2 us at best and pays 1-3 us for ~1000 same-address tickets). */
typedef struct b200_doorbell {
  void* ticket;
  void* flag_host_mapped;
  unsigned long long seq;
} b200_doorbell;
/* `kernel` is one of B200_COPY/MUL/ADD/TRIAD/NSTREAM/INIT; a, b, c as in the launchers above (unused
 * ones may be NULL); s0 is the scalar (mul/triad/nstream) or init_a, s1/s2 are init_b/init_c. n > 0. */
int b200_launch_doorbell(int kernel, void* a, void* b, void* c, size_t n, int elem_bytes, double s0, double s1,
                         double s2, const b200_launch_cfg* cfg, const b200_doorbell* bell, void* stream);
int b200_launch_dot_doorbell(const void* a, const void* b, size_t n, int elem_bytes, void* workspace, void* result_dev,
                             void* result_host_mapped, const b200_launch_cfg* cfg, const b200_doorbell* bell,
                             void* stream);

/* Peer memory across PROCESSES (one process per GPU, e.g. under torchrun / mpirun): every rank creates
 * its mailbox, exports it, exchanges the 64-byte handles by whatever means it has (torch.distributed,
 * MPI, a file), opens its peers' handles and passes the resulting pointers -- its own mailbox at index
 * `rank` -- to b200_launch_dot_allreduce_p2p or b200_join_group.  Thin wrappers of cudaMalloc and
 * cudaIpc{GetMemHandle,OpenMemHandle,CloseMemHandle}; the GPUs must have peer access. */
#define B200_IPC_HANDLE_BYTES 64
int b200_dot_mailbox_create(void** mailbox);   /* on the current device, zeroed */
int b200_dot_mailbox_destroy(void* mailbox);
int b200_ipc_export(void* dev_ptr, void* handle_out /* B200_IPC_HANDLE_BYTES */);
int b200_ipc_open(const void* handle, void** dev_ptr);
int b200_ipc_close(void* dev_ptr);

/* ------------------------------------------------------------------------------------------
 * LEVEL 2 -- the Stream<T> object: three device-resident arrays of n elements, sharded
 * contiguously over `ngpus` devices first_device .. first_device+ngpus-1, blocking calls
 * ("These must be blocking calls", src/Stream.h:36).  One caller thread at a time.
 * ------------------------------------------------------------------------------------------ */
typedef struct b200_ctx b200_ctx;

/* Replaces the model constructor (CUDAStream.cu:80-149) minus its init_arrays call, which the
 * host class issues itself.  ngpus = 0 reads BABELSTREAM_B200_NGPUS (default 1).  Returns NULL on
 * failure (see b200_last_error / *status). */
b200_ctx* b200_create(int first_device, int ngpus, size_t n, int elem_bytes, int* status);
/* Same, with options applied BEFORE the arrays are allocated: `options` is a NULL-terminated array of
 * "key=value" strings (keys of b200_set_option) or NULL; they override the environment.  The only way
 * besides BABELSTREAM_B200_MEM to select mem=managed or mem=pagefault. */
b200_ctx* b200_create_opts(int first_device, int ngpus, size_t n, int elem_bytes, const char* const* options,
                           int* status);
void b200_destroy(b200_ctx* ctx);

/* One process per GPU: makes a single-GPU context rank `rank` of a group of `world` contexts living in
 * other processes, each holding its own contiguous shard.  From then on b200_dot returns the sum over
 * ALL ranks, computed by the fused peer-memory all-reduce inside the dot kernel (no NCCL, no extra
 * launch); every rank must call b200_dot the same number of times.  mailboxes[g] is rank g's mailbox as
 * mapped into this process (see b200_ipc_open); the array is copied.  world = 1 leaves the group.
 * The call counter restarts at every join, so the mailboxes of a NEW group must be fresh (zeroed) ones. */
int b200_join_group(b200_ctx* ctx, int world, int rank, void* const* mailboxes);

/* "key=value" tuning / behaviour options, also read from the environment at create time as
 * BABELSTREAM_B200_<KEY>.  Keys: contract=fast|off, copy_variant=auto|ldst|bulk|tma,
 * dot_variant=auto|two_pass|ticket, dot_combine=auto|p2p|nccl|host (multi-GPU dot: fused peer-memory
 * all-reduce inside the dot kernel | one ncclAllReduce | rank-order sum on the host; auto = host: inside one
 * process the caller is the only consumer of the sum and polls the G doorbells itself -- process groups joined with
 * b200_join_group always use the fused all-reduce), timing=0|1,
 * sync=auto|stream|spin|doorbell (how a blocking call waits: cudaStreamSynchronize | cudaStreamQuery loop |
 * the kernel rings a host flag, see b200_doorbell; auto = doorbell for dot, cudaStreamSynchronize for the rest),
 * mem=device|managed|pagefault (allocation time only: b200_create_opts or the environment; managed = the reference's
 * MEM=MANAGED build, src/cuda/model.cmake:6-10: one cudaMallocManaged range per array, each shard's
 * preferred location its GPU, shard boundaries on 2 MiB; b200_get_arrays then returns the ranges themselves
 * after prefetching them to the CPU, "Unified memory: return pointers to device memory" CUDAStream.cu:217-221;
 * pagefault = the reference's MEM=PAGEFAULT build, CUDAStream.cu:39-40: one 2 MiB-aligned system allocation
 * per array that the GPUs reach through page faults (HMM/ATS); B200_EUNSUPPORTED where
 * cudaDevAttrPageableMemoryAccess is 0),
 * launch=graph|direct (graph, the default: every call replays a one-kernel CUDA graph captured on first use --
 * cudaGraphLaunch + wait returns 2.0-2.5 us sooner than <<<>>> + wait on B200),
 * sweep=alternate|forward (alternate, the default: consecutive kernel calls sweep the arrays in opposite directions,
 * see B200_LAUNCH_REVERSE),
 * pin=auto|off (create time only, ngpus > 1: the caller's thread drives GPU 0, one worker thread drives each other
 * GPU; auto, the default, pins every worker to a core of its own among the CPUs local to its GPU),
 * tune=<kernel>:<field>=<v>,...;...   (fields: vec unroll block grid hint stages chunk keep flags) */
int b200_set_option(b200_ctx* ctx, const char* key, const char* value);
int b200_describe(b200_ctx* ctx, char* buf, int buflen); /* human-readable configuration, for the model's banner */

/* Makes the FIRST call of every kernel as cheap as the later ones: runs copy, mul, add, triad, dot, nstream once in each
 * sweep direction with this scalar, so that their code is loaded (CUDA loads a kernel lazily at its first launch:
 * 1.5-4 ms per elementwise kernel, 37 ms for dot) and their launch plans are captured.  The driver's Classic order
 * drops the first iteration from its statistics, but --order Isolated (src/main.cpp:168-175) averages over ALL calls of
 * a kernel; the model's constructor therefore calls this.  The arrays' contents are undefined afterwards: call it before
 * init_arrays / write_arrays. */
int b200_warmup(b200_ctx* ctx, double scalar);

int b200_init_arrays(b200_ctx* ctx, double init_a, double init_b, double init_c); /* Stream::init_arrays */
int b200_copy(b200_ctx* ctx);                                                     /* Stream::copy     c = a */
int b200_mul(b200_ctx* ctx, double scalar);                                       /* Stream::mul      b = s*c */
int b200_add(b200_ctx* ctx);                                                      /* Stream::add      c = a+b */
int b200_triad(b200_ctx* ctx, double scalar);                                     /* Stream::triad    a = b+s*c */
int b200_nstream(b200_ctx* ctx, double scalar);                                   /* Stream::nstream  a += b+s*c */
int b200_dot(b200_ctx* ctx, double* result);                                      /* Stream::dot      sum a*b; value of type T widened to double */
int b200_run(b200_ctx* ctx, int kernel, double scalar, double* result);           /* dispatch by kernel id */
/* `count` blocking calls in the given order, each launched and waited for before the next -- the driver's loop over
 * bench[] (src/main.cpp:150-165) in one call, for hosts with a costly foreign-function call.  *last_dot: the last dot. */
int b200_run_sequence(b200_ctx* ctx, const int* kernels, int count, double scalar, double* last_dot);

/* Stream::get_arrays (CUDAStream.cu:213-235): gathers all shards into library-owned host buffers
 * (pinned when they fit) that stay valid until b200_destroy. */
int b200_get_arrays(b200_ctx* ctx, const void** a, const void** b, const void** c);
/* Same gather into caller buffers of n*elem_bytes each (any may be NULL). */
int b200_read_arrays(b200_ctx* ctx, void* a, void* b, void* c);
/* Scatter caller host buffers to the shards (test and end-to-end use; any may be NULL). */
int b200_write_arrays(b200_ctx* ctx, const void* a, const void* b, const void* c);

/* Introspection */
int b200_num_gpus(b200_ctx* ctx);
const char* b200_memory_mode(b200_ctx* ctx); /* "DEFAULT" | "MANAGED" | "PAGEFAULT": the reference's "Memory Allocation:" banner line */
size_t b200_array_size(b200_ctx* ctx);
int b200_elem_bytes(b200_ctx* ctx);
/* Device time of the last kernel call in milliseconds, max over GPUs, measured with CUDA events
 * on the launching streams.  Needs option timing=1; returns a negative value otherwise. */
double b200_last_kernel_ms(b200_ctx* ctx);
/* Device pointers and stream of shard g, so that callers can time with their own events. */
int b200_shard_info(b200_ctx* ctx, int g, int* device, void** a, void** b, void** c, size_t* begin, size_t* count,
                    void** stream);

#ifdef __cplusplus
}
#endif
#endif /* B200_STREAM_H */
