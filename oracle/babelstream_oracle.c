/*
 * babelstream_oracle.c -- TEST INFRASTRUCTURE, NOT PRODUCT CODE.
 *
 * A plain-C restatement of the BabelStream kernels as the reference's serial and
 * OpenMP models compute them.  Only tests/, __graft_entry__.smoke() and the
 * cpu_baseline / --impl reference legs of bench.py may load this library, and only
 * as the checker (or the timed CPU baseline) -- never as a fallback for the CUDA path.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks every function here against
 *   (1) the known-answer bit patterns in tests/golden/known_answers.json, produced by
 *       linking the UNMODIFIED /root/reference/src/serial/SerialStream.cpp
 *       (tests/golden/make_golden.py, via oracle/_ref/libref_serial_*.so), and
 *   (2) oracle/_ref/libref_serial_{fast,off}.so directly on seeded random inputs
 *       whenever those artefacts are present.
 *
 * Reference lines restated (all relative to /root/reference):
 *   init   src/serial/SerialStream.cpp:36-46    src/omp/OMPStream.cpp:61-84
 *   copy   src/serial/SerialStream.cpp:56-63    src/omp/OMPStream.cpp:102-122
 *   mul    src/serial/SerialStream.cpp:65-73    src/omp/OMPStream.cpp:124-148
 *   add    src/serial/SerialStream.cpp:75-82    src/omp/OMPStream.cpp:150-173
 *   triad  src/serial/SerialStream.cpp:84-92    src/omp/OMPStream.cpp:175-200
 *   nstream src/serial/SerialStream.cpp:94-102  src/omp/OMPStream.cpp:202-227
 *   dot    src/serial/SerialStream.cpp:104-113  src/omp/OMPStream.cpp:229-250
 *   check_solution  src/main.cpp:331-421        bench[] order/weights src/benchmark.h:31-39
 *
 * Floating-point contraction.  The reference is built with g++ -O3 -march=native, whose
 * default -ffp-contract=fast fuses  b + s*c  into fma(s,c,b)  in triad and nstream
 * (dot is subtler, see oracle_dot below).  This file is compiled with
 * -ffp-contract=off and writes the fused form EXPLICITLY with fma()/fmaf() when `contract` != 0, so that the
 * oracle's rounding does not depend on compiler flags:
 *   contract = 1  ->  the reference's default build      (one rounding in  b + s*c)
 *   contract = 0  ->  the reference built -ffp-contract=off (two roundings)
 *
 * The `scalar` argument is already of type T:  the reference does
 * `const T scalar = startScalar;` (0.4 rounded to T once), SerialStream.cpp:68,87,97.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>
#include <string.h>

#ifdef _OPENMP
#include <omp.h>
#define ORACLE_PARALLEL_FOR _Pragma("omp parallel for")
#else
#define ORACLE_PARALLEL_FOR
#endif

#if defined(__GNUC__)
#pragma GCC optimize("fp-contract=off") /* belt and braces next to the -ffp-contract=off flag */
#endif

typedef long long idx_t; /* the reference indexes with intptr_t */

#define DEFINE_KERNELS(SUF, T, FMA)                                                              \
  void oracle_init_##SUF(T* a, T* b, T* c, idx_t n, T ia, T ib, T ic) {                          \
    ORACLE_PARALLEL_FOR                                                                          \
    for (idx_t i = 0; i < n; i++) { a[i] = ia; b[i] = ib; c[i] = ic; }                           \
  }                                                                                              \
  void oracle_copy_##SUF(const T* a, T* c, idx_t n) {                                            \
    ORACLE_PARALLEL_FOR                                                                          \
    for (idx_t i = 0; i < n; i++) c[i] = a[i];                                                   \
  }                                                                                              \
  void oracle_mul_##SUF(T* b, const T* c, idx_t n, T scalar) {                                   \
    ORACLE_PARALLEL_FOR                                                                          \
    for (idx_t i = 0; i < n; i++) b[i] = scalar * c[i];                                          \
  }                                                                                              \
  void oracle_add_##SUF(const T* a, const T* b, T* c, idx_t n) {                                 \
    ORACLE_PARALLEL_FOR                                                                          \
    for (idx_t i = 0; i < n; i++) c[i] = a[i] + b[i];                                            \
  }                                                                                              \
  void oracle_triad_##SUF(T* a, const T* b, const T* c, idx_t n, T scalar, int contract) {       \
    if (contract) {                                                                              \
      ORACLE_PARALLEL_FOR                                                                        \
      for (idx_t i = 0; i < n; i++) a[i] = FMA(scalar, c[i], b[i]);                              \
    } else {                                                                                     \
      ORACLE_PARALLEL_FOR                                                                        \
      for (idx_t i = 0; i < n; i++) { T p = scalar * c[i]; a[i] = b[i] + p; }                    \
    }                                                                                            \
  }                                                                                              \
  void oracle_nstream_##SUF(T* a, const T* b, const T* c, idx_t n, T scalar, int contract) {     \
    if (contract) {                                                                              \
      ORACLE_PARALLEL_FOR                                                                        \
      for (idx_t i = 0; i < n; i++) { T t = FMA(scalar, c[i], b[i]); a[i] = a[i] + t; }          \
    } else {                                                                                     \
      ORACLE_PARALLEL_FOR                                                                        \
      for (idx_t i = 0; i < n; i++) { T p = scalar * c[i]; T t = b[i] + p; a[i] = a[i] + t; }    \
    }                                                                                            \
  }                                                                                              \
  /* Serial dot, SerialStream.cpp:104-113: one in-order chain  sum += a[i]*b[i].                 \
     With contraction allowed, what GCC 13.3 -O3 makes of this loop (x86-64-v3 here; verified on    \
     oracle/_ref by disassembly and by tests/test_oracle.py) is NOT a chain of FMAs: the            \
     vectoriser multiplies 4 doubles / 8 floats at a time with vmulp[ds] (rounded), folds the        \
     products into `sum` in order with vadds[ds], repeats that once with a 128-bit epilogue, and     \
     only the scalar remainder  n mod (16/sizeof(T))  is contracted to vfmadd231s[ds].  That is      \
     the arithmetic restated here; contract=0 is unfused throughout. */                              \
  T oracle_dot_##SUF(const T* a, const T* b, idx_t n, int contract) {                            \
    T sum = (T)0;                                                                                \
    idx_t head = contract ? n - n % (idx_t)(16 / sizeof(T)) : n;                                 \
    for (idx_t i = 0; i < head; i++) { T p = a[i] * b[i]; sum = sum + p; }                       \
    for (idx_t i = head; i < n; i++) sum = FMA(a[i], b[i], sum);                                 \
    return sum;                                                                                  \
  }                                                                                              \
  /* OpenMP dot: per-thread sequential chains over static chunks, combined in thread order       \
     (OMPStream.cpp:229-250; the combination order of an OpenMP reduction is unspecified, this   \
     is one legal order).  Used only as the timed CPU baseline and in tolerance checks. */       \
  T oracle_dot_omp_##SUF(const T* a, const T* b, idx_t n, int contract) {                        \
    T sum = (T)0;                                                                                \
    if (contract) {                                                                              \
      _Pragma("omp parallel for reduction(+:sum)")                                               \
      for (idx_t i = 0; i < n; i++) sum = FMA(a[i], b[i], sum);                                  \
    } else {                                                                                     \
      _Pragma("omp parallel for reduction(+:sum)")                                               \
      for (idx_t i = 0; i < n; i++) { T p = a[i] * b[i]; sum = sum + p; }                        \
    }                                                                                            \
    return sum;                                                                                  \
  }

DEFINE_KERNELS(f64, double, fma)
DEFINE_KERNELS(f32, float, fmaf)

/* ------------------------------------------------------------------------------------------
 * Benchmark table, src/benchmark.h:18,31-39.  Ids are the reference's BenchId values.
 * ------------------------------------------------------------------------------------------ */
enum { B_COPY = 0, B_MUL = 1, B_ADD = 2, B_TRIAD = 3, B_NSTREAM = 4, B_DOT = 5, B_CLASSIC = 6, B_ALL = 7 };
static const int kRunOrder[6] = { B_COPY, B_MUL, B_ADD, B_TRIAD, B_DOT, B_NSTREAM };
static const int kWeights[6]  = { 2, 2, 3, 3, 2, 4 }; /* in kRunOrder order */

int oracle_num_benchmarks(void) { return 6; }
int oracle_bench_id(int pos)     { return (pos >= 0 && pos < 6) ? kRunOrder[pos] : -1; }
int oracle_bench_weight(int pos) { return (pos >= 0 && pos < 6) ? kWeights[pos] : -1; }

/* run_benchmark(selection, b), src/benchmark.h:62-66 */
int oracle_run_benchmark(int selection, int pos) {
  int id = kRunOrder[pos];
  if (selection == B_ALL) return 1;
  if (selection == B_CLASSIC && id != B_NSTREAM) return 1;
  return selection == id;
}

/* ------------------------------------------------------------------------------------------
 * check_solution, src/main.cpp:331-421.
 *   order: 0 = Classic (each selected kernel once, repeated num_times), 1 = Isolated.
 *   gold scalars are advanced in type T with the SAME expressions as main.cpp; main.cpp is
 *   compiled with the reference's flags too, so `contract` selects fused / unfused gold.
 *   Returns the number of failed checks (0 = pass) and writes the gold values.
 * ------------------------------------------------------------------------------------------ */
#define DEFINE_CHECK(SUF, T, FMA, EPS, ABS)                                                      \
  static int check_one_##SUF(T is, T should, T mrel) {                                           \
    T diff = ABS(is - should);                                                                   \
    T largest = ABS(is) > ABS(should) ? ABS(is) : ABS(should);                                   \
    int same = diff <= largest * mrel;                                                           \
    return (!same || isnan(is)) ? 1 : 0;                                                         \
  }                                                                                              \
  void oracle_gold_##SUF(int selection, int order, idx_t num_times, idx_t array_size,            \
                         int contract, T ia, T ib, T ic, T scalar, T* gold /*[4]: a b c sum*/) { \
    T gA = ia, gB = ib, gC = ic, gS = (T)0;                                                      \
    for (idx_t outer = 0; outer < (order == 0 ? num_times : 6); outer++) {                       \
      for (idx_t inner = 0; inner < (order == 0 ? 6 : num_times); inner++) {                     \
        int pos = (int)(order == 0 ? inner : outer);                                             \
        if (!oracle_run_benchmark(selection, pos)) continue;                                     \
        switch (kRunOrder[pos]) {                                                                \
          case B_COPY: gC = gA; break;                                                           \
          case B_MUL:  gB = scalar * gC; break;                                                  \
          case B_ADD:  gC = gA + gB; break;                                                      \
          case B_TRIAD:                                                                          \
            if (contract) gA = FMA(scalar, gC, gB); else { T p = scalar * gC; gA = gB + p; }     \
            break;                                                                               \
          case B_NSTREAM: {                                                                      \
            T t; if (contract) t = FMA(scalar, gC, gB); else { T p = scalar * gC; t = gB + p; }  \
            gA = gA + t; break; }                                                                \
          case B_DOT: { T p = gA * gB; gS = p * (T)array_size; break; }                          \
        }                                                                                        \
      }                                                                                          \
    }                                                                                            \
    gold[0] = gA; gold[1] = gB; gold[2] = gC; gold[3] = gS;                                      \
  }                                                                                              \
  idx_t oracle_check_solution_##SUF(int selection, int order, idx_t num_times, idx_t array_size, \
                                    int contract, const T* a, const T* b, const T* c, T sum,     \
                                    T ia, T ib, T ic, T scalar, T* gold_out /*[4] or NULL*/) {   \
    T g[4];                                                                                      \
    oracle_gold_##SUF(selection, order, num_times, array_size, contract, ia, ib, ic, scalar, g); \
    if (gold_out) memcpy(gold_out, g, sizeof g);                                                 \
    const T max_rel = (T)EPS * (T)100.0;                                                         \
    const T max_rel_dot = (T)EPS * (T)10000000.0;                                                \
    idx_t failed = 0;                                                                            \
    if (oracle_run_benchmark(selection, 4 /* Dot's position */))                                 \
      failed += check_one_##SUF(sum, g[3], max_rel_dot);                                         \
    for (idx_t i = 0; i < array_size; i++) {                                                     \
      failed += check_one_##SUF(a[i], g[0], max_rel);                                            \
      failed += check_one_##SUF(b[i], g[1], max_rel);                                            \
      failed += check_one_##SUF(c[i], g[2], max_rel);                                            \
    }                                                                                            \
    return failed;                                                                               \
  }

DEFINE_CHECK(f64, double, fma, 2.220446049250313e-16, fabs)
DEFINE_CHECK(f32, float, fmaf, 1.1920928955078125e-07f, fabsf)

/* ------------------------------------------------------------------------------------------
 * One full driver run on host arrays (the loop of src/main.cpp:150-180), used by the tests
 * to produce a/b/c/sum for (selection, order, num_times) and by bench.py's CPU baseline.
 * ------------------------------------------------------------------------------------------ */
#define DEFINE_RUN(SUF, T)                                                                       \
  T oracle_run_##SUF(int selection, int order, idx_t num_times, T* a, T* b, T* c, idx_t n,       \
                     T scalar, int contract, int use_omp_dot) {                                  \
    T sum = (T)0;                                                                                \
    for (idx_t outer = 0; outer < (order == 0 ? num_times : 6); outer++) {                       \
      for (idx_t inner = 0; inner < (order == 0 ? 6 : num_times); inner++) {                     \
        int pos = (int)(order == 0 ? inner : outer);                                             \
        if (!oracle_run_benchmark(selection, pos)) continue;                                     \
        switch (kRunOrder[pos]) {                                                                \
          case B_COPY:    oracle_copy_##SUF(a, c, n); break;                                     \
          case B_MUL:     oracle_mul_##SUF(b, c, n, scalar); break;                              \
          case B_ADD:     oracle_add_##SUF(a, b, c, n); break;                                   \
          case B_TRIAD:   oracle_triad_##SUF(a, b, c, n, scalar, contract); break;               \
          case B_NSTREAM: oracle_nstream_##SUF(a, b, c, n, scalar, contract); break;             \
          case B_DOT:     sum = use_omp_dot ? oracle_dot_omp_##SUF(a, b, n, contract)            \
                                            : oracle_dot_##SUF(a, b, n, contract); break;        \
        }                                                                                        \
      }                                                                                          \
    }                                                                                            \
    return sum;                                                                                  \
  }

DEFINE_RUN(f64, double)
DEFINE_RUN(f32, float)

int oracle_num_threads(void) {
#ifdef _OPENMP
  return omp_get_max_threads();
#else
  return 1;
#endif
}
