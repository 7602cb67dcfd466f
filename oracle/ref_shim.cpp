// ref_shim.cpp -- TEST INFRASTRUCTURE.  A C ABI around the UNMODIFIED reference models.
//
// Compiled by oracle/Makefile together with the reference's own source file, taken where it
// lies under $(BABELSTREAM_ROOT) (default /root/reference):
//     -DSERIAL  + src/serial/SerialStream.cpp   ->  oracle/_ref/libref_serial_{fast,off}.so
//     -DOMP     + src/omp/OMPStream.cpp         ->  oracle/_ref/libref_omp.so
// Nothing from the reference is copied into this repository; this file only *calls* it
// through its public interface (src/Stream.h:29-49, make_stream in src/StreamModels.h:36-99).
//
// The reference models keep a, b, c private and fill them with constants.  To feed seeded
// random inputs through the reference kernels the shim writes through the pointers that
// get_arrays() hands out (serial/omp return the live arrays: SerialStream.cpp:48-54,
// OMPStream.cpp:86-100).
#include <cstring>
#include <memory>

#include "Stream.h"
#include "StreamModels.h"

namespace {
template <class T> struct Handle {
  std::unique_ptr<Stream<T>> s;
  long long n;
};
template <class T> Handle<T>* make(long long n) {
  auto* h = new Handle<T>;
  // same call as src/main.cpp:273-274
  h->s = make_stream<T>(BenchId::All, (intptr_t)n, (size_t)0, startA, startB, startC);
  h->n = n;
  return h;
}
template <class T> void set(Handle<T>* h, const void* a, const void* b, const void* c) {
  T const *pa, *pb, *pc;
  h->s->get_arrays(pa, pb, pc);
  std::memcpy(const_cast<T*>(pa), a, sizeof(T) * h->n);
  std::memcpy(const_cast<T*>(pb), b, sizeof(T) * h->n);
  std::memcpy(const_cast<T*>(pc), c, sizeof(T) * h->n);
}
template <class T> void get(Handle<T>* h, void* a, void* b, void* c) {
  T const *pa, *pb, *pc;
  h->s->get_arrays(pa, pb, pc);
  std::memcpy(a, pa, sizeof(T) * h->n);
  std::memcpy(b, pb, sizeof(T) * h->n);
  std::memcpy(c, pc, sizeof(T) * h->n);
}
}  // namespace

#define DISPATCH(is_float, expr_f, expr_d) do { if (is_float) { expr_f; } else { expr_d; } } while (0)
#define HF ((Handle<float>*)h)
#define HD ((Handle<double>*)h)

extern "C" {
const char* ref_impl_string(void) { return IMPLEMENTATION_STRING; }

void* ref_create(int is_float, long long n) {
  if (is_float) return make<float>(n);
  return make<double>(n);
}
void ref_destroy(void* h, int is_float) { DISPATCH(is_float, delete HF, delete HD); }

void ref_init(void* h, int is_float, double a, double b, double c) {
  DISPATCH(is_float, HF->s->init_arrays((float)a, (float)b, (float)c), HD->s->init_arrays(a, b, c));
}
void ref_set(void* h, int is_float, const void* a, const void* b, const void* c) {
  DISPATCH(is_float, set(HF, a, b, c), set(HD, a, b, c));
}
void ref_get(void* h, int is_float, void* a, void* b, void* c) {
  DISPATCH(is_float, get(HF, a, b, c), get(HD, a, b, c));
}
void ref_copy(void* h, int is_float) { DISPATCH(is_float, HF->s->copy(), HD->s->copy()); }
void ref_mul(void* h, int is_float) { DISPATCH(is_float, HF->s->mul(), HD->s->mul()); }
void ref_add(void* h, int is_float) { DISPATCH(is_float, HF->s->add(), HD->s->add()); }
void ref_triad(void* h, int is_float) { DISPATCH(is_float, HF->s->triad(), HD->s->triad()); }
void ref_nstream(void* h, int is_float) { DISPATCH(is_float, HF->s->nstream(), HD->s->nstream()); }
double ref_dot(void* h, int is_float) {
  if (is_float) return (double)HF->s->dot();
  return HD->s->dot();
}
}
