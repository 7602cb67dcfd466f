// B200Stream.cpp -- host side of the "b200" model; see B200Stream.h.
//
// Interface implemented: /root/reference/src/Stream.h:29-54.  Behaviour mirrored from the
// reference's CUDA model where the driver can observe it:
//   * constructor prints the device banner on stdout regardless of --csv (CUDAStream.cu:94-132),
//     throws std::runtime_error for a bad device index or too little memory (:87-88,135-139),
//     and initialises the arrays (:148);
//   * kernels block until the result is complete (Stream.h:36);
//   * get_arrays hands out model-owned host pointers valid until destruction (:213-235).
#include "B200Stream.h"

#include <cstdio>
#include <cstdlib>

template <class T>
void B200Stream<T>::check(int rc, const char* what)
{
  if (rc == B200_OK) return;
  std::fprintf(stderr, "B200 Error in %s: %s (%d)\n", what, b200_last_error(), rc);
  std::exit(rc < 0 ? -rc : rc);
}

template <class T>
B200Stream<T>::B200Stream(BenchId bs, const intptr_t array_size, const int device_id, T initA, T initB, T initC)
  : array_size(array_size)
{
  (void)bs;
  int status = B200_OK;
  ctx = b200_create(device_id, /*ngpus: from BABELSTREAM_B200_NGPUS*/ 0, (size_t)array_size, (int)sizeof(T), &status);
  if (!ctx)
  {
    if (status == B200_ENODEVICE) throw std::runtime_error(std::string("Invalid device index: ") + b200_last_error());
    if (status == B200_ENOMEM) throw std::runtime_error(std::string("Device does not have enough memory for all buffers: ") + b200_last_error());
    throw std::runtime_error(b200_last_error());
  }

  // Banner, in the style of the reference model
  char name[256] = "";
  b200_device_name(device_id, name, (int)sizeof name);
  int sms = 0, major = 0, minor = 0, l2 = 0, clock_khz = 0, bus = 0;
  size_t mem = 0;
  b200_device_info(device_id, &sms, &major, &minor, &mem, &l2, &clock_khz, &bus);
  std::cout << "B200 model: " << b200_version() << std::endl;
  std::cout << "CUDA Driver: " << getDeviceDriver(device_id) << std::endl;
  // same fields as the reference's line (CUDAStream.cu:103-109): NVML's current memory clock, bus width, ECC
  const int memclock = b200_device_mem_clock_mhz(device_id);
  std::cout << "CUDA Device " << device_id << ": \"" << name << "\" " << sms << " SMs(" << major << "," << minor << ") "
            << "Memory: " << memclock << " MHz x " << bus << "-bit = "
            << 2.0 * memclock * (bus / 8) / 1000.0 << " GB/s PEAK, ECC is " << (b200_device_ecc(device_id) > 0 ? "ON" : "OFF")
            << ", L2 " << (l2 >> 20) << " MiB" << std::endl;
  std::cout << "Memory Allocation: " << b200_memory_mode(ctx) << std::endl;  // CUDAStream.cu:111-120
  char desc[4096] = "";
  if (b200_describe(ctx, desc, (int)sizeof desc) == B200_OK) std::cout << desc;

  // First-use costs (lazy loading of each kernel's code, launch plans) are paid here, not inside the first timed call:
  // --order Isolated averages over every call of a kernel (src/main.cpp:168-175).  Contents undefined until init_arrays.
  check(b200_warmup(ctx, startScalar), "warmup");
  init_arrays(initA, initB, initC);
}

template <class T>
B200Stream<T>::~B200Stream()
{
  b200_destroy(ctx);
}

template <class T>
void B200Stream<T>::init_arrays(T initA, T initB, T initC)
{
  check(b200_init_arrays(ctx, (double)initA, (double)initB, (double)initC), "init_arrays");
}

template <class T>
void B200Stream<T>::get_arrays(T const*& a, T const*& b, T const*& c)
{
  const void *pa, *pb, *pc;
  check(b200_get_arrays(ctx, &pa, &pb, &pc), "get_arrays");
  a = static_cast<T const*>(pa);
  b = static_cast<T const*>(pb);
  c = static_cast<T const*>(pc);
}

template <class T>
void B200Stream<T>::copy()
{
  check(b200_copy(ctx), "copy");
}

// startScalar is a double literal (src/Stream.h:27); the library rounds it to T once, which is the
// serial/OpenMP models' `const T scalar = startScalar;` -- not the CUDA model's double arithmetic
// in float mode.
template <class T>
void B200Stream<T>::mul()
{
  check(b200_mul(ctx, startScalar), "mul");
}

template <class T>
void B200Stream<T>::add()
{
  check(b200_add(ctx), "add");
}

template <class T>
void B200Stream<T>::triad()
{
  check(b200_triad(ctx, startScalar), "triad");
}

template <class T>
void B200Stream<T>::nstream()
{
  check(b200_nstream(ctx, startScalar), "nstream");
}

template <class T>
T B200Stream<T>::dot()
{
  double sum = 0.0;
  check(b200_dot(ctx, &sum), "dot");
  return (T)sum;  // exact: the library widened a value of type T
}

void listDevices(void)
{
  const int count = b200_device_count();
  if (count <= 0)
  {
    std::cerr << "No devices found." << std::endl;
    return;
  }
  std::cout << std::endl << "Devices:" << std::endl;
  for (int i = 0; i < count; i++) std::cout << i << ": " << getDeviceName(i) << std::endl;
  std::cout << std::endl;
}

std::string getDeviceName(const int device)
{
  char name[256] = "";
  if (b200_device_name(device, name, (int)sizeof name) != B200_OK) return std::string("Device name unavailable");
  return std::string(name);
}

std::string getDeviceDriver(const int device)
{
  const int v = b200_driver_version(device);
  return v < 0 ? std::string("Device driver unavailable") : std::to_string(v);
}

template class B200Stream<float>;
template class B200Stream<double>;
