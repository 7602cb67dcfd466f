// latency_probe.cu -- where does the time of a tiny blocking call go?  Host-clock minimum and median of
// launch + wait for: a bare one-block kernel, LEVEL 1 (b200_launch_copy + cudaStreamSynchronize) and
// LEVEL 2 (b200_copy), all on 1024 doubles.  Build: make probe.  Run on the GPU box.
#include <algorithm>
#include <chrono>
#include <cstdio>
#include <vector>

#include <cuda_runtime.h>

#include "b200_stream.h"

__global__ void bare_copy(const double* a, double* c, size_t n) {
  const size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x;
  if (i < n) c[i] = a[i];
}

template <class F>
static void timeit(const char* name, int reps, F f) {
  using clk = std::chrono::steady_clock;
  std::vector<double> t(reps);
  for (int i = 0; i < 50; i++) f();
  for (int i = 0; i < reps; i++) {
    const auto t0 = clk::now();
    f();
    t[i] = std::chrono::duration<double, std::micro>(clk::now() - t0).count();
  }
  std::sort(t.begin(), t.end());
  std::printf("%-46s min %7.3f us   median %7.3f us   p90 %7.3f us\n", name, t[0], t[reps / 2], t[reps * 9 / 10]);
}

int main(int argc, char** argv) {
  const size_t n = argc > 1 ? (size_t)atoll(argv[1]) : 1024;
  const int reps = 3000;
  cudaSetDevice(0);
  cudaSetDeviceFlags(cudaDeviceScheduleSpin);
  cudaStream_t st;
  cudaStreamCreateWithFlags(&st, cudaStreamNonBlocking);
  double *a, *c;
  cudaMalloc(&a, n * 8);
  cudaMalloc(&c, n * 8);
  cudaMemset(a, 0, n * 8);
  const unsigned blocks = (unsigned)((n + 255) / 256);
  timeit("bare kernel <<<>>> + cudaStreamSynchronize", reps, [&] {
    bare_copy<<<blocks, 256, 0, st>>>(a, c, n);
    cudaStreamSynchronize(st);
  });
  timeit("bare kernel, launch only (enqueue cost)", reps, [&] { bare_copy<<<blocks, 256, 0, st>>>(a, c, n); });
  cudaStreamSynchronize(st);
  timeit("LEVEL 1 b200_launch_copy + cudaStreamSynchronize", reps, [&] {
    b200_launch_copy(a, c, n, 8, nullptr, st);
    cudaStreamSynchronize(st);
  });
  timeit("LEVEL 1 b200_launch_copy, launch only", reps, [&] { b200_launch_copy(a, c, n, 8, nullptr, st); });
  cudaStreamSynchronize(st);
  timeit("LEVEL 1 b200_launch_triad + cudaStreamSynchronize", reps, [&] {
    b200_launch_triad(a, c, c, n, 8, 0.4, nullptr, st);
    cudaStreamSynchronize(st);
  });
  int status = 0;
  b200_ctx* ctx = b200_create(0, 1, n, 8, &status);
  if (!ctx) { std::printf("b200_create failed: %s\n", b200_last_error()); return 1; }
  b200_init_arrays(ctx, 0.1, 0.2, 0.0);
  timeit("LEVEL 2 b200_copy (blocking)", reps, [&] { b200_copy(ctx); });
  timeit("LEVEL 2 b200_triad (blocking)", reps, [&] { b200_triad(ctx, 0.4); });
  double r;
  timeit("LEVEL 2 b200_dot (blocking, doorbell)", reps, [&] { b200_dot(ctx, &r); });
  b200_set_option(ctx, "sync", "stream");
  timeit("LEVEL 2 b200_dot (blocking, stream sync)", reps, [&] { b200_dot(ctx, &r); });
  b200_destroy(ctx);
  return 0;
}
