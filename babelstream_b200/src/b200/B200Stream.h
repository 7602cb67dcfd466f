// B200Stream.h -- the BabelStream model "b200": Stream<T> for NVIDIA B200 (sm_100a).
//
// A new MODEL in the sense of the reference's cmake/register_models.cmake: this header is what
// src/StreamModels.h includes when the preprocessor symbol B200 is defined, and B200Stream<T> is
// what make_stream<T>() constructs.  The class implements the eight virtuals of src/Stream.h:37-48
// and holds no CUDA code: every call goes through the C ABI of include/b200_stream.h into
// libb200stream.so (hand-written sm_100a kernels, 1..8 GPUs).
//
// Variant notice (reference LICENSE 3b): results produced with this model are "based on a
// variant of BabelStream".
#pragma once

#include <iostream>
#include <stdexcept>
#include <string>

#include "Stream.h"
#include "b200_stream.h"

#define IMPLEMENTATION_STRING "B200 (sm_100a CUDA, variant of BabelStream)"

template <class T>
class B200Stream : public Stream<T>
{
  b200_ctx* ctx = nullptr;
  intptr_t array_size;

  // Turns a C-ABI status into the reference's error convention: runtime failure -> message on
  // stderr and exit(code), as the CU() macro does (src/cuda/CUDAStream.cu:14-26).
  static void check(int rc, const char* what);

public:
  // Same parameter list as every in-tree model (e.g. src/cuda/CUDAStream.h:43-44): make_stream
  // forwards (BenchId, intptr_t, size_t, double, double, double) from src/main.cpp:273-274.
  // The GPU count is not a driver flag: BABELSTREAM_B200_NGPUS (default 1) selects devices
  // device_id .. device_id + NGPUS - 1.
  B200Stream(BenchId bs, const intptr_t array_size, const int device_id, T initA, T initB, T initC);
  ~B200Stream();

  void copy() override;
  void add() override;
  void mul() override;
  void triad() override;
  void nstream() override;
  T dot() override;

  void get_arrays(T const*& a, T const*& b, T const*& c) override;
  void init_arrays(T initA, T initB, T initC) override;
};
