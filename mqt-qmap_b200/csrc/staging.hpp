// staging.hpp - pinned host arenas and grow-only device arenas of the routing
// batches (CUDA runtime API); the CUDA-free part is flatten.hpp.
#pragma once
#include <cuda_runtime.h>

#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstring>
#include <thread>
#include <vector>

#include "flatten.hpp"

namespace qmg {

struct PinnedBuf {
  uint8_t* p = nullptr;
  size_t cap = 0;
  PinnedBuf() = default;
  PinnedBuf(const PinnedBuf&) = delete;
  PinnedBuf& operator=(const PinnedBuf&) = delete;
  ~PinnedBuf() { release(); }
  void release() {
    if (p != nullptr) {
      cudaFreeHost(p);
      p = nullptr;
    }
    cap = 0;
  }
  // grow-only; contents are not preserved
  bool ensure(size_t n) {
    if (n <= cap) {
      return true;
    }
    release();
    const size_t want = n + n / 4 + 4096;
    if (cudaHostAlloc(reinterpret_cast<void**>(&p), want,
                      cudaHostAllocDefault) != cudaSuccess) {
      cudaGetLastError();
      p = nullptr;
      return false;
    }
    cap = want;
    return true;
  }
};

struct DevArena {
  uint8_t* p = nullptr;
  size_t cap = 0;
  DevArena() = default;
  DevArena(const DevArena&) = delete;
  DevArena& operator=(const DevArena&) = delete;
  ~DevArena() { release(); }
  void release() {
    if (p != nullptr) {
      cudaFree(p);
      p = nullptr;
    }
    cap = 0;
  }
  cudaError_t ensure(size_t n) {
    if (n <= cap) {
      return cudaSuccess;
    }
    release();
    const size_t want = n + n / 8 + 4096;
    const cudaError_t e = cudaMalloc(reinterpret_cast<void**>(&p), want);
    if (e != cudaSuccess) {
      p = nullptr;
      return e;
    }
    cap = want;
    return cudaSuccess;
  }
};

// device + pinned buffers kept between batches of one Architecture so that
// repeated qmg_route_batch calls do not pay cudaMalloc / cudaHostAlloc
struct Workspace {
  DevArena in, out, work;
  PinnedBuf hin, hout;
  size_t memBudget = 0;  // device bytes the search pools may take
  size_t poolCapNodes = size_t{1} << 18;  // automatic pool size, at most
};

// the few cudaDeviceProp fields the launch geometry needs (the query is slow,
// so it is made once per Architecture)
struct DeviceLimits {
  int multiProcessorCount = 0;
  size_t sharedMemPerBlockOptin = 0;
  size_t sharedMemPerMultiprocessor = 0;
};

}  // namespace qmg
