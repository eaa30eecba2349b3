// route_launch.hpp - host-side launch interface of route_kernel<H>.  Every
// heuristic's instantiation is its own translation unit (route_h.cu compiled
// with -DQMG_H=0..4) so that the five are compiled in parallel; api.cu only
// sees these declarations.
#pragma once
#include <cuda_runtime.h>

#include "route_core.hpp"

namespace qmg {

// searches (warps) per CTA.  Measured on config 4 / config 2 (M nodes/s): 1 warp
// per CTA, which makes the shared-memory base a constant, 41.0 / 154 against
// 42.6 / 155 with 4; 12 warps per CTA (one CTA per SM for the fidelity kernel)
// 47.4 / 152 against 45.9 / 153 - kept at 4
#ifndef QMG_WARPS_PER_CTA
#define QMG_WARPS_PER_CTA 4
#endif
constexpr int kWarpsPerCta = QMG_WARPS_PER_CTA;
#ifndef QMG_MIN_CTAS
// 6 CTAs (24 warps, 80 registers/thread) per SM measured best on config 2:
// 3/4/5/6/8 -> 102/153/166/172/166 M nodes/s (round 1)
#define QMG_MIN_CTAS 6
#endif
// upper bound of the per-warp working set (4 warps of a CTA must fit the
// 227 KB of shared memory a CTA can opt into)
constexpr uint32_t kMaxScratchBytes = 56000;
// (the fidelity-aware instantiation runs at most 12 searches per SM on long
// searches - the node pools take the HBM - so it may use more registers)
#ifndef QMG_MIN_CTAS_FID
#define QMG_MIN_CTAS_FID 3
#endif

struct RouteArgs {
  ArchView A;
  CfgView C;
  const InstView* inst;
  const uint32_t* order;
  uint32_t nInst;
  SearchMem* mem;
  uint32_t* ticket;
  Scratch lay;
  uint32_t startFull;
};

// one per translation unit route_h.cu
#define QMG_DECL_ROUTE(H)                                                     \
  cudaError_t routeSetSharedMemoryH##H(int bytes);                            \
  cudaError_t routeLaunchH##H(unsigned grid, size_t smemBytes, cudaStream_t s, \
                              const RouteArgs& a);
QMG_DECL_ROUTE(0)
QMG_DECL_ROUTE(1)
QMG_DECL_ROUTE(2)
QMG_DECL_ROUTE(3)
QMG_DECL_ROUTE(4)
#undef QMG_DECL_ROUTE

}  // namespace qmg
