// route.cuh - the sm_100a routing kernel: a persistent grid of warps, each
// warp owning one resident A* search at a time (route_core.hpp).  Instances
// are handed out through a global atomic counter in a host-chosen order
// (largest first).  Per-warp working set: the layout of the node being
// expanded, the considered-qubit mask, the children of the current chunk and
// the lookahead partials live in shared memory; node pool, open list and the
// layout-keyed hash table live in global memory (L2-resident for the batch
// sizes of BASELINE.json's configs).
#pragma once
#include <cuda_runtime.h>

#include "route_core.hpp"

namespace qmg {

constexpr int kWarpsPerCta = 4;
#ifndef QMG_MIN_CTAS
#define QMG_MIN_CTAS 8
#endif

struct SmemLayout {
  uint32_t bytesPerWarp;
  uint32_t offL, offCurQ, offCurL, offMask, offCh, offLookPen;
};

inline SmemLayout smemLayout(int n) {
  auto al = [](uint32_t v) { return (v + 15U) & ~15U; };
  SmemLayout s{};
  uint32_t o = 0;
  o += al(static_cast<uint32_t>(n) * 2U);  // Q
  s.offL = o;
  o += al(static_cast<uint32_t>(n) * 2U);
  s.offCurQ = o;
  o += al(static_cast<uint32_t>(n) * 2U);
  s.offCurL = o;
  o += al(static_cast<uint32_t>(n) * 2U);
  s.offMask = o;
  o += al((static_cast<uint32_t>(n) + 31U) / 32U * 4U + 4U);
  s.offCh = o;
  o += al(static_cast<uint32_t>(sizeof(Child)) * kChunk);
  s.offLookPen = o;
  o += al(8U * (kChunk / 2) * kLookBlock);
  s.bytesPerWarp = o;
  return s;
}

template <int H>
__global__ void __launch_bounds__(kWarpsPerCta * 32, QMG_MIN_CTAS)
    route_kernel(ArchView A, CfgView C, const InstView* __restrict__ inst,
                 const uint32_t* __restrict__ order, uint32_t nInst,
                 SearchMem* __restrict__ mem, uint32_t* __restrict__ ticket,
                 SmemLayout sl) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = static_cast<int>(threadIdx.x) >> 5;
  const int lane = static_cast<int>(threadIdx.x) & 31;
  const uint32_t gw = blockIdx.x * kWarpsPerCta + warp;
  uint8_t* base = smem + static_cast<size_t>(warp) * sl.bytesPerWarp;
  Scratch S;
  S.Q = reinterpret_cast<int16_t*>(base);
  S.L = reinterpret_cast<int16_t*>(base + sl.offL);
  S.curQ = reinterpret_cast<int16_t*>(base + sl.offCurQ);
  S.curL = reinterpret_cast<int16_t*>(base + sl.offCurL);
  S.consMask = reinterpret_cast<uint32_t*>(base + sl.offMask);
  S.ch = reinterpret_cast<Child*>(base + sl.offCh);
  S.lookPen = reinterpret_cast<double*>(base + sl.offLookPen);
  SearchMem M = mem[gw];
  for (;;) {
    uint32_t t = 0;
    if (lane == 0) {
      t = atomicAdd(ticket, 1U);
    }
    t = __shfl_sync(0xffffffffU, t, 0);
    if (t >= nInst) {
      break;
    }
    const InstView I = inst[order[t]];
    Search<H> s(A, C, I, M, S);
    s.run();
    __syncwarp();
  }
  if (lane == 0) {
    mem[gw].stamp = M.stamp;
  }
}

}  // namespace qmg
