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
#include "route_launch.hpp"

namespace qmg {

template <int H>
__global__ void __launch_bounds__(kWarpsPerCta * 32,
                                  H == 4 ? QMG_MIN_CTAS_FID : QMG_MIN_CTAS)
    route_kernel(ArchView A, CfgView C, const InstView* __restrict__ inst,
                 const uint32_t* __restrict__ order, uint32_t nInst,
                 SearchMem* __restrict__ mem, uint32_t* __restrict__ ticket,
                 Scratch lay, uint32_t startFull) {
  extern __shared__ __align__(16) uint8_t smem[];
  const int warp = kWarpsPerCta == 1 ? 0 : static_cast<int>(threadIdx.x) >> 5;
  const int lane = static_cast<int>(threadIdx.x) & 31;
  const uint32_t gw = blockIdx.x * kWarpsPerCta + warp;
  Scratch S = lay;
  S.base = smem + static_cast<size_t>(warp) * lay.bytes;
  // the layout arrays are moved as 64-bit words: their entries beyond n must
  // be zero (they are never written afterwards)
  for (uint32_t w = kOffVar / 4 + lane; w < lay.bytes / 4; w += 32) {
    reinterpret_cast<uint32_t*>(S.base)[w] = 0;
  }
  // the warp-uniform descriptors are kept in shared memory, not in registers
  InstView* sI = S.inst();
  SearchMem* sM = S.mem();
  {
    const uint32_t* src = reinterpret_cast<const uint32_t*>(mem + gw);
    uint32_t* dst = reinterpret_cast<uint32_t*>(sM);
    for (uint32_t w = lane; w < sizeof(SearchMem) / 4; w += 32) {
      dst[w] = src[w];
    }
  }
  __syncwarp();
  for (;;) {
    uint32_t t = 0;
    if (lane == 0) {
      t = atomicAdd(ticket, 1U);
    }
    t = __shfl_sync(0xffffffffU, t, 0);
    if (t >= nInst) {
      break;
    }
    {
      const uint32_t* src = reinterpret_cast<const uint32_t*>(inst + order[t]);
      uint32_t* dst = reinterpret_cast<uint32_t*>(sI);
      for (uint32_t w = lane; w < sizeof(InstView) / 4; w += 32) {
        dst[w] = __ldg(src + w);
      }
    }
    __syncwarp();
    // the lean variant first; an instance with a long open list or a layer
    // to split is routed again, from its first layer, by the full variant
    // (startFull: the host expects long searches, e.g. fidelity-aware routing
    // on a large architecture)
    bool full = startFull != 0;
    if (!full) {
      Search<H, false> s(A, C, *sI, *sM, S);
      full = !s.run();
      __syncwarp();
    }
    if (full) {
      Search<H, true> s(A, C, *sI, *sM, S);
      s.run();
    }
    __syncwarp();
  }
  if (lane == 0) {
    mem[gw].stamp = sM->stamp;
  }
}

}  // namespace qmg
