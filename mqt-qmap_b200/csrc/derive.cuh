// derive.cuh - per-circuit tables the routing kernel reads that depend only on
// the layers: considered qubits, the chain of layers holding 2q gates
// (Mapper::getNextLayer in closed form) and the splittable flags.  They are
// derived on the device from the CSR layers the host uploads, so they never
// cross PCIe (half of the staging bytes of a typical circuit).
#pragma once
#include <cuda_runtime.h>

#include "route_core.hpp"

namespace qmg {

struct DeriveView {
  uint32_t nLayers;
  uint32_t fidelityAware;
  const uint32_t* g2Off;
  const Gate2* g2;
  const uint32_t* g1Off;
  const Gate1* g1;
  // outputs
  uint16_t* cons;     // 2*g2Off[l] + g1Off[l] is the base of layer l
  uint16_t* consCnt;  // nLayers
  uint32_t* twoQ;     // nLayers (nTwoQ used)
  uint32_t* firstAfter;
  uint8_t* split;
};

// one CTA per circuit
__global__ void __launch_bounds__(128)
    derive_circuit_tables(const DeriveView* __restrict__ views, uint32_t nCirc) {
  const uint32_t ci = blockIdx.x;
  if (ci >= nCirc) {
    return;
  }
  const DeriveView V = views[ci];
  const uint32_t nL = V.nLayers;
  // ---- considered qubits (HeuristicMapper.hpp:228-233) and
  // Mapper::isLayerSplittable (src/sc/Mapper.cpp:283-300), one layer per thread
  for (uint32_t l = threadIdx.x; l < nL; l += blockDim.x) {
    const uint32_t a = V.g2Off[l], b = V.g2Off[l + 1];
    const uint32_t a1 = V.g1Off[l], b1 = V.g1Off[l + 1];
    uint16_t* out = V.cons + (2U * a + a1);
    uint32_t cnt = 0;
    // sorted insertion without duplicates (the lists are a handful long)
    auto put = [&](uint16_t q) {
      uint32_t k = 0;
      while (k < cnt && out[k] < q) {
        ++k;
      }
      if (k < cnt && out[k] == q) {
        return;
      }
      for (uint32_t m = cnt; m > k; --m) {
        out[m] = out[m - 1];
      }
      out[k] = q;
      ++cnt;
    };
    for (uint32_t i = a; i < b; ++i) {
      const Gate2 e = V.g2[i];
      put(e.q1);
      put(e.q2);
    }
    const uint32_t G = b - a, n1 = b1 - a1;
    bool split = G > 1 || n1 > 2;
    if (!split && G == 1) {
      for (uint32_t i = a1; i < b1; ++i) {
        const uint16_t q = V.g1[i].q;
        bool in2 = false;
        for (uint32_t k = 0; k < cnt; ++k) {
          in2 = in2 || out[k] == q;
        }
        split = split || !in2;
      }
    }
    if (V.fidelityAware != 0) {
      for (uint32_t i = a1; i < b1; ++i) {
        put(V.g1[i].q);
      }
    }
    V.consCnt[l] = static_cast<uint16_t>(cnt);
    V.split[l] = split ? 1 : 0;
  }
  // ---- layers holding a 2q gate, ascending, and for every layer the number of
  // such layers up to and including it (Mapper.cpp:437-448), warp 0 scans
  if (threadIdx.x < 32) {
    const uint32_t lane = threadIdx.x;
    uint32_t base = 0;
    for (uint32_t l0 = 0; l0 < nL; l0 += 32) {
      const uint32_t l = l0 + lane;
      const bool has = l < nL && V.g2Off[l + 1] > V.g2Off[l];
      const uint32_t m = __ballot_sync(0xffffffffU, has);
      const uint32_t before = __popc(m & ((1U << lane) - 1U));
      if (l < nL) {
        V.firstAfter[l] = base + before + (has ? 1U : 0U);
        if (has) {
          V.twoQ[base + before] = l;
        }
      }
      base += __popc(m);
    }
  }
}

}  // namespace qmg
