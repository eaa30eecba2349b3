// flat_layers.hpp - Mapper::createLayers (src/sc/Mapper.cpp:138-281) straight
// to the CSR form the routing kernel reads, without per-layer containers.
//
// layers.hpp's LayeredCircuit keeps what the reference keeps per layer (a
// std::map of multiplicities, three std::sets, a vector over all qubits) and
// is what qmg_circuit_split_layer works on; building it costs about 1 us per
// gate.  A routing batch only needs the multiplicity entries in std::map
// order, so for batches of circuits this builder assigns every gate its layer
// (the same rules, :55-136 and :173-233), buckets the gates by layer and
// merges equal qubit pairs inside each bucket.  The result is identical to
// FlatCircuit::fromLayered (tests/test_host.py compares them array by array).
#pragma once
#include <algorithm>
#include <cstdint>
#include <stdexcept>
#include <vector>

#include "route_core.hpp"

namespace qmg {

struct FlatLayers {
  uint32_t nLayers = 0, nLogical = 0;
  std::vector<uint32_t> g2Off, g1Off;
  std::vector<Gate2> g2;
  std::vector<Gate1> g1;
  // the gates of every layer in circuit order (what circuit emission walks)
  std::vector<uint32_t> gateOff, gateIdx;

  // layer index of every gate; returns the number of layers.  Layerings 0, 1
  // and 4 are the ones the heuristic mapper accepts (checkParameters :149-167).
  static uint32_t assignLayers(uint32_t nArchQubits, size_t nGates,
                               const int32_t* ctrl, const int32_t* tgt,
                               int layering, std::vector<uint32_t>& layerOf) {
    layerOf.resize(nGates);
    uint32_t nL = 0;
    constexpr int64_t kUnset = -1;
    std::vector<int64_t> last(nArchQubits, kUnset);
    switch (layering) {
    case 0:  // individual gates
      for (size_t gi = 0; gi < nGates; ++gi) {
        layerOf[gi] = static_cast<uint32_t>(gi);
      }
      nL = static_cast<uint32_t>(nGates);
      break;
    case 1:  // disjoint qubits (processDisjointQubitLayer :55-92)
      for (size_t gi = 0; gi < nGates; ++gi) {
        const int32_t c = ctrl[gi];
        const auto t = static_cast<size_t>(tgt[gi]);
        int64_t idx = 0;
        if (c < 0) {
          idx = last[t] + 1;  // (unset = -1 -> layer 0)
          last[t] = idx;
        } else {
          idx = std::max(last[static_cast<size_t>(c)], last[t]) + 1;
          last[static_cast<size_t>(c)] = idx;
          last[t] = idx;
        }
        layerOf[gi] = static_cast<uint32_t>(idx);
        nL = std::max(nL, static_cast<uint32_t>(idx) + 1);
      }
      break;
    case 4: {  // disjoint 2q blocks (processDisjoint2qBlockLayer :94-136)
      // a 2q gate extends the block of its pair when both qubits sit in the
      // same layer and that layer holds a 2q gate on exactly this pair; 2q
      // gates of one layer are qubit-disjoint up to such blocks, so that is:
      // both qubits' last 2q gate was with each other
      std::vector<int32_t> partner(nArchQubits, -1);
      for (size_t gi = 0; gi < nGates; ++gi) {
        const int32_t c = ctrl[gi];
        const auto t = static_cast<size_t>(tgt[gi]);
        int64_t idx = 0;
        if (c < 0) {
          idx = last[t] == kUnset ? 0 : last[t];
        } else {
          const auto cc = static_cast<size_t>(c);
          const int64_t lc = last[cc], lt = last[t];
          if (lc != kUnset && lt != kUnset) {
            idx = std::max(lc, lt) + 1;
            if (lc == lt && partner[cc] == static_cast<int32_t>(t) &&
                partner[t] == c) {
              --idx;
            }
          } else if (lc != kUnset) {
            idx = lc + 1;
          } else if (lt != kUnset) {
            idx = lt + 1;
          }
          last[cc] = idx;
          last[t] = idx;
          partner[cc] = static_cast<int32_t>(t);
          partner[t] = c;
        }
        layerOf[gi] = static_cast<uint32_t>(idx);
        nL = std::max(nL, static_cast<uint32_t>(idx) + 1);
      }
      break;
    }
    default:
      throw std::invalid_argument(
          "Layering strategy not suitable for heuristic mapper!");
    }
    return nL;
  }

  void build(uint32_t nArchQubits, size_t nGates, const int32_t* ctrl,
             const int32_t* tgt, int layering) {
    nLogical = nArchQubits;
    for (size_t i = 0; i < nGates; ++i) {
      if (tgt[i] < 0 || static_cast<uint32_t>(tgt[i]) >= nArchQubits ||
          ctrl[i] >= static_cast<int32_t>(nArchQubits) || ctrl[i] == tgt[i]) {
        throw std::invalid_argument("gate operand out of range");
      }
    }
    // scratch reused by the calling thread from circuit to circuit (fresh
    // allocations cost page faults that serialise the host threads)
    struct Tmp {
      std::vector<uint32_t> layerOf, fill, keys1, off2, off1, gOff, gIdx;
      std::vector<uint64_t> keys2;
      std::vector<Gate2> g2;
      std::vector<Gate1> g1;
    };
    static thread_local Tmp T;
    const uint32_t nL =
        assignLayers(nArchQubits, nGates, ctrl, tgt, layering, T.layerOf);
    nLayers = nL;
    // ---- bucket the gates by layer (stable)
    T.gOff.assign(static_cast<size_t>(nL) + 1, 0);
    for (size_t gi = 0; gi < nGates; ++gi) {
      ++T.gOff[T.layerOf[gi] + 1];
    }
    for (uint32_t l = 0; l < nL; ++l) {
      T.gOff[l + 1] += T.gOff[l];
    }
    T.gIdx.resize(nGates);
    T.fill.assign(T.gOff.begin(), T.gOff.end() - 1);
    for (size_t gi = 0; gi < nGates; ++gi) {
      T.gIdx[T.fill[T.layerOf[gi]]++] = static_cast<uint32_t>(gi);
    }
    // ---- multiplicity entries per layer: 2q pairs ascending (q1 < q2) with
    // the number of gates in each direction (:237-280), 1q gates per qubit
    T.off2.assign(static_cast<size_t>(nL) + 1, 0);
    T.off1.assign(static_cast<size_t>(nL) + 1, 0);
    T.g2.clear();
    T.g1.clear();
    for (uint32_t l = 0; l < nL; ++l) {
      T.keys2.clear();
      T.keys1.clear();
      for (uint32_t k = T.gOff[l]; k < T.gOff[l + 1]; ++k) {
        const uint32_t gi = T.gIdx[k];
        const int32_t c = ctrl[gi];
        const auto t = static_cast<uint32_t>(tgt[gi]);
        if (c < 0) {
          T.keys1.push_back(t);
        } else {
          const auto cc = static_cast<uint32_t>(c);
          // key = (q1, q2, direction): direction 0 = q1 -> q2 (forward)
          if (cc >= t) {
            T.keys2.push_back((static_cast<uint64_t>(t) << 32) | (cc << 1) | 1U);
          } else {
            T.keys2.push_back((static_cast<uint64_t>(cc) << 32) | (t << 1));
          }
        }
      }
      if (T.keys2.size() > 1) {
        std::sort(T.keys2.begin(), T.keys2.end());
      }
      for (size_t k = 0; k < T.keys2.size();) {
        const uint64_t pairKey = T.keys2[k] >> 1;
        Gate2 e{static_cast<uint16_t>(T.keys2[k] >> 32),
                static_cast<uint16_t>((T.keys2[k] & 0xffffffffU) >> 1), 0, 0};
        while (k < T.keys2.size() && (T.keys2[k] >> 1) == pairKey) {
          if ((T.keys2[k] & 1U) != 0) {
            ++e.rev;
          } else {
            ++e.fwd;
          }
          ++k;
        }
        T.g2.push_back(e);
      }
      if (T.keys1.size() > 1) {
        std::sort(T.keys1.begin(), T.keys1.end());
      }
      for (size_t k = 0; k < T.keys1.size();) {
        Gate1 e{static_cast<uint16_t>(T.keys1[k]), 0};
        const uint32_t q = T.keys1[k];
        while (k < T.keys1.size() && T.keys1[k] == q) {
          ++e.mult;
          ++k;
        }
        T.g1.push_back(e);
      }
      T.off2[l + 1] = static_cast<uint32_t>(T.g2.size());
      T.off1[l + 1] = static_cast<uint32_t>(T.g1.size());
    }
    // exact-size copies
    g2Off = T.off2;
    g1Off = T.off1;
    g2 = T.g2;
    g1 = T.g1;
    gateOff = T.gOff;
    gateIdx = T.gIdx;
  }
};

}  // namespace qmg
