// layers.hpp - host-side layering of a circuit and its flattening to the CSR
// form the routing kernel reads.  Plain C++17 (no CUDA): shared by the product
// library (api.cu) and the host debug build of the search (tests/emu).
//
// Reference semantics: Mapper::createLayers (src/sc/Mapper.cpp:138-281) with
// processDisjointQubitLayer (:55-92) / processDisjoint2qBlockLayer (:94-136),
// Mapper::isLayerSplittable (:283-300), Mapper::splitLayer (:302-435),
// Mapper::getNextLayer (:437-448).
#pragma once
#include <algorithm>
#include <cstdint>
#include <map>
#include <optional>
#include <set>
#include <stdexcept>
#include <utility>
#include <vector>

#include "route_core.hpp"

namespace qmg {

struct LayerData {
  std::vector<uint32_t> gates;  // indices into the circuit's gate list
  std::map<std::pair<uint16_t, uint16_t>, std::pair<uint16_t, uint16_t>> mult2;
  std::vector<uint16_t> mult1;  // per logical qubit
  std::set<uint16_t> active, active1, active2;
};

struct LayeredCircuit {
  uint32_t nQubits = 0;  // architecture qubits (size of the 1q vectors)
  std::vector<int32_t> control, target;  // control -1 => single-qubit gate
  std::vector<LayerData> layers;

  void finishLayer(LayerData& L) const {
    L.mult2.clear();
    L.mult1.assign(nQubits, 0);
    L.active.clear();
    L.active1.clear();
    L.active2.clear();
    for (const auto gi : L.gates) {
      const auto c = control[gi];
      const auto t = static_cast<uint16_t>(target[gi]);
      if (c < 0) {
        L.active.insert(t);
        L.active1.insert(t);
        ++L.mult1[t];
        continue;
      }
      const auto cc = static_cast<uint16_t>(c);
      L.active.insert(cc);
      L.active.insert(t);
      L.active2.insert(cc);
      L.active2.insert(t);
      if (cc >= t) {
        ++L.mult2[{t, cc}].second;
      } else {
        ++L.mult2[{cc, t}].first;
      }
    }
  }

  // layering: 0 individual gates, 1 disjoint qubits, 2 odd gates,
  // 3 qubit triangle, 4 disjoint 2q blocks (include/sc/configuration/
  // Layering.hpp:17-23)
  void build(uint32_t nArchQubits, size_t nGates, const int32_t* ctrl,
             const int32_t* tgt, int layering) {
    nQubits = nArchQubits;
    control.assign(ctrl, ctrl + nGates);
    target.assign(tgt, tgt + nGates);
    for (size_t i = 0; i < nGates; ++i) {
      if (tgt[i] < 0 || static_cast<uint32_t>(tgt[i]) >= nQubits ||
          ctrl[i] >= static_cast<int32_t>(nQubits) || ctrl[i] == tgt[i]) {
        throw std::invalid_argument("gate operand out of range");
      }
    }
    std::vector<std::vector<uint32_t>> ls;
    std::vector<std::optional<size_t>> last(nQubits);
    std::set<int32_t> triangle;
    bool even = true;
    auto slot = [&](size_t idx) -> std::vector<uint32_t>& {
      if (ls.size() <= idx) {
        ls.emplace_back();
      }
      return ls.at(idx);
    };
    for (uint32_t gi = 0; gi < nGates; ++gi) {
      const auto c = control[gi];
      const auto t = static_cast<size_t>(target[gi]);
      switch (layering) {
      case 0:
        ls.push_back({gi});
        break;
      case 1: {
        size_t idx = 0;
        if (c < 0) {
          idx = last[t] ? *last[t] + 1 : 0;
          last[t] = idx;
        } else {
          const auto& lc = last[static_cast<size_t>(c)];
          const auto& lt = last[t];
          if (lc && lt) {
            idx = std::max(*lc, *lt) + 1;
          } else if (lc) {
            idx = *lc + 1;
          } else if (lt) {
            idx = *lt + 1;
          }
          last[static_cast<size_t>(c)] = idx;
          last[t] = idx;
        }
        slot(idx).push_back(gi);
        break;
      }
      case 4: {
        size_t idx = 0;
        if (c < 0) {
          // 1q gates ride along with the last 2q block of their qubit and do
          // not move the qubit's layer pointer
          idx = last[t] ? *last[t] : 0;
        } else {
          const auto& lc = last[static_cast<size_t>(c)];
          const auto& lt = last[t];
          if (lc && lt) {
            idx = std::max(*lc, *lt) + 1;
            if (*lc == *lt) {
              for (const auto gj : ls.at(idx - 1)) {
                if ((control[gj] == c && target[gj] == target[gi]) ||
                    (control[gj] == target[gi] && target[gj] == c)) {
                  --idx;  // same qubit pair: extend the block
                  break;
                }
              }
            }
          } else if (lc) {
            idx = *lc + 1;
          } else if (lt) {
            idx = *lt + 1;
          }
          last[static_cast<size_t>(c)] = idx;
          last[t] = idx;
        }
        slot(idx).push_back(gi);
        break;
      }
      case 2:
        if (even) {
          ls.emplace_back();
        }
        ls.back().push_back(gi);
        even = !even;
        break;
      case 3:
        if (ls.empty()) {
          ls.emplace_back();
        }
        if (c < 0) {
          ls.back().push_back(gi);
        } else {
          triangle.insert(c);
          triangle.insert(target[gi]);
          if (triangle.size() > 3) {
            ls.emplace_back();
            triangle = {c, target[gi]};
          }
          ls.back().push_back(gi);
        }
        break;
      default:
        throw std::invalid_argument("unknown layering");
      }
    }
    layers.clear();
    layers.resize(ls.size());
    for (size_t i = 0; i < ls.size(); ++i) {
      layers[i].gates = std::move(ls[i]);
      finishLayer(layers[i]);
    }
  }

  [[nodiscard]] bool splittable(size_t i) const {
    const auto& L = layers.at(i);
    if (L.mult2.size() > 1 || L.active1.size() > 2) {
      return true;
    }
    if (L.mult2.empty()) {
      return false;
    }
    return std::any_of(L.active1.begin(), L.active1.end(),
                       [&](uint16_t q) { return L.active2.count(q) == 0; });
  }

  // Mapper::splitLayer: 2q pairs alternate (second, fourth, ... stay; first,
  // third, ... move to the new layer index+1), 1q gates follow the 2q gate of
  // their qubit, the remaining 1q qubits alternate starting with "stay".
  void split(size_t index) {
    const LayerData src = layers.at(index);
    std::set<uint16_t> two0, two1;
    bool toFirst = false;
    for (const auto& kv : src.mult2) {
      auto& dst = toFirst ? two0 : two1;
      dst.insert(kv.first.first);
      dst.insert(kv.first.second);
      toFirst = !toFirst;
    }
    std::vector<uint8_t> one0(nQubits, 0);
    toFirst = true;
    for (uint32_t q = 0; q < nQubits; ++q) {
      if (src.mult1[q] == 0) {
        continue;
      }
      if (two0.count(static_cast<uint16_t>(q)) != 0) {
        one0[q] = 1;
      } else if (two1.count(static_cast<uint16_t>(q)) != 0) {
        one0[q] = 0;
      } else {
        one0[q] = toFirst ? 1 : 0;
        toFirst = !toFirst;
      }
    }
    LayerData a, b;
    for (const auto gi : src.gates) {
      const bool first = control[gi] < 0
                             ? one0[static_cast<size_t>(target[gi])] != 0
                             : two0.count(static_cast<uint16_t>(target[gi])) != 0;
      (first ? a : b).gates.push_back(gi);
    }
    finishLayer(a);
    finishLayer(b);
    layers[index] = std::move(a);
    layers.insert(layers.begin() + static_cast<std::ptrdiff_t>(index) + 1,
                  std::move(b));
  }
};

// CSR form of a layered circuit (what CircView points into)
struct FlatCircuit {
  uint32_t nLayers = 0, nLogical = 0;
  std::vector<uint32_t> g2Off, g1Off, consOff, twoQ, firstAfter;
  std::vector<Gate2> g2;
  std::vector<Gate1> g1;
  std::vector<uint16_t> cons;
  std::vector<uint8_t> splittable;
  uint32_t maxG = 1;

  // from caller-provided CSR layers (qmg_layers); considered qubits are the
  // qubits of the 2q gates, plus those of the 1q gates when fidelityAware
  // (HeuristicMapper::getConsideredQubits, HeuristicMapper.hpp:228-233)
  void build(uint32_t nL, uint32_t nLog, const uint32_t* off2, const Gate2* gg2,
             const uint32_t* off1, const Gate1* gg1, bool fidelityAware) {
    nLayers = nL;
    nLogical = nLog;
    g2Off.assign(off2, off2 + nL + 1);
    g2.assign(gg2, gg2 + off2[nL]);
    if (off1 != nullptr) {
      g1Off.assign(off1, off1 + nL + 1);
      g1.assign(gg1, gg1 + off1[nL]);
    } else {
      g1Off.assign(nL + 1, 0);
      g1.clear();
    }
    consOff.assign(1, 0);
    cons.clear();
    splittable.assign(nL, 0);
    twoQ.clear();
    firstAfter.assign(nL, 0);
    for (uint32_t l = 0; l < nL; ++l) {
      if (g2Off[l + 1] > g2Off[l]) {
        twoQ.push_back(l);
      }
      firstAfter[l] = static_cast<uint32_t>(twoQ.size());
    }
    for (uint32_t l = 0; l < nL; ++l) {
      std::set<uint16_t> c, a1, a2;
      for (uint32_t i = g2Off[l]; i < g2Off[l + 1]; ++i) {
        a2.insert(g2[i].q1);
        a2.insert(g2[i].q2);
      }
      for (uint32_t i = g1Off[l]; i < g1Off[l + 1]; ++i) {
        a1.insert(g1[i].q);
      }
      c = a2;
      if (fidelityAware) {
        c.insert(a1.begin(), a1.end());
      }
      cons.insert(cons.end(), c.begin(), c.end());
      consOff.push_back(static_cast<uint32_t>(cons.size()));
      const uint32_t G = g2Off[l + 1] - g2Off[l];
      maxG = std::max(maxG, G);
      bool sp = G > 1 || a1.size() > 2;
      if (!sp && G == 1) {
        sp = std::any_of(a1.begin(), a1.end(),
                         [&](uint16_t q) { return a2.count(q) == 0; });
      }
      splittable[l] = sp ? 1 : 0;
    }
  }

  void fromLayered(const LayeredCircuit& lc, bool fidelityAware) {
    std::vector<uint32_t> off2{0}, off1{0};
    std::vector<Gate2> gg2;
    std::vector<Gate1> gg1;
    for (const auto& L : lc.layers) {
      for (const auto& [e, m] : L.mult2) {
        gg2.push_back({e.first, e.second, m.first, m.second});
      }
      for (uint32_t q = 0; q < lc.nQubits; ++q) {
        if (L.mult1[q] != 0) {
          gg1.push_back({static_cast<uint16_t>(q), L.mult1[q]});
        }
      }
      off2.push_back(static_cast<uint32_t>(gg2.size()));
      off1.push_back(static_cast<uint32_t>(gg1.size()));
    }
    build(static_cast<uint32_t>(lc.layers.size()), lc.nQubits, off2.data(),
          gg2.data(), off1.data(), gg1.data(), fidelityAware);
  }

  [[nodiscard]] CircView view() const {
    CircView v{};
    v.nLayers = nLayers;
    v.nLogical = nLogical;
    v.g2Off = g2Off.data();
    v.g2 = g2.data();
    v.g1Off = g1Off.data();
    v.g1 = g1.data();
    v.consOff = consOff.data();
    v.cons = cons.data();
    v.twoQ = twoQ.data();
    v.firstAfter = firstAfter.data();
    v.nTwoQ = static_cast<uint32_t>(twoQ.size());
    v.splittable = splittable.data();
    return v;
  }
};

// topology part of the device Architecture: incidence lists + adjacency bits
struct ArchTopology {
  int32_t n = 0, nE = 0;
  bool bidirectional = true, unidirectional = true;
  std::vector<uint16_t> edges;
  std::vector<uint32_t> incOff, inc;
  std::vector<uint8_t> adj;

  void build(uint16_t nq, const uint16_t* e, size_t ne) {
    n = nq;
    std::set<std::pair<uint16_t, uint16_t>> cm;
    for (size_t i = 0; i < ne; ++i) {
      if (e[2 * i] >= nq || e[2 * i + 1] >= nq) {
        throw std::invalid_argument("coupling-map entry out of range");
      }
      cm.emplace(e[2 * i], e[2 * i + 1]);
    }
    nE = static_cast<int32_t>(cm.size());
    edges.clear();
    adj.assign(static_cast<size_t>(n) * n, 0);
    std::vector<std::vector<uint32_t>> per(static_cast<size_t>(n));
    bidirectional = true;
    unidirectional = true;
    uint32_t id = 0;
    for (const auto& [a, b] : cm) {
      edges.push_back(a);
      edges.push_back(b);
      adj[static_cast<size_t>(a) * n + b] |= 1U;
      adj[static_cast<size_t>(b) * n + a] |= 2U;
      per[a].push_back(id);
      if (b != a) {
        per[b].push_back(id);
      }
      // Architecture::createDistanceTable :205-217
      if (cm.count({b, a}) == 0) {
        bidirectional = false;
      } else {
        unidirectional = false;
      }
      ++id;
    }
    incOff.assign(1, 0);
    inc.clear();
    for (auto& v : per) {
      std::sort(v.begin(), v.end());
      inc.insert(inc.end(), v.begin(), v.end());
      incOff.push_back(static_cast<uint32_t>(inc.size()));
    }
  }
};

}  // namespace qmg
