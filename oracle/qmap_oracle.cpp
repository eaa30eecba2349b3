// qmap_oracle.cpp - CPU restatement of the reference's heuristic-mapper hot
// path (cost tables + per-layer A* SWAP search).
//
// TEST INFRASTRUCTURE.  Only tests/, __graft_entry__.smoke() and the
// cpu_baseline / --impl reference legs of bench.py may load this library; the
// product (mqt-qmap_b200/) neither links nor imports it.
//
// It is written independently of the CUDA search core: plain C++17 containers
// (std::priority_queue + std::set, as the reference's UniquePriorityQueue does)
// and one function per reference function, each citing the file:line it
// follows (paths relative to the reference tree, cda-tum/mqt-qmap).
//
// PARITY PINNING: this restatement is pinned in tests/test_oracle.py against
//  (1) the golden vectors of the reference's own tests (test/sc/test_general.cpp
//      Dijkstra / edge-skip matrices, test/sc/heuristic/test_heuristic.cpp node
//      counts) restated in tests/golden/reference_tests.json, and
//  (2) per-layer outputs of the reference itself (oracle/_ref, the reference's
//      sources compiled here by oracle/Makefile) committed under
//      tests/golden/route_golden.npz by tests/golden/make_golden.py.
//
// The data model is the C ABI's (include/qmap_b200.h): layers arrive as CSR
// multiplicity entries in std::map<Edge,...> order, results leave as
// qmg_layer_result records, so the same comparison code checks the oracle, the
// compiled reference and the CUDA library.
#include "../include/qmap_b200.h"

#include <algorithm>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <map>
#include <optional>
#include <queue>
#include <set>
#include <string>
#include <utility>
#include <vector>

namespace {

using Matrix = std::vector<std::vector<double>>;
using Edge = std::pair<uint16_t, uint16_t>;
using EdgeSet = std::set<Edge>;  // CouplingMap, include/sc/utils.hpp:38

constexpr double kInf = std::numeric_limits<double>::max();
// include/sc/utils.hpp:28-36
constexpr double kGateCostSwapUni = 34.;  // COST_UNIDIRECTIONAL_SWAP
constexpr double kGateCostSwapBi = 30.;   // COST_BIDIRECTIONAL_SWAP
constexpr double kGateCostReverse = 4.;   // COST_DIRECTION_REVERSE

// --------------------------------------------------------------------------
// Dijkstra::buildTable + Dijkstra::dijkstra, src/sc/utils.cpp:26-89.
// The queue holds pointers to per-vertex records that are overwritten in
// place when a cheaper path is found; a vertex is final once popped.
// --------------------------------------------------------------------------
struct Vertex {
  bool done = false;
  uint16_t at = 0;
  double cost = -1.;
};
struct VertexLater {
  bool operator()(const Vertex* a, const Vertex* b) const {
    return a->cost > b->cost;  // utils.hpp:130-132
  }
};

void shortestPathsFrom(const EdgeSet& cm, std::vector<Vertex>& v, uint16_t src,
                       const Matrix& w) {
  std::priority_queue<Vertex*, std::vector<Vertex*>, VertexLater> pq;
  pq.push(&v[src]);
  while (!pq.empty()) {
    Vertex* cur = pq.top();
    cur->done = true;
    pq.pop();
    const uint16_t here = cur->at;
    for (const auto& e : cm) {  // both directions of every stored edge
      int to = -1;
      if (here == e.first) {
        to = e.second;
      } else if (here == e.second) {
        to = e.first;
      }
      if (to < 0 || v[static_cast<size_t>(to)].done) {
        continue;
      }
      const double c = cur->cost + w[here][static_cast<size_t>(to)];
      Vertex& t = v[static_cast<size_t>(to)];
      if (t.cost < 0 || c < t.cost) {
        t.done = false;
        t.at = static_cast<uint16_t>(to);
        t.cost = c;
        pq.push(&t);
      }
    }
  }
}

Matrix allPairs(const EdgeSet& cm, const Matrix& w) {
  const size_t n = w.size();
  Matrix out(n, std::vector<double>(n, -1.));
  for (size_t i = 0; i < n; ++i) {
    std::vector<Vertex> v(n);
    for (size_t j = 0; j < n; ++j) {
      v[j].at = static_cast<uint16_t>(j);
    }
    v[i].cost = 0.;
    shortestPathsFrom(cm, v, static_cast<uint16_t>(i), w);
    for (size_t j = 0; j < n; ++j) {
      out[i][j] = (i == j) ? 0. : v[j].cost;
    }
  }
  return out;
}

// Dijkstra::buildSingleEdgeSkipTable, utils.cpp:147-183
Matrix skipOneEdge(const Matrix& d, const EdgeSet& cm, double reversal) {
  const size_t n = d.size();
  Matrix t(n, std::vector<double>(n, kInf));
  for (size_t q = 0; q < n; ++q) {
    t[q][q] = 0.;
  }
  for (const auto& [a, b] : cm) {
    for (size_t s = 0; s < n; ++s) {
      for (size_t e = s + 1; e < n; ++e) {
        t[s][e] = std::min(t[s][e], d[s][a] + d[b][e]);
        t[s][e] = std::min(t[s][e], d[s][b] + d[a][e] + reversal);
        if (reversal == 0.) {
          t[e][s] = t[s][e];
        } else {
          t[e][s] = std::min(t[e][s], d[e][a] + d[b][s]);
          t[e][s] = std::min(t[e][s], d[e][b] + d[a][s] + reversal);
        }
      }
    }
  }
  return t;
}

// Dijkstra::buildEdgeSkipTable, utils.cpp:91-145: level k = cheapest cost when
// any k edges of the path are free; stops at the first all-zero level
std::vector<Matrix> skipEdgeLevels(const EdgeSet& cm, const Matrix& w) {
  std::vector<Matrix> lv;
  lv.push_back(allPairs(cm, w));
  const size_t n = w.size();
  for (size_t k = 1; k <= n; ++k) {
    lv.emplace_back(n, std::vector<double>(n, kInf));
    for (size_t q = 0; q < n; ++q) {
      lv[k][q][q] = 0.;
    }
    bool allZero = false;
    for (const auto& [a, b] : cm) {
      for (size_t l = 0; l < k; ++l) {
        allZero = true;  // (re-armed per (edge, l) exactly as the reference)
        for (size_t s = 0; s < n; ++s) {
          for (size_t e = s + 1; e < n; ++e) {
            double& cell = lv[k][s][e];
            cell = std::min(cell, lv[l][s][a] + lv[k - l - 1][b][e]);
            cell = std::min(cell, lv[l][s][b] + lv[k - l - 1][a][e]);
            lv[k][e][s] = cell;
            if (allZero && cell > 0) {
              allZero = false;
            }
          }
        }
      }
    }
    if (allZero) {
      lv.pop_back();
      break;
    }
  }
  return lv;
}

// --------------------------------------------------------------------------
// Architecture (the part the mapper reads), include/sc/Architecture.hpp
// --------------------------------------------------------------------------
struct Arch {
  uint16_t n = 0;
  EdgeSet cm;
  bool bidir = true, unidir = true;
  Matrix dist, distRev;  // distanceTable / distanceTableReversals
  bool fidelity = false;
  std::vector<double> c1;   // singleQubitFidelityCosts
  Matrix c2, cswap;         // twoQubitFidelityCosts / swapFidelityCosts
  std::vector<Matrix> fid;  // fidelityDistanceTables

  // Architecture::createDistanceTable, src/sc/Architecture.cpp:201-230
  void buildDistances() {
    Matrix w(n, std::vector<double>(n, kInf));
    bidir = unidir = true;
    for (const auto& e : cm) {
      if (cm.count({e.second, e.first}) == 0) {
        bidir = false;
        w[e.second][e.first] = kGateCostSwapUni;
        w[e.first][e.second] = kGateCostSwapUni;
      } else {
        unidir = false;
        w[e.first][e.second] = kGateCostSwapBi;
      }
    }
    const Matrix plain = allPairs(cm, w);
    dist = skipOneEdge(plain, cm, 0.);
    distRev = bidir ? dist : skipOneEdge(plain, cm, kGateCostReverse);
  }
  // Architecture.hpp:244-256
  bool connected(int a, int b, bool directed = true) const {
    if (cm.count({static_cast<uint16_t>(a), static_cast<uint16_t>(b)}) != 0) {
      return true;
    }
    return !directed &&
           cm.count({static_cast<uint16_t>(b), static_cast<uint16_t>(a)}) != 0;
  }
  bool edgeBidirectional(int a, int b) const {
    return connected(a, b) && connected(b, a);
  }
  // Architecture.hpp:427-433
  double distance(int a, int b, bool withReversal = true) const {
    return (withReversal ? distRev : dist)[static_cast<size_t>(a)]
                                         [static_cast<size_t>(b)];
  }
  // Architecture.hpp:301-316
  double fidDistance(int a, int b, size_t skip) const {
    if (skip >= fid.size()) {
      return 0.;
    }
    return fid[skip][static_cast<size_t>(a)][static_cast<size_t>(b)];
  }
  // Architecture.cpp:173-186 / :232-294 cost of one SWAP on the edge
  double swapGateCost(int a, int b) const {
    if (bidir) {
      return kGateCostSwapBi;
    }
    if (unidir || !edgeBidirectional(a, b)) {
      return kGateCostSwapUni;
    }
    return kGateCostSwapBi;
  }
};

Arch makeArch(uint16_t n, const uint16_t* edges, size_t nEdges) {
  Arch a;
  a.n = n;
  for (size_t i = 0; i < nEdges; ++i) {
    a.cm.emplace(edges[2 * i], edges[2 * i + 1]);
  }
  a.buildDistances();
  return a;
}

Matrix toMatrix(const double* p, size_t n) {
  Matrix m(n, std::vector<double>(n));
  for (size_t i = 0; i < n; ++i) {
    for (size_t j = 0; j < n; ++j) {
      m[i][j] = p[i * n + j];
    }
  }
  return m;
}
void fromMatrix(const Matrix& m, double* p) {
  const size_t n = m.size();
  for (size_t i = 0; i < n; ++i) {
    for (size_t j = 0; j < n; ++j) {
      p[i * n + j] = m[i][j];
    }
  }
}

// --------------------------------------------------------------------------
// Layers (the products of Mapper::createLayers the search reads,
// include/sc/Mapper.hpp:36-120)
// --------------------------------------------------------------------------
struct Mult {
  uint16_t fwd = 0, rev = 0;
};
struct Layer {
  std::map<Edge, Mult> two;        // twoQubitMultiplicities[layer]
  std::vector<uint16_t> one;       // singleQubitMultiplicities[layer]
  std::set<uint16_t> act, act1, act2;  // activeQubits / 1QGates / 2QGates
};

std::vector<Layer> readLayers(const qmg_layers& L, uint16_t n) {
  std::vector<Layer> out(L.n_layers);
  for (uint32_t l = 0; l < L.n_layers; ++l) {
    Layer& y = out[l];
    y.one.assign(n, 0);
    for (uint32_t i = L.gate2q_offsets[l]; i < L.gate2q_offsets[l + 1]; ++i) {
      const qmg_gate2q& g = L.gates2q[i];
      y.two[{g.q1, g.q2}] = Mult{g.forward, g.reverse};
      y.act.insert(g.q1);
      y.act.insert(g.q2);
      y.act2.insert(g.q1);
      y.act2.insert(g.q2);
    }
    if (L.gate1q_offsets != nullptr) {
      for (uint32_t i = L.gate1q_offsets[l]; i < L.gate1q_offsets[l + 1]; ++i) {
        const qmg_gate1q& g = L.gates1q[i];
        y.one[g.qubit] = g.multiplicity;
        y.act.insert(g.qubit);
        y.act1.insert(g.qubit);
      }
    }
  }
  return out;
}

// --------------------------------------------------------------------------
// HeuristicMapper::Node, include/sc/heuristic/HeuristicMapper.hpp:45-143
// --------------------------------------------------------------------------
struct SNode {
  std::set<Edge> valid;     // validMappedTwoQubitGates
  std::vector<Edge> swaps;  // every entry is a SWAP
  std::vector<int16_t> qubits, locations;
  double fixed = 0., reversals = 0., heur = 0., look = 0.;
  size_t shared = 0, depth = 0, parent = 0, id = 0;
  bool goal = true;  // validMapping
  double total() const { return fixed + reversals + heur + look; }   // :111
  double totalFixed() const { return fixed + reversals + look; }     // :118
};

// identity of a node = its qubits[] (operator<, HeuristicMapper.hpp:438-450)
struct LayoutLess {
  bool operator()(const SNode& x, const SNode& y) const {
    const size_t m = std::min(x.qubits.size(), y.qubits.size());
    for (size_t i = 0; i < m; ++i) {
      if (x.qubits[i] != y.qubits[i]) {
        return x.qubits[i] < y.qubits[i];
      }
    }
    return false;
  }
};
// priority (operator>, HeuristicMapper.hpp:452-483): true = x is served later
struct ServedLater {
  bool operator()(const SNode& x, const SNode& y) const {
    const double fx = x.total(), fy = y.total();
    if (std::abs(fx - fy) > 1e-6) {
      return fx > fy;
    }
    if (x.goal != y.goal) {
      return y.goal;
    }
    const double hx = x.heur + x.look, hy = y.heur + y.look;
    if (std::abs(hx - hy) > 1e-6) {
      return hx > hy;
    }
    if (x.valid.size() != y.valid.size()) {
      return x.valid.size() < y.valid.size();
    }
    return LayoutLess()(x, y);
  }
};

// UniquePriorityQueue, include/sc/heuristic/UniquePriorityQueue.hpp:59-103
class OpenList {
  std::priority_queue<SNode, std::vector<SNode>, ServedLater> heap;
  std::set<SNode, LayoutLess> members;

public:
  void push(const SNode& v) {
    const auto ins = members.insert(v);
    if (ins.second) {
      heap.push(v);
    } else if (ServedLater()(*ins.first, v)) {
      // the queued node of this layout is worse: replace it and rebuild the
      // heap from the membership set (in set order)
      members.erase(ins.first);
      members.insert(v);
      heap = {};
      for (const auto& m : members) {
        heap.push(m);
      }
    }
  }
  void pop() {
    members.erase(heap.top());
    heap.pop();
  }
  const SNode& top() const { return heap.top(); }
  bool empty() const { return heap.empty(); }
  void clear() {
    while (!heap.empty()) {
      pop();
    }
  }
};

// --------------------------------------------------------------------------
// the mapper: state of HeuristicMapper the path touches
// --------------------------------------------------------------------------
struct NoViableMapping {};

struct CallLog {
  qmg_layer_result rec{};
  std::vector<int16_t> qubitsAfter;
  std::vector<Edge> swaps;
  std::vector<Edge> placed;  // (logical, physical) placed by mapUnmappedGates
};

struct Router {
  const Arch& A;
  qmg_config cfg;
  std::vector<Layer> layers;
  std::vector<int16_t> qubits, locations;
  OpenList open;
  size_t nextId = 0;
  bool fidAware = false, tight = true;
  bool record = false;  // final pass: keep one CallLog per aStarMap call
  std::vector<CallLog> calls;
  std::vector<Edge> placedNow;
  uint64_t expandedAll = 0, generatedAll = 0;
  uint32_t splits = 0;
  std::vector<uint32_t> splitLog;  // layers split in the final pass, in order

  Router(const Arch& a, const qmg_config& c) : A(a), cfg(c) {
    // HeuristicMapper::map :51-52, Heuristic.hpp:96-125
    fidAware = cfg.heuristic == QMG_HEUR_FIDELITY_BEST_LOCATION;
    tight = !fidAware;
    qubits.assign(A.n, -1);
    locations.assign(A.n, -1);
  }

  // HeuristicMapper::getConsideredQubits, HeuristicMapper.hpp:228-233
  const std::set<uint16_t>& considered(size_t l) const {
    return fidAware ? layers[l].act : layers[l].act2;
  }
  // Mapper::getNextLayer, src/sc/Mapper.cpp:437-448
  size_t nextTwoQubitLayer(size_t l) const {
    for (size_t k = l + 1; k < layers.size(); ++k) {
      if (!layers[k].two.empty()) {
        return k;
      }
    }
    return std::numeric_limits<size_t>::max();
  }
  // Mapper::isLayerSplittable, Mapper.cpp:283-300
  bool splittable(size_t l) const {
    const Layer& y = layers[l];
    if (y.two.size() > 1 || y.act1.size() > 2) {
      return true;
    }
    if (y.two.empty()) {
      return false;
    }
    for (const auto q : y.act1) {
      if (y.act2.count(q) == 0) {
        return true;
      }
    }
    return false;
  }
  // Mapper::splitLayer, Mapper.cpp:302-435 (multiplicity view)
  void split(size_t l) {
    const Layer src = layers[l];
    Layer keep, moved;  // keep stays at l, moved goes to l + 1
    keep.one.assign(A.n, 0);
    moved.one.assign(A.n, 0);
    bool toKeep = false;
    for (const auto& kv : src.two) {
      Layer& dst = toKeep ? keep : moved;
      dst.two.insert(kv);
      for (const auto q : {kv.first.first, kv.first.second}) {
        dst.act.insert(q);
        dst.act2.insert(q);
      }
      toKeep = !toKeep;
    }
    toKeep = true;
    for (size_t q = 0; q < src.one.size(); ++q) {
      if (src.one[q] == 0) {
        continue;
      }
      const auto qq = static_cast<uint16_t>(q);
      Layer* dst = nullptr;
      if (keep.act2.count(qq) != 0) {
        dst = &keep;
      } else if (moved.act2.count(qq) != 0) {
        dst = &moved;
      } else {
        dst = toKeep ? &keep : &moved;
        toKeep = !toKeep;
      }
      dst->one[q] = src.one[q];
      dst->act.insert(qq);
      dst->act1.insert(qq);
    }
    layers[l] = keep;
    layers.insert(layers.begin() + static_cast<std::ptrdiff_t>(l) + 1, moved);
    ++splits;
    if (record) {
      splitLog.push_back(static_cast<uint32_t>(l));
    }
  }

  // ---- placement of not yet mapped qubits -------------------------------
  void place(uint16_t logical, uint16_t phys) {
    locations[logical] = static_cast<int16_t>(phys);
    qubits[phys] = static_cast<int16_t>(logical);
    placedNow.emplace_back(logical, phys);
  }
  // HeuristicMapper::mapToMinDistance :287-307
  void placeNear(uint16_t anchor, uint16_t target) {
    double best = kInf;
    int pos = -1;
    for (uint16_t p = 0; p < A.n; ++p) {
      if (qubits[p] == -1) {
        const double d = A.distance(locations[anchor], p);
        if (d < best) {
          best = d;
          pos = p;
        }
      }
    }
    if (pos < 0) {
      throw NoViableMapping{};
    }
    place(target, static_cast<uint16_t>(pos));
  }
  // HeuristicMapper::mapUnmappedGates :211-285
  void placeUnmapped(size_t l) {
    const Layer& y = layers[l];
    if (fidAware) {
      for (size_t q = 0; q < y.one.size(); ++q) {
        if (y.one[q] == 0 || locations[q] != -1) {
          continue;
        }
        for (uint16_t p = 0; p < A.n; ++p) {
          if (qubits[p] == -1) {
            place(static_cast<uint16_t>(q), p);
            break;
          }
        }
      }
    }
    for (const auto& kv : y.two) {
      const auto [q1, q2] = kv.first;
      const bool free1 = locations[q1] == -1, free2 = locations[q2] == -1;
      if (free1 && free2) {
        std::optional<Edge> pick;
        for (const auto& e : A.cm) {  // first free edge in set order
          if (qubits[e.first] == -1 && qubits[e.second] == -1) {
            pick = e;
            break;
          }
        }
        if (!pick) {
          double best = std::numeric_limits<int>::max();
          for (uint16_t i = 0; i < A.n; ++i) {
            for (uint16_t j = i + 1; j < A.n; ++j) {
              if (qubits[i] == -1 && qubits[j] == -1) {
                const double d = A.distance(i, j);
                if (d < best) {
                  best = d;
                  pick = Edge{i, j};
                }
              }
            }
          }
        }
        if (!pick) {
          throw NoViableMapping{};
        }
        place(q1, pick->first);
        place(q2, pick->second);
      } else if (free1) {
        placeNear(q2, q1);
      } else if (free2) {
        placeNear(q1, q2);
      }
    }
  }

  // ---- fixed costs -------------------------------------------------------
  // recalculateFixedCostReversals :771-796
  void reversalCost(size_t l, SNode& v) const {
    v.reversals = 0.;
    const auto& two = layers[l].two;
    if (A.bidir || fidAware || v.valid.size() != two.size()) {
      return;
    }
    for (const auto& [e, m] : two) {
      const int p1 = v.locations[e.first], p2 = v.locations[e.second];
      if (!A.connected(p1, p2)) {
        v.reversals += m.fwd * kGateCostReverse;
      } else if (!A.connected(p2, p1)) {
        v.reversals += m.rev * kGateCostReverse;
      }
    }
  }
  // recalculateFixedCost :750-769 with ...NonFidelity :798-819 and
  // ...Fidelity :821-859
  void fixedCost(size_t l, SNode& v) const {
    const Layer& y = layers[l];
    v.valid.clear();
    for (const auto& kv : y.two) {
      const auto [q1, q2] = kv.first;
      if (A.connected(v.locations[q1], v.locations[q2], false)) {
        v.valid.emplace(q1, q2);
      }
    }
    v.fixed = 0.;
    if (fidAware) {
      for (uint16_t q = 0; q < A.n; ++q) {
        if (y.one[q] != 0) {
          v.fixed += y.one[q] * A.c1[static_cast<size_t>(v.locations[q])];
        }
      }
      for (const auto& s : v.swaps) {
        v.fixed += A.cswap[s.first][s.second];
      }
      for (const auto& [e, m] : y.two) {
        if (v.valid.count(e) == 0) {
          continue;
        }
        const auto p1 = static_cast<size_t>(v.locations[e.first]);
        const auto p2 = static_cast<size_t>(v.locations[e.second]);
        v.fixed += (m.fwd * A.c2[p1][p2] + m.rev * A.c2[p2][p1]);
      }
    } else {
      for (const auto& s : v.swaps) {
        v.fixed += A.swapGateCost(s.first, s.second);
      }
    }
    reversalCost(l, v);
  }

  // ---- heuristics --------------------------------------------------------
  // heuristicGateCountMaxDistance :1082-1120
  double hMaxDistance(size_t l, const SNode& v) const {
    if (v.goal) {
      return 0.;
    }
    double h = 0.;
    for (const auto& [e, m] : layers[l].two) {
      const int p1 = v.locations[e.first], p2 = v.locations[e.second];
      if (!A.bidir && v.valid.count(e) != 0) {
        if (!A.connected(p1, p2)) {
          h = std::max(h, static_cast<double>(m.fwd * kGateCostReverse));
        } else if (!A.connected(p2, p1)) {
          h = std::max(h, static_cast<double>(m.rev * kGateCostReverse));
        }
      } else {
        if (m.fwd > 0) {
          h = std::max(h, A.distance(p1, p2));
        }
        if (m.rev > 0) {
          h = std::max(h, A.distance(p2, p1));
        }
      }
    }
    return h;
  }
  // heuristicGateCountSumDistance :1122-1164
  double hSumDistance(size_t l, const SNode& v) const {
    if (v.goal) {
      return 0.;
    }
    double h = 0.;
    for (const auto& [e, m] : layers[l].two) {
      const int p1 = v.locations[e.first], p2 = v.locations[e.second];
      if (!A.bidir && v.valid.count(e) != 0) {
        if (!A.connected(p1, p2)) {
          h += m.fwd * kGateCostReverse;
        } else if (!A.connected(p2, p1)) {
          h += m.rev * kGateCostReverse;
        }
      } else {
        double c = 0.;
        if (m.fwd == 0) {
          c = A.distance(p2, p1);
        } else if (m.rev == 0) {
          c = A.distance(p1, p2);
        } else {
          c = std::max(A.distance(p1, p2), A.distance(p2, p1));
        }
        h += c;
      }
    }
    return h;
  }
  // heuristicGateCountSumDistanceMinusSharedSwaps :1166-1253
  double hSumMinusShared(size_t l, const SNode& v) const {
    if (v.goal) {
      return 0.;
    }
    double h = 0., rev = 0.;
    std::vector<size_t> hops;
    for (const auto& [e, m] : layers[l].two) {
      const int p1 = v.locations[e.first], p2 = v.locations[e.second];
      if (A.unidir) {
        rev += std::min(m.fwd, m.rev) * kGateCostReverse;
      }
      if (v.valid.count(e) != 0) {
        continue;
      }
      double c = 0.;
      if (m.fwd == 0) {
        c = A.distance(p2, p1, false);
      } else if (m.rev == 0) {
        c = A.distance(p1, p2, false);
      } else {
        c = std::min(A.distance(p1, p2, false), A.distance(p2, p1, false));
      }
      h += c;
      hops.push_back(static_cast<size_t>(
          c / (A.unidir ? kGateCostSwapUni : kGateCostSwapBi)));
    }
    std::sort(hops.begin(), hops.end(), std::greater<>());
    size_t share = 0;
    // (the reference's bound is hops.size() - 1 on an unsigned size; hops is
    // never empty here because a non-goal node has an invalid gate)
    for (size_t i = 0; i + 1 < hops.size(); ++i) {
      size_t mine = 0;
      for (size_t j = i + 1; j < hops.size() && mine < hops[i]; ++j) {
        if (hops[j] > 0) {
          ++mine;
          --hops[j];
        }
      }
      share += mine;
    }
    share = v.shared < share ? share - v.shared : 0;
    const double saved = static_cast<double>(
        share * (A.bidir ? kGateCostSwapBi : kGateCostSwapUni));
    return std::max(0., h - saved) + rev;
  }
  // heuristicFidelityBestLocation :1262-1354
  double hFidelity(size_t l, const SNode& v) const {
    const Layer& y = layers[l];
    const size_t skip = considered(l).size() - 1;
    double h = 0., savings = 0.;
    for (uint16_t q = 0; q < A.n; ++q) {
      if (y.one[q] == 0) {
        continue;
      }
      const int here = v.locations[q];
      const double cur = A.c1[static_cast<size_t>(here)];
      double bestGain = 0.;
      for (uint16_t p = 0; p < A.n; ++p) {
        if (A.c1[p] >= cur) {
          continue;
        }
        const double gain =
            y.one[q] * (cur - A.c1[p]) - A.fidDistance(here, p, skip);
        bestGain = std::max(bestGain, gain);
      }
      savings += bestGain;
    }
    for (const auto& [e, m] : y.two) {
      const int l1 = v.locations[e.first], l2 = v.locations[e.second];
      const bool done = v.valid.count(e) != 0;
      double best = kInf;
      for (const auto& [a, b] : A.cm) {
        best = std::min(best, m.fwd * A.c2[a][b] + m.rev * A.c2[b][a] +
                                  A.fidDistance(l1, a, skip) +
                                  A.fidDistance(l2, b, skip));
        best = std::min(best, m.fwd * A.c2[b][a] + m.rev * A.c2[a][b] +
                                  A.fidDistance(l2, a, skip) +
                                  A.fidDistance(l1, b, skip));
      }
      if (done) {
        const auto p1 = static_cast<size_t>(l1), p2 = static_cast<size_t>(l2);
        const double now = (m.fwd * A.c2[p1][p2] + m.rev * A.c2[p2][p1]);
        savings += (now - best);
      } else {
        h += best;
      }
    }
    return h - savings;
  }
  // updateHeuristicCost :1055-1080
  void heuristicCost(size_t l, SNode& v) const {
    v.goal = v.valid.size() == layers[l].two.size();
    switch (cfg.heuristic) {
    case QMG_HEUR_GATE_COUNT_MAX_DISTANCE:
      v.heur = hMaxDistance(l, v);
      break;
    case QMG_HEUR_GATE_COUNT_SUM_DISTANCE:
      v.heur = hSumDistance(l, v);
      break;
    case QMG_HEUR_GATE_COUNT_SUM_DISTANCE_MINUS_SHARED_SWAPS:
      v.heur = hSumMinusShared(l, v);
      break;
    case QMG_HEUR_GATE_COUNT_MAX_DISTANCE_OR_SUM_DISTANCE_MINUS_SHARED_SWAPS:
      v.heur = std::max(hMaxDistance(l, v), hSumMinusShared(l, v));  // :1255
      break;
    default:
      v.heur = hFidelity(l, v);
      break;
    }
  }

  // ---- lookahead ---------------------------------------------------------
  // lookaheadGateCountMaxDistance :1387-1449 / ...SumDistance :1451-1513
  double lookaheadLayer(size_t l, const SNode& v, bool sum) const {
    double pen = 0.;
    for (const auto& [e, m] : layers[l].two) {
      const int l1 = v.locations[e.first], l2 = v.locations[e.second];
      double c = 0.;
      if (l1 == -1 && l2 == -1) {
        continue;
      }
      if (l1 == -1 || l2 == -1) {
        // one end unplaced: nearest free physical qubit
        const int fixedEnd = l1 == -1 ? l2 : l1;
        const bool freeIsFirst = l1 == -1;
        c = kInf;
        for (uint16_t p = 0; p < A.n; ++p) {
          if (v.qubits[p] != -1) {
            continue;
          }
          const int from = freeIsFirst ? p : fixedEnd;
          const int to = freeIsFirst ? fixedEnd : p;
          if (m.fwd > 0) {
            c = std::min(c, A.distance(from, to));
          }
          if (m.rev > 0) {
            c = std::min(c, A.distance(to, from));
          }
        }
      } else {
        c = kInf;
        if (m.fwd > 0) {
          c = std::min(c, A.distance(l1, l2));
        }
        if (m.rev > 0) {
          c = std::min(c, A.distance(l2, l1));
        }
      }
      pen = sum ? pen + c : std::max(pen, c);
    }
    return pen;
  }
  // updateLookaheadPenalty :1356-1385
  void lookahead(size_t l, SNode& v) const {
    v.look = 0.;
    size_t nx = nextTwoQubitLayer(l);
    double factor = cfg.first_lookahead_factor;
    for (uint32_t i = 0; i < cfg.nr_lookaheads; ++i) {
      if (nx == std::numeric_limits<size_t>::max()) {
        break;
      }
      double pen = 0.;
      if (cfg.lookahead_heuristic == QMG_LOOKAHEAD_GATE_COUNT_MAX_DISTANCE) {
        pen = lookaheadLayer(nx, v, false);
      } else if (cfg.lookahead_heuristic ==
                 QMG_LOOKAHEAD_GATE_COUNT_SUM_DISTANCE) {
        pen = lookaheadLayer(nx, v, true);
      }
      v.look += factor * pen;
      factor *= cfg.lookahead_factor;
      nx = nextTwoQubitLayer(nx);
    }
  }

  // ---- node expansion ----------------------------------------------------
  // updateSharedSwaps :975-1053
  void sharedSwap(const Edge& sw, size_t l, SNode& v) const {
    const auto& cons = considered(l);
    const int q1 = v.qubits[sw.first], q2 = v.qubits[sw.second];
    if (q1 == -1 || q2 == -1 || cons.count(static_cast<uint16_t>(q1)) == 0 ||
        cons.count(static_cast<uint16_t>(q2)) == 0) {
      return;
    }
    int partner1 = q1, partner2 = q2;
    for (const auto& kv : layers[l].two) {
      const auto [a, b] = kv.first;
      if (a == q1) {
        partner1 = b;
      } else if (b == q1) {
        partner1 = a;
      }
      if (a == q2) {
        partner2 = b;
      } else if (b == q2) {
        partner2 = a;
      }
    }
    if (partner1 == q1 || partner2 == q2 || partner1 == q2) {
      return;
    }
    const int p3 = v.locations[static_cast<size_t>(partner1)];
    const int p4 = v.locations[static_cast<size_t>(partner2)];
    auto d = [&](int x, int y) {
      if (fidAware) {
        return std::min(A.fidDistance(x, y, 0), A.fidDistance(y, x, 0));
      }
      return std::min(A.distance(x, y, false), A.distance(y, x, false));
    };
    const double before1 = d(sw.first, p3), after1 = d(sw.second, p3);
    const double before2 = d(sw.second, p4), after2 = d(sw.first, p4);
    if (after1 < before1 && after2 < before2) {
      ++v.shared;
    }
  }
  // applySWAP :861-973
  void applySwap(const Edge& sw, size_t l, SNode& v) const {
    const Layer& y = layers[l];
    const int q1 = v.qubits[sw.first], q2 = v.qubits[sw.second];
    sharedSwap(sw, l, v);
    v.qubits[sw.first] = static_cast<int16_t>(q2);
    v.qubits[sw.second] = static_cast<int16_t>(q1);
    if (q1 != -1) {
      v.locations[static_cast<size_t>(q1)] = static_cast<int16_t>(sw.second);
    }
    if (q2 != -1) {
      v.locations[static_cast<size_t>(q2)] = static_cast<int16_t>(sw.first);
    }
    v.swaps.push_back(sw);
    for (const auto& [e, m] : y.two) {
      const int q3 = e.first, q4 = e.second;
      if (q3 != q1 && q3 != q2 && q4 != q1 && q4 != q2) {
        continue;
      }
      const auto p3 = static_cast<uint16_t>(v.locations[e.first]);
      const auto p4 = static_cast<uint16_t>(v.locations[e.second]);
      if (A.connected(p3, p4, false)) {
        if (fidAware && v.valid.count(e) == 0) {
          v.fixed += m.fwd * A.c2[p3][p4] + m.rev * A.c2[p4][p3];
        }
        v.valid.insert(e);
      } else {
        if (fidAware && v.valid.count(e) != 0) {
          auto undo = [&](uint16_t p) {
            return p == sw.first ? sw.second : (p == sw.second ? sw.first : p);
          };
          const uint16_t o3 = undo(p3), o4 = undo(p4);
          v.fixed -= m.fwd * A.c2[o3][o4] + m.rev * A.c2[o4][o3];
        }
        v.valid.erase(e);
      }
    }
    if (fidAware) {
      uint16_t m1 = 0, m2 = 0;
      if (q1 != -1) {
        m1 = y.one[static_cast<size_t>(q1)];
      }
      if (q2 != -1) {
        m2 = y.one[static_cast<size_t>(q2)];
      }
      // (int arithmetic on the promoted uint16 values, as in the reference)
      v.fixed += ((m2 - m1) * A.c1[sw.first] + (m1 - m2) * A.c1[sw.second]);
      v.fixed += A.cswap[sw.first][sw.second];
    } else {
      v.fixed += A.swapGateCost(sw.first, sw.second);
    }
    reversalCost(l, v);
    heuristicCost(l, v);
    if (cfg.lookahead_heuristic != QMG_LOOKAHEAD_NONE) {
      lookahead(l, v);
    }
  }
  // expandNodeAddOneSwap :729-748
  void addChild(const Edge& sw, const SNode& parent, size_t l) {
    SNode c;
    c.id = nextId++;
    c.parent = parent.id;
    c.qubits = parent.qubits;
    c.locations = parent.locations;
    c.swaps = parent.swaps;
    c.valid = parent.valid;
    c.fixed = parent.fixed;
    c.reversals = parent.reversals;
    c.depth = parent.depth + 1;
    c.shared = parent.shared;
    // heur / look / goal start from the Node defaults (0, 0, true)
    if (A.connected(sw.first, sw.second, false)) {
      applySwap(sw, l, c);
    }
    open.push(c);
  }
  // expandNode :698-727
  void expand(const SNode& v, size_t l) {
    std::vector<std::vector<bool>> used(A.n, std::vector<bool>(A.n, false));
    for (const auto q : considered(l)) {
      for (const auto& e : A.cm) {
        if (e.first != v.locations[q] && e.second != v.locations[q]) {
          continue;
        }
        const int q1 = v.qubits[e.first], q2 = v.qubits[e.second];
        if (q1 == -1 || q2 == -1) {
          addChild(e, v, l);
        } else if (!used[static_cast<size_t>(q1)][static_cast<size_t>(q2)]) {
          used[static_cast<size_t>(q1)][static_cast<size_t>(q2)] = true;
          used[static_cast<size_t>(q2)][static_cast<size_t>(q1)] = true;
          addChild(e, v, l);
        }
      }
    }
  }

  // ---- aStarMap :521-696 -------------------------------------------------
  SNode search(size_t l, bool reverse) {
    nextId = 0;
    SNode root;
    root.id = nextId++;
    SNode best;
    bool found = false;
    placeUnmapped(l);
    root.locations = locations;
    root.qubits = qubits;
    fixedCost(l, root);
    heuristicCost(l, root);
    lookahead(l, root);
    open.push(root);

    uint64_t expanded = 0, afterFirst = 0, afterBest = 0, goals = 0,
             goalsAfterBest = 0;
    bool early = false;
    const bool maySplit = cfg.automatic_layer_splits != 0 ? splittable(l) : false;
    while (!open.empty() &&
           (!found || open.top().total() < best.totalFixed())) {
      if (maySplit && expanded >= cfg.automatic_layer_splits_node_limit) {
        split(l);
        // the open list is NOT cleared here: the restarted search inherits it
        return search(reverse ? l + 1 : l, reverse);
      }
      SNode cur = open.top();
      if (cur.goal) {
        ++goals;
        if (!found || cur.totalFixed() < best.totalFixed()) {
          best = cur;
          afterBest = 0;
          goalsAfterBest = 0;
        } else {
          ++goalsAfterBest;
        }
        found = true;
        if (tight) {
          break;
        }
      }
      open.pop();
      expand(cur, l);
      ++expanded;
      if (found) {
        ++afterFirst;
        ++afterBest;
        if (cfg.early_termination != QMG_ET_NONE) {
          uint64_t k = 0;
          switch (cfg.early_termination) {
          case QMG_ET_EXPANDED_NODES:
            k = expanded;
            break;
          case QMG_ET_EXPANDED_NODES_AFTER_FIRST_SOLUTION:
            k = afterFirst;
            break;
          case QMG_ET_EXPANDED_NODES_AFTER_CURRENT_OPTIMAL_SOLUTION:
            k = afterBest;
            break;
          case QMG_ET_SOLUTION_NODES:
            k = goals;
            break;
          case QMG_ET_SOLUTION_NODES_AFTER_CURRENT_OPTIMAL_SOLUTION:
            k = goalsAfterBest;
            break;
          default:
            break;
          }
          if (k >= cfg.early_termination_limit) {
            early = true;
            break;
          }
        }
      }
    }
    if (!found) {
      throw NoViableMapping{};
    }
    expandedAll += expanded;
    generatedAll += nextId;
    if (record) {
      CallLog c;
      c.rec.cost_fixed = best.fixed;
      c.rec.cost_fixed_reversals = best.reversals;
      c.rec.cost_heur = best.heur;
      c.rec.lookahead_penalty = best.look;
      c.rec.expanded_nodes = expanded;
      c.rec.generated_nodes = nextId;
      c.rec.expanded_nodes_after_first_solution = afterFirst;
      c.rec.expanded_nodes_after_optimal_solution = afterBest;
      c.rec.solution_nodes = goals;
      c.rec.solution_nodes_after_optimal_solution = goalsAfterBest;
      c.rec.solution_depth = static_cast<uint32_t>(best.depth);
      c.rec.result_node_id = static_cast<uint32_t>(best.id);
      c.rec.n_swaps = static_cast<uint32_t>(best.swaps.size());
      c.rec.n_placements = static_cast<uint32_t>(placedNow.size());
      c.rec.early_termination = early ? 1 : 0;
      c.rec.status = QMG_OK;
      c.qubitsAfter = best.qubits;
      c.swaps = best.swaps;
      c.placed = placedNow;
      calls.push_back(std::move(c));
    }
    placedNow.clear();
    open.clear();
    return best;
  }

  // pseudoRouteCircuit :309-344 (layers are restored afterwards)
  void pseudoRoute(bool reverse) {
    const auto saved = layers;
    for (size_t i = 0; i < layers.size(); ++i) {
      const size_t l = reverse ? layers.size() - i - 1 : i;
      const SNode r = search(l, reverse);
      qubits = r.qubits;
      locations = r.locations;
    }
    layers = saved;
  }
  // HeuristicMapper::map :74-91 + the layer loop of routeCircuit :346-357
  void run(uint32_t passes) {
    for (uint32_t i = 0; i < passes; ++i) {
      pseudoRoute(false);
      pseudoRoute(true);
    }
    record = true;
    for (size_t l = 0; l < layers.size(); ++l) {
      const SNode r = search(l, false);
      qubits = r.qubits;
      locations = r.locations;
    }
  }
};

}  // namespace

// ==========================================================================
// C entry points (loaded by tests/ and bench.py's CPU legs through ctypes)
// ==========================================================================
#define QO_API extern "C" __attribute__((visibility("default")))

struct qo_arch {
  Arch a;
};

// Architecture(nQ, couplingMap): distance tables are built eagerly
QO_API qo_arch* qo_arch_create(uint16_t n, const uint16_t* edges,
                               size_t nEdges) {
  auto* h = new qo_arch;
  h->a = makeArch(n, edges, nEdges);
  return h;
}
QO_API void qo_arch_destroy(qo_arch* h) { delete h; }

// the cost part of Architecture::createFidelityTable (Architecture.cpp:
// 232-294): the caller passes the -log2 costs, the edge-skip tables are built
// here (:291-293)
QO_API void qo_arch_set_fidelity(qo_arch* h, const double* c1, const double* c2,
                                 const double* cswap) {
  Arch& a = h->a;
  a.c1.assign(c1, c1 + a.n);
  a.c2 = toMatrix(c2, a.n);
  a.cswap = toMatrix(cswap, a.n);
  a.fid = skipEdgeLevels(a.cm, a.cswap);
  a.fidelity = true;
}
QO_API void qo_arch_tables(const qo_arch* h, double* dist, double* distRev) {
  fromMatrix(h->a.dist, dist);
  fromMatrix(h->a.distRev, distRev);
}
QO_API void qo_arch_flags(const qo_arch* h, int* bidirectional,
                          int* unidirectional) {
  *bidirectional = h->a.bidir ? 1 : 0;
  *unidirectional = h->a.unidir ? 1 : 0;
}
QO_API uint32_t qo_arch_fidelity_levels(const qo_arch* h) {
  return static_cast<uint32_t>(h->a.fid.size());
}
QO_API void qo_arch_fidelity_tables(const qo_arch* h, double* out) {
  const size_t nn = static_cast<size_t>(h->a.n) * h->a.n;
  for (size_t k = 0; k < h->a.fid.size(); ++k) {
    fromMatrix(h->a.fid[k], out + k * nn);
  }
}

// stand-alone table builders (the reference's test_general.cpp drives these)
static EdgeSet edgeSet(const uint16_t* edges, size_t nEdges) {
  EdgeSet cm;
  for (size_t i = 0; i < nEdges; ++i) {
    cm.emplace(edges[2 * i], edges[2 * i + 1]);
  }
  return cm;
}
QO_API void qo_build_table(uint16_t n, const uint16_t* edges, size_t nEdges,
                           const double* weights, double* out) {
  fromMatrix(allPairs(edgeSet(edges, nEdges), toMatrix(weights, n)), out);
}
QO_API void qo_single_edge_skip(uint16_t n, const uint16_t* edges,
                                size_t nEdges, const double* dist,
                                double reversalCost, double* out) {
  fromMatrix(skipOneEdge(toMatrix(dist, n), edgeSet(edges, nEdges), reversalCost),
             out);
}
// returns the number of levels; writes at most maxLevels tables
QO_API uint32_t qo_edge_skip_levels(uint16_t n, const uint16_t* edges,
                                    size_t nEdges, const double* weights,
                                    uint32_t maxLevels, double* out) {
  const auto lv = skipEdgeLevels(edgeSet(edges, nEdges), toMatrix(weights, n));
  const size_t nn = static_cast<size_t>(n) * n;
  for (size_t k = 0; k < lv.size() && k < maxLevels; ++k) {
    fromMatrix(lv[k], out + k * nn);
  }
  return static_cast<uint32_t>(lv.size());
}

// Routes one instance (all pseudo passes + the final pass).  `res` uses the
// ABI's result struct; res->layers / layer_qubits / layer_n_swaps have room
// for layerCapacity layers (layer splits add layers).  Returns a qmg_status.
QO_API int qo_route(const qo_arch* h, const qmg_config* cfg,
                    const qmg_instance* inst, qmg_route_result* res,
                    uint32_t layerCapacity) {
  const Arch& a = h->a;
  if (cfg->heuristic == QMG_HEUR_FIDELITY_BEST_LOCATION &&
      (!a.fidelity || cfg->lookahead_heuristic != QMG_LOOKAHEAD_NONE)) {
    return QMG_BAD_ARG;  // HeuristicMapper::checkParameters :149-167
  }
  Router r(a, *cfg);
  r.layers = readLayers(inst->layers, a.n);
  if (inst->initial_locations != nullptr) {
    for (uint16_t q = 0; q < a.n; ++q) {
      const int16_t p = inst->initial_locations[q];
      if (p >= 0) {
        r.locations[q] = p;
        r.qubits[static_cast<size_t>(p)] = static_cast<int16_t>(q);
      }
    }
  }
  int status = QMG_OK;
  try {
    r.run(inst->bidirectional_passes);
  } catch (const NoViableMapping&) {
    status = QMG_NO_VIABLE_MAPPING;
  }
  uint32_t nSwaps = 0, nPlaced = 0;
  const size_t nL = r.calls.size();
  for (size_t l = 0; l < nL; ++l) {
    CallLog& c = r.calls[l];
    c.rec.swap_offset = nSwaps;
    c.rec.placement_offset = nPlaced;
    if (l < layerCapacity) {
      if (res->layers != nullptr) {
        res->layers[l] = c.rec;
      }
      if (res->layer_n_swaps != nullptr) {
        res->layer_n_swaps[l] = static_cast<uint16_t>(c.swaps.size());
      }
      if (res->layer_qubits != nullptr) {
        std::copy(c.qubitsAfter.begin(), c.qubitsAfter.end(),
                  res->layer_qubits + l * a.n);
      }
    }
    for (const auto& s : c.swaps) {
      if (res->swaps != nullptr && nSwaps < res->swap_capacity) {
        res->swaps[2 * nSwaps] = s.first;
        res->swaps[2 * nSwaps + 1] = s.second;
      }
      ++nSwaps;
    }
    for (const auto& p : c.placed) {
      if (res->placements != nullptr && nPlaced < res->placement_capacity) {
        res->placements[nPlaced].logical = p.first;
        res->placements[nPlaced].physical = p.second;
      }
      ++nPlaced;
    }
  }
  res->n_swaps_total = nSwaps;
  res->n_placements_total = nPlaced;
  res->n_layers_done = static_cast<uint32_t>(nL);
  res->n_splits = static_cast<uint32_t>(r.splitLog.size());
  for (size_t k = 0; k < r.splitLog.size(); ++k) {
    if (res->split_layers != nullptr && k < res->split_capacity) {
      res->split_layers[k] = r.splitLog[k];
    }
  }
  res->status = status;
  res->expanded_nodes = r.expandedAll;
  res->generated_nodes = r.generatedAll;
  if (res->final_qubits != nullptr) {
    std::copy(r.qubits.begin(), r.qubits.end(), res->final_qubits);
  }
  if (res->final_locations != nullptr) {
    std::copy(r.locations.begin(), r.locations.end(), res->final_locations);
  }
  return status;
}

QO_API const char* qo_version(void) { return "qmap_oracle 1 (restatement)"; }
