// GpuHeuristicMapper.hpp - the reference-side binding of libqmap_b200.so: a
// drop-in for HeuristicMapper behind the reference's own C++ API.
//
// This file belongs on the REFERENCE's side of the boundary: it includes the
// reference's headers (include/sc/heuristic/HeuristicMapper.hpp), derives from
// HeuristicMapper and is what a maintainer of cda-tum/mqt-qmap would add next
// to it.  bindings/bindings.cpp:69 then constructs GpuHeuristicMapper instead
// of HeuristicMapper; everything else of map() - createLayers, the initial
// layout, routeCircuit's emission of SWAP / CX / H, gate counting, the result
// JSON - stays the reference's code.
//
// Configuration::dataLoggingPath works as in the reference: the device's node
// log is replayed through the reference's own DataLogger, layer by layer.
//
// What changes: the virtual aStarMap (HeuristicMapper.hpp:220) is answered
// from ONE qmg_route_batch call that routes the whole circuit on the device
// (all pseudo-routing passes, placements of mapUnmappedGates, every layer's
// A* search, layer splits).  pseudoRouteCircuit is not virtual
// (HeuristicMapper.hpp:207), so map() hands the base class
// iterativeBidirectionalRoutingPasses = 0 and the device the real count.
//
// Built and tested here by oracle/Makefile (target dropin) against the
// stand-in mqt-core headers; tests/test_gpu_dropin.py runs the reference's own
// map() through it and compares with plain HeuristicMapper call by call.
#pragma once

#include "sc/Architecture.hpp"
#include "sc/heuristic/HeuristicMapper.hpp"
#include "sc/utils.hpp"

#include "qmap_b200.h"

#include <cstdint>
#include <cstring>
#include <string>
#include <vector>

class GpuHeuristicMapper : public HeuristicMapper {
public:
  using HeuristicMapper::HeuristicMapper;
  ~GpuHeuristicMapper() override { release(); }

  void setDevice(int device) { device_ = device; }

  void map(const Configuration& configuration) override {
    devicePasses_ = configuration.iterativeBidirectionalRoutingPasses;
    Configuration base = configuration;
    base.iterativeBidirectionalRoutingPasses = 0;  // (the device runs them)
    routed_ = false;
    HeuristicMapper::map(base);
    results.config.iterativeBidirectionalRoutingPasses = devicePasses_;
  }

protected:
  Node aStarMap(std::size_t layer, bool reverse) override {
    if (reverse) {  // (never asked for: the base runs no pseudo-routing pass)
      return HeuristicMapper::aStarMap(layer, reverse);
    }
    if (!routed_) {
      routeOnDevice();
    }
    const auto n = architecture->getNqubits();
    const qmg_layer_result& r = rec_.at(layer);
    std::vector<std::int16_t> q(layerQ_.begin() + static_cast<std::ptrdiff_t>(layer * n),
                                layerQ_.begin() + static_cast<std::ptrdiff_t>((layer + 1) * n));
    std::vector<std::int16_t> loc(n, DEFAULT_POSITION);
    for (std::uint16_t p = 0; p < n; ++p) {
      if (q[p] >= 0) {
        loc[static_cast<std::size_t>(q[p])] = static_cast<std::int16_t>(p);
      }
    }
    std::vector<Exchange> sw;
    sw.reserve(r.n_swaps);
    for (std::uint32_t k = 0; k < r.n_swaps; ++k) {
      sw.emplace_back(swaps_[2 * (r.swap_offset + k)],
                      swaps_[2 * (r.swap_offset + k) + 1], qc::SWAP);
    }
    Node node(r.result_node_id, 0, q, loc, sw, {}, r.cost_fixed,
              r.cost_fixed_reversals, r.solution_depth);
    node.costHeur = r.cost_heur;
    node.lookaheadPenalty = r.lookahead_penalty;
    node.validMapping = true;
    if (results.config.debug) {  // HeuristicMapper.cpp:642-675
      auto& b = results.layerHeuristicBenchmark.emplace_back();
      b.expandedNodes = r.expanded_nodes;
      b.generatedNodes = r.generated_nodes;
      b.expandedNodesAfterFirstSolution = r.expanded_nodes_after_first_solution;
      b.expandedNodesAfterOptimalSolution = r.expanded_nodes_after_optimal_solution;
      b.solutionNodes = r.solution_nodes;
      b.solutionNodesAfterOptimalSolution = r.solution_nodes_after_optimal_solution;
      b.solutionDepth = r.solution_depth;
      b.earlyTermination = r.early_termination != 0;
      results.heuristicBenchmark.expandedNodes += b.expandedNodes;
      results.heuristicBenchmark.generatedNodes += b.generatedNodes;
      // device time of the launch, spread over the layers by expanded nodes
      const double secs = secondsPerExpandedNode_ * static_cast<double>(b.expandedNodes);
      results.heuristicBenchmark.secondsPerNode += secs;
      if (b.expandedNodes > 0) {
        b.secondsPerNode = secondsPerExpandedNode_;
        b.averageBranchingFactor = static_cast<double>(b.generatedNodes - 1) /
                                   static_cast<double>(b.expandedNodes);
      }
      b.effectiveBranchingFactor =
          computeEffectiveBranchingRate(b.expandedNodes + 1, r.solution_depth);
    }
    if (results.config.dataLoggingEnabled()) {
      logLayer(layer, node);
    }
    return node;
  }

private:
  qmg_arch* arch_ = nullptr;
  int device_ = 0;
  std::size_t devicePasses_ = 0;
  bool routed_ = false;
  double secondsPerExpandedNode_ = 0.;
  std::vector<qmg_layer_result> rec_;
  std::vector<std::int16_t> layerQ_;
  std::vector<std::uint16_t> swaps_;
  // Configuration::dataLoggingPath: every node the device generated in the
  // final pass (qmg_route_result.node_log) and every layer's root layout
  std::vector<qmg_node_log_record> nodeLog_;
  std::vector<std::int16_t> rootQ_;
  std::size_t nodeLogPos_ = 0;

  // what aStarMap writes through the DataLogger for one layer
  // (HeuristicMapper.cpp:541-546, :741-747 logSearchNode for every generated
  // node in generation order; :676-688 logFinalizeLayer), from the device's
  // node log: a node's layout and SWAP list follow from its parent's
  void logLayer(std::size_t layer, const Node& result) {
    const auto n = architecture->getNqubits();
    struct Made {
      std::vector<std::int16_t> qubits;
      std::vector<Exchange> swaps;
    };
    std::vector<Made> made;  // by node id (ids are dense per layer)
    const std::vector<std::int16_t> root(
        rootQ_.begin() + static_cast<std::ptrdiff_t>(layer * n),
        rootQ_.begin() + static_cast<std::ptrdiff_t>((layer + 1) * n));
    for (; nodeLogPos_ < nodeLog_.size() && nodeLog_[nodeLogPos_].layer == layer;
         ++nodeLogPos_) {
      const qmg_node_log_record& r = nodeLog_[nodeLogPos_];
      Made m;
      if (r.root != 0) {
        m.qubits = root;
      } else {
        m = made.at(r.parent);
        std::swap(m.qubits.at(r.swap_first), m.qubits.at(r.swap_second));
        m.swaps.emplace_back(r.swap_first, r.swap_second, qc::SWAP);
      }
      dataLogger->logSearchNode(layer, r.id, r.parent, r.cost_fixed, r.cost_heur,
                                r.lookahead_penalty, m.qubits, r.valid != 0, m.swaps,
                                r.depth);
      if (made.size() <= r.id) {
        made.resize(static_cast<std::size_t>(r.id) + 1);
      }
      made[r.id] = std::move(m);
    }
    qc::CompoundOperation compOp{};
    for (const auto& gate : layers.at(layer)) {
      compOp.emplace_back(gate.op->clone());
    }
    dataLogger->logFinalizeLayer(layer, compOp, singleQubitMultiplicities.at(layer),
                                 twoQubitMultiplicities.at(layer), root, result.id,
                                 result.costFixed, result.costHeur, result.lookaheadPenalty,
                                 result.qubits, result.swaps, result.depth);
  }

  void release() {
    if (arch_ != nullptr) {
      qmg_arch_destroy(arch_);
      arch_ = nullptr;
    }
  }
  static void check(int rc) {
    if (rc == QMG_NO_VIABLE_MAPPING) {
      throw QMAPException("No viable mapping found.");  // HeuristicMapper.cpp:637-639
    }
    if (rc != QMG_OK) {
      throw QMAPException(std::string("qmap_b200: ") + qmg_last_error());
    }
  }
  // (HeuristicMapper.cpp:196-208, findAndSWAP, for both permutations)
  void moveInLayouts(qc::Qubit logical, qc::Qubit physical) {
    for (qc::Permutation* map : {&qcMapped.initialLayout, &qcMapped.outputPermutation}) {
      for (const auto& q : *map) {
        if (q.second == logical) {
          std::swap(map->at(physical), map->at(q.first));
          break;
        }
      }
    }
  }
  static std::vector<double> flat(const Matrix& m) {
    std::vector<double> v;
    v.reserve(m.size() * m.size());
    for (const auto& row : m) {
      v.insert(v.end(), row.begin(), row.end());
    }
    return v;
  }

  // device-resident Architecture: tables are (re)built by the table kernels
  void uploadArchitecture() {
    release();
    const auto n = architecture->getNqubits();
    std::vector<std::uint16_t> edges;
    for (const auto& [a, b] : architecture->getCouplingMap()) {
      edges.push_back(a);
      edges.push_back(b);
    }
    check(qmg_arch_create(n, edges.data(), edges.size() / 2, device_, &arch_));
    if (architecture->isFidelityAvailable()) {
      const auto c2 = flat(architecture->getTwoQubitFidelityCosts());
      const auto cs = flat(architecture->getSwapFidelityCosts());
      check(qmg_arch_set_fidelity_costs(
          arch_, architecture->getSingleQubitFidelityCosts().data(), c2.data(), cs.data()));
    }
  }

  qmg_config deviceConfig() const {
    const Configuration& c = results.config;
    qmg_config g;
    qmg_config_default(&g);
    g.heuristic = static_cast<std::int32_t>(c.heuristic);
    g.lookahead_heuristic = static_cast<std::int32_t>(c.lookaheadHeuristic);
    g.nr_lookaheads = static_cast<std::uint32_t>(c.nrLookaheads);
    g.first_lookahead_factor = c.firstLookaheadFactor;
    g.lookahead_factor = c.lookaheadFactor;
    g.early_termination = static_cast<std::int32_t>(c.earlyTermination);
    g.early_termination_limit = c.earlyTerminationLimit;
    g.automatic_layer_splits = c.automaticLayerSplits ? 1 : 0;
    g.automatic_layer_splits_node_limit = c.automaticLayerSplitsNodeLimit;
    g.layering = static_cast<std::int32_t>(c.layering);
    g.initial_layout = static_cast<std::int32_t>(c.initialLayout);
    g.iterative_bidirectional_routing_passes = static_cast<std::uint32_t>(devicePasses_);
    g.swap_on_first_layer = c.swapOnFirstLayer ? 1 : 0;
    g.debug = 1;
    return g;
  }

  // the whole circuit in one launch: layers as createLayers left them
  // (Mapper.cpp:138-281), start layout as createInitialMapping left it
  void routeOnDevice() {
    uploadArchitecture();
    const auto n = architecture->getNqubits();
    const std::size_t nL = layers.size();
    std::vector<std::uint32_t> off2{0}, off1{0};
    std::vector<qmg_gate2q> g2;
    std::vector<qmg_gate1q> g1;
    for (std::size_t l = 0; l < nL; ++l) {
      for (const auto& [edge, mult] : twoQubitMultiplicities.at(l)) {
        g2.push_back({edge.first, edge.second, mult.first, mult.second});
      }
      const auto& m1 = singleQubitMultiplicities.at(l);
      for (std::uint16_t q = 0; q < m1.size(); ++q) {
        if (m1[q] != 0) {
          g1.push_back({q, m1[q]});
        }
      }
      off2.push_back(static_cast<std::uint32_t>(g2.size()));
      off1.push_back(static_cast<std::uint32_t>(g1.size()));
    }
    qmg_instance inst{};
    inst.layers.n_layers = static_cast<std::uint32_t>(nL);
    inst.layers.n_logical = n;
    inst.layers.gate2q_offsets = off2.data();
    inst.layers.gates2q = g2.data();
    inst.layers.gate1q_offsets = off1.data();
    inst.layers.gates1q = g1.data();
    inst.initial_locations = locations.data();
    inst.bidirectional_passes = static_cast<std::uint32_t>(devicePasses_);
    // room for every layer a split can add: the library then splits layers
    // itself where the reference would (Mapper::splitLayer + restart)
    const std::size_t cap =
        nL + (results.config.automaticLayerSplits ? g2.size() + g1.size() : 0);
    inst.layer_capacity = static_cast<std::uint32_t>(cap);
    const qmg_config cfg = deviceConfig();
    std::size_t swapCap = 16 * cap + 1024;
    const bool logging = results.config.dataLoggingEnabled();
    std::size_t logCap = logging ? (std::size_t{1} << 20) : 0;
    for (int attempt = 0;; ++attempt) {
      rec_.assign(cap, qmg_layer_result{});
      layerQ_.assign(cap * n, -1);
      swaps_.assign(2 * swapCap, 0);
      std::vector<std::int16_t> fq(n), fl(n);
      std::vector<qmg_placement> places(n);
      std::vector<std::uint32_t> splitLog(cap - nL + 1);
      std::vector<qmg_placement> placeLog(n);
      qmg_route_result r{};
      r.layers = rec_.data();
      r.swaps = swaps_.data();
      r.swap_capacity = static_cast<std::uint32_t>(swapCap);
      r.placements = places.data();
      r.placement_capacity = n;
      r.final_qubits = fq.data();
      r.final_locations = fl.data();
      r.layer_qubits = layerQ_.data();
      r.split_layers = splitLog.data();
      r.split_capacity = static_cast<std::uint32_t>(cap - nL);
      r.placement_log = placeLog.data();
      r.placement_log_capacity = n;
      if (logging) {
        nodeLog_.assign(logCap, qmg_node_log_record{});
        rootQ_.assign(cap * n, -1);
        r.node_log = nodeLog_.data();
        r.node_log_capacity = static_cast<std::uint32_t>(logCap);
        r.layer_root_qubits = rootQ_.data();
      }
      qmg_route_stats st{};
      check(qmg_route_batch(arch_, &cfg, 1, &inst, &r, nullptr, &st));
      if (r.status == QMG_BUFFER_TOO_SMALL && attempt < 6) {
        swapCap *= 4;
        continue;
      }
      check(r.status);
      if (logging) {
        if (r.n_node_log > logCap && attempt < 6) {
          logCap = r.n_node_log;  // (the log was cut: once more with room for it)
          continue;
        }
        if (r.n_splits != 0) {
          // (node ids restart when the device splits a layer; the reference
          // writes one more file set per split, DataLogger::splitLayer)
          throw QMAPException("qmap_b200: data logging of runs with automatic "
                              "layer splits is not supported");
        }
        nodeLog_.resize(r.n_node_log);
        nodeLogPos_ = 0;
      }
      secondsPerExpandedNode_ =
          r.expanded_nodes > 0 ? 1e-3 * st.kernel_ms / static_cast<double>(r.expanded_nodes) : 0.;
      // mapUnmappedGates / mapToMinDistance also move the placed qubit in the
      // mapped circuit's initialLayout and outputPermutation (:275-278,
      // :304-306; not for the 1q-gate rule, :213-231): every placement of
      // every pass, in order
      for (std::uint32_t k = 0; k < r.n_placement_log && k < n; ++k) {
        if ((placeLog[k].physical & 0x8000U) == 0) {
          moveInLayouts(placeLog[k].logical, placeLog[k].physical);
        }
      }
      // the host's layers follow the splits of the final pass
      for (std::uint32_t k = 0; k < r.n_splits; ++k) {
        splitLayer(splitLog[k], *architecture);
      }
      break;
    }
    routed_ = true;
  }
};
