// host_api.hpp - host-only parts of the C ABI (include/qmap_b200.h): error
// string, configuration defaults, layered circuits.  Included by api.cu (the
// product library) and by tests/emu/emu.cpp (the host debug build).
#pragma once
#include <cstring>
#include <exception>
#include <memory>
#include <string>
#include <vector>

#include "../../include/qmap_b200.h"
#include "layers.hpp"

namespace qmg {
inline std::string& lastError() {
  thread_local std::string e;
  return e;
}
inline int fail(int code, const std::string& msg) {
  lastError() = msg;
  return code;
}
// factor table of updateLookaheadPenalty (:1361,:1381)
inline std::vector<double> lookaheadFactors(const qmg_config& c, size_t cap) {
  std::vector<double> f;
  double v = c.first_lookahead_factor;
  for (size_t i = 0; i < c.nr_lookaheads && i < cap; ++i) {
    f.push_back(v);
    v *= c.lookahead_factor;
  }
  return f;
}
inline CfgView toCfgView(const qmg_config& c) {
  CfgView v{};
  v.heuristic = c.heuristic;
  v.lookahead = c.lookahead_heuristic;
  v.nrLookaheads = c.nr_lookaheads;
  v.earlyTermination = c.early_termination;
  v.firstFactor = c.first_lookahead_factor;
  v.factor = c.lookahead_factor;
  v.earlyLimit = c.early_termination_limit;
  v.splitLimit = c.automatic_layer_splits_node_limit;
  v.autoSplit = c.automatic_layer_splits;
  // isTight / isFidelityAware, include/sc/configuration/Heuristic.hpp:96-126
  v.fidelityAware = c.heuristic == QMG_HEUR_FIDELITY_BEST_LOCATION ? 1 : 0;
  v.tight = v.fidelityAware ? 0 : 1;
  return v;
}
// HeuristicMapper::checkParameters :149-167 (the parts visible at this level)
inline const char* checkConfig(const qmg_config& c, bool fidelityAvailable) {
  if (c.heuristic < 0 || c.heuristic > 4) {
    return "Unknown heuristic.";
  }
  if (c.lookahead_heuristic < 0 || c.lookahead_heuristic > 2) {
    return "Unknown lookahead heuristic.";
  }
  if (c.layering == QMG_LAYERING_ODD_GATES ||
      c.layering == QMG_LAYERING_QUBIT_TRIANGLE) {
    return "Layering strategy not suitable for heuristic mapper!";
  }
  const bool fid = c.heuristic == QMG_HEUR_FIDELITY_BEST_LOCATION;
  if (fid && !fidelityAvailable) {
    return "Fidelity aware heuristic chosen, but no or insufficient "
           "calibration data available for this architecture!";
  }
  if (fid && c.lookahead_heuristic != QMG_LOOKAHEAD_NONE) {
    return "Fidelity-aware heuristics may only be used with fidelity-aware "
           "lookahead heuristics (or no lookahead)!";
  }
  return nullptr;
}
}  // namespace qmg

struct qmg_circuit {
  qmg::LayeredCircuit lc;
  qmg::FlatCircuit flat;  // with the 1q gates of every layer (fidelity view)
  void reflatten() { flat.fromLayered(lc, true); }
};

#define QMG_API extern "C" __attribute__((visibility("default")))

#ifdef QMG_HOST_API_IMPL
QMG_API const char* qmg_last_error(void) { return qmg::lastError().c_str(); }

QMG_API void qmg_config_default(qmg_config* c) {
  // include/sc/configuration/Configuration.hpp:30-119
  std::memset(c, 0, sizeof(*c));
  c->heuristic = QMG_HEUR_GATE_COUNT_MAX_DISTANCE;
  c->lookahead_heuristic = QMG_LOOKAHEAD_GATE_COUNT_MAX_DISTANCE;
  c->nr_lookaheads = 15;
  c->first_lookahead_factor = 0.75;
  c->lookahead_factor = 0.5;
  c->early_termination = QMG_ET_NONE;
  c->early_termination_limit = 0;
  c->automatic_layer_splits = 1;
  c->automatic_layer_splits_node_limit = 5000;
  c->layering = QMG_LAYERING_INDIVIDUAL_GATES;
  c->initial_layout = QMG_LAYOUT_DYNAMIC;
  c->iterative_bidirectional_routing_passes = 0;
  c->swap_on_first_layer = 0;
  c->add_barriers_between_layers = 0;
  c->add_measurements_to_mapped_circuit = 1;
  c->pre_mapping_optimizations = 1;
  c->post_mapping_optimizations = 1;
}

QMG_API int qmg_circuit_create(uint32_t nArchQubits, size_t nGates,
                               const int32_t* control, const int32_t* target,
                               int32_t layering, qmg_circuit** out) {
  try {
    auto c = std::make_unique<qmg_circuit>();
    c->lc.build(nArchQubits, nGates, control, target, layering);
    c->reflatten();
    *out = c.release();
    return QMG_OK;
  } catch (const std::exception& e) {
    return qmg::fail(QMG_BAD_ARG, e.what());
  }
}
QMG_API int qmg_circuit_layers(const qmg_circuit* c, qmg_layers* out) {
  static_assert(sizeof(qmg_gate2q) == sizeof(qmg::Gate2), "gate2q");
  static_assert(sizeof(qmg_gate1q) == sizeof(qmg::Gate1), "gate1q");
  out->n_layers = c->flat.nLayers;
  out->n_logical = c->flat.nLogical;
  out->gate2q_offsets = c->flat.g2Off.data();
  out->gates2q = reinterpret_cast<const qmg_gate2q*>(c->flat.g2.data());
  out->gate1q_offsets = c->flat.g1Off.data();
  out->gates1q = reinterpret_cast<const qmg_gate1q*>(c->flat.g1.data());
  return QMG_OK;
}
QMG_API int qmg_circuit_layer_gates(const qmg_circuit* c, uint32_t* offsets,
                                    uint32_t* gateIndices) {
  uint32_t k = 0;
  uint32_t l = 0;
  for (const auto& L : c->lc.layers) {
    offsets[l++] = k;
    for (const auto g : L.gates) {
      gateIndices[k++] = g;
    }
  }
  offsets[l] = k;
  return QMG_OK;
}
QMG_API int qmg_circuit_layer_splittable(const qmg_circuit* c, uint32_t layer) {
  if (layer >= c->lc.layers.size()) {
    return 0;
  }
  return c->lc.splittable(layer) ? 1 : 0;
}
QMG_API int qmg_circuit_split_layer(qmg_circuit* c, uint32_t layer) {
  if (layer >= c->lc.layers.size()) {
    return qmg::fail(QMG_BAD_ARG, "layer out of range");
  }
  c->lc.split(layer);
  c->reflatten();
  return QMG_OK;
}
QMG_API void qmg_circuit_destroy(qmg_circuit* c) { delete c; }
#endif
