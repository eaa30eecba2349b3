// host_api.hpp - host-only parts of the C ABI (include/qmap_b200.h): error
// string, configuration defaults, layered circuits.  Included by api.cu (the
// product library) and by tests/emu/emu.cpp (the host debug build).
#pragma once
#include <cstring>
#include <exception>
#include <fstream>
#include <iterator>
#include <atomic>
#include <memory>
#include <mutex>
#include <string>
#include <vector>

#include "../../include/qmap_b200.h"
#include "flat_layers.hpp"
#include "flatten.hpp"
#include "layers.hpp"
#include "qasm_reader.hpp"

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

// createLayers of many circuits at once, CSR only (flat_layers.hpp)
struct qmg_circuit_batch {
  std::vector<qmg::FlatLayers> circ;
};

struct qmg_qasm {
  qmg::QasmCircuit c;
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

QMG_API int qmg_circuit_batch_create(uint32_t nArchQubits, size_t nCircuits,
                                     const uint64_t* gateOffsets,
                                     const int32_t* control,
                                     const int32_t* target, int32_t layering,
                                     qmg_circuit_batch** out) {
  if (out == nullptr || (nCircuits > 0 && gateOffsets == nullptr)) {
    return qmg::fail(QMG_BAD_ARG, "null argument");
  }
  try {
    auto b = std::make_unique<qmg_circuit_batch>();
    b->circ.resize(nCircuits);
    std::atomic<int> bad{0};
    std::string msg;
    std::mutex mu;
    qmg::parallelFor(nCircuits, 16, [&](size_t i) {
      try {
        const uint64_t lo = gateOffsets[i], hi = gateOffsets[i + 1];
        b->circ[i].build(nArchQubits, static_cast<size_t>(hi - lo),
                         control + lo, target + lo, layering);
      } catch (const std::exception& e) {
        std::lock_guard<std::mutex> lk(mu);
        if (bad.exchange(1) == 0) {
          msg = e.what();
        }
      }
    });
    if (bad.load() != 0) {
      return qmg::fail(QMG_BAD_ARG, msg);
    }
    *out = b.release();
    return QMG_OK;
  } catch (const std::exception& e) {
    return qmg::fail(QMG_BAD_ARG, e.what());
  }
}
namespace qmg {
inline void layersOf(const FlatLayers& f, qmg_layers* out) {
  out->n_layers = f.nLayers;
  out->n_logical = f.nLogical;
  out->gate2q_offsets = f.g2Off.data();
  out->gates2q = reinterpret_cast<const qmg_gate2q*>(f.g2.data());
  out->gate1q_offsets = f.g1Off.data();
  out->gates1q = reinterpret_cast<const qmg_gate1q*>(f.g1.data());
}
}  // namespace qmg
QMG_API int qmg_circuit_batch_layers(const qmg_circuit_batch* b, size_t i,
                                     qmg_layers* out) {
  if (b == nullptr || out == nullptr || i >= b->circ.size()) {
    return qmg::fail(QMG_BAD_ARG, "circuit index out of range");
  }
  qmg::layersOf(b->circ[i], out);
  return QMG_OK;
}
QMG_API int qmg_circuit_batch_fill_instances(const qmg_circuit_batch* b,
                                             qmg_instance* instances,
                                             size_t nInstances) {
  if (b == nullptr || instances == nullptr || nInstances > b->circ.size()) {
    return qmg::fail(QMG_BAD_ARG, "more instances than circuits");
  }
  for (size_t i = 0; i < nInstances; ++i) {
    qmg::layersOf(b->circ[i], &instances[i].layers);
  }
  return QMG_OK;
}
QMG_API int qmg_circuit_batch_layer_gates(const qmg_circuit_batch* b, size_t i,
                                          uint32_t* offsets,
                                          uint32_t* gateIndices) {
  if (b == nullptr || i >= b->circ.size()) {
    return qmg::fail(QMG_BAD_ARG, "circuit index out of range");
  }
  const auto& f = b->circ[i];
  std::memcpy(offsets, f.gateOff.data(), f.gateOff.size() * 4);
  std::memcpy(gateIndices, f.gateIdx.data(), f.gateIdx.size() * 4);
  return QMG_OK;
}
QMG_API void qmg_circuit_batch_destroy(qmg_circuit_batch* b) { delete b; }

// Replay of routed instances (the output must be the input circuit
// under the SWAP-induced permutation): starting from start_locations, every
// layer's SWAPs must sit on coupling edges, after them every 2q gate of the
// layer must act on adjacent physical qubits, qubits placed on the way follow
// the placement list, and the layout reached must be final_locations.
// verdict[i]: 1 = valid, 0 = invalid, 2 = not checkable (layers were split by
// the library, outputs missing, or the instance did not finish).
QMG_API int qmg_circuit_batch_verify(const qmg_circuit_batch* b, uint16_t n,
                                     const uint16_t* edges, size_t nEdges,
                                     const uint64_t* gateOffsets,
                                     const int32_t* control,
                                     const int32_t* target,
                                     const qmg_route_result* results,
                                     size_t nInstances,
                                     int32_t fidelityAware, uint8_t* verdict) {
  if (b == nullptr || results == nullptr || verdict == nullptr ||
      nInstances > b->circ.size() || (nEdges > 0 && edges == nullptr)) {
    return qmg::fail(QMG_BAD_ARG, "bad argument");
  }
  std::vector<uint8_t> adj(static_cast<size_t>(n) * n, 0);
  for (size_t e = 0; e < nEdges; ++e) {
    const uint16_t x = edges[2 * e], y = edges[2 * e + 1];
    if (x >= n || y >= n) {
      return qmg::fail(QMG_BAD_ARG, "coupling-map entry out of range");
    }
    adj[static_cast<size_t>(x) * n + y] = 1;
    adj[static_cast<size_t>(y) * n + x] = 1;
  }
  qmg::parallelFor(nInstances, 64, [&](size_t i) {
    const qmg::FlatLayers& f = b->circ[i];
    const qmg_route_result& r = results[i];
    verdict[i] = 2;
    if (r.status != QMG_OK || r.n_splits != 0 || r.n_layers_done != f.nLayers ||
        r.swaps == nullptr || r.start_locations == nullptr ||
        r.final_locations == nullptr ||
        (r.layer_n_swaps == nullptr && r.layers == nullptr) ||
        r.n_swaps_total > r.swap_capacity) {
      return;
    }
    std::vector<int32_t> loc(r.start_locations, r.start_locations + n);
    std::vector<int32_t> at(n, -1);
    bool ok = true;
    for (int q = 0; q < n; ++q) {
      if (loc[q] >= 0) {
        ok = ok && loc[q] < n && at[loc[q]] == -1;
        if (ok) {
          at[loc[q]] = q;
        }
      }
    }
    const int32_t* ctrl = control + gateOffsets[i];
    const int32_t* tgt = target + gateOffsets[i];
    uint32_t sw = 0, pl = 0;
    auto place = [&](int32_t q) {
      // mapUnmappedGates placed q before this layer's search
      while (ok && loc[q] < 0) {
        if (r.placements == nullptr || pl >= r.n_placements_total ||
            pl >= r.placement_capacity) {
          ok = false;
          break;
        }
        const qmg_placement p = r.placements[pl++];
        ok = p.logical < n && p.physical < n && loc[p.logical] < 0 &&
             at[p.physical] == -1;
        if (ok) {
          loc[p.logical] = p.physical;
          at[p.physical] = p.logical;
        }
      }
    };
    for (uint32_t l = 0; l < f.nLayers && ok; ++l) {
      // mapUnmappedGates :211-285: the fidelity-aware mapper places the
      // qubits of the layer's 1q gates first (ascending), then the 2q gates'
      if (fidelityAware != 0) {
        for (uint32_t k = f.g1Off[l]; k < f.g1Off[l + 1] && ok; ++k) {
          place(f.g1[k].q);
        }
      }
      for (uint32_t k = f.g2Off[l]; k < f.g2Off[l + 1] && ok; ++k) {
        const qmg::Gate2 e = f.g2[k];
        // (which of the two is placed first depends on the other's state;
        // either way both are placed before the layer's SWAPs)
        if (loc[e.q1] < 0 || loc[e.q2] < 0) {
          place(loc[e.q1] < 0 && loc[e.q2] >= 0 ? e.q1 : (loc[e.q2] < 0 && loc[e.q1] >= 0 ? e.q2 : e.q1));
          place(e.q2);
          place(e.q1);
        }
      }
      const uint32_t ns = r.layer_n_swaps != nullptr ? r.layer_n_swaps[l]
                                                     : r.layers[l].n_swaps;
      for (uint32_t s = 0; s < ns && ok; ++s, ++sw) {
        if (sw >= r.n_swaps_total) {
          ok = false;
          break;
        }
        const uint16_t x = r.swaps[2 * sw], y = r.swaps[2 * sw + 1];
        ok = x < n && y < n && adj[static_cast<size_t>(x) * n + y] != 0;
        if (ok) {
          const int32_t qx = at[x], qy = at[y];
          at[x] = qy;
          at[y] = qx;
          if (qx >= 0) {
            loc[qx] = y;
          }
          if (qy >= 0) {
            loc[qy] = x;
          }
        }
      }
      for (uint32_t k = f.gateOff[l]; k < f.gateOff[l + 1] && ok; ++k) {
        const uint32_t gi = f.gateIdx[k];
        if (ctrl[gi] >= 0) {
          const int32_t pc = loc[ctrl[gi]], pt = loc[tgt[gi]];
          ok = pc >= 0 && pt >= 0 && adj[static_cast<size_t>(pc) * n + pt] != 0;
        }
      }
    }
    ok = ok && sw == r.n_swaps_total;
    for (int q = 0; q < n && ok; ++q) {
      // (1q-only qubits of the fidelity-aware mapper are placed too; a qubit
      // the replay never needed may still be unplaced here)
      ok = loc[q] < 0 || loc[q] == r.final_locations[q];
    }
    verdict[i] = ok ? 1 : 0;
  });
  return QMG_OK;
}
// ---- OpenQASM 2 reader (qasm_reader.hpp)
QMG_API int qmg_qasm_parse(const char* text, size_t length, qmg_qasm** out) {
  if (text == nullptr || out == nullptr) {
    return qmg::fail(QMG_BAD_ARG, "null argument");
  }
  try {
    auto q = std::make_unique<qmg_qasm>();
    q->c.parse(text, length);
    *out = q.release();
    return QMG_OK;
  } catch (const std::exception& e) {
    return qmg::fail(QMG_BAD_ARG, e.what());
  }
}
QMG_API int qmg_qasm_read_file(const char* path, qmg_qasm** out) {
  if (path == nullptr || out == nullptr) {
    return qmg::fail(QMG_BAD_ARG, "null argument");
  }
  std::ifstream f(path, std::ios::binary);
  if (!f) {
    return qmg::fail(QMG_BAD_ARG, std::string("cannot open ") + path);
  }
  std::string text((std::istreambuf_iterator<char>(f)), std::istreambuf_iterator<char>());
  return qmg_qasm_parse(text.data(), text.size(), out);
}
QMG_API int qmg_qasm_get_info(const qmg_qasm* q, qmg_qasm_info* info) {
  if (q == nullptr || info == nullptr) {
    return qmg::fail(QMG_BAD_ARG, "null argument");
  }
  info->n_qubits = q->c.nQubits;
  info->n_clbits = q->c.nClbits;
  info->n_gates = q->c.kind.size();
  info->n_measurements = q->c.measQubit.size();
  return QMG_OK;
}
QMG_API int qmg_qasm_gates(const qmg_qasm* q, int32_t* control, int32_t* target,
                           uint8_t* kind, double* params, uint8_t* n_params) {
  if (q == nullptr) {
    return qmg::fail(QMG_BAD_ARG, "null argument");
  }
  const size_t g = q->c.kind.size();
  if (g == 0) {
    return QMG_OK;
  }
  if (control != nullptr) {
    std::memcpy(control, q->c.control.data(), g * sizeof(int32_t));
  }
  if (target != nullptr) {
    std::memcpy(target, q->c.target.data(), g * sizeof(int32_t));
  }
  if (kind != nullptr) {
    std::memcpy(kind, q->c.kind.data(), g);
  }
  if (params != nullptr) {
    std::memcpy(params, q->c.params.data(), 3 * g * sizeof(double));
  }
  if (n_params != nullptr) {
    std::memcpy(n_params, q->c.nParams.data(), g);
  }
  return QMG_OK;
}
QMG_API int qmg_qasm_measurements(const qmg_qasm* q, uint32_t* qubit, uint32_t* clbit) {
  if (q == nullptr) {
    return qmg::fail(QMG_BAD_ARG, "null argument");
  }
  const size_t m = q->c.measQubit.size();
  if (m != 0 && qubit != nullptr) {
    std::memcpy(qubit, q->c.measQubit.data(), m * sizeof(uint32_t));
  }
  if (m != 0 && clbit != nullptr) {
    std::memcpy(clbit, q->c.measClbit.data(), m * sizeof(uint32_t));
  }
  return QMG_OK;
}
QMG_API int qmg_qasm_mapper_view(const qmg_qasm* q, uint32_t nPhysical,
                                 int32_t* control, int32_t* target,
                                 int16_t* identity_locations) {
  if (q == nullptr || ((control == nullptr || target == nullptr) && !q->c.kind.empty())) {
    return qmg::fail(QMG_BAD_ARG, "null argument");
  }
  try {
    q->c.mapperView(nPhysical, control, target, identity_locations);
    return QMG_OK;
  } catch (const std::exception& e) {
    return qmg::fail(QMG_BAD_ARG, e.what());
  }
}
QMG_API const char* qmg_qasm_gate_name(uint8_t kind) {
  size_t cnt = 0;
  const char* const* names = qmg::qasmGateNames(&cnt);
  return kind < cnt ? names[kind] : nullptr;
}
QMG_API void qmg_qasm_destroy(qmg_qasm* q) { delete q; }
#endif
