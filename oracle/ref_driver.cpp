// ref_driver.cpp - C-ABI driver around the UNMODIFIED reference sources
// (/root/reference/src/sc/{utils,Architecture,Configuration,Mapper,DataLogger}.cpp
// and src/sc/heuristic/HeuristicMapper.cpp), compiled where they lie against
// the stand-in mqt-core headers in oracle/refshim/.  Output: oracle/_ref/
// libqmapref.so (git-ignored).
//
// TEST INFRASTRUCTURE ONLY: used by tests/, __graft_entry__.smoke() and
// bench.py's cpu_baseline / --impl reference legs to (a) validate the CPU
// restatement in oracle/qmap_oracle.cpp and (b) time the reference on the
// host.  Nothing in the product path links or loads this library.
#include "qasm3/Importer.hpp"
#include "sc/Architecture.hpp"
#include "sc/heuristic/HeuristicMapper.hpp"
#include "sc/utils.hpp"

#include "../include/qmap_b200.h"

// -DQMAP_DROPIN: the same driver around the GPU drop-in (integration/
// GpuHeuristicMapper.hpp) instead of the reference's HeuristicMapper; built
// into oracle/_ref/libqmapdropin.so, which links libqmap_b200.so.  Used by
// tests/test_gpu_dropin.py only.
#ifdef QMAP_DROPIN
#include "../integration/GpuHeuristicMapper.hpp"
using BaseMapper = GpuHeuristicMapper;
// -DQMAP_DROPIN_EMU: the drop-in on the host debug build of the kernel source
// (tests/emu/dropin_shim.cpp) instead of libqmap_b200.so, for containers
// without a GPU; the shim takes the cost tables from the mapper's Architecture
#ifdef QMAP_DROPIN_EMU
extern "C" void qemu_shim_reference_architecture(const void* architecture);
#endif
#else
using BaseMapper = HeuristicMapper;
#endif

#include <chrono>
#include <cstring>
#include <memory>
#include <string>
#include <vector>

namespace {
thread_local std::string gError;

struct CallRecord {
  qmg_layer_result r{};
  std::uint32_t layer = 0;
  std::int32_t reverse = 0;
  std::int32_t finalPass = 0;
  std::int32_t split = 0; // this call recursed after a layer split
};

class TracingMapper : public BaseMapper {
public:
  using BaseMapper::BaseMapper;
  std::vector<CallRecord> calls;
  std::vector<std::int16_t> qubitsBefore; // calls.size() * n
  std::vector<std::int16_t> qubitsAfter;
  std::vector<std::int16_t> locationsAfter;
  std::vector<std::uint16_t> swaps; // pairs
  int depth = 0;
  bool inFinal = false;

  Node aStarMap(std::size_t layer, bool reverse) override {
    const auto before = results.layerHeuristicBenchmark.size();
    const auto qb = qubits;
    const bool fin = results.config.debug || inFinalPass();
    ++depth;
    const auto nestedBefore = calls.size();
    Node res = BaseMapper::aStarMap(layer, reverse);
    --depth;
    CallRecord c;
    c.layer = static_cast<std::uint32_t>(layer);
    c.reverse = reverse ? 1 : 0;
    c.finalPass = fin ? 1 : 0;
    c.split = calls.size() > nestedBefore ? 1 : 0;
    c.r.cost_fixed = res.costFixed;
    c.r.cost_fixed_reversals = res.costFixedReversals;
    c.r.cost_heur = res.costHeur;
    c.r.lookahead_penalty = res.lookaheadPenalty;
    c.r.solution_depth = static_cast<std::uint32_t>(res.depth);
    c.r.result_node_id = static_cast<std::uint32_t>(res.id);
    c.r.swap_offset = static_cast<std::uint32_t>(swaps.size() / 2);
    c.r.n_swaps = static_cast<std::uint32_t>(res.swaps.size());
    for (const auto& s : res.swaps) {
      swaps.push_back(s.first);
      swaps.push_back(s.second);
    }
    if (results.layerHeuristicBenchmark.size() > before && c.split == 0) {
      const auto& b = results.layerHeuristicBenchmark.back();
      c.r.expanded_nodes = b.expandedNodes;
      c.r.generated_nodes = b.generatedNodes;
      c.r.expanded_nodes_after_first_solution =
          b.expandedNodesAfterFirstSolution;
      c.r.expanded_nodes_after_optimal_solution =
          b.expandedNodesAfterOptimalSolution;
      c.r.solution_nodes = b.solutionNodes;
      c.r.solution_nodes_after_optimal_solution =
          b.solutionNodesAfterOptimalSolution;
      c.r.early_termination = b.earlyTermination ? 1 : 0;
    }
    std::uint64_t h = 1469598103934665603ULL;
    for (const auto q : res.qubits) {
      const auto v = static_cast<std::uint16_t>(q);
      h = (h ^ (v & 0xffU)) * 1099511628211ULL;
      h = (h ^ (v >> 8U)) * 1099511628211ULL;
    }
    c.r.layout_hash = h;
    calls.push_back(c);
    qubitsBefore.insert(qubitsBefore.end(), qb.begin(), qb.end());
    qubitsAfter.insert(qubitsAfter.end(), res.qubits.begin(),
                       res.qubits.end());
    locationsAfter.insert(locationsAfter.end(), res.locations.begin(),
                          res.locations.end());
    return res;
  }
  [[nodiscard]] bool inFinalPass() const { return inFinal; }

  void run(const Configuration& cfg) { map(cfg); }
  const std::vector<std::int16_t>& finalQubits() const { return qubits; }
  const std::vector<std::int16_t>& finalLocations() const { return locations; }
  std::size_t nLayers() const { return layers.size(); }
  qc::QuantumComputation& mapped() { return qcMapped; }
};

struct RefArch {
  Architecture arch;
};

struct RefResult {
  std::unique_ptr<TracingMapper> mapper;
  MappingResults results;
  double seconds = 0.;
  std::vector<std::int32_t> mappedOps; // (type, q0, q1) triples; q1=-1
  std::vector<std::int32_t> initialLayout;     // n: physical -> logical or -1
  std::vector<std::int32_t> outputPermutation; // n
  std::string json;
};

Configuration toConfiguration(const qmg_config& c) {
  Configuration cfg;
  cfg.heuristic = static_cast<Heuristic>(c.heuristic);
  cfg.lookaheadHeuristic = static_cast<LookaheadHeuristic>(c.lookahead_heuristic);
  cfg.nrLookaheads = c.nr_lookaheads;
  cfg.firstLookaheadFactor = c.first_lookahead_factor;
  cfg.lookaheadFactor = c.lookahead_factor;
  cfg.earlyTermination = static_cast<EarlyTermination>(c.early_termination);
  cfg.earlyTerminationLimit = c.early_termination_limit;
  cfg.automaticLayerSplits = c.automatic_layer_splits != 0;
  cfg.automaticLayerSplitsNodeLimit = c.automatic_layer_splits_node_limit;
  cfg.layering = static_cast<Layering>(c.layering);
  cfg.initialLayout = static_cast<InitialLayout>(c.initial_layout);
  cfg.iterativeBidirectionalRoutingPasses =
      c.iterative_bidirectional_routing_passes;
  cfg.iterativeBidirectionalRouting =
      c.iterative_bidirectional_routing_passes > 0;
  cfg.swapOnFirstLayer = c.swap_on_first_layer != 0;
  cfg.addBarriersBetweenLayers = c.add_barriers_between_layers != 0;
  cfg.addMeasurementsToMappedCircuit =
      c.add_measurements_to_mapped_circuit != 0;
  cfg.preMappingOptimizations = c.pre_mapping_optimizations != 0;
  cfg.postMappingOptimizations = c.post_mapping_optimizations != 0;
  cfg.debug = c.debug != 0;
  cfg.verbose = c.verbose != 0;
  return cfg;
}

void flatten(const Matrix& m, double* out) {
  const auto n = m.size();
  for (std::size_t i = 0; i < n; ++i) {
    std::memcpy(out + i * n, m[i].data(), n * sizeof(double));
  }
}
Matrix unflatten(const double* in, std::size_t n) {
  Matrix m(n, std::vector<double>(n));
  for (std::size_t i = 0; i < n; ++i) {
    std::memcpy(m[i].data(), in + i * n, n * sizeof(double));
  }
  return m;
}
CouplingMap toCm(const std::uint16_t* edges, std::size_t nE) {
  CouplingMap cm;
  for (std::size_t i = 0; i < nE; ++i) {
    cm.emplace(edges[2 * i], edges[2 * i + 1]);
  }
  return cm;
}

RefResult* runMap(RefArch* a, qc::QuantumComputation&& circ,
                  const qmg_config* c, const char* logPath) {
  auto res = std::make_unique<RefResult>();
  res->mapper = std::make_unique<TracingMapper>(circ, a->arch);
  Configuration cfg = toConfiguration(*c);
  if (logPath != nullptr && logPath[0] != '\0') {
    cfg.dataLoggingPath = logPath;
  }
#ifdef QMAP_DROPIN_EMU
  qemu_shim_reference_architecture(&a->arch);
#endif
  const auto t0 = std::chrono::steady_clock::now();
  res->mapper->run(cfg);
  const auto t1 = std::chrono::steady_clock::now();
  res->seconds = std::chrono::duration<double>(t1 - t0).count();
  res->results = res->mapper->getResults();
  res->json = res->results.json().dump(1);
  const auto n = a->arch.getNqubits();
  auto& mc = res->mapper->mapped();
  res->initialLayout.assign(n, -1);
  res->outputPermutation.assign(n, -1);
  for (const auto& [p, l] : mc.initialLayout) {
    if (p < n) {
      res->initialLayout[p] = static_cast<std::int32_t>(l);
    }
  }
  for (const auto& [p, l] : mc.outputPermutation) {
    if (p < n) {
      res->outputPermutation[p] = static_cast<std::int32_t>(l);
    }
  }
  for (const auto& op : mc) {
    if (op->getType() == qc::Measure) {
      continue;
    }
    std::int32_t q0 = -1;
    std::int32_t q1 = -1;
    if (op->getType() == qc::SWAP) {
      q0 = static_cast<std::int32_t>(op->getTargets().at(0));
      q1 = static_cast<std::int32_t>(op->getTargets().at(1));
    } else if (!op->getControls().empty()) {
      q0 = static_cast<std::int32_t>(op->getControls().begin()->qubit);
      q1 = static_cast<std::int32_t>(op->getTargets().at(0));
    } else {
      q0 = static_cast<std::int32_t>(op->getTargets().at(0));
    }
    res->mappedOps.push_back(static_cast<std::int32_t>(op->getType()));
    res->mappedOps.push_back(q0);
    res->mappedOps.push_back(q1);
  }
  return res.release();
}
} // namespace

#define QREF_TRY try {
#define QREF_CATCH(rv)                                                         \
  }                                                                            \
  catch (const std::exception& e) {                                            \
    gError = e.what();                                                         \
    return rv;                                                                 \
  }

extern "C" {
const char* qref_last_error() { return gError.c_str(); }

// ---- Dijkstra:: table builders (src/sc/utils.cpp:26,91,147) ----
int qref_build_table(std::uint16_t n, const std::uint16_t* edges,
                     std::size_t nE, const double* w, double* out) {
  QREF_TRY
  Matrix d;
  Dijkstra::buildTable(toCm(edges, nE), d, unflatten(w, n));
  flatten(d, out);
  return 0;
  QREF_CATCH(1)
}
int qref_build_single_edge_skip_table(std::uint16_t n, const double* dist,
                                      const std::uint16_t* edges,
                                      std::size_t nE, double rev, double* out) {
  QREF_TRY
  Matrix d;
  Dijkstra::buildSingleEdgeSkipTable(unflatten(dist, n), toCm(edges, nE), rev,
                                     d);
  flatten(d, out);
  return 0;
  QREF_CATCH(1)
}
int qref_build_edge_skip_table(std::uint16_t n, const std::uint16_t* edges,
                               std::size_t nE, const double* w, double* out,
                               std::size_t maxTables, std::size_t* nTables) {
  QREF_TRY
  std::vector<Matrix> t;
  Dijkstra::buildEdgeSkipTable(toCm(edges, nE), t, unflatten(w, n));
  *nTables = t.size();
  for (std::size_t k = 0; k < t.size() && k < maxTables; ++k) {
    flatten(t[k], out + k * static_cast<std::size_t>(n) * n);
  }
  return 0;
  QREF_CATCH(1)
}

// ---- Architecture ----
void* qref_arch_create(std::uint16_t n, const std::uint16_t* edges,
                       std::size_t nE) {
  QREF_TRY
  auto a = std::make_unique<RefArch>();
  a->arch.loadCouplingMap(n, toCm(edges, nE));
  return a.release();
  QREF_CATCH(nullptr)
}
void* qref_arch_load(const char* archFile, const char* calibrationFile) {
  QREF_TRY
  auto a = std::make_unique<RefArch>();
  if (archFile != nullptr && archFile[0] != '\0') {
    a->arch.loadCouplingMap(std::string(archFile));
  }
  if (calibrationFile != nullptr && calibrationFile[0] != '\0') {
    a->arch.loadProperties(std::string(calibrationFile));
  }
  return a.release();
  QREF_CATCH(nullptr)
}
// per-qubit single-qubit error rate and per-directed-edge CX error rate
// (edge order = ascending coupling-map order); NaN = not available
int qref_arch_set_error_rates(void* h, const double* q1err,
                              const double* cxerr) {
  QREF_TRY
  auto* a = static_cast<RefArch*>(h);
  Architecture::Properties props;
  const auto n = a->arch.getNqubits();
  props.setNqubits(n);
  for (std::uint16_t q = 0; q < n; ++q) {
    if (q1err[q] == q1err[q]) {
      props.setSingleQubitErrorRate(q, "x", q1err[q]);
    }
  }
  std::size_t i = 0;
  for (const auto& [e1, e2] : a->arch.getCouplingMap()) {
    if (cxerr[i] == cxerr[i]) {
      props.setTwoQubitErrorRate(e1, e2, cxerr[i]);
    }
    ++i;
  }
  a->arch.loadProperties(props);
  return 0;
  QREF_CATCH(1)
}
void qref_arch_destroy(void* h) { delete static_cast<RefArch*>(h); }
// Architecture::minimumNumberOfSwaps(permutation, limit) (Architecture.cpp:296-385)
long long qref_arch_minimum_number_of_swaps(void* h, const std::uint16_t* perm,
                                            std::size_t n, long long limit) {
  QREF_TRY
  auto* a = static_cast<RefArch*>(h);
  std::vector<std::uint16_t> p(perm, perm + n);
  return static_cast<long long>(a->arch.minimumNumberOfSwaps(p, limit));
  QREF_CATCH(-2)
}
// Architecture::getCouplingLimit() / (qubitChoice) (Architecture.cpp:480-569)
long long qref_arch_coupling_limit(void* h, const std::uint16_t* choice,
                                   std::size_t nChoice) {
  QREF_TRY
  auto* a = static_cast<RefArch*>(h);
  if (choice == nullptr) {
    return static_cast<long long>(a->arch.getCouplingLimit());
  }
  const std::set<std::uint16_t> c(choice, choice + nChoice);
  return static_cast<long long>(a->arch.getCouplingLimit(c));
  QREF_CATCH(-2)
}
int qref_arch_info(void* h, qmg_arch_info* info) {
  auto* a = static_cast<RefArch*>(h);
  std::memset(info, 0, sizeof(*info));
  info->n = a->arch.getNqubits();
  info->n_edges = static_cast<std::uint32_t>(a->arch.getCouplingMap().size());
  info->bidirectional = a->arch.bidirectional() ? 1 : 0;
  info->unidirectional = a->arch.unidirectional() ? 1 : 0;
  info->fidelity_available = a->arch.isFidelityAvailable() ? 1 : 0;
  info->n_fidelity_tables =
      a->arch.isFidelityAvailable()
          ? static_cast<std::uint32_t>(
                a->arch.getFidelityDistanceTables().size())
          : 0U;
  info->device = -1;
  return 0;
}
int qref_arch_edges(void* h, std::uint16_t* out) {
  auto* a = static_cast<RefArch*>(h);
  std::size_t i = 0;
  for (const auto& [e1, e2] : a->arch.getCouplingMap()) {
    out[2 * i] = e1;
    out[2 * i + 1] = e2;
    ++i;
  }
  return 0;
}
int qref_arch_distance_table(void* h, int includeReversal, double* out) {
  QREF_TRY
  flatten(static_cast<RefArch*>(h)->arch.getDistanceTable(includeReversal != 0),
          out);
  return 0;
  QREF_CATCH(1)
}
// which: 0 single-qubit costs (n), 1 two-qubit costs (n*n), 2 swap costs
// (n*n), 3 fidelity distance tables (K*n*n)
int qref_arch_fidelity(void* h, int which, double* out) {
  QREF_TRY
  auto& arch = static_cast<RefArch*>(h)->arch;
  const std::size_t n = arch.getNqubits();
  switch (which) {
  case 0:
    std::memcpy(out, arch.getSingleQubitFidelityCosts().data(),
                n * sizeof(double));
    break;
  case 1:
    flatten(arch.getTwoQubitFidelityCosts(), out);
    break;
  case 2:
    flatten(arch.getSwapFidelityCosts(), out);
    break;
  case 3: {
    const auto& t = arch.getFidelityDistanceTables();
    for (std::size_t k = 0; k < t.size(); ++k) {
      flatten(t[k], out + k * n * n);
    }
    break;
  }
  default:
    gError = "bad selector";
    return 1;
  }
  return 0;
  QREF_CATCH(1)
}

// ---- HeuristicMapper::map ----
void* qref_map_qasm(void* arch, const char* qasmFile, const qmg_config* cfg,
                    const char* dataLoggingPath) {
  QREF_TRY
  auto circ = qasm3::Importer::importf(qasmFile);
  return runMap(static_cast<RefArch*>(arch), std::move(circ), cfg,
                dataLoggingPath);
  QREF_CATCH(nullptr)
}
// gate list: control[i] = -1 for a single-qubit gate (emitted as H)
void* qref_map_gates(void* arch, std::uint32_t nQubits, std::size_t nGates,
                     const std::int32_t* control, const std::int32_t* target,
                     const qmg_config* cfg, const char* dataLoggingPath) {
  QREF_TRY
  qc::QuantumComputation circ(nQubits, nQubits);
  circ.setName("gates");
  for (std::size_t i = 0; i < nGates; ++i) {
    if (control[i] < 0) {
      circ.h(static_cast<qc::Qubit>(target[i]));
    } else {
      circ.cx(qc::Control{static_cast<qc::Qubit>(control[i])},
              static_cast<qc::Qubit>(target[i]));
    }
  }
  return runMap(static_cast<RefArch*>(arch), std::move(circ), cfg,
                dataLoggingPath);
  QREF_CATCH(nullptr)
}
void qref_result_destroy(void* r) { delete static_cast<RefResult*>(r); }

typedef struct {
  double seconds;       // wall time of map()
  double results_time;  // MappingResults::time
  double seconds_per_node;
  std::uint64_t expanded_nodes;
  std::uint64_t generated_nodes;
  std::uint64_t swaps;
  std::uint64_t direction_reverse;
  std::uint64_t output_gates;
  std::uint64_t output_cnots;
  std::uint64_t output_single_qubit_gates;
  std::uint64_t layers;
  std::uint64_t n_calls;      // aStarMap calls recorded (all passes)
  std::uint64_t n_swap_pairs; // total entries in the swaps array / 2
  std::uint64_t n_mapped_ops;
  double total_log_fidelity;
} qref_summary;

int qref_result_summary(void* h, qref_summary* s) {
  auto* r = static_cast<RefResult*>(h);
  s->seconds = r->seconds;
  s->results_time = r->results.time;
  s->seconds_per_node = r->results.heuristicBenchmark.secondsPerNode;
  s->expanded_nodes = r->results.heuristicBenchmark.expandedNodes;
  s->generated_nodes = r->results.heuristicBenchmark.generatedNodes;
  s->swaps = r->results.output.swaps;
  s->direction_reverse = r->results.output.directionReverse;
  s->output_gates = r->results.output.gates;
  s->output_cnots = r->results.output.cnots;
  s->output_single_qubit_gates = r->results.output.singleQubitGates;
  s->layers = r->mapper->nLayers();
  s->n_calls = r->mapper->calls.size();
  s->n_swap_pairs = r->mapper->swaps.size() / 2;
  s->n_mapped_ops = r->mappedOps.size() / 3;
  s->total_log_fidelity = r->results.output.totalLogFidelity;
  return 0;
}
// per aStarMap call, in call-completion order
int qref_result_calls(void* h, qmg_layer_result* out, std::uint32_t* layer,
                      std::int32_t* flags /* bit0 reverse, bit1 split */) {
  auto* r = static_cast<RefResult*>(h);
  for (std::size_t i = 0; i < r->mapper->calls.size(); ++i) {
    const auto& c = r->mapper->calls[i];
    out[i] = c.r;
    layer[i] = c.layer;
    flags[i] = (c.reverse != 0 ? 1 : 0) | (c.split != 0 ? 2 : 0);
  }
  return 0;
}
int qref_result_layouts(void* h, std::int16_t* qubitsBefore,
                        std::int16_t* qubitsAfter, std::int16_t* locationsAfter,
                        std::uint16_t* swaps) {
  auto* r = static_cast<RefResult*>(h);
  auto cp = [](auto* dst, const auto& v) {
    if (dst != nullptr && !v.empty()) {
      std::memcpy(dst, v.data(), v.size() * sizeof(v[0]));
    }
  };
  cp(qubitsBefore, r->mapper->qubitsBefore);
  cp(qubitsAfter, r->mapper->qubitsAfter);
  cp(locationsAfter, r->mapper->locationsAfter);
  cp(swaps, r->mapper->swaps);
  return 0;
}
int qref_result_circuit(void* h, std::int32_t* ops, std::int32_t* initialLayout,
                        std::int32_t* outputPermutation,
                        std::int16_t* finalQubits) {
  auto* r = static_cast<RefResult*>(h);
  if (ops != nullptr) {
    std::memcpy(ops, r->mappedOps.data(),
                r->mappedOps.size() * sizeof(std::int32_t));
  }
  std::memcpy(initialLayout, r->initialLayout.data(),
              r->initialLayout.size() * sizeof(std::int32_t));
  std::memcpy(outputPermutation, r->outputPermutation.data(),
              r->outputPermutation.size() * sizeof(std::int32_t));
  if (finalQubits != nullptr) {
    const auto& q = r->mapper->finalQubits();
    std::memcpy(finalQubits, q.data(), q.size() * sizeof(std::int16_t));
  }
  return 0;
}
const char* qref_result_json(void* h) {
  return static_cast<RefResult*>(h)->json.c_str();
}
}
