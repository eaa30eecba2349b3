/*
 * qmap_b200.h - C ABI of the B200-native heuristic router.
 *
 * One hot path of cda-tum/mqt-qmap is rebuilt behind this boundary: the
 * superconducting heuristic mapper's cost tables and its per-layer A* SWAP
 * search.  Every entry point names the reference interface it replaces
 * (paths relative to the reference tree).  Plain pointers and sizes only; the
 * caller owns every host buffer, the library owns device memory behind the
 * opaque handles.  No exception crosses this boundary: every function returns
 * a qmg_status and qmg_last_error() holds the message (thread-local).
 *
 * There is no CPU fallback: without a CUDA device every compute entry point
 * returns QMG_NO_DEVICE.
 */
#ifndef QMAP_B200_H
#define QMAP_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

typedef enum {
  QMG_OK = 0,
  /* HeuristicMapper.cpp:637-639 "No viable mapping found." */
  QMG_NO_VIABLE_MAPPING = 1,
  /* HeuristicMapper.cpp:563 automatic layer split requested by the search */
  QMG_SPLIT_LIMIT_HIT = 2,
  QMG_BAD_ARG = 3,
  QMG_CUDA_ERROR = 4,
  QMG_OOM = 5,
  /* the per-instance node pool is full (raise qmg_route_options.pool_nodes) */
  QMG_POOL_EXHAUSTED = 6,
  QMG_NO_DEVICE = 7,
  /* a caller-provided output buffer (swaps / placements) is too small */
  QMG_BUFFER_TOO_SMALL = 8
} qmg_status;

/* include/sc/configuration/Heuristic.hpp:17-36 (same order) */
typedef enum {
  QMG_HEUR_GATE_COUNT_MAX_DISTANCE = 0,
  QMG_HEUR_GATE_COUNT_SUM_DISTANCE = 1,
  QMG_HEUR_GATE_COUNT_SUM_DISTANCE_MINUS_SHARED_SWAPS = 2,
  QMG_HEUR_GATE_COUNT_MAX_DISTANCE_OR_SUM_DISTANCE_MINUS_SHARED_SWAPS = 3,
  QMG_HEUR_FIDELITY_BEST_LOCATION = 4
} qmg_heuristic;

/* include/sc/configuration/LookaheadHeuristic.hpp:17-25 */
typedef enum {
  QMG_LOOKAHEAD_NONE = 0,
  QMG_LOOKAHEAD_GATE_COUNT_MAX_DISTANCE = 1,
  QMG_LOOKAHEAD_GATE_COUNT_SUM_DISTANCE = 2
} qmg_lookahead;

/* include/sc/configuration/EarlyTermination.hpp:17-24 */
typedef enum {
  QMG_ET_NONE = 0,
  QMG_ET_EXPANDED_NODES = 1,
  QMG_ET_EXPANDED_NODES_AFTER_FIRST_SOLUTION = 2,
  QMG_ET_EXPANDED_NODES_AFTER_CURRENT_OPTIMAL_SOLUTION = 3,
  QMG_ET_SOLUTION_NODES = 4,
  QMG_ET_SOLUTION_NODES_AFTER_CURRENT_OPTIMAL_SOLUTION = 5
} qmg_early_termination;

/* include/sc/configuration/Layering.hpp:17-23 */
typedef enum {
  QMG_LAYERING_INDIVIDUAL_GATES = 0,
  QMG_LAYERING_DISJOINT_QUBITS = 1,
  QMG_LAYERING_ODD_GATES = 2,
  QMG_LAYERING_QUBIT_TRIANGLE = 3,
  QMG_LAYERING_DISJOINT_2Q_BLOCKS = 4
} qmg_layering;

/* include/sc/configuration/InitialLayout.hpp:20 */
typedef enum {
  QMG_LAYOUT_IDENTITY = 0,
  QMG_LAYOUT_STATIC = 1,
  QMG_LAYOUT_DYNAMIC = 2
} qmg_initial_layout;

/*
 * POD mirror of the fields of Configuration (include/sc/configuration/
 * Configuration.hpp:30-119) the heuristic path reads.  Defaults are the
 * reference's (qmg_config_default()).
 */
typedef struct {
  int32_t heuristic;           /* qmg_heuristic */
  int32_t lookahead_heuristic; /* qmg_lookahead */
  uint32_t nr_lookaheads;
  uint32_t _pad0;
  double first_lookahead_factor;
  double lookahead_factor;
  int32_t early_termination; /* qmg_early_termination */
  int32_t automatic_layer_splits;
  uint64_t early_termination_limit;
  uint64_t automatic_layer_splits_node_limit;
  /* host-side fields (layering / placement / emission) */
  int32_t layering;       /* qmg_layering */
  int32_t initial_layout; /* qmg_initial_layout */
  uint32_t iterative_bidirectional_routing_passes;
  int32_t swap_on_first_layer;
  int32_t add_barriers_between_layers;
  int32_t add_measurements_to_mapped_circuit;
  int32_t pre_mapping_optimizations;
  int32_t post_mapping_optimizations;
  int32_t debug;
  int32_t verbose;
} qmg_config;

void qmg_config_default(qmg_config* cfg);

/* ------------------------------------------------------------------------ */
/* Cost tables                                                              */
/* ------------------------------------------------------------------------ */

/*
 * Stand-alone table kernels.  Matrices are dense row-major n*n doubles, the
 * coupling map is nEdges directed (first, second) pairs in ascending
 * lexicographic order (the iteration order of the reference's
 * std::set<std::pair<uint16_t,uint16_t>>, include/sc/utils.hpp:27-29).
 */

/* Dijkstra::buildTable (src/sc/utils.cpp:26-54): all-pairs path costs over
 * the undirected view of the coupling map with directed weights
 * edgeWeights[from*n+to]; diagonal 0, unreachable -1.
 * exact_association != 0 replays the reference's Dijkstra::dijkstra
 * (:56-89) heap operation by heap operation and returns ITS table bit for
 * bit; for non-uniform weights that table is not always the table of
 * shortest paths (the reference rewrites records that sit in its heap).
 * 0 uses the blocked min-plus Floyd-Warshall kernel: true shortest paths,
 * equal to the reference's table for uniform weights. */
int qmg_build_table(uint16_t n, const uint16_t* edges, size_t nEdges,
                    const double* edgeWeights, int exact_association,
                    double* distanceTable, int device);

/* Dijkstra::buildSingleEdgeSkipTable (src/sc/utils.cpp:147-183) */
int qmg_build_single_edge_skip_table(uint16_t n, const double* distanceTable,
                                     const uint16_t* edges, size_t nEdges,
                                     double reversalCost,
                                     double* edgeSkipDistanceTable, int device);

/* Dijkstra::buildEdgeSkipTable (src/sc/utils.cpp:91-145).  tables receives up
 * to maxTables matrices of n*n doubles; *nTables is the number the build
 * produced (the reference stops at the first all-zero table and drops it);
 * if that is more than maxTables the first maxTables are stored and the call
 * returns QMG_BUFFER_TOO_SMALL. */
int qmg_build_edge_skip_table(uint16_t n, const uint16_t* edges, size_t nEdges,
                              const double* edgeWeights, int exact_association,
                              double* tables, size_t maxTables,
                              size_t* nTables, int device);

/*
 * qmg_build_table_static: qmg_build_table (exact_association = 0, the blocked
 * Floyd-Warshall kernel) for callers that cannot allocate.  It has NO
 * COUNTERPART IN THE REFERENCE and was added:
 * code under CUDA graph capture, tight serving loops and harnesses
 * that forbid cudaMalloc need an entry point that allocates nothing.
 *
 * dEdges (2 * nEdges uint16), dEdgeWeights (n * n doubles, row-major) and
 * dDistanceTable (n * n doubles, output) are DEVICE pointers on `device`; the
 * kernels are queued on `stream` (a cudaStream_t; NULL = the default stream)
 * and the call returns without waiting for them.  The call makes no
 * cudaMalloc / cudaFree and uses no allocating library routine: the padded
 * matrix the kernels work on is a static device array inside the library
 * (QMG_STATIC_TABLE_MAX_QUBITS^2 doubles = 2 MiB, held by every GPU that loads
 * the library until the process exits).  n above QMG_STATIC_TABLE_MAX_QUBITS
 * is rejected with QMG_BAD_ARG.  All calls share that array: a call rewrites
 * every cell it reads, so earlier contents never matter, but calls whose
 * kernels overlap on one GPU (different streams or threads) race on it -
 * serialise them.  The result equals qmg_build_table(..., 0, ...) bit for bit.
 */
#define QMG_STATIC_TABLE_MAX_QUBITS 512
int qmg_build_table_static(uint16_t n, const uint16_t* dEdges, size_t nEdges,
                           const double* dEdgeWeights, double* dDistanceTable,
                           int device, void* stream);

/*
 * Architecture handle: the device-resident form of class Architecture
 * (include/sc/Architecture.hpp:42).  qmg_arch_create replaces
 * Architecture::createDistanceTable (src/sc/Architecture.cpp:201-230);
 * qmg_arch_set_fidelity_costs replaces the table part of
 * Architecture::createFidelityTable (src/sc/Architecture.cpp:232-294): the
 * caller computes the -log2 cost vectors on the host exactly as the reference
 * does (:248-289) and passes them in; the k-edge-skip tables are built on the
 * device.
 */
typedef struct qmg_arch qmg_arch;

int qmg_arch_create(uint16_t n, const uint16_t* edges, size_t nEdges,
                    int device, qmg_arch** out);
int qmg_arch_set_fidelity_costs(qmg_arch* arch,
                                const double* singleQubitFidelityCosts /* n */,
                                const double* twoQubitFidelityCosts /* n*n */,
                                const double* swapFidelityCosts /* n*n */);
void qmg_arch_destroy(qmg_arch* arch);

typedef struct {
  uint16_t n;
  uint16_t _pad;
  uint32_t n_edges;
  int32_t bidirectional;  /* Architecture::bidirectional() */
  int32_t unidirectional; /* Architecture::unidirectional() */
  int32_t fidelity_available;
  uint32_t n_fidelity_tables; /* getFidelityDistanceTables().size() */
  int32_t device;
  float table_build_ms; /* device time spent building the tables */
} qmg_arch_info;

int qmg_arch_get_info(const qmg_arch* arch, qmg_arch_info* info);
/* getDistanceTable(includeReversalCost) (include/sc/Architecture.hpp:258) */
int qmg_arch_get_distance_table(const qmg_arch* arch, int includeReversalCost,
                                double* out /* n*n */);
/* getFidelityDistanceTables() (include/sc/Architecture.hpp:283) */
int qmg_arch_get_fidelity_distance_tables(const qmg_arch* arch,
                                          double* out /* K*n*n */);

/* ------------------------------------------------------------------------ */
/* A* routing                                                               */
/* ------------------------------------------------------------------------ */

/* one entry of TwoQubitMultiplicity (include/sc/Mapper.hpp:33-34): logical
 * pair q1 < q2 with the number of gates q1->q2 (forward) and q2->q1. */
typedef struct {
  uint16_t q1, q2, forward, reverse;
} qmg_gate2q;

/* one non-zero entry of SingleQubitMultiplicity (include/sc/Mapper.hpp:36) */
typedef struct {
  uint16_t qubit, multiplicity;
} qmg_gate1q;

/*
 * The layers of one circuit as produced by Mapper::createLayers
 * (src/sc/Mapper.cpp:138-281), flattened to CSR.  Gates of a layer are in
 * std::map<Edge,...> order (ascending (q1,q2)); 1q entries ascending qubit.
 */
typedef struct {
  uint32_t n_layers;
  uint32_t n_logical; /* logical qubit indices are < n_logical <= n */
  const uint32_t* gate2q_offsets; /* n_layers+1 */
  const qmg_gate2q* gates2q;
  const uint32_t* gate1q_offsets; /* n_layers+1 */
  const qmg_gate1q* gates1q;
} qmg_layers;

/*
 * Layered circuit: Mapper::createLayers (src/sc/Mapper.cpp:138-281) on a gate
 * list over logical qubits (control[i] = -1 for a single-qubit gate), with
 * the multiplicities of :237-280.  Host-only; the CSR view stays valid until
 * the circuit is split or destroyed.
 */
typedef struct qmg_circuit qmg_circuit;
int qmg_circuit_create(uint32_t nArchQubits, size_t nGates,
                       const int32_t* control, const int32_t* target,
                       int32_t layering /* qmg_layering */, qmg_circuit** out);
int qmg_circuit_layers(const qmg_circuit* circ, qmg_layers* out);
/* gate indices of every layer: offsets n_layers+1, indices nGates */
int qmg_circuit_layer_gates(const qmg_circuit* circ, uint32_t* offsets,
                            uint32_t* gateIndices);
/* Mapper::isLayerSplittable (src/sc/Mapper.cpp:283-300) */
int qmg_circuit_layer_splittable(const qmg_circuit* circ, uint32_t layer);
/* Mapper::splitLayer (src/sc/Mapper.cpp:302-435): layer -> layer, layer+1 */
int qmg_circuit_split_layer(qmg_circuit* circ, uint32_t layer);
void qmg_circuit_destroy(qmg_circuit* circ);

/*
 * Mapper::createLayers (src/sc/Mapper.cpp:138-281) for a batch of circuits at
 * once, on the host's threads, producing only the CSR layers a routing batch
 * reads (no per-layer containers; layers of such a circuit cannot be split by
 * qmg_circuit_split_layer - the library splits them itself when the instance
 * has a layer_capacity).  Circuit i is the gates gateOffsets[i] ..
 * gateOffsets[i+1] of control / target (control < 0 = single-qubit gate).
 * The views stay valid until the batch is destroyed.
 */
typedef struct qmg_circuit_batch qmg_circuit_batch;
int qmg_circuit_batch_create(uint32_t nArchQubits, size_t nCircuits,
                             const uint64_t* gateOffsets,
                             const int32_t* control, const int32_t* target,
                             int32_t layering /* qmg_layering */,
                             qmg_circuit_batch** out);
int qmg_circuit_batch_layers(const qmg_circuit_batch* batch, size_t i,
                             qmg_layers* out);
void qmg_circuit_batch_destroy(qmg_circuit_batch* batch);

/*
 * OpenQASM 2 reader for the gate set of the reference's example circuits
 * (qreg / creg, cx, one-qubit gates with parameter expressions, measure,
 * barrier; SURVEY section 8 row f2: inputs of 10^5 .. 10^6 gates).  The
 * reference delegates reading to mqt-core's qasm3::Importer::importf
 * (test/sc/heuristic/test_heuristic.cpp:253) and strips idle qubits in
 * Mapper's constructor (src/sc/Mapper.cpp:33-53); qmg_qasm_mapper_view gives
 * the gate operands as createLayers then reads them (Mapper.cpp:166-172) and
 * the InitialLayout::Identity placement (HeuristicMapper.cpp:176-183), ready
 * for qmg_circuit_create / qmg_circuit_batch_create.  Unsupported gates fail
 * with the reference's message (Mapper.cpp:157-161) in qmg_last_error().
 */
typedef struct qmg_qasm qmg_qasm;
typedef struct {
  uint32_t n_qubits, n_clbits;
  uint64_t n_gates, n_measurements;
} qmg_qasm_info;
int qmg_qasm_read_file(const char* path, qmg_qasm** out);
int qmg_qasm_parse(const char* text, size_t length, qmg_qasm** out);
int qmg_qasm_get_info(const qmg_qasm* q, qmg_qasm_info* info);
/* gate i: control[i] (-1: one-qubit gate), target[i], kind[i] (see
 * qmg_qasm_gate_name), params[3 i .. 3 i + 2] and n_params[i]; any output
 * pointer may be NULL */
int qmg_qasm_gates(const qmg_qasm* q, int32_t* control, int32_t* target,
                   uint8_t* kind, double* params, uint8_t* n_params);
int qmg_qasm_measurements(const qmg_qasm* q, uint32_t* qubit, uint32_t* clbit);
/* operands renumbered over the qubits that are not idle; identity_locations
 * (nPhysical entries, may be NULL): p for every used original index p, else -1 */
int qmg_qasm_mapper_view(const qmg_qasm* q, uint32_t nPhysical, int32_t* control,
                         int32_t* target, int16_t* identity_locations);
const char* qmg_qasm_gate_name(uint8_t kind); /* NULL beyond the last kind */
void qmg_qasm_destroy(qmg_qasm* q);

/* one search unit: a circuit, a start layout and a pass schedule */
typedef struct {
  qmg_layers layers;
  /* start layout: locations[logical] = physical or -1 (Mapper::locations);
   * n entries.  NULL = all unplaced (InitialLayout::Dynamic). */
  const int16_t* initial_locations;
  /* number of forward+backward pseudo-routing passes before the final
   * routing pass (Configuration::iterativeBidirectionalRoutingPasses,
   * HeuristicMapper.cpp:74-89) */
  uint32_t bidirectional_passes;
  /* first layer of the final pass to route and number of layers to route
   * (n_route_layers = 0 means "to the end"); used to resume after a split */
  uint32_t first_layer;
  uint32_t n_route_layers;
  /* capacity, in layers, of the per-layer outputs (layers, layer_qubits,
   * layer_n_swaps); 0 = layers.n_layers.  With a capacity above n_layers the
   * library splits layers itself when a search reaches
   * automatic_layer_splits_node_limit (Mapper::splitLayer, src/sc/Mapper.cpp:
   * 302-435, and the restart on the uncleared open list,
   * HeuristicMapper.cpp:563-584), in pseudo-routing passes too; a circuit never
   * ends up with more layers than it has (distinct) gate entries.  Otherwise a
   * split is handed to the caller as QMG_SPLIT_LIMIT_HIT. */
  uint32_t layer_capacity;
} qmg_instance;

/* instances[i].layers = the layers of circuit i for i < nInstances (the other
 * fields are left as they are) */
int qmg_circuit_batch_fill_instances(const qmg_circuit_batch* batch,
                                     qmg_instance* instances,
                                     size_t nInstances);
/* gate indices of every layer of circuit i (as qmg_circuit_layer_gates) */
int qmg_circuit_batch_layer_gates(const qmg_circuit_batch* batch, size_t i,
                                  uint32_t* offsets, uint32_t* gateIndices);

/* result of one aStarMap call (HeuristicMapper::Node + the counters of
 * MappingResults::LayerHeuristicBenchmarkInfo, include/sc/MappingResults.hpp:
 * 62-93) */
typedef struct {
  double cost_fixed;
  double cost_fixed_reversals;
  double cost_heur;
  double lookahead_penalty;
  uint64_t expanded_nodes;
  uint64_t generated_nodes;
  uint64_t expanded_nodes_after_first_solution;
  uint64_t expanded_nodes_after_optimal_solution;
  uint64_t solution_nodes;
  uint64_t solution_nodes_after_optimal_solution;
  uint64_t layout_hash; /* FNV-1a over the result's qubits[] */
  uint32_t solution_depth;
  uint32_t result_node_id;
  uint32_t swap_offset; /* into qmg_route_result.swaps */
  uint32_t n_swaps;
  uint32_t placement_offset; /* into qmg_route_result.placements */
  uint32_t n_placements;
  int32_t early_termination;
  int32_t status; /* qmg_status of this layer */
} qmg_layer_result;

/* one generated search node as DataLogger::logSearchNode records it
 * (src/sc/DataLogger.cpp:216-269: nodeId; parentId; costFixed; costHeur;
 * lookaheadPenalty; validMapping; depth; qubits; swaps).  The node's layout and
 * SWAP sequence follow from its ancestors and the layer's root layout
 * (qmg_route_result.layer_root_qubits). */
typedef struct {
  double cost_fixed; /* costFixed + costFixedReversals, as logged */
  double cost_heur;
  double lookahead_penalty;
  uint32_t id, parent, depth, layer;
  uint16_t swap_first, swap_second; /* the SWAP that made the node */
  uint16_t valid, root;
} qmg_node_log_record;

/* a logical qubit placed by mapUnmappedGates (HeuristicMapper.cpp:211-307) */
typedef struct {
  uint16_t logical, physical;
} qmg_placement;

/*
 * Caller-allocated output of one instance.  Capacities are in elements; the
 * library fails with QMG_BAD_ARG rather than overrun them.
 */
typedef struct {
  qmg_layer_result* layers; /* capacity n_layers of the instance, or NULL */
  uint16_t* swaps;          /* pairs (first, second); capacity swap_capacity */
  uint32_t swap_capacity;
  uint32_t n_swaps_total;
  qmg_placement* placements; /* capacity n (each logical placed once) */
  uint32_t placement_capacity;
  uint32_t n_placements_total;
  int16_t* final_qubits;    /* n: Mapper::qubits after the last layer */
  int16_t* final_locations; /* n: Mapper::locations after the last layer */
  /* layout before the final routing pass (after pseudo-routing passes) */
  int16_t* start_locations; /* n or NULL */
  /* optional: qubits[] after every layer of the final pass, n_layers*n */
  int16_t* layer_qubits;
  uint32_t n_layers_done; /* layers of the final pass completed */
  int32_t status;         /* QMG_OK, or the status that stopped the instance;
                             on QMG_SPLIT_LIMIT_HIT n_layers_done is the layer
                             to split and final_* hold the layout before it */
  uint64_t expanded_nodes;  /* all passes, incl. pseudo-routing */
  uint64_t generated_nodes; /* all passes */
  /* optional compact per-layer output: number of SWAPs inserted before each
   * layer of the final pass (n_layers entries).  With layers == NULL this is
   * all a caller needs to emit the routed circuit (routeCircuit :369-384). */
  uint16_t* layer_n_swaps;
  /* layers split by the library in the final pass, in the order of the splits:
   * entry k is the index (at that moment) of the layer that became layers
   * index, index+1.  Replaying qmg_circuit_split_layer over this list gives the
   * layer structure the per-layer outputs refer to (n_layers_done layers). */
  uint32_t* split_layers; /* capacity split_capacity, or NULL */
  uint32_t split_capacity;
  uint32_t n_splits;
  /* the final routing pass alone: what the reference reports as
   * heuristicBenchmark.expandedNodes / generatedNodes (its counters are off
   * during pseudoRouteCircuit, HeuristicMapper.cpp:309-344) */
  uint64_t expanded_nodes_final_pass;
  uint64_t generated_nodes_final_pass;
  /* optional (Configuration::dataLoggingPath, DataLogger.cpp:216-269): every
   * node generated in the final routing pass, in generation order, and the
   * root layout (qubits[]) of every layer, n entries per layer.  n_node_log is
   * the number generated (more than node_log_capacity: the log is cut). */
  qmg_node_log_record* node_log;
  uint32_t node_log_capacity;
  uint32_t n_node_log;
  int16_t* layer_root_qubits;
  /* optional: every placement of every pass (the pseudo-routing passes too),
   * in the order mapUnmappedGates made them; `physical` has bit 15 set when
   * the qubit was placed by the 1q-gate rule of the fidelity-aware mapper
   * (HeuristicMapper.cpp:213-231), which does not touch the mapped circuit's
   * initialLayout / outputPermutation (the 2q rules do, :275-306).  Capacity
   * n suffices: a qubit is placed once. */
  qmg_placement* placement_log;
  uint32_t placement_log_capacity;
  uint32_t n_placement_log;
} qmg_route_result;

/*
 * Replay check of routed instances (the reference delegates circuit
 * equivalence to mqt.qcec, test/python/test_heuristic_mapper.py:29-81, which
 * is not part of it): for instance i, starting from results[i].start_locations,
 * the SWAPs of every layer must sit on coupling edges, after them every 2q gate
 * of the layer must act on adjacent physical qubits (placements are consumed in
 * order for qubits placed on the way) and the layout reached must equal
 * final_locations - i.e. the routed circuit is the input circuit under the
 * SWAP-induced permutation.  verdict[i]: 1 valid, 0 invalid, 2 not checkable
 * (instance not finished, layers split by the library, outputs missing).
 * Host only.
 */
int qmg_circuit_batch_verify(const qmg_circuit_batch* batch, uint16_t n,
                             const uint16_t* edges, size_t nEdges,
                             const uint64_t* gateOffsets,
                             const int32_t* control, const int32_t* target,
                             const qmg_route_result* results,
                             size_t nInstances,
                             int32_t fidelityAware /* heuristic is
                             fidelity_best_location: 1q gates place qubits */,
                             uint8_t* verdict);

typedef struct {
  uint32_t pool_nodes;       /* node-pool capacity per resident search, 0=auto */
  uint32_t threads_per_search; /* 32..256, 0 = auto from n */
  uint32_t searches_per_sm;  /* resident searches per SM, 0 = auto */
  /* qmg_batch_create only: keep the compact per-layer outputs (SWAP counts,
   * SWAP lists, layouts) instead of the full qmg_layer_result records; with
   * qmg_route_batch the result pointers decide */
  uint32_t lean_outputs;
  /* cap on the resident searches of the whole grid (rounded down to whole
   * CTAs of 4 searches), 0 = none.  Few searches with large node pools suit
   * instances whose layers need millions of nodes. */
  uint32_t max_searches;
  /* Every search starts in a lean variant of the kernel (flat scan of the
   * open list, no layer splits on the device) and the instance is routed
   * again by the full variant when an open list outgrows 4096 nodes or a layer
   * has to be split; results are identical either way.  1 = use the full
   * variant from the start (long searches expected), 2 = lean first,
   * 0 = auto (full for fidelity-aware routing on more than 32 qubits). */
  uint32_t search_variant;
  /* Node pools come in two tiers: a private pool per resident search
   * (pool_nodes; auto = a share of the memory budget) and a few large pools
   * shared by the whole launch, which a search takes for a layer that outgrows
   * its private pool (the search state moves to the large pool and the
   * search carries on there; results are unchanged).  big_pools: number of large pools, 0 = auto (32 with
   * automatic pool sizes, none with an explicit pool_nodes), 0xFFFFFFFF = none.
   * big_pool_nodes: their capacity in nodes (needed with an explicit
   * pool_nodes, else auto). */
  uint32_t big_pools;
  uint32_t big_pool_nodes;
} qmg_route_options;

typedef struct {
  float kernel_ms;   /* device time of the routing kernel (CUDA events) */
  float h2d_ms;      /* staging */
  float d2h_ms;
  uint32_t grid;     /* CTAs launched */
  uint32_t block;    /* threads per CTA */
  uint64_t h2d_bytes;
  uint64_t d2h_bytes;
  uint64_t expanded_nodes; /* sum over instances and passes */
  uint64_t generated_nodes;
  uint32_t kernel_launches;
  /* qmg_route_batch with pool_nodes == 0: instances whose node pool ran out
   * and that were routed again with fewer resident searches and larger pools
   * (up to three such rounds; an instance still failing keeps
   * QMG_POOL_EXHAUSTED) */
  uint32_t retried_instances;
  /* layers that were searched in one of the launch's large pools */
  uint32_t big_pool_layers;
} qmg_route_stats;

/*
 * Route a batch of independent instances on one device.  For every instance
 * this runs what HeuristicMapper::map does between createInitialMapping and
 * postMappingOptimizations (HeuristicMapper.cpp:74-91): the pseudo-routing
 * passes and the final routing pass, i.e. for each layer mapUnmappedGates
 * (:211-307) followed by the A* search of aStarMap (:521-696) with expandNode
 * (:698-748), applySWAP (:861-973), the heuristics (:1055-1354) and the
 * lookahead penalties (:1356-1513), on an open list with the semantics of
 * UniquePriorityQueue (include/sc/heuristic/UniquePriorityQueue.hpp:59-93).
 */
int qmg_route_batch(const qmg_arch* arch, const qmg_config* cfg,
                    size_t nInstances, const qmg_instance* instances,
                    qmg_route_result* results,
                    const qmg_route_options* options /* NULL = auto */,
                    qmg_route_stats* stats /* NULL = none */);

/*
 * Device-resident variant: stage a batch once, then run it repeatedly with
 * the inputs already in HBM (used to separate kernel time from staging).
 */
typedef struct qmg_batch qmg_batch;
int qmg_batch_create(const qmg_arch* arch, const qmg_config* cfg,
                     size_t nInstances, const qmg_instance* instances,
                     const qmg_route_options* options, qmg_batch** out);
int qmg_batch_run(qmg_batch* batch, qmg_route_stats* stats);
int qmg_batch_fetch(qmg_batch* batch, qmg_route_result* results);
void qmg_batch_destroy(qmg_batch* batch);

/* message of the last failing call on this thread ("" if none) */
const char* qmg_last_error(void);
/* number of visible CUDA devices (0 without a driver) */
int qmg_device_count(void);
const char* qmg_version(void);

#ifdef __cplusplus
}
#endif
#endif /* QMAP_B200_H */
