// emu.cpp - HOST DEBUG BUILD of the routing kernel's source.
//
// TEST INFRASTRUCTURE ONLY.  Compiles mqt-qmap_b200/csrc/route_core.hpp (the
// very code the sm_100a kernel runs, one warp per search) with a 1-lane group
// so that the search logic can be exercised under ASan/UBSan and compared
// with the reference in a container that has no GPU.  It is built into
// tests/emu/libqmg_emu.so, is loaded only by tests/, is not linked into and
// cannot be reached from the product library (libqmap_b200.so), and is never
// timed.  Cost tables are INPUTS here (the table kernels exist only in CUDA).
#define QMG_HOST_API_IMPL
#include "../../mqt-qmap_b200/csrc/host_api.hpp"
#include "../../mqt-qmap_b200/csrc/flatten.hpp"
#include "../../mqt-qmap_b200/csrc/dijkstra_core.hpp"

#include <cstdlib>
#include <memory>
#include <vector>

using namespace qmg;

struct qemu_arch {
  ArchTopology topo;
  std::vector<double> dist, distRev, c1, c2, cswap, fid;
  int nFid = 0;
  bool fidelity = false;
  ArchView view() const {
    ArchView v{};
    v.n = topo.n;
    v.nE = topo.nE;
    v.bidirectional = topo.bidirectional ? 1 : 0;
    v.unidirectional = topo.unidirectional ? 1 : 0;
    v.fidelity = fidelity ? 1 : 0;
    v.nFid = nFid;
    v.edges = topo.edges.data();
    v.incOff = topo.incOff.data();
    v.inc = topo.inc.data();
    v.adj = topo.adj.data();
    v.dist = dist.data();
    v.distRev = distRev.data();
    v.c1 = c1.data();
    v.c2 = c2.data();
    v.cswap = cswap.data();
    v.fid = fid.data();
    return v;
  }
};

QMG_API int qemu_arch_create(uint16_t n, const uint16_t* edges, size_t nEdges,
                             const double* dist, const double* distRev,
                             qemu_arch** out) {
  try {
    auto a = std::make_unique<qemu_arch>();
    a->topo.build(n, edges, nEdges);
    const size_t nn = static_cast<size_t>(n) * n;
    a->dist.assign(dist, dist + nn);
    a->distRev.assign(distRev, distRev + nn);
    *out = a.release();
    return QMG_OK;
  } catch (const std::exception& e) {
    return fail(QMG_BAD_ARG, e.what());
  }
}
QMG_API int qemu_arch_set_fidelity(qemu_arch* a, const double* c1,
                                   const double* c2, const double* cswap,
                                   const double* fid, uint32_t nFid) {
  const size_t n = static_cast<size_t>(a->topo.n);
  a->c1.assign(c1, c1 + n);
  a->c2.assign(c2, c2 + n * n);
  a->cswap.assign(cswap, cswap + n * n);
  a->fid.assign(fid, fid + static_cast<size_t>(nFid) * n * n);
  a->nFid = static_cast<int>(nFid);
  a->fidelity = true;
  return QMG_OK;
}
QMG_API void qemu_arch_destroy(qemu_arch* a) { delete a; }

// Dijkstra::buildTable through the device function of the table kernel K1'
// (dijkstra_core.hpp), one source after the other
QMG_API int qemu_build_table(uint16_t n, const uint16_t* edges, size_t nEdges,
                             const double* w, double* out) {
  try {
    ArchTopology t;
    t.build(n, edges, nEdges);
    const int cap = 2 * t.nE + 2;
    std::vector<uint8_t> visited(n);
    std::vector<uint16_t> heap(static_cast<size_t>(cap));
    for (int src = 0; src < n; ++src) {
      if (!dijkstraReference(n, src, w, t.edges.data(), t.incOff.data(),
                             t.inc.data(), out + static_cast<size_t>(src) * n,
                             visited.data(), heap.data(), cap)) {
        return fail(QMG_BAD_ARG, "heap capacity exceeded");
      }
    }
    return QMG_OK;
  } catch (const std::exception& e) {
    return fail(QMG_BAD_ARG, e.what());
  }
}

namespace {
// the kernel's per-instance dispatch (route.cuh): lean variant first, the full
// one when that gives up (or right away: QMG_EMU_START_FULL=1)
template <int H>
void runBoth(const ArchView& A, const CfgView& C, InstView& I, SearchMem& M,
             Scratch& S, bool startFull) {
  bool full = startFull;
  if (!full) {
    Search<H, false> s(A, C, I, M, S);
    full = !s.run();
  }
  if (full) {
    Search<H, true> s(A, C, I, M, S);
    s.run();
  }
}
}  // namespace

QMG_API int qemu_route_batch(const qemu_arch* arch, const qmg_config* cfg,
                             size_t nInstances, const qmg_instance* instances,
                             qmg_route_result* results, uint32_t poolNodes) {
  static_assert(sizeof(LayerOut) == sizeof(qmg_layer_result), "layer result");
  const char* sfEnv = std::getenv("QMG_EMU_START_FULL");
  const bool startFull = sfEnv != nullptr && sfEnv[0] == '1';
  if (const char* msg = checkConfig(*cfg, arch->fidelity)) {
    return fail(QMG_BAD_ARG, msg);
  }
  const ArchView A = arch->view();
  CfgView C = toCfgView(*cfg);
  const std::vector<double> factors = lookaheadFactors(*cfg, 1U << 20);
  C.factors = factors.data();
  C.nFactors = static_cast<uint32_t>(factors.size());
  const int n = A.n;
  if (poolNodes == 0) {
    poolNodes = 1U << 16;
  }
  uint32_t tableSize = 1;
  while (tableSize < 2 * poolNodes) {
    tableSize <<= 1;
  }
  const uint32_t recBytes = nodeRecBytes(n);
  // QMG_EMU_PRIVATE_POOL=N: a private pool of N nodes and one large pool of
  // poolNodes nodes (the two tiers of the CUDA batches, BigPool)
  BigPool bigPool{};
  std::vector<uint64_t> bigStore, bigTable;
  std::vector<OpenEntry> bigOpen, bigBtop;
  if (const char* e = std::getenv("QMG_EMU_PRIVATE_POOL")) {
    const uint32_t priv = static_cast<uint32_t>(std::max(1, std::atoi(e)));
    bigStore.resize(static_cast<size_t>(poolNodes) * recBytes / 8);
    bigOpen.resize(poolNodes);
    bigTable.assign(tableSize, 0);
    bigBtop.resize(poolNodes / kBlk + 2);
    bigPool.pool = reinterpret_cast<uint8_t*>(bigStore.data());
    bigPool.open = bigOpen.data();
    bigPool.table = bigTable.data();
    bigPool.btop = bigBtop.data();
    bigPool.poolCap = poolNodes;
    bigPool.tableMask = tableSize - 1;
    poolNodes = priv;
    tableSize = 1;
    while (tableSize < 2 * poolNodes) {
      tableSize <<= 1;
    }
  }
  std::vector<uint64_t> poolStore(static_cast<size_t>(poolNodes) * recBytes / 8 + 1);
  std::vector<OpenEntry> open(poolNodes);
  std::vector<uint64_t> table(tableSize, 0);
  Scratch S = scratchLayout(n);
  std::vector<uint64_t> scratch(S.bytes / 8 + 1);
  S.base = reinterpret_cast<uint8_t*>(scratch.data());
  uint32_t stamp = 0;
  for (size_t ii = 0; ii < nInstances; ++ii) {
    const qmg_instance& in = instances[ii];
    qmg_route_result& r = results[ii];
    // the same flattening the CUDA batches use (flatten.hpp)
    const qmg_layers& Ly = in.layers;
    CircSlot slot;
    slot.nLayers = Ly.n_layers;
    slot.nLogical = Ly.n_logical;
    slot.nG2 = Ly.n_layers > 0 ? Ly.gate2q_offsets[Ly.n_layers] : 0;
    slot.nG1 = (Ly.n_layers > 0 && Ly.gate1q_offsets != nullptr)
                   ? Ly.gate1q_offsets[Ly.n_layers]
                   : 0;
    const size_t nLy = Ly.n_layers;
    std::vector<uint32_t> fG2Off(nLy + 1), fG1Off(nLy + 1), fConsOff(nLy + 1),
        fTwoQ(nLy + 1), fFirst(nLy + 1);
    std::vector<Gate2> fG2(slot.nG2 + 1);
    std::vector<Gate1> fG1(slot.nG1 + 1);
    std::vector<uint16_t> fCons(2 * slot.nG2 + slot.nG1 + 1);
    std::vector<uint8_t> fSplit(nLy + 1);
    if (!flattenCircuit(n, C.fidelityAware != 0, Ly.gate2q_offsets,
                        reinterpret_cast<const Gate2*>(Ly.gates2q),
                        Ly.gate1q_offsets,
                        reinterpret_cast<const Gate1*>(Ly.gates1q), slot,
                        fG2Off.data(), fG2.data(), fG1Off.data(), fG1.data(),
                        fConsOff.data(), fCons.data(), fTwoQ.data(),
                        fFirst.data(), fSplit.data())) {
      return fail(QMG_BAD_ARG, "gate operand out of range");
    }
    struct {
      uint32_t maxG;
    } fc{slot.maxG};
    CircView cv{};
    cv.nLayers = slot.nLayers;
    cv.nLogical = slot.nLogical;
    cv.g2Off = fG2Off.data();
    cv.g2 = fG2.data();
    cv.g1Off = fG1Off.data();
    cv.g1 = fG1.data();
    // considered qubits / splittable flags in the layout the device derives
    // (derive.cuh), by the same function
    std::vector<uint16_t> dCons(2 * slot.nG2 + slot.nG1 + 1);
    std::vector<uint16_t> dConsCnt(nLy + 1);
    for (uint32_t l = 0; l < nLy; ++l) {
      deriveLayerTables(fG2Off.data(), fG2.data(), fG1Off.data(), fG1.data(),
                        dCons.data(), dConsCnt.data(), fSplit.data(), l,
                        C.fidelityAware != 0);
    }
    cv.consOff = nullptr;
    cv.cons = dCons.data();
    cv.consCnt = dConsCnt.data();
    cv.twoQ = fTwoQ.data();
    cv.firstAfter = fFirst.data();
    cv.nTwoQ = slot.nTwoQ;
    cv.splittable = fSplit.data();
    std::vector<uint16_t> nsw(static_cast<size_t>(fc.maxG) + 1);
    SearchMem M{};
    M.pool = reinterpret_cast<uint8_t*>(poolStore.data());
    M.open = open.data();
    M.table = table.data();
    std::vector<OpenEntry> btop(poolNodes / kBlk + 2);
    M.btop = btop.data();
    std::vector<PairSlot> pair;
    if (C.heuristic >= 4) {
      const size_t slots = static_cast<size_t>(kPairGates) * A.n * A.n;
      pair.assign(slots, PairSlot{0, 0.});
      M.pair = pair.data();
    }
    M.nsw = nsw.data();
    M.poolCap = poolNodes;
    M.bigHeld = -1;
    M.big = bigPool.poolCap != 0 ? &bigPool : nullptr;
    M.nBig = bigPool.poolCap != 0 ? 1U : 0U;
    M.recBytes = recBytes;
    M.tableMask = tableSize - 1;
    M.stamp = stamp;
    M.maxG = fc.maxG;
    // device-side layer splits: a private, writable copy of the tables
    const uint32_t layerCap = std::max<uint32_t>(in.layer_capacity, slot.nLayers);
    std::vector<uint64_t> privStore;
    if (layerCap > slot.nLayers) {
      M.privLayers = layerCap;
      M.privG2 = static_cast<uint32_t>(slot.nG2) + 1;
      M.privG1 = static_cast<uint32_t>(slot.nG1) + 1;
      privStore.resize(privBytes(M.privLayers, M.privG2, M.privG1) / 8 + 2);
      M.priv = reinterpret_cast<uint8_t*>(privStore.data());
    }
    InstView I{};
    I.circ = cv;
    I.initLoc = in.initial_locations;
    I.passes = in.bidirectional_passes;
    I.firstLayer = in.first_layer;
    I.nRoute = in.n_route_layers;
    I.layerCap = layerCap;
    I.splitLog = r.split_layers;
    I.splitCap = r.split_layers != nullptr ? r.split_capacity : 0;
    I.layers = reinterpret_cast<LayerOut*>(r.layers);
    I.swaps = r.swaps;
    I.swapCap = r.swap_capacity;
    I.places = reinterpret_cast<uint16_t*>(r.placements);
    I.placeCap = r.placement_capacity;
    I.finalQ = r.final_qubits;
    I.finalL = r.final_locations;
    I.startL = r.start_locations;
    I.layerQ = r.layer_qubits;
    I.nodeLog = reinterpret_cast<NodeLogRec*>(r.node_log);
    I.nodeLogCap = r.node_log != nullptr ? r.node_log_capacity : 0;
    I.rootQ = r.layer_root_qubits;
    I.placeLog = reinterpret_cast<uint16_t*>(r.placement_log);
    I.placeLogCap = r.placement_log != nullptr ? r.placement_log_capacity : 0;
    uint32_t summary[8] = {0, 0, 0, 0, 0, 0, 0, 0};
    I.layerNSwaps = r.layer_n_swaps;
    uint64_t counters[4] = {0, 0, 0, 0};
    I.summary = summary;
    I.counters = counters;
    // (a node log is written by the full variant only)
    const bool sf = startFull || I.nodeLog != nullptr || I.placeLog != nullptr;
    switch (C.heuristic) {
    case 0: { runBoth<0>(A, C, I, M, S, sf); break; }
    case 1: { runBoth<1>(A, C, I, M, S, sf); break; }
    case 2: { runBoth<2>(A, C, I, M, S, sf); break; }
    case 3: { runBoth<3>(A, C, I, M, S, sf); break; }
    default: { runBoth<4>(A, C, I, M, S, sf); break; }
    }
    stamp = M.stamp;
    r.n_swaps_total = summary[0];
    r.n_placements_total = summary[1];
    r.n_layers_done = summary[2];
    r.n_splits = summary[5];
    r.n_node_log = summary[6];
    r.n_placement_log = summary[7];
    r.status = static_cast<int32_t>(summary[3]);
    r.expanded_nodes = counters[0];
    r.generated_nodes = counters[1];
    r.expanded_nodes_final_pass = counters[2];
    r.generated_nodes_final_pass = counters[3];
  }
  return QMG_OK;
}
