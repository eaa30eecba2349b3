// api.cu - the C ABI of include/qmap_b200.h on top of the sm_100a kernels
// (tables.cuh, route.cuh).  No CPU fallback: without a CUDA device every
// compute entry point returns QMG_NO_DEVICE.
#define QMG_HOST_API_IMPL
#include "host_api.hpp"
#include "route.cuh"
#include "tables.cuh"

#include <algorithm>
#include <cstdio>
#include <map>
#include <numeric>
#include <tuple>
#include <vector>

using namespace qmg;

namespace {

struct CudaFail {
  int code;
  std::string msg;
};
inline void cuCheck(cudaError_t e, const char* what) {
  if (e != cudaSuccess) {
    const int code = (e == cudaErrorMemoryAllocation) ? QMG_OOM : QMG_CUDA_ERROR;
    throw CudaFail{code, std::string(what) + ": " + cudaGetErrorString(e)};
  }
}
#define CU(x) cuCheck((x), #x)

template <class F> int guarded(F&& f) {
  try {
    return f();
  } catch (const CudaFail& e) {
    return fail(e.code, e.msg);
  } catch (const std::bad_alloc&) {
    return fail(QMG_OOM, "host allocation failed");
  } catch (const std::exception& e) {
    return fail(QMG_BAD_ARG, e.what());
  }
}

int deviceCount() {
  int n = 0;
  if (cudaGetDeviceCount(&n) != cudaSuccess) {
    cudaGetLastError();
    return 0;
  }
  return n;
}
int selectDevice(int device) {
  const int n = deviceCount();
  if (n == 0) {
    return fail(QMG_NO_DEVICE, "no CUDA device available (this library has no "
                               "CPU fallback)");
  }
  if (device < 0 || device >= n) {
    return fail(QMG_BAD_ARG, "device index out of range");
  }
  if (cudaSetDevice(device) != cudaSuccess) {
    return fail(QMG_CUDA_ERROR, "cudaSetDevice failed");
  }
  return QMG_OK;
}

// RAII device buffer
template <class T> struct DevBuf {
  T* p = nullptr;
  size_t count = 0;
  DevBuf() = default;
  DevBuf(const DevBuf&) = delete;
  DevBuf& operator=(const DevBuf&) = delete;
  DevBuf(DevBuf&& o) noexcept : p(o.p), count(o.count) {
    o.p = nullptr;
    o.count = 0;
  }
  DevBuf& operator=(DevBuf&& o) noexcept {
    if (this != &o) {
      release();
      p = o.p;
      count = o.count;
      o.p = nullptr;
      o.count = 0;
    }
    return *this;
  }
  ~DevBuf() { release(); }
  void release() {
    if (p != nullptr) {
      cudaFree(p);
      p = nullptr;
    }
    count = 0;
  }
  void alloc(size_t n) {
    release();
    count = n;
    if (n > 0) {
      CU(cudaMalloc(reinterpret_cast<void**>(&p), n * sizeof(T)));
    }
  }
  void upload(const T* h, size_t n, cudaStream_t s = nullptr) {
    alloc(n);
    if (n > 0) {
      CU(cudaMemcpyAsync(p, h, n * sizeof(T), cudaMemcpyHostToDevice, s));
    }
  }
  void upload(const std::vector<T>& v, cudaStream_t s = nullptr) {
    upload(v.data(), v.size(), s);
  }
};

struct EventTimer {
  cudaEvent_t a = nullptr, b = nullptr;
  cudaStream_t s;
  explicit EventTimer(cudaStream_t st) : s(st) {
    CU(cudaEventCreate(&a));
    CU(cudaEventCreate(&b));
  }
  ~EventTimer() {
    cudaEventDestroy(a);
    cudaEventDestroy(b);
  }
  void start() { CU(cudaEventRecord(a, s)); }
  float stop() {
    CU(cudaEventRecord(b, s));
    CU(cudaEventSynchronize(b));
    float ms = 0;
    CU(cudaEventElapsedTime(&ms, a, b));
    return ms;
  }
};

// --------------------------------------------------------------------------
// table builders on device buffers
// --------------------------------------------------------------------------
struct TopoDev {
  DevBuf<uint16_t> edges;
  DevBuf<uint32_t> incOff, inc, nbrOff;
  DevBuf<uint16_t> nbr;
  DevBuf<uint8_t> adj;
  void upload(const ArchTopology& t) {
    edges.upload(t.edges);
    incOff.upload(t.incOff);
    inc.upload(t.inc);
    adj.upload(t.adj);
    // undirected neighbour lists for the label-correcting kernel
    std::vector<uint32_t> off{0};
    std::vector<uint16_t> nb;
    for (int v = 0; v < t.n; ++v) {
      for (int u = 0; u < t.n; ++u) {
        if (u != v && t.adj[static_cast<size_t>(u) * t.n + v] != 0) {
          nb.push_back(static_cast<uint16_t>(u));
        }
      }
      off.push_back(static_cast<uint32_t>(nb.size()));
    }
    nbrOff.upload(off);
    nbr.upload(nb);
  }
};

// Dijkstra::buildTable on the device; w and out are n*n device matrices
void devBuildTable(int n, const TopoDev& td, const double* w, bool exact,
                   double* out, cudaStream_t s) {
  if (n == 0) {
    return;
  }
  if (exact) {
    const size_t sm = 2 * static_cast<size_t>(n) * sizeof(double);
    sssp_exact_kernel<<<n, 256, sm, s>>>(n, w, td.nbrOff.p, td.nbr.p, out);
    CU(cudaGetLastError());
    return;
  }
  const int nb = (n + kFwTile - 1) / kFwTile;
  const int np = nb * kFwTile;
  DevBuf<double> pad;
  pad.alloc(static_cast<size_t>(np) * np);
  const size_t tot = static_cast<size_t>(np) * np;
  fw_init<<<static_cast<unsigned>((tot + 255) / 256), 256, 0, s>>>(
      pad.p, np, n, w, td.adj.p);
  for (int kb = 0; kb < nb; ++kb) {
    fw_phase1<<<1, 256, 0, s>>>(pad.p, np, kb);
    if (nb > 1) {
      fw_phase2<<<dim3(nb - 1, 2), 256, 0, s>>>(pad.p, np, kb);
      fw_phase3<<<dim3(nb - 1, nb - 1), 256, 0, s>>>(pad.p, np, kb);
    }
  }
  fw_finish<<<static_cast<unsigned>((static_cast<size_t>(n) * n + 255) / 256),
              256, 0, s>>>(pad.p, np, n, out);
  CU(cudaGetLastError());
  CU(cudaStreamSynchronize(s));  // pad is freed on return
}

void devSingleEdgeSkip(int n, const TopoDev& td, int nE, const double* D,
                       double rev, double* out, cudaStream_t s) {
  if (n == 0) {
    return;
  }
  edge_skip1_kernel<<<n, 256, static_cast<size_t>(n) * sizeof(double), s>>>(
      n, D, td.edges.p, nE, rev, out);
  if (rev == 0.) {
    skip_mirror_kernel<<<static_cast<unsigned>(
                             (static_cast<size_t>(n) * n + 255) / 256),
                         256, 0, s>>>(n, out);
  }
  CU(cudaGetLastError());
}

// Dijkstra::buildEdgeSkipTable; returns the tables (device, K*n*n)
size_t devEdgeSkipTables(int n, const TopoDev& td, int nE, const double* w,
                         bool exact, DevBuf<double>& tables, size_t maxTables,
                         cudaStream_t s) {
  const size_t nn = static_cast<size_t>(n) * n;
  // the reference tries k = 1..n and stops at the first all-zero level
  size_t cap = std::min<size_t>(maxTables, static_cast<size_t>(n) + 1);
  cap = std::max<size_t>(cap, 1);
  size_t alloc = std::min<size_t>(cap, 32);
  tables.alloc(alloc * nn);
  devBuildTable(n, td, w, exact, tables.p, s);
  DevBuf<int> flag;
  flag.alloc(1);
  size_t K = 1;
  for (size_t k = 1; k <= static_cast<size_t>(n); ++k) {
    if (K + 1 > alloc) {
      if (alloc >= cap) {
        break;  // caller's capacity reached
      }
      const size_t na = std::min(cap, alloc * 2);
      DevBuf<double> bigger;
      bigger.alloc(na * nn);
      CU(cudaMemcpyAsync(bigger.p, tables.p, K * nn * sizeof(double),
                         cudaMemcpyDeviceToDevice, s));
      CU(cudaStreamSynchronize(s));
      tables = std::move(bigger);
      alloc = na;
    }
    CU(cudaMemsetAsync(flag.p, 0, sizeof(int), s));
    double* cur = tables.p + K * nn;
    edge_skipk_kernel<<<n, 256, 0, s>>>(n, static_cast<int>(k), tables.p,
                                        td.edges.p, nE, cur, flag.p);
    skip_mirror_kernel<<<static_cast<unsigned>((nn + 255) / 256), 256, 0, s>>>(
        n, cur);
    CU(cudaGetLastError());
    int notZero = 0;
    CU(cudaMemcpyAsync(&notZero, flag.p, sizeof(int), cudaMemcpyDeviceToHost,
                       s));
    CU(cudaStreamSynchronize(s));
    if (nE > 0 && notZero == 0) {
      break;  // all-zero level is dropped (utils.cpp:139-143)
    }
    ++K;
  }
  return K;
}

}  // namespace

// ==========================================================================
// stand-alone table entry points
// ==========================================================================
QMG_API int qmg_device_count(void) { return deviceCount(); }
QMG_API const char* qmg_version(void) { return "qmap_b200 0.1 (sm_100a)"; }

QMG_API int qmg_build_table(uint16_t n, const uint16_t* edges, size_t nEdges,
                            const double* edgeWeights, int exact_association,
                            double* distanceTable, int device) {
  if (int rc = selectDevice(device)) {
    return rc;
  }
  return guarded([&] {
    ArchTopology t;
    t.build(n, edges, nEdges);
    TopoDev td;
    td.upload(t);
    const size_t nn = static_cast<size_t>(n) * n;
    DevBuf<double> w, out;
    w.upload(edgeWeights, nn);
    out.alloc(nn);
    devBuildTable(n, td, w.p, exact_association != 0, out.p, nullptr);
    CU(cudaMemcpy(distanceTable, out.p, nn * sizeof(double),
                  cudaMemcpyDeviceToHost));
    return static_cast<int>(QMG_OK);
  });
}

QMG_API int qmg_build_single_edge_skip_table(uint16_t n,
                                             const double* distanceTable,
                                             const uint16_t* edges,
                                             size_t nEdges, double reversalCost,
                                             double* edgeSkipDistanceTable,
                                             int device) {
  if (int rc = selectDevice(device)) {
    return rc;
  }
  return guarded([&] {
    ArchTopology t;
    t.build(n, edges, nEdges);
    TopoDev td;
    td.upload(t);
    const size_t nn = static_cast<size_t>(n) * n;
    DevBuf<double> D, out;
    D.upload(distanceTable, nn);
    out.alloc(nn);
    devSingleEdgeSkip(n, td, t.nE, D.p, reversalCost, out.p, nullptr);
    CU(cudaMemcpy(edgeSkipDistanceTable, out.p, nn * sizeof(double),
                  cudaMemcpyDeviceToHost));
    return static_cast<int>(QMG_OK);
  });
}

QMG_API int qmg_build_edge_skip_table(uint16_t n, const uint16_t* edges,
                                      size_t nEdges, const double* edgeWeights,
                                      int exact_association, double* tables,
                                      size_t maxTables, size_t* nTables,
                                      int device) {
  if (int rc = selectDevice(device)) {
    return rc;
  }
  return guarded([&] {
    ArchTopology t;
    t.build(n, edges, nEdges);
    TopoDev td;
    td.upload(t);
    const size_t nn = static_cast<size_t>(n) * n;
    DevBuf<double> w, tb;
    w.upload(edgeWeights, nn);
    const size_t K = devEdgeSkipTables(n, td, t.nE, w.p, exact_association != 0,
                                       tb, maxTables, nullptr);
    *nTables = K;
    CU(cudaMemcpy(tables, tb.p, std::min(K, maxTables) * nn * sizeof(double),
                  cudaMemcpyDeviceToHost));
    return static_cast<int>(QMG_OK);
  });
}

// ==========================================================================
// Architecture handle
// ==========================================================================
struct qmg_arch {
  ArchTopology topo;
  TopoDev td;
  int device = 0;
  cudaStream_t stream = nullptr;
  DevBuf<double> dist, distRev, c1, c2, cswap, fid;
  std::vector<double> hDist, hDistRev, hFid;
  bool fidelity = false;
  uint32_t nFid = 0;
  float buildMs = 0.F;
  ~qmg_arch() {
    if (stream != nullptr) {
      cudaStreamDestroy(stream);
    }
  }
  ArchView view() const {
    ArchView v{};
    v.n = topo.n;
    v.nE = topo.nE;
    v.bidirectional = topo.bidirectional ? 1 : 0;
    v.unidirectional = topo.unidirectional ? 1 : 0;
    v.fidelity = fidelity ? 1 : 0;
    v.nFid = static_cast<int32_t>(nFid);
    v.edges = td.edges.p;
    v.incOff = td.incOff.p;
    v.inc = td.inc.p;
    v.adj = td.adj.p;
    v.dist = dist.p;
    v.distRev = distRev.p;
    v.c1 = c1.p;
    v.c2 = c2.p;
    v.cswap = cswap.p;
    v.fid = fid.p;
    return v;
  }
};

QMG_API int qmg_arch_create(uint16_t n, const uint16_t* edges, size_t nEdges,
                            int device, qmg_arch** out) {
  if (int rc = selectDevice(device)) {
    return rc;
  }
  return guarded([&] {
    auto a = std::make_unique<qmg_arch>();
    a->device = device;
    a->topo.build(n, edges, nEdges);
    CU(cudaStreamCreate(&a->stream));
    a->td.upload(a->topo);
    const size_t nn = static_cast<size_t>(n) * n;
    // Architecture::createDistanceTable :201-230
    std::vector<double> w(nn, DBL_MAX);
    const auto& t = a->topo;
    for (int e = 0; e < t.nE; ++e) {
      const size_t x = t.edges[2 * e];
      const size_t y = t.edges[2 * e + 1];
      if ((t.adj[x * n + y] & 2U) == 0) {  // twin (y, x) absent
        w[y * n + x] = kCostUniSwap;
        w[x * n + y] = kCostUniSwap;
      } else {
        w[x * n + y] = kCostBiSwap;
      }
    }
    DevBuf<double> dw, simple;
    dw.upload(w, a->stream);
    simple.alloc(nn);
    a->dist.alloc(nn);
    a->distRev.alloc(nn);
    EventTimer tm(a->stream);
    tm.start();
    devBuildTable(n, a->td, dw.p, false, simple.p, a->stream);
    devSingleEdgeSkip(n, a->td, t.nE, simple.p, 0., a->dist.p, a->stream);
    if (t.bidirectional) {
      CU(cudaMemcpyAsync(a->distRev.p, a->dist.p, nn * sizeof(double),
                         cudaMemcpyDeviceToDevice, a->stream));
    } else {
      devSingleEdgeSkip(n, a->td, t.nE, simple.p,
                        static_cast<double>(kCostReverse), a->distRev.p,
                        a->stream);
    }
    a->buildMs = tm.stop();
    a->hDist.resize(nn);
    a->hDistRev.resize(nn);
    CU(cudaMemcpyAsync(a->hDist.data(), a->dist.p, nn * sizeof(double),
                       cudaMemcpyDeviceToHost, a->stream));
    CU(cudaMemcpyAsync(a->hDistRev.data(), a->distRev.p, nn * sizeof(double),
                       cudaMemcpyDeviceToHost, a->stream));
    CU(cudaStreamSynchronize(a->stream));
    *out = a.release();
    return static_cast<int>(QMG_OK);
  });
}

QMG_API int qmg_arch_set_fidelity_costs(qmg_arch* a, const double* c1,
                                        const double* c2, const double* cswap) {
  if (a == nullptr || c1 == nullptr || c2 == nullptr || cswap == nullptr) {
    return fail(QMG_BAD_ARG, "null argument");
  }
  if (int rc = selectDevice(a->device)) {
    return rc;
  }
  return guarded([&] {
    const int n = a->topo.n;
    const size_t nn = static_cast<size_t>(n) * n;
    a->c1.upload(c1, static_cast<size_t>(n), a->stream);
    a->c2.upload(c2, nn, a->stream);
    a->cswap.upload(cswap, nn, a->stream);
    EventTimer tm(a->stream);
    tm.start();
    // Architecture::createFidelityTable :291-293; exact association so that
    // the tables equal the reference's bit for bit
    const size_t K = devEdgeSkipTables(n, a->td, a->topo.nE, a->cswap.p, true,
                                       a->fid, static_cast<size_t>(n) + 1,
                                       a->stream);
    a->buildMs += tm.stop();
    a->nFid = static_cast<uint32_t>(K);
    a->hFid.resize(K * nn);
    CU(cudaMemcpyAsync(a->hFid.data(), a->fid.p, K * nn * sizeof(double),
                       cudaMemcpyDeviceToHost, a->stream));
    CU(cudaStreamSynchronize(a->stream));
    a->fidelity = true;
    return static_cast<int>(QMG_OK);
  });
}

QMG_API void qmg_arch_destroy(qmg_arch* a) {
  if (a != nullptr) {
    cudaSetDevice(a->device);
    delete a;
  }
}

QMG_API int qmg_arch_get_info(const qmg_arch* a, qmg_arch_info* info) {
  if (a == nullptr || info == nullptr) {
    return fail(QMG_BAD_ARG, "null argument");
  }
  std::memset(info, 0, sizeof(*info));
  info->n = static_cast<uint16_t>(a->topo.n);
  info->n_edges = static_cast<uint32_t>(a->topo.nE);
  info->bidirectional = a->topo.bidirectional ? 1 : 0;
  info->unidirectional = a->topo.unidirectional ? 1 : 0;
  info->fidelity_available = a->fidelity ? 1 : 0;
  info->n_fidelity_tables = a->nFid;
  info->device = a->device;
  info->table_build_ms = a->buildMs;
  return QMG_OK;
}
QMG_API int qmg_arch_get_distance_table(const qmg_arch* a,
                                        int includeReversalCost, double* out) {
  if (a == nullptr || out == nullptr) {
    return fail(QMG_BAD_ARG, "null argument");
  }
  const auto& src = includeReversalCost != 0 ? a->hDistRev : a->hDist;
  std::memcpy(out, src.data(), src.size() * sizeof(double));
  return QMG_OK;
}
QMG_API int qmg_arch_get_fidelity_distance_tables(const qmg_arch* a,
                                                  double* out) {
  if (a == nullptr || out == nullptr) {
    return fail(QMG_BAD_ARG, "null argument");
  }
  if (!a->fidelity) {
    return fail(QMG_BAD_ARG, "No fidelity data available.");
  }
  std::memcpy(out, a->hFid.data(), a->hFid.size() * sizeof(double));
  return QMG_OK;
}

// ==========================================================================
// Routing batches
// ==========================================================================
struct qmg_batch {
  const qmg_arch* arch = nullptr;
  CfgView cfg{};
  size_t nInst = 0;
  int n = 0;
  bool wantLayerQ = false;
  // per instance (host)
  std::vector<uint32_t> nLayers, layerOff, swapCap, swapOff;
  uint64_t totLayers = 0, totSwapCap = 0;
  // device inputs
  DevBuf<uint32_t> g2Off, g1Off, consOff, twoQ, firstAfter, order, ticket;
  DevBuf<double> factors;
  DevBuf<Gate2> g2;
  DevBuf<Gate1> g1;
  DevBuf<uint16_t> cons;
  DevBuf<uint8_t> splittable;
  DevBuf<int16_t> initLoc;
  DevBuf<InstView> inst;
  // device working memory
  DevBuf<SearchMem> mem;
  DevBuf<uint8_t> pool;
  DevBuf<OpenEntry> open;
  DevBuf<uint64_t> table;
  DevBuf<uint16_t> nsw;
  // device outputs
  DevBuf<LayerOut> oLayers;
  DevBuf<uint16_t> oSwaps, oPlaces;
  DevBuf<int16_t> oFinalQ, oFinalL, oStartL, oLayerQ;
  DevBuf<uint32_t> oSummary;
  DevBuf<uint64_t> oCounters;
  uint32_t grid = 0, poolNodes = 0;
  SmemLayout sl{};
  size_t smemBytes = 0;
  uint64_t h2dBytes = 0;
  float h2dMs = 0.F;
  uint32_t launches = 0;
};

namespace {

struct CircKey {
  const void* a;
  const void* b;
  const void* c;
  const void* d;
  uint32_t nl;
  bool operator<(const CircKey& o) const {
    return std::tie(a, b, c, d, nl) < std::tie(o.a, o.b, o.c, o.d, o.nl);
  }
};

int batchCreate(const qmg_arch* arch, const qmg_config* cfg, size_t nInst,
                const qmg_instance* instances, const qmg_route_options* opt,
                const qmg_route_result* capacities, qmg_batch** out) {
  if (arch == nullptr || cfg == nullptr || out == nullptr ||
      (nInst > 0 && instances == nullptr)) {
    return fail(QMG_BAD_ARG, "null argument");
  }
  if (const char* msg = checkConfig(*cfg, arch->fidelity)) {
    return fail(QMG_BAD_ARG, msg);
  }
  if (int rc = selectDevice(arch->device)) {
    return rc;
  }
  auto b = std::make_unique<qmg_batch>();
  b->arch = arch;
  b->cfg = toCfgView(*cfg);
  if (cfg->nr_lookaheads > (1U << 20)) {
    return fail(QMG_BAD_ARG, "nr_lookaheads too large");
  }
  b->nInst = nInst;
  const int n = arch->topo.n;
  b->n = n;
  cudaStream_t s = arch->stream;
  EventTimer tm(s);
  tm.start();

  // ---- circuits (shared between instances that point at the same layers)
  std::map<CircKey, uint32_t> seen;
  struct CircPos {
    size_t g2Off, g2, g1Off, g1, consOff, cons, layer;
    uint32_t nLayers, nLogical, maxG;
  };
  std::vector<CircPos> circs;
  std::vector<uint32_t> hG2Off, hG1Off, hConsOff, hTwoQ, hFirst;
  std::vector<Gate2> hG2;
  std::vector<Gate1> hG1;
  std::vector<uint16_t> hCons;
  std::vector<uint8_t> hSplit;
  std::vector<uint32_t> circOf(nInst);
  std::vector<std::pair<size_t, size_t>> twoQPos;
  uint32_t maxG = 1;
  for (size_t i = 0; i < nInst; ++i) {
    const qmg_layers& L = instances[i].layers;
    if (L.n_layers > 0 && (L.gate2q_offsets == nullptr ||
                           (L.gates2q == nullptr &&
                            L.gate2q_offsets[L.n_layers] > 0))) {
      return fail(QMG_BAD_ARG, "instance without layers");
    }
    const CircKey key{L.gate2q_offsets, L.gates2q, L.gate1q_offsets, L.gates1q,
                      L.n_layers};
    auto it = seen.find(key);
    if (it == seen.end()) {
      FlatCircuit fc;
      fc.build(L.n_layers, L.n_logical, L.gate2q_offsets,
               reinterpret_cast<const Gate2*>(L.gates2q), L.gate1q_offsets,
               reinterpret_cast<const Gate1*>(L.gates1q),
               b->cfg.fidelityAware != 0);
      for (const auto& g : fc.g2) {
        if (g.q1 >= n || g.q2 >= n || g.q1 == g.q2) {
          return fail(QMG_BAD_ARG, "2q gate operand out of range");
        }
      }
      for (const auto& g : fc.g1) {
        if (g.q >= n) {
          return fail(QMG_BAD_ARG, "1q gate operand out of range");
        }
      }
      CircPos p{hG2Off.size(), hG2.size(),     hG1Off.size(), hG1.size(),
                hConsOff.size(), hCons.size(), hFirst.size(),  fc.nLayers,
                fc.nLogical,     fc.maxG};
      hG2Off.insert(hG2Off.end(), fc.g2Off.begin(), fc.g2Off.end());
      hG2.insert(hG2.end(), fc.g2.begin(), fc.g2.end());
      hG1Off.insert(hG1Off.end(), fc.g1Off.begin(), fc.g1Off.end());
      hG1.insert(hG1.end(), fc.g1.begin(), fc.g1.end());
      hConsOff.insert(hConsOff.end(), fc.consOff.begin(), fc.consOff.end());
      hCons.insert(hCons.end(), fc.cons.begin(), fc.cons.end());
      hFirst.insert(hFirst.end(), fc.firstAfter.begin(), fc.firstAfter.end());
      twoQPos.push_back({hTwoQ.size(), fc.twoQ.size()});
      hTwoQ.insert(hTwoQ.end(), fc.twoQ.begin(), fc.twoQ.end());
      hSplit.insert(hSplit.end(), fc.splittable.begin(), fc.splittable.end());
      maxG = std::max(maxG, fc.maxG);
      it = seen.emplace(key, static_cast<uint32_t>(circs.size())).first;
      circs.push_back(p);
    }
    circOf[i] = it->second;
  }
  b->g2Off.upload(hG2Off, s);
  b->g2.upload(hG2, s);
  b->g1Off.upload(hG1Off, s);
  b->g1.upload(hG1, s);
  b->consOff.upload(hConsOff, s);
  b->cons.upload(hCons, s);
  b->twoQ.upload(hTwoQ, s);
  b->firstAfter.upload(hFirst, s);
  {
    const std::vector<double> f = lookaheadFactors(*cfg, 1U << 20);
    b->factors.upload(f, s);
    b->cfg.factors = b->factors.p;
    b->cfg.nFactors = static_cast<uint32_t>(f.size());
  }
  b->splittable.upload(hSplit, s);
  b->h2dBytes = hG2Off.size() * 4 + hG2.size() * 8 + hG1Off.size() * 4 +
                hG1.size() * 4 + hConsOff.size() * 4 + hCons.size() * 2 +
                (hTwoQ.size() + hFirst.size()) * 4 + hSplit.size();

  // ---- per-instance inputs / outputs
  b->nLayers.resize(nInst);
  b->layerOff.resize(nInst);
  b->swapCap.resize(nInst);
  b->swapOff.resize(nInst);
  std::vector<int16_t> hInit(nInst * static_cast<size_t>(n), -1);
  for (size_t i = 0; i < nInst; ++i) {
    const auto& c = circs[circOf[i]];
    b->nLayers[i] = c.nLayers;
    b->layerOff[i] = static_cast<uint32_t>(b->totLayers);
    b->totLayers += c.nLayers;
    uint32_t cap = 4U * c.nLayers + 256U;
    if (capacities != nullptr && capacities[i].swaps != nullptr) {
      cap = capacities[i].swap_capacity;
    }
    b->swapCap[i] = cap;
    b->swapOff[i] = static_cast<uint32_t>(b->totSwapCap);
    b->totSwapCap += cap;
    if (capacities != nullptr && capacities[i].layer_qubits != nullptr) {
      b->wantLayerQ = true;
    }
    if (instances[i].initial_locations != nullptr) {
      for (int q = 0; q < n; ++q) {
        const int16_t p = instances[i].initial_locations[q];
        if (p < -1 || p >= n) {
          return fail(QMG_BAD_ARG, "initial location out of range");
        }
        hInit[i * n + q] = p;
      }
    }
  }
  if (b->totLayers > 0xFFFFFFF0ULL || b->totSwapCap > 0x7FFFFFF0ULL) {
    return fail(QMG_BAD_ARG, "batch too large");
  }
  b->initLoc.upload(hInit, s);
  b->h2dBytes += hInit.size() * 2;
  b->oLayers.alloc(std::max<uint64_t>(b->totLayers, 1));
  b->oSwaps.alloc(2 * std::max<uint64_t>(b->totSwapCap, 1));
  b->oPlaces.alloc(2 * std::max<size_t>(nInst * n, 1));
  b->oFinalQ.alloc(std::max<size_t>(nInst * n, 1));
  b->oFinalL.alloc(std::max<size_t>(nInst * n, 1));
  b->oStartL.alloc(std::max<size_t>(nInst * n, 1));
  if (b->wantLayerQ) {
    b->oLayerQ.alloc(std::max<uint64_t>(b->totLayers * n, 1));
  }
  b->oSummary.alloc(8 * std::max<size_t>(nInst, 1));
  b->oCounters.alloc(2 * std::max<size_t>(nInst, 1));

  std::vector<InstView> hInst(nInst);
  for (size_t i = 0; i < nInst; ++i) {
    const auto& c = circs[circOf[i]];
    InstView& I = hInst[i];
    I.circ.nLayers = c.nLayers;
    I.circ.nLogical = c.nLogical;
    I.circ.g2Off = b->g2Off.p + c.g2Off;
    I.circ.g2 = b->g2.p + c.g2;
    I.circ.g1Off = b->g1Off.p + c.g1Off;
    I.circ.g1 = b->g1.p + c.g1;
    I.circ.consOff = b->consOff.p + c.consOff;
    I.circ.cons = b->cons.p + c.cons;
    I.circ.twoQ = b->twoQ.p + twoQPos[circOf[i]].first;
    I.circ.nTwoQ = static_cast<uint32_t>(twoQPos[circOf[i]].second);
    I.circ.firstAfter = b->firstAfter.p + c.layer;
    I.circ.splittable = b->splittable.p + c.layer;
    // a null initial layout means "all unplaced"; the staged copy holds -1s
    I.initLoc = b->initLoc.p + i * n;
    I.passes = instances[i].bidirectional_passes;
    I.firstLayer = instances[i].first_layer;
    I.nRoute = instances[i].n_route_layers;
    I.layers = b->oLayers.p + b->layerOff[i];
    I.swaps = b->oSwaps.p + 2 * static_cast<size_t>(b->swapOff[i]);
    I.swapCap = b->swapCap[i];
    I.places = b->oPlaces.p + 2 * i * n;
    I.placeCap = static_cast<uint32_t>(n);
    I.finalQ = b->oFinalQ.p + i * n;
    I.finalL = b->oFinalL.p + i * n;
    I.startL = b->oStartL.p + i * n;
    I.layerQ = b->wantLayerQ
                   ? b->oLayerQ.p + static_cast<size_t>(b->layerOff[i]) * n
                   : nullptr;
    I.summary = b->oSummary.p + 8 * i;
    I.counters = b->oCounters.p + 2 * i;
  }
  b->inst.upload(hInst, s);
  b->h2dBytes += hInst.size() * sizeof(InstView);
  // largest instances first
  std::vector<uint32_t> hOrder(nInst);
  std::iota(hOrder.begin(), hOrder.end(), 0U);
  std::stable_sort(hOrder.begin(), hOrder.end(), [&](uint32_t x, uint32_t y) {
    const uint64_t wx = static_cast<uint64_t>(b->nLayers[x]) *
                        (1 + 2 * instances[x].bidirectional_passes);
    const uint64_t wy = static_cast<uint64_t>(b->nLayers[y]) *
                        (1 + 2 * instances[y].bidirectional_passes);
    return wx > wy;
  });
  b->order.upload(hOrder, s);
  b->ticket.alloc(1);

  // ---- launch geometry and working memory
  cudaDeviceProp prop{};
  CU(cudaGetDeviceProperties(&prop, arch->device));
  b->sl = smemLayout(n);
  b->smemBytes = static_cast<size_t>(b->sl.bytesPerWarp) * kWarpsPerCta;
  if (b->smemBytes > prop.sharedMemPerBlockOptin) {
    return fail(QMG_BAD_ARG, "architecture too large for the shared-memory "
                             "working set of one search");
  }
  uint32_t warpsPerSm = (opt != nullptr && opt->searches_per_sm != 0)
                            ? opt->searches_per_sm
                            : 32U;
  warpsPerSm = std::max<uint32_t>(kWarpsPerCta, warpsPerSm / kWarpsPerCta *
                                                    kWarpsPerCta);
  const size_t smemPerSm = prop.sharedMemPerMultiprocessor;
  while (warpsPerSm > kWarpsPerCta &&
         (b->smemBytes + 1024) * (warpsPerSm / kWarpsPerCta) > smemPerSm) {
    warpsPerSm -= kWarpsPerCta;
  }
  const uint32_t maxCtas =
      static_cast<uint32_t>(prop.multiProcessorCount) * (warpsPerSm / kWarpsPerCta);
  const uint32_t needCtas = static_cast<uint32_t>(
      std::max<size_t>(1, (nInst + kWarpsPerCta - 1) / kWarpsPerCta));
  b->grid = std::min(maxCtas, needCtas);
  const uint32_t nWarps = b->grid * kWarpsPerCta;
  const uint32_t recBytes =
      static_cast<uint32_t>((sizeof(NodeHdr) + 4U * n + 7U) & ~7U);
  uint32_t poolNodes = (opt != nullptr) ? opt->pool_nodes : 0U;
  if (poolNodes == 0) {
    size_t freeB = 0, totalB = 0;
    CU(cudaMemGetInfo(&freeB, &totalB));
    const size_t budget = std::min<size_t>(freeB / 2, 24ULL << 30);
    const size_t perNode = recBytes + sizeof(OpenEntry) + 4 * sizeof(uint64_t);
    size_t pn = budget / (static_cast<size_t>(nWarps) * perNode);
    pn = std::min<size_t>(pn, 1U << 18);
    pn = std::max<size_t>(pn, 1024);
    poolNodes = static_cast<uint32_t>(pn);
  }
  b->poolNodes = poolNodes;
  uint32_t tableSize = 1;
  while (tableSize < 2 * poolNodes) {
    tableSize <<= 1;
  }
  b->pool.alloc(static_cast<size_t>(nWarps) * poolNodes * recBytes);
  b->open.alloc(static_cast<size_t>(nWarps) * poolNodes);
  b->table.alloc(static_cast<size_t>(nWarps) * tableSize);
  CU(cudaMemsetAsync(b->table.p, 0,
                     static_cast<size_t>(nWarps) * tableSize * sizeof(uint64_t),
                     s));
  b->nsw.alloc(static_cast<size_t>(nWarps) * 32 * maxG);
  std::vector<SearchMem> hMem(nWarps);
  for (uint32_t w = 0; w < nWarps; ++w) {
    SearchMem& M = hMem[w];
    M.pool = b->pool.p + static_cast<size_t>(w) * poolNodes * recBytes;
    M.open = b->open.p + static_cast<size_t>(w) * poolNodes;
    M.table = b->table.p + static_cast<size_t>(w) * tableSize;
    M.nsw = b->nsw.p + static_cast<size_t>(w) * 32 * maxG;
    M.poolCap = poolNodes;
    M.recBytes = recBytes;
    M.tableMask = tableSize - 1;
    M.stamp = 0;
    M.maxG = maxG;
  }
  b->mem.upload(hMem, s);
  {
    const int sb = static_cast<int>(b->smemBytes);
    const auto attr = cudaFuncAttributeMaxDynamicSharedMemorySize;
    CU(cudaFuncSetAttribute(route_kernel<0>, attr, sb));
    CU(cudaFuncSetAttribute(route_kernel<1>, attr, sb));
    CU(cudaFuncSetAttribute(route_kernel<2>, attr, sb));
    CU(cudaFuncSetAttribute(route_kernel<3>, attr, sb));
    CU(cudaFuncSetAttribute(route_kernel<4>, attr, sb));
  }
  b->h2dMs = tm.stop();
  *out = b.release();
  return QMG_OK;
}

int batchRun(qmg_batch* b, qmg_route_stats* stats) {
  if (int rc = selectDevice(b->arch->device)) {
    return rc;
  }
  cudaStream_t s = b->arch->stream;
  CU(cudaMemsetAsync(b->ticket.p, 0, sizeof(uint32_t), s));
  EventTimer tm(s);
  tm.start();
  if (b->nInst > 0) {
    const ArchView av = b->arch->view();
    const uint32_t nI = static_cast<uint32_t>(b->nInst);
    const dim3 blk(kWarpsPerCta * 32);
#define QMG_LAUNCH(HH)                                                         \
  route_kernel<HH><<<b->grid, blk, b->smemBytes, s>>>(                         \
      av, b->cfg, b->inst.p, b->order.p, nI, b->mem.p, b->ticket.p, b->sl)
    switch (b->cfg.heuristic) {
    case 0:
      QMG_LAUNCH(0);
      break;
    case 1:
      QMG_LAUNCH(1);
      break;
    case 2:
      QMG_LAUNCH(2);
      break;
    case 3:
      QMG_LAUNCH(3);
      break;
    default:
      QMG_LAUNCH(4);
      break;
    }
#undef QMG_LAUNCH
    CU(cudaGetLastError());
    ++b->launches;
  }
  const float ms = tm.stop();
  CU(cudaStreamSynchronize(s));
  if (stats != nullptr) {
    stats->kernel_ms = ms;
    stats->h2d_ms = b->h2dMs;
    stats->grid = b->grid;
    stats->block = kWarpsPerCta * 32;
    stats->h2d_bytes = b->h2dBytes;
    stats->kernel_launches = b->launches;
  }
  return QMG_OK;
}

int batchFetch(qmg_batch* b, qmg_route_result* results, qmg_route_stats* stats) {
  if (int rc = selectDevice(b->arch->device)) {
    return rc;
  }
  cudaStream_t s = b->arch->stream;
  const size_t nI = b->nInst;
  const size_t n = static_cast<size_t>(b->n);
  EventTimer tm(s);
  tm.start();
  std::vector<uint32_t> summary(8 * nI);
  std::vector<uint64_t> counters(2 * nI);
  uint64_t bytes = 0;
  auto d2h = [&](void* dst, const void* src, size_t nbytes) {
    if (nbytes > 0 && dst != nullptr) {
      CU(cudaMemcpyAsync(dst, src, nbytes, cudaMemcpyDeviceToHost, s));
      bytes += nbytes;
    }
  };
  d2h(summary.data(), b->oSummary.p, summary.size() * 4);
  d2h(counters.data(), b->oCounters.p, counters.size() * 8);
  CU(cudaStreamSynchronize(s));
  uint64_t exp = 0, gen = 0;
  for (size_t i = 0; i < nI; ++i) {
    qmg_route_result& r = results[i];
    r.n_swaps_total = summary[8 * i];
    r.n_placements_total = summary[8 * i + 1];
    r.n_layers_done = summary[8 * i + 2];
    r.status = static_cast<int32_t>(summary[8 * i + 3]);
    r.expanded_nodes = counters[2 * i];
    r.generated_nodes = counters[2 * i + 1];
    exp += r.expanded_nodes;
    gen += r.generated_nodes;
    d2h(r.layers, b->oLayers.p + b->layerOff[i],
        static_cast<size_t>(b->nLayers[i]) * sizeof(LayerOut));
    const uint32_t ns = std::min(r.n_swaps_total, b->swapCap[i]);
    if (r.swaps != nullptr && r.swap_capacity >= ns) {
      d2h(r.swaps, b->oSwaps.p + 2 * static_cast<size_t>(b->swapOff[i]),
          static_cast<size_t>(ns) * 4);
    } else if (r.swaps != nullptr && r.status == QMG_OK) {
      r.status = QMG_BUFFER_TOO_SMALL;
    }
    const uint32_t np =
        std::min<uint32_t>(r.n_placements_total, static_cast<uint32_t>(n));
    if (r.placements != nullptr && r.placement_capacity >= np) {
      d2h(r.placements, b->oPlaces.p + 2 * i * n, static_cast<size_t>(np) * 4);
    }
    d2h(r.final_qubits, b->oFinalQ.p + i * n, n * 2);
    d2h(r.final_locations, b->oFinalL.p + i * n, n * 2);
    d2h(r.start_locations, b->oStartL.p + i * n, n * 2);
    if (b->wantLayerQ) {
      d2h(r.layer_qubits, b->oLayerQ.p + static_cast<size_t>(b->layerOff[i]) * n,
          static_cast<size_t>(b->nLayers[i]) * n * 2);
    }
  }
  const float ms = tm.stop();
  if (stats != nullptr) {
    stats->d2h_ms = ms;
    stats->d2h_bytes = bytes;
    stats->expanded_nodes = exp;
    stats->generated_nodes = gen;
  }
  return QMG_OK;
}

}  // namespace

QMG_API int qmg_batch_create(const qmg_arch* arch, const qmg_config* cfg,
                             size_t nInstances, const qmg_instance* instances,
                             const qmg_route_options* options,
                             qmg_batch** out) {
  return guarded([&] {
    return batchCreate(arch, cfg, nInstances, instances, options, nullptr, out);
  });
}
QMG_API int qmg_batch_run(qmg_batch* batch, qmg_route_stats* stats) {
  if (batch == nullptr) {
    return fail(QMG_BAD_ARG, "null batch");
  }
  return guarded([&] { return batchRun(batch, stats); });
}
QMG_API int qmg_batch_fetch(qmg_batch* batch, qmg_route_result* results) {
  if (batch == nullptr || results == nullptr) {
    return fail(QMG_BAD_ARG, "null argument");
  }
  return guarded([&] { return batchFetch(batch, results, nullptr); });
}
QMG_API void qmg_batch_destroy(qmg_batch* batch) {
  if (batch != nullptr) {
    cudaSetDevice(batch->arch->device);
    delete batch;
  }
}

QMG_API int qmg_route_batch(const qmg_arch* arch, const qmg_config* cfg,
                            size_t nInstances, const qmg_instance* instances,
                            qmg_route_result* results,
                            const qmg_route_options* options,
                            qmg_route_stats* stats) {
  if (results == nullptr && nInstances > 0) {
    return fail(QMG_BAD_ARG, "null results");
  }
  return guarded([&] {
    qmg_batch* b = nullptr;
    if (int rc = batchCreate(arch, cfg, nInstances, instances, options, results,
                             &b)) {
      return rc;
    }
    std::unique_ptr<qmg_batch> hold(b);
    if (stats != nullptr) {
      std::memset(stats, 0, sizeof(*stats));
    }
    if (int rc = batchRun(b, stats)) {
      return rc;
    }
    return batchFetch(b, results, stats);
  });
}
