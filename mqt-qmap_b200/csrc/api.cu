// api.cu - the C ABI of include/qmap_b200.h on top of the sm_100a kernels
// (tables.cuh, route.cuh).  No CPU fallback: without a CUDA device every
// compute entry point returns QMG_NO_DEVICE.
#define QMG_HOST_API_IMPL
#include "host_api.hpp"
#include "route_launch.hpp"
#include "staging.hpp"
#include "tables.cuh"

#include <algorithm>
#include <cstdio>
#include <map>
#include <memory>
#include <mutex>
#include <numeric>
#include <string>
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
  DevBuf<uint32_t> incOff, inc;
  DevBuf<uint8_t> adj;
  void upload(const ArchTopology& t) {
    edges.upload(t.edges);
    incOff.upload(t.incOff);
    inc.upload(t.inc);
    adj.upload(t.adj);
  }
};

// Dijkstra::buildTable on the device; w and out are n*n device matrices
// exact: replay the reference's Dijkstra (required for non-uniform weights,
// where its table is not the table of shortest paths, see tables.cuh K1');
// otherwise blocked Floyd-Warshall (equal to the reference for uniform weights)
void devBuildTable(int n, int nE, const TopoDev& td, const double* w,
                   bool exact, double* out, cudaStream_t s) {
  if (n == 0) {
    return;
  }
  if (exact) {
    // the reference's own Dijkstra, heap operation by heap operation
    const int heapCap = 2 * nE + 2;
    DevBuf<uint8_t> visited;
    DevBuf<uint16_t> heap;
    DevBuf<int> overflow;
    visited.alloc(static_cast<size_t>(n) * n);
    heap.alloc(static_cast<size_t>(n) * heapCap);
    overflow.alloc(1);
    CU(cudaMemsetAsync(overflow.p, 0, sizeof(int), s));
    dijkstra_reference_kernel<<<(n + 31) / 32, 32, 0, s>>>(
        n, w, td.edges.p, td.incOff.p, td.inc.p, out, visited.p, heap.p,
        heapCap, overflow.p);
    CU(cudaGetLastError());
    int ov = 0;
    CU(cudaMemcpyAsync(&ov, overflow.p, sizeof(int), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));  // (scratch is freed on return)
    if (ov != 0) {
      throw CudaFail{QMG_CUDA_ERROR, "Dijkstra heap capacity exceeded"};
    }
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

// Dijkstra::buildEdgeSkipTable; returns the number of tables the reference
// produces.  At most maxTables of them are kept (device, n*n each, in order);
// level k needs all of T_0..T_{k-1}, so the build itself always holds every
// level and the surplus is dropped at the end.
size_t devEdgeSkipTables(int n, const TopoDev& td, int nE, const double* w,
                         bool exact, DevBuf<double>& tables, size_t maxTables,
                         cudaStream_t s) {
  const size_t nn = static_cast<size_t>(n) * n;
  (void)maxTables;
  // the reference tries k = 1..n and stops at the first all-zero level
  const size_t cap = static_cast<size_t>(n) + 1;
  size_t alloc = std::min<size_t>(cap, 32);
  tables.alloc(alloc * nn);
  devBuildTable(n, nE, td, w, exact, tables.p, s);
  DevBuf<int> flag;
  flag.alloc(1);
  size_t K = 1;
  for (size_t k = 1; k <= static_cast<size_t>(n); ++k) {
    if (K + 1 > alloc) {
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
    devBuildTable(n, t.nE, td, w.p, exact_association != 0, out.p, nullptr);
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
    // *nTables is the true count; a caller whose buffer was smaller gets the
    // first maxTables tables and QMG_BUFFER_TOO_SMALL
    *nTables = K;
    CU(cudaMemcpy(tables, tb.p, std::min(K, maxTables) * nn * sizeof(double),
                  cudaMemcpyDeviceToHost));
    if (K > maxTables) {
      return fail(QMG_BUFFER_TOO_SMALL, "edge-skip tables: buffer holds fewer "
                                        "tables than the build produced");
    }
    return static_cast<int>(QMG_OK);
  });
}

// ==========================================================================
// Table build without any allocation (no counterpart in the reference; added
// for callers that cannot call
// cudaMalloc: CUDA graph capture, tight serving loops, harnesses that forbid
// it).  The Floyd-Warshall kernels (K1) need one padded np x np matrix of
// scratch; here it is a file-scope device array of fixed size that every GPU
// loading the library holds until exit.
// ==========================================================================
namespace {

constexpr int kStaticMaxQubits = QMG_STATIC_TABLE_MAX_QUBITS;  // qmap_b200.h
static_assert(kStaticMaxQubits % kFwTile == 0, "padded size");
// 512 x 512 doubles = 2 MiB per device.  Shared by all calls: a call rewrites
// every cell it reads (fw_static_init), so nothing has to be cleared, but two
// calls whose kernels overlap on one GPU would race on it.
static __device__ __align__(256) double
    gStaticFwScratch[static_cast<size_t>(kStaticMaxQubits) * kStaticMaxQubits];

// padded matrix: 0 on the diagonal, +inf elsewhere
__global__ void fw_static_init(double* __restrict__ d, int np) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<size_t>(np) * np) {
    return;
  }
  const int r = static_cast<int>(idx / np);
  const int c = static_cast<int>(idx % np);
  d[idx] = (r == c) ? 0. : __longlong_as_double(0x7ff0000000000000LL);
}
// both directions of every coupling entry get their weight (the undirected
// view of utils.cpp:67-73); entries out of range are ignored
__global__ void fw_static_edges(double* __restrict__ d, int np, int n,
                                const uint16_t* __restrict__ edges,
                                size_t nEdges, const double* __restrict__ w) {
  const size_t e = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (e >= nEdges) {
    return;
  }
  const int a = edges[2 * e];
  const int b = edges[2 * e + 1];
  if (a >= n || b >= n || a == b) {
    return;
  }
  d[static_cast<size_t>(a) * np + b] = w[static_cast<size_t>(a) * n + b];
  d[static_cast<size_t>(b) * np + a] = w[static_cast<size_t>(b) * n + a];
}

constexpr int kMaxDevices = 64;
std::mutex gStaticMutex;
double* gStaticScratchAddr[kMaxDevices] = {};

// the device address of the scratch array on `device` (which is current);
// looked up once per device - in host code the array's name is only a
// stand-in, never a device address
int staticScratch(int device, double** out) {
  if (device < 0 || device >= kMaxDevices) {
    return fail(QMG_BAD_ARG, "device index out of range");
  }
  std::lock_guard<std::mutex> lk(gStaticMutex);
  if (gStaticScratchAddr[device] == nullptr) {
    void* p = nullptr;
    const cudaError_t e = cudaGetSymbolAddress(&p, gStaticFwScratch);
    if (e != cudaSuccess || p == nullptr) {
      cudaGetLastError();
      return fail(QMG_CUDA_ERROR, std::string("cudaGetSymbolAddress: ") +
                                      cudaGetErrorString(e));
    }
    gStaticScratchAddr[device] = static_cast<double*>(p);
  }
  *out = gStaticScratchAddr[device];
  return QMG_OK;
}

}  // namespace

QMG_API int qmg_build_table_static(uint16_t n, const uint16_t* dEdges,
                                   size_t nEdges, const double* dEdgeWeights,
                                   double* dDistanceTable, int device,
                                   void* stream) {
  if (n > kStaticMaxQubits) {
    return fail(QMG_BAD_ARG,
                "qmg_build_table_static: n = " + std::to_string(n) +
                    " exceeds the static scratch (at most " +
                    std::to_string(kStaticMaxQubits) +
                    " qubits); use qmg_build_table");
  }
  if ((nEdges > 0 && dEdges == nullptr) ||
      (n > 0 && (dEdgeWeights == nullptr || dDistanceTable == nullptr))) {
    return fail(QMG_BAD_ARG, "null argument");
  }
  if (int rc = selectDevice(device)) {
    return rc;
  }
  if (n == 0) {
    return QMG_OK;
  }
  double* pad = nullptr;
  if (int rc = staticScratch(device, &pad)) {
    return rc;
  }
  cudaStream_t s = static_cast<cudaStream_t>(stream);
  const int nb = (n + kFwTile - 1) / kFwTile;
  const int np = nb * kFwTile;
  const size_t tot = static_cast<size_t>(np) * np;
  fw_static_init<<<static_cast<unsigned>((tot + 255) / 256), 256, 0, s>>>(pad,
                                                                          np);
  if (nEdges > 0) {
    fw_static_edges<<<static_cast<unsigned>((nEdges + 255) / 256), 256, 0, s>>>(
        pad, np, n, dEdges, nEdges, dEdgeWeights);
  }
  for (int kb = 0; kb < nb; ++kb) {
    fw_phase1<<<1, 256, 0, s>>>(pad, np, kb);
    if (nb > 1) {
      fw_phase2<<<dim3(nb - 1, 2), 256, 0, s>>>(pad, np, kb);
      fw_phase3<<<dim3(nb - 1, nb - 1), 256, 0, s>>>(pad, np, kb);
    }
  }
  fw_finish<<<static_cast<unsigned>((static_cast<size_t>(n) * n + 255) / 256),
              256, 0, s>>>(pad, np, n, dDistanceTable);
  const cudaError_t e = cudaGetLastError();
  if (e != cudaSuccess) {
    return fail(QMG_CUDA_ERROR, cudaGetErrorString(e));
  }
  return QMG_OK;
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
  // buffers of the last routing batch, reused by the next one (batch.cuh)
  mutable std::unique_ptr<Workspace> wsCache;
  DeviceLimits limits;
  ~qmg_arch() {
    wsCache.reset();
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
    {
      cudaDeviceProp prop{};
      CU(cudaGetDeviceProperties(&prop, device));
      a->limits.multiProcessorCount = prop.multiProcessorCount;
      a->limits.sharedMemPerBlockOptin = prop.sharedMemPerBlockOptin;
      a->limits.sharedMemPerMultiprocessor = prop.sharedMemPerMultiprocessor;
    }
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
    // uniform SWAP costs (all edges bidirectional, or all unidirectional):
    // Floyd-Warshall; mixed 30/34: the reference's Dijkstra replayed
    const bool uniform = t.bidirectional || t.unidirectional;
    devBuildTable(n, t.nE, a->td, dw.p, !uniform, simple.p, a->stream);
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
#include "batch.cuh"
