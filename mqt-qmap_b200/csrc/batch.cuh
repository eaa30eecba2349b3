// batch.cuh - routing batches: staging of the caller's host buffers into one
// pinned input arena (one H2D copy), launch of route_kernel<H>, and the
// two-phase D2H of the results (summary first, then only the bytes in use).
// Included by api.cu after the qmg_arch definition.
#pragma once

#include <chrono>
#include <cstdlib>
#include <memory>
#include <mutex>
#include <unordered_map>

#include "derive.cuh"
#include "staging.hpp"

using qmg_workspace = qmg::Workspace;

struct qmg_batch {
  const qmg_arch* arch = nullptr;
  CfgView cfg{};
  size_t nInst = 0;
  int n = 0;
  bool wantLayers = false, wantLayerQ = false, wantLayerNS = false;
  bool wantNodeLog = false, wantPlaceLog = false;
  size_t oPlaceLog = 0;
  std::vector<uint32_t> logCap;
  std::vector<uint64_t> logOff;
  uint64_t totLogCap = 0;
  size_t oNodeLog = 0, oRootQ = 0;
  // nLayers: layers of the instance's circuit; layerCap: rows of its
  // per-layer outputs (more when the device may split layers)
  std::vector<uint32_t> nLayers, layerCap, swapCap;
  std::vector<uint64_t> layerOff, swapOff, splitOff;
  uint64_t totLayers = 0, totSwapCap = 0, totSplitCap = 0;
  bool deviceSplits = false;
  std::unique_ptr<qmg_workspace> ws;
  // output arena offsets (bytes)
  size_t oSummary = 0, oCounters = 0, oFinalQ = 0, oFinalL = 0, oStartL = 0,
         oPlaces = 0, oSplitLog = 0, oSmallEnd = 0, oLayerNS = 0, oLayers = 0, oLayerQ = 0,
         oSwapsPacked = 0, oSwapsWork = 0, oSwapCounter = 0, outSize = 0;
  // input arena offsets of what the launch needs
  size_t iInst = 0, iOrder = 0, iMem = 0, iTicket = 0, inSize = 0;
  size_t tableOff = 0, tableBytes = 0;
  uint32_t grid = 0, poolNodes = 0, nBig = 0, bigNodes = 0;
  size_t bigTableOff = 0, bigTableBytes = 0, iBig = 0;
  uint32_t startFull = 0;  // route_kernel: full search variant from the start
  Scratch sl{};
  size_t smemBytes = 0;
  uint64_t h2dBytes = 0;
  float h2dMs = 0.F;
  uint32_t launches = 0;
  ~qmg_batch();
};

namespace {

std::mutex gWsMutex;

// QMAP_B200_TRACE=1 prints the host-side stage times of every batch
bool traceOn() {
  static const bool on = std::getenv("QMAP_B200_TRACE") != nullptr;
  return on;
}
struct StageClock {
  std::chrono::steady_clock::time_point t = std::chrono::steady_clock::now();
  std::string log;
  void mark(const char* what) {
    if (!traceOn()) {
      return;
    }
    const auto now = std::chrono::steady_clock::now();
    char buf[64];
    std::snprintf(buf, sizeof(buf), " %s=%.2fms", what,
                  std::chrono::duration<double, std::milli>(now - t).count());
    log += buf;
    t = now;
  }
  void flush(const char* who) {
    if (traceOn()) {
      std::fprintf(stderr, "[qmap_b200] %s:%s\n", who, log.c_str());
    }
  }
};

std::unique_ptr<qmg_workspace> takeWorkspace(const qmg_arch* a) {
  std::lock_guard<std::mutex> lk(gWsMutex);
  if (a->wsCache) {
    return std::move(a->wsCache);
  }
  return std::make_unique<qmg_workspace>();
}
void giveWorkspace(const qmg_arch* a, std::unique_ptr<qmg_workspace> ws) {
  std::lock_guard<std::mutex> lk(gWsMutex);
  if (!a->wsCache) {
    a->wsCache = std::move(ws);
  }
}

struct CircKey {
  const void* a;
  const void* b;
  const void* c;
  const void* d;
  uint32_t nl, nLogical;
  bool operator==(const CircKey& o) const {
    return a == o.a && b == o.b && c == o.c && d == o.d && nl == o.nl &&
           nLogical == o.nLogical;
  }
};
struct CircKeyHash {
  size_t operator()(const CircKey& k) const {
    uint64_t h = reinterpret_cast<uintptr_t>(k.a) * 0x9E3779B97F4A7C15ULL;
    h ^= reinterpret_cast<uintptr_t>(k.b) + 0xBF58476D1CE4E5B9ULL + (h << 6) +
         (h >> 2);
    h ^= reinterpret_cast<uintptr_t>(k.c) * 0x94D049BB133111EBULL + (h >> 29);
    return static_cast<size_t>(h ^ (h >> 32) ^ k.nl ^ (k.nLogical * 0x9E3779B9U));
  }
};

}  // namespace

qmg_batch::~qmg_batch() {
  if (ws && arch != nullptr) {
    giveWorkspace(arch, std::move(ws));
  }
}

namespace {

template <class T> T* at(uint8_t* base, size_t off) {
  return reinterpret_cast<T*>(base + off);
}

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
  if (cfg->nr_lookaheads > (1U << 20)) {
    return fail(QMG_BAD_ARG, "nr_lookaheads too large");
  }
  if (int rc = selectDevice(arch->device)) {
    return rc;
  }
  auto b = std::make_unique<qmg_batch>();
  b->arch = arch;
  b->cfg = toCfgView(*cfg);
  b->nInst = nInst;
  const int n = arch->topo.n;
  b->n = n;
  b->ws = takeWorkspace(arch);
  qmg_workspace& W = *b->ws;
  cudaStream_t s = arch->stream;
  const auto t0 = std::chrono::steady_clock::now();
  const bool fid = b->cfg.fidelityAware != 0;
  StageClock sc;

  // ---- unique circuits (instances that point at the same layers share one):
  // open-addressing table keyed by the caller's array pointers
  std::vector<CircSlot> circs;
  std::vector<const qmg_layers*> circSrc;
  std::vector<CircKey> circKey;
  std::vector<uint32_t> circOf(nInst);
  size_t tabSize = 64;
  while (tabSize < 2 * nInst) {
    tabSize <<= 1;
  }
  std::vector<uint32_t> tab(tabSize, 0);  // circuit index + 1
  for (size_t i = 0; i < nInst; ++i) {
    const qmg_layers& L = instances[i].layers;
    if (L.n_layers > 0 &&
        (L.gate2q_offsets == nullptr ||
         (L.gate1q_offsets != nullptr && L.gates1q == nullptr &&
          L.gate1q_offsets[L.n_layers] > 0))) {
      return fail(QMG_BAD_ARG, "instance without layers");
    }
    const CircKey key{L.gate2q_offsets, L.gates2q, L.gate1q_offsets, L.gates1q,
                      L.n_layers, L.n_logical};
    size_t slot = CircKeyHash()(key) & (tabSize - 1);
    uint32_t hit = 0;
    while ((hit = tab[slot]) != 0 && !(circKey[hit - 1] == key)) {
      slot = (slot + 1) & (tabSize - 1);
    }
    if (hit == 0) {
      CircSlot c;
      c.nLayers = L.n_layers;
      c.nLogical = L.n_logical;
      c.nG2 = L.n_layers > 0 ? L.gate2q_offsets[L.n_layers] : 0;
      c.nG1 = (L.n_layers > 0 && L.gate1q_offsets != nullptr)
                  ? L.gate1q_offsets[L.n_layers]
                  : 0;
      if (c.nG2 > 0 && L.gates2q == nullptr) {
        return fail(QMG_BAD_ARG, "instance without layers");
      }
      circs.push_back(c);
      circSrc.push_back(&L);
      circKey.push_back(key);
      hit = static_cast<uint32_t>(circs.size());
      tab[slot] = hit;
    }
    circOf[i] = hit - 1;
  }

  sc.mark("dedupe");
  // ---- per-instance sizes
  b->nLayers.resize(nInst);
  b->layerCap.resize(nInst);
  b->splitOff.resize(nInst);
  b->layerOff.resize(nInst);
  b->swapCap.resize(nInst);
  b->swapOff.resize(nInst);
  b->logCap.assign(nInst, 0);
  b->logOff.assign(nInst, 0);
  for (size_t i = 0; i < nInst; ++i) {
    const auto& c = circs[circOf[i]];
    b->nLayers[i] = c.nLayers;
    b->layerCap[i] = std::max(c.nLayers, instances[i].layer_capacity);
    b->deviceSplits = b->deviceSplits || b->layerCap[i] > c.nLayers;
    b->layerOff[i] = b->totLayers;
    b->totLayers += b->layerCap[i];
    b->splitOff[i] = b->totSplitCap;
    b->totSplitCap += b->layerCap[i] - c.nLayers;
    uint32_t cap = 4U * b->layerCap[i] + 256U;
    if (capacities != nullptr) {
      if (capacities[i].swaps != nullptr) {
        cap = capacities[i].swap_capacity;
      }
      b->wantLayers = b->wantLayers || capacities[i].layers != nullptr;
      b->wantLayerQ = b->wantLayerQ || capacities[i].layer_qubits != nullptr;
      b->wantLayerNS = b->wantLayerNS || capacities[i].layer_n_swaps != nullptr;
      b->wantPlaceLog = b->wantPlaceLog || capacities[i].placement_log != nullptr;
      if (capacities[i].node_log != nullptr && capacities[i].node_log_capacity > 0) {
        b->wantNodeLog = true;
        b->logCap[i] = capacities[i].node_log_capacity;
      }
    }
    b->logOff[i] = b->totLogCap;
    b->totLogCap += b->logCap[i];
    b->swapCap[i] = cap;
    b->swapOff[i] = b->totSwapCap;
    b->totSwapCap += cap;
  }
  if (capacities == nullptr) {
    // device-resident batches keep the full records unless told otherwise
    b->wantLayers = opt == nullptr || opt->lean_outputs == 0;
  }
  if (b->totLayers > 0xFFFFFFF0ULL || b->totSwapCap > 0x3FFFFFF0ULL) {
    return fail(QMG_BAD_ARG, "batch too large");
  }
  if (capacities == nullptr) {
    b->wantLayerNS = true;
  }

  sc.mark("sizes");
  // ---- launch geometry
  const DeviceLimits& prop = arch->limits;
  b->sl = scratchLayout(n);
  if (b->sl.bytes > kMaxScratchBytes) {
    return fail(QMG_BAD_ARG, "architecture too large for the shared-memory "
                             "working set of one search");
  }
  b->smemBytes = static_cast<size_t>(b->sl.bytes) * kWarpsPerCta;
  if (b->smemBytes > prop.sharedMemPerBlockOptin) {
    return fail(QMG_BAD_ARG, "architecture too large for the shared-memory "
                             "working set of one search");
  }
  uint32_t warpsPerSm = (opt != nullptr && opt->searches_per_sm != 0)
                            ? opt->searches_per_sm
                            : static_cast<uint32_t>(QMG_MIN_CTAS) * kWarpsPerCta;
  // more than the register file admits would only queue idle CTAs
  warpsPerSm = std::min<uint32_t>(
      warpsPerSm,
      static_cast<uint32_t>(cfg->heuristic >= 4 ? QMG_MIN_CTAS_FID : QMG_MIN_CTAS) *
          kWarpsPerCta);
  warpsPerSm = std::max<uint32_t>(kWarpsPerCta,
                                  warpsPerSm / kWarpsPerCta * kWarpsPerCta);
  while (warpsPerSm > kWarpsPerCta &&
         (b->smemBytes + 1024) * (warpsPerSm / kWarpsPerCta) >
             prop.sharedMemPerMultiprocessor) {
    warpsPerSm -= kWarpsPerCta;
  }
  const uint32_t maxCtas = static_cast<uint32_t>(prop.multiProcessorCount) *
                           (warpsPerSm / kWarpsPerCta);
  const uint32_t needCtas = static_cast<uint32_t>(
      std::max<size_t>(1, (nInst + kWarpsPerCta - 1) / kWarpsPerCta));
  b->grid = std::min(maxCtas, needCtas);
  if (opt != nullptr && opt->max_searches != 0) {
    b->grid = std::max<uint32_t>(
        1, std::min(b->grid, opt->max_searches / kWarpsPerCta));
  }
  const uint32_t nWarps = b->grid * kWarpsPerCta;
  const uint32_t recBytes = nodeRecBytes(n);
  uint32_t poolNodes = (opt != nullptr) ? opt->pool_nodes : 0U;
  uint32_t nBig = 0, bigNodes = 0;
  if (poolNodes == 0) {
    if (W.memBudget == 0) {
      size_t freeB = 0, totalB = 0;
      CU(cudaMemGetInfo(&freeB, &totalB));
      W.memBudget = std::min<size_t>(freeB / 2, 16ULL << 30);
      // QMAP_B200_POOL_MB: the first-tier budget of the node pools, up to 80 %
      // of the free HBM (long searches want most of it; tests use a tiny one
      // to make the pools run out)
      if (const char* e = std::getenv("QMAP_B200_POOL_MB")) {
        const long mb = std::atol(e);
        if (mb > 0) {
          W.memBudget = std::min<size_t>(freeB / 10 * 8, static_cast<size_t>(mb) << 20);
          W.poolCapNodes = size_t{1} << 24;
        }
      }
    }
    const size_t budget = W.memBudget;
    const size_t perNode = recBytes + sizeof(OpenEntry) + 4 * sizeof(uint64_t);
    // Two tiers: every resident search owns a private pool sized for the
    // common case (a share of the budget), the rest of the budget is a few
    // large pools that searches take for the layers that outgrow the private
    // one (BigPool, route_core.hpp).  QMAP_B200_BIG_POOLS / _PRIVATE_PCT tune
    // the split (0 large pools = the whole budget in private pools).
    nBig = (opt != nullptr && opt->big_pools != 0) ? opt->big_pools : 32U;
    size_t privatePct = 40;
    if (const char* e = std::getenv("QMAP_B200_BIG_POOLS")) {
      nBig = static_cast<uint32_t>(std::max(0L, std::atol(e)));
    }
    if (const char* e = std::getenv("QMAP_B200_PRIVATE_PCT")) {
      privatePct = static_cast<size_t>(std::min(100L, std::max(1L, std::atol(e))));
    }
    if (opt != nullptr && opt->big_pools == 0xFFFFFFFFU) {
      nBig = 0;
    }
    nBig = std::min<uint32_t>(nBig, std::max<uint32_t>(1, nWarps / 8));
    size_t pn = budget / (static_cast<size_t>(nWarps) * perNode);
    if (nBig > 0) {
      const size_t pnShare =
          std::max<size_t>(1024, std::min<size_t>(pn * privatePct / 100, W.poolCapNodes));
      const size_t used = pnShare * nWarps * perNode;
      const size_t bn = used < budget ? (budget - used) / (nBig * perNode) : 0;
      if (bn >= 4 * pnShare) {
        pn = pnShare;
        bigNodes = static_cast<uint32_t>(std::min<size_t>(bn, kMaxPoolNodes));
      } else {
        nBig = 0;  // (a budget too small for two tiers)
      }
    }
    pn = std::min<size_t>(pn, W.poolCapNodes);
    pn = std::max<size_t>(pn, 1024);
    poolNodes = static_cast<uint32_t>(std::min<size_t>(pn, kMaxPoolNodes));
  } else if (opt->big_pools != 0 && opt->big_pools != 0xFFFFFFFFU &&
             opt->big_pool_nodes != 0) {
    nBig = opt->big_pools;
    bigNodes = std::min<uint32_t>(opt->big_pool_nodes, kMaxPoolNodes);
  }
  if (nBig == 0 || bigNodes == 0) {
    nBig = 0;
    bigNodes = 0;
  }
  if (poolNodes > kMaxPoolNodes) {
    return fail(QMG_BAD_ARG, "pool_nodes exceeds 2^26 - 1 nodes per search");
  }
  b->poolNodes = poolNodes;
  b->nBig = nBig;
  b->bigNodes = bigNodes;
  {
    const uint32_t v = opt != nullptr ? opt->search_variant : 0U;
    b->startFull = v == 1 || (v == 0 && cfg->heuristic >= 4 && n > 32) ? 1U : 0U;
    if (b->wantNodeLog || b->wantPlaceLog) {
      b->startFull = 1;  // (the logs are written by the full variant)
    }
  }
  uint32_t tableSize = 1;
  while (tableSize < 2 * poolNodes) {
    tableSize <<= 1;
  }

  sc.mark("geometry");
  // ---- input arena layout: circuits first (each contiguous), then the
  // per-instance sections
  Layout li;
  for (auto& c : circs) {
    c.placeUpload(li);
  }
  const size_t circBytes = li.size;
  const size_t sInit = li.add(nInst * n * 2),
               sFactors = li.add(static_cast<size_t>(cfg->nr_lookaheads) * 8);
  const size_t sDerive = li.add(circs.size() * sizeof(DeriveView));
  b->iInst = li.add(nInst * sizeof(InstView));
  b->iOrder = li.add(nInst * 4);
  b->iMem = li.add(static_cast<size_t>(nWarps) * sizeof(SearchMem));
  b->iBig = li.add(static_cast<size_t>(nBig) * sizeof(BigPool));
  b->inSize = li.size;
  // device-only tail of the input arena: the tables derived on the device
  Layout ld;
  ld.size = b->inSize;
  for (auto& c : circs) {
    c.placeDerived(ld);
  }
  const size_t inTotal = ld.size;

  // ---- output arena layout
  Layout lo;
  b->oSwapCounter = lo.add(16);  // [0] packed-SWAP counter, [1] ticket
  b->oSummary = lo.add(nInst * 8 * 4);
  b->oCounters = lo.add(nInst * 4 * 8);
  b->oFinalQ = lo.add(nInst * n * 2);
  b->oFinalL = lo.add(nInst * n * 2);
  b->oStartL = lo.add(nInst * n * 2);
  b->oPlaces = lo.add(nInst * n * 4);
  b->oSplitLog = lo.add(b->totSplitCap * 4);
  b->oPlaceLog = lo.add(b->wantPlaceLog ? nInst * n * 4 : 0);
  b->oSmallEnd = lo.size;
  b->oLayerNS = lo.add(b->wantLayerNS ? b->totLayers * 2 : 0);
  b->oLayers = lo.add(b->wantLayers ? b->totLayers * sizeof(LayerOut) : 0);
  b->oLayerQ = lo.add(b->wantLayerQ ? b->totLayers * n * 2 : 0);
  b->oNodeLog = lo.add(b->wantNodeLog ? b->totLogCap * sizeof(NodeLogRec) : 0);
  b->oRootQ = lo.add(b->wantNodeLog ? b->totLayers * n * 2 : 0);
  b->oSwapsPacked = lo.add(b->totSwapCap * 4);
  const size_t pinnedOut = lo.size;
  b->oSwapsWork = lo.add(b->totSwapCap * 4);
  b->outSize = lo.size;

  // ---- work arena layout
  Layout lw;
  const size_t wPool = lw.add(static_cast<size_t>(nWarps) * poolNodes * recBytes);
  const size_t wOpen =
      lw.add(static_cast<size_t>(nWarps) * poolNodes * sizeof(OpenEntry));
  const size_t wTable = lw.add(static_cast<size_t>(nWarps) * tableSize * 8);
  // memo of the fidelity-aware best-edge term (SearchMem::pairVal); its tags
  // follow the hash tables and are cleared with them
  const size_t pairSlots =
      cfg->heuristic >= 4 ? static_cast<size_t>(kPairGates) * n * n : 0;
  const size_t wPair =
      lw.add(static_cast<size_t>(nWarps) * pairSlots * sizeof(PairSlot));
  const size_t wPairEnd = lw.size;
  const size_t btopPerWarp = poolNodes / kBlk + 2;
  const size_t wBtop =
      lw.add(static_cast<size_t>(nWarps) * btopPerWarp * sizeof(OpenEntry));
  b->tableOff = wTable;
  b->tableBytes = wPairEnd - wTable;
  uint32_t bigTableSize = 1;
  while (nBig > 0 && bigTableSize < 2 * bigNodes) {
    bigTableSize <<= 1;
  }
  const size_t bigBtop = bigNodes / kBlk + 2;
  const size_t wBigPool = lw.add(static_cast<size_t>(nBig) * bigNodes * recBytes);
  const size_t wBigOpen = lw.add(static_cast<size_t>(nBig) * bigNodes * sizeof(OpenEntry));
  const size_t wBigTable = lw.add(nBig > 0 ? static_cast<size_t>(nBig) * bigTableSize * 8 : 0);
  const size_t wBigBtop = lw.add(static_cast<size_t>(nBig) * bigBtop * sizeof(OpenEntry));
  b->bigTableOff = wBigTable;
  b->bigTableBytes = nBig > 0 ? static_cast<size_t>(nBig) * bigTableSize * 8 : 0;

  if (!W.hin.ensure(b->inSize) || !W.hout.ensure(pinnedOut)) {
    return fail(QMG_OOM, "pinned host allocation failed");
  }
  CU(W.in.ensure(inTotal));
  CU(W.out.ensure(b->outSize));
  uint8_t* H = W.hin.p;
  uint8_t* D = W.in.p;

  sc.mark("alloc");
  // ---- flatten the circuits into the pinned arena (host threads); the arena
  // is cut into ranges of circuits and the thread that completes a range
  // starts its host-to-device copy, so the copies overlap the flattening
  const size_t nC = circs.size();
  const size_t nChunks =
      std::min<size_t>(16, std::max<size_t>(1, circBytes >> 22));
  std::vector<uint32_t> chunkOf(nC);
  std::vector<size_t> chunkBegin(nChunks + 1, circBytes);
  std::unique_ptr<std::atomic<uint32_t>[]> chunkLeft(
      new std::atomic<uint32_t>[nChunks]);
  {
    std::vector<uint32_t> cnt(nChunks, 0);
    const size_t per = (circBytes + nChunks - 1) / nChunks;
    size_t k = 0;
    chunkBegin[0] = 0;
    for (size_t ci = 0; ci < nC; ++ci) {
      const size_t at0 = circs[ci].bG2Off;
      while (k + 1 < nChunks && at0 >= (k + 1) * per) {
        ++k;
        chunkBegin[k] = at0;
      }
      chunkOf[ci] = static_cast<uint32_t>(k);
      ++cnt[k];
    }
    for (size_t q = k + 1; q < nChunks; ++q) {
      chunkBegin[q] = circBytes;
    }
    for (size_t q = 0; q < nChunks; ++q) {
      chunkLeft[q].store(cnt[q]);
    }
  }
  std::atomic<int> bad{0}, copyErr{0};
  const int device = arch->device;
  parallelFor(nC, 16, [&](size_t ci) {
    CircSlot& c = circs[ci];
    const qmg_layers& L = *circSrc[ci];
    if (!stageCircuit(n, L.gate2q_offsets,
                      reinterpret_cast<const Gate2*>(L.gates2q),
                      L.gate1q_offsets,
                      reinterpret_cast<const Gate1*>(L.gates1q), c,
                      at<uint32_t>(H, c.bG2Off), at<Gate2>(H, c.bG2),
                      at<uint32_t>(H, c.bG1Off), at<Gate1>(H, c.bG1))) {
      bad.store(1);
    }
    DeriveView& dv = at<DeriveView>(H, sDerive)[ci];
    dv.nLayers = c.nLayers;
    dv.fidelityAware = fid ? 1U : 0U;
    dv.g2Off = at<uint32_t>(D, c.bG2Off);
    dv.g2 = at<Gate2>(D, c.bG2);
    dv.g1Off = at<uint32_t>(D, c.bG1Off);
    dv.g1 = at<Gate1>(D, c.bG1);
    dv.cons = at<uint16_t>(D, c.dCons);
    dv.consCnt = at<uint16_t>(D, c.dConsCnt);
    dv.twoQ = at<uint32_t>(D, c.dTwoQ);
    dv.firstAfter = at<uint32_t>(D, c.dFirst);
    dv.split = at<uint8_t>(D, c.dSplit);
    const uint32_t k = chunkOf[ci];
    if (chunkLeft[k].fetch_sub(1) == 1) {
      const size_t lo = chunkBegin[k], hi = chunkBegin[k + 1];
      if (hi > lo) {
        if (cudaSetDevice(device) != cudaSuccess ||
            cudaMemcpyAsync(D + lo, H + lo, hi - lo, cudaMemcpyHostToDevice,
                            s) != cudaSuccess) {
          copyErr.store(1);
        }
      }
    }
  });
  if (bad.load() != 0) {
    return fail(QMG_BAD_ARG, "malformed layers (offsets or gate operands)");
  }
  if (copyErr.load() != 0) {
    return fail(QMG_CUDA_ERROR, cudaGetErrorString(cudaGetLastError()));
  }
  sc.mark("flatten");
  uint32_t maxG = 1;
  for (const auto& c : circs) {
    maxG = std::max(maxG, c.maxG);
  }
  const size_t wNsw = lw.add(static_cast<size_t>(nWarps) * 32 * maxG * 2);
  // private, writable circuit tables of every resident search (layers split
  // on the device); too large a batch falls back to QMG_SPLIT_LIMIT_HIT
  uint32_t privLayers = 0, privG2 = 0, privG1 = 0;
  size_t privStride = 0, wPriv = 0;
  if (b->deviceSplits) {
    for (size_t i = 0; i < nInst; ++i) {
      privLayers = std::max(privLayers, b->layerCap[i]);
    }
    for (const auto& c : circs) {
      privG2 = std::max<uint32_t>(privG2, static_cast<uint32_t>(c.nG2) + 1U);
      privG1 = std::max<uint32_t>(privG1, static_cast<uint32_t>(c.nG1) + 1U);
    }
    privStride = (privBytes(privLayers, privG2, privG1) + 255U) & ~size_t{255};
    if (privStride * nWarps > (4ULL << 30)) {
      b->deviceSplits = false;
    } else {
      wPriv = lw.add(privStride * nWarps);
    }
  }
  CU(W.work.ensure(lw.size));
  uint8_t* Wk = W.work.p;
  uint8_t* O = W.out.p;

  sc.mark("work-alloc");
  // ---- factors, per-instance inputs and views
  {
    const std::vector<double> f = lookaheadFactors(*cfg, 1U << 20);
    if (!f.empty()) {
      std::memcpy(H + sFactors, f.data(), f.size() * 8);
    }
    b->cfg.factors = at<double>(D, sFactors);
    b->cfg.nFactors = static_cast<uint32_t>(f.size());
  }
  std::atomic<int> badInit{0};
  parallelFor(nInst, 256, [&](size_t i) {
    const auto& c = circs[circOf[i]];
    int16_t* init = at<int16_t>(H, sInit) + i * n;
    if (instances[i].initial_locations != nullptr) {
      // locations must be injective: two logical qubits on one physical
      // qubit would give the search layouts that are not permutations
      uint64_t used[16] = {};  // n <= 1024 (uint16 qubit ids; checked below)
      for (int q = 0; q < n; ++q) {
        const int16_t p = instances[i].initial_locations[q];
        if (p < -1 || p >= n) {
          badInit.store(1);
        } else if (p >= 0 && n <= 1024) {
          if ((used[p >> 6] >> (p & 63)) & 1U) {
            badInit.store(2);
          }
          used[p >> 6] |= uint64_t{1} << (p & 63);
        }
        init[q] = p;
      }
    } else {
      std::fill(init, init + n, static_cast<int16_t>(-1));
    }
    InstView& I = at<InstView>(H, b->iInst)[i];
    std::memset(&I, 0, sizeof(I));
    I.circ.nLayers = c.nLayers;
    I.circ.nLogical = c.nLogical;
    I.circ.g2Off = at<uint32_t>(D, c.bG2Off);
    I.circ.g2 = at<Gate2>(D, c.bG2);
    I.circ.g1Off = at<uint32_t>(D, c.bG1Off);
    I.circ.g1 = at<Gate1>(D, c.bG1);
    I.circ.consOff = nullptr;
    I.circ.cons = at<uint16_t>(D, c.dCons);
    I.circ.consCnt = at<uint16_t>(D, c.dConsCnt);
    I.circ.twoQ = at<uint32_t>(D, c.dTwoQ);
    I.circ.firstAfter = at<uint32_t>(D, c.dFirst);
    I.circ.nTwoQ = c.nTwoQ;
    I.circ.splittable = at<uint8_t>(D, c.dSplit);
    I.initLoc = at<int16_t>(D, sInit) + i * n;
    I.passes = instances[i].bidirectional_passes;
    I.firstLayer = instances[i].first_layer;
    I.nRoute = instances[i].n_route_layers;
    I.layerCap = b->layerCap[i];
    I.splitCap = b->layerCap[i] - c.nLayers;
    I.splitLog = at<uint32_t>(O, b->oSplitLog) + b->splitOff[i];
    I.layers = b->wantLayers ? at<LayerOut>(O, b->oLayers) + b->layerOff[i]
                             : nullptr;
    I.layerNSwaps = b->wantLayerNS ? at<uint16_t>(O, b->oLayerNS) + b->layerOff[i]
                                   : nullptr;
    I.layerQ = b->wantLayerQ
                   ? at<int16_t>(O, b->oLayerQ) + b->layerOff[i] * n
                   : nullptr;
    I.swaps = at<uint16_t>(O, b->oSwapsWork) + 2 * b->swapOff[i];
    I.swapCap = b->swapCap[i];
    I.swapsPacked = at<uint16_t>(O, b->oSwapsPacked);
    I.swapCounter = at<uint32_t>(O, b->oSwapCounter);
    I.places = at<uint16_t>(O, b->oPlaces) + 2 * i * n;
    I.placeCap = static_cast<uint32_t>(n);
    I.finalQ = at<int16_t>(O, b->oFinalQ) + i * n;
    I.finalL = at<int16_t>(O, b->oFinalL) + i * n;
    I.startL = at<int16_t>(O, b->oStartL) + i * n;
    if (b->wantNodeLog && b->logCap[i] > 0) {
      I.nodeLog = at<NodeLogRec>(O, b->oNodeLog) + b->logOff[i];
      I.nodeLogCap = b->logCap[i];
      I.rootQ = at<int16_t>(O, b->oRootQ) + b->layerOff[i] * n;
    }
    if (b->wantPlaceLog) {
      I.placeLog = at<uint16_t>(O, b->oPlaceLog) + 2 * i * n;
      I.placeLogCap = static_cast<uint32_t>(n);
    }
    I.summary = at<uint32_t>(O, b->oSummary) + 8 * i;
    I.counters = at<uint64_t>(O, b->oCounters) + 4 * i;
  });
  if (badInit.load() != 0) {
    return fail(QMG_BAD_ARG, badInit.load() == 2
                                 ? "two logical qubits on one physical qubit in "
                                   "initial_locations"
                                 : "initial location out of range");
  }
  {
    // largest instances first
    uint32_t* order = at<uint32_t>(H, b->iOrder);
    std::iota(order, order + nInst, 0U);
    bool uniform = true;
    for (size_t i = 1; i < nInst && uniform; ++i) {
      uniform = b->nLayers[i] == b->nLayers[0] &&
                instances[i].bidirectional_passes ==
                    instances[0].bidirectional_passes;
    }
    if (!uniform) {
      std::stable_sort(order, order + nInst, [&](uint32_t x, uint32_t y) {
        const uint64_t wx = static_cast<uint64_t>(b->nLayers[x]) *
                            (1 + 2 * instances[x].bidirectional_passes);
        const uint64_t wy = static_cast<uint64_t>(b->nLayers[y]) *
                            (1 + 2 * instances[y].bidirectional_passes);
        return wx > wy;
      });
    }
  }
  {
    SearchMem* mem = at<SearchMem>(H, b->iMem);
    for (uint32_t w = 0; w < nWarps; ++w) {
      SearchMem& M = mem[w];
      M = SearchMem{};
      M.pool = Wk + wPool + static_cast<size_t>(w) * poolNodes * recBytes;
      M.open = at<OpenEntry>(Wk, wOpen) + static_cast<size_t>(w) * poolNodes;
      M.table = at<uint64_t>(Wk, wTable) + static_cast<size_t>(w) * tableSize;
      M.btop = at<OpenEntry>(Wk, wBtop) + static_cast<size_t>(w) * btopPerWarp;
      if (pairSlots != 0) {
        M.pair = at<PairSlot>(Wk, wPair) + static_cast<size_t>(w) * pairSlots;
      }
      M.nsw = at<uint16_t>(Wk, wNsw) + static_cast<size_t>(w) * 32 * maxG;
      M.poolCap = poolNodes;
      M.bigHeld = -1;
      M.big = nBig > 0 ? at<BigPool>(D, b->iBig) : nullptr;
      M.nBig = nBig;
      M.recBytes = recBytes;
      M.tableMask = tableSize - 1;
      M.stamp = 0;
      M.maxG = maxG;
      if (b->deviceSplits) {
        M.priv = Wk + wPriv + static_cast<size_t>(w) * privStride;
        M.privLayers = privLayers;
        M.privG2 = privG2;
        M.privG1 = privG1;
      }
    }
  }
  for (uint32_t k = 0; k < nBig; ++k) {
    BigPool& P = at<BigPool>(H, b->iBig)[k];
    P = BigPool{};
    P.pool = Wk + wBigPool + static_cast<size_t>(k) * bigNodes * recBytes;
    P.open = at<OpenEntry>(Wk, wBigOpen) + static_cast<size_t>(k) * bigNodes;
    P.table = at<uint64_t>(Wk, wBigTable) + static_cast<size_t>(k) * bigTableSize;
    P.btop = at<OpenEntry>(Wk, wBigBtop) + static_cast<size_t>(k) * bigBtop;
    P.poolCap = bigNodes;
    P.tableMask = bigTableSize - 1;
  }
  sc.mark("views");
  // (the circuit ranges are already on their way)
  CU(cudaMemcpyAsync(D + circBytes, H + circBytes, b->inSize - circBytes,
                     cudaMemcpyHostToDevice, s));
  if (!circs.empty()) {
    derive_circuit_tables<<<static_cast<unsigned>(circs.size()), 128, 0, s>>>(
        at<DeriveView>(D, sDerive), static_cast<uint32_t>(circs.size()));
    CU(cudaGetLastError());
  }
  CU(cudaMemsetAsync(Wk + b->tableOff, 0, b->tableBytes, s));
  if (b->bigTableBytes != 0) {
    CU(cudaMemsetAsync(Wk + b->bigTableOff, 0, b->bigTableBytes, s));
  }
  b->h2dBytes = b->inSize;
  {
    const int sb = static_cast<int>(b->smemBytes);
    CU(routeSetSharedMemoryH0(sb));
    CU(routeSetSharedMemoryH1(sb));
    CU(routeSetSharedMemoryH2(sb));
    CU(routeSetSharedMemoryH3(sb));
    CU(routeSetSharedMemoryH4(sb));
  }
  sc.mark("attrs");
  CU(cudaStreamSynchronize(s));
  sc.mark("h2d-sync");
  sc.flush("batch_create");
  b->h2dMs = std::chrono::duration<float, std::milli>(
                 std::chrono::steady_clock::now() - t0)
                 .count();
  *out = b.release();
  return QMG_OK;
}

int batchRun(qmg_batch* b, qmg_route_stats* stats) {
  if (int rc = selectDevice(b->arch->device)) {
    return rc;
  }
  cudaStream_t s = b->arch->stream;
  qmg_workspace& W = *b->ws;
  // packed-SWAP counter and instance ticket
  CU(cudaMemsetAsync(W.out.p + b->oSwapCounter, 0, 16, s));
  EventTimer tm(s);
  tm.start();
  if (b->nInst > 0) {
    RouteArgs ra{};
    ra.A = b->arch->view();
    ra.C = b->cfg;
    ra.inst = at<InstView>(W.in.p, b->iInst);
    ra.order = at<uint32_t>(W.in.p, b->iOrder);
    ra.nInst = static_cast<uint32_t>(b->nInst);
    ra.mem = at<SearchMem>(W.in.p, b->iMem);
    ra.ticket = at<uint32_t>(W.out.p, b->oSwapCounter) + 1;
    ra.lay = b->sl;
    ra.startFull = b->startFull;
    const size_t sm = b->smemBytes;
    switch (b->cfg.heuristic) {
    case 0:
      CU(routeLaunchH0(b->grid, sm, s, ra));
      break;
    case 1:
      CU(routeLaunchH1(b->grid, sm, s, ra));
      break;
    case 2:
      CU(routeLaunchH2(b->grid, sm, s, ra));
      break;
    case 3:
      CU(routeLaunchH3(b->grid, sm, s, ra));
      break;
    default:
      CU(routeLaunchH4(b->grid, sm, s, ra));
      break;
    }
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
  qmg_workspace& W = *b->ws;
  const size_t nI = b->nInst;
  const size_t n = static_cast<size_t>(b->n);
  const auto t0 = std::chrono::steady_clock::now();
  uint8_t* Hh = W.hout.p;
  StageClock sc;
  const uint8_t* O = W.out.p;
  // phase 1: summaries, counters, layouts, placements
  CU(cudaMemcpyAsync(Hh, O, b->oSmallEnd, cudaMemcpyDeviceToHost, s));
  CU(cudaStreamSynchronize(s));
  sc.mark("d2h-small");
  // phase 2: per-layer outputs and the SWAPs actually produced
  const uint32_t packed = *at<uint32_t>(Hh, b->oSwapCounter);
  const size_t end2 = b->oSwapsPacked + static_cast<size_t>(packed) * 4;
  if (end2 > b->oSmallEnd) {
    CU(cudaMemcpyAsync(Hh + b->oSmallEnd, O + b->oSmallEnd, end2 - b->oSmallEnd,
                       cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
  }
  sc.mark("d2h-bulk");
  const uint32_t* summary = at<uint32_t>(Hh, b->oSummary);
  const uint64_t* counters = at<uint64_t>(Hh, b->oCounters);
  std::atomic<uint64_t> exp{0}, gen{0};
  parallelFor(nI, 512, [&](size_t i) {
    qmg_route_result& r = results[i];
    r.n_swaps_total = summary[8 * i];
    r.n_placements_total = summary[8 * i + 1];
    r.n_layers_done = summary[8 * i + 2];
    r.n_splits = summary[8 * i + 5];
    if (r.split_layers != nullptr) {
      const uint32_t k = std::min(
          {r.n_splits, r.split_capacity, b->layerCap[i] - b->nLayers[i]});
      std::memcpy(r.split_layers, at<uint32_t>(Hh, b->oSplitLog) + b->splitOff[i],
                  static_cast<size_t>(k) * 4);
    }
    r.status = static_cast<int32_t>(summary[8 * i + 3]);
    r.expanded_nodes = counters[4 * i];
    r.generated_nodes = counters[4 * i + 1];
    r.expanded_nodes_final_pass = counters[4 * i + 2];
    r.generated_nodes_final_pass = counters[4 * i + 3];
    exp.fetch_add(r.expanded_nodes, std::memory_order_relaxed);
    gen.fetch_add(r.generated_nodes, std::memory_order_relaxed);
    const size_t nl = std::min(
        b->layerCap[i], std::max(b->nLayers[i], r.n_layers_done));
    if (r.layers != nullptr && b->wantLayers) {
      std::memcpy(r.layers, at<LayerOut>(Hh, b->oLayers) + b->layerOff[i],
                  nl * sizeof(LayerOut));
    }
    if (r.layer_n_swaps != nullptr && b->wantLayerNS) {
      std::memcpy(r.layer_n_swaps, at<uint16_t>(Hh, b->oLayerNS) + b->layerOff[i],
                  nl * 2);
    }
    if (r.layer_qubits != nullptr && b->wantLayerQ) {
      std::memcpy(r.layer_qubits, at<int16_t>(Hh, b->oLayerQ) + b->layerOff[i] * n,
                  nl * n * 2);
    }
    r.n_node_log = summary[8 * i + 6];
    r.n_placement_log = summary[8 * i + 7];
    if (b->wantPlaceLog && r.placement_log != nullptr) {
      const size_t k = std::min<size_t>({r.n_placement_log, n, r.placement_log_capacity});
      std::memcpy(r.placement_log, at<uint16_t>(Hh, b->oPlaceLog) + 2 * i * n, k * 4);
    }
    if (b->wantNodeLog && b->logCap[i] > 0 && r.node_log != nullptr) {
      const size_t k = std::min<size_t>({r.n_node_log, b->logCap[i], r.node_log_capacity});
      std::memcpy(r.node_log, at<NodeLogRec>(Hh, b->oNodeLog) + b->logOff[i],
                  k * sizeof(NodeLogRec));
      if (r.layer_root_qubits != nullptr) {
        std::memcpy(r.layer_root_qubits, at<int16_t>(Hh, b->oRootQ) + b->layerOff[i] * n,
                    nl * n * 2);
      }
    }
    if (r.swaps != nullptr) {
      const uint32_t ns = r.n_swaps_total;
      if (ns <= b->swapCap[i] && ns <= r.swap_capacity) {
        std::memcpy(r.swaps,
                    at<uint32_t>(Hh, b->oSwapsPacked) + summary[8 * i + 4],
                    static_cast<size_t>(ns) * 4);
      } else if (r.status == QMG_OK) {
        r.status = QMG_BUFFER_TOO_SMALL;
      }
    }
    const uint32_t np =
        std::min<uint32_t>(r.n_placements_total, static_cast<uint32_t>(n));
    if (r.placements != nullptr && r.placement_capacity >= np) {
      std::memcpy(r.placements, at<uint16_t>(Hh, b->oPlaces) + 2 * i * n,
                  static_cast<size_t>(np) * 4);
    }
    if (r.final_qubits != nullptr) {
      std::memcpy(r.final_qubits, at<int16_t>(Hh, b->oFinalQ) + i * n, n * 2);
    }
    if (r.final_locations != nullptr) {
      std::memcpy(r.final_locations, at<int16_t>(Hh, b->oFinalL) + i * n, n * 2);
    }
    if (r.start_locations != nullptr) {
      std::memcpy(r.start_locations, at<int16_t>(Hh, b->oStartL) + i * n, n * 2);
    }
  });
  sc.mark("scatter");
  sc.flush("batch_fetch");
  if (stats != nullptr) {
    stats->d2h_ms = std::chrono::duration<float, std::milli>(
                        std::chrono::steady_clock::now() - t0)
                        .count();
    stats->d2h_bytes = std::max(end2, b->oSmallEnd);
    stats->big_pool_layers = at<uint32_t>(Hh, b->oSwapCounter)[2];
    stats->expanded_nodes = exp.load();
    stats->generated_nodes = gen.load();
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
    if (int rc = batchFetch(b, results, stats)) {
      return rc;
    }
    if (options != nullptr && options->pool_nodes != 0) {
      return static_cast<int>(QMG_OK);
    }
    // ---- instances whose node pool ran out: again, with fewer resident
    // searches and pools 16 times as large, out of most of the free HBM
    uint64_t pool = b->poolNodes;
    const size_t perNode = nodeRecBytes(b->n) + sizeof(OpenEntry) +
                           4 * sizeof(uint64_t) + 1;
    bool retried = false;
    for (int tier = 0; tier < 3; ++tier) {
      std::vector<size_t> idx;
      for (size_t i = 0; i < nInstances; ++i) {
        if (results[i].status == QMG_POOL_EXHAUSTED) {
          idx.push_back(i);
        }
      }
      if (idx.empty()) {
        break;
      }
      retried = true;
      hold->ws->work.release();
      hold.reset();
      size_t freeB = 0, totalB = 0;
      CU(cudaMemGetInfo(&freeB, &totalB));
      const size_t budget = freeB / 10 * 8;
      pool = std::min<uint64_t>(pool * 16, kMaxPoolNodes);
      pool = std::min<uint64_t>(pool, budget / (kWarpsPerCta * perNode));
      qmg_route_options o2{};
      if (options != nullptr) {
        o2 = *options;
      }
      o2.pool_nodes = static_cast<uint32_t>(pool);
      o2.max_searches = static_cast<uint32_t>(std::max<size_t>(
          kWarpsPerCta,
          std::min<size_t>(budget / (pool * perNode), size_t{1} << 20)));
      std::vector<qmg_instance> inst2(idx.size());
      std::vector<qmg_route_result> res2(idx.size());
      for (size_t j = 0; j < idx.size(); ++j) {
        inst2[j] = instances[idx[j]];
        res2[j] = results[idx[j]];
      }
      qmg_batch* b2 = nullptr;
      if (int rc = batchCreate(arch, cfg, idx.size(), inst2.data(), &o2,
                               res2.data(), &b2)) {
        return rc;
      }
      hold.reset(b2);
      qmg_route_stats st2{};
      if (int rc = batchRun(b2, &st2)) {
        return rc;
      }
      if (int rc = batchFetch(b2, res2.data(), &st2)) {
        return rc;
      }
      for (size_t j = 0; j < idx.size(); ++j) {
        results[idx[j]] = res2[j];
      }
      if (stats != nullptr) {
        stats->kernel_ms += st2.kernel_ms;
        stats->h2d_ms += st2.h2d_ms;
        stats->d2h_ms += st2.d2h_ms;
        stats->h2d_bytes += st2.h2d_bytes;
        stats->d2h_bytes += st2.d2h_bytes;
        stats->kernel_launches += st2.kernel_launches;
        stats->retried_instances += static_cast<uint32_t>(idx.size());
        stats->big_pool_layers += st2.big_pool_layers;
      }
    }
    if (retried) {
      hold->ws->work.release();  // (most of the HBM: not kept between calls)
      hold->ws->memBudget = 0;
      if (stats != nullptr) {
        stats->expanded_nodes = 0;
        stats->generated_nodes = 0;
        for (size_t i = 0; i < nInstances; ++i) {
          stats->expanded_nodes += results[i].expanded_nodes;
          stats->generated_nodes += results[i].generated_nodes;
        }
      }
    }
    return static_cast<int>(QMG_OK);
  });
}
