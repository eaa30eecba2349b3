// route_core.hpp - the A* SWAP search of one routing instance, written once
// against a small "lane group" abstraction.
//
// Compiled by nvcc with a 32-lane warp group (route.cu: the shipped sm_100a
// kernel, one warp per resident search) and by g++ with a 1-lane group
// (tests/emu: a host build of the SAME source used only to debug the search
// logic under ASan/UBSan where no GPU exists; it is not part of the product
// library and nothing in the product path can reach it).
//
// Reference semantics reproduced (paths relative to the reference tree):
//   aStarMap                src/sc/heuristic/HeuristicMapper.cpp:521-696
//   expandNode              :698-727   expandNodeAddOneSwap :729-748
//   applySWAP               :861-973   updateSharedSwaps    :975-1053
//   recalculateFixedCost*   :750-859   heuristics           :1055-1354
//   lookahead               :1356-1513 mapUnmappedGates     :211-307
//   node order / identity   include/sc/heuristic/HeuristicMapper.hpp:438-483
//   open list               include/sc/heuristic/UniquePriorityQueue.hpp:59-93
#pragma once
#include <stdint.h>
#include <float.h>
#include <math.h>
#include <string.h>

#if defined(__CUDACC__)
#define QMG_HD __host__ __device__ __forceinline__
#define QMG_D __device__ __forceinline__
#define QMG_NOUNROLL _Pragma("unroll 1")
#define QMG_PRAGMA_(x) _Pragma(#x)
#define QMG_UNROLL(n) QMG_PRAGMA_(unroll n)
// (measured: out-of-line reductions cost 6 % on config 2, so the default is
// inline; -DQMG_OUTLINE_REDUCTIONS keeps the experiment buildable)
#if defined(QMG_OUTLINE_REDUCTIONS)
#define QMG_HD_RED inline __host__ __device__ __noinline__
#else
#define QMG_HD_RED QMG_HD
#endif
#define QMG_COLD inline __host__ __device__ __noinline__
#else
#define QMG_COLD inline
#define QMG_HD_RED inline
#define QMG_HD inline
#define QMG_D inline
#define QMG_NOUNROLL
#define QMG_UNROLL(n)
#endif

namespace qmg {

constexpr double kCostUniSwap = 34.0;  // Architecture.hpp:37 COST_UNIDIRECTIONAL_SWAP
constexpr double kCostBiSwap = 30.0;   // Architecture.hpp:39 COST_BIDIRECTIONAL_SWAP
constexpr uint32_t kCostReverse = 4U;  // Architecture.hpp:40 COST_DIRECTION_REVERSE
constexpr double kTol = 1e-6;          // HeuristicMapper.hpp:463,473
constexpr uint32_t kNone = 0xFFFFFFFFU;
constexpr int kChunk = 32;    // children scored per pass
constexpr int kLookBlock = 16; // lookahead layers scored per pass
// lookahead scratch: a (kChunk/2) x kLookBlock block of terms
constexpr int kLookPenDoubles = (kChunk / 2) * kLookBlock;
#ifndef QMG_SCAN_MAX
#define QMG_SCAN_MAX 96
#endif
// open-list length up to which push() scans (tests build the host variant
// with a small value so that short golden searches use the hash table)
constexpr uint32_t kScanMax = QMG_SCAN_MAX;
// open lists longer than kBlkMin keep the top of every kBlk-entry block, so
// that a pop reads the block tops and re-scans one block instead of the list
#ifndef QMG_BLK
#define QMG_BLK 256
#endif
#ifndef QMG_BLK_MIN
#define QMG_BLK_MIN 4096
#endif
// The kernel is bound by instruction fetch, so the
// unrolling / in-flight factors below are kept as small as the latency they
// hide allows.  Measured on config 4, 32768 instances per step (M expanded
// nodes/s): 65.4 with (block 512, flight 4, scans 4/3, wide 4, two inlined
// commit paths in push) -> 75.8 (one commit path, block 256, flight 2) -> 79.4
// (scans 2/1) -> 82.4 (global-address hint, glob()) -> 83.5 (wide 2); at 16384
// instances per step: 70.7 -> 75.3 (QMG_TABLE_FROM_START).
// unrolling of the fidelity heuristic's scans over physical qubits (1q gates)
// and coupling edges (2q gates; only run when the memo misses)
#ifndef QMG_SCAN1_UNROLL
#define QMG_SCAN1_UNROLL 2
#endif
#ifndef QMG_SCAN2_UNROLL
#define QMG_SCAN2_UNROLL 1
#endif
// open-list entries in flight per lane when block tops / blocks are scanned
#ifndef QMG_WIDE
#define QMG_WIDE 2
#endif
// the full search variant indexes the open list by the layout hash table from
// the first node of a layer on (the lean variant scans open lists of up to
// QMG_SCAN_MAX nodes): all children then take pushChunk, and pushOne with its
// copies of the layout comparison and the record write leaves the hot code
#ifndef QMG_TABLE_FROM_START
#define QMG_TABLE_FROM_START 1
#endif
// layout comparisons in flight per lane in pushChunk's look-up
#ifndef QMG_FLIGHT
#define QMG_FLIGHT 2
#endif
constexpr uint32_t kPairGates = 2;
constexpr uint32_t kBlk = QMG_BLK;  // (tests build the host variant with 4 / 8)
constexpr uint32_t kBlkMin = QMG_BLK_MIN;
constexpr bool kTableFromStart = QMG_TABLE_FROM_START != 0;

enum Status : int32_t {
  ST_OK = 0,
  ST_NO_VIABLE = 1,
  ST_SPLIT = 2,
  // internal: the lean variant of the search met a long open list or a layer
  // to split; the instance is routed again by the full variant (never
  // reported to the caller)
  ST_NEEDS_FULL = 100,
  ST_BAD_ARG = 3,
  ST_POOL = 6,
  ST_BUFFER = 8,
};

struct Gate2 {
  uint16_t q1, q2, fwd, rev;
};
struct Gate1 {
  uint16_t q, mult;
};

// device-resident Architecture (include/sc/Architecture.hpp:42)
struct ArchView {
  int32_t n, nE;
  int32_t bidirectional, unidirectional, fidelity, nFid;
  const uint16_t* edges;   // 2*nE, ascending (first, second)
  const uint32_t* incOff;  // n+1
  const uint32_t* inc;     // ids of the directed edges touching p, ascending
  const uint8_t* adj;      // n*n; bit0: (r,c) in coupling map, bit1: (c,r)
  const double* dist;      // getDistanceTable(false)
  const double* distRev;   // getDistanceTable(true)
  const double* c1;        // singleQubitFidelityCosts
  const double* c2;        // twoQubitFidelityCosts
  const double* cswap;     // swapFidelityCosts
  const double* fid;       // fidelityDistanceTables, nFid*n*n
};

struct CfgView {
  int32_t heuristic, lookahead;
  uint32_t nrLookaheads;
  int32_t earlyTermination;
  double firstFactor, factor;
  uint64_t earlyLimit;
  uint64_t splitLimit;
  int32_t autoSplit;
  int32_t tight, fidelityAware;
  // factors[i] = firstFactor * factor^i built by repeated multiplication as
  // updateLookaheadPenalty does (:1361,:1381); min(nrLookaheads, cap) entries
  const double* factors;
  uint32_t nFactors;
};

// layers of one circuit (Mapper::createLayers, src/sc/Mapper.cpp:138-281)
struct CircView {
  uint32_t nLayers, nLogical;
  const uint32_t* g2Off;
  const Gate2* g2;
  const uint32_t* g1Off;
  const Gate1* g1;
  // getConsideredQubits(layer), ascending: either a CSR (consOff) or, when
  // consOff is null, at most 2G+G1 entries from 2*g2Off[l]+g1Off[l] with
  // their number in consCnt[l] (the layout the device derives itself)
  const uint32_t* consOff;
  const uint16_t* cons;
  const uint16_t* consCnt;
  // Mapper::getNextLayer (src/sc/Mapper.cpp:437-448) in closed form: twoQ is
  // the ascending list of layers holding a 2q gate, firstAfter[l] the index in
  // twoQ of the first such layer > l
  const uint32_t* twoQ;
  const uint32_t* firstAfter;
  uint32_t nTwoQ;
  const uint8_t* splittable;  // Mapper::isLayerSplittable
};

struct LayerOut {  // == qmg_layer_result
  double cost_fixed, cost_fixed_reversals, cost_heur, lookahead_penalty;
  uint64_t expanded, generated, expAfterFirst, expAfterOpt, solNodes,
      solAfterOpt, layoutHash;
  uint32_t depth, resultId, swapOff, nSwaps, placeOff, nPlace;
  int32_t early, status;
};

// one generated search node as DataLogger::logSearchNode records it
// (src/sc/DataLogger.cpp:216-269), == qmg_node_log_record.  The node's layout
// and SWAP sequence follow from its ancestors' records and the layer's root
// layout (InstView::rootQ).
struct NodeLogRec {
  double cost_fixed;  // costFixed + costFixedReversals
  double cost_heur, lookahead_penalty;
  uint32_t id, parent, depth, layer;
  uint16_t swap_first, swap_second;  // the SWAP that made the node
  uint16_t valid, root;
};
static_assert(sizeof(NodeLogRec) == 48, "NodeLogRec layout");

struct InstView {
  CircView circ;
  const int16_t* initLoc;  // n or null
  uint32_t passes;         // bidirectional pseudo-routing passes
  uint32_t firstLayer, nRoute;
  // capacity (in layers) of the per-layer outputs; layers split on the device
  // (Mapper::splitLayer) need layerCap > circ.nLayers
  uint32_t layerCap;
  uint32_t splitCap;
  uint32_t* splitLog;  // layer index of every split of the final pass, in order
  // outputs
  LayerOut* layers;
  uint16_t* swaps;
  uint32_t swapCap;
  uint16_t* places;
  uint32_t placeCap;
  int16_t* finalQ;
  int16_t* finalL;
  int16_t* startL;
  int16_t* layerQ;
  uint16_t* layerNSwaps;  // optional compact per-layer SWAP counts
  // packed SWAP output: at the end of the instance its SWAPs are appended to
  // swapsPacked at an offset taken from the batch-wide counter
  uint16_t* swapsPacked;
  uint32_t* swapCounter;
  // [0]=nSwapsTotal [1]=nPlaceTotal [2]=nLayersDone [3]=status
  // [4]=offset of the instance's SWAPs in swapsPacked [5]=layer splits of the
  // final pass
  uint32_t* summary;
  // [0]=expanded [1]=generated (all passes) [2], [3] = the same for the final
  // routing pass only (what the reference's heuristicBenchmark counts: it
  // switches the counters off inside pseudoRouteCircuit, :309-344)
  uint64_t* counters;
  // optional log of every node generated in the final pass, in generation
  // order ([6] of summary = their number), and the root layout of every layer
  NodeLogRec* nodeLog;
  uint32_t nodeLogCap;
  int16_t* rootQ;  // layerCap * n
  // optional: every placement of every pass (pseudo-routing passes too), in
  // order ([7] of summary = their number); the physical qubit has bit 15 set
  // when the qubit was placed by the 1q-gate rule of the fidelity-aware
  // mapper (:213-231), which - unlike the 2q rules - leaves the circuit's
  // initialLayout / outputPermutation alone
  uint16_t* placeLog;
  uint32_t placeLogCap;
};

// entries of one layout array in a node record: n rounded up to a multiple of
// 4, so that both arrays start on 8-byte boundaries and are moved as 64-bit
// words
QMG_HD uint32_t padN(int n) { return (static_cast<uint32_t>(n) + 3U) & ~3U; }

struct NodeHdr {  // 72 bytes, followed by qubits[padN(n)], locations[padN(n)]
  double fixed, rev, heur, look;
  uint64_t hash;
  uint32_t parent, shared;
  uint32_t depth;
  uint16_t swA, swB;
  uint16_t nValid, flags;  // flags bit0 valid, bit1 open
  uint32_t openPos;
  uint32_t id;
  // a node that outlives a layer split (the restarted call inherits the open
  // list) was scored against the unsplit layer: epoch = number of splits
  // before the node was made; extra = entries of the reference's
  // validMappedTwoQubitGates set that belong to gates no longer in the layer
  // (children copy the set and only update gates of the current layer,
  // HeuristicMapper.cpp:729-748, 861-973)
  uint16_t epoch, extra;
};
static_assert(sizeof(NodeHdr) == 72, "NodeHdr layout");
// bytes of one node record
inline uint32_t nodeRecBytes(int n) {
  return static_cast<uint32_t>(sizeof(NodeHdr)) + 4U * padN(n);
}

struct OpenEntry {  // 32 bytes
  double f, h2;
  uint64_t hash;
  uint32_t idx;
  uint16_t nValid;
  uint16_t valid;
};

// entry of the layout-keyed hash table (SearchMem::table), 64 bits:
// [stamp:24][hash bits 50..63:14][pool index:26].  stamp identifies the
// running aStarMap call (0 = never written); the hash bits spare the probe a
// dependent read of the node record; a match is always confirmed by comparing
// the layouts.
constexpr uint32_t kStampMask = 0xFFFFFFU;
constexpr uint32_t kTabIdxMask = 0x3FFFFFFU;
constexpr uint32_t kMaxPoolNodes = kTabIdxMask;  // per search
QMG_HD uint64_t tabPack(uint32_t stamp, uint64_t hash, uint32_t idx) {
  return (static_cast<uint64_t>(stamp & kStampMask) << 40) |
         (((hash >> 50) & 0x3FFFULL) << 26) | idx;
}
QMG_HD uint32_t tabStamp(uint64_t e) { return static_cast<uint32_t>(e >> 40); }
QMG_HD uint32_t tabIdx(uint64_t e) {
  return static_cast<uint32_t>(e) & kTabIdxMask;
}
QMG_HD bool tabHashEq(uint64_t e, uint64_t hash) {
  return ((e >> 26) & 0x3FFFULL) == ((hash >> 50) & 0x3FFFULL);
}

// one slot of the fidelity best-edge memo: the value and the aStarMap call
// (pairKey) that computed it, read with one 16-byte load
struct alignas(16) PairSlot {
  uint64_t tag;
  double val;
};

struct Child {
  double fixed, rev, heur, look;
  uint64_t hash;
  uint32_t shared;
  uint16_t a, b;
  uint16_t nValid, valid;
  uint32_t pad;
};

// A large node pool shared by all searches of a launch.  Every resident
// search owns a small private pool (SearchMem::pool ...); a search whose layer
// outgrows it takes one of the few large pools for the rest of that layer
// (Search::migrateToBig) and carries on there with its state.  The
// private pools can then be sized for the common case instead of the worst
// one, which is what lets more searches stay resident per SM.
struct BigPool {
  uint8_t* pool;
  OpenEntry* open;
  uint64_t* table;
  OpenEntry* btop;
  uint32_t poolCap, tableMask;
  uint32_t lock;   // 0 free, 1 held (atomicCAS)
  uint32_t stamp;  // table stamp of the last holder (24 bits used)
};

// per resident search, in global memory
struct SearchMem {
  uint8_t* pool;
  OpenEntry* open;
  uint64_t* table;
  // tops of the kBlk-entry blocks of `open` (kept once the open list is long,
  // see Search::blkEnter): poolCap / kBlk + 2 entries
  OpenEntry* btop;
  // fidelity-aware scoring, full variant: memo of the best-edge term of the
  // layer's first kPairGates 2q gates, keyed on the locations (l1, l2) of the
  // gate's qubits: kPairGates * n * n slots; tag = pairKey() of the aStarMap
  // call that filled the slot.  Null = no memo.
  PairSlot* pair;
  uint64_t reserved0;
  uint16_t* nsw;  // scratch for heuristic 2/3: kChunk * maxG
  uint32_t poolCap, recBytes, tableMask, stamp;
  uint32_t maxG;
  // private, writable copy of the instance's circuit tables (layers split on
  // the device); null = the device never splits
  uint32_t privLayers, privG2, privG1;
  uint8_t* priv;
  uint32_t epoch;       // splits of the layer being searched so far
  uint32_t childExtra;  // NodeHdr::extra of the children being generated
  // large pools of the launch (null / 0: none) and, while one is held, its
  // index and the private pool's fields it replaced
  BigPool* big;
  uint32_t nBig;
  int32_t bigHeld;     // -1: the search runs in its private pool
  uint32_t tabStampV;  // stamp written into table entries by this call
  uint32_t poolCap0, tableMask0;
  uint8_t* pool0;
  OpenEntry* open0;
  uint64_t* table0;
  OpenEntry* btop0;
  uint64_t bigUses;    // (statistics) layers this search ran in a large pool
};

// per resident search, in shared memory (device) / one byte buffer (host).
// Fixed-size sections come first so that their offsets are compile-time
// constants (immediates of the shared-memory instructions); the
// n-dependent sections follow at multiples of one stride:
// [ch][lookPen][InstView][SearchMem] [Q][L][curQ][curL][consMask]
constexpr uint32_t kAl16(uint32_t v) { return (v + 15U) & ~15U; }
constexpr uint32_t kOffCh = 0;
constexpr uint32_t kOffLook = kOffCh + kAl16(sizeof(Child) * kChunk);
constexpr uint32_t kOffInst = kOffLook + kAl16(8U * kLookPenDoubles);
constexpr uint32_t kOffMem = kOffInst + kAl16(sizeof(InstView));
constexpr uint32_t kOffVar = kOffMem + kAl16(sizeof(SearchMem));
struct Scratch {
  uint8_t* base;
  uint32_t stride;  // bytes of one n-entry int16 section (16-byte aligned)
  uint32_t bytes;
  QMG_HD Child* ch() const { return reinterpret_cast<Child*>(base + kOffCh); }
  // kLookPenDoubles doubles; also output staging
  QMG_HD double* lookPen() const {
    return reinterpret_cast<double*>(base + kOffLook);
  }
  QMG_HD InstView* inst() const {
    return reinterpret_cast<InstView*>(base + kOffInst);
  }
  QMG_HD SearchMem* mem() const {
    return reinterpret_cast<SearchMem*>(base + kOffMem);
  }
  // layout of the node being expanded
  QMG_HD int16_t* Q() const {
    return reinterpret_cast<int16_t*>(base + kOffVar);
  }
  QMG_HD int16_t* L() const {
    return reinterpret_cast<int16_t*>(base + kOffVar + stride);
  }
  // Mapper::qubits / Mapper::locations of the instance
  QMG_HD int16_t* curQ() const {
    return reinterpret_cast<int16_t*>(base + kOffVar + 2 * stride);
  }
  QMG_HD int16_t* curL() const {
    return reinterpret_cast<int16_t*>(base + kOffVar + 3 * stride);
  }
  QMG_HD uint32_t* consMask() const {
    return reinterpret_cast<uint32_t*>(base + kOffVar + 4 * stride);
  }
};

// section sizes for an n-qubit architecture; base is set by the owner of the
// buffer
inline Scratch scratchLayout(int n) {
  Scratch s{};
  s.base = nullptr;
  s.stride = kAl16(static_cast<uint32_t>(n) * 2U + 2U);
  s.bytes = kOffVar + 4 * s.stride +
            kAl16((static_cast<uint32_t>(n) + 31U) / 32U * 4U + 4U);
  return s;
}

QMG_HD uint32_t mix32(uint32_t x) {
  x ^= x >> 16;
  x *= 0x7feb352dU;
  x ^= x >> 15;
  x *= 0x846ca68bU;
  x ^= x >> 16;
  return x;
}
// layout hash: sum over physical qubits of a 64-bit mix of (p, qubits[p]);
// updated incrementally per SWAP.  Also reported as
// qmg_layer_result.layout_hash.
QMG_HD uint64_t cellHash(int p, int q) {
  const uint32_t x = (static_cast<uint32_t>(p) << 16) | static_cast<uint16_t>(q);
  const uint32_t lo = mix32(x + 0x9E3779B9U);
  const uint32_t hi = mix32(x ^ 0x85EBCA6BU);
  return (static_cast<uint64_t>(hi) << 32) | lo;
}

// ---------------------------------------------------------------------------
// lane groups
// ---------------------------------------------------------------------------
#if defined(__CUDA_ARCH__)
struct Group {
  static constexpr int W = 32;
  int lane;
  QMG_D Group() : lane(static_cast<int>(threadIdx.x) & 31) {}
  QMG_D void sync() const { __syncwarp(); }
  QMG_D uint32_t ballot(bool p) const { return __ballot_sync(0xffffffffU, p); }
  QMG_D bool any(bool p) const { return __any_sync(0xffffffffU, p); }
  template <class T> QMG_D T bcast(T v, int src) const {
    return __shfl_sync(0xffffffffU, v, src);
  }
  template <class T> QMG_D T down(T v, int d) const {
    return __shfl_down_sync(0xffffffffU, v, d);
  }
  template <class T> QMG_D T xorv(T v, int d) const {
    return __shfl_xor_sync(0xffffffffU, v, d);
  }
  // lanes (of those with p set) that hold the same value; 0 for a lane
  // without p.  Called by the whole group.
  QMG_D uint32_t matchAny(bool p, uint64_t v) const {
    const uint32_t mask = __ballot_sync(0xffffffffU, p);
    uint32_t r = 0;
    if (p) {
      r = __match_any_sync(mask, v);
    }
    return r;
  }
};
#else
struct Group {
  static constexpr int W = 1;
  int lane = 0;
  void sync() const {}
  uint32_t ballot(bool p) const { return p ? 1U : 0U; }
  bool any(bool p) const { return p; }
  template <class T> T bcast(T v, int) const { return v; }
  template <class T> T down(T v, int) const { return v; }
  template <class T> T xorv(T v, int) const { return v; }
  uint32_t matchAny(bool p, uint64_t) const { return p ? 1U : 0U; }
};
#endif

template <class T> QMG_HD T ldg(const T* p) {
#if defined(__CUDA_ARCH__)
  return __ldg(p);
#else
  return *p;
#endif
}
// circuit tables are read with plain (coherent) loads: they may be the
// search's private copy, which the search rewrites when it splits a layer
template <class T> QMG_HD T ldc(const T* p) { return *p; }
// node pools, open lists and tables are reached through pointers kept in
// shared memory, which the compiler must treat as generic addresses; they
// always point into the launch's arenas in global memory
template <class T> QMG_HD T* glob(T* p) {
#if defined(__CUDA_ARCH__) && !defined(QMG_NO_GLOBAL_HINT)
  __builtin_assume(__isGlobal(p));
#endif
  return p;
}
QMG_HD Gate2 ldGate(const Gate2* p) {
#if defined(__CUDA_ARCH__)
  const uint2 v = *reinterpret_cast<const uint2*>(p);
  Gate2 g;
  g.q1 = static_cast<uint16_t>(v.x & 0xffffU);
  g.q2 = static_cast<uint16_t>(v.x >> 16);
  g.fwd = static_cast<uint16_t>(v.y & 0xffffU);
  g.rev = static_cast<uint16_t>(v.y >> 16);
  return g;
#else
  return *p;
#endif
}

QMG_HD int ctz32(uint32_t v) {
#if defined(__CUDA_ARCH__)
  return __ffs(static_cast<int>(v)) - 1;
#else
  return __builtin_ctz(v);
#endif
}
QMG_HD int popc32(uint32_t v) {
#if defined(__CUDA_ARCH__)
  return __popc(v);
#else
  return __builtin_popcount(v);
#endif
}
QMG_HD Gate1 ldGate1(const Gate1* p) {
#if defined(__CUDA_ARCH__)
  const uint32_t v = *reinterpret_cast<const uint32_t*>(p);
  Gate1 g;
  g.q = static_cast<uint16_t>(v & 0xffffU);
  g.mult = static_cast<uint16_t>(v >> 16);
  return g;
#else
  return *p;
#endif
}

// 64-bit words of int16 arrays (8-byte aligned)
QMG_HD uint64_t ld64(const int16_t* p) {
#if defined(__CUDA_ARCH__)
  return *reinterpret_cast<const uint64_t*>(p);
#else
  uint64_t v;
  memcpy(&v, p, 8);
  return v;
#endif
}
QMG_HD void st64(int16_t* p, uint64_t v) {
#if defined(__CUDA_ARCH__)
  *reinterpret_cast<uint64_t*>(p) = v;
#else
  memcpy(p, &v, 8);
#endif
}

QMG_HD_RED double gminV(double v) {
  const Group g;
  for (int d = Group::W / 2; d > 0; d >>= 1) {
    const double o = g.xorv(v, d);
    v = o < v ? o : v;
  }
  return v;
}
QMG_HD_RED double gmaxV(double v) {
  const Group g;
  for (int d = Group::W / 2; d > 0; d >>= 1) {
    const double o = g.xorv(v, d);
    v = o > v ? o : v;
  }
  return v;
}
QMG_HD_RED uint64_t gsum64V(uint64_t v) {
  const Group g;
  for (int d = Group::W / 2; d > 0; d >>= 1) {
    v += g.xorv(v, d);
  }
  return v;
}
QMG_HD_RED uint32_t gmaxU32V(uint32_t v) {
  const Group g;
  for (int d = Group::W / 2; d > 0; d >>= 1) {
    const uint32_t o = g.xorv(v, d);
    v = o > v ? o : v;
  }
  return v;
}
QMG_HD_RED uint64_t gmaxU64V(uint64_t v) {
  const Group g;
  for (int d = Group::W / 2; d > 0; d >>= 1) {
    const uint64_t o = g.xorv(v, d);
    v = o > v ? o : v;
  }
  return v;
}
QMG_HD double gmin(const Group&, double v) { return gminV(v); }
QMG_HD double gmax(const Group&, double v) { return gmaxV(v); }
QMG_HD uint64_t gsum64(const Group&, uint64_t v) { return gsum64V(v); }
// (value, index) argmin; ties -> smallest index ("first strict minimum")
QMG_HD void gargmin(const Group& g, double& v, uint32_t& i) {
  for (int d = Group::W / 2; d > 0; d >>= 1) {
    const double ov = g.xorv(v, d);
    const uint32_t oi = g.xorv(i, d);
    if (ov < v || (ov == v && oi < i)) {
      v = ov;
      i = oi;
    }
  }
}

// ---------------------------------------------------------------------------
// per-layer derived tables and layer splits
// ---------------------------------------------------------------------------
// getConsideredQubits(l) (HeuristicMapper.hpp:228-233: the qubits of the 2q
// gates; of all gates for the fidelity-aware heuristic), ascending and unique,
// written at cons + 2*g2Off[l] + g1Off[l]; Mapper::isLayerSplittable
// (src/sc/Mapper.cpp:283-300).  Executed by ONE thread.
QMG_HD void deriveLayerTables(const uint32_t* g2Off, const Gate2* g2,
                              const uint32_t* g1Off, const Gate1* g1,
                              uint16_t* cons, uint16_t* consCnt,
                              uint8_t* splittable, uint32_t l, bool fidAware) {
  const uint32_t a = g2Off[l], b = g2Off[l + 1];
  const uint32_t a1 = g1Off[l], b1 = g1Off[l + 1];
  uint16_t* out = cons + (2U * a + a1);
  uint32_t cnt = 0;
  // sorted insertion without duplicates (the lists are a handful long)
  auto put = [&](uint16_t q) {
    uint32_t k = 0;
    while (k < cnt && out[k] < q) {
      ++k;
    }
    if (k < cnt && out[k] == q) {
      return;
    }
    for (uint32_t m = cnt; m > k; --m) {
      out[m] = out[m - 1];
    }
    out[k] = q;
    ++cnt;
  };
  for (uint32_t i = a; i < b; ++i) {
    const Gate2 e = g2[i];
    put(e.q1);
    put(e.q2);
  }
  const uint32_t G = b - a, n1 = b1 - a1;
  bool split = G > 1 || n1 > 2;
  if (!split && G == 1) {
    for (uint32_t i = a1; i < b1; ++i) {
      const uint16_t q = g1[i].q;
      bool in2 = false;
      for (uint32_t k = 0; k < cnt; ++k) {
        in2 = in2 || out[k] == q;
      }
      split = split || !in2;
    }
  }
  if (fidAware) {
    for (uint32_t i = a1; i < b1; ++i) {
      put(g1[i].q);
    }
  }
  consCnt[l] = static_cast<uint16_t>(cnt);
  splittable[l] = split ? 1 : 0;
}

// the writable copy of one instance's circuit tables, carved from
// SearchMem::priv for capacities (layers, 2q gate entries, 1q gate entries)
struct PrivView {
  uint32_t* active;  // != 0 while the search reads the private copy
  CircView* saved;   // the shared view, to restore after the pass
  uint32_t *g2Off, *g1Off, *twoQ, *firstAfter;
  Gate2 *g2, *tmp2;
  Gate1 *g1, *tmp1;
  uint16_t *cons, *consCnt;
  uint8_t* splittable;
  uint32_t* ranges;  // [2e], [2e+1]: 2q gate range of the layer at epoch e
  size_t bytes;
};
constexpr uint32_t kMaxEpochs = 64;
QMG_HD PrivView privView(uint8_t* base, uint32_t capL, uint32_t capG2,
                         uint32_t capG1) {
  PrivView v;
  size_t o = 0;
  auto take = [&](size_t bytes) {
    uint8_t* p = base + o;
    o += (bytes + 15U) & ~static_cast<size_t>(15U);
    return p;
  };
  v.active = reinterpret_cast<uint32_t*>(take(16));
  v.saved = reinterpret_cast<CircView*>(take(sizeof(CircView)));
  v.g2Off = reinterpret_cast<uint32_t*>(take(4U * (capL + 2U)));
  v.g1Off = reinterpret_cast<uint32_t*>(take(4U * (capL + 2U)));
  v.twoQ = reinterpret_cast<uint32_t*>(take(4U * (capL + 1U)));
  v.firstAfter = reinterpret_cast<uint32_t*>(take(4U * (capL + 1U)));
  v.g2 = reinterpret_cast<Gate2*>(take(sizeof(Gate2) * capG2));
  v.tmp2 = reinterpret_cast<Gate2*>(take(sizeof(Gate2) * capG2));
  v.g1 = reinterpret_cast<Gate1*>(take(sizeof(Gate1) * capG1));
  v.tmp1 = reinterpret_cast<Gate1*>(take(sizeof(Gate1) * capG1));
  v.cons = reinterpret_cast<uint16_t*>(take(2U * (2U * capG2 + capG1)));
  v.consCnt = reinterpret_cast<uint16_t*>(take(2U * (capL + 1U)));
  v.splittable = take(capL + 1U);
  v.ranges = reinterpret_cast<uint32_t*>(take(8U * kMaxEpochs));
  v.bytes = o;
  return v;
}

inline size_t privBytes(uint32_t capL, uint32_t capG2, uint32_t capG1) {
  // (only the offsets are wanted: any aligned address will do)
  return privView(reinterpret_cast<uint8_t*>(static_cast<uintptr_t>(4096)), capL,
                  capG2, capG1)
      .bytes;
}

// switch the search to a private copy of its circuit tables
QMG_COLD void privActivate(CircView* cv, uint8_t* priv, uint32_t capL,
                           uint32_t capG2, uint32_t capG1, Group g) {
  const PrivView p = privView(priv, capL, capG2, capG1);
  const uint32_t nL = cv->nLayers;
  const uint32_t nG2 = cv->g2Off[nL], nG1 = cv->g1Off[nL];
  for (uint32_t i = g.lane; i <= nL; i += Group::W) {
    p.g2Off[i] = cv->g2Off[i];
    p.g1Off[i] = cv->g1Off[i];
  }
  for (uint32_t i = g.lane; i < nL; i += Group::W) {
    p.firstAfter[i] = cv->firstAfter[i];
    p.consCnt[i] = cv->consCnt[i];
    p.splittable[i] = cv->splittable[i];
  }
  for (uint32_t i = g.lane; i < cv->nTwoQ; i += Group::W) {
    p.twoQ[i] = cv->twoQ[i];
  }
  for (uint32_t i = g.lane; i < nG2; i += Group::W) {
    p.g2[i] = cv->g2[i];
  }
  for (uint32_t i = g.lane; i < nG1; i += Group::W) {
    p.g1[i] = cv->g1[i];
  }
  for (uint32_t i = g.lane; i < 2U * nG2 + nG1; i += Group::W) {
    p.cons[i] = cv->cons[i];
  }
  g.sync();
  if (g.lane == 0) {
    *p.saved = *cv;
    cv->g2Off = p.g2Off;
    cv->g1Off = p.g1Off;
    cv->g2 = p.g2;
    cv->g1 = p.g1;
    cv->cons = p.cons;
    cv->consCnt = p.consCnt;
    cv->twoQ = p.twoQ;
    cv->firstAfter = p.firstAfter;
    cv->splittable = p.splittable;
    *p.active = 1;
  }
  g.sync();
}

// back to the shared tables (pseudoRouteCircuit restores the layers it split,
// HeuristicMapper.cpp:309-344)
QMG_COLD void privRestore(CircView* cv, uint8_t* priv, uint32_t capL,
                          uint32_t capG2, uint32_t capG1, Group g) {
  const PrivView p = privView(priv, capL, capG2, capG1);
  g.sync();
  if (g.lane == 0) {
    *cv = *p.saved;
    *p.active = 0;
  }
  g.sync();
}

// Mapper::splitLayer (src/sc/Mapper.cpp:302-435) on the private copy: layer l
// keeps every second 2q gate (the 2nd, 4th, ...), the others move to a new
// layer l+1; a 1q gate follows the 2q gate on its qubit, the remaining ones
// alternate (first one stays).  Rare (once per splitLimit expansions at most),
// so one lane does it.
QMG_COLD void privSplit(CircView* cv, uint8_t* priv, uint32_t capL,
                        uint32_t capG2, uint32_t capG1, uint32_t l,
                        bool fidAware, Group g) {
  const PrivView p = privView(priv, capL, capG2, capG1);
  g.sync();
  if (g.lane == 0) {
    const uint32_t nL = cv->nLayers;
    const uint32_t lo = p.g2Off[l], hi = p.g2Off[l + 1];
    const uint32_t lo1 = p.g1Off[l], hi1 = p.g1Off[l + 1];
    const uint32_t G = hi - lo, G1 = hi1 - lo1;
    const uint32_t nKeep = G / 2U;
    for (uint32_t i = 0; i < G; ++i) {
      p.tmp2[i] = p.g2[lo + i];
    }
    for (uint32_t i = 0; i < G; ++i) {
      const uint32_t dst = (i & 1U) != 0 ? i / 2U : nKeep + i / 2U;
      p.g2[lo + dst] = p.tmp2[i];
    }
    // side of every 1q gate: 0 stays, 1 moves (kept in the layer's cons
    // region, which is rebuilt below)
    uint16_t* side = p.cons + (2U * lo + lo1);
    bool stay = true;
    uint32_t nKeep1 = 0;
    for (uint32_t i = 0; i < G1; ++i) {
      p.tmp1[i] = p.g1[lo1 + i];
      const uint16_t q = p.tmp1[i].q;
      bool inKeep = false, inMoved = false;
      for (uint32_t k = 0; k < G; ++k) {
        const bool hit = p.tmp2[k].q1 == q || p.tmp2[k].q2 == q;
        if ((k & 1U) != 0) {
          inKeep = inKeep || hit;
        } else {
          inMoved = inMoved || hit;
        }
      }
      bool toKeep;
      if (inKeep) {
        toKeep = true;
      } else if (inMoved) {
        toKeep = false;
      } else {
        toKeep = stay;
        stay = !stay;
      }
      side[i] = toKeep ? 0 : 1;
      nKeep1 += toKeep ? 1U : 0U;
    }
    uint32_t k0 = 0, k1 = 0;
    for (uint32_t i = 0; i < G1; ++i) {
      if (side[i] == 0) {
        p.g1[lo1 + k0++] = p.tmp1[i];
      } else {
        p.g1[lo1 + nKeep1 + k1++] = p.tmp1[i];
      }
    }
    // insert layer l+1
    for (uint32_t k = nL; k > l; --k) {
      p.g2Off[k + 1] = p.g2Off[k];
      p.g1Off[k + 1] = p.g1Off[k];
    }
    p.g2Off[l + 1] = lo + nKeep;
    p.g1Off[l + 1] = lo1 + nKeep1;
    for (uint32_t k = nL; k > l + 1; --k) {
      p.consCnt[k] = p.consCnt[k - 1];
      p.splittable[k] = p.splittable[k - 1];
    }
    deriveLayerTables(p.g2Off, p.g2, p.g1Off, p.g1, p.cons, p.consCnt,
                      p.splittable, l, fidAware);
    deriveLayerTables(p.g2Off, p.g2, p.g1Off, p.g1, p.cons, p.consCnt,
                      p.splittable, l + 1, fidAware);
    // Mapper::getNextLayer (Mapper.cpp:437-448) in closed form
    uint32_t t = 0;
    for (uint32_t k = 0; k <= nL; ++k) {
      const bool has = p.g2Off[k + 1] > p.g2Off[k];
      if (has) {
        p.twoQ[t] = k;
        ++t;
      }
      p.firstAfter[k] = t;
    }
    cv->nTwoQ = t;
    cv->nLayers = nL + 1;
  }
  g.sync();
}

// ---------------------------------------------------------------------------
// the search of one instance
// ---------------------------------------------------------------------------
// H = qmg_heuristic (0..4): one kernel instantiation per heuristic keeps the
// resident code of a launch small (the instruction cache, not the data path,
// bounds this kernel)
// kFull = false is the variant the common case runs: open lists are scanned
// flat and layers are never split on the device, which keeps its resident
// code and register use small (the kernel is instruction-fetch bound).  It
// gives up with ST_NEEDS_FULL as soon as an open list outgrows kBlkMin or a
// layer has to be split; the caller then routes the instance again with
// kFull = true (per-block tops of the open list, Mapper::splitLayer and the
// restart on the stale open list).
template <int H, bool kFull> struct Search {
  static constexpr bool kFid = (H == 4);
  static constexpr bool kTight = !kFid;  // isTight, Heuristic.hpp:96-110
  const ArchView& A;
  const CfgView& C;
  InstView& I;  // (the circuit view is re-pointed when a layer is split)
  SearchMem& M;
  Scratch& S;
  Group g;
  int n;

  // state of the running aStarMap call
  uint32_t layer = 0;
  uint32_t nPool = 0, nOpen = 0, nextId = 0;
  uint32_t G = 0;  // 2q gates in the layer
  const Gate2* gates = nullptr;
  const Gate1* gates1 = nullptr;
  uint32_t G1 = 0;
  const uint16_t* cons = nullptr;
  uint32_t nCons = 0;
  uint32_t fidK = 0;  // table index |considered|-1 (kNone => cost 0)
  uint32_t nLook = 0;  // lookahead layers available from this layer
  const uint32_t* lookLayers = nullptr;
  int16_t* pQ = nullptr;  // layout the scoring functions read
  int16_t* pL = nullptr;
  bool tableMode = false;
  bool blockMode = false;  // M.btop holds the top of every open-list block
  bool maskReady = false;
  bool allowSplit = false;  // this aStarMap call may stop at the split limit
  uint64_t totExpanded = 0, totGenerated = 0;
  uint32_t nLog = 0;      // nodes logged so far (InstView::nodeLog)
  uint32_t parentId = 0;  // Node::id of the node being expanded
  uint64_t pseudoExpanded = 0, pseudoGenerated = 0;  // of the pseudo passes

  QMG_HD Search(const ArchView& a, const CfgView& c, InstView& i,
                SearchMem& m, Scratch& s)
      : A(a), C(c), I(i), M(m), S(s), n(a.n) {}

  // ---- table accessors (include/sc/Architecture.hpp:244-256,301-316,427-433)
  QMG_HD bool adjU(int a, int b) const { return ldg(A.adj + a * n + b) != 0; }
  QMG_HD bool adjD(int a, int b) const {
    return (ldg(A.adj + a * n + b) & 1U) != 0;
  }
  QMG_HD double dRev(int a, int b) const { return ldg(A.distRev + a * n + b); }
  QMG_HD double dPlain(int a, int b) const { return ldg(A.dist + a * n + b); }
  QMG_HD double fidDist(int a, int b, uint32_t k) const {
    if (k >= static_cast<uint32_t>(A.nFid)) {
      return 0.;
    }
    return ldg(A.fid + (static_cast<size_t>(k) * n + a) * n + b);
  }
  QMG_HD double swapCostGC(int a, int b) const {
    // applySWAP :957-964 / recalculateFixedCostNonFidelity :807-815
    if (A.bidirectional) {
      return kCostBiSwap;
    }
    if (A.unidirectional || ldg(A.adj + a * n + b) != 3U) {
      return kCostUniSwap;
    }
    return kCostBiSwap;
  }

  QMG_HD NodeHdr* hdr(uint32_t idx) const {
    return reinterpret_cast<NodeHdr*>(glob(M.pool) +
                                      static_cast<size_t>(idx) * M.recBytes);
  }
  QMG_HD int16_t* nodeQ(uint32_t idx) const {
    return reinterpret_cast<int16_t*>(
        glob(M.pool) + static_cast<size_t>(idx) * M.recBytes + sizeof(NodeHdr));
  }
  QMG_HD int16_t* nodeL(uint32_t idx) const { return nodeQ(idx) + padN(n); }
  QMG_HD bool considered(int q) const {
    return (S.consMask()[q >> 5] >> (q & 31)) & 1U;
  }

  // ------------------------------------------------------------------------
  // scoring of one node: the parent layout (S.Q() / S.L()) with the SWAP on the
  // physical edge (a, b) applied; a == b scores the layout itself (root).
  // Executed by ONE lane (gate-count heuristics).
  // ------------------------------------------------------------------------
  struct ParentCosts {
    double fixed;
    uint32_t shared;
  };

  QMG_HD int locAfter(int q, int q1, int q2, int a, int b) const {
    if (q == q1) {
      return b;
    }
    if (q == q2) {
      return a;
    }
    return pL[q];
  }

  // updateSharedSwaps :975-1053 (state before the swap)
  QMG_HD uint32_t sharedSwapDelta(int a, int b, int q1, int q2) const {
    if (q1 < 0 || q2 < 0 || !considered(q1) || !considered(q2)) {
      return 0;
    }
    int p1 = q1;
    int p2 = q2;
    QMG_NOUNROLL
    for (uint32_t i = 0; i < G; ++i) {
      const Gate2 e = ldGate(gates + i);
      if (e.q1 == q1) {
        p1 = e.q2;
      } else if (e.q2 == q1) {
        p1 = e.q1;
      }
      if (e.q1 == q2) {
        p2 = e.q2;
      } else if (e.q2 == q2) {
        p2 = e.q1;
      }
    }
    if (p1 != q1 && p2 != q2 && p1 != q2) {
      const int physQ3 = pL[p1];
      const int physQ4 = pL[p2];
      double e1b, e1n, e2b, e2n;
      if (kFid) {
        e1b = fmin(fidDist(a, physQ3, 0), fidDist(physQ3, a, 0));
        e1n = fmin(fidDist(b, physQ3, 0), fidDist(physQ3, b, 0));
        e2b = fmin(fidDist(b, physQ4, 0), fidDist(physQ4, b, 0));
        e2n = fmin(fidDist(a, physQ4, 0), fidDist(physQ4, a, 0));
      } else {
        e1b = fmin(dPlain(a, physQ3), dPlain(physQ3, a));
        e1n = fmin(dPlain(b, physQ3), dPlain(physQ3, b));
        e2b = fmin(dPlain(b, physQ4), dPlain(physQ4, b));
        e2n = fmin(dPlain(a, physQ4), dPlain(physQ4, a));
      }
      if (e1n < e1b && e2n < e2b) {
        return 1;
      }
    }
    return 0;
  }

  // heuristicGateCountSumDistanceMinusSharedSwaps :1166-1253 given the
  // per-gate swap counts in nsw[0..cnt) (unsorted)
  QMG_HD double sharedSwapReduction(uint16_t* nsw, uint32_t cnt,
                                    uint32_t shared) const {
    // std::sort(..., greater): insertion sort on small per-lane arrays
    QMG_NOUNROLL
    for (uint32_t i = 1; i < cnt; ++i) {
      const uint16_t v = nsw[i];
      uint32_t j = i;
      QMG_NOUNROLL
      while (j > 0 && nsw[j - 1] < v) {
        nsw[j] = nsw[j - 1];
        --j;
      }
      nsw[j] = v;
    }
    uint64_t maxShared = 0;
    QMG_NOUNROLL
    for (uint32_t i = 0; i + 1 < cnt; ++i) {
      uint32_t edgeShared = 0;
      QMG_NOUNROLL
      for (uint32_t j = i + 1; j < cnt && edgeShared < nsw[i]; ++j) {
        if (nsw[j] > 0) {
          ++edgeShared;
          --nsw[j];
        }
      }
      maxShared += edgeShared;
    }
    if (shared < maxShared) {
      maxShared -= shared;
    } else {
      maxShared = 0;
    }
    return static_cast<double>(
        maxShared * (A.bidirectional ? static_cast<uint64_t>(30)
                                     : static_cast<uint64_t>(34)));
  }

  QMG_HD void scoreGateCount(int a, int b, const ParentCosts& pc, Child& out,
                             uint16_t* nsw) const {
    const bool root = (a == b);
    const int q1 = root ? -2 : pQ[a];
    const int q2 = root ? -2 : pQ[b];
    out.a = static_cast<uint16_t>(a);
    out.b = static_cast<uint16_t>(b);
    constexpr int h = H;
    // Node::sharedSwaps is only read by heuristics 2 and 3 (:1237-1241)
    out.shared = pc.shared;
    if (!root && (h == 2 || h == 3)) {
      out.shared += sharedSwapDelta(a, b, q1, q2);
    }
    out.fixed = root ? 0. : pc.fixed + swapCostGC(a, b);

    const bool wantMax = (h == 0 || h == 3);
    const bool wantSum = (h == 1);
    const bool wantShared = (h == 2 || h == 3);
    uint32_t nValid = (root || !kFull) ? 0U : M.childExtra;
    uint32_t revFixed = 0;  // recalculateFixedCostReversals :771-796
    double hMax = 0.;
    double hSum = 0.;
    double hShared = 0.;
    double costReversals = 0.;
    uint32_t cnt = 0;
    QMG_NOUNROLL
    for (uint32_t i = 0; i < G; ++i) {
      const Gate2 e = ldGate(gates + i);
      const int p1 = locAfter(e.q1, q1, q2, a, b);
      const int p2 = locAfter(e.q2, q1, q2, a, b);
      const uint32_t ad = ldg(A.adj + p1 * n + p2);
      const bool valid = ad != 0;
      nValid += valid ? 1U : 0U;
      if (!A.bidirectional) {
        if (valid) {
          uint32_t r = 0;
          if ((ad & 1U) == 0) {
            r = e.fwd * kCostReverse;
          } else if ((ad & 2U) == 0) {
            r = e.rev * kCostReverse;
          }
          revFixed += r;
          if (wantMax) {
            hMax = fmax(hMax, static_cast<double>(r));
          }
          if (wantSum) {
            hSum += r;
          }
        } else {
          // only read when stale entries make the set as large as the layer
          // although this gate is not adjacent (:771-796 then charges it)
          revFixed += e.fwd * kCostReverse;
        }
      }
      if (wantShared && A.unidirectional) {
        costReversals += (e.fwd < e.rev ? e.fwd : e.rev) * kCostReverse;
      }
      if (!valid) {
        if (wantMax) {
          if (e.fwd > 0) {
            hMax = fmax(hMax, dRev(p1, p2));
          }
          if (e.rev > 0) {
            hMax = fmax(hMax, dRev(p2, p1));
          }
        }
        if (wantSum) {
          double sc;
          if (e.fwd == 0) {
            sc = dRev(p2, p1);
          } else if (e.rev == 0) {
            sc = dRev(p1, p2);
          } else {
            sc = fmax(dRev(p1, p2), dRev(p2, p1));
          }
          hSum += sc;
        }
        if (wantShared) {
          double sc;
          if (e.fwd == 0) {
            sc = dPlain(p2, p1);
          } else if (e.rev == 0) {
            sc = dPlain(p1, p2);
          } else {
            sc = fmin(dPlain(p1, p2), dPlain(p2, p1));
          }
          hShared += sc;
          nsw[cnt++] = static_cast<uint16_t>(static_cast<uint64_t>(
              sc / (A.unidirectional ? kCostUniSwap : kCostBiSwap)));
        }
      }
    }
    // (on a bidirectional architecture the reference also feeds validly
    // mapped gates through the distance branch, :1108-1116; adjacent qubits
    // have edge-skip distance 0, so skipping them here is equivalent)
    const bool validMapping = (nValid == G);
    out.nValid = static_cast<uint16_t>(nValid);
    out.valid = validMapping ? 1 : 0;
    out.rev = 0.;
    if (!A.bidirectional && validMapping) {
      out.rev = static_cast<double>(revFixed);
    }
    double heur = 0.;
    if (!validMapping) {
      double hs = 0.;
      if (wantShared) {
        const double red = sharedSwapReduction(nsw, cnt, out.shared);
        hs = fmax(0., hShared - red) + costReversals;
      }
      if (h == 0) {
        heur = hMax;
      } else if (h == 1) {
        heur = hSum;
      } else if (h == 2) {
        heur = hs;
      } else {
        heur = fmax(hMax, hs);
      }
    }
    out.heur = heur;
  }

  // the contribution of ONE gate of a later layer to the lookahead penalty
  // (:1387-1513) under the layout "parent with swap (a,b)" (a < 0: the parent
  // itself).  Returns false when both operands are unplaced (no penalty).
  QMG_HD bool lookaheadGateCost(const Gate2& e, int a, int b, int q1, int q2,
                                double& cost) const {
    const int loc1 = locAfter(e.q1, q1, q2, a, b);
    const int loc2 = locAfter(e.q2, q1, q2, a, b);
    if (loc1 < 0 && loc2 < 0) {
      return false;
    }
    if (loc1 < 0 || loc2 < 0) {
      // the placed partner; the other one may go to any free qubit
      const bool firstFree = loc1 < 0;
      const int loc = firstFree ? loc2 : loc1;
      double mn = DBL_MAX;
      QMG_NOUNROLL
      for (int j = 0; j < n; ++j) {
        const int qj = (j == a) ? q2 : ((j == b) ? q1 : pQ[j]);
        if (qj != -1) {
          continue;
        }
        // :1405-1412 (loc1 free) and :1421-1428 (loc2 free)
        if (e.fwd > 0) {
          mn = fmin(mn, firstFree ? dRev(j, loc) : dRev(loc, j));
        }
        if (e.rev > 0) {
          mn = fmin(mn, firstFree ? dRev(loc, j) : dRev(j, loc));
        }
      }
      cost = mn;
    } else {
      cost = DBL_MAX;
      if (e.fwd > 0) {
        cost = fmin(cost, dRev(loc1, loc2));
      }
      if (e.rev > 0) {
        cost = fmin(cost, dRev(loc2, loc1));
      }
    }
    return true;
  }

  // lookaheadGateCountMaxDistance / SumDistance :1387-1513 for one later
  // layer and one child layout.  Executed by ONE lane.
  QMG_HD double lookaheadPenalty(uint32_t ll, int a, int b, int q1,
                                 int q2) const {
    const uint32_t lo = ldc(I.circ.g2Off + ll);
    const uint32_t hi = ldc(I.circ.g2Off + ll + 1);
    const bool isMax = (C.lookahead == 1);
    double penalty = 0.;
    QMG_NOUNROLL
    for (uint32_t i = lo; i < hi; ++i) {
      const Gate2 e = ldGate(I.circ.g2 + i);
      double cost;
      if (!lookaheadGateCost(e, a, b, q1, q2, cost)) {
        continue;
      }
      if (isMax) {
        penalty = fmax(penalty, cost);
      } else {
        penalty += cost;
      }
    }
    return penalty;
  }

  // updateLookaheadPenalty :1356-1385 for the cnt children in S.ch() (or the
  // root when rootOnly).  Items (child, lookahead layer) are spread over the
  // lanes (16 lookahead layers x 16 children per pass); the factor-weighted
  // sum is then accumulated per child in layer order with the reference's
  // operation order (acc += factor_i * penalty_i).
  QMG_HD void scoreLookahead(int cnt, bool rootOnly) {
    if (C.lookahead == 0 || nLook == 0) {
      QMG_NOUNROLL
      for (int c = g.lane; c < cnt; c += Group::W) {
        S.ch()[c].look = 0.;
      }
      g.sync();
      return;
    }
    constexpr int kHalf = kChunk / 2;
    QMG_NOUNROLL
    for (int c0 = 0; c0 < cnt; c0 += kHalf) {
      const int nc = (cnt - c0) < kHalf ? (cnt - c0) : kHalf;
      QMG_NOUNROLL
      for (int c = g.lane; c < nc; c += Group::W) {
        S.ch()[c0 + c].look = 0.;
      }
      QMG_NOUNROLL
      for (uint32_t s0 = 0; s0 < nLook; s0 += kLookBlock) {
        const uint32_t nb =
            (nLook - s0) < static_cast<uint32_t>(kLookBlock)
                ? (nLook - s0)
                : static_cast<uint32_t>(kLookBlock);
        g.sync();
        QMG_NOUNROLL
        for (int it = g.lane; it < nc * kLookBlock; it += Group::W) {
          const int c = it / kLookBlock;
          const uint32_t sl = static_cast<uint32_t>(it % kLookBlock);
          if (sl >= nb) {
            continue;
          }
          int a = -1, b = -1, q1 = -2, q2 = -2;
          if (!rootOnly) {
            const Child& ch = S.ch()[c0 + c];
            a = ch.a;
            b = ch.b;
            q1 = pQ[a];
            q2 = pQ[b];
          }
          // factor_i * penalty_i (rounded as in the reference, then summed
          // in layer order below)
          S.lookPen()[it] =
              ldg(C.factors + s0 + sl) *
              lookaheadPenalty(ldc(lookLayers + s0 + sl), a, b, q1, q2);
        }
        g.sync();
        QMG_NOUNROLL
        for (int c = g.lane; c < nc; c += Group::W) {
          double acc = S.ch()[c0 + c].look;
          QMG_NOUNROLL
          for (uint32_t sl = 0; sl < nb; ++sl) {
            acc += S.lookPen()[c * kLookBlock + sl];
          }
          S.ch()[c0 + c].look = acc;
        }
      }
    }
    g.sync();
  }

  // ------------------------------------------------------------------------
  // fidelity-aware scoring (recalculateFixedCostFidelity :821-859, the
  // fidelity branches of applySWAP :893-952, heuristicFidelityBestLocation
  // :1262-1354).
  // ------------------------------------------------------------------------
  // identifies the running aStarMap call (and its restarts after a split)
  QMG_HD uint64_t pairKey() const {
    return (static_cast<uint64_t>(M.stamp) << 8) | (M.epoch + 1U);
  }
  // Scores the cnt nodes of S.ch() (root: the layout itself), ONE CHILD PER
  // LANE: everything that is a chain of dependent table reads (validity of the
  // gates, fixed-cost deltas, memo look-ups) runs for all children at once;
  // the scans over physical qubits / coupling edges are done by the whole
  // group for one child at a time, and only for children that need them (the
  // best-edge term is usually found in the memo).  Sums are accumulated per
  // child in the reference's order.
  QMG_HD void scoreFidelityChunk(int cnt, bool root, const ParentCosts& pc) {
    QMG_NOUNROLL
    for (int c0 = 0; c0 < cnt; c0 += Group::W) {
      const int c = c0 + g.lane;
      const bool act = c < cnt;
      int a = -1, b = -1, q1 = -2, q2 = -2;
      if (act && !root) {
        a = S.ch()[c].a;
        b = S.ch()[c].b;
        q1 = pQ[a];
        q2 = pQ[b];
      }
      double fixed = root ? 0. : pc.fixed;
      uint32_t nValid = (root || !kFull) ? 0U : M.childExtra;
      if (act) {
        if (root) {
          QMG_NOUNROLL
          for (uint32_t i = 0; i < G1; ++i) {
            const Gate1 e = ldGate1(gates1 + i);
            fixed += e.mult * ldg(A.c1 + pL[e.q]);
          }
        }
        QMG_NOUNROLL
        for (uint32_t i = 0; i < G; ++i) {
          const Gate2 e = ldGate(gates + i);
          const int p3 = locAfter(e.q1, q1, q2, a, b);
          const int p4 = locAfter(e.q2, q1, q2, a, b);
          const bool valid = adjU(p3, p4);
          nValid += valid ? 1U : 0U;
          if (root) {
            if (valid) {
              fixed += (e.fwd * ldg(A.c2 + p3 * n + p4) +
                        e.rev * ldg(A.c2 + p4 * n + p3));
            }
            continue;
          }
          const bool touched =
              (e.q1 == q1 || e.q1 == q2 || e.q2 == q1 || e.q2 == q2);
          if (!touched) {
            continue;
          }
          const int o3 = pL[e.q1];
          const int o4 = pL[e.q2];
          const bool before = adjU(o3, o4);
          if (valid && !before) {
            fixed += e.fwd * ldg(A.c2 + p3 * n + p4) +
                     e.rev * ldg(A.c2 + p4 * n + p3);
          } else if (!valid && before) {
            fixed -= e.fwd * ldg(A.c2 + o3 * n + o4) +
                     e.rev * ldg(A.c2 + o4 * n + o3);
          }
        }
        if (!root) {
          // :934-952
          int m1 = 0;
          int m2 = 0;
          QMG_NOUNROLL
          for (uint32_t i = 0; i < G1; ++i) {
            const Gate1 e = ldGate1(gates1 + i);
            if (e.q == q1) {
              m1 = e.mult;
            }
            if (e.q == q2) {
              m2 = e.mult;
            }
          }
          fixed += ((m2 - m1) * ldg(A.c1 + a) + (m1 - m2) * ldg(A.c1 + b));
          fixed += ldg(A.cswap + a * n + b);
        }
      }

      // heuristicFidelityBestLocation
      double costHeur = 0.;
      double savings = 0.;
      QMG_NOUNROLL
      for (uint32_t i = 0; i < G1; ++i) {
        const Gate1 e = ldGate1(gates1 + i);
        int loc = 0;
        double curr = 0.;
        if (act) {
          loc = locAfter(e.q, q1, q2, a, b);
          curr = ldg(A.c1 + loc);
        }
        double best = 0.;
        uint32_t m = g.ballot(act);
        QMG_NOUNROLL
        while (m != 0) {
          const int src = ctz32(m);
          m &= m - 1;
          const int sLoc = g.bcast(loc, src);
          const double sCurr = g.bcast(curr, src);
          double v = 0.;
          QMG_UNROLL(QMG_SCAN1_UNROLL)
          for (int p = g.lane; p < n; p += Group::W) {
            const double cp = ldg(A.c1 + p);
            if (cp >= sCurr) {
              continue;
            }
            const double cur = e.mult * (sCurr - cp) - fidDist(sLoc, p, fidK);
            v = fmax(v, cur);
          }
          v = gmax(g, v);
          if (g.lane == src) {
            best = v;
          }
        }
        savings += best;
      }
      QMG_NOUNROLL
      for (uint32_t i = 0; i < G; ++i) {
        const Gate2 e = ldGate(gates + i);
        int l1 = 0, l2 = 0;
        if (act) {
          l1 = locAfter(e.q1, q1, q2, a, b);
          l2 = locAfter(e.q2, q1, q2, a, b);
        }
        double sc = DBL_MAX;
        // the minimum over the coupling edges depends on (l1, l2) only (gate
        // multiplicities and table index are fixed for the call): long
        // searches meet the same pair of locations in many nodes, so the full
        // variant keeps the value per pair - the same double, computed once
        bool hit = false;
        size_t slot = 0;
        const bool memo = kFull && M.pair != nullptr && i < kPairGates;
        const uint64_t key = pairKey();
        if (memo && act) {
          slot = (static_cast<size_t>(i) * n + l1) * n + l2;
          const PairSlot ps = glob(M.pair)[slot];
          if (ps.tag == key) {
            sc = ps.val;
            hit = true;
          }
        }
        uint32_t m = g.ballot(act && !hit);
        QMG_NOUNROLL
        while (m != 0) {
          const int src = ctz32(m);
          m &= m - 1;
          const int s1 = g.bcast(l1, src);
          const int s2 = g.bcast(l2, src);
          double v = DBL_MAX;
          // (iterations are independent gathers: unrolled so that they overlap)
          QMG_UNROLL(QMG_SCAN2_UNROLL)
          for (int k = g.lane; k < A.nE; k += Group::W) {
            const int q3 = ldg(A.edges + 2 * k);
            const int q4 = ldg(A.edges + 2 * k + 1);
            const double c34 = ldg(A.c2 + q3 * n + q4);
            const double c43 = ldg(A.c2 + q4 * n + q3);
            const double v1 = e.fwd * c34 + e.rev * c43 +
                              fidDist(s1, q3, fidK) + fidDist(s2, q4, fidK);
            v = fmin(v, v1);
            const double v2 = e.fwd * c43 + e.rev * c34 +
                              fidDist(s2, q3, fidK) + fidDist(s1, q4, fidK);
            v = fmin(v, v2);
          }
          v = gmin(g, v);
          if (g.lane == src) {
            sc = v;
            if (memo) {
              PairSlot ps;
              ps.tag = key;
              ps.val = v;
              glob(M.pair)[slot] = ps;
            }
          }
        }
        if (act) {
          if (adjU(l1, l2)) {
            const double currEdge = (e.fwd * ldg(A.c2 + l1 * n + l2) +
                                     e.rev * ldg(A.c2 + l2 * n + l1));
            savings += (currEdge - sc);
          } else {
            costHeur += sc;
          }
        }
      }
      if (act) {
        Child& out = S.ch()[c];
        out.fixed = fixed;
        out.rev = 0.;
        out.heur = costHeur - savings;
        out.shared = 0;
        out.nValid = static_cast<uint16_t>(nValid);
        out.valid = (nValid == G) ? 1 : 0;
      }
      // (memo slots written above are read by later passes of this loop)
      g.sync();
    }
  }

  // ------------------------------------------------------------------------
  // open list: UniquePriorityQueue semantics on (open[], pool, table[])
  // ------------------------------------------------------------------------
  // node order, HeuristicMapper.hpp:452-483, first four keys.
  // returns -1 if x has priority, +1 if y has priority, 0 if tied
  QMG_HD static int cmp4(const OpenEntry& x, const OpenEntry& y) {
    if (fabs(x.f - y.f) > kTol) {
      return x.f > y.f ? 1 : -1;
    }
    if (x.valid != y.valid) {
      return y.valid ? 1 : -1;
    }
    if (fabs(x.h2 - y.h2) > kTol) {
      return x.h2 > y.h2 ? 1 : -1;
    }
    if (x.nValid != y.nValid) {
      return x.nValid < y.nValid ? 1 : -1;
    }
    return 0;
  }
  // what the open-list order needs of the search, by value (the out-of-line
  // helpers of long open lists must not take `this`: that would force the
  // whole Search object into local memory)
  struct OpenCtx {
    const uint8_t* pool;
    uint32_t recBytes;
    int n;
  };
  QMG_HD OpenCtx octx() const { return OpenCtx{M.pool, M.recBytes, n}; }
  QMG_HD static const int16_t* nodeQOf(const OpenCtx& c, uint32_t idx) {
    return reinterpret_cast<const int16_t*>(
        glob(c.pool) + static_cast<size_t>(idx) * c.recBytes + sizeof(NodeHdr));
  }
  QMG_HD static const NodeHdr* hdrOf(const OpenCtx& c, uint32_t idx) {
    return reinterpret_cast<const NodeHdr*>(glob(c.pool) +
                                            static_cast<size_t>(idx) * c.recBytes);
  }
  // 4 consecutive qubits[] entries as one integer whose unsigned order is the
  // lexicographic order of the (signed) entries
  QMG_HD static uint64_t lexChunkOf(int n, const int16_t* q, int i0) {
    uint64_t k = 0;
    for (int t = 0; t < 4; ++t) {
      const int i = i0 + t;
      const uint32_t v =
          i < n ? (static_cast<uint32_t>(static_cast<int32_t>(q[i]) + 0x8000) &
                   0xffffU)
                : 0U;
      k = (k << 16) | v;
    }
    return k;
  }
  QMG_HD uint64_t lexChunk(const int16_t* q, int i0) const {
    return lexChunkOf(n, q, i0);
  }
  // full order incl. the lexicographic tie-break (:438-450, :482): among
  // tied nodes the lexicographically LARGEST qubits[] is on top
  QMG_HD static bool hasPriorityOf(const OpenCtx& c, const OpenEntry& x,
                                   const OpenEntry& y) {
    const int r = cmp4(x, y);
    if (r != 0) {
      return r < 0;
    }
    if (x.idx == y.idx) {
      return false;
    }
    const int16_t* qx = nodeQOf(c, x.idx);
    const int16_t* qy = nodeQOf(c, y.idx);
    QMG_NOUNROLL
    for (int i = 0; i < c.n; i += 4) {
      const uint64_t kx = lexChunkOf(c.n, qx, i);
      const uint64_t ky = lexChunkOf(c.n, qy, i);
      if (kx != ky) {
        return kx > ky;
      }
    }
    return false;
  }
  QMG_HD bool hasPriority(const OpenEntry& x, const OpenEntry& y) const {
    return hasPriorityOf(octx(), x, y);
  }

  QMG_HD uint32_t gmaxU32(uint32_t v) const { return gmaxU32V(v); }
  QMG_HD uint64_t gmaxU64(uint64_t v) const { return gmaxU64V(v); }

  // position of the top of the open list (group-uniform result).  Every lane
  // reduces its strided share with the pairwise order; the lanes' winners are
  // then narrowed key by key (cost band, validity, heuristic band, valid-gate
  // count, lexicographic layout), which selects the same node as the pairwise
  // order whenever that order is a strict weak order.
  // kWide: QMG_WIDE entries per lane are loaded before they are compared, so that
  // their memory latencies overlap (long lists: the entries come from L2/HBM)
  template <bool kWide = false>
  QMG_HD static uint32_t topOfS(const OpenCtx& c, const OpenEntry* arr,
                                uint32_t lo, uint32_t hi) {
    const Group g;
    const int n = c.n;
    OpenEntry best;
    uint32_t bestPos = kNone;
    best.f = DBL_MAX;
    best.h2 = DBL_MAX;
    best.hash = 0;
    best.idx = kNone;
    best.nValid = 0;
    best.valid = 0;
    uint32_t i = lo + g.lane;
    if (kWide) {
      constexpr uint32_t kIn = QMG_WIDE;  // entries in flight per lane
      QMG_NOUNROLL
      for (; i + (kIn - 1) * Group::W < hi; i += kIn * Group::W) {
        OpenEntry e[kIn];
        for (uint32_t u = 0; u < kIn; ++u) {
          e[u] = glob(arr)[i + u * Group::W];
        }
        for (uint32_t u = 0; u < kIn; ++u) {
          if (bestPos == kNone || hasPriorityOf(c, e[u], best)) {
            best = e[u];
            bestPos = i + u * Group::W;
          }
        }
      }
    }
    QMG_NOUNROLL
    for (; i < hi; i += Group::W) {
      const OpenEntry e = glob(arr)[i];
      if (bestPos == kNone || hasPriorityOf(c, e, best)) {
        best = e;
        bestPos = i;
      }
    }
    if (Group::W == 1) {
      return bestPos;
    }
    bool cand = bestPos != kNone;
    const double fmn = gmin(g, cand ? best.f : DBL_MAX);
    cand = cand && !(fabs(best.f - fmn) > kTol);
    uint32_t m = g.ballot(cand);
    if (popc32(m) > 1) {
      const uint32_t mv = g.ballot(cand && best.valid != 0);
      if (mv != 0) {
        cand = cand && best.valid != 0;
      }
      const double hmn = gmin(g, cand ? best.h2 : DBL_MAX);
      cand = cand && !(fabs(best.h2 - hmn) > kTol);
      const uint32_t nv = gmaxU32V(cand ? best.nValid + 1U : 0U);
      cand = cand && (best.nValid + 1U == nv);
      m = g.ballot(cand);
      const int16_t* q = cand ? nodeQOf(c, best.idx) : nullptr;
      QMG_NOUNROLL
      for (int i = 0; i < n && popc32(m) > 1; i += 4) {
        const uint64_t k = cand ? lexChunkOf(n, q, i) : 0ULL;
        const uint64_t kmx = gmaxU64V(k);
        cand = cand && k == kmx;
        m = g.ballot(cand);
      }
    }
    const int src = ctz32(m);
    return g.bcast(bestPos, src);
  }

  // ---- long open lists: tops per block (M.btop); out of line and static so
  // that neither the resident code nor the registers of the short-list path
  // change
  QMG_COLD static uint32_t blkTopOf(OpenCtx c, const OpenEntry* arr, uint32_t lo,
                                    uint32_t hi) {
    return topOfS<true>(c, arr, lo, hi);
  }
  QMG_COLD static void blkRescan(OpenCtx c, const OpenEntry* open,
                                 OpenEntry* btop, uint32_t nOpen, uint32_t b) {
    const Group g;
    const uint32_t lo = b * kBlk;
    const uint32_t hi = lo + kBlk < nOpen ? lo + kBlk : nOpen;
    const uint32_t p = blkTopOf(c, open, lo, hi);
    if (g.lane == 0) {
      OpenEntry e = glob(open)[p];
      e.hash = p;  // (block tops carry their position, not the layout hash)
      glob(btop)[b] = e;
    }
    g.sync();
  }
  QMG_COLD static void blkEnter(OpenCtx c, const OpenEntry* open, OpenEntry* btop,
                                uint32_t nOpen) {
    const Group g;
    g.sync();
    const uint32_t nb = (nOpen + kBlk - 1) / kBlk;
    QMG_NOUNROLL
    for (uint32_t b = 0; b < nb; ++b) {
      blkRescan(c, open, btop, nOpen, b);
    }
  }
  // the top of the open list: its entry, with the position in .hash
  QMG_COLD static OpenEntry blkTop(OpenCtx c, const OpenEntry* btop,
                                   uint32_t nOpen) {
    const uint32_t nb = (nOpen + kBlk - 1) / kBlk;
    const uint32_t b = blkTopOf(c, btop, 0, nb);
    return glob(btop)[b];
  }
  // the entry e at pos was appended (fresh: it opened a new block) or replaced
  // by a better one
  QMG_COLD static void blkChanged(OpenCtx c, OpenEntry* btop, OpenEntry e,
                                  uint32_t pos, bool fresh) {
    const Group g;
    if (g.lane == 0) {
      const uint32_t b = pos / kBlk;
      bool take = fresh;
      if (!take) {
        const OpenEntry t = glob(btop)[b];
        take = t.idx == e.idx || hasPriorityOf(c, e, t);
      }
      if (take) {
        e.hash = pos;  // (block tops carry their position, not the hash)
        glob(btop)[b] = e;
      }
    }
    g.sync();
  }
  // the entry at pos was removed and the one at `last` (= nOpen now) moved
  // into its place
  QMG_COLD static void blkRemoved(OpenCtx c, const OpenEntry* open,
                                  OpenEntry* btop, uint32_t nOpen, uint32_t pos,
                                  uint32_t movedIdx) {
    const uint32_t last = nOpen;
    const uint32_t bp = pos / kBlk;
    const uint32_t bl = last / kBlk;
    if (bp * kBlk < nOpen) {
      blkRescan(c, open, btop, nOpen, bp);
    }
    if (bl != bp && bl * kBlk < nOpen && btop[bl].idx == movedIdx) {
      blkRescan(c, open, btop, nOpen, bl);
    }
  }

  // the entry on top of the open list and its position
  QMG_HD OpenEntry topEntry(uint32_t& pos) const {
    if (kFull && blockMode) {
      const OpenEntry e = blkTop(octx(), M.btop, nOpen);
      pos = static_cast<uint32_t>(e.hash);
      return e;
    }
    pos = topOfS(octx(), M.open, 0, nOpen);
    return glob(M.open)[pos];
  }

  // pop: take the entry at pos (node idx, goal flag valid) off the open list
  QMG_HD void removeOpen(uint32_t pos, uint32_t idx, bool valid) {
    // group-uniform; lane 0 writes
    const uint32_t last = nOpen - 1;
    uint32_t moved = kNone;
    if (pos != last) {
      const OpenEntry e = glob(M.open)[last];
      moved = e.idx;
      if (g.lane == 0) {
        glob(M.open)[pos] = e;
        hdr(e.idx)->openPos = pos;
      }
    }
    if (g.lane == 0) {
      // flags: bit0 goal, bit1 open (cleared)
      hdr(idx)->flags = static_cast<uint16_t>(valid ? 1U : 0U);
    }
    nOpen = last;
    g.sync();
    if (kFull && blockMode) {
      blkRemoved(octx(), M.open, M.btop, nOpen, pos, moved);
    }
  }

  // write a node record: parent layout (pQ/pL) with the child's swap.  The
  // parent's arrays are copied as 64-bit words (the scratch arrays are 8-byte
  // aligned and zero beyond n), then the four entries the SWAP changes are
  // stored individually.
  QMG_HD void writeNode(uint32_t idx, const Child& ch, uint32_t parentIdx,
                        uint32_t depth, uint32_t id, uint32_t openPos,
                        bool root) {
    int16_t* q = nodeQ(idx);
    int16_t* l = nodeL(idx);
    if (!kFull) {
      // (short searches on small architectures: one entry per lane is fewer
      // instructions than the word copy + patch; records of the lean variant
      // are never compared word by word)
      const int a = ch.a;
      const int b = ch.b;
      const int q1 = root ? -2 : pQ[a];
      const int q2 = root ? -2 : pQ[b];
      QMG_NOUNROLL
      for (int i = g.lane; i < n; i += Group::W) {
        int16_t vq = pQ[i];
        int16_t vl = pL[i];
        if (!root) {
          if (i == a) {
            vq = static_cast<int16_t>(q2);
          } else if (i == b) {
            vq = static_cast<int16_t>(q1);
          }
          if (i == q1) {
            vl = static_cast<int16_t>(b);
          } else if (i == q2) {
            vl = static_cast<int16_t>(a);
          }
        }
        q[i] = vq;
        l[i] = vl;
      }
    } else {
      QMG_NOUNROLL
      for (int i0 = 4 * g.lane; i0 < n; i0 += 4 * Group::W) {
        st64(q + i0, ld64(pQ + i0));
        st64(l + i0, ld64(pL + i0));
      }
    }
    if (kFull && !root) {
      g.sync();  // (orders the word stores before the entry stores)
      const int a = ch.a;
      const int b = ch.b;
      const int q1 = pQ[a];
      const int q2 = pQ[b];
      if (g.lane == 0) {
        q[a] = static_cast<int16_t>(q2);
        q[b] = static_cast<int16_t>(q1);
        if (q1 >= 0) {
          l[q1] = static_cast<int16_t>(b);
        }
        if (q2 >= 0) {
          l[q2] = static_cast<int16_t>(a);
        }
      }
    }
    if (g.lane == 0) {
      NodeHdr* h = hdr(idx);
      h->fixed = ch.fixed;
      h->rev = ch.rev;
      h->heur = ch.heur;
      h->look = ch.look;
      h->hash = ch.hash;
      h->parent = parentIdx;
      h->shared = ch.shared;
      h->depth = depth;
      h->swA = ch.a;
      h->swB = ch.b;
      h->nValid = ch.nValid;
      h->flags = static_cast<uint16_t>((ch.valid ? 1U : 0U) | 2U);
      h->openPos = openPos;
      h->id = id;
      if (kFull) {
        h->epoch = static_cast<uint16_t>(M.epoch);
        h->extra = static_cast<uint16_t>(root ? 0U : M.childExtra);
      }
    }
  }

  QMG_HD static OpenEntry entryOf(const Child& ch, uint32_t idx) {
    OpenEntry e;
    e.f = ch.fixed + ch.rev + ch.heur + ch.look;  // Node::getTotalCost
    e.h2 = ch.heur + ch.look;
    e.hash = ch.hash;
    e.idx = idx;
    e.nValid = ch.nValid;
    e.valid = ch.valid;
    return e;
  }

  // this lane's share of comparing the stored node idx with the layout
  // "parent with the SWAP on (a, b)" (a < 0: the parent itself): true if any
  // of the lane's entries differ.  Word by word: stored XOR parent must be 0
  // except at the entries a and b, where it must be pQ[a] XOR pQ[b].
  QMG_HD bool layoutDiffers(uint32_t idx, int a, int b) const {
    const int16_t* q = nodeQ(idx);
    if (!kFull) {
      bool df = false;
      QMG_NOUNROLL
      for (int i0 = 0; i0 < n; i0 += Group::W) {
        const int i = i0 + g.lane;
        if (i < n) {
          int16_t v = pQ[i];
          if (a >= 0) {
            if (i == a) {
              v = pQ[b];
            } else if (i == b) {
              v = pQ[a];
            }
          }
          df = df || (q[i] != v);
        }
      }
      return df;
    }
    uint32_t d = 0;
    if (a >= 0) {
      d = (static_cast<uint32_t>(static_cast<uint16_t>(pQ[a])) ^
           static_cast<uint32_t>(static_cast<uint16_t>(pQ[b])));
    }
    bool diff = false;
    QMG_NOUNROLL
    for (int i0 = 4 * g.lane; i0 < n; i0 += 4 * Group::W) {
      const uint64_t x = ld64(q + i0) ^ ld64(pQ + i0);
      uint32_t lo = static_cast<uint32_t>(x);
      uint32_t hi = static_cast<uint32_t>(x >> 32);
      const uint32_t da = static_cast<uint32_t>(a - i0);
      const uint32_t db = static_cast<uint32_t>(b - i0);
      if (a >= 0 && da < 4U) {
        const uint32_t v = d << ((da & 1U) << 4);
        lo ^= (da & 2U) ? 0U : v;
        hi ^= (da & 2U) ? v : 0U;
      }
      if (a >= 0 && db < 4U) {
        const uint32_t v = d << ((db & 1U) << 4);
        lo ^= (db & 2U) ? 0U : v;
        hi ^= (db & 2U) ? v : 0U;
      }
      diff = diff || ((lo | hi) != 0U);
    }
    return diff;
  }
  // does the stored node idx have the layout "parent with ch's swap"?
  QMG_HD bool sameLayout(uint32_t idx, const Child& ch, bool root) const {
    // (one vote for the whole layout: equal hashes almost always mean equal
    // layouts, and the loads of all iterations overlap this way)
    return !g.any(layoutDiffers(idx, root ? -1 : ch.a, root ? -1 : ch.b));
  }

  QMG_HD uint32_t homeSlot(uint64_t h) const {
    return static_cast<uint32_t>(h ^ (h >> 32)) & M.tableMask;
  }
  QMG_HD uint32_t curStamp() const { return M.tabStampV & kStampMask; }

  // switch from scanning the open list to the layout-keyed hash table
  QMG_HD void buildTable() {
    g.sync();
    if (g.lane == 0) {
      const uint32_t cur = curStamp();
      QMG_NOUNROLL
      for (uint32_t i = 0; i < nOpen; ++i) {
        const OpenEntry e = glob(M.open)[i];
        uint32_t slot = homeSlot(e.hash);
        QMG_NOUNROLL
        while (tabStamp(glob(M.table)[slot]) == cur) {
          slot = (slot + 1) & M.tableMask;
        }
        glob(M.table)[slot] = tabPack(cur, e.hash, e.idx);
      }
    }
    tableMode = true;
    g.sync();
  }

  // DataLogger::logSearchNode for a generated node (final pass, full variant)
  QMG_HD void logNode(const Child& ch, uint32_t id, uint32_t depth, bool root) {
    if (!kFull || !record || I.nodeLog == nullptr) {
      return;
    }
    if (g.lane == 0 && nLog < I.nodeLogCap) {
      NodeLogRec r;
      r.cost_fixed = ch.fixed + ch.rev;
      r.cost_heur = ch.heur;
      r.lookahead_penalty = ch.look;
      r.id = id;
      r.parent = root ? 0U : parentId;
      r.depth = depth;
      r.layer = layer;
      r.swap_first = root ? static_cast<uint16_t>(0) : ch.a;
      r.swap_second = root ? static_cast<uint16_t>(0) : ch.b;
      r.valid = ch.valid;
      r.root = root ? 1 : 0;
      I.nodeLog[nLog] = r;
    }
    if (root && I.rootQ != nullptr) {
      QMG_NOUNROLL
      for (int i = g.lane; i < n; i += Group::W) {
        I.rootQ[static_cast<size_t>(layer) * n + i] = pQ[i];
      }
    }
    ++nLog;
  }

  // UniquePriorityQueue::push :59-82 for one node.  Group-uniform.  Returns
  // false when the node pool is exhausted.  Only nodes in the open list take
  // part in the layout identity (pop erases the membership, :84-93), so short
  // open lists are scanned directly; longer ones are indexed by the hash table.
  QMG_HD bool push(const Child& ch, uint32_t parentIdx, uint32_t depth,
                   bool root) {
    const uint32_t id = nextId++;
    logNode(ch, id, depth, root);
    return pushOne(ch, parentIdx, depth, root, id);
  }
  QMG_HD bool pushOne(const Child& ch, uint32_t parentIdx, uint32_t depth,
                      bool root, uint32_t id) {
    const uint64_t h = ch.hash;
    uint32_t found = kNone;  // pool index of an OPEN node with this layout
    uint32_t slot = 0;
    // (kAlwaysTable: the scan below is not part of the full variant's code)
    constexpr bool kAlwaysTable = kFull && kTableFromStart;
    if (!kAlwaysTable && !tableMode && nOpen > kScanMax) {
      buildTable();
    }
    if (!kAlwaysTable && !tableMode) {
      QMG_NOUNROLL
      for (uint32_t i0 = 0; i0 < nOpen && found == kNone; i0 += Group::W) {
        const uint32_t i = i0 + g.lane;
        uint32_t m = g.ballot(i < nOpen && glob(M.open)[i].hash == h);
        QMG_NOUNROLL
        while (m != 0 && found == kNone) {
          const int src = ctz32(m);
          m &= m - 1;
          const uint32_t idx = glob(M.open)[i0 + src].idx;
          if (sameLayout(idx, ch, root)) {
            found = idx;
          }
        }
      }
    } else {
      const uint32_t cur = curStamp();
      slot = homeSlot(h);
      for (;;) {
        const uint64_t e = glob(M.table)[slot];
        if (tabStamp(e) != cur) {
          break;  // free slot
        }
        const uint32_t idx = tabIdx(e);
        if (tabHashEq(e, h) && idx < nPool && sameLayout(idx, ch, root)) {
          found = idx;
          break;
        }
        slot = (slot + 1) & M.tableMask;
      }
      if (found != kNone && (hdr(found)->flags & 2U) == 0) {
        found = kNone;  // closed: the slot is re-pointed to the new node below
      }
    }
    // one commit path for both outcomes (a second inlined copy of writeNode
    // and blkChanged costs instruction-cache space, which is what bounds the
    // kernel): either the record of the open duplicate is overwritten in
    // place, or a new record, open entry and table entry are made
    uint32_t idx = found;
    uint32_t pos = 0;
    const bool fresh = found == kNone;
    if (!fresh) {
      // layout is in the open list: keep the better one (:63-79)
      pos = hdr(found)->openPos;
      const OpenEntry ex = glob(M.open)[pos];
      if (!(cmp4(entryOf(ch, found), ex) < 0)) {
        return true;
      }
      // existing > new: replace in place (the old node was never expanded,
      // so nothing refers to its record)
      g.sync();
    } else {
      if (nPool >= M.poolCap) {
        return false;
      }
      idx = nPool++;
      pos = nOpen++;
    }
    writeNode(idx, ch, parentIdx, depth, id, pos, root);
    const OpenEntry ne = entryOf(ch, idx);
    if (g.lane == 0) {
      glob(M.open)[pos] = ne;
      if (fresh && tableMode) {
        glob(M.table)[slot] = tabPack(curStamp(), h, idx);
      }
    }
    g.sync();
    if (kFull && blockMode) {
      blkChanged(octx(), M.btop, ne, pos, fresh && pos % kBlk == 0);
    }
    return true;
  }

  // push() for the cnt children in S.ch() of one parent, long open lists
  // (hash-table mode): the look-ups - table probe, layout comparison, the
  // record of an open duplicate - are done for all children at once, ONE CHILD
  // PER LANE, so that their memory latencies overlap; the children are then
  // committed one after the other in generation order, which keeps node ids,
  // pool and open-list positions exactly as push() child by child gives them.
  // Siblings that could see each other's insertion (equal layout hash, or the
  // same free table slot) take the one-by-one path at their turn.
  QMG_HD bool pushChunk(int cnt, uint32_t parentIdx, uint32_t depth) {
    const uint32_t cur = curStamp();
    enum : uint32_t { kSkip = 0, kReplace = 1, kNew = 2, kSlow = 3 };
    QMG_NOUNROLL
    for (int c0 = 0; c0 < cnt; c0 += Group::W) {
      const int c = c0 + g.lane;
      const bool act = c < cnt;
      uint64_t h = 0;
      int a = 0, b = 0;
      if (act) {
        h = S.ch()[c].hash;
        a = S.ch()[c].a;
        b = S.ch()[c].b;
      }
      uint32_t slot = homeSlot(h);
      uint32_t found = kNone;
      bool done = !act;
      // ---- look-up: probe to the next candidate, confirm candidates by
      // comparing layouts (the whole group per candidate, several in flight)
      QMG_NOUNROLL
      for (;;) {
        uint32_t cand = kNone;
        if (!done) {
          QMG_NOUNROLL
          for (;;) {
            const uint64_t e = glob(M.table)[slot];
            if (tabStamp(e) != cur) {
              done = true;  // free slot: not in the table
              break;
            }
            if (tabHashEq(e, h) && tabIdx(e) < nPool) {
              cand = tabIdx(e);
              break;
            }
            slot = (slot + 1) & M.tableMask;
          }
        }
        uint32_t m = g.ballot(cand != kNone);
        if (m == 0) {
          break;
        }
        QMG_NOUNROLL
        while (m != 0) {
          constexpr int kFlight = QMG_FLIGHT;
          int src[kFlight];
          bool diff[kFlight];
          for (int k = 0; k < kFlight; ++k) {
            src[k] = -1;
            if (m != 0) {
              src[k] = ctz32(m);
              m &= m - 1;
            }
          }
          for (int k = 0; k < kFlight; ++k) {
            diff[k] = false;
            if (src[k] >= 0) {
              diff[k] = layoutDiffers(g.bcast(cand, src[k]), g.bcast(a, src[k]),
                                      g.bcast(b, src[k]));
            }
          }
          for (int k = 0; k < kFlight; ++k) {
            if (src[k] >= 0) {
              const bool same = !g.any(diff[k]);
              if (g.lane == src[k]) {
                if (same) {
                  found = cand;
                  done = true;
                } else {
                  slot = (slot + 1) & M.tableMask;
                }
              }
            }
          }
        }
      }
      // ---- what happens to the child, from the duplicate's record
      uint32_t dec = kSkip;
      uint32_t pos = 0;
      bool freeSlot = false;
      if (act) {
        dec = kNew;
        freeSlot = found == kNone;
        if (found != kNone) {
          const NodeHdr* hf = hdr(found);
          const uint32_t fl = hf->flags;
          // (closed: the slot is re-pointed to the new node)
          if ((fl & 2U) != 0) {
            // the open-list keys of the duplicate, as entryOf() made them
            OpenEntry ex;
            ex.f = hf->fixed + hf->rev + hf->heur + hf->look;
            ex.h2 = hf->heur + hf->look;
            ex.hash = 0;
            ex.idx = found;
            ex.nValid = hf->nValid;
            ex.valid = static_cast<uint16_t>(fl & 1U);
            pos = hf->openPos;
            dec = cmp4(entryOf(S.ch()[c], found), ex) < 0 ? kReplace : kSkip;
          }
        }
      }
      // ---- siblings that may interact: equal hashes, or two new nodes that
      // found the same free slot
      const uint32_t below = g.lane == 0 ? 0U : (0xffffffffU >> (32 - g.lane));
      {
        const uint32_t same = g.matchAny(act, h);
        if ((same & below) != 0) {
          dec = kSlow;
        }
        const bool wantsFree = act && dec == kNew && freeSlot;
        const uint32_t sameSlot = g.matchAny(wantsFree, static_cast<uint64_t>(slot));
        if (wantsFree && (sameSlot & below) != 0) {
          dec = kSlow;
        }
      }
      // ---- commit in generation order
      bool reprobe = false;
      const int nj = (cnt - c0) < Group::W ? (cnt - c0) : Group::W;
      QMG_NOUNROLL
      for (int j = 0; j < nj; ++j) {
        const Child ch = S.ch()[c0 + j];
        const uint32_t id = nextId++;
        logNode(ch, id, depth, false);
        const uint32_t d = g.bcast(dec, j);
        if (d == kSkip) {
          continue;
        }
        if (d == kSlow) {
          if (!pushOne(ch, parentIdx, depth, false, id)) {
            return false;
          }
          reprobe = true;  // it may have taken a free slot a later sibling holds
          continue;
        }
        // (one commit path for kReplace and kNew, see pushOne)
        const bool fresh = d != kReplace;
        uint32_t idx = g.bcast(found, j);
        uint32_t npos = g.bcast(pos, j);
        uint32_t sl = g.bcast(slot, j);
        if (!fresh) {
          g.sync();
        } else {
          if (nPool >= M.poolCap) {
            return false;
          }
          if (reprobe && g.bcast(freeSlot, j)) {
            QMG_NOUNROLL
            while (tabStamp(glob(M.table)[sl]) == cur) {
              sl = (sl + 1) & M.tableMask;
            }
          }
          idx = nPool++;
          npos = nOpen++;
        }
        writeNode(idx, ch, parentIdx, depth, id, npos, false);
        const OpenEntry ne = entryOf(ch, idx);
        if (g.lane == 0) {
          glob(M.open)[npos] = ne;
          if (fresh) {
            glob(M.table)[sl] = tabPack(cur, ch.hash, idx);
          }
        }
        g.sync();
        if (blockMode) {
          blkChanged(octx(), M.btop, ne, npos, fresh && npos % kBlk == 0);
        }
      }
    }
    return true;
  }

  // score the cnt nodes in S.ch() (their a/b are set; root: the layout itself):
  // costs, layout hash and lookahead penalty
  QMG_HD void evaluate(int cnt, bool root, const ParentCosts& pc,
                       uint64_t parentHash) {
    g.sync();
    if (kFid) {
      scoreFidelityChunk(cnt, root, pc);
    } else {
      QMG_NOUNROLL
      for (int c = g.lane; c < cnt; c += Group::W) {
        Child out;
        const int a = root ? 0 : S.ch()[c].a;
        const int b = root ? 0 : S.ch()[c].b;
        scoreGateCount(a, b, pc, out,
                       M.nsw + static_cast<size_t>(g.lane) * M.maxG);
        S.ch()[c].fixed = out.fixed;
        S.ch()[c].rev = out.rev;
        S.ch()[c].heur = out.heur;
        S.ch()[c].shared = out.shared;
        S.ch()[c].nValid = out.nValid;
        S.ch()[c].valid = out.valid;
        if (!root) {
          const int q1 = pQ[a];
          const int q2 = pQ[b];
          S.ch()[c].hash = parentHash - cellHash(a, q1) - cellHash(b, q2) +
                         cellHash(a, q2) + cellHash(b, q1);
        }
      }
    }
    if (kFid && !root) {
      QMG_NOUNROLL
      for (int c = g.lane; c < cnt; c += Group::W) {
        const int a = S.ch()[c].a;
        const int b = S.ch()[c].b;
        const int q1 = pQ[a];
        const int q2 = pQ[b];
        S.ch()[c].hash = parentHash - cellHash(a, q1) - cellHash(b, q2) +
                       cellHash(a, q2) + cellHash(b, q1);
      }
    }
    g.sync();
  }

  // expandNode :698-727, resumable.  The candidate order (considered qubit
  // ascending, then coupling edge ascending) and the usedSwaps rule are
  // reproduced by a closed form: a swap between two occupied physical qubits
  // is generated only from the smallest considered qubit of the pair and only
  // on the smaller of the (up to two) directed edges between them.  Fills
  // S.ch()[0..cnt) with up to kChunk swaps and sets enumDone when the parent is
  // exhausted.
  uint32_t enumCi = 0, enumK = 0;
  bool enumDone = true;

  QMG_HD int gatherChunk() {
    int cnt = 0;
    QMG_NOUNROLL
    while (enumCi < nCons) {
      const int q = ldc(cons + enumCi);
      const int p = pL[q];
      const uint32_t lo = p < 0 ? 0U : ldg(A.incOff + p);
      const uint32_t hi = p < 0 ? 0U : ldg(A.incOff + p + 1);
      QMG_NOUNROLL
      while (lo + enumK < hi) {
        const uint32_t k = lo + enumK + g.lane;
        bool keep = false;
        int ea = 0;
        int eb = 0;
        if (k < hi) {
          const uint32_t e = ldg(A.inc + k);
          ea = ldg(A.edges + 2 * e);
          eb = ldg(A.edges + 2 * e + 1);
          const int qa = pQ[ea];
          const int qb = pQ[eb];
          if (qa == -1 || qb == -1) {
            keep = true;
          } else {
            const int other = (qa == q) ? qb : qa;
            const bool otherFirst = considered(other) && other < q;
            const bool twinFirst =
                (eb < ea) && ((ldg(A.adj + ea * n + eb) & 2U) != 0);
            keep = !otherFirst && !twinFirst;
          }
        }
        const uint32_t bal = g.ballot(keep);
        const int add = popc32(bal);
        if (cnt + add > kChunk) {
          return cnt;  // resume with this block in the next chunk
        }
        if (keep) {
          const int rank = popc32(bal & ((1U << g.lane) - 1U));
          S.ch()[cnt + rank].a = static_cast<uint16_t>(ea);
          S.ch()[cnt + rank].b = static_cast<uint16_t>(eb);
        }
        cnt += add;
        enumK += Group::W;
      }
      ++enumCi;
      enumK = 0;
    }
    enumDone = true;
    return cnt;
  }


  // ------------------------------------------------------------------------
  // mapUnmappedGates :211-285 / mapToMinDistance :287-307 on S.curQ() / S.curL()
  // ------------------------------------------------------------------------
  uint32_t nPlace = 0;  // running count into I.places
  bool record = false;

  uint32_t nPlaceLog = 0;
  QMG_HD void place(int logical, int phys, bool oneQubitRule = false) {
    if (kFull && I.placeLog != nullptr) {
      if (g.lane == 0 && nPlaceLog < I.placeLogCap) {
        I.placeLog[2 * nPlaceLog] = static_cast<uint16_t>(logical);
        I.placeLog[2 * nPlaceLog + 1] =
            static_cast<uint16_t>(phys | (oneQubitRule ? 0x8000 : 0));
      }
      ++nPlaceLog;
    }
    if (g.lane == 0) {
      S.curQ()[phys] = static_cast<int16_t>(logical);
      S.curL()[logical] = static_cast<int16_t>(phys);
      if (record && I.places != nullptr && nPlace < I.placeCap) {
        I.places[2 * nPlace] = static_cast<uint16_t>(logical);
        I.places[2 * nPlace + 1] = static_cast<uint16_t>(phys);
      }
    }
    if (record) {
      ++nPlace;
    }
    g.sync();
  }

  QMG_HD int firstFree() const {
    QMG_NOUNROLL
    for (int p0 = 0; p0 < n; p0 += Group::W) {
      const int p = p0 + g.lane;
      const uint32_t bal = g.ballot(p < n && S.curQ()[p] == -1);
      if (bal != 0) {
        return p0 + ctz32(bal);
      }
    }
    return -1;
  }

  QMG_HD bool mapToMinDistance(int source, int target) {
    const int src = S.curL()[source];
    double best = DBL_MAX;
    uint32_t pos = kNone;
    QMG_NOUNROLL
    for (int i = g.lane; i < n; i += Group::W) {
      if (S.curQ()[i] == -1) {
        const double d = dRev(src, i);
        if (d < best) {
          best = d;
          pos = static_cast<uint32_t>(i);
        }
      }
    }
    gargmin(g, best, pos);
    if (pos == kNone) {
      return false;
    }
    place(target, static_cast<int>(pos));
    return true;
  }

  QMG_HD bool mapUnmappedGates() {
    g.sync();
    if (kFid) {
      QMG_NOUNROLL
      for (uint32_t i = 0; i < G1; ++i) {
        const Gate1 e = ldGate1(gates1 + i);
        if (S.curL()[e.q] == -1) {
          const int p = firstFree();
          if (p >= 0) {
            place(e.q, p, true);
          }
        }
      }
    }
    QMG_NOUNROLL
    for (uint32_t i = 0; i < G; ++i) {
      const Gate2 e = ldGate(gates + i);
      const int l1 = S.curL()[e.q1];
      const int l2 = S.curL()[e.q2];
      if (l1 == -1 && l2 == -1) {
        // first coupling edge (ascending) with both ends free
        int ca = -1;
        int cb = -1;
        QMG_NOUNROLL
        for (int k0 = 0; k0 < A.nE; k0 += Group::W) {
          const int k = k0 + g.lane;
          bool ok = false;
          if (k < A.nE) {
            ok = S.curQ()[ldg(A.edges + 2 * k)] == -1 &&
                 S.curQ()[ldg(A.edges + 2 * k + 1)] == -1;
          }
          const uint32_t bal = g.ballot(ok);
          if (bal != 0) {
            const int kk = k0 + ctz32(bal);
            ca = ldg(A.edges + 2 * kk);
            cb = ldg(A.edges + 2 * kk + 1);
            break;
          }
        }
        if (ca < 0) {
          // two free qubits with minimal distance, first strict minimum in
          // (i, j) order
          double best = 2147483647.0;
          uint32_t code = kNone;
          QMG_NOUNROLL
          for (int i = g.lane; i < n; i += Group::W) {
            if (S.curQ()[i] != -1) {
              continue;
            }
            QMG_NOUNROLL
            for (int j = i + 1; j < n; ++j) {
              if (S.curQ()[j] != -1) {
                continue;
              }
              const double d = dRev(i, j);
              const uint32_t cd = static_cast<uint32_t>(i) * n + j;
              if (d < best || (d == best && cd < code)) {
                best = d;
                code = cd;
              }
            }
          }
          gargmin(g, best, code);
          if (code == kNone) {
            return false;
          }
          ca = static_cast<int>(code / n);
          cb = static_cast<int>(code % n);
        }
        place(e.q1, ca);
        place(e.q2, cb);
      } else if (l1 == -1) {
        if (!mapToMinDistance(e.q2, e.q1)) {
          return false;
        }
      } else if (l2 == -1) {
        if (!mapToMinDistance(e.q1, e.q2)) {
          return false;
        }
      }
    }
    return true;
  }

  // ------------------------------------------------------------------------
  // aStarMap :521-696
  // ------------------------------------------------------------------------
  struct LayerStats {
    uint64_t expanded = 0, expAfterFirst = 0, expAfterOpt = 0, solNodes = 0,
             solAfterOpt = 0;
    uint32_t result = kNone;  // pool index; kNone with rootResult = the root
    bool rootResult = false;
    int32_t early = 0;
  };

  QMG_HD void setLayer(uint32_t l) {
    layer = l;
    const uint32_t lo = ldc(I.circ.g2Off + l);
    G = ldc(I.circ.g2Off + l + 1) - lo;
    gates = I.circ.g2 + lo;
    const uint32_t lo1 = ldc(I.circ.g1Off + l);
    G1 = ldc(I.circ.g1Off + l + 1) - lo1;
    gates1 = I.circ.g1 + lo1;
    if (I.circ.consOff != nullptr) {
      const uint32_t loc = ldc(I.circ.consOff + l);
      nCons = ldc(I.circ.consOff + l + 1) - loc;
      cons = I.circ.cons + loc;
    } else {
      nCons = ldc(I.circ.consCnt + l);
      cons = I.circ.cons + (2U * lo + lo1);
    }
    fidK = nCons == 0 ? kNone : nCons - 1;
    const uint32_t fa = ldc(I.circ.firstAfter + l);
    const uint32_t avail = I.circ.nTwoQ - fa;
    nLook = C.lookahead == 0 ? 0U
                             : (avail < C.nFactors ? avail : C.nFactors);
    lookLayers = I.circ.twoQ + fa;
    maskReady = false;
  }

  // bit mask of getConsideredQubits(layer); built on the first expansion
  QMG_HD void buildMask() {
    g.sync();
    QMG_NOUNROLL
    for (int w = g.lane; w < (n + 31) / 32; w += Group::W) {
      S.consMask()[w] = 0;
    }
    g.sync();
    if (g.lane == 0) {
      QMG_NOUNROLL
      for (uint32_t i = 0; i < nCons; ++i) {
        const int q = ldc(cons + i);
        S.consMask()[q >> 5] |= 1U << (q & 31);
      }
    }
    maskReady = true;
    g.sync();
  }

  // a fresh stamp for the entries this aStarMap call writes into the hash
  // table it uses (24 bits; when they wrap the table is cleared).  The private
  // table counts with the search's own M.stamp, a large pool's table with the
  // pool's counter (its holders change).  Group-uniform.
  QMG_HD void newTableStamp() {
    const bool inBig = kFull && M.bigHeld >= 0;
    uint32_t s = inBig ? M.big[M.bigHeld].stamp + 1U : M.stamp;
    g.sync();
    if ((s & kStampMask) == 0) {
      QMG_NOUNROLL
      for (uint32_t i = g.lane; i <= M.tableMask; i += Group::W) {
        glob(M.table)[i] = 0;
      }
      ++s;
    }
    if (g.lane == 0) {
      if (inBig) {
        M.big[M.bigHeld].stamp = s;
      } else {
        M.stamp = s;
      }
      M.tabStampV = s;
    }
    g.sync();
  }

  // take one of the launch's large pools (waiting for one to come free) and
  // run in it until releaseBig(); false: the launch has none
  QMG_HD bool acquireBig() {
    if (!kFull || M.big == nullptr || M.nBig == 0 || M.bigHeld >= 0) {
      return false;
    }
    g.sync();
    int32_t got = -1;
#if defined(__CUDA_ARCH__)
    for (;;) {
      if (g.lane == 0) {
        const uint32_t i0 = static_cast<uint32_t>(M.stamp * 2654435761U) % M.nBig;
        for (uint32_t k = 0; k < M.nBig; ++k) {
          const uint32_t i = (i0 + k) % M.nBig;
          if (atomicCAS(&M.big[i].lock, 0U, 1U) == 0U) {
            got = static_cast<int32_t>(i);
            break;
          }
        }
        if (got < 0) {
          __nanosleep(2000);
        } else {
          __threadfence();
        }
      }
      if (g.bcast(got, 0) >= 0) {
        break;
      }
    }
#else
    for (uint32_t k = 0; k < M.nBig; ++k) {
      // (single-threaded host build: nobody will release one)
      const uint32_t i = (static_cast<uint32_t>(M.stamp * 2654435761U) + k) % M.nBig;
      if (M.big[i].lock == 0U) {
        M.big[i].lock = 1U;
        got = static_cast<int32_t>(i);
        break;
      }
    }
#endif
    got = g.bcast(got, 0);
    if (got < 0) {
      return false;
    }
    if (g.lane == 0) {
      const BigPool b = M.big[got];
      M.pool0 = M.pool;
      M.open0 = M.open;
      M.table0 = M.table;
      M.btop0 = M.btop;
      M.poolCap0 = M.poolCap;
      M.tableMask0 = M.tableMask;
      M.pool = b.pool;
      M.open = b.open;
      M.table = b.table;
      M.btop = b.btop;
      M.poolCap = b.poolCap;
      M.tableMask = b.tableMask;
      M.bigHeld = got;
      ++M.bigUses;
      if (I.swapCounter != nullptr) {
        // ([2] of the batch-wide counters: layers searched in a large pool)
#if defined(__CUDA_ARCH__)
        atomicAdd(I.swapCounter + 2, 1U);
#else
        ++I.swapCounter[2];
#endif
      }
    }
    g.sync();
    return true;
  }
  // The private pool is about to fill up: the search state moves to one of
  // the launch's large pools - node records and open list keep their indices,
  // the hash table and the block tops are rebuilt - and the search goes on
  // there until the layer is done.  No-op without large pools.
  QMG_HD void migrateToBig() {
    if (!kFull || M.bigHeld >= 0 || !acquireBig()) {
      return;
    }
    const uint64_t* src = reinterpret_cast<const uint64_t*>(M.pool0);
    uint64_t* dst = reinterpret_cast<uint64_t*>(M.pool);
    const size_t words = static_cast<size_t>(nPool) * M.recBytes / 8U;
    QMG_NOUNROLL
    for (size_t w = g.lane; w < words; w += Group::W) {
      dst[w] = src[w];
    }
    QMG_NOUNROLL
    for (uint32_t i = g.lane; i < nOpen; i += Group::W) {
      glob(M.open)[i] = M.open0[i];
    }
    g.sync();
    newTableStamp();
    if (tableMode) {
      buildTable();
    }
    if (blockMode) {
      blkEnter(octx(), M.open, M.btop, nOpen);
    }
    g.sync();
  }
  QMG_HD void releaseBig() {
    g.sync();
    if (g.lane == 0) {
      const int32_t i = M.bigHeld;
      M.pool = M.pool0;
      M.open = M.open0;
      M.table = M.table0;
      M.btop = M.btop0;
      M.poolCap = M.poolCap0;
      M.tableMask = M.tableMask0;
      M.bigHeld = -1;
#if defined(__CUDA_ARCH__)
      __threadfence();
      atomicExch(&M.big[i].lock, 0U);
#else
      M.big[i].lock = 0U;
#endif
    }
    g.sync();
  }

  // resume: the call restarts on a layer that was just split and inherits the
  // open list of the interrupted call (:575-583 - `nodes` is not cleared)
  QMG_HD int astar(LayerStats& st, bool resume) {
    nextId = 0;
    g.sync();
    if (!resume) {
      nPool = 0;
      nOpen = 0;
      tableMode = kFull && kTableFromStart;
      blockMode = false;
      if (g.lane == 0) {
        ++M.stamp;  // (M lives in shared memory on the device)
        M.epoch = 0;
      }
    } else if (kFull && g.lane == 0) {
      M.epoch = M.epoch + 1 < kMaxEpochs ? M.epoch + 1 : kMaxEpochs - 1;
    }
    g.sync();
    if (!resume) {
      newTableStamp();
    }
    if (kFull && g.lane == 0) {
      M.childExtra = 0;
      if (M.priv != nullptr) {
        uint32_t* r =
            privView(M.priv, M.privLayers, M.privG2, M.privG1).ranges + 2U * M.epoch;
        r[0] = static_cast<uint32_t>(gates - I.circ.g2);
        r[1] = r[0] + G;
      }
    }
    g.sync();
    // without a private circuit copy (SearchMem::priv) layers are only split
    // in the final pass, by the host, which then resumes; a pseudo-routing
    // pass searches on
    const bool splittable =
        allowSplit && C.autoSplit != 0 && ldc(I.circ.splittable + layer) != 0;
    // one loop iteration = evaluate one chunk of nodes (first the root from
    // the mapper's current layout, then the children of the node being
    // expanded) and push them; when a parent is exhausted the next node is
    // taken from the open list.  Scoring, push and selection have a single
    // call site each, which keeps the kernel's resident code small.
    bool rootMode = true;
    bool found = false;
    double bestFixedTotal = 0.;
    uint32_t parent = kNone;
    uint32_t depth = 0;
    uint64_t parentHash = 0;
    ParentCosts pc;
    pc.fixed = 0.;
    pc.shared = 0;
    pQ = S.curQ();
    pL = S.curL();
    enumDone = true;
    // the root's lookahead penalty is only reported, never compared, when the
    // root is a goal of a tight heuristic
    const bool rootLookWanted =
        record && (I.layers != nullptr || (kFull && I.nodeLog != nullptr));
    QMG_NOUNROLL
    for (;;) {
      // room for a chunk of new nodes: a private pool that is about to fill
      // up hands the search over to one of the launch's large pools (the
      // state moves, nothing is repeated)
      if (kFull && nPool + static_cast<uint32_t>(kChunk) > M.poolCap &&
          M.bigHeld < 0 && M.big != nullptr) {
        migrateToBig();
      }
      const int cnt = rootMode ? 1 : gatherChunk();
      if (cnt > 0) {
        evaluate(cnt, rootMode, pc, parentHash);
        bool goalRoot = false;
        if (rootMode) {
          goalRoot = S.ch()[0].valid != 0 && kTight && !resume;
        }
        if (!goalRoot || rootLookWanted) {
          scoreLookahead(cnt, rootMode);
        }
        if (rootMode) {
          if (goalRoot) {
            // the root is a goal and the heuristic is tight: the first pop
            // ends the search (:585-599) - no node record is needed
            logNode(S.ch()[0], 0, 0, true);
            nextId = 1;
            st.solNodes = 1;
            st.rootResult = true;
            return ST_OK;
          }
          uint64_t hsh = 0;
          QMG_NOUNROLL
          for (int i = g.lane; i < n; i += Group::W) {
            hsh += cellHash(i, pQ[i]);
          }
          hsh = gsum64(g, hsh);
          if (g.lane == 0) {
            S.ch()[0].hash = hsh;
            S.ch()[0].a = 0;
            S.ch()[0].b = 0;
          }
          g.sync();
        }
        if (kFull && tableMode && !rootMode) {
          if (!pushChunk(cnt, parent, depth)) {
            return ST_POOL;
          }
        } else {
          QMG_NOUNROLL
          for (int c = 0; c < cnt; ++c) {
            const Child ch = S.ch()[c];
            if (!push(ch, parent, depth, rootMode)) {
              return ST_POOL;
            }
          }
        }
      }
      if (!rootMode) {
        if (!enumDone) {
          continue;  // more children of this parent
        }
        ++st.expanded;
        if (found) {
          ++st.expAfterFirst;
          ++st.expAfterOpt;
          if (C.earlyTermination != 0) {
            uint64_t v = 0;
            switch (C.earlyTermination) {
            case 1:
              v = st.expanded;
              break;
            case 2:
              v = st.expAfterFirst;
              break;
            case 3:
              v = st.expAfterOpt;
              break;
            case 4:
              v = st.solNodes;
              break;
            case 5:
              v = st.solAfterOpt;
              break;
            default:
              break;
            }
            if (v >= C.earlyLimit) {
              st.early = 1;
              break;
            }
          }
        }
      }
      rootMode = false;
      // ---- next node (:560-602)
      if (nOpen == 0) {
        break;
      }
      if (!kFull) {
        if (nOpen > kBlkMin) {
          return ST_NEEDS_FULL;
        }
      } else if (!blockMode && nOpen > kBlkMin) {
        blkEnter(octx(), M.open, M.btop, nOpen);
        blockMode = true;
      }
      uint32_t pos = 0;
      const OpenEntry top = topEntry(pos);
      if (found && !(top.f < bestFixedTotal)) {
        break;
      }
      if (splittable && st.expanded >= C.splitLimit) {
        return ST_SPLIT;
      }
      const uint32_t cur = top.idx;
      const NodeHdr* h = hdr(cur);
      if (top.valid) {
        ++st.solNodes;
        const double fixedTotal = h->fixed + h->rev + h->look;
        if (!found || fixedTotal < bestFixedTotal) {
          st.result = cur;
          bestFixedTotal = fixedTotal;
          st.expAfterOpt = 0;
          st.solAfterOpt = 0;
        } else {
          ++st.solAfterOpt;
        }
        found = true;
        if (kTight) {
          break;
        }
      }
      removeOpen(pos, cur, top.valid != 0);
      // ---- stage the node to expand
      if (!maskReady) {
        buildMask();
      }
      pQ = S.Q();
      pL = S.L();
      const int16_t* nq = nodeQ(cur);
      const int16_t* nl = nodeL(cur);
      QMG_NOUNROLL
      for (int i = g.lane; i < n; i += Group::W) {
        pQ[i] = nq[i];
        pL[i] = nl[i];
      }
      pc.fixed = h->fixed;
      pc.shared = h->shared;
      if (kFull) {
        uint32_t ex = h->extra;
        if (h->epoch != M.epoch) {
          ex += staleExtra(cur, h->epoch);
        }
        g.sync();
        if (g.lane == 0) {
          M.childExtra = ex;
        }
      }
      parentHash = h->hash;
      depth = h->depth + 1;
      parent = cur;
      parentId = h->id;
      enumCi = 0;
      enumK = 0;
      enumDone = false;
      g.sync();
    }
    if (!found) {
      return ST_NO_VIABLE;
    }
    return ST_OK;
  }


  // one aStarMap call incl. mapUnmappedGates; adopts the result layout
  // node idx was made ep splits ago: the gates its layer had then and the
  // current one has not, adjacent in the node's layout (these stay in the
  // validMappedTwoQubitGates set its children copy)
  QMG_HD uint32_t staleExtra(uint32_t idx, uint32_t ep) const {
    const uint32_t* r =
        privView(M.priv, M.privLayers, M.privG2, M.privG1).ranges + 2U * ep;
    const uint32_t lo = r[0], hi = r[1];
    const uint32_t cl = static_cast<uint32_t>(gates - I.circ.g2), ch = cl + G;
    const int16_t* nl = nodeL(idx);
    uint64_t cnt = 0;
    QMG_NOUNROLL
    for (uint32_t i = lo + g.lane; i < hi; i += Group::W) {
      if (i < cl || i >= ch) {
        const Gate2 e = ldGate(I.circ.g2 + i);
        cnt += adjU(nl[e.q1], nl[e.q2]) ? 1U : 0U;
      }
    }
    return static_cast<uint32_t>(gsum64(g, cnt));
  }

  // Mapper::splitLayer for the layer being searched
  QMG_HD void splitOnDevice(uint32_t l, bool finalPass) {
    const PrivView p = privView(M.priv, M.privLayers, M.privG2, M.privG1);
    if (*p.active == 0) {
      privActivate(&I.circ, M.priv, M.privLayers, M.privG2, M.privG1, g);
    }
    privSplit(&I.circ, M.priv, M.privLayers, M.privG2, M.privG1, l, kFid, g);
    if (finalPass && g.lane == 0) {
      const uint32_t k = I.summary[5];
      if (I.splitLog != nullptr && k < I.splitCap) {
        I.splitLog[k] = l;
      }
      I.summary[5] = k + 1;
    }
    g.sync();
  }
  QMG_HD void restoreSharedCircuit() {
    if (kFull && M.priv != nullptr) {
      const PrivView p = privView(M.priv, M.privLayers, M.privG2, M.privG1);
      if (*p.active != 0) {
        privRestore(&I.circ, M.priv, M.privLayers, M.privG2, M.privG1, g);
      }
    }
  }

  QMG_HD int routeLayer(uint32_t l, bool finalPass, bool reverse,
                        uint32_t& swapTotal) {
    const int rc = routeLayerIn(l, finalPass, reverse, swapTotal);
    // (a large pool is held for one layer: its result has been adopted)
    if (kFull && M.bigHeld >= 0) {
      releaseBig();
    }
    return rc;
  }
  QMG_HD int routeLayerIn(uint32_t l, bool finalPass, bool reverse,
                          uint32_t& swapTotal) {
    const bool wantRec = finalPass && I.layers != nullptr;
    if (!kFid && !wantRec && !(kFull && I.nodeLog != nullptr)) {
      // a layer without 2q gates needs no search when no record of it is
      // wanted: nothing to place, the root (1 generated node) is the result
      if (ldc(I.circ.g2Off + l + 1) == ldc(I.circ.g2Off + l)) {
        totGenerated += 1;
        if (finalPass) {
          if (I.layerNSwaps != nullptr && g.lane == 0) {
            I.layerNSwaps[l] = 0;
          }
          if (I.layerQ != nullptr) {
            QMG_NOUNROLL
            for (int i = g.lane; i < n; i += Group::W) {
              I.layerQ[static_cast<size_t>(l) * n + i] = S.curQ()[i];
            }
          }
        }
        return ST_OK;
      }
    }
    setLayer(l);
    const uint32_t placeBefore = nPlace;
    record = finalPass;
    if (!mapUnmappedGates()) {
      // no free physical qubit (pair) within reach: the reference's search
      // then runs dry ("No viable mapping found.", :637-639)
      return ST_NO_VIABLE;
    }
    LayerStats st;
    int rc;
    bool resume = false;
    QMG_NOUNROLL
    for (;;) {
      uint32_t room = 0;
      if (M.priv != nullptr) {
        room = M.privLayers;
        if (finalPass && I.layerCap < room) {
          room = I.layerCap;
        }
      }
      const bool devSplit = I.circ.nLayers < room;
      allowSplit = finalPass || devSplit;
      st = LayerStats{};
      rc = astar(st, resume);
      if (rc == ST_POOL) {
        // (the full variant moves to a large pool before its private pool
        // fills up, see astar; the lean one hands the instance over)
        if (!kFull) {
          return ST_NEEDS_FULL;
        }
      }
      if (rc != ST_SPLIT || !devSplit) {
        break;
      }
      if (!kFull) {
        return ST_NEEDS_FULL;
      }
      // the interrupted call leaves no trace in the counters (:575-583)
      splitOnDevice(l, finalPass);
      if (reverse) {
        ++l;
      }
      setLayer(l);
      if (!mapUnmappedGates()) {
        return ST_NO_VIABLE;
      }
      resume = true;
    }
    totExpanded += st.expanded;
    totGenerated += nextId;
    if (rc != ST_OK) {
      return rc;
    }
    g.sync();
    // the 15 eight-byte fields of LayerOut are staged in shared memory and
    // written by 15 lanes as one coalesced 120-byte store
    uint64_t* stage = reinterpret_cast<uint64_t*>(S.lookPen());
    uint32_t d = 0;
    if (st.rootResult) {
      if (wantRec && g.lane == 0) {
        const Child& r = S.ch()[0];
        LayerOut o{};
        o.cost_fixed = r.fixed;
        o.cost_fixed_reversals = r.rev;
        o.cost_heur = r.heur;
        o.lookahead_penalty = r.look;
        o.layoutHash = 0;
        o.depth = 0;
        o.resultId = 0;
        *reinterpret_cast<LayerOut*>(stage) = o;
      }
      if (finalPass && I.layerQ != nullptr) {
        QMG_NOUNROLL
        for (int i = g.lane; i < n; i += Group::W) {
          I.layerQ[static_cast<size_t>(l) * n + i] = S.curQ()[i];
        }
      }
    } else {
      const uint32_t res = st.result;
      const NodeHdr* h = hdr(res);
      const int16_t* rq = nodeQ(res);
      const int16_t* rl = nodeL(res);
      d = h->depth;
      QMG_NOUNROLL
      for (int i = g.lane; i < n; i += Group::W) {
        const int16_t vq = rq[i];
        S.curQ()[i] = vq;
        S.curL()[i] = rl[i];
        if (finalPass && I.layerQ != nullptr) {
          I.layerQ[static_cast<size_t>(l) * n + i] = vq;
        }
      }
      if (wantRec && g.lane == 0) {
        LayerOut o{};
        o.cost_fixed = h->fixed;
        o.cost_fixed_reversals = h->rev;
        o.cost_heur = h->heur;
        o.lookahead_penalty = h->look;
        o.layoutHash = h->hash;
        o.depth = d;
        o.resultId = h->id;
        *reinterpret_cast<LayerOut*>(stage) = o;
      }
      if (finalPass && g.lane == 0) {
        // swaps root -> result
        if (I.swaps != nullptr && swapTotal + d <= I.swapCap) {
          uint32_t cur = res;
          QMG_NOUNROLL
          for (uint32_t k = d; k > 0; --k) {
            const NodeHdr* c = hdr(cur);
            I.swaps[2 * (swapTotal + k - 1)] = c->swA;
            I.swaps[2 * (swapTotal + k - 1) + 1] = c->swB;
            cur = c->parent;
          }
        }
      }
    }
    if (finalPass && I.layerNSwaps != nullptr && g.lane == 0) {
      I.layerNSwaps[l] = static_cast<uint16_t>(d > 0xffffU ? 0xffffU : d);
    }
    if (finalPass && I.layers == nullptr) {
      swapTotal += d;
      if (I.swaps != nullptr && swapTotal > I.swapCap) {
        return ST_BUFFER;
      }
    } else if (finalPass) {
      if (g.lane == 0) {
        LayerOut& o = *reinterpret_cast<LayerOut*>(stage);
        o.expanded = st.expanded;
        o.generated = nextId;
        o.expAfterFirst = st.expAfterFirst;
        o.expAfterOpt = st.expAfterOpt;
        o.solNodes = st.solNodes;
        o.solAfterOpt = st.solAfterOpt;
        o.swapOff = swapTotal;
        o.nSwaps = d;
        o.placeOff = placeBefore;
        o.nPlace = nPlace - placeBefore;
        o.early = st.early;
        o.status = ST_OK;
      }
      g.sync();
      uint64_t* dst = reinterpret_cast<uint64_t*>(I.layers + l);
      QMG_NOUNROLL
      for (int k = g.lane; k < static_cast<int>(sizeof(LayerOut) / 8);
           k += Group::W) {
        dst[k] = stage[k];
      }
      swapTotal += d;
      if (I.swaps != nullptr && swapTotal > I.swapCap) {
        return ST_BUFFER;
      }
    }
    g.sync();
    return ST_OK;
  }

  // HeuristicMapper::map :74-91 for one instance.  Returns false (kFull =
  // false only) when the instance has to be routed again by the full variant;
  // nothing that a second run does not overwrite has been written then.
  QMG_HD bool run() {
    QMG_NOUNROLL
    for (int i = g.lane; i < n; i += Group::W) {
      S.curQ()[i] = -1;
      S.curL()[i] = -1;
    }
    if (g.lane == 0) {
      I.summary[5] = 0;
      if (M.priv != nullptr) {
        *privView(M.priv, M.privLayers, M.privG2, M.privG1).active = 0;
      }
    }
    g.sync();
    if (I.initLoc != nullptr) {
      QMG_NOUNROLL
      for (int i = g.lane; i < n; i += Group::W) {
        const int16_t p = I.initLoc[i];
        S.curL()[i] = p;
        if (p >= 0) {
          S.curQ()[p] = static_cast<int16_t>(i);
        }
      }
    }
    g.sync();
    uint32_t swapTotal = 0;
    int rc = ST_OK;
    uint32_t doneLayers = 0;
    QMG_NOUNROLL
    for (uint32_t pass = 0; pass < I.passes && rc == ST_OK; ++pass) {
      // (the layer count grows when a layer is split; every pseudo-routing
      // pass ends by restoring the layers, :309-344)
      QMG_NOUNROLL
      for (uint32_t i = 0; i < I.circ.nLayers && rc == ST_OK; ++i) {
        rc = routeLayer(i, false, false, swapTotal);
      }
      restoreSharedCircuit();
      QMG_NOUNROLL
      for (uint32_t i = 0; i < I.circ.nLayers && rc == ST_OK; ++i) {
        rc = routeLayer(I.circ.nLayers - 1 - i, false, true, swapTotal);
      }
      restoreSharedCircuit();
    }
    if (rc == ST_OK) {
      if (I.startL != nullptr) {
        QMG_NOUNROLL
        for (int i = g.lane; i < n; i += Group::W) {
          I.startL[i] = S.curL()[i];
        }
      }
      const uint32_t first = I.firstLayer;
      doneLayers = first;
      pseudoExpanded = totExpanded;
      pseudoGenerated = totGenerated;
      QMG_NOUNROLL
      for (uint32_t l = first;; ++l) {
        const uint32_t nL = I.circ.nLayers;
        const uint32_t last =
            (I.nRoute == 0 || first + I.nRoute > nL) ? nL : first + I.nRoute;
        if (l >= last) {
          break;
        }
        rc = routeLayer(l, true, false, swapTotal);
        if (rc != ST_OK) {
          break;
        }
        ++doneLayers;
      }
      restoreSharedCircuit();
    }
    g.sync();
    if (!kFull && rc == ST_NEEDS_FULL) {
      return false;
    }
    QMG_NOUNROLL
    for (int i = g.lane; i < n; i += Group::W) {
      I.finalQ[i] = S.curQ()[i];
      I.finalL[i] = S.curL()[i];
    }
    uint32_t packedOff = 0;
    if (I.swapsPacked != nullptr && I.swaps != nullptr) {
      const uint32_t ns = swapTotal <= I.swapCap ? swapTotal : 0U;
      if (g.lane == 0 && ns > 0) {
#if defined(__CUDA_ARCH__)
        packedOff = atomicAdd(I.swapCounter, ns);
#else
        packedOff = *I.swapCounter;
        *I.swapCounter += ns;
#endif
      }
      packedOff = g.bcast(packedOff, 0);
      const uint32_t* src = reinterpret_cast<const uint32_t*>(I.swaps);
      uint32_t* dst = reinterpret_cast<uint32_t*>(I.swapsPacked) + packedOff;
      QMG_NOUNROLL
      for (uint32_t k = g.lane; k < ns; k += Group::W) {
        dst[k] = src[k];
      }
    }
    if (g.lane == 0) {
      I.summary[4] = packedOff;
      I.summary[0] = swapTotal;
      I.summary[1] = nPlace;
      I.summary[2] = doneLayers;
      I.summary[3] = static_cast<uint32_t>(rc);
      I.summary[6] = nLog;
      I.summary[7] = nPlaceLog;
      I.counters[0] = totExpanded;
      I.counters[1] = totGenerated;
      I.counters[2] = totExpanded - pseudoExpanded;
      I.counters[3] = totGenerated - pseudoGenerated;
    }
    g.sync();
    return true;
  }
};

}  // namespace qmg
