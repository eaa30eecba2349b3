// flatten.hpp - host-side preparation of the routing batches that needs no
// CUDA: a tiny parallel-for over host threads, the byte layout helper of the
// staging arenas, and the flattening of the caller's CSR layers into the
// arrays the kernel reads (considered qubits, list of 2q layers, splittable
// flags).  Reference semantics: Mapper::createLayers' consumers in
// src/sc/heuristic/HeuristicMapper.cpp (considered qubits :520-533) and
// Mapper::isLayerSplittable (src/sc/Mapper.cpp:283-300).
#pragma once
#include <algorithm>
#include <atomic>
#include <cstdint>
#include <cstdlib>
#include <cstring>
#include <thread>
#include <vector>

#include "route_core.hpp"

namespace qmg {

// host threads used for staging: QMAP_B200_THREADS, else the hardware
// concurrency (one process per GPU shares the host: launchers should set the
// variable to cores / ranks)
inline unsigned hostThreads() {
  static const unsigned cached = [] {
    unsigned t = 0;
    if (const char* e = std::getenv("QMAP_B200_THREADS")) {
      t = static_cast<unsigned>(std::strtoul(e, nullptr, 10));
    }
    if (t == 0) {
      t = std::thread::hardware_concurrency();
    }
    if (t == 0) {
      t = 4;
    }
    return std::min(t, 64U);
  }();
  return cached;
}

// f(i) for i in [0, n) on up to hostThreads() threads, dynamic chunks
template <class F> void parallelFor(size_t n, size_t grain, F&& f) {
  if (n == 0) {
    return;
  }
  const unsigned nt = static_cast<unsigned>(
      std::min<size_t>(hostThreads(), (n + grain - 1) / grain));
  if (nt <= 1) {
    for (size_t i = 0; i < n; ++i) {
      f(i);
    }
    return;
  }
  std::atomic<size_t> next{0};
  auto work = [&]() {
    for (;;) {
      const size_t b = next.fetch_add(grain);
      if (b >= n) {
        return;
      }
      const size_t e = std::min(n, b + grain);
      for (size_t i = b; i < e; ++i) {
        f(i);
      }
    }
  };
  std::vector<std::thread> th;
  th.reserve(nt - 1);
  for (unsigned t = 1; t < nt; ++t) {
    th.emplace_back(work);
  }
  work();
  for (auto& t : th) {
    t.join();
  }
}

// byte layout of a sequence of typed sections, each 16-byte aligned
struct Layout {
  size_t size = 0;
  size_t add(size_t bytes) {
    const size_t off = size;
    size = (size + bytes + 15) & ~static_cast<size_t>(15);
    return off;
  }
};

// where one circuit's arrays live.  Uploaded part (byte offsets into the
// staging arena; each circuit's sections are contiguous, so a range of
// circuits is one range of bytes and can be copied to the device as soon as
// it is filled): the CSR layers as the caller gave them.  Derived part (byte
// offsets into a device-only arena): the tables derive.cuh computes.
struct CircSlot {
  uint32_t nLayers = 0, nLogical = 0;
  size_t nG2 = 0, nG1 = 0;
  size_t bG2Off = 0, bG1Off = 0, bG2 = 0, bG1 = 0, bEnd = 0;
  size_t dCons = 0, dConsCnt = 0, dTwoQ = 0, dFirst = 0, dSplit = 0;
  uint32_t nTwoQ = 0, maxG = 1;
  void placeUpload(Layout& lay) {
    const size_t nOff = static_cast<size_t>(nLayers) + 1;
    bG2Off = lay.add(nOff * 4);
    bG1Off = lay.add(nOff * 4);
    bG2 = lay.add(nG2 * sizeof(Gate2));
    bG1 = lay.add(nG1 * sizeof(Gate1));
    bEnd = lay.size;
  }
  void placeDerived(Layout& lay) {
    dCons = lay.add((2 * nG2 + nG1) * 2);
    dConsCnt = lay.add(static_cast<size_t>(nLayers) * 2);
    dTwoQ = lay.add(static_cast<size_t>(nLayers) * 4);
    dFirst = lay.add(static_cast<size_t>(nLayers) * 4);
    dSplit = lay.add(nLayers);
  }
};

// copy one circuit's CSR layers into the staging arena, validating them on the
// way (monotone offsets, operands in range); fills nTwoQ and maxG.
inline bool stageCircuit(int n, const uint32_t* off2, const Gate2* g2,
                         const uint32_t* off1, const Gate1* g1, CircSlot& s,
                         uint32_t* oG2Off, Gate2* oG2, uint32_t* oG1Off,
                         Gate1* oG1) {
  const uint32_t nL = s.nLayers;
  std::memcpy(oG2Off, off2, (static_cast<size_t>(nL) + 1) * 4);
  if (s.nG2 > 0) {
    std::memcpy(oG2, g2, s.nG2 * sizeof(Gate2));
  }
  if (off1 != nullptr) {
    std::memcpy(oG1Off, off1, (static_cast<size_t>(nL) + 1) * 4);
    if (s.nG1 > 0) {
      std::memcpy(oG1, g1, s.nG1 * sizeof(Gate1));
    }
  } else {
    std::memset(oG1Off, 0, (static_cast<size_t>(nL) + 1) * 4);
  }
  uint32_t ntq = 0, maxG = 1;
  bool ok = nL == 0 || (oG2Off[0] == 0 && oG1Off[0] == 0);
  for (uint32_t l = 0; l < nL; ++l) {
    ok = ok && oG2Off[l + 1] >= oG2Off[l] && oG1Off[l + 1] >= oG1Off[l];
    const uint32_t G = oG2Off[l + 1] - oG2Off[l];
    maxG = std::max(maxG, G);
    ntq += G > 0 ? 1U : 0U;
  }
  ok = ok && (nL == 0 || (oG2Off[nL] == s.nG2 && oG1Off[nL] == s.nG1));
  uint32_t worst = 0;
  bool same = false;
  for (size_t i = 0; i < s.nG2; ++i) {
    worst = std::max<uint32_t>(worst, std::max(oG2[i].q1, oG2[i].q2));
    same = same || oG2[i].q1 == oG2[i].q2;
  }
  for (size_t i = 0; i < s.nG1; ++i) {
    worst = std::max<uint32_t>(worst, oG1[i].q);
  }
  s.nTwoQ = ntq;
  s.maxG = maxG;
  return ok && !same && (s.nG2 + s.nG1 == 0 || worst < static_cast<uint32_t>(n));
}

// derive the kernel's per-layer tables for one circuit.  Returns false when a
// gate operand is out of range.
inline bool flattenCircuit(int n, bool fidelityAware, const uint32_t* off2,
                           const Gate2* g2, const uint32_t* off1,
                           const Gate1* g1, CircSlot& s, uint32_t* oG2Off,
                           Gate2* oG2, uint32_t* oG1Off, Gate1* oG1,
                           uint32_t* oConsOff, uint16_t* oCons, uint32_t* oTwoQ,
                           uint32_t* oFirstAfter, uint8_t* oSplit) {
  const uint32_t nL = s.nLayers;
  std::memcpy(oG2Off, off2, (static_cast<size_t>(nL) + 1) * 4);
  if (s.nG2 > 0) {
    std::memcpy(oG2, g2, s.nG2 * sizeof(Gate2));
  }
  if (off1 != nullptr) {
    std::memcpy(oG1Off, off1, (static_cast<size_t>(nL) + 1) * 4);
    if (s.nG1 > 0) {
      std::memcpy(oG1, g1, s.nG1 * sizeof(Gate1));
    }
  } else {
    std::memset(oG1Off, 0, (static_cast<size_t>(nL) + 1) * 4);
  }
  uint32_t nc = 0, ntq = 0, maxG = 1;
  std::vector<uint16_t> tmp;
  oConsOff[0] = 0;
  for (uint32_t l = 0; l < nL; ++l) {
    const uint32_t a = off2[l], b = off2[l + 1];
    const uint32_t a1 = off1 != nullptr ? off1[l] : 0;
    const uint32_t b1 = off1 != nullptr ? off1[l + 1] : 0;
    const uint32_t G = b - a;
    maxG = std::max(maxG, G);
    if (G > 0) {
      oTwoQ[ntq++] = l;
    }
    oFirstAfter[l] = ntq;
    for (uint32_t i = a; i < b; ++i) {
      if (g2[i].q1 >= n || g2[i].q2 >= n || g2[i].q1 == g2[i].q2) {
        return false;
      }
    }
    for (uint32_t i = a1; i < b1; ++i) {
      if (g1[i].q >= n) {
        return false;
      }
    }
    const uint32_t n1 = b1 - a1;
    bool split;
    if (G == 1 && (n1 == 0 || !fidelityAware) && n1 <= 2) {
      // common case (individual gates): two considered qubits
      const uint16_t x = std::min(g2[a].q1, g2[a].q2);
      const uint16_t y = std::max(g2[a].q1, g2[a].q2);
      oCons[nc++] = x;
      oCons[nc++] = y;
      split = false;
      for (uint32_t i = a1; i < b1; ++i) {
        split = split || (g1[i].q != x && g1[i].q != y);
      }
    } else {
      tmp.clear();
      for (uint32_t i = a; i < b; ++i) {
        tmp.push_back(g2[i].q1);
        tmp.push_back(g2[i].q2);
      }
      std::sort(tmp.begin(), tmp.end());
      tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
      // Mapper::isLayerSplittable :283-300
      split = G > 1 || n1 > 2;
      if (!split && G == 1) {
        for (uint32_t i = a1; i < b1; ++i) {
          split = split || !std::binary_search(tmp.begin(), tmp.end(), g1[i].q);
        }
      }
      if (fidelityAware) {
        for (uint32_t i = a1; i < b1; ++i) {
          tmp.push_back(g1[i].q);
        }
        std::sort(tmp.begin(), tmp.end());
        tmp.erase(std::unique(tmp.begin(), tmp.end()), tmp.end());
      }
      for (const auto q : tmp) {
        oCons[nc++] = q;
      }
    }
    oSplit[l] = split ? 1 : 0;
    oConsOff[l + 1] = nc;
  }
  s.nTwoQ = ntq;
  s.maxG = maxG;
  return true;
}

}  // namespace qmg
