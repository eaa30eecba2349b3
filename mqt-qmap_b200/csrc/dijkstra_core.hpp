// dijkstra_core.hpp - Dijkstra::dijkstra (src/sc/utils.cpp:56-89) as the
// reference really behaves, for one source.
//
// The reference keeps POINTERS to per-vertex records in a std::priority_queue
// and overwrites a record in place when it finds a cheaper path (:80-82), so
// keys change while they sit in the heap, the heap property breaks, and a
// vertex can be popped (and frozen, :62) before its label is final.  For
// non-uniform weights the table it returns is therefore NOT the table of true
// shortest paths; routing decisions downstream depend on the table as built,
// so it is reproduced literally: the same binary heap (libstdc++'s
// std::push_heap / std::pop_heap sift sequences on an array of vertex ids),
// the same relaxation order (coupling entries ascending, :67), the same
// strict-improvement test.
//
// Compiled by nvcc (tables.cuh: one thread per source) and by g++ (tests/emu,
// the host debug build that is compared with the compiled reference).
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define QMG_DJ_HD __host__ __device__ inline
#else
#define QMG_DJ_HD inline
#endif

namespace qmg {

// NodeComparator (include/sc/utils.hpp:130-132): "x after y"
QMG_DJ_HD bool djLater(const double* cost, uint16_t x, uint16_t y) {
  return cost[x] > cost[y];
}

// std::__push_heap: value climbs from hole towards top
QMG_DJ_HD void djSiftUp(uint16_t* heap, const double* cost, int hole, int top,
                        uint16_t value) {
  int parent = (hole - 1) / 2;
  while (hole > top && djLater(cost, heap[parent], value)) {
    heap[hole] = heap[parent];
    hole = parent;
    parent = (hole - 1) / 2;
  }
  heap[hole] = value;
}

// std::__adjust_heap on heap[0..len) with the hole at 0
QMG_DJ_HD void djAdjust(uint16_t* heap, const double* cost, int len,
                        uint16_t value) {
  int hole = 0;
  int child = 0;
  while (child < (len - 1) / 2) {
    child = 2 * (child + 1);
    if (djLater(cost, heap[child], heap[child - 1])) {
      --child;
    }
    heap[hole] = heap[child];
    hole = child;
  }
  if ((len & 1) == 0 && child == (len - 2) / 2) {
    child = 2 * (child + 1);
    heap[hole] = heap[child - 1];
    hole = child - 1;
  }
  djSiftUp(heap, cost, hole, 0, value);
}

// One single-source run.  cost[n] receives the labels (-1 = unreached, the
// source keeps 0), visited[n] and heap[heapCap] are scratch; heapCap >=
// 2 * nE + 2 can never overflow (a vertex is pushed only on a strict
// improvement found from a neighbour's first pop).  incOff / inc: ids of the
// directed coupling entries touching a vertex, ascending (= the order in
// which :67 meets them); w[from * n + to] = edgeWeights.  Returns false if
// the heap capacity was too small.
QMG_DJ_HD bool dijkstraReference(int n, int src, const double* w,
                                 const uint16_t* edges, const uint32_t* incOff,
                                 const uint32_t* inc, double* cost,
                                 uint8_t* visited, uint16_t* heap,
                                 int heapCap) {
  for (int v = 0; v < n; ++v) {
    cost[v] = -1.;
    visited[v] = 0;
  }
  cost[src] = 0.;
  int len = 0;
  heap[len++] = static_cast<uint16_t>(src);
  while (len > 0) {
    const uint16_t cur = heap[0];
    visited[cur] = 1;
    // queue.pop(): std::pop_heap + pop_back
    if (len > 1) {
      const uint16_t value = heap[len - 1];
      heap[len - 1] = heap[0];
      djAdjust(heap, cost, len - 1, value);
    }
    --len;
    const double base = cost[cur];
    for (uint32_t k = incOff[cur]; k < incOff[cur + 1]; ++k) {
      const uint32_t e = inc[k];
      const uint16_t a = edges[2 * e];
      const uint16_t b = edges[2 * e + 1];
      const uint16_t to = (a == cur) ? b : a;
      if (visited[to] != 0) {
        continue;
      }
      const double c = base + w[static_cast<size_t>(cur) * n + to];
      if (cost[to] < 0. || c < cost[to]) {
        cost[to] = c;
        if (len >= heapCap) {
          return false;
        }
        // queue.push(): push_back + std::push_heap
        heap[len] = to;
        djSiftUp(heap, cost, len, 0, to);
        ++len;
      }
    }
  }
  return true;
}

}  // namespace qmg
