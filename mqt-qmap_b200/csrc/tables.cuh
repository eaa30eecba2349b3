// tables.cuh - sm_100a kernels for the Architecture cost tables.
//
//   K1  fw_phase{1,2,3}     blocked min-plus Floyd-Warshall (Dijkstra::buildTable,
//                           src/sc/utils.cpp:26-89).  32x32 f64 tiles staged in
//                           shared memory; min-plus has no MMA form, so no
//                           tensor cores.
//                           Exact shortest paths; equal to the reference's
//                           table for UNIFORM weights only (see K1').
//   K1' dijkstra_reference_kernel
//                           the reference's Dijkstra::dijkstra as it really
//                           behaves (dijkstra_core.hpp): its heap of pointers
//                           to records that are rewritten in place can freeze
//                           a vertex before its label is final, so for
//                           non-uniform weights (mixed 30/34 SWAP costs,
//                           fidelity costs) its table differs from the true
//                           shortest paths.  One thread per source replays the
//                           same heap operations; bit-exact for any weights.
//   K2  edge_skip1_kernel   Dijkstra::buildSingleEdgeSkipTable (utils.cpp:147-183)
//   K2' edge_skipk_kernel   one level of Dijkstra::buildEdgeSkipTable (:91-145)
#pragma once
#include <cuda_runtime.h>
#include <float.h>
#include <stdint.h>

#include "dijkstra_core.hpp"

namespace qmg {

constexpr int kFwTile = 32;

__device__ __forceinline__ double dmin2(double a, double b) {
  return b < a ? b : a;  // std::min(a, b)
}

// dist is an np x np matrix (np multiple of 32), +inf where no path is known.
__global__ void __launch_bounds__(kFwTile* kFwTile / 4)
    fw_phase1(double* __restrict__ d, int np, int kb) {
  __shared__ double t[kFwTile][kFwTile + 1];
  const int tx = threadIdx.x & 31;
  const int ty0 = threadIdx.x >> 5;  // 0..7
  const int base = kb * kFwTile;
  for (int r = ty0; r < kFwTile; r += 8) {
    t[r][tx] = d[static_cast<size_t>(base + r) * np + base + tx];
  }
  __syncthreads();
  for (int k = 0; k < kFwTile; ++k) {
    for (int r = ty0; r < kFwTile; r += 8) {
      const double v = t[r][k] + t[k][tx];
      if (v < t[r][tx]) {
        t[r][tx] = v;
      }
    }
    __syncthreads();
  }
  for (int r = ty0; r < kFwTile; r += 8) {
    d[static_cast<size_t>(base + r) * np + base + tx] = t[r][tx];
  }
}

// blockIdx.x: tile index j != kb; blockIdx.y: 0 = row tile (kb, j), 1 = column
// tile (j, kb)
__global__ void __launch_bounds__(kFwTile* kFwTile / 4)
    fw_phase2(double* __restrict__ d, int np, int kb) {
  __shared__ double diag[kFwTile][kFwTile + 1];
  __shared__ double t[kFwTile][kFwTile + 1];
  int j = blockIdx.x;
  if (j >= kb) {
    ++j;
  }
  const bool rowTile = blockIdx.y == 0;
  const int tx = threadIdx.x & 31;
  const int ty0 = threadIdx.x >> 5;
  const int base = kb * kFwTile;
  const int r0 = rowTile ? base : j * kFwTile;
  const int c0 = rowTile ? j * kFwTile : base;
  for (int r = ty0; r < kFwTile; r += 8) {
    diag[r][tx] = d[static_cast<size_t>(base + r) * np + base + tx];
    t[r][tx] = d[static_cast<size_t>(r0 + r) * np + c0 + tx];
  }
  __syncthreads();
  for (int k = 0; k < kFwTile; ++k) {
    for (int r = ty0; r < kFwTile; r += 8) {
      const double v = rowTile ? diag[r][k] + t[k][tx] : t[r][k] + diag[k][tx];
      if (v < t[r][tx]) {
        t[r][tx] = v;
      }
    }
    __syncthreads();
  }
  for (int r = ty0; r < kFwTile; r += 8) {
    d[static_cast<size_t>(r0 + r) * np + c0 + tx] = t[r][tx];
  }
}

// all tiles (i, j) with i != kb and j != kb
__global__ void __launch_bounds__(kFwTile* kFwTile / 4)
    fw_phase3(double* __restrict__ d, int np, int kb) {
  __shared__ double colT[kFwTile][kFwTile + 1];  // (i, kb)
  __shared__ double rowT[kFwTile][kFwTile + 1];  // (kb, j)
  int i = blockIdx.y;
  int j = blockIdx.x;
  if (i >= kb) {
    ++i;
  }
  if (j >= kb) {
    ++j;
  }
  const int tx = threadIdx.x & 31;
  const int ty0 = threadIdx.x >> 5;
  const int base = kb * kFwTile;
  for (int r = ty0; r < kFwTile; r += 8) {
    colT[r][tx] = d[static_cast<size_t>(i * kFwTile + r) * np + base + tx];
    rowT[r][tx] = d[static_cast<size_t>(base + r) * np + j * kFwTile + tx];
  }
  __syncthreads();
  double acc[4];
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    acc[q] = d[static_cast<size_t>(i * kFwTile + ty0 + 8 * q) * np +
               j * kFwTile + tx];
  }
#pragma unroll 8
  for (int k = 0; k < kFwTile; ++k) {
    const double b = rowT[k][tx];
#pragma unroll
    for (int q = 0; q < 4; ++q) {
      acc[q] = dmin2(acc[q], colT[ty0 + 8 * q][k] + b);
    }
  }
#pragma unroll
  for (int q = 0; q < 4; ++q) {
    d[static_cast<size_t>(i * kFwTile + ty0 + 8 * q) * np + j * kFwTile + tx] =
        acc[q];
  }
}

// np x np padded matrix from the n x n weights: w where the pair is adjacent in
// the undirected view of the coupling map (utils.cpp:67-73), 0 on the
// diagonal, +inf elsewhere
__global__ void fw_init(double* __restrict__ d, int np, int n,
                        const double* __restrict__ w,
                        const uint8_t* __restrict__ adj) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<size_t>(np) * np) {
    return;
  }
  const int r = static_cast<int>(idx / np);
  const int c = static_cast<int>(idx % np);
  double v = __longlong_as_double(0x7ff0000000000000LL);
  if (r == c) {
    v = 0.;
  } else if (r < n && c < n && adj[static_cast<size_t>(r) * n + c] != 0) {
    v = w[static_cast<size_t>(r) * n + c];
  }
  d[idx] = v;
}
// n x n result: unreachable (>= DBL_MAX) -> -1 (utils.cpp:32,50)
__global__ void fw_finish(const double* __restrict__ d, int np, int n,
                          double* __restrict__ out) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<size_t>(n) * n) {
    return;
  }
  const int r = static_cast<int>(idx / n);
  const int c = static_cast<int>(idx % n);
  const double v = d[static_cast<size_t>(r) * np + c];
  out[idx] = (r == c) ? 0. : (v >= DBL_MAX ? -1. : v);
}

// K1': one thread per source qubit runs the reference's Dijkstra literally
// (dijkstra_core.hpp); out row = the labels, scratch = visited flags and heap
__global__ void __launch_bounds__(32)
    dijkstra_reference_kernel(int n, const double* __restrict__ w,
                              const uint16_t* __restrict__ edges,
                              const uint32_t* __restrict__ incOff,
                              const uint32_t* __restrict__ inc,
                              double* __restrict__ out,
                              uint8_t* __restrict__ visited,
                              uint16_t* __restrict__ heap, int heapCap,
                              int* __restrict__ overflow) {
  const int src = blockIdx.x * blockDim.x + threadIdx.x;
  if (src >= n) {
    return;
  }
  double* cost = out + static_cast<size_t>(src) * n;
  if (!dijkstraReference(n, src, w, edges, incOff, inc, cost,
                         visited + static_cast<size_t>(src) * n,
                         heap + static_cast<size_t>(src) * heapCap, heapCap)) {
    *overflow = 1;
  }
}

// K2: one CTA per source row q1; threads over q2.  Computes
//   up[q1][q2] = min_e min(D[q1][e1] + D[e2][q2], D[q1][e2] + D[e1][q2] + rev)
// for every ordered pair; the host-side launcher mirrors the upper triangle
// when rev == 0 (utils.cpp:166-169) with skip_mirror_kernel.
__global__ void __launch_bounds__(256)
    edge_skip1_kernel(int n, const double* __restrict__ D,
                      const uint16_t* __restrict__ edges, int nE, double rev,
                      double* __restrict__ out) {
  extern __shared__ double row[];  // D[q1][*]
  const int q1 = blockIdx.x;
  for (int v = threadIdx.x; v < n; v += blockDim.x) {
    row[v] = D[static_cast<size_t>(q1) * n + v];
  }
  __syncthreads();
  for (int q2 = threadIdx.x; q2 < n; q2 += blockDim.x) {
    double best = DBL_MAX;
    if (q2 == q1) {
      best = 0.;
    } else {
      for (int e = 0; e < nE; ++e) {
        const int e1 = edges[2 * e];
        const int e2 = edges[2 * e + 1];
        best = dmin2(best, row[e1] + D[static_cast<size_t>(e2) * n + q2]);
        best = dmin2(best, row[e2] + D[static_cast<size_t>(e1) * n + q2] + rev);
      }
    }
    out[static_cast<size_t>(q1) * n + q2] = best;
  }
}
__global__ void skip_mirror_kernel(int n, double* __restrict__ t) {
  const size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
  if (idx >= static_cast<size_t>(n) * n) {
    return;
  }
  const int r = static_cast<int>(idx / n);
  const int c = static_cast<int>(idx % n);
  if (r > c) {
    t[idx] = t[static_cast<size_t>(c) * n + r];
  }
}

// K2': level k of buildEdgeSkipTable.  tables = k matrices T_0..T_{k-1}
// (n*n each, consecutive); writes the upper triangle of T_k, mirrors are done
// by skip_mirror_kernel.  notZero is set when any entry is > 0 (:128-130).
__global__ void __launch_bounds__(256)
    edge_skipk_kernel(int n, int k, const double* __restrict__ tables,
                      const uint16_t* __restrict__ edges, int nE,
                      double* __restrict__ out, int* __restrict__ notZero) {
  const int q1 = blockIdx.x;
  const size_t nn = static_cast<size_t>(n) * n;
  bool nz = false;
  for (int q2 = threadIdx.x; q2 < n; q2 += blockDim.x) {
    double best = DBL_MAX;
    if (q2 == q1) {
      best = 0.;
    } else if (q2 > q1) {
      for (int e = 0; e < nE; ++e) {
        const int e1 = edges[2 * e];
        const int e2 = edges[2 * e + 1];
        for (int l = 0; l < k; ++l) {
          const double* Tl = tables + static_cast<size_t>(l) * nn;
          const double* Tr = tables + static_cast<size_t>(k - l - 1) * nn;
          best = dmin2(best, Tl[static_cast<size_t>(q1) * n + e1] +
                                 Tr[static_cast<size_t>(e2) * n + q2]);
          best = dmin2(best, Tl[static_cast<size_t>(q1) * n + e2] +
                                 Tr[static_cast<size_t>(e1) * n + q2]);
        }
      }
      nz = nz || best > 0.;
    }
    out[static_cast<size_t>(q1) * n + q2] = best;
  }
  if (nz) {
    *notZero = 1;
  }
}

}  // namespace qmg
