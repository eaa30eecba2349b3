// route_h.cu - the instantiation of route_kernel<QMG_H> (route.cuh) and its
// launcher; compiled once per heuristic (-DQMG_H=0..4).
#include "route.cuh"
#include "route_launch.hpp"

#ifndef QMG_H
#error "compile with -DQMG_H=<heuristic 0..4>"
#endif

#define QMG_CAT2(a, b) a##b
#define QMG_CAT(a, b) QMG_CAT2(a, b)

namespace qmg {

cudaError_t QMG_CAT(routeSetSharedMemoryH, QMG_H)(int bytes) {
  return cudaFuncSetAttribute(route_kernel<QMG_H>,
                              cudaFuncAttributeMaxDynamicSharedMemorySize,
                              bytes);
}

cudaError_t QMG_CAT(routeLaunchH, QMG_H)(unsigned grid, size_t smemBytes,
                                         cudaStream_t s, const RouteArgs& a) {
  route_kernel<QMG_H><<<grid, kWarpsPerCta * 32, smemBytes, s>>>(
      a.A, a.C, a.inst, a.order, a.nInst, a.mem, a.ticket, a.lay, a.startFull);
  return cudaGetLastError();
}

}  // namespace qmg
